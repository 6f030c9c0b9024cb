# full-size configs[3] / configs[4] on one GPU (the scaling origin of the spp split): every sample of the named profile
R=gpurun_out/r06
mkdir -p $R
timeout 400 python bench.py --workload foliage_1080p_256spp_4b --steps 2 --warmup 3 --cpu-seconds 8 > $R/bench_foliage_256spp.json 2> $R/bench_foliage.err
timeout 500 python bench.py --workload city_4k_1024spp_8b --steps 2 --warmup 3 --no-cpu-baseline > $R/bench_city_1024spp.json 2> $R/bench_city.err
cut -c1-300 $R/bench_foliage_256spp.json $R/bench_city_1024spp.json
