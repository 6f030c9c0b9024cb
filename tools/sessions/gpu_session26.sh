R=gpurun_out/r03
mkdir -p $R
timeout 600 python bench.py --workload foliage_1080p_256spp_4b --spp 8 --steps 1 --warmup 1 --no-cpu-baseline > $R/foliage_8spp_counts.json 2> $R/foliage_8spp_counts.err
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_extend|k_shadow' -s 10 -c 10 -o $R/foliage_full -f python bench.py --workload foliage_1080p_256spp_4b --spp 8 --steps 1 --warmup 1 --no-cpu-baseline > $R/ncu_foliage.log 2>&1
timeout 600 python bench.py --workload city_4k_1024spp_8b --spp 4 --steps 1 --warmup 1 --no-cpu-baseline > $R/city_4spp_counts.json 2> $R/city_4spp_counts.err
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k_extend|k_shadow' -s 18 -c 18 -o $R/city_full -f python bench.py --workload city_4k_1024spp_8b --spp 4 --steps 1 --warmup 1 --no-cpu-baseline > $R/ncu_city.log 2>&1
