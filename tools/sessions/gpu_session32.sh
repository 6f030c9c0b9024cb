R=gpurun_out/r04d
mkdir -p $R
timeout 1500 python -m pytest tests -m gpu -x -q > $R/pytest.log 2>&1; echo "pytest rc=$?" >> $R/pytest.log
timeout 900 python bench.py > $R/bench_console_1080p_128spp.json 2> $R/bench_console.err
timeout 900 python bench.py --workload foliage_1080p_256spp_4b --spp 64 --steps 2 --warmup 3 --cpu-seconds 8 > $R/bench_foliage_64spp.json 2> $R/bench_foliage.err
timeout 1200 python bench.py --workload city_4k_1024spp_8b --spp 16 --steps 3 --warmup 3 --no-cpu-baseline > $R/bench_city_16spp.json 2> $R/bench_city.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $R/launches_console_8spp.csv python bench.py --spp 8 --steps 1 --warmup 1 --no-cpu-baseline > $R/ncu_launches.log 2>&1
nvidia-smi --query-gpu=memory.used,memory.total --format=csv > $R/mem.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_extend|k_shade|k_shadow' -s 18 -c 18 -o $R/console_full -f python bench.py --spp 8 --steps 1 --warmup 1 --no-cpu-baseline > $R/ncu_full.log 2>&1
