mkdir -p gpurun_out/r02
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02/pytest.log
timeout 900 python bench.py > gpurun_out/r02/bench_console_1080p_128spp.json 2> gpurun_out/r02/bench_console.err
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02/bench_reference_arm.json 2> gpurun_out/r02/bench_reference.err
timeout 900 python bench.py --workload foliage_1080p_256spp_4b --spp 64 --steps 2 --warmup 3 --cpu-seconds 8 > gpurun_out/r02/bench_foliage_64spp.json 2> gpurun_out/r02/bench_foliage.err
timeout 1200 python bench.py --workload city_4k_1024spp_8b --spp 16 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r02/bench_city_16spp.json 2> gpurun_out/r02/bench_city.err
PT_B200_ACCEL=lbvh timeout 1200 python bench.py --workload city_4k_1024spp_8b --spp 16 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r02/bench_city_16spp_lbvh.json 2> gpurun_out/r02/bench_city_lbvh.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02/launches_console_8spp.csv python bench.py --spp 8 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r02/ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_extend|k_shade|k_shadow' -s 18 -c 18 -o gpurun_out/r02/console_full -f python bench.py --spp 8 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r02/ncu_full.log 2>&1
