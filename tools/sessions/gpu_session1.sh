set -x
mkdir -p gpurun_out/s1
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/s1/gpu.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/s1/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/s1/pytest.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/s1/bench_console.json 2> gpurun_out/s1/bench_console.err
timeout 900 python bench.py --workload foliage_1080p_256spp_4b --spp 64 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/s1/bench_foliage.json 2> gpurun_out/s1/bench_foliage.err
timeout 1200 python bench.py --workload city_4k_1024spp_8b --spp 16 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/s1/bench_city.json 2> gpurun_out/s1/bench_city.err
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_shade|k_shadow' -s 12 -c 12 -o gpurun_out/s1/shade_shadow_console -f python bench.py --spp 8 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/s1/ncu1.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k_extend|k_shade|k_shadow' -s 0 -c 9 -o gpurun_out/s1/city_1spp -f python bench.py --workload city_4k_1024spp_8b --spp 2 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/s1/ncu2.log 2>&1
ls -la gpurun_out/s1
