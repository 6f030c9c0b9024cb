mkdir -p gpurun_out/s3
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/s3/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/s3/pytest.log
OUT=s3 VARIANTS="base2 stk64 hitq stream sh6 sh5 hitq_sh6 all" bash tools/gpu_ab.sh
PT_B200_ACCEL=lbvh timeout 900 python bench.py --workload city_4k_1024spp_8b --spp 8 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/s3/bench_city_lbvh.json 2> gpurun_out/s3/bench_city_lbvh.err
PT_B200_ACCEL=lbvh timeout 900 python bench.py --spp 32 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/s3/bench_console_lbvh.json 2> gpurun_out/s3/bench_console_lbvh.err
