mkdir -p gpurun_out/s5
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/s5/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/s5/pytest.log
OUT=s5 VARIANTS="base4 compact0" bash tools/gpu_ab.sh
timeout 900 python bench.py > gpurun_out/s5/bench_console.json 2> gpurun_out/s5/bench_console.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/s5/launches_console_8spp.csv python bench.py --spp 8 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/s5/ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_extend|k_shade|k_shadow' -s 18 -c 9 -o gpurun_out/s5/console_full -f python bench.py --spp 8 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/s5/ncu_full.log 2>&1
