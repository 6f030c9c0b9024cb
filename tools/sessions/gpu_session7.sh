mkdir -p gpurun_out/s7
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/s7/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/s7/pytest.log
OUT=s7 VARIANTS="base6 sync" bash tools/gpu_ab.sh
