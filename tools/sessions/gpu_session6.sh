mkdir -p gpurun_out/s6
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/s6/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/s6/pytest.log
OUT=s6 VARIANTS="base5 compact0" bash tools/gpu_ab.sh
