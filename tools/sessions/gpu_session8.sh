mkdir -p gpurun_out/s8
OUT=s8 VARIANTS="base7 extahead extahead6 pfl1" bash tools/gpu_ab.sh
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/s8/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/s8/pytest.log
