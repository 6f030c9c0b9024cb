mkdir -p gpurun_out/s2
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/s2/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/s2/pytest.log
OUT=s2 VARIANTS="base smstack upop1 upop2 ext8 sh6 leaf8 refill28 upop2_ext8_sh6" bash tools/gpu_ab.sh
