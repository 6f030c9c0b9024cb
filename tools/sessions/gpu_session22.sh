mkdir -p gpurun_out/s22
timeout 1200 python -m pytest tests -m gpu -x -q -k "textur or aov or debug or configs or image or head" > gpurun_out/s22/pytest.log 2>&1; echo "rc=$?" >> gpurun_out/s22/pytest.log
VARIANTS="notexmod texmod notexmod texmod" OUT=s22 bash tools/gpu_ab.sh
