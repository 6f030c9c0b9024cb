mkdir -p gpurun_out/s9
OUT=s9 VARIANTS="base8 defer smext smboth" bash tools/gpu_ab.sh
PT_B200_LIB=$PWD/path-tracer_b200/ab/libpt_smboth.so timeout 900 python -m pytest tests -m gpu -x -q -k "image or partition or closest_hit_random or full_size" > gpurun_out/s9/pytest_smboth.log 2>&1; echo "pytest rc=$?" >> gpurun_out/s9/pytest_smboth.log
PT_B200_LIB=$PWD/path-tracer_b200/ab/libpt_defer.so timeout 900 python -m pytest tests -m gpu -x -q -k "image or partition or closest_hit_random or full_size" > gpurun_out/s9/pytest_defer.log 2>&1; echo "pytest rc=$?" >> gpurun_out/s9/pytest_defer.log
