mkdir -p gpurun_out/s11
OUT=s11 VARIANTS="base9 shpf shpfl1" bash tools/gpu_ab.sh
