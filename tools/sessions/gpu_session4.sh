mkdir -p gpurun_out/s4
OUT=s4 VARIANTS="base3 stk64 hitq stream sh6 sh5 shade5 all" bash tools/gpu_ab.sh
