R=gpurun_out/r03
mkdir -p $R
nvidia-smi -L > $R/gpus_8.txt
for N in 8 4; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 3 --warmup 3 --no-cpu-baseline > $R/bench_console_${N}gpu.json 2> $R/bench_console_${N}gpu.err
done
