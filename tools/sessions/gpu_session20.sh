R=gpurun_out/r03
mkdir -p $R
nvidia-smi -L > $R/gpus_2.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline > $R/bench_console_2gpu.json 2> $R/bench_console_2gpu.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 1 --warmup 1 > $R/bench_reference_2gpu.json 2> $R/bench_reference_2gpu.err
timeout 300 python - > $R/lbvh_build_twice_b.json 2> $R/lbvh_build_twice_b.err <<'PY'
import json, sys
sys.path.insert(0, ".")
import __graft_entry__
pkg = __graft_entry__.load_package()
pkg.set_accel_builder("lbvh")
out = {}
for recipe, n, seed in (("city", 20000000, 5), ("console", 1000000, 3)):
    out[recipe] = []
    for i in range(3):
        sc = pkg.Scene.synthetic(recipe, n, seed)
        out[recipe].append(round(sc.accel_info()["build_ms"], 3))
        sc.close()
print(json.dumps(dict(lbvh_build_ms=out)))
PY
timeout 600 python -m pytest tests -m gpu -x -q -k "lbvh" > $R/pytest_lbvh.log 2>&1; echo "rc=$?" >> $R/pytest_lbvh.log
