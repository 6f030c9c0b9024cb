"""Freezes the canonical per-ray work of the bench workloads into profiles/canonical_counts.json.

Uses the oracle's canonical binned-SAH BVH2 counter (oracle/canonical.cpp) on a 1-spp trace of every `stride`-th
pixel of each workload.  Run on the CPU: python tools/canonical_counts.py [workload ...]
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__

pkg = __graft_entry__.load_package()
import bench
import oracle_lib

out_path = os.path.join(ROOT, "profiles", "canonical_counts.json")
res = json.load(open(out_path)) if os.path.exists(out_path) else {}
for name in (sys.argv[1:] or ["console_small", "console_1080p_128spp_5b"]):
    recipe, n_tris, seed, w, h, spp, bounces, tm = bench.WORKLOADS[name]
    t0 = time.time()
    sa = pkg.Scene.synthetic(recipe, n_tris, seed, device=-1).arrays()
    osc = oracle_lib.OracleScene(sa)
    stride = max(1, (w * h) // 200000)
    c = osc.canonical_counts(oracle_lib.make_profile(w, h, 1, bounces), stride, threads=os.cpu_count() or 8)
    c["pixel_stride"] = stride
    c["bytes_per_ray_closest"] = bench.ray_bytes(c["n_node_closest"], c["n_tri_closest"])
    c["bytes_per_ray_any"] = bench.ray_bytes(c["n_node_any"], c["n_tri_any"])
    c["flops_per_ray_closest"] = 30 * c["n_node_closest"] + 52 * c["n_tri_closest"]
    res[name] = c
    print(name, c, "%.1fs" % (time.time() - t0), flush=True)
json.dump(res, open(out_path, "w"), indent=1, sort_keys=True)
