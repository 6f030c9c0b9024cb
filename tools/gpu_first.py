"""First GPU contact: closest-hit parity + image parity on the test scenes, then a timing probe."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__

pkg = __graft_entry__.load_package()
import numpy as np

import isf_py
import oracle_lib
import rays as R

print("devices", pkg.device_count(), flush=True)
G = os.path.join(ROOT, "tests/golden/scenes")
res = {}
for name in sorted(os.listdir(G)):
    path = os.path.join(G, name, "scene.isf")
    sa = isf_py.load_isf(path)
    osc = oracle_lib.OracleScene(sa)
    gsc = pkg.Scene.load(path)
    rr = R.random_rays(sa, 1 << 18)
    t0 = time.time(); g = gsc.trace_closest(rr); t1 = time.time()
    o = osc.trace_closest(rr, threads=16)
    cmp = R.compare_hits(g, o)
    bounces = 0 if name == "white_furnace_direct" else 4
    prof = pkg.Profile(width=256, height=256, samples=16, bounces=bounces)
    rend = pkg.Renderer(profile=prof)
    img = rend.render(gsc, keep_accum=True)
    oacc = osc.render_accum(oracle_lib.make_profile(256, 256, 16, bounces), rng_mode=1, threads=16)
    gacc = rend.last_accum
    diff = np.abs(gacc - oacc)
    rel = float(np.sqrt(np.mean((gacc - oacc) ** 2)) / max(1e-12, np.mean(oacc)))
    close = float(np.mean(diff <= 1e-3 * np.maximum(np.abs(oacc), 1.0)))
    res[name] = dict(trace=cmp, trace_s=t1 - t0, rel_rmse=rel, frac_close=close, stats=rend.last_stats)
    print(name, json.dumps(res[name]), flush=True)
    pkg.write_png(os.path.join(ROOT, "gpurun_out", name + ".png"), img)

for recipe, ntri in (("console", 1000000),):
    t0 = time.time(); sc = pkg.Scene.synthetic(recipe, ntri, 3); t1 = time.time()
    print(recipe, sc.info(), "build_s", t1 - t0, flush=True)
    prof = pkg.Profile(width=1920, height=1080, samples=8, bounces=5)
    rend = pkg.Renderer(profile=prof)
    for it in range(2):
        img = rend.render(sc)
        st = rend.last_stats
        print(recipe, "Msamples/s", st["samples"] / st["ms_total"] / 1e3, "Mrays/s", (st["extend_rays"] + st["shadow_rays"]) / st["ms_total"] / 1e3, st, flush=True)
    pkg.write_png(os.path.join(ROOT, "gpurun_out", recipe + ".png"), img)
json.dump(res, open(os.path.join(ROOT, "gpurun_out", "first.json"), "w"), indent=1)
