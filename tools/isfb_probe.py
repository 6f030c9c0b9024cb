"""Times the binary side-car on a full-size scene: python tools/isfb_probe.py [workload] [dir]
build (generator + host SAH + upload) vs save vs load (read + validate stored tree + upload), and checks that the
reloaded scene renders the same accumulation buffer."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__
pkg = __graft_entry__.load_package()
import numpy as np
import bench

name = sys.argv[1] if len(sys.argv) > 1 else "city_4k_1024spp_8b"
out_dir = sys.argv[2] if len(sys.argv) > 2 else "/tmp"
recipe, n_tris, seed, w, h, spp, bounces, tm = bench.WORKLOADS[name]
t0 = time.time()
a = pkg.Scene.synthetic(recipe, n_tris, seed)
t_build = time.time() - t0
path = os.path.join(out_dir, name + ".isfb")
t0 = time.time()
a.save_binary(path, with_accel=True)
t_save = time.time() - t0
t0 = time.time()
b = pkg.Scene.load(path)
t_load = time.time() - t0
prof = pkg.Profile(width=480, height=270, samples=2, bounces=bounces, tonemap=tm)
ra, rb = pkg.Renderer(profile=prof), pkg.Renderer(profile=prof)
ra.render(a, keep_accum=True)
rb.render(b, keep_accum=True)
print(json.dumps(dict(workload=name, n_tris=a.info()["n_tris"], n_nodes=a.info()["n_nodes"], file_gb=round(os.path.getsize(path) / 1e9, 3),
                      synth_plus_build_s=round(t_build, 2), sah_build_ms=round(a.accel_info()["build_ms"], 1),
                      save_s=round(t_save, 2), load_s=round(t_load, 2), reload_build_ms=b.accel_info()["build_ms"],
                      same_nodes=a.info()["n_nodes"] == b.info()["n_nodes"],
                      same_image=bool(np.allclose(ra.last_accum, rb.last_accum, rtol=1e-4, atol=1e-5)))))
os.remove(path)
