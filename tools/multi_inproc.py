"""In-process multi-GPU render probe: python tools/multi_inproc.py [workload] [spp per GPU] [repeats]
Replicates the scene on every visible device (pt_scene_replicate), renders one frame of spp*G samples with
pt_render_multi and prints wall-clock and per-replica device times beside the one-device pt_render of spp samples
(weak scaling, like bench.py --gpus G) and of spp*G samples (the same frame on one device: image comparison)."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import __graft_entry__
pkg = __graft_entry__.load_package()
import bench
name = sys.argv[1] if len(sys.argv) > 1 else "console_1080p_128spp_5b"
recipe, n_tris, seed, w, h, spp, bounces, tm = bench.WORKLOADS[name]
spp = int(sys.argv[2]) if len(sys.argv) > 2 else spp
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
G = pkg.device_count()
t0 = time.perf_counter()
sc = pkg.Scene.synthetic(recipe, n_tris, seed)
t1 = time.perf_counter()
replicas = [sc] + [sc.replicate(d) for d in range(1, G)]
t2 = time.perf_counter()
one = pkg.Profile(width=w, height=h, samples=spp, bounces=bounces, tonemap=tm)
allp = pkg.Profile(width=w, height=h, samples=spp * G, bounces=bounces, tonemap=tm)
rend1, rendG = pkg.Renderer(profile=one), pkg.Renderer(profile=allp)
out = dict(workload=name, n_gpus=G, spp_per_gpu=spp, scene_s=round(t1 - t0, 3), replicate_s=round(t2 - t1, 3))
wall1, wallm, dev = [], [], []
for it in range(reps + 1):  # first iteration: warm-up (wave buffers are allocated lazily)
    a = time.perf_counter(); rend1.render(sc); b = time.perf_counter()
    rgb, _, stats = pkg.render_multi(replicas, allp); c = time.perf_counter()
    if it:
        wall1.append(b - a); wallm.append(c - b); dev.append([round(s["ms_total"], 2) for s in stats])
_, accum, _ = pkg.render_multi(replicas, allp, keep_accum=True)
ref = rendG.render(sc, keep_accum=True)
n1, nG = w * h * spp, w * h * spp * G
out.update(one_gpu_wall_ms=round(1e3 * min(wall1), 2), multi_wall_ms=round(1e3 * min(wallm), 2), replica_device_ms=dev[-1],
           one_gpu_msamples=round(n1 / min(wall1) / 1e6, 1), multi_msamples=round(nG / min(wallm) / 1e6, 1),
           speedup_weak=round((nG / min(wallm)) / (n1 / min(wall1)), 3),
           accum_rel_err_vs_one_device=float(np.abs(accum - rendG.last_accum).max() / max(1e-30, np.abs(rendG.last_accum).max())),
           rgb_max_level_diff=int(np.abs(rgb.astype(np.int16) - ref.astype(np.int16)).max()))
print(json.dumps(out))
