"""Quick per-kernel timing probe: python tools/perf_probe.py [workload] [spp]   (env PT_B200_VARIANT selects A/B kernels)"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__
pkg = __graft_entry__.load_package()
import torch
import bench
name = sys.argv[1] if len(sys.argv) > 1 else "console_1080p_128spp_5b"
recipe, n_tris, seed, w, h, spp, bounces, tm = bench.WORKLOADS[name]
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 16
sc = pkg.Scene.synthetic(recipe, n_tris, seed)
print(json.dumps(dict(workload=name, accel=sc.accel_info(), nodes=sc.info()["n_nodes"])), flush=True)
prof = pkg.Profile(width=w, height=h, samples=spp, bounces=bounces, tonemap=tm)
acc = torch.zeros(h, w, 3, device="cuda")
for timing in (False, True):
    pkg.set_kernel_timing(timing)
    for it in range(3):
        acc.zero_()
        st = pkg.render_accum_device(sc, prof, acc.data_ptr())
    rays = st["extend_rays"] + st["shadow_rays"]
    print(json.dumps(dict(variant=os.environ.get("PT_B200_VARIANT", "0"), timing=timing, ms=round(st["ms_total"], 3),
                          msamples=round(st["samples"] / st["ms_total"] / 1e3, 1), mrays=round(rays / st["ms_total"] / 1e3, 1),
                          extend=round(st["ms_extend"], 3), shade=round(st["ms_shade"], 3), shadow=round(st["ms_shadow"], 3),
                          other=round(st["ms_other"], 3), checksum=float(acc.double().sum()))), flush=True)
