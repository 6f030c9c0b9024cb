#!/usr/bin/env python
"""Oracle-vs-oracle noise floor of the converged-image comparison (tests/test_gpu_parity.py::
test_image_converged_4096spp): rel-RMSE between two 48x48 / 4096 spp renders of the CPU oracle with different Philox
seeds, per reference scene.  CPU only.  Output: profiles/noise_floor_4096spp.json (quoted in DESIGN.md)."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__  # noqa: E402

__graft_entry__.load_package()
import isf_py  # noqa: E402
import oracle_lib  # noqa: E402

SCENES = ["alpha_transparency", "cube", "head", "reflection", "spheres", "white_furnace_direct", "white_furnace_indirect"]


def rel_rmse(a, b):
    return float(np.sqrt(np.mean((a.astype(np.float64) - b) ** 2)) / max(1e-12, np.mean(b)))


def main():
    th = os.cpu_count() or 8
    out = {}
    for name in SCENES:
        sa = isf_py.load_isf(os.path.join(ROOT, "tests", "golden", "scenes", name, "scene.isf"))
        osc = oracle_lib.OracleScene(sa)
        prof = oracle_lib.make_profile(48, 48, 4096, 0 if name == "white_furnace_direct" else 4)
        a = osc.render_accum(prof, rng_mode=1, seed=2, threads=th) / 4096.0
        b = osc.render_accum(prof, rng_mode=1, seed=3, threads=th) / 4096.0
        clip = lambda x: np.minimum(x, 50.0)
        out[name] = {"floor_clip50": rel_rmse(clip(b), clip(a)), "floor_raw": rel_rmse(b, a), "mean": float(a.mean())}
        print(name, out[name], flush=True)
    with open(os.path.join(ROOT, "profiles", "noise_floor_4096spp.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
