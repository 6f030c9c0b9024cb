"""Debug: per-vertex path records GPU vs oracle."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__
pkg = __graft_entry__.load_package()
import numpy as np
import isf_py, oracle_lib
G = os.path.join(ROOT, "tests/golden/scenes")
names = sys.argv[1:] or sorted(os.listdir(G))
for name in names:
    path = os.path.join(G, name, "scene.isf")
    sa = isf_py.load_isf(path)
    osc = oracle_lib.OracleScene(sa)
    gsc = pkg.Scene.load(path)
    b = 0 if name == "white_furnace_direct" else 4
    prof = pkg.Profile(width=128, height=128, samples=1, bounces=b)
    pix = np.arange(128 * 128, dtype=np.uint32)
    gv, gc = gsc.trace_paths(prof, pix)
    ov, oc = osc.trace_paths(prof.c, pix)
    same_cnt = gc == oc
    bit = np.array([gv[i].tobytes() == ov[i].tobytes() for i in range(len(pix))])
    print(name, "paths", len(pix), "count mismatch", int((~same_cnt).sum()), "bit-exact paths", int(bit.sum()), flush=True)
    shown = 0
    for i in np.nonzero(~bit)[0]:
        for k in range(max(gc[i], oc[i])):
            if gv[i, k].tobytes() != ov[i, k].tobytes():
                if shown < 6:
                    print("  px", i % 128, i // 128, "first diff at bounce", k, "counts", gc[i], oc[i])
                    for f in gv.dtype.names:
                        if np.asarray(gv[i, k][f]).tobytes() != np.asarray(ov[i, k][f]).tobytes():
                            print("     ", f, "gpu", gv[i, k][f], "oracle", ov[i, k][f])
                    shown += 1
                break
    # wavefront image vs path records: final radiance
    rend = pkg.Renderer(profile=prof)
    rend.render(gsc, keep_accum=True)
    fin = np.array([gv[i, gc[i] - 1]["radiance"] for i in range(len(pix))]).reshape(128, 128, 3)
    d = np.abs(fin - rend.last_accum)
    print("   wavefront vs record kernel: max abs diff", float(d.max()), "rel", float(d.max() / max(1e-9, np.abs(fin).max())))
