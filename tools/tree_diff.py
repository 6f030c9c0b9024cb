"""Tree-independence probe: python tools/tree_diff.py [workload] [n_pixels]
Traces the same seeded paths through trees from different builders / leaf sizes and reports every path whose
per-bounce record (prim_id, t) differs, with the Moller-Trumbore result of the ray against both candidate triangles
(pt_triangle_intersect), so that a difference can be attributed to a culled box (one tree never tested the nearer
triangle) rather than to the intersection arithmetic."""
import ctypes as C, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import __graft_entry__
pkg = __graft_entry__.load_package()
import bench

name = sys.argv[1] if len(sys.argv) > 1 else "city_4k_1024spp_8b"
n_pix = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 20
recipe, n_tris, seed, w, h, spp, bounces, tm = bench.WORKLOADS[name]
prof = pkg.Profile(width=w, height=h, samples=1, bounces=bounces, tonemap=tm)
rng = np.random.default_rng(7)
pixels = rng.choice(w * h, size=min(n_pix, w * h), replace=False).astype(np.uint32)

TREES = [("sah", 2), ("sah", 8), ("lbvh", 2), ("lbvh", 4), ("sah", 1)]
if os.environ.get("TREES"):
    TREES = [(t.split(":")[0], int(t.split(":")[1])) for t in os.environ["TREES"].split(",")]


def tri_table(sc):
    """prim_id -> triangle index (or -1 for a sphere) and the flat positions."""
    from importlib import import_module
    abi = import_module(pkg.__name__ + ".abi")
    api = import_module(pkg.__name__ + ".api")
    d = abi.pt_scene_desc()
    api._check(api.lib().pt_scene_get_desc(sc.h, C.byref(d)))
    pos = np.ctypeslib.as_array(d.tri_pos, shape=(d.n_tris * 9,)).reshape(-1, 9)
    ids = []
    for i in range(d.n_models):
        m = d.models[i]
        if m.type == 0:
            ids.append(np.arange(m.first_tri, m.first_tri + m.n_tris, dtype=np.int64))
        else:
            ids.append(np.array([-1], np.int64))
    return np.concatenate(ids), pos


ref = None
for accel, leaf in TREES:
    os.environ["PT_B200_LEAF_TRIS"] = str(leaf)
    pkg.set_accel_builder(accel)
    sc = pkg.Scene.synthetic(recipe, n_tris, seed)
    verts, counts = sc.trace_paths(prof, pixels, sample=0, seed=1)
    tag = f"{accel}-leaf{leaf}"
    if ref is None:
        ref = (tag, verts, counts, sc)
        table, pos = tri_table(sc)
        print(json.dumps(dict(tree=tag, paths=len(pixels), vertices=int(counts.sum()), info=sc.accel_info())), flush=True)
        continue
    rtag, rv, rc, rsc = ref
    k = np.arange(verts.shape[1])[None, :]
    live = k < np.minimum(counts, rc)[:, None]
    neq = ((verts["prim_id"] != rv["prim_id"]) | (verts["t"].view(np.uint32) != rv["t"].view(np.uint32))) & live
    bad = np.nonzero(neq.any(axis=1) | (counts != rc))[0]
    print(json.dumps(dict(tree=tag, vs=rtag, differing_paths=int(len(bad)), info=sc.accel_info())), flush=True)
    for p in bad[:24]:
        b = int(np.argmax(neq[p])) if neq[p].any() else int(min(counts[p], rc[p]))
        a, r = verts[p, b], rv[p, b]
        ray = np.concatenate([r["ray_o"], r["ray_d"]]).astype(np.float32)
        same_ray = bool((a["ray_o"] == r["ray_o"]).all() and (a["ray_d"] == r["ray_d"]).all())
        row = dict(path=int(p), bounce=b, same_ray=same_ray, ref=dict(prim=int(r["prim_id"]), t=float(r["t"])),
                   got=dict(prim=int(a["prim_id"]), t=float(a["t"])), ray=[float(x).hex() for x in ray])
        for key, pid in (("mt_ref_prim", int(r["prim_id"])), ("mt_got_prim", int(a["prim_id"]))):
            if pid >= 0 and table[pid] >= 0:
                hit = pkg.triangle_intersect(ray[None, :], pos[table[pid]][None, :])[0]
                row[key] = dict(side=int(hit["prim_id"]), t=float(hit["t"]), u=float(hit["u"]), v=float(hit["v"]),
                                w=float(1.0 - hit["u"] - hit["v"]))
                tri = pos[table[pid]].reshape(3, 3)
                row[key]["extent"] = float(np.abs(tri.max(0) - tri.min(0)).max())
                row[key]["tri"] = [float(x).hex() for x in pos[table[pid]]]
        print(json.dumps(row), flush=True)
    sc.close()
