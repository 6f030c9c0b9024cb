"""CPU probe behind DESIGN §2.1: how far outside a triangle's exact bounding box can a ray pass and still be accepted
by the f32 Moller-Trumbore test?  python tools/leak_probe.py [n_rays]
Traces edge-grazing rays (tests/rays.py) with the oracle's brute force, keeps the accepted (ray, triangle) pairs and
finds, per pair, the smallest pad e for which the ray meets box(triangle) grown by e (f64 slab test, bisection)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import conftest  # noqa: F401  (puts the package on the path)
import numpy as np
import oracle_lib
import rays as R

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 19
mode = sys.argv[2] if len(sys.argv) > 2 else "walls"
if mode == "walls":
    sa, targets, cam, rng = R.edge_grazing_scene(n_walls=1500)
else:  # triangles of every shape and orientation: needles with the sharp corner at v0 / v1 / v2, obtuse, tiny, large
    import path_tracer_b200 as pt
    rng = np.random.Generator(np.random.PCG64(5))
    unit = lambda k: (lambda v: v / np.linalg.norm(v, axis=1, keepdims=True))(rng.normal(size=(k, 3)))
    m = 3000
    v0 = rng.uniform(-50, 50, size=(m, 3))
    a = unit(m)
    b = np.cross(a, unit(m)); b /= np.linalg.norm(b, axis=1, keepdims=True)
    L1 = 10 ** rng.uniform(-1.3, 1.3, size=(m, 1))
    e1 = a * L1
    e2 = a * L1 * rng.uniform(-0.5, 1.5, size=(m, 1)) + b * 10 ** rng.uniform(-2, 1, size=(m, 1))
    tris = np.stack([v0, v0 + e1, v0 + e2], 1)
    perm = rng.integers(0, 3, size=m)  # which vertex is stored first
    tris = np.stack([tris[np.arange(m), (perm + k) % 3] for k in range(3)], 1)
    sa = pt.abi.SceneArrays()
    sa.tri_pos = tris.reshape(m, 9).astype(np.float32)
    sa.tri_nrm = np.tile(np.array([0, 0, 1] * 3, np.float32), (m, 1))
    sa.tri_uv = np.tile(np.array([0, 0, 1, 0, 0, 1], np.float32), (m, 1))
    sa.models.append(dict(type=0, first_tri=0, n_tris=m, center=[0, 0, 0], radius=0.0))
    sa.materials.append(dict(albedo=[1, 1, 1], emissive=[0, 0, 0], opacity=1.0, metalness=0.0, roughness=1.0, ior=1.0))
    cam = np.array([70.0, 45.0, 85.0])
    sa.camera[12:15] = cam.astype(np.float32)
    targets = []
    for t in tris:
        nrm = np.cross(t[1] - t[0], t[2] - t[0])
        for i, j, k in ((0, 1, 2), (1, 2, 0), (2, 0, 1)):
            out = np.cross(t[j] - t[i], nrm)
            out /= np.linalg.norm(out)
            if np.dot(out, t[k] - t[i]) > 0:
                out = -out
            targets.append((t[i], t[j], out))
rays = R.edge_grazing_rays(targets, cam, rng, n)
ref = oracle_lib.OracleScene(sa).trace_closest(rays, brute=True, threads=os.cpu_count())
hit = ref["prim_id"] >= 0
r = rays[hit].astype(np.float64)
tri = sa.tri_pos[ref["prim_id"][hit]].astype(np.float64).reshape(-1, 3, 3)
o, d = r[:, :3], r[:, 3:]
lo, hi = tri.min(1), tri.max(1)


def misses(e):
    with np.errstate(divide="ignore", invalid="ignore"):
        t0, t1 = (lo - e[:, None] - o) / d, (hi + e[:, None] - o) / d
    tn = np.nanmax(np.minimum(t0, t1), axis=1)
    tf = np.nanmin(np.maximum(t0, t1), axis=1)
    return tn > tf


need = np.zeros(len(r))
out = misses(need)
a, b = np.zeros(len(r)), np.full(len(r), 1.0)
for _ in range(60):
    m = 0.5 * (a + b)
    mm = misses(m)
    a, b = np.where(mm, m, a), np.where(mm, b, m)
need = np.where(out, b, 0.0)
e1, e2 = tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0]
l1, l2, l3 = (np.linalg.norm(x, axis=1) for x in (e1, e2, e2 - e1))
ratio = np.clip((l1 + l2) / l3, 1, 4096)
allp = sa.tri_pos.reshape(-1, 3).astype(np.float64)
slo, shi = allp.min(0), allp.max(0)
v0 = tri[:, 0]
far = np.maximum(np.abs(v0 - slo), np.abs(v0 - shi))
D = np.sqrt(np.maximum((far ** 2).sum(1), ((v0 - cam) ** 2).sum(1)))  # scene_host.hpp origin_reach(v0)
eps = 2.0 ** -24
tv = np.linalg.norm(o - v0, axis=1)
given = 2.4e-7 * D * ratio
print(f"accepted pairs {len(r)}, of which the ray misses the exact box: {int(out.sum())}")
if not out.any():
    print("  (axis-aligned flat geometry is the case that matters: an oblique triangle's box is fat in every direction)")
    sys.exit(0)
for name, unit in (("eps*reach(v0)*ratio", eps * D * ratio), ("eps*|o-v0|*ratio", eps * tv * ratio)):
    x = (need / unit)[out]
    print(f"  pad needed in units of {name:20s}: max {x.max():8.3f}  p99.9 {np.percentile(x, 99.9):8.3f}  p99 {np.percentile(x, 99):8.3f}")
print(f"  absolute: max needed {need.max():.3e}; smallest pad given (4 eps * reach * ratio) {given.min():.3e}")
print(f"  largest needed / given {np.max(need / given):.3f}; pairs the pad would lose: {int((need > given).sum())}")
