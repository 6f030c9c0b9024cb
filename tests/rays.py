"""Seeded ray batches for the closest-hit parity tests (scheme of the reference's fixture generator,
tests/moller_trumbore/gen_tests.py:86-94,117-118: origins and targets on a sphere around the scene)."""
import numpy as np


def scene_bounds(sa):
    pts = [sa.tri_pos.reshape(-1, 3)] if len(sa.tri_pos) else []
    for m in sa.models:
        if m["type"] == 1:
            c, r = np.asarray(m["center"], np.float32), np.float32(m["radius"])
            pts.append(np.stack([c - r, c + r]))
    p = np.concatenate(pts)
    lo, hi = p.min(0), p.max(0)
    c = 0.5 * (lo + hi)
    return c.astype(np.float32), np.float32(0.5 * np.linalg.norm(hi - lo))


def _unit(rng, n):
    v = rng.normal(size=(n, 3))
    return v / np.linalg.norm(v, axis=1, keepdims=True)


def random_rays(sa, n, seed=0xB2000001):
    """o = c + 1.5 R s1, d = normalize(c + 1.5 R s2 - o), s1/s2 uniform on the unit sphere; float32."""
    rng = np.random.Generator(np.random.PCG64(seed))
    c, R = scene_bounds(sa)
    o = c + 1.5 * R * _unit(rng, n)
    t = c + 1.5 * R * _unit(rng, n) * rng.uniform(0, 1, size=(n, 1)) ** (1 / 3)
    d = t - o
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return np.concatenate([o, d], 1).astype(np.float32)


def compare_hits(gpu, ref):
    """Returns dict(n, id_mismatch, max_rel_t, max_abs_uv); t/u/v compared where ids agree and hit."""
    same = gpu["prim_id"] == ref["prim_id"]
    hit = same & (ref["prim_id"] >= 0)
    out = dict(n=len(ref), id_mismatch=int((~same).sum()), hits=int((ref["prim_id"] >= 0).sum()))
    if hit.any():
        out["max_rel_t"] = float(np.max(np.abs(gpu["t"][hit] - ref["t"][hit]) / np.abs(ref["t"][hit])))
        out["max_abs_uv"] = float(max(np.max(np.abs(gpu["u"][hit] - ref["u"][hit])), np.max(np.abs(gpu["v"][hit] - ref["v"][hit]))))
        out["bitexact_tuv"] = bool(np.array_equal(gpu["t"][hit], ref["t"][hit]) and np.array_equal(gpu["u"][hit], ref["u"][hit])
                                   and np.array_equal(gpu["v"][hit], ref["v"][hit]))
    return out


# ---- box margins: rays that graze triangle edges (test_gpu_parity.py, test_oracle_kat.py)
def edge_grazing_scene(n_walls=1500, seed=21):
    """Thin tall wall quads (1 x 16, the sharp corner of each triangle at v0) scattered through a 100-unit world and
    seen from ~150 units away, with rays aimed within a few 1e-4 of their edges.  At that distance the f32
    Moller-Trumbore test (triangle.rs:37-82) accepts rays that pass up to ~1e-4 OUTSIDE a thin triangle's far edge, so
    a tree whose leaf boxes hug the exact triangle drops hits that brute force finds (seen once in 3 M city paths)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    pos, targets = [], []
    for i in range(n_walls):
        base = np.array([rng.uniform(-50, 50), -2.0, rng.uniform(-50, 50)])
        if i % 2 == 0:  # axis-aligned wall: its box is flat
            w = np.array([1.0, 0, 0]) if i % 4 == 0 else np.array([0, 0, 1.0])
        else:
            a = rng.uniform(0, 2 * np.pi)
            w = np.array([np.cos(a), 0, np.sin(a)])
        h = np.array([0, rng.uniform(8, 16), 0])
        p0, p1, p2, p3 = base, base + h, base + w + h, base + w
        pos.append(np.concatenate([p0, p1, p2]))  # e1 = h, e2 = w + h: sharp corner at v0
        pos.append(np.concatenate([p0, p2, p3]))
        wn, hn = w / np.linalg.norm(w), h / np.linalg.norm(h)
        dn = np.cross(np.cross(w, h), w + h)
        dn /= np.linalg.norm(dn)
        for (a, b, out) in ((p1, p2, hn), (p0, p2, dn), (p0, p1, -wn), (p3, p2, wn)):  # top, diagonal, both sides
            targets.append((a, b, out))
    import path_tracer_b200 as pt
    sa = pt.abi.SceneArrays()
    sa.tri_pos = np.asarray(pos, np.float64).astype(np.float32)
    n = len(sa.tri_pos)
    sa.tri_nrm = np.tile(np.array([0, 0, 1] * 3, np.float32), (n, 1))
    sa.tri_uv = np.tile(np.array([0, 0, 1, 0, 0, 1], np.float32), (n, 1))
    sa.models.append(dict(type=0, first_tri=0, n_tris=n, center=[0, 0, 0], radius=0.0))
    sa.materials.append(dict(albedo=[1, 1, 1], emissive=[0, 0, 0], opacity=1.0, metalness=0.0, roughness=1.0, ior=1.0))
    cam = np.array([70.0, 45.0, 85.0])
    sa.camera[12:15] = cam.astype(np.float32)
    return sa, targets, cam, rng


def edge_grazing_rays(targets, cam, rng, n):
    pick = rng.integers(0, len(targets), size=n)
    s = rng.uniform(0.02, 0.98, size=n)
    off = rng.uniform(-3e-4, 3e-4, size=n)
    o = np.empty((n, 3))
    o[: n // 2] = cam  # primary-like
    o[n // 2:] = rng.uniform(-60, 60, size=(n - n // 2, 3)) * np.array([1, 0.3, 1]) + np.array([0, 20, 0])
    t = np.empty((n, 3))
    for k in range(n):
        a, b, out = targets[pick[k]]
        t[k] = a + (b - a) * s[k] + out * off[k]
    d = t - o
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return np.concatenate([o, d], 1).astype(np.float32)
