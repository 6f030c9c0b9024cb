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
