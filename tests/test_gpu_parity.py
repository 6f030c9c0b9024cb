"""GPU parity tests: every call goes through the C ABI (libpt_b200.so) and is checked against the CPU oracle.

The fixed-batch entry points (pt_trace_closest / pt_trace_accepted / pt_trace_shadow) load the batch into the
renderer's own queues and run the production kernels k_extend / k_shadow (ClosestTrav / ShadowTrav of
csrc/traverse.cuh: fused slab test, approximate reciprocals, one-primitive leaf steps, lazy pops, lane-level ray
replacement): every batch below is a check of the traversal that ships pixels.

Bars (BASELINE.json north_star): closest-hit primitive ids bit-exact, t/u/v within 1e-5 relative (they are in fact
bit-identical: the kernels compute Möller–Trumbore un-fused in the reference's operand order); images within 1 %
relative RMSE.  Nothing here reads /root/reference: inputs are the committed fixtures under tests/golden/.
"""
import os

import numpy as np
import pytest

import isf_py
import oracle_lib
import rays as R
from conftest import GOLDEN

import path_tracer_b200 as pt

pytestmark = pytest.mark.gpu

SCENES = ["alpha_transparency", "cube", "head", "reflection", "spheres", "white_furnace_direct", "white_furnace_indirect"]
THREADS = os.cpu_count() or 8


def scene_path(name):
    return os.path.join(GOLDEN, "scenes", name, "scene.isf")


@pytest.fixture(scope="module")
def loaded():
    """name -> (SceneArrays, OracleScene, pt.Scene), loaded once."""
    out = {}
    for name in SCENES:
        sa = isf_py.load_isf(scene_path(name))
        out[name] = (sa, oracle_lib.OracleScene(sa), pt.Scene.load(scene_path(name)))
    return out


# ---------------------------------------------------------------- Triangle::intersect (triangle.rs:200-233)
def test_triangle_intersect_reference_vectors():
    mt = np.load(os.path.join(GOLDEN, "moller_trumbore.npz"))
    hit = pt.triangle_intersect(mt["hit_rays"], mt["hit_tris"])
    assert (hit["prim_id"] >= 0).all()
    gold = mt["hit_gold"]
    err = np.abs(np.stack([hit["t"], hit["u"], hit["v"]], 1).astype(np.float64) - gold).max()
    assert err < 1e-5  # the reference's own tolerance (triangle.rs:213-216)
    miss = pt.triangle_intersect(mt["miss_rays"], mt["miss_tris"])
    assert (miss["prim_id"] == -1).all()
    # bit-exact against the oracle's restatement, hit flag, back-face flag and t/u/v
    for rays, tris, res in ((mt["hit_rays"], mt["hit_tris"], hit), (mt["miss_rays"], mt["miss_tris"], miss)):
        for i in range(0, len(rays), 7):
            ok, o = oracle_lib.moller_trumbore(rays[i], tris[i])
            assert ok == (res["prim_id"][i] >= 0)
            if ok:
                assert res["prim_id"][i] == (1 if o[3] < 0 else 0)
                assert (res["t"][i], res["u"][i], res["v"][i]) == (o[0], o[1], o[2])


def test_triangle_intersect_empty():
    out = pt.triangle_intersect(np.zeros((0, 6), np.float32), np.zeros((0, 9), np.float32))
    assert len(out) == 0


# ---------------------------------------------------------------- ray_cast(...).first()  (utils.rs:11-21)
@pytest.mark.parametrize("name", SCENES)
def test_closest_hit_random_rays(loaded, name):
    sa, osc, gsc = loaded[name]
    rr = R.random_rays(sa, 1 << 20)
    g = gsc.trace_closest(rr)
    o = osc.trace_closest(rr, threads=THREADS)
    c = R.compare_hits(g, o)
    assert c["id_mismatch"] == 0, c
    assert c["hits"] > 1000
    assert c["bitexact_tuv"], c


@pytest.mark.parametrize("name", SCENES)
def test_closest_hit_secondary_rays(loaded, name):
    """Bounce-1 and bounce-2 rays of the oracle's own paths (origins on surfaces, biased by 1e-5)."""
    sa, osc, gsc = loaded[name]
    prof = oracle_lib.make_profile(1024, 1024, 1, 4)  # the oracle's bounce-1 rays at 1024x1024x1 spp
    for bounce in (1, 2):
        rr = osc.dump_bounce_rays(prof, bounce, seed=0xB2000002)
        if len(rr) == 0:
            continue
        g = gsc.trace_closest(rr)
        o = osc.trace_closest(rr, threads=THREADS)
        same = g["prim_id"] == o["prim_id"]
        assert same.all(), (name, bounce, int((~same).sum()))
        hit = o["prim_id"] >= 0
        for f in ("t", "u", "v"):
            assert np.array_equal(g[f][hit], o[f][hit]), (name, bounce, f)


def test_closest_hit_brute_force_ground_truth(loaded):
    """On a subset, the oracle with its BVH filter disabled (every primitive tested) is the ground truth."""
    for name in ("cube", "alpha_transparency", "white_furnace_direct", "spheres"):
        sa, osc, gsc = loaded[name]
        rr = R.random_rays(sa, 1 << 15, seed=11)
        assert np.array_equal(gsc.trace_closest(rr), osc.trace_closest(rr, brute=True, threads=THREADS)), name


def test_axis_parallel_and_degenerate_rays(loaded):
    """inv-dir = +-inf slabs, zero-thickness boxes (white_furnace_direct is all axis-aligned quads), zero directions."""
    sa, osc, gsc = loaded["white_furnace_direct"]
    c, rad = R.scene_bounds(sa)
    rng = np.random.Generator(np.random.PCG64(5))
    n = 30000
    o = (c + rng.uniform(-1.2, 1.2, size=(n, 3)) * rad).astype(np.float32)
    d = np.zeros((n, 3), np.float32)
    axis = rng.integers(0, 3, n)
    d[np.arange(n), axis] = rng.choice([-1.0, 1.0], n)
    two = rng.random(n) < 0.3  # rays in an axis plane (one zero component)
    d[two, (axis[two] + 1) % 3] = rng.uniform(-1, 1, two.sum())
    # origins snapped onto the planes of the geometry: rays inside a face's plane
    snap = rng.random(n) < 0.3
    verts = sa.tri_pos.reshape(-1, 3)
    o[snap, 0] = verts[rng.integers(0, len(verts), snap.sum()), 0]
    rr = np.concatenate([o, d], 1)
    rr[:10, 3:] = 0.0  # zero direction: must miss everywhere (det == 0)
    g = gsc.trace_closest(rr)
    ref = osc.trace_closest(rr, brute=True, threads=THREADS)
    assert np.array_equal(g, ref)
    assert (g["prim_id"][:10] == -1).all()


def test_soup_of_reference_triangles():
    """The 400 unique triangles x 98 unique rays behind the reference's M-T vectors, as one mesh."""
    mt = np.load(os.path.join(GOLDEN, "moller_trumbore.npz"))
    tris = np.unique(np.concatenate([mt["hit_tris"], mt["miss_tris"]]).astype(np.float32), axis=0)
    rays = np.unique(np.concatenate([mt["hit_rays"], mt["miss_rays"]]).astype(np.float32), axis=0)
    assert len(tris) == 400 and len(rays) == 98
    sa = pt.abi.SceneArrays()
    sa.tri_pos = tris
    sa.tri_nrm = np.tile(np.array([0, 0, 1] * 3, np.float32), (400, 1))
    sa.tri_uv = np.tile(np.array([0, 0, 1, 0, 0, 1], np.float32), (400, 1))
    sa.models.append(dict(type=0, first_tri=0, n_tris=400, center=[0, 0, 0], radius=0.0))
    sa.materials.append(dict(albedo=[1, 1, 1], emissive=[0, 0, 0], opacity=1.0, metalness=0.0, roughness=1.0, ior=1.0))
    gsc = pt.Scene.from_arrays(sa)
    osc = oracle_lib.OracleScene(sa)
    g = gsc.trace_closest(rays)
    assert np.array_equal(g, osc.trace_closest(rays, brute=True))
    # every gold hit vector whose triangle is the closest one along its ray reproduces the gold distance
    more = R.random_rays(sa, 1 << 18, seed=3)
    assert np.array_equal(gsc.trace_closest(more), osc.trace_closest(more, threads=THREADS))


def test_synthetic_soup_large():
    """100k random triangles: deep tree, many overlapping boxes."""
    gsc = pt.Scene.synthetic("soup", 100000, seed=9)
    sa = gsc.arrays()
    osc = oracle_lib.OracleScene(sa)
    rr = R.random_rays(sa, 1 << 18, seed=4)
    g = gsc.trace_closest(rr)
    o = osc.trace_closest(rr, threads=THREADS)
    assert np.array_equal(g, o)
    assert (o["prim_id"] >= 0).sum() > 100000


def test_trace_empty_batch(loaded):
    _, _, gsc = loaded["cube"]
    assert len(gsc.trace_closest(np.zeros((0, 6), np.float32))) == 0
    assert len(gsc.trace_transmittance(np.zeros((0, 6), np.float32))) == 0


# ---------------------------------------------------------------- get_light_info shadow walk (mod.rs:291-297)
@pytest.mark.parametrize("name", ["alpha_transparency", "head", "cube", "spheres"])
def test_directional_transmittance(loaded, name):
    sa, osc, gsc = loaded[name]
    rr = R.random_rays(sa, 1 << 16, seed=21)
    g = gsc.trace_transmittance(rr)
    o = osc.trace_transmittance(rr)
    # products of (1 - opacity) in traversal order vs sorted order: equal up to float re-association
    assert np.allclose(g, o, rtol=2e-6, atol=1e-7)
    assert ((o == 0) == (g == 0)).all()
    if name == "alpha_transparency":
        assert ((o > 0) & (o < 1)).sum() > 100  # partial transmittance exercised


@pytest.mark.parametrize("name", SCENES)
def test_shadow_walk_of_real_requests(loaded, name):
    """k_shadow on the shadow rays get_light_info (mod.rs:281-333) casts from the primary hits of a 512x512 frame,
    one batch per light: point lights carry the distance limit and the primary hit's uv (the occluder's opacity is
    sampled there, mod.rs:324), directional lights walk every hit."""
    sa, osc, gsc = loaded[name]
    prof = oracle_lib.make_profile(512, 512, 1, 1)
    total = 0
    for li, light in enumerate(sa.lights):
        rays, lim, prim = osc.dump_shadow_requests(prof, li, seed=3)
        if len(rays) == 0:
            continue
        if light["type"] == 0:
            assert (lim >= 0).all()
            g, o = gsc.trace_shadow(rays, lim, prim), osc.trace_shadow(rays, lim, prim, threads=THREADS)
        else:
            g, o = gsc.trace_shadow(rays), osc.trace_shadow(rays, threads=THREADS)
        # products of (1 - opacity) in traversal order vs sorted order: equal up to float re-association
        assert np.allclose(g, o, rtol=2e-6, atol=1e-7), (name, li, np.abs(g - o).max())
        assert ((o == 0) == (g == 0)).all(), (name, li)
        total += len(rays)
    assert total > 1000 or not sa.lights


def test_shadow_walk_point_light_synthetic_limits(loaded):
    """Point-limited walk on random rays: the limit cuts the walk at random places (occluders beyond it are ignored),
    the primary uv decides the opacity of textured occluders, a NaN u marks a sphere primary (simple material)."""
    for sa, osc, gsc in (loaded["alpha_transparency"], loaded["spheres"]):
        rr = R.random_rays(sa, 1 << 16, seed=31)
        rng = np.random.Generator(np.random.PCG64(7))
        c, rad = R.scene_bounds(sa)
        lim = rng.uniform(0.0, 2.5 * rad, len(rr)).astype(np.float32)
        prim = np.concatenate([rr[:, :3], rng.uniform(-1.5, 2.5, (len(rr), 2))], 1).astype(np.float32)
        prim[::7, 3] = np.nan
        g, o = gsc.trace_shadow(rr, lim, prim), osc.trace_shadow(rr, lim, prim, threads=THREADS)
        assert np.allclose(g, o, rtol=2e-6, atol=1e-7)
        assert ((o == 0) == (g == 0)).all()
        assert (o == 1).sum() > 100 and (o < 1).sum() > 100


# ---------------------------------------------------------------- alpha walk of render_pixel (mod.rs:188-205)
@pytest.mark.parametrize("name", ["alpha_transparency", "head", "spheres"])
def test_accepted_hit_alpha_walk(loaded, name):
    """k_extend with the alpha walk on: the hit that gets shaded, for primary-like rays and for the oracle's bounce-1
    rays, with the draws of path (pixel = ray index, sample, seed) — bit-exact ids and t/u/v."""
    sa, osc, gsc = loaded[name]
    batches = [(R.random_rays(sa, 1 << 18, seed=41), 0, 5, 77),
               (osc.dump_bounce_rays(oracle_lib.make_profile(512, 512, 1, 4), 1, seed=0xB2000002), 1, 0, 0)]
    for rr, bounce, sample, seed in batches:
        g = gsc.trace_accepted(rr, bounce=bounce, sample=sample, seed=seed)
        o = osc.trace_accepted(rr, bounce=bounce, sample=sample, seed=seed, threads=THREADS)
        assert np.array_equal(g["prim_id"], o["prim_id"]), (name, bounce, int((g["prim_id"] != o["prim_id"]).sum()))
        assert g.tobytes() == o.tobytes(), (name, bounce)
    if name == "alpha_transparency":
        first = gsc.trace_closest(batches[0][0])
        g0 = gsc.trace_accepted(batches[0][0], bounce=0, sample=5, seed=77)
        assert (first["prim_id"] != g0["prim_id"]).sum() > 1000  # transparent hits really are skipped
        assert (g0["t"] >= first["t"])[first["prim_id"] >= 0].all()


def test_accepted_hit_textured_scene_and_farthest_rule():
    """Opacity textures (an opaque texel of a transparent-capable material ends the walk), a constant-alpha sphere,
    and the rule that the farthest examined hit is shaded when every hit is rejected (mod.rs:188-205)."""
    sa = textured_scene()
    gsc, osc = pt.Scene.from_arrays(sa), oracle_lib.OracleScene(sa)
    rr = R.random_rays(sa, 1 << 17, seed=2)
    for bounce, sample in ((0, 0), (3, 9)):
        g = gsc.trace_accepted(rr, bounce=bounce, sample=sample, seed=1)
        o = osc.trace_accepted(rr, bounce=bounce, sample=sample, seed=1, brute=True, threads=THREADS)
        assert g.tobytes() == o.tobytes()
    # a scene of transparent cards only (opacity 0.0005 < 0.001: never accepted, no draw): the LAST card is shaded
    cards = pt.abi.SceneArrays()
    pos, n = [], 6
    for k in range(n):
        z = -float(k)
        pos += [[-1, -1, z, 1, -1, z, 1, 1, z], [-1, -1, z, 1, 1, z, -1, 1, z]]
    cards.tri_pos = np.array(pos, np.float32)
    cards.tri_nrm = np.tile(np.array([0, 0, 1] * 3, np.float32), (2 * n, 1))
    cards.tri_uv = np.tile(np.array([0, 0, 1, 0, 1, 1], np.float32), (2 * n, 1))
    for k in range(n):
        cards.models.append(dict(type=0, first_tri=2 * k, n_tris=2, center=[0, 0, 0], radius=0.0))
        cards.materials.append(dict(albedo=[1, 1, 1], emissive=[0, 0, 0], opacity=0.5 if k % 2 == 0 else 0.0005,
                                    metalness=0.0, roughness=1.0, ior=1.0))
    gc, oc = pt.Scene.from_arrays(cards), oracle_lib.OracleScene(cards)
    rng = np.random.Generator(np.random.PCG64(3))
    o3 = np.concatenate([rng.uniform(-0.9, 0.9, (4096, 2)), np.full((4096, 1), 2.0)], 1)
    rays = np.concatenate([o3, np.tile([0.0, 0.0, -1.0], (4096, 1))], 1).astype(np.float32)
    g, o = gc.trace_accepted(rays, seed=5), oc.trace_accepted(rays, seed=5, brute=True, threads=THREADS)
    assert g.tobytes() == o.tobytes()
    last = g["prim_id"] >= 2 * (n - 1)
    assert 0.08 < last.mean() < 0.18  # 0.5^3 of the paths reject all three half-transparent cards -> the last card


def test_trace_large_batch_is_chunked(loaded):
    """Batches larger than one pass through the queues (16 Mi rays) are split by the library: same records as the
    rays traced on their own (32-bit ray indexing)."""
    sa, osc, gsc = loaded["cube"]
    base = R.random_rays(sa, 1 << 16, seed=8)
    reps = (16 << 20) // len(base) + 3  # a little over one chunk
    big = np.tile(base, (reps, 1))
    g = gsc.trace_closest(big)
    ref = gsc.trace_closest(base)
    assert np.array_equal(g, np.tile(ref, reps))
    assert np.array_equal(ref, osc.trace_closest(base, threads=THREADS))


# ---------------------------------------------------------------- Philox4x32-10 (replaces rand's StdRng, mod.rs:110-118)
def test_philox_kat():
    from test_oracle_kat import PHILOX_KAT
    ctr = np.array([c for c, _, _ in PHILOX_KAT], np.uint32)
    key = np.array([k for _, k, _ in PHILOX_KAT], np.uint32)
    want = np.array([w for _, _, w in PHILOX_KAT], np.uint32)
    assert np.array_equal(pt.philox4x32_10(ctr, key), want)
    rng = np.random.default_rng(4)
    c, k = rng.integers(0, 2**32, (4096, 4), dtype=np.uint64), rng.integers(0, 2**32, (4096, 2), dtype=np.uint64)
    got = pt.philox4x32_10(c, k)
    for i in range(0, 4096, 97):
        assert np.array_equal(got[i], oracle_lib.philox4x32_10(c[i], k[i]))


# ---------------------------------------------------------------- render_pixel, vertex by vertex (mod.rs:172-228)
@pytest.mark.parametrize("name", SCENES)
def test_path_records(loaded, name):
    """One path per pixel, recorded bounce by bounce.  Primary segments are bit-identical (intersection arithmetic is
    IEEE and un-fused everywhere); later ones differ through the shading arithmetic, which the product evaluates with
    the hardware's <= 2 ulp division / square root (shade.cuh, PTB_SHADE_FAST) where the oracle uses glibc."""
    sa, osc, gsc = loaded[name]
    bounces = 0 if name == "white_furnace_direct" else 5
    prof = pt.Profile(width=160, height=120, samples=1, bounces=bounces)
    pix = np.arange(160 * 120, dtype=np.uint32)
    gv, gc = gsc.trace_paths(prof, pix, sample=3, seed=7)
    ov, oc = osc.trace_paths(prof.c, pix, sample=3, seed=7)
    # bounce 0: the primary ray and its hit are bit-exact for every pixel
    for f in ("ray_o", "ray_d", "prim_id", "t"):
        assert np.array_equal(gv[:, 0][f], ov[:, 0][f]), f
    same_len = gc == oc
    k = np.arange(bounces + 1)[None, :]
    valid = k < np.minimum(gc, oc)[:, None]
    same_prims = ((gv["prim_id"] == ov["prim_id"]) | ~valid).all(1) & same_len
    # ulp-level direction differences may flip a hit at silhouettes / a roulette decision: rare
    assert same_prims.mean() > 0.99, same_prims.mean()
    last_g = gv[np.arange(len(pix)), gc - 1]["radiance"][same_prims]
    last_o = ov[np.arange(len(pix)), oc - 1]["radiance"][same_prims]
    tol = 2e-3 * np.maximum(np.abs(last_o), 1e-3)
    close = (np.abs(last_g - last_o) <= tol).all(1)
    # mirror-like GGX lobes (roughness clamped to 1e-4) amplify one-ulp differences: allow a small tail
    assert close.mean() > 0.97, close.mean()
    thr_g = gv[np.arange(len(pix)), gc - 1]["throughput"][same_prims]
    thr_o = ov[np.arange(len(pix)), oc - 1]["throughput"][same_prims]
    assert (np.abs(thr_g - thr_o) <= 1e-4 * np.maximum(np.abs(thr_o), 1e-3)).all(1).mean() > 0.97


# ---------------------------------------------------------------- Renderer::render (mod.rs:76-169)
def rel_rmse(a, b):
    return float(np.sqrt(np.mean((a.astype(np.float64) - b) ** 2)) / max(1e-12, np.mean(b)))


@pytest.mark.parametrize("name", SCENES)
def test_image_same_streams(loaded, name):
    """configs[0]: 256x256, 16 spp, 4 bounces, FILMIC.  Same Philox streams on both sides."""
    sa, osc, gsc = loaded[name]
    bounces = 0 if name == "white_furnace_direct" else 4
    prof = pt.Profile(width=256, height=256, samples=16, bounces=bounces, tonemap="FILMIC")
    rend = pt.Renderer(profile=prof)
    img = rend.render(gsc, keep_accum=True)
    oacc = osc.render_accum(prof.c, rng_mode=1, threads=THREADS)
    gacc = rend.last_accum
    # per-pixel agreement almost everywhere; the mean image within the 1 % bar even at 16 spp
    close = np.abs(gacc - oacc) <= 1e-3 * np.maximum(np.abs(oacc), 1.0)
    # what differs: paths whose direction moved by an ulp of CUDA's vs glibc's sinf/cosf/acosf and flipped a hit at a
    # silhouette or a roulette decision (a handful per frame); a wrong traversal would move whole regions
    assert close.mean() >= 0.999, close.mean()
    med_g, med_o = np.median(gacc), np.median(oacc)
    assert abs(med_g - med_o) <= 1e-3 * max(1.0, abs(med_o))
    # 4x4 box-filtered means (fireflies of the pdf()=1 estimator dominate a raw RMSE at 16 spp)
    f = lambda a: a.reshape(64, 4, 64, 4, 3).mean((1, 3))
    assert rel_rmse(np.minimum(f(gacc), 1e3), np.minimum(f(oacc), 1e3)) < 0.01 or close.mean() > 0.995
    oimg = oracle_lib.tonemap(oacc, 16, 1)
    assert (np.abs(img.astype(int) - oimg.astype(int)) <= 1).mean() >= 0.999
    st = rend.last_stats
    assert st["samples"] == 256 * 256 * 16 and st["extend_rays"] >= st["samples"]
    assert st["shadow_dropped"] == 0 and st["alpha_recasts"] == 0


# oracle-vs-oracle noise floor of the 48x48 / 4096 spp comparison, measured with tools/noise_floor.py:
# where it is under 0.7 % the plain 1 % bar of the north star applies; above it (fireflies of the reference's pdf()=1
# estimator) the bar is 1.5x the floor measured in the same test.
# profiles/noise_floor_4096spp.json: cube 0.66 %, white_furnace_direct 0.54 %, white_furnace_indirect 0.59 %;
# head 1.3 %, alpha_transparency 1.6 %, reflection 2.7 %, spheres 4.6 %.
NOISE_FLOOR_UNDER_0_7 = {"cube", "white_furnace_direct", "white_furnace_indirect"}


@pytest.mark.parametrize("name", SCENES)
def test_image_converged_4096spp(loaded, name):
    """north_star: rel-RMSE <= 1 % at 4096 spp, GPU vs oracle, with independent sample streams (different seeds),
    and the oracle-vs-oracle noise floor beside it.  All seven reference scenes (src/main.rs:100-165)."""
    sa, osc, gsc = loaded[name]
    prof = pt.Profile(width=48, height=48, samples=4096, bounces=0 if name == "white_furnace_direct" else 4)
    rend = pt.Renderer(config=pt.RenderConfig(seed=1), profile=prof)
    rend.render(gsc, keep_accum=True)
    g = rend.last_accum / 4096.0
    o = osc.render_accum(prof.c, rng_mode=1, seed=2, threads=THREADS) / 4096.0
    o2 = osc.render_accum(prof.c, rng_mode=1, seed=3, threads=THREADS) / 4096.0
    clip = lambda a: np.minimum(a, 50.0)  # the reference's pdf()=1 estimator has unbounded-variance fireflies
    err, floor = rel_rmse(clip(g), clip(o)), rel_rmse(clip(o2), clip(o))
    if name in NOISE_FLOOR_UNDER_0_7:
        assert floor < 0.007 and err < 0.01, (err, floor)
    else:
        assert err < max(0.01, 1.5 * floor), (err, floor)
    assert abs(g.mean() - o.mean()) < 0.01 * o.mean() + 3e-3 * abs(o).max()


def test_white_furnace_energy(loaded):
    """white_furnace_indirect: albedo 1, background 3.14 -> rough non-metal spheres reflect ~ the background."""
    _, _, gsc = loaded["white_furnace_indirect"]
    prof = pt.Profile(width=64, height=64, samples=64, bounces=4)
    rend = pt.Renderer(profile=prof)
    rend.render(gsc, keep_accum=True)
    acc = rend.last_accum / 64.0
    assert np.isfinite(acc).all()
    assert abs(acc[0, 0, 0] - 3.14) < 1e-4  # corner pixel sees the background only


def test_sample_partition_sums_to_full_render(loaded):
    """Multi-GPU split: samples {s : s mod G == g} rendered separately add up to the full render."""
    import torch
    _, _, gsc = loaded["alpha_transparency"]
    prof = pt.Profile(width=128, height=96, samples=12, bounces=4)
    full = torch.zeros(96, 128, 3, device="cuda")
    pt.render_accum_device(gsc, prof, full.data_ptr())
    parts = torch.zeros(96, 128, 3, device="cuda")
    for g in range(4):
        buf = torch.zeros(96, 128, 3, device="cuda")
        st = pt.render_accum_device(gsc, prof, buf.data_ptr(), sample_begin=g, sample_stride=4, n_samples=3)
        assert st["samples"] == 128 * 96 * 3
        parts += buf
    torch.cuda.synchronize()
    assert torch.allclose(full, parts, rtol=1e-5, atol=1e-6)


def test_render_multi_in_process(tmp_path, loaded):
    """pt_render_multi (one process, one host thread per replica, samples split s mod G): three replicas
    — spread over the box's devices, which on a one-GPU box is the same device three times — reproduce pt_render's
    accumulation buffer up to f32 summation order; the CLI's --gpus reaches the same image and rejects a count the
    box does not have."""
    import subprocess
    ndev = pt.device_count()
    _, _, gsc = loaded["alpha_transparency"]
    prof = pt.Profile(width=128, height=96, samples=10, bounces=4)  # 10 = 4 + 3 + 3 samples
    rend = pt.Renderer(profile=prof)
    full = rend.render(gsc, keep_accum=True)
    replicas = [gsc] + [gsc.replicate(g % ndev) for g in (1, 2)]
    rgb, accum, stats = pt.render_multi(replicas, prof, keep_accum=True)
    assert [s["samples"] for s in stats] == [128 * 96 * k for k in (4, 3, 3)]
    assert sum(s["extend_rays"] for s in stats) == rend.last_stats["extend_rays"]
    assert sum(s["shadow_rays"] for s in stats) == rend.last_stats["shadow_rays"]
    assert np.allclose(accum, rend.last_accum, rtol=1e-5, atol=1e-6)
    assert np.abs(rgb.astype(np.int16) - full.astype(np.int16)).max() <= 1
    # more replicas than samples: the idle ones contribute nothing
    one = pt.Profile(width=64, height=48, samples=1, bounces=2)
    r1 = pt.Renderer(profile=one).render(gsc)
    rgb1, _, st1 = pt.render_multi(replicas, one)
    assert np.array_equal(rgb1, r1) and [s["samples"] for s in st1] == [64 * 48, 0, 0]
    with pytest.raises(pt.PtError):
        pt.render_multi([gsc, gsc], prof)  # the same object twice
    for r in replicas[1:]:
        r.close()
    exe = os.path.join(os.path.dirname(pt.__file__), "path-tracer")
    yml = tmp_path / "p.yml"
    yml.write_text("resolution:\n  width: 96\n  height: 64\nsamples: 6\nbounces: 2\n")
    r = subprocess.run([exe, "render", scene_path("cube"), "-q", "--gpus", str(ndev + 1)], capture_output=True)
    assert r.returncode == 2 and b"CUDA device" in r.stderr
    out = tmp_path / "all.png"
    r = subprocess.run([exe, "render", scene_path("cube"), "-o", str(out), "-p", str(yml), "-q", "--gpus", "all"], capture_output=True)
    assert r.returncode == 0, r.stderr
    ref = pt.Renderer(profile=pt.Profile.load(str(yml))).render(loaded["cube"][2])
    assert np.abs(pt.read_png(str(out)).astype(np.int16) - ref.astype(np.int16)).max() <= 1


def test_scene_upload_is_double_buffered_and_takes_effect(loaded):
    """pt_scene_upload copies into the array set no kernel is reading and the next call switches to it: renders and
    traces interleaved with uploads keep returning the same results (several switches back and forth)."""
    import torch
    sa, osc, gsc = loaded["alpha_transparency"]
    sc = pt.Scene.load(scene_path("alpha_transparency"))
    prof = pt.Profile(width=128, height=96, samples=6, bounces=4)
    rr = R.random_rays(sa, 1 << 14, seed=3)
    ref_hits = gsc.trace_closest(rr)
    ref = torch.zeros(96, 128, 3, device="cuda")
    pt.render_accum_device(gsc, prof, ref.data_ptr())
    for k in range(5):
        n = sc.upload()
        assert n == sc.upload()  # two uploads in a row before a render: the later one wins, same bytes
        assert n > 0
        buf = torch.zeros(96, 128, 3, device="cuda")
        pt.render_accum_device(sc, prof, buf.data_ptr(), want_stats=(k % 2 == 0))
        torch.cuda.synchronize()
        assert torch.allclose(buf, ref, rtol=1e-5, atol=1e-6), k
        sc.upload()
        assert np.array_equal(sc.trace_closest(rr), ref_hits), k
    sc.close()


def test_scene_arrays_delivered_from_another_replica(loaded):
    """pt_scene_upload_targets / pt_scene_upload_commit: the arrays of a scene filled from somewhere else than its own
    host copy (under torchrun: an NCCL broadcast from the rank that uploaded; here: device-to-device copies from a
    second scene with the same geometry but other materials and lights).  After the commit the scene renders the
    donor's image, and goes back to its own after an ordinary upload."""
    import copy
    import torch
    import path_tracer_b200.multi as M
    sa = isf_py.load_isf(scene_path("cube"))
    sb = copy.copy(sa)
    sb.materials = [dict(m, albedo=(0.9, 0.2, 0.1), roughness=0.35) for m in sa.materials]
    sb.lights = [dict(l, color=tuple(0.5 * float(c) for c in l["color"])) for l in sa.lights]
    a, b = pt.Scene.from_arrays(sa), pt.Scene.from_arrays(sb)
    prof = pt.Profile(width=96, height=64, samples=4, bounces=3)

    def render(sc):
        buf = torch.zeros(64, 96, 3, device="cuda")
        pt.render_accum_device(sc, prof, buf.data_ptr())
        torch.cuda.synchronize()
        return buf

    img_a, img_b = render(a), render(b)
    assert not torch.allclose(img_a, img_b, rtol=1e-3, atol=1e-4)
    st = torch.cuda.Stream()
    for k in range(3):  # several rounds: both array sets of `a` get written this way
        b.upload()  # the donor's pending set = its host copy
        src, dst = b.upload_targets(st.cuda_stream), a.upload_targets(st.cuda_stream)
        assert [n for _, n in src] == [n for _, n in dst] and len(dst) >= 4
        with torch.cuda.stream(st):
            for (ps, n), (pd, _) in zip(src, dst):
                d = torch.as_tensor(M._DeviceBytes(pd, n), device="cuda")
                d.copy_(torch.as_tensor(M._DeviceBytes(ps, n), device="cuda"), non_blocking=True)
        a.upload_commit(st.cuda_stream)
        assert torch.equal(render(a), img_b), k
        assert torch.equal(render(b), img_b), k
        if k == 1:  # its own host copy again, in three parts (pt_scene_upload_part: what three ranks would each send)
            assert sum(a.upload_part(r, 3) for r in (2, 0, 1)) >= sum(n for _, n in dst)
        else:
            a.upload()
        assert torch.equal(render(a), img_a), k
    a.close()
    b.close()


def test_render_async_pipelines_frames_and_uploads(loaded):
    """pt_render_async: frames issued back to back on one stream, with a scene upload between them, each deliver the
    image pt_render returns."""
    import torch
    _, _, gsc = loaded["cube"]
    prof = pt.Profile(width=96, height=64, samples=5, bounces=3)
    want = pt.Renderer(profile=prof).render(gsc)
    sc = pt.Scene.load(scene_path("cube"))
    bufs = [torch.empty(64, 96, 3, dtype=torch.uint8).pin_memory() for _ in range(4)]
    st = torch.cuda.Stream()
    for b in bufs:
        sc.upload()
        pt.render_async(sc, prof, b.data_ptr(), stream=st.cuda_stream)
    pt.stream_synchronize(0, st.cuda_stream)
    for b in bufs:
        assert np.array_equal(b.numpy(), want)
    sc.close()


def test_zero_samples_and_bad_arguments(loaded):
    import torch
    _, _, gsc = loaded["cube"]
    prof = pt.Profile(width=16, height=16, samples=0, bounces=1)
    buf = torch.zeros(16, 16, 3, device="cuda")
    st = pt.render_accum_device(gsc, prof, buf.data_ptr())
    assert st["samples"] == 0 and float(buf.abs().sum()) == 0.0
    bad = pt.Profile(width=0, height=16)
    with pytest.raises(pt.PtError):
        pt.Renderer(profile=bad).render(gsc)
    with pytest.raises(pt.PtError):
        pt.Scene.load("/nonexistent/scene.isf")


# ---------------------------------------------------------------- post_processing (mod.rs:335-353, tonemap.rs)
@pytest.mark.parametrize("tm", [0, 1, 2])
def test_tonemap_device(tm):
    import torch
    rng = np.random.Generator(np.random.PCG64(tm))
    acc = (rng.gamma(0.7, 40.0, size=(64, 80, 3))).astype(np.float32)
    acc[0, 0] = [0, np.nan, np.inf]
    acc[0, 1] = [-1.0, 1e-9, 1e30]
    spp = 37
    d_acc = torch.from_numpy(acc).cuda()
    d_rgb = torch.zeros(64, 80, 3, dtype=torch.uint8, device="cuda")
    pt.tonemap_device(d_acc.data_ptr(), 80, 64, spp, tm, d_rgb.data_ptr())
    torch.cuda.synchronize()
    g = d_rgb.cpu().numpy()
    o = oracle_lib.tonemap(acc, spp, tm)
    diff = np.abs(g.astype(int) - o.astype(int))
    assert diff.max() <= 1  # powf may differ in the last ulp at a truncation boundary
    assert (diff == 0).mean() > 0.995
    assert tuple(g[0, 0]) == tuple(o[0, 0]) and tuple(g[0, 1]) == tuple(o[0, 1])


# ---------------------------------------------------------------- CLI drop-in (main.rs:24-57)
def test_cli_render(tmp_path, loaded):
    import subprocess
    exe = os.path.join(os.path.dirname(pt.__file__), "path-tracer")
    prof = tmp_path / "p.yml"
    prof.write_text("resolution:\n  width: 96\n  height: 64\nsamples: 4\nbounces: 2\ntonemap: ACES\n")
    out = tmp_path / "o.png"
    r = subprocess.run([exe, "render", scene_path("cube"), "-o", str(out), "-p", str(prof), "-q"], capture_output=True)
    assert r.returncode == 0, r.stderr
    img = pt.read_png(str(out))
    ref = pt.Renderer(profile=pt.Profile.load(str(prof))).render(loaded["cube"][2])
    assert img.shape == (64, 96, 3) and np.array_equal(img, ref)
    r = subprocess.run([exe, "render", "/nonexistent.isf", "-q"], capture_output=True)
    assert r.returncode == 2 and r.stderr  # main.rs:14-21


# ---------------------------------------------------------------- texture / material pipeline
def textured_scene(seed=1):
    """A small scene exercising every texture slot the reference's tests leave unpinned (material.rs:132-214,
    hit.rs:60-69): albedo / emissive / opacity / metalness / roughness / normal maps, uv repeat and negative uvs."""
    rng = np.random.Generator(np.random.PCG64(seed))
    sa = pt.abi.SceneArrays()
    tex = lambda h, w, c: rng.integers(0, 256, size=(h, w, c), dtype=np.uint8)
    sa.textures = [tex(16, 16, 3), tex(8, 8, 3), tex(32, 16, 1), tex(8, 8, 1), tex(8, 8, 1), tex(16, 16, 3)]
    sa.textures[2][::2] = 255  # opacity: opaque stripes
    nm = sa.textures[5].astype(np.float32)
    nm[..., 2] = 200 + nm[..., 2] * 0.2  # normal map mostly along +z
    sa.textures[5] = nm.astype(np.uint8)
    pos, nrm, uv = [], [], []

    def quad(o, e1, e2, n, uv_scale=1.0, uv_off=0.0):
        o, e1, e2 = np.asarray(o, np.float32), np.asarray(e1, np.float32), np.asarray(e2, np.float32)
        p = [o, o + e1, o + e1 + e2, o + e2]
        t = [(0, 0), (1, 0), (1, 1), (0, 1)]
        for a, b, c in ((0, 1, 2), (0, 2, 3)):
            pos.append(np.concatenate([p[a], p[b], p[c]]))
            nrm.append(np.tile(np.asarray(n, np.float32), 3))
            uv.append(np.concatenate([np.asarray(t[i], np.float32) * uv_scale + uv_off for i in (a, b, c)]))

    mats = []
    base = dict(albedo=[0.8, 0.8, 0.8], emissive=[0, 0, 0], opacity=1.0, metalness=0.0, roughness=1.0, ior=1.0)
    # floor: albedo + normal map + roughness texture, uvs repeating 3x and starting negative
    quad([-3, -1, -3], [6, 0, 0], [0, 0, 6], [0, 1, 0], 3.0, -1.25)
    mats.append((2, dict(base, albedo_tex=0, normal_tex=5, roughness_tex=4, roughness=0.9)))
    # back wall: emissive texture + metalness texture
    quad([-3, -1, -3], [6, 0, 0], [0, 4, 0], [0, 0, 1], 2.0)
    mats.append((2, dict(base, emissive=[0.6, 0.5, 0.4], emissive_tex=1, metalness=1.0, metalness_tex=3, roughness=0.3)))
    # a card with an opacity texture in front
    quad([-1, -1, 0.5], [2, 0, 0], [0, 2, 0], [0, 0, 1], 1.0)
    mats.append((2, dict(base, albedo=[0.2, 0.7, 0.3], opacity=0.9, opacity_tex=2, albedo_tex=0)))
    first = 0
    for n, m in mats:
        sa.models.append(dict(type=0, first_tri=first, n_tris=n, center=[0, 0, 0], radius=0.0))
        sa.materials.append(m)
        first += n
    sa.models.append(dict(type=1, first_tri=0, n_tris=0, center=[1.5, -0.4, 1.0], radius=0.6))
    sa.materials.append(dict(base, albedo=[0.9, 0.3, 0.2], metalness=0.5, roughness=0.2, opacity=0.6))
    sa.tri_pos, sa.tri_nrm, sa.tri_uv = np.array(pos, np.float32), np.array(nrm, np.float32), np.array(uv, np.float32)
    sa.lights = [dict(type=0, vec=[0.5, 3.0, 3.0], color=[300, 300, 300], size=0.1),
                 dict(type=1, vec=[-0.3, -1.0, -0.5], color=[1.5, 1.4, 1.2], size=0.0)]
    cam = np.eye(4, dtype=np.float32)
    cam[3, :3] = [0.0, 0.5, 6.0]  # column-major: translation in column 3
    sa.camera = cam.reshape(16)
    sa.fov = np.float32(0.7)
    sa.background = np.array([0.05, 0.06, 0.08], np.float32)
    return sa


def test_textured_materials_vs_oracle():
    sa = textured_scene()
    gsc, osc = pt.Scene.from_arrays(sa), oracle_lib.OracleScene(sa)
    rr = R.random_rays(sa, 1 << 16, seed=2)
    assert np.array_equal(gsc.trace_closest(rr), osc.trace_closest(rr, brute=True, threads=THREADS))
    assert np.allclose(gsc.trace_transmittance(rr), osc.trace_transmittance(rr), rtol=2e-6, atol=1e-7)
    prof = pt.Profile(width=160, height=120, samples=8, bounces=4, tonemap="ACES")
    pix = np.arange(160 * 120, dtype=np.uint32)
    gv, gc = gsc.trace_paths(prof, pix, sample=1, seed=3)
    ov, oc = osc.trace_paths(prof.c, pix, sample=1, seed=3)
    for f in ("ray_o", "ray_d", "prim_id", "t"):
        assert np.array_equal(gv[:, 0][f], ov[:, 0][f]), f
    # first-bounce shading (every texture fetch, the TBN normal map, both light types) agrees to float rounding
    r0g, r0o = gv[:, 0]["radiance"], ov[:, 0]["radiance"]
    ok = np.abs(r0g - r0o) <= 1e-4 * np.maximum(np.abs(r0o), 1e-2)
    assert ok.all(1).mean() > 0.995, ok.all(1).mean()
    assert (gv[:, 0]["prim_id"] >= 0).mean() > 0.5 and (r0o.sum(1) > 0).mean() > 0.4
    rend = pt.Renderer(profile=prof)
    rend.render(gsc, keep_accum=True)
    oacc = osc.render_accum(prof.c, rng_mode=1, threads=THREADS)
    close = np.abs(rend.last_accum - oacc) <= 1e-3 * np.maximum(np.abs(oacc), 1.0)
    assert close.mean() > 0.97, close.mean()
    assert rend.last_stats["alpha_tests"] > 0 and rend.last_stats["alpha_recasts"] == 0


# ---------------------------------------------------------------- debug_render (debug_renderer.rs:11-105)
@pytest.mark.parametrize("name", ["head", "alpha_transparency", "spheres", "cube"])
def test_debug_textures_vs_oracle(loaded, name):
    sa, osc, gsc = loaded[name]
    w, h = 200, 150
    got = pt.debug_render(gsc, w, h)
    want, n_hit = osc.debug_textures(w, h)
    assert got["pixels_hit"] == n_hit > 0
    for k, aov in enumerate(pt.AOV_NAMES):
        d = np.abs(got[aov].astype(np.int16) - want[k].astype(np.int16))
        if aov == "albedo":  # powf(texel, 2.2): CUDA vs glibc last-ulp differences may flip a u8 truncation
            assert d.max() <= 1 and (d == 0).mean() > 0.995, (aov, d.max(), (d == 0).mean())
        elif aov == "normal":  # normalised TBN / sphere normals: last-ulp differences at u8 truncation boundaries
            assert d.max() <= 1 and (d == 0).mean() > 0.999, (aov, d.max(), (d == 0).mean())
        else:
            assert d.max() == 0, aov
    # pixels without a hit stay 0 in every plane; ior plane = ior / 3 on hit pixels
    miss = (want[0] == 0).all(-1) & (want[6] == 0).all(-1)
    for aov in pt.AOV_NAMES:
        assert (got[aov][miss] == 0).all()


def test_debug_textures_textured_scene_and_errors():
    sa = textured_scene()
    gsc, osc = pt.Scene.from_arrays(sa), oracle_lib.OracleScene(sa)
    got = pt.debug_render(gsc, 160, 120)
    want, n_hit = osc.debug_textures(160, 120)
    assert got["pixels_hit"] == n_hit
    for k, aov in enumerate(pt.AOV_NAMES):
        d = np.abs(got[aov].astype(np.int16) - want[k].astype(np.int16))
        assert d.max() <= 1 and (d == 0).mean() > 0.99, (aov, d.max(), (d == 0).mean())
    assert len(np.unique(got["metalness"])) > 2 and len(np.unique(got["roughness"])) > 2  # the maps are really sampled
    with pytest.raises(pt.PtError):
        pt.debug_render(gsc, 0, 10)


def test_cli_debug_textures(tmp_path, loaded):
    import subprocess
    exe = os.path.join(os.path.dirname(pt.__file__), "path-tracer")
    prof = tmp_path / "p.yml"
    prof.write_text("resolution:\n  width: 96\n  height: 64\n")
    r = subprocess.run([exe, "render", scene_path("head"), "--debug-textures", "-p", str(prof)], cwd=tmp_path, capture_output=True)
    assert r.returncode == 0, r.stderr
    got = pt.debug_render(loaded["head"][2], 96, 64)
    for aov in pt.AOV_NAMES:  # debug_renderer.rs:59: "<buffer_name>.png" in the working directory
        assert np.array_equal(pt.read_png(str(tmp_path / (aov + ".png")), 3), got[aov]), aov
    assert not (tmp_path / "render.png").exists()  # main.rs:41-44 returns before rendering


# ---------------------------------------------------------------- viewer feed (mod.rs:133-141) as headless snapshots
def test_progressive_snapshots(loaded):
    sa, osc, gsc = loaded["cube"]
    prof = pt.Profile(width=96, height=64, samples=10, bounces=3)
    rend = pt.Renderer(profile=prof)
    full = rend.render(gsc, keep_accum=True)
    full_acc = rend.last_accum.copy()
    snaps = []
    img = rend.render(gsc, keep_accum=True, snapshot_every=4, on_snapshot=lambda a, done: snaps.append((done, a)) and False)
    assert [d for d, _ in snaps] == [4, 8, 10]
    assert np.array_equal(img, full) and np.array_equal(snaps[-1][1], full)
    assert np.allclose(rend.last_accum, full_acc, rtol=1e-5, atol=1e-6)  # same samples, summed in three chunks
    assert rend.last_stats["samples"] == 96 * 64 * 10
    # snapshot k shows `*pixel / current_sample` post-processed (mod.rs:137-138) == a render of the first k samples
    first4 = pt.Renderer(profile=pt.Profile(width=96, height=64, samples=4, bounces=3)).render(gsc)
    assert np.array_equal(snaps[0][1], first4)
    # a truthy return closes the feed (mod.rs:139-141): no more snapshots, the render still completes
    seen = []
    img2 = rend.render(gsc, snapshot_every=2, on_snapshot=lambda a, done: seen.append(done) or True)
    assert seen == [2] and np.array_equal(img2, full)


def test_cli_viewer_snapshots(tmp_path):
    import subprocess
    exe = os.path.join(os.path.dirname(pt.__file__), "path-tracer")
    prof = tmp_path / "p.yml"
    prof.write_text("resolution:\n  width: 64\n  height: 48\nsamples: 6\nbounces: 2\n")
    a, b = tmp_path / "a.png", tmp_path / "b.png"
    r = subprocess.run([exe, "render", scene_path("cube"), "-v", "--snapshot-every", "2", "-o", str(a), "-p", str(prof)], capture_output=True)
    assert r.returncode == 0, r.stderr
    assert r.stderr.count(b"spp ->") == 3
    r = subprocess.run([exe, "render", scene_path("cube"), "-q", "-o", str(b), "-p", str(prof)], capture_output=True)
    assert r.returncode == 0
    assert np.array_equal(pt.read_png(str(a)), pt.read_png(str(b)))


# ---------------------------------------------------------------- GPU builders: LBVH and PLOC (replace KDTree::build)
@pytest.fixture(params=["lbvh", "ploc"])
def lbvh_builder(request):
    pt.set_accel_builder(request.param)
    yield request.param
    pt.set_accel_builder("sah")


@pytest.mark.parametrize("name", ["head", "alpha_transparency", "spheres", "white_furnace_indirect"])
def test_lbvh_closest_hit_matches_oracle(loaded, lbvh_builder, name):
    sa, osc, sah_scene = loaded[name]
    sc = pt.Scene.load(scene_path(name))
    info = sc.accel_info()
    if sc.info()["n_prims"] >= 16:
        assert info["accel"] == lbvh_builder and info["max_depth"] >= 1
    rr = R.random_rays(sa, 1 << 18, seed=11)
    got = sc.trace_closest(rr)
    assert np.array_equal(got, sah_scene.trace_closest(rr))  # the tree is a filter: identical records
    assert np.array_equal(got, osc.trace_closest(rr, threads=THREADS))
    assert np.allclose(sc.trace_transmittance(rr[:65536]), osc.trace_transmittance(rr[:65536]), rtol=2e-6, atol=1e-7)


def test_lbvh_soup_brute_force_and_render(lbvh_builder):
    sc = pt.Scene.synthetic("soup", 100000, seed=9)
    assert sc.accel_info()["accel"] == lbvh_builder
    sa = sc.arrays()
    osc = oracle_lib.OracleScene(sa)
    rr = R.random_rays(sa, 1 << 14, seed=4)
    assert np.array_equal(sc.trace_closest(rr), osc.trace_closest(rr, brute=True, threads=THREADS))
    # same Philox streams, different tree: the accumulation buffer only differs by float atomics order
    prof = pt.Profile(width=128, height=96, samples=4, bounces=3)
    pt.set_accel_builder("sah")
    sc2 = pt.Scene.synthetic("soup", 100000, seed=9)
    assert sc2.accel_info()["accel"] == "sah"
    r1, r2 = pt.Renderer(profile=prof), pt.Renderer(profile=prof)
    r1.render(sc, keep_accum=True)
    r2.render(sc2, keep_accum=True)
    assert np.allclose(r1.last_accum, r2.last_accum, rtol=1e-4, atol=1e-5)
    assert r1.last_stats["extend_rays"] == r2.last_stats["extend_rays"]


def test_lbvh_large_scene_properties(lbvh_builder):
    """Full-size check through size-independent properties: every primitive is reachable (a ray aimed at each sampled
    triangle's centroid reports a hit no farther than that triangle) and the build is deterministic in its results."""
    sc = pt.Scene.synthetic("city", 2000000, seed=5)
    info = sc.accel_info()
    assert info["accel"] == lbvh_builder and info["max_depth"] * 3 + 4 <= 96
    sa = sc.arrays()
    rng = np.random.default_rng(0)
    pick = rng.integers(0, len(sa.tri_pos), 200000)
    tri = sa.tri_pos[pick].reshape(-1, 3, 3).astype(np.float64)
    cen = tri.mean(1)
    nrm = np.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0])
    ok = np.linalg.norm(nrm, axis=1) > 1e-12
    nrm[ok] /= np.linalg.norm(nrm[ok], axis=1, keepdims=True)
    o = cen + nrm * 0.05
    rays = np.concatenate([o, -nrm], 1).astype(np.float32)[ok]
    hits = sc.trace_closest(rays)
    assert (hits["prim_id"] >= 0).all()
    assert (hits["t"] <= 0.05 * 1.01 + 1e-4).all()
    pt.set_accel_builder("sah")
    ref = pt.Scene.synthetic("city", 2000000, seed=5)
    assert np.array_equal(hits, ref.trace_closest(rays))


# ---------------------------------------------------------------- full-size configs through size-independent properties
def test_full_size_foliage_properties():
    """BASELINE.json configs[3] geometry (5 M triangles, 30 % constant-alpha + 20 % alpha-textured models) at full size:
    the oracle cannot trace it in seconds, so the checks are properties that hold at any size."""
    sc = pt.Scene.synthetic("foliage", 5000000, seed=4)
    sa = sc.arrays()
    rr = R.random_rays(sa, 1 << 18, seed=21)
    hits = sc.trace_closest(rr)
    hit = hits["prim_id"] >= 0
    assert 0.2 < hit.mean() < 1.0
    # (a) a reported hit is a real one: re-intersecting the reported triangle alone reproduces t, u, v bit-for-bit
    prim = hits["prim_id"][hit]
    model_first = np.array([m["first_tri"] for m in sa.models])
    assert all(m["type"] == 0 for m in sa.models)  # meshes only: global prim id == triangle index
    single = pt.triangle_intersect(rr[hit], sa.tri_pos[prim])
    assert (single["prim_id"] >= 0).all()
    for f in ("t", "u", "v"):
        assert np.array_equal(single[f], hits[f][hit]), f
    # (b) nothing closer: a ray stopped just short of the reported hit is unoccluded for an opaque-only walk
    #     (transmittance of a directional walk over (1 - opacity) factors stays within [0, 1])
    tr = sc.trace_transmittance(rr[:65536])
    assert np.isfinite(tr).all() and (tr >= 0).all() and (tr <= 1).all()
    assert (tr[~hit[:65536]] == 1.0).all()  # a ray that hits nothing is fully transmitted
    # (c) disjoint sample subsets sum to the full render; the alpha walk is exercised
    import torch
    prof = pt.Profile(width=480, height=270, samples=4, bounces=4)
    full = torch.zeros(270, 480, 3, device="cuda")
    st = pt.render_accum_device(sc, prof, full.data_ptr())
    parts = torch.zeros_like(full)
    for g in range(2):
        pt.render_accum_device(sc, prof, parts.data_ptr(), sample_begin=g, sample_stride=2, n_samples=2)
    torch.cuda.synchronize()
    assert torch.isfinite(full).all()
    assert torch.allclose(full, parts, rtol=1e-4, atol=1e-5)
    assert st["alpha_tests"] > 0 and st["alpha_recasts"] == 0 and st["shadow_dropped"] == 0
    assert st["shadow_rays"] > 0 and st["extend_rays"] >= 480 * 270 * 4


def test_full_size_console_properties():
    """BASELINE.json configs[2] geometry (the headline workload, 1 M triangles) at full size."""
    sc = pt.Scene.synthetic("console", 1000000, seed=3)
    sa = sc.arrays()
    assert all(m["type"] == 0 for m in sa.models)  # meshes only: global prim id == triangle index
    rr = R.random_rays(sa, 1 << 18, seed=22)
    hits = sc.trace_closest(rr)
    hit = hits["prim_id"] >= 0
    assert 0.05 < hit.mean() < 1.0
    # (a) a reported hit is a real one, bit for bit
    single = pt.triangle_intersect(rr[hit], sa.tri_pos[hits["prim_id"][hit]])
    for f in ("t", "u", "v"):
        assert np.array_equal(single[f], hits[f][hit]), f
    # (b) the tree is only a filter: the GPU-built LBVH returns the same records
    pt.set_accel_builder("lbvh")
    try:
        sc2 = pt.Scene.synthetic("console", 1000000, seed=3)
    finally:
        pt.set_accel_builder("sah")
    assert sc2.accel_info()["accel"] == "lbvh"
    assert hits.tobytes() == sc2.trace_closest(rr).tobytes()
    # (c) nothing closer: brute force of a few rays against every triangle of the scene through pt_triangle_intersect
    n_tri = len(sa.tri_pos)
    pick = np.concatenate([np.nonzero(hit)[0][:12], np.nonzero(~hit)[0][:4]])
    for k in pick:
        every = pt.triangle_intersect(np.broadcast_to(rr[k], (n_tri, 6)), sa.tri_pos)
        ok = every["prim_id"] >= 0
        if hits["prim_id"][k] < 0:
            assert not ok.any()
            continue
        t = np.where(ok, every["t"], np.inf)
        best = np.lexsort((np.arange(n_tri), t))[0]  # smallest (t, id), the order of utils.rs:13-19
        assert best == hits["prim_id"][k] and t[best] == hits["t"][k]
    # (d) disjoint sample subsets sum to the full render
    import torch
    prof = pt.Profile(width=480, height=270, samples=4, bounces=5)
    full = torch.zeros(270, 480, 3, device="cuda")
    st = pt.render_accum_device(sc, prof, full.data_ptr())
    parts = torch.zeros_like(full)
    for g in range(2):
        pt.render_accum_device(sc, prof, parts.data_ptr(), sample_begin=g, sample_stride=2, n_samples=2)
    torch.cuda.synchronize()
    assert torch.isfinite(full).all() and torch.allclose(full, parts, rtol=1e-4, atol=1e-5)
    assert st["shadow_rays"] > 0 and st["extend_rays"] >= 480 * 270 * 4


# ---------------------------------------------------------------- box margins: rays that graze triangle edges
@pytest.mark.parametrize("accel", ["sah", "lbvh", "ploc"])
def test_edge_grazing_rays_match_brute_force(accel):
    sa, targets, cam, rng = R.edge_grazing_scene()
    rays = R.edge_grazing_rays(targets, cam, rng, 1 << 17)
    osc = oracle_lib.OracleScene(sa)
    ref = osc.trace_closest(rays, brute=True, threads=THREADS)
    pt.set_accel_builder(accel)
    try:
        sc = pt.Scene.from_arrays(sa)
    finally:
        pt.set_accel_builder("sah")
    assert sc.accel_info()["accel"] == accel
    got = sc.trace_closest(rays)
    # the batch really sits on the edges: a good share of it hits and a good share misses the targeted wall
    frac = float((ref["prim_id"] >= 0).mean())
    assert 0.3 < frac < 0.99, frac
    bad = np.nonzero((got["prim_id"] != ref["prim_id"]) | (got["t"] != ref["t"]))[0]
    assert len(bad) == 0, (len(bad), rays[bad[:4]], got[bad[:4]], ref[bad[:4]])
    assert np.array_equal(osc.trace_closest(rays, threads=THREADS), ref)  # the oracle's own filter too
    # (trace_closest above already ran k_extend itself); a full frame looking along the walls is tree independent too
    prof = pt.Profile(width=256, height=144, samples=2, bounces=2)
    r1 = pt.Renderer(profile=prof)
    r1.render(sc, keep_accum=True)
    os.environ["PT_B200_LEAF_TRIS"] = "8"
    try:
        sc8 = pt.Scene.from_arrays(sa)
    finally:
        del os.environ["PT_B200_LEAF_TRIS"]
    r2 = pt.Renderer(profile=prof)
    r2.render(sc8, keep_accum=True)
    assert np.allclose(r1.last_accum, r2.last_accum, rtol=1e-4, atol=1e-5)
