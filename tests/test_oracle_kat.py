"""Pins the CPU oracle (oracle/) against the reference's own known answers (CPU only, no GPU).

  * the 3012 hit + 3012 miss Möller–Trumbore vectors of the reference
    (tests/moller_trumbore/{hit,miss}_tests.yml, asserted by src/scene/internal/triangle.rs:200-233,
    tolerance 1e-5 absolute, triangle.rs:213-216) — committed as tests/golden/moller_trumbore.npz;
  * the seven SHA-1 hashes of the raw 800x600 RGB8 image (src/main.rs:100-165) through the oracle's
    ChaCha12/PCG32 emulation of rand 0.8 StdRng — committed as tests/golden/sha1_kat.json.
"""
import hashlib
import json
import os

import numpy as np
import pytest

import isf_py
import oracle_lib
from conftest import GOLDEN

# Scenes whose reference hash the oracle reproduces bit-for-bit: the whole integrator (raygen, all-hits ray_cast,
# alpha walk, shadow transmittance, Cook-Torrance, cutoff/roulette, FILMIC + gamma + u8), the ChaCha12/PCG32 stream
# emulation, the isf/PNG decoding and glibc's libm all agree with the reference's Rust binary on these.
SHA1_MATCH = ["cube", "reflection", "head", "spheres", "alpha_transparency", "white_furnace_indirect"]
# The one known miss.  white_furnace_direct is built from axis-aligned quads lit by an axis-parallel directional
# light at 0 bounces: every triangle AABB is flat and every shadow ray has two zero direction components, i.e. the
# measure-zero slab-test cases (0 * inf) whose outcome is decided inside the third-party kdtree-ray crate, whose
# source is not under /root/reference.  The oracle implements brute-force ray_cast semantics there.
SHA1_KNOWN_MISS = ["white_furnace_direct"]


@pytest.fixture(scope="module")
def mt():
    return np.load(os.path.join(GOLDEN, "moller_trumbore.npz"))


def test_mt_hit_vectors(mt):
    rays, tris, gold = mt["hit_rays"], mt["hit_tris"], mt["hit_gold"]
    assert len(rays) == 3012
    worst = 0.0
    for r, t, g in zip(rays, tris, gold):
        hit, out = oracle_lib.moller_trumbore(r, t)
        assert hit
        err = np.abs(out[:3].astype(np.float64) - g).max()
        worst = max(worst, err)
    assert worst < 1e-5  # triangle.rs:213-216


def test_mt_miss_vectors(mt):
    rays, tris = mt["miss_rays"], mt["miss_tris"]
    assert len(rays) == 3012
    for r, t in zip(rays, tris):
        hit, _ = oracle_lib.moller_trumbore(r, t)
        assert not hit


def _render_sha1(name, kat):
    p = kat["profile"]
    bounces = kat.get("bounces_override", {}).get(name, p["bounces"])
    sa = isf_py.load_isf(os.path.join(GOLDEN, "scenes", name, "scene.isf"))
    sc = oracle_lib.OracleScene(sa)
    prof = oracle_lib.make_profile(p["width"], p["height"], p["samples"], bounces, 1)
    accum = sc.render_accum(prof, rng_mode=0, threads=os.cpu_count() or 8)
    img = oracle_lib.tonemap(accum, p["samples"], 1)
    return hashlib.sha1(img.tobytes()).hexdigest()


@pytest.fixture(scope="module")
def kat():
    with open(os.path.join(GOLDEN, "sha1_kat.json")) as f:
        return json.load(f)


@pytest.mark.parametrize("name", SHA1_MATCH)
def test_sha1_image_kat(name, kat):
    """src/main.rs:100-165: the oracle's image is bit-identical to the reference's."""
    assert _render_sha1(name, kat) == kat["scenes"][name]


@pytest.mark.parametrize("name", SHA1_KNOWN_MISS)
def test_sha1_known_miss_is_still_a_miss(name, kat):
    """Documents the unpinned scene; if this starts matching, move it to SHA1_MATCH."""
    assert _render_sha1(name, kat) != kat["scenes"][name]


def test_bvh_filter_equals_brute_force():
    """The oracle's all-hits BVH filter (stand-in for kdtree-ray) must not change ray_cast results."""
    import rays as R
    for name in ("cube", "alpha_transparency", "white_furnace_direct", "spheres", "reflection"):
        sa = isf_py.load_isf(os.path.join(GOLDEN, "scenes", name, "scene.isf"))
        sc = oracle_lib.OracleScene(sa)
        rr = R.random_rays(sa, 20000, seed=7)
        a = sc.trace_closest(rr, brute=False)
        b = sc.trace_closest(rr, brute=True)
        assert np.array_equal(a, b), name


def test_bvh_filter_keeps_edge_grazing_hits():
    """Rays aimed within a few 1e-4 of the edges of thin triangles 150 units away: f32 Moller-Trumbore accepts rays
    that pass slightly outside such a triangle, and the filter's boxes carry a margin for exactly that
    (oracle/scene.cpp prim_box).  Without the margin this batch loses hits."""
    import rays as R
    sa, targets, cam, rng = R.edge_grazing_scene()
    rr = R.edge_grazing_rays(targets, cam, rng, 1 << 16)
    sc = oracle_lib.OracleScene(sa)
    b = sc.trace_closest(rr, brute=True)
    assert np.array_equal(sc.trace_closest(rr, brute=False), b)
    assert 0.3 < float((b["prim_id"] >= 0).mean()) < 0.99
