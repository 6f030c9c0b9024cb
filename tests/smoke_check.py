"""__graft_entry__.smoke(): one small invocation of the hot path on cuda:0, checked against the CPU oracle."""
import os

import numpy as np

import isf_py
import oracle_lib
import rays as R

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def run(pkg):
    if pkg.device_count() < 1:
        raise RuntimeError("smoke() needs a CUDA device: libpt_b200 has no CPU fallback")
    path = os.path.join(GOLDEN, "scenes", "alpha_transparency", "scene.isf")
    sa = isf_py.load_isf(path)
    osc = oracle_lib.OracleScene(sa)
    gsc = pkg.Scene.load(path)
    rr = R.random_rays(sa, 1 << 14)
    g, o = gsc.trace_closest(rr), osc.trace_closest(rr)  # through k_extend, the kernel that renders
    assert np.array_equal(g, o), "closest-hit batch differs from the oracle"
    ga, oa = gsc.trace_accepted(rr, bounce=0, sample=1, seed=2), osc.trace_accepted(rr, bounce=0, sample=1, seed=2)
    assert ga.tobytes() == oa.tobytes(), "alpha-walk batch differs from the oracle"
    gt, ot = gsc.trace_shadow(rr), osc.trace_shadow(rr)  # through k_shadow
    assert np.allclose(gt, ot, rtol=2e-6, atol=1e-7) and ((gt == 0) == (ot == 0)).all(), "shadow batch differs"
    prof = pkg.Profile(width=64, height=64, samples=4, bounces=4)
    rend = pkg.Renderer(profile=prof)
    img = rend.render(gsc, keep_accum=True)
    oacc = osc.render_accum(prof.c, rng_mode=1)
    close = np.abs(rend.last_accum - oacc) <= 1e-3 * np.maximum(np.abs(oacc), 1.0)
    assert close.mean() >= 0.999, "render differs from the oracle: %.4f of the pixels agree" % close.mean()
    assert img.shape == (64, 64, 3)
    st = rend.last_stats
    print("smoke ok: %d rays bit-exact, %.1f%% pixels identical to the oracle, %d kernel launches, %.3f ms"
          % (len(rr), 100 * close.mean(), st["kernel_launches"], st["ms_total"]))
