"""CPU-only checks of the drop-in boundary: the C ABI library loads and exports every symbol include/pt_b200.h
declares, the host-side loaders (isf JSON, profile YAML, PNG) agree with independent Python readers, and compute
entry points fail loudly (no CPU fallback) when no CUDA device exists."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import isf_py
from conftest import GOLDEN, ROOT

import path_tracer_b200 as pt

SCENES = sorted(os.listdir(os.path.join(GOLDEN, "scenes")))
HAS_GPU = pt.device_count() > 0


def test_header_symbols_exported():
    hdr = open(os.path.join(ROOT, "include", "pt_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = set(re.findall(r"\b(pt_[a-z0-9_]+)\s*\(", hdr))
    assert len(names) >= 20
    L = pt.lib()
    for n in names:
        assert hasattr(L, n), n
    assert names == set(pt.api.EXPORTS), names ^ set(pt.api.EXPORTS)


def test_struct_layouts_match_header():
    """sizeof() of every ctypes mirror equals the C compiler's (compiled from the header itself)."""
    src = '#include "pt_b200.h"\n#include <stdio.h>\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu",' \
          "sizeof(pt_material),sizeof(pt_model),sizeof(pt_light),sizeof(pt_texture),sizeof(pt_scene_desc)," \
          "sizeof(pt_profile),sizeof(pt_hit),sizeof(pt_stats),sizeof(pt_path_vertex),sizeof(pt_status));return 0;}"
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "s.c"), "w").write(src)
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), "-o", os.path.join(d, "s"), os.path.join(d, "s.c")])
        sizes = [int(x) for x in subprocess.check_output([os.path.join(d, "s")]).split()]
    a = pt.abi
    mine = [C.sizeof(t) for t in (a.pt_material, a.pt_model, a.pt_light, a.pt_texture, a.pt_scene_desc, a.pt_profile,
                                  a.pt_hit, a.pt_stats, a.pt_path_vertex)] + [4]
    assert mine == sizes
    assert a.HIT_DTYPE.itemsize == C.sizeof(a.pt_hit) and a.PATH_VERTEX_DTYPE.itemsize == C.sizeof(a.pt_path_vertex)


def test_profile_defaults_and_yaml(tmp_path):
    p = pt.Profile()  # config/profile.rs:34-40, resolution.rs:9-16
    assert (p.width, p.height, p.bounces, p.samples, p.tonemap) == (1920, 1080, 4, 64, 1)
    f = tmp_path / "p.yml"
    f.write_text("resolution:\n  width: 800\n  height: 600\nbounces: 7\nsamples: 128\nbrdf: COOK_TORRANCE\ntonemap: REINHARD\n")
    p = pt.Profile.load(str(f))
    assert (p.width, p.height, p.bounces, p.samples, p.tonemap) == (800, 600, 7, 128, 0)
    f.write_text("samples: 3\n")
    p = pt.Profile.load(str(f))
    assert (p.width, p.height, p.bounces, p.samples, p.tonemap) == (1920, 1080, 4, 3, 1)
    f.write_text("tonemap: ACES # comment\nbounces: 0\n")
    p = pt.Profile.load(str(f))
    assert (p.tonemap, p.bounces) == (2, 0)
    for bad in ("tonemap: NOPE\n", "brdf: LAMBERT\n", "resolution:\n  width: 5\n", "samples: many\n"):
        f.write_text(bad)
        with pytest.raises(pt.PtError):
            pt.Profile.load(str(f))
    with pytest.raises(pt.PtError) as e:
        pt.Profile.load(str(tmp_path / "missing.yml"))
    assert e.value.code == -2


@pytest.mark.parametrize("name", SCENES)
def test_isf_loader_matches_python_reader(name):
    """The C++ isf loader (scene/isf.rs schema + serde defaults + texture bank) vs the independent Python one."""
    path = os.path.join(GOLDEN, "scenes", name, "scene.isf")
    sc = pt.Scene.load(path, device=-1)  # host-only handle: loaders need no GPU
    a, b = sc.arrays(), isf_py.load_isf(path)
    assert np.array_equal(a.tri_pos, b.tri_pos) and np.array_equal(a.tri_nrm, b.tri_nrm) and np.array_equal(a.tri_uv, b.tri_uv)
    assert len(a.models) == len(b.models)
    for ma, mb in zip(a.models, b.models):
        assert ma["type"] == mb["type"] and ma["n_tris"] == mb["n_tris"] and ma["first_tri"] == mb["first_tri"]
        if ma["type"] == 1:
            assert np.allclose(ma["center"], np.asarray(mb["center"], np.float32), rtol=0, atol=0) and ma["radius"] == np.float32(mb["radius"])
    for ma, mb in zip(a.materials, b.materials):
        for k in ("albedo", "emissive"):
            assert np.array_equal(np.asarray(ma[k], np.float32), np.asarray(mb[k], np.float32)), k
        for k in ("opacity", "metalness", "roughness", "ior"):
            assert np.float32(ma[k]) == np.float32(mb[k]), k
        for k in ("albedo_tex", "emissive_tex", "opacity_tex", "metalness_tex", "roughness_tex", "normal_tex"):
            assert (ma[k] < 0) == (mb[k] < 0), k
            if ma[k] >= 0:
                assert np.array_equal(a.textures[ma[k]], b.textures[mb[k]]), k
    assert len(a.lights) == len(b.lights)
    for la, lb in zip(a.lights, b.lights):
        assert la["type"] == lb["type"]
        assert np.array_equal(np.asarray(la["vec"], np.float32), np.asarray(lb["vec"], np.float32))
        assert np.array_equal(np.asarray(la["color"], np.float32), np.asarray(lb["color"], np.float32))
    assert np.array_equal(a.camera, b.camera) and a.fov == b.fov and np.array_equal(a.background, b.background)
    assert sc.info()["n_prims"] == b.n_prims()


def test_isf_errors(tmp_path):
    with pytest.raises(pt.PtError) as e:
        pt.Scene.load(str(tmp_path / "nope.isf"), device=-1)
    assert e.value.code == -2
    f = tmp_path / "bad.isf"
    f.write_text("{ not json")
    with pytest.raises(pt.PtError) as e:
        pt.Scene.load(str(f), device=-1)
    assert e.value.code == -3
    f.write_text('{"models": [], "lights": [], "background": [0,0,0]}')  # camera missing
    with pytest.raises(pt.PtError):
        pt.Scene.load(str(f), device=-1)


def test_png_roundtrip(tmp_path):
    rng = np.random.Generator(np.random.PCG64(1))
    img = rng.integers(0, 256, size=(37, 53, 3), dtype=np.uint8)
    p = str(tmp_path / "a.png")
    pt.write_png(p, img)
    assert np.array_equal(pt.read_png(p, 3), img)
    from PIL import Image
    assert np.array_equal(np.asarray(Image.open(p).convert("RGB")), img)
    gray = rng.integers(0, 256, size=(20, 31), dtype=np.uint8)
    Image.fromarray(gray, "L").save(str(tmp_path / "g.png"))
    assert np.array_equal(pt.read_png(str(tmp_path / "g.png"), 1)[:, :, 0], gray)
    ref = os.path.join(GOLDEN, "scenes", "head", "albedo_tex_0.png")
    assert np.array_equal(pt.read_png(ref, 3), np.asarray(Image.open(ref).convert("RGB")))


def test_accel_builder_selection_host_only():
    pt.set_accel_builder("lbvh")
    pt.set_accel_builder("ploc")
    pt.set_accel_builder("auto")
    pt.set_accel_builder("sah")
    with pytest.raises(pt.PtError):
        pt.set_accel_builder(7)
    sc = pt.Scene.synthetic("soup", 1000, seed=1, device=-1)
    assert sc.accel_info()["accel"] == "sah"  # host-only scenes build no tree


def test_synthetic_scenes_host_only():
    for recipe, n in (("console", 20000), ("foliage", 20000), ("city", 20000), ("soup", 5000)):
        sc = pt.Scene.synthetic(recipe, n, seed=3, device=-1)
        info = sc.info()
        assert 0.5 * n <= info["n_tris"] <= 1.5 * n, (recipe, info)
        sa = sc.arrays()
        assert np.isfinite(sa.tri_pos).all() and np.isfinite(sa.tri_nrm).all()
        sc2 = pt.Scene.synthetic(recipe, n, seed=3, device=-1)
        assert np.array_equal(sc2.arrays().tri_pos, sa.tri_pos)  # deterministic in (recipe, n, seed)
    with pytest.raises(pt.PtError):
        pt.Scene.synthetic("nope", 10, device=-1)


def test_scene_replicate_host_only_and_multi_errors():
    """pt_scene_replicate copies the parsed scene (texture descriptors re-pointed at the copy's own texels);
    pt_render_multi refuses host-only / duplicate / missing replicas instead of falling back to the CPU."""
    src = pt.Scene.load(os.path.join(GOLDEN, "scenes", "alpha_transparency", "scene.isf"), device=-1)
    a = src.arrays()
    rep = src.replicate(-1)
    src.close()  # the replica must not point into the source
    b = rep.arrays()
    for k in ("tri_pos", "tri_nrm", "tri_uv"):
        assert np.array_equal(getattr(a, k), getattr(b, k)), k
    assert len(a.textures) == len(b.textures) > 0
    for ta, tb in zip(a.textures, b.textures):
        assert np.array_equal(ta, tb)
    assert rep.info() == pt.Scene.load(os.path.join(GOLDEN, "scenes", "alpha_transparency", "scene.isf"), device=-1).info()
    prof = pt.Profile(width=8, height=8, samples=2)
    with pytest.raises(pt.PtError) as e:
        pt.render_multi([rep], prof)
    assert e.value.code == -4  # host-only: no CPU fallback
    with pytest.raises(pt.PtError):
        pt.render_multi([], prof)
    if not HAS_GPU:
        with pytest.raises(pt.PtError) as e:
            rep.replicate(0)
        assert e.value.code == -4


def test_cli_errors_exit_code_2(tmp_path):
    exe = os.path.join(os.path.dirname(pt.__file__), "path-tracer")
    for args in ([], ["bogus"], ["render"], ["render", "/nonexistent/scene.isf", "-q"], ["render", "x.isf", "--nope"],
                 ["render", "x.isf", "--gpus"], ["render", "/nonexistent/scene.isf", "-q", "--gpus", "all"]):
        r = subprocess.run([exe] + args, capture_output=True)
        assert r.returncode == 2, (args, r.returncode)  # main.rs:14-21
        assert r.stderr
    bad = tmp_path / "p.yml"
    bad.write_text("tonemap: NOPE\n")
    r = subprocess.run([exe, "render", os.path.join(GOLDEN, "scenes", "cube", "scene.isf"), "-p", str(bad)], capture_output=True)
    assert r.returncode == 2


@pytest.mark.skipif(HAS_GPU, reason="checks the no-device error path")
def test_no_cpu_fallback_without_device():
    path = os.path.join(GOLDEN, "scenes", "cube", "scene.isf")
    with pytest.raises(pt.PtError) as e:
        pt.Scene.load(path, device=0)
    assert e.value.code == -4
    host = pt.Scene.load(path, device=-1)
    with pytest.raises(pt.PtError) as e:
        host.trace_closest(np.zeros((4, 6), np.float32))
    assert e.value.code == -4
    with pytest.raises(pt.PtError):
        pt.Renderer(profile=pt.Profile(width=8, height=8, samples=1)).render(host)
    with pytest.raises(pt.PtError):
        pt.triangle_intersect(np.zeros((1, 6), np.float32), np.zeros((1, 9), np.float32))
    with pytest.raises(pt.PtError) as e:
        pt.debug_render(host, 8, 8)
    assert e.value.code == -4
    with pytest.raises(pt.PtError) as e:
        pt.Renderer(profile=pt.Profile(width=8, height=8, samples=2)).render(host, snapshot_every=1, on_snapshot=lambda a, d: False)
    assert e.value.code == -4
