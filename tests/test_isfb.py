"""The binary side-car of an isf scene (csrc/isf_binary.cpp): an independent Python reader of the documented layout
agrees with the scene the JSON loader produced, a stored BVH is checked before use, damaged files are rejected, and
(GPU) a scene loaded from the side-car traces and renders exactly like the JSON one without building a tree."""
import os
import struct
import subprocess

import numpy as np
import pytest

from conftest import GOLDEN, ROOT

import path_tracer_b200 as pt

SCENES = sorted(os.listdir(os.path.join(GOLDEN, "scenes")))
EXE = os.path.join(ROOT, "path-tracer_b200", "path-tracer")
HEADER = struct.Struct("<4sIIIQIIQQIIiI")  # 64 bytes, isf_binary.cpp


def align16(n):
    return (n + 15) & ~15


def read_isfb(path):
    """Independent reader of the layout documented at the top of isf_binary.cpp."""
    b = open(path, "rb").read()
    assert HEADER.size == 64
    (magic, version, flags, n_models, n_tris, n_tex, n_lights, n_nodes, n_leaf, kind, rev, depth, root) = HEADER.unpack_from(b, 0)
    assert magic == b"ISFB" and version == 1
    off = 64
    view = np.frombuffer(b, "<f4", 20, off)
    off = align16(off + 80)
    out = dict(flags=flags, kind=kind, rev=rev, depth=depth, camera=view[:16], fov=view[16], background=view[17:20])
    model_dt = np.dtype([("type", "<i4"), ("first_tri", "<u4"), ("n_tris", "<u4"), ("center", "<f4", 3), ("radius", "<f4")])
    mat_dt = np.dtype([("albedo", "<f4", 3), ("emissive", "<f4", 3), ("opacity", "<f4"), ("metalness", "<f4"),
                       ("roughness", "<f4"), ("ior", "<f4"), ("tex", "<i4", 6)])
    light_dt = np.dtype([("type", "<i4"), ("vec", "<f4", 3), ("color", "<f4", 3), ("size", "<f4")])
    assert model_dt.itemsize == 28 and mat_dt.itemsize == 64 and light_dt.itemsize == 32
    out["models"] = np.frombuffer(b, model_dt, n_models, off)
    off = align16(off + 28 * n_models)
    out["materials"] = np.frombuffer(b, mat_dt, n_models, off)
    off = align16(off + 64 * n_models)
    out["lights"] = np.frombuffer(b, light_dt, n_lights, off)
    off = align16(off + 32 * n_lights)
    for key, per in (("pos", 9), ("nrm", 9), ("uv", 6)):
        out[key] = np.frombuffer(b, "<f4", n_tris * per, off).reshape(-1, per)
        off = align16(off + 4 * n_tris * per)
    out["textures"] = []
    for _ in range(n_tex):
        w, h, ch, zero = struct.unpack_from("<IIII", b, off)
        assert zero == 0
        off += 16
        out["textures"].append(np.frombuffer(b, np.uint8, w * h * ch, off).reshape(h, w, ch))
        off = align16(off + w * h * ch)
    out["accel_offset"] = off
    if flags & 1:
        out["nodes"] = np.frombuffer(b, "<u4", n_nodes * 32, off).reshape(-1, 32)
        off += 128 * n_nodes
        out["leaf_tris"] = np.frombuffer(b, "<u4", n_leaf, off)
        off = align16(off + 4 * n_leaf)
    assert off == len(b), (off, len(b))
    return out


def same_scene(a, b):
    assert np.array_equal(a.tri_pos, b.tri_pos) and np.array_equal(a.tri_nrm, b.tri_nrm) and np.array_equal(a.tri_uv, b.tri_uv)
    assert a.models == b.models and a.materials == b.materials and a.lights == b.lights
    assert len(a.textures) == len(b.textures)
    for x, y in zip(a.textures, b.textures):
        assert x.shape == y.shape and np.array_equal(x, y)
    assert np.array_equal(a.camera, b.camera) and a.fov == b.fov and np.array_equal(a.background, b.background)


@pytest.mark.parametrize("scene", SCENES)
def test_pack_round_trip(tmp_path, scene):
    src = os.path.join(GOLDEN, "scenes", scene, "scene.isf")
    host = pt.Scene.load(src, device=-1)
    out = str(tmp_path / "scene.isfb")
    host.save_binary(out, with_accel=True)
    ref = host.arrays()
    f = read_isfb(out)
    # the file holds what the JSON loader produced, bit for bit
    assert np.array_equal(f["pos"], ref.tri_pos) and np.array_equal(f["nrm"], ref.tri_nrm) and np.array_equal(f["uv"], ref.tri_uv)
    assert np.array_equal(f["camera"], ref.camera) and f["fov"] == ref.fov and np.array_equal(f["background"], ref.background)
    assert [int(m["type"]) for m in f["models"]] == [m["type"] for m in ref.models]
    assert [int(m["n_tris"]) for m in f["models"]] == [m["n_tris"] for m in ref.models]
    assert [float(m["roughness"]) for m in f["materials"]] == [np.float32(m["roughness"]) for m in ref.materials]
    assert len(f["textures"]) == len(ref.textures)
    for x, y in zip(f["textures"], ref.textures):
        assert np.array_equal(x, y)
    # the stored tree: leaves are a permutation of the triangles, the root is node 0
    assert f["flags"] & 1 and f["kind"] == 0 and f["depth"] >= 1
    assert sorted(f["leaf_tris"].tolist()) == list(range(len(ref.tri_pos)))
    # reading it back gives the same scene, and re-saving it the same bytes (the stored tree is adopted, not rebuilt)
    back = pt.Scene.load(out, device=-1)
    same_scene(back.arrays(), ref)
    assert back.info()["n_nodes"] == len(f["nodes"])
    out2 = str(tmp_path / "again.isfb")
    back.save_binary(out2, with_accel=True)
    assert open(out, "rb").read() == open(out2, "rb").read()
    # without the tree
    bare = str(tmp_path / "bare.isfb")
    host.save_binary(bare, with_accel=False)
    g = read_isfb(bare)
    assert g["flags"] == 0 and "nodes" not in g
    assert pt.Scene.load(bare, device=-1).info()["n_nodes"] == 0


def test_damaged_files_are_rejected(tmp_path):
    src = os.path.join(GOLDEN, "scenes", "cube", "scene.isf")
    good = str(tmp_path / "cube.isfb")
    pt.Scene.load(src, device=-1).save_binary(good)
    raw = bytearray(open(good, "rb").read())
    f = read_isfb(good)

    def load(mutated):
        p = str(tmp_path / "bad.isfb")
        open(p, "wb").write(bytes(mutated))
        return pt.Scene.load(p, device=-1)

    for cut in (10, 64, 200, len(raw) // 2, len(raw) - 1):  # truncation anywhere
        with pytest.raises(pt.PtError):
            load(raw[:cut])
    bad = bytearray(raw)
    struct.pack_into("<I", bad, 4, 99)  # version
    with pytest.raises(pt.PtError, match="version"):
        load(bad)
    bad = bytearray(raw)
    struct.pack_into("<Q", bad, 16, 1 << 40)  # triangle count far beyond the file
    with pytest.raises(pt.PtError, match="truncated"):
        load(bad)
    # a tree that points outside the node array, shares a child, loses a triangle or has a box that is too small
    a = f["accel_offset"]
    child0 = a + 96  # node 0, child[0]
    bad = bytearray(raw)
    struct.pack_into("<I", bad, child0, 0x00FFFFFF)
    with pytest.raises(pt.PtError, match="rejected"):
        load(bad)
    bad = bytearray(raw)
    c0, c1 = struct.unpack_from("<II", raw, child0)
    struct.pack_into("<II", bad, child0, c0, c0)
    with pytest.raises(pt.PtError, match="rejected"):
        load(bad)
    bad = bytearray(raw)
    leaf_off = a + 128 * len(f["nodes"])
    struct.pack_into("<I", bad, leaf_off, int(f["leaf_tris"][1]))  # duplicate triangle in the leaf order
    with pytest.raises(pt.PtError, match="permutation"):
        load(bad)
    bad = bytearray(raw)
    hi_x0 = a + 16  # node 0: lo_x[4], hi_x[4], ...
    lo_x0 = struct.unpack_from("<f", raw, a)[0]
    struct.pack_into("<f", bad, hi_x0, lo_x0)  # squash child 0's box along x
    with pytest.raises(pt.PtError, match="rejected"):
        load(bad)
    # a tree written under other margin rules is ignored, not an error: the scene loads and is rebuilt when needed
    bad = bytearray(raw)
    struct.pack_into("<I", bad, 52, 1)  # accel_rev
    sc = load(bad)
    assert sc.info()["n_nodes"] == 0 and sc.info()["n_tris"] == 12


def test_cli_pack_and_render_args(tmp_path):
    src = os.path.join(GOLDEN, "scenes", "spheres", "scene.isf")
    out = str(tmp_path / "s.isfb")
    r = subprocess.run([EXE, "pack", src, out], capture_output=True)
    assert r.returncode == 0, r.stderr
    f = read_isfb(out)
    assert f["flags"] & 1
    r = subprocess.run([EXE, "pack", src, str(tmp_path / "n.isfb"), "--no-accel"], capture_output=True)
    assert r.returncode == 0 and read_isfb(str(tmp_path / "n.isfb"))["flags"] == 0
    assert subprocess.run([EXE, "pack", src], capture_output=True).returncode == 2
    assert subprocess.run([EXE, "pack", str(tmp_path / "missing.isf"), out], capture_output=True).returncode == 2


# ---------------------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
@pytest.mark.parametrize("scene", ["head", "spheres", "alpha_transparency"])
def test_isfb_scene_traces_like_json_scene(tmp_path, scene):
    import rays as raygen
    src = os.path.join(GOLDEN, "scenes", scene, "scene.isf")
    a = pt.Scene.load(src)
    out = str(tmp_path / "scene.isfb")
    a.save_binary(out)
    b = pt.Scene.load(out)
    assert b.accel_info()["build_ms"] == 0.0 and b.info()["n_nodes"] == a.info()["n_nodes"]
    rays = raygen.random_rays(a.arrays(), 1 << 16, seed=5)
    ha, hb = a.trace_closest(rays), b.trace_closest(rays)
    assert ha.tobytes() == hb.tobytes()
    assert (ha["prim_id"] >= 0).any()
    assert np.array_equal(a.trace_transmittance(rays), b.trace_transmittance(rays))
    prof = pt.Profile(width=96, height=64, samples=4, bounces=3)
    ra = pt.Renderer(profile=prof).render(a)
    rb = pt.Renderer(profile=prof).render(b)
    assert np.array_equal(ra, rb)


@pytest.mark.gpu
@pytest.mark.parametrize("builder", ["lbvh", "ploc"])
def test_isfb_keeps_an_lbvh_tree(tmp_path, builder):
    pt.set_accel_builder(builder)
    try:
        a = pt.Scene.synthetic("console", 60000, seed=3)
    finally:
        pt.set_accel_builder("sah")
    assert a.accel_info()["accel"] == builder
    out = str(tmp_path / "console.isfb")
    a.save_binary(out)
    b = pt.Scene.load(out)
    assert b.accel_info()["accel"] == builder and b.accel_info()["build_ms"] == 0.0
    assert b.accel_info()["max_depth"] == a.accel_info()["max_depth"]
    prof = pt.Profile(width=128, height=72, samples=2, bounces=3)
    assert np.array_equal(pt.Renderer(profile=prof).render(a), pt.Renderer(profile=prof).render(b))
