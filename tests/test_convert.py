"""`path-tracer convert` / pt_convert_gltf (reference src/scene/gltf.rs:146-265): glTF 2.0 -> isf, host code, no GPU.

The reference parses through the un-vendored easy-gltf crate, so there is no golden vector: the expected scene is
computed here from the glTF 2.0 specification (node TRS composition, accessor decoding, material defaults) and the
mapping rules of gltf.rs (texture file families, spot -> point, size 0.1, colour * intensity)."""
import base64
import json
import os
import struct
import subprocess

import numpy as np
import pytest
from PIL import Image

import isf_py
import path_tracer_b200 as pt

EXE = os.path.join(os.path.dirname(pt.__file__), "path-tracer")


def quat_matrix(q):
    x, y, z, w = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def trs(t=(0, 0, 0), q=(0, 0, 0, 1), s=(1, 1, 1)):
    m = np.eye(4)
    m[:3, :3] = quat_matrix(q) @ np.diag(s)
    m[:3, 3] = t
    return m


def build_gltf(tmp_path, glb=False, embed=False):
    """A quad (u16 indices) under a rotated+scaled parent/child pair, an un-indexed triangle with u8-normalised uvs and
    no normals, two materials with PNG textures, a camera and three KHR lights.  Returns (path, expectations)."""
    pos = np.array([[-1, 0, -1], [1, 0, -1], [1, 0, 1], [-1, 0, 1]], np.float32)
    nrm = np.tile(np.array([[0, 1, 0]], np.float32), (4, 1))
    uv = np.array([[0, 0], [1, 0], [1, 1], [0, 1]], np.float32)
    idx = np.array([0, 1, 2, 0, 2, 3], np.uint16)
    tpos = np.array([[0, 0, 0], [2, 0, 0], [0, 2, 0]], np.float32)
    tuv = np.array([[0, 0], [255, 0], [0, 128]], np.uint8)
    chunks, views, accessors = [], [], []

    def add(arr, ctype, atype, normalized=False, stride=None):
        raw = arr.tobytes()
        off = sum(len(c) for c in chunks)
        pad = (-len(raw)) % 4
        chunks.append(raw + b"\0" * pad)
        v = {"buffer": 0, "byteOffset": off, "byteLength": len(raw)}
        if stride:
            v["byteStride"] = stride
        views.append(v)
        a = {"bufferView": len(views) - 1, "componentType": ctype, "count": len(arr), "type": atype}
        if normalized:
            a["normalized"] = True
        accessors.append(a)
        return len(accessors) - 1

    a_pos, a_nrm, a_uv = add(pos, 5126, "VEC3"), add(nrm, 5126, "VEC3"), add(uv, 5126, "VEC2")
    a_idx = add(idx, 5123, "SCALAR")
    a_tpos = add(tpos, 5126, "VEC3")
    tuv_padded = np.zeros((3, 4), np.uint8)
    tuv_padded[:, :2] = tuv
    a_tuv = add(tuv_padded, 5121, "VEC2", normalized=True, stride=4)
    accessors[a_tuv]["count"] = 3
    rng = np.random.default_rng(5)
    base_img = rng.integers(0, 256, (6, 5, 4), dtype=np.uint8)
    mr_img = rng.integers(0, 256, (4, 4, 3), dtype=np.uint8)
    nrm_img = rng.integers(0, 256, (3, 7, 3), dtype=np.uint8)
    Image.fromarray(base_img, "RGBA").save(tmp_path / "base.png")
    Image.fromarray(mr_img, "RGB").save(tmp_path / "mr.png")
    Image.fromarray(nrm_img, "RGB").save(tmp_path / "n.png")
    images = []
    for f in ("base.png", "mr.png", "n.png"):
        if embed or glb:
            raw = (tmp_path / f).read_bytes()
            off = sum(len(c) for c in chunks)
            chunks.append(raw + b"\0" * ((-len(raw)) % 4))
            views.append({"buffer": 0, "byteOffset": off, "byteLength": len(raw)})
            images.append({"bufferView": len(views) - 1, "mimeType": "image/png"})
        else:
            images.append({"uri": f})
    blob = b"".join(chunks)
    q = (0.0, np.sin(0.35), 0.0, np.cos(0.35))
    cam_t = (0.5, 1.5, 6.0)
    doc = {
        "asset": {"version": "2.0"}, "scene": 0,
        "scenes": [{"nodes": [0, 3, 4, 5, 6]}, {"nodes": []}],
        "nodes": [
            {"translation": [1, 2, 3], "rotation": list(q), "scale": [2, 1, 0.5], "children": [1]},
            {"matrix": [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0.25, 0, -1, 1], "mesh": 0, "children": [2]},
            {"mesh": 1, "translation": [0, 1, 0]},
            {"camera": 0, "translation": list(cam_t)},
            {"translation": [4, 5, 6], "extensions": {"KHR_lights_punctual": {"light": 0}}},
            {"rotation": [np.sin(0.5), 0, 0, np.cos(0.5)], "extensions": {"KHR_lights_punctual": {"light": 1}}},
            {"translation": [-1, -2, -3], "extensions": {"KHR_lights_punctual": {"light": 2}}},
        ],
        "meshes": [
            {"primitives": [{"attributes": {"POSITION": a_pos, "NORMAL": a_nrm, "TEXCOORD_0": a_uv}, "indices": a_idx, "material": 0}]},
            {"primitives": [{"attributes": {"POSITION": a_tpos, "TEXCOORD_0": a_tuv}, "material": 1}, {"attributes": {"POSITION": a_tpos}}]},
        ],
        "materials": [
            {"pbrMetallicRoughness": {"baseColorFactor": [0.5, 0.25, 0.125, 0.75], "baseColorTexture": {"index": 0},
                                      "metallicFactor": 0.3, "roughnessFactor": 0.6, "metallicRoughnessTexture": {"index": 1}},
             "normalTexture": {"index": 2}, "emissiveFactor": [0.1, 0.2, 0.3], "emissiveTexture": {"index": 2}},
            {"pbrMetallicRoughness": {"baseColorTexture": {"index": 0}}},
        ],
        "textures": [{"source": 0}, {"source": 1}, {"source": 2}],
        "images": images,
        "cameras": [{"type": "perspective", "perspective": {"yfov": 0.6, "znear": 0.05, "zfar": 500.0, "aspectRatio": 1.5}}],
        "extensions": {"KHR_lights_punctual": {"lights": [
            {"type": "point", "color": [1, 0.5, 0.25], "intensity": 40.0},
            {"type": "directional", "intensity": 2.0},
            {"type": "spot", "color": [0, 1, 0], "spot": {"outerConeAngle": 0.5}}]}},
        "accessors": accessors, "bufferViews": views,
    }
    if glb:
        doc["buffers"] = [{"byteLength": len(blob)}]
        js = json.dumps(doc).encode()
        js += b" " * ((-len(js)) % 4)
        blob_p = blob + b"\0" * ((-len(blob)) % 4)
        total = 12 + 8 + len(js) + 8 + len(blob_p)
        path = tmp_path / "scene.glb"
        path.write_bytes(b"glTF" + struct.pack("<II", 2, total) + struct.pack("<I", len(js)) + b"JSON" + js +
                         struct.pack("<I", len(blob_p)) + b"BIN\0" + blob_p)
    else:
        if embed:
            doc["buffers"] = [{"byteLength": len(blob), "uri": "data:application/octet-stream;base64," + base64.b64encode(blob).decode()}]
        else:
            (tmp_path / "geo data.bin").write_bytes(blob)
            doc["buffers"] = [{"byteLength": len(blob), "uri": "geo%20data.bin"}]
        path = tmp_path / "scene.gltf"
        path.write_text(json.dumps(doc))
    m0 = trs((1, 2, 3), q, (2, 1, 0.5))
    m1 = m0 @ np.array([[1, 0, 0, 0.25], [0, 1, 0, 0], [0, 0, 1, -1], [0, 0, 0, 1]], float)
    m2 = m1 @ trs((0, 1, 0))
    exp = dict(m1=m1, m2=m2, pos=pos, nrm=nrm, uv=uv, idx=idx, tpos=tpos, tuv=tuv.astype(np.float32) / 255, cam=trs(cam_t),
               base=base_img, mr=mr_img, nimg=nrm_img, dir_q=(np.sin(0.5), 0, 0, np.cos(0.5)))
    return path, exp


def check_scene(out_dir, exp):
    raw = json.loads((out_dir / "scene.isf").read_text())
    assert [m["type"] for m in raw["models"]] == ["Mesh"] * 3
    assert raw["background"] == [0.0, 0.0, 0.0]
    # quad: world-space positions, normals as vectors (not re-normalised), uvs, index order
    tri = raw["models"][0]["triangles"]
    assert len(tri) == 2
    P = (exp["m1"] @ np.c_[exp["pos"], np.ones(4)].T).T[:, :3]
    N = (exp["m1"][:3, :3] @ exp["nrm"].T).T
    for t in range(2):
        for k in range(3):
            v = exp["idx"][t * 3 + k]
            assert np.allclose(tri[t][k]["position"], P[v], rtol=1e-5, atol=1e-6)
            assert np.allclose(tri[t][k]["normal"], N[v], rtol=1e-5, atol=1e-6)
            assert np.allclose(tri[t][k]["tex_coords"], exp["uv"][v])
    # un-indexed triangle: normalised u8 uvs, zero normals (missing attribute)
    tri1 = raw["models"][1]["triangles"]
    P1 = (exp["m2"] @ np.c_[exp["tpos"], np.ones(3)].T).T[:, :3]
    assert len(tri1) == 1
    for k in range(3):
        assert np.allclose(tri1[0][k]["position"], P1[k], rtol=1e-5, atol=1e-6)
        assert tri1[0][k]["normal"] == [0.0, 0.0, 0.0]
        assert np.allclose(tri1[0][k]["tex_coords"], exp["tuv"][k], atol=1e-7)
    # materials (gltf.rs:79-130) and texture families (gltf.rs:47-77)
    m0, m1, m2 = (raw["models"][i]["material"] for i in range(3))
    assert np.allclose(m0["albedo"]["factor"], [0.5, 0.25, 0.125]) and m0["albedo"]["texture"] == "albedo_tex_0.png"
    assert m0["opacity"] == {"factor": 0.75, "texture": "alpha_tex_0.png"}
    assert np.isclose(m0["metalness"]["factor"], 0.3) and np.isclose(m0["roughness"]["factor"], 0.6)
    assert {m0["metalness"]["texture"], m0["roughness"]["texture"]} == {"gray_tex_0.png", "gray_tex_1.png"}
    assert m0["normal_texture"] == m0["emissive"]["texture"] == "vec_tex_0.png"  # same image: one file
    assert np.allclose(m0["emissive"]["factor"], [0.1, 0.2, 0.3]) and m0["ior"] == 1.0
    assert m1["albedo"] == {"factor": [1.0, 1.0, 1.0], "texture": "albedo_tex_0.png"}  # de-duplicated image
    assert m1["metalness"] == {"factor": 1.0, "texture": None} and m1["normal_texture"] is None
    assert m2["albedo"] == {"factor": [1.0, 1.0, 1.0], "texture": None} and m2["opacity"] == {"factor": 1.0, "texture": None}
    assert np.array_equal(np.asarray(Image.open(out_dir / "albedo_tex_0.png")), exp["base"][:, :, :3])
    assert np.array_equal(np.asarray(Image.open(out_dir / "alpha_tex_0.png")), exp["base"][:, :, 3])
    assert np.array_equal(np.asarray(Image.open(out_dir / m0["metalness"]["texture"])), exp["mr"][:, :, 2])
    assert np.array_equal(np.asarray(Image.open(out_dir / m0["roughness"]["texture"])), exp["mr"][:, :, 1])
    assert np.array_equal(np.asarray(Image.open(out_dir / "vec_tex_0.png")), exp["nimg"])
    # camera (gltf.rs:206-221): column-major node transform
    cam = raw["camera"]
    assert np.allclose(np.array(cam["transform"]).T, exp["cam"], atol=1e-6)
    assert np.isclose(cam["fov"], 0.6) and np.isclose(cam["znear"], 0.05) and cam["zfar"] == 500.0
    # lights (gltf.rs:236-264)
    l0, l1, l2 = raw["lights"]
    assert l0["type"] == "Point" and np.allclose(l0["position"], [4, 5, 6]) and np.allclose(l0["color"], [40, 20, 10]) and np.isclose(l0["size"], 0.1)
    d = quat_matrix(exp["dir_q"]) @ np.array([0, 0, -1.0])
    assert l1["type"] == "Directional" and np.allclose(l1["direction"], d, atol=1e-6) and np.allclose(l1["color"], [2, 2, 2])
    assert l2["type"] == "Point" and np.allclose(l2["position"], [-1, -2, -3]) and np.allclose(l2["color"], [0, 1, 0])
    # and the result is a loadable isf scene for both loaders
    sa = isf_py.load_isf(str(out_dir / "scene.isf"))
    sc = pt.Scene.load(str(out_dir / "scene.isf"), device=-1)
    got = sc.arrays()
    assert got.tri_pos.shape == (4, 9) and np.array_equal(got.tri_pos, sa.tri_pos) and np.array_equal(got.tri_uv, sa.tri_uv)
    assert len(got.textures) == len(sa.textures) >= 5


@pytest.mark.parametrize("variant", ["gltf", "gltf_embedded", "glb"])
def test_convert_cli(tmp_path, variant):
    path, exp = build_gltf(tmp_path, glb=variant == "glb", embed=variant == "gltf_embedded")
    out = tmp_path / "out" / "nested"
    r = subprocess.run([EXE, "convert", str(path), str(out)], capture_output=True)
    assert r.returncode == 0, r.stderr
    check_scene(out, exp)


def test_convert_api_and_errors(tmp_path):
    path, exp = build_gltf(tmp_path)
    pt.convert_gltf_to_isf(str(path), str(tmp_path / "o2"))
    check_scene(tmp_path / "o2", exp)
    # output exists but is a file (gltf.rs:153-155), missing input, no camera, progressive JPEG: all exit code 2 (main.rs:14-21)
    (tmp_path / "afile").write_text("x")
    doc = json.loads(path.read_text())
    doc["nodes"][3].pop("camera")
    (tmp_path / "nocam.gltf").write_text(json.dumps(doc))
    doc = json.loads(path.read_text())
    Image.fromarray(np.zeros((16, 16, 3), np.uint8), "RGB").save(tmp_path / "prog.jpg", progressive=True)
    doc["images"][0] = {"uri": "prog.jpg"}
    (tmp_path / "jpeg.gltf").write_text(json.dumps(doc))
    for args, needle in (([str(path), str(tmp_path / "afile")], b"is not a directory"),
                         ([str(tmp_path / "missing.gltf"), str(tmp_path / "o3")], b"cannot open"),
                         ([str(tmp_path / "nocam.gltf"), str(tmp_path / "o4")], b"No camera found"),
                         ([str(tmp_path / "jpeg.gltf"), str(tmp_path / "o5")], b"progressive"),
                         ([str(path)], b"required arguments")):
        r = subprocess.run([EXE, "convert"] + args, capture_output=True)
        assert r.returncode == 2 and needle in r.stderr, (args, r.stderr)
    with pytest.raises(pt.PtError):
        pt.convert_gltf_to_isf(str(tmp_path / "missing.gltf"), str(tmp_path / "o6"))


def smooth_image(h, w, seed):
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w]
    img = np.stack([127 + 120 * np.sin(x / 9.0 + seed), 127 + 120 * np.cos(y / 7.0), 127 + 100 * np.sin((x + y) / 11.0)], -1)
    return np.clip(img + rng.normal(0, 4, img.shape), 0, 255).astype(np.uint8)


@pytest.mark.parametrize("subsampling,quality,size", [(0, 95, (40, 56)), (1, 90, (33, 47)), (2, 85, (50, 37)), (2, 60, (16, 16))])
def test_jpeg_decode_matches_pil(tmp_path, subsampling, quality, size):
    """Baseline JPEG textures (image::open accepts them in texture_bank.rs:33,49): decoded pixels agree with libjpeg's
    to within IDCT / chroma-upsampling differences."""
    h, w = size
    img = smooth_image(h, w, subsampling + quality)
    p = tmp_path / "t.jpg"
    Image.fromarray(img, "RGB").save(p, quality=quality, subsampling=subsampling)
    got = pt.read_png(str(p), 3).astype(np.int16)
    want = np.asarray(Image.open(p).convert("RGB")).astype(np.int16)
    assert got.shape == want.shape
    d = np.abs(got - want)
    if subsampling == 0:
        assert d.max() <= 2 and d.mean() < 0.5, (d.max(), d.mean())
    else:  # linear vs libjpeg's integer "fancy" chroma upsampling
        assert d.mean() < 1.0 and d.max() <= 6, (d.mean(), d.max())
    gray = pt.read_png(str(p), 1)[:, :, 0].astype(np.int16)
    g32 = got.astype(np.int64)
    luma = (2126 * g32[:, :, 0] + 7152 * g32[:, :, 1] + 722 * g32[:, :, 2]) // 10000
    assert np.array_equal(gray, luma)


def test_jpeg_gray_restart_and_errors(tmp_path):
    g = smooth_image(45, 61, 3)[:, :, 0]
    p = tmp_path / "g.jpg"
    Image.fromarray(g, "L").save(p, quality=92)
    got = pt.read_png(str(p), 1)[:, :, 0].astype(np.int16)
    want = np.asarray(Image.open(p)).astype(np.int16)
    assert np.abs(got - want).max() <= 2
    rgb = pt.read_png(str(p), 3)
    assert (rgb[:, :, 0] == rgb[:, :, 1]).all() and (rgb[:, :, 1] == rgb[:, :, 2]).all()
    # restart intervals (DRI/RSTn): same pixels as the file without them
    img = smooth_image(48, 64, 9)
    a, b = tmp_path / "a.jpg", tmp_path / "b.jpg"
    Image.fromarray(img, "RGB").save(a, quality=90, subsampling=2)
    Image.fromarray(img, "RGB").save(b, quality=90, subsampling=2, restart_marker_blocks=2)
    assert b.read_bytes().count(b"\xff\xdd") >= 1
    assert np.array_equal(pt.read_png(str(a), 3), pt.read_png(str(b), 3))
    prog = tmp_path / "p.jpg"
    Image.fromarray(img, "RGB").save(prog, progressive=True)
    with pytest.raises(pt.PtError) as e:
        pt.read_png(str(prog), 3)
    assert "progressive" in str(e.value)
    trunc = tmp_path / "t.jpg"
    trunc.write_bytes(a.read_bytes()[:40])
    with pytest.raises(pt.PtError):
        pt.read_png(str(trunc), 3)


def test_convert_with_jpeg_texture(tmp_path):
    path, exp = build_gltf(tmp_path)
    doc = json.loads(path.read_text())
    Image.fromarray(smooth_image(24, 32, 1), "RGB").save(tmp_path / "base.jpg", quality=95, subsampling=0)
    doc["images"][0] = {"uri": "base.jpg"}
    (tmp_path / "j.gltf").write_text(json.dumps(doc))
    out = tmp_path / "oj"
    pt.convert_gltf_to_isf(str(tmp_path / "j.gltf"), str(out))
    alb = np.asarray(Image.open(out / "albedo_tex_0.png")).astype(np.int16)
    want = np.asarray(Image.open(tmp_path / "base.jpg").convert("RGB")).astype(np.int16)
    assert alb.shape == want.shape and np.abs(alb - want).max() <= 3 and np.abs(alb - want).mean() < 0.5
    assert (np.asarray(Image.open(out / "alpha_tex_0.png")) == 255).all()  # a JPEG has no alpha: opaque
