"""Independent (Python json + Pillow) isf loader used by the tests.

Follows scene/isf.rs:5-142 (schema + serde defaults) and scene/internal/texture_bank.rs:21-51
(into_rgb8 / into_luma8) of the reference.  It deliberately shares no code with the product's
C++ loader (path-tracer_b200/csrc/isf_loader.cpp) so the two can be checked against each other.
"""
import json
import os

import numpy as np

from path_tracer_b200.abi import (PT_LIGHT_DIRECTIONAL, PT_LIGHT_POINT, PT_MODEL_MESH, PT_MODEL_SPHERE, SceneArrays)


def _material(mat, root, tex_cache, textures):
    def tex(path, channels):
        if path is None:
            return -1
        from PIL import Image
        full = os.path.realpath(os.path.join(root, path))
        key = (full, channels)
        if key not in tex_cache:
            img = Image.open(full)
            if channels == 3:
                arr = np.asarray(img.convert("RGB"), np.uint8)
            else:
                if img.mode not in ("L",):
                    raise NotImplementedError("luma conversion of non-gray textures")
                arr = np.asarray(img, np.uint8)[:, :, None]
            tex_cache[key] = len(textures)
            textures.append(np.ascontiguousarray(arr))
        return tex_cache[key]

    out = {}
    alb = mat["albedo"]  # required (isf.rs:81)
    out["albedo"] = alb.get("factor", [1.0, 1.0, 1.0])
    out["albedo_tex"] = tex(alb.get("texture"), 3)
    em = mat.get("emissive")
    if em is None:  # derived Default: factor = [0,0,0]
        out["emissive"], out["emissive_tex"] = [0.0, 0.0, 0.0], -1
    else:
        out["emissive"] = em.get("factor", [1.0, 1.0, 1.0])
        out["emissive_tex"] = tex(em.get("texture"), 3)
    op = mat.get("opacity") or {}
    out["opacity"] = op.get("factor", 1.0)
    out["opacity_tex"] = tex(op.get("texture"), 1)
    me = mat.get("metalness")
    if me is None:  # derived Default: 0.0
        out["metalness"], out["metalness_tex"] = 0.0, -1
    else:
        out["metalness"] = me.get("factor", 1.0)
        out["metalness_tex"] = tex(me.get("texture"), 1)
    ro = mat.get("roughness") or {}
    out["roughness"] = ro.get("factor", 1.0)
    out["roughness_tex"] = tex(ro.get("texture"), 1)
    out["ior"] = mat.get("ior", 1.0)
    out["normal_tex"] = tex(mat.get("normal_texture"), 3)
    return out


def load_isf(path):
    with open(path) as f:
        doc = json.load(f)
    root = os.path.dirname(os.path.abspath(path))
    sa = SceneArrays()
    pos, nrm, uv = [], [], []
    tex_cache = {}
    ntri = 0
    for m in doc["models"]:
        if m["type"] == "Sphere":
            sa.models.append(dict(type=PT_MODEL_SPHERE, first_tri=0, n_tris=0, center=m["center"], radius=m["radius"]))
        elif m["type"] == "Mesh":
            tris = m["triangles"]
            for t in tris:
                pos.append([c for v in t for c in v["position"]])
                nrm.append([c for v in t for c in v["normal"]])
                uv.append([c for v in t for c in v["tex_coords"]])
            sa.models.append(dict(type=PT_MODEL_MESH, first_tri=ntri, n_tris=len(tris), center=[0, 0, 0], radius=0.0))
            ntri += len(tris)
        else:
            raise ValueError("unknown model type " + str(m["type"]))
        sa.materials.append(_material(m["material"], root, tex_cache, sa.textures))
    sa.tri_pos = np.asarray(pos, np.float64).astype(np.float32).reshape(-1, 9)
    sa.tri_nrm = np.asarray(nrm, np.float64).astype(np.float32).reshape(-1, 9)
    sa.tri_uv = np.asarray(uv, np.float64).astype(np.float32).reshape(-1, 6)
    for l in doc["lights"]:
        if l["type"] == "Point":
            sa.lights.append(dict(type=PT_LIGHT_POINT, vec=l["position"], color=l["color"], size=l["size"]))
        else:
            sa.lights.append(dict(type=PT_LIGHT_DIRECTIONAL, vec=l["direction"], color=l["color"], size=0.0))
    cam = doc["camera"]
    sa.camera = np.asarray(cam["transform"], np.float64).astype(np.float32).reshape(16)  # [c][r] -> column-major
    sa.fov = np.float32(cam["fov"])
    sa.background = np.asarray(doc["background"], np.float64).astype(np.float32)
    return sa
