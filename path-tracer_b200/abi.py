"""ctypes mirror of include/pt_b200.h (structs only; no compute).

Also holds `SceneArrays`, the flat numpy form of an isf scene (scene/isf.rs:5-142 of the
reference) from which a `pt_scene_desc` is built for `pt_scene_create`.
"""
import ctypes as C

import numpy as np

PT_MODEL_MESH, PT_MODEL_SPHERE = 0, 1
PT_LIGHT_POINT, PT_LIGHT_DIRECTIONAL = 0, 1
PT_TONEMAP_REINHARD, PT_TONEMAP_FILMIC, PT_TONEMAP_ACES = 0, 1, 2
TONEMAPS = {"REINHARD": 0, "FILMIC": 1, "ACES": 2}
PT_BRDF_COOK_TORRANCE = 0


class pt_material(C.Structure):
    _fields_ = [("albedo", C.c_float * 3), ("emissive", C.c_float * 3), ("opacity", C.c_float),
                ("metalness", C.c_float), ("roughness", C.c_float), ("ior", C.c_float),
                ("albedo_tex", C.c_int32), ("emissive_tex", C.c_int32), ("opacity_tex", C.c_int32),
                ("metalness_tex", C.c_int32), ("roughness_tex", C.c_int32), ("normal_tex", C.c_int32)]


class pt_model(C.Structure):
    _fields_ = [("type", C.c_int32), ("first_tri", C.c_uint32), ("n_tris", C.c_uint32),
                ("center", C.c_float * 3), ("radius", C.c_float)]


class pt_light(C.Structure):
    _fields_ = [("type", C.c_int32), ("vec", C.c_float * 3), ("color", C.c_float * 3), ("size", C.c_float)]


class pt_texture(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("channels", C.c_uint32),
                ("data", C.POINTER(C.c_uint8))]


class pt_scene_desc(C.Structure):
    _fields_ = [("n_models", C.c_uint32), ("models", C.POINTER(pt_model)), ("materials", C.POINTER(pt_material)),
                ("n_tris", C.c_uint64), ("tri_pos", C.POINTER(C.c_float)), ("tri_nrm", C.POINTER(C.c_float)),
                ("tri_uv", C.POINTER(C.c_float)), ("n_textures", C.c_uint32), ("textures", C.POINTER(pt_texture)),
                ("n_lights", C.c_uint32), ("lights", C.POINTER(pt_light)), ("camera", C.c_float * 16),
                ("fov", C.c_float), ("background", C.c_float * 3)]


class pt_profile(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("bounces", C.c_uint32), ("samples", C.c_uint32),
                ("brdf", C.c_int32), ("tonemap", C.c_int32)]


class pt_hit(C.Structure):
    _fields_ = [("prim_id", C.c_int32), ("t", C.c_float), ("u", C.c_float), ("v", C.c_float)]


HIT_DTYPE = np.dtype([("prim_id", "<i4"), ("t", "<f4"), ("u", "<f4"), ("v", "<f4")])


class pt_path_vertex(C.Structure):
    _fields_ = [("ray_o", C.c_float * 3), ("ray_d", C.c_float * 3), ("prim_id", C.c_int32), ("t", C.c_float),
                ("throughput", C.c_float * 3), ("radiance", C.c_float * 3)]


PATH_VERTEX_DTYPE = np.dtype([("ray_o", "<f4", 3), ("ray_d", "<f4", 3), ("prim_id", "<i4"), ("t", "<f4"),
                              ("throughput", "<f4", 3), ("radiance", "<f4", 3)])


class pt_stats(C.Structure):
    _fields_ = [("samples", C.c_uint64), ("extend_rays", C.c_uint64), ("shadow_rays", C.c_uint64),
                ("alpha_recasts", C.c_uint64), ("kernel_launches", C.c_uint64), ("ms_total", C.c_float),
                ("ms_extend", C.c_float), ("ms_shade", C.c_float), ("ms_shadow", C.c_float), ("ms_other", C.c_float),
                ("extend_launches", C.c_uint64), ("shade_launches", C.c_uint64), ("shadow_launches", C.c_uint64),
                ("alpha_tests", C.c_uint64), ("shadow_dropped", C.c_uint64)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


class SceneArrays:
    """Flat scene: lists of model/material/light dicts + float32 triangle arrays + uint8 textures."""

    def __init__(self):
        self.models = []      # dict(type, first_tri, n_tris, center, radius)
        self.materials = []   # dict(albedo, emissive, opacity, metalness, roughness, ior, *_tex)
        self.tri_pos = np.zeros((0, 9), np.float32)
        self.tri_nrm = np.zeros((0, 9), np.float32)
        self.tri_uv = np.zeros((0, 6), np.float32)
        self.textures = []    # np.uint8 arrays (H, W, channels)
        self.lights = []      # dict(type, vec, color, size)
        self.camera = np.eye(4, dtype=np.float32).reshape(16)  # column-major
        self.fov = np.float32(0.4)
        self.background = np.zeros(3, np.float32)
        self._keep = []

    def n_prims(self):
        return sum(2 if m["type"] == PT_MODEL_SPHERE else m["n_tris"] for m in self.models)

    def to_desc(self):
        """Build a pt_scene_desc whose pointers stay valid while `self` is alive."""
        d = pt_scene_desc()
        n = len(self.models)
        models = (pt_model * max(n, 1))()
        mats = (pt_material * max(n, 1))()
        for i, (m, ma) in enumerate(zip(self.models, self.materials)):
            models[i].type = m["type"]
            models[i].first_tri = m.get("first_tri", 0)
            models[i].n_tris = m.get("n_tris", 0)
            models[i].center = (C.c_float * 3)(*[float(x) for x in m.get("center", (0, 0, 0))])
            models[i].radius = float(m.get("radius", 0.0))
            mats[i].albedo = (C.c_float * 3)(*[float(x) for x in ma["albedo"]])
            mats[i].emissive = (C.c_float * 3)(*[float(x) for x in ma["emissive"]])
            for k in ("opacity", "metalness", "roughness", "ior"):
                setattr(mats[i], k, float(ma[k]))
            for k in ("albedo_tex", "emissive_tex", "opacity_tex", "metalness_tex", "roughness_tex", "normal_tex"):
                setattr(mats[i], k, int(ma.get(k, -1)))
        pos = np.ascontiguousarray(self.tri_pos, np.float32).reshape(-1, 9)
        nrm = np.ascontiguousarray(self.tri_nrm, np.float32).reshape(-1, 9)
        uv = np.ascontiguousarray(self.tri_uv, np.float32).reshape(-1, 6)
        texs = (pt_texture * max(len(self.textures), 1))()
        tex_keep = []
        for i, t in enumerate(self.textures):
            t = np.ascontiguousarray(t, np.uint8)
            if t.ndim == 2:
                t = t[:, :, None]
            tex_keep.append(t)
            texs[i].height, texs[i].width, texs[i].channels = t.shape
            texs[i].data = t.ctypes.data_as(C.POINTER(C.c_uint8))
        lights = (pt_light * max(len(self.lights), 1))()
        for i, l in enumerate(self.lights):
            lights[i].type = l["type"]
            lights[i].vec = (C.c_float * 3)(*[float(x) for x in l["vec"]])
            lights[i].color = (C.c_float * 3)(*[float(x) for x in l["color"]])
            lights[i].size = float(l.get("size", 0.0))
        d.n_models, d.models, d.materials = n, models, mats
        d.n_tris = pos.shape[0]
        d.tri_pos, d.tri_nrm, d.tri_uv = _fp(pos), _fp(nrm), _fp(uv)
        d.n_textures, d.textures = len(self.textures), texs
        d.n_lights, d.lights = len(self.lights), lights
        d.camera = (C.c_float * 16)(*[float(x) for x in np.asarray(self.camera, np.float32).reshape(16)])
        d.fov = float(self.fov)
        d.background = (C.c_float * 3)(*[float(x) for x in self.background])
        self._keep = [models, mats, pos, nrm, uv, texs, tex_keep, lights]
        return d
