"""Host-side mirror of the reference's render interface over the C ABI (include/pt_b200.h).

Reference surface mirrored (file:line relative to the reference root):
  Profile / Profile::load        src/config/profile.rs:10-40, src/config/resolution.rs:4-16
  load_internal(path) -> Scene   src/scene/mod.rs:16-22
  Renderer::new / render         src/renderer/mod.rs:61-73, 76-169
  ray_cast                       src/renderer/utils.rs:11-21 (first hit of the sorted list)
There is no CPU fallback: every compute call goes through libpt_b200.so and raises PtError when the
library or a CUDA device is missing.
"""
import ctypes as C
import os

import numpy as np

from .abi import (HIT_DTYPE, PATH_VERTEX_DTYPE, TONEMAPS, SceneArrays, pt_hit, pt_profile, pt_scene_desc, pt_stats)

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

EXPORTS = ["pt_last_error", "pt_device_count", "pt_profile_default", "pt_profile_load_yaml", "pt_scene_load_isf",
           "pt_scene_create", "pt_scene_synthetic", "pt_scene_destroy", "pt_set_accel_builder", "pt_scene_accel_info", "pt_scene_upload", "pt_scene_upload_targets", "pt_scene_upload_commit", "pt_scene_upload_part", "pt_set_kernel_timing", "pt_scene_get_desc", "pt_scene_info",
           "pt_trace_closest", "pt_trace_closest_device", "pt_trace_closest_device_noalloc", "pt_trace_accepted", "pt_trace_shadow", "pt_trace_transmittance",
           "pt_triangle_intersect", "pt_philox4x32_10",
           "pt_trace_paths", "pt_render_accum_device", "pt_aov_name", "pt_debug_textures", "pt_render_progressive", "pt_convert_gltf", "pt_scene_save_binary",
           "pt_tonemap_device", "pt_render", "pt_render_async", "pt_stream_synchronize", "pt_scene_replicate", "pt_render_multi", "pt_write_png", "pt_read_png", "pt_free"]


SNAPSHOT_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.POINTER(C.c_uint8), C.c_uint32, C.c_uint32, C.c_uint32)
AOV_NAMES = ("normal", "albedo", "opacity", "metalness", "roughness", "emissive", "ior")  # debug_renderer.rs:80-102


class PtError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("pt_b200 error %d: %s" % (code, msg))
        self.code = code


def lib():
    """dlopen path-tracer_b200/libpt_b200.so (built in-tree by csrc/Makefile); fails loudly when missing."""
    global _LIB
    if _LIB is None:
        so = os.environ.get("PT_B200_LIB") or os.path.join(_HERE, "libpt_b200.so")  # override: A/B builds only
        if not os.path.exists(so):
            raise PtError(-4, "libpt_b200.so not built: run `python __graft_entry__.py` (no CPU fallback exists)")
        L = C.CDLL(so)
        for name in EXPORTS:
            getattr(L, name)  # AttributeError if a declared symbol is missing
        L.pt_last_error.restype = C.c_char_p
        vp, u32, u64 = C.c_void_p, C.c_uint32, C.c_uint64
        L.pt_profile_default.argtypes = [C.POINTER(pt_profile)]
        L.pt_profile_load_yaml.argtypes = [C.c_char_p, C.POINTER(pt_profile)]
        L.pt_scene_load_isf.argtypes = [C.c_char_p, C.c_int, C.POINTER(vp)]
        L.pt_scene_create.argtypes = [C.POINTER(pt_scene_desc), C.c_int, C.POINTER(vp)]
        L.pt_scene_synthetic.argtypes = [C.c_char_p, u64, u64, C.c_int, C.POINTER(vp)]
        L.pt_scene_destroy.argtypes = [vp]
        L.pt_scene_destroy.restype = None
        L.pt_scene_upload.argtypes = [vp, vp, C.POINTER(u64)]
        L.pt_scene_upload_targets.argtypes = [vp, vp, C.POINTER(vp), C.POINTER(u64), u32, C.POINTER(u32)]
        L.pt_scene_upload_commit.argtypes = [vp, vp]
        L.pt_scene_upload_part.argtypes = [vp, u32, u32, C.POINTER(u64)]
        L.pt_set_kernel_timing.argtypes = [C.c_int]
        L.pt_set_kernel_timing.restype = None
        L.pt_scene_get_desc.argtypes = [vp, C.POINTER(pt_scene_desc)]
        L.pt_scene_info.argtypes = [vp, C.POINTER(u64), C.POINTER(u64), C.POINTER(u64), C.POINTER(u64)]
        L.pt_trace_closest.argtypes = [vp, vp, C.c_size_t, vp]
        L.pt_trace_closest_device.argtypes = [vp, vp, C.c_size_t, vp, vp]
        L.pt_trace_closest_device_noalloc.argtypes = [vp, vp, C.c_size_t, vp, vp]
        L.pt_trace_transmittance.argtypes = [vp, vp, C.c_size_t, vp]
        L.pt_trace_accepted.argtypes = [vp, vp, C.c_size_t, u32, u32, u64, vp]
        L.pt_trace_shadow.argtypes = [vp, vp, vp, vp, C.c_size_t, vp]
        L.pt_triangle_intersect.argtypes = [vp, C.c_int, vp, vp, C.c_size_t, vp]
        L.pt_philox4x32_10.argtypes = [C.c_int, vp, vp, C.c_size_t, vp]
        L.pt_trace_paths.argtypes = [vp, C.POINTER(pt_profile), vp, C.c_size_t, u32, u64, vp, vp]
        L.pt_render_accum_device.argtypes = [vp, C.POINTER(pt_profile), u32, u32, u32, u64, vp, vp, C.POINTER(pt_stats)]
        L.pt_tonemap_device.argtypes = [vp, u32, u32, u32, C.c_int32, vp, C.c_int, vp]
        L.pt_render.argtypes = [vp, C.POINTER(pt_profile), u64, vp, vp, C.POINTER(pt_stats)]
        L.pt_render_async.argtypes = [vp, C.POINTER(pt_profile), u64, vp, vp]
        L.pt_stream_synchronize.argtypes = [C.c_int, vp]
        L.pt_scene_replicate.argtypes = [vp, C.c_int, C.POINTER(vp)]
        L.pt_render_multi.argtypes = [C.POINTER(vp), u32, C.POINTER(pt_profile), u64, vp, vp, C.POINTER(pt_stats)]
        L.pt_write_png.argtypes = [C.c_char_p, vp, u32, u32]
        L.pt_read_png.argtypes = [C.c_char_p, u32, C.POINTER(u32), C.POINTER(u32), C.POINTER(C.POINTER(C.c_uint8))]
        L.pt_set_accel_builder.argtypes = [C.c_int]
        L.pt_scene_accel_info.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_double), C.POINTER(C.c_int)]
        L.pt_convert_gltf.argtypes = [C.c_char_p, C.c_char_p]
        L.pt_scene_save_binary.argtypes = [vp, C.c_char_p, C.c_int]
        L.pt_aov_name.argtypes = [C.c_int]
        L.pt_aov_name.restype = C.c_char_p
        L.pt_debug_textures.argtypes = [vp, u32, u32, vp, C.POINTER(u64)]
        L.pt_render_progressive.argtypes = [vp, C.POINTER(pt_profile), u64, u32, SNAPSHOT_FN, vp, vp, vp, C.POINTER(pt_stats)]
        L.pt_free.argtypes = [vp]
        L.pt_free.restype = None
        _LIB = L
    return _LIB


def _check(rc):
    if rc != 0:
        raise PtError(rc, (lib().pt_last_error() or b"").decode("utf-8", "replace"))


def device_count():
    return lib().pt_device_count()


ACCEL_SAH, ACCEL_LBVH, ACCEL_AUTO, ACCEL_PLOC = 0, 1, 2, 3
TRACE_NOALLOC_MAX_RAYS = 65536  # PT_TRACE_NOALLOC_MAX_RAYS of include/pt_b200.h


def set_accel_builder(accel):
    """Acceleration-structure builder of the scenes created from now on: "sah" (host binned SAH, default), "lbvh"
    (GPU Morton/radix-sort LBVH), "ploc" (GPU Morton order + nearest-neighbour agglomeration) or "auto".  Replaces KDTree::build (scene/internal/model.rs:96, internal/mod.rs:42)."""
    if isinstance(accel, str):
        accel = {"sah": ACCEL_SAH, "lbvh": ACCEL_LBVH, "auto": ACCEL_AUTO, "ploc": ACCEL_PLOC}[accel]
    _check(lib().pt_set_accel_builder(accel))


class Profile:
    """Render profile (config/profile.rs:10-25).  Defaults: 1920x1080, 4 bounces, 64 samples, FILMIC."""

    def __init__(self, width=None, height=None, bounces=None, samples=None, tonemap=None):
        self.c = pt_profile()
        lib().pt_profile_default(C.byref(self.c))
        if width is not None:
            self.c.width = width
        if height is not None:
            self.c.height = height
        if bounces is not None:
            self.c.bounces = bounces
        if samples is not None:
            self.c.samples = samples
        if tonemap is not None:
            self.c.tonemap = TONEMAPS[tonemap] if isinstance(tonemap, str) else int(tonemap)

    @classmethod
    def load(cls, path):
        p = cls()
        _check(lib().pt_profile_load_yaml(os.fsencode(path), C.byref(p.c)))
        return p

    width = property(lambda s: s.c.width)
    height = property(lambda s: s.c.height)
    bounces = property(lambda s: s.c.bounces)
    samples = property(lambda s: s.c.samples)
    tonemap = property(lambda s: s.c.tonemap)


class Scene:
    """Immutable device-resident scene (scene::internal::Scene, scene/internal/mod.rs:26-51)."""

    def __init__(self, handle, keep=None):
        self.h = handle
        self._keep = keep

    @classmethod
    def load(cls, path, device=0):
        h = C.c_void_p()
        _check(lib().pt_scene_load_isf(os.fsencode(path), device, C.byref(h)))
        return cls(h)

    @classmethod
    def from_arrays(cls, arrays, device=0):
        d = arrays.to_desc()
        h = C.c_void_p()
        _check(lib().pt_scene_create(C.byref(d), device, C.byref(h)))
        return cls(h)

    @classmethod
    def synthetic(cls, recipe, n_tris, seed=0, device=0):
        h = C.c_void_p()
        _check(lib().pt_scene_synthetic(recipe.encode(), n_tris, seed, device, C.byref(h)))
        return cls(h)

    def close(self):
        if self.h:
            lib().pt_scene_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def replicate(self, device):
        """A copy of this scene on another device (-1: host-only); reuses the parsed arrays and the built tree."""
        h = C.c_void_p()
        _check(lib().pt_scene_replicate(self.h, device, C.byref(h)))
        return Scene(h)

    def upload(self, stream=None):
        """Re-send the scene's device arrays from pinned host memory (async on `stream`); returns bytes copied."""
        n = C.c_uint64()
        _check(lib().pt_scene_upload(self.h, stream, C.byref(n)))
        return n.value

    def upload_part(self, part, parts):
        """upload() of one part of `parts` of every array (+ the remainder behind the last shard); returns bytes copied."""
        n = C.c_uint64()
        _check(lib().pt_scene_upload_part(self.h, part, parts, C.byref(n)))
        return n.value

    def upload_targets(self, stream=None):
        """[(device pointer, bytes)] of the array set no kernel is reading (the one a pending upload() wrote); `stream`
        is made to wait until it may be written / read.  Same order on every replica of a scene."""
        n = C.c_uint32()
        _check(lib().pt_scene_upload_targets(self.h, stream, None, None, 0, C.byref(n)))
        ptrs, sizes = (C.c_void_p * n.value)(), (C.c_uint64 * n.value)()
        _check(lib().pt_scene_upload_targets(self.h, stream, ptrs, sizes, n.value, C.byref(n)))
        return [(int(ptrs[i] or 0), int(sizes[i])) for i in range(n.value)]

    def upload_commit(self, stream=None):
        """The arrays of upload_targets() are complete once the work enqueued on `stream` so far has run."""
        _check(lib().pt_scene_upload_commit(self.h, stream))

    def info(self):
        a, b, c, d = C.c_uint64(), C.c_uint64(), C.c_uint64(), C.c_uint64()
        _check(lib().pt_scene_info(self.h, C.byref(a), C.byref(b), C.byref(c), C.byref(d)))
        return dict(n_tris=a.value, n_prims=b.value, n_nodes=c.value, device_bytes=d.value)

    def accel_info(self):
        a, ms, d = C.c_int(), C.c_double(), C.c_int()
        _check(lib().pt_scene_accel_info(self.h, C.byref(a), C.byref(ms), C.byref(d)))
        return dict(accel={0: "sah", 1: "lbvh", 3: "ploc"}[a.value], build_ms=ms.value, max_depth=d.value)

    def save_binary(self, path, with_accel=True):
        """Write the scene as an isfb file (raw arrays, optionally with its BVH); Scene.load reads it back."""
        _check(lib().pt_scene_save_binary(self.h, os.fsencode(path), 1 if with_accel else 0))

    def arrays(self):
        """Host copy of the flat scene as a SceneArrays (used to hand the same scene to a checker)."""
        d = pt_scene_desc()
        _check(lib().pt_scene_get_desc(self.h, C.byref(d)))
        sa = SceneArrays()
        nt = d.n_tris
        f = lambda p, n: np.ctypeslib.as_array(p, shape=(n,)).copy() if n else np.zeros(0, np.float32)
        sa.tri_pos = f(d.tri_pos, nt * 9).reshape(-1, 9)
        sa.tri_nrm = f(d.tri_nrm, nt * 9).reshape(-1, 9)
        sa.tri_uv = f(d.tri_uv, nt * 6).reshape(-1, 6)
        for i in range(d.n_models):
            m, ma = d.models[i], d.materials[i]
            sa.models.append(dict(type=m.type, first_tri=m.first_tri, n_tris=m.n_tris, center=list(m.center), radius=m.radius))
            md = dict(albedo=list(ma.albedo), emissive=list(ma.emissive))
            for k in ("opacity", "metalness", "roughness", "ior", "albedo_tex", "emissive_tex", "opacity_tex",
                      "metalness_tex", "roughness_tex", "normal_tex"):
                md[k] = getattr(ma, k)
            sa.materials.append(md)
        for i in range(d.n_textures):
            t = d.textures[i]
            sa.textures.append(np.ctypeslib.as_array(t.data, shape=(t.height, t.width, t.channels)).copy())
        for i in range(d.n_lights):
            l = d.lights[i]
            sa.lights.append(dict(type=l.type, vec=list(l.vec), color=list(l.color), size=l.size))
        sa.camera = np.array(list(d.camera), np.float32)
        sa.fov = np.float32(d.fov)
        sa.background = np.array(list(d.background), np.float32)
        return sa

    # ray_cast(scene, ray).first()  (renderer/utils.rs:11-21)
    def trace_closest(self, rays):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        out = np.zeros(rays.shape[0], HIT_DTYPE)
        _check(lib().pt_trace_closest(self.h, rays.ctypes.data, rays.shape[0], out.ctypes.data))
        return out

    # device pointers (e.g. torch_tensor.data_ptr()) on the scene's device: n*6 float32 rays in, n pt_hit (16 B) out;
    # asynchronous on `stream` (a raw cudaStream_t value, None = default stream)
    def trace_closest_device(self, rays_ptr, n, out_ptr, stream=None):
        _check(lib().pt_trace_closest_device(self.h, rays_ptr, n, out_ptr, stream))

    # the same with the wave scratch in the library's static device array: no allocation, n <= TRACE_NOALLOC_MAX_RAYS
    def trace_closest_device_noalloc(self, rays_ptr, n, out_ptr, stream=None):
        _check(lib().pt_trace_closest_device_noalloc(self.h, rays_ptr, n, out_ptr, stream))

    def trace_transmittance(self, rays):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        out = np.zeros(rays.shape[0], np.float32)
        _check(lib().pt_trace_transmittance(self.h, rays.ctypes.data, rays.shape[0], out.ctypes.data))
        return out

    # the hit render_pixel shades after its alpha walk (renderer/mod.rs:188-205); ray i = path (pixel i, sample, seed)
    def trace_accepted(self, rays, bounce=0, sample=0, seed=0):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        out = np.zeros(rays.shape[0], HIT_DTYPE)
        _check(lib().pt_trace_accepted(self.h, rays.ctypes.data, rays.shape[0], bounce, sample, seed, out.ctypes.data))
        return out

    # get_light_info's shadow walk (renderer/mod.rs:281-333): directional (limits None) or point-limited
    def trace_shadow(self, rays, limits=None, primary=None):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        out = np.zeros(rays.shape[0], np.float32)
        lp = pp = None
        if limits is not None:
            limits = np.ascontiguousarray(limits, np.float32).reshape(-1)
            primary = np.ascontiguousarray(primary, np.float32).reshape(-1, 5)
            assert len(limits) == len(primary) == len(rays)
            lp, pp = limits.ctypes.data, primary.ctypes.data
        _check(lib().pt_trace_shadow(self.h, rays.ctypes.data, lp, pp, rays.shape[0], out.ctypes.data))
        return out

    # render_pixel (renderer/mod.rs:172-228) with a per-bounce record
    def trace_paths(self, profile, pixels, sample=0, seed=0):
        """Returns (vertices[n_paths, bounces+1] of PATH_VERTEX_DTYPE, counts[n_paths])."""
        pixels = np.ascontiguousarray(pixels, np.uint32)
        out = np.zeros((len(pixels), profile.bounces + 1), PATH_VERTEX_DTYPE)
        counts = np.zeros(len(pixels), np.uint32)
        _check(lib().pt_trace_paths(self.h, C.byref(profile.c), pixels.ctypes.data, len(pixels), sample, seed,
                                    out.ctypes.data, counts.ctypes.data))
        return out, counts


def triangle_intersect(rays, tris, device=0):
    """Triangle::intersect (scene/internal/triangle.rs:37-82) on (ray, triangle) pairs; prim_id -1 miss / 0 front / 1 back."""
    rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
    tris = np.ascontiguousarray(tris, np.float32).reshape(-1, 9)
    assert len(rays) == len(tris)
    out = np.zeros(len(rays), HIT_DTYPE)
    _check(lib().pt_triangle_intersect(None, device, rays.ctypes.data, tris.ctypes.data, len(rays), out.ctypes.data))
    return out


def philox4x32_10(counters, keys, device=0):
    """The device Philox4x32-10 on (counter[4], key[2]) pairs -> uint32[n, 4] (known-answer hook)."""
    c = np.ascontiguousarray(counters, np.uint32).reshape(-1, 4)
    k = np.ascontiguousarray(keys, np.uint32).reshape(-1, 2)
    assert len(c) == len(k)
    out = np.zeros((len(c), 4), np.uint32)
    _check(lib().pt_philox4x32_10(device, c.ctypes.data, k.ctypes.data, len(c), out.ctypes.data))
    return out


def convert_gltf_to_isf(input_path, output_dir):
    """scene::convert_gltf_to_isf (src/scene/gltf.rs:146-204): writes output_dir/scene.isf + texture PNGs."""
    _check(lib().pt_convert_gltf(os.fsencode(input_path), os.fsencode(output_dir)))


def load_internal(path, device=0):
    """scene::load_internal (src/scene/mod.rs:16-22)."""
    return Scene.load(path, device)


class RenderConfig:
    """Subset of config::RenderConfig (src/config/mod.rs:23-42) that affects rendering."""

    def __init__(self, quiet=True, viewer=False, seed=0):
        self.quiet, self.viewer, self.seed = quiet, viewer, seed


class Renderer:
    """Renderer::new(&config, profile) / render(&scene) -> RgbImage (src/renderer/mod.rs:61-169)."""

    def __init__(self, config=None, profile=None):
        self.config = config or RenderConfig()
        self.profile = profile or Profile()
        self.last_stats = None
        self.last_accum = None

    def render(self, scene, keep_accum=False, snapshot_every=0, on_snapshot=None):
        """on_snapshot(rgb[h, w, 3] copy, samples_done) -> truthy to stop watching: the headless form of the viewer
        feed (renderer/mod.rs:133-141), called after every `snapshot_every` samples per pixel."""
        p = self.profile.c
        rgb = np.zeros((p.height, p.width, 3), np.uint8)
        accum = np.zeros((p.height, p.width, 3), np.float32) if keep_accum else None
        st = pt_stats()
        if on_snapshot is not None and snapshot_every > 0:
            def _cb(user, img, w, h, done):
                arr = np.ctypeslib.as_array(img, shape=(h, w, 3)).copy()
                return 1 if on_snapshot(arr, done) else 0
            cb = SNAPSHOT_FN(_cb)
            _check(lib().pt_render_progressive(scene.h, C.byref(p), self.config.seed, snapshot_every, cb, None,
                                               rgb.ctypes.data, accum.ctypes.data if keep_accum else None, C.byref(st)))
        else:
            _check(lib().pt_render(scene.h, C.byref(p), self.config.seed, rgb.ctypes.data,
                                   accum.ctypes.data if keep_accum else None, C.byref(st)))
        self.last_stats = st.as_dict()
        self.last_accum = accum
        return rgb


def render_async(scene, profile, rgb8_host_ptr, seed=0, stream=None):
    """pt_render_async: enqueue one frame (render + tonemap + copy into host memory at rgb8_host_ptr) on `stream`."""
    _check(lib().pt_render_async(scene.h, C.byref(profile.c), seed, rgb8_host_ptr, stream))


def stream_synchronize(device=0, stream=None):
    _check(lib().pt_stream_synchronize(device, stream))


def render_multi(replicas, profile, seed=0, keep_accum=False):
    """pt_render_multi: one frame split by sample index over `replicas` (Scene objects, one per device) in this
    process; returns (rgb8[h, w, 3], accum or None, [per-replica stats])."""
    p = profile.c
    n = len(replicas)
    handles = (C.c_void_p * n)(*[r.h for r in replicas])
    rgb = np.zeros((p.height, p.width, 3), np.uint8)
    accum = np.zeros((p.height, p.width, 3), np.float32) if keep_accum else None
    st = (pt_stats * n)()
    _check(lib().pt_render_multi(handles, n, C.byref(p), seed, rgb.ctypes.data,
                                 accum.ctypes.data if keep_accum else None, st))
    return rgb, accum, [s.as_dict() for s in st]


def debug_render(scene, width, height):
    """debug_render (src/renderer/debug_renderer.rs:11-62): dict AOV name -> rgb8[h, w, 3], plus the hit-pixel count
    under the key "pixels_hit" (the reference writes `<name>.png` for each plane, and nothing when no ray hits)."""
    planes = np.zeros((len(AOV_NAMES), height, width, 3), np.uint8)
    n = C.c_uint64()
    _check(lib().pt_debug_textures(scene.h, width, height, planes.ctypes.data, C.byref(n)))
    out = {lib().pt_aov_name(i).decode(): planes[i] for i in range(len(AOV_NAMES))}
    out["pixels_hit"] = n.value
    return out


def render_accum_device(scene, profile, accum_ptr, sample_begin=0, sample_stride=1, n_samples=None, seed=0, stream=None,
                        want_stats=True):
    """pt_render_accum_device on a raw device pointer (e.g. torch_tensor.data_ptr()); returns stats dict or None."""
    n = profile.samples if n_samples is None else n_samples
    st = pt_stats()
    _check(lib().pt_render_accum_device(scene.h, C.byref(profile.c), sample_begin, sample_stride, n, seed, accum_ptr,
                                        stream, C.byref(st) if want_stats else None))
    return st.as_dict() if want_stats else None


def set_kernel_timing(on):
    lib().pt_set_kernel_timing(1 if on else 0)


def tonemap_device(accum_ptr, width, height, total_samples, tonemap, rgb8_ptr, device=0, stream=None):
    _check(lib().pt_tonemap_device(accum_ptr, width, height, total_samples, tonemap, rgb8_ptr, device, stream))


def write_png(path, rgb):
    rgb = np.ascontiguousarray(rgb, np.uint8)
    _check(lib().pt_write_png(os.fsencode(path), rgb.ctypes.data, rgb.shape[1], rgb.shape[0]))


def read_png(path, channels=3):
    w, h = C.c_uint32(), C.c_uint32()
    p = C.POINTER(C.c_uint8)()
    _check(lib().pt_read_png(os.fsencode(path), channels, C.byref(w), C.byref(h), C.byref(p)))
    arr = np.ctypeslib.as_array(p, shape=(h.value, w.value, channels)).copy()
    lib().pt_free(p)
    return arr
