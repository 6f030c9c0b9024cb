"""ctypes binding of oracle/_build/liboracle.so (CPU oracle; test infrastructure only)."""
import ctypes as C
import os
import subprocess

import numpy as np

from path_tracer_b200.abi import HIT_DTYPE, PATH_VERTEX_DTYPE, pt_hit, pt_profile, pt_scene_desc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(ROOT, "oracle", "_build", "liboracle.so")
        if not os.path.exists(so):
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")])
        L = C.CDLL(so)
        L.oracle_scene_create.restype = C.c_void_p
        L.oracle_scene_create.argtypes = [C.POINTER(pt_scene_desc)]
        L.oracle_scene_destroy.argtypes = [C.c_void_p]
        L.oracle_scene_synthetic.restype = C.c_void_p
        L.oracle_scene_synthetic.argtypes = [C.c_char_p, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64)]
        L.oracle_moller_trumbore.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.oracle_trace_closest.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int, C.c_int]
        L.oracle_trace_all.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        L.oracle_trace_transmittance.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int]
        L.oracle_trace_accepted.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_uint32, C.c_uint32, C.c_uint64,
                                            C.c_void_p, C.c_int, C.c_int]
        L.oracle_trace_shadow.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p,
                                          C.c_int, C.c_int]
        L.oracle_philox4x32_10.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.oracle_philox4x32_10.restype = None
        L.oracle_dump_shadow_requests.argtypes = [C.c_void_p, C.POINTER(pt_profile), C.c_uint64, C.c_uint32, C.c_void_p,
                                                  C.c_void_p, C.c_void_p, C.c_void_p]
        L.oracle_render_accum.restype = C.c_double
        L.oracle_render_accum.argtypes = [C.c_void_p, C.POINTER(pt_profile), C.c_int, C.c_uint64, C.c_uint32,
                                          C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.oracle_tonemap.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int32, C.c_void_p]
        L.oracle_dump_bounce_rays.argtypes = [C.c_void_p, C.POINTER(pt_profile), C.c_uint64, C.c_uint32, C.c_void_p,
                                              C.c_void_p]
        L.oracle_trace_paths.argtypes = [C.c_void_p, C.POINTER(pt_profile), C.c_void_p, C.c_size_t, C.c_uint32,
                                         C.c_uint64, C.c_int, C.c_void_p, C.c_void_p]
        L.oracle_canonical_counts.argtypes = [C.c_void_p, C.POINTER(pt_profile), C.c_uint32, C.c_uint64, C.c_int, C.c_void_p]
        L.oracle_debug_textures.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]
        _LIB = L
    return _LIB


def make_profile(width, height, samples, bounces, tonemap=1):
    return pt_profile(width, height, bounces, samples, 0, tonemap)


class OracleScene:
    def __init__(self, arrays):
        self.arrays = arrays
        self._desc = arrays.to_desc()
        self.h = lib().oracle_scene_create(C.byref(self._desc))

    @classmethod
    def synthetic(cls, recipe, n_tris, seed=0):
        """The product's procedural generator compiled into the oracle (oracle/synth_glue.cpp): the same scene as
        pt_scene_synthetic(recipe, n_tris, seed) without loading libpt_b200.so."""
        self = cls.__new__(cls)
        self.arrays = self._desc = None
        n = C.c_uint64()
        self.h = lib().oracle_scene_synthetic(recipe.encode(), n_tris, seed, C.byref(n))
        if not self.h:
            raise ValueError("oracle_scene_synthetic(%r, %d) failed" % (recipe, n_tris))
        self.n_tris = n.value
        return self

    def __del__(self):
        if getattr(self, "h", None):
            lib().oracle_scene_destroy(self.h)
            self.h = None

    def trace_closest(self, rays, brute=False, threads=8):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        out = np.zeros(rays.shape[0], HIT_DTYPE)
        lib().oracle_trace_closest(self.h, rays.ctypes.data, rays.shape[0], out.ctypes.data, int(brute), threads)
        return out

    def trace_all(self, ray, brute=True, max_out=64):
        ray = np.ascontiguousarray(ray, np.float32).reshape(6)
        out = np.zeros(max_out, HIT_DTYPE)
        n = lib().oracle_trace_all(self.h, ray.ctypes.data, out.ctypes.data, max_out, int(brute))
        return out[:min(n, max_out)]

    def trace_transmittance(self, rays, brute=False):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        out = np.zeros(rays.shape[0], np.float32)
        lib().oracle_trace_transmittance(self.h, rays.ctypes.data, rays.shape[0], out.ctypes.data, int(brute))
        return out

    def trace_accepted(self, rays, bounce=0, sample=0, seed=0, brute=False, threads=8):
        """The hit the alpha walk (mod.rs:188-205) accepts for ray i = path (pixel i, sample, seed) at `bounce`."""
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        out = np.zeros(rays.shape[0], HIT_DTYPE)
        lib().oracle_trace_accepted(self.h, rays.ctypes.data, rays.shape[0], bounce, sample, seed, out.ctypes.data,
                                    int(brute), threads)
        return out

    def trace_shadow(self, rays, limits=None, primary=None, brute=False, threads=8):
        """get_light_info's shadow walk with unit light colour; limits/primary as pt_trace_shadow."""
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        out = np.zeros(rays.shape[0], np.float32)
        lp = pp = None
        if limits is not None:
            limits = np.ascontiguousarray(limits, np.float32).reshape(-1)
            primary = np.ascontiguousarray(primary, np.float32).reshape(-1, 5)
            lp, pp = limits.ctypes.data, primary.ctypes.data
        lib().oracle_trace_shadow(self.h, rays.ctypes.data, lp, pp, rays.shape[0], out.ctypes.data, int(brute), threads)
        return out

    def dump_shadow_requests(self, profile, light, seed=0):
        """The shadow rays get_light_info casts toward light `light` from the primary hits of a 1-spp trace:
        (rays[n, 6], limits[n] (-1 = directional), primary[n, 5])."""
        n = profile.width * profile.height
        rays = np.zeros((n, 6), np.float32)
        limits = np.zeros(n, np.float32)
        prim = np.zeros((n, 5), np.float32)
        valid = np.zeros(n, np.uint8)
        lib().oracle_dump_shadow_requests(self.h, C.byref(profile), seed, light, rays.ctypes.data, limits.ctypes.data,
                                          prim.ctypes.data, valid.ctypes.data)
        k = valid != 0
        return rays[k], limits[k], prim[k]

    def render_accum(self, profile, rng_mode=1, seed=0, sample_begin=0, sample_stride=1, n_samples=None, threads=8,
                     brute=False, return_counters=False):
        n_samples = profile.samples if n_samples is None else n_samples
        accum = np.zeros((profile.height, profile.width, 3), np.float32)
        cnt = np.zeros(2, np.uint64)
        ms = lib().oracle_render_accum(self.h, C.byref(profile), rng_mode, seed, sample_begin, sample_stride, n_samples,
                                       threads, int(brute), accum.ctypes.data, cnt.ctypes.data)
        if return_counters:
            return accum, ms, cnt
        return accum

    def trace_paths(self, profile, pixels, sample=0, seed=0, brute=False):
        pixels = np.ascontiguousarray(pixels, np.uint32)
        out = np.zeros((len(pixels), profile.bounces + 1), PATH_VERTEX_DTYPE)
        counts = np.zeros(len(pixels), np.uint32)
        lib().oracle_trace_paths(self.h, C.byref(profile), pixels.ctypes.data, len(pixels), sample, seed, int(brute),
                                 out.ctypes.data, counts.ctypes.data)
        return out, counts

    def canonical_counts(self, profile, pixel_stride=1, seed=0, threads=8):
        """Canonical BVH2 work per ray: dict with n_node/n_tri means for path segments and shadow rays."""
        out = np.zeros(6, np.uint64)
        lib().oracle_canonical_counts(self.h, C.byref(profile), pixel_stride, seed, threads, out.ctypes.data)
        rc, nc, pc, ra, na, pa = [int(x) for x in out]
        return dict(rays_closest=rc, n_node_closest=nc / max(rc, 1), n_tri_closest=pc / max(rc, 1),
                    rays_any=ra, n_node_any=na / max(ra, 1), n_tri_any=pa / max(ra, 1))

    def debug_textures(self, width, height):
        """debug_renderer.rs:11-105: (planes[7, h, w, 3] uint8 in AOV order, pixels that hit)."""
        planes = np.zeros((7, height, width, 3), np.uint8)
        n = lib().oracle_debug_textures(self.h, width, height, planes.ctypes.data)
        return planes, n

    def dump_bounce_rays(self, profile, bounce, seed=0):
        n = profile.width * profile.height
        rays = np.zeros((n, 6), np.float32)
        valid = np.zeros(n, np.uint8)
        lib().oracle_dump_bounce_rays(self.h, C.byref(profile), seed, bounce, rays.ctypes.data, valid.ctypes.data)
        return rays[valid != 0]


def tonemap(accum, total_samples, tonemap_type=1):
    accum = np.ascontiguousarray(accum, np.float32)
    h, w = accum.shape[:2]
    out = np.zeros((h, w, 3), np.uint8)
    lib().oracle_tonemap(accum.ctypes.data, w, h, total_samples, tonemap_type, out.ctypes.data)
    return out


def philox4x32_10(ctr4, key2):
    c = np.ascontiguousarray(ctr4, np.uint32)
    k = np.ascontiguousarray(key2, np.uint32)
    out = np.zeros(4, np.uint32)
    lib().oracle_philox4x32_10(c.ctypes.data, k.ctypes.data, out.ctypes.data)
    return out


def moller_trumbore(ray6, tri9):
    ray6 = np.ascontiguousarray(ray6, np.float32)
    tri9 = np.ascontiguousarray(tri9, np.float32)
    out = np.zeros(4, np.float32)
    hit = lib().oracle_moller_trumbore(ray6.ctypes.data, tri9.ctypes.data, out.ctypes.data)
    return bool(hit), out
