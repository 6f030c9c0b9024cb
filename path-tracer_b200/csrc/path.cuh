// path.cuh — per-path device functions shared by the wavefront kernels (kernels.cu) and the one-thread-per-path
// record kernel (k_trace_paths).  Each function is the sm_100a equivalent of one piece of the reference's
// render_pixel loop (file:line relative to the reference root):
//   generate_ray        renderer/mod.rs:107-124 + scene/internal/camera.rs:36-48
//   shade_begin         mod.rs:184-186 (miss), hit.rs:55-137 (surface), material_sample.rs, mod.rs:245 (emissive)
//   light_request       mod.rs:248-262 + the un-shadowed part of get_light_info mod.rs:281-333
//   (the traversals themselves - ray_cast, the alpha walk, the shadow walk - are in traverse.cuh)
//   shade_finish        mod.rs:265-275 (GGX sample + throughput), :219-221 (cutoff), utils.rs:23-31 (roulette)
#pragma once
#include "shade.cuh"

namespace ptb {

constexpr float NORMAL_BIAS = 0.00001f;  // renderer/mod.rs:59

__device__ __forceinline__ RngKey make_key(const WaveParams &P, uint32_t slot) {
  uint32_t npix = P.width * P.height;
  RngKey k;
  k.pixel = slot % npix;
  k.sample = P.sample_first + (slot / npix) * P.sample_stride;
  k.seed_lo = P.seed_lo;
  k.seed_hi = P.seed_hi;
  return k;
}

__device__ __forceinline__ Ray generate_ray(const DevScene &S, const WaveParams &P, const RngKey &key) {
  uint32_t x = key.pixel % P.width, y = key.pixel / P.width;
  uint4 rb = rng_block(key, 0, 0);
  float wf = (float)P.width, hf = (float)P.height;
  float ratio = wf / hf;
  float tn = S.tan_half_fov;  // (fov / 2).tan(), evaluated once on the host with the C library's tanf
  float sx = (float)x + u32_to_unit(rb.x);
  sx = sx / wf * 2.0f - 1.0f;
  sx *= tn * ratio;
  float sy = (float)y + u32_to_unit(rb.y);
  sy = 1.0f - sy / hf * 2.0f;
  sy *= tn;
  V3 dir = normalize3(mk3(sx, sy, -1.0f));
  const float *m = S.camera;
  V3 c0 = mk3(m[0], m[1], m[2]), c1 = mk3(m[4], m[5], m[6]), c2 = mk3(m[8], m[9], m[10]), c3 = mk3(m[12], m[13], m[14]);
  Ray r;
  r.d = ((c0 * dir.x + c1 * dir.y) + c2 * dir.z) + c3 * 0.0f;  // Matrix4 * (dir, 0); not re-normalised
  r.o = c3;
  return r;
}

// Opacity of an examined hit (Hit::get_material_sample(...).opacity, hit.rs:84-91).
__device__ __forceinline__ float hit_opacity(const DevScene &S, int idx, float bu, float bv) {
  uint32_t i = (uint32_t)(idx & HIT_INDEX_MASK);
  if (i >= S.n_tris) return S.materials[S.spheres[(i - S.n_tris) >> 1].model].opacity;
  const float4 *q = reinterpret_cast<const float4 *>(S.shade + i);
  float4 q2 = __ldg(q + 2), q3 = __ldg(q + 3);
  float u0x = q2.y, u0y = q2.z, u1x = q2.w, u1y = q3.x, u2x = q3.y, u2y = q3.z;
  float tu = (u0x + bu * (u1x - u0x)) + bv * (u2x - u0x);
  float tv = (u0y + bu * (u1y - u0y)) + bv * (u2y - u0y);
  return eval_opacity(S, __float_as_int(q3.w), false, tu, tv);
}

struct ShadeCtx {
  bool lit;  // the ray hit a surface (lights are evaluated, a bounce may follow)
  Surface sf;
  Brdf brdf;
  V3 view;
};

__device__ __forceinline__ void shade_begin(const DevScene &S, const Ray &r, float dist, float bu, float bv, int idx,
                                            V3 thr, V3 &rad, ShadeCtx &c) {
  c.lit = false;
  if (idx < 0) {
    rad = rad + mul_el(thr, mk3(S.background[0], S.background[1], S.background[2]));  // mod.rs:184-186
    return;
  }
  c.lit = true;
  c.sf = make_surface(S, r, dist, bu, bv, idx);
  MatSample ms = material_sample(S, c.sf.model, c.sf.is_sphere, c.sf.tu, c.sf.tv);
  c.brdf = make_brdf(ms);
  c.view = -1.0f * r.d;
  rad = rad + mul_el(thr, ms.emissive);  // mod.rs:245
}

// One light of mod.rs:248-262.  Returns false when the contribution is identically zero (no shadow ray needed);
// else the shadow ray (so, sd), the point-light distance (`limit`, -1 = directional) and the un-shadowed
// contribution c = throughput * eval_direct * L.
__device__ __forceinline__ bool light_request(const DevScene &S, const ShadeCtx &c, V3 thr, uint32_t li, V3 &so, V3 &sd,
                                              V3 &contrib, float &limit) {
  const DevLight L = S.lights[li];
  V3 lc = mk3(L.color[0], L.color[1], L.color[2]);
  V3 dir;
  limit = -1.0f;
  if (L.type == PT_LIGHT_DIRECTIONAL) {
    dir = mk3(L.vec[0], L.vec[1], L.vec[2]);
  } else {
    V3 dv = c.sf.position - mk3(L.vec[0], L.vec[1], L.vec[2]);
    // the distance limit of the shadow walk stays IEEE (it decides which occluders count, mod.rs:320-323)
    limit = length3(dv);
#if PTB_SHADE_FAST
    dir = dv * sh_div(1.0f, limit);
#else
    dir = normalize3(dv);
#endif
    float dissipation = 4.0f * PTB_PI * limit * limit;
    lc = sh_vdiv(lc, dissipation);
  }
  sd = -1.0f * dir;
  so = c.sf.position + c.sf.normal * NORMAL_BIAS;
  contrib = mul_el(mul_el(thr, eval_direct(c.brdf, c.sf.shading_n, c.view, sd)), lc);
  bool lzero = lc.x == 0.0f && lc.y == 0.0f && lc.z == 0.0f;
  bool czero = contrib.x == 0.0f && contrib.y == 0.0f && contrib.z == 0.0f;
  return !lzero && !czero;
}

// Returns true when the path continues with ray `next` and the updated throughput.  The cutoff and the roulette
// also run on the last iteration (they rescale the throughput there without any effect on the colour), exactly as
// render_pixel does, so that recorded throughputs compare bit-for-bit with the reference's.
__device__ __forceinline__ bool shade_finish(const WaveParams &P, const RngKey &key, uint32_t bounce, ShadeCtx &c,
                                             V3 &thr, Ray &next) {
  if (!c.lit) return false;  // miss: render_pixel returned at mod.rs:184-186
  const bool more = bounce < P.bounces;
  uint4 rb = rng_block(key, bounce + 1, 0);
  if (more) {  // mod.rs:265-275
    next.o = c.sf.position + c.sf.normal * NORMAL_BIAS;
    next.d = brdf_sample(c.brdf, c.sf.shading_n, c.view, u32_to_unit(rb.x), u32_to_unit(rb.y));
    V3 w = eval_indirect(c.brdf, c.sf.shading_n, c.view, next.d);  // `/ brdf.pdf()` with pdf() == 1: x / 1.0 == x exactly
    thr = mul_el(thr, w);
  }
  if (dot3(thr, thr) < 0.00001f) return false;  // mod.rs:219-221
  if (bounce > 3) {                             // utils.rs:23-31
    float p = fmaxf(fmaxf(thr.x, thr.y), thr.z);
    thr = thr * sh_div(1.0f, p);
    if (u32_to_unit(rb.z) > p) return false;
  }
  return more;
}

}  // namespace ptb
