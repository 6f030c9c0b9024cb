// shade.cuh — device material lookup, Cook-Torrance GGX, Philox RNG, tonemap.
//   Material::get_*      reference scene/internal/material.rs:115-218
//   MaterialSample       reference renderer/material_sample.rs:20-40
//   Hit::new_triangle    reference renderer/hit.rs:100-137 (deferred here from the traversal)
//   Hit::get_normal      reference renderer/hit.rs:55-82
//   CookTorrance         reference renderer/brdf/cook_torrance.rs:19-183
//   transform_to_world   reference renderer/brdf/mod.rs:35-48
//   reflection           reference renderer/utils.rs:34-36
//   tonemap              reference renderer/tonemap.rs:15-54
// Operand order follows the Rust source so that, given equal random numbers, a path computes the
// same f32 values as the reference up to the last-ulp differences of sinf/cosf/acosf/powf.
#pragma once
#include "isect.cuh"

namespace ptb {

#define PTB_PI 3.14159265358979323846f

// Arithmetic of the SHADING functions (Cook-Torrance, light set-up, GGX sampling).  PTB_SHADE_FAST = 0 evaluates them
// with IEEE division / square root and libm-style sinf / cosf / acosf in the reference's operand order (bit-identical
// to the CPU oracle up to the last ulp of the transcendental functions).  PTB_SHADE_FAST = 1 uses the hardware
// approximations (div.approx / rsqrt.approx / sqrt.approx, <= 2 ulp) and an algebraically equivalent form of the GGX
// sample that needs no inverse trigonometry.  Intersection arithmetic (Moller-Trumbore, spheres, raygen, hit points,
// ray origins) is never touched by this switch: hit ids and t/u/v stay bit-exact; only shading VALUES move, by ~1e-7
// relative, far inside the image tolerance of the north star (1 % rel-RMSE) and of tests/test_gpu_parity.py.
// Why: the 34 IEEE divisions and 10 square roots of the shading code expand to ~600 of k_shade's ~2900 instructions
// and sinf / cosf / acosf with their large-argument slow paths to another ~350; the kernel's top stall was
// instruction fetch (52 KB of code against a 32 KB instruction cache).
#ifndef PTB_SHADE_FAST
#define PTB_SHADE_FAST 1
#endif
__device__ __forceinline__ float sh_div(float a, float b) {
#if PTB_SHADE_FAST
  return __fdividef(a, b);
#else
  return a / b;
#endif
}
__device__ __forceinline__ float sh_sqrt(float a) {
#if PTB_SHADE_FAST
  float r;
  asm("sqrt.approx.f32 %0, %1;" : "=f"(r) : "f"(a));
  return r;
#else
  return sqrtf(a);
#endif
}

__device__ __forceinline__ V3 sh_vdiv(V3 a, float s) { return {sh_div(a.x, s), sh_div(a.y, s), sh_div(a.z, s)}; }
__device__ __forceinline__ V3 sh_normalize(V3 a) {
#if PTB_SHADE_FAST
  return a * rsqrtf(dot3(a, a));
#else
  return a * (1.0f / length3(a));
#endif
}

// ---------------- Philox4x32-10 (counter-based; replaces rand's ChaCha12 StdRng) ----------------
// key = (pixel, sample), counter = (block, tag, seed_lo, seed_hi).  Draw names:
//   tag 0 block 0: [jitter x, jitter y, -, -]
//   tag b+1 block 0: [r1, r2, russian roulette, -] of bounce b
//   tag 0x80000000 | (b+1), block = global prim id: [alpha draw of that examined hit at bounce b, -, -, -]
// The alpha walk (mod.rs:188-205) gives every examined hit with 0.001 < opacity < 1 its own uniform draw; the
// reference takes it from the path's sequential stream, here it is addressed by the hit's primitive id, so the
// accept / reject decision of a hit does not depend on the order in which hits are found and the walk needs ONE
// traversal (traverse.cuh).  Same distribution of the accepted hit; the CPU oracle addresses its Philox draws the
// same way (oracle/rng.hpp).
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k0, lo1, hi0 ^ c.w ^ k1, lo0);
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return c;
}
__device__ __forceinline__ float u32_to_unit(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }

struct RngKey {
  uint32_t pixel, sample, seed_lo, seed_hi;
};
__device__ __forceinline__ uint4 rng_block(const RngKey &k, uint32_t tag, uint32_t block) {
  return philox4x32_10(make_uint4(block, tag, k.seed_lo, k.seed_hi), k.pixel, k.sample);
}
__device__ __forceinline__ float rng_alpha(const RngKey &k, uint32_t bounce, uint32_t prim_id) {
  return u32_to_unit(rng_block(k, 0x80000000u | (bounce + 1u), prim_id).x);
}

// ---------------- textures / materials ----------------
// material.rs:115-130: nearest texel, repeat, truncation toward zero (Rust `as i64` saturates).
// `(c as i64).rem_euclid(n)` (material.rs:121-128): `as i64` truncates toward zero, saturates and maps NaN to 0.
// Coordinates that fit 32 bits (all but absurd uvs) take a 32-bit remainder; the 64-bit one is kept out of line.
__device__ __noinline__ uint32_t texel_index_wide(float c, uint32_t n) {
  long long i = __float2ll_rz(c);  // saturates like `as i64`
  long long r = i % (long long)n;
  if (r < 0) r += (long long)n;
  return (uint32_t)r;
}
#ifndef PTB_FAST_TEXMOD
#define PTB_FAST_TEXMOD 1
#endif
__device__ __forceinline__ uint32_t texel_index(float c, uint32_t n) {
  if (c != c) return 0u;
#if PTB_FAST_TEXMOD
  // |i| < 2^21: the quotient estimate floor(float(i) * rcp(n)) is off by at most one (float(i) is exact, the product
  // is within 2^21 * 2^-22 of i / n for a reciprocal up to 4 ulp out), so one conditional add and one conditional
  // subtract give exactly rem_euclid: ~8 instructions instead of the ~30 of a remainder by a run-time divisor.
  if (fabsf(c) < 2097152.0f) {
    const int i = __float2int_rz(c);
    float inv;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv) : "f"((float)n));
    const int q = __float2int_rd((float)i * inv);
    int r = i - q * (int)n;
    if (r < 0) r += (int)n;
    if (r >= (int)n) r -= (int)n;
    return (uint32_t)r;
  }
#endif
  if (fabsf(c) < 2147483520.0f) {
    const int i = __float2int_rz(c);
    int r = i % (int)n;  // n < 2^31 (a texture side)
    if (r < 0) r += (int)n;
    return (uint32_t)r;
  }
  return texel_index_wide(c, n);
}
// PTB_TEX_INLINE = 0 keeps ONE out-of-line copy of the texel addressing per kernel instead of one per texture slot
// (six in k_shade, -430 instructions).  Measured (profiles/r07/ab_s1.jsonl): the call costs more than the smaller
// footprint gains wherever textures are actually fetched (k_shade +19 % on the console scene, +8 % on foliage), so
// the default stays inline.
#ifndef PTB_TEX_INLINE
#define PTB_TEX_INLINE 1
#endif
#if PTB_TEX_INLINE
__device__ __forceinline__
#else
__device__ __noinline__
#endif
const uint8_t *get_pixel(const DevScene &S, int tex, float u, float v) {
  const DevTexture t = S.textures[tex];
  float cx = u * (float)t.width, cy = v * (float)t.height;
  const uint32_t x = texel_index(cx, t.width), y = texel_index(cy, t.height);
  return S.texels + t.offset + ((size_t)y * t.width + (size_t)x) * t.channels;
}

struct MatSample {
  float metalness, roughness, opacity;
  V3 albedo, emissive;
};

// `texel as f32 / 255.` (material.rs:152-214) from the second half of the scene's 512-entry table (k_init_srgb_lut
// fills it with that very IEEE division): bit-identical, one load instead of a division sequence per texel
__device__ __forceinline__ float unit_of(const DevScene &S, uint8_t texel) { return __ldg(S.srgb_lut + 256 + texel); }

// the rarely used texture slots share one out-of-line copy of the texel addressing; albedo and opacity (every textured
// hit / every examined transparent hit) keep theirs inline
__device__ __noinline__ const uint8_t *get_pixel_cold(const DevScene &S, int tex, float u, float v) {
  return get_pixel(S, tex, u, v);
}

__device__ __forceinline__ float eval_opacity(const DevScene &S, int model, bool simple, float u, float v) {
  const DevMaterial &m = S.materials[model];
  if (!simple && m.opacity_tex >= 0) return unit_of(S, get_pixel(S, m.opacity_tex, u, v)[0]) * m.opacity;
  return m.opacity;
}

__device__ __forceinline__ MatSample material_sample(const DevScene &S, int model, bool simple, float u, float v) {
  const DevMaterial &m = S.materials[model];
  MatSample r;
  V3 af = mk3(m.albedo[0], m.albedo[1], m.albedo[2]);
  V3 ef = mk3(m.emissive[0], m.emissive[1], m.emissive[2]);
  r.metalness = m.metalness;
  float rough = m.roughness;
  r.albedo = af;
  r.emissive = ef;
  r.opacity = m.opacity;
  if (!simple) {
    if (m.metalness_tex >= 0) r.metalness = unit_of(S, get_pixel_cold(S, m.metalness_tex, u, v)[0]) * m.metalness;
    if (m.roughness_tex >= 0) rough = unit_of(S, get_pixel_cold(S, m.roughness_tex, u, v)[0]) * m.roughness;
    if (m.albedo_tex >= 0) {
      const uint8_t *p = get_pixel(S, m.albedo_tex, u, v);
      // powf(texel / 255, 2.2) through a 256-entry table filled by k_init_srgb_lut with the very same expression
      r.albedo = mul_el(mk3(__ldg(S.srgb_lut + p[0]), __ldg(S.srgb_lut + p[1]), __ldg(S.srgb_lut + p[2])), af);
    }
    if (m.opacity_tex >= 0) r.opacity = unit_of(S, get_pixel_cold(S, m.opacity_tex, u, v)[0]) * m.opacity;
    if (m.emissive_tex >= 0) {
      const uint8_t *p = get_pixel_cold(S, m.emissive_tex, u, v);
      r.emissive = mul_el(mk3(unit_of(S, p[0]), unit_of(S, p[1]), unit_of(S, p[2])), ef);
    }
  }
  r.roughness = fmaxf(rough, 0.0001f);
  return r;
}

// ---------------- surface reconstruction from a HitRec ----------------
struct Surface {
  V3 position;
  V3 normal;       // "geometric" normal of the reference = interpolated vertex normal (un-normalised)
  V3 shading_n;    // Hit::get_normal
  float tu, tv;    // interpolated texture coordinates
  int model;
  bool is_sphere;
};

// hit.rs:60-69 + :116-127.  e1,e2 = triangle edges, d1,d2 = uv deltas.
__device__ __forceinline__ V3 normal_from_map(V3 hit_normal, V3 e1, V3 e2, float d1x, float d1y, float d2x, float d2y,
                                              const uint8_t *p) {
  float f = sh_div(1.0f, d1x * d2y - d2x * d1y);
  V3 tangent = sh_normalize(mk3(f * (d2y * e1.x - d1y * e2.x), f * (d2y * e1.y - d1y * e2.y), f * (d2y * e1.z - d1y * e2.z)));
  V3 nm = mk3(sh_div((float)p[0], 127.5f) - 1.0f, sh_div((float)p[1], 127.5f) - 1.0f, sh_div((float)p[2], 127.5f) - 1.0f);
  V3 bit = cross3(hit_normal, tangent);
  V3 w = (tangent * nm.x + bit * nm.y) + hit_normal * nm.z;
  return sh_normalize(w);
}

__device__ __forceinline__ Surface make_surface(const DevScene &S, const Ray &r, float dist, float bu, float bv, int idx) {
  Surface sf;
  uint32_t i = (uint32_t)(idx & HIT_INDEX_MASK);
  if (i >= S.n_tris) {
    uint32_t k = (i - S.n_tris) >> 1, is_exit = (i - S.n_tris) & 1u;
    const DevSphere sp = S.spheres[k];
    // model.rs:42-62: the hit point is recomputed from the quadratic root, not from `dist`
    V3 c = mk3(sp.cx, sp.cy, sp.cz);
    V3 oc = r.o - c;
    float a = dot3(r.d, r.d);
    float b = 2.0f * dot3(oc, r.d);
    float cc = dot3(oc, oc) - sp.radius * sp.radius;
    float sq = sqrtf(b * b - 4.0f * a * cc);
    float t = is_exit ? (-b + sq) / (2.0f * a) : (-b - sq) / (2.0f * a);
    V3 p = r.o + r.d * t;
    V3 n = normalize3(p - c);
    sf.position = p;
    sf.normal = is_exit ? -n : n;
    sf.shading_n = sf.normal;
    sf.tu = sf.tv = 0.f;
    sf.model = sp.model;
    sf.is_sphere = true;
    return sf;
  }
  const DevTriShade &ts = S.shade[i];
  const float4 *q = reinterpret_cast<const float4 *>(&ts);
  float4 q0 = __ldg(q), q1 = __ldg(q + 1), q2 = __ldg(q + 2), q3 = __ldg(q + 3);
  V3 n0 = mk3(q0.x, q0.y, q0.z), n1 = mk3(q0.w, q1.x, q1.y), n2 = mk3(q1.z, q1.w, q2.x);
  float u0x = q2.y, u0y = q2.z, u1x = q2.w, u1y = q3.x, u2x = q3.y, u2y = q3.z;
  sf.model = __float_as_int(q3.w);
  sf.is_sphere = false;
  sf.position = r.o + r.d * dist;
  sf.normal = ((1.0f - bu - bv) * n0 + bu * n1) + bv * n2;
  sf.tu = (u0x + bu * (u1x - u0x)) + bv * (u2x - u0x);
  sf.tv = (u0y + bu * (u1y - u0y)) + bv * (u2y - u0y);
  V3 n = sf.normal;
  const DevMaterial &m = S.materials[sf.model];
  if (m.normal_tex >= 0 && S.tri_pos9 != nullptr) {  // hit.rs:60-69 (tangent: hit.rs:116-127)
    const float *pp = S.tri_pos9 + (size_t)i * 9;
    V3 e1 = mk3(pp[3] - pp[0], pp[4] - pp[1], pp[5] - pp[2]);
    V3 e2 = mk3(pp[6] - pp[0], pp[7] - pp[1], pp[8] - pp[2]);
    n = normal_from_map(sf.normal, e1, e2, u1x - u0x, u1y - u0y, u2x - u0x, u2y - u0y,
                        get_pixel_cold(S, m.normal_tex, sf.tu, sf.tv));
  }
  if (idx & HIT_BACKFACE) n = -n;
  sf.shading_n = n;
  return sf;
}

// ---------------- Cook-Torrance (cook_torrance.rs) ----------------
__device__ __forceinline__ float powi5(float a) {  // llvm.powi(x,5): r=a; a*=a; a*=a; r*=a
  float a2 = a * a;
  float a4 = a2 * a2;
  return a * a4;
}

struct Brdf {
  float metalness, roughness;
  V3 albedo, emissive, f0, m;
};

__device__ __forceinline__ Brdf make_brdf(const MatSample &s) {
  Brdf b;
  b.metalness = s.metalness;
  b.roughness = s.roughness;
  b.albedo = s.albedo;
  b.emissive = s.emissive;
  b.f0 = mk3(0.04f, 0.04f, 0.04f) * (1.0f - s.metalness) + s.albedo * s.metalness;
  b.m = mk3(0.f, 0.f, 0.f);
  return b;
}
__device__ __forceinline__ V3 fresnel(const Brdf &b, float c) {
  return b.f0 + mk3(1.0f - b.f0.x, 1.0f - b.f0.y, 1.0f - b.f0.z) * powi5(1.0f - c);
}
__device__ __forceinline__ float g1(float ndv, float k) { return sh_div(ndv, ndv * (1.0f - k) + k); }
__device__ __forceinline__ float geometry(const Brdf &b, V3 n, V3 v, V3 l) {
  float ndv = fmaxf(dot3(n, v), 0.0f), ndl = fmaxf(dot3(n, l), 0.0f);
  float a1 = b.roughness + 1.0f;
  float k = (a1 * a1) / 8.0f;  // exact either way: a division by a power of two
  return g1(ndv, k) * g1(ndl, k);
}
__device__ __forceinline__ float ndf(const Brdf &b, V3 n, V3 h) {
  float a = b.roughness * b.roughness, a2 = a * a;
  float ndh = fmaxf(dot3(n, h), 0.0f);
  float denom = (ndh * ndh) * (a2 - 1.0f) + 1.0f;
  denom = PTB_PI * denom * denom;
  return sh_div(a2, denom);
}
__device__ __forceinline__ V3 diffuse(const Brdf &b, V3 ks, V3 n, V3 l) {
  V3 kd = mk3(1.0f - ks.x, 1.0f - ks.y, 1.0f - ks.z) * (1.0f - b.metalness);
  V3 dif = sh_vdiv(mul_el(kd, b.albedo), PTB_PI);
  return dif * fmaxf(dot3(n, l), 0.0f);
}
__device__ __forceinline__ V3 transform_to_world(V3 vec, V3 n) {
  V3 nt;
#if PTB_SHADE_FAST
  if (fabsf(n.x) > fabsf(n.y)) nt = mk3(n.z, 0.0f, -n.x) * rsqrtf(n.x * n.x + n.z * n.z);
  else nt = mk3(0.0f, -n.z, n.y) * rsqrtf(n.y * n.y + n.z * n.z);
#else
  if (fabsf(n.x) > fabsf(n.y)) nt = mk3(n.z, 0.0f, -n.x) / sqrtf(n.x * n.x + n.z * n.z);
  else nt = mk3(0.0f, -n.z, n.y) / sqrtf(n.y * n.y + n.z * n.z);
#endif
  V3 nb = cross3(n, nt);
  return mk3(vec.x * nb.x + vec.y * n.x + vec.z * nt.x, vec.x * nb.y + vec.y * n.y + vec.z * nt.y,
             vec.x * nb.z + vec.y * n.z + vec.z * nt.z);
}
__device__ __forceinline__ V3 reflection(V3 i, V3 n) { return (2.0f * fmaxf(dot3(i, n), 0.0f)) * n - i; }

__device__ __forceinline__ V3 brdf_sample(Brdf &b, V3 n, V3 v, float r1, float r2) {
  float a = b.roughness * b.roughness, a2 = a * a;
#if PTB_SHADE_FAST
  // cook_torrance.rs:119-133 with theta eliminated: cos^2 = (1 - r1) / den, hence sin^2 = 1 - cos^2 = r1 a2 / den
  // (no cancellation, no acos / sin / cos of theta); phi = 2 pi r2 through sincospi (exact range reduction)
  const float inv_den = sh_div(1.0f, r1 * (a2 - 1.0f) + 1.0f);
  const float ct = sh_sqrt((1.0f - r1) * inv_den), st = sh_sqrt((r1 * a2) * inv_den);
  float sp, cp;
  sincospif(2.0f * r2, &sp, &cp);
  V3 ml = sh_normalize(mk3(st * cp, ct, st * sp));
#else
  float theta = acosf(sqrtf((1.0f - r1) / (r1 * (a2 - 1.0f) + 1.0f)));
  float phi = 2.0f * PTB_PI * r2;
  float st = sinf(theta);
  V3 ml = normalize3(mk3(st * cosf(phi), cosf(theta), st * sinf(phi)));
#endif
  b.m = sh_normalize(transform_to_world(ml, n));
  return sh_normalize(reflection(v, b.m));
}
__device__ __forceinline__ V3 eval_direct(const Brdf &b, V3 n, V3 v, V3 l) {
  V3 h = sh_normalize(v + l);
  float d = ndf(b, n, h);
  V3 f = fresnel(b, fmaxf(dot3(h, v), 0.0f));
  float g = geometry(b, n, v, l);
  V3 spec = sh_vdiv((d * f) * g, fmaxf((4.0f * fmaxf(dot3(n, v), 0.0f)) * fmaxf(dot3(n, l), 0.0f), 0.0001f));
  spec = spec * fmaxf(dot3(n, l), 0.0f);
  return (diffuse(b, f, n, l) + spec) + b.emissive;
}
__device__ __forceinline__ V3 eval_indirect(const Brdf &b, V3 n, V3 v, V3 l) {
  V3 h = sh_normalize(v + l);
  V3 f = fresnel(b, fmaxf(dot3(h, v), 0.0f));
  float g = geometry(b, n, v, l);
  V3 spec = mk3(0.f, 0.f, 0.f);
  if (dot3(n, l) > 0.0f) {
    float num = fabsf(dot3(v, b.m));
    float den = fabsf(dot3(v, n)) * fabsf(dot3(b.m, n));
    spec = (f * g) * sh_div(num, den);
  }
  return diffuse(b, f, n, l) + spec;
}

// ---------------- tonemap + post_processing (tonemap.rs, mod.rs:335-353) ----------------
__device__ __forceinline__ float clamp01_nan(float x) { return x != x ? x : (x < 0.0f ? 0.0f : (x > 1.0f ? 1.0f : x)); }
__device__ __forceinline__ float tonemap1(int type, float c) {
  if (type == 0) return c / (c + 1.0f);
  if (type == 1) {
    float x = fmaxf(c - 0.004f, 0.0f);
    float num = x * (6.2f * x + 0.5f);
    float den = x * (6.2f * x + 1.7f) + 0.06f;
    return num / den;
  }
  float num = c * (2.51f * c + 0.03f);
  float den = c * (2.43f * c + 0.59f) + 0.14f;
  return clamp01_nan(num / den);
}
__device__ __forceinline__ uint8_t to_u8(float f) {  // Rust `as u8`: saturating, NaN -> 0
  if (f != f) return 0;
  if (f <= 0.0f) return 0;
  if (f >= 255.0f) return 255;
  return (uint8_t)f;
}

}  // namespace ptb
