// oracle/shading.hpp — TEST INFRASTRUCTURE (CPU oracle), not product code.
// Material lookup, Cook-Torrance and tonemap restated from the reference:
//   Material::get_*        scene/internal/material.rs:115-218
//   MaterialSample         renderer/material_sample.rs:20-40
//   Hit::get_normal        renderer/hit.rs:55-82
//   CookTorrance           renderer/brdf/cook_torrance.rs:19-183
//   transform_to_world     renderer/brdf/mod.rs:35-48
//   reflection / RR        renderer/utils.rs:23-36
//   tonemap                renderer/tonemap.rs:15-54
#pragma once
#include "rng.hpp"
#include "scene.hpp"

namespace orc {

static const float PI_F = 3.14159265358979323846f;  // std::f32::consts::PI

struct MaterialSample {
  float metalness, roughness;
  V3 albedo;
  float opacity;
  V3 emissive;
  float ior;
};

// material.rs:115-130: nearest texel, repeat, `as i64` truncation, no V flip.
static inline const uint8_t *get_pixel(const Scene &s, int32_t tex, V2 uv) {
  const pt_texture &t = s.textures[tex];
  float cx = uv.x * (float)t.width;
  float cy = uv.y * (float)t.height;
  int64_t x = rem_euclid(f32_as_i64(cx), (int64_t)t.width);
  int64_t y = rem_euclid(f32_as_i64(cy), (int64_t)t.height);
  return t.data + ((size_t)y * t.width + (size_t)x) * t.channels;
}

static inline MaterialSample material_sample(const Scene &s, int32_t model, bool simple, V2 uv) {
  const pt_material &m = s.materials[model];
  MaterialSample r;
  V3 af = v3(m.albedo[0], m.albedo[1], m.albedo[2]);
  V3 ef = v3(m.emissive[0], m.emissive[1], m.emissive[2]);
  if (simple) {  // material_sample.rs:31-40 (spheres)
    r.metalness = m.metalness;
    r.roughness = rmax(m.roughness, 0.0001f);
    r.albedo = af;
    r.opacity = m.opacity;
    r.emissive = ef;
    r.ior = m.ior;
    return r;
  }
  r.metalness = m.metalness_tex >= 0 ? (float)get_pixel(s, m.metalness_tex, uv)[0] / 255.0f * m.metalness : m.metalness;
  float rough = m.roughness_tex >= 0 ? (float)get_pixel(s, m.roughness_tex, uv)[0] / 255.0f * m.roughness : m.roughness;
  r.roughness = rmax(rough, 0.0001f);
  if (m.albedo_tex >= 0) {
    const uint8_t *p = get_pixel(s, m.albedo_tex, uv);
    r.albedo = mul_el(v3(std::pow((float)p[0] / 255.0f, 2.2f), std::pow((float)p[1] / 255.0f, 2.2f),
                         std::pow((float)p[2] / 255.0f, 2.2f)),
                      af);
  } else {
    r.albedo = af;
  }
  r.opacity = m.opacity_tex >= 0 ? (float)get_pixel(s, m.opacity_tex, uv)[0] / 255.0f * m.opacity : m.opacity;
  if (m.emissive_tex >= 0) {
    const uint8_t *p = get_pixel(s, m.emissive_tex, uv);
    r.emissive = mul_el(v3((float)p[0] / 255.0f, (float)p[1] / 255.0f, (float)p[2] / 255.0f), ef);
  } else {
    r.emissive = ef;
  }
  r.ior = m.ior;
  return r;
}

// Hit::get_material_sample (hit.rs:84-91): the variant of the HIT decides simple/textured.
static inline MaterialSample hit_material_sample(const Scene &s, const Hit &h, int32_t model) {
  return material_sample(s, model, h.is_sphere, h.tex_coords);
}

// hit.rs:55-82
static inline V3 hit_shading_normal(const Scene &s, const Hit &h, int32_t model) {
  if (h.is_sphere) return h.normal;
  const pt_material &m = s.materials[model];
  V3 n = h.normal;
  if (m.normal_tex >= 0) {
    const uint8_t *p = get_pixel(s, m.normal_tex, h.tex_coords);
    V3 nm = v3((float)p[0] / 127.5f - 1.0f, (float)p[1] / 127.5f - 1.0f, (float)p[2] / 127.5f - 1.0f);
    V3 bit = cross(h.normal, h.tangent);
    // Matrix3::from_cols(T, B, N) * nm = (T*x + B*y) + N*z
    V3 w = (h.tangent * nm.x + bit * nm.y) + h.normal * nm.z;
    n = normalize(w);
  }
  return h.is_backface ? -n : n;
}

// brdf/mod.rs:35-48
static inline V3 transform_to_world(V3 vec, V3 n) {
  V3 nt;
  if (std::fabs(n.x) > std::fabs(n.y)) nt = v3(n.z, 0.0f, -n.x) / std::sqrt(n.x * n.x + n.z * n.z);
  else nt = v3(0.0f, -n.z, n.y) / std::sqrt(n.y * n.y + n.z * n.z);
  V3 nb = cross(n, nt);
  return v3(vec.x * nb.x + vec.y * n.x + vec.z * nt.x, vec.x * nb.y + vec.y * n.y + vec.z * nt.y,
            vec.x * nb.z + vec.y * n.z + vec.z * nt.z);
}

// utils.rs:34-36
static inline V3 reflection(V3 i, V3 n) { return (2.0f * rmax(dot(i, n), 0.0f)) * n - i; }

struct CookTorrance {
  float metalness, roughness;
  V3 albedo, emissive, f0, m;  // m = microfacet normal

  explicit CookTorrance(const MaterialSample &s)
      : metalness(s.metalness), roughness(s.roughness), albedo(s.albedo), emissive(s.emissive) {
    f0 = v3(0.04f, 0.04f, 0.04f) * (1.0f - metalness) + albedo * metalness;  // :180-182
    m = v3(0, 0, 0);
  }
  V3 fresnel(float c) const {  // :143-147
    return f0 + v3(1.0f - f0.x, 1.0f - f0.y, 1.0f - f0.z) * powi(1.0f - c, 5);
  }
  static float g1(float ndv, float k) { return ndv / (ndv * (1.0f - k) + k); }  // :149-154
  float geometry(V3 n, V3 v, V3 l) const {                                       // :156-165
    float ndv = rmax(dot(n, v), 0.0f), ndl = rmax(dot(n, l), 0.0f);
    float k = powi(roughness + 1.0f, 2) / 8.0f;
    return g1(ndv, k) * g1(ndl, k);
  }
  float ndf(V3 n, V3 h) const {  // :167-178
    float a = roughness * roughness, a2 = a * a;
    float ndh = rmax(dot(n, h), 0.0f);
    float denom = (ndh * ndh) * (a2 - 1.0f) + 1.0f;
    denom = PI_F * denom * denom;
    return a2 / denom;
  }
  V3 diffuse(V3 ks, V3 n, V3 l) const {  // :107-117
    V3 kd = v3(1.0f - ks.x, 1.0f - ks.y, 1.0f - ks.z) * (1.0f - metalness);
    V3 dif = mul_el(kd, albedo) / PI_F;
    return dif * rmax(dot(n, l), 0.0f);
  }
  V3 sample(V3 n, V3 v, float r1, float r2) {  // :20-32, :119-141
    float a = roughness * roughness, a2 = a * a;
    float theta = std::acos(std::sqrt((1.0f - r1) / (r1 * (a2 - 1.0f) + 1.0f)));
    float phi = 2.0f * PI_F * r2;
    float st = std::sin(theta);
    V3 ml = normalize(v3(st * std::cos(phi), std::cos(theta), st * std::sin(phi)));
    m = normalize(transform_to_world(ml, n));
    return normalize(reflection(v, m));
  }
  V3 eval_direct(V3 n, V3 v, V3 l) const {  // :34-58
    V3 h = normalize(v + l);
    float d = ndf(n, h);
    V3 f = fresnel(rmax(dot(h, v), 0.0f));
    float g = geometry(n, v, l);
    V3 spec = ((d * f) * g) / rmax((4.0f * rmax(dot(n, v), 0.0f)) * rmax(dot(n, l), 0.0f), 0.0001f);
    spec = spec * rmax(dot(n, l), 0.0f);
    return (diffuse(f, n, l) + spec) + emissive;
  }
  V3 eval_indirect(V3 n, V3 v, V3 l) const {  // :60-86
    V3 h = normalize(v + l);
    V3 f = fresnel(rmax(dot(h, v), 0.0f));
    float g = geometry(n, v, l);
    V3 spec = v3(0, 0, 0);
    if (dot(n, l) > 0.0f) {
      float num = std::fabs(dot(v, m));
      float den = std::fabs(dot(v, n)) * std::fabs(dot(m, n));
      spec = (f * g) * (num / den);
    }
    return diffuse(f, n, l) + spec;
  }
};

// tonemap.rs:15-54 + post_processing renderer/mod.rs:335-353
static inline V3 tonemap(int32_t type, V3 c) {
  if (type == PT_TONEMAP_REINHARD) return div_el(c, c + v3(1, 1, 1));
  if (type == PT_TONEMAP_FILMIC) {
    V3 x = c - v3(0.004f, 0.004f, 0.004f);
    x = v3(rmax(x.x, 0.0f), rmax(x.y, 0.0f), rmax(x.z, 0.0f));
    V3 num = mul_el(x, 6.2f * x + v3(0.5f, 0.5f, 0.5f));
    V3 den = mul_el(x, 6.2f * x + v3(1.7f, 1.7f, 1.7f)) + v3(0.06f, 0.06f, 0.06f);
    return div_el(num, den);
  }
  V3 num = mul_el(c, 2.51f * c + v3(0.03f, 0.03f, 0.03f));
  V3 den = mul_el(c, 2.43f * c + v3(0.59f, 0.59f, 0.59f)) + v3(0.14f, 0.14f, 0.14f);
  V3 r = div_el(num, den);
  auto clamp01 = [](float x) { return x != x ? x : (x < 0.0f ? 0.0f : (x > 1.0f ? 1.0f : x)); };
  return v3(clamp01(r.x), clamp01(r.y), clamp01(r.z));
}

static inline void post_process(int32_t type, V3 c, uint8_t out[3]) {
  V3 t = tonemap(type, c);
  const float g = 1.0f / 2.2f;
  out[0] = f32_as_u8(std::pow(t.x, g) * 255.0f);
  out[1] = f32_as_u8(std::pow(t.y, g) * 255.0f);
  out[2] = f32_as_u8(std::pow(t.z, g) * 255.0f);
}

}  // namespace orc
