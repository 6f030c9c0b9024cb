// oracle/scene.hpp — TEST INFRASTRUCTURE (CPU oracle), not product code.
// Host scene + intersection restated from the reference:
//   Triangle::intersect   scene/internal/triangle.rs:37-82
//   Model::intersect      scene/internal/model.rs:23-73
//   Hit::new_triangle     renderer/hit.rs:100-137
//   ray_cast              renderer/utils.rs:11-21  (all hits, stable sort by dist)
// The KD-tree (third-party kdtree-ray ^1.2.0, not in /root/reference) is only a
// conservative candidate filter, so ray_cast == brute force over all primitives in
// (model, triangle) order; the BVH below is an all-hits filter validated against it.
#pragma once
#include <algorithm>
#include <cstdint>
#include <vector>

#include "../include/pt_b200.h"
#include "vecmath.hpp"

namespace orc {

struct CanonBvh;  // canonical.hpp

struct Hit {
  float dist;
  V3 position;
  V3 normal;   // interpolated (triangle, un-normalised) or geometric (sphere)
  V3 tangent;
  V2 tex_coords;
  bool is_backface;
  bool is_sphere;
  int32_t prim_id;  // global id (pt_hit.prim_id convention)
  int32_t model;
  float u, v;       // barycentrics (triangle only)
};

struct Scene {
  std::vector<pt_model> models;
  std::vector<pt_material> materials;
  std::vector<float> pos, nrm, uv;
  std::vector<pt_texture> textures;
  std::vector<std::vector<uint8_t>> texdata;
  std::vector<pt_light> lights;
  float camera[16];
  float fov;
  V3 background;
  std::vector<int32_t> prim_base;  // first global prim id of model m
  std::vector<int32_t> tri_model;  // model of triangle i
  int32_t n_prims = 0;

  // all-hits BVH over primitives (triangles + spheres)
  struct Node {
    float lo[3], hi[3];
    int32_t left, right;  // inner: children; leaf: left = ~first, right = count
  };
  std::vector<Node> nodes;
  std::vector<int32_t> bvh_prims;  // >=0: triangle index; <0: ~model (sphere)
  CanonBvh *canon = nullptr;       // canonical work counter (built on demand, canonical.hpp)

  V3 P(uint32_t tri, int k) const { return {pos[tri * 9 + k * 3], pos[tri * 9 + k * 3 + 1], pos[tri * 9 + k * 3 + 2]}; }
  V3 N(uint32_t tri, int k) const { return {nrm[tri * 9 + k * 3], nrm[tri * 9 + k * 3 + 1], nrm[tri * 9 + k * 3 + 2]}; }
  V2 T(uint32_t tri, int k) const { return {uv[tri * 6 + k * 2], uv[tri * 6 + k * 2 + 1]}; }
};

// triangle.rs:37-82.  Returns false on a miss; t,u,v,det out.
static inline bool moller_trumbore(V3 o, V3 d, V3 p0, V3 p1, V3 p2, float &t, float &u, float &v, float &det) {
  V3 v0v1 = p1 - p0;
  V3 v0v2 = p2 - p0;
  V3 pvec = cross(d, v0v2);
  det = dot(v0v1, pvec);
  if (std::fabs(det) < 0.000001f) return false;
  float invdet = 1.0f / det;
  V3 tvec = o - p0;
  u = dot(tvec, pvec) * invdet;
  if (!(u >= 0.0f && u <= 1.0f)) return false;  // !(0.0..=1.).contains(&u): NaN rejects
  V3 qvec = cross(tvec, v0v1);
  v = dot(d, qvec) * invdet;
  if (v < 0.0f || u + v > 1.0f) return false;
  t = dot(v0v2, qvec) * invdet;
  if (t < 0.000001f) return false;
  return true;
}

// hit.rs:100-137
static inline void fill_triangle_hit(const Scene &s, uint32_t tri, V3 o, V3 d, float t, float u, float v, float det,
                                     Hit &h) {
  h.dist = t;
  h.position = o + d * t;
  h.normal = ((1.0f - u - v) * s.N(tri, 0) + u * s.N(tri, 1)) + v * s.N(tri, 2);
  V2 t0 = s.T(tri, 0), t1 = s.T(tri, 1), t2 = s.T(tri, 2);
  h.tex_coords = (t0 + u * (t1 - t0)) + v * (t2 - t0);
  V3 e1 = s.P(tri, 1) - s.P(tri, 0);
  V3 e2 = s.P(tri, 2) - s.P(tri, 0);
  V2 d1 = t1 - t0, d2 = t2 - t0;
  float f = 1.0f / (d1.x * d2.y - d2.x * d1.y);
  h.tangent = normalize(v3(f * (d2.y * e1.x - d1.y * e2.x), f * (d2.y * e1.y - d1.y * e2.y),
                           f * (d2.y * e1.z - d1.y * e2.z)));
  h.is_backface = det < 0.0f;
  h.is_sphere = false;
  h.u = u;
  h.v = v;
}

// model.rs:26-64.  Appends 0, 1 or 2 hits (entry first).
static inline void sphere_hits(const Scene &s, int32_t m, V3 o, V3 d, std::vector<Hit> &out) {
  const pt_model &mo = s.models[m];
  V3 c = v3(mo.center[0], mo.center[1], mo.center[2]);
  float radius = mo.radius;
  V3 oc = o - c;
  float a = dot(d, d);
  float b = 2.0f * dot(oc, d);
  float cc = dot(oc, oc) - radius * radius;
  float disc = b * b - 4.0f * a * cc;
  if (disc < 0.0f) return;
  float t1 = (-b - std::sqrt(disc)) / (2.0f * a);
  float t2 = (-b + std::sqrt(disc)) / (2.0f * a);
  if (t2 < 0.0f) return;
  Hit h2{};
  V3 p2 = o + d * t2;
  h2.normal = -normalize(p2 - c);
  h2.dist = magnitude(p2 - o);
  h2.position = p2;
  h2.is_sphere = true;
  h2.is_backface = false;
  h2.model = m;
  h2.prim_id = s.prim_base[m] + 1;
  h2.tangent = v3(0, 0, 0);
  h2.tex_coords = {0, 0};
  if (t1 < 0.0f) {
    out.push_back(h2);
    return;
  }
  Hit h1 = h2;
  V3 p1 = o + d * t1;
  h1.normal = normalize(p1 - c);
  h1.dist = magnitude(p1 - o);
  h1.position = p1;
  h1.prim_id = s.prim_base[m];
  out.push_back(h1);
  out.push_back(h2);
}

static inline void try_triangle(const Scene &s, uint32_t tri, V3 o, V3 d, std::vector<Hit> &out) {
  float t, u, v, det;
  if (!moller_trumbore(o, d, s.P(tri, 0), s.P(tri, 1), s.P(tri, 2), t, u, v, det)) return;
  Hit h{};
  fill_triangle_hit(s, tri, o, d, t, u, v, det, h);
  h.model = s.tri_model[tri];
  h.prim_id = s.prim_base[h.model] + (int32_t)(tri - s.models[h.model].first_tri);
  out.push_back(h);
}

void build_bvh(Scene &s);
// utils.rs:11-21.  brute=true tests every primitive; otherwise the BVH filter is used.
void ray_cast(const Scene &s, V3 o, V3 d, std::vector<Hit> &out, bool brute);

}  // namespace orc
