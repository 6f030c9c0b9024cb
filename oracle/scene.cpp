// oracle/scene.cpp — TEST INFRASTRUCTURE (CPU oracle), not product code.
// All-hits BVH filter + ray_cast (renderer/utils.rs:11-21).  The filter plays the role of
// the reference's two-level kdtree-ray candidate search: it may only drop primitives whose
// (inflated) bounding box the ray misses, so the sorted hit list equals brute force.
#include "scene.hpp"

#include <cfloat>
#include <numeric>

namespace orc {

namespace {
struct Box {
  float lo[3], hi[3];
  void reset() {
    for (int k = 0; k < 3; ++k) { lo[k] = FLT_MAX; hi[k] = -FLT_MAX; }
  }
  void grow(const float *p) {
    for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], p[k]); hi[k] = std::max(hi[k], p[k]); }
  }
  void grow(const Box &b) {
    for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], b.lo[k]); hi[k] = std::max(hi[k], b.hi[k]); }
  }
};

// Exact extent of a primitive (pad = 0) or its extent plus what the f32 intersection arithmetic can accept beyond
// it: Moller-Trumbore's barycentrics are a few ulp of |o - v0| off, which in space is a few ulp of the origin
// distance D outside the edges through v0 and (|e1| + |e2|) / |e2 - e1| times that outside the opposite edge; the
// sphere discriminant is a few ulp of |o - c|^2 off.  D = reach(v0 or centre) bounds the distance from any ray origin
// (the camera or a point on a surface, i.e. inside the scene box `all`) to the primitive; all == nullptr: exact extent.
static float reach(const float *p, const Box &all, const float *cam) {
  float f2 = 0, c2 = 0;
  for (int k = 0; k < 3; ++k) {
    float f = std::max(std::fabs(p[k] - all.lo[k]), std::fabs(p[k] - all.hi[k])), c = p[k] - cam[k];
    f2 += f * f;
    c2 += c * c;
  }
  float D = std::sqrt(std::max(f2, c2));
  return D < 1e30f ? D : 1.0f;
}

Box prim_box(const Scene &s, int32_t prim, const Box *all) {
  Box b;
  b.reset();
  if (prim >= 0) {
    const float *p = &s.pos[(size_t)prim * 9];
    for (int k = 0; k < 3; ++k) b.grow(p + k * 3);
    if (!all) return b;
    const float D = reach(p, *all, &s.camera[12]);
    float l1 = 0, l2 = 0, l3 = 0;
    for (int k = 0; k < 3; ++k) {
      float e1 = p[3 + k] - p[k], e2 = p[6 + k] - p[k], e3 = e2 - e1;
      l1 += e1 * e1; l2 += e2 * e2; l3 += e3 * e3;
    }
    // a triangle whose doubled area is below 1e-6 / 8 is never accepted (|det| < 1e-6, triangle.rs:49) for |d| <= 8
    float e1[3] = {p[3] - p[0], p[4] - p[1], p[5] - p[2]}, e2[3] = {p[6] - p[0], p[7] - p[1], p[8] - p[2]};
    float cx = e1[1] * e2[2] - e1[2] * e2[1], cy = e1[2] * e2[0] - e1[0] * e2[2], cz = e1[0] * e2[1] - e1[1] * e2[0];
    float ratio = 1.0f;
    if (std::sqrt(cx * cx + cy * cy + cz * cz) * 8.0f >= 1e-6f)
      ratio = std::min(std::max((std::sqrt(l1) + std::sqrt(l2)) / std::max(std::sqrt(l3), 1e-30f), 1.0f), 4096.0f);
    float pad = 2.4e-7f * D * ratio;
    for (int k = 0; k < 3; ++k) { b.lo[k] -= pad; b.hi[k] += pad; }
  } else {
    const pt_model &m = s.models[~prim];
    float r = std::fabs(m.radius);
    if (all) {
      const float D = reach(m.center, *all, &s.camera[12]);
      r = std::sqrt(r * r + 1.2e-7f * D * D) * 1.0001f + 1e-6f;
    }
    for (int k = 0; k < 3; ++k) { b.lo[k] = m.center[k] - r; b.hi[k] = m.center[k] + r; }
  }
  return b;
}

int32_t build_rec(Scene &s, std::vector<Box> &boxes, std::vector<float> &cent, int32_t first, int32_t count) {
  Box b, cb;
  b.reset();
  cb.reset();
  for (int32_t i = first; i < first + count; ++i) {
    int32_t slot = i;
    b.grow(boxes[slot]);
    cb.grow(&cent[(size_t)slot * 3]);
  }
  int32_t idx = (int32_t)s.nodes.size();
  s.nodes.push_back({});
  Scene::Node n;
  for (int k = 0; k < 3; ++k) {
    // conservative inflation: a few ulps relative + a tiny absolute term
    float e = std::max(std::fabs(b.lo[k]), std::fabs(b.hi[k])) * 4e-6f + 1e-7f;
    n.lo[k] = b.lo[k] - e;
    n.hi[k] = b.hi[k] + e;
  }
  int axis = 0;
  float ext[3] = {cb.hi[0] - cb.lo[0], cb.hi[1] - cb.lo[1], cb.hi[2] - cb.lo[2]};
  if (ext[1] > ext[axis]) axis = 1;
  if (ext[2] > ext[axis]) axis = 2;
  if (count <= 4 || !(ext[axis] > 0.0f)) {
    n.left = ~first;
    n.right = count;
    s.nodes[idx] = n;
    return idx;
  }
  // median split on the centroid along the widest axis: permute (prims, boxes, cent) together
  int32_t mid = first + count / 2;
  std::vector<int32_t> order(count);
  std::iota(order.begin(), order.end(), first);
  std::nth_element(order.begin(), order.begin() + count / 2, order.end(),
                   [&](int32_t a, int32_t c) { return cent[(size_t)a * 3 + axis] < cent[(size_t)c * 3 + axis]; });
  std::vector<int32_t> p(count);
  std::vector<Box> bx(count);
  std::vector<float> ce((size_t)count * 3);
  for (int32_t i = 0; i < count; ++i) {
    p[i] = s.bvh_prims[order[i]];
    bx[i] = boxes[order[i]];
    for (int k = 0; k < 3; ++k) ce[(size_t)i * 3 + k] = cent[(size_t)order[i] * 3 + k];
  }
  for (int32_t i = 0; i < count; ++i) {
    s.bvh_prims[first + i] = p[i];
    boxes[first + i] = bx[i];
    for (int k = 0; k < 3; ++k) cent[(size_t)(first + i) * 3 + k] = ce[(size_t)i * 3 + k];
  }
  n.left = build_rec(s, boxes, cent, first, mid - first);
  n.right = build_rec(s, boxes, cent, mid, first + count - mid);
  s.nodes[idx] = n;
  return idx;
}

// Conservative slab test, no tmax (the reference's traversal has none).  NaN from 0*inf is
// ignored by fmin/fmax, which widens the interval.
inline bool ray_box(const Scene::Node &n, const float o[3], const float inv[3]) {
  float tn = 0.0f, tf = FLT_MAX;
  for (int k = 0; k < 3; ++k) {
    float t0 = (n.lo[k] - o[k]) * inv[k];
    float t1 = (n.hi[k] - o[k]) * inv[k];
    float a = std::fmin(t0, t1), b = std::fmax(t0, t1);
    if (a != a || b != b) continue;  // o on the slab plane with d==0: treat as inside
    tn = std::fmax(tn, a);
    tf = std::fmin(tf, b);
  }
  return tn <= tf * 1.00001f + 1e-6f;
}
}  // namespace

void build_bvh(Scene &s) {
  s.nodes.clear();
  s.bvh_prims.clear();
  for (size_t m = 0; m < s.models.size(); ++m) {
    const pt_model &mo = s.models[m];
    if (mo.type == PT_MODEL_SPHERE) s.bvh_prims.push_back(~(int32_t)m);
    else
      for (uint32_t t = 0; t < mo.n_tris; ++t) s.bvh_prims.push_back((int32_t)(mo.first_tri + t));
  }
  size_t n = s.bvh_prims.size();
  if (n == 0) return;
  std::vector<Box> boxes(n);
  std::vector<float> cent(n * 3);
  Box all;
  all.reset();
  for (size_t i = 0; i < n; ++i) all.grow(prim_box(s, s.bvh_prims[i], nullptr));
  for (size_t i = 0; i < n; ++i) {
    boxes[i] = prim_box(s, s.bvh_prims[i], &all);
    for (int k = 0; k < 3; ++k) cent[i * 3 + k] = 0.5f * (boxes[i].lo[k] + boxes[i].hi[k]);
  }
  s.nodes.reserve(n);
  build_rec(s, boxes, cent, 0, (int32_t)n);
}

void ray_cast(const Scene &s, V3 o, V3 d, std::vector<Hit> &out, bool brute) {
  out.clear();
  if (brute || s.nodes.empty()) {
    for (size_t m = 0; m < s.models.size(); ++m) {
      const pt_model &mo = s.models[m];
      if (mo.type == PT_MODEL_SPHERE) sphere_hits(s, (int32_t)m, o, d, out);
      else
        for (uint32_t t = 0; t < mo.n_tris; ++t) try_triangle(s, mo.first_tri + t, o, d, out);
    }
  } else {
    float oo[3] = {o.x, o.y, o.z};
    float inv[3] = {1.0f / d.x, 1.0f / d.y, 1.0f / d.z};
    int32_t stack[128];
    int sp = 0;
    stack[sp++] = 0;
    while (sp) {
      const Scene::Node &n = s.nodes[stack[--sp]];
      if (!ray_box(n, oo, inv)) continue;
      if (n.left < 0) {
        int32_t first = ~n.left;
        for (int32_t i = first; i < first + n.right; ++i) {
          int32_t p = s.bvh_prims[i];
          if (p >= 0) try_triangle(s, (uint32_t)p, o, d, out);
          else sphere_hits(s, ~p, o, d, out);
        }
      } else {
        stack[sp++] = n.left;
        stack[sp++] = n.right;
      }
    }
  }
  // res.sort_by(dist) is stable over (model, triangle) push order == (dist, prim_id) order.
  std::sort(out.begin(), out.end(), [](const Hit &a, const Hit &b) {
    if (a.dist != b.dist) return a.dist < b.dist;
    return a.prim_id < b.prim_id;
  });
}

}  // namespace orc
