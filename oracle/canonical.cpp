// oracle/canonical.cpp — TEST / MEASUREMENT INFRASTRUCTURE (CPU oracle), not product code.  See canonical.hpp.
#include "canonical.hpp"

#include <algorithm>
#include <cmath>

namespace orc {
namespace {

struct Box {
  float lo[3], hi[3];
  void reset() {
    for (int k = 0; k < 3; ++k) { lo[k] = FLT_MAX; hi[k] = -FLT_MAX; }
  }
  void grow(const Box &b) {
    for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], b.lo[k]); hi[k] = std::max(hi[k], b.hi[k]); }
  }
  float half_area() const {
    float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    if (dx < 0 || dy < 0 || dz < 0) return 0.f;
    return dx * dy + dy * dz + dz * dx;
  }
};

constexpr int NBINS = 16;
constexpr int LEAF_MAX = 4;

struct Builder {
  const Scene &s;
  CanonBvh &out;
  std::vector<Box> box;  // per slot of out.prims
  std::vector<float> cen;

  int32_t rec(int32_t first, int32_t count) {
    Box b, cb;
    b.reset();
    cb.reset();
    for (int32_t i = first; i < first + count; ++i) {
      b.grow(box[i]);
      for (int k = 0; k < 3; ++k) {
        cb.lo[k] = std::min(cb.lo[k], cen[(size_t)i * 3 + k]);
        cb.hi[k] = std::max(cb.hi[k], cen[(size_t)i * 3 + k]);
      }
    }
    int32_t idx = (int32_t)out.nodes.size();
    out.nodes.push_back({});
    CanonBvh::Node n;
    for (int k = 0; k < 3; ++k) { n.lo[k] = b.lo[k]; n.hi[k] = b.hi[k]; }
    int best_axis = -1, best_split = -1;
    float best_cost = FLT_MAX;
    if (count > LEAF_MAX) {
      for (int axis = 0; axis < 3; ++axis) {
        float lo = cb.lo[axis], ext = cb.hi[axis] - cb.lo[axis];
        if (!(ext > 0.f)) continue;
        float scale = (float)NBINS / ext;
        Box bb[NBINS];
        int bc[NBINS] = {0};
        for (int q = 0; q < NBINS; ++q) bb[q].reset();
        for (int32_t i = first; i < first + count; ++i) {
          int q = std::min(NBINS - 1, std::max(0, (int)((cen[(size_t)i * 3 + axis] - lo) * scale)));
          bb[q].grow(box[i]);
          bc[q]++;
        }
        float la[NBINS], ra[NBINS];
        int lc[NBINS], rc[NBINS];
        Box acc;
        acc.reset();
        int cnt = 0;
        for (int q = 0; q < NBINS; ++q) { acc.grow(bb[q]); cnt += bc[q]; la[q] = acc.half_area(); lc[q] = cnt; }
        acc.reset();
        cnt = 0;
        for (int q = NBINS - 1; q >= 0; --q) { acc.grow(bb[q]); cnt += bc[q]; ra[q] = acc.half_area(); rc[q] = cnt; }
        for (int q = 0; q < NBINS - 1; ++q) {
          if (lc[q] == 0 || rc[q + 1] == 0) continue;
          float cost = la[q] * (float)lc[q] + ra[q + 1] * (float)rc[q + 1];
          if (cost < best_cost) { best_cost = cost; best_axis = axis; best_split = q; }
        }
      }
    }
    if (count <= LEAF_MAX) {
      n.left = ~first;
      n.right = count;
      out.nodes[idx] = n;
      return idx;
    }
    int32_t mid;
    if (best_axis < 0) {
      mid = first + count / 2;
    } else {
      float lo = cb.lo[best_axis], scale = (float)NBINS / (cb.hi[best_axis] - cb.lo[best_axis]);
      // stable three-array partition by bin
      std::vector<int32_t> order(count);
      for (int32_t i = 0; i < count; ++i) order[i] = first + i;
      auto it = std::stable_partition(order.begin(), order.end(), [&](int32_t i) {
        int q = std::min(NBINS - 1, std::max(0, (int)((cen[(size_t)i * 3 + best_axis] - lo) * scale)));
        return q <= best_split;
      });
      mid = first + (int32_t)(it - order.begin());
      std::vector<int32_t> p(count);
      std::vector<Box> bx(count);
      std::vector<float> ce((size_t)count * 3);
      for (int32_t i = 0; i < count; ++i) {
        p[i] = out.prims[order[i]];
        bx[i] = box[order[i]];
        for (int k = 0; k < 3; ++k) ce[(size_t)i * 3 + k] = cen[(size_t)order[i] * 3 + k];
      }
      for (int32_t i = 0; i < count; ++i) {
        out.prims[first + i] = p[i];
        box[first + i] = bx[i];
        for (int k = 0; k < 3; ++k) cen[(size_t)(first + i) * 3 + k] = ce[(size_t)i * 3 + k];
      }
      if (mid == first || mid == first + count) mid = first + count / 2;
    }
    n.left = rec(first, mid - first);
    n.right = rec(mid, first + count - mid);
    out.nodes[idx] = n;
    return idx;
  }
};

// slab test against [0, t_max]; returns entry parameter or FLT_MAX on a miss
inline float slab(const CanonBvh::Node &n, const float o[3], const float inv[3], float t_max) {
  float tn = 0.0f, tf = t_max;
  for (int k = 0; k < 3; ++k) {
    float t0 = (n.lo[k] - o[k]) * inv[k], t1 = (n.hi[k] - o[k]) * inv[k];
    float a = std::fmin(t0, t1), b = std::fmax(t0, t1);
    if (a != a || b != b) continue;
    tn = std::fmax(tn, a);
    tf = std::fmin(tf, b);
  }
  return tn <= tf * 1.00001f + 1e-6f ? tn : FLT_MAX;
}

// ray parameter of the nearest intersection of primitive `p` (FLT_MAX = none)
inline float prim_t(const Scene &s, int32_t p, V3 o, V3 d) {
  if (p >= 0) {
    float t, u, v, det;
    if (moller_trumbore(o, d, s.P(p, 0), s.P(p, 1), s.P(p, 2), t, u, v, det)) return t;
    return FLT_MAX;
  }
  const pt_model &mo = s.models[~p];
  V3 oc = o - v3(mo.center[0], mo.center[1], mo.center[2]);
  float a = dot(d, d), b = 2.0f * dot(oc, d), c = dot(oc, oc) - mo.radius * mo.radius;
  float disc = b * b - 4.0f * a * c;
  if (!(disc >= 0.0f)) return FLT_MAX;
  float sq = std::sqrt(disc);
  float t1 = (-b - sq) / (2.0f * a), t2 = (-b + sq) / (2.0f * a);
  if (t2 < 0.0f) return FLT_MAX;
  return t1 < 0.0f ? t2 : t1;
}

inline bool prim_opaque(const Scene &s, int32_t p) {
  int32_t m = p >= 0 ? s.tri_model[p] : ~p;
  const pt_material &ma = s.materials[m];
  return ma.opacity >= 1.0f && (p < 0 || ma.opacity_tex < 0);
}

}  // namespace

void CanonBvh::build(const Scene &s) {
  nodes.clear();
  prims.clear();
  for (size_t m = 0; m < s.models.size(); ++m) {
    const pt_model &mo = s.models[m];
    if (mo.type == PT_MODEL_SPHERE) prims.push_back(~(int32_t)m);
    else
      for (uint32_t t = 0; t < mo.n_tris; ++t) prims.push_back((int32_t)(mo.first_tri + t));
  }
  size_t n = prims.size();
  if (n == 0) return;
  Builder B{s, *this, {}, {}};
  B.box.resize(n);
  B.cen.resize(n * 3);
  for (size_t i = 0; i < n; ++i) {
    Box b;
    b.reset();
    int32_t p = prims[i];
    if (p >= 0) {
      for (int k = 0; k < 3; ++k)
        for (int a = 0; a < 3; ++a) {
          float x = s.pos[(size_t)p * 9 + k * 3 + a];
          b.lo[a] = std::min(b.lo[a], x);
          b.hi[a] = std::max(b.hi[a], x);
        }
    } else {
      const pt_model &mo = s.models[~p];
      for (int a = 0; a < 3; ++a) { b.lo[a] = mo.center[a] - mo.radius; b.hi[a] = mo.center[a] + mo.radius; }
    }
    B.box[i] = b;
    for (int k = 0; k < 3; ++k) B.cen[i * 3 + k] = 0.5f * (b.lo[k] + b.hi[k]);
  }
  nodes.reserve(2 * n);
  B.rec(0, (int32_t)n);
}

void CanonBvh::count_closest(const Scene &s, V3 o, V3 d, CanonCounts &c) const {
  ++c.rays_closest;
  if (nodes.empty()) return;
  const float oo[3] = {o.x, o.y, o.z}, inv[3] = {1.0f / d.x, 1.0f / d.y, 1.0f / d.z};
  float best = FLT_MAX;
  struct E { int32_t node; float t; };
  E stack[128];
  int sp = 0;
  if (slab(nodes[0], oo, inv, best) == FLT_MAX) return;
  stack[sp++] = {0, 0.0f};
  while (sp) {
    E e = stack[--sp];
    if (e.t > best) continue;
    const Node &n = nodes[e.node];
    if (n.left < 0) {
      int32_t first = ~n.left;
      for (int32_t i = first; i < first + n.right; ++i) {
        ++c.prims_closest;
        float t = prim_t(s, prims[i], o, d);
        if (t < best) best = t;
      }
      continue;
    }
    ++c.nodes_closest;  // one 64-byte inner node: two child boxes + two references
    float tl = slab(nodes[n.left], oo, inv, best), tr = slab(nodes[n.right], oo, inv, best);
    if (tl <= tr) {
      if (tr != FLT_MAX) stack[sp++] = {n.right, tr};
      if (tl != FLT_MAX) stack[sp++] = {n.left, tl};
    } else {
      if (tl != FLT_MAX) stack[sp++] = {n.left, tl};
      stack[sp++] = {n.right, tr};
    }
  }
}

void CanonBvh::count_any(const Scene &s, V3 o, V3 d, float t_max, CanonCounts &c) const {
  ++c.rays_any;
  if (nodes.empty()) return;
  const float oo[3] = {o.x, o.y, o.z}, inv[3] = {1.0f / d.x, 1.0f / d.y, 1.0f / d.z};
  int32_t stack[128];
  int sp = 0;
  if (slab(nodes[0], oo, inv, t_max) == FLT_MAX) return;
  stack[sp++] = 0;
  while (sp) {
    const Node &n = nodes[stack[--sp]];
    if (n.left < 0) {
      int32_t first = ~n.left;
      for (int32_t i = first; i < first + n.right; ++i) {
        ++c.prims_any;
        float t = prim_t(s, prims[i], o, d);
        if (t <= t_max && prim_opaque(s, prims[i])) return;  // fully blocked: an ideal walk stops here
      }
      continue;
    }
    ++c.nodes_any;
    float tl = slab(nodes[n.left], oo, inv, t_max), tr = slab(nodes[n.right], oo, inv, t_max);
    if (tr != FLT_MAX) stack[sp++] = n.right;
    if (tl != FLT_MAX) stack[sp++] = n.left;
  }
}

}  // namespace orc
