// isect.cuh — device intersection primitives and the BVH4 node test.
//   Möller–Trumbore      reference scene/internal/triangle.rs:37-82   (bit-exact: un-fused, cgmath order)
//   sphere quadratic     reference scene/internal/model.rs:26-64
// The traversals built from these (traverse.cuh) replace renderer/utils.rs:11-21.  The box culling is a
// conservative filter only: results must equal brute force over all primitives with the reference's stable
// (dist, push order) ordering == lexicographic (dist, global prim id).
#pragma once
#include <math_constants.h>

#include "device_scene.cuh"

namespace ptb {

struct V3 {
  float x, y, z;
};
__device__ __forceinline__ V3 mk3(float x, float y, float z) { return V3{x, y, z}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ V3 operator*(float s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
__device__ __forceinline__ V3 operator/(V3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
__device__ __forceinline__ V3 operator-(V3 a) { return {-a.x, -a.y, -a.z}; }
__device__ __forceinline__ V3 mul_el(V3 a, V3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
// cgmath: dot = (x*x' + y*y') + z*z'
__device__ __forceinline__ float dot3(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
__device__ __forceinline__ V3 cross3(V3 a, V3 b) {
  return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
__device__ __forceinline__ float length3(V3 a) { return sqrtf(dot3(a, a)); }
__device__ __forceinline__ V3 normalize3(V3 a) { return a * (1.0f / length3(a)); }

struct Ray {
  V3 o, d;
};

// Candidate key ordering of the reference's stable sort: (dist, global prim id).
__device__ __forceinline__ bool key_less(float t, int id, float t2, int id2) { return t < t2 || (t == t2 && id < id2); }

struct TraceState {
  float best_t;
  int best_id;
  float u, v;
  int idx;        // HitRec.idx encoding
};

// triangle.rs:37-82 with precomputed edges.  Returns true and t,u,v,det on a hit.
__device__ __forceinline__ bool moller_trumbore(const Ray &r, V3 v0, V3 e1, V3 e2, float &t, float &u, float &v,
                                                float &det) {
  V3 pvec = cross3(r.d, e2);
  det = dot3(e1, pvec);
  if (fabsf(det) < 0.000001f) return false;
  float invdet = __frcp_rn(det);  // correctly rounded 1/det == Rust's `1. / det`
  V3 tvec = r.o - v0;
  u = dot3(tvec, pvec) * invdet;
  if (!(u >= 0.0f && u <= 1.0f)) return false;
  V3 qvec = cross3(tvec, e1);
  v = dot3(r.d, qvec) * invdet;
  if (v < 0.0f || u + v > 1.0f) return false;
  t = dot3(e2, qvec) * invdet;
  if (t < 0.000001f) return false;
  return true;
}

// model.rs:26-64.  n_hits in {0,1,2}; dist[0] is the entry (or the only, exit) hit.
// exit_only = 1 when the origin is inside (only the exit hit exists).
__device__ __forceinline__ int sphere_hits(const Ray &r, const DevSphere &s, float dist[2], float tpar[2], int &exit_only) {
  V3 c = mk3(s.cx, s.cy, s.cz);
  V3 oc = r.o - c;
  float a = dot3(r.d, r.d);
  float b = 2.0f * dot3(oc, r.d);
  float cc = dot3(oc, oc) - s.radius * s.radius;
  float disc = b * b - 4.0f * a * cc;
  if (disc < 0.0f || disc != disc) return 0;
  float sq = sqrtf(disc);
  float t1 = (-b - sq) / (2.0f * a);
  float t2 = (-b + sq) / (2.0f * a);
  if (t2 < 0.0f || t2 != t2) return 0;
  V3 p2 = r.o + r.d * t2;
  float d2 = length3(p2 - r.o);
  if (t1 < 0.0f) {
    exit_only = 1;
    dist[0] = d2;
    tpar[0] = t2;
    return 1;
  }
  exit_only = 0;
  V3 p1 = r.o + r.d * t1;
  dist[0] = length3(p1 - r.o);
  tpar[0] = t1;
  dist[1] = d2;
  tpar[1] = t2;
  return 2;
}

// Reciprocal direction for the slab test, clamped to a finite magnitude so 0*inf never appears;
// a clamped slab is [-huge, +huge] when the origin is inside it (conservative).
#ifndef PTB_FAST_INV
#define PTB_FAST_INV 1
#endif
__device__ __forceinline__ float safe_inv(float d) {
  float ad = fabsf(d);
#if PTB_FAST_INV
  // the slab test is a filter with margins of tens of ulp (scene_host.hpp): a 1-ulp reciprocal is enough, and three
  // IEEE divisions per ray were 2.5 % of k_extend's instructions
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(d));
  return ad < 1e-18f ? copysignf(1e18f, d) : r;
#else
  return ad < 1e-18f ? copysignf(1e18f, d) : 1.0f / d;
#endif
}

struct RayPre {
  float ix, iy, iz;     // safe reciprocals
  float prune_scale;    // slab-parameter bound = best key * prune_scale (non-unit directions, rounding margin)
};

__device__ __forceinline__ RayPre precompute(const Ray &r) {
  RayPre p;
  p.ix = safe_inv(r.d.x);
  p.iy = safe_inv(r.d.y);
  p.iz = safe_inv(r.d.z);
  // Triangle keys are the ray parameter t, sphere keys are Euclidean distances t*|d|: a bound that is
  // valid for both is key*max(1, 1/|d|); 1.0002 absorbs rounding differences between slab and M-T.
#if PTB_FAST_INV
  const float l2 = dot3(r.d, r.d);
  float s = l2 > 0.0f ? fmaxf(1.0f, rsqrtf(l2)) : CUDART_INF_F;  // 2-ulp rsqrt against a 2e-4 allowance
#else
  float len = length3(r.d);
  float s = len > 0.0f ? fmaxf(1.0f, 1.0f / len) : CUDART_INF_F;
#endif
  p.prune_scale = s * 1.0002f;
  return p;
}

// Slab test of child k of a node; returns the entry parameter or +inf when the ray misses the box.
// (plane - o) * inv is evaluated as one fused multiply-add
// fma(plane, inv, -o*inv).  The box test is only a conservative filter (boxes are inflated by
// 4e-6 relative, ~33 ulp, and the exit is widened below), so its rounding may differ from the un-fused arithmetic that
// Moller-Trumbore must use; the absolute error of either form is ~1 ulp of the coordinate magnitude / |d|.
struct NodeOrder {
  float oix, oiy, oiz;  // -o * inv
};
__device__ __forceinline__ NodeOrder node_order(const Ray &r, const RayPre &p) {
  NodeOrder n;
  n.oix = -(r.o.x * p.ix);
  n.oiy = -(r.o.y * p.iy);
  n.oiz = -(r.o.z * p.iz);
  return n;
}
#ifndef PTB_SLAB_FOLD
#define PTB_SLAB_FOLD 1
#endif
// The any-hit walk of k_shadow keeps the two comparisons: folded, it measured +4 % on console and +2 % on foliage
// (profiles/r07/ab_s18_node_step_diet.jsonl, ab_s20_...jsonl); k_extend -2 % folded.
#ifndef PTB_SHADOW_SLAB_FOLD
#define PTB_SHADOW_SLAB_FOLD 0
#endif
__device__ __forceinline__ float fmin_nan(float a, float b) {
  float r;
  asm("min.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
}
// The slab test on the six plane parameters of a box: entry parameter, or +inf when the ray misses the box within
// [0, t_max].  Empty child slots of the float node carry NaN planes (bvh_build.cpp, lbvh.cu): every t is NaN, fminf /
// fmaxf then leave tf = NaN and the comparison is false, so no test of the child reference is needed.
template <bool FOLD = (PTB_SLAB_FOLD != 0)>
__device__ __forceinline__ bool slab_test(float t0x, float t1x, float t0y, float t1y, float t0z, float t1z, float t_max,
                                          float &tn) {
  tn = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fmaxf(fminf(t0z, t1z), 0.0f));
  const float tf = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fmaxf(t0z, t1z));
  // FOLD: one comparison - the widened exit and the pruning bound are merged with a NaN-propagating minimum first (an
  // empty slot's NaN exit must survive the merge for the comparison to fail)
  if (FOLD) return tn <= fmin_nan(__fmaf_rn(tf, 1.000002f, 1e-7f), t_max);
  return tn <= __fmaf_rn(tf, 1.000002f, 1e-7f) && tn <= t_max;
}
__device__ __forceinline__ float slab_fma(float lox, float loy, float loz, float hix, float hiy, float hiz, const RayPre &p,
                                          const NodeOrder &n, float t_max, uint32_t child) {
  (void)child;
  float tn;
  const bool hit = slab_test(__fmaf_rn(lox, p.ix, n.oix), __fmaf_rn(hix, p.ix, n.oix), __fmaf_rn(loy, p.iy, n.oiy),
                             __fmaf_rn(hiy, p.iy, n.oiy), __fmaf_rn(loz, p.iz, n.oiz), __fmaf_rn(hiz, p.iz, n.oiz), t_max, tn);
  return hit ? tn : CUDART_INF_F;
}
// Centre / half-extent form (PTB_CNODE): per axis tm = fma(c, inv, -o inv), p = h * inv, near = tm - |p|, far = tm + |p|.
// Two roundings per parameter instead of one (~1.2e-7 relative, against the 2e-6 widening of the exit); the host
// rounds h up so that [c - h, c + h] contains the builder's box (render.cu).  NaN centres mark empty slots.
template <bool FOLD = (PTB_SLAB_FOLD != 0)>
__device__ __forceinline__ bool cslab_test(float cx, float cy, float cz, float hx, float hy, float hz, const RayPre &p,
                                           const NodeOrder &n, float t_max, float &tn) {
  const float mx = __fmaf_rn(cx, p.ix, n.oix), my = __fmaf_rn(cy, p.iy, n.oiy), mz = __fmaf_rn(cz, p.iz, n.oiz);
  const float px = fabsf(hx * p.ix), py = fabsf(hy * p.iy), pz = fabsf(hz * p.iz);
  tn = fmaxf(fmaxf(mx - px, my - py), fmaxf(mz - pz, 0.0f));
  const float tf = fminf(fminf(mx + px, my + py), mz + pz);
  if (FOLD) return tn <= fmin_nan(__fmaf_rn(tf, 1.000002f, 1e-7f), t_max);
  return tn <= __fmaf_rn(tf, 1.000002f, 1e-7f) && tn <= t_max;
}
__device__ __forceinline__ float cslab(float cx, float cy, float cz, float hx, float hy, float hz, const RayPre &p,
                                       const NodeOrder &n, float t_max) {
  float tn;
  return cslab_test(cx, cy, cz, hx, hy, hz, p, n, t_max, tn) ? tn : CUDART_INF_F;
}

// The same test as a predicate (any-hit walks need no entry distance).  Conservative like slab_fma: a ray with
// non-finite slab parameters may be let into a box it misses, never kept out of one it hits.
template <bool FOLD = (PTB_SLAB_FOLD != 0)>
__device__ __forceinline__ bool slab_hit(float lox, float loy, float loz, float hix, float hiy, float hiz, const RayPre &p,
                                         const NodeOrder &n, float t_max) {
  float tn;
  return slab_test<FOLD>(__fmaf_rn(lox, p.ix, n.oix), __fmaf_rn(hix, p.ix, n.oix), __fmaf_rn(loy, p.iy, n.oiy),
                   __fmaf_rn(hiy, p.iy, n.oiy), __fmaf_rn(loz, p.iz, n.oiz), __fmaf_rn(hiz, p.iz, n.oiz), t_max, tn);
}

// One node = four 256-bit read-only loads (LDG.E.256): a lane whose ray is alone in its node touches the node's
// cache line four times instead of seven (the L1 processes one line per load instruction per cycle, and incoherent
// rays make every lane's node a different line: the traversal kernels are bound by exactly that).
struct NodeData {
  float4 lo_x, hi_x, lo_y, hi_y, lo_z, hi_z;
  uint4 ch;
};
__device__ __forceinline__ void ldg256(const void *p, float4 &a, float4 &b) {
  asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
               : "l"(p));
}
__device__ __forceinline__ NodeData load_node(const DevNode *nodes, uint32_t idx) {
  const char *p = reinterpret_cast<const char *>(nodes + idx);
  NodeData n;
  ldg256(p, n.lo_x, n.hi_x);
  ldg256(p + 32, n.lo_y, n.hi_y);
  ldg256(p + 64, n.lo_z, n.hi_z);
  n.ch = __ldg(reinterpret_cast<const uint4 *>(p + 96));
  return n;
}

// The quantised node (DevQNode): two 256-bit loads.  Plane byte q of child K decodes to origin + q * step; its slab
// parameter (origin + q step - o) * inv is evaluated as fma(1 + q 2^-15, A, B) with A = (step 2^15) * inv and
// B = (origin - o) * inv - A per axis and node, and the float 1 + q 2^-15 built from the byte by one byte permute
// (0x3F80qq00): no integer-to-float conversion, one PRMT per plane.  Rounding: 2^-8 step (scene_host.hpp keeps a guard
// of 0.05 step) plus the rounding of o * inv that the float node had as well.
struct QNodeData {
  float ox, oy, oz, sx, sy, sz;
  uint32_t qlx, qhx, qly, qhy, qlz, qhz;
  uint4 ch;
};
__device__ __forceinline__ QNodeData load_qnode(const DevQNode *nodes, uint32_t idx) {
  const char *p = reinterpret_cast<const char *>(nodes + idx);
  QNodeData n;
  uint32_t a0, a1, a2, a3, a4, a5;
  asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(a0), "=r"(a1), "=r"(a2), "=r"(a3), "=r"(a4), "=r"(a5), "=r"(n.qlx), "=r"(n.qhx)
               : "l"(p));
  asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(n.qly), "=r"(n.qhy), "=r"(n.qlz), "=r"(n.qhz), "=r"(n.ch.x), "=r"(n.ch.y), "=r"(n.ch.z), "=r"(n.ch.w)
               : "l"(p + 32));
  n.ox = __uint_as_float(a0);
  n.oy = __uint_as_float(a1);
  n.oz = __uint_as_float(a2);
  n.sx = __uint_as_float(a3);
  n.sy = __uint_as_float(a4);
  n.sz = __uint_as_float(a5);
  return n;
}
struct QNodeRay {  // A and B of the three axes for one (ray, node) pair
  float ax, ay, az, bx, by, bz;
  uint32_t one;  // DevScene::f32_one_bits
};
__device__ __forceinline__ QNodeRay qnode_ray(const QNodeData &n, const RayPre &p, const NodeOrder &o, uint32_t one) {
  QNodeRay r;
  r.one = one;
  r.ax = n.sx * p.ix;
  r.ay = n.sy * p.iy;
  r.az = n.sz * p.iz;
  r.bx = __fmaf_rn(n.ox, p.ix, o.oix) - r.ax;
  r.by = __fmaf_rn(n.oy, p.iy, o.oiy) - r.ay;
  r.bz = __fmaf_rn(n.oz, p.iz, o.oiz) - r.az;
  return r;
}
template <int K>
__device__ __forceinline__ float qplane(uint32_t w, uint32_t one) {  // 1 + q 2^-15, q = byte K of w
  return __uint_as_float(__byte_perm(w, one, 0x7604u | (K << 4)));
}
template <int K>
__device__ __forceinline__ bool qslab_test(const QNodeData &n, const QNodeRay &r, float t_max, float &tn) {
  return slab_test(__fmaf_rn(qplane<K>(n.qlx, r.one), r.ax, r.bx), __fmaf_rn(qplane<K>(n.qhx, r.one), r.ax, r.bx),
                   __fmaf_rn(qplane<K>(n.qly, r.one), r.ay, r.by), __fmaf_rn(qplane<K>(n.qhy, r.one), r.ay, r.by),
                   __fmaf_rn(qplane<K>(n.qlz, r.one), r.az, r.bz), __fmaf_rn(qplane<K>(n.qhz, r.one), r.az, r.bz), t_max, tn);
}
template <int K>
__device__ __forceinline__ float qslab(const QNodeData &n, const QNodeRay &r, float t_max) {
  float tn;
  return qslab_test<K>(n, r, t_max, tn) ? tn : CUDART_INF_F;
}

// Prefetch of the record behind a child reference into L1 (an inner node's line, or the first line of a leaf's first
// triangle): issued when the reference is decided, so that the load of the next visit finds the line on its way.
__device__ __forceinline__ void prefetch_l1(const void *p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
__device__ __forceinline__ void prefetch_ref(const DevScene &S, uint32_t ref) {
  if (!(ref & 0x80000000u)) {
#if PTB_QNODE
    prefetch_l1(S.qnodes + ref);
#elif PTB_CNODE
    prefetch_l1(S.cnodes + ref);
#else
    prefetch_l1(S.nodes + ref);
#endif
  } else if (!(ref & 0x40000000u)) {
    prefetch_l1(S.tris + ((ref & 0x3FFFFFFFu) >> 3));
  }
}

#define PTB_STACK 96

#ifndef PTB_CSWAP_SELECT
#define PTB_CSWAP_SELECT 1
#endif
#if PTB_CSWAP_SELECT
// branch-free exchange: one comparison, min / max of the keys, two selects of the references (the entry distances
// are never NaN here: misses are +inf)
#define PTB_CSWAP(ta, ra, tb, rb)                    \
  {                                                  \
    const bool _p = tb < ta;                         \
    const float _lo = fminf(ta, tb), _hi = fmaxf(ta, tb); \
    const uint32_t _rl = _p ? rb : ra, _rh = _p ? ra : rb; \
    ta = _lo; tb = _hi; ra = _rl; rb = _rh;          \
  }
#else
#define PTB_CSWAP(ta, ra, tb, rb) \
  if (tb < ta) {                  \
    float _t = ta; ta = tb; tb = _t; \
    uint32_t _r = ra; ra = rb; rb = _r; \
  }
#endif

}  // namespace ptb
