// traverse.cuh — resumable (one step at a time) BVH4 traversals for the persistent kernels.
// A lane owns one ray; `step()` visits one node or one leaf and returns true when the query is complete, so a warp
// can replace finished rays lane by lane instead of waiting for its longest ray (Aila & Laine 2009, "terminated ray
// replacement").  Same arithmetic as trace_closest / trace_transmittance (isect.cuh, path.cuh): results are identical.
//   ClosestTrav  renderer/utils.rs:11-21 (smallest (dist, prim id) key above (t_lo, id_lo))
//   ShadowTrav   renderer/mod.rs:281-333 (product of (1 - opacity) over every hit, early out at 0)
#pragma once
#include "path.cuh"

#ifndef PTB_STK64
#define PTB_STK64 1  // measured: k_extend -3 % (console, city, foliage)
#endif

namespace ptb {

// The lane's closest-hit traversal stack (local memory): (entry distance, child reference) pairs, kept outside the
// traversal struct so that the scalar state stays in registers.  PTB_STK64 packs a pair into one 64-bit word
// (one STL.64 / LDL.64 per push / pop instead of two 32-bit accesses).
struct ClosestStack {
#if PTB_STK64
  uint2 e[PTB_STACK];
  __device__ __forceinline__ void put(int i, float t, uint32_t r) { e[i] = make_uint2(__float_as_uint(t), r); }
  __device__ __forceinline__ void get(int i, float &t, uint32_t &r) const {
    const uint2 v = e[i];
    t = __uint_as_float(v.x);
    r = v.y;
  }
#else
  float t_[PTB_STACK];
  uint32_t r_[PTB_STACK];
  __device__ __forceinline__ void put(int i, float t, uint32_t r) { t_[i] = t; r_[i] = r; }
  __device__ __forceinline__ void get(int i, float &t, uint32_t &r) const { t = t_[i]; r = r_[i]; }
#endif
};

#ifndef PTB_MERGED_POP
#define PTB_MERGED_POP 0
#endif

struct ClosestTrav {
  Ray r;
  RayPre pre;
  TraceState st;
  float t_prune, t_lo;
  int id_lo;
  uint32_t cur;
  int sp;

  __device__ __forceinline__ void begin(const Ray &ray, float tlo, int idlo) {
    r = ray;
    pre = precompute(ray);
    st.best_t = CUDART_INF_F;
    st.best_id = 0x7FFFFFFF;
    st.idx = -1;
    st.u = st.v = 0.f;
    st.flags = 0;
    t_prune = CUDART_INF_F;
    t_lo = tlo;
    id_lo = idlo;
    cur = 0;
    sp = 0;
  }
  // restart behind the current best hit (alpha walk), keeping the ray
  __device__ __forceinline__ void restart_behind() {
    t_lo = st.best_t;
    id_lo = st.best_id;
    st.best_t = CUDART_INF_F;
    st.best_id = 0x7FFFFFFF;
    st.idx = -1;
    st.u = st.v = 0.f;
    st.flags = 0;
    t_prune = CUDART_INF_F;
    cur = 0;
    sp = 0;
  }
  static constexpr uint32_t DONE = 0xFFFFFFFFu;
  // PTB_MERGED_POP: node_step / leaf_step leave NEED_POP instead of popping, and the kernel pops once per iteration
  // for every lane that needs it: the pop loop (entries behind the current best hit are skipped lazily) then runs with
  // the lanes of both paths together instead of once per call site with 2-5 lanes each.
  static constexpr uint32_t NEED_POP = 0xFFFFFFFEu;
  __device__ __forceinline__ bool at_inner() const { return !(cur & 0x80000000u); }
  __device__ __forceinline__ bool at_leaf() const { return (cur & 0x80000000u) && cur < NEED_POP; }
  __device__ __forceinline__ void after(const ClosestStack &stk) {
#if PTB_MERGED_POP
    (void)stk;
    cur = NEED_POP;
#else
    pop(stk);
#endif
  }

  __device__ __forceinline__ void pop(const ClosestStack &stk) {
    while (sp > 0) {
      --sp;
      float t;
      uint32_t ref;
      stk.get(sp, t, ref);
      if (t <= t_prune) { cur = ref; return; }
    }
    cur = DONE;
  }
  // Precondition: at_inner().  Visits one node: cur becomes the nearest child hit, or the next stack entry.
  __device__ __forceinline__ void node_step(const DevScene &S, const NodeOrder &no, ClosestStack &stk) {
    const NodeData nd = load_node(S.nodes, cur);
    const uint4 ch = nd.ch;
    float t0 = slab_fma(nd.lo_x.x, nd.lo_y.x, nd.lo_z.x, nd.hi_x.x, nd.hi_y.x, nd.hi_z.x, pre, no, t_prune, ch.x);
    float t1 = slab_fma(nd.lo_x.y, nd.lo_y.y, nd.lo_z.y, nd.hi_x.y, nd.hi_y.y, nd.hi_z.y, pre, no, t_prune, ch.y);
    float t2 = slab_fma(nd.lo_x.z, nd.lo_y.z, nd.lo_z.z, nd.hi_x.z, nd.hi_y.z, nd.hi_z.z, pre, no, t_prune, ch.z);
    float t3 = slab_fma(nd.lo_x.w, nd.lo_y.w, nd.lo_z.w, nd.hi_x.w, nd.hi_y.w, nd.hi_z.w, pre, no, t_prune, ch.w);
    uint32_t r0 = ch.x, r1 = ch.y, r2 = ch.z, r3 = ch.w;
    PTB_CSWAP(t0, r0, t1, r1) PTB_CSWAP(t2, r2, t3, r3) PTB_CSWAP(t0, r0, t2, r2) PTB_CSWAP(t1, r1, t3, r3)
    PTB_CSWAP(t1, r1, t2, r2)
    if (t3 < CUDART_INF_F) stk.put(sp++, t3, r3);
    if (t2 < CUDART_INF_F) stk.put(sp++, t2, r2);
    if (t1 < CUDART_INF_F) stk.put(sp++, t1, r1);
    if (t0 < CUDART_INF_F) cur = r0;
    else after(stk);
  }
  // Tests ONE primitive of leaf `ref` (uniform cost per step).  Returns true when the leaf has more primitives, with
  // `ref` advanced to the rest (first+1, count-1 == ref + 7).
  __device__ __forceinline__ bool leaf_step_ref(const DevScene &S, uint32_t &ref) {
    if (ref & 0x40000000u) {
      leaf_closest(S, ref, r, t_lo, id_lo, st, t_prune, pre);
      return false;
    }
    leaf_closest(S, ref & ~7u, r, t_lo, id_lo, st, t_prune, pre);  // count-1 = 0: the first primitive only
    if (!(ref & 7u)) return false;
    ref += 7u;
    return true;
  }
  // Precondition: at_leaf().  A triangle leaf with more primitives left stays current.
  __device__ __forceinline__ void leaf_step(const DevScene &S, const ClosestStack &stk) {
    if (!leaf_step_ref(S, cur)) after(stk);
  }
};

struct ShadowTrav {
  Ray r;
  RayPre pre;
  float t_max, limit, pu, pv, tr;
  V3 pos;
  uint32_t cur;
  int sp;

  // limit < 0: directional; else point light at distance `limit` from the primary hit `p` (see trace_transmittance)
  __device__ __forceinline__ void begin(const Ray &ray, float lim, V3 p, float u, float v) {
    r = ray;
    pre = precompute(ray);
    limit = lim;
    pos = p;
    pu = u;
    pv = v;
    t_max = CUDART_INF_F;
    if (lim >= 0.0f) {
#if PTB_FAST_INV
      // a culling bound with a 2e-4 allowance: approximate square roots and division are enough (the exact
      // `distance > limit` test of mod.rs:319-330 is applied per hit in leaf_step)
      const V3 b = ray.o - p;
      const float l2 = dot3(ray.d, ray.d);
      float bias;
      asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(bias) : "f"(dot3(b, b)));
      t_max = l2 > 0.f ? (lim + bias * 1.000001f) * rsqrtf(l2) * 1.0002f + 1e-6f : CUDART_INF_F;
#else
      float len = length3(ray.d);
      float bias = length3(ray.o - p);
      t_max = len > 0.f ? (lim + bias) / len * 1.0002f + 1e-6f : CUDART_INF_F;
#endif
    }
    tr = 1.0f;
    cur = 0;
    sp = 0;
  }
  static constexpr uint32_t DONE = 0xFFFFFFFFu;
  __device__ __forceinline__ bool at_inner() const { return !(cur & 0x80000000u); }
  __device__ __forceinline__ bool at_leaf() const { return (cur & 0x80000000u) && cur != DONE; }
  __device__ __forceinline__ void pop(const uint32_t *stk) { cur = (tr == 0.0f || sp == 0) ? DONE : stk[--sp]; }

  __device__ __forceinline__ void node_step(const DevScene &S, const NodeOrder &no, uint32_t *stk) {
    const NodeData nd = load_node(S.nodes, cur);
    const uint4 ch = nd.ch;
    if (slab_fma(nd.lo_x.x, nd.lo_y.x, nd.lo_z.x, nd.hi_x.x, nd.hi_y.x, nd.hi_z.x, pre, no, t_max, ch.x) < CUDART_INF_F) stk[sp++] = ch.x;
    if (slab_fma(nd.lo_x.y, nd.lo_y.y, nd.lo_z.y, nd.hi_x.y, nd.hi_y.y, nd.hi_z.y, pre, no, t_max, ch.y) < CUDART_INF_F) stk[sp++] = ch.y;
    if (slab_fma(nd.lo_x.z, nd.lo_y.z, nd.lo_z.z, nd.hi_x.z, nd.hi_y.z, nd.hi_z.z, pre, no, t_max, ch.z) < CUDART_INF_F) stk[sp++] = ch.z;
    if (slab_fma(nd.lo_x.w, nd.lo_y.w, nd.lo_z.w, nd.hi_x.w, nd.hi_y.w, nd.hi_z.w, pre, no, t_max, ch.w) < CUDART_INF_F) stk[sp++] = ch.w;
    pop(stk);
  }
  // One primitive per step (see ClosestTrav::leaf_step).
  __device__ __forceinline__ void leaf_step(const DevScene &S, const uint32_t *stk) {
    const bool point = limit >= 0.0f;
    if (cur & 0x40000000u) {
      const DevSphere s = S.spheres[cur & 0x3FFFFFFFu];
      float dist[2], tp[2];
      int exit_only = 0;
      int nh = sphere_hits(r, s, dist, tp, exit_only);
      for (int k = 0; k < nh; ++k) {
        if (point) {
          V3 p = r.o + r.d * tp[k];
          if (length3(p - pos) > limit) continue;
        }
        float op = point ? eval_opacity(S, s.model, pu != pu, pu, pv) : S.materials[s.model].opacity;
        tr *= 1.0f - op;
      }
      pop(stk);
      return;
    }
    const float4 *tp = reinterpret_cast<const float4 *>(S.tris + ((cur & 0x3FFFFFFFu) >> 3));
    float4 A = __ldg(tp), B = __ldg(tp + 1), C = __ldg(tp + 2);
    float t, u, v, det;
    if (moller_trumbore(r, mk3(A.x, A.y, A.z), mk3(B.x, B.y, B.z), mk3(C.x, C.y, C.z), t, u, v, det)) {
      int tri = __float_as_int(B.w);
      bool counted = true;
      float op;
      if (point) {
        V3 p = r.o + r.d * t;
        counted = !(length3(p - pos) > limit);
        op = counted ? eval_opacity(S, S.shade[tri].model, pu != pu, pu, pv) : 0.0f;
      } else if (__float_as_uint(C.w) & TRI_FLAG_ALPHA) {
        op = hit_opacity(S, tri, u, v);
      } else {
        op = S.materials[S.shade[tri].model].opacity;
      }
      if (counted) tr *= 1.0f - op;
    }
    if ((cur & 7u) && tr != 0.0f) cur += 7u;
    else pop(stk);
  }
};

}  // namespace ptb
