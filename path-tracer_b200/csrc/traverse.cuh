// traverse.cuh — the BVH4 traversals: ONE implementation of each query, used by every kernel.
//   ClosestTrav  renderer/utils.rs:11-21 (first element of the sorted all-hits list) and, with `alpha`, the hit that
//                the alpha walk of render_pixel accepts (renderer/mod.rs:188-205)
//   ShadowTrav   renderer/mod.rs:281-333 (product of (1 - opacity) over every hit, early out at 0)
// Both are resumable: a lane owns one ray and `node_step` / `leaf_step` visit one node / test one primitive, so a
// warp of a persistent kernel can replace finished rays lane by lane instead of waiting for its longest ray (Aila &
// Laine 2009, "terminated ray replacement").  trace_closest / trace_transmittance at the end of the file run the same
// structs to completion for the one-thread-per-path kernels (k_trace_paths, k_debug_aov): there is no second traversal.
#pragma once
#include "path.cuh"

namespace ptb {

// The lane's closest-hit traversal stack (local memory): packed (entry distance, child reference) words, one
// STL.64 / LDL.64 per push / pop (measured -3 % against two 32-bit arrays), kept outside the traversal struct so
// that the scalar state stays in registers.  Three more words hold the farthest REJECTED hit of the alpha walk
// ((t, prim id) (u, v) (idx, -)): it is the hit that is shaded when no hit is accepted (mod.rs:188-205 leaves `hit`
// at the last examined element), touched only when a transparent hit is rejected.
struct ClosestStack {
  uint2 e[PTB_STACK + 1 + 3];  // sentinel + entries + the farthest rejected hit
  __device__ __forceinline__ void put(int i, float t, uint32_t r) { e[i] = make_uint2(__float_as_uint(t), r); }
  __device__ __forceinline__ void get(int i, float &t, uint32_t &r) const {
    const uint2 v = e[i];
    t = __uint_as_float(v.x);
    r = v.y;
  }
};
constexpr int FAR_SLOT = PTB_STACK + 1;

// One examined hit of the alpha walk (mod.rs:196-204): accepted when opacity >= 1, or when opacity > 0.001 and the
// hit's own draw is below it.  Out of line: it is reached by a few lanes at a time and holds a texture fetch and ten
// Philox rounds that the traversal loop should neither carry in its registers nor in its instruction footprint.
__device__ __noinline__ bool alpha_accepts(const DevScene &S, uint32_t pixel, uint32_t sample, uint32_t seed_lo,
                                           uint32_t seed_hi, uint32_t bounce, int idx, int prim_id, float u, float v) {
  const float op = hit_opacity(S, idx, u, v);
  if (op >= 1.0f) return true;
  if (!(op > 0.001f)) return false;
  RngKey k;
  k.pixel = pixel;
  k.sample = sample;
  k.seed_lo = seed_lo;
  k.seed_hi = seed_hi;
  return rng_alpha(k, bounce, (uint32_t)prim_id) < op;
}

// Where the lazy pop (skip stack entries behind the current best hit) runs in the persistent kernel (k_extend<POP>):
//   0  inline at its two call sites (after a node visit that entered no child, after the last primitive of a leaf)
//   1  one site per loop iteration for every lane that needs a pop, looping until a live entry is found
//   2  one site, ONE stack entry per iteration: a lane that draws a stale entry tries again next iteration (measured:
//      console +6 %, foliage -5 %, city +3 %; not used)
// -DPTB_POP_MODE=<n> forces one mode for every scene (A/B builds).
#ifdef PTB_POP_MODE
#define PTB_POP_MODE_FORCED 1
#endif

// PTB_STACK_PTR 1: the traversals keep the address of the next free stack word instead of an index (a push is one
// predicated STL + one predicated add, without the index -> address LEA of every push and pop).
#ifndef PTB_STACK_PTR
#define PTB_STACK_PTR 1
#endif
// PTB_SORT3 1: the four children are ordered by a three-exchange network instead of the full five-exchange sort: the
// nearest child is exact (it is visited next), two entered children are pushed in exact order, three are pushed with
// the nearer of the two pair-minima on top.  Results do not depend on the visiting order (candidate keys decide).
// Measured (profiles/r07/ab_s18_node_step_diet.jsonl): k_extend -0.6 % on console, +8 % on foliage, +3.4 % on city -
// the ten instructions it saves per node visit buy nothing (the kernel waits on the L1 data pipe, not on issue
// slots), the worse visiting order costs node visits.  Off.
#ifndef PTB_SORT3
#define PTB_SORT3 0
#endif

struct ClosestTrav {
  Ray r;
  RayPre pre;
  TraceState st;   // nearest accepted hit so far
  float t_prune;   // slab-parameter bound derived from it
  uint32_t cur;
  static constexpr uint32_t DONE = 0xFFFFFFFFu;
  static constexpr uint32_t NEED_POP = 0xFFFFFFFEu;  // PTB_POP_MODE != 0: the kernel's pop site takes over
#if PTB_STACK_PTR
  uint2 *top;      // next free stack word
  __device__ __forceinline__ void stack_reset(ClosestStack &stk) {
    stk.e[0] = make_uint2(0u, DONE);  // sentinel: passes every pruning test and ends the walk, so no pop tests for "empty"
    top = stk.e + 1;
  }
  __device__ __forceinline__ void push(ClosestStack &, float t, uint32_t ref) { *top++ = make_uint2(__float_as_uint(t), ref); }
  __device__ __forceinline__ void take(const ClosestStack &, float &t, uint32_t &ref) {
    const uint2 v = *--top;
    t = __uint_as_float(v.x);
    ref = v.y;
  }
#else
  int sp;
  __device__ __forceinline__ void stack_reset(ClosestStack &stk) {
    stk.put(0, 0.0f, DONE);
    sp = 1;
  }
  __device__ __forceinline__ void push(ClosestStack &stk, float t, uint32_t ref) { stk.put(sp++, t, ref); }
  __device__ __forceinline__ void take(const ClosestStack &stk, float &t, uint32_t &ref) { stk.get(--sp, t, ref); }
#endif

  __device__ __forceinline__ void begin(const Ray &ray, ClosestStack &stk, bool alpha) {
    r = ray;
    pre = precompute(ray);
    st.best_t = CUDART_INF_F;
    st.best_id = 0x7FFFFFFF;
    st.idx = -1;
    st.u = st.v = 0.f;
    t_prune = CUDART_INF_F;
    cur = 0;
    stack_reset(stk);
    if (alpha) stk.put(FAR_SLOT, -1.0f, 0xFFFFFFFFu);  // no rejected hit yet
  }
  __device__ __forceinline__ bool at_inner() const { return !(cur & 0x80000000u); }
  __device__ __forceinline__ bool at_leaf() const { return (cur & 0x80000000u) && cur < NEED_POP; }
  // what node_step / leaf_step do when the current subtree is exhausted
  __device__ __forceinline__ void after(const ClosestStack &stk, bool deferred) {
    if (deferred) cur = NEED_POP;
    else pop(stk);
  }
  // PTB_POP_MODE 2: examine one stack entry
  __device__ __forceinline__ void pop_one(const ClosestStack &stk) {
    float t;
    uint32_t ref;
    take(stk, t, ref);
    if (t <= t_prune) cur = ref;
  }

  // next stack entry that can still hold a nearer hit (entries behind the current best hit are skipped lazily)
  __device__ __forceinline__ void pop(const ClosestStack &stk) {
    for (;;) {  // ends at the sentinel at the latest (t = 0, ref = DONE)
      float t;
      uint32_t ref;
      take(stk, t, ref);
      if (t <= t_prune) {
        cur = ref;
        return;
      }
    }
  }
  // Precondition: at_inner().  Visits one node: cur becomes the nearest child hit, or the next stack entry.
  __device__ __forceinline__ void node_step(const DevScene &S, const NodeOrder &no, ClosestStack &stk, bool deferred = false) {
#if PTB_QNODE
    const QNodeData nd = load_qnode(S.qnodes, cur);
    const uint4 ch = nd.ch;
    const QNodeRay nr = qnode_ray(nd, pre, no, S.f32_one_bits);
    float t0 = qslab<0>(nd, nr, t_prune), t1 = qslab<1>(nd, nr, t_prune);
    float t2 = qslab<2>(nd, nr, t_prune), t3 = qslab<3>(nd, nr, t_prune);
#elif PTB_CNODE
    const NodeData nd = load_node(S.cnodes, cur);
    const uint4 ch = nd.ch;
    float t0 = cslab(nd.lo_x.x, nd.lo_y.x, nd.lo_z.x, nd.hi_x.x, nd.hi_y.x, nd.hi_z.x, pre, no, t_prune);
    float t1 = cslab(nd.lo_x.y, nd.lo_y.y, nd.lo_z.y, nd.hi_x.y, nd.hi_y.y, nd.hi_z.y, pre, no, t_prune);
    float t2 = cslab(nd.lo_x.z, nd.lo_y.z, nd.lo_z.z, nd.hi_x.z, nd.hi_y.z, nd.hi_z.z, pre, no, t_prune);
    float t3 = cslab(nd.lo_x.w, nd.lo_y.w, nd.lo_z.w, nd.hi_x.w, nd.hi_y.w, nd.hi_z.w, pre, no, t_prune);
#else
    const NodeData nd = load_node(S.nodes, cur);
    const uint4 ch = nd.ch;
    float t0 = slab_fma(nd.lo_x.x, nd.lo_y.x, nd.lo_z.x, nd.hi_x.x, nd.hi_y.x, nd.hi_z.x, pre, no, t_prune, ch.x);
    float t1 = slab_fma(nd.lo_x.y, nd.lo_y.y, nd.lo_z.y, nd.hi_x.y, nd.hi_y.y, nd.hi_z.y, pre, no, t_prune, ch.y);
    float t2 = slab_fma(nd.lo_x.z, nd.lo_y.z, nd.lo_z.z, nd.hi_x.z, nd.hi_y.z, nd.hi_z.z, pre, no, t_prune, ch.z);
    float t3 = slab_fma(nd.lo_x.w, nd.lo_y.w, nd.lo_z.w, nd.hi_x.w, nd.hi_y.w, nd.hi_z.w, pre, no, t_prune, ch.w);
#endif
    uint32_t r0 = ch.x, r1 = ch.y, r2 = ch.z, r3 = ch.w;
#if PTB_SORT3
    PTB_CSWAP(t0, r0, t1, r1) PTB_CSWAP(t2, r2, t3, r3) PTB_CSWAP(t0, r0, t2, r2)
    if (t3 < CUDART_INF_F) push(stk, t3, r3);
    if (t1 < CUDART_INF_F) push(stk, t1, r1);
    if (t2 < CUDART_INF_F) push(stk, t2, r2);  // the larger of the two pair-minima: never behind both of the others
#else
    PTB_CSWAP(t0, r0, t1, r1) PTB_CSWAP(t2, r2, t3, r3) PTB_CSWAP(t0, r0, t2, r2) PTB_CSWAP(t1, r1, t3, r3)
    PTB_CSWAP(t1, r1, t2, r2)
#if PTB_PREFETCH_CUR
    if (t0 < CUDART_INF_F) prefetch_ref(S, r0);
#endif
#if PTB_PREFETCH_PUSH
    if (t1 < CUDART_INF_F) prefetch_ref(S, r1);
#endif
    if (t3 < CUDART_INF_F) push(stk, t3, r3);
    if (t2 < CUDART_INF_F) push(stk, t2, r2);
    if (t1 < CUDART_INF_F) push(stk, t1, r1);
#endif
    if (t0 < CUDART_INF_F) cur = r0;
    else after(stk, deferred);
  }

  // One candidate hit with key (t, id).  Without `alpha` the smallest key wins (ray_cast's first element).  With it,
  // a hit of a possibly transparent primitive is examined at once — its decision depends only on the hit itself
  // (shade.cuh, "Draw names") — and a rejected one only competes for "farthest rejected hit".
  template <typename KeyFn>
  __device__ __forceinline__ void consider(const DevScene &S, ClosestStack &stk, float t, int id, float u, float v,
                                           int idx, uint32_t flags, bool alpha, uint32_t bounce, KeyFn key_of,
                                           uint32_t &n_alpha) {
    if (!key_less(t, id, st.best_t, st.best_id)) return;
    if (alpha && (flags & TRI_FLAG_ALPHA)) {
      ++n_alpha;
      const RngKey k = key_of();
      if (!alpha_accepts(S, k.pixel, k.sample, k.seed_lo, k.seed_hi, bounce, idx, id, u, v)) {
        float ft;
        uint32_t fid;
        stk.get(FAR_SLOT, ft, fid);
        if (key_less(ft, (int)fid, t, id)) {
          stk.put(FAR_SLOT, t, (uint32_t)id);
          stk.put(FAR_SLOT + 1, u, __float_as_uint(v));
          stk.put(FAR_SLOT + 2, __int_as_float(idx), 0u);
        }
        return;
      }
    }
    st.best_t = t;
    st.best_id = id;
    st.u = u;
    st.v = v;
    st.idx = idx;
    t_prune = t * pre.prune_scale;
  }

  // Precondition: at_leaf().  Tests ONE primitive of the current leaf (uniform cost per step); a triangle leaf with
  // more primitives left stays current (first+1, count-1 == ref + 7).
  template <typename KeyFn>
  __device__ __forceinline__ void leaf_step(const DevScene &S, ClosestStack &stk, bool alpha, uint32_t bounce,
                                            KeyFn key_of, uint32_t &n_alpha, bool deferred = false) {
    if (cur & 0x40000000u) {
      const uint32_t si = cur & 0x3FFFFFFFu;
      const DevSphere sph = S.spheres[si];
      float dist[2], tp[2];
      int exit_only = 0;
      const int n = sphere_hits(r, sph, dist, tp, exit_only);
      for (int k = 0; k < n; ++k) {
        const int is_exit = (n == 1) ? exit_only : k;
        consider(S, stk, dist[k], sph.prim_id + is_exit, 0.f, 0.f, (int)(S.n_tris + 2u * si + (uint32_t)is_exit), sph.flags,
                 alpha, bounce, key_of, n_alpha);
      }
      after(stk, deferred);
      return;
    }
    const float4 *tp = reinterpret_cast<const float4 *>(S.tris + ((cur & 0x3FFFFFFFu) >> 3));
    const float4 A = __ldg(tp), B = __ldg(tp + 1), C = __ldg(tp + 2);
    float t, u, v, det;
    if (moller_trumbore(r, mk3(A.x, A.y, A.z), mk3(B.x, B.y, B.z), mk3(C.x, C.y, C.z), t, u, v, det))
      consider(S, stk, t, __float_as_int(A.w), u, v, __float_as_int(B.w) | (det < 0.0f ? HIT_BACKFACE : 0),
               __float_as_uint(C.w), alpha, bounce, key_of, n_alpha);
    if (cur & 7u) cur += 7u;
    else after(stk, deferred);
  }

  // leaf_step on a leaf reference held outside `cur` (PTB_SPECULATE, k_extend): tests ONE primitive of `ref`; `ref` moves
  // on to the leaf's next primitive or becomes NEED_POP when the leaf is exhausted.  Does not touch `cur` or the stack
  // position (consider() only reads / writes the farthest-rejected record).
  template <typename KeyFn>
  __device__ __forceinline__ void leaf_test(const DevScene &S, ClosestStack &stk, uint32_t &ref, bool alpha, uint32_t bounce,
                                            KeyFn key_of, uint32_t &n_alpha) {
    if (ref & 0x40000000u) {
      const uint32_t si = ref & 0x3FFFFFFFu;
      const DevSphere sph = S.spheres[si];
      float dist[2], tp[2];
      int exit_only = 0;
      const int n = sphere_hits(r, sph, dist, tp, exit_only);
      for (int k = 0; k < n; ++k) {
        const int is_exit = (n == 1) ? exit_only : k;
        consider(S, stk, dist[k], sph.prim_id + is_exit, 0.f, 0.f, (int)(S.n_tris + 2u * si + (uint32_t)is_exit), sph.flags,
                 alpha, bounce, key_of, n_alpha);
      }
      ref = NEED_POP;
      return;
    }
    const float4 *tp = reinterpret_cast<const float4 *>(S.tris + ((ref & 0x3FFFFFFFu) >> 3));
    const float4 A = __ldg(tp), B = __ldg(tp + 1), C = __ldg(tp + 2);
    float t, u, v, det;
    if (moller_trumbore(r, mk3(A.x, A.y, A.z), mk3(B.x, B.y, B.z), mk3(C.x, C.y, C.z), t, u, v, det))
      consider(S, stk, t, __float_as_int(A.w), u, v, __float_as_int(B.w) | (det < 0.0f ? HIT_BACKFACE : 0),
               __float_as_uint(C.w), alpha, bounce, key_of, n_alpha);
    if (ref & 7u) ref += 7u;
    else ref = NEED_POP;
  }

  // After DONE: when nothing was accepted, the farthest rejected hit is the one the reference shades.
  __device__ __forceinline__ void finish(const ClosestStack &stk, bool alpha) {
    if (!alpha || st.idx >= 0) return;
    float ft, fu;
    uint32_t fid, fv, fi, unused;
    stk.get(FAR_SLOT, ft, fid);
    if (!(ft >= 0.0f)) return;
    stk.get(FAR_SLOT + 1, fu, fv);
    float fidx;
    stk.get(FAR_SLOT + 2, fidx, unused);
    fi = __float_as_uint(fidx);
    st.best_t = ft;
    st.best_id = (int)fid;
    st.u = fu;
    st.v = __uint_as_float(fv);
    st.idx = (int)fi;
  }
};

#ifndef PTB_SHADOW_ORDERED
// 1: k_shadow visits the children of a node nearest-first.  Measured (profiles/r07/ab_s1.jsonl): k_shadow +19 % on the
// console scene, +7 % on foliage, +10 % on city — the sorting network costs more than finding an occluder earlier
// saves (most shadow rays of these scenes are unoccluded and visit every node on their way anyway).  Off.
#define PTB_SHADOW_ORDERED 0
#endif
#ifndef PTB_SHADOW_NOPUSH
// 1: the last entered child becomes current without going through the stack (one STL + one LDL fewer per node visit).
// Measured (profiles/r07/ab_s2.jsonl): k_shadow +7 % on console, +5 % on foliage, +5 % on city - the extra live
// register costs 22 B more spills at 64 registers than the stack round trip it saves.  Off.
#define PTB_SHADOW_NOPUSH 0
#endif

#ifndef PTB_SHADOW_CACHE
// 1: every lane of k_shadow remembers the last exactly-opaque triangle that ended one of its walks and tests the next
// shadow ray against it before traversing: neighbouring shaded points are usually shadowed by the same occluder, and
// one hit on an opaque primitive already makes the product of (1 - opacity) zero (mod.rs:291-297, 319-329), whatever
// else lies on the ray - the result is unchanged, the traversal is skipped.
// Measured (profiles/r07/ab_s11_shadow_cache.jsonl): k_shadow +10 % on console, +5 % on foliage, +-0 on city - too few
// shadow rays of these scenes are occluded for the extra test (and 22 B more spills) to pay.  Off.
#define PTB_SHADOW_CACHE 0
#endif
// k_shadow keeps the index: the pointer form costs it 62 bytes of spills at 64 registers and measured +3 ... +4 %
// (profiles/r07/ab_s20_centre_half_extent_shadow_stack.jsonl).
#ifndef PTB_SHADOW_STACK_PTR
#define PTB_SHADOW_STACK_PTR 0
#endif
// Prefetches (isect.cuh prefetch_ref): PTB_PREFETCH_CUR - the nearest child, as soon as the sort has decided it (its
// load is some thirty instructions and a warp vote away); PTB_PREFETCH_PUSH - the second nearest child, which goes to
// the stack; PTB_SHADOW_PREFETCH - every child the shadow walk pushes (all of them are visited unless the ray is
// occluded first).  Measured (profiles/r07/ab_s21_prefetch_grid_constant_wave.jsonl): k_extend +2 % (CUR), +1 % (PUSH),
// +4.5 % (both) on console, +2.8 % / +2 % / +4.6 % on foliage, +2.8 % / - / +5.5 % on city; k_shadow +11 % on console,
// +13 % on foliage, +11 % on city: every extra request through the L1 data pipe costs more than the latency it
// hides.  All off.
#ifndef PTB_PREFETCH_CUR
#define PTB_PREFETCH_CUR 0
#endif
#ifndef PTB_PREFETCH_PUSH
#define PTB_PREFETCH_PUSH 0
#endif
#ifndef PTB_SHADOW_PREFETCH
#define PTB_SHADOW_PREFETCH 0
#endif

struct ShadowTrav {
  Ray r;
  RayPre pre;
  float t_max, limit, pu, pv, tr;
  V3 pos;
  uint32_t cur;
  static constexpr uint32_t DONE = 0xFFFFFFFFu;
#if PTB_SHADOW_STACK_PTR
  uint32_t *top;  // next free stack word
  __device__ __forceinline__ void stack_reset(uint32_t *stk) {
    stk[0] = DONE;  // sentinel: popping it ends the walk
    top = stk + 1;
  }
  __device__ __forceinline__ void push(uint32_t *, uint32_t ref) { *top++ = ref; }
  __device__ __forceinline__ uint32_t take(const uint32_t *) { return *--top; }
#else
  int sp;
  __device__ __forceinline__ void stack_reset(uint32_t *stk) {
    stk[0] = DONE;
    sp = 1;
  }
  __device__ __forceinline__ void push(uint32_t *stk, uint32_t ref) { stk[sp++] = ref; }
  __device__ __forceinline__ uint32_t take(const uint32_t *stk) { return stk[--sp]; }
#endif
  __device__ __forceinline__ void shadow_push(const DevScene &S, uint32_t *stk, uint32_t ref) {
    push(stk, ref);
#if PTB_SHADOW_PREFETCH
    prefetch_ref(S, ref);
#else
    (void)S;
#endif
  }
  uint32_t occluder;  // leaf-order index of the opaque triangle that ended this walk (valid when it did)

  // limit < 0: directional light (every hit on the ray, occluder sampled at its own uv, mod.rs:291-297)
  // limit >= 0: point light at distance `limit` from the primary hit `p`; occluders farther than the light are
  // ignored and the occluder's material is sampled with the PRIMARY hit's uv/variant (mod.rs:319-329, u NaN = sphere)
  __device__ __forceinline__ void begin(const Ray &ray, float lim, V3 p, float u, float v, uint32_t *stk) {
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
    stack_reset(stk);
  }
  __device__ __forceinline__ bool at_inner() const { return !(cur & 0x80000000u); }
  __device__ __forceinline__ bool at_leaf() const { return (cur & 0x80000000u) && cur != DONE; }
  __device__ __forceinline__ void pop(const uint32_t *stk) { cur = tr == 0.0f ? DONE : take(stk); }
  // After begin(): if the ray hits opaque triangle `tri` (leaf order) within the light's reach, the walk is over.
  __device__ __forceinline__ void try_cached(const DevScene &S, uint32_t tri) {
    const float4 *tp = reinterpret_cast<const float4 *>(S.tris + tri);
    const float4 A = __ldg(tp), B = __ldg(tp + 1), C = __ldg(tp + 2);
    if (!(__float_as_uint(C.w) & TRI_FLAG_OPAQUE1)) return;
    float t, u, v, det;
    if (!moller_trumbore(r, mk3(A.x, A.y, A.z), mk3(B.x, B.y, B.z), mk3(C.x, C.y, C.z), t, u, v, det)) return;
    if (limit >= 0.0f) {
      const V3 p = r.o + r.d * t;
      if (length3(p - pos) > limit) return;
    }
    tr = 0.0f;
    cur = DONE;
  }

  __device__ __forceinline__ void node_step(const DevScene &S, const NodeOrder &no, uint32_t *stk) {
#if PTB_QNODE
    const QNodeData nd = load_qnode(S.qnodes, cur);
    const uint4 ch = nd.ch;
    const QNodeRay nr = qnode_ray(nd, pre, no, S.f32_one_bits);
    float tn;
    if (qslab_test<0>(nd, nr, t_max, tn)) push(stk, ch.x);
    if (qslab_test<1>(nd, nr, t_max, tn)) push(stk, ch.y);
    if (qslab_test<2>(nd, nr, t_max, tn)) push(stk, ch.z);
    if (qslab_test<3>(nd, nr, t_max, tn)) push(stk, ch.w);
    pop(stk);
    return;
#elif PTB_CNODE
    const NodeData nd = load_node(S.cnodes, cur);
    const uint4 ch = nd.ch;
    float tn;
    if (cslab_test<(PTB_SHADOW_SLAB_FOLD != 0)>(nd.lo_x.x, nd.lo_y.x, nd.lo_z.x, nd.hi_x.x, nd.hi_y.x, nd.hi_z.x, pre, no, t_max, tn)) push(stk, ch.x);
    if (cslab_test<(PTB_SHADOW_SLAB_FOLD != 0)>(nd.lo_x.y, nd.lo_y.y, nd.lo_z.y, nd.hi_x.y, nd.hi_y.y, nd.hi_z.y, pre, no, t_max, tn)) push(stk, ch.y);
    if (cslab_test<(PTB_SHADOW_SLAB_FOLD != 0)>(nd.lo_x.z, nd.lo_y.z, nd.lo_z.z, nd.hi_x.z, nd.hi_y.z, nd.hi_z.z, pre, no, t_max, tn)) push(stk, ch.z);
    if (cslab_test<(PTB_SHADOW_SLAB_FOLD != 0)>(nd.lo_x.w, nd.lo_y.w, nd.lo_z.w, nd.hi_x.w, nd.hi_y.w, nd.hi_z.w, pre, no, t_max, tn)) push(stk, ch.w);
    pop(stk);
    return;
#else
    const NodeData nd = load_node(S.nodes, cur);
    const uint4 ch = nd.ch;
#endif
#if PTB_QNODE || PTB_CNODE
#elif PTB_SHADOW_ORDERED
    float t0 = slab_fma(nd.lo_x.x, nd.lo_y.x, nd.lo_z.x, nd.hi_x.x, nd.hi_y.x, nd.hi_z.x, pre, no, t_max, ch.x);
    float t1 = slab_fma(nd.lo_x.y, nd.lo_y.y, nd.lo_z.y, nd.hi_x.y, nd.hi_y.y, nd.hi_z.y, pre, no, t_max, ch.y);
    float t2 = slab_fma(nd.lo_x.z, nd.lo_y.z, nd.lo_z.z, nd.hi_x.z, nd.hi_y.z, nd.hi_z.z, pre, no, t_max, ch.z);
    float t3 = slab_fma(nd.lo_x.w, nd.lo_y.w, nd.lo_z.w, nd.hi_x.w, nd.hi_y.w, nd.hi_z.w, pre, no, t_max, ch.w);
    uint32_t r0 = ch.x, r1 = ch.y, r2 = ch.z, r3 = ch.w;
    PTB_CSWAP(t0, r0, t1, r1) PTB_CSWAP(t2, r2, t3, r3) PTB_CSWAP(t0, r0, t2, r2) PTB_CSWAP(t1, r1, t3, r3)
    PTB_CSWAP(t1, r1, t2, r2)
    if (t3 < CUDART_INF_F) push(stk, r3);
    if (t2 < CUDART_INF_F) push(stk, r2);
    if (t1 < CUDART_INF_F) push(stk, r1);
    if (t0 < CUDART_INF_F) push(stk, r0);
#elif !PTB_SHADOW_NOPUSH
    if (slab_hit<(PTB_SHADOW_SLAB_FOLD != 0)>(nd.lo_x.x, nd.lo_y.x, nd.lo_z.x, nd.hi_x.x, nd.hi_y.x, nd.hi_z.x, pre, no, t_max)) shadow_push(S, stk, ch.x);
    if (slab_hit<(PTB_SHADOW_SLAB_FOLD != 0)>(nd.lo_x.y, nd.lo_y.y, nd.lo_z.y, nd.hi_x.y, nd.hi_y.y, nd.hi_z.y, pre, no, t_max)) shadow_push(S, stk, ch.y);
    if (slab_hit<(PTB_SHADOW_SLAB_FOLD != 0)>(nd.lo_x.z, nd.lo_y.z, nd.lo_z.z, nd.hi_x.z, nd.hi_y.z, nd.hi_z.z, pre, no, t_max)) shadow_push(S, stk, ch.z);
    if (slab_hit<(PTB_SHADOW_SLAB_FOLD != 0)>(nd.lo_x.w, nd.lo_y.w, nd.lo_z.w, nd.hi_x.w, nd.hi_y.w, nd.hi_z.w, pre, no, t_max)) shadow_push(S, stk, ch.w);
#else
    // every child the ray enters is visited (in no particular order): the last one found becomes current directly,
    // only the others go through the local-memory stack (one STL + one LDL fewer per node visit)
    const bool h0 = slab_fma(nd.lo_x.x, nd.lo_y.x, nd.lo_z.x, nd.hi_x.x, nd.hi_y.x, nd.hi_z.x, pre, no, t_max, ch.x) < CUDART_INF_F;
    const bool h1 = slab_fma(nd.lo_x.y, nd.lo_y.y, nd.lo_z.y, nd.hi_x.y, nd.hi_y.y, nd.hi_z.y, pre, no, t_max, ch.y) < CUDART_INF_F;
    const bool h2 = slab_fma(nd.lo_x.z, nd.lo_y.z, nd.lo_z.z, nd.hi_x.z, nd.hi_y.z, nd.hi_z.z, pre, no, t_max, ch.z) < CUDART_INF_F;
    const bool h3 = slab_fma(nd.lo_x.w, nd.lo_y.w, nd.lo_z.w, nd.hi_x.w, nd.hi_y.w, nd.hi_z.w, pre, no, t_max, ch.w) < CUDART_INF_F;
    uint32_t nxt = DONE;
    if (h0) nxt = ch.x;
    if (h1) {
      if (nxt != DONE) push(stk, nxt);
      nxt = ch.y;
    }
    if (h2) {
      if (nxt != DONE) push(stk, nxt);
      nxt = ch.z;
    }
    if (h3) {
      if (nxt != DONE) push(stk, nxt);
      nxt = ch.w;
    }
    if (nxt != DONE) {
      cur = nxt;
      return;
    }
#endif
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
      const uint32_t flags = __float_as_uint(C.w);
      bool counted = true;
      if (point) {
        V3 p = r.o + r.d * t;
        counted = !(length3(p - pos) > limit);
      }
      if (counted) {
        if (flags & TRI_FLAG_OPAQUE1) {
          tr = 0.0f;  // opacity is exactly 1 whatever uv it is sampled at: tr * (1 - 1); no material fetch
          occluder = (cur & 0x3FFFFFFFu) >> 3;
        } else {
          const int tri = __float_as_int(B.w);
          float op;
          if (point) op = eval_opacity(S, S.shade[tri].model, pu != pu, pu, pv);
          else if (flags & TRI_FLAG_ALPHA) op = hit_opacity(S, tri, u, v);
          else op = S.materials[S.shade[tri].model].opacity;
          tr *= 1.0f - op;
        }
      }
    }
    if ((cur & 7u) && tr != 0.0f) cur += 7u;
    else pop(stk);
  }
};

// ---- the same traversals run to completion by one thread (k_trace_paths, k_debug_aov) ----
// Closest hit of ray_cast's sorted list (alpha = false) or the hit the alpha walk accepts (alpha = true, draws of
// path `key` at `bounce`).  st.idx < 0: miss.  n_alpha counts the opacity evaluations.
__device__ __forceinline__ void trace_closest(const DevScene &S, const Ray &r, bool alpha, const RngKey &key,
                                              uint32_t bounce, TraceState &st, uint32_t &n_alpha) {
  ClosestStack stk;
  ClosestTrav tv;
  tv.begin(r, stk, alpha);
  const NodeOrder no = node_order(tv.r, tv.pre);
  auto key_of = [&]() { return key; };
  while (tv.cur != ClosestTrav::DONE) {
    if (tv.at_inner()) tv.node_step(S, no, stk);
    else tv.leaf_step(S, stk, alpha, bounce, key_of, n_alpha);
  }
  tv.finish(stk, alpha);
  st = tv.st;
}

__device__ __forceinline__ float trace_transmittance(const DevScene &S, const Ray &r, float limit, V3 pos, float pu,
                                                     float pv) {
  uint32_t stk[PTB_STACK + 1];
  ShadowTrav tv;
  tv.begin(r, limit, pos, pu, pv, stk);
  const NodeOrder no = node_order(tv.r, tv.pre);
  while (tv.cur != ShadowTrav::DONE) {
    if (tv.at_inner()) tv.node_step(S, no, stk);
    else tv.leaf_step(S, stk);
  }
  return tv.tr;
}

}  // namespace ptb
