// kernels.cu — sm_100a wavefront kernels: raygen, extend (+alpha test), shade, shadow, resolve, tonemap,
// and the fixed-batch trace entry points.  Compiled with --fmad=false (see Makefile).
//
// Reference functions replaced (file:line relative to the reference root):
//   k_raygen   renderer/mod.rs:107-124, scene/internal/camera.rs:36-48
//   k_extend   renderer/utils.rs:11-21 + alpha walk renderer/mod.rs:188-205
//   k_shade    renderer/mod.rs:184-186,207-225,230-278; brdf/cook_torrance.rs; utils.rs:23-31
//   k_shadow   renderer/mod.rs:281-333
//   k_resolve  renderer/mod.rs:130 ; k_tonemap renderer/mod.rs:150-162,335-353
#include <atomic>
#include <cstdlib>
#include <mutex>

#include "kernels.cuh"
#include "traverse.cuh"

#ifndef PTB_EXTEND_MINB
// 64 registers / 8 CTAs per SM (28 B spilled) since the alpha walk lost its restart state (round 2): k_extend -2.5 %
// (console), -3 % (foliage) against 72 registers / 7 CTAs (profiles/r07/ab_s2.jsonl); in round 1 it was +-0
#define PTB_EXTEND_MINB 8
#endif
#ifndef PTB_SHADE_MINB
// Round 1 (IEEE shading arithmetic, 3 360 instructions): 64 registers / 8 CTAs per SM won (resident warps beat
// registers: 5 / 6 / 7 / 8 / 9 CTAs = +9 % / ref / -8 % / -11 % / -8 % on console).  With PTB_SHADE_FAST the kernel is a
// third shorter and the 140 B it spills at 64 registers dominate: 72 registers / 7 CTAs per SM is -36 % on console
// (3.10 -> 1.99 ms per 16 spp), +-1 % on foliage and city (profiles/r07/ab_s2.jsonl).
#define PTB_SHADE_MINB 7
#endif
#ifndef PTB_SHADOW_MINB
// 90 -> 72 registers: k_shadow -2 % (console) ... -18 % (foliage); 72 -> 64 registers / 8 CTAs per SM (40 B spilled):
// another -1 % (console) ... -4.5 % (foliage)
#define PTB_SHADOW_MINB 8
#endif
// 1: wave-state loads/stores use the streaming (evict-first) cache policy so that they do not displace BVH nodes
// (measured: k_shade -10 %, k_extend -2 % on the console workload).
#ifndef PTB_STREAM
#define PTB_STREAM 1
#endif
// ordered hit list between k_extend and k_shade (k_hit_list below); read by k_extend, k_shade and the launchers
#ifndef PTB_HIT_LIST
#define PTB_HIT_LIST 1
#endif
// instrumented build: k_extend / k_shadow count their node visits and primitive tests (printed by
// $PT_B200_DUMP_MARKS); never on in the shipped library (two more live registers per lane)
#ifndef PTB_COUNT_STEPS
#define PTB_COUNT_STEPS 0
#endif

namespace ptb {

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }
__device__ __forceinline__ uint32_t lanemask_lt() {
  uint32_t m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

// Warp-ballot work fetch: one atomic per warp hands out 32 consecutive queue positions.
__device__ __forceinline__ uint32_t fetch_work(uint32_t *cursor) {
  uint32_t base = 0;
  if (lane_id() == 0) base = atomicAdd(cursor, 32u);
  return __shfl_sync(0xffffffffu, base, 0) + lane_id();
}

// Warp-aggregated append: one atomic per warp for all lanes with `want`.
__device__ __forceinline__ uint32_t warp_append(uint32_t *counter, bool want) {
  uint32_t mask = __ballot_sync(0xffffffffu, want);
  if (mask == 0) return 0;
  int leader = __ffs(mask) - 1;
  uint32_t base = 0;
  if ((int)lane_id() == leader) base = atomicAdd(counter, (uint32_t)__popc(mask));
  base = __shfl_sync(0xffffffffu, base, leader);
  return base + (uint32_t)__popc(mask & lanemask_lt());
}

__device__ __forceinline__ void stat_add(uint32_t *counters, int which, unsigned long long v) {
  atomicAdd(reinterpret_cast<unsigned long long *>(counters + which), v);
}

// ------------------------------------------------------------------------------------------------
// Fixed ray batches (pt_trace_closest / pt_trace_accepted / pt_trace_shadow): the batch is loaded into the wave
// queues and traced by the SAME kernels that render (k_extend, k_shadow) - the parity harness runs the production
// traversal, not a side kernel.  Queue position q of the chunk holds ray `slot_base + q`; the slot doubles as the
// "pixel" of the RNG key in alpha mode.
__global__ void __launch_bounds__(256) k_load_rays(const float *__restrict__ rays, uint32_t n, uint32_t slot_base,
                                                   const __grid_constant__ WaveState W, uint32_t buf) {
  const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q == 0) {
    W.counters[CTR_QUEUE(buf)] = n;
    W.counters[CTR_WORK_EXTEND] = 0;
  }
  if (q >= n) return;
  const float *r = rays + (size_t)q * 6;
  W.ray_o[buf][q] = make_float4(r[0], r[1], r[2], 0.f);
  W.ray_d[buf][q] = make_float4(r[3], r[4], r[5], __int_as_float((int)(slot_base + q)));
}

// Shadow-queue entries of a batch: limits[q] >= 0 = point light at that distance from the primary hit prim[q] =
// (position xyz, uv of the primary hit; u NaN = the primary hit is a sphere); limits == nullptr: directional.
// The contribution is (1, 1, 1), so k_shadow leaves the transmittance itself in radiance[q].x.
__global__ void __launch_bounds__(256) k_load_shadow(const float *__restrict__ rays, const float *__restrict__ limits,
                                                     const float *__restrict__ prim, uint32_t n, WaveState W,
                                                     uint32_t which) {
  const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q == 0) {
    W.counters[CTR_SHADOW(which)] = n;
    W.counters[CTR_WORK_SHADOW] = 0;
  }
  if (q >= n) return;
  const float *r = rays + (size_t)q * 6;
  const float *p = prim ? prim + (size_t)q * 5 : nullptr;
  W.sh_o[q] = make_float4(r[0], r[1], r[2], limits ? limits[q] : -1.0f);
  W.sh_d[q] = make_float4(r[3], r[4], r[5], __int_as_float((int)q));
  W.sh_c[q] = make_float4(1.f, 1.f, 1.f, p ? p[3] : 0.f);
  W.sh_p[q] = make_float4(p ? p[0] : 0.f, p ? p[1] : 0.f, p ? p[2] : 0.f, p ? p[4] : 0.f);
  W.radiance[q] = make_float4(0.f, 0.f, 0.f, 0.f);
}
__global__ void __launch_bounds__(256) k_store_transmittance(WaveState W, uint32_t n, float *__restrict__ out) {
  const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q < n) out[q] = W.radiance[q].x;
}

// ------------------------------------------------------------------------------------------------
// Order of the (implicit) bounce-0 queue.  PTB_TILE_RAYGEN: consecutive queue positions walk 8x4 pixel tiles of one
// sample instead of image rows, so the 32 primary rays of a warp - and, because the queues keep their order through
// k_hit_list / k_shade, the bounce-1 rays that start where those hit - cover a compact patch of the image: fewer
// distinct BVH nodes and texels per warp (k_extend -1.8 % console, -2 % foliage, -7.3 % city, profiles/r07/ab_s3.jsonl).
// Pixels outside the largest 8x4-aligned rectangle keep row order.  The slot (pixel + sample * W*H) is what names the
// path, so results do not depend on this order.
#ifndef PTB_TILE_RAYGEN
#define PTB_TILE_RAYGEN 1
#endif
__device__ __forceinline__ uint32_t queue_to_slot(const WaveParams &P, uint32_t q) {
#if PTB_TILE_RAYGEN
  const uint32_t W = P.width, H = P.height, npix = W * H;
  const uint32_t k = q / npix, p = q - k * npix;
  const uint32_t W8 = W & ~7u, H4 = H & ~3u, n_main = W8 * H4;
  uint32_t x, y;
  if (p < n_main) {
    const uint32_t tile = p >> 5, in = p & 31u, tiles_x = W8 >> 3;
    const uint32_t ty = tile / tiles_x, tx = tile - ty * tiles_x;
    x = tx * 8u + (in & 7u);
    y = ty * 4u + (in >> 3);
  } else {
    uint32_t r = p - n_main;
    const uint32_t ws = W - W8;  // right strip (ws columns x H4 rows), then the bottom rows
    if (r < ws * H4) {
      y = r / ws;
      x = W8 + (r - y * ws);
    } else {
      r -= ws * H4;
      y = r / W;
      x = r - y * W;
      y += H4;
    }
  }
  return k * npix + y * W + x;
#else
  (void)P;
  return q;
#endif
}

// raygen (mod.rs:107-124): one thread per queue position of buffer 0 (already compact); zeroes the path's radiance sum.
// The throughput of a primary ray is 1 by definition and is not stored: k_hit_list / k_shade know it at bounce 0.
// (Regenerating the whole primary ray from q in k_extend / k_shade instead of storing it was measured: the 2 ms of this
// kernel go away, and come back as +1.4 ms in k_extend - ray generation inside its divergent refill path - and +0.4 ms
// in k_shade: no gain, profiles/r07/bench_s8_implicit_queue.json.)
__global__ void __launch_bounds__(256) k_raygen(const __grid_constant__ DevScene S, const __grid_constant__ WaveParams P,
                                                WaveState W) {
  const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q == 0) {
    W.counters[CTR_QUEUE(0)] = P.n_paths;
    W.counters[CTR_WORK_EXTEND] = 0;
  }
  if (q >= P.n_paths) return;
  const uint32_t slot = queue_to_slot(P, q);
  const Ray r = generate_ray(S, P, make_key(P, slot));
  __stcs(&W.ray_o[0][q], make_float4(r.o.x, r.o.y, r.o.z, 0.f));
  __stcs(&W.ray_d[0][q], make_float4(r.d.x, r.d.y, r.d.z, __int_as_float((int)slot)));
  __stcs(&W.radiance[slot], make_float4(0.f, 0.f, 0.f, 0.f));
}

// Lane-level work fetch: every idle lane of the warp takes the next queue position.  A warp reserves queue positions
// `chunk` at a time and always has ONE reservation in flight: the atomic that reserves the following chunk is issued
// when the current chunk is opened and its result is only read when that chunk runs out, so its round trip to L2
// (the largest single stall of k_shadow and 5 % of k_extend's in profiles/r01) overlaps with traversal work; as soon
// as the following chunk's base has arrived, the ray records it covers are prefetched into L2.
// `drained` becomes true (warp-uniformly) once the reservations have passed n.
// Measured (console / city / foliage, profiles/r02): k_shadow -26 % / +5 % / +1 %, k_extend +19 % / +6 % / +7 % (with or
// without spills: per-warp reservations widen the window of queue positions in flight across the GPU fourfold, and
// closest-hit traversal loses node locality in L1/L2), so the shadow kernel fetches ahead and the extend kernel does not.
#ifndef PTB_FETCH_AHEAD_EXTEND
#define PTB_FETCH_AHEAD_EXTEND 0
#endif
#ifdef PTB_FETCH_AHEAD_SHADOW
#define PTB_FETCH_AHEAD_SHADOW_FORCED 1
#endif
#ifndef PTB_FETCH_AHEAD_SHADOW
#define PTB_FETCH_AHEAD_SHADOW 1
#endif
#ifndef PTB_PF_L1
#define PTB_PF_L1 0
#endif
__device__ __forceinline__ void prefetch_l2(const void *p) {
#if PTB_PF_L1
  asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
#else
  asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
#endif
}

struct WorkFetch {
  uint32_t next, end;  // unconsumed part of the current reservation (warp-uniform)
  uint32_t ahead;      // lane 0: first position of the reservation requested last
  uint32_t chunk;      // queue positions per reservation
  bool ahead_fresh;    // the records of the `ahead` reservation have not been prefetched yet
  bool drained;

  __device__ __forceinline__ void init(uint32_t *cursor, uint32_t n) {
    const uint32_t total_warps = gridDim.x * (blockDim.x >> 5);
    chunk = n >= total_warps * 256u ? 64u : 32u;  // short queues: small chunks, so that every SM gets some work
    next = end = 0;
    ahead = 0;
    drained = false;
    ahead_fresh = true;
    if (lane_id() == 0) ahead = atomicAdd(cursor, chunk);
  }
  // pf(position) prefetches the records of one queue position.
  template <typename Prefetch>
  __device__ __forceinline__ uint32_t refill(uint32_t *cursor, bool idle, uint32_t n, Prefetch pf) {
    const uint32_t m = __ballot_sync(0xffffffffu, idle);
    uint32_t idx = 0xFFFFFFFFu;
    if (m && !drained) {
      const uint32_t cnt = (uint32_t)__popc(m), rank = (uint32_t)__popc(m & lanemask_lt()), avail = end - next;
      if (cnt < avail) {
        if (idle) idx = next + rank;
        next += cnt;
        if (next >= n) drained = true;  // every later position, reserved or not, is past the queue
        else if (ahead_fresh) {  // the reservation requested at the last switch has arrived by now: warm its records
          const uint32_t nb = __shfl_sync(0xffffffffu, ahead, 0);
          for (uint32_t k = lane_id(); k < chunk; k += 32u)
            if (nb + k < n) pf(nb + k);
          ahead_fresh = false;
        }
      } else {  // the current reservation runs out: continue in the one requested earlier, request another
        const uint32_t nb = __shfl_sync(0xffffffffu, ahead, 0);
        if (idle) idx = rank < avail ? next + rank : nb + (rank - avail);
        next = nb + (cnt - avail);
        end = nb + chunk;
        if (nb >= n) {
          drained = true;
        } else {
          if (lane_id() == 0) ahead = atomicAdd(cursor, chunk);
          ahead_fresh = true;
        }
      }
    }
    return idx < n ? idx : 0xFFFFFFFFu;
  }
};

// Synchronous variant: one atomic per refill, its result needed at once.
__device__ __forceinline__ uint32_t refill(uint32_t *cursor, bool idle, uint32_t n, bool &drained) {
  uint32_t m = __ballot_sync(0xffffffffu, idle);
  uint32_t idx = 0xFFFFFFFFu;
  if (m && !drained) {
    int leader = __ffs(m) - 1;
    uint32_t cnt = (uint32_t)__popc(m), base = 0;
    if ((int)lane_id() == leader) base = atomicAdd(cursor, cnt);
    base = __shfl_sync(0xffffffffu, base, leader);
    if (idle) idx = base + (uint32_t)__popc(m & lanemask_lt());
    if (base + cnt >= n) drained = true;
  }
  return idx < n ? idx : 0xFFFFFFFFu;
}

__device__ __forceinline__ void add_radiance(float4 *radiance, uint32_t slot, V3 c) {
  float *dst = reinterpret_cast<float *>(radiance + slot);
  if (c.x != 0.0f) atomicAdd(dst + 0, c.x);
  if (c.y != 0.0f) atomicAdd(dst + 1, c.y);
  if (c.z != 0.0f) atomicAdd(dst + 2, c.z);
}

// wave-state accesses: plain, or streaming (PTB_STREAM)
__device__ __forceinline__ float4 ld_wave(const float4 *p) {
#if PTB_STREAM
  return __ldcs(p);
#else
  return *p;
#endif
}
__device__ __forceinline__ void st_wave(float4 *p, float4 v) {
#if PTB_STREAM
  __stcs(p, v);
#else
  *p = v;
#endif
}

// Thresholds re-tuned on the round's last kernels (profiles/r07/ab_s24_thresholds.jsonl): refill at 16 busy lanes
// instead of 18: k_extend -0.7 % (console), -0.8 % (foliage), k_shadow -1 %; 20 is +2 %; leaf threshold 10 / 14: +-0.5 %.
#ifndef PTB_REFILL_T
#define PTB_REFILL_T 16
#endif
#ifndef PTB_LEAF_T
#define PTB_LEAF_T 12
#endif
// PTB_SPECULATE 1: speculative traversal in k_extend (see the kernel).  Not offered for k_shadow: its result is a
// product over the occluders, so the order in which a lane meets them - which would then depend on the lane's warp
// mates - shows in the last bit of the transmittance; the closest-hit walk is decided by exact keys and stays
// reproducible.  Measured (profiles/r07/ab_s32_speculative_traversal.jsonl; all 105 GPU tests green on the variant,
// identical checksums): k_extend +2.3 % on console, -0.2 % on foliage, +2 % on city - more lanes per node visit do
// not shorten a walk that waits for its slowest lane's memory, and the parked leaf costs wasted visits.  Off.
#ifndef PTB_SPECULATE
#define PTB_SPECULATE 0
#endif
#ifndef PTB_REFILL_T_SHADOW
#define PTB_REFILL_T_SHADOW PTB_REFILL_T
#endif
constexpr int REFILL_THRESHOLD = PTB_REFILL_T;  // refill when fewer lanes than this still own a ray
constexpr int REFILL_THRESHOLD_SHADOW = PTB_REFILL_T_SHADOW;
constexpr int LEAF_THRESHOLD = PTB_LEAF_T;      // run the (divergent) primitive test once this many lanes wait at a leaf

// extend: the hit that render_pixel shades (closest hit, alpha walk included) of every queued ray.  Persistent warps
// with lane-level ray replacement; inside a warp, lanes that reach a leaf wait until LEAF_THRESHOLD lanes have one (or
// nobody is at an inner node any more), so that node visits and primitive tests each run with most lanes active
// ("while-while").  alpha_on = 0 gives the first element of ray_cast's list instead (pt_trace_closest); trace_out,
// when not null, also receives the public hit record of every queue position (the fixed-batch parity harness).
// POP = where the lazy pop runs (traverse.cuh, PTB_POP_MODE): 0 inline at its two call sites, 1 at one site per loop
// iteration.  Measured (profiles/r07/ab_s14_pop_modes.jsonl): mode 1 is +0.8 % on console (5.4 node visits per ray: a
// leaf popped after a node step joins the primitive phase one iteration later), -3 % on foliage and city (21 - 23
// node visits per ray, many stale stack entries: the pop loop was 15 % of foliage's warp instructions at 3 lanes).
// launch_extend picks by tree depth.
template <int POP>
__global__ void __launch_bounds__(128, PTB_EXTEND_MINB) k_extend(const __grid_constant__ DevScene S, const __grid_constant__ WaveParams P,
                                                const __grid_constant__ WaveState W, uint32_t bounce, uint32_t alpha_on, pt_hit *__restrict__ trace_out) {
  const uint32_t in = bounce & 1u, outb = in ^ 1u;
  const uint32_t n = W.counters[CTR_QUEUE(in)];
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    W.counters[CTR_QUEUE(outb)] = 0;  // the queues k_shade(bounce) appends to
    W.counters[CTR_SHADOW(outb)] = 0;
    W.counters[CTR_HITS] = 0;  // filled by k_hit_list after this kernel
    stat_add(W.counters, CTR_STAT_EXTEND, n);
  }
  constexpr uint32_t NONE = 0xFFFFFFFFu;
  const bool alpha = alpha_on != 0;
  uint32_t n_alpha = 0;
#if PTB_COUNT_STEPS
  uint32_t cnt_nodes = 0, cnt_prims = 0;
#endif
  ClosestTrav tv;
  NodeOrder no{};
  ClosestStack stk;
  uint32_t q = NONE;
  tv.cur = ClosestTrav::DONE;
#if PTB_SPECULATE
  uint32_t held = ClosestTrav::NEED_POP;  // the parked leaf (NEED_POP: none); empty again whenever a ray ends
#endif
#if PTB_FETCH_AHEAD_EXTEND
  WorkFetch wf;
  wf.init(&W.counters[CTR_WORK_EXTEND], n);
  const float4 *pf_o = W.ray_o[in], *pf_d = W.ray_d[in];
#define drained wf.drained
#else
  bool drained = false;
#endif
  for (;;) {
#if PTB_FETCH_AHEAD_EXTEND
    uint32_t nq = wf.refill(&W.counters[CTR_WORK_EXTEND], q == NONE, n, [&](uint32_t p) {
      prefetch_l2(pf_o + p);
      prefetch_l2(pf_d + p);
    });
#else
    uint32_t nq = refill(&W.counters[CTR_WORK_EXTEND], q == NONE, n, drained);
#endif
    if (nq != NONE) {
      q = nq;
      const float4 o4 = ld_wave(&W.ray_o[in][q]), d4 = ld_wave(&W.ray_d[in][q]);
      Ray r;
      r.o = mk3(o4.x, o4.y, o4.z);
      r.d = mk3(d4.x, d4.y, d4.z);
      tv.begin(r, stk, alpha);
      no = node_order(tv.r, tv.pre);
    }
    if (__ballot_sync(0xffffffffu, q != NONE) == 0) break;
#if PTB_SPECULATE
    // Speculative traversal (Aila & Laine 2009): a lane that arrives at a leaf while the warp is not ready for a
    // primitive phase parks the leaf in `held` and goes on with its stack instead of idling through the others' node
    // visits; a primitive phase tests the parked leaf first.  The nodes it visits meanwhile may turn out to lie behind a
    // hit of the parked leaf (wasted visits, same result: candidate keys decide, not the order).  Pops at one site.
    for (;;) {
      const bool has = q != NONE;
#if PTB_COUNT_STEPS
      if (has && tv.at_inner()) ++cnt_nodes;
#endif
      if (has && tv.at_inner()) tv.node_step(S, no, stk, true);
      const bool cur_leaf = has && tv.at_leaf(), held_leaf = has && held != ClosestTrav::NEED_POP;
      const uint32_t m_leaf = __ballot_sync(0xffffffffu, cur_leaf || held_leaf);
      const uint32_t m_inner = __ballot_sync(0xffffffffu, has && tv.at_inner());
      if (__popc(m_leaf) >= LEAF_THRESHOLD || m_inner == 0) {
        if (cur_leaf || held_leaf) {
#if PTB_COUNT_STEPS
          ++cnt_prims;
#endif
          uint32_t ref = held_leaf ? held : tv.cur;
          tv.leaf_test(S, stk, ref, alpha, bounce,
                       [&]() { return make_key(P, (uint32_t)__float_as_int(ld_wave(&W.ray_d[in][q]).w)); }, n_alpha);
          if (held_leaf) held = ref;
          else tv.cur = ref;
        }
      } else if (cur_leaf && !held_leaf) {
        held = tv.cur;
        tv.cur = ClosestTrav::NEED_POP;
      }
      if (has && tv.cur == ClosestTrav::NEED_POP) tv.pop(stk);
      if (has && tv.cur == ClosestTrav::DONE && held == ClosestTrav::NEED_POP) {
        tv.finish(stk, alpha);
        const bool hit = tv.st.idx >= 0;
        st_wave(&W.hit[q], hit ? make_float4(tv.st.best_t, tv.st.u, tv.st.v, __int_as_float(tv.st.idx))
                               : make_float4(0.f, 0.f, 0.f, __int_as_float(-1)));
        if (trace_out) {
          pt_hit h;
          h.prim_id = hit ? tv.st.best_id : -1;
          h.t = hit ? tv.st.best_t : 0.f;
          h.u = hit ? tv.st.u : 0.f;
          h.v = hit ? tv.st.v : 0.f;
          trace_out[q] = h;
        }
        q = NONE;
      }
      const uint32_t busy = __ballot_sync(0xffffffffu, q != NONE);
      if (busy == 0 || (!drained && __popc(busy) < REFILL_THRESHOLD)) break;
    }
    continue;
#endif
    for (;;) {
      const bool has = q != NONE;
#if PTB_COUNT_STEPS
      if (has && tv.at_inner()) ++cnt_nodes;
#endif
      if (has && tv.at_inner()) tv.node_step(S, no, stk, POP != 0);
      const bool leaf = has && tv.at_leaf();
      const uint32_t m_leaf = __ballot_sync(0xffffffffu, leaf);
      const uint32_t m_inner = __ballot_sync(0xffffffffu, has && tv.at_inner());
#if PTB_COUNT_STEPS
      if (leaf && (__popc(m_leaf) >= LEAF_THRESHOLD || m_inner == 0)) ++cnt_prims;
#endif
      if (leaf && (__popc(m_leaf) >= LEAF_THRESHOLD || m_inner == 0))
        tv.leaf_step(S, stk, alpha, bounce,
                     [&]() {  // the path's slot names its RNG key; re-read from the queue record when a draw is needed
                       return make_key(P, (uint32_t)__float_as_int(ld_wave(&W.ray_d[in][q]).w));
                     },
                     n_alpha, POP != 0);
      if (POP == 1 && has && tv.cur == ClosestTrav::NEED_POP) tv.pop(stk);
      if (POP == 2 && has && tv.cur == ClosestTrav::NEED_POP) tv.pop_one(stk);
      if (has && tv.cur == ClosestTrav::DONE) {
        tv.finish(stk, alpha);
        const bool hit = tv.st.idx >= 0;
        st_wave(&W.hit[q], hit ? make_float4(tv.st.best_t, tv.st.u, tv.st.v, __int_as_float(tv.st.idx))
                               : make_float4(0.f, 0.f, 0.f, __int_as_float(-1)));
        if (trace_out) {
          pt_hit h;
          h.prim_id = hit ? tv.st.best_id : -1;
          h.t = hit ? tv.st.best_t : 0.f;
          h.u = hit ? tv.st.u : 0.f;
          h.v = hit ? tv.st.v : 0.f;
          trace_out[q] = h;
        }
        q = NONE;
      }
      const uint32_t busy = __ballot_sync(0xffffffffu, q != NONE);
      if (busy == 0 || (!drained && __popc(busy) < REFILL_THRESHOLD)) break;
    }
  }
#undef drained
  n_alpha = __reduce_add_sync(0xffffffffu, n_alpha);
  if (lane_id() == 0 && n_alpha) stat_add(W.counters, CTR_STAT_ALPHA, n_alpha);
#if PTB_COUNT_STEPS
  cnt_nodes = __reduce_add_sync(0xffffffffu, cnt_nodes);
  cnt_prims = __reduce_add_sync(0xffffffffu, cnt_prims);
  if (lane_id() == 0) {
    stat_add(W.counters, CTR_STAT_EXT_NODES, cnt_nodes);
    stat_add(W.counters, CTR_STAT_EXT_PRIMS, cnt_prims);
  }
#endif
}

// Warp-aggregated append to the path queue AND the shadow queue filled with it: one 64-bit atomic per warp reserves
// slots in both (low word = path queue, high word = shadow queue).  Returns the lane's positions.
__device__ __forceinline__ void warp_append2(uint32_t *pair, bool want_path, bool want_shadow, uint32_t &pos_path,
                                             uint32_t &pos_shadow) {
  const uint32_t mp = __ballot_sync(0xffffffffu, want_path), ms = __ballot_sync(0xffffffffu, want_shadow);
  pos_path = pos_shadow = 0;
  if ((mp | ms) == 0) return;
  unsigned long long base = 0;
  if (lane_id() == 0)
    base = atomicAdd(reinterpret_cast<unsigned long long *>(pair),
                     (unsigned long long)__popc(mp) | ((unsigned long long)__popc(ms) << 32));
  base = __shfl_sync(0xffffffffu, base, 0);
  pos_path = (uint32_t)base + (uint32_t)__popc(mp & lanemask_lt());
  pos_shadow = (uint32_t)(base >> 32) + (uint32_t)__popc(ms & lanemask_lt());
}

// The same reservation for a whole CTA: the warps add their counts into shared memory and ONE thread issues the
// global atomic.  Every warp of every resident CTA reserving on its own is one atomic per 32 shaded paths on a
// single 64-bit word; those serialise in the L2 slice that owns the word, and k_shade at bounce 0 (all lanes alive,
// every warp arriving at the atomic at the same rate) ran at exactly that rate: 1.1 ns per reservation, or 2.1 ns
// when the word happened to live in a slow slice - the address, hence the wave size and the GPU, decided whether
// the kernel took 2.8 or 5.4 ms (profiles/r07/counter_address_s6d.txt).  Four times fewer atomics put the
// reservation rate below the kernel's memory time again.  Must be called by all threads of the CTA.
#ifndef PTB_SHADE_CTA_APPEND
#define PTB_SHADE_CTA_APPEND 1
#endif
struct CtaAppend {  // shared-memory scratch of cta_append2, double-buffered by call parity (two barriers per call)
  unsigned long long cnt[2][4], base[2];
};
__device__ __forceinline__ void cta_append2(uint32_t *pair, bool want_path, bool want_shadow, uint32_t &pos_path,
                                            uint32_t &pos_shadow, CtaAppend &sh, uint32_t &call) {
  const uint32_t mp = __ballot_sync(0xffffffffu, want_path), ms = __ballot_sync(0xffffffffu, want_shadow);
  const uint32_t par = call++ & 1u, w = threadIdx.x >> 5;
  if (lane_id() == 0) sh.cnt[par][w] = (unsigned long long)__popc(mp) | ((unsigned long long)__popc(ms) << 32);
  __syncthreads();
  unsigned long long off = 0ull, total = 0ull;
#pragma unroll
  for (uint32_t j = 0; j < 4; ++j) {
    const unsigned long long c = sh.cnt[par][j];
    if (j < w) off += c;
    total += c;
  }
  if (threadIdx.x == 0) sh.base[par] = total ? atomicAdd(reinterpret_cast<unsigned long long *>(pair), total) : 0ull;
  __syncthreads();
  const unsigned long long base = sh.base[par] + off;
  pos_path = (uint32_t)base + (uint32_t)__popc(mp & lanemask_lt());
  pos_shadow = (uint32_t)(base >> 32) + (uint32_t)__popc(ms & lanemask_lt());
}

// PTB_HIT_LIST (on: console frame -5.9 %, foliage -0.3 ... -1.2 %, profiles/r06/ab_hitlist.jsonl): a scan pass between
// k_extend and k_shade lists the queue positions that hit something, in queue order up to 256-position chunks, so
// that every k_shade warp runs with 32 paths to shade instead of hits interleaved with misses (15 / 12 of 32 lanes
// on bounces 1 / 2 of the console scene).  Misses get their background term in the scan pass.  The list is a compact
// u32 array (WaveState::hit_list: 4 bytes per hit, written and read coalesced); its length is counters[CTR_HITS],
// reset by k_extend.
#if PTB_HIT_LIST
__global__ void __launch_bounds__(256) k_hit_list(const __grid_constant__ DevScene S, const __grid_constant__ WaveParams P,
                                                  const __grid_constant__ WaveState W, uint32_t bounce) {
  const uint32_t in = bounce & 1u;
  const uint32_t n = W.counters[CTR_QUEUE(in)];
  const V3 bgc = mk3(S.background[0], S.background[1], S.background[2]);
  const bool bg = bgc.x != 0.0f || bgc.y != 0.0f || bgc.z != 0.0f;
  constexpr uint32_t K = 8;  // positions per lane and iteration: one reservation atomic per 256 queue positions
  const uint32_t warps = gridDim.x * (blockDim.x >> 5), warp = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  for (uint32_t base = warp * (32u * K); base < n; base += warps * (32u * K)) {
    int idx[K];
#pragma unroll
    for (uint32_t k = 0; k < K; ++k) {
      const uint32_t q = base + k * 32u + lane_id();
      idx[k] = q < n ? __float_as_int(reinterpret_cast<const float *>(&W.hit[q])[3]) : -1;
    }
    uint32_t m[K], total = 0;
#pragma unroll
    for (uint32_t k = 0; k < K; ++k) {
      m[k] = __ballot_sync(0xffffffffu, idx[k] >= 0);
      total += (uint32_t)__popc(m[k]);
      const uint32_t q = base + k * 32u + lane_id();
      if (bg && q < n && idx[k] < 0) {  // mod.rs:184-186: color + throughput * background, and the path ends
        const uint32_t slot = (uint32_t)__float_as_int(__ldcs(&reinterpret_cast<const float *>(&W.ray_d[in][q])[3]));
        V3 t = mk3(1.f, 1.f, 1.f);  // bounce 0: the throughput of a primary ray is not stored
        if (bounce != 0) {
          const float4 t4 = ld_wave(&W.thr[in][q]);
          t = mk3(t4.x, t4.y, t4.z);
        }
        add_radiance(W.radiance, slot, mul_el(t, bgc));
      }
    }
    if (total == 0) continue;
    uint32_t pos = 0;
    if (lane_id() == 0) pos = atomicAdd(&W.counters[CTR_HITS], total);
    pos = __shfl_sync(0xffffffffu, pos, 0);
#pragma unroll
    for (uint32_t k = 0; k < K; ++k) {
      if (idx[k] >= 0) __stcs(&W.hit_list[pos + (uint32_t)__popc(m[k] & lanemask_lt())], base + k * 32u + lane_id());
      pos += (uint32_t)__popc(m[k]);
    }
  }
}
#endif

// shade: miss/background, emissive, next-event setup (shadow queue), GGX sample, throughput, cutoff, RR,
// and compaction of the surviving paths into the other queue buffer.  Queue positions are assigned statically
// (warp-strided): no work-fetch atomic; the last light's shadow entry and the next ray share one reservation atomic.
__global__ void __launch_bounds__(128, PTB_SHADE_MINB) k_shade(const __grid_constant__ DevScene S,
                                                               const __grid_constant__ WaveParams P, const __grid_constant__ WaveState W,
                                                               uint32_t bounce) {
  const uint32_t in = bounce & 1u, outb = in ^ 1u;
#if PTB_HIT_LIST
  const uint32_t n = W.counters[CTR_HITS];
#else
  const uint32_t n = W.counters[CTR_QUEUE(in)];
#endif
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    W.counters[CTR_WORK_EXTEND] = 0;
    // the shadow kernel's cursor is reset here, not in k_extend: k_shadow(bounce - 1) may still be running beside
    // k_extend(bounce) on another stream (render.cu), but it has always finished before this kernel starts
    W.counters[CTR_WORK_SHADOW] = 0;
  }
  uint32_t *pair = &W.counters[CTR_QUEUE(outb)];
  const uint32_t warps = gridDim.x * (blockDim.x >> 5), warp = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
#if PTB_SHADE_CTA_APPEND
  __shared__ CtaAppend s_app;
  uint32_t app_call = 0;
  // every warp runs the same number of rounds (the CTA synchronises inside cta_append2); warps past the end idle
  const uint32_t rounds = (n + warps * 32u - 1u) / (warps * 32u);
  for (uint32_t round = 0; round < rounds; ++round) {
    const uint32_t base = (round * warps + warp) * 32u;
#define PTB_APPEND2(a, b, c, d) cta_append2(pair, a, b, c, d, s_app, app_call)
#else
  for (uint32_t base = warp * 32u; base < n; base += warps * 32u) {
#define PTB_APPEND2(a, b, c, d) warp_append2(pair, a, b, c, d)
#endif
#if PTB_HIT_LIST
    const bool active = base + lane_id() < n;
    uint32_t q = 0;
    if (active) q = __ldcs(&W.hit_list[base + lane_id()]);
#else
    const uint32_t q = base + lane_id();
    const bool active = q < n;
#endif
    Ray r{};
    V3 thr{}, rad = mk3(0.f, 0.f, 0.f);
    ShadeCtx ctx{};
    uint32_t slot = 0;
    if (active) {
      const float4 o4 = ld_wave(&W.ray_o[in][q]), d4 = ld_wave(&W.ray_d[in][q]), h4 = ld_wave(&W.hit[q]);
      slot = (uint32_t)__float_as_int(d4.w);
      r.o = mk3(o4.x, o4.y, o4.z);
      r.d = mk3(d4.x, d4.y, d4.z);
      thr = mk3(1.f, 1.f, 1.f);  // bounce 0: not stored by k_raygen
      if (bounce != 0) {
        const float4 t4 = ld_wave(&W.thr[in][q]);
        thr = mk3(t4.x, t4.y, t4.z);
      }
      shade_begin(S, r, h4.x, h4.y, h4.z, __float_as_int(h4.w), thr, rad, ctx);
      add_radiance(W.radiance, slot, rad);  // background (miss) or emissive: usually zero, then no memory traffic
    }
    // next-event estimation set-up: one shadow-queue entry per light (mod.rs:248-262); the LAST light's reservation
    // is fused with the path-queue reservation below
    const uint32_t nl = S.n_lights;
    bool want = false;
    V3 so{}, sd{}, c{};
    float limit = -1.0f;
    for (uint32_t li = 0; li < nl; ++li) {
      want = false;
      if (ctx.lit) want = light_request(S, ctx, thr, li, so, sd, c, limit);
      if (li + 1 == nl) break;
      uint32_t pp, ps;
      PTB_APPEND2(false, want, pp, ps);
      if (want && ps >= W.shadow_capacity) atomicAdd(&W.counters[CTR_DROPPED], 1u);
      if (want && ps < W.shadow_capacity) {
        st_wave(&W.sh_o[ps], make_float4(so.x, so.y, so.z, limit));
        st_wave(&W.sh_d[ps], make_float4(sd.x, sd.y, sd.z, __int_as_float((int)slot)));
        st_wave(&W.sh_c[ps], make_float4(c.x, c.y, c.z, ctx.sf.is_sphere ? __int_as_float(0x7fc00000) : ctx.sf.tu));
        st_wave(&W.sh_p[ps], make_float4(ctx.sf.position.x, ctx.sf.position.y, ctx.sf.position.z, ctx.sf.tv));
      }
    }
    const float pu = ctx.sf.is_sphere ? __int_as_float(0x7fc00000) : ctx.sf.tu, pv = ctx.sf.tv;
    const V3 ppos = ctx.sf.position;
    Ray next{};
    bool alive = false;
    if (active) alive = shade_finish(P, make_key(P, slot), bounce, ctx, thr, next);
    uint32_t pp, ps;
    PTB_APPEND2(alive, want, pp, ps);
    if (want && ps >= W.shadow_capacity) atomicAdd(&W.counters[CTR_DROPPED], 1u);
    if (want && ps < W.shadow_capacity) {
      st_wave(&W.sh_o[ps], make_float4(so.x, so.y, so.z, limit));
      st_wave(&W.sh_d[ps], make_float4(sd.x, sd.y, sd.z, __int_as_float((int)slot)));
      st_wave(&W.sh_c[ps], make_float4(c.x, c.y, c.z, pu));
      st_wave(&W.sh_p[ps], make_float4(ppos.x, ppos.y, ppos.z, pv));
    }
    if (alive) {
      st_wave(&W.ray_o[outb][pp], make_float4(next.o.x, next.o.y, next.o.z, 0.f));
      st_wave(&W.ray_d[outb][pp], make_float4(next.d.x, next.d.y, next.d.z, __int_as_float((int)slot)));
      st_wave(&W.thr[outb][pp], make_float4(thr.x, thr.y, thr.z, 0.f));
    }
  }
#undef PTB_APPEND2
}

// One thread follows one whole path (render_pixel, mod.rs:172-228) and records every bounce: the per-vertex
// parity harness (pt_trace_paths).  Lights are evaluated in file order with the shadow walk inline.
__global__ void __launch_bounds__(128) k_trace_paths(const __grid_constant__ DevScene S,
                                                     const __grid_constant__ WaveParams P,
                                                     const uint32_t *__restrict__ pixels, uint32_t n_paths,
                                                     pt_path_vertex *__restrict__ out, uint32_t *__restrict__ counts) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_paths) return;
  RngKey key;
  key.pixel = pixels[i];
  key.sample = P.sample_first;
  key.seed_lo = P.seed_lo;
  key.seed_hi = P.seed_hi;
  Ray r = generate_ray(S, P, key);
  V3 thr = mk3(1.f, 1.f, 1.f), rad = mk3(0.f, 0.f, 0.f);
  pt_path_vertex *rec = out + (size_t)i * (P.bounces + 1);
  uint32_t nv = 0, n_alpha = 0;
  for (uint32_t bounce = 0; bounce <= P.bounces; ++bounce) {
    TraceState st;
    trace_closest(S, r, S.any_alpha != 0, key, bounce, st, n_alpha);
    ShadeCtx ctx;
    shade_begin(S, r, st.best_t, st.u, st.v, st.idx, thr, rad, ctx);
    if (ctx.lit) {
      for (uint32_t li = 0; li < S.n_lights; ++li) {
        V3 so, sd, c;
        float limit;
        if (!light_request(S, ctx, thr, li, so, sd, c, limit)) continue;
        Ray sr;
        sr.o = so;
        sr.d = sd;
        float tr = trace_transmittance(S, sr, limit, ctx.sf.position,
                                       ctx.sf.is_sphere ? __int_as_float(0x7fc00000) : ctx.sf.tu, ctx.sf.tv);
        if (tr != 0.0f) rad = rad + c * tr;
      }
    }
    Ray next{};
    bool alive = shade_finish(P, key, bounce, ctx, thr, next);
    pt_path_vertex v;
    v.ray_o[0] = r.o.x; v.ray_o[1] = r.o.y; v.ray_o[2] = r.o.z;
    v.ray_d[0] = r.d.x; v.ray_d[1] = r.d.y; v.ray_d[2] = r.d.z;
    v.prim_id = st.idx < 0 ? -1 : st.best_id;
    v.t = st.idx < 0 ? 0.f : st.best_t;
    v.throughput[0] = thr.x; v.throughput[1] = thr.y; v.throughput[2] = thr.z;
    v.radiance[0] = rad.x; v.radiance[1] = rad.y; v.radiance[2] = rad.z;
    rec[nv++] = v;
    if (!alive) break;
    r = next;
  }
  counts[i] = nv;
}

// Triangle::intersect (scene/internal/triangle.rs:37-82) on independent (ray, triangle) pairs: the entry point
// behind the reference's own hit/miss unit vectors (triangle.rs:200-233).
__global__ void __launch_bounds__(256) k_triangle_intersect(const float *__restrict__ rays, const float *__restrict__ tris,
                                                            uint32_t n, pt_hit *__restrict__ out) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float *q = rays + (size_t)i * 6, *p = tris + (size_t)i * 9;
  Ray r;
  r.o = mk3(q[0], q[1], q[2]);
  r.d = mk3(q[3], q[4], q[5]);
  V3 v0 = mk3(p[0], p[1], p[2]);
  V3 e1 = mk3(p[3] - p[0], p[4] - p[1], p[5] - p[2]);
  V3 e2 = mk3(p[6] - p[0], p[7] - p[1], p[8] - p[2]);
  float t = 0.f, u = 0.f, v = 0.f, det = 0.f;
  bool hit = moller_trumbore(r, v0, e1, e2, t, u, v, det);
  pt_hit h;
  h.prim_id = hit ? (det < 0.0f ? 1 : 0) : -1;  // 0 front face, 1 back face (is_backface = det < 0)
  h.t = hit ? t : 0.f;
  h.u = hit ? u : 0.f;
  h.v = hit ? v : 0.f;
  out[i] = h;
}

// The device generator on caller-chosen (counter, key) pairs: the known-answer hook (pt_philox4x32_10).
__global__ void __launch_bounds__(256) k_philox(const uint32_t *__restrict__ ctr, const uint32_t *__restrict__ key, uint32_t n,
                                                uint32_t *__restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint4 r = philox4x32_10(make_uint4(ctr[i * 4], ctr[i * 4 + 1], ctr[i * 4 + 2], ctr[i * 4 + 3]), key[i * 2], key[i * 2 + 1]);
  out[i * 4] = r.x;
  out[i * 4 + 1] = r.y;
  out[i * 4 + 2] = r.z;
  out[i * 4 + 3] = r.w;
}

// debug_render / render_debug_pixels (renderer/debug_renderer.rs:11-105): one pixel-centre ray per pixel, the first
// element of ray_cast (no alpha walk), seven AOV planes: normal, albedo, opacity, metalness, roughness, emissive, ior.
// planes: 7 * W*H*3 bytes (caller-zeroed: pixels without a hit stay 0); n_hit counts the pixels that hit something.
__global__ void __launch_bounds__(128) k_debug_aov(const __grid_constant__ DevScene S, uint32_t width, uint32_t height,
                                                   uint8_t *__restrict__ planes, uint32_t *n_hit) {
  const uint32_t npix = width * height;
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  bool got = false;
  if (p < npix) {
    const uint32_t x = p % width, y = p / width;
    const float wf = (float)width, hf = (float)height, ratio = wf / hf, tn = S.tan_half_fov;
    float sx = (float)x + 0.5f;  // debug_renderer.rs:24-30
    sx = sx / wf * 2.0f - 1.0f;
    sx *= tn * ratio;
    float sy = (float)y + 0.5f;
    sy = 1.0f - sy / hf * 2.0f;
    sy *= tn;
    V3 dir = normalize3(mk3(sx, sy, -1.0f));
    const float *m = S.camera;
    V3 c0 = mk3(m[0], m[1], m[2]), c1 = mk3(m[4], m[5], m[6]), c2 = mk3(m[8], m[9], m[10]), c3 = mk3(m[12], m[13], m[14]);
    Ray r;
    r.d = ((c0 * dir.x + c1 * dir.y) + c2 * dir.z) + c3 * 0.0f;
    r.o = c3;
    TraceState st;
    uint32_t unused = 0;
    trace_closest(S, r, false, RngKey{}, 0u, st, unused);
    if (st.idx >= 0) {
      got = true;
      const Surface sf = make_surface(S, r, st.best_t, st.u, st.v, st.idx);
      const MatSample ms = material_sample(S, sf.model, sf.is_sphere, sf.tu, sf.tv);
      const float ior = S.materials[sf.model].ior;
      const size_t plane = (size_t)npix * 3;
      auto put = [&](int k, V3 c) {  // `(pixel.x * 255.) as u8`, debug_renderer.rs:53-57
        uint8_t *o = planes + plane * k + (size_t)p * 3;
        o[0] = to_u8(c.x * 255.0f);
        o[1] = to_u8(c.y * 255.0f);
        o[2] = to_u8(c.z * 255.0f);
      };
      const V3 one = mk3(1.0f, 1.0f, 1.0f);
      put(0, sf.shading_n * 0.5f + mk3(0.5f, 0.5f, 0.5f));
      put(1, ms.albedo);
      put(2, one * ms.opacity);
      put(3, one * ms.metalness);
      put(4, one * ms.roughness);
      put(5, ms.emissive);
      put(6, one * ior / 3.0f);
    }
  }
  const uint32_t m = __ballot_sync(0xffffffffu, got);
  if (lane_id() == 0 && m) atomicAdd(n_hit, (uint32_t)__popc(m));
}

// shadow: transmittance of every queued shadow ray; adds contribution * transmittance to the path's radiance.
// Same lane-level ray replacement and node / primitive phases as k_extend.
// FETCH = the asynchronous work fetch with record prefetch (WorkFetch above): -32 % on the console scene, whose shadow
// walks are short (5.9 node visits) and whose queue records come from DRAM; +6 % on foliage (20.7 node visits per ray:
// refills are rare and the chunked reservation only costs balance), profiles/r07/ab_s22_occupancy_fetch_ahead.jsonl.
// launch_shadow picks by tree depth, like launch_extend; -DPTB_FETCH_AHEAD_SHADOW=<0|1> forces one.
template <int FETCH>
__global__ void __launch_bounds__(128, PTB_SHADOW_MINB) k_shadow(const __grid_constant__ DevScene S, const __grid_constant__ WaveState W, uint32_t bounce) {
  uint32_t n = W.counters[CTR_SHADOW((bounce & 1u) ^ 1u)];
  if (n > W.shadow_capacity) n = W.shadow_capacity;
  if (blockIdx.x == 0 && threadIdx.x == 0) stat_add(W.counters, CTR_STAT_SHADOW, n);
  constexpr uint32_t NONE = 0xFFFFFFFFu;
  ShadowTrav tv;
  NodeOrder no{};
  uint32_t stk[PTB_STACK + 1];  // + the sentinel
  uint32_t q = NONE, slot = 0;
  V3 c{};
  tv.cur = ShadowTrav::DONE;
#if PTB_SHADOW_CACHE
  uint32_t last_occ = NONE;  // the opaque triangle that shadowed this lane's previous ray
#endif
#if PTB_COUNT_STEPS
  uint32_t cnt_nodes = 0, cnt_prims = 0;
#endif
  WorkFetch wf;
  bool drained_sync = false;
  if (FETCH) wf.init(&W.counters[CTR_WORK_SHADOW], n);
  for (;;) {
    uint32_t nq;
    if (FETCH)
      nq = wf.refill(&W.counters[CTR_WORK_SHADOW], q == NONE, n, [&](uint32_t p) {
        prefetch_l2(W.sh_o + p);
        prefetch_l2(W.sh_d + p);
        prefetch_l2(W.sh_c + p);
        prefetch_l2(W.sh_p + p);
      });
    else
      nq = refill(&W.counters[CTR_WORK_SHADOW], q == NONE, n, drained_sync);
    const bool drained = FETCH ? wf.drained : drained_sync;
    if (nq != NONE) {
      q = nq;
      float4 o4 = ld_wave(&W.sh_o[q]), d4 = ld_wave(&W.sh_d[q]), c4 = ld_wave(&W.sh_c[q]), p4 = ld_wave(&W.sh_p[q]);
      Ray r;
      r.o = mk3(o4.x, o4.y, o4.z);
      r.d = mk3(d4.x, d4.y, d4.z);
      slot = (uint32_t)__float_as_int(d4.w);
      c = mk3(c4.x, c4.y, c4.z);
      tv.begin(r, o4.w, mk3(p4.x, p4.y, p4.z), c4.w, p4.w, stk);
      no = node_order(tv.r, tv.pre);
#if PTB_SHADOW_CACHE
      tv.occluder = NONE;
      if (last_occ != NONE) tv.try_cached(S, last_occ);
#endif
    }
    if (__ballot_sync(0xffffffffu, q != NONE) == 0) break;
    for (;;) {
      const bool has = q != NONE;
#if PTB_COUNT_STEPS
      if (has && tv.at_inner()) ++cnt_nodes;
#endif
      if (has && tv.at_inner()) tv.node_step(S, no, stk);
      const bool leaf = has && tv.at_leaf();
      const uint32_t m_leaf = __ballot_sync(0xffffffffu, leaf);
      const uint32_t m_inner = __ballot_sync(0xffffffffu, has && tv.at_inner());
#if PTB_COUNT_STEPS
      if (leaf && (__popc(m_leaf) >= LEAF_THRESHOLD || m_inner == 0)) ++cnt_prims;
#endif
      if (leaf && (__popc(m_leaf) >= LEAF_THRESHOLD || m_inner == 0)) tv.leaf_step(S, stk);
      if (has && tv.cur == ShadowTrav::DONE) {
#if PTB_SHADOW_CACHE
        if (tv.occluder != NONE) last_occ = tv.occluder;  // a cache hit leaves it as it is
#endif
        if (tv.tr != 0.0f) {
          float *dst = reinterpret_cast<float *>(W.radiance + slot);
          atomicAdd(dst + 0, c.x * tv.tr);
          atomicAdd(dst + 1, c.y * tv.tr);
          atomicAdd(dst + 2, c.z * tv.tr);
        }
        q = NONE;
      }
      const uint32_t busy = __ballot_sync(0xffffffffu, q != NONE);
      if (busy == 0 || (!drained && __popc(busy) < REFILL_THRESHOLD_SHADOW)) break;
    }
  }
#if PTB_COUNT_STEPS
  cnt_nodes = __reduce_add_sync(0xffffffffu, cnt_nodes);
  cnt_prims = __reduce_add_sync(0xffffffffu, cnt_prims);
  if (lane_id() == 0) {
    stat_add(W.counters, CTR_STAT_SH_NODES, cnt_nodes);
    stat_add(W.counters, CTR_STAT_SH_PRIMS, cnt_prims);
  }
#endif
}

// resolve: *pixel += color (mod.rs:130) for the wave's samples of each pixel, in sample order.
__global__ void __launch_bounds__(256) k_resolve(const __grid_constant__ WaveParams P, WaveState W, float *__restrict__ accum) {
  uint32_t npix = P.width * P.height;
  uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= npix) return;
  float sx = accum[p * 3 + 0], sy = accum[p * 3 + 1], sz = accum[p * 3 + 2];
  for (uint32_t k = 0; k < P.spp_in_wave; ++k) {
    float4 c = W.radiance[p + k * npix];
    sx += c.x;
    sy += c.y;
    sz += c.z;
  }
  accum[p * 3 + 0] = sx;
  accum[p * 3 + 1] = sy;
  accum[p * 3 + 2] = sz;
}

__global__ void __launch_bounds__(256) k_tonemap(const float *__restrict__ accum, uint32_t npix, float total_samples,
                                                 int tonemap, uint8_t *__restrict__ rgb8) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npix * 3) return;
  float c = accum[i] / total_samples;
  float t = tonemap1(tonemap, c);
  rgb8[i] = to_u8(powf(t, 1.0f / 2.2f) * 255.0f);
}

// Sum of two accumulation buffers (pt_render_multi: the per-device partial sums are added on the root device in
// replica order, so the result does not depend on which device finished first).
__global__ void __launch_bounds__(256) k_accum_add(float *__restrict__ dst, const float *__restrict__ src, uint32_t n) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) dst[i] += src[i];
}

// sRGB -> linear table: the albedo texel conversion of material.rs:132-146, evaluated once per texel value.
// Second half: texel / 255 (the metalness / roughness / opacity / emissive scalings of material.rs:152-214).
__global__ void k_init_srgb_lut(float *lut) {
  lut[threadIdx.x] = powf((float)threadIdx.x / 255.0f, 2.2f);
  lut[256 + threadIdx.x] = (float)threadIdx.x / 255.0f;
}

// ------------------------------------------------------------------------------------------------
// host-side launchers
// Per-device launch geometry, cached on first use.  pt_render_multi launches from one host thread per device, so
// the cache is indexed by the current device and its entries are atomics (two threads may fill one entry with the
// same value).
constexpr int MAX_DEVICES = 64;
static int current_device() {
  int dev = 0;
  cudaGetDevice(&dev);
  return dev >= 0 && dev < MAX_DEVICES ? dev : 0;
}
static int sm_count() {
  static std::atomic<int> cache[MAX_DEVICES];
  const int dev = current_device();
  int n = cache[dev].load(std::memory_order_relaxed);
  if (n == 0) {
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
    cache[dev].store(n, std::memory_order_relaxed);
  }
  return n;
}
constexpr int PERSIST_CTAS_PER_SM = 8;
// Grid of a persistent kernel: exactly the CTAs that are resident at once (occupancy API), so that statically
// partitioned kernels have no second round of CTAs.
template <typename K>
static int resident_grid(K kernel, int threads) {
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, 0) != cudaSuccess || per_sm <= 0) per_sm = 4;
  return sm_count() * per_sm;
}

void launch_load_rays(const float *rays, uint32_t n, uint32_t slot_base, const WaveState &W, uint32_t buf, cudaStream_t st) {
  k_load_rays<<<(n + 255) / 256, 256, 0, st>>>(rays, n, slot_base, W, buf);
}
void launch_load_shadow(const float *rays, const float *limits, const float *prim, uint32_t n, const WaveState &W,
                        uint32_t which, cudaStream_t st) {
  k_load_shadow<<<(n + 255) / 256, 256, 0, st>>>(rays, limits, prim, n, W, which);
}
void launch_store_transmittance(const WaveState &W, uint32_t n, float *out, cudaStream_t st) {
  k_store_transmittance<<<(n + 255) / 256, 256, 0, st>>>(W, n, out);
}
void launch_trace_paths(const DevScene &S, const WaveParams &P, const uint32_t *pixels, uint32_t n, pt_path_vertex *out,
                        uint32_t *counts, cudaStream_t st) {
  k_trace_paths<<<(n + 127) / 128, 128, 0, st>>>(S, P, pixels, n, out, counts);
}
void launch_philox(const uint32_t *ctr, const uint32_t *key, uint32_t n, uint32_t *out, cudaStream_t st) {
  k_philox<<<(n + 255) / 256, 256, 0, st>>>(ctr, key, n, out);
}
void launch_triangle_intersect(const float *rays, const float *tris, uint32_t n, pt_hit *out, cudaStream_t st) {
  k_triangle_intersect<<<(n + 255) / 256, 256, 0, st>>>(rays, tris, n, out);
}
void launch_debug_aov(const DevScene &S, uint32_t width, uint32_t height, uint8_t *planes, uint32_t *n_hit, cudaStream_t st) {
  k_debug_aov<<<(width * height + 127) / 128, 128, 0, st>>>(S, width, height, planes, n_hit);
}
void launch_init_srgb_lut(float *lut, cudaStream_t st) { k_init_srgb_lut<<<1, 256, 0, st>>>(lut); }
void launch_raygen(const DevScene &S, const WaveParams &P, const WaveState &W, cudaStream_t st) {
  k_raygen<<<(P.n_paths + 255) / 256, 256, 0, st>>>(S, P, W);
}
void launch_extend(const DevScene &S, const WaveParams &P, const WaveState &W, uint32_t bounce, bool alpha,
                   pt_hit *trace_out, cudaStream_t st) {
  // deep trees (long traversals with many stale stack entries) take the merged pop site; PTB_POP_MODE forces one
  const int grid = sm_count() * PERSIST_CTAS_PER_SM;
#if defined(PTB_POP_MODE_FORCED)
  k_extend<PTB_POP_MODE><<<grid, 128, 0, st>>>(S, P, W, bounce, alpha ? 1u : 0u, trace_out);
#else
  if (S.tree_depth >= 16) k_extend<1><<<grid, 128, 0, st>>>(S, P, W, bounce, alpha ? 1u : 0u, trace_out);
  else k_extend<0><<<grid, 128, 0, st>>>(S, P, W, bounce, alpha ? 1u : 0u, trace_out);
#endif
}
void launch_shade(const DevScene &S, const WaveParams &P, const WaveState &W, uint32_t bounce, cudaStream_t st) {
  static std::atomic<int> cache[MAX_DEVICES];
  const int dev = current_device();
  int grid = cache[dev].load(std::memory_order_relaxed);
  if (!grid) {
    grid = resident_grid(k_shade, 128);
    cache[dev].store(grid, std::memory_order_relaxed);
  }
  k_shade<<<grid, 128, 0, st>>>(S, P, W, bounce);
}
// the scan pass between k_extend and k_shade (a no-op build without PTB_HIT_LIST)
void launch_hit_list(const DevScene &S, const WaveParams &P, const WaveState &W, uint32_t bounce, cudaStream_t st) {
#if PTB_HIT_LIST
  k_hit_list<<<sm_count() * 8, 256, 0, st>>>(S, P, W, bounce);
#else
  (void)S; (void)P; (void)W; (void)bounce; (void)st;
#endif
}
int shade_launches() { return 1 + PTB_HIT_LIST; }
void launch_shadow(const DevScene &S, const WaveState &W, uint32_t bounce, cudaStream_t st) {
  const int grid = sm_count() * PERSIST_CTAS_PER_SM;
#if defined(PTB_FETCH_AHEAD_SHADOW_FORCED)
  k_shadow<PTB_FETCH_AHEAD_SHADOW><<<grid, 128, 0, st>>>(S, W, bounce);
#else
  if (S.tree_depth >= 16) k_shadow<0><<<grid, 128, 0, st>>>(S, W, bounce);  // long walks: plain work fetch
  else k_shadow<1><<<grid, 128, 0, st>>>(S, W, bounce);
#endif
}
void launch_resolve(const WaveParams &P, const WaveState &W, float *accum, cudaStream_t st) {
  uint32_t npix = P.width * P.height;
  k_resolve<<<(npix + 255) / 256, 256, 0, st>>>(P, W, accum);
}
void launch_accum_add(float *dst, const float *src, uint32_t n, cudaStream_t st) {
  k_accum_add<<<sm_count() * 8, 256, 0, st>>>(dst, src, n);
}
void launch_tonemap(const float *accum, uint32_t npix, uint32_t total_samples, int tonemap, uint8_t *rgb8, cudaStream_t st) {
  k_tonemap<<<(npix * 3 + 255) / 256, 256, 0, st>>>(accum, npix, (float)total_samples, tonemap, rgb8);
}

// ------------------------------------------------------------------------------------------------
// pt_trace_closest_device_noalloc: the closest-hit batch of pt_trace_closest_device for callers that must not allocate
// (CUDA graph capture, serving loops).  k_extend needs wave scratch per ray - the queue record (origin, direction +
// slot) it reads, the internal hit record it writes beside the public one - plus the queue counters; here that
// scratch is a file-scope device array instead of the scene's cudaMalloc'ed wave arena: 48 bytes per ray for at most
// PT_TRACE_NOALLOC_MAX_RAYS rays + 256 bytes of counters = 3 MiB + 256 B, held by every GPU that loads the library
// until the process exits.  Every call (any scene, any stream) uses the same array, so calls that overlap on the GPU
// race; calls on one stream are ordered.  The array keeps the previous call's contents: the counters are cleared on
// the stream before the batch is loaded, the ray records of [0, n) are rewritten, the rest is never read.
constexpr size_t TRACE_STATIC_CTR_BYTES = 256;  // CTR_COUNT uint32 counters, padded to the array's alignment
static_assert(CTR_COUNT * sizeof(uint32_t) <= TRACE_STATIC_CTR_BYTES, "counter block too small");
constexpr size_t TRACE_STATIC_BYTES = TRACE_STATIC_CTR_BYTES + 3 * sizeof(float4) * (size_t)PT_TRACE_NOALLOC_MAX_RAYS;
static __device__ __align__(256) unsigned char g_trace_scratch[TRACE_STATIC_BYTES];

// In host code `g_trace_scratch` is only a handle of the symbol: its device address comes from cudaGetSymbolAddress,
// looked up once per device with that device current (each device has its own instance of the array).
static cudaError_t trace_scratch_address(unsigned char **out) {
  static std::mutex mu;
  static unsigned char *addr[MAX_DEVICES];
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= MAX_DEVICES) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> hold(mu);
  if (!addr[dev]) {
    void *p = nullptr;
    e = cudaGetSymbolAddress(&p, g_trace_scratch);
    if (e != cudaSuccess) return e;
    if (!p) return cudaErrorInvalidSymbol;
    addr[dev] = static_cast<unsigned char *>(p);
  }
  *out = addr[dev];
  return cudaSuccess;
}

// Current device = the scene's.  No allocation: one memset node, k_load_rays and the production k_extend on `st`.
cudaError_t launch_trace_closest_static(const DevScene &S, const float *rays_dev, uint32_t n, pt_hit *out_dev, cudaStream_t st) {
  if (n == 0) return cudaSuccess;
  if (n > (uint32_t)PT_TRACE_NOALLOC_MAX_RAYS) return cudaErrorInvalidValue;
  unsigned char *base = nullptr;
  cudaError_t e = trace_scratch_address(&base);
  if (e != cudaSuccess) return e;
  float4 *rec = reinterpret_cast<float4 *>(base + TRACE_STATIC_CTR_BYTES);
  WaveState W{};
  W.counters = reinterpret_cast<uint32_t *>(base);
  W.ray_o[0] = W.ray_o[1] = rec;
  W.ray_d[0] = W.ray_d[1] = rec + (size_t)PT_TRACE_NOALLOC_MAX_RAYS;
  W.hit = rec + 2 * (size_t)PT_TRACE_NOALLOC_MAX_RAYS;
  W.capacity = (uint32_t)PT_TRACE_NOALLOC_MAX_RAYS;
  WaveParams P{};
  P.width = 0x80000000u;  // make_key: pixel = slot (not used without the alpha walk)
  P.height = 1;
  P.spp_in_wave = 1;
  P.sample_stride = 1;
  P.n_paths = n;
  e = cudaMemsetAsync(base, 0, TRACE_STATIC_CTR_BYTES, st);  // queue sizes, work cursor, statistics of the last call
  if (e != cudaSuccess) return e;
  launch_load_rays(rays_dev, n, 0, W, 0, st);
  launch_extend(S, P, W, 0, false, out_dev, st);
  return cudaGetLastError();
}

}  // namespace ptb
