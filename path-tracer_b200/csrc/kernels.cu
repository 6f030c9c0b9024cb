// kernels.cu — sm_100a wavefront kernels: raygen, extend (+alpha test), shade, shadow, resolve, tonemap,
// and the fixed-batch trace entry points.  Compiled with --fmad=false (see Makefile).
//
// Reference functions replaced (file:line relative to the reference root):
//   k_raygen   renderer/mod.rs:107-124, scene/internal/camera.rs:36-48
//   k_extend   renderer/utils.rs:11-21 + alpha walk renderer/mod.rs:188-205
//   k_shade    renderer/mod.rs:184-186,207-225,230-278; brdf/cook_torrance.rs; utils.rs:23-31
//   k_shadow   renderer/mod.rs:281-333
//   k_resolve  renderer/mod.rs:130 ; k_tonemap renderer/mod.rs:150-162,335-353
#include "kernels.cuh"
#include "shade.cuh"

namespace ptb {

constexpr float NORMAL_BIAS = 0.00001f;  // renderer/mod.rs:59

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

// ------------------------------------------------------------------------------------------------
// Fixed ray batch -> first element of ray_cast's sorted list (pt_trace_closest).
__global__ void __launch_bounds__(128) k_trace_closest(const __grid_constant__ DevScene S, const float *__restrict__ rays,
                                                       uint32_t n, pt_hit *__restrict__ out, uint32_t *cursor) {
  for (;;) {
    uint32_t q = fetch_work(cursor);
    if (__all_sync(0xffffffffu, q >= n)) break;
    if (q >= n) continue;
    Ray r;
    r.o = mk3(rays[q * 6 + 0], rays[q * 6 + 1], rays[q * 6 + 2]);
    r.d = mk3(rays[q * 6 + 3], rays[q * 6 + 4], rays[q * 6 + 5]);
    TraceState st;
    trace_closest(S, r, -1.0f, -1, st);
    pt_hit h;
    h.prim_id = st.idx < 0 ? -1 : st.best_id;
    h.t = st.idx < 0 ? 0.f : st.best_t;
    h.u = st.u;
    h.v = st.v;
    out[q] = h;
  }
}

// All-hits transmittance walk shared by k_shadow and k_trace_transmittance.
//   limit < 0: directional light (every hit on the ray, occluder sampled at its own uv, mod.rs:291-297)
//   limit >= 0: point light at distance `limit` from `pos`; occluders farther than the light are ignored and
//   the occluder's material is sampled with the PRIMARY hit's uv/variant (mod.rs:319-329, puv.x NaN = sphere).
__device__ __forceinline__ float trace_transmittance(const DevScene &S, const Ray &r, float limit, V3 pos, float pu,
                                                     float pv) {
  const RayPre pre = precompute(r);
  const bool point = limit >= 0.0f;
  float t_max = CUDART_INF_F;
  if (point) {
    float len = length3(r.d);
    float bias = length3(r.o - pos);
    t_max = len > 0.f ? (limit + bias) / len * 1.0002f + 1e-6f : CUDART_INF_F;
  }
  float tr = 1.0f;
  uint32_t stk[PTB_STACK];
  int sp = 0;
  uint32_t cur = 0;
  for (;;) {
    if (!(cur & 0x80000000u)) {
      const float4 *np = reinterpret_cast<const float4 *>(S.nodes + cur);
      float4 lx = __ldg(np), ly = __ldg(np + 1), lz = __ldg(np + 2);
      float4 hx = __ldg(np + 3), hy = __ldg(np + 4), hz = __ldg(np + 5);
      uint4 ch = __ldg(reinterpret_cast<const uint4 *>(np + 6));
      if (ch.x != 0xFFFFFFFFu && slab(lx.x, ly.x, lz.x, hx.x, hy.x, hz.x, r, pre, t_max) < CUDART_INF_F && sp < PTB_STACK) stk[sp++] = ch.x;
      if (ch.y != 0xFFFFFFFFu && slab(lx.y, ly.y, lz.y, hx.y, hy.y, hz.y, r, pre, t_max) < CUDART_INF_F && sp < PTB_STACK) stk[sp++] = ch.y;
      if (ch.z != 0xFFFFFFFFu && slab(lx.z, ly.z, lz.z, hx.z, hy.z, hz.z, r, pre, t_max) < CUDART_INF_F && sp < PTB_STACK) stk[sp++] = ch.z;
      if (ch.w != 0xFFFFFFFFu && slab(lx.w, ly.w, lz.w, hx.w, hy.w, hz.w, r, pre, t_max) < CUDART_INF_F && sp < PTB_STACK) stk[sp++] = ch.w;
    } else if (cur & 0x40000000u) {
      const DevSphere s = S.spheres[cur & 0x3FFFFFFFu];
      float dist[2], tp[2];
      int exit_only = 0;
      int nh = sphere_hits(r, s, dist, tp, exit_only);
      for (int k = 0; k < nh; ++k) {
        if (point) {
          V3 p = r.o + r.d * tp[k];
          if (length3(p - pos) > limit) continue;
        }
        // sphere occluder: own variant is "simple"; with a point light the primary's variant decides,
        // but a sphere's material has no texture-dependent difference unless the primary is a triangle
        float op = point ? eval_opacity(S, s.model, pu != pu, pu, pv) : S.materials[s.model].opacity;
        tr *= 1.0f - op;
      }
    } else {
      uint32_t first = (cur & 0x3FFFFFFFu) >> 3, count = (cur & 7u) + 1u;
      for (uint32_t i = 0; i < count; ++i) {
        const float4 *tp = reinterpret_cast<const float4 *>(S.tris + first + i);
        float4 A = __ldg(tp), B = __ldg(tp + 1), C = __ldg(tp + 2);
        float t, u, v, det;
        if (!moller_trumbore(r, mk3(A.x, A.y, A.z), mk3(B.x, B.y, B.z), mk3(C.x, C.y, C.z), t, u, v, det)) continue;
        int tri = __float_as_int(B.w);
        float op;
        if (point) {
          V3 p = r.o + r.d * t;
          if (length3(p - pos) > limit) continue;
          op = eval_opacity(S, S.shade[tri].model, pu != pu, pu, pv);
        } else if (__float_as_uint(C.w) & TRI_FLAG_ALPHA) {
          op = hit_opacity(S, tri, u, v);
        } else {
          op = S.materials[S.shade[tri].model].opacity;
        }
        tr *= 1.0f - op;
        if (tr == 0.0f) return 0.0f;
      }
    }
    if (tr == 0.0f) return 0.0f;
    if (sp == 0) break;
    cur = stk[--sp];
  }
  return tr;
}

__global__ void __launch_bounds__(128) k_trace_transmittance(const __grid_constant__ DevScene S,
                                                             const float *__restrict__ rays, uint32_t n,
                                                             float *__restrict__ out, uint32_t *cursor) {
  for (;;) {
    uint32_t q = fetch_work(cursor);
    if (__all_sync(0xffffffffu, q >= n)) break;
    if (q >= n) continue;
    Ray r;
    r.o = mk3(rays[q * 6 + 0], rays[q * 6 + 1], rays[q * 6 + 2]);
    r.d = mk3(rays[q * 6 + 3], rays[q * 6 + 4], rays[q * 6 + 5]);
    out[q] = trace_transmittance(S, r, -1.0f, mk3(0, 0, 0), 0.f, 0.f);
  }
}

// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ RngKey make_key(const WaveParams &P, uint32_t slot) {
  uint32_t npix = P.width * P.height;
  RngKey k;
  k.pixel = slot % npix;
  k.sample = P.sample_first + (slot / npix) * P.sample_stride;
  k.seed_lo = P.seed_lo;
  k.seed_hi = P.seed_hi;
  return k;
}

// raygen: one thread per slot; writes queue buffer 0 in slot order (already compact).
__global__ void __launch_bounds__(256) k_raygen(const __grid_constant__ DevScene S, const __grid_constant__ WaveParams P,
                                                WaveState W) {
  uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
  if (slot == 0) {
    W.counters[CTR_QUEUE0] = P.n_paths;
    W.counters[CTR_WORK_EXTEND] = 0;
  }
  if (slot >= P.n_paths) return;
  RngKey key = make_key(P, slot);
  uint32_t x = key.pixel % P.width, y = key.pixel / P.width;
  uint4 rb = rng_block(key, 0, 0);
  float wf = (float)P.width, hf = (float)P.height;
  float ratio = wf / hf;
  float tn = tanf(S.fov / 2.0f);
  float sx = (float)x + u32_to_unit(rb.x);
  sx = sx / wf * 2.0f - 1.0f;
  sx *= tn * ratio;
  float sy = (float)y + u32_to_unit(rb.y);
  sy = 1.0f - sy / hf * 2.0f;
  sy *= tn;
  V3 dir = normalize3(mk3(sx, sy, -1.0f));
  const float *m = S.camera;
  V3 c0 = mk3(m[0], m[1], m[2]), c1 = mk3(m[4], m[5], m[6]), c2 = mk3(m[8], m[9], m[10]), c3 = mk3(m[12], m[13], m[14]);
  V3 d = ((c0 * dir.x + c1 * dir.y) + c2 * dir.z) + c3 * 0.0f;
  W.ray_o[0][slot] = make_float4(c3.x, c3.y, c3.z, 0.f);
  W.ray_d[0][slot] = make_float4(d.x, d.y, d.z, __int_as_float((int)slot));
  W.thr[0][slot] = make_float4(1.f, 1.f, 1.f, 0.f);
  W.radiance[slot] = make_float4(0.f, 0.f, 0.f, 0.f);
}

// extend: closest accepted hit of every queued ray.  The alpha walk of render_pixel (mod.rs:188-205) is done
// here by re-casting from the rejected hit's key: accept the first hit with opacity >= 1 or a winning draw,
// else the farthest examined hit.
__global__ void __launch_bounds__(128) k_extend(const __grid_constant__ DevScene S, const __grid_constant__ WaveParams P,
                                                WaveState W, uint32_t bounce) {
  const uint32_t in = bounce & 1u, outb = in ^ 1u;
  const uint32_t n = W.counters[CTR_QUEUE0 + in];
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    W.counters[CTR_QUEUE0 + outb] = 0;
    W.counters[CTR_SHADOW] = 0;
    W.counters[CTR_WORK_SHADE] = 0;
    stat_add(W.counters, CTR_STAT_EXTEND, n);
  }
  uint32_t recasts = 0;
  for (;;) {
    uint32_t q = fetch_work(&W.counters[CTR_WORK_EXTEND]);
    if (__all_sync(0xffffffffu, q >= n)) break;
    if (q >= n) continue;
    float4 o4 = W.ray_o[in][q], d4 = W.ray_d[in][q];
    Ray r;
    r.o = mk3(o4.x, o4.y, o4.z);
    r.d = mk3(d4.x, d4.y, d4.z);
    TraceState st;
    trace_closest(S, r, -1.0f, -1, st);
    if (S.any_alpha && st.idx >= 0 && (st.flags & TRI_FLAG_ALPHA)) {
      RngKey key = make_key(P, (uint32_t)__float_as_int(d4.w));
      uint32_t j = 0;
      for (;;) {
        float op = hit_opacity(S, st.idx, st.u, st.v);
        if (op >= 1.0f) break;
        if (op > 0.001f) {
          if (rng_alpha(key, bounce, j++) < op) break;
        }
        TraceState nx;
        trace_closest(S, r, st.best_t, st.best_id, nx);
        ++recasts;
        if (nx.idx < 0) break;  // nothing behind: the last examined hit is shaded
        st = nx;
      }
    }
    W.hit[q] = make_float4(st.best_t, st.u, st.v, __int_as_float(st.best_id));
    W.hit_idx[q] = st.idx;
  }
  recasts = __reduce_add_sync(0xffffffffu, recasts);
  if (lane_id() == 0 && recasts) stat_add(W.counters, CTR_STAT_RECAST, recasts);
}

// shade: miss/background, emissive, next-event setup (shadow queue), GGX sample, throughput, cutoff, RR,
// and compaction of the surviving paths into the other queue buffer.
__global__ void __launch_bounds__(128) k_shade(const __grid_constant__ DevScene S, const __grid_constant__ WaveParams P,
                                               WaveState W, uint32_t bounce) {
  const uint32_t in = bounce & 1u, outb = in ^ 1u;
  const uint32_t n = W.counters[CTR_QUEUE0 + in];
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    W.counters[CTR_WORK_EXTEND] = 0;
    W.counters[CTR_WORK_SHADOW] = 0;
  }
  for (;;) {
    uint32_t q = fetch_work(&W.counters[CTR_WORK_SHADE]);
    if (__all_sync(0xffffffffu, q >= n)) break;
    const bool active = q < n;
    bool alive = false;
    bool lit = false;  // hit a surface (lights are evaluated)
    Ray r{};
    V3 thr{}, rad{}, view{};
    Surface sf{};
    Brdf brdf{};
    uint32_t slot = 0;
    if (active) {
      float4 o4 = W.ray_o[in][q], d4 = W.ray_d[in][q], t4 = W.thr[in][q], h4 = W.hit[q];
      int idx = W.hit_idx[q];
      slot = (uint32_t)__float_as_int(d4.w);
      r.o = mk3(o4.x, o4.y, o4.z);
      r.d = mk3(d4.x, d4.y, d4.z);
      thr = mk3(t4.x, t4.y, t4.z);
      float4 r4 = W.radiance[slot];
      rad = mk3(r4.x, r4.y, r4.z);
      if (idx < 0) {
        rad = rad + mul_el(thr, mk3(S.background[0], S.background[1], S.background[2]));  // mod.rs:184-186
      } else {
        lit = true;
        sf = make_surface(S, r, h4.x, h4.y, h4.z, idx);
        MatSample ms = material_sample(S, sf.model, sf.is_sphere, sf.tu, sf.tv);
        brdf = make_brdf(ms);
        view = -1.0f * r.d;
        rad = rad + mul_el(thr, ms.emissive);  // mod.rs:245
      }
    }
    // next-event estimation set-up: one shadow-queue entry per light (mod.rs:248-262)
    for (uint32_t li = 0; li < S.n_lights; ++li) {
      bool want = false;
      V3 so{}, sd{}, c{};
      float limit = -1.0f;
      if (lit) {
        const DevLight L = S.lights[li];
        V3 lc = mk3(L.color[0], L.color[1], L.color[2]);
        V3 dir;
        if (L.type == PT_LIGHT_DIRECTIONAL) {
          dir = mk3(L.vec[0], L.vec[1], L.vec[2]);
        } else {
          V3 dv = sf.position - mk3(L.vec[0], L.vec[1], L.vec[2]);
          limit = length3(dv);
          dir = normalize3(dv);
          float dissipation = 4.0f * PTB_PI * limit * limit;
          lc = lc / dissipation;
        }
        sd = -1.0f * dir;
        so = sf.position + sf.normal * NORMAL_BIAS;
        c = mul_el(mul_el(thr, eval_direct(brdf, sf.shading_n, view, sd)), lc);
        bool lzero = lc.x == 0.0f && lc.y == 0.0f && lc.z == 0.0f;
        bool czero = c.x == 0.0f && c.y == 0.0f && c.z == 0.0f;
        want = !lzero && !czero;
      }
      uint32_t pos = warp_append(&W.counters[CTR_SHADOW], want);
      if (want && pos < W.shadow_capacity) {
        W.sh_o[pos] = make_float4(so.x, so.y, so.z, limit);
        W.sh_d[pos] = make_float4(sd.x, sd.y, sd.z, __int_as_float((int)slot));
        W.sh_c[pos] = make_float4(c.x, c.y, c.z, sf.is_sphere ? __int_as_float(0x7fc00000) : sf.tu);
        W.sh_p[pos] = make_float4(sf.position.x, sf.position.y, sf.position.z, sf.tv);
      }
    }
    V3 no{}, nd{};
    if (lit) {
      alive = true;
      if (bounce < P.bounces) {  // mod.rs:265-275
        RngKey key = make_key(P, slot);
        uint4 rb = rng_block(key, bounce + 1, 0);
        no = sf.position + sf.normal * NORMAL_BIAS;
        nd = brdf_sample(brdf, sf.shading_n, view, u32_to_unit(rb.x), u32_to_unit(rb.y));
        V3 w = eval_indirect(brdf, sf.shading_n, view, nd) / 1.0f;
        thr = mul_el(thr, w);
        if (dot3(thr, thr) < 0.00001f) alive = false;  // mod.rs:219-221
        if (alive && bounce > 3) {                        // utils.rs:23-31
          float p = fmaxf(fmaxf(thr.x, thr.y), thr.z);
          thr = thr * (1.0f / p);
          if (u32_to_unit(rb.z) > p) alive = false;
        }
      } else {
        alive = false;  // last iteration of `for bounce in 0..=bounces`
      }
    }
    if (active) W.radiance[slot] = make_float4(rad.x, rad.y, rad.z, 0.f);
    uint32_t pos = warp_append(&W.counters[CTR_QUEUE0 + outb], alive);
    if (alive) {
      W.ray_o[outb][pos] = make_float4(no.x, no.y, no.z, 0.f);
      W.ray_d[outb][pos] = make_float4(nd.x, nd.y, nd.z, __int_as_float((int)slot));
      W.thr[outb][pos] = make_float4(thr.x, thr.y, thr.z, 0.f);
    }
  }
}

// shadow: transmittance of every queued shadow ray; adds contribution * transmittance to the path's radiance.
__global__ void __launch_bounds__(128) k_shadow(const __grid_constant__ DevScene S, WaveState W) {
  uint32_t n = W.counters[CTR_SHADOW];
  if (n > W.shadow_capacity) n = W.shadow_capacity;
  if (blockIdx.x == 0 && threadIdx.x == 0) stat_add(W.counters, CTR_STAT_SHADOW, n);
  for (;;) {
    uint32_t q = fetch_work(&W.counters[CTR_WORK_SHADOW]);
    if (__all_sync(0xffffffffu, q >= n)) break;
    if (q >= n) continue;
    float4 o4 = W.sh_o[q], d4 = W.sh_d[q], c4 = W.sh_c[q], p4 = W.sh_p[q];
    Ray r;
    r.o = mk3(o4.x, o4.y, o4.z);
    r.d = mk3(d4.x, d4.y, d4.z);
    float tr = trace_transmittance(S, r, o4.w, mk3(p4.x, p4.y, p4.z), c4.w, p4.w);
    if (tr != 0.0f) {
      float *dst = reinterpret_cast<float *>(W.radiance + (uint32_t)__float_as_int(d4.w));
      atomicAdd(dst + 0, c4.x * tr);
      atomicAdd(dst + 1, c4.y * tr);
      atomicAdd(dst + 2, c4.z * tr);
    }
  }
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

// ------------------------------------------------------------------------------------------------
// host-side launchers
static int g_sm_count = 0;
static int sm_count() {
  if (g_sm_count == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_sm_count, cudaDevAttrMultiProcessorCount, dev);
    if (g_sm_count <= 0) g_sm_count = 148;
  }
  return g_sm_count;
}
constexpr int PERSIST_CTAS_PER_SM = 8;  // 8 x 128 threads = 1024 resident threads per SM

void launch_trace_closest(const DevScene &S, const float *rays, uint32_t n, pt_hit *out, uint32_t *cursor, cudaStream_t st) {
  cudaMemsetAsync(cursor, 0, sizeof(uint32_t), st);
  k_trace_closest<<<sm_count() * PERSIST_CTAS_PER_SM, 128, 0, st>>>(S, rays, n, out, cursor);
}
void launch_trace_transmittance(const DevScene &S, const float *rays, uint32_t n, float *out, uint32_t *cursor, cudaStream_t st) {
  cudaMemsetAsync(cursor, 0, sizeof(uint32_t), st);
  k_trace_transmittance<<<sm_count() * PERSIST_CTAS_PER_SM, 128, 0, st>>>(S, rays, n, out, cursor);
}
void launch_raygen(const DevScene &S, const WaveParams &P, const WaveState &W, cudaStream_t st) {
  k_raygen<<<(P.n_paths + 255) / 256, 256, 0, st>>>(S, P, W);
}
void launch_extend(const DevScene &S, const WaveParams &P, const WaveState &W, uint32_t bounce, cudaStream_t st) {
  k_extend<<<sm_count() * PERSIST_CTAS_PER_SM, 128, 0, st>>>(S, P, W, bounce);
}
void launch_shade(const DevScene &S, const WaveParams &P, const WaveState &W, uint32_t bounce, cudaStream_t st) {
  k_shade<<<sm_count() * PERSIST_CTAS_PER_SM, 128, 0, st>>>(S, P, W, bounce);
}
void launch_shadow(const DevScene &S, const WaveState &W, cudaStream_t st) {
  k_shadow<<<sm_count() * PERSIST_CTAS_PER_SM, 128, 0, st>>>(S, W);
}
void launch_resolve(const WaveParams &P, const WaveState &W, float *accum, cudaStream_t st) {
  uint32_t npix = P.width * P.height;
  k_resolve<<<(npix + 255) / 256, 256, 0, st>>>(P, W, accum);
}
void launch_tonemap(const float *accum, uint32_t npix, uint32_t total_samples, int tonemap, uint8_t *rgb8, cudaStream_t st) {
  k_tonemap<<<(npix * 3 + 255) / 256, 256, 0, st>>>(accum, npix, (float)total_samples, tonemap, rgb8);
}

}  // namespace ptb
