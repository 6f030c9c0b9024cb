// lbvh.cu — GPU acceleration-structure builds for large meshes: 60-bit Morton codes, a hand-written LSD radix sort,
// then either the Karras 2012 radix-tree hierarchy (LBVH) or PLOC's nearest-neighbour agglomeration, a bottom-up box fit and a level-by-level collapse into the same 128-byte 4-wide
// nodes the host SAH builder emits (bvh_build.cpp).  Replaces KDTree::build of the reference's third-party kdtree-ray
// crate (call sites scene/internal/model.rs:96, scene/internal/mod.rs:42) when PT_ACCEL_LBVH is selected.
// The tree is only a conservative candidate filter (results equal brute force; tests/test_gpu_parity.py), so the two
// builders are interchangeable: SAH gives faster traversal, LBVH builds a 20 M-triangle scene in tens of milliseconds.
#include <cuda_runtime.h>

#include <math_constants.h>

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "host_api.hpp"

namespace ptb {
namespace {

#define LB_TRY(expr)                                                                                \
  do {                                                                                              \
    cudaError_t _e = (expr);                                                                        \
    if (_e != cudaSuccess) {                                                                        \
      set_error("CUDA error %s at %s:%d (%s)", cudaGetErrorString(_e), __FILE__, __LINE__, #expr);  \
      return PT_ERR_CUDA;                                                                           \
    }                                                                                               \
  } while (0)

struct DevBuf {  // frees every temporary on scope exit
  std::vector<void *> ptrs;
  ~DevBuf() {
    for (void *p : ptrs) cudaFree(p);
  }
  template <typename T>
  cudaError_t alloc(T **out, size_t count) {
    void *p = nullptr;
    cudaError_t e = cudaMalloc(&p, (count ? count : 1) * sizeof(T));
    if (e == cudaSuccess) ptrs.push_back(p);
    *out = reinterpret_cast<T *>(p);
    return e;
  }
};

// ---- order-preserving float <-> uint for atomicMin/Max on bounds ----
__device__ __forceinline__ uint32_t f2o(float f) {
  uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float o2f(uint32_t o) {
  uint32_t u = (o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o;
#ifdef __CUDA_ARCH__
  return __uint_as_float(u);
#else
  float f;
  memcpy(&f, &u, 4);
  return f;
#endif
}

// ---- 1. primitive boxes (triangles then spheres) + centroid bounds + scene box ----
// Grid-stride: every thread keeps running bounds in registers; one warp shuffle reduction, one shared-memory
// reduction per block and 12 atomics per block at the end (one atomic set per warp of 32 primitives, all on the same
// cache line, cost more than the rest of the kernel on 20 M triangles).
__global__ void __launch_bounds__(256) k_prim_boxes(const float *__restrict__ pos9, uint32_t nt,
                                                    const float4 *__restrict__ spheres, uint32_t ns,
                                                    float4 *__restrict__ blo, float4 *__restrict__ bhi,
                                                    uint32_t *__restrict__ cbounds /* [0..5] centroid, [8..13] box */) {
  const uint32_t n = nt + ns;
  float acc[12];  // centroid min xyz, centroid max xyz, box min xyz, box max xyz
  for (int k = 0; k < 3; ++k) {
    acc[k] = acc[6 + k] = FLT_MAX;
    acc[3 + k] = acc[9 + k] = -FLT_MAX;
  }
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    float lo[3], hi[3];
    if (i < nt) {
      const float *p = pos9 + (size_t)i * 9;
      for (int k = 0; k < 3; ++k) {
        lo[k] = fminf(fminf(p[k], p[3 + k]), p[6 + k]);
        hi[k] = fmaxf(fmaxf(p[k], p[3 + k]), p[6 + k]);
      }
    } else {
      const float4 s = spheres[i - nt];
      const float r = fabsf(s.w);  // exact extent; the margins are added by k_pad_boxes once the scene box is known
      lo[0] = s.x - r; lo[1] = s.y - r; lo[2] = s.z - r;
      hi[0] = s.x + r; hi[1] = s.y + r; hi[2] = s.z + r;
    }
    blo[i] = make_float4(lo[0], lo[1], lo[2], 0.f);
    bhi[i] = make_float4(hi[0], hi[1], hi[2], 0.f);
    for (int k = 0; k < 3; ++k) {
      const float c = 0.5f * (lo[k] + hi[k]);
      acc[k] = fminf(acc[k], c);
      acc[3 + k] = fmaxf(acc[3 + k], c);
      acc[6 + k] = fminf(acc[6 + k], lo[k]);
      acc[9 + k] = fmaxf(acc[9 + k], hi[k]);
    }
  }
  __shared__ float red[8][12];
  for (int k = 0; k < 12; ++k) {
    const bool is_min = k < 3 || (k >= 6 && k < 9);
    float v = acc[k];
    for (int o = 16; o; o >>= 1) {
      const float w = __shfl_xor_sync(0xffffffffu, v, o);
      v = is_min ? fminf(v, w) : fmaxf(v, w);
    }
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5][k] = v;
  }
  __syncthreads();
  if (threadIdx.x < 12) {
    const int k = threadIdx.x;
    const bool is_min = k < 3 || (k >= 6 && k < 9);
    float v = red[0][k];
    for (int w = 1; w < 8; ++w) v = is_min ? fminf(v, red[w][k]) : fmaxf(v, red[w][k]);
    uint32_t *dst = &cbounds[k < 6 ? k : k + 2];
    if (is_min) {
      if (v < FLT_MAX) atomicMin(dst, f2o(v));
    } else if (v > -FLT_MAX) {
      atomicMax(dst, f2o(v));
    }
  }
}

struct SceneBox { float lo[3], hi[3], cam[3], scale; };
__device__ __forceinline__ SceneBox scene_box(const uint32_t *__restrict__ cbounds, float4 cam) {
  return SceneBox{{o2f(cbounds[8]), o2f(cbounds[9]), o2f(cbounds[10])},
                  {o2f(cbounds[11]), o2f(cbounds[12]), o2f(cbounds[13])},
                  {cam.x, cam.y, cam.z}, cam.w};  // w = bvh_margin_scale()
}

// ---- 1b. margins of the primitive boxes (scene_host.hpp): what the f32 intersection arithmetic can leak ----
__global__ void __launch_bounds__(256) k_pad_boxes(const float *__restrict__ pos9, uint32_t nt,
                                                   const float4 *__restrict__ spheres, uint32_t ns,
                                                   float4 *__restrict__ blo, float4 *__restrict__ bhi,
                                                   const uint32_t *__restrict__ cbounds, float4 cam) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nt + ns) return;
  const SceneBox sb = scene_box(cbounds, cam);
  if (i < nt) {
    float p[9];
    for (int k = 0; k < 9; ++k) p[k] = pos9[(size_t)i * 9 + k];
    const float e = triangle_leak_pad(p, sb.scale * origin_reach(p, sb.lo, sb.hi, sb.cam));
    float4 lo = blo[i], hi = bhi[i];
    blo[i] = make_float4(lo.x - e, lo.y - e, lo.z - e, 0.f);
    bhi[i] = make_float4(hi.x + e, hi.y + e, hi.z + e, 0.f);
  } else {
    const float4 s = spheres[i - nt];
    const float c[3] = {s.x, s.y, s.z};
    const float r = sphere_box_radius(s.w, sb.scale * origin_reach(c, sb.lo, sb.hi, sb.cam));
    blo[i] = make_float4(s.x - r, s.y - r, s.z - r, 0.f);
    bhi[i] = make_float4(s.x + r, s.y + r, s.z + r, 0.f);
  }
}

__device__ __forceinline__ uint32_t expand10(uint32_t v) {  // 10 bits -> every third bit
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}

// ---- 2. 60-bit Morton code of the box centroid (20 bits per axis) as two 30-bit halves ----
// 30 bits (a 1024^3 grid over the scene box) put dozens of triangles of a detailed object into one cell when the
// scene box is much larger than the object (the console scene: a 1 M-triangle object on a ground plane), and the
// order inside a cell is then arbitrary - for the radix tree and for PLOC's neighbour search alike.  hi = the top 10
// bits of every axis interleaved, lo = the next 10: sorting by (hi, lo) is two stable passes of the 32-bit sort.
__global__ void __launch_bounds__(256) k_morton(const float4 *__restrict__ blo, const float4 *__restrict__ bhi, uint32_t n,
                                                const uint32_t *__restrict__ cbounds, uint32_t *__restrict__ key_hi,
                                                uint32_t *__restrict__ key_lo, uint32_t *__restrict__ work_keys,
                                                uint32_t *__restrict__ vals) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 lo = blo[i], hi = bhi[i];
  const float c[3] = {0.5f * (lo.x + hi.x), 0.5f * (lo.y + hi.y), 0.5f * (lo.z + hi.z)};
  uint32_t q[3];
  for (int k = 0; k < 3; ++k) {
    const float mn = o2f(cbounds[k]), mx = o2f(cbounds[3 + k]);
    const float ext = mx - mn;
    float t = ext > 0.f ? (c[k] - mn) / ext : 0.f;
    t = fminf(fmaxf(t * 1048576.f, 0.f), 1048575.f);
    q[k] = (uint32_t)t;
  }
  const uint32_t h = (expand10(q[0] >> 10) << 2) | (expand10(q[1] >> 10) << 1) | expand10(q[2] >> 10);
  const uint32_t l = (expand10(q[0] & 1023u) << 2) | (expand10(q[1] & 1023u) << 1) | expand10(q[2] & 1023u);
  key_hi[i] = h;
  key_lo[i] = l;
  work_keys[i] = l;  // the first sort pass orders by the low half
  vals[i] = i;
}
__global__ void __launch_bounds__(256) k_gather_u32(const uint32_t *__restrict__ src, const uint32_t *__restrict__ idx, uint32_t n,
                                                    uint32_t *__restrict__ dst) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = src[idx[i]];
}

// ---- 3. LSD radix sort of (key, value) pairs, 8 bits per pass ----
constexpr int RS_THREADS = 256, RS_ITEMS = 16, RS_TILE = RS_THREADS * RS_ITEMS;

__global__ void __launch_bounds__(RS_THREADS) k_radix_hist(const uint32_t *__restrict__ keys, uint32_t n, int shift,
                                                          uint32_t *__restrict__ counts /* [256][blocks] */) {
  __shared__ uint32_t h[256];
  h[threadIdx.x] = 0;
  __syncthreads();
  const uint32_t base = blockIdx.x * RS_TILE;
  for (int k = 0; k < RS_ITEMS; ++k) {
    const uint32_t i = base + k * RS_THREADS + threadIdx.x;
    if (i < n) atomicAdd(&h[(keys[i] >> shift) & 255u], 1u);
  }
  __syncthreads();
  counts[threadIdx.x * gridDim.x + blockIdx.x] = h[threadIdx.x];
}

// Stable scatter: the tile is processed in rounds of 256 keys in input order; within a round a key's rank among the
// equal digits is (keys of earlier warps) + (earlier lanes of its own warp, via __match_any_sync).
__global__ void __launch_bounds__(RS_THREADS) k_radix_scatter(const uint32_t *__restrict__ keys_in,
                                                             const uint32_t *__restrict__ vals_in, uint32_t n, int shift,
                                                             const uint32_t *__restrict__ offsets /* scanned counts */,
                                                             uint32_t *__restrict__ keys_out, uint32_t *__restrict__ vals_out) {
  __shared__ uint32_t running[256];              // next output position of each digit for this block
  __shared__ uint32_t wcount[RS_THREADS / 32][256];  // per warp, per digit: keys of this round
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  running[threadIdx.x] = offsets[threadIdx.x * gridDim.x + blockIdx.x];
  const uint32_t base = blockIdx.x * RS_TILE;
  for (int k = 0; k < RS_ITEMS; ++k) {
    for (int w = 0; w < RS_THREADS / 32; ++w) wcount[w][threadIdx.x] = 0;
    __syncthreads();
    const uint32_t i = base + k * RS_THREADS + threadIdx.x;
    const bool live = i < n;
    uint32_t key = 0, val = 0, digit = 256 + lane;  // dead lanes match nobody
    if (live) {
      key = keys_in[i];
      val = vals_in[i];
      digit = (key >> shift) & 255u;
    }
    const uint32_t peers = __match_any_sync(0xffffffffu, digit);
    const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
    if (live && rank == 0) wcount[warp][digit] = __popc(peers);
    __syncthreads();
    if (live) {
      uint32_t pos = running[digit] + rank;
      for (uint32_t w = 0; w < warp; ++w) pos += wcount[w][digit];
      keys_out[pos] = key;
      vals_out[pos] = val;
    }
    __syncthreads();
    uint32_t tot = 0;
    for (int w = 0; w < RS_THREADS / 32; ++w) tot += wcount[w][threadIdx.x];
    running[threadIdx.x] += tot;
    __syncthreads();
  }
}

// ---- exclusive scan of a uint32 array (three phases; the block sums are scanned by one block) ----
constexpr int SC_THREADS = 256, SC_ITEMS = 8, SC_TILE = SC_THREADS * SC_ITEMS;

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t *total) {  // 256 threads
  __shared__ uint32_t wsum[SC_THREADS / 32];
  __shared__ uint32_t tot;
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  uint32_t inc = v;
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if ((int)lane >= o) inc += t;
  }
  if (lane == 31) wsum[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    uint32_t s = lane < SC_THREADS / 32 ? wsum[lane] : 0;
    uint32_t si = s;
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t t = __shfl_up_sync(0xffffffffu, si, o);
      if ((int)lane >= o) si += t;
    }
    if (lane < SC_THREADS / 32) wsum[lane] = si - s;
    if (lane == 31) tot = si;
  }
  __syncthreads();
  const uint32_t r = wsum[warp] + inc - v;
  if (total) *total = tot;
  __syncthreads();
  return r;
}

__global__ void __launch_bounds__(SC_THREADS) k_scan_tiles(const uint32_t *__restrict__ in, uint32_t *__restrict__ out,
                                                          uint32_t n, uint32_t *__restrict__ tile_sums) {
  const uint32_t base = blockIdx.x * SC_TILE + threadIdx.x * SC_ITEMS;
  uint32_t v[SC_ITEMS], s = 0;
  for (int k = 0; k < SC_ITEMS; ++k) {
    v[k] = base + k < n ? in[base + k] : 0u;
    s += v[k];
  }
  uint32_t total;
  uint32_t ex = block_exclusive_scan(s, &total);
  for (int k = 0; k < SC_ITEMS; ++k) {
    if (base + k < n) out[base + k] = ex;
    ex += v[k];
  }
  if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}
__global__ void __launch_bounds__(SC_THREADS) k_scan_sums(uint32_t *sums, uint32_t n, uint32_t *grand_total) {
  uint32_t carry = 0;  // one block walks the tile sums
  for (uint32_t base = 0; base < n; base += SC_THREADS) {
    const uint32_t i = base + threadIdx.x;
    const uint32_t v = i < n ? sums[i] : 0u;
    uint32_t total;
    const uint32_t ex = block_exclusive_scan(v, &total);
    if (i < n) sums[i] = carry + ex;
    carry += total;
  }
  if (threadIdx.x == 0 && grand_total) *grand_total = carry;
}
__global__ void __launch_bounds__(SC_THREADS) k_scan_add(uint32_t *out, uint32_t n, const uint32_t *__restrict__ tile_sums) {
  const uint32_t base = blockIdx.x * SC_TILE + threadIdx.x * SC_ITEMS;
  const uint32_t add = tile_sums[blockIdx.x];
  for (int k = 0; k < SC_ITEMS; ++k)
    if (base + k < n) out[base + k] += add;
}

// out[i] = sum of in[0..i); out may alias in.  tmp: at least ceil(n / SC_TILE) + 1 uints; grand total -> tmp[tiles].
int exclusive_scan_u32(const uint32_t *in, uint32_t *out, uint32_t n, uint32_t *tmp, cudaStream_t st) {
  if (n == 0) return PT_OK;
  const uint32_t tiles = (n + SC_TILE - 1) / SC_TILE;
  k_scan_tiles<<<tiles, SC_THREADS, 0, st>>>(in, out, n, tmp);
  k_scan_sums<<<1, SC_THREADS, 0, st>>>(tmp, tiles, tmp + tiles);
  k_scan_add<<<tiles, SC_THREADS, 0, st>>>(out, n, tmp);
  LB_TRY(cudaGetLastError());
  return PT_OK;
}

int radix_sort_pairs(uint32_t *keys, uint32_t *vals, uint32_t *keys_alt, uint32_t *vals_alt, uint32_t n, int bits,
                     uint32_t *counts, uint32_t *scan_tmp, cudaStream_t st) {
  const uint32_t blocks = (n + RS_TILE - 1) / RS_TILE;
  uint32_t *ki = keys, *vi = vals, *ko = keys_alt, *vo = vals_alt;
  int passes = 0;
  for (int shift = 0; shift < bits; shift += 8, ++passes) {
    k_radix_hist<<<blocks, RS_THREADS, 0, st>>>(ki, n, shift, counts);
    int rc = exclusive_scan_u32(counts, counts, 256u * blocks, scan_tmp, st);
    if (rc) return rc;
    k_radix_scatter<<<blocks, RS_THREADS, 0, st>>>(ki, vi, n, shift, counts, ko, vo);
    std::swap(ki, ko);
    std::swap(vi, vo);
  }
  LB_TRY(cudaGetLastError());
  if (passes & 1) {  // result must end in keys/vals
    LB_TRY(cudaMemcpyAsync(keys, ki, n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
    LB_TRY(cudaMemcpyAsync(vals, vi, n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
  }
  return PT_OK;
}

// ---- 4. Karras 2012: radix tree over the sorted codes (ties broken by position) ----
// Node references: bit31 = 1 -> sorted primitive position (a radix-tree leaf); else internal node index.
constexpr uint32_t PRIM = 0x80000000u;

__device__ __forceinline__ int delta(const uint32_t *__restrict__ khi, const uint32_t *__restrict__ klo, int n, int i, int j) {
  if (j < 0 || j >= n) return -1;
  const uint32_t a = khi[i], b = khi[j];
  if (a != b) return __clz(a ^ b);
  const uint32_t c = klo[i], d = klo[j];
  if (c != d) return 32 + __clz(c ^ d);
  return 64 + __clz((uint32_t)i ^ (uint32_t)j);
}

__global__ void __launch_bounds__(256) k_radix_tree(const uint32_t *__restrict__ keys, const uint32_t *__restrict__ keys_lo, int n,
                                                    uint32_t *__restrict__ left,
                                                    uint32_t *__restrict__ right, uint32_t *__restrict__ rlo,
                                                    uint32_t *__restrict__ rhi, uint32_t *__restrict__ parent_int,
                                                    uint32_t *__restrict__ parent_prim) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n - 1) return;
  const int d = delta(keys, keys_lo, n, i, i + 1) - delta(keys, keys_lo, n, i, i - 1) >= 0 ? 1 : -1;
  const int dmin = delta(keys, keys_lo, n, i, i - d);
  int lmax = 2;
  while (delta(keys, keys_lo, n, i, i + lmax * d) > dmin) lmax *= 2;
  int l = 0;
  for (int t = lmax >> 1; t >= 1; t >>= 1)
    if (delta(keys, keys_lo, n, i, i + (l + t) * d) > dmin) l += t;
  const int j = i + l * d;
  const int dnode = delta(keys, keys_lo, n, i, j);
  int s = 0;
  for (int t = (l + 1) >> 1;; t = (t + 1) >> 1) {
    if (delta(keys, keys_lo, n, i, i + (s + t) * d) > dnode) s += t;
    if (t == 1) break;
  }
  const int gamma = i + s * d + min(d, 0);
  const int lo = min(i, j), hi = max(i, j);
  const uint32_t lref = lo == gamma ? (PRIM | (uint32_t)gamma) : (uint32_t)gamma;
  const uint32_t rref = hi == gamma + 1 ? (PRIM | (uint32_t)(gamma + 1)) : (uint32_t)(gamma + 1);
  left[i] = lref;
  right[i] = rref;
  rlo[i] = (uint32_t)lo;
  rhi[i] = (uint32_t)hi;
  if (lref & PRIM) parent_prim[gamma] = (uint32_t)i; else parent_int[gamma] = (uint32_t)i;
  if (rref & PRIM) parent_prim[gamma + 1] = (uint32_t)i; else parent_int[gamma + 1] = (uint32_t)i;
  if (i == 0) parent_int[0] = 0xFFFFFFFFu;
}

// ---- 5. bottom-up box fit: the second thread to reach a node merges its children ----
__global__ void __launch_bounds__(256) k_fit(int n, const uint32_t *__restrict__ sorted_prim, const float4 *__restrict__ blo,
                                             const float4 *__restrict__ bhi, const uint32_t *__restrict__ left,
                                             const uint32_t *__restrict__ right, const uint32_t *__restrict__ parent_int,
                                             const uint32_t *__restrict__ parent_prim, float4 *__restrict__ nlo,
                                             float4 *__restrict__ nhi, uint32_t *__restrict__ visits) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t node = parent_prim[i];
  while (node != 0xFFFFFFFFu) {
    if (atomicAdd(&visits[node], 1u) == 0) return;  // first arrival: the sibling subtree is not finished yet
    __threadfence();
    float4 lo[2], hi[2];
    const uint32_t ch[2] = {left[node], right[node]};
    for (int k = 0; k < 2; ++k) {
      if (ch[k] & PRIM) {
        const uint32_t p = sorted_prim[ch[k] & ~PRIM];
        lo[k] = blo[p];
        hi[k] = bhi[p];
      } else {
        lo[k] = __ldcg(&nlo[ch[k]]);
        hi[k] = __ldcg(&nhi[ch[k]]);
      }
    }
    __stcg(&nlo[node], make_float4(fminf(lo[0].x, lo[1].x), fminf(lo[0].y, lo[1].y), fminf(lo[0].z, lo[1].z), 0.f));
    __stcg(&nhi[node], make_float4(fmaxf(hi[0].x, hi[1].x), fmaxf(hi[0].y, hi[1].y), fmaxf(hi[0].z, hi[1].z), 0.f));
    __threadfence();
    node = parent_int[node];
  }
}

// ---- 4b/5b. PLOC (Meister & Bittner 2018, "Parallel Locally-Ordered Clustering"): bottom-up agglomeration in Morton
// order.  Every cluster looks for the neighbour within PLOC_R positions whose union with it has the smallest surface
// area; mutual nearest neighbours merge; the survivors are compacted and the search repeats until one cluster is
// left.  Unlike the radix tree, which splits where the Morton codes differ, the merges minimise the very quantity
// the SAH measures, so the tree traces like a top-down SAH tree while the build stays a handful of streaming passes.
// Output: the same arrays the radix-tree path produces (left / right / rlo / rhi / boxes, root = internal node 0, leaf
// references = positions in a primitive order in which every subtree is contiguous), so collapse and leaf ordering
// are shared.
constexpr uint32_t NONE32 = 0xFFFFFFFFu;
// Search radius 8; $PT_B200_PLOC_R = 16 | 32 for A/B runs.  Measured frame time against the host SAH tree
// (profiles/r07/builders_s7.jsonl, radius 8 / 16 / 32): console +8 / +14 / +19 %, foliage +17 / +18 / +22 %, city
// -15 / -12 / -2 %: on these scenes a wider search does not buy better trees, only a longer build.
int ploc_radius() {
  static const int r = [] {
    const char *e = std::getenv("PT_B200_PLOC_R");
    const int v = e ? std::atoi(e) : 8;
    return v == 16 || v == 32 ? v : 8;
  }();
  return r;
}

__global__ void __launch_bounds__(256) k_ploc_init(uint32_t n, const uint32_t *__restrict__ sorted_prim,
                                                   const float4 *__restrict__ blo, const float4 *__restrict__ bhi,
                                                   uint32_t *__restrict__ cref, float4 *__restrict__ clo, float4 *__restrict__ chi) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t p = sorted_prim[i];
  cref[i] = PRIM | i;
  clo[i] = blo[p];
  chi[i] = bhi[p];
}

__device__ __forceinline__ float union_half_area(float4 alo, float4 ahi, float4 blo, float4 bhi) {
  const float dx = fmaxf(ahi.x, bhi.x) - fminf(alo.x, blo.x), dy = fmaxf(ahi.y, bhi.y) - fminf(alo.y, blo.y),
              dz = fmaxf(ahi.z, bhi.z) - fminf(alo.z, blo.z);
  return dx * dy + dy * dz + dz * dx;
}

// nearest neighbour of every cluster within +-PLOC_R positions (ties: the lower position)
template <int PLOC_R>
__global__ void __launch_bounds__(256) k_ploc_nn(uint32_t c, const float4 *__restrict__ clo, const float4 *__restrict__ chi,
                                                 uint32_t *__restrict__ nn) {
  __shared__ float4 slo[256 + 2 * PLOC_R], shi[256 + 2 * PLOC_R];
  const int base = (int)(blockIdx.x * 256u) - PLOC_R;
  for (int k = threadIdx.x; k < 256 + 2 * PLOC_R; k += 256) {
    const int j = base + k;
    if (j >= 0 && j < (int)c) {
      slo[k] = clo[j];
      shi[k] = chi[j];
    }
  }
  __syncthreads();
  const int i = (int)(blockIdx.x * 256u + threadIdx.x);
  if (i >= (int)c) return;
  const float4 mlo = slo[threadIdx.x + PLOC_R], mhi = shi[threadIdx.x + PLOC_R];
  float best = CUDART_INF_F;
  int bj = -1;
  for (int d = -PLOC_R; d <= PLOC_R; ++d) {
    const int j = i + d;
    if (d == 0 || j < 0 || j >= (int)c) continue;
    const float a = union_half_area(mlo, mhi, slo[threadIdx.x + PLOC_R + d], shi[threadIdx.x + PLOC_R + d]);
    if (a < best || bj < 0) {  // bj < 0: NaN / inf areas still yield a neighbour
      best = a;
      bj = j;
    }
  }
  nn[i] = (uint32_t)bj;
}

struct PlocTree {
  uint32_t *left, *right, *cnt, *par_i, *par_p;  // par_*: parent internal node | side << 31
  float4 *nlo, *nhi;
  uint32_t n;  // primitives
};
__device__ __forceinline__ uint32_t ref_count(const PlocTree &T, uint32_t ref) { return (ref & PRIM) ? 1u : T.cnt[ref]; }

// mutual nearest neighbours merge into a new internal node (the lower position keeps the merged cluster)
__global__ void __launch_bounds__(256) k_ploc_merge(uint32_t c, uint32_t *__restrict__ cref, float4 *__restrict__ clo,
                                                    float4 *__restrict__ chi, const uint32_t *__restrict__ nn, PlocTree T,
                                                    uint32_t *__restrict__ merge_ctr, uint32_t *__restrict__ keep) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= c) return;
  const uint32_t j = nn[i];
  if (nn[j] != i) {
    keep[i] = 1u;
    return;
  }
  if (i > j) {
    keep[i] = 0u;
    return;
  }
  const uint32_t k = atomicAdd(merge_ctr, 1u);
  const uint32_t idx = (T.n - 2u) - k;  // n - 1 merges in total: the last one is internal node 0, the root
  const uint32_t a = cref[i], b = cref[j];
  const float4 alo = clo[i], ahi = chi[i], blo_ = clo[j], bhi_ = chi[j];
  const float4 lo = make_float4(fminf(alo.x, blo_.x), fminf(alo.y, blo_.y), fminf(alo.z, blo_.z), 0.f);
  const float4 hi = make_float4(fmaxf(ahi.x, bhi_.x), fmaxf(ahi.y, bhi_.y), fmaxf(ahi.z, bhi_.z), 0.f);
  T.left[idx] = a;
  T.right[idx] = b;
  T.nlo[idx] = lo;
  T.nhi[idx] = hi;
  T.cnt[idx] = ref_count(T, a) + ref_count(T, b);
  if (a & PRIM) T.par_p[a & ~PRIM] = idx; else T.par_i[a] = idx;
  if (b & PRIM) T.par_p[b & ~PRIM] = idx | 0x80000000u; else T.par_i[b] = idx | 0x80000000u;
  cref[i] = idx;
  clo[i] = lo;
  chi[i] = hi;
  keep[i] = 1u;
}

__global__ void __launch_bounds__(256) k_ploc_compact(uint32_t c, const uint32_t *__restrict__ keep,
                                                      const uint32_t *__restrict__ pos, const uint32_t *__restrict__ cref,
                                                      const float4 *__restrict__ clo, const float4 *__restrict__ chi,
                                                      uint32_t *__restrict__ cref2, float4 *__restrict__ clo2,
                                                      float4 *__restrict__ chi2) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= c || !keep[i]) return;
  const uint32_t p = pos[i];
  cref2[p] = cref[i];
  clo2[p] = clo[i];
  chi2[p] = chi[i];
}

// Depth-first position of a subtree's first primitive: the sizes of the left siblings on the way to the root.
__device__ __forceinline__ uint32_t dfs_offset(const PlocTree &T, uint32_t parent) {
  uint32_t off = 0;
  while (parent != NONE32) {
    const uint32_t idx = parent & 0x7FFFFFFFu;
    if (parent & 0x80000000u) off += ref_count(T, T.left[idx]);
    parent = T.par_i[idx];
  }
  return off;
}
__global__ void __launch_bounds__(256) k_ploc_leaf_pos(PlocTree T, const uint32_t *__restrict__ sorted_in,
                                                       uint32_t *__restrict__ sorted_out, uint32_t *__restrict__ newpos) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= T.n) return;
  const uint32_t off = dfs_offset(T, T.par_p[i]);
  newpos[i] = off;
  sorted_out[off] = sorted_in[i];
}
__global__ void __launch_bounds__(256) k_ploc_node_range(PlocTree T, uint32_t *__restrict__ rlo, uint32_t *__restrict__ rhi) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i + 1 >= T.n) return;
  const uint32_t off = dfs_offset(T, T.par_i[i]);
  rlo[i] = off;
  rhi[i] = off + T.cnt[i] - 1u;
}
__global__ void __launch_bounds__(256) k_ploc_fix_refs(PlocTree T, const uint32_t *__restrict__ newpos) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i + 1 >= T.n) return;
  const uint32_t l = T.left[i], r = T.right[i];
  if (l & PRIM) T.left[i] = PRIM | newpos[l & ~PRIM];
  if (r & PRIM) T.right[i] = PRIM | newpos[r & ~PRIM];
}

// ---- 6. collapse to 4-wide nodes, one tree level per launch ----
struct CollapseArgs {
  const uint32_t *left, *right, *rlo, *rhi;
  const float4 *nlo, *nhi;       // internal boxes
  const float4 *blo, *bhi;       // primitive boxes (original primitive order)
  const uint32_t *sorted_prim;   // sorted position -> primitive (triangles < nt <= spheres)
  const uint32_t *tri_rank;      // sorted position -> number of triangles before it (n + 1 entries)
  uint32_t nt;
  uint32_t max_leaf;             // triangles per leaf (bvh_max_leaf_tris)
  Bvh4Node *nodes;
  uint32_t *node_count;
  uint32_t node_cap;
  const uint32_t *cbounds;       // scene box (k_prim_boxes) for the slab-plane margin
  float4 cam;                    // camera position, w = bvh_margin_scale()
};

// A radix-tree node is emitted as one leaf when it holds <= max_leaf primitives, all triangles (spheres sit alone).
__device__ __forceinline__ bool is_leaf_ref(const CollapseArgs &A, uint32_t ref) {
  if (ref & PRIM) return true;
  const uint32_t lo = A.rlo[ref], hi = A.rhi[ref], cnt = hi - lo + 1;
  return cnt <= A.max_leaf && A.tri_rank[hi + 1] - A.tri_rank[lo] == cnt;
}
__device__ __forceinline__ void ref_box(const CollapseArgs &A, uint32_t ref, float4 &lo, float4 &hi) {
  if (ref & PRIM) {
    const uint32_t p = A.sorted_prim[ref & ~PRIM];
    lo = A.blo[p];
    hi = A.bhi[p];
  } else {
    lo = A.nlo[ref];
    hi = A.nhi[ref];
  }
}
__device__ __forceinline__ float half_area(float4 lo, float4 hi) {
  const float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
  return dx * dy + dy * dz + dz * dx;
}
__device__ __forceinline__ uint32_t leaf_child(const CollapseArgs &A, uint32_t ref) {
  uint32_t lo, cnt;
  if (ref & PRIM) {
    lo = ref & ~PRIM;
    cnt = 1;
    const uint32_t p = A.sorted_prim[lo];
    if (p >= A.nt) return BVH_LEAF | BVH_SPHERE | (p - A.nt);
  } else {
    lo = A.rlo[ref];
    cnt = A.rhi[ref] - lo + 1;
  }
  return BVH_LEAF | (A.tri_rank[lo] << 3) | (cnt - 1);
}

__global__ void __launch_bounds__(128) k_collapse(const CollapseArgs A, const uint2 *__restrict__ work_in, uint32_t n_in,
                                                  uint2 *__restrict__ work_out, uint32_t *__restrict__ n_out) {
  const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= n_in) return;
  const uint2 item = work_in[w];  // x = radix-tree node (internal, not a leaf), y = 4-wide node index
  const SceneBox sb = scene_box(A.cbounds, A.cam);
  const float S = sb.scale * coordinate_scale(sb.lo, sb.hi, sb.cam);
  uint32_t kids[4];
  int nk = 2;
  kids[0] = A.left[item.x];
  kids[1] = A.right[item.x];
  while (nk < 4) {  // open the inner child with the largest box, as the host builder does
    int best = -1;
    float ba = -1.f;
    for (int k = 0; k < nk; ++k) {
      if (is_leaf_ref(A, kids[k])) continue;
      float4 lo, hi;
      ref_box(A, kids[k], lo, hi);
      const float a = half_area(lo, hi);
      if (a > ba) { ba = a; best = k; }
    }
    if (best < 0) break;
    const uint32_t c = kids[best];
    kids[best] = A.left[c];
    kids[nk++] = A.right[c];
  }
  Bvh4Node nd;
  for (int k = 0; k < 4; ++k) {
    nd.lo_x[k] = nd.lo_y[k] = nd.lo_z[k] = CUDART_NAN_F;  // empty slot: NaN planes fail every slab test
    nd.hi_x[k] = nd.hi_y[k] = nd.hi_z[k] = CUDART_NAN_F;
    nd.child[k] = BVH_EMPTY;
    nd.pad[k] = 0;
  }
  for (int k = 0; k < nk; ++k) {
    float4 lo, hi;
    ref_box(A, kids[k], lo, hi);
    // slab-plane margin, identical to bvh_build.cpp (scene_host.hpp plane_pad)
    const float ex = plane_pad(lo.x, hi.x, S), ey = plane_pad(lo.y, hi.y, S), ez = plane_pad(lo.z, hi.z, S);
    nd.lo_x[k] = lo.x - ex; nd.hi_x[k] = hi.x + ex;
    nd.lo_y[k] = lo.y - ey; nd.hi_y[k] = hi.y + ey;
    nd.lo_z[k] = lo.z - ez; nd.hi_z[k] = hi.z + ez;
    if (is_leaf_ref(A, kids[k])) {
      nd.child[k] = leaf_child(A, kids[k]);
    } else {
      const uint32_t id4 = atomicAdd(A.node_count, 1u);
      nd.child[k] = id4;
      if (id4 < A.node_cap) work_out[atomicAdd(n_out, 1u)] = make_uint2(kids[k], id4);
    }
  }
  A.nodes[item.y] = nd;
}

__global__ void __launch_bounds__(256) k_leaf_order(uint32_t n, uint32_t nt, const uint32_t *__restrict__ sorted_prim,
                                                    const uint32_t *__restrict__ tri_rank, uint32_t *__restrict__ leaf_tris,
                                                    uint32_t *__restrict__ is_tri) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (is_tri) {
    is_tri[i] = sorted_prim[i] < nt ? 1u : 0u;
    return;
  }
  const uint32_t p = sorted_prim[i];
  if (p < nt) leaf_tris[tri_rank[i]] = p;
}

}  // namespace

// Builds `out` for scene `s` on the current CUDA device.  Returns PT_OK, or PT_ERR_INVALID with *too_deep = 1 when the
// radix tree is deeper than the traversal stack allows (the caller then falls back to the host SAH build).
int build_bvh4_lbvh(const HostScene &s, Bvh4 &out, float *build_ms, int *too_deep, int method) {
  if (too_deep) *too_deep = 0;
  const bool ploc = method == 1;
  const uint32_t nt = (uint32_t)s.n_tris(), ns = (uint32_t)s.sphere_models.size(), n = nt + ns;
  if (n < 16) {
    set_error("LBVH build needs at least 16 primitives");
    return PT_ERR_INVALID;
  }
  cudaStream_t st = nullptr;
  DevBuf B;
  float *d_pos = nullptr;
  float4 *d_sph = nullptr, *blo = nullptr, *bhi = nullptr, *nlo = nullptr, *nhi = nullptr;
  uint32_t *cb = nullptr, *keys = nullptr, *vals = nullptr, *keys2 = nullptr, *vals2 = nullptr, *counts = nullptr;
  uint32_t *key_hi = nullptr, *key_lo = nullptr;  // unsorted Morton halves per primitive
  uint32_t *scan_tmp = nullptr, *left = nullptr, *right = nullptr, *rlo = nullptr, *rhi = nullptr, *par_i = nullptr;
  uint32_t *par_p = nullptr, *visits = nullptr, *tri_rank = nullptr, *leaf_tris = nullptr, *ctr = nullptr;
  uint2 *work[2] = {nullptr, nullptr};
  Bvh4Node *nodes = nullptr;
  // PLOC only: cluster arrays (ping-pong), nearest neighbours, keep flags / positions, subtree sizes, new leaf positions
  uint32_t *cref[2] = {nullptr, nullptr}, *nn = nullptr, *keep = nullptr, *kpos = nullptr, *cnt = nullptr, *newpos = nullptr;
  float4 *clo[2] = {nullptr, nullptr}, *chi[2] = {nullptr, nullptr};
  const uint32_t rs_blocks = (n + RS_TILE - 1) / RS_TILE;
  const uint32_t node_cap = n;  // every 4-wide node has >= 2 children, so there are fewer nodes than leaves
  LB_TRY(B.alloc(&d_pos, (size_t)nt * 9));
  LB_TRY(B.alloc(&d_sph, ns));
  LB_TRY(B.alloc(&blo, n));
  LB_TRY(B.alloc(&bhi, n));
  LB_TRY(B.alloc(&nlo, n));
  LB_TRY(B.alloc(&nhi, n));
  LB_TRY(B.alloc(&cb, 16));
  LB_TRY(B.alloc(&keys, n));
  LB_TRY(B.alloc(&vals, n));
  LB_TRY(B.alloc(&keys2, n));
  LB_TRY(B.alloc(&key_hi, n));
  LB_TRY(B.alloc(&key_lo, n));
  LB_TRY(B.alloc(&vals2, n));
  LB_TRY(B.alloc(&counts, 256 * (size_t)rs_blocks));
  LB_TRY(B.alloc(&scan_tmp, (256 * (size_t)rs_blocks + n) / SC_TILE + 16));
  LB_TRY(B.alloc(&left, n));
  LB_TRY(B.alloc(&right, n));
  LB_TRY(B.alloc(&rlo, n));
  LB_TRY(B.alloc(&rhi, n));
  LB_TRY(B.alloc(&par_i, n));
  LB_TRY(B.alloc(&par_p, n));
  LB_TRY(B.alloc(&visits, n));
  LB_TRY(B.alloc(&tri_rank, (size_t)n + 1));
  LB_TRY(B.alloc(&leaf_tris, nt));
  LB_TRY(B.alloc(&ctr, 8));
  LB_TRY(B.alloc(&work[0], n));
  LB_TRY(B.alloc(&work[1], n));
  LB_TRY(B.alloc(&nodes, node_cap));
  if (ploc) {
    for (int k = 0; k < 2; ++k) {
      LB_TRY(B.alloc(&cref[k], n));
      LB_TRY(B.alloc(&clo[k], n));
      LB_TRY(B.alloc(&chi[k], n));
    }
    LB_TRY(B.alloc(&nn, n));
    LB_TRY(B.alloc(&keep, (size_t)n + 1));
    LB_TRY(B.alloc(&kpos, (size_t)n + 1));
    LB_TRY(B.alloc(&cnt, n));
    LB_TRY(B.alloc(&newpos, n));
  }
  LB_TRY(cudaMemcpyAsync(d_pos, s.pos.data(), (size_t)nt * 9 * sizeof(float), cudaMemcpyHostToDevice, st));
  if (ns) {
    std::vector<float4> sph(ns);
    for (uint32_t k = 0; k < ns; ++k) {
      const pt_model &m = s.models[s.sphere_models[k]];
      sph[k] = make_float4(m.center[0], m.center[1], m.center[2], m.radius);
    }
    LB_TRY(cudaMemcpy(d_sph, sph.data(), ns * sizeof(float4), cudaMemcpyHostToDevice));
  }
  cudaEvent_t e0, e1;
  LB_TRY(cudaEventCreate(&e0));
  LB_TRY(cudaEventCreate(&e1));
  LB_TRY(cudaEventRecord(e0, st));
  // [0..5] centroid bounds, [8..13] scene box, both as order-preserving uints (min slots start at the maximum)
  const uint32_t init_cb[16] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u, 0u, 0u, 0u,
                                0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u, 0u, 0u, 0u};
  const float4 cam = make_float4(s.camera[12], s.camera[13], s.camera[14], bvh_margin_scale());
  LB_TRY(cudaMemcpyAsync(cb, init_cb, sizeof init_cb, cudaMemcpyHostToDevice, st));
  const uint32_t g256 = (n + 255) / 256;
  k_prim_boxes<<<std::min(g256, 148u * 16u), 256, 0, st>>>(d_pos, nt, d_sph, ns, blo, bhi, cb);
  k_morton<<<g256, 256, 0, st>>>(blo, bhi, n, cb, key_hi, key_lo, keys, vals);
  k_pad_boxes<<<g256, 256, 0, st>>>(d_pos, nt, d_sph, ns, blo, bhi, cb, cam);  // symmetric: centroids unchanged
  // order by (hi, lo): sort by the low half, then a stable sort by the high half
  int rc = radix_sort_pairs(keys, vals, keys2, vals2, n, 32, counts, scan_tmp, st);
  if (rc) return rc;
  k_gather_u32<<<g256, 256, 0, st>>>(key_hi, vals, n, keys);
  if ((rc = radix_sort_pairs(keys, vals, keys2, vals2, n, 32, counts, scan_tmp, st))) return rc;
  k_gather_u32<<<g256, 256, 0, st>>>(key_lo, vals, n, keys2);  // keys = sorted high halves, keys2 = their low halves
  uint32_t *sorted = vals;  // sorted position -> primitive
  if (ploc) {
    PlocTree T{left, right, cnt, par_i, par_p, nlo, nhi, n};
    LB_TRY(cudaMemsetAsync(par_i, 0xFF, (size_t)n * sizeof(uint32_t), st));
    LB_TRY(cudaMemsetAsync(par_p, 0xFF, (size_t)n * sizeof(uint32_t), st));
    LB_TRY(cudaMemsetAsync(ctr + 2, 0, sizeof(uint32_t), st));  // merge counter
    k_ploc_init<<<g256, 256, 0, st>>>(n, vals, blo, bhi, cref[0], clo[0], chi[0]);
    uint32_t c = n;
    int iters = 0;
    while (c > 1) {
      if (++iters > 512) {  // every pass merges at least the globally closest pair; hundreds of passes mean a degenerate input
        if (too_deep) *too_deep = 1;
        set_error("PLOC did not converge");
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        return PT_ERR_INVALID;
      }
      const uint32_t gc = (c + 255) / 256;
      if (ploc_radius() == 8) k_ploc_nn<8><<<gc, 256, 0, st>>>(c, clo[0], chi[0], nn);
      else if (ploc_radius() == 32) k_ploc_nn<32><<<gc, 256, 0, st>>>(c, clo[0], chi[0], nn);
      else k_ploc_nn<16><<<gc, 256, 0, st>>>(c, clo[0], chi[0], nn);
      k_ploc_merge<<<gc, 256, 0, st>>>(c, cref[0], clo[0], chi[0], nn, T, ctr + 2, keep);
      LB_TRY(cudaMemsetAsync(keep + c, 0, sizeof(uint32_t), st));
      if ((rc = exclusive_scan_u32(keep, kpos, c + 1, scan_tmp, st))) return rc;
      k_ploc_compact<<<gc, 256, 0, st>>>(c, keep, kpos, cref[0], clo[0], chi[0], cref[1], clo[1], chi[1]);
      LB_TRY(cudaMemcpyAsync(&c, kpos + c, sizeof c, cudaMemcpyDeviceToHost, st));
      LB_TRY(cudaStreamSynchronize(st));
      std::swap(cref[0], cref[1]);
      std::swap(clo[0], clo[1]);
      std::swap(chi[0], chi[1]);
    }
    // depth-first primitive order (every subtree contiguous), node ranges, leaf references in the new order
    k_ploc_leaf_pos<<<g256, 256, 0, st>>>(T, vals, vals2, newpos);
    k_ploc_node_range<<<g256, 256, 0, st>>>(T, rlo, rhi);
    k_ploc_fix_refs<<<g256, 256, 0, st>>>(T, newpos);
    sorted = vals2;
  }
  // triangle ranks in sorted order (n + 1 entries: the last one is the triangle count)
  k_leaf_order<<<g256, 256, 0, st>>>(n, nt, sorted, nullptr, nullptr, tri_rank);
  LB_TRY(cudaMemsetAsync(tri_rank + n, 0, sizeof(uint32_t), st));
  if ((rc = exclusive_scan_u32(tri_rank, tri_rank, n + 1, scan_tmp, st))) return rc;
  k_leaf_order<<<g256, 256, 0, st>>>(n, nt, sorted, tri_rank, leaf_tris, nullptr);
  if (!ploc) {
    k_radix_tree<<<g256, 256, 0, st>>>(keys, keys2, (int)n, left, right, rlo, rhi, par_i, par_p);
    LB_TRY(cudaMemsetAsync(visits, 0, (size_t)n * sizeof(uint32_t), st));
    k_fit<<<g256, 256, 0, st>>>((int)n, vals, blo, bhi, left, right, par_i, par_p, nlo, nhi, visits);
  }
  LB_TRY(cudaGetLastError());
  // level-by-level collapse; the root of the radix tree (internal node 0) becomes 4-wide node 0
  CollapseArgs A{left, right, rlo, rhi, nlo, nhi, blo, bhi, sorted, tri_rank, nt, bvh_max_leaf_tris(), nodes, ctr, node_cap, cb, cam};
  const uint32_t one = 1;
  LB_TRY(cudaMemcpyAsync(ctr, &one, sizeof one, cudaMemcpyHostToDevice, st));  // node_count = 1 (the root)
  const uint2 root_item = make_uint2(0u, 0u);
  LB_TRY(cudaMemcpyAsync(work[0], &root_item, sizeof root_item, cudaMemcpyHostToDevice, st));
  uint32_t n_in = 1;
  int depth = 0;
  // A root that is itself one leaf (<= 4 triangles) cannot occur: n >= 8.
  while (n_in) {
    ++depth;
    if (depth * 3 + 4 > 96) {  // traversal stack bound (PTB_STACK)
      if (too_deep) *too_deep = 1;
      set_error("LBVH too deep for the traversal stack");
      cudaEventDestroy(e0);
      cudaEventDestroy(e1);
      return PT_ERR_INVALID;
    }
    LB_TRY(cudaMemsetAsync(ctr + 1, 0, sizeof(uint32_t), st));
    k_collapse<<<(n_in + 127) / 128, 128, 0, st>>>(A, work[0], n_in, work[1], ctr + 1);
    LB_TRY(cudaMemcpyAsync(&n_in, ctr + 1, sizeof n_in, cudaMemcpyDeviceToHost, st));
    LB_TRY(cudaStreamSynchronize(st));
    std::swap(work[0], work[1]);
  }
  LB_TRY(cudaEventRecord(e1, st));
  LB_TRY(cudaEventSynchronize(e1));
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  if (build_ms) *build_ms = ms;
  uint32_t n_nodes = 0;
  LB_TRY(cudaMemcpy(&n_nodes, ctr, sizeof n_nodes, cudaMemcpyDeviceToHost));
  if (n_nodes > node_cap) {
    set_error("LBVH node pool overflow");
    return PT_ERR_INVALID;
  }
  out.nodes.resize(n_nodes);
  out.leaf_tris.resize(nt);
  LB_TRY(cudaMemcpy(out.nodes.data(), nodes, (size_t)n_nodes * sizeof(Bvh4Node), cudaMemcpyDeviceToHost));
  LB_TRY(cudaMemcpy(out.leaf_tris.data(), leaf_tris, (size_t)nt * sizeof(uint32_t), cudaMemcpyDeviceToHost));
  out.max_depth = depth;
  return PT_OK;
}

}  // namespace ptb
