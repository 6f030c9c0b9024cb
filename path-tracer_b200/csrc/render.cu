// render.cu — device scene upload, wave scheduling and the C ABI for the traced/rendered paths.
// Replaces Renderer::render (reference renderer/mod.rs:76-169): instead of `samples` rayon fork-joins over
// pixels, whole waves of (pixel, sample) paths advance bounce by bounce through persistent kernels whose
// queue sizes stay on the device (no host synchronisation inside a frame).
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "host_api.hpp"
#include "kernels.cuh"

namespace ptb {

thread_local std::string g_last_error;
void set_error(const char *fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_last_error = buf;
}

#define CUDA_TRY(expr)                                                                  \
  do {                                                                                  \
    cudaError_t _e = (expr);                                                            \
    if (_e != cudaSuccess) {                                                            \
      set_error("CUDA error %s at %s:%d (%s)", cudaGetErrorString(_e), __FILE__, __LINE__, #expr); \
      return PT_ERR_CUDA;                                                               \
    }                                                                                   \
  } while (0)

struct DeviceGuard {
  int prev = -1;
  bool ok = false;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    ok = cudaSetDevice(dev) == cudaSuccess;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

}  // namespace ptb

using namespace ptb;

struct pt_scene {
  HostScene host;
  Bvh4 bvh;
  int device = 0;
  DevScene dev{};
  std::vector<void *> allocs;
  struct Staged {  // device array + its pinned host image (pt_scene_upload re-sends it)
    void *dev, *host;
    size_t bytes;
    void *dev2;          // second device copy: the target of the next pt_scene_upload (allocated on its first call)
    const void **field;  // the DevScene member that points at `dev`
  };
  std::vector<Staged> staged;
  // pt_scene_upload is double-buffered: it copies into the set of arrays the kernels are NOT reading, on its own
  // stream, and the next render / trace call switches to that set - so the upload of the next frame's scene overlaps
  // the rendering of the current one.  use_ev[k]: the last call that read set k; upload_ev: the pending upload.
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t upload_ev = nullptr, use_ev[2] = {nullptr, nullptr};
  int cur_set = 0;
  bool upload_pending = false;
  std::vector<cudaEvent_t> events;  // pool for per-kernel timing (pt_set_kernel_timing)
  uint64_t device_bytes = 0;
  double build_ms = 0;
  int accel = PT_ACCEL_SAH;
  bool accel_preloaded = false;  // bvh came validated from an isfb file: build_device_scene only uploads it
  // wave buffers, grown lazily and reused between renders (guarded: one render at a time per scene)
  std::mutex mu;
  WaveState wave{};
  std::vector<void *> wave_allocs;
  // Recorded at the end of every call that used the wave buffers, waited on at the start of the next one: calls on
  // one scene from different streams are serialised on the device (the host side is serialised by `mu`).
  cudaEvent_t busy = nullptr;
  // second stream + events of the overlapped schedule (k_shadow of bounce b beside k_extend of bounce b + 1)
  cudaStream_t side = nullptr;
  cudaEvent_t ev_shade = nullptr, ev_shadow = nullptr;
  // frame buffers of pt_render / pt_render_multi and staging buffers of the host-pointer trace entry points, kept
  // between calls (cudaFree synchronises the device: not once per frame)
  struct FrameBuf {
    void *p = nullptr;
    size_t bytes = 0;
  } multi_accum, multi_tmp, multi_rgb, trace_in, trace_aux, trace_out;
  std::mutex multi_mu;  // held by pt_render_multi over every replica of the call (taken in address order)
};

namespace {

// `out` must be a member of sc->dev (it is re-pointed when pt_scene_upload switches array sets).
template <typename T>
int upload(pt_scene *sc, const std::vector<T> &v, const T **out) {
  size_t bytes = std::max<size_t>(v.size(), 1) * sizeof(T);
  void *p = nullptr;
  CUDA_TRY(cudaMalloc(&p, bytes));
  sc->allocs.push_back(p);
  sc->device_bytes += bytes;
  if (!v.empty()) {
    void *hp = nullptr;
    CUDA_TRY(cudaMallocHost(&hp, v.size() * sizeof(T)));
    std::memcpy(hp, v.data(), v.size() * sizeof(T));
    sc->staged.push_back({p, hp, v.size() * sizeof(T), nullptr, reinterpret_cast<const void **>(out)});
    CUDA_TRY(cudaMemcpy(p, hp, v.size() * sizeof(T), cudaMemcpyHostToDevice));
  }
  *out = reinterpret_cast<const T *>(p);
  return PT_OK;
}

bool material_may_be_transparent(const pt_material &m) { return !(m.opacity_tex < 0 && m.opacity >= 1.0f); }
uint32_t material_flags(const pt_material &m) {
  return (material_may_be_transparent(m) ? TRI_FLAG_ALPHA : 0u) | (m.opacity_tex < 0 && m.opacity == 1.0f ? TRI_FLAG_OPAQUE1 : 0u);
}

int g_accel = -1;  // -1: not set through the API -> $PT_B200_ACCEL or SAH
int accel_choice() {
  if (g_accel >= 0) return g_accel;
  if (const char *e = std::getenv("PT_B200_ACCEL")) {
    if (!std::strcmp(e, "lbvh")) return PT_ACCEL_LBVH;
    if (!std::strcmp(e, "ploc")) return PT_ACCEL_PLOC;
    if (!std::strcmp(e, "auto")) return PT_ACCEL_AUTO;
  }
  return PT_ACCEL_SAH;
}

int build_device_scene(pt_scene *sc) {
  HostScene &h = sc->host;
  unsigned hw = std::max(1u, std::thread::hardware_concurrency());
  int accel = accel_choice();
  const uint64_t n_prims_geom = h.n_tris() + h.sphere_models.size();
  if (accel == PT_ACCEL_AUTO) accel = n_prims_geom >= (4ull << 20) ? PT_ACCEL_PLOC : PT_ACCEL_SAH;
  if (n_prims_geom < 16) accel = PT_ACCEL_SAH;  // the LBVH root must not be a single leaf (leaves hold <= 8 triangles)
  bool built = sc->accel_preloaded;
  if (!built) sc->accel = PT_ACCEL_SAH;
  if (!built && (accel == PT_ACCEL_LBVH || accel == PT_ACCEL_PLOC)) {
    float ms = 0;
    int too_deep = 0;
    int rc = build_bvh4_lbvh(h, sc->bvh, &ms, &too_deep, accel == PT_ACCEL_PLOC ? 1 : 0);
    if (rc == PT_OK) {
      built = true;
      sc->accel = accel;
      sc->build_ms = ms;
    } else if (!too_deep) {
      return rc;
    }  // too deep for the traversal stack (pathological duplicates): fall through to the SAH builder
  }
  if (!built) {
    auto t0 = std::chrono::steady_clock::now();
    build_bvh4(h, sc->bvh, (int)std::min(hw, 32u));
    sc->build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  }
  if (sc->bvh.max_depth * 3 + 4 > 96) {
    set_error("acceleration structure too deep (%d levels) for the traversal stack", sc->bvh.max_depth);
    return PT_ERR_INVALID;
  }
  const uint32_t nt = (uint32_t)h.n_tris();
  bool any_alpha = false, any_normal_map = false;
  for (const pt_material &m : h.materials) {
    any_alpha |= material_may_be_transparent(m);
    any_normal_map |= m.normal_tex >= 0;
  }
  // one extra record at the end: the degenerate triangle behind the empty child slots of the quantised nodes
  // (all-zero edges: |det| < 1e-6 rejects it for every ray)
  const size_t n_leaf_tris = sc->bvh.leaf_tris.size();
  std::vector<DevTri> tris(n_leaf_tris + 1);
  tris[n_leaf_tris].a = tris[n_leaf_tris].b = tris[n_leaf_tris].c = make_float4(0.f, 0.f, 0.f, 0.f);
  for (size_t i = 0; i < n_leaf_tris; ++i) {
    uint32_t t = sc->bvh.leaf_tris[i];
    const float *p = &h.pos[(size_t)t * 9];
    int model = h.tri_model[t];
    int prim = h.prim_base[model] + (int)(t - h.models[model].first_tri);
    uint32_t flags = material_flags(h.materials[model]);
    DevTri d;
    d.a = make_float4(p[0], p[1], p[2], 0.f);
    d.b = make_float4(p[3] - p[0], p[4] - p[1], p[5] - p[2], 0.f);  // v0v1, triangle.rs:43
    d.c = make_float4(p[6] - p[0], p[7] - p[1], p[8] - p[2], 0.f);  // v0v2, triangle.rs:44
    std::memcpy(&d.a.w, &prim, 4);
    std::memcpy(&d.b.w, &t, 4);
    std::memcpy(&d.c.w, &flags, 4);
    tris[i] = d;
  }
  std::vector<DevTriShade> shade(nt);
  for (uint32_t t = 0; t < nt; ++t) {
    DevTriShade s;
    std::memcpy(s.n0, &h.nrm[(size_t)t * 9], 36);
    std::memcpy(s.uv0, &h.uv[(size_t)t * 6], 24);
    s.model = h.tri_model[t];
    shade[t] = s;
  }
  std::vector<DevSphere> spheres;
  for (int m : h.sphere_models) {
    const pt_model &mo = h.models[m];
    DevSphere s{mo.center[0], mo.center[1], mo.center[2], mo.radius, m, h.prim_base[m], material_flags(h.materials[m]), 0u};
    spheres.push_back(s);
  }
  std::vector<DevMaterial> mats(h.materials.size());
  for (size_t i = 0; i < mats.size(); ++i) {
    static_assert(sizeof(DevMaterial) == sizeof(pt_material), "layout");
    std::memcpy(&mats[i], &h.materials[i], sizeof(pt_material));
  }
  std::vector<DevTexture> texs;
  std::vector<uint8_t> texels;
  for (size_t i = 0; i < h.textures.size(); ++i) {
    DevTexture t{(unsigned long long)texels.size(), h.textures[i].width, h.textures[i].height, h.textures[i].channels, 0};
    texels.insert(texels.end(), h.texdata[i].begin(), h.texdata[i].end());
    while (texels.size() % 16) texels.push_back(0);
    texs.push_back(t);
  }
  std::vector<DevLight> lights;
  for (const pt_light &l : h.lights) {
    DevLight d{l.type, {l.vec[0], l.vec[1], l.vec[2]}, {l.color[0], l.color[1], l.color[2]}, 0.f};
    lights.push_back(d);
  }
  DevScene &D = sc->dev;
  int rc;
  static_assert(sizeof(DevNode) == sizeof(Bvh4Node), "node layout");
  static_assert(sizeof(DevQNode) == sizeof(Bvh4QNode), "quantised node layout");
  D.nodes = nullptr;
  D.cnodes = nullptr;
  D.qnodes = nullptr;
#if defined(PTB_BOTH_NODE_FORMATS)
  const bool want_float = true, want_c = true, want_q = true;  // A/B builds: one host library under kernels of any kind
#elif PTB_QNODE
  const bool want_float = false, want_c = false, want_q = true;
#elif PTB_CNODE
  const bool want_float = false, want_c = true, want_q = false;
#else
  const bool want_float = true, want_c = false, want_q = false;
#endif
  if (want_float)
    if ((rc = upload(sc, sc->bvh.nodes, reinterpret_cast<const Bvh4Node **>(&D.nodes)))) return rc;
  if (want_c) {
    // centre / half-extent per child and axis, the half-extent rounded up so that [c - h, c + h] contains [lo, hi]
    std::vector<Bvh4Node> cn(sc->bvh.nodes);
    for (Bvh4Node &nd : cn) {
      float *lo[3] = {nd.lo_x, nd.lo_y, nd.lo_z}, *hi[3] = {nd.hi_x, nd.hi_y, nd.hi_z};
      for (int a = 0; a < 3; ++a)
        for (int k = 0; k < 4; ++k) {
          const double l = lo[a][k], h = hi[a][k];
          if (!(l <= h)) continue;  // empty slot: NaN planes stay NaN
          const float c = (float)(0.5 * (l + h));
          const double hd = std::max(h - (double)c, (double)c - l);
          float he = (float)hd;
          if ((double)he < hd) he = std::nextafterf(he, INFINITY);
          lo[a][k] = c;
          hi[a][k] = he;
        }
    }
    if ((rc = upload(sc, cn, reinterpret_cast<const Bvh4Node **>(&D.cnodes)))) return rc;
  }
  if (want_q) {
    if (n_leaf_tris >= (1u << 27)) {
      set_error("too many triangles for the leaf reference encoding");
      return PT_ERR_INVALID;
    }
    std::vector<Bvh4QNode> qn;
    quantise_bvh4(sc->bvh, BVH_LEAF | ((uint32_t)n_leaf_tris << 3), (int)std::min(hw, 32u), qn);
    if ((rc = upload(sc, qn, reinterpret_cast<const Bvh4QNode **>(&D.qnodes)))) return rc;
  }
  if ((rc = upload(sc, tris, &D.tris))) return rc;
  if ((rc = upload(sc, shade, &D.shade))) return rc;
  if ((rc = upload(sc, spheres, &D.spheres))) return rc;
  if ((rc = upload(sc, mats, &D.materials))) return rc;
  if ((rc = upload(sc, texs, &D.textures))) return rc;
  if ((rc = upload(sc, texels, &D.texels))) return rc;
  if ((rc = upload(sc, lights, &D.lights))) return rc;
  D.tri_pos9 = nullptr;
  if (any_normal_map)
    if ((rc = upload(sc, h.pos, &D.tri_pos9))) return rc;
  {
    float *lut = nullptr;
    CUDA_TRY(cudaMalloc(&lut, 512 * sizeof(float)));
    sc->allocs.push_back(lut);
    launch_init_srgb_lut(lut, nullptr);
    CUDA_TRY(cudaGetLastError());
    D.srgb_lut = lut;
  }
  D.n_tris = nt;
  D.n_spheres = (uint32_t)spheres.size();
  D.n_lights = (uint32_t)lights.size();
  D.n_nodes = (uint32_t)sc->bvh.nodes.size();
  D.root_is_valid = 1;
  D.tree_depth = (uint32_t)std::max(0, sc->bvh.max_depth);
  D.any_alpha = any_alpha ? 1u : 0u;
  D.f32_one_bits = 0x3F800000u;
  std::memcpy(D.camera, h.camera, sizeof D.camera);
  D.fov = h.fov;
  D.tan_half_fov = std::tan(h.fov / 2.0f);  // f32 tanf of the C library == Rust's f32::tan
  std::memcpy(D.background, h.background, sizeof D.background);
  CUDA_TRY(cudaEventCreateWithFlags(&sc->busy, cudaEventDisableTiming));
  return PT_OK;
}

void free_wave(pt_scene *sc) {
  for (void *p : sc->wave_allocs) cudaFree(p);
  sc->wave_allocs.clear();
  sc->wave = WaveState{};
}

// The wave state lives in ONE device allocation carved into its arrays (2 MiB-aligned pieces), counters first.
constexpr size_t WAVE_ALIGN = (size_t)2 << 20;
size_t wave_piece(size_t count, size_t elem) { return (std::max<size_t>(count, 1) * elem + WAVE_ALIGN - 1) & ~(WAVE_ALIGN - 1); }

int ensure_wave(pt_scene *sc, uint32_t n_paths) {
  // queue positions are 32-bit and the shadow queue holds up to n_lights entries per path
  const uint64_t shadow_cap64 = (uint64_t)n_paths * std::max(1u, sc->dev.n_lights);
  if (shadow_cap64 > 0xFFFFFF00ull) {
    set_error("%u paths x %u lights do not fit the 32-bit shadow queue (render a smaller frame or fewer lights)", n_paths,
              sc->dev.n_lights);
    return PT_ERR_INVALID;
  }
  const uint32_t shadow_cap = (uint32_t)shadow_cap64;
  if (sc->wave.capacity >= n_paths && sc->wave.shadow_capacity >= shadow_cap) return PT_OK;
  free_wave(sc);
  const size_t bytes = wave_piece(CTR_COUNT, sizeof(uint32_t)) + 8 * wave_piece(n_paths, sizeof(float4)) +
                       wave_piece(n_paths, sizeof(uint32_t)) + 4 * wave_piece(shadow_cap, sizeof(float4));
  void *mem = nullptr;
  if (cudaMalloc(&mem, bytes) != cudaSuccess) {
    cudaGetLastError();
    set_error("out of device memory: %.1f GB of wave state for %u paths", bytes / 1e9, n_paths);
    return PT_ERR_NOMEM;
  }
  sc->wave_allocs.push_back(mem);
  char *base = static_cast<char *>(mem);
  size_t off = 0;
  auto take = [&](size_t count, size_t elem) {
    char *p = base + off;
    off += wave_piece(count, elem);
    return p;
  };
  WaveState &W = sc->wave;
  // $PT_B200_CTR_OFFSET moves the counters inside their 2 MiB piece: the A/B knob behind the note at cta_append2
  // (kernels.cu) - which L2 slice owns the queue counters decided k_shade's speed while every warp reserved on its own
  size_t ctr_off = 0;
  if (const char *e = std::getenv("PT_B200_CTR_OFFSET")) ctr_off = ((size_t)std::strtoull(e, nullptr, 0) & ~(size_t)255) % (WAVE_ALIGN - 4096);
  W.counters = reinterpret_cast<uint32_t *>(take(CTR_COUNT, sizeof(uint32_t)) + ctr_off);
  for (int b = 0; b < 2; ++b) {
    W.ray_o[b] = reinterpret_cast<float4 *>(take(n_paths, sizeof(float4)));
    W.ray_d[b] = reinterpret_cast<float4 *>(take(n_paths, sizeof(float4)));
    W.thr[b] = reinterpret_cast<float4 *>(take(n_paths, sizeof(float4)));
  }
  W.hit = reinterpret_cast<float4 *>(take(n_paths, sizeof(float4)));
  W.hit_list = reinterpret_cast<uint32_t *>(take(n_paths, sizeof(uint32_t)));
  W.radiance = reinterpret_cast<float4 *>(take(n_paths, sizeof(float4)));
  W.sh_o = reinterpret_cast<float4 *>(take(shadow_cap, sizeof(float4)));
  W.sh_d = reinterpret_cast<float4 *>(take(shadow_cap, sizeof(float4)));
  W.sh_c = reinterpret_cast<float4 *>(take(shadow_cap, sizeof(float4)));
  W.sh_p = reinterpret_cast<float4 *>(take(shadow_cap, sizeof(float4)));
  W.capacity = n_paths;
  W.shadow_capacity = shadow_cap;
  return PT_OK;
}

// grows a cached frame buffer of the scene (current device = the scene's)
cudaError_t frame_buf(pt_scene::FrameBuf &b, size_t bytes) {
  if (b.bytes >= bytes) return cudaSuccess;
  cudaFree(b.p);
  b = pt_scene::FrameBuf{};
  cudaError_t e = cudaMalloc(&b.p, bytes);
  if (e == cudaSuccess) b.bytes = bytes;
  return e;
}

// Called with sc->mu held at the start of every call that launches kernels on the scene: a pending pt_scene_upload
// becomes visible (the stream waits for the copy, the DevScene pointers switch to the freshly written set).
void apply_pending_upload(pt_scene *sc, cudaStream_t st) {
  if (!sc->upload_pending) return;
  cudaStreamWaitEvent(st, sc->upload_ev, 0);
  for (auto &e : sc->staged) {
    std::swap(e.dev, e.dev2);
    *e.field = e.dev;
  }
  sc->cur_set ^= 1;
  sc->upload_pending = false;
}
// ... and at its end: the array set it read stays untouched until these kernels have finished.
void mark_scene_use(pt_scene *sc, cudaStream_t st) {
  if (sc->use_ev[sc->cur_set]) cudaEventRecord(sc->use_ev[sc->cur_set], st);
}

// Host-side mutual exclusion on the scene's wave buffers plus device-side ordering between streams: the stream waits
// for the previous user of the buffers before it touches them, and records itself as that user on the way out.
struct WaveLock {
  pt_scene *sc;
  cudaStream_t st;
  std::unique_lock<std::mutex> lk;
  WaveLock(pt_scene *s, cudaStream_t stream) : sc(s), st(stream), lk(s->mu) {
    if (sc->busy) cudaStreamWaitEvent(st, sc->busy, 0);
    apply_pending_upload(sc, st);
  }
  ~WaveLock() {
    if (sc->busy) cudaEventRecord(sc->busy, st);
    mark_scene_use(sc, st);
  }
};

// $PT_B200_OVERLAP (default on): k_shadow(b) only adds to the radiance sums, so it runs on a second stream beside
// k_hit_list / k_extend of bounce b + 1 and the tails of the persistent kernels overlap; k_shade(b + 1), which refills
// the shadow queue, waits for it.  Off while per-kernel timing is requested (one event timeline).
bool overlap_enabled() {
  static const int on = [] {
    const char *e = std::getenv("PT_B200_OVERLAP");
    return e ? std::atoi(e) : 1;
  }();
  return on != 0;
}

int no_device() {
  set_error("scene was created host-only (device -1): no CPU fallback exists for tracing or rendering");
  return PT_ERR_CUDA;
}

int check_profile(const pt_profile *p) {
  if (!p || p->width == 0 || p->height == 0) {
    set_error("profile resolution must be non-zero");
    return PT_ERR_INVALID;
  }
  if ((uint64_t)p->width * p->height > (1ull << 28)) {
    set_error("resolution too large");
    return PT_ERR_INVALID;
  }
  if (p->tonemap < 0 || p->tonemap > 2 || p->brdf != PT_BRDF_COOK_TORRANCE) {
    set_error("unknown tonemap or brdf");
    return PT_ERR_INVALID;
  }
  return PT_OK;
}

}  // namespace

namespace ptb {
int scene_from_host(HostScene &&host, int device, pt_scene **out, Bvh4 *accel, int accel_kind) {
  if (!out) return PT_ERR_INVALID;
  *out = nullptr;
  std::string err = host.finalize();
  if (!err.empty()) {
    set_error("%s", err.c_str());
    return PT_ERR_INVALID;
  }
  if (accel && accel_kind >= 0) {
    err = validate_bvh4(host, *accel);
    if (!err.empty()) {
      set_error("stored acceleration structure rejected: %s", err.c_str());
      return PT_ERR_PARSE;
    }
  }
  auto adopt = [&](pt_scene *sc) {
    if (accel && accel_kind >= 0) {
      sc->bvh = std::move(*accel);
      sc->accel = accel_kind == PT_ACCEL_LBVH || accel_kind == PT_ACCEL_PLOC ? accel_kind : PT_ACCEL_SAH;
      sc->accel_preloaded = true;
    }
  };
  if (device == -1) {  // host-only scene: loaders / generators / introspection without a GPU
    pt_scene *sc = new pt_scene();
    sc->host = std::move(host);
    sc->device = -1;
    adopt(sc);
    *out = sc;
    return PT_OK;
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    set_error("no CUDA device available (this library has no CPU fallback)");
    return PT_ERR_CUDA;
  }
  if (device < 0 || device >= ndev) {
    set_error("device %d out of range (%d devices)", device, ndev);
    return PT_ERR_INVALID;
  }
  DeviceGuard g(device);
  pt_scene *sc = new pt_scene();
  sc->host = std::move(host);
  sc->device = device;
  adopt(sc);
  int rc = build_device_scene(sc);
  if (rc != PT_OK) {
    pt_scene_destroy(sc);
    return rc;
  }
  *out = sc;
  return PT_OK;
}
}  // namespace ptb

extern "C" {

const char *pt_last_error(void) { return g_last_error.c_str(); }

int pt_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

int pt_scene_create(const pt_scene_desc *desc, int device, pt_scene **out) {
  if (!desc || !out) {
    set_error("null argument");
    return PT_ERR_INVALID;
  }
  if (desc->n_models && (!desc->models || !desc->materials)) {
    set_error("null model/material arrays");
    return PT_ERR_INVALID;
  }
  return abi_guard([&]() -> int {
    HostScene h;
    h.from_desc(*desc);
    return scene_from_host(std::move(h), device, out);
  });
}

void pt_scene_destroy(pt_scene *sc) {
  if (!sc) return;
  if (sc->device < 0) {
    delete sc;
    return;
  }
  DeviceGuard g(sc->device);
  free_wave(sc);
  for (pt_scene::FrameBuf *b : {&sc->multi_accum, &sc->multi_tmp, &sc->multi_rgb, &sc->trace_in, &sc->trace_aux, &sc->trace_out})
    cudaFree(b->p);
  if (sc->busy) cudaEventDestroy(sc->busy);
  if (sc->upload_ev) cudaEventDestroy(sc->upload_ev);
  for (int k = 0; k < 2; ++k)
    if (sc->use_ev[k]) cudaEventDestroy(sc->use_ev[k]);
  if (sc->copy_stream) cudaStreamDestroy(sc->copy_stream);
  if (sc->ev_shade) cudaEventDestroy(sc->ev_shade);
  if (sc->ev_shadow) cudaEventDestroy(sc->ev_shadow);
  if (sc->side) cudaStreamDestroy(sc->side);
  for (void *p : sc->allocs) cudaFree(p);
  for (auto &s : sc->staged) cudaFreeHost(s.host);
  for (cudaEvent_t e : sc->events) cudaEventDestroy(e);
  delete sc;
}

// The second array set, its stream and events (first upload of a scene; sc->mu held, current device = the scene's).
static int ensure_upload_sets(pt_scene *sc) {
  if (sc->copy_stream) return PT_OK;
  CUDA_TRY(cudaStreamCreateWithFlags(&sc->copy_stream, cudaStreamNonBlocking));
  CUDA_TRY(cudaEventCreateWithFlags(&sc->upload_ev, cudaEventDisableTiming));
  for (int k = 0; k < 2; ++k) CUDA_TRY(cudaEventCreateWithFlags(&sc->use_ev[k], cudaEventDisableTiming));
  for (auto &e : sc->staged) {
    CUDA_TRY(cudaMalloc(&e.dev2, e.bytes));
    sc->allocs.push_back(e.dev2);  // freed with the scene whichever set is current by then
    sc->device_bytes += e.bytes;
  }
  return PT_OK;
}

int pt_scene_upload_targets(pt_scene *sc, void *stream, void **ptrs, uint64_t *bytes, uint32_t capacity, uint32_t *count) {
  if (!sc || !count || (capacity && (!ptrs || !bytes))) {
    set_error("null argument");
    return PT_ERR_INVALID;
  }
  if (sc->device < 0) return no_device();
  DeviceGuard g(sc->device);
  std::lock_guard<std::mutex> lock(sc->mu);
  int rc = ensure_upload_sets(sc);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  // e.dev2 is the set that is not current - also while an upload into it is pending (the switch happens at the next
  // render): wait for its last readers and for that upload
  CUDA_TRY(cudaStreamWaitEvent(st, sc->use_ev[sc->cur_set ^ 1], 0));
  if (sc->upload_pending) CUDA_TRY(cudaStreamWaitEvent(st, sc->upload_ev, 0));
  uint32_t n = 0;
  for (auto &e : sc->staged) {
    if (n < capacity) {
      ptrs[n] = e.dev2;
      bytes[n] = e.bytes;
    }
    ++n;
  }
  *count = n;
  return PT_OK;
}

int pt_scene_upload_commit(pt_scene *sc, void *stream) {
  if (!sc) {
    set_error("null argument");
    return PT_ERR_INVALID;
  }
  if (sc->device < 0) return no_device();
  DeviceGuard g(sc->device);
  std::lock_guard<std::mutex> lock(sc->mu);
  if (!sc->copy_stream) {
    set_error("pt_scene_upload_commit without pt_scene_upload_targets");
    return PT_ERR_INVALID;
  }
  // a pending upload's event is replaced by one that is later on `stream`, which waited for it in _targets
  CUDA_TRY(cudaEventRecord(sc->upload_ev, (cudaStream_t)stream));
  sc->upload_pending = true;
  return PT_OK;
}

static int upload_impl(pt_scene *sc, uint32_t part, uint32_t parts, uint64_t *bytes_out);

int pt_scene_upload(pt_scene *sc, void *stream, uint64_t *bytes_out) {
  (void)stream;  // kept for the ABI: the copies run on the scene's own copy stream (see the header)
  return upload_impl(sc, 0, 1, bytes_out);
}

int pt_scene_upload_part(pt_scene *sc, uint32_t part, uint32_t parts, uint64_t *bytes_out) {
  if (parts == 0 || part >= parts) {
    set_error("pt_scene_upload_part: part %u of %u", part, parts);
    return PT_ERR_INVALID;
  }
  return upload_impl(sc, part, parts, bytes_out);
}

static int upload_impl(pt_scene *sc, uint32_t part, uint32_t parts, uint64_t *bytes_out) {
  if (!sc) {
    set_error("null argument");
    return PT_ERR_INVALID;
  }
  if (sc->device < 0) return no_device();
  DeviceGuard g(sc->device);
  std::lock_guard<std::mutex> lock(sc->mu);
  {
    int rc = ensure_upload_sets(sc);
    if (rc) return rc;
  }
  // the set the kernels are not reading; wait for the last kernels that did read it (two frames ago)
  const int target = sc->cur_set ^ 1;
  CUDA_TRY(cudaStreamWaitEvent(sc->copy_stream, sc->use_ev[target], 0));
  uint64_t total = 0;
  for (auto &e : sc->staged) {
    if (parts == 1) {
      CUDA_TRY(cudaMemcpyAsync(e.dev2, e.host, e.bytes, cudaMemcpyHostToDevice, sc->copy_stream));
      total += e.bytes;
      continue;
    }
    const size_t shard = (e.bytes / parts) & ~(size_t)255, tail = (size_t)parts * shard;
    char *d = static_cast<char *>(e.dev2);
    const char *h = static_cast<const char *>(e.host);
    if (shard) CUDA_TRY(cudaMemcpyAsync(d + part * shard, h + part * shard, shard, cudaMemcpyHostToDevice, sc->copy_stream));
    if (e.bytes > tail) CUDA_TRY(cudaMemcpyAsync(d + tail, h + tail, e.bytes - tail, cudaMemcpyHostToDevice, sc->copy_stream));
    total += shard + (e.bytes - tail);
  }
  CUDA_TRY(cudaEventRecord(sc->upload_ev, sc->copy_stream));
  sc->upload_pending = true;
  if (bytes_out) *bytes_out = total;
  return PT_OK;
}

static int g_kernel_timing = 0;
void pt_set_kernel_timing(int on) { g_kernel_timing = on; }

int pt_scene_get_desc(const pt_scene *sc, pt_scene_desc *out) {
  if (!sc || !out) return PT_ERR_INVALID;
  sc->host.to_desc(*out);
  return PT_OK;
}

int pt_set_accel_builder(int accel) {
  if (accel < PT_ACCEL_SAH || accel > PT_ACCEL_PLOC) {
    set_error("unknown acceleration-structure builder %d", accel);
    return PT_ERR_INVALID;
  }
  g_accel = accel;
  return PT_OK;
}

int pt_scene_accel_info(const pt_scene *sc, int *accel, double *build_ms, int *max_depth) {
  if (!sc) return PT_ERR_INVALID;
  if (accel) *accel = sc->accel;
  if (build_ms) *build_ms = sc->build_ms;
  if (max_depth) *max_depth = sc->bvh.max_depth;
  return PT_OK;
}

int pt_scene_save_binary(const pt_scene *sc, const char *path, int with_accel) {
  if (!sc || !path) {
    set_error("null argument");
    return PT_ERR_INVALID;
  }
  if (!with_accel) return save_isfb_file(path, sc->host, nullptr, -1);
  if (!sc->bvh.nodes.empty()) return save_isfb_file(path, sc->host, &sc->bvh, sc->accel);
  // host-only scene (device -1): the SAH builder is host code, so the tree can be written without a GPU
  Bvh4 bvh;
  build_bvh4(sc->host, bvh, (int)std::min(std::max(1u, std::thread::hardware_concurrency()), 32u));
  if (bvh.max_depth * 3 + 4 > 96) {
    set_error("acceleration structure too deep (%d levels) for the traversal stack", bvh.max_depth);
    return PT_ERR_INVALID;
  }
  return save_isfb_file(path, sc->host, &bvh, PT_ACCEL_SAH);
}

int pt_scene_info(const pt_scene *sc, uint64_t *n_tris, uint64_t *n_prims, uint64_t *n_nodes, uint64_t *device_bytes) {
  if (!sc) return PT_ERR_INVALID;
  if (n_tris) *n_tris = sc->host.n_tris();
  if (n_prims) *n_prims = (uint64_t)sc->host.n_prims;
  if (n_nodes) *n_nodes = sc->bvh.nodes.size();
  if (device_bytes) *device_bytes = sc->device_bytes;
  return PT_OK;
}

// ---- fixed ray batches through the production kernels ----
// Rays per pass through the wave queues: bounds the wave memory a batch needs (~200 B per ray) whatever n is.
static constexpr size_t TRACE_CHUNK = 16u << 20;

// rays_dev / out_dev: device pointers on the scene's device.  alpha = false: ray_cast(...).first(); alpha = true: the
// hit the alpha walk accepts with the draws of path (pixel = ray index, sample, seed) at `bounce`.
static int trace_closest_dev(pt_scene *sc, const float *rays_dev, size_t n, size_t index_base, pt_hit *out_dev, bool alpha,
                             uint32_t bounce, uint32_t sample, uint64_t seed, cudaStream_t st) {
  WaveLock lock(sc, st);
  int rc = ensure_wave(sc, (uint32_t)std::min(n, TRACE_CHUNK));
  if (rc) return rc;
  const WaveState &W = sc->wave;
  for (size_t b = 0; b < n; b += TRACE_CHUNK) {
    const uint32_t c = (uint32_t)std::min(TRACE_CHUNK, n - b);
    WaveParams P{};
    P.width = 0x80000000u;  // make_key: pixel = slot, sample = sample_first
    P.height = 1;
    P.bounces = bounce;
    P.spp_in_wave = 1;
    P.sample_first = sample;
    P.sample_stride = 1;
    P.seed_lo = (uint32_t)seed;
    P.seed_hi = (uint32_t)(seed >> 32);
    P.n_paths = c;
    launch_load_rays(rays_dev + b * 6, c, (uint32_t)(index_base + b), W, bounce & 1u, st);
    launch_extend(sc->dev, P, W, bounce, alpha, out_dev + b, st);
  }
  CUDA_TRY(cudaGetLastError());
  return PT_OK;
}

int pt_trace_closest_device(const pt_scene *sc, const float *rays_dev, size_t n, pt_hit *out_dev, void *stream) {
  if (!sc || (n && (!rays_dev || !out_dev))) {
    set_error("bad argument");
    return PT_ERR_INVALID;
  }
  if (n == 0) return PT_OK;
  if (sc->device < 0) return no_device();
  DeviceGuard g(sc->device);
  return trace_closest_dev(const_cast<pt_scene *>(sc), rays_dev, n, 0, out_dev, false, 0, 0, 0, (cudaStream_t)stream);
}

// The same query without any allocation: the wave scratch is the static device array of kernels.cu (3 MiB + 256 B per
// GPU) instead of the scene's wave arena, so n is limited and overlapping calls race (include/pt_b200.h).
int pt_trace_closest_device_noalloc(const pt_scene *sc, const float *rays_dev, size_t n, pt_hit *out_dev, void *stream) {
  if (!sc || (n && (!rays_dev || !out_dev))) {
    set_error("bad argument");
    return PT_ERR_INVALID;
  }
  if (n > (size_t)PT_TRACE_NOALLOC_MAX_RAYS) {
    set_error("pt_trace_closest_device_noalloc: n = %zu exceeds PT_TRACE_NOALLOC_MAX_RAYS = %d (the static scratch array); "
              "split the batch or use pt_trace_closest_device", n, PT_TRACE_NOALLOC_MAX_RAYS);
    return PT_ERR_INVALID;
  }
  if (n == 0) return PT_OK;
  if (sc->device < 0) return no_device();
  DeviceGuard g(sc->device);
  pt_scene *msc = const_cast<pt_scene *>(sc);
  WaveLock lock(msc, (cudaStream_t)stream);  // ordering against pt_scene_upload and the scene's other users; no allocation
  CUDA_TRY(launch_trace_closest_static(msc->dev, rays_dev, (uint32_t)n, out_dev, (cudaStream_t)stream));
  return PT_OK;
}

// Host-pointer batches: staged chunk by chunk through buffers cached on the scene.
static int trace_closest_host(const pt_scene *csc, const float *rays, size_t n, pt_hit *out, bool alpha, uint32_t bounce,
                              uint32_t sample, uint64_t seed) {
  pt_scene *sc = const_cast<pt_scene *>(csc);
  if (!sc || (n && (!rays || !out)) || (alpha && n > 0x7FFFFFFFull)) {
    set_error("bad argument");
    return PT_ERR_INVALID;
  }
  if (n == 0) return PT_OK;
  if (sc->device < 0) return no_device();
  DeviceGuard g(sc->device);
  std::lock_guard<std::mutex> hold(sc->multi_mu);  // the staging buffers
  const size_t cap = std::min(n, TRACE_CHUNK);
  CUDA_TRY(frame_buf(sc->trace_in, cap * 6 * sizeof(float)));
  CUDA_TRY(frame_buf(sc->trace_out, cap * sizeof(pt_hit)));
  float *d_rays = static_cast<float *>(sc->trace_in.p);
  pt_hit *d_out = static_cast<pt_hit *>(sc->trace_out.p);
  for (size_t b = 0; b < n; b += TRACE_CHUNK) {
    const size_t c = std::min(TRACE_CHUNK, n - b);
    CUDA_TRY(cudaMemcpy(d_rays, rays + b * 6, c * 6 * sizeof(float), cudaMemcpyHostToDevice));
    int rc = trace_closest_dev(sc, d_rays, c, b, d_out, alpha, bounce, sample, seed, nullptr);
    if (rc) return rc;
    CUDA_TRY(cudaMemcpy(out + b, d_out, c * sizeof(pt_hit), cudaMemcpyDeviceToHost));
  }
  return PT_OK;
}

int pt_trace_closest(const pt_scene *sc, const float *rays, size_t n, pt_hit *out) {
  return trace_closest_host(sc, rays, n, out, false, 0, 0, 0);
}

int pt_trace_accepted(const pt_scene *sc, const float *rays, size_t n, uint32_t bounce, uint32_t sample, uint64_t seed,
                      pt_hit *out) {
  return trace_closest_host(sc, rays, n, out, true, bounce, sample, seed);
}

int pt_trace_shadow(const pt_scene *csc, const float *rays, const float *limits, const float *primary, size_t n,
                    float *out) {
  pt_scene *sc = const_cast<pt_scene *>(csc);
  if (!sc || (n && (!rays || !out)) || ((limits == nullptr) != (primary == nullptr))) {
    set_error("bad argument");
    return PT_ERR_INVALID;
  }
  if (n == 0) return PT_OK;
  if (sc->device < 0) return no_device();
  DeviceGuard g(sc->device);
  std::lock_guard<std::mutex> hold(sc->multi_mu);
  const size_t cap = std::min(n, TRACE_CHUNK);
  CUDA_TRY(frame_buf(sc->trace_in, cap * 6 * sizeof(float)));
  CUDA_TRY(frame_buf(sc->trace_out, cap * sizeof(float)));
  if (limits) CUDA_TRY(frame_buf(sc->trace_aux, cap * 6 * sizeof(float)));
  float *d_rays = static_cast<float *>(sc->trace_in.p), *d_out = static_cast<float *>(sc->trace_out.p);
  float *d_lim = limits ? static_cast<float *>(sc->trace_aux.p) : nullptr, *d_prim = limits ? d_lim + cap : nullptr;
  for (size_t b = 0; b < n; b += TRACE_CHUNK) {
    const uint32_t c = (uint32_t)std::min(TRACE_CHUNK, n - b);
    CUDA_TRY(cudaMemcpy(d_rays, rays + b * 6, (size_t)c * 6 * sizeof(float), cudaMemcpyHostToDevice));
    if (limits) {
      CUDA_TRY(cudaMemcpy(d_lim, limits + b, (size_t)c * sizeof(float), cudaMemcpyHostToDevice));
      CUDA_TRY(cudaMemcpy(d_prim, primary + b * 5, (size_t)c * 5 * sizeof(float), cudaMemcpyHostToDevice));
    }
    {
      WaveLock lock(sc, nullptr);
      int rc = ensure_wave(sc, c);
      if (rc) return rc;
      launch_load_shadow(d_rays, d_lim, d_prim, c, sc->wave, 1u, nullptr);
      launch_shadow(sc->dev, sc->wave, 0u, nullptr);  // bounce 0 reads the shadow queue filled "with buffer 1"
      launch_store_transmittance(sc->wave, c, d_out, nullptr);
      CUDA_TRY(cudaGetLastError());
    }
    CUDA_TRY(cudaMemcpy(out + b, d_out, (size_t)c * sizeof(float), cudaMemcpyDeviceToHost));
  }
  return PT_OK;
}

int pt_trace_transmittance(const pt_scene *sc, const float *rays, size_t n, float *out) {
  return pt_trace_shadow(sc, rays, nullptr, nullptr, n, out);
}

int pt_triangle_intersect(const pt_scene *sc, int device, const float *rays, const float *tris, size_t n, pt_hit *out) {
  if ((n && (!rays || !tris || !out)) || n > 0x7FFFFFFFull) {
    set_error("bad argument");
    return PT_ERR_INVALID;
  }
  if (n == 0) return PT_OK;
  if (sc) device = sc->device;
  if (device < 0 || device >= pt_device_count()) {
    set_error("no CUDA device %d (this library has no CPU fallback)", device);
    return PT_ERR_CUDA;
  }
  DeviceGuard g(device);
  float *d_rays = nullptr, *d_tris = nullptr;
  pt_hit *d_out = nullptr;
  cudaError_t e = cudaMalloc(&d_rays, n * 6 * sizeof(float));
  if (e == cudaSuccess) e = cudaMalloc(&d_tris, n * 9 * sizeof(float));
  if (e == cudaSuccess) e = cudaMalloc(&d_out, n * sizeof(pt_hit));
  if (e == cudaSuccess) e = cudaMemcpy(d_rays, rays, n * 6 * sizeof(float), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(d_tris, tris, n * 9 * sizeof(float), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    launch_triangle_intersect(d_rays, d_tris, (uint32_t)n, d_out, nullptr);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpy(out, d_out, n * sizeof(pt_hit), cudaMemcpyDeviceToHost);
  cudaFree(d_rays);
  cudaFree(d_tris);
  cudaFree(d_out);
  CUDA_TRY(e);
  return PT_OK;
}

int pt_philox4x32_10(int device, const uint32_t *counters, const uint32_t *keys, size_t n, uint32_t *out) {
  if ((n && (!counters || !keys || !out)) || n > (1u << 24)) {
    set_error("bad argument");
    return PT_ERR_INVALID;
  }
  if (n == 0) return PT_OK;
  if (device < 0 || device >= pt_device_count()) {
    set_error("no CUDA device %d (this library has no CPU fallback)", device);
    return PT_ERR_CUDA;
  }
  DeviceGuard g(device);
  uint32_t *d_c = nullptr, *d_k = nullptr, *d_o = nullptr;
  cudaError_t e = cudaMalloc(&d_c, n * 16);
  if (e == cudaSuccess) e = cudaMalloc(&d_k, n * 8);
  if (e == cudaSuccess) e = cudaMalloc(&d_o, n * 16);
  if (e == cudaSuccess) e = cudaMemcpy(d_c, counters, n * 16, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(d_k, keys, n * 8, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    launch_philox(d_c, d_k, (uint32_t)n, d_o, nullptr);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpy(out, d_o, n * 16, cudaMemcpyDeviceToHost);
  cudaFree(d_c);
  cudaFree(d_k);
  cudaFree(d_o);
  CUDA_TRY(e);
  return PT_OK;
}

int pt_trace_paths(const pt_scene *sc, const pt_profile *prof, const uint32_t *pixels, size_t n_paths, uint32_t sample,
                   uint64_t seed, pt_path_vertex *out, uint32_t *counts) {
  if (!sc || (n_paths && (!pixels || !out || !counts)) || n_paths > 0x7FFFFFFFull) {
    set_error("bad argument");
    return PT_ERR_INVALID;
  }
  int rc = check_profile(prof);
  if (rc) return rc;
  if (n_paths == 0) return PT_OK;
  if (sc->device < 0) return no_device();
  const uint64_t npix = (uint64_t)prof->width * prof->height;
  for (size_t i = 0; i < n_paths; ++i)
    if (pixels[i] >= npix) {
      set_error("pixel index %u out of range", pixels[i]);
      return PT_ERR_INVALID;
    }
  DeviceGuard g(sc->device);
  WaveParams P{};
  P.width = prof->width;
  P.height = prof->height;
  P.bounces = prof->bounces;
  P.spp_in_wave = 1;
  P.sample_first = sample;
  P.sample_stride = 1;
  P.seed_lo = (uint32_t)seed;
  P.seed_hi = (uint32_t)(seed >> 32);
  P.n_paths = (uint32_t)n_paths;
  const size_t nv = n_paths * ((size_t)prof->bounces + 1);
  uint32_t *d_pix = nullptr, *d_cnt = nullptr;
  pt_path_vertex *d_out = nullptr;
  cudaError_t e = cudaMalloc(&d_pix, n_paths * sizeof(uint32_t));
  if (e == cudaSuccess) e = cudaMalloc(&d_cnt, n_paths * sizeof(uint32_t));
  if (e == cudaSuccess) e = cudaMalloc(&d_out, nv * sizeof(pt_path_vertex));
  if (e == cudaSuccess) e = cudaMemcpy(d_pix, pixels, n_paths * sizeof(uint32_t), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemset(d_out, 0, nv * sizeof(pt_path_vertex));
  if (e == cudaSuccess) {
    pt_scene *msc = const_cast<pt_scene *>(sc);
    std::lock_guard<std::mutex> lock(msc->mu);
    apply_pending_upload(msc, nullptr);
    launch_trace_paths(sc->dev, P, d_pix, (uint32_t)n_paths, d_out, d_cnt, nullptr);
    mark_scene_use(msc, nullptr);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpy(out, d_out, nv * sizeof(pt_path_vertex), cudaMemcpyDeviceToHost);
  if (e == cudaSuccess) e = cudaMemcpy(counts, d_cnt, n_paths * sizeof(uint32_t), cudaMemcpyDeviceToHost);
  cudaFree(d_pix);
  cudaFree(d_cnt);
  cudaFree(d_out);
  CUDA_TRY(e);
  return PT_OK;
}

int pt_render_accum_device(const pt_scene *csc, const pt_profile *prof, uint32_t sample_begin, uint32_t sample_stride,
                           uint32_t n_samples, uint64_t seed, float *accum_dev, void *stream, pt_stats *stats_out) {
  pt_scene *sc = const_cast<pt_scene *>(csc);
  if (!sc || !accum_dev) {
    set_error("null argument");
    return PT_ERR_INVALID;
  }
  int rc = check_profile(prof);
  if (rc) return rc;
  if (sample_stride == 0) sample_stride = 1;
  if (sc->device < 0) return no_device();
  DeviceGuard g(sc->device);
  cudaStream_t st = (cudaStream_t)stream;
  WaveLock lock(sc, st);
  const uint64_t npix = (uint64_t)prof->width * prof->height;
  // wave size: as many samples per pixel as fit ~256 Mi paths.  Large waves keep the late-bounce queues (a few % of the
  // paths) big enough to fill 148 SMs and amortise every launch's tail: on the console workload 8 Mi -> 64 Mi paths is
  // +21 % throughput, 64 Mi -> 128 Mi -> 256 Mi another +2.2 % / +3.5 % (kernel time 93.9 / 91.9 / 90.6 ms per frame).
  // State is ~292 B per path (78 GB at 256 Mi: this is what 180 GB of HBM are for), capped to half of the free device
  // memory and to the paths the call actually renders.
  uint64_t target_paths = 256ull << 20;
  if (const char *e = std::getenv("PT_B200_WAVE_PATHS")) target_paths = std::max<uint64_t>(1, std::strtoull(e, nullptr, 0));
  {
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) {
      uint64_t per_path = 148 + 64ull * std::max(1u, sc->dev.n_lights);  // 8 float4 arrays + the hit list + the shadow queue
      uint64_t have = (uint64_t)sc->wave.capacity * per_path;  // already allocated wave buffers count as free
      target_paths = std::min<uint64_t>(target_paths, std::max<uint64_t>(npix, (free_b + have) / 2 / per_path));
    }
  }
  // queue positions are 32-bit: the shadow queue holds up to n_lights entries per path
  target_paths = std::min<uint64_t>(target_paths, 0xFFFFFF00ull / std::max(1u, sc->dev.n_lights));
  uint32_t spw = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(n_samples ? n_samples : 1, target_paths / npix));
  if (n_samples == 0) {
    if (stats_out) std::memset(stats_out, 0, sizeof *stats_out);
    return PT_OK;
  }
  if ((rc = ensure_wave(sc, (uint32_t)(npix * spw)))) return rc;
  WaveState &W = sc->wave;
  CUDA_TRY(cudaMemsetAsync(W.counters, 0, CTR_COUNT * sizeof(uint32_t), st));
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (stats_out) {
    CUDA_TRY(cudaEventCreate(&e0));
    CUDA_TRY(cudaEventCreate(&e1));
    CUDA_TRY(cudaEventRecord(e0, st));
  }
  uint64_t launches = 0, n_ext = 0, n_shade = 0, n_shadow = 0;
  // optional per-kernel timing: one event after every launch, classes 0 extend, 1 shade, 2 shadow, 3 other
  const bool timing = stats_out && g_kernel_timing;
  std::vector<int> marks;
  auto mark = [&](int cls) {
    if (!timing) return;
    if (marks.size() >= sc->events.size()) {
      cudaEvent_t e = nullptr;
      cudaEventCreate(&e);
      sc->events.push_back(e);
    }
    cudaEventRecord(sc->events[marks.size()], st);
    marks.push_back(cls);
  };
  mark(3);
  bool overlap = overlap_enabled() && !timing && sc->dev.n_lights > 0;
  if (overlap && !sc->side) {
    if (cudaStreamCreateWithFlags(&sc->side, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&sc->ev_shade, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&sc->ev_shadow, cudaEventDisableTiming) != cudaSuccess) {
      cudaGetLastError();
      overlap = false;
    }
  }
  bool shadow_pending = false;  // a k_shadow on the side stream that the main stream has not waited for yet
  for (uint32_t done = 0; done < n_samples; done += spw) {
    WaveParams P;
    P.width = prof->width;
    P.height = prof->height;
    P.bounces = prof->bounces;
    P.spp_in_wave = std::min(spw, n_samples - done);
    P.sample_first = sample_begin + done * sample_stride;
    P.sample_stride = sample_stride;
    P.seed_lo = (uint32_t)seed;
    P.seed_hi = (uint32_t)(seed >> 32);
    P.n_paths = (uint32_t)(npix * P.spp_in_wave);
    launch_raygen(sc->dev, P, W, st);
    mark(3);
    ++launches;
    for (uint32_t b = 0; b <= prof->bounces; ++b) {
      launch_extend(sc->dev, P, W, b, sc->dev.any_alpha != 0, nullptr, st);
      mark(0);
      launch_hit_list(sc->dev, P, W, b, st);
      mark(4);  // the scan pass: reported with the shade class, listed on its own by $PT_B200_DUMP_MARKS
      if (shadow_pending) {  // k_shade refills the shadow queue the previous bounce's k_shadow reads
        CUDA_TRY(cudaStreamWaitEvent(st, sc->ev_shadow, 0));
        shadow_pending = false;
      }
      launch_shade(sc->dev, P, W, b, st);
      mark(1);
      launches += 1 + shade_launches();
      ++n_ext;
      ++n_shade;
      if (sc->dev.n_lights) {
        if (overlap) {
          CUDA_TRY(cudaEventRecord(sc->ev_shade, st));
          CUDA_TRY(cudaStreamWaitEvent(sc->side, sc->ev_shade, 0));
          launch_shadow(sc->dev, W, b, sc->side);
          CUDA_TRY(cudaEventRecord(sc->ev_shadow, sc->side));
          shadow_pending = true;
        } else {
          launch_shadow(sc->dev, W, b, st);
        }
        mark(2);
        ++launches;
        ++n_shadow;
      }
    }
    if (shadow_pending) {  // the radiance sums are complete only after the last k_shadow
      CUDA_TRY(cudaStreamWaitEvent(st, sc->ev_shadow, 0));
      shadow_pending = false;
    }
    launch_resolve(P, W, accum_dev, st);
    mark(3);
    ++launches;
  }
  CUDA_TRY(cudaGetLastError());
  if (stats_out) {
    CUDA_TRY(cudaEventRecord(e1, st));
    CUDA_TRY(cudaEventSynchronize(e1));
    std::memset(stats_out, 0, sizeof *stats_out);
    cudaEventElapsedTime(&stats_out->ms_total, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    uint32_t ctr[CTR_COUNT];
    CUDA_TRY(cudaMemcpy(ctr, W.counters, sizeof ctr, cudaMemcpyDeviceToHost));
    auto u64 = [&](int i) { return (uint64_t)ctr[i] | ((uint64_t)ctr[i + 1] << 32); };
    stats_out->samples = npix * n_samples;
    stats_out->extend_rays = u64(CTR_STAT_EXTEND);
    stats_out->shadow_rays = u64(CTR_STAT_SHADOW);
    stats_out->alpha_recasts = 0;  // the alpha walk is single-pass (traverse.cuh): nothing is re-cast
    stats_out->alpha_tests = u64(CTR_STAT_ALPHA);
    stats_out->shadow_dropped = ctr[CTR_DROPPED];
    stats_out->kernel_launches = launches;
    stats_out->extend_launches = n_ext;
    stats_out->shade_launches = n_shade;
    stats_out->shadow_launches = n_shadow;
    if (timing) {
      float acc[5] = {0, 0, 0, 0, 0};
      const bool dump = std::getenv("PT_B200_DUMP_MARKS") != nullptr;
      static const char *cls[5] = {"extend", "shade", "shadow", "other", "hit_list"};
      for (size_t i = 1; i < marks.size(); ++i) {
        float ms = 0;
        cudaEventElapsedTime(&ms, sc->events[i - 1], sc->events[i]);
        acc[marks[i]] += ms;
        if (dump) std::fprintf(stderr, "mark %zu %s %.4f ms\n", i, cls[marks[i]], ms);
      }
      acc[1] += acc[4];
      if (dump && (u64(CTR_STAT_EXT_NODES) || u64(CTR_STAT_SH_NODES)))  // a -DPTB_COUNT_STEPS build
        std::fprintf(stderr, "steps: extend %.3f nodes %.3f prims per ray; shadow %.3f nodes %.3f prims per ray\n",
                     (double)u64(CTR_STAT_EXT_NODES) / std::max<uint64_t>(1, u64(CTR_STAT_EXTEND)),
                     (double)u64(CTR_STAT_EXT_PRIMS) / std::max<uint64_t>(1, u64(CTR_STAT_EXTEND)),
                     (double)u64(CTR_STAT_SH_NODES) / std::max<uint64_t>(1, u64(CTR_STAT_SHADOW)),
                     (double)u64(CTR_STAT_SH_PRIMS) / std::max<uint64_t>(1, u64(CTR_STAT_SHADOW)));
      stats_out->ms_extend = acc[0];
      stats_out->ms_shade = acc[1];
      stats_out->ms_shadow = acc[2];
      stats_out->ms_other = acc[3];
    }
  }
  return PT_OK;
}

int pt_tonemap_device(const float *accum_dev, uint32_t width, uint32_t height, uint32_t total_samples, int32_t tonemap,
                      uint8_t *rgb8_dev, int device, void *stream) {
  if (!accum_dev || !rgb8_dev || width == 0 || height == 0 || tonemap < 0 || tonemap > 2) {
    set_error("bad argument");
    return PT_ERR_INVALID;
  }
  DeviceGuard g(device);
  launch_tonemap(accum_dev, width * height, total_samples, tonemap, rgb8_dev, (cudaStream_t)stream);
  CUDA_TRY(cudaGetLastError());
  return PT_OK;
}

int pt_render(const pt_scene *csc, const pt_profile *prof, uint64_t seed, uint8_t *rgb8_out, float *accum_out,
              pt_stats *stats_out) {
  pt_scene *sc = const_cast<pt_scene *>(csc);
  if (!sc || !rgb8_out) {
    set_error("null argument");
    return PT_ERR_INVALID;
  }
  int rc = check_profile(prof);
  if (rc) return rc;
  if (sc->device < 0) return no_device();
  DeviceGuard g(sc->device);
  const size_t npix = (size_t)prof->width * prof->height;
  // the frame buffers are cached on the scene (cudaMalloc / cudaFree once per frame cost ~1 ms and a device sync)
  std::lock_guard<std::mutex> hold(sc->multi_mu);
  if (frame_buf(sc->multi_accum, npix * 3 * sizeof(float)) != cudaSuccess || frame_buf(sc->multi_rgb, npix * 3) != cudaSuccess) {
    cudaGetLastError();
    set_error("out of device memory");
    return PT_ERR_NOMEM;
  }
  float *d_accum = static_cast<float *>(sc->multi_accum.p);
  uint8_t *d_rgb = static_cast<uint8_t *>(sc->multi_rgb.p);
  CUDA_TRY(cudaMemsetAsync(d_accum, 0, npix * 3 * sizeof(float), nullptr));
  pt_stats local;
  rc = pt_render_accum_device(sc, prof, 0, 1, prof->samples, seed, d_accum, nullptr, stats_out ? stats_out : &local);
  if (rc == PT_OK) rc = pt_tonemap_device(d_accum, prof->width, prof->height, prof->samples, prof->tonemap, d_rgb, sc->device, nullptr);
  if (rc != PT_OK) return rc;
  CUDA_TRY(cudaMemcpy(rgb8_out, d_rgb, npix * 3, cudaMemcpyDeviceToHost));
  if (accum_out) CUDA_TRY(cudaMemcpy(accum_out, d_accum, npix * 3 * sizeof(float), cudaMemcpyDeviceToHost));
  return PT_OK;
}

int pt_render_async(const pt_scene *csc, const pt_profile *prof, uint64_t seed, uint8_t *rgb8_out, void *stream) {
  pt_scene *sc = const_cast<pt_scene *>(csc);
  if (!sc || !rgb8_out) {
    set_error("null argument");
    return PT_ERR_INVALID;
  }
  int rc = check_profile(prof);
  if (rc) return rc;
  if (sc->device < 0) return no_device();
  DeviceGuard g(sc->device);
  cudaStream_t st = (cudaStream_t)stream;
  const size_t npix = (size_t)prof->width * prof->height;
  std::lock_guard<std::mutex> hold(sc->multi_mu);
  if (frame_buf(sc->multi_accum, npix * 3 * sizeof(float)) != cudaSuccess || frame_buf(sc->multi_rgb, npix * 3) != cudaSuccess) {
    cudaGetLastError();
    set_error("out of device memory");
    return PT_ERR_NOMEM;
  }
  float *d_accum = static_cast<float *>(sc->multi_accum.p);
  uint8_t *d_rgb = static_cast<uint8_t *>(sc->multi_rgb.p);
  // the frame buffers are reused by the next call: stream order (same stream) or the scene's busy event (another
  // stream) keeps a frame's tonemap + copy ahead of the next frame's memset
  if (sc->busy) CUDA_TRY(cudaStreamWaitEvent(st, sc->busy, 0));
  CUDA_TRY(cudaMemsetAsync(d_accum, 0, npix * 3 * sizeof(float), st));
  rc = pt_render_accum_device(sc, prof, 0, 1, prof->samples, seed, d_accum, st, nullptr);
  if (rc == PT_OK) rc = pt_tonemap_device(d_accum, prof->width, prof->height, prof->samples, prof->tonemap, d_rgb, sc->device, st);
  if (rc != PT_OK) return rc;
  CUDA_TRY(cudaMemcpyAsync(rgb8_out, d_rgb, npix * 3, cudaMemcpyDeviceToHost, st));
  {
    std::lock_guard<std::mutex> lock(sc->mu);
    if (sc->busy) CUDA_TRY(cudaEventRecord(sc->busy, st));  // now covers the tonemap and the copy as well
  }
  return PT_OK;
}

int pt_stream_synchronize(int device, void *stream) {
  DeviceGuard g(device);
  CUDA_TRY(cudaStreamSynchronize((cudaStream_t)stream));
  return PT_OK;
}

int pt_scene_replicate(const pt_scene *src, int device, pt_scene **out) {
  if (!src || !out) {
    set_error("null argument");
    return PT_ERR_INVALID;
  }
  *out = nullptr;
  int ndev = 0;
  if (device != -1) {
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
      cudaGetLastError();
      set_error("no CUDA device available (this library has no CPU fallback)");
      return PT_ERR_CUDA;
    }
    if (device < 0 || device >= ndev) {
      set_error("device %d out of range (%d devices)", device, ndev);
      return PT_ERR_INVALID;
    }
  }
  pt_scene *sc = new pt_scene();
  sc->host = src->host;
  std::string err = sc->host.finalize();  // re-points the texture descriptors at the copy's own texel vectors
  if (!err.empty()) {
    delete sc;
    set_error("%s", err.c_str());
    return PT_ERR_INVALID;
  }
  sc->device = device;
  if (src->device >= 0 || src->accel_preloaded) {  // the source holds a tree (built or validated at load): reuse it
    sc->bvh = src->bvh;
    sc->accel = src->accel;
    sc->accel_preloaded = true;
  }
  if (device == -1) {
    *out = sc;
    return PT_OK;
  }
  DeviceGuard g(device);
  int rc = build_device_scene(sc);
  if (rc != PT_OK) {
    pt_scene_destroy(sc);
    return rc;
  }
  *out = sc;
  return PT_OK;
}

int pt_render_multi(const pt_scene *const *replicas, uint32_t n_replicas, const pt_profile *prof, uint64_t seed,
                    uint8_t *rgb8_out, float *accum_out, pt_stats *stats_out) {
  if (!replicas || n_replicas == 0 || !rgb8_out) {
    set_error("null argument");
    return PT_ERR_INVALID;
  }
  int rc = check_profile(prof);
  if (rc) return rc;
  for (uint32_t g = 0; g < n_replicas; ++g) {
    if (!replicas[g]) {
      set_error("replica %u is null", g);
      return PT_ERR_INVALID;
    }
    if (replicas[g]->device < 0) return no_device();
    for (uint32_t k = 0; k < g; ++k)
      if (replicas[k] == replicas[g]) {
        set_error("replica %u is the same scene object as replica %u (each needs its own wave buffers)", g, k);
        return PT_ERR_INVALID;
      }
  }
  const uint32_t G = n_replicas;
  const size_t npix = (size_t)prof->width * prof->height, nf = npix * 3;
  std::vector<pt_scene *> order;
  for (uint32_t g = 0; g < G; ++g) order.push_back(const_cast<pt_scene *>(replicas[g]));
  std::sort(order.begin(), order.end());
  std::vector<std::unique_lock<std::mutex>> held;
  for (pt_scene *r : order) held.emplace_back(r->multi_mu);
  std::vector<float *> accum(G, nullptr);
  std::vector<int> rcs(G, PT_OK);
  std::vector<std::string> errs(G);
  std::vector<pt_stats> stats(G);
  // one host thread per replica (the rayon pool of mod.rs:106 becomes one wave scheduler per device); replica g
  // renders the samples s = g (mod G) of the full frame, so the union of samples is the same for every G
  auto work = [&](uint32_t g) {
    const pt_scene *sc = replicas[g];
    DeviceGuard dg(sc->device);
    const uint32_t mine = prof->samples > g ? (prof->samples - g + G - 1) / G : 0;
    pt_scene *msc = const_cast<pt_scene *>(sc);
    if (frame_buf(msc->multi_accum, nf * sizeof(float)) == cudaSuccess) accum[g] = static_cast<float *>(msc->multi_accum.p);
    if (!accum[g] || cudaMemsetAsync(accum[g], 0, nf * sizeof(float), nullptr) != cudaSuccess) {
      cudaGetLastError();
      rcs[g] = PT_ERR_NOMEM;
      errs[g] = "out of device memory";
      return;
    }
    rcs[g] = pt_render_accum_device(sc, prof, g, G, mine, seed, accum[g], nullptr, &stats[g]);
    if (rcs[g] != PT_OK) errs[g] = g_last_error;  // thread-local: hand it to the calling thread
    else if (cudaDeviceSynchronize() != cudaSuccess) {
      rcs[g] = PT_ERR_CUDA;
      errs[g] = "device synchronisation failed after rendering";
    }
  };
  {
    std::vector<std::thread> threads;
    for (uint32_t g = 1; g < G; ++g) threads.emplace_back(work, g);
    work(0);
    for (auto &t : threads) t.join();
  }
  for (uint32_t g = 0; g < G; ++g)
    if (rcs[g] != PT_OK) {
      set_error("replica %u (device %d): %s", g, replicas[g]->device, errs[g].c_str());
      return rcs[g];
    }
  // reduce on replica 0's device in replica order: peer copies (NVLink when the devices are peers) + one add each
  const int root = replicas[0]->device;
  DeviceGuard dg(root);
  pt_scene *rsc = const_cast<pt_scene *>(replicas[0]);
  cudaError_t e = cudaSuccess;
  if (G > 1) e = frame_buf(rsc->multi_tmp, nf * sizeof(float));
  if (e == cudaSuccess) e = frame_buf(rsc->multi_rgb, nf);
  float *tmp = static_cast<float *>(rsc->multi_tmp.p);
  uint8_t *d_rgb = static_cast<uint8_t *>(rsc->multi_rgb.p);
  for (uint32_t g = 1; g < G && e == cudaSuccess; ++g) {
    const int dev = replicas[g]->device;
    if (dev != root) {
      int can = 0;
      if (cudaDeviceCanAccessPeer(&can, root, dev) == cudaSuccess && can)
        if (cudaDeviceEnablePeerAccess(dev, 0) != cudaSuccess) cudaGetLastError();  // already enabled is fine
      e = cudaMemcpyPeer(tmp, root, accum[g], dev, nf * sizeof(float));
    } else {
      e = cudaMemcpy(tmp, accum[g], nf * sizeof(float), cudaMemcpyDeviceToDevice);
    }
    if (e != cudaSuccess) break;
    launch_accum_add(accum[0], tmp, (uint32_t)nf, nullptr);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) {
    launch_tonemap(accum[0], (uint32_t)npix, prof->samples, prof->tonemap, d_rgb, nullptr);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpy(rgb8_out, d_rgb, nf, cudaMemcpyDeviceToHost);
  if (e == cudaSuccess && accum_out) e = cudaMemcpy(accum_out, accum[0], nf * sizeof(float), cudaMemcpyDeviceToHost);
  CUDA_TRY(e);
  if (stats_out) std::memcpy(stats_out, stats.data(), G * sizeof(pt_stats));
  return PT_OK;
}

const char *pt_aov_name(int aov) {
  static const char *names[PT_AOV_COUNT] = {"normal", "albedo", "opacity", "metalness", "roughness", "emissive", "ior"};
  return aov >= 0 && aov < PT_AOV_COUNT ? names[aov] : nullptr;
}

int pt_debug_textures(const pt_scene *sc, uint32_t width, uint32_t height, uint8_t *planes_out, uint64_t *pixels_hit) {
  if (!sc || !planes_out || width == 0 || height == 0 || (uint64_t)width * height > (1ull << 28)) {
    set_error("bad argument");
    return PT_ERR_INVALID;
  }
  if (sc->device < 0) return no_device();
  DeviceGuard g(sc->device);
  const size_t bytes = (size_t)width * height * 3 * PT_AOV_COUNT;
  uint8_t *d_planes = nullptr;
  uint32_t *d_hit = nullptr;
  CUDA_TRY(cudaMalloc(&d_planes, bytes));
  cudaError_t e = cudaMalloc(&d_hit, sizeof(uint32_t));
  if (e == cudaSuccess) e = cudaMemset(d_planes, 0, bytes);
  if (e == cudaSuccess) e = cudaMemset(d_hit, 0, sizeof(uint32_t));
  if (e == cudaSuccess) {
    pt_scene *msc = const_cast<pt_scene *>(sc);
    std::lock_guard<std::mutex> lock(msc->mu);
    apply_pending_upload(msc, nullptr);
    launch_debug_aov(sc->dev, width, height, d_planes, d_hit, nullptr);
    mark_scene_use(msc, nullptr);
    e = cudaGetLastError();
  }
  uint32_t n_hit = 0;
  if (e == cudaSuccess) e = cudaMemcpy(planes_out, d_planes, bytes, cudaMemcpyDeviceToHost);
  if (e == cudaSuccess) e = cudaMemcpy(&n_hit, d_hit, sizeof n_hit, cudaMemcpyDeviceToHost);
  cudaFree(d_planes);
  cudaFree(d_hit);
  CUDA_TRY(e);
  if (pixels_hit) *pixels_hit = n_hit;
  return PT_OK;
}

int pt_render_progressive(const pt_scene *sc, const pt_profile *prof, uint64_t seed, uint32_t snapshot_every,
                          pt_snapshot_fn on_snapshot, void *user, uint8_t *rgb8_out, float *accum_out,
                          pt_stats *stats_out) {
  if (!sc || !rgb8_out) {
    set_error("null argument");
    return PT_ERR_INVALID;
  }
  int rc = check_profile(prof);
  if (rc) return rc;
  if (sc->device < 0) return no_device();
  if (snapshot_every == 0 || !on_snapshot) snapshot_every = prof->samples ? prof->samples : 1;
  DeviceGuard g(sc->device);
  const size_t npix = (size_t)prof->width * prof->height;
  float *d_accum = nullptr;
  uint8_t *d_rgb = nullptr;
  CUDA_TRY(cudaMalloc(&d_accum, npix * 3 * sizeof(float)));
  if (cudaMalloc(&d_rgb, npix * 3) != cudaSuccess) {
    cudaFree(d_accum);
    set_error("out of device memory");
    return PT_ERR_NOMEM;
  }
  cudaMemset(d_accum, 0, npix * 3 * sizeof(float));
  pt_stats total{}, part{};
  bool watching = on_snapshot != nullptr;
  cudaError_t e = cudaSuccess;
  uint32_t done = 0;
  do {  // at least one pass so that samples == 0 still yields the (NaN -> 0) image the reference would produce
    const uint32_t chunk = std::min(snapshot_every, prof->samples - done);
    rc = pt_render_accum_device(sc, prof, done, 1, chunk, seed, d_accum, nullptr, &part);
    if (rc != PT_OK) break;
    done += chunk;
    total.samples += part.samples;
    total.extend_rays += part.extend_rays;
    total.shadow_rays += part.shadow_rays;
    total.alpha_recasts += part.alpha_recasts;
    total.alpha_tests += part.alpha_tests;
    total.shadow_dropped += part.shadow_dropped;
    total.kernel_launches += part.kernel_launches + 1;
    total.extend_launches += part.extend_launches;
    total.shade_launches += part.shade_launches;
    total.shadow_launches += part.shadow_launches;
    total.ms_total += part.ms_total;
    total.ms_extend += part.ms_extend;
    total.ms_shade += part.ms_shade;
    total.ms_shadow += part.ms_shadow;
    total.ms_other += part.ms_other;
    const bool last = done >= prof->samples;
    if (!last && !watching) continue;
    // `*pixel / current_sample` then post_processing: what the viewer shows (mod.rs:137-138), and the final pass
    rc = pt_tonemap_device(d_accum, prof->width, prof->height, last ? prof->samples : done, prof->tonemap, d_rgb,
                           sc->device, nullptr);
    if (rc != PT_OK) break;
    e = cudaMemcpy(rgb8_out, d_rgb, npix * 3, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) break;
    // a non-zero return is the viewer window being closed (mod.rs:139-141): rendering goes on without snapshots
    if (watching && on_snapshot(user, rgb8_out, prof->width, prof->height, done) != 0) watching = false;
  } while (done < prof->samples);
  if (rc == PT_OK && e == cudaSuccess && accum_out)
    e = cudaMemcpy(accum_out, d_accum, npix * 3 * sizeof(float), cudaMemcpyDeviceToHost);
  cudaFree(d_accum);
  cudaFree(d_rgb);
  if (rc != PT_OK) return rc;
  CUDA_TRY(e);
  if (stats_out) *stats_out = total;
  return PT_OK;
}

}  // extern "C"
