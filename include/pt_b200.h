/* pt_b200.h — C ABI of the B200-native wavefront path tracer.
 *
 * This is the drop-in boundary for the hot path of flomonster/path-tracer
 * (reference = /root/reference, Rust, CPU only).  The reference has no FFI; its hot
 * path sits behind in-process Rust calls.  Every entry point below cites the
 * reference interface it replaces (file:line relative to the reference root).
 * A Rust host binds these with an `extern "C"` block.
 *
 * Conventions: every function returns PT_OK (0) or a negative pt_status; it never
 * throws or aborts across the ABI.  pt_last_error() returns a thread-local message.
 * Handles are opaque and caller-owned.  There is NO CPU fallback: a call that needs
 * the GPU fails with PT_ERR_CUDA when no device is usable.
 */
#ifndef PT_B200_H
#define PT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  PT_OK = 0,
  PT_ERR_INVALID = -1, /* bad argument / malformed scene              */
  PT_ERR_IO = -2,      /* file missing / unreadable (reference: exit 2) */
  PT_ERR_PARSE = -3,   /* JSON / YAML / PNG syntax (reference: exit 2)  */
  PT_ERR_CUDA = -4,    /* CUDA runtime error or no device               */
  PT_ERR_NOMEM = -5
} pt_status;

/* ---- flat scene description (what scene/isf.rs:5-142 deserialises into) ---- */

enum { PT_MODEL_MESH = 0, PT_MODEL_SPHERE = 1 };   /* scene/isf.rs:31-43 */
enum { PT_LIGHT_POINT = 0, PT_LIGHT_DIRECTIONAL = 1 }; /* scene/isf.rs:58-75 */
enum { PT_TONEMAP_REINHARD = 0, PT_TONEMAP_FILMIC = 1, PT_TONEMAP_ACES = 2 }; /* renderer/tonemap.rs:4-13 */
enum { PT_BRDF_COOK_TORRANCE = 0 };                /* renderer/brdf/mod.rs:50-55 */

/* One material per model (scene/internal/model.rs:10-21, material.rs:10-25).
 * Texture slots index pt_scene_desc.textures; -1 = no texture. */
typedef struct {
  float albedo[3];
  float emissive[3];
  float opacity;
  float metalness;
  float roughness;
  float ior;
  int32_t albedo_tex;    /* 3-channel, sRGB (material.rs:132-146) */
  int32_t emissive_tex;  /* 3-channel (material.rs:189-201)        */
  int32_t opacity_tex;   /* 1-channel (material.rs:207-214)        */
  int32_t metalness_tex; /* 1-channel (material.rs:152-159)        */
  int32_t roughness_tex; /* 1-channel (material.rs:165-172)        */
  int32_t normal_tex;    /* 3-channel (material.rs:178-187)        */
} pt_material;

/* Model i owns material i.  Mesh: triangles [first_tri, first_tri+n_tris) of the
 * tri_* arrays, in file order.  Sphere: center/radius (model.rs:26-64). */
typedef struct {
  int32_t type;
  uint32_t first_tri;
  uint32_t n_tris;
  float center[3];
  float radius;
} pt_model;

typedef struct {
  int32_t type;
  float vec[3]; /* Point: position; Directional: direction (light.rs:6-17) */
  float color[3];
  float size;   /* unused by the reference (renderer/mod.rs:304)          */
} pt_light;

/* 8-bit texture, row-major, channels = 3 (into_rgb8) or 1 (into_luma8)
 * (scene/internal/texture_bank.rs:21-51). */
typedef struct {
  uint32_t width, height, channels;
  const uint8_t *data;
} pt_texture;

typedef struct {
  uint32_t n_models;
  const pt_model *models;
  const pt_material *materials; /* n_models entries */
  uint64_t n_tris;
  const float *tri_pos; /* n_tris*9: v0.xyz v1.xyz v2.xyz            */
  const float *tri_nrm; /* n_tris*9                                   */
  const float *tri_uv;  /* n_tris*6                                   */
  uint32_t n_textures;
  const pt_texture *textures;
  uint32_t n_lights;
  const pt_light *lights;
  float camera[16];     /* column-major transform[c][r] (camera.rs:26) */
  float fov;            /* vertical, radians (camera.rs:12)            */
  float background[3];
} pt_scene_desc;

/* Profile (config/profile.rs:10-40, config/resolution.rs:4-16). */
typedef struct {
  uint32_t width, height; /* default 1920x1080 */
  uint32_t bounces;       /* default 4         */
  uint32_t samples;       /* default 64        */
  int32_t brdf;           /* PT_BRDF_*         */
  int32_t tonemap;        /* PT_TONEMAP_*, default FILMIC */
} pt_profile;

/* Closest hit record.  prim_id is the global primitive id: models in file order, a
 * mesh's triangles contiguous in file order, a sphere takes two ids (entry, exit) —
 * the order in which renderer/utils.rs:11-21 pushes hits before its stable sort.
 * prim_id = -1 on a miss.  t is Hit::get_dist (hit.rs:39-45). */
typedef struct {
  int32_t prim_id;
  float t, u, v;
} pt_hit;

/* One bounce of one path (render_pixel, renderer/mod.rs:172-228), as recorded by pt_trace_paths. */
typedef struct {
  float ray_o[3], ray_d[3]; /* the segment cast at this bounce (mod.rs:182)                     */
  int32_t prim_id;          /* hit accepted by the alpha walk (mod.rs:188-205), -1 = miss        */
  float t;                  /* its Hit::get_dist                                                 */
  float throughput[3];      /* after this bounce's update (mod.rs:271-274, roulette included)    */
  float radiance[3];        /* `color` after this bounce (emissive + shadowed direct light)      */
} pt_path_vertex;

/* Counters of one render (reference-equivalent ray_cast invocations). */
typedef struct {
  uint64_t samples;       /* W*H*spp rendered by this call                    */
  uint64_t extend_rays;   /* one per path segment (renderer/mod.rs:182)       */
  uint64_t shadow_rays;   /* one per light per shaded hit (mod.rs:291,319), skipped-as-zero excluded */
  uint64_t alpha_recasts; /* continuation re-casts: always 0 (the alpha walk is one traversal); kept for layout */
  uint64_t kernel_launches;
  float ms_total;         /* raygen -> accumulation resident, CUDA events     */
  float ms_extend, ms_shade, ms_shadow, ms_other; /* filled only while pt_set_kernel_timing(1) */
  uint64_t extend_launches, shade_launches, shadow_launches;
  uint64_t alpha_tests;    /* opacity evaluations of the alpha walk (renderer/mod.rs:190-204), an implementation count */
  uint64_t shadow_dropped; /* shadow rays that did not fit their queue: always 0, anything else is a bug        */
} pt_stats;

typedef struct pt_scene pt_scene; /* immutable device-resident scene */

/* ---- error reporting ---- */
const char *pt_last_error(void);
/* Number of usable CUDA devices (0 if none); never fails. */
int pt_device_count(void);

/* ---- profile: Profile::load / Default (config/profile.rs:27-40) ---- */
void pt_profile_default(pt_profile *out);
int pt_profile_load_yaml(const char *path, pt_profile *out);

/* ---- scene: load_internal (scene/mod.rs:16-22) + Scene::load (internal/mod.rs:35-51) ---- */
/* Parses scene.isf (JSON), decodes PNG textures relative to its directory, flattens,
 * builds the acceleration structure and uploads everything to `device`. */
int pt_scene_load_isf(const char *path, int device, pt_scene **out);
/* Same, from caller-provided flat arrays (copied; caller keeps ownership). */
int pt_scene_create(const pt_scene_desc *desc, int device, pt_scene **out);
/* Procedural scenes for the large configs: recipe in {"console","foliage","city","soup"},
 * n_tris is a target triangle count. */
int pt_scene_synthetic(const char *recipe, uint64_t n_tris, uint64_t seed, int device, pt_scene **out);
void pt_scene_destroy(pt_scene *scene);
/* Acceleration-structure builder used by the scene constructors above (process-wide; default PT_ACCEL_SAH, or the
 * value of $PT_B200_ACCEL = sah|lbvh|auto).  Replaces KDTree::build of the kdtree-ray crate (call sites
 * scene/internal/model.rs:96, scene/internal/mod.rs:42).  Either tree is a conservative candidate filter: traced
 * results are identical.  SAH: binned-SAH build on the host cores (best traversal speed).  LBVH: Morton-code radix
 * sort + Karras hierarchy on the GPU (fastest build, for large meshes).  PLOC: the same Morton order, but the hierarchy
 * is built bottom-up by merging mutual nearest neighbours (smallest surface area of the union) — SAH-like traversal
 * speed at GPU build speed.  AUTO: PLOC from 4 M primitives up, SAH below.  ($PT_B200_ACCEL = sah|lbvh|ploc|auto.) */
enum { PT_ACCEL_SAH = 0, PT_ACCEL_LBVH = 1, PT_ACCEL_AUTO = 2, PT_ACCEL_PLOC = 3 };
int pt_set_accel_builder(int accel);
/* What the scene was built with: *accel = PT_ACCEL_SAH, PT_ACCEL_LBVH or PT_ACCEL_PLOC, *build_ms = acceleration-structure build
 * time (host wall clock for SAH; device time for LBVH), *max_depth = levels of 4-wide nodes. */
int pt_scene_accel_info(const pt_scene *scene, int *accel, double *build_ms, int *max_depth);
/* Re-sends every device array of the scene (nodes, triangles, shading records, materials, texels, lights) from
 * its pinned host image: the host->device leg of a call that starts from host memory (Scene::load's result living on
 * the host).  Asynchronous and double-buffered: the copies go, on the scene's own copy stream, into a second set of
 * device arrays (allocated on the first call) that no kernel is reading, and the next render / trace call on the
 * scene waits for them and switches to that set - so the upload for frame k+1 overlaps the rendering of frame k.
 * `stream` is unused (kept for source compatibility).  *bytes_out = bytes copied. */
int pt_scene_upload(pt_scene *scene, void *stream, uint64_t *bytes_out);
/* The same refresh for a scene replica whose arrays arrive from somewhere else than this process's host copy - the
 * multi-GPU form of Scene::load's result reaching every device: ONE rank uploads from host memory (pt_scene_upload),
 * the others receive the arrays from that GPU over NVLink (an NCCL broadcast, a peer copy), because eight uploads of
 * the same scene share one host's PCIe / memory bandwidth and NVSwitch does not (path-tracer_b200/multi.py
 * distribute_scene).
 * pt_scene_upload_targets: the arrays of the set no kernel is reading (the set a pt_scene_upload just wrote, if one is
 * pending), as device pointers and sizes in bytes, in the same order on every replica of a scene; at most `capacity`
 * entries are written, *count = how many there are.  `stream` is made to wait until the set may be written or read
 * (the kernels of two frames ago, a pending upload): enqueue the transfer on it.
 * pt_scene_upload_commit: the arrays are complete once the work enqueued on `stream` so far has run; the next render
 * / trace call on the scene waits for it and switches to that set, exactly as after pt_scene_upload. */
int pt_scene_upload_targets(pt_scene *scene, void *stream, void **ptrs, uint64_t *bytes, uint32_t capacity,
                            uint32_t *count);
int pt_scene_upload_commit(pt_scene *scene, void *stream);
/* pt_scene_upload of one part of `parts`: of every array, bytes [part * shard, (part + 1) * shard) with
 * shard = (size / parts) rounded down to a multiple of 256, plus the remainder behind the last shard (sent by every
 * part).  With one part per rank followed by an in-place all-gather of the first parts * shard bytes of each array
 * (pt_scene_upload_targets, then pt_scene_upload_commit) the scene crosses the host's PCIe once per frame, not once per
 * GPU, and NVSwitch replicates it.  *bytes_out = bytes this call copied. */
int pt_scene_upload_part(pt_scene *scene, uint32_t part, uint32_t parts, uint64_t *bytes_out);
/* Introspection (host copies; pointers valid until pt_scene_destroy). */
int pt_scene_get_desc(const pt_scene *scene, pt_scene_desc *out);
int pt_scene_info(const pt_scene *scene, uint64_t *n_tris, uint64_t *n_prims, uint64_t *n_nodes,
                  uint64_t *device_bytes);

/* ---- ray_cast (renderer/utils.rs:11-21): first element of the sorted hit list ---- */
/* The fixed-batch entry points below load the batch into the renderer's own ray queues and trace it with the SAME
 * kernels that render (k_extend / k_shadow): what they return is what the hot path computes.  Batches of any size
 * are processed in chunks of 16 Mi rays.  Calls on one scene are serialised (host mutex + a device event), so they
 * may be issued from several threads / streams. */
/* rays: n*6 host floats (origin xyz, direction xyz); out: n host pt_hit. */
int pt_trace_closest(const pt_scene *scene, const float *rays, size_t n, pt_hit *out);
/* Same on device pointers (rays_dev, out_dev on the scene's device), on `stream`
 * (a cudaStream_t passed as void*; NULL = default stream). */
int pt_trace_closest_device(const pt_scene *scene, const float *rays_dev, size_t n, pt_hit *out_dev,
                            void *stream);
/* pt_trace_closest_device for callers that cannot allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same inputs, same results (bit-identical: the same k_extend kernel), but the per-ray wave scratch lives in a static
 * device array of the library - 48 bytes per ray for PT_TRACE_NOALLOC_MAX_RAYS rays + 256 B = 3 MiB + 256 B held on
 * every GPU that loads the library until exit - so the call makes no cudaMalloc / cudaFree / allocating library call.
 * n > PT_TRACE_NOALLOC_MAX_RAYS fails with PT_ERR_INVALID (pt_last_error names the limit); nothing is launched.
 * All calls on a GPU, whatever the scene, share the one array: calls on the same stream are ordered, calls that
 * overlap on the GPU (different streams, no dependency between them) race and return garbage.  Asynchronous on
 * `stream` like pt_trace_closest_device. */
#define PT_TRACE_NOALLOC_MAX_RAYS 65536
int pt_trace_closest_device_noalloc(const pt_scene *scene, const float *rays_dev, size_t n, pt_hit *out_dev,
                                    void *stream);
/* The hit render_pixel shades after its alpha walk (renderer/mod.rs:188-205): the first hit of the sorted list with
 * opacity >= 1, or with opacity > 0.001 and a uniform draw below it; the farthest hit when none is accepted.  Ray i
 * uses the draws of path (pixel = i, sample, seed) at `bounce` (the Philox addressing of DESIGN.md).  n < 2^31. */
int pt_trace_accepted(const pt_scene *scene, const float *rays, size_t n, uint32_t bounce, uint32_t sample,
                      uint64_t seed, pt_hit *out);
/* Shadow walk of get_light_info (renderer/mod.rs:281-333): out[i] = product over the hits of ray i of (1 - opacity).
 * limits == primary == NULL: directional light (mod.rs:291-297: every hit counts, opacity sampled at the occluder's
 * own uv).  Otherwise a point light (mod.rs:301-330): limits[i] = |hit - light|, primary = n*5 floats (position xyz of
 * the primary hit, its texture coordinates u v; u = NaN when the primary hit is a sphere): occluders farther than
 * limits[i] from the primary hit are ignored and the occluder's opacity is sampled with the PRIMARY hit's
 * uv/variant (mod.rs:324). */
int pt_trace_shadow(const pt_scene *scene, const float *rays, const float *limits, const float *primary, size_t n,
                    float *out);
/* pt_trace_shadow(scene, rays, NULL, NULL, n, out). */
int pt_trace_transmittance(const pt_scene *scene, const float *rays, size_t n, float *out);

/* ---- Triangle::intersect (scene/internal/triangle.rs:37-82) on n independent (ray, triangle) pairs ---- */
/* rays: n*6 host floats; tris: n*9 host floats (v0 v1 v2); out[i].prim_id = -1 miss, 0 front-face hit,
 * 1 back-face hit (Hit::Triangle.is_backface); t,u,v as tested by triangle.rs:200-233. */
int pt_triangle_intersect(const pt_scene *scene_or_null, int device, const float *rays, const float *tris, size_t n,
                          pt_hit *out);

/* ---- the random source: StdRng::seed_from_u64 + gen::<f32>() of renderer/mod.rs:110-118 is replaced by the
 * counter-based Philox4x32-10 (north star); this evaluates the device generator on n caller-chosen (counter[4],
 * key[2]) pairs so that it can be checked against the published Random123 known-answer vectors.  Host pointers. */
int pt_philox4x32_10(int device, const uint32_t *counters, const uint32_t *keys, size_t n, uint32_t *out);

/* ---- render_pixel (renderer/mod.rs:172-228), one recorded path per requested pixel ---- */
/* Traces sample `sample` of each pixel in pixels[0..n_paths) with the profile's resolution/bounces and records
 * every bounce: out holds n_paths*(bounces+1) vertices (path i at out[i*(bounces+1)]), counts[i] = vertices
 * written.  Host pointers.  Lights are evaluated in file order. */
int pt_trace_paths(const pt_scene *scene, const pt_profile *profile, const uint32_t *pixels, size_t n_paths,
                   uint32_t sample, uint64_t seed, pt_path_vertex *out, uint32_t *counts);

/* Per-kernel CUDA-event timing inside pt_render_accum_device (one event per launch; replaces the reference's only
 * timer, the whole-second "Done: {}s" of renderer/mod.rs:103,164-166).  Off by default. */
void pt_set_kernel_timing(int on);

/* ---- Renderer::render (renderer/mod.rs:76-169) ---- */
/* Renders samples {sample_begin + k*sample_stride | k in [0,n_samples)} of every pixel
 * and ADDS their radiance into accum_dev (device, W*H*3 floats, row-major i=x+y*W,
 * caller-zeroed).  Multi-GPU drivers give each rank a disjoint sample subset and then
 * sum the buffers (one NCCL reduce).  Asynchronous on `stream` until stats are read:
 * if stats_out != NULL the call synchronises and fills it. */
int pt_render_accum_device(const pt_scene *scene, const pt_profile *profile, uint32_t sample_begin,
                           uint32_t sample_stride, uint32_t n_samples, uint64_t seed, float *accum_dev,
                           void *stream, pt_stats *stats_out);
/* post_processing (renderer/mod.rs:150-162,335-353): mean = accum/total_samples,
 * tonemap, gamma 1/2.2, saturating u8.  accum_dev: W*H*3 floats, rgb8_dev: W*H*3 bytes. */
int pt_tonemap_device(const float *accum_dev, uint32_t width, uint32_t height, uint32_t total_samples,
                      int32_t tonemap, uint8_t *rgb8_dev, int device, void *stream);
/* Whole call a user makes: render all profile->samples on the scene's device, tonemap,
 * copy back.  rgb8_out: W*H*3 host bytes (row-major RGB); accum_out optional W*H*3
 * host floats (un-normalised sums); stats_out optional. */
int pt_render(const pt_scene *scene, const pt_profile *profile, uint64_t seed, uint8_t *rgb8_out,
              float *accum_out, pt_stats *stats_out);

/* pt_render without the wait: the render, the tonemap and the copy of the image into rgb8_out (host memory; pinned
 * memory if the copy is to be asynchronous) are enqueued on `stream` and the call returns.  rgb8_out is valid after
 * pt_stream_synchronize(device, stream) (or the caller's own cudaStreamSynchronize).  Frames issued back to back on
 * one stream run in order, and a pt_scene_upload issued between them overlaps the frame before it: the pipelined form
 * of a host loop that streams scene updates (bench.py's end-to-end leg at one GPU). */
int pt_render_async(const pt_scene *scene, const pt_profile *profile, uint64_t seed, uint8_t *rgb8_out, void *stream);
int pt_stream_synchronize(int device, void *stream);

/* ---- Renderer::render over several GPUs of one box, in one process ---- */
/* A copy of `scene` resident on `device` (-1: host-only copy).  Nothing is parsed again and the acceleration
 * structure of the source (built or loaded with it) is reused, so a replica costs one upload.  The reference shares
 * one `&Scene` between its rayon workers (renderer/mod.rs:76,106); a GPU needs the scene in its own HBM. */
int pt_scene_replicate(const pt_scene *scene, int device, pt_scene **out);
/* pt_render split by sample index: replica g of n_replicas renders the samples {s : s mod n_replicas == g} of the
 * full frame on its own device from its own host thread (what rayon's pool is to mod.rs:105-106); the partial sums
 * are copied device-to-device onto replica 0's device (NVLink when the devices are peers) and added there in replica
 * order, then post-processed once.  The set of samples, hence the image up to f32 summation order, is the same for
 * every n_replicas.  replicas: distinct scene objects (two may share a device); rgb8_out / accum_out as pt_render;
 * stats_out: optional array of n_replicas entries (per-replica counts and device time).  The one-process-per-GPU
 * form of the same split (torchrun + one NCCL reduce) is built from pt_render_accum_device. */
int pt_render_multi(const pt_scene *const *replicas, uint32_t n_replicas, const pt_profile *profile, uint64_t seed,
                    uint8_t *rgb8_out, float *accum_out, pt_stats *stats_out);

/* ---- debug_render (renderer/debug_renderer.rs:11-105): `render --debug-textures` ---- */
/* One pixel-centre ray per pixel, first element of ray_cast, material sampled at the hit.  planes_out holds
 * PT_AOV_COUNT consecutive row-major RGB8 images of width*height pixels (host memory), each channel
 * `(value * 255.) as u8`; pixels whose ray hits nothing are 0.  *pixels_hit = rays that hit (the reference
 * writes no PNG at all when it is 0).  ior is stored as ior/3 (debug_renderer.rs:102). */
enum { PT_AOV_NORMAL = 0, PT_AOV_ALBEDO, PT_AOV_OPACITY, PT_AOV_METALNESS, PT_AOV_ROUGHNESS, PT_AOV_EMISSIVE,
       PT_AOV_IOR, PT_AOV_COUNT };
const char *pt_aov_name(int aov); /* "normal", "albedo", ...: the PNG base names of debug_renderer.rs:59 */
int pt_debug_textures(const pt_scene *scene, uint32_t width, uint32_t height, uint8_t *planes_out,
                      uint64_t *pixels_hit);

/* ---- progressive output: the viewer feed of Renderer::render (renderer/mod.rs:133-141, viewer.rs) ---- */
/* The SFML window itself is unavailable (no graphics stack); its content is delivered headless instead: after
 * every `snapshot_every` samples per pixel the running mean (`*pixel / current_sample`, mod.rs:137) is
 * post-processed and handed to on_snapshot(user, rgb8, width, height, samples_done) (rgb8 = rgb8_out, valid during
 * the call).  A non-zero return stops further snapshots (the window was closed, mod.rs:139-141); rendering
 * continues.  The final image is identical to pt_render's.  snapshot_every = 0 or on_snapshot = NULL: no snapshots. */
typedef int (*pt_snapshot_fn)(void *user, const uint8_t *rgb8, uint32_t width, uint32_t height,
                              uint32_t samples_done);
int pt_render_progressive(const pt_scene *scene, const pt_profile *profile, uint64_t seed, uint32_t snapshot_every,
                          pt_snapshot_fn on_snapshot, void *user, uint8_t *rgb8_out, float *accum_out,
                          pt_stats *stats_out);

/* ---- convert_gltf_to_isf (scene/gltf.rs:146-204): `path-tracer convert <input> <output_dir>` ---- */
/* Reads a glTF 2.0 file (.gltf with external / base64 buffers, or .glb), writes <output_dir>/scene.isf and the
 * texture PNGs next to it (albedo_tex_N / alpha_tex_N / gray_tex_N / vec_tex_N, gltf.rs:47-77).  Host code only (no
 * GPU).  Images: PNG and baseline JPEG (progressive JPEG is rejected). */
int pt_convert_gltf(const char *input, const char *output_dir);

/* ---- binary side-car of an isf scene ("isfb"; no reference counterpart: scene/isf.rs is JSON only) ---- */
/* Writes the flat scene of `scene` (decoded textures included) as raw little-endian arrays; with_accel != 0 also
 * stores the scene's 4-wide BVH (a host-only scene, device -1, gets the host SAH tree built for the purpose).
 * pt_scene_load_isf recognises such a file by its magic, checks a stored tree against the scene (every reference in
 * range, leaves partition the triangles, child boxes contain their contents) and skips the build.  Layout:
 * path-tracer_b200/csrc/isf_binary.cpp. */
int pt_scene_save_binary(const pt_scene *scene, const char *path, int with_accel);

/* ---- image output: RgbImage::save (main.rs:50) ---- */
int pt_write_png(const char *path, const uint8_t *rgb8, uint32_t width, uint32_t height);
/* Texture decode used by the texture bank (texture_bank.rs:33,49): PNG or baseline JPEG, chosen by the file's
 * signature as image::open does; channels_out in {1,3}.
 * *data_out is malloc'ed; free with pt_free. */
int pt_read_png(const char *path, uint32_t channels_out, uint32_t *width, uint32_t *height,
                uint8_t **data_out);
void pt_free(void *p);

#ifdef __cplusplus
}
#endif
#endif /* PT_B200_H */
