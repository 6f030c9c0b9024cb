// device_scene.cuh — HBM layout of the immutable scene and of the wavefront path state.
// All kernels are compiled with --fmad=false: the intersection and shading arithmetic must
// round exactly like the reference's un-fused f32 Rust code (cgmath operand order).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ptb {

// 48-byte intersection record, leaf order: three 128-bit loads per triangle test.
//   a = (v0.xyz, global prim id)   b = (e1.xyz, original triangle index)   c = (e2.xyz, flags)
// e1 = v1-v0 and e2 = v2-v0 are the f32 differences of triangle.rs:43-44, computed once.
struct DevTri {
  float4 a, b, c;
};
constexpr uint32_t TRI_FLAG_ALPHA = 1u;    // material can yield opacity < 1 (mod.rs:198-204)
constexpr uint32_t TRI_FLAG_OPAQUE1 = 2u;  // opacity is exactly 1 with no opacity texture: a shadow walk that meets
                                           // this primitive multiplies by (1 - 1) = 0 without reading the material

// 64-byte shading record, original triangle order (touched only for examined hits).
struct DevTriShade {
  float n0[3], n1[3], n2[3];
  float uv0[2], uv1[2], uv2[2];
  int32_t model;
};

struct DevSphere {
  float cx, cy, cz, radius;
  int32_t model;
  int32_t prim_id;  // entry id; exit = prim_id + 1
  uint32_t flags;
  uint32_t pad;
};

struct DevTexture {
  unsigned long long offset;  // into the texel pool
  uint32_t width, height, channels, pad;
};

struct DevMaterial {
  float albedo[3], emissive[3];
  float opacity, metalness, roughness, ior;
  int32_t albedo_tex, emissive_tex, opacity_tex, metalness_tex, roughness_tex, normal_tex;
};

struct DevLight {
  int32_t type;
  float vec[3];
  float color[3];
  float pad;
};

struct DevNode {  // == ptb::Bvh4Node, one 128-byte line = four 32-byte chunks: x planes, y planes, z planes, children
  float4 lo_x, hi_x, lo_y, hi_y, lo_z, hi_z;
  uint4 child;
  uint4 pad;
};

// PTB_QNODE 1: the traversal kernels read the 64-byte quantised node (== ptb::Bvh4QNode, scene_host.hpp) - two
// 256-bit loads per node visit instead of four; 0: the 128-byte float node.
// Measured (profiles/r07/ab_s19_quantised_nodes.jsonl, all 104 GPU parity tests green on the quantised nodes):
// k_extend +22 % on console, +10 % on foliage, +30 % on city; k_shadow +7 % / +5 % / +13 %.  Halving the node's load
// instructions does not pay for the 24 byte permutes + 9 per-node set-up operations of the decode, three more spilled
// registers and the looser boxes (steps of 1/250 of the parent's extent).  Off; the float node stays.
#ifndef PTB_QNODE
#define PTB_QNODE 0
#endif
struct DevQNode {
  float ox, oy, oz, sx, sy, sz;             // origin, step * 2^15 per axis
  uint32_t qlx, qhx, qly, qhy, qlz, qhz;    // byte k = child slot k
  uint4 child;
};
static_assert(sizeof(DevQNode) == 64, "quantised node layout");

// PTB_CNODE 1: the float node stores centre / half-extent per child and axis instead of lo / hi (same 128-byte layout:
// the "lo" chunks hold the centres, the "hi" chunks the half-extents): the slab parameters of a child are
// tm -+ |h * inv| with tm = fma(c, inv, -o * inv) - multiply-adds and adds on the FMA pipe in place of the six
// min / max of the lo / hi form, which run on the half-rate ALU pipe.
// Measured (profiles/r07/ab_s20_centre_half_extent_shadow_stack.jsonl): k_extend -0.7 % on console, -0.5 % on foliage,
// +0.4 % on city; k_shadow +1.5 ... +3 %.  The ALU pipe is the busiest unit (60 % of its peak, ncu) but not the limit:
// the kernels wait for the slowest lane's node line (L2 latency).  Off; the lo / hi node stays.
#ifndef PTB_CNODE
#define PTB_CNODE 0
#endif

struct DevScene {
  const DevNode *nodes;    // float nodes, lo / hi planes (PTB_QNODE 0 && PTB_CNODE 0 builds)
  const DevNode *cnodes;   // float nodes, centre / half-extent (PTB_CNODE 1)
  const DevQNode *qnodes;  // quantised nodes
  const DevTri *tris;
  const DevTriShade *shade;
  const DevSphere *spheres;
  const DevMaterial *materials;
  const DevTexture *textures;
  const uint8_t *texels;
  const DevLight *lights;
  const float *tri_pos9;  // original-order positions; only uploaded when a normal map exists
  const float *srgb_lut;  // [0,256): powf(i / 255, 2.2), the sRGB -> linear step of material.rs:132-146; [256,512): i / 255
  uint32_t n_tris, n_spheres, n_lights, n_nodes;
  uint32_t root_is_valid;
  uint32_t tree_depth;  // levels of 4-wide nodes (host-side launch heuristics only)
  uint32_t any_alpha;  // some material may be non-opaque
  uint32_t f32_one_bits;  // 0x3F800000, kept in the parameter bank so that the plane decode (isect.cuh qplane) reads it as a constant-bank operand and its byte selector stays an immediate
  float camera[16];
  float fov;
  float tan_half_fov;  // (fov / 2).tan() of mod.rs:117,121, evaluated once on the host
  float background[3];
};

// Internal hit record.  idx: -1 miss; else bits[29:0] = triangle index, or n_tris + 2*sphere + exit;
// bit 30 = triangle back face (det < 0).  h = (dist, u, v, global prim id bits).
struct HitRec {
  float4 h;
  int32_t idx;
};
constexpr int32_t HIT_BACKFACE = 0x40000000;
constexpr int32_t HIT_INDEX_MASK = 0x3FFFFFFF;

// One wave of paths, SoA.  Queue position q (compacted every bounce) owns ray/throughput;
// `slot` = pixel + k*W*H (k-th sample of the wave) owns the radiance sum and the RNG key.
struct WaveState {
  float4 *ray_o[2];   // xyz origin
  float4 *ray_d[2];   // xyz direction, w = slot (int bits)
  float4 *thr[2];     // xyz throughput
  float4 *hit;        // (dist, u, v, HitRec.idx bits; idx < 0 = miss)
  uint32_t *hit_list; // queue positions that hit something, in queue order (k_hit_list -> k_shade)
  float4 *radiance;   // per slot: xyz sum of this path's contributions (zeroed by raygen, atomically added to)
  // shadow queue (one entry per light per shaded hit with a non-zero unshadowed contribution)
  float4 *sh_o;       // xyz origin, w = point-light distance (or -1 for directional)
  float4 *sh_d;       // xyz direction, w = slot
  float4 *sh_c;       // rgb unshadowed contribution, w = primary uv.x (NaN: primary hit is a sphere)
  float4 *sh_p;       // xyz primary hit position, w = primary uv.y
  uint32_t *counters; // see CTR_* below
  uint32_t capacity;       // paths per wave
  uint32_t shadow_capacity;
};

// counters[] layout (uint32): queue sizes are produced on the device and consumed by the next persistent kernel, so
// the host never synchronises inside a frame.  Path queue buffer b and the shadow queue filled together with it
// share one aligned 64-bit word (low = path queue size, high = shadow queue size) so that k_shade reserves its slots
// in both with ONE atomic.
enum {
  CTR_QUEUE0 = 0,   // size of path queue buffer 0        } one 64-bit word
  CTR_SHADOW0 = 1,  // shadow rays emitted with buffer 0  }
  CTR_QUEUE1 = 2,   // size of path queue buffer 1        } one 64-bit word
  CTR_SHADOW1 = 3,  // shadow rays emitted with buffer 1  }
  CTR_WORK_EXTEND = 4,  // persistent work-fetch cursors
  CTR_WORK_SHADOW = 5,
  CTR_HITS = 6,         // length of the ordered hit list (k_hit_list), reset by k_extend
  CTR_DROPPED = 7,      // shadow-queue entries that did not fit the queue (must stay 0; reported through pt_stats)
  CTR_STAT_EXTEND = 8,  // 64-bit statistics start here (pairs)
  CTR_STAT_SHADOW = 10,
  CTR_STAT_ALPHA = 12,  // opacity evaluations of the alpha walk
  // -DPTB_COUNT_STEPS builds only (tools/build_variant.sh steps -DPTB_COUNT_STEPS=1): node visits / primitive tests
  CTR_STAT_EXT_NODES = 16,
  CTR_STAT_EXT_PRIMS = 18,
  CTR_STAT_SH_NODES = 20,
  CTR_STAT_SH_PRIMS = 22,
  CTR_COUNT = 24
};
#define CTR_QUEUE(b) (2u * (b))
#define CTR_SHADOW(b) (2u * (b) + 1u)

struct WaveParams {
  uint32_t width, height;
  uint32_t bounces;
  uint32_t spp_in_wave;    // samples of every pixel in this wave
  uint32_t sample_first;   // sample index of wave-sample 0
  uint32_t sample_stride;  // distance between consecutive wave samples
  uint32_t seed_lo, seed_hi;
  uint32_t n_paths;        // width*height*spp_in_wave
};

}  // namespace ptb
