// scene_host.hpp — host-side flat scene, the replacement of scene::internal::Scene
// (reference scene/internal/mod.rs:26-51).  Owns copies of everything in a pt_scene_desc.
#pragma once
#include <cmath>
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/pt_b200.h"

namespace ptb {

struct HostScene {
  std::vector<pt_model> models;
  std::vector<pt_material> materials;
  std::vector<float> pos, nrm, uv;  // 9 / 9 / 6 floats per triangle, file order
  std::vector<pt_texture> textures;  // .data points into texdata
  std::vector<std::vector<uint8_t>> texdata;
  std::vector<pt_light> lights;
  float camera[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
  float fov = 0.4f;
  float background[3] = {0, 0, 0};

  // derived by finalize()
  std::vector<int32_t> prim_base;   // first global prim id of model m (sphere: 2 ids)
  std::vector<int32_t> tri_model;   // model of triangle i
  std::vector<int32_t> sphere_models;  // models that are spheres, in file order
  int64_t n_prims = 0;

  uint64_t n_tris() const { return pos.size() / 9; }
  // Validates ranges and fills the derived tables.  Returns "" or an error message.
  std::string finalize();
  void from_desc(const pt_scene_desc &d);
  void to_desc(pt_scene_desc &d) const;
  int add_texture(uint32_t w, uint32_t h, uint32_t ch, std::vector<uint8_t> &&data);
};

// ---- acceleration structure: 4-wide BVH, 128-byte cache-line-aligned nodes ----
// Child slot k of a node: box (lo_x[k],lo_y[k],lo_z[k])-(hi_x[k],hi_y[k],hi_z[k]) and child[k]:
//   0xFFFFFFFF            empty slot
//   bit31=0               inner node index
//   bit31=1, bit30=0      triangle leaf: bits[29:3] = first slot in the leaf-ordered triangle array,
//                         bits[2:0] = count-1
//   bit31=1, bit30=1      sphere leaf: bits[29:0] = index into the sphere array
struct alignas(128) Bvh4Node {
  float lo_x[4], hi_x[4], lo_y[4], hi_y[4], lo_z[4], hi_z[4];  // one 32-byte chunk per axis (256-bit loads)
  uint32_t child[4];
  uint32_t pad[4];
};
static_assert(sizeof(Bvh4Node) == 128, "node must be one cache line");

constexpr uint32_t BVH_EMPTY = 0xFFFFFFFFu;
constexpr uint32_t BVH_LEAF = 0x80000000u;
constexpr uint32_t BVH_SPHERE = 0x40000000u;

struct Bvh4 {
  std::vector<Bvh4Node> nodes;      // nodes[0] = root
  std::vector<uint32_t> leaf_tris;  // leaf order -> original triangle index
  uint32_t root_ref = BVH_EMPTY;    // a leaf ref when the scene has a single leaf
  int max_depth = 0;
};

// ---- box margins shared by both builders (host SAH and device LBVH) --------------------------------------------
// The contract of the traversal is brute-force semantics: a box cull must never drop a primitive that the reference's
// f32 arithmetic (triangle.rs:37-82, model.rs:26-64) would accept.  That arithmetic accepts rays that miss the exact
// primitive by a rounding-sized margin which grows with the distance |o - v0| to the ray origin, so the margins are
// scaled by the scene, not by the local coordinate:
//   D(p) = max(|p - farthest corner of the scene box|, |p - camera|)   bounds |o - p| for every origin the renderer
//                                                      uses (the camera or a point on a surface), p = v0 / centre
//   S = largest |coordinate| of the scene box and the camera
#if defined(__CUDACC__)
#define PTB_HD __host__ __device__
#else
#define PTB_HD
#endif
PTB_HD inline float origin_reach(const float p[3], const float lo[3], const float hi[3], const float cam[3]) {
  float f2 = 0.f, c2 = 0.f;
  for (int k = 0; k < 3; ++k) {
    const float f = fmaxf(fabsf(p[k] - lo[k]), fabsf(p[k] - hi[k])), c = p[k] - cam[k];
    f2 += f * f;
    c2 += c * c;
  }
  const float D = sqrtf(fmaxf(f2, c2));
  return D == D && D < 1e30f ? D : 1.0f;
}
PTB_HD inline float coordinate_scale(const float lo[3], const float hi[3], const float cam[3]) {
  float S = 0.f;
  for (int k = 0; k < 3; ++k) S = fmaxf(S, fmaxf(fmaxf(fabsf(lo[k]), fabsf(hi[k])), fabsf(cam[k])));
  return S < 1e30f ? S : 1.0f;
}
// Moller-Trumbore's barycentrics carry an error of a few ulp of |o - v0| |d| |e| / det; in space that lets a ray pass
// a few ulp of D outside the edges through v0, and (|e1| + |e2|) / |e2 - e1| times as much outside the edge opposite
// v0 (a thin triangle with its sharp corner at v0 leaks most).  p = the 9 position floats of the triangle.
PTB_HD inline float triangle_leak_pad(const float *p, float D) {
  float l1 = 0.f, l2 = 0.f, l3 = 0.f, e1[3], e2[3];
  for (int k = 0; k < 3; ++k) {
    e1[k] = p[3 + k] - p[k];
    e2[k] = p[6 + k] - p[k];
    const float c = e2[k] - e1[k];
    l1 += e1[k] * e1[k];
    l2 += e2[k] * e2[k];
    l3 += c * c;
  }
  // |det| = |e1 . (d x e2)| <= |d| |e1 x e2| and triangle.rs:49 rejects |det| < 1e-6: a triangle whose doubled area
  // is below 1e-6 / 8 cannot be hit by any ray with |d| <= 8 (the renderer's are unit length), so it needs no
  // leak margin however thin it is (sphere poles and other collapsed triangles would otherwise get the largest one)
  const float cx = e1[1] * e2[2] - e1[2] * e2[1], cy = e1[2] * e2[0] - e1[0] * e2[2], cz = e1[0] * e2[1] - e1[1] * e2[0];
  float ratio = 1.0f;
  if (sqrtf(cx * cx + cy * cy + cz * cz) * 8.0f >= 1e-6f) {
    ratio = (sqrtf(l1) + sqrtf(l2)) / fmaxf(sqrtf(l3), 1e-30f);
    ratio = fminf(fmaxf(ratio, 1.0f), 4096.0f);
  }
  return 2.4e-7f * D * ratio;
}
// The sphere test's discriminant carries an error of a few ulp of |o - c|^2: a tangent ray at distance h is accepted
// while h^2 <= r^2 + that error.
PTB_HD inline float sphere_box_radius(float radius, float D) {
  const float r = fabsf(radius);
  return sqrtf(r * r + 1.2e-7f * D * D) * 1.0001f + 1e-6f;
}
// Slab planes of a node box: the fused slab test (isect.cuh slab_fma) evaluates fma(plane, inv, -(o * inv)).  The
// roundings that scale with t are absorbed by the widened exit (tf * 1.000002) and by 4e-6 |plane|; the rounding of
// the product o * inv is ~1 ulp of |o| in space whatever t is (a wall at x = 0.01 seen from x = 50), hence 8 ulp of S.
PTB_HD inline float plane_pad(float lo, float hi, float S) {
  return fmaxf(fabsf(lo), fabsf(hi)) * 4e-6f + S * 4.8e-7f + 1e-7f;
}

// A/B knob: $PT_B200_MARGIN_SCALE multiplies D and S (0 = the coordinate-relative rules of revision 1).
float bvh_margin_scale();

// ---- the node the traversal kernels read: 64 bytes, child boxes quantised to 8 bits per plane -------------------
// Plane q of an axis decodes to origin + q * step (q in 0..255); `scale` stores step * 2^15 because the kernels build
// the float 1 + q * 2^-15 from the byte with one byte permute and evaluate the slab parameter of the plane as
// fma(1 + q 2^-15, scale * inv, (origin - o) * inv - scale * inv) (isect.cuh).  q_lo is rounded down and q_hi up with
// a guard of 0.05 step (the rounding of that expression is 2^-8 step), so the decoded box always contains the float
// box the builders produced (margins above) - quantisation only ever makes the filter less tight.
// Words q[0..5] = lo_x, hi_x, lo_y, hi_y, lo_z, hi_z; byte k of a word = child slot k.  An empty slot has
// lo = 255 > hi = 0 on every axis and child = `empty_ref`: a leaf holding one degenerate triangle that nothing can
// hit (appended to the leaf-ordered triangle array), so a ray that passes an inverted box through rounding (flat
// nodes, extreme directions) costs one rejected primitive test instead of a bad reference.
struct alignas(64) Bvh4QNode {
  float origin[3];
  float scale[3];
  uint32_t q[6];
  uint32_t child[4];
};
static_assert(sizeof(Bvh4QNode) == 64, "quantised node is half a cache line");
void quantise_bvh4(const Bvh4 &bvh, uint32_t empty_ref, int threads, std::vector<Bvh4QNode> &out);

// Binned-SAH BVH2 build over triangles + spheres, collapsed to 4-wide.  Boxes carry the margins above
// so a box cull can never drop a primitive Möller–Trumbore would accept.
void build_bvh4(const HostScene &s, Bvh4 &out, int threads);
// Largest number of triangles either builder puts into one leaf (default 2; $PT_B200_LEAF_TRIS in 1..8 for A/B runs).
uint32_t bvh_max_leaf_tris();

// GPU LBVH build (lbvh.cu) on the current CUDA device: Morton codes + radix sort + Karras hierarchy + bottom-up fit,
// collapsed to the same 4-wide nodes.  *build_ms = device time of the build (CUDA events, upload excluded).
// Returns PT_ERR_INVALID with *too_deep = 1 when the tree exceeds the traversal stack (caller falls back to SAH).
// method 0: Karras radix tree over the Morton order (LBVH).  method 1: PLOC — mutual-nearest-neighbour agglomeration
// in Morton order, minimising the surface area of every merge: SAH-like trees at GPU build speed.
int build_bvh4_lbvh(const HostScene &s, Bvh4 &out, float *build_ms, int *too_deep, int method = 0);

// Revision of the box-inflation / node-layout rules both builders follow; an isfb accel section written under another
// revision is ignored on load (isf_binary.cpp).
constexpr uint32_t BVH_ACCEL_REV = 2;
// isf_binary.cpp: the binary side-car of an isf scene, optionally with its BVH.
bool is_isfb_file(const char *path);
int save_isfb_file(const char *path, const HostScene &s, const Bvh4 *accel, int accel_kind);
// *accel_kind = -1 when the file carries no usable accel section; `accel` may be null (scene only).
int load_isfb_file(const char *path, HostScene &out, Bvh4 *accel, int *accel_kind);
// "" when `b` is a complete, in-range, box-consistent BVH of `s` (s.finalize() must have run).
std::string validate_bvh4(const HostScene &s, const Bvh4 &b);

}  // namespace ptb
