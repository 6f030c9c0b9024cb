// scene_host.hpp — host-side flat scene, the replacement of scene::internal::Scene
// (reference scene/internal/mod.rs:26-51).  Owns copies of everything in a pt_scene_desc.
#pragma once
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

// Binned-SAH BVH2 build over triangles + spheres, collapsed to 4-wide.  Boxes are inflated
// conservatively so a box cull can never drop a primitive Möller–Trumbore would accept.
void build_bvh4(const HostScene &s, Bvh4 &out, int threads);
// Largest number of triangles either builder puts into one leaf (default 2; $PT_B200_LEAF_TRIS in 1..8 for A/B runs).
uint32_t bvh_max_leaf_tris();

// GPU LBVH build (lbvh.cu) on the current CUDA device: Morton codes + radix sort + Karras hierarchy + bottom-up fit,
// collapsed to the same 4-wide nodes.  *build_ms = device time of the build (CUDA events, upload excluded).
// Returns PT_ERR_INVALID with *too_deep = 1 when the tree exceeds the traversal stack (caller falls back to SAH).
int build_bvh4_lbvh(const HostScene &s, Bvh4 &out, float *build_ms, int *too_deep);

}  // namespace ptb
