// oracle/canonical.hpp — TEST / MEASUREMENT INFRASTRUCTURE (CPU oracle), not product code.
// The canonical work counter of SURVEY.md §8(d): a deterministic binned-SAH BVH2 (16 bins, <= 4 primitives per
// leaf) traversed front-to-back.  It freezes the per-ray means n_node (inner nodes visited) and n_tri (primitives
// tested) that define the ALGORITHMIC bytes B_ray = 64*n_node + 48*n_tri + 48 of a ray, independently of
// whatever tree the CUDA kernels use (a worse product tree cannot inflate the roofline fraction).
//   closest: ordered traversal with t_max shrinking to the closest hit   (path segments, utils.rs:11-21)
//   any    : traversal that stops at the first opaque occluder closer than t_max (shadow rays, mod.rs:281-333;
//            the reference walks all hits, an ideal implementation stops at the first opaque one)
#pragma once
#include <cfloat>

#include "scene.hpp"

namespace orc {

struct CanonCounts {
  uint64_t rays_closest = 0, nodes_closest = 0, prims_closest = 0;
  uint64_t rays_any = 0, nodes_any = 0, prims_any = 0;
};

struct CanonBvh {
  struct Node {
    float lo[3], hi[3];
    int32_t left, right;  // inner: child indices; leaf: left = ~first, right = count
  };
  std::vector<Node> nodes;
  std::vector<int32_t> prims;  // >= 0 triangle index, < 0: ~model (sphere)
  void build(const Scene &s);
  // Counts one closest-hit query into c.
  void count_closest(const Scene &s, V3 o, V3 d, CanonCounts &c) const;
  // Counts one any-hit query limited to parameter t_max (FLT_MAX = unlimited).
  void count_any(const Scene &s, V3 o, V3 d, float t_max, CanonCounts &c) const;
};

}  // namespace orc
