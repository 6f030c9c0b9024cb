// bvh_build.cpp — host acceleration-structure build.  Replaces the reference's two-level
// third-party KD-tree (KDTree::build call sites: scene/internal/model.rs:96, internal/mod.rs:42)
// with ONE flattened 4-wide BVH over all triangles and spheres of the scene: binned-SAH BVH2
// (16 bins, <=4 triangles per leaf, spheres alone in a leaf) collapsed to 128-byte 4-wide nodes.
// Nothing here is taken from the reference: its tree lives in the kdtree-ray crate.
#include <algorithm>
#include <atomic>
#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <thread>

#include "scene_host.hpp"

namespace ptb {
namespace {

struct Box {
  float lo[3], hi[3];
  void reset() {
    for (int k = 0; k < 3; ++k) { lo[k] = FLT_MAX; hi[k] = -FLT_MAX; }
  }
  void grow(const Box &b) {
    for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], b.lo[k]); hi[k] = std::max(hi[k], b.hi[k]); }
  }
  void grow_pt(const float *p) {
    for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], p[k]); hi[k] = std::max(hi[k], p[k]); }
  }
  float half_area() const {
    float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    if (dx < 0 || dy < 0 || dz < 0) return 0.f;
    return dx * dy + dy * dz + dz * dx;
  }
};

struct Node2 {
  Box box;
  uint32_t left, right;  // inner: child node indices; leaf: left = first, right = count | LEAF2
};
constexpr uint32_t LEAF2 = 0x80000000u;
constexpr int NBINS = 16;

struct Builder {
  const HostScene &s;
  uint32_t nt;
  std::vector<Box> pbox;       // per primitive (triangles then spheres)
  std::vector<uint32_t> order;  // permutation being partitioned
  std::vector<Node2> nodes;
  std::atomic<uint32_t> next{0};
  std::atomic<int> live_threads{1};
  int max_threads;
  uint32_t max_leaf = bvh_max_leaf_tris();

  Builder(const HostScene &sc, int threads) : s(sc), nt((uint32_t)sc.n_tris()), max_threads(std::max(1, threads)) {}

  bool is_sphere(uint32_t prim) const { return prim >= nt; }

  void centroid(uint32_t prim, float c[3]) const {
    const Box &b = pbox[prim];
    for (int k = 0; k < 3; ++k) c[k] = 0.5f * (b.lo[k] + b.hi[k]);
  }

  uint32_t build(uint32_t first, uint32_t count, int depth) {
    uint32_t idx = next.fetch_add(1);
    Box box, cbox;
    box.reset();
    cbox.reset();
    bool has_sphere = false;
    for (uint32_t i = first; i < first + count; ++i) {
      uint32_t p = order[i];
      box.grow(pbox[p]);
      float c[3];
      centroid(p, c);
      cbox.grow_pt(c);
      has_sphere |= is_sphere(p);
    }
    Node2 n;
    n.box = box;
    if (count == 1 || (count <= max_leaf && !has_sphere)) {
      n.left = first;
      n.right = count | LEAF2;
      nodes[idx] = n;
      return idx;
    }
    // binned SAH over the three axes
    int best_axis = -1, best_split = -1;
    float best_cost = FLT_MAX;
    for (int axis = 0; axis < 3; ++axis) {
      float lo = cbox.lo[axis], ext = cbox.hi[axis] - cbox.lo[axis];
      if (!(ext > 0.f)) continue;
      float scale = (float)NBINS / ext;
      Box bb[NBINS];
      uint32_t bc[NBINS] = {0};
      for (int b = 0; b < NBINS; ++b) bb[b].reset();
      for (uint32_t i = first; i < first + count; ++i) {
        uint32_t p = order[i];
        float c = 0.5f * (pbox[p].lo[axis] + pbox[p].hi[axis]);
        int b = std::min(NBINS - 1, std::max(0, (int)((c - lo) * scale)));
        bb[b].grow(pbox[p]);
        bc[b]++;
      }
      float la[NBINS], ra[NBINS];
      uint32_t lc[NBINS], rc[NBINS];
      Box acc;
      acc.reset();
      uint32_t cnt = 0;
      for (int b = 0; b < NBINS; ++b) { acc.grow(bb[b]); cnt += bc[b]; la[b] = acc.half_area(); lc[b] = cnt; }
      acc.reset();
      cnt = 0;
      for (int b = NBINS - 1; b >= 0; --b) { acc.grow(bb[b]); cnt += bc[b]; ra[b] = acc.half_area(); rc[b] = cnt; }
      for (int b = 0; b < NBINS - 1; ++b) {
        if (lc[b] == 0 || rc[b + 1] == 0) continue;
        float cost = la[b] * (float)lc[b] + ra[b + 1] * (float)rc[b + 1];
        if (cost < best_cost) { best_cost = cost; best_axis = axis; best_split = b; }
      }
    }
    uint32_t mid;
    if (best_axis < 0) {
      mid = first + count / 2;  // all centroids coincide: split by index
    } else {
      float lo = cbox.lo[best_axis], scale = (float)NBINS / (cbox.hi[best_axis] - cbox.lo[best_axis]);
      auto it = std::partition(order.begin() + first, order.begin() + first + count, [&](uint32_t p) {
        float c = 0.5f * (pbox[p].lo[best_axis] + pbox[p].hi[best_axis]);
        int b = std::min(NBINS - 1, std::max(0, (int)((c - lo) * scale)));
        return b <= best_split;
      });
      mid = (uint32_t)(it - order.begin());
      if (mid == first || mid == first + count) mid = first + count / 2;
    }
    uint32_t lcount = mid - first, rcount = first + count - mid;
    uint32_t l, r;
    if (count > 200000 && live_threads.load() < max_threads) {
      live_threads.fetch_add(1);
      std::thread th([&]() { l = build(first, lcount, depth + 1); live_threads.fetch_sub(1); });
      r = build(mid, rcount, depth + 1);
      th.join();
    } else {
      l = build(first, lcount, depth + 1);
      r = build(mid, rcount, depth + 1);
    }
    n.left = l;
    n.right = r;
    nodes[idx] = n;
    return idx;
  }
};

}  // namespace

// Triangles per leaf.  Measured on B200 (profiles/r02, k_extend ms at 16 / 4 / 8 spp on console / city / foliage):
// 8: 6.75 / 24.7 / 48.5, 4: 5.92 / 19.1 / 40.4, 3: 5.24 / 16.4 / 38.3, 2: 5.22 / 16.5 / 36.6 — the traversal tests one
// primitive per step, so a large leaf only trades cheap box tests for Moller-Trumbore tests that mostly miss.
uint32_t bvh_max_leaf_tris() {
  if (const char *e = std::getenv("PT_B200_LEAF_TRIS")) return (uint32_t)std::min(8, std::max(1, std::atoi(e)));
  return 2;
}

float bvh_margin_scale() {
  if (const char *e = std::getenv("PT_B200_MARGIN_SCALE")) return std::max(0.0f, (float)std::atof(e));
  return 1.0f;
}

namespace {

inline Box inflate(const Box &b, float S) {
  Box r;
  for (int k = 0; k < 3; ++k) {
    float e = plane_pad(b.lo[k], b.hi[k], S);
    r.lo[k] = b.lo[k] - e;
    r.hi[k] = b.hi[k] + e;
  }
  return r;
}

}  // namespace

void build_bvh4(const HostScene &s, Bvh4 &out, int threads) {
  Builder B(s, threads);
  const uint32_t nt = B.nt, ns = (uint32_t)s.sphere_models.size();
  const uint32_t n = nt + ns;
  B.pbox.resize(n);
  Box scene;
  scene.reset();
  for (uint32_t t = 0; t < nt; ++t) {
    Box b;
    b.reset();
    for (int k = 0; k < 3; ++k) b.grow_pt(&s.pos[(size_t)t * 9 + k * 3]);
    B.pbox[t] = b;
    scene.grow(b);
  }
  for (uint32_t k = 0; k < ns; ++k) {
    const pt_model &m = s.models[s.sphere_models[k]];
    float r = std::fabs(m.radius);
    Box b;
    for (int a = 0; a < 3; ++a) { b.lo[a] = m.center[a] - r; b.hi[a] = m.center[a] + r; }
    B.pbox[nt + k] = b;
    scene.grow(b);
  }
  // margins (scene_host.hpp): primitive boxes grow by what the f32 intersection arithmetic can leak
  const float *cam = &s.camera[12];
  const float ms = bvh_margin_scale();
  const float S = ms * coordinate_scale(scene.lo, scene.hi, cam);
  for (uint32_t t = 0; t < nt; ++t) {
    const float *p = &s.pos[(size_t)t * 9];
    const float e = triangle_leak_pad(p, ms * origin_reach(p, scene.lo, scene.hi, cam));
    for (int a = 0; a < 3; ++a) { B.pbox[t].lo[a] -= e; B.pbox[t].hi[a] += e; }
  }
  for (uint32_t k = 0; k < ns; ++k) {
    const pt_model &m = s.models[s.sphere_models[k]];
    const float r = sphere_box_radius(m.radius, ms * origin_reach(m.center, scene.lo, scene.hi, cam));
    for (int a = 0; a < 3; ++a) { B.pbox[nt + k].lo[a] = m.center[a] - r; B.pbox[nt + k].hi[a] = m.center[a] + r; }
  }
  B.order.resize(n);
  for (uint32_t i = 0; i < n; ++i) B.order[i] = i;
  B.nodes.resize(2 * (size_t)n);
  uint32_t root = B.build(0, n, 0);

  // ---- collapse BVH2 -> BVH4 (DFS order so near children sit close in memory) ----
  out.nodes.clear();
  out.leaf_tris.clear();
  out.leaf_tris.reserve(nt);
  out.max_depth = 0;
  auto leaf_ref = [&](const Node2 &nd) -> uint32_t {
    uint32_t first = nd.left, count = nd.right & ~LEAF2;
    if (count == 1 && B.is_sphere(B.order[first])) return BVH_LEAF | BVH_SPHERE | (B.order[first] - nt);
    uint32_t slot = (uint32_t)out.leaf_tris.size();
    for (uint32_t i = 0; i < count; ++i) out.leaf_tris.push_back(B.order[first + i]);
    return BVH_LEAF | (slot << 3) | (count - 1);
  };
  const Node2 &rn = B.nodes[root];
  if (rn.right & LEAF2) {  // whole scene in one leaf: wrap it in a single node
    Bvh4Node nd{};
    for (int k = 0; k < 4; ++k) {
      nd.lo_x[k] = nd.lo_y[k] = nd.lo_z[k] = NAN;  // empty slot: NaN planes fail every slab test (isect.cuh slab_fma)
      nd.hi_x[k] = nd.hi_y[k] = nd.hi_z[k] = NAN;
      nd.child[k] = BVH_EMPTY;
    }
    Box b = inflate(rn.box, S);
    nd.lo_x[0] = b.lo[0]; nd.lo_y[0] = b.lo[1]; nd.lo_z[0] = b.lo[2];
    nd.hi_x[0] = b.hi[0]; nd.hi_y[0] = b.hi[1]; nd.hi_z[0] = b.hi[2];
    nd.child[0] = leaf_ref(rn);
    out.nodes.push_back(nd);
    out.max_depth = 1;
    return;
  }
  struct Work { uint32_t n2; uint32_t n4; int depth; };
  std::vector<Work> stack;
  out.nodes.emplace_back();
  stack.push_back({root, 0, 1});
  while (!stack.empty()) {
    Work w = stack.back();
    stack.pop_back();
    out.max_depth = std::max(out.max_depth, w.depth);
    uint32_t kids[4];
    int nk = 0;
    kids[nk++] = B.nodes[w.n2].left;
    kids[nk++] = B.nodes[w.n2].right;
    while (nk < 4) {  // open the inner child with the largest area
      int best = -1;
      float ba = -1.f;
      for (int k = 0; k < nk; ++k) {
        const Node2 &c = B.nodes[kids[k]];
        if (c.right & LEAF2) continue;
        float a = c.box.half_area();
        if (a > ba) { ba = a; best = k; }
      }
      if (best < 0) break;
      uint32_t c = kids[best];
      kids[best] = B.nodes[c].left;
      kids[nk++] = B.nodes[c].right;
    }
    Bvh4Node nd{};
    for (int k = 0; k < 4; ++k) {
      nd.lo_x[k] = nd.lo_y[k] = nd.lo_z[k] = NAN;  // empty slot: NaN planes fail every slab test (isect.cuh slab_fma)
      nd.hi_x[k] = nd.hi_y[k] = nd.hi_z[k] = NAN;
      nd.child[k] = BVH_EMPTY;
    }
    for (int k = 0; k < nk; ++k) {
      const Node2 &c = B.nodes[kids[k]];
      Box b = inflate(c.box, S);
      nd.lo_x[k] = b.lo[0]; nd.lo_y[k] = b.lo[1]; nd.lo_z[k] = b.lo[2];
      nd.hi_x[k] = b.hi[0]; nd.hi_y[k] = b.hi[1]; nd.hi_z[k] = b.hi[2];
      if (c.right & LEAF2) {
        nd.child[k] = leaf_ref(c);
      } else {
        uint32_t id4 = (uint32_t)out.nodes.size();
        out.nodes.emplace_back();
        nd.child[k] = id4;
        stack.push_back({kids[k], id4, w.depth + 1});
      }
    }
    out.nodes[w.n4] = nd;
  }
}

// Bvh4Node -> Bvh4QNode (scene_host.hpp).  Per axis: step = (extent of the union of the child boxes) / 250 rounded up,
// origin two steps below the union, so every plane lands in 1..254 after the guard.
void quantise_bvh4(const Bvh4 &bvh, uint32_t empty_ref, int threads, std::vector<Bvh4QNode> &out) {
  const size_t n = bvh.nodes.size();
  out.assign(n, Bvh4QNode{});
  auto work = [&](size_t begin, size_t end) {
    for (size_t i = begin; i < end; ++i) {
      const Bvh4Node &nd = bvh.nodes[i];
      Bvh4QNode qn{};
      const float *lo[3] = {nd.lo_x, nd.lo_y, nd.lo_z}, *hi[3] = {nd.hi_x, nd.hi_y, nd.hi_z};
      for (int a = 0; a < 3; ++a) {
        double nlo = INFINITY, nhi = -INFINITY;
        for (int k = 0; k < 4; ++k)
          if (nd.child[k] != BVH_EMPTY) {
            nlo = std::min(nlo, (double)lo[a][k]);
            nhi = std::max(nhi, (double)hi[a][k]);
          }
        if (!(nlo <= nhi)) nlo = nhi = 0.0;  // no children (cannot happen in a built tree)
        const double mag = std::max(std::fabs(nlo), std::fabs(nhi));
        double step_d = std::max((nhi - nlo) / 250.0, mag * 0x1p-18 + 1e-30);  // >= 32 ulp of the coordinate
        float step = std::nextafterf((float)step_d, INFINITY);
        float origin = std::nextafterf((float)(nlo - 2.0 * (double)step), -INFINITY);
        uint32_t wlo = 0, whi = 0;
        int tries = 0;
        for (bool fits = false; !fits;) {  // repeats only if a plane left 0..255 (never, given the two spare steps)
          fits = ++tries > 200;  // non-finite boxes (rejected by HostScene::finalize) must not hang the build
          if (fits) break;
          fits = true;
          wlo = whi = 0;
          for (int k = 0; k < 4; ++k) {
            int ql = 255, qh = 0;
            if (nd.child[k] != BVH_EMPTY) {
              ql = (int)std::floor(((double)lo[a][k] - (double)origin) / (double)step - 0.05);
              qh = (int)std::ceil(((double)hi[a][k] - (double)origin) / (double)step + 0.05);
              if (ql < 0 || qh > 255) fits = false;
              // the decoded planes must enclose the float box with most of the guard left
              if ((double)origin + ql * (double)step > (double)lo[a][k] - 0.04 * (double)step ||
                  (double)origin + qh * (double)step < (double)hi[a][k] + 0.04 * (double)step)
                fits = false;
            }
            wlo |= (uint32_t)(ql & 255) << (8 * k);
            whi |= (uint32_t)(qh & 255) << (8 * k);
          }
          if (!fits) {
            step = std::nextafterf(step * 1.01f, INFINITY);
            origin = std::nextafterf((float)(nlo - 2.0 * (double)step), -INFINITY);
          }
        }
        qn.origin[a] = origin;
        qn.scale[a] = step * 32768.0f;  // exact (power of two)
        qn.q[2 * a] = wlo;
        qn.q[2 * a + 1] = whi;
      }
      for (int k = 0; k < 4; ++k) qn.child[k] = nd.child[k] == BVH_EMPTY ? empty_ref : nd.child[k];
      out[i] = qn;
    }
  };
  const int nt = (int)std::max<size_t>(1, std::min<size_t>((size_t)std::max(1, threads), n / 4096 + 1));
  std::vector<std::thread> pool;
  for (int t = 0; t < nt; ++t) pool.emplace_back(work, n * t / nt, n * (t + 1) / nt);
  for (auto &th : pool) th.join();
}


}  // namespace ptb
