// isf_binary.cpp — "isfb": the binary side-car of an isf scene.
//
// The reference stores a scene as one JSON document (scene/isf.rs:5-142) and parses it with serde_json
// (scene/mod.rs:16-22); a 20 M-triangle scene is ~6 GB of decimal text.  An isfb file holds the same flat scene as raw
// little-endian arrays, decoded textures included, so a load is one sequential read.  It may also carry the 4-wide
// BVH of the scene (either builder's), checked structurally on load, so that the 4 s host SAH build of a 20 M-triangle
// scene is paid once.
//
// Layout (all little-endian, sections padded to 16 bytes):
//   header  64 B : "ISFB", u32 version, u32 flags (bit0 = accel present), u32 n_models, u64 n_tris, u32 n_textures,
//                  u32 n_lights, u64 n_nodes, u64 n_leaf_tris, u32 accel_kind, u32 accel_rev, i32 max_depth, u32 root_ref
//   view         : camera[16] f32, fov f32, background[3] f32
//   models       : n_models x pt_model (28 B)        materials : n_models x pt_material (64 B)
//   lights       : n_lights x pt_light (32 B)
//   tri_pos      : n_tris x 9 f32    tri_nrm : n_tris x 9 f32    tri_uv : n_tris x 6 f32
//   textures     : per texture { u32 width, height, channels, 0 } + width*height*channels bytes
//   accel        : n_nodes x 128 B nodes, n_leaf_tris x u32 (only when flags bit0)
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "host_api.hpp"

namespace ptb {
namespace {

constexpr uint32_t ISFB_MAGIC = 0x42465349u;  // "ISFB"
constexpr uint32_t ISFB_VERSION = 1;

struct Header {
  uint32_t magic, version, flags, n_models;
  uint64_t n_tris;
  uint32_t n_textures, n_lights;
  uint64_t n_nodes, n_leaf_tris;
  uint32_t accel_kind, accel_rev;
  int32_t max_depth;
  uint32_t root_ref;
};
static_assert(sizeof(Header) == 64, "isfb header is 64 bytes");
static_assert(sizeof(pt_model) == 28 && sizeof(pt_material) == 64 && sizeof(pt_light) == 32, "isfb record sizes");

struct Writer {
  FILE *f;
  uint64_t off = 0;
  bool ok = true;
  void put(const void *p, size_t n) {
    if (n && std::fwrite(p, 1, n, f) != n) ok = false;
    off += n;
  }
  void pad() {
    static const char z[16] = {0};
    put(z, (16 - off % 16) % 16);
  }
};

struct Reader {
  FILE *f;
  uint64_t off = 0, size = 0;
  bool get(void *p, size_t n) {
    if (n > size - off) return false;
    if (n && std::fread(p, 1, n, f) != n) return false;
    off += n;
    return true;
  }
  bool pad() {
    char z[16];
    return get(z, (16 - off % 16) % 16);
  }
};

}  // namespace

bool is_isfb_file(const char *path) {
  FILE *f = std::fopen(path, "rb");
  if (!f) return false;
  uint32_t m = 0;
  bool ok = std::fread(&m, 4, 1, f) == 1 && m == ISFB_MAGIC;
  std::fclose(f);
  return ok;
}

int save_isfb_file(const char *path, const HostScene &s, const Bvh4 *accel, int accel_kind) {
  FILE *f = std::fopen(path, "wb");
  if (!f) {
    set_error("cannot create '%s'", path);
    return PT_ERR_IO;
  }
  Writer w{f};
  Header h{};
  h.magic = ISFB_MAGIC;
  h.version = ISFB_VERSION;
  h.n_models = (uint32_t)s.models.size();
  h.n_tris = s.n_tris();
  h.n_textures = (uint32_t)s.textures.size();
  h.n_lights = (uint32_t)s.lights.size();
  if (accel && !accel->nodes.empty()) {
    h.flags |= 1u;
    h.n_nodes = accel->nodes.size();
    h.n_leaf_tris = accel->leaf_tris.size();
    h.accel_kind = (uint32_t)accel_kind;
    h.accel_rev = BVH_ACCEL_REV;
    h.max_depth = accel->max_depth;
    h.root_ref = accel->root_ref;
  }
  w.put(&h, sizeof h);
  w.put(s.camera, sizeof s.camera);
  w.put(&s.fov, 4);
  w.put(s.background, sizeof s.background);
  w.pad();
  w.put(s.models.data(), s.models.size() * sizeof(pt_model));
  w.pad();
  w.put(s.materials.data(), s.materials.size() * sizeof(pt_material));
  w.pad();
  w.put(s.lights.data(), s.lights.size() * sizeof(pt_light));
  w.pad();
  w.put(s.pos.data(), s.pos.size() * 4);
  w.pad();
  w.put(s.nrm.data(), s.nrm.size() * 4);
  w.pad();
  w.put(s.uv.data(), s.uv.size() * 4);
  w.pad();
  for (size_t i = 0; i < s.textures.size(); ++i) {
    uint32_t th[4] = {s.textures[i].width, s.textures[i].height, s.textures[i].channels, 0};
    w.put(th, sizeof th);
    w.put(s.texdata[i].data(), s.texdata[i].size());
    w.pad();
  }
  if (h.flags & 1u) {
    w.put(accel->nodes.data(), accel->nodes.size() * sizeof(Bvh4Node));
    w.put(accel->leaf_tris.data(), accel->leaf_tris.size() * 4);
    w.pad();
  }
  bool ok = w.ok;
  ok &= std::fclose(f) == 0;
  if (!ok) {
    set_error("write error on '%s'", path);
    return PT_ERR_IO;
  }
  return PT_OK;
}

int load_isfb_file(const char *path, HostScene &out, Bvh4 *accel, int *accel_kind) {
  FILE *f = std::fopen(path, "rb");
  if (!f) {
    set_error("cannot open '%s'", path);
    return PT_ERR_IO;
  }
  struct Closer {
    FILE *f;
    ~Closer() { std::fclose(f); }
  } closer{f};
  Reader r{f};
  std::fseek(f, 0, SEEK_END);
  r.size = (uint64_t)std::ftell(f);
  std::fseek(f, 0, SEEK_SET);
  auto bad = [&](const char *what) {
    set_error("'%s': %s", path, what);
    return PT_ERR_PARSE;
  };
  Header h;
  if (!r.get(&h, sizeof h) || h.magic != ISFB_MAGIC) return bad("not an isfb file");
  if (h.version != ISFB_VERSION) return bad("unsupported isfb version");
  // every count is checked against the file size before anything is allocated
  const uint64_t tri_bytes = h.n_tris * 96;
  if (h.n_tris > r.size / 96 || (uint64_t)h.n_models * 92 > r.size || (uint64_t)h.n_lights * 32 > r.size ||
      (uint64_t)h.n_textures * 16 > r.size || h.n_nodes > r.size / 128 || h.n_leaf_tris > r.size / 4 ||
      tri_bytes > r.size)
    return bad("truncated (counts exceed the file size)");
  out = HostScene();
  if (!r.get(out.camera, sizeof out.camera) || !r.get(&out.fov, 4) || !r.get(out.background, sizeof out.background) ||
      !r.pad())
    return bad("truncated view section");
  out.models.resize(h.n_models);
  out.materials.resize(h.n_models);
  out.lights.resize(h.n_lights);
  if (!r.get(out.models.data(), out.models.size() * sizeof(pt_model)) || !r.pad() ||
      !r.get(out.materials.data(), out.materials.size() * sizeof(pt_material)) || !r.pad() ||
      !r.get(out.lights.data(), out.lights.size() * sizeof(pt_light)) || !r.pad())
    return bad("truncated model/material/light section");
  out.pos.resize(h.n_tris * 9);
  out.nrm.resize(h.n_tris * 9);
  out.uv.resize(h.n_tris * 6);
  if (!r.get(out.pos.data(), out.pos.size() * 4) || !r.pad() || !r.get(out.nrm.data(), out.nrm.size() * 4) ||
      !r.pad() || !r.get(out.uv.data(), out.uv.size() * 4) || !r.pad())
    return bad("truncated triangle section");
  for (uint32_t i = 0; i < h.n_textures; ++i) {
    uint32_t th[4];
    if (!r.get(th, sizeof th)) return bad("truncated texture header");
    const uint64_t bytes = (uint64_t)th[0] * th[1] * th[2];
    if ((th[2] != 1 && th[2] != 3) || !th[0] || !th[1] || bytes > r.size - r.off) return bad("bad texture header");
    std::vector<uint8_t> px(bytes);
    if (!r.get(px.data(), bytes) || !r.pad()) return bad("truncated texture");
    out.add_texture(th[0], th[1], th[2], std::move(px));
  }
  if (accel_kind) *accel_kind = -1;
  if ((h.flags & 1u) && accel && h.accel_rev == BVH_ACCEL_REV) {
    accel->nodes.resize(h.n_nodes);
    accel->leaf_tris.resize(h.n_leaf_tris);
    if (!r.get(accel->nodes.data(), h.n_nodes * sizeof(Bvh4Node)) || !r.get(accel->leaf_tris.data(), h.n_leaf_tris * 4))
      return bad("truncated accel section");
    accel->max_depth = h.max_depth;
    accel->root_ref = h.root_ref;
    if (accel_kind) *accel_kind = (int)h.accel_kind;
  }  // an accel section written under other box-inflation rules (accel_rev) is ignored: the scene is rebuilt
  return PT_OK;
}

// Structural check of a BVH read from a file: every reference in range, every node reached exactly once, the leaf
// slots partition the leaf-ordered triangle array, leaf_tris is a permutation of the triangles, every sphere is
// referenced exactly once, the depth fits the traversal stack, and every child box contains what hangs below it
// (triangle vertices, sphere extents, grandchild boxes) — the property the traversal's culling relies on.
std::string validate_bvh4(const HostScene &s, const Bvh4 &b) {
  const uint64_t nt = s.n_tris(), ns = s.sphere_models.size(), nn = b.nodes.size();
  if (nn == 0) return "no nodes";
  if (b.leaf_tris.size() != nt) return "leaf triangle count differs from the scene";
  std::vector<uint8_t> seen_tri(nt, 0), seen_slot(nt, 0), seen_node(nn, 0), seen_sphere(ns, 0);
  for (uint32_t t : b.leaf_tris) {
    if (t >= nt || seen_tri[t]) return "leaf order is not a permutation of the triangles";
    seen_tri[t] = 1;
  }
  struct Item { uint32_t node; int depth; float lo[3], hi[3]; };
  std::vector<Item> stack;
  Item root{0, 1, {-INFINITY, -INFINITY, -INFINITY}, {INFINITY, INFINITY, INFINITY}};
  stack.push_back(root);
  seen_node[0] = 1;
  int max_depth = 0;
  uint64_t slots = 0, spheres = 0;
  while (!stack.empty()) {
    Item it = stack.back();
    stack.pop_back();
    max_depth = std::max(max_depth, it.depth);
    const Bvh4Node &n = b.nodes[it.node];
    for (int k = 0; k < 4; ++k) {
      const uint32_t c = n.child[k];
      if (c == BVH_EMPTY) continue;
      const float lo[3] = {n.lo_x[k], n.lo_y[k], n.lo_z[k]}, hi[3] = {n.hi_x[k], n.hi_y[k], n.hi_z[k]};
      for (int a = 0; a < 3; ++a) {
        if (!(lo[a] <= hi[a])) return "child box is empty or NaN";
        if (!(lo[a] >= it.lo[a] && hi[a] <= it.hi[a])) return "child box is not inside its parent's";
      }
      auto inside = [&](const float *p, float r) {
        for (int a = 0; a < 3; ++a)
          if (!(p[a] - r >= lo[a] && p[a] + r <= hi[a])) return false;
        return true;
      };
      if (!(c & BVH_LEAF)) {
        if (c >= nn || seen_node[c]) return "inner reference out of range or shared";
        seen_node[c] = 1;
        Item ch{c, it.depth + 1, {lo[0], lo[1], lo[2]}, {hi[0], hi[1], hi[2]}};
        stack.push_back(ch);
      } else if (c & BVH_SPHERE) {
        const uint32_t si = c & 0x3FFFFFFFu;
        if (si >= ns || seen_sphere[si]) return "sphere reference out of range or repeated";
        seen_sphere[si] = 1;
        ++spheres;
        const pt_model &m = s.models[s.sphere_models[si]];
        if (!inside(m.center, std::fabs(m.radius))) return "sphere outside its leaf box";
      } else {
        const uint64_t first = (c & 0x3FFFFFFFu) >> 3, count = (c & 7u) + 1;
        if (first + count > nt) return "triangle leaf out of range";
        for (uint64_t i = first; i < first + count; ++i) {
          if (seen_slot[i]) return "triangle slot referenced twice";
          seen_slot[i] = 1;
          const float *p = &s.pos[(size_t)b.leaf_tris[i] * 9];
          for (int v = 0; v < 3; ++v)
            if (!inside(p + 3 * v, 0.f)) return "triangle outside its leaf box";
        }
        slots += count;
      }
    }
  }
  if (slots != nt) return "leaves do not cover every triangle";
  if (spheres != ns) return "leaves do not cover every sphere";
  for (uint64_t i = 0; i < nn; ++i)
    if (!seen_node[i]) return "unreachable node";
  if (max_depth != b.max_depth) return "recorded depth differs from the tree";
  return "";
}

}  // namespace ptb
