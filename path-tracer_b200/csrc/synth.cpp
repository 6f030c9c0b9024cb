// synth.cpp — procedural scenes for the large benchmark configs (BASELINE.json configs[2..4]).  The
// reference's "PS5" scene (README.md:15) is not distributed and isf-as-JSON is impractical at 1-20 M
// triangles, so these are generated in memory with a fixed SplitMix64 stream:
//   console  PS5-class: two curved thin white panels, glossy black body, blue emissive strips, ground
//            plane, one 1024^2 albedo texture, one point light, background 0.
//   foliage  displaced blobs + alpha "cards": 30 % of models opacity factor 0.3/0.5, 20 % with a 256^2
//            opacity texture; one directional + one point light; background 0.051.
//   city     heightfield + boxes + tessellated spheres, mixed metal/rough; one directional light; background 0.5.
//   soup     uniformly random triangles in [-1,1]^3 (traversal stress / parity batches).
#include <cmath>
#include <cstring>
#include <functional>

#include "host_api.hpp"

namespace ptb {
namespace {

struct Rng {
  uint64_t s;
  uint64_t next() {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
  }
  float uni() { return (float)(next() >> 40) * (1.0f / 16777216.0f); }
  float range(float a, float b) { return a + (b - a) * uni(); }
};

struct P3 {
  float x, y, z;
};
P3 operator+(P3 a, P3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
P3 operator-(P3 a, P3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
P3 operator*(P3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
P3 cross(P3 a, P3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
float dot(P3 a, P3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
P3 norm(P3 a) {
  float l = std::sqrt(dot(a, a));
  return l > 0 ? a * (1.0f / l) : P3{0, 1, 0};
}

pt_material mat(float r, float g, float b, float metal, float rough) {
  pt_material m{};
  m.albedo[0] = r; m.albedo[1] = g; m.albedo[2] = b;
  m.opacity = 1.f; m.metalness = metal; m.roughness = rough; m.ior = 1.f;
  m.albedo_tex = m.emissive_tex = m.opacity_tex = m.metalness_tex = m.roughness_tex = m.normal_tex = -1;
  return m;
}

struct Mesh {
  HostScene &s;
  uint32_t first;
  explicit Mesh(HostScene &sc) : s(sc), first((uint32_t)sc.n_tris()) {}
  void tri(P3 a, P3 b, P3 c, P3 na, P3 nb, P3 nc, float ua, float va, float ub, float vb, float uc, float vc) {
    float p[9] = {a.x, a.y, a.z, b.x, b.y, b.z, c.x, c.y, c.z};
    float n[9] = {na.x, na.y, na.z, nb.x, nb.y, nb.z, nc.x, nc.y, nc.z};
    float t[6] = {ua, va, ub, vb, uc, vc};
    s.pos.insert(s.pos.end(), p, p + 9);
    s.nrm.insert(s.nrm.end(), n, n + 9);
    s.uv.insert(s.uv.end(), t, t + 6);
  }
  // parametric grid: f(u,v) -> position; normals from finite differences (flip to invert)
  void grid(int nu, int nv, const std::function<P3(float, float)> &f, bool flip = false, bool closed_u = false) {
    std::vector<P3> P((size_t)(nu + 1) * (nv + 1)), N(P.size());
    for (int j = 0; j <= nv; ++j)
      for (int i = 0; i <= nu; ++i) P[(size_t)j * (nu + 1) + i] = f((float)i / nu, (float)j / nv);
    const float e = 1e-3f;
    for (int j = 0; j <= nv; ++j)
      for (int i = 0; i <= nu; ++i) {
        float u = (float)i / nu, v = (float)j / nv;
        P3 du = f(std::fmin(u + e, closed_u ? u + e : 1.f), v) - f(std::fmax(u - e, closed_u ? u - e : 0.f), v);
        P3 dv = f(u, std::fmin(v + e, 1.f)) - f(u, std::fmax(v - e, 0.f));
        P3 n = norm(cross(du, dv));
        N[(size_t)j * (nu + 1) + i] = flip ? n * -1.f : n;
      }
    for (int j = 0; j < nv; ++j)
      for (int i = 0; i < nu; ++i) {
        size_t a = (size_t)j * (nu + 1) + i, b = a + 1, c = a + nu + 1, d = c + 1;
        float u0 = (float)i / nu, u1 = (float)(i + 1) / nu, v0 = (float)j / nv, v1 = (float)(j + 1) / nv;
        if (!flip) {
          tri(P[a], P[b], P[d], N[a], N[b], N[d], u0, v0, u1, v0, u1, v1);
          tri(P[a], P[d], P[c], N[a], N[d], N[c], u0, v0, u1, v1, u0, v1);
        } else {
          tri(P[a], P[d], P[b], N[a], N[d], N[b], u0, v0, u1, v1, u1, v0);
          tri(P[a], P[c], P[d], N[a], N[c], N[d], u0, v0, u0, v1, u1, v1);
        }
      }
  }
  void box(P3 lo, P3 hi) {
    P3 c[8];
    for (int i = 0; i < 8; ++i) c[i] = {(i & 1) ? hi.x : lo.x, (i & 2) ? hi.y : lo.y, (i & 4) ? hi.z : lo.z};
    const int f[6][4] = {{0, 2, 6, 4}, {1, 5, 7, 3}, {0, 4, 5, 1}, {2, 3, 7, 6}, {0, 1, 3, 2}, {4, 6, 7, 5}};
    const P3 n[6] = {{-1, 0, 0}, {1, 0, 0}, {0, -1, 0}, {0, 1, 0}, {0, 0, -1}, {0, 0, 1}};
    for (int k = 0; k < 6; ++k) {
      tri(c[f[k][0]], c[f[k][1]], c[f[k][2]], n[k], n[k], n[k], 0, 0, 1, 0, 1, 1);
      tri(c[f[k][0]], c[f[k][2]], c[f[k][3]], n[k], n[k], n[k], 0, 0, 1, 1, 0, 1);
    }
  }
  void finish(const pt_material &m) {
    pt_model mo{};
    mo.type = PT_MODEL_MESH;
    mo.first_tri = first;
    mo.n_tris = (uint32_t)s.n_tris() - first;
    if (mo.n_tris == 0) return;
    s.models.push_back(mo);
    s.materials.push_back(m);
  }
};

void look_at(HostScene &s, P3 eye, P3 target, float fov) {
  P3 fwd = norm(target - eye);
  P3 right = norm(cross(fwd, P3{0, 1, 0}));
  P3 up = cross(right, fwd);
  float m[16] = {right.x, right.y, right.z, 0, up.x, up.y, up.z, 0, -fwd.x, -fwd.y, -fwd.z, 0, eye.x, eye.y, eye.z, 1};
  std::memcpy(s.camera, m, sizeof m);
  s.fov = fov;
}

int checker_texture(HostScene &s, uint32_t size, uint32_t channels, Rng &r, bool binary) {
  std::vector<uint8_t> d((size_t)size * size * channels);
  uint32_t cell = size / 16;
  for (uint32_t y = 0; y < size; ++y)
    for (uint32_t x = 0; x < size; ++x) {
      bool on = ((x / cell) + (y / cell)) & 1;
      for (uint32_t c = 0; c < channels; ++c) {
        uint8_t v = binary ? (on ? 255 : 40) : (uint8_t)((on ? 200 : 120) + (int)(r.next() % 40));
        d[((size_t)y * size + x) * channels + c] = v;
      }
    }
  return s.add_texture(size, size, channels, std::move(d));
}

int grid_dim(uint64_t tris) { return std::max(2, (int)std::sqrt((double)tris / 2.0)); }

// ---- console: PS5-class stand-in ----
void make_console(HostScene &s, uint64_t n_tris, Rng &r) {
  uint64_t per_panel_side = n_tris * 18 / 100, body = n_tris * 26 / 100;
  int tex = checker_texture(s, 1024, 3, r, false);
  for (int side = 0; side < 2; ++side) {
    float zoff = side ? 0.62f : -0.62f;
    for (int face = 0; face < 2; ++face) {  // outer and inner skin of a thin curved panel
      Mesh m(s);
      int n = grid_dim(per_panel_side);
      float th = face ? 0.03f : 0.f;
      float sgn = side ? 1.f : -1.f;
      m.grid(n, n, [=](float u, float v) {
        float x = (u - 0.5f) * 4.2f, y = (v - 0.5f) * 6.0f;
        float bulge = 0.35f * std::cos(u * 3.14159f - 1.5708f) + 0.15f * std::sin(v * 3.14159f);
        float flare = 0.25f * (v - 0.5f) * (v - 0.5f) * 4.f;
        return P3{x * (1.f + 0.08f * flare), y, zoff + sgn * (bulge - th + flare * 0.3f)};
      }, (side == 1) == (face == 1));
      pt_material ma = mat(0.8f, 0.8f, 0.8f, 0.f, 0.4f);
      if (face == 0) ma.albedo_tex = tex;
      m.finish(ma);
    }
  }
  {  // glossy black body: rounded slab
    Mesh m(s);
    int nu = grid_dim(body) * 2, nv = std::max(2, grid_dim(body) / 2);
    m.grid(nu, nv, [](float u, float v) {
      float a = u * 6.2831853f;
      float ca = std::cos(a), sa = std::sin(a);
      float e = 0.25f;  // superellipse cross-section
      float cx = std::copysign(std::pow(std::fabs(ca), e), ca) * 1.9f, cz = std::copysign(std::pow(std::fabs(sa), e), sa) * 0.5f;
      float y = (v - 0.5f) * 5.6f;
      float taper = 1.f - 0.1f * (v - 0.5f) * (v - 0.5f) * 4.f;
      return P3{cx * taper, y, cz * taper};
    }, false, true);
    m.finish(mat(0.02f, 0.02f, 0.02f, 0.f, 0.15f));
  }
  for (int k = 0; k < 2; ++k) {  // blue emissive strips
    Mesh m(s);
    float z = k ? 0.56f : -0.56f;
    int n = grid_dim(n_tris / 100);
    m.grid(n, n, [=](float u, float v) { return P3{(u - 0.5f) * 0.06f - 1.95f + 3.9f * 0.f, (v - 0.5f) * 5.0f, z + (k ? 0.03f : -0.03f) * u}; });
    pt_material ma = mat(0.1f, 0.2f, 0.9f, 0.f, 0.5f);
    ma.emissive[0] = 0.1f; ma.emissive[1] = 0.3f; ma.emissive[2] = 2.0f;
    m.finish(ma);
  }
  {  // ground plane (2 triangles)
    Mesh m(s);
    m.tri({-30, -3.01f, -30}, {30, -3.01f, 30}, {30, -3.01f, -30}, {0, 1, 0}, {0, 1, 0}, {0, 1, 0}, 0, 0, 1, 1, 1, 0);
    m.tri({-30, -3.01f, -30}, {-30, -3.01f, 30}, {30, -3.01f, 30}, {0, 1, 0}, {0, 1, 0}, {0, 1, 0}, 0, 0, 0, 1, 1, 1);
    m.finish(mat(0.5f, 0.5f, 0.5f, 0.f, 0.6f));
  }
  pt_light l{};
  l.type = PT_LIGHT_POINT;
  l.vec[0] = 6.f; l.vec[1] = 8.f; l.vec[2] = 9.f;
  l.color[0] = l.color[1] = l.color[2] = 1000.f;
  l.size = 0.1f;
  s.lights.push_back(l);
  s.background[0] = s.background[1] = s.background[2] = 0.f;
  look_at(s, {7.5f, 2.5f, 16.f}, {0, 0, 0}, 0.3995965f);
}

// blob: displaced uv-sphere
void add_blob(Mesh &m, P3 c, float rad, int n, float amp, float fx, float fy) {
  m.grid(2 * n, n, [=](float u, float v) {
    float a = u * 6.2831853f, b = v * 3.14159265f;
    float rr = rad * (1.f + amp * std::sin(fx * a) * std::sin(fy * b));
    return P3{c.x + rr * std::sin(b) * std::cos(a), c.y + rr * std::cos(b), c.z + rr * std::sin(b) * std::sin(a)};
  }, true, true);
}

void make_foliage(HostScene &s, uint64_t n_tris, Rng &r) {
  const int n_models = 240;
  int otex = checker_texture(s, 256, 1, r, true);
  uint64_t per_model = std::max<uint64_t>(64, n_tris / n_models);
  {
    Mesh g(s);
    g.tri({-60, 0, -60}, {60, 0, 60}, {60, 0, -60}, {0, 1, 0}, {0, 1, 0}, {0, 1, 0}, 0, 0, 1, 1, 1, 0);
    g.tri({-60, 0, -60}, {-60, 0, 60}, {60, 0, 60}, {0, 1, 0}, {0, 1, 0}, {0, 1, 0}, 0, 0, 0, 1, 1, 1);
    g.finish(mat(0.35f, 0.3f, 0.2f, 0.f, 0.9f));
  }
  for (int k = 0; k < n_models; ++k) {
    Mesh m(s);
    float roll = r.uni();
    P3 base{r.range(-40, 40), 0.f, r.range(-40, 40)};
    if (k % 2 == 0) {  // a few displaced blobs (bush)
      int blobs = 4;
      int n = std::max(2, (int)std::sqrt((double)per_model / blobs / 4.0));
      for (int b = 0; b < blobs; ++b)
        add_blob(m, base + P3{r.range(-2, 2), r.range(1, 6), r.range(-2, 2)}, r.range(1.0f, 2.5f), n, 0.12f, 5.f + b, 7.f);
    } else {  // leaf cards: many small quads
      uint64_t cards = per_model / 2;
      for (uint64_t c = 0; c < cards; ++c) {
        P3 o = base + P3{r.range(-4, 4), r.range(0.5f, 9), r.range(-4, 4)};
        P3 a = norm(P3{r.range(-1, 1), r.range(-1, 1), r.range(-1, 1)}) * r.range(0.2f, 0.6f);
        P3 bb = norm(cross(a, P3{r.range(-1, 1), r.range(-1, 1), r.range(-1, 1)})) * r.range(0.2f, 0.6f);
        P3 n = norm(cross(a, bb));
        m.tri(o, o + a, o + a + bb, n, n, n, 0, 0, 1, 0, 1, 1);
        m.tri(o, o + a + bb, o + bb, n, n, n, 0, 0, 1, 1, 0, 1);
      }
    }
    pt_material ma = mat(r.range(0.1f, 0.4f), r.range(0.4f, 0.8f), r.range(0.1f, 0.3f), 0.f, r.range(0.4f, 0.9f));
    if (roll < 0.3f) ma.opacity = (k & 2) ? 0.3f : 0.5f;
    else if (roll < 0.5f) ma.opacity_tex = otex;
    m.finish(ma);
  }
  pt_light d{};
  d.type = PT_LIGHT_DIRECTIONAL;
  P3 dir = norm(P3{-0.4f, -1.f, -0.3f});
  d.vec[0] = dir.x; d.vec[1] = dir.y; d.vec[2] = dir.z;
  d.color[0] = d.color[1] = d.color[2] = 2.f;
  s.lights.push_back(d);
  pt_light p{};
  p.type = PT_LIGHT_POINT;
  p.vec[0] = 10; p.vec[1] = 25; p.vec[2] = 20;
  p.color[0] = p.color[1] = p.color[2] = 4000.f;
  p.size = 0.1f;
  s.lights.push_back(p);
  s.background[0] = s.background[1] = s.background[2] = 0.051f;
  look_at(s, {55, 30, 75}, {0, 4, 0}, 0.5f);
}

void make_city(HostScene &s, uint64_t n_tris, Rng &r) {
  {  // heightfield terrain: 40 % of the budget
    Mesh m(s);
    int n = grid_dim(n_tris * 40 / 100);
    m.grid(n, n, [](float u, float v) {
      float x = (u - 0.5f) * 100.f, z = (v - 0.5f) * 100.f;
      float y = 1.5f * std::sin(x * 0.21f) * std::cos(z * 0.17f) + 0.4f * std::sin(x * 1.3f + z * 0.9f);
      return P3{x, y - 2.f, z};
    }, true);
    m.finish(mat(0.45f, 0.45f, 0.42f, 0.f, 0.8f));
  }
  uint64_t box_tris = n_tris * 10 / 100, sphere_tris = n_tris * 50 / 100;
  const int n_groups = 16;
  for (int g = 0; g < n_groups; ++g) {  // buildings
    Mesh m(s);
    uint64_t boxes = box_tris / 12 / n_groups;
    for (uint64_t b = 0; b < boxes; ++b) {
      float x = r.range(-48, 48), z = r.range(-48, 48), w = r.range(0.3f, 1.5f), d = r.range(0.3f, 1.5f), h = r.range(1, 14);
      m.box({x - w, -2, z - d}, {x + w, h, z + d});
    }
    m.finish(mat(r.range(0.3f, 0.9f), r.range(0.3f, 0.9f), r.range(0.3f, 0.9f), (g & 1) ? 1.f : 0.f, r.range(0.05f, 0.9f)));
  }
  for (int g = 0; g < n_groups; ++g) {  // tessellated spheres (domes, tanks)
    Mesh m(s);
    int per_group = 24;
    int n = std::max(2, (int)std::sqrt((double)sphere_tris / n_groups / per_group / 4.0));
    for (int k = 0; k < per_group; ++k)
      add_blob(m, {r.range(-45, 45), r.range(0, 10), r.range(-45, 45)}, r.range(0.8f, 3.f), n, 0.f, 1.f, 1.f);
    m.finish(mat(r.range(0.5f, 1.f), r.range(0.5f, 1.f), r.range(0.5f, 1.f), (g % 3 == 0) ? 1.f : 0.f, r.range(0.02f, 0.7f)));
  }
  pt_light d{};
  d.type = PT_LIGHT_DIRECTIONAL;
  P3 dir = norm(P3{0.5f, -1.f, 0.35f});
  d.vec[0] = dir.x; d.vec[1] = dir.y; d.vec[2] = dir.z;
  d.color[0] = d.color[1] = d.color[2] = 3.f;
  s.lights.push_back(d);
  s.background[0] = s.background[1] = s.background[2] = 0.5f;
  look_at(s, {70, 45, 85}, {0, 2, 0}, 0.6f);
}

void make_soup(HostScene &s, uint64_t n_tris, Rng &r) {
  Mesh m(s);
  for (uint64_t t = 0; t < n_tris; ++t) {
    P3 c{r.range(-1, 1), r.range(-1, 1), r.range(-1, 1)};
    float sz = 2.0f / std::cbrt((float)n_tris + 1.f);
    P3 a = c + P3{r.range(-sz, sz), r.range(-sz, sz), r.range(-sz, sz)};
    P3 b = c + P3{r.range(-sz, sz), r.range(-sz, sz), r.range(-sz, sz)};
    P3 cc = c + P3{r.range(-sz, sz), r.range(-sz, sz), r.range(-sz, sz)};
    P3 n = norm(cross(b - a, cc - a));
    m.tri(a, b, cc, n, n, n, 0, 0, 1, 0, 0, 1);
  }
  m.finish(mat(0.7f, 0.7f, 0.7f, 0.f, 0.5f));
  pt_light p{};
  p.type = PT_LIGHT_POINT;
  p.vec[0] = 3; p.vec[1] = 4; p.vec[2] = 5;
  p.color[0] = p.color[1] = p.color[2] = 300.f;
  s.lights.push_back(p);
  s.background[0] = s.background[1] = s.background[2] = 0.2f;
  look_at(s, {0, 0, 6}, {0, 0, 0}, 0.5f);
}

}  // namespace

int make_synthetic(const char *recipe, uint64_t n_tris, uint64_t seed, HostScene &out) {
  if (!recipe || n_tris < 16 || n_tris > (1ull << 26)) {
    set_error("synthetic scene: bad recipe or triangle count");
    return PT_ERR_INVALID;
  }
  Rng r{seed * 0x9E3779B97F4A7C15ull + 0xB200ull};
  std::string name(recipe);
  if (name == "console") make_console(out, n_tris, r);
  else if (name == "foliage") make_foliage(out, n_tris, r);
  else if (name == "city") make_city(out, n_tris, r);
  else if (name == "soup") make_soup(out, n_tris, r);
  else {
    set_error("unknown synthetic recipe `%s` (console|foliage|city|soup)", recipe);
    return PT_ERR_INVALID;
  }
  return PT_OK;
}

}  // namespace ptb

// The oracle links this file for the generators alone (oracle/Makefile: "same generator code compiled into the
// oracle"), without the device-side scene constructor.
#ifndef PTB_SYNTH_GENERATOR_ONLY
extern "C" int pt_scene_synthetic(const char *recipe, uint64_t n_tris, uint64_t seed, int device, pt_scene **out) {
  if (!out) {
    ptb::set_error("null argument");
    return PT_ERR_INVALID;
  }
  return ptb::abi_guard([&]() -> int {
    ptb::HostScene h;
    int rc = ptb::make_synthetic(recipe, n_tris, seed, h);
    if (rc != PT_OK) return rc;
    return ptb::scene_from_host(std::move(h), device, out);
  });
}
#endif
