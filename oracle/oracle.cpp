// oracle/oracle.cpp — TEST INFRASTRUCTURE: CPU restatement of the reference's hot path.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library; the product (path-tracer_b200/) never links or calls it.
//
// Parity pinning: Möller–Trumbore is pinned by the reference's 3012 hit + 3012 miss vectors
// (tests/moller_trumbore/*.yml, tolerance of triangle.rs:213-216); the whole integrator is
// checked against the seven SHA-1 image hashes of src/main.rs:100-165 through the ChaCha12
// (rand 0.8 StdRng) emulation — see tests/test_oracle_kat.py for which of them reproduce.
//
// Restates: Renderer::render renderer/mod.rs:76-169, render_pixel :172-228,
// compute_radiance :230-278, get_light_info :281-333, post_processing :335-353.
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <thread>

#include "canonical.hpp"
#include "shading.hpp"

using namespace orc;

namespace {

const float NORMAL_BIAS = 0.00001f;  // renderer/mod.rs:59

struct Counters {
  uint64_t extend = 0, shadow = 0;
  bool canon_on = false;  // also run the canonical work counter on every ray (canonical.hpp)
  CanonCounts canon;
};

// get_light_info, renderer/mod.rs:281-333
void light_info(const Scene &s, const pt_light &light, const Hit &hit, bool brute, std::vector<Hit> &tmp, V3 &radiance,
                V3 &direction, Counters &cnt) {
  V3 lc = v3(light.color[0], light.color[1], light.color[2]);
  V3 ori = hit.position + hit.normal * NORMAL_BIAS;
  if (light.type == PT_LIGHT_DIRECTIONAL) {
    V3 dir = v3(light.vec[0], light.vec[1], light.vec[2]);
    V3 sdir = -1.0f * dir;
    V3 color = lc;
    ray_cast(s, ori, sdir, tmp, brute);
    ++cnt.shadow;
    if (cnt.canon_on) s.canon->count_any(s, ori, sdir, FLT_MAX, cnt.canon);
    for (const Hit &sh : tmp) {
      MaterialSample ms = hit_material_sample(s, sh, sh.model);
      color = color * (1.0f - ms.opacity);
      if (sum(color) == 0.0f) break;
    }
    radiance = color;
    direction = dir;
    return;
  }
  V3 lp = v3(light.vec[0], light.vec[1], light.vec[2]);
  V3 dir = hit.position - lp;
  float dist = magnitude(dir);
  dir = normalize(dir);
  V3 sdir = -1.0f * dir;
  float dissipation = 4.0f * PI_F * dist * dist;
  V3 ld = lc / dissipation;
  ray_cast(s, ori, sdir, tmp, brute);
  ++cnt.shadow;
  if (cnt.canon_on) s.canon->count_any(s, ori, sdir, dist, cnt.canon);
  for (const Hit &sh : tmp) {
    if (magnitude(sh.position - hit.position) > dist) break;
    // quirk (mod.rs:324): the PRIMARY hit's variant/uv with the occluder's material
    MaterialSample ms = hit_material_sample(s, hit, sh.model);
    ld = ld * (1.0f - ms.opacity);
    if (sum(ld) == 0.0f) break;
  }
  radiance = ld;
  direction = dir;
}

// render_pixel, renderer/mod.rs:172-228 (+ compute_radiance :230-278 inlined)
struct Recorder {  // optional per-bounce record (pt_path_vertex), filled by render_pixel
  pt_path_vertex *out = nullptr;
  uint32_t n = 0;
  void put(V3 ro, V3 rd, int32_t prim, float t, V3 thr, V3 color) {
    pt_path_vertex v;
    v.ray_o[0] = ro.x; v.ray_o[1] = ro.y; v.ray_o[2] = ro.z;
    v.ray_d[0] = rd.x; v.ray_d[1] = rd.y; v.ray_d[2] = rd.z;
    v.prim_id = prim;
    v.t = t;
    v.throughput[0] = thr.x; v.throughput[1] = thr.y; v.throughput[2] = thr.z;
    v.radiance[0] = color.x; v.radiance[1] = color.y; v.radiance[2] = color.z;
    out[n++] = v;
  }
};

V3 render_pixel(const Scene &s, const pt_profile &prof, V3 ro, V3 rd, PathRng &rng, bool brute, std::vector<Hit> &hits,
                std::vector<Hit> &tmp, Counters &cnt, Recorder *rec = nullptr) {
  V3 color = v3(0, 0, 0), thr = v3(1, 1, 1);
  for (uint32_t bounce = 0; bounce < prof.bounces + 1; ++bounce) {
    ray_cast(s, ro, rd, hits, brute);
    ++cnt.extend;
    if (cnt.canon_on) s.canon->count_closest(s, ro, rd, cnt.canon);
    if (hits.empty()) {
      if (rec) rec->put(ro, rd, -1, 0.0f, thr, color + mul_el(thr, s.background));
      return color + mul_el(thr, s.background);
    }
    const Hit *hit = nullptr;
    MaterialSample mat{};
    V3 normal{};
    for (const Hit &h : hits) {  // alpha walk, mod.rs:188-205
      mat = hit_material_sample(s, h, h.model);
      normal = hit_shading_normal(s, h, h.model);
      hit = &h;
      float op = mat.opacity;
      if (op >= 1.0f) break;
      if (op > 0.001f) {
        if (rng.alpha(bounce, (uint32_t)h.prim_id) < op) break;
      }
    }
    V3 view = -1.0f * rd;
    CookTorrance brdf(mat);
    color = color + mul_el(thr, mat.emissive);
    for (const pt_light &light : s.lights) {
      V3 lr, ldir;
      light_info(s, light, *hit, brute, tmp, lr, ldir, cnt);
      if (lr.x == 0.0f && lr.y == 0.0f && lr.z == 0.0f) continue;
      V3 rl = -1.0f * ldir;
      color = color + mul_el(mul_el(thr, brdf.eval_direct(normal, view, rl)), lr);
    }
    const V3 seg_o = ro, seg_d = rd;
    const int32_t seg_prim = hit->prim_id;
    const float seg_t = hit->dist;
    if (bounce < prof.bounces) {
      V3 no = hit->position + hit->normal * NORMAL_BIAS;
      float r1 = rng.bsdf(bounce, 0);
      float r2 = rng.bsdf(bounce, 1);
      V3 nd = brdf.sample(normal, view, r1, r2);
      V3 w = brdf.eval_indirect(normal, view, nd) / 1.0f;
      thr = mul_el(thr, w);
      ro = no;
      rd = nd;
    } else {
      ro = v3(0, 0, 0);
      rd = v3(0, 0, 0);
    }
    bool stop = dot(thr, thr) < 0.00001f;
    if (!stop && bounce > 3) {  // russian_roulette, utils.rs:23-31
      float p = rmax(rmax(thr.x, thr.y), thr.z);
      thr = thr * (1.0f / p);
      if (rng.rr(bounce) > p) stop = true;
    }
    if (rec) rec->put(seg_o, seg_d, seg_prim, seg_t, thr, color);
    if (stop) return color;
  }
  return color;
}

// raygen, renderer/mod.rs:107-124 + camera.rs:36-48
void raygen(const Scene &s, const pt_profile &prof, uint32_t x, uint32_t y, PathRng &rng, V3 &o, V3 &d) {
  float wf = (float)prof.width, hf = (float)prof.height;
  float ratio = wf / hf;
  float sx = (float)x + rng.jitter(0);
  sx = sx / wf * 2.0f - 1.0f;
  sx *= std::tan(s.fov / 2.0f) * ratio;
  float sy = (float)y + rng.jitter(1);
  sy = 1.0f - sy / hf * 2.0f;
  sy *= std::tan(s.fov / 2.0f);
  V3 dir = normalize(v3(sx, sy, -1.0f));
  const float *m = s.camera;  // column-major: column c at m[c*4..]
  V3 c0 = v3(m[0], m[1], m[2]), c1 = v3(m[4], m[5], m[6]), c2 = v3(m[8], m[9], m[10]), c3 = v3(m[12], m[13], m[14]);
  d = ((c0 * dir.x + c1 * dir.y) + c2 * dir.z) + c3 * 0.0f;
  o = c3;
}

}  // namespace

extern "C" {

// triangle.rs:37-82 on one (ray, triangle) pair.  out4 = t,u,v,det.  Returns 1 on hit.
int oracle_moller_trumbore(const float *ray6, const float *tri9, float *out4) {
  float t = 0, u = 0, v = 0, det = 0;
  bool hit = moller_trumbore(v3(ray6[0], ray6[1], ray6[2]), v3(ray6[3], ray6[4], ray6[5]), v3(tri9[0], tri9[1], tri9[2]),
                             v3(tri9[3], tri9[4], tri9[5]), v3(tri9[6], tri9[7], tri9[8]), t, u, v, det);
  out4[0] = t; out4[1] = u; out4[2] = v; out4[3] = det;
  return hit ? 1 : 0;
}

void *oracle_scene_create(const pt_scene_desc *d) {
  Scene *s = new Scene();
  s->models.assign(d->models, d->models + d->n_models);
  s->materials.assign(d->materials, d->materials + d->n_models);
  s->pos.assign(d->tri_pos, d->tri_pos + d->n_tris * 9);
  s->nrm.assign(d->tri_nrm, d->tri_nrm + d->n_tris * 9);
  s->uv.assign(d->tri_uv, d->tri_uv + d->n_tris * 6);
  s->texdata.resize(d->n_textures);
  for (uint32_t i = 0; i < d->n_textures; ++i) {
    const pt_texture &t = d->textures[i];
    s->texdata[i].assign(t.data, t.data + (size_t)t.width * t.height * t.channels);
    pt_texture c = t;
    c.data = s->texdata[i].data();
    s->textures.push_back(c);
  }
  s->lights.assign(d->lights, d->lights + d->n_lights);
  std::memcpy(s->camera, d->camera, sizeof s->camera);
  s->fov = d->fov;
  s->background = v3(d->background[0], d->background[1], d->background[2]);
  s->tri_model.resize(d->n_tris);
  int32_t id = 0;
  for (uint32_t m = 0; m < d->n_models; ++m) {
    s->prim_base.push_back(id);
    if (s->models[m].type == PT_MODEL_SPHERE) id += 2;
    else {
      for (uint32_t t = 0; t < s->models[m].n_tris; ++t) s->tri_model[s->models[m].first_tri + t] = (int32_t)m;
      id += (int32_t)s->models[m].n_tris;
    }
  }
  s->n_prims = id;
  build_bvh(*s);
  return s;
}

void oracle_scene_destroy(void *p) {
  Scene *s = (Scene *)p;
  if (s) delete s->canon;
  delete s;
}

// ray_cast(...).first(): closest hit of each ray.  brute != 0 bypasses the BVH filter.
int oracle_trace_closest(const void *sp, const float *rays, size_t n, pt_hit *out, int brute, int threads) {
  const Scene &s = *(const Scene *)sp;
  if (threads < 1) threads = 1;
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t)
    pool.emplace_back([&, t]() {
      std::vector<Hit> hits;
      for (size_t i = t; i < n; i += threads) {
        const float *r = rays + i * 6;
        ray_cast(s, v3(r[0], r[1], r[2]), v3(r[3], r[4], r[5]), hits, brute != 0);
        if (hits.empty()) out[i] = pt_hit{-1, 0, 0, 0};
        else out[i] = pt_hit{hits[0].prim_id, hits[0].dist, hits[0].u, hits[0].v};
      }
    });
  for (auto &t : pool) t.join();
  return 0;
}

// The hit render_pixel shades after its alpha walk (mod.rs:188-205), for ray i with the Philox draws of path
// (pixel = i, sample, seed) at `bounce`: the checker of pt_trace_accepted.
int oracle_trace_accepted(const void *sp, const float *rays, size_t n, uint32_t bounce, uint32_t sample, uint64_t seed,
                          pt_hit *out, int brute, int threads) {
  const Scene &s = *(const Scene *)sp;
  if (threads < 1) threads = 1;
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t)
    pool.emplace_back([&, t]() {
      std::vector<Hit> hits;
      for (size_t i = t; i < n; i += threads) {
        const float *r = rays + i * 6;
        ray_cast(s, v3(r[0], r[1], r[2]), v3(r[3], r[4], r[5]), hits, brute != 0);
        if (hits.empty()) {
          out[i] = pt_hit{-1, 0, 0, 0};
          continue;
        }
        PathRng rng;
        rng.init_philox((uint32_t)i, sample, seed);
        const Hit *hit = nullptr;
        for (const Hit &h : hits) {
          hit = &h;
          float op = hit_material_sample(s, h, h.model).opacity;
          if (op >= 1.0f) break;
          if (op > 0.001f) {
            if (rng.alpha(bounce, (uint32_t)h.prim_id) < op) break;
          }
        }
        out[i] = pt_hit{hit->prim_id, hit->dist, hit->u, hit->v};
      }
    });
  for (auto &t : pool) t.join();
  return 0;
}

// Shadow walk of get_light_info (mod.rs:281-333) with unit light colour: the checker of pt_trace_shadow.
// limits == nullptr: directional (every hit, occluder sampled at its own uv).  Else point light: limits[i] = distance
// of the light from the primary hit, primary = n*5 floats (position, uv of the primary hit; u NaN = sphere).
int oracle_trace_shadow(const void *sp, const float *rays, const float *limits, const float *primary, size_t n, float *out,
                        int brute, int threads) {
  const Scene &s = *(const Scene *)sp;
  if (threads < 1) threads = 1;
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t)
    pool.emplace_back([&, t]() {
      std::vector<Hit> hits;
      for (size_t i = t; i < n; i += threads) {
        const float *r = rays + i * 6;
        ray_cast(s, v3(r[0], r[1], r[2]), v3(r[3], r[4], r[5]), hits, brute != 0);
        float tr = 1.0f;
        if (!limits) {
          for (const Hit &h : hits) {
            tr *= 1.0f - hit_material_sample(s, h, h.model).opacity;
            if (tr == 0.0f) break;
          }
        } else {
          const float *p = primary + i * 5;
          Hit prim{};
          prim.position = v3(p[0], p[1], p[2]);
          prim.is_sphere = !(p[3] == p[3]);
          prim.tex_coords = {p[3], p[4]};
          for (const Hit &h : hits) {
            if (magnitude(h.position - prim.position) > limits[i]) break;  // mod.rs:320-323
            tr *= 1.0f - hit_material_sample(s, prim, h.model).opacity;     // mod.rs:324
            if (tr == 0.0f) break;
          }
        }
        out[i] = tr;
      }
    });
  for (auto &t : pool) t.join();
  return 0;
}

// Philox4x32-10 on one (counter, key): the known-answer hook of tests/test_oracle_kat.py (Random123 vectors).
void oracle_philox4x32_10(const uint32_t *ctr4, const uint32_t *key2, uint32_t *out4) {
  uint32_t c[4] = {ctr4[0], ctr4[1], ctr4[2], ctr4[3]};
  philox4x32_10(c, key2[0], key2[1]);
  for (int i = 0; i < 4; ++i) out4[i] = c[i];
}

// Full sorted hit list of ONE ray (debug / tie analysis).  Returns the number of hits.
int oracle_trace_all(const void *sp, const float *ray6, pt_hit *out, int max_out, int brute) {
  const Scene &s = *(const Scene *)sp;
  std::vector<Hit> hits;
  ray_cast(s, v3(ray6[0], ray6[1], ray6[2]), v3(ray6[3], ray6[4], ray6[5]), hits, brute != 0);
  for (int i = 0; i < (int)hits.size() && i < max_out; ++i) out[i] = pt_hit{hits[i].prim_id, hits[i].dist, hits[i].u, hits[i].v};
  return (int)hits.size();
}

// Directional shadow transmittance (renderer/mod.rs:291-297) with unit light colour.
int oracle_trace_transmittance(const void *sp, const float *rays, size_t n, float *out, int brute) {
  const Scene &s = *(const Scene *)sp;
  std::vector<Hit> hits;
  for (size_t i = 0; i < n; ++i) {
    const float *r = rays + i * 6;
    ray_cast(s, v3(r[0], r[1], r[2]), v3(r[3], r[4], r[5]), hits, brute != 0);
    float tr = 1.0f;
    for (const Hit &h : hits) {
      tr *= 1.0f - hit_material_sample(s, h, h.model).opacity;
      if (tr == 0.0f) break;
    }
    out[i] = tr;
  }
  return 0;
}

// Renderer::render.  Adds samples {sample_begin + k*sample_stride, k<n_samples} (0-based sample
// index s; the reference's current_sample is s+1) into accum (W*H*3, caller-zeroed).
// rng_mode 0 = ChaCha12 seeded (s+1) + i*profile.samples (mod.rs:110-112), 1 = Philox.
// counters_out (optional): [extend_rays, shadow_rays].  Returns elapsed milliseconds.
double oracle_render_accum(const void *sp, const pt_profile *prof, int rng_mode, uint64_t seed, uint32_t sample_begin,
                           uint32_t sample_stride, uint32_t n_samples, int threads, int brute, float *accum,
                           uint64_t *counters_out) {
  const Scene &s = *(const Scene *)sp;
  if (threads < 1) threads = 1;
  const uint32_t W = prof->width, H = prof->height;
  const uint64_t npix = (uint64_t)W * H;
  std::vector<Counters> cnts(threads);
  auto t0 = std::chrono::steady_clock::now();
  for (uint32_t k = 0; k < n_samples; ++k) {  // one fork-join per sample pass, like mod.rs:105-106
    const uint32_t sidx = sample_begin + k * sample_stride;
    std::atomic<uint64_t> next{0};
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t)
      pool.emplace_back([&, t]() {
        std::vector<Hit> hits, tmp;
        PathRng rng;
        const uint64_t CH = 1024;
        for (;;) {
          uint64_t b = next.fetch_add(CH);
          if (b >= npix) break;
          uint64_t e = std::min(npix, b + CH);
          for (uint64_t i = b; i < e; ++i) {
            uint32_t x = (uint32_t)(i % W), y = (uint32_t)(i / W);
            if (rng_mode == RNG_CHACHA) rng.init_chacha((uint64_t)(sidx + 1) + i * (uint64_t)prof->samples);
            else rng.init_philox((uint32_t)i, sidx, seed);
            V3 o, d;
            raygen(s, *prof, x, y, rng, o, d);
            V3 c = render_pixel(s, *prof, o, d, rng, brute != 0, hits, tmp, cnts[t]);
            accum[i * 3 + 0] += c.x;
            accum[i * 3 + 1] += c.y;
            accum[i * 3 + 2] += c.z;
          }
        }
      });
    for (auto &t : pool) t.join();
  }
  auto t1 = std::chrono::steady_clock::now();
  if (counters_out) {
    counters_out[0] = counters_out[1] = 0;
    for (auto &c : cnts) { counters_out[0] += c.extend; counters_out[1] += c.shadow; }
  }
  return std::chrono::duration<double, std::milli>(t1 - t0).count();
}

// Canonical per-ray work of a profile: traces sample 0 of every `pixel_stride`-th pixel and counts
// the canonical BVH2 traversal of every path segment (closest hit) and every shadow ray (any hit).
// out6 = [rays_closest, nodes_closest, prims_closest, rays_any, nodes_any, prims_any].
int oracle_canonical_counts(void *sp, const pt_profile *prof, uint32_t pixel_stride, uint64_t seed, int threads,
                            uint64_t *out6) {
  Scene &s = *(Scene *)sp;
  if (!s.canon) {
    s.canon = new CanonBvh();
    s.canon->build(s);
  }
  if (threads < 1) threads = 1;
  if (pixel_stride < 1) pixel_stride = 1;
  const uint64_t npix = (uint64_t)prof->width * prof->height;
  std::vector<Counters> cnts(threads);
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t)
    pool.emplace_back([&, t]() {
      std::vector<Hit> hits, tmp;
      cnts[t].canon_on = true;
      for (uint64_t i = (uint64_t)t * pixel_stride; i < npix; i += (uint64_t)threads * pixel_stride) {
        PathRng rng;
        rng.init_philox((uint32_t)i, 0, seed);
        V3 o, d;
        raygen(s, *prof, (uint32_t)(i % prof->width), (uint32_t)(i / prof->width), rng, o, d);
        render_pixel(s, *prof, o, d, rng, false, hits, tmp, cnts[t]);
      }
    });
  for (auto &t : pool) t.join();
  for (int k = 0; k < 6; ++k) out6[k] = 0;
  for (auto &c : cnts) {
    out6[0] += c.canon.rays_closest; out6[1] += c.canon.nodes_closest; out6[2] += c.canon.prims_closest;
    out6[3] += c.canon.rays_any; out6[4] += c.canon.nodes_any; out6[5] += c.canon.prims_any;
  }
  return 0;
}

// render_pixel with a per-bounce record (Philox streams), the checker of pt_trace_paths.
int oracle_trace_paths(const void *sp, const pt_profile *prof, const uint32_t *pixels, size_t n_paths, uint32_t sample,
                       uint64_t seed, int brute, pt_path_vertex *out, uint32_t *counts) {
  const Scene &s = *(const Scene *)sp;
  std::vector<Hit> hits, tmp;
  Counters cnt;
  std::memset(out, 0, n_paths * (prof->bounces + 1) * sizeof(pt_path_vertex));
  for (size_t i = 0; i < n_paths; ++i) {
    PathRng rng;
    rng.init_philox(pixels[i], sample, seed);
    V3 o, d;
    raygen(s, *prof, pixels[i] % prof->width, pixels[i] / prof->width, rng, o, d);
    Recorder rec;
    rec.out = out + i * (prof->bounces + 1);
    render_pixel(s, *prof, o, d, rng, brute != 0, hits, tmp, cnt, &rec);
    counts[i] = rec.n;
  }
  return 0;
}

// Final pass of Renderer::render (mod.rs:150-162) + post_processing (:335-353).
int oracle_tonemap(const float *accum, uint32_t width, uint32_t height, uint32_t total_samples, int32_t tonemap_type,
                   uint8_t *rgb8) {
  const float n = (float)total_samples;
  for (uint64_t i = 0; i < (uint64_t)width * height; ++i) {
    V3 c = v3(accum[i * 3], accum[i * 3 + 1], accum[i * 3 + 2]) / n;
    post_process(tonemap_type, c, rgb8 + i * 3);
  }
  return 0;
}

// debug_render + render_debug_pixels, renderer/debug_renderer.rs:11-105: one pixel-centre ray per pixel, the first
// element of ray_cast, seven AOV planes in the order normal, albedo, opacity, metalness, roughness, emissive, ior.
// planes: 7 * W*H*3 bytes, row-major (the reference's x*H+y buffer index is used for both the write and the read,
// debug_renderer.rs:38,52, so image[(x,y)] is simply pixel (x,y)); pixels without a hit stay 0.  Returns the number of
// pixels that hit something (the reference writes no file at all when it is 0).
int oracle_debug_textures(const void *sp, uint32_t W, uint32_t H, uint8_t *planes) {
  const Scene &s = *(const Scene *)sp;
  const size_t plane = (size_t)W * H * 3;
  std::memset(planes, 0, plane * 7);
  std::vector<Hit> hits;
  const float wf = (float)W, hf = (float)H, ratio = wf / hf;
  int n_hit = 0;
  auto put = [&](int k, size_t px, V3 c) {  // `(pixel.x * 255.) as u8` (debug_renderer.rs:53-57)
    uint8_t *o = planes + plane * k + px * 3;
    const float v[3] = {c.x * 255.0f, c.y * 255.0f, c.z * 255.0f};
    for (int i = 0; i < 3; ++i) o[i] = !(v[i] == v[i]) || v[i] <= 0.0f ? 0 : v[i] >= 255.0f ? 255 : (uint8_t)v[i];
  };
  for (uint32_t y = 0; y < H; ++y)
    for (uint32_t x = 0; x < W; ++x) {
      float sx = (float)x + 0.5f;
      sx = sx / wf * 2.0f - 1.0f;
      sx *= std::tan(s.fov / 2.0f) * ratio;
      float sy = (float)y + 0.5f;
      sy = 1.0f - sy / hf * 2.0f;
      sy *= std::tan(s.fov / 2.0f);
      V3 dir = normalize(v3(sx, sy, -1.0f));
      const float *m = s.camera;
      V3 c0 = v3(m[0], m[1], m[2]), c1 = v3(m[4], m[5], m[6]), c2 = v3(m[8], m[9], m[10]), c3 = v3(m[12], m[13], m[14]);
      V3 d = ((c0 * dir.x + c1 * dir.y) + c2 * dir.z) + c3 * 0.0f;
      ray_cast(s, c3, d, hits, false);
      if (hits.empty()) continue;
      ++n_hit;
      const Hit &h = hits[0];
      MaterialSample mat = hit_material_sample(s, h, h.model);
      V3 n = hit_shading_normal(s, h, h.model);
      const size_t px = (size_t)y * W + x;
      const V3 one = v3(1.0f, 1.0f, 1.0f);
      put(0, px, n * 0.5f + v3(0.5f, 0.5f, 0.5f));
      put(1, px, mat.albedo);
      put(2, px, one * mat.opacity);
      put(3, px, one * mat.metalness);
      put(4, px, one * mat.roughness);
      put(5, px, mat.emissive);
      put(6, px, one * mat.ior / 3.0f);
    }
  return n_hit;
}

// Dumps the bounce-`bounce` rays (origin, direction) of a 1-spp Philox trace: the secondary-ray
// batches of SURVEY.md §8(c).  rays_out: W*H*6 floats; valid_out: W*H bytes.
int oracle_dump_bounce_rays(const void *sp, const pt_profile *prof, uint64_t seed, uint32_t bounce_wanted,
                            float *rays_out, uint8_t *valid_out);
}

// The shadow requests of get_light_info (mod.rs:281-333) toward light `light` from the primary hits (first element of
// ray_cast, no alpha walk) of a 1-spp Philox trace: ray (origin = hit + normal * bias, direction = -light dir),
// limit (point: |hit - light|, directional: -1), primary (position, uv; u NaN for a sphere hit).
extern "C" int oracle_dump_shadow_requests(const void *sp, const pt_profile *prof, uint64_t seed, uint32_t light,
                                           float *rays_out, float *limits_out, float *primary_out, uint8_t *valid_out) {
  const Scene &s = *(const Scene *)sp;
  const uint32_t W = prof->width, H = prof->height;
  std::vector<Hit> hits;
  for (uint64_t i = 0; i < (uint64_t)W * H; ++i) {
    valid_out[i] = 0;
    if (light >= s.lights.size()) continue;
    PathRng rng;
    rng.init_philox((uint32_t)i, 0, seed);
    V3 o, d;
    raygen(s, *prof, (uint32_t)(i % W), (uint32_t)(i / W), rng, o, d);
    ray_cast(s, o, d, hits, false);
    if (hits.empty()) continue;
    const Hit &h = hits[0];
    const pt_light &L = s.lights[light];
    V3 ori = h.position + h.normal * NORMAL_BIAS;
    V3 sdir;
    float limit = -1.0f;
    if (L.type == PT_LIGHT_DIRECTIONAL) {
      sdir = -1.0f * v3(L.vec[0], L.vec[1], L.vec[2]);
    } else {
      V3 dir = h.position - v3(L.vec[0], L.vec[1], L.vec[2]);
      limit = magnitude(dir);
      sdir = -1.0f * normalize(dir);
    }
    float r[6] = {ori.x, ori.y, ori.z, sdir.x, sdir.y, sdir.z};
    std::memcpy(rays_out + i * 6, r, sizeof r);
    limits_out[i] = limit;
    float p[5] = {h.position.x, h.position.y, h.position.z, h.is_sphere ? NAN : h.tex_coords.x, h.tex_coords.y};
    std::memcpy(primary_out + i * 5, p, sizeof p);
    valid_out[i] = 1;
  }
  return 0;
}

// Implemented separately to keep render_pixel identical to the reference's control flow.
int oracle_dump_bounce_rays(const void *sp, const pt_profile *prof, uint64_t seed, uint32_t bounce_wanted,
                            float *rays_out, uint8_t *valid_out) {
  const Scene &s = *(const Scene *)sp;
  const uint32_t W = prof->width, H = prof->height;
  const int threads = (int)std::max(1u, std::thread::hardware_concurrency());
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t)
    pool.emplace_back([&, t]() {
      std::vector<Hit> hits;
      for (uint64_t i = t; i < (uint64_t)W * H; i += threads) {
        PathRng rng;
        rng.init_philox((uint32_t)i, 0, seed);
        V3 o, d;
        raygen(s, *prof, (uint32_t)(i % W), (uint32_t)(i / W), rng, o, d);
        valid_out[i] = 0;
        for (uint32_t bounce = 0; bounce <= bounce_wanted; ++bounce) {
          if (bounce == bounce_wanted) {
            float r[6] = {o.x, o.y, o.z, d.x, d.y, d.z};
            std::memcpy(rays_out + i * 6, r, sizeof r);
            valid_out[i] = 1;
            break;
          }
          ray_cast(s, o, d, hits, false);
          if (hits.empty()) break;
          const Hit &h = hits[0];
          MaterialSample mat = hit_material_sample(s, h, h.model);
          V3 n = hit_shading_normal(s, h, h.model);
          CookTorrance brdf(mat);
          V3 view = -1.0f * d;
          V3 nd = brdf.sample(n, view, rng.bsdf(bounce, 0), rng.bsdf(bounce, 1));
          o = h.position + h.normal * NORMAL_BIAS;
          d = nd;
          if (!(d.x == d.x && d.y == d.y && d.z == d.z)) break;
        }
      }
    });
  for (auto &t : pool) t.join();
  return 0;
}
