// scene_host.cpp — flat host scene bookkeeping (Scene::load, reference scene/internal/mod.rs:35-51).
#include "scene_host.hpp"

#include <cstring>

namespace ptb {

int HostScene::add_texture(uint32_t w, uint32_t h, uint32_t ch, std::vector<uint8_t> &&data) {
  texdata.push_back(std::move(data));
  pt_texture t{w, h, ch, nullptr};
  textures.push_back(t);
  return (int)textures.size() - 1;
}

void HostScene::from_desc(const pt_scene_desc &d) {
  models.assign(d.models, d.models + d.n_models);
  materials.assign(d.materials, d.materials + d.n_models);
  pos.assign(d.tri_pos, d.tri_pos + d.n_tris * 9);
  nrm.assign(d.tri_nrm, d.tri_nrm + d.n_tris * 9);
  uv.assign(d.tri_uv, d.tri_uv + d.n_tris * 6);
  textures.clear();
  texdata.clear();
  for (uint32_t i = 0; i < d.n_textures; ++i) {
    const pt_texture &t = d.textures[i];
    std::vector<uint8_t> v(t.data, t.data + (size_t)t.width * t.height * t.channels);
    add_texture(t.width, t.height, t.channels, std::move(v));
  }
  lights.assign(d.lights, d.lights + d.n_lights);
  std::memcpy(camera, d.camera, sizeof camera);
  fov = d.fov;
  std::memcpy(background, d.background, sizeof background);
}

void HostScene::to_desc(pt_scene_desc &d) const {
  d.n_models = (uint32_t)models.size();
  d.models = models.data();
  d.materials = materials.data();
  d.n_tris = n_tris();
  d.tri_pos = pos.data();
  d.tri_nrm = nrm.data();
  d.tri_uv = uv.data();
  d.n_textures = (uint32_t)textures.size();
  d.textures = textures.data();
  d.n_lights = (uint32_t)lights.size();
  d.lights = lights.data();
  std::memcpy(d.camera, camera, sizeof camera);
  d.fov = fov;
  std::memcpy(d.background, background, sizeof background);
}

std::string HostScene::finalize() {
  if (models.size() != materials.size()) return "materials must have one entry per model";
  if (models.empty()) return "scene has no models";  // kdtree-ray asserts a non-empty shape list
  for (size_t i = 0; i < textures.size(); ++i) {
    textures[i].data = texdata[i].data();
    if (textures[i].width == 0 || textures[i].height == 0) return "texture with zero size";
    if (textures[i].channels != 1 && textures[i].channels != 3) return "texture channels must be 1 or 3";
  }
  const uint64_t nt = n_tris();
  if (nrm.size() != nt * 9 || uv.size() != nt * 6) return "triangle arrays have inconsistent sizes";
  if (nt >= (1ull << 27)) return "too many triangles (limit 2^27)";
  tri_model.assign(nt, -1);
  prim_base.clear();
  sphere_models.clear();
  int64_t id = 0;
  for (size_t m = 0; m < models.size(); ++m) {
    const pt_model &mo = models[m];
    prim_base.push_back((int32_t)id);
    if (mo.type == PT_MODEL_SPHERE) {
      sphere_models.push_back((int32_t)m);
      id += 2;
    } else if (mo.type == PT_MODEL_MESH) {
      if (mo.n_tris == 0) return "mesh model with no triangles";  // kdtree-ray build panics
      if ((uint64_t)mo.first_tri + mo.n_tris > nt) return "model triangle range out of bounds";
      for (uint32_t t = 0; t < mo.n_tris; ++t) {
        if (tri_model[mo.first_tri + t] != -1) return "overlapping model triangle ranges";
        tri_model[mo.first_tri + t] = (int32_t)m;
      }
      id += mo.n_tris;
    } else {
      return "unknown model type";
    }
    const pt_material &ma = materials[m];
    const int32_t slots[6] = {ma.albedo_tex, ma.emissive_tex, ma.opacity_tex, ma.metalness_tex, ma.roughness_tex, ma.normal_tex};
    const uint32_t want[6] = {3, 3, 1, 1, 1, 3};
    for (int k = 0; k < 6; ++k) {
      if (slots[k] < 0) continue;
      if ((size_t)slots[k] >= textures.size()) return "material texture index out of range";
      if (textures[slots[k]].channels != want[k]) return "material texture has the wrong channel count";
    }
  }
  for (uint64_t t = 0; t < nt; ++t)
    if (tri_model[t] < 0) return "triangle not owned by any model";
  if (id >= (1ll << 31)) return "too many primitives";
  n_prims = id;
  for (const pt_light &l : lights)
    if (l.type != PT_LIGHT_POINT && l.type != PT_LIGHT_DIRECTIONAL) return "unknown light type";
  return "";
}

}  // namespace ptb
