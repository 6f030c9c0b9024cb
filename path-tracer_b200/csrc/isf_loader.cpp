// isf_loader.cpp — scene.isf reader.  Mirrors load_internal (reference scene/mod.rs:16-22), the serde schema
// and defaults of scene/isf.rs:5-142, Scene::load (scene/internal/mod.rs:35-51) and the path-keyed
// TextureBank (scene/internal/texture_bank.rs:21-51: RGB textures via into_rgb8, gray via into_luma8).
#include <climits>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <map>
#include <sstream>

#include "host_api.hpp"
#include "json.hpp"

namespace ptb {
namespace {

struct Fail {
  std::string msg;
};

const Json &field(const Json &o, const char *key, const char *ctx) {
  const Json *v = o.get(key);
  if (!v) throw Fail{std::string("missing field `") + key + "` in " + ctx};
  return *v;
}
float f32(const Json &v, const char *ctx) {
  if (v.kind != Json::Number) throw Fail{std::string("invalid type: expected f32 in ") + ctx};
  return (float)v.num;  // decimal -> f64 -> f32, like serde_json
}
template <int N>
void farray(const Json &v, float *out, const char *ctx) {
  if (v.kind != Json::Array || (int)v.arr.size() != N)
    throw Fail{std::string("invalid length: expected an array of length ") + std::to_string(N) + " in " + ctx};
  for (int i = 0; i < N; ++i) out[i] = f32(v.arr[i], ctx);
}
const char *opt_string(const Json &o, const char *key) {
  const Json *v = o.get(key);
  if (!v || v->is_null()) return nullptr;
  if (v->kind != Json::String) throw Fail{std::string("invalid type: expected a string for `") + key + "`"};
  return v->str.c_str();
}

struct TextureBank {
  std::string root;
  HostScene &scene;
  std::map<std::string, int> rgb, gray;
  int get(const char *rel, uint32_t channels) {
    if (!rel) return -1;
    std::string joined = rel[0] == '/' ? std::string(rel) : root + "/" + rel;
    char real[PATH_MAX];
    if (!realpath(joined.c_str(), real)) throw Fail{"Invalid path: texture `" + joined + "` not found"};
    auto &cache = channels == 3 ? rgb : gray;
    auto it = cache.find(real);
    if (it != cache.end()) return it->second;
    uint32_t w = 0, h = 0;
    std::vector<uint8_t> data;
    if (png_read(real, channels, &w, &h, data) != PT_OK) throw Fail{"cannot decode texture `" + std::string(real) + "`"};
    int id = scene.add_texture(w, h, channels, std::move(data));
    cache[real] = id;
    return id;
  }
};

pt_material parse_material(const Json &m, TextureBank &bank) {
  if (m.kind != Json::Object) throw Fail{"invalid type: material must be a map"};
  pt_material r{};
  const Json &alb = field(m, "albedo", "Material");  // required, isf.rs:81
  r.albedo[0] = r.albedo[1] = r.albedo[2] = 1.0f;    // #[serde(default = "one")]
  if (const Json *f = alb.get("factor")) farray<3>(*f, r.albedo, "albedo.factor");
  r.albedo_tex = bank.get(opt_string(alb, "texture"), 3);
  r.emissive_tex = -1;
  if (const Json *em = m.get("emissive")) {  // present: factor defaults to one(); absent: derived Default = 0
    r.emissive[0] = r.emissive[1] = r.emissive[2] = 1.0f;
    if (const Json *f = em->get("factor")) farray<3>(*f, r.emissive, "emissive.factor");
    r.emissive_tex = bank.get(opt_string(*em, "texture"), 3);
  }
  r.opacity = 1.0f;
  r.opacity_tex = -1;
  if (const Json *op = m.get("opacity")) {
    if (const Json *f = op->get("factor")) r.opacity = f32(*f, "opacity.factor");
    r.opacity_tex = bank.get(opt_string(*op, "texture"), 1);
  }
  r.metalness = 0.0f;  // derived Default
  r.metalness_tex = -1;
  if (const Json *me = m.get("metalness")) {
    r.metalness = 1.0f;  // #[serde(default = "One::one")]
    if (const Json *f = me->get("factor")) r.metalness = f32(*f, "metalness.factor");
    r.metalness_tex = bank.get(opt_string(*me, "texture"), 1);
  }
  r.roughness = 1.0f;
  r.roughness_tex = -1;
  if (const Json *ro = m.get("roughness")) {
    if (const Json *f = ro->get("factor")) r.roughness = f32(*f, "roughness.factor");
    r.roughness_tex = bank.get(opt_string(*ro, "texture"), 1);
  }
  r.ior = 1.0f;
  if (const Json *i = m.get("ior")) r.ior = f32(*i, "ior");
  r.normal_tex = bank.get(opt_string(m, "normal_texture"), 3);
  return r;
}

void parse_scene(const Json &doc, const std::string &root, HostScene &out) {
  if (doc.kind != Json::Object) throw Fail{"invalid type: expected struct Scene"};
  TextureBank bank{root, out, {}, {}};
  const Json &models = field(doc, "models", "Scene");
  if (models.kind != Json::Array) throw Fail{"invalid type: `models` must be a sequence"};
  for (const Json &m : models.arr) {
    const Json &type = field(m, "type", "Model");
    if (type.kind != Json::String) throw Fail{"invalid type: model tag must be a string"};
    pt_model mo{};
    if (type.str == "Sphere") {
      mo.type = PT_MODEL_SPHERE;
      mo.radius = f32(field(m, "radius", "Sphere"), "radius");
      farray<3>(field(m, "center", "Sphere"), mo.center, "center");
    } else if (type.str == "Mesh") {
      mo.type = PT_MODEL_MESH;
      const Json &tris = field(m, "triangles", "Mesh");
      if (tris.kind != Json::Array) throw Fail{"invalid type: `triangles` must be a sequence"};
      mo.first_tri = (uint32_t)out.n_tris();
      mo.n_tris = (uint32_t)tris.arr.size();
      for (const Json &t : tris.arr) {
        if (t.kind != Json::Array || t.arr.size() != 3) throw Fail{"invalid length: a triangle is 3 vertices"};
        float p[9], n[9], uv[6];
        for (int k = 0; k < 3; ++k) {
          const Json &v = t.arr[k];
          farray<3>(field(v, "position", "Vertex"), p + k * 3, "position");
          farray<3>(field(v, "normal", "Vertex"), n + k * 3, "normal");
          farray<2>(field(v, "tex_coords", "Vertex"), uv + k * 2, "tex_coords");
        }
        out.pos.insert(out.pos.end(), p, p + 9);
        out.nrm.insert(out.nrm.end(), n, n + 9);
        out.uv.insert(out.uv.end(), uv, uv + 6);
      }
    } else {
      throw Fail{"unknown variant `" + type.str + "`, expected `Sphere` or `Mesh`"};
    }
    out.models.push_back(mo);
    out.materials.push_back(parse_material(field(m, "material", "Model"), bank));
  }
  const Json &cam = field(doc, "camera", "Scene");
  const Json &tr = field(cam, "transform", "Camera");
  if (tr.kind != Json::Array || tr.arr.size() != 4) throw Fail{"invalid length: camera transform is 4x4"};
  for (int c = 0; c < 4; ++c) farray<4>(tr.arr[c], out.camera + c * 4, "camera.transform");
  out.fov = f32(field(cam, "fov", "Camera"), "fov");
  (void)f32(field(cam, "zfar", "Camera"), "zfar");
  (void)f32(field(cam, "znear", "Camera"), "znear");
  const Json &lights = field(doc, "lights", "Scene");
  if (lights.kind != Json::Array) throw Fail{"invalid type: `lights` must be a sequence"};
  for (const Json &l : lights.arr) {
    const Json &type = field(l, "type", "Light");
    pt_light li{};
    if (type.kind == Json::String && type.str == "Point") {
      li.type = PT_LIGHT_POINT;
      farray<3>(field(l, "position", "Point"), li.vec, "position");
      farray<3>(field(l, "color", "Point"), li.color, "color");
      li.size = f32(field(l, "size", "Point"), "size");
    } else if (type.kind == Json::String && type.str == "Directional") {
      li.type = PT_LIGHT_DIRECTIONAL;
      farray<3>(field(l, "direction", "Directional"), li.vec, "direction");
      farray<3>(field(l, "color", "Directional"), li.color, "color");
    } else {
      throw Fail{"unknown light variant, expected `Point` or `Directional`"};
    }
    out.lights.push_back(li);
  }
  farray<3>(field(doc, "background", "Scene"), out.background, "background");
}

}  // namespace

int load_isf_file(const char *path, HostScene &out) {
  std::ifstream f(path, std::ios::binary);
  if (!f) {
    set_error("%s: No such file or directory (os error 2)", path);
    return PT_ERR_IO;
  }
  std::stringstream ss;
  ss << f.rdbuf();
  std::string text = ss.str();
  Json doc;
  std::string err;
  if (!json_parse(text.data(), text.size(), doc, err)) {
    set_error("%s", err.c_str());
    return PT_ERR_PARSE;
  }
  std::string p(path);
  size_t slash = p.find_last_of('/');
  std::string root = slash == std::string::npos ? "." : p.substr(0, slash);
  try {
    parse_scene(doc, root, out);
  } catch (const Fail &e) {
    set_error("%s", e.msg.c_str());
    return e.msg.rfind("Invalid path", 0) == 0 || e.msg.rfind("cannot decode", 0) == 0 ? PT_ERR_IO : PT_ERR_PARSE;
  }
  return PT_OK;
}

}  // namespace ptb

extern "C" int pt_scene_load_isf(const char *path, int device, pt_scene **out) {
  if (!path || !out) {
    ptb::set_error("null argument");
    return PT_ERR_INVALID;
  }
  return ptb::abi_guard([&]() -> int {
    ptb::HostScene h;
    if (ptb::is_isfb_file(path)) {  // binary side-car (isf_binary.cpp), told apart by its magic
      ptb::Bvh4 accel;
      int kind = -1;
      int rc = ptb::load_isfb_file(path, h, &accel, &kind);
      if (rc != PT_OK) return rc;
      return ptb::scene_from_host(std::move(h), device, out, kind >= 0 ? &accel : nullptr, kind);
    }
    int rc = ptb::load_isf_file(path, h);
    if (rc != PT_OK) return rc;
    return ptb::scene_from_host(std::move(h), device, out);
  });
}
