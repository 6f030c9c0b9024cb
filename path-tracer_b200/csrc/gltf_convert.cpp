// gltf_convert.cpp — `path-tracer convert <in.gltf|in.glb> <out_dir>`: glTF 2.0 -> isf.
// Mirrors scene::convert_gltf_to_isf (reference src/scene/gltf.rs:146-204) and its From impls (:206-265):
//   * scene 0 only; the first camera met in the node traversal (perspective only, gltf.rs:165-169, :206-221);
//   * one isf Mesh per glTF primitive, vertices in world space (node transforms applied), TEXCOORD_0, triangles only;
//   * material mapping of convert_material (gltf.rs:79-130): base colour rgb -> albedo, base colour alpha -> opacity,
//     metallicRoughness B -> metalness / G -> roughness, emissive, normal map, ior = 1;
//   * texture files named like ReverseTextureBank (gltf.rs:47-77): albedo_tex_N.png (RGB of the base colour image),
//     alpha_tex_N.png (its alpha as 8-bit gray), gray_tex_N.png, vec_tex_N.png, N = how many textures of that
//     family existed before;
//   * KHR_lights_punctual: directional -> Directional, point and spot -> Point with size 0.1 (gltf.rs:236-264),
//     colour * intensity.
// The reference delegates parsing to the easy-gltf crate (^1.1.4, source not vendored): what that crate does is
// restated from the glTF 2.0 specification.  Not on the accelerated path: plain host C++.
// Images: PNG (png.cpp) and baseline JPEG (jpeg.cpp).  Limits (an error, exit code 2): progressive JPEG and other
// image formats, sparse accessors, non-triangle primitive modes, orthographic cameras, scenes without a camera (the
// reference panics there).
#include <sys/stat.h>

#include <cerrno>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <string>
#include <vector>

#include "host_api.hpp"
#include "json.hpp"

namespace ptb {
namespace {

struct Fail {
  std::string msg;
};
[[noreturn]] void fail(const std::string &m) { throw Fail{m}; }

struct Mat4 {  // column-major, m[c][r] like cgmath::Matrix4
  float m[4][4];
};
Mat4 identity() {
  Mat4 r{};
  for (int i = 0; i < 4; ++i) r.m[i][i] = 1.f;
  return r;
}
Mat4 mul(const Mat4 &a, const Mat4 &b) {  // a * b
  Mat4 r{};
  for (int c = 0; c < 4; ++c)
    for (int row = 0; row < 4; ++row) {
      float s = 0.f;
      for (int k = 0; k < 4; ++k) s += a.m[k][row] * b.m[c][k];
      r.m[c][row] = s;
    }
  return r;
}
void xform(const Mat4 &M, const float v[3], float w, float out[3]) {
  for (int r = 0; r < 3; ++r) out[r] = M.m[0][r] * v[0] + M.m[1][r] * v[1] + M.m[2][r] * v[2] + M.m[3][r] * w;
}

std::vector<uint8_t> read_file(const std::string &path) {
  std::ifstream f(path, std::ios::binary);
  if (!f) fail(path + ": cannot open");
  return std::vector<uint8_t>((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
}

std::vector<uint8_t> base64_decode(const std::string &s, size_t from) {
  std::vector<uint8_t> out;
  uint32_t acc = 0;
  int bits = 0;
  for (size_t i = from; i < s.size(); ++i) {
    const char c = s[i];
    int v;
    if (c >= 'A' && c <= 'Z') v = c - 'A';
    else if (c >= 'a' && c <= 'z') v = c - 'a' + 26;
    else if (c >= '0' && c <= '9') v = c - '0' + 52;
    else if (c == '+' || c == '-') v = 62;
    else if (c == '/' || c == '_') v = 63;
    else continue;  // '=', whitespace
    acc = (acc << 6) | (uint32_t)v;
    bits += 6;
    if (bits >= 8) {
      bits -= 8;
      out.push_back((uint8_t)(acc >> bits));
    }
  }
  return out;
}

std::string percent_decode(const std::string &s) {
  std::string o;
  for (size_t i = 0; i < s.size(); ++i) {
    if (s[i] == '%' && i + 2 < s.size() && std::isxdigit((unsigned char)s[i + 1]) && std::isxdigit((unsigned char)s[i + 2])) {
      o.push_back((char)std::strtol(s.substr(i + 1, 2).c_str(), nullptr, 16));
      i += 2;
    } else o.push_back(s[i]);
  }
  return o;
}

double num(const Json *j, double dflt) { return j && j->kind == Json::Number ? j->num : dflt; }
int64_t idx(const Json *j) { return j && j->kind == Json::Number ? (int64_t)j->num : -1; }

struct Doc {
  Json root;
  std::string dir;
  std::vector<std::vector<uint8_t>> buffers;
  std::vector<uint8_t> glb_bin;
  bool has_glb_bin = false;

  const Json &arr(const char *key, int64_t i) const {
    const Json *a = root.get(key);
    if (!a || a->kind != Json::Array || i < 0 || (size_t)i >= a->arr.size()) fail(std::string("glTF: bad index into ") + key);
    return a->arr[(size_t)i];
  }
  size_t count(const char *key) const {
    const Json *a = root.get(key);
    return a && a->kind == Json::Array ? a->arr.size() : 0;
  }
  std::vector<uint8_t> load_uri(const std::string &uri) const {
    if (uri.rfind("data:", 0) == 0) {
      size_t comma = uri.find(',');
      if (comma == std::string::npos) fail("glTF: malformed data URI");
      return base64_decode(uri, comma + 1);
    }
    return read_file(dir + "/" + percent_decode(uri));
  }
  void load_buffers() {
    for (size_t i = 0; i < count("buffers"); ++i) {
      const Json &b = arr("buffers", (int64_t)i);
      const Json *uri = b.get("uri");
      if (uri && uri->kind == Json::String) buffers.push_back(load_uri(uri->str));
      else if (i == 0 && has_glb_bin) buffers.push_back(glb_bin);
      else fail("glTF: buffer without uri");
    }
  }
  // bytes of a bufferView
  void view(int64_t vi, const uint8_t *&p, size_t &len, size_t &stride) const {
    const Json &v = arr("bufferViews", vi);
    int64_t b = idx(v.get("buffer"));
    if (b < 0 || (size_t)b >= buffers.size()) fail("glTF: bufferView.buffer out of range");
    size_t off = (size_t)num(v.get("byteOffset"), 0);
    len = (size_t)num(v.get("byteLength"), 0);
    stride = (size_t)num(v.get("byteStride"), 0);
    if (off + len > buffers[(size_t)b].size()) fail("glTF: bufferView exceeds its buffer");
    p = buffers[(size_t)b].data() + off;
  }
  // accessor -> count x ncomp floats (integers converted; normalized integers mapped per the glTF specification)
  std::vector<float> read_floats(int64_t ai, int want_comp) const {
    const Json &a = arr("accessors", ai);
    if (a.get("sparse")) fail("glTF: sparse accessors are not supported");
    const Json *ty = a.get("type");
    static const std::map<std::string, int> ncomp = {{"SCALAR", 1}, {"VEC2", 2}, {"VEC3", 3}, {"VEC4", 4}, {"MAT4", 16}};
    if (!ty || ty->kind != Json::String || !ncomp.count(ty->str)) fail("glTF: unsupported accessor type");
    const int nc = ncomp.at(ty->str);
    if (want_comp && nc < want_comp) fail("glTF: accessor has too few components");
    const int ct = (int)num(a.get("componentType"), 0);
    const size_t csz = ct == 5120 || ct == 5121 ? 1 : ct == 5122 || ct == 5123 ? 2 : ct == 5125 || ct == 5126 ? 4 : 0;
    if (!csz) fail("glTF: unsupported accessor componentType");
    const size_t count = (size_t)num(a.get("count"), 0);
    const bool normalized = a.get("normalized") && a.get("normalized")->kind == Json::Bool && a.get("normalized")->b;
    const int outc = want_comp ? want_comp : nc;
    std::vector<float> out(count * (size_t)outc, 0.f);
    if (idx(a.get("bufferView")) < 0) return out;  // all zeros per the specification
    const uint8_t *p;
    size_t len, stride;
    view(idx(a.get("bufferView")), p, len, stride);
    const size_t off = (size_t)num(a.get("byteOffset"), 0);
    if (!stride) stride = csz * (size_t)nc;
    if (count && off + (count - 1) * stride + csz * (size_t)nc > len) fail("glTF: accessor exceeds its bufferView");
    for (size_t i = 0; i < count; ++i)
      for (int c = 0; c < outc; ++c) {
        const uint8_t *q = p + off + i * stride + (size_t)c * csz;
        float v;
        switch (ct) {
          case 5120: { int8_t x; std::memcpy(&x, q, 1); v = normalized ? std::fmax((float)x / 127.f, -1.f) : (float)x; break; }
          case 5121: { uint8_t x = *q; v = normalized ? (float)x / 255.f : (float)x; break; }
          case 5122: { int16_t x; std::memcpy(&x, q, 2); v = normalized ? std::fmax((float)x / 32767.f, -1.f) : (float)x; break; }
          case 5123: { uint16_t x; std::memcpy(&x, q, 2); v = normalized ? (float)x / 65535.f : (float)x; break; }
          case 5125: { uint32_t x; std::memcpy(&x, q, 4); v = (float)x; break; }
          default: std::memcpy(&v, q, 4);
        }
        out[i * (size_t)outc + (size_t)c] = v;
      }
    return out;
  }
  std::vector<uint32_t> read_indices(int64_t ai) const {
    const Json &a = arr("accessors", ai);
    const int ct = (int)num(a.get("componentType"), 0);
    if (ct != 5121 && ct != 5123 && ct != 5125) fail("glTF: index accessor must be unsigned");
    std::vector<float> dummy;
    const size_t count = (size_t)num(a.get("count"), 0);
    std::vector<uint32_t> out(count, 0);
    if (idx(a.get("bufferView")) < 0) return out;
    const uint8_t *p;
    size_t len, stride;
    view(idx(a.get("bufferView")), p, len, stride);
    const size_t off = (size_t)num(a.get("byteOffset"), 0), csz = ct == 5121 ? 1 : ct == 5123 ? 2 : 4;
    if (!stride) stride = csz;
    if (count && off + (count - 1) * stride + csz > len) fail("glTF: index accessor exceeds its bufferView");
    for (size_t i = 0; i < count; ++i) {
      const uint8_t *q = p + off + i * stride;
      if (ct == 5121) out[i] = *q;
      else if (ct == 5123) { uint16_t x; std::memcpy(&x, q, 2); out[i] = x; }
      else std::memcpy(&out[i], q, 4);
    }
    return out;
  }
  // encoded bytes of image i
  std::vector<uint8_t> image_bytes(int64_t ii, std::string &name) const {
    const Json &im = arr("images", ii);
    const Json *uri = im.get("uri");
    if (uri && uri->kind == Json::String) {
      name = uri->str.rfind("data:", 0) == 0 ? "image " + std::to_string(ii) : uri->str;
      return load_uri(uri->str);
    }
    name = "image " + std::to_string(ii);
    const uint8_t *p;
    size_t len, stride;
    view(idx(im.get("bufferView")), p, len, stride);
    return std::vector<uint8_t>(p, p + len);
  }
};

Mat4 node_local(const Json &n) {
  Mat4 M = identity();
  if (const Json *m = n.get("matrix")) {
    if (m->kind == Json::Array && m->arr.size() == 16)
      for (int i = 0; i < 16; ++i) M.m[i / 4][i % 4] = (float)m->arr[(size_t)i].num;
    return M;
  }
  float t[3] = {0, 0, 0}, q[4] = {0, 0, 0, 1}, s[3] = {1, 1, 1};
  auto rd = [](const Json *a, float *o, size_t k) {
    if (a && a->kind == Json::Array && a->arr.size() == k)
      for (size_t i = 0; i < k; ++i) o[i] = (float)a->arr[i].num;
  };
  rd(n.get("translation"), t, 3);
  rd(n.get("rotation"), q, 4);
  rd(n.get("scale"), s, 3);
  const float x = q[0], y = q[1], z = q[2], w = q[3];
  const float R[3][3] = {{1 - 2 * (y * y + z * z), 2 * (x * y + z * w), 2 * (x * z - y * w)},   // column 0
                         {2 * (x * y - z * w), 1 - 2 * (x * x + z * z), 2 * (y * z + x * w)},   // column 1
                         {2 * (x * z + y * w), 2 * (y * z - x * w), 1 - 2 * (x * x + y * y)}};  // column 2
  for (int c = 0; c < 3; ++c)
    for (int r = 0; r < 3; ++r) M.m[c][r] = R[c][r] * s[c];
  for (int r = 0; r < 3; ++r) M.m[3][r] = t[r];
  return M;
}

// shortest decimal that parses back to the same f32 (what serde_json's ryu formatting guarantees)
std::string f32s(float v) {
  if (!std::isfinite(v)) return "null";  // serde_json writes null for non-finite floats
  char buf[64];
  for (int p = 1; p <= 9; ++p) {
    std::snprintf(buf, sizeof buf, "%.*g", p, (double)v);
    if (std::strtof(buf, nullptr) == v) break;
  }
  std::string s = buf;
  if (s.find_first_of(".eEn") == std::string::npos) s += ".0";
  return s;
}
std::string vec_s(const float *v, int n) {
  std::string s = "[";
  for (int i = 0; i < n; ++i) s += (i ? "," : "") + f32s(v[i]);
  return s + "]";
}
std::string json_str(const std::string &v) {
  std::string s = "\"";
  for (char c : v) {
    if (c == '"' || c == '\\') { s += '\\'; s += c; }
    else if ((unsigned char)c < 0x20) { char b[8]; std::snprintf(b, sizeof b, "\\u%04x", c); s += b; }
    else s += c;
  }
  return s + "\"";
}

struct TextureBank {  // ReverseTextureBank, gltf.rs:18-77: one counter per texture family, images deduplicated
  std::string out_dir;
  const Doc *doc;
  std::map<int64_t, std::string> albedo, alpha, rgb;
  std::map<std::pair<int64_t, int>, std::string> gray;  // (image, channel)
  std::map<int64_t, std::vector<uint8_t>> rgba_cache;
  std::map<int64_t, std::pair<uint32_t, uint32_t>> dims;

  const std::vector<uint8_t> &rgba(int64_t image) {
    auto it = rgba_cache.find(image);
    if (it != rgba_cache.end()) return it->second;
    std::string name;
    std::vector<uint8_t> bytes = doc->image_bytes(image, name);
    uint32_t w = 0, h = 0;
    std::vector<uint8_t> px;
    if (image_decode(bytes.data(), bytes.size(), name.c_str(), 4, &w, &h, px) != PT_OK) fail(pt_last_error());
    dims[image] = {w, h};
    return rgba_cache[image] = std::move(px);
  }
  int64_t image_of(const Json *texinfo) const {
    if (!texinfo || texinfo->kind != Json::Object) return -1;
    int64_t t = idx(texinfo->get("index"));
    if (t < 0) return -1;
    return idx(doc->arr("textures", t).get("source"));
  }
  void save(const std::string &file, int64_t image, int mode /* 0 rgb, 1 alpha, 2 green, 3 blue */) {
    const std::vector<uint8_t> &px = rgba(image);
    const uint32_t w = dims[image].first, h = dims[image].second;
    std::vector<uint8_t> o;
    const size_t n = (size_t)w * h;
    if (mode == 0) {
      o.resize(n * 3);
      for (size_t i = 0; i < n; ++i) std::memcpy(&o[i * 3], &px[i * 4], 3);
    } else {
      const int ch = mode == 1 ? 3 : mode == 2 ? 1 : 2;
      o.resize(n);
      for (size_t i = 0; i < n; ++i) o[i] = px[i * 4 + (size_t)ch];
    }
    if (png_write_ch((out_dir + "/" + file).c_str(), o.data(), w, h, mode == 0 ? 3 : 1) != PT_OK) fail(pt_last_error());
  }
  std::string albedo_path(int64_t image) {
    auto it = albedo.find(image);
    if (it != albedo.end()) return it->second;
    std::string f = "albedo_tex_" + std::to_string(albedo.size()) + ".png";
    save(f, image, 0);
    return albedo[image] = f;
  }
  std::string alpha_path(int64_t image) {
    auto it = alpha.find(image);
    if (it != alpha.end()) return it->second;
    std::string f = "alpha_tex_" + std::to_string(alpha.size()) + ".png";
    save(f, image, 1);
    return alpha[image] = f;
  }
  std::string rgb_path(int64_t image) {
    auto it = rgb.find(image);
    if (it != rgb.end()) return it->second;
    std::string f = "vec_tex_" + std::to_string(rgb.size()) + ".png";
    save(f, image, 0);
    return rgb[image] = f;
  }
  std::string gray_path(int64_t image, int channel_mode) {
    auto key = std::make_pair(image, channel_mode);
    auto it = gray.find(key);
    if (it != gray.end()) return it->second;
    std::string f = "gray_tex_" + std::to_string(gray.size()) + ".png";
    save(f, image, channel_mode);
    return gray[key] = f;
  }
};

std::string tex_field(const std::string &path) { return path.empty() ? "null" : json_str(path); }

// convert_material, gltf.rs:79-130
std::string material_json(const Doc &d, int64_t mi, TextureBank &bank) {
  float base[4] = {1, 1, 1, 1}, emissive[3] = {0, 0, 0}, metal = 1.f, rough = 1.f;
  std::string albedo_tex, alpha_tex, emissive_tex, metal_tex, rough_tex, normal_tex;
  if (mi >= 0) {
    const Json &m = d.arr("materials", mi);
    auto rd = [](const Json *a, float *o, size_t k) {
      if (a && a->kind == Json::Array && a->arr.size() == k)
        for (size_t i = 0; i < k; ++i) o[i] = (float)a->arr[i].num;
    };
    if (const Json *pbr = m.get("pbrMetallicRoughness")) {
      rd(pbr->get("baseColorFactor"), base, 4);
      metal = (float)num(pbr->get("metallicFactor"), 1.0);
      rough = (float)num(pbr->get("roughnessFactor"), 1.0);
      int64_t bi = bank.image_of(pbr->get("baseColorTexture"));
      if (bi >= 0) {
        albedo_tex = bank.albedo_path(bi);
        alpha_tex = bank.alpha_path(bi);
      }
      int64_t mr = bank.image_of(pbr->get("metallicRoughnessTexture"));
      if (mr >= 0) {
        metal_tex = bank.gray_path(mr, 3);  // metalness lives in B
        rough_tex = bank.gray_path(mr, 2);  // roughness in G
      }
    }
    rd(m.get("emissiveFactor"), emissive, 3);
    int64_t ei = bank.image_of(m.get("emissiveTexture"));
    if (ei >= 0) emissive_tex = bank.rgb_path(ei);
    int64_t ni = bank.image_of(m.get("normalTexture"));
    if (ni >= 0) normal_tex = bank.rgb_path(ni);
  }
  std::string s = "{\"albedo\":{\"factor\":" + vec_s(base, 3) + ",\"texture\":" + tex_field(albedo_tex) + "},";
  s += "\"emissive\":{\"factor\":" + vec_s(emissive, 3) + ",\"texture\":" + tex_field(emissive_tex) + "},";
  s += "\"opacity\":{\"factor\":" + f32s(base[3]) + ",\"texture\":" + tex_field(alpha_tex) + "},";
  s += "\"metalness\":{\"factor\":" + f32s(metal) + ",\"texture\":" + tex_field(metal_tex) + "},";
  s += "\"roughness\":{\"factor\":" + f32s(rough) + ",\"texture\":" + tex_field(rough_tex) + "},";
  s += "\"ior\":1.0,\"normal_texture\":" + tex_field(normal_tex) + "}";
  return s;
}

struct Output {
  std::string models, lights, camera;
  bool have_camera = false;
  size_t n_models = 0, n_lights = 0;
};

void emit_primitive(const Doc &d, const Json &prim, const Mat4 &M, TextureBank &bank, Output &out) {
  const int mode = (int)num(prim.get("mode"), 4);
  if (mode != 4) fail("glTF: only TRIANGLES primitives are supported (mode " + std::to_string(mode) + ")");
  const Json *attrs = prim.get("attributes");
  if (!attrs || idx(attrs->get("POSITION")) < 0) fail("glTF: primitive without POSITION");
  std::vector<float> pos = d.read_floats(idx(attrs->get("POSITION")), 3);
  const size_t nv = pos.size() / 3;
  std::vector<float> nrm(nv * 3, 0.f), uv(nv * 2, 0.f);  // easy-gltf leaves missing attributes at zero
  if (idx(attrs->get("NORMAL")) >= 0) {
    nrm = d.read_floats(idx(attrs->get("NORMAL")), 3);
    if (nrm.size() != nv * 3) fail("glTF: NORMAL count differs from POSITION");
  }
  if (idx(attrs->get("TEXCOORD_0")) >= 0) {
    uv = d.read_floats(idx(attrs->get("TEXCOORD_0")), 2);
    if (uv.size() != nv * 2) fail("glTF: TEXCOORD_0 count differs from POSITION");
  }
  std::vector<uint32_t> indices;
  if (idx(prim.get("indices")) >= 0) indices = d.read_indices(idx(prim.get("indices")));
  else {
    indices.resize(nv);
    for (size_t i = 0; i < nv; ++i) indices[i] = (uint32_t)i;
  }
  for (float &x : std::vector<float>()) (void)x;
  // world space: positions as points, normals as vectors (w = 0), not re-normalised
  for (size_t i = 0; i < nv; ++i) {
    float p[3], n[3];
    xform(M, &pos[i * 3], 1.f, p);
    xform(M, &nrm[i * 3], 0.f, n);
    std::memcpy(&pos[i * 3], p, 12);
    std::memcpy(&nrm[i * 3], n, 12);
  }
  std::string s = "{\"type\":\"Mesh\",\"triangles\":[";
  const size_t ntri = indices.size() / 3;
  s.reserve(ntri * 400 + 512);
  for (size_t t = 0; t < ntri; ++t) {
    s += t ? ",[" : "[";
    for (int k = 0; k < 3; ++k) {
      const uint32_t v = indices[t * 3 + (size_t)k];
      if (v >= nv) fail("glTF: vertex index out of range");
      s += k ? ",{" : "{";
      s += "\"position\":" + vec_s(&pos[(size_t)v * 3], 3) + ",\"normal\":" + vec_s(&nrm[(size_t)v * 3], 3) +
           ",\"tex_coords\":" + vec_s(&uv[(size_t)v * 2], 2) + "}";
    }
    s += "]";
  }
  s += "],\"material\":" + material_json(d, idx(prim.get("material")), bank) + "}";
  out.models += (out.n_models++ ? "," : "") + s;
}

void walk(const Doc &d, int64_t ni, const Mat4 &parent, TextureBank &bank, Output &out, int depth) {
  if (depth > 256) fail("glTF: node hierarchy too deep (cycle?)");
  const Json &n = d.arr("nodes", ni);
  const Mat4 M = mul(parent, node_local(n));
  if (idx(n.get("camera")) >= 0 && !out.have_camera) {  // scenes[0].cameras[0], gltf.rs:169
    const Json &cam = d.arr("cameras", idx(n.get("camera")));
    const Json *persp = cam.get("perspective");
    if (!persp) fail("Orthographic camera not supported");  // gltf.rs:211-213
    float cols[16];
    for (int c = 0; c < 4; ++c)
      for (int r = 0; r < 4; ++r) cols[c * 4 + r] = M.m[c][r];
    std::string t = "[";
    for (int c = 0; c < 4; ++c) t += (c ? "," : "") + vec_s(&cols[c * 4], 4);
    t += "]";
    // an infinite far plane has no JSON number: the largest finite f32 stands in for it
    const float zfar = persp->get("zfar") ? (float)num(persp->get("zfar"), 0) : 3.4028235e38f;
    out.camera = "{\"transform\":" + t + ",\"fov\":" + f32s((float)num(persp->get("yfov"), 0.0)) + ",\"zfar\":" + f32s(zfar) +
                 ",\"znear\":" + f32s((float)num(persp->get("znear"), 0.0)) + "}";
    out.have_camera = true;
  }
  if (idx(n.get("mesh")) >= 0) {
    const Json &mesh = d.arr("meshes", idx(n.get("mesh")));
    const Json *prims = mesh.get("primitives");
    if (prims && prims->kind == Json::Array)
      for (const Json &p : prims->arr) emit_primitive(d, p, M, bank, out);
  }
  if (const Json *ext = n.get("extensions"))
    if (const Json *kl = ext->get("KHR_lights_punctual")) {
      const int64_t li = idx(kl->get("light"));
      const Json *rext = d.root.get("extensions");
      const Json *rk = rext ? rext->get("KHR_lights_punctual") : nullptr;
      const Json *ls = rk ? rk->get("lights") : nullptr;
      if (!ls || ls->kind != Json::Array || li < 0 || (size_t)li >= ls->arr.size()) fail("glTF: bad KHR_lights_punctual light index");
      const Json &L = ls->arr[(size_t)li];
      float color[3] = {1, 1, 1};
      if (const Json *c = L.get("color"))
        if (c->kind == Json::Array && c->arr.size() == 3)
          for (int i = 0; i < 3; ++i) color[i] = (float)c->arr[(size_t)i].num;
      const float intensity = (float)num(L.get("intensity"), 1.0);
      for (float &c : color) c *= intensity;  // (color * intensity), gltf.rs:244,253,262
      const Json *ty = L.get("type");
      std::string s;
      if (ty && ty->kind == Json::String && ty->str == "directional") {
        const float fwd[3] = {0, 0, -1};  // lights shine down their local -Z
        float dir[3];
        xform(M, fwd, 0.f, dir);
        const float len = std::sqrt(dir[0] * dir[0] + dir[1] * dir[1] + dir[2] * dir[2]);
        if (len > 0.f)
          for (float &x : dir) x /= len;
        s = "{\"type\":\"Directional\",\"direction\":" + vec_s(dir, 3) + ",\"color\":" + vec_s(color, 3) + "}";
      } else {  // point, and spot -> point (gltf.rs:246-264)
        const float pos[3] = {M.m[3][0], M.m[3][1], M.m[3][2]};
        s = "{\"type\":\"Point\",\"position\":" + vec_s(pos, 3) + ",\"color\":" + vec_s(color, 3) + ",\"size\":0.1}";
      }
      out.lights += (out.n_lights++ ? "," : "") + s;
    }
  if (const Json *ch = n.get("children"))
    if (ch->kind == Json::Array)
      for (const Json &c : ch->arr) walk(d, idx(&c), M, bank, out, depth + 1);
}

}  // namespace

int convert_gltf_to_isf(const char *input, const char *output_dir) {
  try {
    // gltf.rs:150-155
    struct stat sb;
    if (stat(output_dir, &sb) != 0) {
      std::string acc;
      for (const char *p = output_dir;; ++p) {  // create_dir_all
        if (*p == '/' || *p == 0) {
          if (!acc.empty() && mkdir(acc.c_str(), 0777) != 0 && errno != EEXIST) fail(acc + ": " + std::strerror(errno));
          if (*p == 0) break;
        }
        acc.push_back(*p);
      }
    } else if (!S_ISDIR(sb.st_mode)) {
      fail(std::string("'") + output_dir + "' is not a directory");
    }
    Doc d;
    std::string in = input;
    size_t slash = in.find_last_of('/');
    d.dir = slash == std::string::npos ? "." : in.substr(0, slash);
    std::vector<uint8_t> file = read_file(in);
    std::string json_text;
    if (file.size() >= 12 && !std::memcmp(file.data(), "glTF", 4)) {  // GLB container
      size_t p = 12;
      while (p + 8 <= file.size()) {
        uint32_t len, type;
        std::memcpy(&len, &file[p], 4);
        std::memcpy(&type, &file[p + 4], 4);
        if (p + 8 + (size_t)len > file.size()) fail(in + ": truncated GLB chunk");
        if (type == 0x4E4F534Au) json_text.assign((const char *)&file[p + 8], len);
        else if (type == 0x004E4942u && !d.has_glb_bin) {
          d.glb_bin.assign(&file[p + 8], &file[p + 8] + len);
          d.has_glb_bin = true;
        }
        p += 8 + (size_t)len;
      }
      if (json_text.empty()) fail(in + ": GLB without a JSON chunk");
    } else {
      json_text.assign((const char *)file.data(), file.size());
    }
    std::string err;
    if (!json_parse(json_text.data(), json_text.size(), d.root, err)) fail(in + ": " + err);
    d.load_buffers();
    if (d.count("scenes") == 0) fail("No scenes found in gltf file");  // gltf.rs:160-162
    TextureBank bank;
    bank.out_dir = output_dir;
    bank.doc = &d;
    Output out;
    const Json &scene0 = d.arr("scenes", 0);
    if (const Json *nodes = scene0.get("nodes"))
      if (nodes->kind == Json::Array)
        for (const Json &n : nodes->arr) walk(d, idx(&n), identity(), bank, out, 0);
    if (!out.have_camera) fail("No camera found");  // the reference panics here (gltf.rs:165-168)
    std::string isf = "{\"models\":[" + out.models + "],\"camera\":" + out.camera + ",\"lights\":[" + out.lights +
                      "],\"background\":[0.0,0.0,0.0]}";
    const std::string path = std::string(output_dir) + "/scene.isf";
    FILE *f = std::fopen(path.c_str(), "wb");
    if (!f) fail(path + ": cannot open for writing");
    const bool ok = std::fwrite(isf.data(), 1, isf.size(), f) == isf.size();
    if (std::fclose(f) != 0 || !ok) fail(path + ": write failed");
    return PT_OK;
  } catch (const Fail &e) {
    set_error("%s", e.msg.c_str());
    return PT_ERR_PARSE;
  } catch (const std::exception &e) {
    set_error("convert: %s", e.what());
    return PT_ERR_INVALID;
  }
}

}  // namespace ptb

extern "C" int pt_convert_gltf(const char *input, const char *output_dir) {
  if (!input || !output_dir) {
    ptb::set_error("null argument");
    return PT_ERR_INVALID;
  }
  return ptb::abi_guard([&]() -> int { return ptb::convert_gltf_to_isf(input, output_dir); });
}
