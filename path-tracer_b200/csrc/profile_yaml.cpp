// profile_yaml.cpp — render profile reader.  Mirrors Profile::load + serde defaults of the reference
// (config/profile.rs:10-40, config/resolution.rs:4-16, brdf/mod.rs:50-55, tonemap.rs:4-13): a YAML mapping
// with optional keys resolution{width,height}, bounces (4), samples (64), brdf (COOK_TORRANCE),
// tonemap (FILMIC); unknown keys are ignored, unknown enum strings and malformed values are errors.
#include <cctype>
#include <cstdlib>
#include <fstream>
#include <sstream>
#include <vector>

#include "host_api.hpp"

namespace {

std::string trim(const std::string &s) {
  size_t a = 0, b = s.size();
  while (a < b && std::isspace((unsigned char)s[a])) ++a;
  while (b > a && std::isspace((unsigned char)s[b - 1])) --b;
  return s.substr(a, b - a);
}
std::string strip_comment(const std::string &s) {
  bool sq = false, dq = false;
  for (size_t i = 0; i < s.size(); ++i) {
    if (s[i] == '\'' && !dq) sq = !sq;
    else if (s[i] == '"' && !sq) dq = !dq;
    else if (s[i] == '#' && !sq && !dq && (i == 0 || std::isspace((unsigned char)s[i - 1]))) return s.substr(0, i);
  }
  return s;
}
std::string unquote(const std::string &s) {
  if (s.size() >= 2 && ((s.front() == '"' && s.back() == '"') || (s.front() == '\'' && s.back() == '\''))) return s.substr(1, s.size() - 2);
  return s;
}
bool parse_uint(const std::string &v, uint64_t &out) {
  std::string t = unquote(trim(v));
  if (t.empty()) return false;
  size_t i = t[0] == '+' ? 1 : 0;
  if (i >= t.size()) return false;
  uint64_t r = 0;
  for (; i < t.size(); ++i) {
    if (t[i] == '_') continue;
    if (!std::isdigit((unsigned char)t[i])) return false;
    r = r * 10 + (uint64_t)(t[i] - '0');
    if (r > 0xFFFFFFFFull) return false;
  }
  out = r;
  return true;
}
struct KV {
  int indent;
  std::string key, value;
};
bool split_kv(const std::string &line, std::string &k, std::string &v) {
  size_t c = line.find(':');
  if (c == std::string::npos) return false;
  if (c + 1 < line.size() && !std::isspace((unsigned char)line[c + 1])) return false;
  k = unquote(trim(line.substr(0, c)));
  v = trim(line.substr(c + 1));
  return !k.empty();
}
// {a: 1, b: 2}
bool parse_flow_map(const std::string &v, std::vector<KV> &out) {
  std::string t = trim(v);
  if (t.size() < 2 || t.front() != '{' || t.back() != '}') return false;
  std::string body = t.substr(1, t.size() - 2);
  std::stringstream ss(body);
  std::string item;
  while (std::getline(ss, item, ',')) {
    if (trim(item).empty()) continue;
    std::string k, val;
    if (!split_kv(trim(item) + " ", k, val)) return false;
    out.push_back({0, k, val});
  }
  return true;
}

}  // namespace

extern "C" {

void pt_profile_default(pt_profile *out) {
  if (!out) return;
  out->width = 1920;   // resolution.rs:9-16
  out->height = 1080;
  out->bounces = 4;    // profile.rs:34-36
  out->samples = 64;   // profile.rs:38-40
  out->brdf = PT_BRDF_COOK_TORRANCE;
  out->tonemap = PT_TONEMAP_FILMIC;
}

int pt_profile_load_yaml(const char *path, pt_profile *out) {
  if (!path || !out) {
    ptb::set_error("null argument");
    return PT_ERR_INVALID;
  }
  std::ifstream f(path);
  if (!f) {
    ptb::set_error("%s: No such file or directory (os error 2)", path);
    return PT_ERR_IO;
  }
  pt_profile p;
  pt_profile_default(&p);
  std::vector<KV> top;
  std::vector<std::vector<KV>> children;
  std::string line;
  int lineno = 0;
  bool any = false;
  while (std::getline(f, line)) {
    ++lineno;
    std::string s = strip_comment(line);
    if (trim(s).empty() || trim(s) == "---" || trim(s) == "...") continue;
    if (s.find('\t') != std::string::npos && trim(s.substr(0, s.find_first_not_of(" \t"))).empty() && s[0] == '\t') {
      ptb::set_error("found character that cannot start any token at line %d (tab indentation)", lineno);
      return PT_ERR_PARSE;
    }
    int indent = (int)s.find_first_not_of(' ');
    std::string k, v;
    if (!split_kv(trim(s), k, v)) {
      ptb::set_error("invalid type: expected struct Profile (could not parse line %d)", lineno);
      return PT_ERR_PARSE;
    }
    any = true;
    if (indent == 0) {
      top.push_back({0, k, v});
      children.emplace_back();
      if (!v.empty() && v[0] == '{' && !parse_flow_map(v, children.back())) {
        ptb::set_error("invalid flow mapping at line %d", lineno);
        return PT_ERR_PARSE;
      }
    } else {
      if (top.empty() || !top.back().value.empty()) {
        ptb::set_error("mapping values are not allowed in this context at line %d", lineno);
        return PT_ERR_PARSE;
      }
      children.back().push_back({indent, k, v});
    }
  }
  if (!any) {
    ptb::set_error("invalid type: unit value, expected struct Profile");
    return PT_ERR_PARSE;
  }
  for (size_t i = 0; i < top.size(); ++i) {
    const std::string &k = top[i].key, &v = top[i].value;
    uint64_t u = 0;
    if (k == "resolution") {
      bool hw = false, hh = false;
      if (children[i].empty() && !(trim(v).empty())) {
        ptb::set_error("resolution: invalid type, expected struct Resolution");
        return PT_ERR_PARSE;
      }
      for (const KV &c : children[i]) {
        if (c.key == "width" || c.key == "height") {
          if (!parse_uint(c.value, u)) {
            ptb::set_error("resolution.%s: invalid value `%s`, expected u32", c.key.c_str(), c.value.c_str());
            return PT_ERR_PARSE;
          }
          if (c.key == "width") { p.width = (uint32_t)u; hw = true; }
          else { p.height = (uint32_t)u; hh = true; }
        }
      }
      if (!hw || !hh) {
        ptb::set_error("resolution: missing field `%s`", hw ? "height" : "width");
        return PT_ERR_PARSE;
      }
    } else if (k == "bounces" || k == "samples") {
      if (!parse_uint(v, u)) {
        ptb::set_error("%s: invalid value `%s`, expected usize", k.c_str(), v.c_str());
        return PT_ERR_PARSE;
      }
      (k == "bounces" ? p.bounces : p.samples) = (uint32_t)u;
    } else if (k == "brdf") {
      if (unquote(v) != "COOK_TORRANCE") {
        ptb::set_error("brdf: unknown variant `%s`, expected `COOK_TORRANCE`", unquote(v).c_str());
        return PT_ERR_PARSE;
      }
      p.brdf = PT_BRDF_COOK_TORRANCE;
    } else if (k == "tonemap") {
      std::string t = unquote(v);
      if (t == "REINHARD") p.tonemap = PT_TONEMAP_REINHARD;
      else if (t == "FILMIC") p.tonemap = PT_TONEMAP_FILMIC;
      else if (t == "ACES") p.tonemap = PT_TONEMAP_ACES;
      else {
        ptb::set_error("tonemap: unknown variant `%s`, expected one of `REINHARD`, `FILMIC`, `ACES`", t.c_str());
        return PT_ERR_PARSE;
      }
    }
  }
  *out = p;
  return PT_OK;
}
}
