// json.cpp — recursive-descent JSON parser (see json.hpp).
#include "json.hpp"

#include <cstdlib>
#include <cstring>

namespace ptb {

const Json *Json::get(const char *key) const {
  if (kind != Object) return nullptr;
  const Json *found = nullptr;
  for (const auto &kv : obj)
    if (kv.first == key) found = &kv.second;
  return found;
}

namespace {
struct P {
  const char *s, *e, *p;
  std::string err;
  int depth = 0;
  bool fail(const char *m) {
    if (err.empty()) {
      size_t line = 1, col = 1;
      for (const char *q = s; q < p && q < e; ++q) {
        if (*q == '\n') { ++line; col = 1; } else ++col;
      }
      err = std::string(m) + " at line " + std::to_string(line) + " column " + std::to_string(col);
    }
    return false;
  }
  void ws() {
    while (p < e && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) ++p;
  }
  bool lit(const char *w) {
    size_t n = std::strlen(w);
    if ((size_t)(e - p) < n || std::memcmp(p, w, n) != 0) return fail("expected value");
    p += n;
    return true;
  }
  bool string(std::string &out) {
    if (p >= e || *p != '"') return fail("expected string");
    ++p;
    out.clear();
    while (p < e && *p != '"') {
      unsigned char c = (unsigned char)*p++;
      if (c < 0x20) return fail("control character in string");
      if (c != '\\') { out.push_back((char)c); continue; }
      if (p >= e) return fail("EOF in string");
      char x = *p++;
      switch (x) {
        case '"': out.push_back('"'); break;
        case '\\': out.push_back('\\'); break;
        case '/': out.push_back('/'); break;
        case 'b': out.push_back('\b'); break;
        case 'f': out.push_back('\f'); break;
        case 'n': out.push_back('\n'); break;
        case 'r': out.push_back('\r'); break;
        case 't': out.push_back('\t'); break;
        case 'u': {
          if (e - p < 4) return fail("bad \\u escape");
          unsigned cp = 0;
          for (int i = 0; i < 4; ++i) {
            char h = *p++;
            cp <<= 4;
            if (h >= '0' && h <= '9') cp |= h - '0';
            else if (h >= 'a' && h <= 'f') cp |= h - 'a' + 10;
            else if (h >= 'A' && h <= 'F') cp |= h - 'A' + 10;
            else return fail("bad \\u escape");
          }
          if (cp < 0x80) out.push_back((char)cp);
          else if (cp < 0x800) { out.push_back((char)(0xC0 | (cp >> 6))); out.push_back((char)(0x80 | (cp & 0x3F))); }
          else { out.push_back((char)(0xE0 | (cp >> 12))); out.push_back((char)(0x80 | ((cp >> 6) & 0x3F))); out.push_back((char)(0x80 | (cp & 0x3F))); }
          break;
        }
        default: return fail("bad escape");
      }
    }
    if (p >= e) return fail("EOF in string");
    ++p;
    return true;
  }
  bool value(Json &v) {
    if (++depth > 128) return fail("recursion limit exceeded");
    ws();
    if (p >= e) return fail("EOF while parsing a value");
    bool ok = true;
    char c = *p;
    if (c == '{') {
      v.kind = Json::Object;
      ++p;
      ws();
      if (p < e && *p == '}') { ++p; }
      else {
        for (;;) {
          ws();
          std::string k;
          if (!string(k)) { ok = false; break; }
          ws();
          if (p >= e || *p != ':') { ok = fail("expected `:`"); break; }
          ++p;
          v.obj.emplace_back(std::move(k), Json());
          if (!value(v.obj.back().second)) { ok = false; break; }
          ws();
          if (p < e && *p == ',') { ++p; continue; }
          if (p < e && *p == '}') { ++p; break; }
          ok = fail("expected `,` or `}`");
          break;
        }
      }
    } else if (c == '[') {
      v.kind = Json::Array;
      ++p;
      ws();
      if (p < e && *p == ']') { ++p; }
      else {
        for (;;) {
          v.arr.emplace_back();
          if (!value(v.arr.back())) { ok = false; break; }
          ws();
          if (p < e && *p == ',') { ++p; continue; }
          if (p < e && *p == ']') { ++p; break; }
          ok = fail("expected `,` or `]`");
          break;
        }
      }
    } else if (c == '"') {
      v.kind = Json::String;
      ok = string(v.str);
    } else if (c == 't') {
      v.kind = Json::Bool; v.b = true; ok = lit("true");
    } else if (c == 'f') {
      v.kind = Json::Bool; v.b = false; ok = lit("false");
    } else if (c == 'n') {
      v.kind = Json::Null; ok = lit("null");
    } else if (c == '-' || (c >= '0' && c <= '9')) {
      const char *q = p;
      if (*q == '-') ++q;
      if (q >= e || *q < '0' || *q > '9') ok = fail("invalid number");
      else {
        while (q < e && ((*q >= '0' && *q <= '9') || *q == '.' || *q == 'e' || *q == 'E' || *q == '+' || *q == '-')) ++q;
        std::string tmp(p, q);
        char *endp = nullptr;
        v.num = std::strtod(tmp.c_str(), &endp);
        if (endp != tmp.c_str() + tmp.size()) ok = fail("invalid number");
        v.kind = Json::Number;
        p = q;
      }
    } else {
      ok = fail("expected value");
    }
    --depth;
    return ok;
  }
};
}  // namespace

bool json_parse(const char *text, size_t len, Json &out, std::string &err) {
  P ps{text, text + len, text, {}, 0};
  if (!ps.value(out)) { err = ps.err; return false; }
  ps.ws();
  if (ps.p != ps.e) { ps.fail("trailing characters"); err = ps.err; return false; }
  return true;
}

}  // namespace ptb
