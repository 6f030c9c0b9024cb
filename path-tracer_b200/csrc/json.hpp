// json.hpp — minimal JSON DOM (RFC 8259) for the isf scene format; numbers are parsed with strtod to a
// double and narrowed to f32 by the caller, the same decimal -> f64 -> f32 path serde_json takes.
#pragma once
#include <map>
#include <memory>
#include <string>
#include <vector>

namespace ptb {
struct Json {
  enum Kind { Null, Bool, Number, String, Array, Object } kind = Null;
  bool b = false;
  double num = 0;
  std::string str;
  std::vector<Json> arr;
  std::vector<std::pair<std::string, Json>> obj;  // insertion order kept; duplicate keys: last wins on lookup

  const Json *get(const char *key) const;
  bool is_null() const { return kind == Null; }
};
// Parses `text`; on failure returns false and fills err with "message at line L column C".
bool json_parse(const char *text, size_t len, Json &out, std::string &err);
}  // namespace ptb
