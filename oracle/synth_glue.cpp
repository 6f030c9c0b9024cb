// oracle/synth_glue.cpp — TEST INFRASTRUCTURE: the product's procedural scene generators (path-tracer_b200/csrc/
// synth.cpp, host code, no CUDA) compiled into the oracle, so that the CPU arm of bench.py builds the very same
// synthetic scenes without loading libpt_b200.so.  The generator is input data, not the path under
// test: the oracle's ray_cast / render_pixel never call into the product.
#include <cstdarg>
#include <cstdio>
#include <string>

#include "../path-tracer_b200/csrc/host_api.hpp"

namespace ptb {
static thread_local std::string g_oracle_synth_error;
void set_error(const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_oracle_synth_error = buf;
}
}  // namespace ptb

extern "C" {
void *oracle_scene_create(const pt_scene_desc *d);

// recipe in {console, foliage, city, soup}; returns an oracle scene handle (oracle_scene_destroy) or null.
void *oracle_scene_synthetic(const char *recipe, uint64_t n_tris, uint64_t seed, uint64_t *n_tris_out) {
  ptb::HostScene h;
  if (ptb::make_synthetic(recipe, n_tris, seed, h) != PT_OK) return nullptr;
  if (!h.finalize().empty()) return nullptr;
  pt_scene_desc d;
  h.to_desc(d);
  if (n_tris_out) *n_tris_out = d.n_tris;
  return oracle_scene_create(&d);  // copies everything it needs
}
}
