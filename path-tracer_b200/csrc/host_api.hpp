// host_api.hpp — internal glue shared by the host translation units of libpt_b200.so.
#pragma once
#include <exception>
#include <new>
#include <string>

#include "scene_host.hpp"

namespace ptb {
void set_error(const char *fmt, ...);
// Nothing may throw across the C ABI (include/pt_b200.h): entry points that parse untrusted files or allocate
// host memory in proportion to them run their body through this.
template <typename F>
int abi_guard(F &&body) noexcept {
  try {
    return body();
  } catch (const std::bad_alloc &) {
    set_error("out of host memory");
    return PT_ERR_NOMEM;
  } catch (const std::exception &e) {
    set_error("internal error: %s", e.what());
    return PT_ERR_INVALID;
  } catch (...) {
    set_error("internal error");
    return PT_ERR_INVALID;
  }
}
// Finalises `host`, builds the BVH, uploads to `device`.  Takes ownership of host.  With `accel` (read from an isfb
// file, kind = PT_ACCEL_*) the tree is validated against the scene and used instead of building one.
int scene_from_host(HostScene &&host, int device, pt_scene **out, Bvh4 *accel = nullptr, int accel_kind = -1);
// isf_loader.cpp: scene/mod.rs:16-22 + isf.rs serde schema.  Returns PT_OK or an error (message set).
int load_isf_file(const char *path, HostScene &out);
// png.cpp
int png_read(const char *path, uint32_t channels_out, uint32_t *w, uint32_t *h, std::vector<uint8_t> &data);
int png_decode(const uint8_t *bytes, size_t size, const char *name, uint32_t channels_out, uint32_t *w, uint32_t *h,
               std::vector<uint8_t> &data);
// jpeg.cpp: baseline JPEG; image_decode picks PNG or JPEG from the signature (like image::open)
int jpeg_decode(const uint8_t *bytes, size_t size, const char *name, uint32_t channels_out, uint32_t *w, uint32_t *h,
                std::vector<uint8_t> &data);
int image_decode(const uint8_t *bytes, size_t size, const char *name, uint32_t channels_out, uint32_t *w, uint32_t *h,
                 std::vector<uint8_t> &data);
int png_write(const char *path, const uint8_t *rgb8, uint32_t w, uint32_t h);
int png_write_ch(const char *path, const uint8_t *pixels, uint32_t w, uint32_t h, uint32_t channels);
// gltf_convert.cpp: scene::convert_gltf_to_isf (src/scene/gltf.rs:146-204)
int convert_gltf_to_isf(const char *input, const char *output_dir);
// synth.cpp
int make_synthetic(const char *recipe, uint64_t n_tris, uint64_t seed, HostScene &out);
}  // namespace ptb
