// png.cpp — PNG decode/encode on zlib.  Stands in for the `image` crate at the two places the render path
// touches it: texture loading (reference scene/internal/texture_bank.rs:33 into_rgb8, :49 into_luma8) and
// RgbImage::save (reference src/main.rs:50).  Non-interlaced images, bit depths 1-8 (16-bit: high byte).
#include <zlib.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "host_api.hpp"

namespace ptb {
namespace {

uint32_t be32(const uint8_t *p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

int paeth(int a, int b, int c) {
  int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
  if (pa <= pb && pa <= pc) return a;
  return pb <= pc ? b : c;
}

}  // namespace

int png_read(const char *path, uint32_t channels_out, uint32_t *w_out, uint32_t *h_out, std::vector<uint8_t> &out) {
  FILE *f = std::fopen(path, "rb");
  if (!f) {
    set_error("%s: cannot open", path);
    return PT_ERR_IO;
  }
  std::vector<uint8_t> file;
  uint8_t buf[65536];
  size_t n;
  while ((n = std::fread(buf, 1, sizeof buf, f)) > 0) file.insert(file.end(), buf, buf + n);
  std::fclose(f);
  return image_decode(file.data(), file.size(), path, channels_out, w_out, h_out, out);  // PNG, or JPEG by signature
}

// channels_out: 1 (into_luma8), 3 (into_rgb8) or 4 (into_rgba8; alpha 255 when the file has none).
int png_decode(const uint8_t *bytes, size_t size, const char *path, uint32_t channels_out, uint32_t *w_out,
               uint32_t *h_out, std::vector<uint8_t> &out) {
  struct View {  // keeps the body below identical for file and in-memory sources
    const uint8_t *p;
    size_t n;
    size_t size() const { return n; }
    const uint8_t *data() const { return p; }
    const uint8_t &operator[](size_t i) const { return p[i]; }
  } file{bytes, size};
  static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
  if (file.size() < 8 || std::memcmp(file.data(), sig, 8) != 0) {
    set_error("%s: not a PNG file", path);
    return PT_ERR_PARSE;
  }
  uint32_t w = 0, h = 0;
  int depth = 0, ctype = -1, interlace = 0;
  std::vector<uint8_t> idat, plte;
  size_t p = 8;
  bool end = false;
  while (!end && p + 12 <= file.size()) {
    uint32_t len = be32(&file[p]);
    const uint8_t *type = &file[p + 4];
    if (p + 12 + (size_t)len > file.size()) break;
    const uint8_t *data = &file[p + 8];
    if (!std::memcmp(type, "IHDR", 4) && len >= 13) {
      w = be32(data);
      h = be32(data + 4);
      depth = data[8];
      ctype = data[9];
      interlace = data[12];
    } else if (!std::memcmp(type, "PLTE", 4)) {
      plte.assign(data, data + len);
    } else if (!std::memcmp(type, "IDAT", 4)) {
      idat.insert(idat.end(), data, data + len);
    } else if (!std::memcmp(type, "IEND", 4)) {
      end = true;
    }
    p += 12 + (size_t)len;
  }
  int nch = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : ctype == 6 ? 4 : 0;
  if (w == 0 || h == 0 || nch == 0 || interlace != 0 || !(depth == 1 || depth == 2 || depth == 4 || depth == 8 || depth == 16) ||
      (depth < 8 && ctype != 0 && ctype != 3) || (ctype == 3 && (depth == 16 || plte.empty()))) {
    set_error("%s: unsupported PNG (color type %d, depth %d, interlace %d)", path, ctype, depth, interlace);
    return PT_ERR_PARSE;
  }
  // untrusted header: cap the sides so that (stride + 1) * h cannot overflow or ask for an absurd allocation
  if (w > 65536u || h > 65536u || (uint64_t)w * h * (uint64_t)nch * (depth > 8 ? 2 : 1) > (1ull << 31)) {
    set_error("%s: PNG too large (%u x %u)", path, w, h);
    return PT_ERR_PARSE;
  }
  const size_t bpp_bits = (size_t)nch * depth;
  const size_t stride = ((size_t)w * bpp_bits + 7) / 8;
  const size_t bpp = std::max<size_t>(1, bpp_bits / 8);
  std::vector<uint8_t> raw((stride + 1) * h);
  uLongf rawlen = (uLongf)raw.size();
  int zr = uncompress(raw.data(), &rawlen, idat.data(), (uLong)idat.size());
  if (zr != Z_OK || rawlen != raw.size()) {
    set_error("%s: corrupt PNG data (zlib %d)", path, zr);
    return PT_ERR_PARSE;
  }
  // unfilter in place
  std::vector<uint8_t> img(stride * h);
  for (uint32_t y = 0; y < h; ++y) {
    const uint8_t *src = &raw[(stride + 1) * y];
    uint8_t ft = src[0];
    ++src;
    uint8_t *dst = &img[stride * y];
    const uint8_t *up = y ? &img[stride * (y - 1)] : nullptr;
    for (size_t x = 0; x < stride; ++x) {
      int a = x >= bpp ? dst[x - bpp] : 0, b = up ? up[x] : 0, c = (up && x >= bpp) ? up[x - bpp] : 0;
      int v = src[x];
      switch (ft) {
        case 0: break;
        case 1: v += a; break;
        case 2: v += b; break;
        case 3: v += (a + b) / 2; break;
        case 4: v += paeth(a, b, c); break;
        default:
          set_error("%s: bad PNG filter %d", path, ft);
          return PT_ERR_PARSE;
      }
      dst[x] = (uint8_t)v;
    }
  }
  // expand to 8-bit RGB triplets / gray, then convert
  out.assign((size_t)w * h * channels_out, 0);
  for (uint32_t y = 0; y < h; ++y) {
    const uint8_t *row = &img[stride * y];
    for (uint32_t x = 0; x < w; ++x) {
      uint8_t s[4] = {0, 0, 0, 255};
      for (int c = 0; c < nch; ++c) {
        if (depth == 8) s[c] = row[(size_t)x * nch + c];
        else if (depth == 16) s[c] = row[((size_t)x * nch + c) * 2];
        else {
          size_t bit = (size_t)x * depth;
          int v = (row[bit / 8] >> (8 - depth - (bit % 8))) & ((1 << depth) - 1);
          s[c] = ctype == 3 ? (uint8_t)v : (uint8_t)(v * 255 / ((1 << depth) - 1));
        }
      }
      uint8_t r, g, b;
      bool gray_src = false;
      if (ctype == 3) {
        size_t i = (size_t)s[0] * 3;
        if (i + 2 < plte.size()) { r = plte[i]; g = plte[i + 1]; b = plte[i + 2]; }
        else { r = g = b = 0; }
      } else if (ctype == 0 || ctype == 4) {
        r = g = b = s[0];
        gray_src = true;
      } else {
        r = s[0]; g = s[1]; b = s[2];
      }
      uint8_t *o = &out[((size_t)y * w + x) * channels_out];
      if (channels_out == 3) {
        o[0] = r; o[1] = g; o[2] = b;
      } else if (channels_out == 4) {
        o[0] = r; o[1] = g; o[2] = b;
        o[3] = ctype == 6 ? s[3] : ctype == 4 ? s[1] : 255;
      } else {
        // image crate rgb_to_luma for u8: (2126 r + 7152 g + 722 b) / 10000; gray sources pass through
        o[0] = gray_src ? r : (uint8_t)((2126u * r + 7152u * g + 722u * b) / 10000u);
      }
    }
  }
  *w_out = w;
  *h_out = h;
  return PT_OK;
}

int png_write(const char *path, const uint8_t *rgb8, uint32_t w, uint32_t h) { return png_write_ch(path, rgb8, w, h, 3); }

// 8-bit RGB (channels 3) or grayscale (channels 1) PNG.
int png_write_ch(const char *path, const uint8_t *rgb8, uint32_t w, uint32_t h, uint32_t channels) {
  const size_t row = (size_t)w * channels;
  std::vector<uint8_t> raw((row + 1) * h);
  for (uint32_t y = 0; y < h; ++y) {
    raw[(row + 1) * y] = 0;
    std::memcpy(&raw[(row + 1) * y + 1], rgb8 + (size_t)y * row, row);
  }
  uLongf clen = compressBound((uLong)raw.size());
  std::vector<uint8_t> comp(clen);
  if (compress2(comp.data(), &clen, raw.data(), (uLong)raw.size(), 6) != Z_OK) {
    set_error("zlib compress failed");
    return PT_ERR_IO;
  }
  FILE *f = std::fopen(path, "wb");
  if (!f) {
    set_error("%s: cannot open for writing", path);
    return PT_ERR_IO;
  }
  auto chunk = [&](const char *type, const uint8_t *data, uint32_t len) {
    uint8_t hdr[8] = {(uint8_t)(len >> 24), (uint8_t)(len >> 16), (uint8_t)(len >> 8), (uint8_t)len,
                      (uint8_t)type[0], (uint8_t)type[1], (uint8_t)type[2], (uint8_t)type[3]};
    std::fwrite(hdr, 1, 8, f);
    if (len) std::fwrite(data, 1, len, f);
    uLong crc = crc32(0L, hdr + 4, 4);
    if (len) crc = crc32(crc, data, len);
    uint8_t c[4] = {(uint8_t)(crc >> 24), (uint8_t)(crc >> 16), (uint8_t)(crc >> 8), (uint8_t)crc};
    std::fwrite(c, 1, 4, f);
  };
  static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
  std::fwrite(sig, 1, 8, f);
  uint8_t ihdr[13] = {(uint8_t)(w >> 24), (uint8_t)(w >> 16), (uint8_t)(w >> 8), (uint8_t)w,
                      (uint8_t)(h >> 24), (uint8_t)(h >> 16), (uint8_t)(h >> 8), (uint8_t)h, 8,
                      (uint8_t)(channels == 1 ? 0 : 2), 0, 0, 0};
  chunk("IHDR", ihdr, 13);
  chunk("IDAT", comp.data(), (uint32_t)clen);
  chunk("IEND", nullptr, 0);
  bool ok = std::fclose(f) == 0;
  if (!ok) {
    set_error("%s: write failed", path);
    return PT_ERR_IO;
  }
  return PT_OK;
}

}  // namespace ptb

extern "C" {
int pt_write_png(const char *path, const uint8_t *rgb8, uint32_t width, uint32_t height) {
  if (!path || !rgb8 || !width || !height) {
    ptb::set_error("bad argument");
    return PT_ERR_INVALID;
  }
  return ptb::png_write(path, rgb8, width, height);
}
int pt_read_png(const char *path, uint32_t channels_out, uint32_t *width, uint32_t *height, uint8_t **data_out) {
  if (!path || !width || !height || !data_out || (channels_out != 1 && channels_out != 3)) {
    ptb::set_error("bad argument");
    return PT_ERR_INVALID;
  }
  return ptb::abi_guard([&]() -> int {
    std::vector<uint8_t> d;
    int rc = ptb::png_read(path, channels_out, width, height, d);
    if (rc != PT_OK) return rc;
    *data_out = (uint8_t *)std::malloc(d.size() ? d.size() : 1);
    if (!*data_out) return PT_ERR_NOMEM;
    std::memcpy(*data_out, d.data(), d.size());
    return PT_OK;
  });
}
void pt_free(void *p) { std::free(p); }
}
