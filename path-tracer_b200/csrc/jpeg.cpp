// jpeg.cpp — baseline / extended-sequential JPEG decoder (ITU-T T.81, 8-bit, Huffman) for textures.
// Stands in for the `image` crate's JPEG path at the two places the reference reaches it: texture loading
// (scene/internal/texture_bank.rs:33,49 `image::open(..).into_rgb8()/into_luma8()`) and the glTF converter
// (scene/gltf.rs through easy-gltf).  Host code, off the accelerated path.  Supported: 1 (gray) or 3 (YCbCr, or RGB
// when an Adobe APP14 segment says so) components, any sampling factors up to 4x4, restart intervals, 8/16-bit
// quantisation tables.  Not supported (an error): progressive (SOF2), lossless, arithmetic coding, 12-bit, CMYK.
// Sub-sampled chroma is interpolated linearly between sample centres and the inverse DCT is evaluated in double
// precision, so decoded pixels can differ by a level or two from decoders that use libjpeg's integer IDCT and integer
// "fancy" upsampling (parity unpinned: no reference test reads a JPEG).
#include <cmath>
#include <cstring>
#include <string>
#include <vector>

#include "host_api.hpp"

namespace ptb {
namespace {

struct Huff {
  // canonical code tables: for each code length, the first code, the first symbol index and the number of codes
  int mincode[17], maxcode[18], valptr[17];
  uint8_t vals[256];
  bool present = false;
};

struct Component {
  int id = 0, h = 1, v = 1, tq = 0, td = 0, ta = 0;
  int pred = 0;
  int bw = 0, bh = 0;  // size in blocks (padded to whole MCUs)
  std::vector<uint8_t> plane;  // bw*8 x bh*8 samples
};

struct Reader {
  const uint8_t *p;
  size_t n, pos = 0;
  uint32_t acc = 0;
  int bits = 0;
  bool hit_marker = false;
  int get_bit() {
    if (bits == 0) {
      uint8_t b = 0;
      if (!hit_marker && pos < n) {
        b = p[pos++];
        if (b == 0xFF) {
          uint8_t b2 = pos < n ? p[pos] : 0;
          if (b2 == 0) ++pos;  // stuffed zero
          else {               // a marker ends the entropy-coded segment: feed zeros from here on
            hit_marker = true;
            --pos;
            b = 0;
          }
        }
      }
      acc = b;
      bits = 8;
    }
    --bits;
    return (int)((acc >> bits) & 1u);
  }
  int get_bits(int k) {
    int v = 0;
    for (int i = 0; i < k; ++i) v = (v << 1) | get_bit();
    return v;
  }
  void reset_bits() {
    bits = 0;
    hit_marker = false;
  }
};

int decode_symbol(Reader &r, const Huff &h) {
  int code = 0;
  for (int len = 1; len <= 16; ++len) {
    code = (code << 1) | r.get_bit();
    if (h.maxcode[len] >= 0 && code <= h.maxcode[len] && code >= h.mincode[len]) return h.vals[h.valptr[len] + code - h.mincode[len]];
  }
  return -1;
}
int extend(int v, int t) { return t == 0 ? 0 : (v < (1 << (t - 1)) ? v - (1 << t) + 1 : v); }

const int ZIGZAG[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                        41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                        30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

void idct8x8(const int *coef, const uint16_t *q, uint8_t *out, int stride) {
  static double C[8][8];
  static bool init = false;
  if (!init) {
    for (int x = 0; x < 8; ++x)
      for (int u = 0; u < 8; ++u) C[x][u] = (u == 0 ? std::sqrt(0.125) : 0.5) * std::cos((2 * x + 1) * u * M_PI / 16.0);
    init = true;
  }
  double f[64], tmp[64];
  for (int i = 0; i < 64; ++i) f[i] = (double)coef[i] * (double)q[i];
  for (int y = 0; y < 8; ++y)      // rows: tmp[y][x] = sum_u C[x][u] f[y][u]
    for (int x = 0; x < 8; ++x) {
      double s = 0;
      for (int u = 0; u < 8; ++u) s += C[x][u] * f[y * 8 + u];
      tmp[y * 8 + x] = s;
    }
  for (int x = 0; x < 8; ++x)      // columns
    for (int y = 0; y < 8; ++y) {
      double s = 0;
      for (int v = 0; v < 8; ++v) s += C[y][v] * tmp[v * 8 + x];
      int p = (int)std::lround(s + 128.0);
      out[y * stride + x] = (uint8_t)(p < 0 ? 0 : p > 255 ? 255 : p);
    }
}

uint8_t clamp8(double v) {
  int p = (int)std::lround(v);
  return (uint8_t)(p < 0 ? 0 : p > 255 ? 255 : p);
}

}  // namespace

int jpeg_decode(const uint8_t *bytes, size_t size, const char *name, uint32_t channels_out, uint32_t *w_out,
                uint32_t *h_out, std::vector<uint8_t> &out) {
  auto bad = [&](const char *why) {
    set_error("%s: %s", name, why);
    return PT_ERR_PARSE;
  };
  if (size < 4 || bytes[0] != 0xFF || bytes[1] != 0xD8) return bad("not a JPEG file");
  uint16_t qt[4][64] = {};
  bool have_qt[4] = {false, false, false, false};
  Huff dc[4], ac[4];
  std::vector<Component> comps;
  int width = 0, height = 0, restart = 0, adobe_transform = -1;
  bool have_sof = false, done = false;
  size_t pos = 2;
  while (!done && pos + 4 <= size) {
    if (bytes[pos] != 0xFF) return bad("corrupt JPEG (marker expected)");
    while (pos < size && bytes[pos] == 0xFF) ++pos;  // fill bytes
    if (pos >= size) break;
    const uint8_t m = bytes[pos++];
    if (m == 0xD9) break;
    if (m == 0x01 || (m >= 0xD0 && m <= 0xD7)) continue;
    if (pos + 2 > size) return bad("truncated JPEG");
    const size_t len = ((size_t)bytes[pos] << 8) | bytes[pos + 1];
    if (len < 2 || pos + len > size) return bad("truncated JPEG segment");
    const uint8_t *d = bytes + pos + 2;
    const size_t dl = len - 2;
    if (m == 0xDB) {  // DQT
      size_t i = 0;
      while (i < dl) {
        const int pq = d[i] >> 4, tq = d[i] & 15;
        ++i;
        if (tq > 3 || i + (pq ? 128 : 64) > dl) return bad("bad quantisation table");
        for (int k = 0; k < 64; ++k) {
          uint16_t v = pq ? (uint16_t)((d[i] << 8) | d[i + 1]) : d[i];
          i += pq ? 2 : 1;
          qt[tq][ZIGZAG[k]] = v;
        }
        have_qt[tq] = true;
      }
    } else if (m == 0xC0 || m == 0xC1) {  // SOF0 / SOF1
      if (dl < 6 || d[0] != 8) return bad("only 8-bit JPEG is supported");
      height = (d[1] << 8) | d[2];
      width = (d[3] << 8) | d[4];
      const int nf = d[5];
      if ((nf != 1 && nf != 3) || dl < 6 + 3 * (size_t)nf) return bad("unsupported JPEG component count (gray or 3 components only)");
      comps.resize((size_t)nf);
      for (int c = 0; c < nf; ++c) {
        comps[(size_t)c].id = d[6 + 3 * c];
        comps[(size_t)c].h = d[7 + 3 * c] >> 4;
        comps[(size_t)c].v = d[7 + 3 * c] & 15;
        comps[(size_t)c].tq = d[8 + 3 * c] & 3;
        if (comps[(size_t)c].h < 1 || comps[(size_t)c].h > 4 || comps[(size_t)c].v < 1 || comps[(size_t)c].v > 4) return bad("bad sampling factors");
      }
      have_sof = true;
    } else if (m == 0xC2 || (m >= 0xC3 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC)) {
      return bad("unsupported JPEG coding process (progressive / lossless / arithmetic); re-encode as baseline JPEG or PNG");
    } else if (m == 0xC4) {  // DHT
      size_t i = 0;
      while (i + 17 <= dl) {
        const int tc = d[i] >> 4, th = d[i] & 15;
        if (tc > 1 || th > 3) return bad("bad Huffman table id");
        Huff &h = tc ? ac[th] : dc[th];
        int counts[17] = {0}, total = 0;
        for (int l = 1; l <= 16; ++l) {
          counts[l] = d[i + (size_t)l];
          total += counts[l];
        }
        i += 17;
        if (total > 256 || i + (size_t)total > dl) return bad("bad Huffman table");
        std::memcpy(h.vals, d + i, (size_t)total);
        i += (size_t)total;
        int code = 0, k = 0;
        for (int l = 1; l <= 16; ++l) {
          h.valptr[l] = k;
          h.mincode[l] = code;
          code += counts[l];
          k += counts[l];
          h.maxcode[l] = counts[l] ? code - 1 : -1;
          code <<= 1;
        }
        h.present = true;
      }
    } else if (m == 0xDD) {  // DRI
      if (dl >= 2) restart = (d[0] << 8) | d[1];
    } else if (m == 0xEE) {  // Adobe APP14
      if (dl >= 12 && !std::memcmp(d, "Adobe", 5)) adobe_transform = d[11];
    } else if (m == 0xDA) {  // SOS: the entropy-coded data follows
      if (!have_sof) return bad("SOS before SOF");
      if (dl < 1) return bad("truncated SOS header");
      const int ns = d[0];
      if (ns != (int)comps.size() || dl < 1 + 2 * (size_t)ns + 3) return bad("non-interleaved scans are not supported");
      for (int k = 0; k < ns; ++k) {
        const int cid = d[1 + 2 * k];
        Component *cp = nullptr;
        for (Component &c : comps)
          if (c.id == cid) cp = &c;
        if (!cp) return bad("scan names an unknown component");
        cp->td = d[2 + 2 * k] >> 4;
        cp->ta = d[2 + 2 * k] & 15;
        if (cp->td > 3 || cp->ta > 3 || !dc[cp->td].present || !ac[cp->ta].present || !have_qt[cp->tq]) return bad("missing Huffman or quantisation table");
      }
      int hmax = 1, vmax = 1;
      for (const Component &c : comps) {
        hmax = std::max(hmax, c.h);
        vmax = std::max(vmax, c.v);
      }
      if (width <= 0 || height <= 0) return bad("bad JPEG dimensions");
      if (comps.size() == 1) comps[0].h = comps[0].v = hmax = vmax = 1;  // a single-component scan is never interleaved
      const int mcux = (width + 8 * hmax - 1) / (8 * hmax), mcuy = (height + 8 * vmax - 1) / (8 * vmax);
      for (Component &c : comps) {
        c.bw = mcux * c.h;
        c.bh = mcuy * c.v;
        c.plane.assign((size_t)c.bw * 8 * (size_t)c.bh * 8, 0);
        c.pred = 0;
      }
      Reader r{bytes, size};
      r.pos = pos + len;
      int until_restart = restart;
      for (int my = 0; my < mcuy; ++my)
        for (int mx = 0; mx < mcux; ++mx) {
          if (restart && until_restart == 0) {  // RSTn: byte-align, skip the marker, reset the predictors
            r.reset_bits();
            while (r.pos + 1 < size && !(bytes[r.pos] == 0xFF && bytes[r.pos + 1] >= 0xD0 && bytes[r.pos + 1] <= 0xD7)) ++r.pos;
            if (r.pos + 1 < size) r.pos += 2;
            for (Component &c : comps) c.pred = 0;
            until_restart = restart;
          }
          for (Component &c : comps)
            for (int by = 0; by < c.v; ++by)
              for (int bx = 0; bx < c.h; ++bx) {
                int coef[64] = {0};
                const int t = decode_symbol(r, dc[c.td]);
                if (t < 0 || t > 11) return bad("corrupt JPEG data (DC)");
                c.pred += extend(r.get_bits(t), t);
                coef[0] = c.pred;
                for (int k = 1; k < 64;) {
                  const int rs = decode_symbol(r, ac[c.ta]);
                  if (rs < 0) return bad("corrupt JPEG data (AC)");
                  const int run = rs >> 4, sz = rs & 15;
                  if (sz == 0) {
                    if (run == 15) { k += 16; continue; }
                    break;  // end of block
                  }
                  k += run;
                  if (k > 63) return bad("corrupt JPEG data (run past the block)");
                  coef[ZIGZAG[k]] = extend(r.get_bits(sz), sz);
                  ++k;
                }
                const int px = (mx * c.h + bx) * 8, py = (my * c.v + by) * 8;
                idct8x8(coef, qt[c.tq], &c.plane[(size_t)py * (size_t)c.bw * 8 + (size_t)px], c.bw * 8);
              }
          if (restart) --until_restart;
        }
      // colour conversion with replicated chroma
      out.assign((size_t)width * (size_t)height * channels_out, 255);
      const bool rgb_direct = comps.size() == 3 && adobe_transform == 0;
      for (int y = 0; y < height; ++y)
        for (int x = 0; x < width; ++x) {
          double s[3] = {0, 128, 128};
          for (size_t c = 0; c < comps.size(); ++c) {
            const Component &cp = comps[c];
            const int stride = cp.bw * 8;
            if (cp.h == hmax && cp.v == vmax) {
              s[c] = cp.plane[(size_t)y * (size_t)stride + (size_t)x];
              continue;
            }
            // subsampled plane: linear interpolation between sample centres (libjpeg's "fancy" triangle filter up to
            // its integer rounding); the plane's valid extent is the subsampled image size
            const int cw = (width * cp.h + hmax - 1) / hmax, ch = (height * cp.v + vmax - 1) / vmax;
            const double u = (x + 0.5) * cp.h / hmax - 0.5, v = (y + 0.5) * cp.v / vmax - 0.5;
            int x0 = (int)std::floor(u), y0 = (int)std::floor(v);
            const double fx = u - x0, fy = v - y0;
            auto at = [&](int xx, int yy) {
              xx = xx < 0 ? 0 : xx >= cw ? cw - 1 : xx;
              yy = yy < 0 ? 0 : yy >= ch ? ch - 1 : yy;
              return (double)cp.plane[(size_t)yy * (size_t)stride + (size_t)xx];
            };
            s[c] = (at(x0, y0) * (1 - fx) + at(x0 + 1, y0) * fx) * (1 - fy) + (at(x0, y0 + 1) * (1 - fx) + at(x0 + 1, y0 + 1) * fx) * fy;
          }
          uint8_t R, G, B;
          if (comps.size() == 1) R = G = B = clamp8(s[0]);
          else if (rgb_direct) { R = clamp8(s[0]); G = clamp8(s[1]); B = clamp8(s[2]); }
          else {
            R = clamp8(s[0] + 1.402 * (s[2] - 128.0));
            G = clamp8(s[0] - 0.344136 * (s[1] - 128.0) - 0.714136 * (s[2] - 128.0));
            B = clamp8(s[0] + 1.772 * (s[1] - 128.0));
          }
          uint8_t *o = &out[((size_t)y * (size_t)width + (size_t)x) * channels_out];
          if (channels_out == 1) o[0] = comps.size() == 1 ? R : (uint8_t)((2126u * R + 7152u * G + 722u * B) / 10000u);  // image crate luma weights
          else { o[0] = R; o[1] = G; o[2] = B; }
        }
      *w_out = (uint32_t)width;
      *h_out = (uint32_t)height;
      done = true;
    }
    pos += len;
  }
  if (!done) return bad("JPEG without image data");
  return PT_OK;
}

// Texture decode by content: PNG or JPEG (image::open guesses the format the same way, from the signature).
int image_decode(const uint8_t *bytes, size_t size, const char *name, uint32_t channels_out, uint32_t *w, uint32_t *h,
                 std::vector<uint8_t> &out) {
  if (size >= 2 && bytes[0] == 0xFF && bytes[1] == 0xD8) return jpeg_decode(bytes, size, name, channels_out, w, h, out);
  return png_decode(bytes, size, name, channels_out, w, h, out);
}

}  // namespace ptb
