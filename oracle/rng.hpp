// oracle/rng.hpp — TEST INFRASTRUCTURE (CPU oracle), not product code.
// Two random sources behind one "named draw" interface:
//  * ChaCha12 + PCG32 seed expansion = rand 0.8 `StdRng::seed_from_u64` + `gen::<f32>()`
//    (third-party crate `rand ^0.8.5`, Cargo.toml:20; call sites renderer/mod.rs:110-118,201,
//    brdf/cook_torrance.rs:123-124, renderer/utils.rs:30).  Published algorithm restated
//    (SURVEY.md Appendix B); draws are consumed strictly in call order.
//  * Philox4x32-10 keyed by (pixel, sample): the counter-based generator the CUDA path
//    uses; draws are addressed by name so both sides agree without sharing state.
#pragma once
#include <cstdint>
#include <cstring>

namespace orc {

static inline float u32_to_unit_f32(uint32_t x) {  // rand 0.8 Standard for f32
  return (float)(x >> 8) * (1.0f / 16777216.0f);
}

struct ChaCha12 {
  uint32_t key[8];
  uint64_t counter = 0;
  uint32_t buf[16];
  int idx = 16;

  static inline uint32_t rotl(uint32_t v, int n) { return (v << n) | (v >> (32 - n)); }
  void seed_from_u64(uint64_t state) {
    const uint64_t MUL = 6364136223846793005ULL, INC = 11634580027462260723ULL;
    for (int i = 0; i < 8; ++i) {
      state = state * MUL + INC;
      uint32_t xorshifted = (uint32_t)(((state >> 18) ^ state) >> 27);
      uint32_t rot = (uint32_t)(state >> 59);
      key[i] = (xorshifted >> rot) | (xorshifted << ((32 - rot) & 31));
    }
    counter = 0;
    idx = 16;
  }
  void block() {
    uint32_t s[16] = {0x61707865u, 0x3320646eu, 0x79622d32u, 0x6b206574u};
    for (int i = 0; i < 8; ++i) s[4 + i] = key[i];
    s[12] = (uint32_t)counter;
    s[13] = (uint32_t)(counter >> 32);
    s[14] = 0;
    s[15] = 0;
    uint32_t x[16];
    std::memcpy(x, s, sizeof x);
#define ORC_QR(a, b, c, d)                    \
  x[a] += x[b]; x[d] = rotl(x[d] ^ x[a], 16); \
  x[c] += x[d]; x[b] = rotl(x[b] ^ x[c], 12); \
  x[a] += x[b]; x[d] = rotl(x[d] ^ x[a], 8);  \
  x[c] += x[d]; x[b] = rotl(x[b] ^ x[c], 7);
    for (int r = 0; r < 6; ++r) {  // 12 rounds = 6 double rounds
      ORC_QR(0, 4, 8, 12) ORC_QR(1, 5, 9, 13) ORC_QR(2, 6, 10, 14) ORC_QR(3, 7, 11, 15)
      ORC_QR(0, 5, 10, 15) ORC_QR(1, 6, 11, 12) ORC_QR(2, 7, 8, 13) ORC_QR(3, 4, 9, 14)
    }
#undef ORC_QR
    for (int i = 0; i < 16; ++i) buf[i] = x[i] + s[i];
    ++counter;
    idx = 0;
  }
  uint32_t next_u32() {
    if (idx >= 16) block();
    return buf[idx++];
  }
};

// Philox4x32-10 (Salmon et al. 2011), same constants as curand_philox4x32_x.h.
static inline void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
  for (int r = 0; r < 10; ++r) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c[0];
    uint64_t p1 = (uint64_t)0xCD9E8D57u * c[2];
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
    uint32_t n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
    uint32_t n3 = (uint32_t)p0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
}

enum RngMode { RNG_CHACHA = 0, RNG_PHILOX = 1 };

// Named draws.  Philox addressing (must match path-tracer_b200/csrc/shade.cuh):
//   key = (pixel, sample), counter = (block, tag, seed_lo, seed_hi)
//   jitter          : tag 0, block 0 -> out[0], out[1]
//   bounce b        : tag b+1, block 0 -> out[0]=r1, out[1]=r2, out[2]=rr
//   alpha draw of the examined hit with global primitive id p at bounce b: tag 0x80000000|(b+1), block p -> out[0]
// The reference's alpha walk (mod.rs:188-205) draws from the path's sequential stream for every examined hit with
// 0.001 < opacity < 1 (the ChaCha mode below does exactly that).  In Philox mode the draw is addressed by the hit's
// primitive id instead of its position in the list: every examined hit still gets its own independent uniform draw,
// so the accepted hit has the same distribution, and the decision for one hit no longer depends on the hits before
// it (which lets the CUDA path decide during ONE traversal).  tests/test_oracle_kat.py checks the two modes against
// each other on the converged alpha_transparency image.
struct PathRng {
  RngMode mode;
  ChaCha12 cc;
  uint32_t pixel = 0, sample = 0, seed_lo = 0, seed_hi = 0;
  uint32_t cache[4] = {0, 0, 0, 0};
  uint32_t cache_tag = 0xffffffffu, cache_block = 0xffffffffu;

  void init_chacha(uint64_t seed) {
    mode = RNG_CHACHA;
    cc.seed_from_u64(seed);
  }
  void init_philox(uint32_t px, uint32_t s, uint64_t seed) {
    mode = RNG_PHILOX;
    pixel = px; sample = s;
    seed_lo = (uint32_t)seed; seed_hi = (uint32_t)(seed >> 32);
    cache_tag = cache_block = 0xffffffffu;
  }
  float philox(uint32_t tag, uint32_t block, int lane) {
    if (tag != cache_tag || block != cache_block) {
      cache[0] = block; cache[1] = tag; cache[2] = seed_lo; cache[3] = seed_hi;
      philox4x32_10(cache, pixel, sample);
      cache_tag = tag; cache_block = block;
    }
    return u32_to_unit_f32(cache[lane]);
  }
  float jitter(int axis) { return mode == RNG_CHACHA ? u32_to_unit_f32(cc.next_u32()) : philox(0, 0, axis); }
  float alpha(uint32_t bounce, uint32_t prim_id) {
    if (mode == RNG_CHACHA) return u32_to_unit_f32(cc.next_u32());
    return philox(0x80000000u | (bounce + 1), prim_id, 0);
  }
  float bsdf(uint32_t bounce, int which) {
    return mode == RNG_CHACHA ? u32_to_unit_f32(cc.next_u32()) : philox(bounce + 1, 0, which);
  }
  float rr(uint32_t bounce) { return mode == RNG_CHACHA ? u32_to_unit_f32(cc.next_u32()) : philox(bounce + 1, 0, 2); }
};

}  // namespace orc
