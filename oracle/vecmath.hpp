// oracle/vecmath.hpp — TEST INFRASTRUCTURE (CPU oracle), not product code.
// f32 vector algebra in cgmath-0.18 operand order:
//   dot(a,b)   = (a.x*b.x + a.y*b.y) + a.z*b.z
//   cross(a,b) = (a.y*b.z - a.z*b.y, a.z*b.x - a.x*b.z, a.x*b.y - a.y*b.x)
//   normalize  = v * (1 / sqrt(dot(v,v)))
// Compile with -ffp-contract=off: Rust never contracts to FMA.
#pragma once
#include <cmath>
#include <cstdint>

namespace orc {

struct V2 {
  float x, y;
};
struct V3 {
  float x, y, z;
};

static inline V3 v3(float x, float y, float z) { return V3{x, y, z}; }
static inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
static inline V3 operator*(float s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
static inline V3 operator/(V3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
static inline V3 operator-(V3 a) { return {-a.x, -a.y, -a.z}; }
static inline V3 mul_el(V3 a, V3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
static inline V3 div_el(V3 a, V3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
static inline float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
static inline V3 cross(V3 a, V3 b) {
  return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
static inline float magnitude(V3 a) { return std::sqrt(dot(a, a)); }
static inline V3 normalize(V3 a) { return a * (1.0f / magnitude(a)); }
static inline float sum(V3 a) { return (a.x + a.y) + a.z; }

static inline V2 operator+(V2 a, V2 b) { return {a.x + b.x, a.y + b.y}; }
static inline V2 operator-(V2 a, V2 b) { return {a.x - b.x, a.y - b.y}; }
static inline V2 operator*(float s, V2 a) { return {s * a.x, s * a.y}; }

// Rust f32::max / f32::min: return the non-NaN operand.
static inline float rmax(float a, float b) { return std::fmax(a, b); }
static inline float rmin(float a, float b) { return std::fmin(a, b); }

// llvm.powi.f32 (compiler-rt __powisf2, square-and-multiply).
static inline float powi(float a, int b) {
  const bool recip = b < 0;
  float r = 1.0f;
  while (true) {
    if (b & 1) r *= a;
    b /= 2;
    if (b == 0) break;
    a *= a;
  }
  return recip ? 1.0f / r : r;
}

// Rust `f as i64`: truncate toward zero, saturate, NaN -> 0.
static inline int64_t f32_as_i64(float f) {
  if (f != f) return 0;
  if (f >= 9.2233720368547758e18f) return INT64_MAX;
  if (f <= -9.2233720368547758e18f) return INT64_MIN;
  return (int64_t)f;
}
static inline int64_t rem_euclid(int64_t a, int64_t m) {
  int64_t r = a % m;
  return r < 0 ? r + m : r;
}
// Rust `f as u8`.
static inline uint8_t f32_as_u8(float f) {
  if (f != f) return 0;
  if (f <= 0.0f) return 0;
  if (f >= 255.0f) return 255;
  return (uint8_t)f;
}

}  // namespace orc
