// host_common.hpp — host-side helpers shared by the scene builder and the scene compiler.
// Product code (no oracle dependency).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/b200rt.h"

namespace b200rt {

struct D3 {
  double x = 0, y = 0, z = 0;
  D3() {}
  D3(double a, double b, double c) : x(a), y(b), z(c) {}
  explicit D3(const double* p) : x(p[0]), y(p[1]), z(p[2]) {}
  explicit D3(const float* p) : x(p[0]), y(p[1]), z(p[2]) {}
  double operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
inline D3 operator+(D3 a, D3 b) { return D3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline D3 operator-(D3 a, D3 b) { return D3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline D3 operator*(D3 a, double s) { return D3(a.x * s, a.y * s, a.z * s); }
inline D3 operator*(double s, D3 a) { return D3(a.x * s, a.y * s, a.z * s); }
inline D3 operator/(D3 a, double s) { return D3(a.x / s, a.y / s, a.z / s); }
inline double dot(D3 a, D3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline D3 cross(D3 l, D3 r) { return D3(l.y * r.z - l.z * r.y, l.z * r.x - l.x * r.z, l.x * r.y - l.y * r.x); }
inline double length(D3 a) { return std::sqrt(dot(a, a)); }
inline D3 unit(D3 a) { double l = length(a); return l > 0 ? a / l : a; }

// Row-major 3x4 affine matrix (the 4x4 of transform.rs without its constant last row).
struct Affine {
  double m[12];
  static Affine identity() {
    Affine a;
    const double I[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
    std::memcpy(a.m, I, sizeof(I));
    return a;
  }
  D3 point(D3 p) const {
    return D3(m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3], m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7],
              m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11]);
  }
  D3 dir(D3 p) const {
    return D3(m[0] * p.x + m[1] * p.y + m[2] * p.z, m[4] * p.x + m[5] * p.y + m[6] * p.z,
              m[8] * p.x + m[9] * p.y + m[10] * p.z);
  }
  Affine operator*(const Affine& b) const {  // this * b
    Affine r;
    for (int i = 0; i < 3; ++i) {
      for (int j = 0; j < 4; ++j) {
        double s = 0;
        for (int k = 0; k < 3; ++k) s += m[i * 4 + k] * b.m[k * 4 + j];
        if (j == 3) s += m[i * 4 + 3];
        r.m[i * 4 + j] = s;
      }
    }
    return r;
  }
  bool inverse(Affine& out) const {
    double a = m[0], b = m[1], c = m[2], d = m[4], e = m[5], f = m[6], g = m[8], h = m[9], i = m[10];
    double det = a * (e * i - f * h) - b * (d * i - f * g) + c * (d * h - e * g);
    if (det == 0. || !std::isfinite(det)) return false;
    double id = 1.0 / det;
    double r[9] = {(e * i - f * h) * id, (c * h - b * i) * id, (b * f - c * e) * id,
                   (f * g - d * i) * id, (a * i - c * g) * id, (c * d - a * f) * id,
                   (d * h - e * g) * id, (b * g - a * h) * id, (a * e - b * d) * id};
    for (int row = 0; row < 3; ++row) {
      out.m[row * 4 + 0] = r[row * 3 + 0];
      out.m[row * 4 + 1] = r[row * 3 + 1];
      out.m[row * 4 + 2] = r[row * 3 + 2];
      out.m[row * 4 + 3] = -(r[row * 3 + 0] * m[3] + r[row * 3 + 1] * m[7] + r[row * 3 + 2] * m[11]);
    }
    return true;
  }
  // linear part diagonal with positive entries: axis-aligned shapes stay axis-aligned
  bool is_positive_diagonal() const {
    return m[1] == 0 && m[2] == 0 && m[4] == 0 && m[6] == 0 && m[8] == 0 && m[9] == 0 && m[0] > 0 && m[5] > 0 && m[10] > 0;
  }
  // linear part = s * R with R orthonormal, s > 0 (returns s, or 0 when not a similarity)
  double similarity_scale() const {
    D3 c0(m[0], m[4], m[8]), c1(m[1], m[5], m[9]), c2(m[2], m[6], m[10]);
    double s0 = length(c0), s1 = length(c1), s2 = length(c2);
    double s = (s0 + s1 + s2) / 3;
    if (!(s > 0)) return 0;
    double tol = 1e-9 * s;
    if (std::fabs(s0 - s) > tol || std::fabs(s1 - s) > tol || std::fabs(s2 - s) > tol) return 0;
    if (std::fabs(dot(c0, c1)) > tol * s || std::fabs(dot(c0, c2)) > tol * s || std::fabs(dot(c1, c2)) > tol * s) return 0;
    if (dot(cross(c0, c1), c2) <= 0) return 0;
    return s;
  }
};

// Scene-construction RNG (PCG32, O'Neill 2014): the scene generators are deterministic in `seed`.
struct Pcg32 {
  uint64_t state, inc;
  explicit Pcg32(uint64_t seed, uint64_t seq = 54u) {
    state = 0; inc = (seq << 1u) | 1u;
    next_u32(); state += seed; next_u32();
  }
  uint32_t next_u32() {
    uint64_t old = state;
    state = old * 6364136223846793005ULL + inc;
    uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u), rot = (uint32_t)(old >> 59u);
    return (xs >> rot) | (xs << ((-rot) & 31));
  }
  double next() { return (next_u32() >> 8) * (1.0 / 16777216.0); }
  double range(double lo, double hi) { return lo + (hi - lo) * next(); }
};

}  // namespace b200rt

// The builder behind b200rt_builder_* : accumulates the flat tables of b200rt_scene_desc.
struct b200rt_builder {
  std::vector<b200rt_prim> prims;
  std::vector<float> tri_normals;
  std::vector<b200rt_material> materials;
  std::vector<b200rt_texture> textures;
  std::vector<b200rt_image> images;
  std::vector<uint8_t> image_data;
  std::vector<b200rt_instance> instances;
  std::vector<int32_t> lights;
  std::vector<float> light_scales;  // one per entry of `lights`
  b200rt_perlin perlin;
  bool has_perlin = false;
  uint32_t flags = 0;
  std::vector<b200rt::Affine> xf_stack;  // composed matrices
  std::vector<int32_t> xf_instance;      // instance id of each stack level
  std::string err;
};
