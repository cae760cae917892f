// device_math.cuh — float3 / double3 helpers, the counter-based RNG and the sampling transforms.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define DEV __device__ __forceinline__
#define PI_F 3.14159265358979323846f
#define PI_D 3.14159265358979323846

// ---- float3 ---------------------------------------------------------------
DEV float3 f3(float x, float y, float z) { return make_float3(x, y, z); }
DEV float3 f3(float4 v) { return make_float3(v.x, v.y, v.z); }
DEV float3 operator+(float3 a, float3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
DEV float3 operator-(float3 a, float3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
DEV float3 operator-(float3 a) { return f3(-a.x, -a.y, -a.z); }
DEV float3 operator*(float3 a, float s) { return f3(a.x * s, a.y * s, a.z * s); }
DEV float3 operator*(float s, float3 a) { return f3(a.x * s, a.y * s, a.z * s); }
DEV float3 operator*(float3 a, float3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }  // Vec3::elemul (vec3.rs:66-72)
DEV float dot(float3 a, float3 b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z)); }
DEV float3 cross(float3 l, float3 r) { return f3(l.y * r.z - l.z * r.y, l.z * r.x - l.x * r.z, l.x * r.y - l.y * r.x); }
DEV float length(float3 a) { return sqrtf(dot(a, a)); }
DEV float3 normalize(float3 a) { return a * rsqrtf(dot(a, a)); }
DEV float comp(float3 v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : v.z); }
DEV float maxcomp(float3 v) { return fmaxf(v.x, fmaxf(v.y, v.z)); }

// ---- double3 (exact candidate confirmation; B200 issues FP64 at half the FP32 rate) ------------
struct d3 { double x, y, z; };
DEV d3 D3(double x, double y, double z) { d3 r; r.x = x; r.y = y; r.z = z; return r; }
DEV d3 D3(float3 v) { return D3((double)v.x, (double)v.y, (double)v.z); }
DEV d3 D3(float4 v) { return D3((double)v.x, (double)v.y, (double)v.z); }
DEV d3 operator+(d3 a, d3 b) { return D3(a.x + b.x, a.y + b.y, a.z + b.z); }
DEV d3 operator-(d3 a, d3 b) { return D3(a.x - b.x, a.y - b.y, a.z - b.z); }
DEV d3 operator-(d3 a) { return D3(-a.x, -a.y, -a.z); }
DEV d3 operator*(d3 a, double s) { return D3(a.x * s, a.y * s, a.z * s); }
DEV d3 operator*(double s, d3 a) { return D3(a.x * s, a.y * s, a.z * s); }
DEV d3 operator/(d3 a, double s) { return D3(a.x / s, a.y / s, a.z / s); }
// No FMA contraction here: the oracle (and Rust) evaluate x*x + y*y + z*z with separate roundings.
DEV double ddot(d3 a, d3 b) { return __dadd_rn(__dadd_rn(__dmul_rn(a.x, b.x), __dmul_rn(a.y, b.y)), __dmul_rn(a.z, b.z)); }
DEV d3 dcross(d3 l, d3 r) {
  return D3(__dsub_rn(__dmul_rn(l.y, r.z), __dmul_rn(l.z, r.y)), __dsub_rn(__dmul_rn(l.z, r.x), __dmul_rn(l.x, r.z)),
            __dsub_rn(__dmul_rn(l.x, r.y), __dmul_rn(l.y, r.x)));
}
DEV d3 dat(d3 o, double t, d3 d) { return D3(__dadd_rn(o.x, __dmul_rn(t, d.x)), __dadd_rn(o.y, __dmul_rn(t, d.y)), __dadd_rn(o.z, __dmul_rn(t, d.z))); }
DEV double dcomp(d3 v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : v.z); }
DEV float3 to_f3(d3 v) { return f3((float)v.x, (float)v.y, (float)v.z); }

// ---- Philox4x32-10 (Salmon et al., SC'11), keyed (seed) with counter (pixel, sample, bounce, stream) ----------
DEV uint4 philox4x32(uint4 c, uint2 k) {
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += 0x9E3779B9u;
    k.y += 0xBB67AE85u;
  }
  return c;
}
// uniform in (0,1), 23 bits, exactly representable in f32 and f64
DEV float u01(uint32_t x) { return ((float)(x >> 9) + 0.5f) * (1.0f / 8388608.0f); }

// The draw ConstantMedium::hit takes inside `hit` (medium.rs:37-38), as a pure function of (ray seed, prim index).
DEV uint32_t pcg_hash(uint32_t v) {
  uint32_t state = v * 747796405u + 2891336453u;
  uint32_t word = ((state >> ((state >> 28u) + 4u)) ^ state) * 277803737u;
  return (word >> 22u) ^ word;
}
DEV double medium_xi(uint32_t seed, uint32_t prim) {
  uint32_t h = pcg_hash(seed ^ (prim * 0x9E3779B9u));
  return ((double)(h >> 9) + 0.5) * (1.0 / 8388608.0);
}

DEV float medium_xi_f32(uint32_t seed, uint32_t prim) {  // the same draw, exactly representable in f32
  uint32_t h = pcg_hash(seed ^ (prim * 0x9E3779B9u));
  return ((float)(h >> 9) + 0.5f) * (1.0f / 8388608.0f);
}

// ---- sampling transforms (distribution-equal to the reference's rejection loops, vec3.rs:111-162) -------------
// uniform on the unit sphere == Vec3::random_in_unit_sphere() (Marsaglia, returns a point ON the sphere)
DEV float3 sample_unit_sphere(float u1, float u2) {
  float z = 1.0f - 2.0f * u1;
  float r = sqrtf(fmaxf(0.0f, 1.0f - z * z));
  float s, c;
  sincospif(2.0f * u2, &s, &c);
  return f3(r * c, r * s, z);
}
// uniform in the unit disk == Vec3::random_in_unit_disk()
DEV float2 sample_unit_disk(float u1, float u2) {
  float r = sqrtf(u1), s, c;
  sincospif(2.0f * u2, &s, &c);
  return make_float2(r * c, r * s);
}
DEV void onb_from_w(float3 w, float3& u, float3& v) {
  float3 a = fabsf(w.x) > 0.9f ? f3(0, 1, 0) : f3(1, 0, 0);
  v = normalize(cross(w, a));
  u = cross(w, v);
}
DEV float3 sample_cosine_hemisphere(float3 n, float u1, float u2) {
  float3 u, v;
  onb_from_w(n, u, v);
  float s, c;
  sincospif(2.0f * u1, &s, &c);
  float r = sqrtf(u2);
  return (c * r) * u + (s * r) * v + sqrtf(fmaxf(0.0f, 1.0f - u2)) * n;
}
