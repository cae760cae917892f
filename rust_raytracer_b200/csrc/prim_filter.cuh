// prim_filter.cuh — f32 candidate filters: each returns true only when the primitive CERTAINLY does not
// produce a hit in [t_min, t_best] (forward error bounds of the f32 evaluation are subtracted before every
// comparison, u = 2^-24).  Survivors are decided by prim_exact.cuh.  This is where the FP32 pipe does the bulk
// of the leaf work; the bounds are generous on purpose (a false "maybe" only costs one exact test).
#pragma once
#include "device_math.cuh"

// Per-ray constants shared by all tests.
struct RayF {
  float3 o, d, inv;  // inv = 1/d with |d| clamped away from 0 (no NaN in the slab test)
  float a;           // d . d
  float time;
};

// Approximate unit operations (MUFU.RCP / MUFU.RSQ, <= 2 ulp) for the FILTERS only: every comparison below carries
// an explicit error margin several ulp wide, and a wrong "maybe" only costs one exact test.  The IEEE-rounded
// forms cost ~8 instructions each and sat on every ray's critical path (three reciprocals in make_rayf).
DEV float rcp_fast(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
DEV float rsqrt_fast(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
DEV float sqrt_fast(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
// 1/d for the slab tests.  A (near-)zero component gets +-2^100: a power of two, so plane * inv and o * inv are exact and
// fmaf(plane, inv, -o * inv) has the exact sign of plane - o: on that axis the slab test degenerates to the exact
// inside / outside test of the origin and needs no rounding slack (make_rayt).
#define B200RT_RCP_CLAMP 1e-30f
DEV float safe_rcp(float d) { return fabsf(d) > B200RT_RCP_CLAMP ? rcp_fast(d) : copysignf(0x1p100f, d); }

// Filter verdicts
#define FILTER_MAYBE 0        // the exact test decides
#define FILTER_REJECT 1       // certain miss
#define FILTER_MEDIUM_HIT 2   // ConstantMedium only: certain scatter before the boundary exit / the best hit, origin inside

// ConstantMedium::hit (medium.rs:25-53) in f32.  [t_in, t_out] is the boundary interval (errors e_in, e_out):
// distance_inside <= (min(t_out, t_best) - max(t_in, t_min, 0)) * |d|; the scattering distance
// hit_distance = -1/density * ln(xi) is known to ~3 ulp, so when it exceeds that upper bound the medium is a
// certain miss and the f64 boundary roots / log never run.  Conversely, when the ray certainly starts inside
// (t_in < t_min) and hit_distance is certainly below the LOWER bound of distance_inside, the outcome is a hit at
// t = max(t_min, 0) + hit_distance / |d| whatever the exact boundary roots are, and certainly closer than the best
// hit so far: no decision depends on its low bits, so t itself is taken in f32 (north_star: t within 1e-5) and the
// whole f64 stage (boundary roots, f64 log / sqrt / division) is skipped for half of all book-2 rays.
DEV int medium_far_reject(const RayF& r, float t_in, float e_in, float t_out, float e_out, float neg_inv_density, float xi, float t_min, float t_best, float* t_hit) {
  const float lo = fmaxf(fmaxf(t_in - e_in, t_min), 0.f);
  const float hi = fminf(t_out + e_out, t_best);
  if (!(r.a <= 3.0e38f)) return FILTER_MAYBE;  // |d|^2 overflowed: no f32 verdict
  const float len = sqrt_fast(r.a);  // <= 2 ulp, inside the 2e-6 margins below
  float upper = fmaxf(hi - lo, 0.f) + 2e-7f * (fabsf(hi) + fabsf(lo));
  upper = upper * len * (1.0f + 2e-6f);
  const float hd = neg_inv_density * logf(xi);  // >= 0
  if (hd * (1.0f - 2e-6f) > upper) return FILTER_REJECT;  // NaN/inf upper (t_best = inf, unbounded boundary) never rejects
  if (t_in + e_in < t_min) {
    const float lo1 = fmaxf(t_min, 0.f);                                    // exact: t1 after both clamps
    const float hi1 = fminf(t_out - e_out, t_best * (1.0f - 2.4e-7f));      // lower bound of min(t_out, t_max); t_best is rounded up
    const float lower = ((hi1 - lo1) - 2e-7f * (fabsf(hi1) + fabsf(lo1))) * len * (1.0f - 2e-6f);
    if (hi1 > lo1 * (1.0f + 1e-6f) + 1e-30f && hd * (1.0f + 2e-6f) < lower) {
      *t_hit = lo1 + hd / sqrtf(r.a);  // medium.rs:52 in f32: ~4 ulp, against the 1e-5 tolerance on t
      return FILTER_MEDIUM_HIT;
    }
  }
  return FILTER_MAYBE;
}

// sphere.rs:24-43 in f32: disc error <= ~20u * a * (|oc|^2 + r^2)
// eps: absolute error of the baked world-space coordinates of an instanced primitive (0 when not instanced):
// the exact test then runs in OBJECT space on the original values, so the filter must tolerate the baking error.
// medium = true: the sphere is a ConstantMedium boundary (both roots matter, then the distance test).
DEV int sphere_reject(const RayF& r, float3 c, float radius, float t_min, float t_best, float eps, bool medium = false, float neg_inv_density = 0.f, float xi = 0.5f, float* t_hit = nullptr) {
  float3 oc = r.o - c;
  float hb = dot(oc, r.d), oc2 = dot(oc, oc), r2 = radius * radius;
  float cc = oc2 - r2;
  float disc = hb * hb - r.a * cc;
  float E = 2e-6f * r.a * (oc2 + r2);
  if (eps > 0.f) E += r.a * eps * (8.2f * sqrt_fast(oc2) + 4.0f * fabsf(radius));
  // An overflowed intermediate (coordinates around 1e10 and beyond: a * c and half_b^2 pass 3.4e38) must not decide anything:
  // with the product a * c at +inf and half_b * half_b exact inside the FMA, disc comes out -inf for a ray that hits.
  // The f64 stage has the range.
  if (!(fabsf(disc) <= 3.0e38f) || !(E <= 3.0e38f)) return false;
  if (disc < -E) return true;
  if (disc > 4.0f * E) {
    const float rs = rsqrt_fast(disc), s = disc * rs, inva = rcp_fast(r.a);  // s, inva: <= 3 ulp
    float dn = 6.2e-7f * sqrt_fast(oc2 * r.a) + 1.01f * E * rs;
    if (eps > 0.f) dn += 2.05f * eps * sqrt_fast(r.a);
    dn *= inva;
    float t0 = (-hb - s) * inva, t1 = (-hb + s) * inva;
    float e0 = dn + 1e-6f * fabsf(t0), e1 = dn + 1e-6f * fabsf(t1);
    if (t0 - e0 > t_best) return true;                       // nearest root already beyond the best hit
    if (t1 + e1 < t_min) return true;                        // sphere entirely behind t_min
    if (medium) return medium_far_reject(r, t0, e0, t1, e1, neg_inv_density, xi, t_min, t_best, t_hit);
    if (t0 + e0 < t_min && t1 - e1 > t_best) return true;    // origin inside, exit beyond the best hit
  }
  return false;
}

// rectangle.rs:15-26 in f32
DEV bool rect_reject(const RayF& r, int axis, float a0, float b0, float a1, float b1, float k, float t_min, float t_best, float eps) {
  const int ia = axis == 0 ? 1 : 0, ib = axis == 2 ? 1 : 2;
  const float inv = comp(r.inv, axis);
  float t = (k - comp(r.o, axis)) * inv;
  float e = 6e-7f * fabsf(t) + 2.0f * eps * fabsf(inv);
  if (t + e < t_min || t - e > t_best) return true;
  float oa = comp(r.o, ia), da = comp(r.d, ia), ob = comp(r.o, ib), db = comp(r.d, ib);
  float x = fmaf(t, da, oa), y = fmaf(t, db, ob);
  float ex = 4e-7f * (fabsf(oa) + fabsf(t * da)) + e * fabsf(da) + 2.0f * eps;
  float ey = 4e-7f * (fabsf(ob) + fabsf(t * db)) + e * fabsf(db) + 2.0f * eps;
  if (x < a0 - ex || x > a1 + ex || y < b0 - ey || y > b1 + ey) return true;
  return false;  // includes the NaN / parallel case: the exact test decides
}

// Cube (cube.rs:63-66): the six rectangle hits are exactly the slab entry and exit points.
// `cand` (optional) returns the faces that can produce the closest valid hit: bits 0..5 = sides whose plane
// parameter lies within the f32 error of the slab ENTRY point, bits 8..13 = sides near the EXIT point (side
// numbering of cube_side: XY@min.z, XY@max.z, XZ@min.y, XZ@max.y, YZ@min.x, YZ@max.x).  A face whose plane hit
// is neither is out of its rectangle bounds for certain, so the exact test may skip it.
DEV int box_reject(const RayF& r, float3 mn, float3 mx, float t_min, float t_best, float eps, uint32_t* cand = nullptr, bool medium = false, float neg_inv_density = 0.f,
                   float xi = 0.5f, float* t_hit = nullptr) {
  float3 t0 = (mn - r.o) * r.inv, t1 = (mx - r.o) * r.inv;
  float tn = fmaxf(fmaxf(fminf(t0.x, t1.x), fminf(t0.y, t1.y)), fminf(t0.z, t1.z));
  float tf = fminf(fminf(fmaxf(t0.x, t1.x), fmaxf(t0.y, t1.y)), fmaxf(t0.z, t1.z));
  float es = eps > 0.f ? 2.0f * eps * fmaxf(fabsf(r.inv.x), fmaxf(fabsf(r.inv.y), fabsf(r.inv.z))) : 0.f;
  float en = 1e-6f * fabsf(tn) + es, ef = 1e-6f * fabsf(tf) + es;
  if (tn - en > tf + ef) return true;
  if (tf + ef < t_min) return true;
  if (tn - en > t_best) return true;
  if (medium) return medium_far_reject(r, tn, en, tf, ef, neg_inv_density, xi, t_min, t_best, t_hit);  // ConstantMedium boundary
  if (tn + en < t_min && tf - ef > t_best) return true;
  if (cand) {
    const float ts[6] = {t0.z, t1.z, t0.y, t1.y, t0.x, t1.x};
    uint32_t cn = 0, cf = 0;
#pragma unroll
    for (int s = 0; s < 6; ++s) {
      const float e = 1e-6f * fabsf(ts[s]) + es;
      const bool in_range = (ts[s] + e >= t_min) && (ts[s] - e <= t_best);
      const bool near_n = fabsf(ts[s] - tn) <= 2.0f * (e + en), near_f = fabsf(ts[s] - tf) <= 2.0f * (e + ef);
      if (in_range && near_n) cn |= 1u << s;
      else if (in_range && near_f) cf |= 1u << s;
    }
    if (!(cn | cf)) return true;
    *cand = cn | (cf << 8);
  }
  return false;
}
// f32 plane parameter of one cube side (for ordering decisions in the exact stage)
DEV float box_side_t(const RayF& r, float3 mn, float3 mx, int side) {
  const int axis = 2 - (side >> 1);
  const float k = (side & 1) ? comp(mx, axis) : comp(mn, axis);
  return (k - comp(r.o, axis)) * comp(r.inv, axis);
}

// mesh.rs:56-101 in f32 with condition-number-scaled margins (q = |d||e0||e1| / |det|)
DEV bool tri_reject(const RayF& r, float3 v0, float3 v1, float3 v2, float t_min, float t_best, float eps) {
  float3 e0 = v1 - v0, e1 = v2 - v0;
  float3 s0 = cross(r.d, e1);
  float dd = dot(s0, e0);
  float e0_2 = dot(e0, e0), e1_2 = dot(e1, e1);
  float scale2 = r.a * e0_2 * e1_2;
  if (!(dd * dd > 1e-8f * scale2)) return false;  // grazing or degenerate: let the exact test decide
  float inv = rcp_fast(dd);
  float cond = sqrt_fast(scale2) * fabsf(inv);
  float rel = eps * rsqrt_fast(fminf(e0_2, e1_2));  // baking error relative to the shorter edge
  if (cond * rel > 0.05f) return false;
  float q = cond * (1.5e-6f + 8.0f * rel);
  float3 dv = r.o - v0;
  float dv2 = dot(dv, dv);
  float b1 = dot(dv, s0) * inv;
  float db1 = q * (sqrt_fast(dv2 * rcp_fast(e0_2)) + fabsf(b1));
  if (b1 < -db1 || b1 > 1.0f + db1) return true;
  float3 s1 = cross(dv, e0);
  float b2 = dot(r.d, s1) * inv;
  float db2 = q * (sqrt_fast(dv2 * rcp_fast(e1_2)) + fabsf(b2));
  if (b2 < -db2 || b1 + b2 > 1.0f + db1 + db2) return true;
  float t = dot(e1, s1) * inv;
  float dt = q * (sqrt_fast(dv2 * rcp_fast(r.a)) + fabsf(t));
  if (t + dt < t_min || t - dt > t_best) return true;
  return false;
}
