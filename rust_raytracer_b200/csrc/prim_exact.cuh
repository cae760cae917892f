// prim_exact.cuh — double-precision primitive tests that follow the reference's arithmetic operation by
// operation (no FMA contraction), evaluated from the f32 scene/ray values.  They DECIDE every accepted hit:
// the f32 code in prim_filter.cuh only throws away candidates that certainly miss.  (north_star: "f32 on
// device, f64 only where parity demands it"; on B200 FP64 issues at half the FP32 rate, so confirming the few
// surviving candidates per ray in f64 is affordable and gives |dt|/t ~ 1e-15 against the oracle.)
#pragma once
#include "device_math.cuh"
#include "device_scene.h"
#include "../../include/b200rt.h"

#define DINF (__longlong_as_double(0x7ff0000000000000LL))

struct RayD { d3 o, d; double time; };

// objects/sphere.rs:24-43 — nearest root in the CLOSED interval [t_min, t_max]; tangent rays hit (disc < 0 misses)
DEV bool sphere_exact(const RayD& r, d3 c, double radius, double t_min, double t_max, double& t) {
  d3 oc = r.o - c;
  double a = ddot(r.d, r.d);
  double half_b = ddot(oc, r.d);
  double cc = __dsub_rn(ddot(oc, oc), __dmul_rn(radius, radius));
  double disc = __dsub_rn(__dmul_rn(half_b, half_b), __dmul_rn(a, cc));
  if (disc < 0.) return false;
  double sqrt_d = sqrt(disc);
  double root = __ddiv_rn(__dsub_rn(-half_b, sqrt_d), a);
  if (!(root >= t_min && root <= t_max)) {
    root = __ddiv_rn(__dadd_rn(-half_b, sqrt_d), a);
    if (!(root >= t_min && root <= t_max)) return false;
  }
  t = root;
  return true;
}

// objects/rectangle.rs:15-34,53-72,90-109.  axis: 0 = YZ (k = x), 1 = XZ (k = y), 2 = XY (k = z).
// Deviation: a ray parallel to the plane gives t = inf/NaN upstream and falls through both
// rejections; here a non-finite t is a miss.
DEV bool rect_exact(const RayD& r, int axis, double a0, double b0, double a1, double b1, double k, double t_min, double t_max, double& t, double& pa, double& pb) {
  const int ia = axis == 0 ? 1 : 0, ib = axis == 2 ? 1 : 2;
  double tt = __ddiv_rn(__dsub_rn(k, dcomp(r.o, axis)), dcomp(r.d, axis));
  if (tt < t_min || tt > t_max) return false;
  if (!isfinite(tt)) return false;
  double a = __dadd_rn(dcomp(r.o, ia), __dmul_rn(tt, dcomp(r.d, ia)));
  double b = __dadd_rn(dcomp(r.o, ib), __dmul_rn(tt, dcomp(r.d, ib)));
  if (a < a0 || a > a1 || b < b0 || b > b1) return false;
  t = tt; pa = a; pb = b;
  return true;
}

// The rectangle `side` (0..5) of Cube::new (objects/cube.rs:17-54): XY@min.z, XY@max.z, XZ@min.y, XZ@max.y, YZ@min.x, YZ@max.x
DEV void cube_side(int side, d3 mn, d3 mx, int& axis, double& a0, double& b0, double& a1, double& b1, double& k) {
  axis = 2 - (side >> 1);
  const int ia = axis == 0 ? 1 : 0, ib = axis == 2 ? 1 : 2;
  a0 = dcomp(mn, ia); b0 = dcomp(mn, ib); a1 = dcomp(mx, ia); b1 = dcomp(mx, ib);
  k = (side & 1) ? dcomp(mx, axis) : dcomp(mn, axis);
}
// objects/cube.rs:63-66 -> hit.rs:56-67: linear closest hit over the six sides (a later side wins an exact tie)
DEV bool cube_exact(const RayD& r, d3 mn, d3 mx, double t_min, double t_max, double& t, int& side_out) {
  double closest = t_max;
  bool any = false;
  for (int s = 0; s < 6; ++s) {
    int axis; double a0, b0, a1, b1, k, tt, pa, pb;
    cube_side(s, mn, mx, axis, a0, b0, a1, b1, k);
    if (rect_exact(r, axis, a0, b0, a1, b1, k, t_min, closest, tt, pa, pb)) { closest = tt; side_out = s; any = true; }
  }
  t = closest;
  return any;
}

// The same closest-of-six result from the candidate faces only (prim_filter.cuh box_reject): entry-side
// candidates first; an exit-side candidate is skipped when its f32 plane parameter is certainly beyond the hit
// already accepted.  The accept rule (smaller t, or equal t and a later side) is cube.rs:63-66 -> hit.rs:56-67
// made order-independent, so skipping certain misses cannot change the winner.
template <typename SideLB>
DEV bool cube_exact_cand(const RayD& r, d3 mn, d3 mx, double t_min, double t_max, uint32_t cand, SideLB side_lower_bound32, double& t, int& side_out) {
  double best = t_max;
  int side = -1;
  uint32_t m = cand & 0x3fu;
  for (int pass = 0; pass < 2; ++pass) {
    while (m) {
      const int s = __ffs((int)m) - 1;
      m &= m - 1u;
      if (pass == 1 && side >= 0) {
        if ((double)side_lower_bound32(s) > best) continue;
      }
      int axis; double a0, b0, a1, b1, k, tt, pa, pb;
      cube_side(s, mn, mx, axis, a0, b0, a1, b1, k);
      if (rect_exact(r, axis, a0, b0, a1, b1, k, t_min, t_max, tt, pa, pb)) {
        if (side < 0 || tt < best || (tt == best && s > side)) { best = tt; side = s; }
      }
    }
    m = (cand >> 8) & 0x3fu;
  }
  if (side < 0) return false;
  t = best; side_out = side;
  return true;
}

// objects/mesh.rs:56-101 (Moeller-Trumbore, two-sided, inclusive bounds)
DEV bool tri_exact(const RayD& r, d3 pa, d3 pb, d3 pc, double t_min, double t_max, double& t, double& b1, double& b2) {
  d3 e0 = pb - pa, e1 = pc - pa;
  d3 s0 = dcross(r.d, e1);
  double dd = ddot(s0, e0);
  if (dd == 0.0) return false;
  double div = __ddiv_rn(1.0, dd);
  d3 d = r.o - pa;
  double bb1 = __dmul_rn(ddot(d, s0), div);
  if (bb1 < 0.0 || bb1 > 1.0) return false;
  d3 s1 = dcross(d, e0);
  double bb2 = __dmul_rn(ddot(r.d, s1), div);
  if (bb2 < 0.0 || __dadd_rn(bb1, bb2) > 1.0) return false;
  double tt = __dmul_rn(ddot(e1, s1), div);
  if (tt < t_min || tt > t_max) return false;
  if (!(tt == tt)) return false;
  t = tt; b1 = bb1; b2 = bb2;
  return true;
}

// objects/medium.rs:25-53 with a Sphere (kind 0) or Cube (kind 1) boundary; xi is the injected draw.
DEV bool medium_exact(const RayD& r, int kind, d3 a, d3 b, double radius, double neg_inv_density, double xi, double t_min, double t_max, double& t) {
  double t1, t2;
  int side;
  if (kind == 0) {
    if (!sphere_exact(r, a, radius, -DINF, DINF, t1)) return false;
    if (!sphere_exact(r, a, radius, __dadd_rn(t1, 0.0001), DINF, t2)) return false;
  } else {
    if (!cube_exact(r, a, b, -DINF, DINF, t1, side)) return false;
    if (!cube_exact(r, a, b, __dadd_rn(t1, 0.0001), DINF, t2, side)) return false;
  }
  t1 = fmax(t1, t_min);
  t2 = fmin(t2, t_max);
  if (t1 >= t2) return false;
  t1 = fmax(t1, 0.);
  double ray_length = sqrt(ddot(r.d, r.d));
  double distance_inside = __dmul_rn(__dsub_rn(t2, t1), ray_length);
  double hit_distance = __dmul_rn(neg_inv_density, log(xi));
  if (hit_distance > distance_inside) return false;
  t = __dadd_rn(t1, __ddiv_rn(hit_distance, ray_length));
  return true;
}

// MovingSphere (absent upstream): centre(time) = c0 + ((time - time0) / (time1 - time0)) * (c1 - c0)
DEV d3 moving_center(d3 c0, d3 c1, double time0, double time1, double time) {
  double f = __ddiv_rn(__dsub_rn(time, time0), __dsub_rn(time1, time0));
  d3 dc = c1 - c0;
  return D3(__dadd_rn(c0.x, __dmul_rn(f, dc.x)), __dadd_rn(c0.y, __dmul_rn(f, dc.y)), __dadd_rn(c0.z, __dmul_rn(f, dc.z)));
}
