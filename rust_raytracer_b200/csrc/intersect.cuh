// intersect.cuh — World::hit on the device (world.rs:27-29 -> bvh.rs:86-102 -> primitive hit): BVH4 traversal
// with a per-thread stack, f32 filters + f64 confirmation per leaf primitive, and the HitRecord "finish"
// (hit.rs:17-48) with the Transform::hit normal/front_face behaviour (transform.rs:151-165).
#pragma once
#include "prim_exact.cuh"
#include "prim_filter.cuh"

struct Ray {
  float3 o;
  float time;
  float3 d;
  float tmax;
  uint32_t mseed;
};

struct HitRaw {
  double t;
  int prim;  // device (leaf-order) primitive index, -1 = miss
  int aux;   // Cube side
};

struct HitInfo {  // HitRecord (hit.rs:7-14)
  float t;
  float3 p;
  float3 n;  // unit, facing the ray
  int front_face;
  float u, v;
  int material;
  int prim_orig;
  int type;
};

DEV RayF make_rayf(const Ray& r) {
  RayF f;
  f.o = r.o; f.d = r.d; f.time = r.time;
  f.inv = f3(safe_rcp(r.d.x), safe_rcp(r.d.y), safe_rcp(r.d.z));
  f.a = dot(r.d, r.d);
  return f;
}
DEV RayD make_rayd(const Ray& r) {
  RayD d;
  d.o = D3(r.o); d.d = D3(r.d); d.time = (double)r.time;
  return d;
}

// One leaf primitive: filter in f32, confirm in f64.  Updates h when the primitive hits in [t_min, h.t].
DEV void test_prim(const DevScene& S, const Ray& ray, const RayF& rf, int k, float t_min, float& t_best32, HitRaw& h) {
  const int4 meta = __ldg(&S.prim_meta[k]);
  const int type = meta.x & 0xff;
  const float4 A = __ldg(&S.prim_data[3 * k]);
  double t;
  int aux = 0;
  switch (type) {
    case B200RT_PRIM_SPHERE: {
      if (sphere_reject(rf, f3(A), A.w, t_min, t_best32)) return;
      RayD rd = make_rayd(ray);
      if (!sphere_exact(rd, D3(A), (double)A.w, (double)t_min, h.t, t)) return;
      break;
    }
    case B200RT_PRIM_MOVING_SPHERE: {
      const float4 B = __ldg(&S.prim_data[3 * k + 1]);
      const float4 C = __ldg(&S.prim_data[3 * k + 2]);
      RayD rd = make_rayd(ray);
      d3 c = moving_center(D3(A), D3(B), (double)C.x, (double)C.y, rd.time);
      if (!sphere_exact(rd, c, (double)A.w, (double)t_min, h.t, t)) return;
      break;
    }
    case B200RT_PRIM_XY_RECT: case B200RT_PRIM_XZ_RECT: case B200RT_PRIM_YZ_RECT: {
      const float4 B = __ldg(&S.prim_data[3 * k + 1]);
      const int axis = type == B200RT_PRIM_XY_RECT ? 2 : (type == B200RT_PRIM_XZ_RECT ? 1 : 0);
      if (rect_reject(rf, axis, A.x, A.y, A.z, A.w, B.x, t_min, t_best32)) return;
      RayD rd = make_rayd(ray);
      double pa, pb;
      if (!rect_exact(rd, axis, (double)A.x, (double)A.y, (double)A.z, (double)A.w, (double)B.x, (double)t_min, h.t, t, pa, pb)) return;
      break;
    }
    case B200RT_PRIM_BOX: {
      const float4 B = __ldg(&S.prim_data[3 * k + 1]);
      if (box_reject(rf, f3(A), f3(B), t_min, t_best32)) return;
      RayD rd = make_rayd(ray);
      if (!cube_exact(rd, D3(A), D3(B), (double)t_min, h.t, t, aux)) return;
      break;
    }
    case B200RT_PRIM_TRIANGLE: {
      const float4 B = __ldg(&S.prim_data[3 * k + 1]);
      const float4 C = __ldg(&S.prim_data[3 * k + 2]);
      if (tri_reject(rf, f3(A), f3(B), f3(C), t_min, t_best32)) return;
      RayD rd = make_rayd(ray);
      double b1, b2;
      if (!tri_exact(rd, D3(A), D3(B), D3(C), (double)t_min, h.t, t, b1, b2)) return;
      break;
    }
    case B200RT_PRIM_MEDIUM_SPHERE: {
      const float4 B = __ldg(&S.prim_data[3 * k + 1]);
      RayD rd = make_rayd(ray);
      if (!medium_exact(rd, 0, D3(A), D3(A), (double)A.w, (double)B.x, medium_xi(ray.mseed, (uint32_t)meta.y), (double)t_min, h.t, t)) return;
      break;
    }
    case B200RT_PRIM_MEDIUM_BOX: {
      const float4 B = __ldg(&S.prim_data[3 * k + 1]);
      if (box_reject(rf, f3(A), f3(B), -3.0e38f, t_best32)) return;
      RayD rd = make_rayd(ray);
      if (!medium_exact(rd, 1, D3(A), D3(B), 0.0, (double)B.w, medium_xi(ray.mseed, (uint32_t)meta.y), (double)t_min, h.t, t)) return;
      break;
    }
    default: return;
  }
  h.t = t; h.prim = k; h.aux = aux;
  t_best32 = __double2float_ru(t);
}

// Closest hit in [t_min, ray.tmax] (ANY = false) or any hit (ANY = true: shadow rays, returns at the first hit).
template <bool ANY>
DEV void trace(const DevScene& S, const Ray& ray, float t_min, HitRaw& h) {
  h.t = (double)ray.tmax; h.prim = -1; h.aux = 0;
  float t_best32 = ray.tmax;
  const RayF rf = make_rayf(ray);
  if (S.flat) {  // hit.rs:56-67: the brute-force arm (data/scene_200_no_bvh.json)
    for (int k = 0; k < S.n_prims; ++k) {
      test_prim(S, ray, rf, k, t_min, t_best32, h);
      if (ANY && h.prim >= 0) return;
    }
    return;
  }
  uint2 stack[B200RT_STACK];
  int sp = 0;
  uint32_t cur = 0;  // root node
  for (;;) {
    if (!(cur & B200RT_LEAF_BIT)) {
      const float4* nd = S.nodes + 8 * (size_t)cur;
      const float4 lox = __ldg(nd + 0), loy = __ldg(nd + 1), loz = __ldg(nd + 2);
      const float4 hix = __ldg(nd + 3), hiy = __ldg(nd + 4), hiz = __ldg(nd + 5);
      const int4 ref = __ldg((const int4*)(nd + 6));
      float tn[4];
      uint32_t rr[4] = {(uint32_t)ref.x, (uint32_t)ref.y, (uint32_t)ref.z, (uint32_t)ref.w};
      const float lx[4] = {lox.x, lox.y, lox.z, lox.w}, ly[4] = {loy.x, loy.y, loy.z, loy.w}, lz[4] = {loz.x, loz.y, loz.z, loz.w};
      const float hx[4] = {hix.x, hix.y, hix.z, hix.w}, hy[4] = {hiy.x, hiy.y, hiy.z, hiy.w}, hz[4] = {hiz.x, hiz.y, hiz.z, hiz.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float t0x = (lx[i] - rf.o.x) * rf.inv.x, t1x = (hx[i] - rf.o.x) * rf.inv.x;
        float t0y = (ly[i] - rf.o.y) * rf.inv.y, t1y = (hy[i] - rf.o.y) * rf.inv.y;
        float t0z = (lz[i] - rf.o.z) * rf.inv.z, t1z = (hz[i] - rf.o.z) * rf.inv.z;
        float n = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fmaxf(fminf(t0z, t1z), t_min));
        float f = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fminf(fmaxf(t0z, t1z), t_best32));
        bool hit = (n <= f * 1.000001f) && (rr[i] != (uint32_t)B200RT_CHILD_EMPTY);
        tn[i] = hit ? n : 3.0e38f;
      }
      // sort the four (tn, ref) pairs ascending (5-comparator network)
#define CSWAP(a, b) if (tn[b] < tn[a]) { float tt = tn[a]; tn[a] = tn[b]; tn[b] = tt; uint32_t tr = rr[a]; rr[a] = rr[b]; rr[b] = tr; }
      CSWAP(0, 1) CSWAP(2, 3) CSWAP(0, 2) CSWAP(1, 3) CSWAP(1, 2)
#undef CSWAP
      if (tn[3] < 3.0e38f) stack[sp++] = make_uint2(rr[3], __float_as_uint(tn[3]));
      if (tn[2] < 3.0e38f) stack[sp++] = make_uint2(rr[2], __float_as_uint(tn[2]));
      if (tn[1] < 3.0e38f) stack[sp++] = make_uint2(rr[1], __float_as_uint(tn[1]));
      if (tn[0] < 3.0e38f) { cur = rr[0]; continue; }
    } else {
      const int first = (int)(cur & 0x07FFFFFFu), count = (int)((cur >> 27) & 15u) + 1;
      for (int k = first; k < first + count; ++k) test_prim(S, ray, rf, k, t_min, t_best32, h);
      if (ANY && h.prim >= 0) return;
    }
    // pop the next entry that can still beat the best hit
    bool found = false;
    while (sp > 0) {
      uint2 e = stack[--sp];
      if (__uint_as_float(e.y) <= t_best32) { cur = e.x; found = true; break; }
    }
    if (!found) return;
  }
}

DEV float3 mat3_mul(const float4* rows, float3 v) { return f3(dot(f3(__ldg(rows)), v), dot(f3(__ldg(rows + 1)), v), dot(f3(__ldg(rows + 2)), v)); }
DEV float3 mat3_tmul(const float4* rows, float3 v) {
  float3 r0 = f3(__ldg(rows)), r1 = f3(__ldg(rows + 1)), r2 = f3(__ldg(rows + 2));
  return f3(r0.x * v.x + r1.x * v.y + r2.x * v.z, r0.y * v.x + r1.y * v.y + r2.y * v.z, r0.z * v.x + r1.z * v.y + r2.z * v.z);
}

DEV void sphere_uv(d3 n, float& u, float& v) {  // sphere.rs:16-20
  double theta = acos(-n.y), phi = atan2(-n.z, n.x) + PI_D;
  u = (float)(phi * (1.0 / PI_D) * 0.5);
  v = (float)(theta * (1.0 / PI_D));
}

// HitRecord::new (hit.rs:17-39) for the winning primitive; WANT_UV = false skips the sphere uv (acos/atan2).
DEV void finish_hit(const DevScene& S, const Ray& ray, const HitRaw& h, HitInfo& out, const bool WANT_UV) {
  const int k = h.prim;
  const int4 meta = __ldg(&S.prim_meta[k]);
  const int type = meta.x & 0xff;
  out.type = type; out.material = meta.x >> 8; out.prim_orig = meta.y; out.t = (float)h.t;
  out.u = 0.f; out.v = 0.f;
  const RayD rd = make_rayd(ray);
  const d3 p = dat(rd.o, h.t, rd.d);
  out.p = to_f3(p);
  const float4 A = __ldg(&S.prim_data[3 * k]);
  d3 on;  // outward normal in WORLD space (object space == world space unless instanced)
  switch (type) {
    case B200RT_PRIM_SPHERE: on = (p - D3(A)) / (double)A.w; break;
    case B200RT_PRIM_MOVING_SPHERE: {
      const float4 B = __ldg(&S.prim_data[3 * k + 1]);
      const float4 C = __ldg(&S.prim_data[3 * k + 2]);
      on = (p - moving_center(D3(A), D3(B), (double)C.x, (double)C.y, rd.time)) / (double)A.w;
      break;
    }
    case B200RT_PRIM_XY_RECT: case B200RT_PRIM_XZ_RECT: case B200RT_PRIM_YZ_RECT: {
      const int axis = type == B200RT_PRIM_XY_RECT ? 2 : (type == B200RT_PRIM_XZ_RECT ? 1 : 0);
      const int ia = axis == 0 ? 1 : 0, ib = axis == 2 ? 1 : 2;
      on = D3(axis == 0 ? 1. : 0., axis == 1 ? 1. : 0., axis == 2 ? 1. : 0.);
      out.u = (float)((dcomp(p, ia) - (double)A.x) / ((double)A.z - (double)A.x));  // rectangle.rs:32,70,107
      out.v = (float)((dcomp(p, ib) - (double)A.y) / ((double)A.w - (double)A.y));
      break;
    }
    case B200RT_PRIM_BOX: {
      const float4 B = __ldg(&S.prim_data[3 * k + 1]);
      int axis; double a0, b0, a1, b1, kk;
      cube_side(h.aux, D3(A), D3(B), axis, a0, b0, a1, b1, kk);
      const int ia = axis == 0 ? 1 : 0, ib = axis == 2 ? 1 : 2;
      on = D3(axis == 0 ? 1. : 0., axis == 1 ? 1. : 0., axis == 2 ? 1. : 0.);
      out.u = (float)((dcomp(p, ia) - a0) / (a1 - a0));
      out.v = (float)((dcomp(p, ib) - b0) / (b1 - b0));
      break;
    }
    case B200RT_PRIM_TRIANGLE: {  // mesh.rs:102-104: normalised barycentric blend of the vertex normals
      const float4 B = __ldg(&S.prim_data[3 * k + 1]);
      const float4 C = __ldg(&S.prim_data[3 * k + 2]);
      double t, b1 = 0, b2 = 0;
      tri_exact(rd, D3(A), D3(B), D3(C), -DINF, DINF, t, b1, b2);
      const double b0 = 1.0 - b1 - b2;
      const float4* nn = S.tri_normals + 3 * (size_t)meta.w;
      d3 n = b0 * D3(__ldg(nn)) + b1 * D3(__ldg(nn + 1)) + b2 * D3(__ldg(nn + 2));
      on = n / sqrt(ddot(n, n));  // OBJECT space when instanced (handled below)
      break;
    }
    default: on = D3(1., 0., 0.); break;  // ConstantMedium: arbitrary normal (medium.rs:47)
  }
  if (meta.z < 0) {
    const bool front = ddot(rd.d, on) < 0.;
    d3 n = front ? on : -on;
    n = n / sqrt(ddot(n, n));  // hit.rs:33
    out.front_face = front ? 1 : 0;
    out.n = to_f3(n);
    if (WANT_UV && (type == B200RT_PRIM_SPHERE || type == B200RT_PRIM_MOVING_SPHERE)) sphere_uv(on, out.u, out.v);
    return;
  }
  // Transform::hit (transform.rs:151-165): the child record is built against the OBJECT-space ray d' = M^-1 d,
  // then n_out = M * n' (not the inverse transpose) and front_face = d' . n_out < 0.
  const float4* Mrows = S.instances + 6 * (size_t)meta.z;
  const float3 dprime = mat3_mul(Mrows + 3, ray.d);
  float3 nobj;
  if (type == B200RT_PRIM_TRIANGLE || type == B200RT_PRIM_MEDIUM_SPHERE || type == B200RT_PRIM_MEDIUM_BOX) nobj = to_f3(on);
  else if (type == B200RT_PRIM_SPHERE || type == B200RT_PRIM_MOVING_SPHERE) nobj = normalize(mat3_tmul(Mrows, to_f3(on)));  // similarity: R^T n
  else nobj = to_f3(on);  // axis shapes under a positive diagonal matrix keep their axis normal
  if (WANT_UV && (type == B200RT_PRIM_SPHERE || type == B200RT_PRIM_MOVING_SPHERE)) sphere_uv(D3(nobj), out.u, out.v);
  const bool front_obj = dot(dprime, nobj) < 0.f;
  const float3 nff = front_obj ? nobj : -nobj;
  const float3 outward = mat3_mul(Mrows, nff);
  const bool front = dot(dprime, outward) < 0.f;
  out.front_face = front ? 1 : 0;
  out.n = front ? normalize(outward) : -normalize(outward);
}
