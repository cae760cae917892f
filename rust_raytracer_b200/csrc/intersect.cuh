// intersect.cuh — World::hit on the device (world.rs:27-29 -> bvh.rs:86-102 -> primitive hit): BVH4 traversal
// with a per-thread stack, f32 filters + f64 confirmation per leaf primitive, and the HitRecord "finish"
// (hit.rs:17-48) with the Transform::hit normal/front_face behaviour (transform.rs:151-165).
#pragma once
#include "prim_exact.cuh"
#include "prim_filter.cuh"

// Debug build (-DB200RT_COUNTERS, tools/dbg): per-ray work statistics of the traversal.
#ifdef B200RT_COUNTERS
__device__ unsigned long long g_dbg[64];
__device__ unsigned int* g_dbg_steps;  // per queue position: HitRaw::dbg (tools/dbg/dump_rays.py)
#define DBG_INC(i) atomicAdd(&g_dbg[(i)], 1ull)
#define DBG_RAY(x) x
#else
#define DBG_INC(i)
#define DBG_RAY(x)
#endif
// counter slots: 0 rays, 1 inner nodes visited, 2 leaves visited, 4 pops culled, 5 to_object, 6 child boxes slab-tested
// (non-empty slots of the visited nodes), 7 HitRecord::new, 8 sphere uv, 9 primary rays generated,
// 10+type primitive tests (f32 filter calls), 20+type f64 decisions, 30+type accepted,
// 40+material type scatter calls, 48+texture type get_color calls, 56 light samples
// (bench.py turns these into SURVEY.md section 8(d)'s algorithmic flops)

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
#ifdef B200RT_COUNTERS
  unsigned int dbg;  // this ray's traversal: inner nodes | leaves << 10 | f32 filter calls << 16 | f64 decisions << 24
#endif
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

// Transform::hit (transform.rs:152-155): the ray in OBJECT space, o' = M^-1 * o (point), d' = M^-1 * d (direction),
// evaluated in f64 with the reference's operation order (nalgebra mat-vec, vec3.rs:174-184).
DEV d3 affine_point(const double* m, d3 p) {
  return D3(__dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(m[0], p.x), __dmul_rn(m[1], p.y)), __dmul_rn(m[2], p.z)), m[3]),
            __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(m[4], p.x), __dmul_rn(m[5], p.y)), __dmul_rn(m[6], p.z)), m[7]),
            __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(m[8], p.x), __dmul_rn(m[9], p.y)), __dmul_rn(m[10], p.z)), m[11]));
}
DEV d3 affine_dir(const double* m, d3 p) {
  return D3(__dadd_rn(__dadd_rn(__dmul_rn(m[0], p.x), __dmul_rn(m[1], p.y)), __dmul_rn(m[2], p.z)),
            __dadd_rn(__dadd_rn(__dmul_rn(m[4], p.x), __dmul_rn(m[5], p.y)), __dmul_rn(m[6], p.z)),
            __dadd_rn(__dadd_rn(__dmul_rn(m[8], p.x), __dmul_rn(m[9], p.y)), __dmul_rn(m[10], p.z)));
}
DEV RayD to_object(const DevScene& S, int inst, const RayD& w) {
  const double* minv = S.instances_d + 24 * (size_t)inst + 12;
  RayD r;
  r.o = affine_point(minv, w.o); r.d = affine_dir(minv, w.d); r.time = w.time;
  return r;
}

// One leaf primitive: filter in f32 on the world-space (baked) record, confirm in f64 on the ORIGINAL record
// (object space for instanced primitives, exactly like Transform::hit).  Updates h when the primitive hits in
// [t_min, h.t].
//
// Code size is a first-order cost here (the SM's L1.5 instruction cache is 32 KB and divergent warps walk
// different parts of this function), so every heavy f64 sequence exists exactly ONCE: one quadratic-root site
// serves Sphere, MovingSphere and the sphere ConstantMedium boundary; one rectangle site serves the three
// rectangles, the Cube faces and the box ConstantMedium boundary; the medium tail (medium.rs:33-52) is shared.
//
// `objf`: Transform::hit in full generality (transform.rs:151-165: any rotation, scale: Vec3, any child).  An instanced
// primitive whose shape does not survive the transform (a rotated Cube / rectangle / box medium, a sphere under a
// non-uniform scale) keeps its OBJECT-space record and is filtered on the object-space ray: the caller (test_prim) has
// then replaced `rf` by that ray and `eps` by its rounding error; the f64 decision runs in object space for every
// instanced primitive anyway.
DEV void test_prim_body(const DevScene& S, const Ray& ray, const RayF& rf, const float eps, const int4 meta, int k, float t_min, float& t_best32, HitRaw& h) {
  const float4* __restrict__ prim_data = S.prim_data;
  const int type = meta.x & B200RT_PRIM_TYPE_MASK;
  const bool inst = meta.z >= 0;
  const float4 A = __ldg(&prim_data[3 * k]);
  const float4 WB = __ldg(&prim_data[3 * k + 1]);
  const bool is_medium = type == B200RT_PRIM_MEDIUM_SPHERE || type == B200RT_PRIM_MEDIUM_BOX;
  const bool is_rect = type == B200RT_PRIM_XY_RECT || type == B200RT_PRIM_XZ_RECT || type == B200RT_PRIM_YZ_RECT;
  const int rect_axis = type == B200RT_PRIM_XY_RECT ? 2 : (type == B200RT_PRIM_XZ_RECT ? 1 : 0);
  DBG_INC(10 + type);
  DBG_RAY(h.dbg += 1u << 16);
  uint32_t cand = 0x3fu;  // Cube faces to confirm: bits 0..5 near the slab entry, bits 8..13 near the exit
  int verdict = FILTER_MAYBE;
  float t_medium = 0.f;
  // ---- f32 filter (certain misses only)
  if (type == B200RT_PRIM_SPHERE || type == B200RT_PRIM_MEDIUM_SPHERE) {
    const float xi = is_medium ? medium_xi_f32(ray.mseed, (uint32_t)meta.y) : 0.5f;
    verdict = sphere_reject(rf, f3(A), A.w, t_min, t_best32, eps, is_medium, WB.x, xi, &t_medium);
    if (verdict == FILTER_REJECT) return;
  } else if (type == B200RT_PRIM_BOX || type == B200RT_PRIM_MEDIUM_BOX) {
    const float xi = is_medium ? medium_xi_f32(ray.mseed, (uint32_t)meta.y) : 0.5f;
    verdict = box_reject(rf, f3(A), f3(WB), t_min, t_best32, eps, is_medium ? nullptr : &cand, is_medium, WB.w, xi, &t_medium);
    if (verdict == FILTER_REJECT) return;
  } else if (is_rect) {
    if (rect_reject(rf, rect_axis, A.x, A.y, A.z, A.w, WB.x, t_min, t_best32, eps)) return;
    cand = 1u << (2 * (2 - rect_axis));  // the rectangle is the "min" side of a degenerate box on its axis
  } else if (type == B200RT_PRIM_TRIANGLE) {
    const float4 C = __ldg(&prim_data[3 * k + 2]);
    if (tri_reject(rf, f3(A), f3(WB), f3(C), t_min, t_best32, eps)) return;
  }  // moving spheres go straight to the exact test
  if (verdict == FILTER_MEDIUM_HIT) {  // certain interior scatter, certainly the closest hit so far: t in f32
    DBG_INC(30 + type);
    h.t = (double)t_medium; h.prim = k; h.aux = 0;
    t_best32 = t_medium;
    return;
  }
  // ---- f64 decision
  RayD rd = make_rayd(ray);
  const float4* rec = prim_data + 3 * (size_t)k;
  if (inst) { rd = to_object(S, meta.z, rd); rec = S.prim_obj + 3 * (size_t)k; DBG_INC(5); }
  DBG_INC(20 + type);
  DBG_RAY(h.dbg += 1u << 24);
  const float4 EA = inst ? __ldg(rec) : A;
  const float4 EB = inst ? __ldg(rec + 1) : WB;
  const double tmin_d = (double)t_min, tmax_d = h.t;
  double t_in = 0., t_out = 0.;  // surfaces: t_in is the hit; media: the boundary interval
  int aux = 0;
  if (type == B200RT_PRIM_SPHERE || type == B200RT_PRIM_MOVING_SPHERE || type == B200RT_PRIM_MEDIUM_SPHERE) {
    // objects/sphere.rs:24-43 — nearest root in the CLOSED interval [t_min, t_max]; tangent rays hit.
    // ConstantMedium (medium.rs:26-31): first root in (-inf, inf), then the next one in [first + 0.0001, inf).
    d3 c = D3(EA);
    if (type == B200RT_PRIM_MOVING_SPHERE) {
      const float4 C = __ldg(rec + 2);
      c = moving_center(c, D3(EB), (double)C.x, (double)C.y, rd.time);
    }
    const double radius = (double)EA.w;
    const d3 oc = rd.o - c;
    const double a = ddot(rd.d, rd.d);
    const double half_b = ddot(oc, rd.d);
    const double cc = __dsub_rn(ddot(oc, oc), __dmul_rn(radius, radius));
    const double disc = __dsub_rn(__dmul_rn(half_b, half_b), __dmul_rn(a, cc));
    if (disc < 0.) return;
    const double sqrt_d = sqrt(disc);
    double lo = is_medium ? -DINF : tmin_d;
    const double hi = is_medium ? DINF : tmax_d;
    int found = 0;
#pragma unroll 1
    for (int i = 0; i < 2; ++i) {
      const double root = __ddiv_rn(i == 0 ? __dsub_rn(-half_b, sqrt_d) : __dadd_rn(-half_b, sqrt_d), a);
      if (root >= lo && root <= hi) {
        if (found == 0) { t_in = root; found = 1; if (!is_medium) break; lo = __dadd_rn(root, 0.0001); }
        else { t_out = root; found = 2; }
      }
    }
    if (found == 0 || (is_medium && found != 2)) return;
  } else if (type == B200RT_PRIM_TRIANGLE) {
    const float4 C = __ldg(rec + 2);
    double b1, b2;
    if (!tri_exact(rd, D3(EA), D3(EB), D3(C), tmin_d, tmax_d, t_in, b1, b2)) return;
  } else {
    // Rectangles (rectangle.rs:15-34,53-72,90-109), Cube = closest of six rectangles (cube.rs:63-66), and the box
    // ConstantMedium boundary: one pass over the candidate faces per boundary hit wanted.
    d3 mn = D3(EA), mx = D3(EB);
    if (is_rect) {  // A = (a0, b0, a1, b1), B.x = k
      const double kk = (double)EB.x;
      if (rect_axis == 2) { mn = D3((double)EA.x, (double)EA.y, kk); mx = D3((double)EA.z, (double)EA.w, kk); }
      else if (rect_axis == 1) { mn = D3((double)EA.x, kk, (double)EA.y); mx = D3((double)EA.z, kk, (double)EA.w); }
      else { mn = D3(kk, (double)EA.x, (double)EA.y); mx = D3(kk, (double)EA.z, (double)EA.w); }
    }
    const float es = inst ? 2.0f * eps * fmaxf(fabsf(rf.inv.x), fmaxf(fabsf(rf.inv.y), fabsf(rf.inv.z))) : 0.f;
    auto lower = [&](int s) { const float ts = box_side_t(rf, f3(A), f3(WB), s); return ts - (2e-6f * fabsf(ts) + es); };
    double lo = is_medium ? -DINF : tmin_d;
    const double hi = is_medium ? DINF : tmax_d;
#pragma unroll 1
    for (int pass = 0; pass < (is_medium ? 2 : 1); ++pass) {
      double tt;
      if (!cube_exact_cand(rd, mn, mx, lo, hi, cand, lower, tt, aux)) return;
      if (pass == 0) { t_in = tt; lo = __dadd_rn(tt, 0.0001); } else t_out = tt;
    }
    if (type != B200RT_PRIM_BOX) aux = 0;  // only the Cube records which side was hit (finish_hit)
  }
  double t = t_in;
  if (is_medium) {  // medium.rs:33-52; xi is the injected draw
    const double neg_inv_density = type == B200RT_PRIM_MEDIUM_SPHERE ? (double)EB.x : (double)EB.w;
    double t1 = fmax(t_in, tmin_d);
    const double t2 = fmin(t_out, tmax_d);
    if (t1 >= t2) return;
    t1 = fmax(t1, 0.);
    const double ray_length = sqrt(ddot(rd.d, rd.d));
    const double distance_inside = __dmul_rn(__dsub_rn(t2, t1), ray_length);
    const double hit_distance = __dmul_rn(neg_inv_density, log(medium_xi(ray.mseed, (uint32_t)meta.y)));
    if (hit_distance > distance_inside) return;
    t = __dadd_rn(t1, __ddiv_rn(hit_distance, ray_length));
    aux = 0;
  }
  DBG_INC(30 + type);
  h.t = t; h.prim = k; h.aux = aux;
  t_best32 = __double2float_ru(t);
}

// The object-space arm, out of line: a second copy of the leaf test that only scenes with such primitives ever fetch
// (inlined, it cost book-2 — which has none — 9 %: +8 % instructions and +50 % spill traffic in k_wf_extend at 64
// registers).  Everything goes in and out BY VALUE, so the caller keeps its hit record and ray constants in registers
// and only pays a predicate on its hot path.  The ray is taken to object space in f64 and rounded once, so the filters
// see an origin that is off by <= 1 ulp per component — the same error class as a baked coordinate, handed to them as `eps`.
struct ObjfHit { double t; int aux; int hit; };
__device__ __noinline__ ObjfHit test_prim_objf(const float4* __restrict__ prim_data, const float4* __restrict__ prim_obj, const double* __restrict__ instances_d,
                                               float ox, float oy, float oz, float dx, float dy, float dz, float time, uint32_t mseed, int4 meta, int k, float t_min,
                                               float t_best32, double t_best) {
  DevScene S{};  // the leaf test reads nothing else of the scene
  S.prim_data = prim_data; S.prim_obj = prim_obj; S.instances_d = instances_d;
  Ray ray;
  ray.o = f3(ox, oy, oz); ray.d = f3(dx, dy, dz); ray.time = time; ray.tmax = t_best32; ray.mseed = mseed;
  const RayD rd = to_object(S, meta.z, make_rayd(ray));
  RayF rf;
  rf.o = to_f3(rd.o); rf.d = to_f3(rd.d); rf.time = time;
  rf.inv = f3(safe_rcp(rf.d.x), safe_rcp(rf.d.y), safe_rcp(rf.d.z));
  rf.a = dot(rf.d, rf.d);
  const float eps = 1.2e-7f * fmaxf(fabsf(rf.o.x), fmaxf(fabsf(rf.o.y), fabsf(rf.o.z))) + 1e-30f;
  HitRaw h;
  h.t = t_best; h.prim = -1; h.aux = 0;
  DBG_RAY(h.dbg = 0);
  test_prim_body(S, ray, rf, eps, meta, k, t_min, t_best32, h);
  ObjfHit r;
  r.t = h.t; r.aux = h.aux; r.hit = h.prim >= 0;
  return r;
}

// One leaf primitive.  OBJF = false: the scene holds no object-space-filtered primitive (DevScene::has_objf, decided by the
// host per launch), and the kernel carries neither the branch nor the register constraints of the call (the hot kernels
// are instantiated both ways: even the never-taken call cost book-2's extend 5.7 %).
template <bool OBJF>
DEV void test_prim(const DevScene& S, const Ray& ray, const RayF& rf, int k, float t_min, float& t_best32, HitRaw& h) {
  const int4 meta = __ldg(&S.prim_meta[k]);
  if (OBJF && (meta.x & B200RT_PRIM_OBJF)) {
    const ObjfHit r = test_prim_objf(S.prim_data, S.prim_obj, S.instances_d, ray.o.x, ray.o.y, ray.o.z, ray.d.x, ray.d.y, ray.d.z, ray.time, ray.mseed, meta, k, t_min, t_best32, h.t);
    if (r.hit) { h.t = r.t; h.prim = k; h.aux = r.aux; t_best32 = __double2float_ru(r.t); }
    return;
  }
  test_prim_body(S, ray, rf, meta.z >= 0 ? __ldg(&S.inst_eps[meta.z]) : 0.f, meta, k, t_min, t_best32, h);
}

#define B200RT_TRAV_DONE 0xFFFFFFFFu  // leaf bit set with an impossible count: never a real child reference

// Traversal stack.  north_star asks for a shared-memory short stack; it is implemented (the first B200RT_SSTACK
// entries of every thread in shared memory, entry-major so a warp's 8-byte accesses are conflict-free, deeper
// entries overflowing to a per-thread local array) and was measured on book-2: 8 / 16 / 24 shared entries cost
// 66.7 / 70.1 / 82.4 ms of extend per step against 56.2 ms for the plain local-memory stack — the carve-out takes
// L1 capacity away from the node / primitive / spill traffic, and the two-way push/pop adds instructions.  The
// default is therefore 0 (local memory, L1-resident); profiles/r1_experiments.md.
#ifndef B200RT_SSTACK
#define B200RT_SSTACK 0
#endif
#define B200RT_BLOCK 128  // every kernel that traverses runs 128-thread blocks
struct TStack {
#if B200RT_SSTACK > 0
  uint2* sm;  // &shared[threadIdx.x]; entry k at sm[k * B200RT_BLOCK]
#endif
  uint2 loc[B200RT_STACK - B200RT_SSTACK];
#ifndef B200RT_STACK_TOPREG  // 1: the top entry lives in registers, so a pop's local-memory load feeds the NEXT pop, not this one
#define B200RT_STACK_TOPREG 0
#endif
#if B200RT_STACK_TOPREG && B200RT_SSTACK == 0
  uint2 top;  // valid iff sp > 0; loc[k] holds entry k for k < sp - 1
  DEV void push(int& sp, uint2 v) {
    if (sp > 0) loc[sp - 1] = top;
    top = v;
    ++sp;
  }
  DEV uint2 pop(int& sp) {
    --sp;
    const uint2 r = top;
    if (sp > 0) top = loc[sp - 1];
    return r;
  }
#else
  DEV void push(int& sp, uint2 v) {
#if B200RT_SSTACK > 0
    if (sp < B200RT_SSTACK) sm[sp * B200RT_BLOCK] = v; else loc[sp - B200RT_SSTACK] = v;
#else
    loc[sp] = v;
#endif
    ++sp;
  }
  DEV uint2 pop(int& sp) {
    --sp;
#if B200RT_SSTACK > 0
    return sp < B200RT_SSTACK ? sm[sp * B200RT_BLOCK] : loc[sp - B200RT_SSTACK];
#else
    return loc[sp];
#endif
  }
#endif
};

DEV uint32_t pop_next(TStack& stack, int& sp, float t_best32) {
  while (sp > 0) {
    const uint2 e = stack.pop(sp);
    if (__uint_as_float(e.y) <= t_best32) return e.x;
    DBG_INC(4);
  }
  return B200RT_TRAV_DONE;
}

#ifndef B200RT_NODE_FAST
#define B200RT_NODE_FAST 1
#endif
#ifndef B200RT_NODE_FULLSORT
#define B200RT_NODE_FULLSORT 0
#endif

// Per-ray constants of the node test.  t = plane * inv - o * inv is one FFMA per plane; the near / far plane of
// every axis is chosen by ADDRESS (the node stores lo.x|hi.x|lo.y|hi.y|lo.z|hi.z as six float4 rows, so the
// direction's sign bit picks row 2a or row 2a + 1), which removes the per-child min/max pairs.  The rounding error of
// o * inv (the only term the fused form adds over (plane - o) * inv) is covered so the
// test stays conservative through one per-ray slack term on the comparison (2 * max|o * inv| * 2^-22).
struct RayT {
  float3 inv, oi;  // 1/d, o/d
  float slack;     // 2 * max_axis |o/d| * 2^-22: absolute error budget of (near - far) from rounding o/d
  const float4* nodes;   // S.nodes, laundered through a register pair (see node_step)
};
DEV RayT make_rayt(const DevScene& S, const RayF& rf) {
  RayT t;
  asm("mov.b64 %0, %1;" : "=l"(t.nodes) : "l"(S.nodes));
  t.inv = rf.inv;
  t.oi = rf.o * rf.inv;
  // 2 * 2^-22 * max |o/d| over the axes whose reciprocal is a rounded one (a clamped axis is exact, see safe_rcp: without
  // this an axis-parallel ray would carry a slack of ~1e26 and lose the slab culling of the other two axes altogether)
#ifndef B200RT_SLACK_PER_AXIS
#define B200RT_SLACK_PER_AXIS 1
#endif
#if B200RT_SLACK_PER_AXIS
  const float sx = fabsf(rf.d.x) > B200RT_RCP_CLAMP ? fabsf(t.oi.x) : 0.f, sy = fabsf(rf.d.y) > B200RT_RCP_CLAMP ? fabsf(t.oi.y) : 0.f,
              sz = fabsf(rf.d.z) > B200RT_RCP_CLAMP ? fabsf(t.oi.z) : 0.f;
  t.slack = 4.7683716e-7f * fmaxf(sx, fmaxf(sy, sz));
#else
  t.slack = 4.7683716e-7f * fmaxf(fabsf(t.oi.x), fmaxf(fabsf(t.oi.y), fabsf(t.oi.z)));
#endif
  return t;
}

// BVH4 node step: slab-tests the four children of inner node `cur` against [t_min, t_best32], pushes the far
// hits and returns the nearest child reference (or the next stack entry when no child is hit).
DEV uint32_t node_step(const DevScene& S, const RayF& rf, const RayT& rt, uint32_t cur, float t_min, float t_best32, TStack& stack, int& sp) {
  DBG_INC(1);
#ifdef B200RT_COUNTERS
  { const int4 rfc = __ldg(reinterpret_cast<const int4*>(rt.nodes + (cur << 3)) + 6);
    atomicAdd(&g_dbg[6], (unsigned long long)((rfc.x != B200RT_CHILD_EMPTY) + (rfc.y != B200RT_CHILD_EMPTY) + (rfc.z != B200RT_CHILD_EMPTY) + (rfc.w != B200RT_CHILD_EMPTY))); }
#endif
  float tn[4];
#if B200RT_NODE_FAST
  // 32-bit row indices off a base pointer the compiler cannot see through: the near row of an axis is
  // c8 + (sign bit of 1/d) (one LEA.HI), the far row its xor 1, the address one IMAD.WIDE (FMA pipe) with the axis
  // in the load's immediate offset — instead of select + or + add + add-with-carry per row against a uniform base.
  const uint32_t c8 = cur << 3;
  uint32_t sx, sy, sz;  // opaque sign bits: keeps the far index an independent 32-bit sum (one IADD3 + one IMAD.WIDE)
  asm("shr.u32 %0, %1, 31;" : "=r"(sx) : "r"(__float_as_uint(rt.inv.x)));
  asm("shr.u32 %0, %1, 31;" : "=r"(sy) : "r"(__float_as_uint(rt.inv.y)));
  asm("shr.u32 %0, %1, 31;" : "=r"(sz) : "r"(__float_as_uint(rt.inv.z)));
  const float4 nxv = __ldg(rt.nodes + (c8 + sx)), fxv = __ldg(rt.nodes + (c8 + 1u - sx));
  const float4 nyv = __ldg(rt.nodes + (c8 + 2u + sy)), fyv = __ldg(rt.nodes + (c8 + 3u - sy));
  const float4 nzv = __ldg(rt.nodes + (c8 + 4u + sz)), fzv = __ldg(rt.nodes + (c8 + 5u - sz));
  const int4 ref = __ldg(reinterpret_cast<const int4*>(rt.nodes + c8) + 6);
  uint32_t rr[4] = {(uint32_t)ref.x, (uint32_t)ref.y, (uint32_t)ref.z, (uint32_t)ref.w};
  const float ax[4] = {nxv.x, nxv.y, nxv.z, nxv.w}, ay[4] = {nyv.x, nyv.y, nyv.z, nyv.w}, az[4] = {nzv.x, nzv.y, nzv.z, nzv.w};
  const float bx[4] = {fxv.x, fxv.y, fxv.z, fxv.w}, by[4] = {fyv.x, fyv.y, fyv.z, fyv.w}, bz[4] = {fzv.x, fzv.y, fzv.z, fzv.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float n = fmaxf(fmaxf(fmaf(ax[i], rt.inv.x, -rt.oi.x), fmaf(ay[i], rt.inv.y, -rt.oi.y)), fmaxf(fmaf(az[i], rt.inv.z, -rt.oi.z), t_min));
    const float f = fminf(fminf(fmaf(bx[i], rt.inv.x, -rt.oi.x), fmaf(by[i], rt.inv.y, -rt.oi.y)), fminf(fmaf(bz[i], rt.inv.z, -rt.oi.z), t_best32));
    tn[i] = (n <= fmaf(f, 1.000001f, rt.slack)) ? n : 3.0e38f;  // empty slots hold an inverted infinite box: never hit (inf <= -inf is false)
  }
#else
  const float4* nd = S.nodes + 8 * (size_t)cur;
  const float4 lox = __ldg(nd + 0), loy = __ldg(nd + 2), loz = __ldg(nd + 4);
  const float4 hix = __ldg(nd + 1), hiy = __ldg(nd + 3), hiz = __ldg(nd + 5);
  const int4 ref = __ldg((const int4*)(nd + 6));
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
#endif
  // Nearest child into slot 0 (three comparators); the full 5-comparator sort of the rest only matters when three
  // or four children are hit (rare: ~1.4 hits per visited node on book-2), and pop_next culls by the stored tn anyway.
#define CSWAP(a, b) if (tn[b] < tn[a]) { float tt = tn[a]; tn[a] = tn[b]; tn[b] = tt; uint32_t tr = rr[a]; rr[a] = rr[b]; rr[b] = tr; }
  CSWAP(0, 1) CSWAP(2, 3) CSWAP(0, 2)
#if B200RT_NODE_FULLSORT
  CSWAP(1, 3) CSWAP(1, 2)
#endif
#undef CSWAP
  if (tn[3] < 3.0e38f) stack.push(sp, make_uint2(rr[3], __float_as_uint(tn[3])));
  if (tn[2] < 3.0e38f) stack.push(sp, make_uint2(rr[2], __float_as_uint(tn[2])));
  if (tn[1] < 3.0e38f) stack.push(sp, make_uint2(rr[1], __float_as_uint(tn[1])));
#ifndef B200RT_NODE_PREFETCH  // 1: the second-nearest child's node (one 128 B line) is pulled towards L1 when it is pushed
#define B200RT_NODE_PREFETCH 0
#endif
#if B200RT_NODE_PREFETCH
  if (tn[1] < 3.0e38f && !(rr[1] & B200RT_LEAF_BIT)) asm volatile("prefetch.global.L1 [%0];" ::"l"(rt.nodes + (rr[1] << 3)));
#endif
  if (tn[0] < 3.0e38f) return rr[0];
  return pop_next(stack, sp, t_best32);
}

#ifndef B200RT_TRAV_WW
#define B200RT_TRAV_WW 1
#endif


// Closest hit in [t_min, ray.tmax] (ANY = false) or any hit (ANY = true: shadow rays, returns at the first hit).
// "while-while" traversal: every lane first descends inner nodes until it holds a leaf (or is done), then the
// warp runs the (expensive, f64-confirmed) leaf tests together.  Primitives outside the BVH — the scene-enclosing
// ones the host kept out of it, or ALL of them for the flat-list arm (hit.rs:56-67, n_bvh_prims = 0) — are fed
// through the same leaf code as pseudo-leaves of up to 16 after the traversal, when t_best already culls most of
// their work: test_prim is instantiated once per kernel.
//
// `lanes` != 0: the caller computed, with the warp CONVERGED (a __ballot_sync over the full mask), exactly which
// lanes call trace() together.  Those lanes then wait for each other at the end of the BVH so that the out-of-BVH
// tail (book-2's fog: an f64 medium test for half of the rays) runs once per warp with all of them converged
// instead of once per group of lanes that happen to leave the BVH together.  Every lane of `lanes` executes the
// same number of __all_sync votes: one per trip of the outer loop until the vote passes, and the vote passes for
// all of them in the same trip.  Callers that sit in divergent code (megakernel path loop, SPPM bounce loops) pass
// 0: no vote, no waiting (__activemask() is not a safe substitute there: it may name lanes that never arrive).
template <bool ANY, bool OBJF = true>
DEV void trace(const DevScene& S, const Ray& ray, float t_min, HitRaw& h, const unsigned lanes = 0u) {
  h.t = (double)ray.tmax; h.prim = -1; h.aux = 0;
  DBG_RAY(h.dbg = 0);
  float t_best32 = ray.tmax;
  const RayF rf = make_rayf(ray);
  const RayT rt = make_rayt(S, rf);
  DBG_INC(0);
  TStack stack;
#if B200RT_SSTACK > 0
  __shared__ uint2 s_stack[B200RT_SSTACK * B200RT_BLOCK];
  stack.sm = s_stack + threadIdx.x;
#endif
  int sp = 0;
  int tail = S.n_bvh_prims;                          // next primitive outside the BVH
  const bool wait_tail = !ANY && lanes != 0u && !S.flat && S.n_bvh_prims < S.n_prims;  // warp-uniform
  uint32_t cur = S.flat ? B200RT_TRAV_DONE : 0u;  // root node
  for (;;) {
#if B200RT_TRAV_WW
    while (!(cur & B200RT_LEAF_BIT)) { cur = node_step(S, rf, rt, cur, t_min, t_best32, stack, sp); DBG_RAY(h.dbg += 1u); }
#else
    if (!(cur & B200RT_LEAF_BIT)) { cur = node_step(S, rf, rt, cur, t_min, t_best32, stack, sp); continue; }
#endif
    if (wait_tail) {  // the lanes that finished the BVH wait here until every lane of `lanes` has
      const bool all_done = __all_sync(lanes, cur == B200RT_TRAV_DONE);
      if (cur == B200RT_TRAV_DONE && !all_done) continue;
    }
    if (cur == B200RT_TRAV_DONE) {
      if (tail >= S.n_prims) break;
      const int c = min(16, S.n_prims - tail);
      cur = B200RT_LEAF_BIT | ((uint32_t)(c - 1) << 27) | (uint32_t)tail;
      tail += c;
    }
    const int first = (int)(cur & 0x07FFFFFFu), count = (int)((cur >> 27) & 15u) + 1;
    DBG_INC(2);
    DBG_RAY(h.dbg += 1u << 10);
    for (int k = first; k < first + count; ++k) test_prim<OBJF>(S, ray, rf, k, t_min, t_best32, h);
    if (ANY && h.prim >= 0) return;
    cur = pop_next(stack, sp, t_best32);
  }
}

DEV void sphere_uv(d3 n, float& u, float& v) {  // sphere.rs:16-20
  DBG_INC(8);
  double theta = acos(-n.y), phi = atan2(-n.z, n.x) + PI_D;
  u = (float)(phi * (1.0 / PI_D) * 0.5);
  v = (float)(theta * (1.0 / PI_D));
}

// HitRecord::new (hit.rs:17-39) for the winning primitive; WANT_UV = false skips the sphere uv (acos/atan2).
// Instanced primitives follow Transform::hit (transform.rs:151-165): the record is built in OBJECT space, then
// p = M p', n_out = M n' (not the inverse transpose) and front_face = d' . n_out < 0 with the OBJECT-space d'.
DEV void finish_hit(const DevScene& S, const Ray& ray, const HitRaw& h, HitInfo& out, const bool WANT_UV) {
  const int k = h.prim;
  const int4 meta = __ldg(&S.prim_meta[k]);
  const int type = meta.x & B200RT_PRIM_TYPE_MASK;
  const bool inst = meta.z >= 0;
  DBG_INC(7);
  out.type = type; out.material = meta.x >> 8; out.prim_orig = meta.y; out.t = (float)h.t;
  out.u = 0.f; out.v = 0.f;
  RayD rd = make_rayd(ray);
  const float4* rec = S.prim_data + 3 * (size_t)k;
  if (inst) { rd = to_object(S, meta.z, rd); rec = S.prim_obj + 3 * (size_t)k; }
  const d3 p = dat(rd.o, h.t, rd.d);
  const float4 A = __ldg(rec);
  d3 on;  // outward normal in the primitive's own space
  switch (type) {
    // (p - c) / r: |r| drops out below, its SIGN does not (sphere.rs:44 — a negative radius turns the normal inwards,
    // the hollow-glass-sphere idiom)
    case B200RT_PRIM_SPHERE: on = p - D3(A); if (A.w < 0.f) on = -on; break;
    case B200RT_PRIM_MOVING_SPHERE: {
      const float4 B = __ldg(rec + 1), C = __ldg(rec + 2);
      on = p - moving_center(D3(A), D3(B), (double)C.x, (double)C.y, rd.time);
      if (A.w < 0.f) on = -on;
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
      const float4 B = __ldg(rec + 1);
      int axis; double a0, b0, a1, b1, kk;
      cube_side(h.aux, D3(A), D3(B), axis, a0, b0, a1, b1, kk);
      const int ia = axis == 0 ? 1 : 0, ib = axis == 2 ? 1 : 2;
      on = D3(axis == 0 ? 1. : 0., axis == 1 ? 1. : 0., axis == 2 ? 1. : 0.);
      out.u = (float)((dcomp(p, ia) - a0) / (a1 - a0));
      out.v = (float)((dcomp(p, ib) - b0) / (b1 - b0));
      break;
    }
    case B200RT_PRIM_TRIANGLE: {  // mesh.rs:102-104: normalised barycentric blend of the vertex normals
      const float4 B = __ldg(rec + 1), C = __ldg(rec + 2);
      double t, b1 = 0, b2 = 0;
      tri_exact(rd, D3(A), D3(B), D3(C), -DINF, DINF, t, b1, b2);
      const double b0 = __dsub_rn(__dsub_rn(1.0, b1), b2);
      const float4* nn = S.tri_normals + 3 * (size_t)meta.w;
      on = b0 * D3(__ldg(nn)) + b1 * D3(__ldg(nn + 1)) + b2 * D3(__ldg(nn + 2));
      break;
    }
    default: on = D3(1., 0., 0.); break;  // ConstantMedium: arbitrary normal (medium.rs:47)
  }
  // `on` is the outward normal up to a positive scale.  The cancellation-prone parts (p, p - c, the sign tests of
  // hit.rs:25 / transform.rs:160) stay in f64; Vec3::unit (hit.rs:33) is taken in f32 after the final rounding
  // (2e-7 relative against the 1e-5 tolerance; three f64 divisions and a square root less per hit).
  if (WANT_UV && (type == B200RT_PRIM_SPHERE || type == B200RT_PRIM_MOVING_SPHERE)) sphere_uv(on * (1.0 / fabs((double)A.w)), out.u, out.v);
  const bool front = ddot(rd.d, on) < 0.;  // hit.rs:25
  const d3 n = front ? on : -on;            // hit.rs:26-33
  if (!inst) {
    out.front_face = front ? 1 : 0;
    out.n = normalize(to_f3(n));
    out.p = to_f3(p);
    return;
  }
  const double* m = S.instances_d + 24 * (size_t)meta.z;
  const d3 outward = affine_dir(m, n);  // transform.rs:158
  out.p = to_f3(affine_point(m, p));    // transform.rs:159
  const bool ff = ddot(rd.d, outward) < 0.;  // transform.rs:160 -> hit.rs:41-48, against the OBJECT-space ray
  out.front_face = ff ? 1 : 0;
  out.n = normalize(to_f3(ff ? outward : -outward));
}
