// sppm.cuh — the reference's SPPM integrator (integrator/photon_mapper.rs:17-366, light.rs:18-235, material.rs:27-45)
// on the device: SURVEY.md section 8(f) rank 1.  Four stages per render:
//   k_sppm_photons : AllLights::emit + the photon walk of generate_photon_map (Russian roulette of scatter_photon),
//                    one persistent thread per photon over the same traversal as the path kernels; photons are
//                    appended to the global / caustic maps with warp-aggregated atomics
//   k_sppm_grid_*  : the kd-tree (kd-tree 0.4.1) becomes a uniform grid built by a counting sort on the device
//                    (cell histogram -> one-block scan -> scatter of photon indices)
//   k_sppm_eye     : update_sppm — one jittered eye ray per pixel through the specular chain to the first diffuse
//                    hit, then SPPM::update against both maps: exact k-nearest (expanding cell shells, max-heap in
//                    local memory) the first time, radius gather afterwards
//   k_sppm_final   : impl Integrator for SPPMIntegrator — emission + the pixel's caustic and global estimates at the
//                    first diffuse hit
// Everything stochastic is keyed Philox (photon, iteration, bounce) / (pixel, iteration, bounce) / (pixel, sample,
// bounce): parity with the oracle restatement is statistical, as for the path integrators.
#pragma once
#include "wavefront.cuh"

#define SPPM_MAX_K 128      // heap capacity of the k-nearest estimate (100 / 50 upstream)
#define SPPM_MAX_BOUNCES 4096
#ifndef SPPM_REFILL_LANES
#define SPPM_REFILL_LANES 8  // idle lanes of a photon warp that trigger a work fetch
#endif

struct SppmConst {
  int iterations, photons_per_iter, k_global, k_caustic;
  double alpha;  // f64: (ALPHA * photons as f64) as usize (photon_mapper.rs:59) truncates, 0.7f * 10 would give 6
  float light_scale;
};
struct SppmGrid {  // one photon map, cell-ordered: what a gather tests (the position) apart from what an accepted photon adds
  const float4* pos;   // (pos, 0) per photon
  const float4* attr;  // 2 x float4 per photon: (power, 0) (normal, 0) — the photon's direction is not used by any bsdf here
  const uint32_t* cell_start;  // res^3 + 1
  uint32_t n;
  int res;
  float3 lo;
  float inv_cell, cell;
};
// A photon map as the concatenation of up to 16 segments: the emission-ordered buffers of the devices that traced the
// photons (one segment on one device).  The grid build reads them in place — over NVLink peer mappings when they live on
// other GPUs — and writes the cell-ordered copy locally: the all-gather of the map is fused into the counting sort.
#define SPPM_MAX_SEGS 16
struct SppmSegs {
  const float4* p[SPPM_MAX_SEGS];
  uint32_t start[SPPM_MAX_SEGS + 1];  // photon i lives in segment s with start[s] <= i < start[s + 1]
  int n;
};
DEV const float4* sppm_seg_photon(const SppmSegs& S, uint32_t i) {
  int s = 0;
  while (s + 1 < S.n && i >= S.start[s + 1]) ++s;
  return S.p[s] + 4 * (size_t)(i - S.start[s]);
}
struct SppmPixels {
  float4* g;    // global: flux rgb, radius2
  float4* c;    // caustic
  uint2* n;     // photons (global, caustic)
};
struct SppmEye {  // what the eye pass hands from its trace kernel to the gather / k-nearest kernels
  float4* vp;          // per pixel: the visible point (position, 1) or (., 0) when the eye path found no diffuse surface
  float4* vb;          // per pixel: bsdf at the visible point
  uint32_t* knn_list;  // (pixel | map << 31) whose statistics are still empty: the k-nearest kernel's work list
  uint32_t* knn_n;
};
struct SppmCounters {
  unsigned int n_global, n_caustic, overflow, cursor;  // cursor: next photon id of the pass (per-lane work fetch)
  unsigned long long photon_rays, eye_rays;
  unsigned int lo[3], hi[3];  // ordered-uint bounds of the stored photon positions
};

enum { SPPM_DIFFUSE = 0, SPPM_SPECULAR = 1, SPPM_ABSORB = 2 };

DEV unsigned int sppm_ord(float f) { unsigned int u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__host__ __device__ inline float sppm_unord(unsigned int u) {
  u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
  float f;
#ifdef __CUDA_ARCH__
  f = __uint_as_float(u);
#else
  std::memcpy(&f, &u, 4);
#endif
  return f;
}

// Material::scatter (material.rs:104-212) with its Interaction: the new direction and the attenuation f.
// u1,u2 drive the direction, u3 the dielectric choice.  Metal fuzz / Lambertian use the same transforms as shade_hit.
DEV int sppm_scatter(const DevScene& S, const float4 mat, const HitInfo& h, const Ray& ray, float u1, float u2, float u3, float3& dir, float3& f) {
  const int mtype = __float_as_int(mat.x), tex = __float_as_int(mat.y);
  if (mtype == B200RT_MAT_DIFFUSE_LIGHT) {  // material.rs:198-208: bsdf 1/pi, scatters like Lambertian
    dir = h.n + sample_unit_sphere(u1, u2);
    if (fabsf(dir.x) < 1e-8f && fabsf(dir.y) < 1e-8f && fabsf(dir.z) < 1e-8f) dir = h.n;
    f = f3(1, 1, 1) * (1.0f / PI_F);
    return SPPM_DIFFUSE;
  }
  f = texture_color(S, tex, h);
  if (mtype == B200RT_MAT_LAMBERTIAN) {
    dir = h.n + sample_unit_sphere(u1, u2);
    if (fabsf(dir.x) < 1e-8f && fabsf(dir.y) < 1e-8f && fabsf(dir.z) < 1e-8f) dir = h.n;
    return SPPM_DIFFUSE;
  }
  if (mtype == B200RT_MAT_ISOTROPIC) { dir = sample_unit_sphere(u1, u2); return SPPM_DIFFUSE; }
  if (mtype == B200RT_MAT_METAL) {
    dir = reflect3(normalize(ray.d), h.n) + mat.z * sample_unit_sphere(u1, u2);
    return dot(dir, h.n) > 0.f ? SPPM_SPECULAR : SPPM_ABSORB;
  }
  const float ratio = h.front_face ? 1.0f / mat.z : mat.z;  // Dielectric (Reflect / Refract are both "specular" for the maps)
  const float3 ud = normalize(ray.d);
  const float cos_theta = fminf(dot(-ud, h.n), 1.0f);
  const float sin_theta = sqrtf(fmaxf(0.f, 1.0f - cos_theta * cos_theta));
  dir = (ratio * sin_theta > 1.0f || schlick(cos_theta, ratio) > u3) ? reflect3(ud, h.n) : refract3(ud, h.n, ratio);
  return SPPM_SPECULAR;
}
// Material::bsdf at a gather point (material.rs:105-107, 128-130, 158-160, 198-200)
DEV float3 sppm_bsdf(const DevScene& S, const float4 mat, const HitInfo& h) {
  return __float_as_int(mat.x) == B200RT_MAT_DIFFUSE_LIGHT ? f3(1, 1, 1) * (1.0f / PI_F) : texture_color(S, __float_as_int(mat.y), h);
}

DEV float3 sppm_light_flux(const DevScene& S, int li) {  // the DiffuseLight colour (light.rs:77, 143) x the light's own scale (light.rs:71,131)
  const float4 L0 = __ldg(&S.lights[3 * li]), L1 = __ldg(&S.lights[3 * li + 1]);
  const float own_scale = __ldg(&S.lights[3 * li + 2]).z;
  const float4 lmat = __ldg(&S.materials[__float_as_int(L0.w)]);
  HitInfo lh;
  lh.p = f3(L1); lh.n = f3(0, 1, 0); lh.u = 0.f; lh.v = 0.f; lh.t = 0.f; lh.front_face = 1;
  return texture_color(S, __float_as_int(lmat.y), lh) * own_scale;
}
// AllLights::emit (light.rs:214-219) -> SphereDiffuseLight::emit (97-102) / XZRectLight::emit (157-165)
DEV void sppm_emit(const DevScene& S, const SppmConst& sp, uint4 r0, uint4 r1, Ray& ray, float3& power) {
  float tot = 0.f;
  for (int i = 0; i < S.n_lights; ++i) tot += length(sppm_light_flux(S, i) * sp.light_scale);
  float pick = u01(r0.x) * tot;
  int li = 0;
  for (; li + 1 < S.n_lights; ++li) {
    const float w = length(sppm_light_flux(S, li) * sp.light_scale);
    if (pick < w) break;
    pick -= w;
  }
  const float4 L0 = __ldg(&S.lights[3 * li]), L1 = __ldg(&S.lights[3 * li + 1]);
  const float3 flux = sppm_light_flux(S, li);
  if (__float_as_int(L0.x) == 0) {
    const float3 n = sample_unit_sphere(u01(r0.y), u01(r0.z));
    float3 d = sample_unit_sphere(u01(r1.x), u01(r1.y));
    if (!(dot(d, n) > 0.f)) d = -d;
    ray.o = f3(L1) + n * (L1.w + 0.0001f); ray.d = d;
    power = flux * sp.light_scale;
  } else {  // rectangle lights emit towards the negative axis (upstream only has the XZ light with n = (0,-1,0))
    const float4 L2 = __ldg(&S.lights[3 * li + 2]);
    const int axis = __float_as_int(L0.z), ia = axis == 0 ? 1 : 0, ib = axis == 2 ? 1 : 2;
    float c[3];
    c[ia] = L1.x + (L1.z - L1.x) * u01(r0.y); c[ib] = L1.y + (L1.w - L1.y) * u01(r0.z); c[axis] = L2.x;
    const float3 n = f3(axis == 0 ? -1.f : 0.f, axis == 1 ? -1.f : 0.f, axis == 2 ? -1.f : 0.f);
    float3 d = sample_unit_sphere(u01(r1.x), u01(r1.y));
    if (!(dot(d, n) > 0.f)) d = -d;
    ray.o = f3(c[0], c[1], c[2]); ray.d = d;
    power = flux * (sp.light_scale * fmaxf(dot(n, d), 0.f));
  }
  ray.time = 0.f; ray.tmax = __int_as_float(0x7f800000); ray.mseed = r1.w;
}

DEV void sppm_store(float4* photons, unsigned int* counter, unsigned int cap, unsigned int* overflow, bool pred, float3 p, float3 power, float3 d, float3 n) {
  const unsigned int at = warp_append(counter, pred);
  if (!pred) return;
  if (at >= cap) { atomicAdd(overflow, 1u); return; }
  float4* o = photons + 4 * (size_t)at;
  o[0] = make_float4(p.x, p.y, p.z, 0.f); o[1] = make_float4(power.x, power.y, power.z, 0.f);
  o[2] = make_float4(d.x, d.y, d.z, 0.f); o[3] = make_float4(n.x, n.y, n.z, 0.f);
}

// SPPMIntegrator::generate_photon_map (photon_mapper.rs:238-281)
template <bool OBJF>
// [id_first, id_first + id_count): the photon ids this launch emits (a shard of the iteration's photons on one of several GPUs)
// q_ray != nullptr: the walks do not start at a light but continue the *q_n records of a wavefront queue (ray stream q_ray, state
// stream q_st; dead records skipped) — the tail of the wavefront photon pass below: the few photons that outlive its bulk.
__global__ void __launch_bounds__(128) k_sppm_photons(DevScene S, SppmConst sp, uint2 key, int iter, float4* pg, float4* pc, unsigned int cap, SppmCounters* cnt,
                                                      unsigned int id_first, unsigned int id_count, const float4* __restrict__ q_ray, const float4* __restrict__ q_st,
                                                      const unsigned int* __restrict__ q_n) {
  unsigned long long n_rays = 0;
  unsigned int lo[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, hi[3] = {0u, 0u, 0u};
  // Per-lane work fetch: a photon walk lasts 1 to dozens of bounces (2.9 on average in the Cornell box), so a warp that
  // waits for its longest photon runs at 5.6 of 32 lanes (ncu, profiles/r1_sppm_full_summary.txt).  Lanes whose photon
  // has died take the next photon ids from a global cursor (one warp-aggregated atomic) as soon as SPPM_REFILL_LANES of
  // them are idle.  Every draw is keyed (photon id, iteration, bounce), so which lane walks which photon changes nothing
  // but the order of the records in the map.  (The path tracer's extend kernel does NOT do this: there the queue order
  // is what keeps a warp coherent; here there is no order to lose.)
  const unsigned lane = threadIdx.x & 31;
  const unsigned int n_photons = q_ray ? *q_n : id_count;  // the cursor counts inside the shard (or the queue)
  unsigned int id = 0;
  int bounce = 0;
  bool alive = false, exhausted = false, has_specular = false, has_diffuse = false;
  Ray ray;
  float3 power = f3(0, 0, 0);
  ray.o = ray.d = f3(0, 0, 0); ray.time = 0.f; ray.tmax = 0.f; ray.mseed = 0u;
  for (;;) {
    const unsigned idle = __ballot_sync(0xffffffffu, !alive);
    if (!exhausted && (__popc(idle) >= SPPM_REFILL_LANES || idle == 0xffffffffu)) {
      const int leader = __ffs((int)idle) - 1;
      unsigned int base = 0;
      if ((int)lane == leader) base = atomicAdd(&cnt->cursor, (unsigned int)__popc(idle));
      base = __shfl_sync(0xffffffffu, base, leader);
      if (!alive) {
        id = base + __popc(idle & ((1u << lane) - 1u));
        const bool have = id < n_photons;
        id += id_first;  // the global photon id keys every draw
        if (have && q_ray) {  // continue a queued walk
          const size_t at = 2 * (size_t)(id - id_first);
          const float4 o = __ldcs(q_ray + at), d = __ldcs(q_ray + at + 1);
          if (!(d.x == 0.f && d.y == 0.f && d.z == 0.f)) {
            const float4 sa = __ldcs(q_st + at), sb = __ldcs(q_st + at + 1);
            ray.o = f3(o); ray.time = o.w; ray.d = f3(d); ray.tmax = __int_as_float(0x7f800000); ray.mseed = __float_as_uint(d.w);
            power = f3(sa);
            has_specular = (__float_as_uint(sa.w) & 1u) != 0u; has_diffuse = (__float_as_uint(sa.w) & 2u) != 0u;
            id = __float_as_uint(sb.x); bounce = (int)__float_as_uint(sb.y);
            alive = true;
          }
        } else if (have) {
          sppm_emit(S, sp, philox4x32(make_uint4(id, (uint32_t)iter, 0u, 0x50u), key), philox4x32(make_uint4(id, (uint32_t)iter, 0u, 0x51u), key), ray, power);
          alive = true; bounce = 1; has_specular = false; has_diffuse = false;
        }
      }
      exhausted = base + (unsigned int)__popc(idle) >= n_photons;  // warp-uniform
    }
    if (!__any_sync(0xffffffffu, alive)) break;  // only reached with the cursor exhausted
    {
      bool st_g = false, st_c = false;
      float3 hp = f3(0, 0, 0), hn = f3(0, 0, 0), hd = f3(0, 0, 0), hpow = f3(0, 0, 0);
      if (alive) {
        HitRaw hr;
        trace<false, OBJF>(S, ray, 0.0001f, hr);  // photon_mapper.rs:246
        n_rays++;
        if (hr.prim < 0) alive = false;
        else {
          const int4 meta = __ldg(&S.prim_meta[hr.prim]);
          const float4 mat = __ldg(&S.materials[meta.x >> 8]);
          HitInfo h;
          finish_hit(S, ray, hr, h, texture_needs_uv(S, __float_as_int(mat.y)));
          const uint4 r = philox4x32(make_uint4(id, (uint32_t)iter, (uint32_t)bounce, 0x50u), key);
          float3 dir, f;
          int inter = sppm_scatter(S, mat, h, ray, u01(r.x), u01(r.y), u01(r.z), dir, f);
          float3 new_power = power;
          if (inter != SPPM_ABSORB) {  // scatter_photon (material.rs:27-45): Russian roulette on max(f)
            const float hmax = maxcomp(f);
            if (u01(r.w) > hmax) inter = SPPM_ABSORB;
            else new_power = power * (f * (1.0f / hmax));
          }
          if (inter == SPPM_DIFFUSE) {
            st_g = true; st_c = !has_diffuse && has_specular;
            hp = h.p; hn = h.n; hd = ray.d; hpow = power;
            has_diffuse = true;
          } else if (inter == SPPM_ABSORB) alive = false;
          else has_specular = true;
          if (alive && bounce >= SPPM_MAX_BOUNCES) alive = false;  // upstream loops until absorbed or lost; the cap only guards the kernel
          if (alive) {
            ray.o = h.p; ray.d = dir; ray.tmax = __int_as_float(0x7f800000);
            ray.mseed = philox4x32(make_uint4(id, (uint32_t)iter, (uint32_t)bounce, 0x51u), key).x;
            power = new_power;
            ++bounce;
          }
        }
      }
      sppm_store(pg, &cnt->n_global, cap, &cnt->overflow, st_g, hp, hpow, hd, hn);
      sppm_store(pc, &cnt->n_caustic, cap, &cnt->overflow, st_c, hp, hpow, hd, hn);
      if (st_g) {
        lo[0] = min(lo[0], sppm_ord(hp.x)); lo[1] = min(lo[1], sppm_ord(hp.y)); lo[2] = min(lo[2], sppm_ord(hp.z));
        hi[0] = max(hi[0], sppm_ord(hp.x)); hi[1] = max(hi[1], sppm_ord(hp.y)); hi[2] = max(hi[2], sppm_ord(hp.z));
      }
    }
  }
  for (int a = 0; a < 3; ++a) {
    for (int o = 16; o > 0; o >>= 1) { lo[a] = min(lo[a], __shfl_down_sync(0xffffffffu, lo[a], o)); hi[a] = max(hi[a], __shfl_down_sync(0xffffffffu, hi[a], o)); }
    if (lane == 0) { atomicMin(&cnt->lo[a], lo[a]); atomicMax(&cnt->hi[a], hi[a]); }
  }
  for (int o = 16; o > 0; o >>= 1) n_rays += __shfl_down_sync(0xffffffffu, n_rays, o);
  if (lane == 0 && n_rays) atomicAdd(&cnt->photon_rays, n_rays);
}

// ---- the photon pass through the wavefront queues (round 2) --------------------------------------------------------
// k_sppm_photons above follows one photon per thread through emission, traversal, hit record, texture, scatter and roulette:
// 150 registers, and its warps spend 5 issue slots waiting for instructions for every one they use (stall no_instruction 5.0
// per issue, profiles/r1_sppm_photons_refill_full.txt) — divergent lanes walk a code footprint no instruction cache holds.
// Here the photons ride the path tracer's queues instead: k_sppm_ph_generate emits (AllLights::emit), the SAME k_wf_extend
// traverses (t_min 0.0001, every hit "goes on" — DiffuseLight scatters photons, material.rs:198-208), and k_sppm_ph_shade
// streams the queue in order: hit record, scatter_photon with its Russian roulette (material.rs:27-45), photon records
// appended to the maps, the survivor stored at offset[bin] + rank of the next set (an absorbed photon leaves a dead record).
// State stream of a photon: (power rgb, flags: 1 = has_specular, 2 = has_diffuse) (photon id, bounce, 0, 0).
// Every draw is keyed (photon id, iteration, bounce) exactly as in k_sppm_photons: the stored photons are the same set.
__global__ void __launch_bounds__(256) k_sppm_ph_generate(DevScene S, WfBuffers B, PathMap M, SppmConst sp, uint2 key, int iter, unsigned int id_first) {
  const unsigned int n_new = wf_new_paths(B.cnt, M, B.pool);
  const unsigned long long base = B.cnt->next_path;
  const unsigned int off = B.cnt->n_ext_next;
  for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_new; i += gridDim.x * blockDim.x) {
    const uint32_t id = id_first + (uint32_t)(base + i);
    Ray r;
    float3 power;
    sppm_emit(S, sp, philox4x32(make_uint4(id, (uint32_t)iter, 0u, 0x50u), key), philox4x32(make_uint4(id, (uint32_t)iter, 0u, 0x51u), key), r, power);
    float4* ray = B.ray + 2 * (size_t)(off + i);
    float4* st = B.st + 2 * (size_t)(off + i);
    WF_ST(ray + 0, make_float4(r.o.x, r.o.y, r.o.z, r.time));
    WF_ST(ray + 1, make_float4(r.d.x, r.d.y, r.d.z, __uint_as_float(r.mseed)));
    WF_ST(st + 0, make_float4(power.x, power.y, power.z, __uint_as_float(0u)));
    WF_ST(st + 1, make_float4(__uint_as_float(id), __uint_as_float(1u), 0.f, 0.f));
  }
}
// k_wf_control + a log of the queue length per wavefront iteration (the host sizes the next pass's launch count from it)
__global__ void k_sppm_ph_control(WfBuffers B, PathMap M, unsigned int* n_log, int j) {
  for (unsigned int k = threadIdx.x; k < B.n_bins; k += blockDim.x) B.bin_count[k] = 0;
  if (threadIdx.x != 0) return;
  WfCounters* c = B.cnt;
  const unsigned int n_new = wf_new_paths(c, M, B.pool);
  c->next_path += n_new;
  c->n_ext = c->n_ext_next + n_new;
  c->n_ext_next = 0;
  c->ext_cursor = 0;
  c->extend_rays += c->n_ext;
  n_log[j] = c->n_ext;
}
__global__ void __launch_bounds__(128, 4) k_sppm_ph_shade(DevScene S, WfBuffers B, uint2 key, int iter, float4* pg, float4* pc, unsigned int cap, SppmCounters* cnt) {
  __shared__ unsigned int s_off[WF_SMEM_BINS + 1];
  const bool off_in_smem = B.n_bins <= WF_SMEM_BINS;
  if (off_in_smem) for (unsigned int k = threadIdx.x; k <= B.n_bins; k += blockDim.x) s_off[k] = B.bin_off[k];
  __syncthreads();
  const unsigned int n_ext = B.cnt->n_ext;
  const unsigned int n_work = (n_ext + 31u) & ~31u;  // whole warps stay in the loop (warp-aggregated appends)
  unsigned int lo[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, hi[3] = {0u, 0u, 0u};
  for (unsigned int gi = blockIdx.x * blockDim.x + threadIdx.x; gi < n_work; gi += gridDim.x * blockDim.x) {
    bool st_g = false, st_c = false, alive = false;
    float3 hp = f3(0, 0, 0), hn = f3(0, 0, 0), hd = f3(0, 0, 0), hpow = f3(0, 0, 0), dir = f3(0, 0, 0), new_power = f3(0, 0, 0);
    uint32_t id = 0, bounce = 0, flags = 0, next_seed = 0;
    unsigned int rank = WF_NO_RANK;
    int hprim = WF_PRIM_DEAD;
    float time = 0.f;
    if (gi < n_ext) {
      const uint4 hh = WF_LD(B.hit + gi);
      hprim = wf_unpack_prim(hh.z);
      rank = hh.w;
      if (hprim >= 0) {
        const float4 o = WF_LD(B.ray + 2 * (size_t)gi), d = WF_LD(B.ray + 2 * (size_t)gi + 1);
        const float4 sa = WF_LD(B.st + 2 * (size_t)gi), sb = WF_LD(B.st + 2 * (size_t)gi + 1);
        Ray ray;
        ray.o = f3(o); ray.time = o.w; ray.d = f3(d); ray.tmax = __int_as_float(0x7f800000); ray.mseed = __float_as_uint(d.w);
        time = o.w;
        const float3 power = f3(sa);
        flags = __float_as_uint(sa.w); id = __float_as_uint(sb.x); bounce = __float_as_uint(sb.y);
        HitRaw hr;
        hr.t = __longlong_as_double((long long)(((unsigned long long)hh.y << 32) | hh.x)); hr.prim = hprim; hr.aux = wf_unpack_aux(hh.z);
        const int4 meta = __ldg(&S.prim_meta[hprim]);
        const float4 mat = __ldg(&S.materials[meta.x >> 8]);
        HitInfo h;
        finish_hit(S, ray, hr, h, texture_needs_uv(S, __float_as_int(mat.y)));
        const uint4 r = philox4x32(make_uint4(id, (uint32_t)iter, bounce, 0x50u), key);
        float3 f;
        int inter = sppm_scatter(S, mat, h, ray, u01(r.x), u01(r.y), u01(r.z), dir, f);
        new_power = power;
        if (inter != SPPM_ABSORB) {  // scatter_photon (material.rs:27-45): Russian roulette on max(f)
          const float hmax = maxcomp(f);
          if (u01(r.w) > hmax) inter = SPPM_ABSORB;
          else new_power = power * (f * (1.0f / hmax));
        }
        alive = inter != SPPM_ABSORB;
        if (inter == SPPM_DIFFUSE) {
          st_g = true; st_c = !(flags & 2u) && (flags & 1u);
          hp = h.p; hn = h.n; hd = ray.d; hpow = power;
          flags |= 2u;
        } else if (inter == SPPM_SPECULAR) flags |= 1u;
        if (alive && bounce >= SPPM_MAX_BOUNCES) alive = false;  // upstream loops until absorbed or lost; the cap only guards the kernel
        if (alive && rank != WF_NO_RANK) {  // the next record, at the place extend reserved in the next set
          next_seed = philox4x32(make_uint4(id, (uint32_t)iter, bounce, 0x51u), key).x;
          const unsigned int bin = 1u + __umulhi((unsigned int)hprim, B.bin_mul);
          const size_t dst = 2 * (size_t)((off_in_smem ? s_off[bin] : __ldg(&B.bin_off[bin])) + rank);
          WF_ST(B.ray_n + dst, make_float4(h.p.x, h.p.y, h.p.z, time));
          WF_ST(B.ray_n + dst + 1, make_float4(dir.x, dir.y, dir.z, __uint_as_float(next_seed)));
          WF_ST(B.st_n + dst, make_float4(new_power.x, new_power.y, new_power.z, __uint_as_float(flags)));
          WF_ST(B.st_n + dst + 1, make_float4(__uint_as_float(id), __uint_as_float(bounce + 1u), 0.f, 0.f));
        }
      }
      if (!alive && rank != WF_NO_RANK && hprim >= 0) {  // predicted to go on, absorbed here: a dead record extend will skip
        const unsigned int bin = 1u + __umulhi((unsigned int)hprim, B.bin_mul);
        const size_t dst = 2 * (size_t)((off_in_smem ? s_off[bin] : __ldg(&B.bin_off[bin])) + rank);
        WF_ST(B.ray_n + dst, make_float4(0.f, 0.f, 0.f, 0.f));
        WF_ST(B.ray_n + dst + 1, make_float4(0.f, 0.f, 0.f, 0.f));
      }
    }
    sppm_store(pg, &cnt->n_global, cap, &cnt->overflow, st_g, hp, hpow, hd, hn);
    sppm_store(pc, &cnt->n_caustic, cap, &cnt->overflow, st_c, hp, hpow, hd, hn);
    if (st_g) {
      lo[0] = min(lo[0], sppm_ord(hp.x)); lo[1] = min(lo[1], sppm_ord(hp.y)); lo[2] = min(lo[2], sppm_ord(hp.z));
      hi[0] = max(hi[0], sppm_ord(hp.x)); hi[1] = max(hi[1], sppm_ord(hp.y)); hi[2] = max(hi[2], sppm_ord(hp.z));
    }
  }
  for (int a = 0; a < 3; ++a) {
    for (int o = 16; o > 0; o >>= 1) { lo[a] = min(lo[a], __shfl_down_sync(0xffffffffu, lo[a], o)); hi[a] = max(hi[a], __shfl_down_sync(0xffffffffu, hi[a], o)); }
    if ((threadIdx.x & 31) == 0 && lo[a] <= hi[a]) { atomicMin(&cnt->lo[a], lo[a]); atomicMax(&cnt->hi[a], hi[a]); }
  }
}

// ---- the photon grid: counting sort by cell ----------------------------------------------------------------------
DEV int3 sppm_cell_of(float3 p, float3 lo, float inv_cell, int res) {
  return make_int3(min(res - 1, max(0, (int)floorf((p.x - lo.x) * inv_cell))), min(res - 1, max(0, (int)floorf((p.y - lo.y) * inv_cell))),
                   min(res - 1, max(0, (int)floorf((p.z - lo.z) * inv_cell))));
}
DEV uint32_t sppm_cell_id(int3 c, int res) { return (uint32_t)((c.z * res + c.y) * res + c.x); }

__global__ void k_sppm_grid_count(SppmSegs photons, uint32_t n, float3 lo, float inv_cell, int res, uint32_t* cell_count) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    atomicAdd(&cell_count[sppm_cell_id(sppm_cell_of(f3(*sppm_seg_photon(photons, i)), lo, inv_cell, res), res)], 1u);
}
// Exclusive scan of n_cells counts into cell_start (n_cells + 1 entries); cursor = copy of the starts.  Three small
// launches over 4096-cell tiles (tile sums -> scan of the <= 512 tile sums -> scan inside every tile); the first
// version walked all cells from one block with a 1.5 KB stride per thread: 340 us per map at 389k cells.
#define SPPM_SCAN_TILE 4096
DEV uint32_t sppm_block_exclusive_scan(uint32_t v, uint32_t* ws /* 33 */, uint32_t& total) {  // blockDim.x == 1024
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t inc = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= (unsigned)d) inc += o; }
  if (lane == 31) ws[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = ws[lane], winc = w;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t o = __shfl_up_sync(0xffffffffu, winc, d); if (lane >= (unsigned)d) winc += o; }
    ws[lane] = winc - w;
    if (lane == 31) ws[32] = winc;
  }
  __syncthreads();
  total = ws[32];
  return ws[warp] + inc - v;
}
__global__ void __launch_bounds__(1024) k_sppm_scan_tile_sums(const uint32_t* __restrict__ cell_count, uint32_t n_cells, uint32_t* __restrict__ tile_sum) {
  __shared__ uint32_t ws[33];
  const uint32_t c0 = blockIdx.x * SPPM_SCAN_TILE + 4u * threadIdx.x;
  uint32_t v = 0;
#pragma unroll
  for (uint32_t k = 0; k < 4; ++k) if (c0 + k < n_cells) v += cell_count[c0 + k];
  uint32_t total;
  sppm_block_exclusive_scan(v, ws, total);
  if (threadIdx.x == 0) tile_sum[blockIdx.x] = total;
}
__global__ void __launch_bounds__(1024) k_sppm_scan_tile_offsets(uint32_t* tile_sum, uint32_t n_tiles, uint32_t* cell_start, uint32_t n_cells) {  // one block, n_tiles <= 1024
  __shared__ uint32_t ws[33];
  const uint32_t v = threadIdx.x < n_tiles ? tile_sum[threadIdx.x] : 0u;
  uint32_t total;
  const uint32_t ex = sppm_block_exclusive_scan(v, ws, total);
  if (threadIdx.x < n_tiles) tile_sum[threadIdx.x] = ex;
  if (threadIdx.x == 0) cell_start[n_cells] = total;
}
__global__ void __launch_bounds__(1024) k_sppm_scan_cells(const uint32_t* __restrict__ cell_count, uint32_t n_cells, const uint32_t* __restrict__ tile_off,
                                                          uint32_t* __restrict__ cell_start, uint32_t* __restrict__ cursor) {
  __shared__ uint32_t ws[33];
  const uint32_t c0 = blockIdx.x * SPPM_SCAN_TILE + 4u * threadIdx.x;
  uint32_t c[4], v = 0;
#pragma unroll
  for (uint32_t k = 0; k < 4; ++k) { c[k] = c0 + k < n_cells ? cell_count[c0 + k] : 0u; v += c[k]; }
  uint32_t total;
  uint32_t run = tile_off[blockIdx.x] + sppm_block_exclusive_scan(v, ws, total);
#pragma unroll
  for (uint32_t k = 0; k < 4; ++k) {
    if (c0 + k < n_cells) { cell_start[c0 + k] = run; cursor[c0 + k] = run; }
    run += c[k];
  }
}
// The photon records themselves are moved into cell order, split into what a gather tests (16-byte positions: the run of a
// grid row is one contiguous read) and what an accepted photon adds (power, normal: one 32-byte sector).
__global__ void k_sppm_grid_fill(SppmSegs photons, uint32_t n, float3 lo, float inv_cell, int res, uint32_t* cursor, float4* __restrict__ pos_out,
                                 float4* __restrict__ attr_out) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float4* src = sppm_seg_photon(photons, i);
    const float4 a = src[0], b = src[1], d = src[3];
    const uint32_t at = atomicAdd(&cursor[sppm_cell_id(sppm_cell_of(f3(a), lo, inv_cell, res), res)], 1u);
    pos_out[at] = a;
    attr_out[2 * (size_t)at] = b; attr_out[2 * (size_t)at + 1] = d;
  }
}

// ---- SPPM::update (photon_mapper.rs:42-64) over the grid ---------------------------------------------------------
// The eye pass is three kernels (round 2; the first version followed one pixel per thread through the trace AND both map
// look-ups: 6.9 active lanes per instruction, 14 % issue-active, the k-nearest of the first iteration alone 45 of 232 ms):
//   k_sppm_eye_trace : one thread per pixel, the eye ray through the specular chain to its first diffuse hit; stores the visible
//                      point (position, bsdf) and lists the (pixel, map) pairs whose statistics are still empty
//   k_sppm_gather    : SPPM_GATHER_LANES lanes per pixel share the radius gather — the cells a pixel's sphere overlaps are
//                      walked as x-runs (the photons of consecutive cells of one grid row are one contiguous run of the
//                      cell-ordered array), one run per lane at a time
//   k_sppm_knn       : one warp per listed (pixel, map): exact k-nearest by growing cubes of cells, the squared distances of
//                      the cube's photons kept in shared memory, the k-th smallest found by bisection on the float bits
// Squared distance of a photon, one rounding sequence at every site (the k-nearest kernel compares distances computed in
// different passes bit for bit).
DEV float sppm_d2(float4 pos, float3 p) {
  const float dx = __fsub_rn(pos.x, p.x), dy = __fsub_rn(pos.y, p.y), dz = __fsub_rn(pos.z, p.z);
  return __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmul_rn(dx, dx)));
}
// Vec3::elemul(rec.mat.bsdf(photon.direction, rec), photon.power) * (1 - disk_factor) (photon_mapper.rs:72-74, 90-92)
DEV float3 sppm_photon_flux(const SppmGrid& G, uint32_t i, float4 pos4, float3 p, float3 bsdf) {
  const float4* at = G.attr + 2 * (size_t)i;
  const float3 power = f3(__ldg(at)), nrm = f3(__ldg(at + 1));
  const float3 d = f3(pos4) - p;
  const float disk = fabsf(dot(nrm, d * rsqrtf(dot(d, d))));
  return bsdf * power * (1.0f - disk);
}

// SPPMIntegrator::update_sppm (photon_mapper.rs:282-305), first half: one jittered eye ray per pixel and iteration
// [pix_first, pix_first + pix_count): the pixels whose statistics this device owns
template <bool OBJF>
__global__ void __launch_bounds__(128) k_sppm_eye_trace(DevScene S, RenderConst P, int iter, SppmPixels px, SppmEye E, SppmCounters* cnt, uint32_t pix_first, uint32_t pix_count) {
  unsigned long long n_rays = 0;
  const uint32_t n_work = (pix_count + 31u) & ~31u;  // whole warps walk the loop (warp-aggregated appends)
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_work; i += gridDim.x * blockDim.x) {
    const bool active = i < pix_count;
    const uint32_t pixel = pix_first + i;
    bool found = false;
    float3 hp = f3(0, 0, 0), hb = f3(0, 0, 0);
    if (active) {
      Ray ray = camera_ray(P, pixel, (uint32_t)iter);
      for (int bounce = 1; bounce <= SPPM_MAX_BOUNCES; ++bounce) {
        HitRaw hr;
        trace<false, OBJF>(S, ray, 0.001f, hr);  // photon_mapper.rs:290
        n_rays++;
        if (hr.prim < 0) break;
        const int4 meta = __ldg(&S.prim_meta[hr.prim]);
        const float4 mat = __ldg(&S.materials[meta.x >> 8]);
        HitInfo h;
        finish_hit(S, ray, hr, h, texture_needs_uv(S, __float_as_int(mat.y)));
        const uint4 r = philox4x32(make_uint4(pixel, (uint32_t)iter, (uint32_t)bounce, 0x60u), P.key);
        float3 dir, f;
        const int inter = sppm_scatter(S, mat, h, ray, u01(r.x), u01(r.y), u01(r.z), dir, f);
        if (inter == SPPM_DIFFUSE) { found = true; hp = h.p; hb = sppm_bsdf(S, mat, h); break; }
        if (inter == SPPM_ABSORB) break;
        ray.o = h.p; ray.d = dir; ray.tmax = __int_as_float(0x7f800000); ray.mseed = r.w;
      }
      E.vp[pixel] = make_float4(hp.x, hp.y, hp.z, found ? 1.f : 0.f);
      if (found) E.vb[pixel] = make_float4(hb.x, hb.y, hb.z, 0.f);
    }
    uint2 n = make_uint2(1u, 1u);
    if (found) n = px.n[pixel];
    const unsigned int at_c = warp_append(E.knn_n, found && n.y == 0u);  // caustic first, like upstream
    if (found && n.y == 0u) E.knn_list[at_c] = pixel | 0x80000000u;
    const unsigned int at_g = warp_append(E.knn_n, found && n.x == 0u);
    if (found && n.x == 0u) E.knn_list[at_g] = pixel;
  }
  for (int o = 16; o > 0; o >>= 1) n_rays += __shfl_down_sync(0xffffffffu, n_rays, o);
  if ((threadIdx.x & 31) == 0 && n_rays) atomicAdd(&cnt->eye_rays, n_rays);
}

// PhotonMap::estimate_flux_within_radius (photon_mapper.rs:99-114: squared distance < radius^2) + SPPM::update (49-64),
// SPPM_GATHER_LANES lanes per pixel.  The photons of the cells [a.x, b.x] of one grid row are the contiguous run
// cell_start[row + a.x] .. cell_start[row + b.x + 1] of the cell-ordered array.  The lanes first fetch the bounds of up to
// SPPM_GATHER_LANES rows at once (one row per lane: the dependent cell_start loads of all rows are in flight together), then
// walk the rows one after the other with the lanes striding over the row's photons.  (One row per lane, the first version, ran
// at 9 of 32 lanes: photons sit on surfaces, so 1-3 of a pixel's 9-16 rows hold all of them — ~24 per surface cell here.)
#ifndef SPPM_GATHER_LANES
#define SPPM_GATHER_LANES 8
#endif
#ifndef SPPM_GATHER_CLIP
#define SPPM_GATHER_CLIP 1
#endif
DEV void sppm_gather_map(const SppmGrid& G, bool on, float3 p, float3 bsdf, float radius2, unsigned sub, float3& flux, uint32_t& count) {
  flux = f3(0, 0, 0); count = 0;
  const float r = sqrtf(radius2), rr = r * r;
  const int3 a = sppm_cell_of(p - f3(r, r, r), G.lo, G.inv_cell, G.res), b = sppm_cell_of(p + f3(r, r, r), G.lo, G.inv_cell, G.res);
  const int ny = b.y - a.y + 1, rows = (on && G.n != 0) ? ny * (b.z - a.z + 1) : 0;  // uniform over the pixel's lanes; the shuffles below are reached by every lane
  const unsigned lane = threadIdx.x & 31u, group_base = lane & ~(unsigned)(SPPM_GATHER_LANES - 1);
  const int max_rows = __reduce_max_sync(0xffffffffu, rows);  // the groups of a warp walk their row batches together
  for (int r0 = 0; r0 < max_rows; r0 += SPPM_GATHER_LANES) {
    uint32_t j0 = 0, j1 = 0;
    const int mine = r0 + (int)sub;
    if (mine < rows) {
      // the cells of this row the SPHERE reaches, not the box: the row's distance from the point in y and z leaves
      // sqrt(r^2 - dy^2 - dz^2) along x (cell faces moved out by 1e-4 cell: a photon is binned with one float rounding)
      const int y = a.y + mine % ny, z = a.z + mine / ny;
      const float slack = 1e-4f * G.cell;
      const float ylo = G.lo.y + (float)y * G.cell, zlo = G.lo.z + (float)z * G.cell;
      const float dy = fmaxf(0.f, fmaxf(ylo - slack - p.y, p.y - (ylo + G.cell + slack)));
      const float dz = fmaxf(0.f, fmaxf(zlo - slack - p.z, p.z - (zlo + G.cell + slack)));
#if SPPM_GATHER_CLIP
      const float rem = rr - dy * dy - dz * dz;
#else
      const float rem = 1e30f + 0.f * (dy + dz);  // A/B: the whole box (tools/dbg/sppm_clip_check.py compares the per-pixel photon counts)
#endif
      if (rem >= 0.f) {
        const float dx = sqrtf(rem) + slack;
        const int xa = max(a.x, min(G.res - 1, max(0, (int)floorf((p.x - dx - G.lo.x) * G.inv_cell))));
        const int xb = min(b.x, min(G.res - 1, max(0, (int)floorf((p.x + dx - G.lo.x) * G.inv_cell))));
        const uint32_t base = (uint32_t)((z * G.res + y) * G.res);
        if (xa <= xb) { j0 = __ldg(&G.cell_start[base + (uint32_t)xa]); j1 = __ldg(&G.cell_start[base + (uint32_t)xb + 1u]); }
      }
    }
#pragma unroll
    for (int k = 0; k < SPPM_GATHER_LANES; ++k) {
      const uint32_t e = __shfl_sync(0xffffffffu, j1, group_base + k);
      for (uint32_t j = __shfl_sync(0xffffffffu, j0, group_base + k) + sub; j < e; j += SPPM_GATHER_LANES) {
        const float4 pos = __ldg(G.pos + j);
        if (sppm_d2(pos, p) < rr) { ++count; flux = flux + sppm_photon_flux(G, j, pos, p, bsdf); }
      }
    }
  }
#pragma unroll
  for (int o = SPPM_GATHER_LANES / 2; o > 0; o >>= 1) {
    flux.x += __shfl_xor_sync(0xffffffffu, flux.x, o); flux.y += __shfl_xor_sync(0xffffffffu, flux.y, o); flux.z += __shfl_xor_sync(0xffffffffu, flux.z, o);
    count += __shfl_xor_sync(0xffffffffu, count, o);
  }
}
DEV void sppm_update_gathered(float3 flux, uint32_t m, double alpha, float4& st, uint32_t& photons) {  // photon_mapper.rs:55-63
  const uint32_t prev = photons;
  photons += (uint32_t)(alpha * (double)m);
  const float frac = (float)((double)photons / (double)(prev + m));
  st = make_float4((st.x + flux.x) * frac, (st.y + flux.y) * frac, (st.z + flux.z) * frac, st.w * frac);
}
__global__ void __launch_bounds__(256) k_sppm_gather(SppmGrid Gg, SppmGrid Gc, SppmPixels px, SppmEye E, double alpha, uint32_t pix_first, uint32_t pix_count) {
  const unsigned sub = threadIdx.x % SPPM_GATHER_LANES;
  const uint32_t per_pass = gridDim.x * blockDim.x / SPPM_GATHER_LANES;
  const uint32_t n_work = (pix_count + (32u / SPPM_GATHER_LANES) - 1u) / (32u / SPPM_GATHER_LANES) * (32u / SPPM_GATHER_LANES);  // whole warps (shuffles)
  for (uint32_t i = (blockIdx.x * blockDim.x + threadIdx.x) / SPPM_GATHER_LANES; i < n_work; i += per_pass) {
    const uint32_t pixel = pix_first + i;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < pix_count) v = __ldg(&E.vp[pixel]);
    const bool valid = v.w != 0.f;  // uniform over the pixel's lanes
    uint2 n = make_uint2(0u, 0u);
    float3 bsdf = f3(0, 0, 0);
    if (valid) { n = px.n[pixel]; bsdf = f3(__ldg(&E.vb[pixel])); }
    const float3 p = f3(v);
    // a map whose statistics are still empty (n == 0) is the k-nearest kernel's; the shuffles below are reached by all lanes
    float4 c = make_float4(0.f, 0.f, 0.f, 0.f), g = c;
    if (valid && n.y) c = px.c[pixel];
    if (valid && n.x) g = px.g[pixel];
    float3 fc, fg;
    uint32_t mc, mg;
    sppm_gather_map(Gc, valid && n.y != 0u, p, bsdf, c.w, sub, fc, mc);
    sppm_gather_map(Gg, valid && n.x != 0u, p, bsdf, g.w, sub, fg, mg);
    if (sub == 0 && valid) {
      if (n.y) { sppm_update_gathered(fc, mc, alpha, c, n.y); px.c[pixel] = c; reinterpret_cast<uint32_t*>(&px.n[pixel])[1] = n.y; }
      if (n.x) { sppm_update_gathered(fg, mg, alpha, g, n.x); px.g[pixel] = g; reinterpret_cast<uint32_t*>(&px.n[pixel])[0] = n.x; }
    }
  }
}

// PhotonMap::estimate_flux_by_count (photon_mapper.rs:77-96): the k nearest photons, exactly — one warp per (pixel, map).
// Cube s = the cells within Chebyshev distance s of the point's cell; every photon outside it lies more than s cells away
// along some axis.  The warp grows s (from the cell table alone) until the cube holds k photons, selects the k-th smallest
// squared distance among them — an upper bound of the true one — and, if that reaches beyond s cells, jumps to the cube that
// covers it and selects again: then the k nearest of the cube are the k nearest of the map.  Selection: the cube's squared
// distances go to shared memory and a quickselect (count against a pivot, compact the side that holds the k-th) narrows them
// down; a cube with more photons than the buffer holds is first narrowed by value — a 2048-bin histogram of the float bits
// over the current interval, refined until the bin with the k-th fits.  The flux of the photons up to the k-th distance is
// what kd_tree.nearests(k) returns, its largest squared distance the initial radius^2 (photon_mapper.rs:80-95).
// (First version: bisection on the float bits — 31 counting passes per point, over the grid itself for a large cube, and a
// fill + count per shell: 19.7 ms for the first iteration at the reference configuration.)
#define SPPM_KNN_CAP 2048
struct SppmCube { int x0, x1, y0, y1, z0, z1; };
DEV SppmCube sppm_cube(const SppmGrid& G, int3 c0, int s) {
  SppmCube q;
  q.x0 = max(0, c0.x - s); q.x1 = min(G.res - 1, c0.x + s); q.y0 = max(0, c0.y - s); q.y1 = min(G.res - 1, c0.y + s);
  q.z0 = max(0, c0.z - s); q.z1 = min(G.res - 1, c0.z + s);
  return q;
}
// Every photon of the cube, the whole warp on one grid row at a time: the lanes fetch the bounds of 32 rows together, then
// stride over the photons of each non-empty row (the x-run of a row is contiguous in the cell-ordered array).  f(j, position,
// ordinal) with ordinal = the photon's place in this enumeration.  (One row per lane, the first version, ran at 2-5 of 32
// lanes: photons sit on surfaces, so a few of a cube's rows hold all of them.)  Called with the warp converged.
template <class F>
DEV void sppm_for_each_in_cube(const SppmGrid& G, const SppmCube& q, unsigned lane, F f) {
  const int ny = q.y1 - q.y0 + 1, rows = ny * (q.z1 - q.z0 + 1);
  uint32_t run = 0;
  for (int r0 = 0; r0 < rows; r0 += 32) {
    uint32_t j0 = 0, j1 = 0;
    const int mine = r0 + (int)lane;
    if (mine < rows) {
      const uint32_t base = (uint32_t)(((q.z0 + mine / ny) * G.res + q.y0 + mine % ny) * G.res);
      j0 = __ldg(&G.cell_start[base + (uint32_t)q.x0]); j1 = __ldg(&G.cell_start[base + (uint32_t)q.x1 + 1u]);
    }
    unsigned nonempty = __ballot_sync(0xffffffffu, j1 > j0);
    while (nonempty) {
      const int k = __ffs((int)nonempty) - 1;
      nonempty &= nonempty - 1u;
      const uint32_t b0 = __shfl_sync(0xffffffffu, j0, k), b1 = __shfl_sync(0xffffffffu, j1, k);
      for (uint32_t j = b0 + lane; j < b1; j += 32) f(j, __ldg(G.pos + j), run + (j - b0));
      run += b1 - b0;
    }
  }
}
// The same enumeration restricted to the cells a SPHERE around p reaches (radius^2 = rr): per row, the x-range left by the
// row's distance from p in y and z (cell faces moved out by 1e-4 cell, as in the radius gather).  The inner loop has a
// warp-uniform trip count, so f may vote: f(j, position, valid).
template <class F>
DEV void sppm_for_each_in_ball(const SppmGrid& G, const SppmCube& q, float3 p, float rr, unsigned lane, F f) {
  const int ny = q.y1 - q.y0 + 1, rows = ny * (q.z1 - q.z0 + 1);
  const float slack = 1e-4f * G.cell;
  for (int r0 = 0; r0 < rows; r0 += 32) {
    uint32_t j0 = 0, j1 = 0;
    const int mine = r0 + (int)lane;
    if (mine < rows) {
      const int y = q.y0 + mine % ny, z = q.z0 + mine / ny;
      const float ylo = G.lo.y + (float)y * G.cell, zlo = G.lo.z + (float)z * G.cell;
      const float dy = fmaxf(0.f, fmaxf(ylo - slack - p.y, p.y - (ylo + G.cell + slack)));
      const float dz = fmaxf(0.f, fmaxf(zlo - slack - p.z, p.z - (zlo + G.cell + slack)));
      const float rem = rr - dy * dy - dz * dz;
      if (rem >= 0.f) {
        const float dx = sqrtf(rem) + slack;
        const int xa = max(q.x0, min(G.res - 1, max(0, (int)floorf((p.x - dx - G.lo.x) * G.inv_cell))));
        const int xb = min(q.x1, min(G.res - 1, max(0, (int)floorf((p.x + dx - G.lo.x) * G.inv_cell))));
        const uint32_t base = (uint32_t)((z * G.res + y) * G.res);
        if (xa <= xb) { j0 = __ldg(&G.cell_start[base + (uint32_t)xa]); j1 = __ldg(&G.cell_start[base + (uint32_t)xb + 1u]); }
      }
    }
    unsigned nonempty = __ballot_sync(0xffffffffu, j1 > j0);
    while (nonempty) {
      const int k = __ffs((int)nonempty) - 1;
      nonempty &= nonempty - 1u;
      const uint32_t b0 = __shfl_sync(0xffffffffu, j0, k), b1 = __shfl_sync(0xffffffffu, j1, k);
      for (uint32_t jb = b0; jb < b1; jb += 32) {
        const uint32_t j = jb + lane;
        const bool valid = j < b1;
        f(j, valid ? __ldg(G.pos + j) : make_float4(0.f, 0.f, 0.f, 0.f), valid);
      }
    }
  }
}
DEV uint32_t sppm_warp_sum(uint32_t v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// photons in the cube (all lanes) and this lane's exclusive share of them, from the cell table alone
DEV uint32_t sppm_cube_count(const SppmGrid& G, const SppmCube& q, unsigned lane, uint32_t& before_me) {
  const int ny = q.y1 - q.y0 + 1, rows = ny * (q.z1 - q.z0 + 1);
  uint32_t mine = 0;
  for (int row = (int)lane; row < rows; row += 32) {
    const uint32_t base = (uint32_t)(((q.z0 + row / ny) * G.res + q.y0 + row % ny) * G.res);
    mine += __ldg(&G.cell_start[base + (uint32_t)q.x1 + 1u]) - __ldg(&G.cell_start[base + (uint32_t)q.x0]);
  }
  uint32_t incl = mine;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= (unsigned)d) incl += o; }
  before_me = incl - mine;
  return __shfl_sync(0xffffffffu, incl, 31);
}
// the want-th smallest (1-based) of keys[0, n), 1 <= want <= n: quickselect in place down to one key per lane, then a rank by shuffles
DEV uint32_t sppm_select_smem(uint32_t* keys, uint32_t n, uint32_t want, unsigned lane) {
  while (n > 32u) {
    const uint32_t a = keys[0], b = keys[n >> 1], c = keys[n - 1u];
    const uint32_t pivot = max(min(a, b), min(max(a, b), c));  // median of three
    uint32_t below = 0, equal = 0;
    for (uint32_t i = lane; i < n; i += 32) { const uint32_t v = keys[i]; below += v < pivot ? 1u : 0u; equal += v == pivot ? 1u : 0u; }
    below = sppm_warp_sum(below); equal = sppm_warp_sum(equal);
    if (want > below && want <= below + equal) return pivot;
    const bool take_low = want <= below;
    if (!take_low) want -= below + equal;
    uint32_t out = 0;
    for (uint32_t i0 = 0; i0 < n; i0 += 32) {  // warp-uniform trip count; writes land at or below the positions already read
      const uint32_t i = i0 + lane;
      const uint32_t v = i < n ? keys[i] : 0u;
      const bool keep = i < n && (take_low ? v < pivot : v > pivot);
      const unsigned m = __ballot_sync(0xffffffffu, keep);
      __syncwarp();
      if (keep) keys[out + __popc(m & ((1u << lane) - 1u))] = v;
      out += __popc(m);
    }
    __syncwarp();
    n = out;
  }
  // at most 32 left: every lane ranks its key among them (ties broken by lane), the lane of rank want - 1 holds the answer
  const uint32_t v = lane < n ? keys[lane] : 0xffffffffu;
  uint32_t rank = 0;
#pragma unroll
  for (int o = 0; o < 32; ++o) {
    const uint32_t u = __shfl_sync(0xffffffffu, v, o);
    rank += (u < v || (u == v && (unsigned)o < lane)) ? 1u : 0u;
  }
  const unsigned holder = __ballot_sync(0xffffffffu, lane < n && rank == want - 1u);
  return __shfl_sync(0xffffffffu, v, __ffs((int)holder) - 1);
}
// the want-th smallest squared distance (as float bits) of the n_in photons of cube q; `before_me` from sppm_cube_count
DEV uint32_t sppm_select_cube(const SppmGrid& G, const SppmCube& q, float3 p, uint32_t* keys, uint32_t n_in, uint32_t before_me, uint32_t want, unsigned lane) {
  __syncwarp();
  if (n_in <= SPPM_KNN_CAP) {
    sppm_for_each_in_cube(G, q, lane, [&](uint32_t, float4 pos, uint32_t at) { keys[at] = __float_as_uint(sppm_d2(pos, p)); });
    __syncwarp();
    return sppm_select_smem(keys, n_in, want, lane);
  }
  // too many for the buffer: narrow the interval [lo, hi] of float bits that holds the want-th, 2048 bins at a time
  uint32_t lo = 0u, hi = 0x7f800000u;
  for (;;) {
    const unsigned long long span = (unsigned long long)(hi - lo) + 1ull;
    for (uint32_t i = lane; i < SPPM_KNN_CAP; i += 32) keys[i] = 0u;
    __syncwarp();
    sppm_for_each_in_cube(G, q, lane, [&](uint32_t, float4 pos, uint32_t) {
      const uint32_t key = __float_as_uint(sppm_d2(pos, p));
      if (key >= lo && key <= hi) atomicAdd(&keys[(uint32_t)(((unsigned long long)(key - lo) * SPPM_KNN_CAP) / span)], 1u);
    });
    __syncwarp();
    // the bin that holds the want-th of [lo, hi]: every lane owns 64 consecutive bins
    uint32_t mine = 0;
    for (uint32_t i = 0; i < SPPM_KNN_CAP / 32; ++i) mine += keys[lane * (SPPM_KNN_CAP / 32) + i];
    uint32_t incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= (unsigned)d) incl += o; }
    const unsigned owner_mask = __ballot_sync(0xffffffffu, incl >= want);
    const int owner = __ffs((int)owner_mask) - 1;  // first lane whose inclusive sum reaches want (exists: want <= photons in [lo, hi])
    uint32_t bin = 0, before = 0, in_bin = 0;
    if ((int)lane == owner) {
      uint32_t run = incl - mine;
      for (uint32_t i = 0; i < SPPM_KNN_CAP / 32; ++i) {
        const uint32_t c = keys[lane * (SPPM_KNN_CAP / 32) + i];
        if (run + c >= want) { bin = lane * (SPPM_KNN_CAP / 32) + i; before = run; in_bin = c; break; }
        run += c;
      }
    }
    bin = __shfl_sync(0xffffffffu, bin, owner); before = __shfl_sync(0xffffffffu, before, owner); in_bin = __shfl_sync(0xffffffffu, in_bin, owner);
    // keys of bin b: key - lo in [ceil(b span / 2048), ceil((b + 1) span / 2048) - 1]
    const uint32_t new_lo = lo + (uint32_t)(((unsigned long long)bin * span + SPPM_KNN_CAP - 1) / SPPM_KNN_CAP);
    const uint32_t new_hi = lo + (uint32_t)((((unsigned long long)bin + 1ull) * span + SPPM_KNN_CAP - 1) / SPPM_KNN_CAP) - 1u;
    lo = new_lo; hi = new_hi; want -= before;
    __syncwarp();
    if (lo == hi) return lo;
    if (in_bin <= SPPM_KNN_CAP) {  // collect the bin's keys and finish in shared memory
      uint32_t n = 0;  // warp-uniform running count: the lanes append through ballots, a cube row at a time would diverge, so use an atomic cursor
      __shared__ uint32_t s_cursor[4];
      if (lane == 0) s_cursor[threadIdx.x >> 5] = 0u;
      __syncwarp();
      sppm_for_each_in_cube(G, q, lane, [&](uint32_t, float4 pos, uint32_t) {
        const uint32_t key = __float_as_uint(sppm_d2(pos, p));
        if (key >= lo && key <= hi) keys[atomicAdd(&s_cursor[threadIdx.x >> 5], 1u)] = key;
      });
      __syncwarp();
      n = s_cursor[threadIdx.x >> 5];
      __syncwarp();
      return sppm_select_smem(keys, n, want, lane);
    }
  }
}
__global__ void __launch_bounds__(128) k_sppm_knn(SppmGrid Gg, SppmGrid Gc, SppmPixels px, SppmEye E, int k_global, int k_caustic) {
  __shared__ uint32_t s_keys[4][SPPM_KNN_CAP];
  const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const uint32_t n_list = *E.knn_n;
  uint32_t* keys = s_keys[w];
  for (uint32_t e = blockIdx.x * 4u + w; e < n_list; e += gridDim.x * 4u) {
    const uint32_t entry = E.knn_list[e], pixel = entry & 0x7fffffffu;
    const bool caustic = (entry >> 31) != 0u;
    const SppmGrid G = caustic ? Gc : Gg;
    const uint32_t k = (uint32_t)(caustic ? k_caustic : k_global);
    const float3 p = f3(__ldg(&E.vp[pixel])), bsdf = f3(__ldg(&E.vb[pixel]));
    float3 flux = f3(0, 0, 0);
    uint32_t tau = 0u;  // bits of the k-th smallest squared distance (non-negative floats order like their bits)
    if (G.n != 0) {
      const int3 c0 = sppm_cell_of(p, G.lo, G.inv_cell, G.res);
      SppmCube q;
      uint32_t n_in = 0, before_me = 0;
      int s = 0;
      bool whole = false;
      for (;; ++s) {  // the smallest cube with k photons
        q = sppm_cube(G, c0, s);
        whole = q.x0 == 0 && q.y0 == 0 && q.z0 == 0 && q.x1 == G.res - 1 && q.y1 == G.res - 1 && q.z1 == G.res - 1;
        n_in = sppm_cube_count(G, q, lane, before_me);
        if (n_in >= k || whole) break;
      }
      // the k-th smallest squared distance inside this cube: the true one if it lies within s cells, else an upper bound of it
      tau = sppm_select_cube(G, q, p, keys, n_in, before_me, n_in < k ? n_in : k, lane);
      for (;;) {
        const float reach = (float)s * G.cell;
        if (whole || tau <= __float_as_uint(reach * reach)) break;  // no photon outside the cube is nearer than s cells
        // the cube that reaches as far as the bound holds the k nearest of the map, all of them inside the bound's sphere:
        // only the cells that sphere touches are read, and only the distances up to the bound are kept
        const float bound = __uint_as_float(tau);
        const int s_need = (int)ceilf(sqrtf(bound) * G.inv_cell * 1.0001f);
        s = max(s + 1, min(s_need, G.res));
        q = sppm_cube(G, c0, s);
        whole = q.x0 == 0 && q.y0 == 0 && q.z0 == 0 && q.x1 == G.res - 1 && q.y1 == G.res - 1 && q.z1 == G.res - 1;
        __syncwarp();
        uint32_t n2 = 0;  // warp-uniform
        sppm_for_each_in_ball(G, q, p, bound, lane, [&](uint32_t, float4 pos, bool valid) {
          const uint32_t key = __float_as_uint(sppm_d2(pos, p));
          const bool keep = valid && key <= tau;
          const unsigned m = __ballot_sync(0xffffffffu, keep);
          const uint32_t at = n2 + __popc(m & ((1u << lane) - 1u));
          if (keep && at < SPPM_KNN_CAP) keys[at] = key;
          n2 += __popc(m);
        });
        __syncwarp();
        if (n2 <= SPPM_KNN_CAP) tau = sppm_select_smem(keys, n2, k, lane);  // n2 >= k: the k of the smaller cube are among them
        else { n_in = sppm_cube_count(G, q, lane, before_me); tau = sppm_select_cube(G, q, p, keys, n_in, before_me, k, lane); }
      }
      sppm_for_each_in_ball(G, q, p, __uint_as_float(tau), lane, [&](uint32_t j, float4 pos, bool valid) {
        if (valid && __float_as_uint(sppm_d2(pos, p)) <= tau) flux = flux + sppm_photon_flux(G, j, pos, p, bsdf);
      });
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) { flux.x += __shfl_xor_sync(0xffffffffu, flux.x, o); flux.y += __shfl_xor_sync(0xffffffffu, flux.y, o); flux.z += __shfl_xor_sync(0xffffffffu, flux.z, o); }
      __syncwarp();
    }
    if (lane == 0) {  // SPPM::update, first time (photon_mapper.rs:50-54)
      const float4 st = make_float4(flux.x, flux.y, flux.z, __uint_as_float(tau));
      if (caustic) { px.c[pixel] = st; reinterpret_cast<uint32_t*>(&px.n[pixel])[1] = k; }
      else { px.g[pixel] = st; reinterpret_cast<uint32_t*>(&px.n[pixel])[0] = k; }
    }
  }
}

// impl Integrator for SPPMIntegrator (photon_mapper.rs:327-365) under Camera::capture_image
DEV float3 sppm_adjust(float4 st, float n_emitted) {  // adjust_flux (photon_mapper.rs:117-119): 0 / 0 stays NaN like upstream
  const float d = PI_F * st.w * n_emitted;
  return f3(st.x / d, st.y / d, st.z / d);
}
template <bool OBJF>
__global__ void __launch_bounds__(128) k_sppm_final(DevScene S, RenderConst P, PathMap M, SppmPixels px, float n_emitted, float* __restrict__ accum, SppmCounters* cnt) {
  unsigned long long n_rays = 0;
  for (unsigned long long id = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; id < M.total; id += (unsigned long long)gridDim.x * blockDim.x) {
    uint32_t pixel, sample;
    path_to_pixel(M, P, id, pixel, sample);
    Ray ray = camera_ray(P, pixel, sample);
    float3 throughput = f3(1, 1, 1), radiance = f3(0, 0, 0);
    int depth = P.max_depth;
    for (int bounce = 1;; ++bounce) {
      HitRaw hr;
      trace<false, OBJF>(S, ray, P.t_min, hr);
      n_rays++;
      if (hr.prim < 0 || depth <= 0) break;
      depth--;
      const int4 meta = __ldg(&S.prim_meta[hr.prim]);
      const float4 mat = __ldg(&S.materials[meta.x >> 8]);
      HitInfo h;
      finish_hit(S, ray, hr, h, texture_needs_uv(S, __float_as_int(mat.y)));
      if (__float_as_int(mat.x) == B200RT_MAT_DIFFUSE_LIGHT) radiance = radiance + throughput * texture_color(S, __float_as_int(mat.y), h);
      const uint4 r = philox4x32(make_uint4(pixel, sample, (uint32_t)bounce, 0x70u), P.key);
      float3 dir, f;
      const int inter = sppm_scatter(S, mat, h, ray, u01(r.x), u01(r.y), u01(r.z), dir, f);
      if (inter == SPPM_DIFFUSE) {
        radiance = radiance + throughput * sppm_adjust(px.c[pixel], n_emitted) + throughput * sppm_adjust(px.g[pixel], n_emitted);
        break;
      }
      if (inter == SPPM_ABSORB) break;
      throughput = throughput * f;
      ray.o = h.p; ray.d = dir; ray.tmax = __int_as_float(0x7f800000); ray.mseed = r.w;
    }
    // NaN estimates (a pixel whose map was empty) poison the pixel upstream and tonemap to 0; here they are dropped by accum_add
    accum_add(accum, pixel, radiance);
  }
  for (int o = 16; o > 0; o >>= 1) n_rays += __shfl_down_sync(0xffffffffu, n_rays, o);
  if ((threadIdx.x & 31) == 0 && n_rays) atomicAdd(&cnt->eye_rays, n_rays);
}
