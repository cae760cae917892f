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
struct SppmGrid {  // one photon map
  const float4* photons;  // 4 x float4 per photon: (pos, 0) (power, 0) (direction, 0) (normal, 0)
  const uint32_t* cell_start;  // res^3 + 1
  const uint32_t* index;       // photon indices sorted by cell
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

DEV float3 sppm_light_flux(const DevScene& S, int li) {  // the DiffuseLight colour (light.rs:77, 143)
  const float4 L0 = __ldg(&S.lights[3 * li]), L1 = __ldg(&S.lights[3 * li + 1]);
  const float4 lmat = __ldg(&S.materials[__float_as_int(L0.w)]);
  HitInfo lh;
  lh.p = f3(L1); lh.n = f3(0, 1, 0); lh.u = 0.f; lh.v = 0.f; lh.t = 0.f; lh.front_face = 1;
  return texture_color(S, __float_as_int(lmat.y), lh);
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
__global__ void __launch_bounds__(128) k_sppm_photons(DevScene S, SppmConst sp, uint2 key, int iter, float4* pg, float4* pc, unsigned int cap, SppmCounters* cnt,
                                                      unsigned int id_first, unsigned int id_count) {
  unsigned long long n_rays = 0;
  unsigned int lo[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, hi[3] = {0u, 0u, 0u};
  // Per-lane work fetch: a photon walk lasts 1 to dozens of bounces (2.9 on average in the Cornell box), so a warp that
  // waits for its longest photon runs at 5.6 of 32 lanes (ncu, profiles/r1_sppm_full_summary.txt).  Lanes whose photon
  // has died take the next photon ids from a global cursor (one warp-aggregated atomic) as soon as SPPM_REFILL_LANES of
  // them are idle.  Every draw is keyed (photon id, iteration, bounce), so which lane walks which photon changes nothing
  // but the order of the records in the map.  (The path tracer's extend kernel does NOT do this: there the queue order
  // is what keeps a warp coherent; here there is no order to lose.)
  const unsigned lane = threadIdx.x & 31;
  const unsigned int n_photons = id_count;  // the cursor counts inside the shard
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
        if (have) {
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
// The photon records themselves are moved into cell order (`sorted`), so that a gather over a cell reads one contiguous
// run of 64-byte records instead of chasing indices through the emission-ordered array; index[] becomes the identity
// over the sorted array (the consumers keep their cell_start -> index -> photon walk).
__global__ void k_sppm_grid_fill(SppmSegs photons, uint32_t n, float3 lo, float inv_cell, int res, uint32_t* cursor, uint32_t* index,
                                 float4* __restrict__ sorted) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float4* src = sppm_seg_photon(photons, i);
    const float4 a = src[0], b = src[1], c = src[2], d = src[3];
    const uint32_t pos = atomicAdd(&cursor[sppm_cell_id(sppm_cell_of(f3(a), lo, inv_cell, res), res)], 1u);
    float4* dst = sorted + 4 * (size_t)pos;
    dst[0] = a; dst[1] = b; dst[2] = c; dst[3] = d;
    index[pos] = pos;
  }
}

// ---- SPPM::update (photon_mapper.rs:42-64) over the grid ---------------------------------------------------------
// Vec3::elemul(rec.mat.bsdf(photon.direction, rec), photon.power) * (1 - disk_factor) (photon_mapper.rs:72-74, 90-92)
DEV float3 sppm_photon_flux(const SppmGrid& G, uint32_t i, float3 p, float3 bsdf) {
  const float4* ph = G.photons + 4 * (size_t)i;
  const float3 pos = f3(__ldg(ph)), power = f3(__ldg(ph + 1)), nrm = f3(__ldg(ph + 3));
  const float3 d = pos - p;
  const float disk = fabsf(dot(nrm, d * rsqrtf(dot(d, d))));
  return bsdf * power * (1.0f - disk);
}
DEV void sppm_heap_push(float* hd, uint32_t* hi, int& n, int k, float d, uint32_t idx) {
  if (n < k) {  // sift up
    int c = n++;
    while (c > 0) {
      const int par = (c - 1) >> 1;
      if (hd[par] >= d) break;
      hd[c] = hd[par]; hi[c] = hi[par]; c = par;
    }
    hd[c] = d; hi[c] = idx;
  } else if (d < hd[0]) {  // replace the farthest, sift down
    int c = 0;
    for (;;) {
      int l = 2 * c + 1, r = l + 1, m = l;
      if (l >= n) break;
      if (r < n && hd[r] > hd[l]) m = r;
      if (hd[m] <= d) break;
      hd[c] = hd[m]; hi[c] = hi[m]; c = m;
    }
    hd[c] = d; hi[c] = idx;
  }
}
DEV void sppm_visit_cell(const SppmGrid& G, int x, int y, int z, float3 p, float* hd, uint32_t* hi, int& n, int k) {
  const uint32_t c = sppm_cell_id(make_int3(x, y, z), G.res);
  for (uint32_t j = __ldg(&G.cell_start[c]), e = __ldg(&G.cell_start[c + 1]); j < e; ++j) {
    const uint32_t i = __ldg(&G.index[j]);
    const float3 d = f3(__ldg(G.photons + 4 * (size_t)i)) - p;
    sppm_heap_push(hd, hi, n, k, dot(d, d), i);
  }
}
// PhotonMap::estimate_flux_by_count (photon_mapper.rs:77-96): the k nearest photons, exactly
DEV void sppm_estimate_by_count(const SppmGrid& G, float3 p, float3 bsdf, int k, float3& flux, float& radius2) {
  float hd[SPPM_MAX_K];
  uint32_t hi[SPPM_MAX_K];
  int n = 0;
  flux = f3(0, 0, 0); radius2 = 0.f;
  if (G.n == 0) return;
  const int3 c0 = sppm_cell_of(p, G.lo, G.inv_cell, G.res);
  for (int s = 0; s < G.res; ++s) {  // shell s: cells at Chebyshev distance s from c0
    const int z0 = max(0, c0.z - s), z1 = min(G.res - 1, c0.z + s), y0 = max(0, c0.y - s), y1 = min(G.res - 1, c0.y + s);
    for (int z = z0; z <= z1; ++z)
      for (int y = y0; y <= y1; ++y) {
        if (abs(z - c0.z) == s || abs(y - c0.y) == s) {
          for (int x = max(0, c0.x - s); x <= min(G.res - 1, c0.x + s); ++x) sppm_visit_cell(G, x, y, z, p, hd, hi, n, k);
        } else {
          if (c0.x - s >= 0) sppm_visit_cell(G, c0.x - s, y, z, p, hd, hi, n, k);
          if (s > 0 && c0.x + s < G.res) sppm_visit_cell(G, c0.x + s, y, z, p, hd, hi, n, k);
        }
      }
    // every photon not visited yet lies at least s cells away along some axis
    const float reach = (float)s * G.cell;
    if (n == k && hd[0] <= reach * reach) break;
  }
  for (int j = 0; j < n; ++j) { radius2 = fmaxf(radius2, hd[j]); flux = flux + sppm_photon_flux(G, hi[j], p, bsdf); }
}
// PhotonMap::estimate_flux_within_radius (photon_mapper.rs:99-114): squared distance < radius^2
DEV void sppm_estimate_within(const SppmGrid& G, float3 p, float3 bsdf, float radius2, float3& flux, uint32_t& count) {
  flux = f3(0, 0, 0); count = 0;
  if (G.n == 0) return;
  const float r = sqrtf(radius2);
  const int3 a = sppm_cell_of(p - f3(r, r, r), G.lo, G.inv_cell, G.res), b = sppm_cell_of(p + f3(r, r, r), G.lo, G.inv_cell, G.res);
  for (int z = a.z; z <= b.z; ++z)
    for (int y = a.y; y <= b.y; ++y)
      for (int x = a.x; x <= b.x; ++x) {
        const uint32_t c = sppm_cell_id(make_int3(x, y, z), G.res);
        for (uint32_t j = __ldg(&G.cell_start[c]), e = __ldg(&G.cell_start[c + 1]); j < e; ++j) {
          const uint32_t i = __ldg(&G.index[j]);
          const float3 d = f3(__ldg(G.photons + 4 * (size_t)i)) - p;
          if (dot(d, d) < r * r) { ++count; flux = flux + sppm_photon_flux(G, i, p, bsdf); }
        }
      }
}
DEV void sppm_update(const SppmGrid& G, float3 p, float3 bsdf, int k_init, double alpha, float4& st, uint32_t& photons) {
  if (photons == 0) {
    photons = (uint32_t)k_init;
    float3 flux; float r2;
    sppm_estimate_by_count(G, p, bsdf, k_init, flux, r2);
    st = make_float4(flux.x, flux.y, flux.z, r2);
  } else {
    float3 flux; uint32_t m;
    sppm_estimate_within(G, p, bsdf, st.w, flux, m);
    const uint32_t prev = photons;
    photons += (uint32_t)(alpha * (double)m);
    const float frac = (float)((double)photons / (double)(prev + m));
    st = make_float4((st.x + flux.x) * frac, (st.y + flux.y) * frac, (st.z + flux.z) * frac, st.w * frac);
  }
}

// SPPMIntegrator::update_sppm (photon_mapper.rs:282-305): one jittered eye ray per pixel and iteration
template <bool OBJF>
// [pix_first, pix_first + pix_count): the pixels whose statistics this device owns
__global__ void __launch_bounds__(128) k_sppm_eye(DevScene S, RenderConst P, SppmConst sp, int iter, SppmGrid Gg, SppmGrid Gc, SppmPixels px, SppmCounters* cnt,
                                                  uint32_t pix_first, uint32_t pix_count) {
  const uint32_t n_pix = pix_first + pix_count;
  unsigned long long n_rays = 0;
  for (uint32_t pixel = pix_first + blockIdx.x * blockDim.x + threadIdx.x; pixel < n_pix; pixel += gridDim.x * blockDim.x) {
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
      if (inter == SPPM_DIFFUSE) {
        const float3 bsdf = sppm_bsdf(S, mat, h);
        float4 c = px.c[pixel], g = px.g[pixel];
        uint2 n = px.n[pixel];
        sppm_update(Gc, h.p, bsdf, sp.k_caustic, sp.alpha, c, n.y);  // caustic first, like upstream
        sppm_update(Gg, h.p, bsdf, sp.k_global, sp.alpha, g, n.x);
        px.c[pixel] = c; px.g[pixel] = g; px.n[pixel] = n;
        break;
      }
      if (inter == SPPM_ABSORB) break;
      ray.o = h.p; ray.d = dir; ray.tmax = __int_as_float(0x7f800000); ray.mseed = r.w;
    }
  }
  for (int o = 16; o > 0; o >>= 1) n_rays += __shfl_down_sync(0xffffffffu, n_rays, o);
  if ((threadIdx.x & 31) == 0 && n_rays) atomicAdd(&cnt->eye_rays, n_rays);
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
