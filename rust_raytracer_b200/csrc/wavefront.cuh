// wavefront.cuh — the wavefront pipeline over SoA float4 ray queues (north_star): generate -> extend ->
// per-material shade -> light-sample/shadow, with queue compaction by warp-aggregated atomics and in-place
// regeneration of terminated paths so the pool stays full.
//
// Per path slot (HBM, SoA, one float4/uint4 per array => fully coalesced 16-byte accesses):
//   ray_o   (o.xyz, time)            ray_d  (d.xyz, t_max)
//   hit     (t as 2 x u32 [f64], device prim, aux)
//   st_a    (throughput.rgb, prev_pdf_b)
//   st_b    (pixel, sample, bounce, medium seed)
// Queues hold slot indices (u32).  Shadow requests are stored by queue position: sh_o, sh_d, sh_c(contrib.rgb, pixel).
#pragma once
#include "kernels.cuh"

#define WF_NCLASS 6  // 0 miss, 1 Lambertian, 2 Metal, 3 Dielectric, 4 DiffuseLight, 5 Isotropic
#define WF_MISS 0

struct WfCounters {
  unsigned long long next_path;     // next path id to generate
  unsigned long long extend_rays, shadow_rays, paths_done;
  unsigned int n_ext;               // size of the current extend queue
  unsigned int n_ext_next;          // size of the queue being filled by shade/generate
  unsigned int n_free;              // slots released by shade this iteration
  unsigned int n_shadow;            // shadow requests this iteration
  unsigned int n_class[WF_NCLASS];  // per-material shade queues
  unsigned int ext_cursor;          // work-fetch cursor of the persistent extend kernel
  unsigned int shadow_cursor;
  unsigned int n_new;               // paths to generate this iteration
  unsigned int pad;
  unsigned long long gen_base;      // first path id of this iteration's generation
};

struct WfBuffers {
  float4* ray_o; float4* ray_d; uint4* hit; float4* st_a; uint4* st_b;
  unsigned int* q_ext; unsigned int* q_ext_next; unsigned int* q_free; unsigned int* q_class;  // q_class: WF_NCLASS x pool
  float4* sh_o; float4* sh_d; float4* sh_c; unsigned int* sh_seed;
  WfCounters* cnt;
  unsigned int pool;
};

// Append one element per participating lane with a single atomic per warp (warp-aggregated compaction).
DEV unsigned int warp_append(unsigned int* counter, bool pred) {
  const unsigned mask = __ballot_sync(__activemask(), pred);
  if (!pred) return 0xffffffffu;
  const unsigned lane = threadIdx.x & 31;
  const int leader = __ffs(mask) - 1;
  unsigned int base = 0;
  if ((int)lane == leader) base = atomicAdd(counter, (unsigned)__popc(mask));
  base = __shfl_sync(mask, base, leader);
  return base + __popc(mask & ((1u << lane) - 1u));
}

// Between stages: one thread turns the counters over (no host round trip inside an iteration).
// phase 0: before generate  (decide how many paths to start, from the free list or the whole pool on the first pass)
__global__ void k_wf_control_generate(WfBuffers B, PathMap M, int first_pass) {
  WfCounters* c = B.cnt;
  unsigned long long remaining = M.total - c->next_path;
  unsigned int avail = first_pass ? B.pool : c->n_free;
  unsigned int n_new = (unsigned int)(remaining < (unsigned long long)avail ? remaining : (unsigned long long)avail);
  c->n_new = n_new;
  c->gen_base = c->next_path;
  c->next_path += n_new;
}
// phase 1: before extend (the queue filled by shade + generate becomes the current one; class queues reset)
__global__ void k_wf_control_extend(WfBuffers B) {
  WfCounters* c = B.cnt;
  c->n_ext = c->n_ext_next;
  c->n_ext_next = 0;
  c->n_free = 0;
  c->n_shadow = 0;
  c->ext_cursor = 0;
  c->shadow_cursor = 0;
  c->extend_rays += c->n_ext;
  for (int i = 0; i < WF_NCLASS; ++i) c->n_class[i] = 0;
}
__global__ void k_wf_control_shadow(WfBuffers B) { B.cnt->shadow_rays += B.cnt->n_shadow; }

// generate: Camera::get_ray for new paths (camera.rs:57-64, 97-99) into free slots
__global__ void __launch_bounds__(256) k_wf_generate(WfBuffers B, RenderConst P, PathMap M, int first_pass) {
  const unsigned int n_new = B.cnt->n_new;
  const unsigned long long base = B.cnt->gen_base;
  for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_new; i += gridDim.x * blockDim.x) {
    const unsigned int slot = first_pass ? i : B.q_free[i];
    uint32_t pixel, sample;
    const bool valid = true;
    path_to_pixel(M, P, base + i, pixel, sample);
    {
      Ray r = camera_ray(P, pixel, sample);
      B.ray_o[slot] = make_float4(r.o.x, r.o.y, r.o.z, r.time);
      B.ray_d[slot] = make_float4(r.d.x, r.d.y, r.d.z, r.tmax);
      B.st_a[slot] = make_float4(1.f, 1.f, 1.f, -1.f);
      B.st_b[slot] = make_uint4(pixel, sample, 0u, r.mseed);
    }
    unsigned int pos = warp_append(&B.cnt->n_ext_next, valid);
    if (valid) B.q_ext_next[pos] = slot;
  }
}

DEV int class_of_material(const DevScene& S, int device_prim) {
  const int4 meta = __ldg(&S.prim_meta[device_prim]);
  return 1 + __float_as_int(__ldg(&S.materials[meta.x >> 8]).x);
}

// extend: persistent warps fetch 32 rays at a time from the queue (warp-level work fetch), traverse, store the
// hit and bin the slot by material class for the shade kernels.
#ifndef B200RT_EXTEND_MINBLOCKS
#define B200RT_EXTEND_MINBLOCKS 6
#endif
__global__ void __launch_bounds__(128, B200RT_EXTEND_MINBLOCKS) k_wf_extend(DevScene S, WfBuffers B, float t_min) {
  const unsigned int n = B.cnt->n_ext;
  const unsigned lane = threadIdx.x & 31;
  for (;;) {
    unsigned int base = 0;
    if (lane == 0) base = atomicAdd(&B.cnt->ext_cursor, 32u);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (base >= n) break;
    const unsigned int i = base + lane;
    const bool active = i < n;
    int cls = WF_MISS;
    unsigned int slot = 0;
    if (active) {
      slot = B.q_ext[i];
      const float4 o = B.ray_o[slot], d = B.ray_d[slot];
      Ray r;
      r.o = f3(o); r.time = o.w; r.d = f3(d); r.tmax = d.w; r.mseed = B.st_b[slot].w;
      HitRaw h;
      trace<false>(S, r, t_min, h);
      const unsigned long long tb = (unsigned long long)__double_as_longlong(h.t);
      B.hit[slot] = make_uint4((unsigned int)tb, (unsigned int)(tb >> 32), (unsigned int)h.prim, (unsigned int)h.aux);
      cls = h.prim < 0 ? WF_MISS : class_of_material(S, h.prim);
    }
    // bin by material class: the leader lane of every class present in the warp issues its atomic in the same
    // instruction, so the warp pays one atomic round trip instead of one per class
    const unsigned peers = __match_any_sync(0xffffffffu, active ? (unsigned)cls : (unsigned)WF_NCLASS);
    const int leader = __ffs((int)peers) - 1;
    unsigned int cbase = 0;
    if (active && (int)lane == leader) cbase = atomicAdd(&B.cnt->n_class[cls], (unsigned)__popc(peers));
    cbase = __shfl_sync(0xffffffffu, cbase, leader);
    if (active) B.q_class[(size_t)cls * B.pool + cbase + __popc(peers & ((1u << lane) - 1u))] = slot;
  }
}

// shade: ONE launch over all per-material queues.  Each queue is padded to a multiple of 32 in the index space, so
// every warp works on a single material class (coherent branches) and the six classes run concurrently.
__global__ void __launch_bounds__(128) k_wf_shade(DevScene S, WfBuffers B, RenderConst P, float* __restrict__ accum, int only_cls) {
  unsigned int off[WF_NCLASS + 1];
  off[0] = 0;
#pragma unroll
  for (int c = 0; c < WF_NCLASS; ++c) off[c + 1] = off[c] + ((only_cls < 0 || only_cls == c) ? ((B.cnt->n_class[c] + 31u) & ~31u) : 0u);
  for (unsigned int gi = blockIdx.x * blockDim.x + threadIdx.x; gi < off[WF_NCLASS]; gi += gridDim.x * blockDim.x) {
    int cls = 0;
#pragma unroll
    for (int c = 1; c < WF_NCLASS; ++c) cls += (gi >= off[c]) ? 1 : 0;
    const unsigned int i = gi - off[cls];
    const unsigned int* q = B.q_class + (size_t)cls * B.pool;
    const bool active = i < B.cnt->n_class[cls];
    bool cont = false, shadow = false;
    unsigned int slot = 0;
    Ray ray;
    PathState ps;
    ShadowReq sr;
    sr.valid = false;
    if (active) {
      slot = q[i];
      const float4 o = B.ray_o[slot], d = B.ray_d[slot], sa = B.st_a[slot];
      const uint4 sb = B.st_b[slot], hh = B.hit[slot];
      ray.o = f3(o); ray.time = o.w; ray.d = f3(d); ray.tmax = d.w; ray.mseed = sb.w;
      ps.throughput = f3(sa); ps.prev_pdf_b = sa.w; ps.radiance = f3(0, 0, 0);
      ps.pixel = sb.x; ps.sample = sb.y; ps.bounce = (int)sb.z;
      if (cls == WF_MISS) {
        ps.radiance = ps.throughput * background_color(P, ray.d);
      } else if (ps.bounce < P.max_depth) {  // photon_mapper.rs:336-339: the depth budget is checked after the hit
        HitRaw h;
        h.t = __longlong_as_double((long long)(((unsigned long long)hh.y << 32) | hh.x)); h.prim = (int)hh.z; h.aux = (int)hh.w;
        cont = shade_hit(S, P, ray, h, ps, sr);
        shadow = sr.valid;
      }
      accum_add(accum, ps.pixel, ps.radiance);
      if (cont) {
        B.ray_o[slot] = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.time);
        B.ray_d[slot] = make_float4(ray.d.x, ray.d.y, ray.d.z, __int_as_float(0x7f800000));
        B.st_a[slot] = make_float4(ps.throughput.x, ps.throughput.y, ps.throughput.z, ps.prev_pdf_b);
        B.st_b[slot] = make_uint4(ps.pixel, ps.sample, (unsigned int)ps.bounce, ray.mseed);
      }
    }
    unsigned int pe = warp_append(&B.cnt->n_ext_next, active && cont);
    if (active && cont) B.q_ext_next[pe] = slot;
    unsigned int pf = warp_append(&B.cnt->n_free, active && !cont);
    if (active && !cont) B.q_free[pf] = slot;
    unsigned int pq = warp_append(&B.cnt->n_shadow, shadow);
    if (shadow) {
      B.sh_o[pq] = make_float4(sr.ray.o.x, sr.ray.o.y, sr.ray.o.z, sr.ray.time);
      B.sh_d[pq] = make_float4(sr.ray.d.x, sr.ray.d.y, sr.ray.d.z, sr.ray.tmax);
      B.sh_c[pq] = make_float4(sr.contrib.x, sr.contrib.y, sr.contrib.z, __uint_as_float(ps.pixel));
      B.sh_seed[pq] = sr.ray.mseed;
    }
  }
}

// shadow: any-hit traversal of the light-sample rays; unoccluded samples add their contribution
__global__ void __launch_bounds__(128) k_wf_shadow(DevScene S, WfBuffers B, float t_min, float* __restrict__ accum) {
  const unsigned int n = B.cnt->n_shadow;
  const unsigned lane = threadIdx.x & 31;
  for (;;) {
    unsigned int base = 0;
    if (lane == 0) base = atomicAdd(&B.cnt->shadow_cursor, 32u);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (base >= n) break;
    const unsigned int i = base + lane;
    if (i < n) {
      const float4 o = B.sh_o[i], d = B.sh_d[i], c = B.sh_c[i];
      Ray r;
      r.o = f3(o); r.time = o.w; r.d = f3(d); r.tmax = d.w; r.mseed = B.sh_seed[i];
      HitRaw h;
      trace<true>(S, r, t_min, h);
      if (h.prim < 0) accum_add(accum, __float_as_uint(c.w), f3(c));
    }
  }
}
