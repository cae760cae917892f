// wavefront.cuh — the wavefront pipeline over SoA float4 ray queues (north_star): generate -> extend ->
// per-material shade -> light-sample/shadow, with stream compaction by warp-aggregated atomics.
//
// The queues hold the ray PAYLOAD, not indices, as one 96-byte record (three 32-byte DRAM sectors) per position:
//   sector 0: (o.xyz, time) (d.xyz, medium seed bits; t_max of an extend ray is always +inf)   read by extend + shade
//   sector 1: (throughput.rgb, prev_pdf_b) (pixel, sample, bounce, 0)                           read by shade
//   sector 2: (t as 2 x u32 [f64], device prim, aux) (spare)                                    written by extend
// Shade gathers whole records (every sector it touches is fully used: the SoA version of this layout read 154 B
// per ray from DRAM for 84 B of payload), compacts the surviving paths straight into the NEXT set (one
// warp-aggregated atomic per warp, then 64 contiguous bytes per lane); generate appends freshly started paths
// behind them, so extend and generate touch memory strictly in order.  The two sets are swapped by the host
// between iterations.  Shadow requests are stored by queue position: sh_o, sh_d, sh_c, sh_seed.
#pragma once
#include "kernels.cuh"

// Shade work is binned by WHERE the ray ended: bin 0 = miss, bins 1.. = the hit primitive's index in BVH leaf order
// (spatially coherent, and one primitive has one material) scaled to WF_NBINS - 1 bins.  Shade walks the bins in
// order, so (a) its warps are material-coherent and (b) the survivors land in the next queue grouped by origin
// region and by expected traversal cost (fog-born rays together, ground-box rays together, ...), which is what
// keeps the extend warps converged.  Bin storage is WF_NBINS x pool indices: HBM is plentiful (180 GB).
#define WF_REC 6  // float4 per queue record (96 bytes)
#ifndef WF_NBINS
#define WF_NBINS 256
#endif

struct WfCounters {
  unsigned long long next_path;     // next path id to generate
  unsigned long long extend_rays, shadow_rays, paths_done;
  unsigned int n_ext;               // rays in the current set (extend + shade work list)
  unsigned int n_ext_next;          // survivors appended to the next set by shade
  unsigned int n_shadow;            // shadow requests this iteration
  unsigned int ext_cursor;          // work-fetch cursor of the persistent extend kernel
  unsigned int shadow_cursor;
  unsigned int n_work;              // shade index space: sum of the bin sizes, each padded to a multiple of 32
  unsigned int pad[2];
};

struct WfBuffers {
  float4* rec;    // current set: WF_REC float4 per queue position
  float4* rec_n;  // next set
  unsigned int* q_bin;    // WF_NBINS x pool queue positions
  unsigned int* bin_count;  // WF_NBINS
  unsigned int* bin_off;    // WF_NBINS + 1: start of every bin in the padded shade index space
  unsigned int bin_mul;     // bin = 1 + umulhi(prim, bin_mul)
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

// How many paths this iteration starts: the room left behind the survivors, capped by the paths that remain.
// generate and the control kernel evaluate the same expression on the same (not yet updated) counters.
DEV unsigned int wf_new_paths(const WfCounters* c, const PathMap& M, unsigned int pool) {
  const unsigned long long remaining = M.total - c->next_path;
  const unsigned int room = pool - c->n_ext_next;
  return (unsigned int)(remaining < (unsigned long long)room ? remaining : (unsigned long long)room);
}

// generate: Camera::get_ray for new paths (camera.rs:57-64, 97-99), appended behind the survivors of the last
// shade pass.  Path ids are consecutive, so a warp's primary rays cover an 8x4 pixel tile (coherent traversal).
__global__ void __launch_bounds__(256) k_wf_generate(WfBuffers B, RenderConst P, PathMap M) {
  const unsigned int n_new = wf_new_paths(B.cnt, M, B.pool);
  const unsigned long long base = B.cnt->next_path;
  const unsigned int off = B.cnt->n_ext_next;
  for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_new; i += gridDim.x * blockDim.x) {
    uint32_t pixel, sample;
    path_to_pixel(M, P, base + i, pixel, sample);
    const Ray r = camera_ray(P, pixel, sample);
    float4* rec = B.rec + (size_t)WF_REC * (off + i);
    rec[0] = make_float4(r.o.x, r.o.y, r.o.z, r.time);
    rec[1] = make_float4(r.d.x, r.d.y, r.d.z, __uint_as_float(r.mseed));
    rec[2] = make_float4(1.f, 1.f, 1.f, -1.f);
    rec[3] = make_float4(__uint_as_float(pixel), __uint_as_float(sample), 0.f, 0.f);
  }
}

// One thread turns the counters over between generate and extend (no host round trip inside an iteration).
__global__ void k_wf_control(WfBuffers B, PathMap M) {
  WfCounters* c = B.cnt;
  const unsigned int n_new = wf_new_paths(c, M, B.pool);
  c->next_path += n_new;
  c->n_ext = c->n_ext_next + n_new;
  c->n_ext_next = 0;
  c->n_shadow = 0;
  c->ext_cursor = 0;
  c->shadow_cursor = 0;
  c->extend_rays += c->n_ext;
  for (int i = 0; i < WF_NBINS; ++i) B.bin_count[i] = 0;
}
// after extend: the padded prefix sum of the bin sizes (shade's index space)
__global__ void k_wf_control_shade(WfBuffers B) {
  unsigned int off = 0;
  for (int i = 0; i < WF_NBINS; ++i) { B.bin_off[i] = off; off += (B.bin_count[i] + 31u) & ~31u; }
  B.bin_off[WF_NBINS] = off;
  B.cnt->n_work = off;
}
__global__ void k_wf_control_shadow(WfBuffers B) { B.cnt->shadow_rays += B.cnt->n_shadow; }

// extend: persistent warps fetch 32 consecutive rays at a time (warp-level work fetch: coalesced 512-byte
// reads of ray_o / ray_d), traverse, store the hit and bin the queue position for shade.
// (Measured and rejected: per-lane ray replacement a la Aila & Laine's dynamic fetch.  Refilling idle lanes from
// the queue mixes rays of different bins in one warp and loses the converged out-of-BVH tail: 60 ms -> 74..99 ms
// of extend per step on book-2 for refill thresholds 16..28, profiles/r1_experiments.md.)
#ifndef B200RT_EXTEND_MINBLOCKS
#define B200RT_EXTEND_MINBLOCKS 8
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
    unsigned int bin = 0;
    if (active) {
      float4* rec = B.rec + (size_t)WF_REC * i;
      const float4 o = rec[0], d = rec[1];
      Ray r;
      r.o = f3(o); r.time = o.w; r.d = f3(d); r.tmax = __int_as_float(0x7f800000); r.mseed = __float_as_uint(d.w);
      HitRaw h;
      trace<false>(S, r, t_min, h);
      const unsigned long long tb = (unsigned long long)__double_as_longlong(h.t);
      rec[4] = make_float4(__uint_as_float((unsigned int)tb), __uint_as_float((unsigned int)(tb >> 32)), __int_as_float(h.prim), __int_as_float(h.aux));
      rec[5] = make_float4(0.f, 0.f, 0.f, 0.f);  // the whole 32-byte sector is written: no read-modify-write
      bin = h.prim < 0 ? 0u : 1u + __umulhi((unsigned int)h.prim, B.bin_mul);
    }
    // the leader lane of every bin present in the warp issues its atomic in the same instruction: one atomic
    // round trip per warp, whatever the number of bins
    const unsigned peers = __match_any_sync(0xffffffffu, active ? bin : (unsigned)WF_NBINS);
    const int leader = __ffs((int)peers) - 1;
    unsigned int cbase = 0;
    if (active && (int)lane == leader) cbase = atomicAdd(&B.bin_count[bin], (unsigned)__popc(peers));
    cbase = __shfl_sync(0xffffffffu, cbase, leader);
    if (active) B.q_bin[(size_t)bin * B.pool + cbase + __popc(peers & ((1u << lane) - 1u))] = i;
  }
}

// shade: ONE launch over all bins.  Each bin is padded to a multiple of 32 in the index space, so every warp
// works on a single bin (one or a few neighbouring primitives: coherent material branches).  Survivors are
// written to the next set at their compacted position.
#ifndef B200RT_SHADE_MINBLOCKS
#define B200RT_SHADE_MINBLOCKS 4
#endif
template <int INTEG>
__global__ void __launch_bounds__(128, B200RT_SHADE_MINBLOCKS) k_wf_shade(DevScene S, WfBuffers B, RenderConst P, float* __restrict__ accum) {
  __shared__ unsigned int s_off[WF_NBINS + 1];
  __shared__ unsigned int s_cnt[WF_NBINS];
  for (int k = threadIdx.x; k <= WF_NBINS; k += blockDim.x) { s_off[k] = B.bin_off[k]; if (k < WF_NBINS) s_cnt[k] = B.bin_count[k]; }
  __syncthreads();
  const unsigned int n_work = s_off[WF_NBINS];
  const unsigned int stride = gridDim.x * blockDim.x;
  // queue position of work item g (bins are 32-aligned in the index space, so the lookup is warp-uniform)
  auto fetch_src = [&](unsigned int g, bool& act, unsigned int& src) {
    act = false; src = 0u;
    if (g >= n_work) return;
    int lo = 0, hi = WF_NBINS;  // the bin with s_off[bin] <= g < s_off[bin + 1]
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (g >= s_off[mid]) lo = mid; else hi = mid; }
    const unsigned int i = g - s_off[lo];
    act = i < s_cnt[lo];
    if (act) src = B.q_bin[(size_t)lo * B.pool + i];
  };
  // Software pipeline over the grid-stride loop: the gathers are latency-bound (two dependent levels, random
  // 16-byte reads), so the payload of the NEXT work item and the queue index of the one after it are in flight
  // while the current item is shaded.
  const unsigned int g0 = blockIdx.x * blockDim.x + threadIdx.x;
  bool a1, a2;
  unsigned int s1, s2;
  float4 n_o, n_d, n_sa, n_sb, n_hh;
  fetch_src(g0, a1, s1);
  fetch_src(g0 + stride, a2, s2);
  auto load_rec = [&](unsigned int src) { const float4* r = B.rec + (size_t)WF_REC * src; n_o = r[0]; n_d = r[1]; n_sa = r[2]; n_sb = r[3]; n_hh = r[4]; };
  if (a1) load_rec(s1);
  for (unsigned int gi = g0; gi < n_work; gi += stride) {
    const bool active = a1;
    const float4 o = n_o, d = n_d, sa = n_sa;
    const uint4 sb = make_uint4(__float_as_uint(n_sb.x), __float_as_uint(n_sb.y), __float_as_uint(n_sb.z), 0u);
    const uint4 hh = make_uint4(__float_as_uint(n_hh.x), __float_as_uint(n_hh.y), __float_as_uint(n_hh.z), __float_as_uint(n_hh.w));
    a1 = a2; s1 = s2;
    if (a1) load_rec(s1);
    fetch_src(gi + 2u * stride, a2, s2);
    bool cont = false, shadow = false;
    Ray ray;
    PathState ps;
    ShadowReq sr;
    sr.valid = false;
    if (active) {
      ray.o = f3(o); ray.time = o.w; ray.d = f3(d); ray.tmax = __int_as_float(0x7f800000); ray.mseed = __float_as_uint(d.w);
      ps.throughput = f3(sa); ps.prev_pdf_b = sa.w; ps.radiance = f3(0, 0, 0);
      ps.pixel = sb.x; ps.sample = sb.y; ps.bounce = (int)sb.z;
      if ((int)hh.z < 0) {  // miss
        ps.radiance = ps.throughput * background_color(P, ray.d);
      } else if (ps.bounce < P.max_depth) {  // photon_mapper.rs:336-339: the depth budget is checked after the hit
        HitRaw h;
        h.t = __longlong_as_double((long long)(((unsigned long long)hh.y << 32) | hh.x)); h.prim = (int)hh.z; h.aux = (int)hh.w;
        cont = shade_hit<INTEG>(S, P, ray, h, ps, sr);
        shadow = sr.valid;
      }
      accum_add(accum, ps.pixel, ps.radiance);
    }
    const unsigned int pe = warp_append(&B.cnt->n_ext_next, active && cont);
    if (active && cont) {
      float4* out = B.rec_n + (size_t)WF_REC * pe;
      out[0] = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.time);
      out[1] = make_float4(ray.d.x, ray.d.y, ray.d.z, __uint_as_float(ray.mseed));
      out[2] = make_float4(ps.throughput.x, ps.throughput.y, ps.throughput.z, ps.prev_pdf_b);
      out[3] = make_float4(__uint_as_float(ps.pixel), __uint_as_float(ps.sample), __uint_as_float((unsigned int)ps.bounce), 0.f);
    }
    if (INTEG != B200RT_INTEGRATOR_NEE) continue;  // only the NEE integrator produces shadow requests
    const unsigned int pq = warp_append(&B.cnt->n_shadow, shadow);
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
