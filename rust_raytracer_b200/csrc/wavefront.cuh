// wavefront.cuh — the wavefront pipeline over SoA float4 ray queues (north_star): generate -> extend ->
// per-material shade -> light-sample/shadow, with stream compaction by warp-aggregated atomics.
//
// The queues hold the ray PAYLOAD, not indices, as three streams per queue position (80 bytes):
//   ray   32 B: (o.xyz, time) (d.xyz, medium seed bits; t_max of an extend ray is always +inf)   read by extend + shade
//   state 32 B: (throughput.rgb, prev_pdf_b) (pixel, sample, bounce, 0)                           read by shade
//   hit   16 B: (t as 2 x u32 [f64], device prim + 2 | Cube side << 28, rank in bin)             extend -> shade
// Extend, shade and generate all read and write their streams strictly in order (every fetched DRAM sector is
// used whole: extend reads nothing but the ray stream); the only scattered traffic is the two 32-byte stores of
// a surviving path into its bin of the NEXT ray / state set (stores do not stall the warp the way the gathers of
// an index-sorted shade did: 154 B read per ray for 84 B of payload, latency-bound at 28 % issue).  The two sets
// are swapped by the host between iterations.  Shadow requests are stored by queue position: sh_o, sh_d, sh_c, sh_seed.
#pragma once
#include "kernels.cuh"

// The NEXT queue is ordered by WHERE each ray starts: bin = the hit primitive's index in BVH leaf order (spatially
// coherent) scaled to n_bins - 1 bins.  Extend knows the bin of every ray that will go on (hit, and the material
// scatters) and takes its rank inside the bin with one warp-aggregated atomic; a one-thread kernel turns the bin
// sizes into offsets; shade then walks the current queue IN ORDER (pure streaming reads of whole records) and
// writes each surviving path at offset[bin] + rank of the next set.  Extend warps therefore receive rays that left
// neighbouring primitives, with similar traversal cost (fog-born rays together, ground-box rays together, ...).
// A path that was predicted to go on but ended in shade (metal absorption, zero pdf, depth budget) leaves a dead
// record (zero direction) that extend skips.
// The queue streams are touched once per iteration and never fit L2 (GBs): streaming (evict-first) loads and
// stores keep them from displacing the scene tables there.
#ifndef B200RT_WF_STREAMING
#define B200RT_WF_STREAMING 1
#endif
#if B200RT_WF_STREAMING
#define WF_LD(p) __ldcs(p)
#define WF_ST(p, v) __stcs(p, v)
#else
#define WF_LD(p) (*(p))
#define WF_ST(p, v) (*(p) = (v))
#endif
#define WF_POS_BYTES 144  // device bytes per queue position: 2 sets x (ray 32 + state 32) + hit 16
#define WF_MAX_BINS 16384    // capacity of the bin tables; the number of bins in use is WfBuffers::n_bins
#define WF_SMEM_BINS 2048    // shade keeps the offsets in shared memory up to this many bins
#define WF_NO_BIN 0xffffffffu
#define WF_MAX_PIPES 4       // concurrent wavefront pipelines per context (api_render.cuh)

struct WfCounters {
  unsigned long long next_path;     // next path id to generate
  unsigned long long extend_rays, shadow_rays, paths_done;
  unsigned long long dead_rays;     // dead records met by extend (not rays: subtracted from the statistics)
  unsigned int n_ext;               // rays in the current set (extend + shade work list)
  unsigned int n_ext_next;          // survivors appended to the next set by shade
  unsigned int n_shadow;            // shadow requests this iteration
  unsigned int ext_cursor;          // work-fetch cursor of the persistent extend kernel
  unsigned int shadow_cursor;
  unsigned int pad;
};

struct WfBuffers {
  float4* ray;    // current set: 2 float4 per queue position
  float4* ray_n;  // next set
  float4* st;     // path state, current set: 2 float4 per queue position
  float4* st_n;
  uint4* hit;     // hit of the current set, written by extend: 1 uint4 per queue position
  unsigned int* bin_count;  // n_bins: continuing rays per bin (filled by extend)
  unsigned int* bin_off;    // n_bins + 1: start of every bin in the next set
  unsigned int bin_mul;     // bin = 1 + umulhi(prim, bin_mul)
  unsigned int n_bins;      // multiple of 1024 (the control kernels run one 1024-thread block), <= WF_MAX_BINS
  // Bins >= split_bin are filled from both ends by shade (k_wf_shade<INTEG, true>): the rays expected to traverse for long from the
  // front of the bin's range, the others and the dead records from the back (split_cnt[2 bin], [2 bin + 1]); extend then meets
  // warps of one kind.  split_box: every bin, "long" = the ray crosses the scene's dense box [split_lo, split_hi] (a 300,000-
  // triangle mesh in a room of a dozen walls: scene_compile.cpp dense_box).  Otherwise the bins of the primitives OUTSIDE the BVH
  // (book-2's fog) by whether the ray reaches the BVH root (measured and left off).  split_bin == n_bins: off.
  unsigned int* split_cnt;
  unsigned int split_bin;
  int split_box;
  float3 split_lo, split_hi;
  float4* sh_o; float4* sh_d; float4* sh_c; unsigned int* sh_seed;
  WfCounters* cnt;
  unsigned int pool;
};

// Append one element per participating lane with a single atomic per warp (warp-aggregated compaction).
// Must be called with the whole warp converged (every call site sits at the top level of a loop that whole warps
// walk together): the vote names all 32 lanes instead of guessing them with __activemask().
DEV unsigned int warp_append(unsigned int* counter, bool pred) {
  const unsigned mask = __ballot_sync(0xffffffffu, pred);
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
    float4* ray = B.ray + 2 * (size_t)(off + i);
    float4* st = B.st + 2 * (size_t)(off + i);
    WF_ST(ray + 0, make_float4(r.o.x, r.o.y, r.o.z, r.time));
    WF_ST(ray + 1, make_float4(r.d.x, r.d.y, r.d.z, __uint_as_float(r.mseed)));
    WF_ST(st + 0, make_float4(1.f, 1.f, 1.f, -1.f));
    WF_ST(st + 1, make_float4(__uint_as_float(pixel), __uint_as_float(sample), 0.f, 0.f));
  }
}

// One block turns the counters over between generate and extend (no host round trip inside an iteration).
__global__ void k_wf_control(WfBuffers B, PathMap M) {  // <<<1, WF_CTRL_THREADS>>>
  for (unsigned int k = threadIdx.x; k < B.n_bins; k += blockDim.x) B.bin_count[k] = 0;
  if (B.split_bin < B.n_bins) for (unsigned int k = 2u * B.split_bin + threadIdx.x; k < 2u * B.n_bins; k += blockDim.x) B.split_cnt[k] = 0;
  if (threadIdx.x != 0) return;
  WfCounters* c = B.cnt;
  const unsigned int n_new = wf_new_paths(c, M, B.pool);
  c->next_path += n_new;
  c->n_ext = c->n_ext_next + n_new;
  c->n_ext_next = 0;
  c->n_shadow = 0;
  c->ext_cursor = 0;
  c->shadow_cursor = 0;
  c->extend_rays += c->n_ext;
}
// after extend: bin sizes -> offsets in the next set; their total is the number of records shade will write.
// One block of WF_CTRL_THREADS threads: every thread owns n_bins / WF_CTRL_THREADS consecutive bins, the per-thread sums
// are scanned in shared memory (Hillis-Steele).  256 threads: with two pipelines in flight this kernel is launched
// while the other pipeline's persistent extend blocks fill the SMs and a 1024-thread block needs a whole SM to drain
// first (measured: 128 / 256 / 1024 threads make no difference on book-2, 1475.5 / 1475.8 / 1475.4 Mpaths/s).
#ifndef WF_CTRL_THREADS
#define WF_CTRL_THREADS 256  // power of two, divides 1024 (n_bins is a multiple of 1024)
#endif
__global__ void __launch_bounds__(WF_CTRL_THREADS) k_wf_control_shade(WfBuffers B) {  // <<<1, WF_CTRL_THREADS>>>
  __shared__ unsigned int sc[2][WF_CTRL_THREADS];
  const int t = threadIdx.x;
  const unsigned int per = B.n_bins / WF_CTRL_THREADS, first = (unsigned int)t * per;
  unsigned int mine = 0;
  for (unsigned int k = 0; k < per; ++k) mine += B.bin_count[first + k];
  sc[0][t] = mine;
  __syncthreads();
  int cur = 0;
  for (int d = 1; d < WF_CTRL_THREADS; d <<= 1) {
    sc[cur ^ 1][t] = sc[cur][t] + (t >= d ? sc[cur][t - d] : 0u);
    cur ^= 1;
    __syncthreads();
  }
  const unsigned int incl = sc[cur][t];
  unsigned int run = incl - mine;
  for (unsigned int k = 0; k < per; ++k) { B.bin_off[first + k] = run; run += B.bin_count[first + k]; }
  if (t == WF_CTRL_THREADS - 1) { B.bin_off[B.n_bins] = incl; B.cnt->n_ext_next = incl; }
}
__global__ void k_wf_control_shadow(WfBuffers B) { B.cnt->shadow_rays += B.cnt->n_shadow; }

// extend: persistent warps fetch 32 consecutive rays at a time (warp-level work fetch: one coalesced 1 KB read of
// the ray stream), traverse, store the hit and rank the queue position inside its bin for shade.
// (Measured and rejected: per-lane ray replacement a la Aila & Laine's dynamic fetch.  Refilling idle lanes from
// the queue mixes rays of different bins in one warp and loses the converged out-of-BVH tail: 60 ms -> 74..99 ms
// of extend per step on book-2 for refill thresholds 16..28, profiles/r1_experiments.md.  k_wf_extend_dyn below is the
// third implementation, for scenes with a dense box: correct, 8 % slower than this kernel there, off by default.)
#ifndef B200RT_EXTEND_BATCHES  // 32-ray batches a warp takes per cursor atomic (the later ones prefetched into L2)
#define B200RT_EXTEND_BATCHES 1
#endif
#ifndef B200RT_EXTEND_MINBLOCKS
#define B200RT_EXTEND_MINBLOCKS 8
#endif
#define WF_PRIM_DEAD (-2)
#define WF_NO_RANK 0xffffffffu
// hit.z: device primitive (-2 dead, -1 miss, else < 2^27) biased by 2, Cube side in the top bits
DEV unsigned int wf_pack_prim(int prim, int aux) { return (unsigned int)(prim + 2) | ((unsigned int)aux << 28); }
DEV int wf_unpack_prim(unsigned int w) { return (int)(w & 0x0FFFFFFFu) - 2; }
DEV int wf_unpack_aux(unsigned int w) { return (int)(w >> 28); }
template <bool OBJF>
__global__ void __launch_bounds__(128, B200RT_EXTEND_MINBLOCKS) k_wf_extend(DevScene S, WfBuffers B, float t_min, int light_scatters) {
  const unsigned int n = B.cnt->n_ext;
  const unsigned lane = threadIdx.x & 31;
  for (;;) {
    unsigned int base0 = 0;
    if (lane == 0) base0 = atomicAdd(&B.cnt->ext_cursor, 32u * B200RT_EXTEND_BATCHES);
    base0 = __shfl_sync(0xffffffffu, base0, 0);
    if (base0 >= n) break;
#if B200RT_EXTEND_BATCHES > 1
#pragma unroll 1
   for (unsigned int hb = 0; hb < B200RT_EXTEND_BATCHES; ++hb) {
    const unsigned int base = base0 + 32u * hb;
    if (base >= n) break;
    if (hb + 1 < B200RT_EXTEND_BATCHES && base + 32u + lane < n) asm volatile("prefetch.global.L2 [%0];" ::"l"(B.ray + 2 * (size_t)(base + 32u + lane)));
#else
   {
    const unsigned int base = base0;
#endif
    const unsigned int i = base + lane;
    const bool active = i < n;
    const float4* ray = B.ray + 2 * (size_t)i;
    unsigned int bin = WF_NO_BIN;  // "does not go on"
    HitRaw h;
    h.t = 0.; h.prim = WF_PRIM_DEAD; h.aux = 0;
    DBG_RAY(h.dbg = 0);
    bool dead = false;
    float4 o = make_float4(0.f, 0.f, 0.f, 0.f), d = o;
    if (active) {
      o = WF_LD(ray + 0); d = WF_LD(ray + 1);
      dead = d.x == 0.f && d.y == 0.f && d.z == 0.f;
    }
    // the lanes that traverse, taken with the warp converged: trace() lets exactly these wait for each other
    const unsigned tracing = __ballot_sync(0xffffffffu, active && !dead);
    if (active && !dead) {
      Ray r;
      r.o = f3(o); r.time = o.w; r.d = f3(d); r.tmax = __int_as_float(0x7f800000); r.mseed = __float_as_uint(d.w);
      trace<false, OBJF>(S, r, t_min, h, tracing);
      if (h.prim >= 0) {  // will the path go on?  Every material scatters except DiffuseLight (material.rs:201-212 behind a flag)
        const int4 meta = __ldg(&S.prim_meta[h.prim]);
        const int mtype = __float_as_int(__ldg(&S.materials[meta.x >> 8]).x);
        if (mtype != B200RT_MAT_DIFFUSE_LIGHT || light_scatters) bin = 1u + __umulhi((unsigned int)h.prim, B.bin_mul);
      }
    }
    // rank inside the bin: the leader lane of every bin present in the warp issues its atomic in the same
    // instruction, so the warp pays one atomic round trip whatever the number of bins
    const unsigned peers = __match_any_sync(0xffffffffu, bin);
    const int leader = __ffs((int)peers) - 1;
    const bool goes_on = bin != WF_NO_BIN;
    unsigned int rank = 0;
    if (goes_on && (int)lane == leader) rank = atomicAdd(&B.bin_count[bin], (unsigned)__popc(peers));
    rank = __shfl_sync(0xffffffffu, rank, leader) + __popc(peers & ((1u << lane) - 1u));
    const unsigned n_dead = __popc(__ballot_sync(0xffffffffu, dead));
    if (n_dead && lane == 0) atomicAdd(&B.cnt->dead_rays, (unsigned long long)n_dead);
    DBG_RAY(if (active && g_dbg_steps) g_dbg_steps[i] = dead ? 0xffffffffu : h.dbg);
    if (active) {
      const unsigned long long tb = (unsigned long long)__double_as_longlong(h.t);
      WF_ST(B.hit + i, make_uint4((unsigned int)tb, (unsigned int)(tb >> 32), wf_pack_prim(h.prim, h.aux), goes_on ? rank : WF_NO_RANK));
    }
   }
  }
}

// extend with per-lane ray replacement, for scenes with a dense box (a mesh in a room: WfBuffers::split_box).  There the rays that
// cross the box take 12 to 40 node steps each and a warp of k_wf_extend runs as long as its deepest lane (6 of 32 lanes per
// instruction, profiles/r2_mesh_split_extend_shade_full.txt).  Here a lane that has finished its ray waits only until `thresh`
// lanes of its warp are idle; those lanes then store their hits, take their ranks and fetch the next rays of the queue together,
// outside the node loop, and start at the root while the others go on.  On book-2 (5 node steps per ray, an out-of-BVH tail that
// wants the warp converged) the same idea lost 20-80 % twice (profiles/r1_experiments.md); the refill costs ~200 instructions per
// group of >= thresh rays, which is noise against 2,500 per box-crossing ray and ruin against 700.  Scenes whose primitives are
// all inside the BVH only (no tail after the traversal); the host falls back to k_wf_extend otherwise.
// MEASURED (profiles/r2_experiments.md section 12): 197 ms against k_wf_extend's 182 on the mesh config.  The replacement buys 5 %
// (209 ms when the same code refills only whole warps) and the per-lane state carried across the refill branch costs 15 %
// (270 B of spills at 72 registers).  Launched only with B200RT_WF_DYN=1; kept with its parity test as the record of the attempt.
#ifndef B200RT_EXTEND_DYN_MINBLOCKS
#define B200RT_EXTEND_DYN_MINBLOCKS 7
#endif
template <bool OBJF>
__global__ void __launch_bounds__(128, B200RT_EXTEND_DYN_MINBLOCKS) k_wf_extend_dyn(DevScene S, WfBuffers B, float t_min, int light_scatters, int thresh) {
  const unsigned int n = B.cnt->n_ext;
  const unsigned lane = threadIdx.x & 31;
  const unsigned FULL = 0xffffffffu;
  unsigned int idx = 0xffffffffu;  // queue position of the ray this lane holds; none
  uint32_t cur = B200RT_TRAV_DONE;
  bool more = true;                // the queue still has rays (warp-uniform)
  // the ray, once: RayF plus the two per-ray terms of RayT that are not in it (Ray and RayT are re-assembled from these where
  // trace()'s callees want them: a second copy of o, d and 1/d kept across the refill branch costs the registers the node step needs)
  RayF rf;
  float3 oi = f3(0.f, 0.f, 0.f);
  float slack = 0.f;
  unsigned int mseed = 0u;
  HitRaw h;
  TStack stack;
  int sp = 0;
  float t_best32 = 0.f;
  rf.o = rf.d = rf.inv = f3(0.f, 0.f, 0.f); rf.a = 0.f; rf.time = 0.f;
  h.t = 0.; h.prim = -1; h.aux = 0;
  for (;;) {
    const unsigned idle = __ballot_sync(FULL, cur == B200RT_TRAV_DONE);
    if (idle == FULL || (more && (int)__popc(idle) >= thresh)) {
      bool more_l = more;
      if (cur == B200RT_TRAV_DONE) {  // exactly the lanes of `idle`
        // (1) finished rays: will the path go on?  bin, rank inside the bin (one atomic per bin present), hit record
        const bool has = idx != 0xffffffffu;
        unsigned int bin = WF_NO_BIN;
        if (has && h.prim >= 0) {
          const int4 meta = __ldg(&S.prim_meta[h.prim]);
          const int mtype = __float_as_int(__ldg(&S.materials[meta.x >> 8]).x);
          if (mtype != B200RT_MAT_DIFFUSE_LIGHT || light_scatters) bin = 1u + __umulhi((unsigned int)h.prim, B.bin_mul);
        }
        const bool goes_on = bin != WF_NO_BIN;
        const unsigned peers = __match_any_sync(idle, goes_on ? bin : (0x80000000u | lane));
        const int leader = __ffs((int)peers) - 1;
        unsigned int rank = 0;
        if (goes_on && (int)lane == leader) rank = atomicAdd(&B.bin_count[bin], (unsigned)__popc(peers));
        rank = __shfl_sync(idle, rank, leader) + __popc(peers & ((1u << lane) - 1u));
        if (has) {
          const unsigned long long tb = (unsigned long long)__double_as_longlong(h.t);
          WF_ST(B.hit + idx, make_uint4((unsigned int)tb, (unsigned int)(tb >> 32), wf_pack_prim(h.prim, h.aux), goes_on ? rank : WF_NO_RANK));
          idx = 0xffffffffu;
        }
        // (2) the next rays of the queue, one per idle lane
        if (more) {
          const unsigned k = __popc(idle);
          const int first = __ffs((int)idle) - 1;
          unsigned int base = 0;
          if ((int)lane == first) base = atomicAdd(&B.cnt->ext_cursor, k);
          base = __shfl_sync(idle, base, first);
          more_l = base + k < n;
          const unsigned int i = base + __popc(idle & ((1u << lane) - 1u));
          bool dead = false;
          if (i < n) {
            const float4 o = WF_LD(B.ray + 2 * (size_t)i), d = WF_LD(B.ray + 2 * (size_t)i + 1);
            dead = d.x == 0.f && d.y == 0.f && d.z == 0.f;
            if (dead) {  // nothing to trace, nothing goes on
              WF_ST(B.hit + i, make_uint4(0u, 0u, wf_pack_prim(WF_PRIM_DEAD, 0), WF_NO_RANK));
            } else {
              idx = i;
              Ray ray;
              ray.o = f3(o); ray.time = o.w; ray.d = f3(d); ray.tmax = __int_as_float(0x7f800000); ray.mseed = __float_as_uint(d.w);
              rf = make_rayf(ray);
              const RayT rt = make_rayt(S, rf);
              oi = rt.oi; slack = rt.slack; mseed = ray.mseed;
              h.t = (double)ray.tmax; h.prim = -1; h.aux = 0;
              t_best32 = ray.tmax;
              sp = 0;
              cur = 0u;  // root
              DBG_INC(0);
            }
          }
          const unsigned n_dead = __popc(__ballot_sync(idle, dead));
          if (n_dead && (int)lane == first) atomicAdd(&B.cnt->dead_rays, (unsigned long long)n_dead);
        }
      }
      more = __shfl_sync(FULL, more_l, __ffs((int)idle) - 1) != 0;
      if (!__any_sync(FULL, cur != B200RT_TRAV_DONE)) {
        if (!more) break;
        continue;  // a group of dead records only: fetch again
      }
    }
    {
      RayT rt;
      asm("mov.b64 %0, %1;" : "=l"(rt.nodes) : "l"(S.nodes));
      rt.inv = rf.inv; rt.oi = oi; rt.slack = slack;
      while (!(cur & B200RT_LEAF_BIT)) cur = node_step(S, rf, rt, cur, t_min, t_best32, stack, sp);
    }
    if (cur != B200RT_TRAV_DONE) {
      Ray ray;
      ray.o = rf.o; ray.d = rf.d; ray.time = rf.time; ray.tmax = __int_as_float(0x7f800000); ray.mseed = mseed;
      const int first = (int)(cur & 0x07FFFFFFu), count = (int)((cur >> 27) & 15u) + 1;
      DBG_INC(2);
      for (int k = first; k < first + count; ++k) test_prim<OBJF>(S, ray, rf, k, t_min, t_best32, h);
      cur = pop_next(stack, sp, t_best32);
    }
  }
}

// ---- bulk asynchronous copies (TMA, 1-D): global -> shared memory behind an mbarrier
#ifndef B200RT_SHADE_BULK  // > 0: stages of the shade kernel's input tiles kept in flight by cp.async.bulk (0: plain loads + L2 prefetch)
#define B200RT_SHADE_BULK 0
#endif
DEV uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
DEV void mbar_init(uint32_t bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory"); }
DEV void mbar_expect_tx(uint32_t bar, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory"); }
DEV void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar, unsigned long long policy) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar), "l"(policy) : "memory");
}
DEV void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile("{\n.reg .pred P1;\nLAB_WAIT:\nmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n@P1 bra DONE;\nbra LAB_WAIT;\nDONE:\n}" ::"r"(bar), "r"(parity) : "memory");
}

// shade: walks the current queue in order (streaming reads of whole records) and stores every surviving path at
// its place in the next set.
#ifndef B200RT_SHADE_MINBLOCKS_PATH  // the PATH instantiation fits 96 registers without spilling (5 blocks/SM) ...
#define B200RT_SHADE_MINBLOCKS_PATH 5
#endif
#ifndef B200RT_SHADE_MINBLOCKS  // ... MIXTURE / NEE need 120 / 126
#define B200RT_SHADE_MINBLOCKS 4
#endif
// Does the ray reach any child box of the BVH root within [t_min, inf)?  (The first node step of trace(), without the stack.)
DEV bool wf_ray_hits_root(const DevScene& S, const Ray& r, float t_min) {
  const float3 inv = f3(safe_rcp(r.d.x), safe_rcp(r.d.y), safe_rcp(r.d.z));
  const float4 lox = __ldg(S.nodes + 0), hix = __ldg(S.nodes + 1), loy = __ldg(S.nodes + 2), hiy = __ldg(S.nodes + 3), loz = __ldg(S.nodes + 4), hiz = __ldg(S.nodes + 5);
  const float lx[4] = {lox.x, lox.y, lox.z, lox.w}, hx[4] = {hix.x, hix.y, hix.z, hix.w}, ly[4] = {loy.x, loy.y, loy.z, loy.w},
              hy[4] = {hiy.x, hiy.y, hiy.z, hiy.w}, lz[4] = {loz.x, loz.y, loz.z, loz.w}, hz[4] = {hiz.x, hiz.y, hiz.z, hiz.w};
  bool any = false;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float t0x = (lx[i] - r.o.x) * inv.x, t1x = (hx[i] - r.o.x) * inv.x, t0y = (ly[i] - r.o.y) * inv.y, t1y = (hy[i] - r.o.y) * inv.y;
    const float t0z = (lz[i] - r.o.z) * inv.z, t1z = (hz[i] - r.o.z) * inv.z;
    const float n = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fmaxf(fminf(t0z, t1z), t_min));
    const float f = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fmaxf(t0z, t1z));
    any = any || n <= f * 1.00001f;  // an ordering hint, not a decision: empty slots (inverted infinite boxes) never pass
  }
  return any;
}

// Does the ray cross the box within [t_min, inf)?  An ordering hint like wf_ray_hits_root, not a decision.
DEV bool wf_ray_hits_box(const float3 lo, const float3 hi, const Ray& r, float t_min) {
  const float3 inv = f3(safe_rcp(r.d.x), safe_rcp(r.d.y), safe_rcp(r.d.z));
  const float t0x = (lo.x - r.o.x) * inv.x, t1x = (hi.x - r.o.x) * inv.x, t0y = (lo.y - r.o.y) * inv.y, t1y = (hi.y - r.o.y) * inv.y;
  const float t0z = (lo.z - r.o.z) * inv.z, t1z = (hi.z - r.o.z) * inv.z;
  const float n = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fmaxf(fminf(t0z, t1z), t_min));
  const float f = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fmaxf(t0z, t1z));
  return n <= f * 1.00001f;
}

#ifndef B200RT_SHADE_COOP_PERLIN  // 1: NOISE textures evaluated by the whole warp, one hit at a time (perlin_turb_warp): measured -0.65 %, profiles/r2_experiments.md section 7
#define B200RT_SHADE_COOP_PERLIN 0
#endif
// SPLIT: the two-ended bins of WfBuffers::split_bin (a separate instantiation: the plain one is the kernel the book scenes run)
template <int INTEG, bool SPLIT>
__global__ void __launch_bounds__(128, INTEG == B200RT_INTEGRATOR_PATH ? B200RT_SHADE_MINBLOCKS_PATH : B200RT_SHADE_MINBLOCKS)
k_wf_shade(DevScene S, WfBuffers B, RenderConst P, float* __restrict__ accum) {
  __shared__ unsigned int s_off[WF_SMEM_BINS + 1];
  const bool off_in_smem = B.n_bins <= WF_SMEM_BINS;
  if (off_in_smem) for (unsigned int k = threadIdx.x; k <= B.n_bins; k += blockDim.x) s_off[k] = B.bin_off[k];
  __syncthreads();
  const unsigned int n_ext = B.cnt->n_ext;
  const unsigned int n_work = (n_ext + 31u) & ~31u;  // whole warps stay in the loop (warp-aggregated appends below)
  const unsigned int stride = gridDim.x * blockDim.x;
  const unsigned int g0 = blockIdx.x * blockDim.x + threadIdx.x;
  float4 n_o, n_d, n_sa, n_sb;
  uint4 n_hh;
  auto load_rec = [&](unsigned int src) {
    const float4* r = B.ray + 2 * (size_t)src;
    const float4* t = B.st + 2 * (size_t)src;
    n_o = WF_LD(r + 0); n_d = WF_LD(r + 1); n_sa = WF_LD(t + 0); n_sb = WF_LD(t + 1); n_hh = WF_LD(B.hit + src);
  };
#ifndef B200RT_SHADE_L2PF  // iterations ahead whose records shade pulls into L2 (15.9 -> 14.5 ms of shade per step at 1; 14.7 at 2)
#define B200RT_SHADE_L2PF 1
#endif
#ifndef B200RT_SHADE_PF_LEVEL
#define B200RT_SHADE_PF_LEVEL "L2"
#endif
#ifndef B200RT_SHADE_PREFETCH  // measured: with streaming reads the explicit prefetch only costs registers (18.3 ms vs 17.2 ms at 96 regs)
#define B200RT_SHADE_PREFETCH 0
#endif
#if B200RT_SHADE_BULK
  // Input tiles of 128 queue positions (ray 4 KB | state 4 KB | hit 2 KB) are copied to shared memory by the TMA unit,
  // B200RT_SHADE_BULK tiles ahead of the one being shaded: the loads occupy no registers and no warp waits on DRAM.
  __shared__ __align__(128) unsigned char s_tile[B200RT_SHADE_BULK][10240];
  __shared__ __align__(8) unsigned long long s_bar[B200RT_SHADE_BULK];
  const unsigned int n_tiles = (n_ext + 127u) >> 7;
  unsigned long long l2_evict_first;
  asm("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(l2_evict_first));
  auto issue_tile = [&](unsigned int tile, int st) {  // one thread
    const unsigned int first = tile << 7, cnt = min(128u, n_ext - first);
    const uint32_t bar = smem_u32(&s_bar[st]), dst = smem_u32(&s_tile[st][0]);
    mbar_expect_tx(bar, cnt * 80u);
    bulk_g2s(dst, B.ray + 2 * (size_t)first, cnt * 32u, bar, l2_evict_first);
    bulk_g2s(dst + 4096u, B.st + 2 * (size_t)first, cnt * 32u, bar, l2_evict_first);
    bulk_g2s(dst + 8192u, B.hit + first, cnt * 16u, bar, l2_evict_first);
  };
  if (threadIdx.x == 0) {
    for (int st = 0; st < B200RT_SHADE_BULK; ++st) mbar_init(smem_u32(&s_bar[st]), 1u);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0)
    for (int st = 0; st < B200RT_SHADE_BULK; ++st) { const unsigned int tile = blockIdx.x + st * gridDim.x; if (tile < n_tiles) issue_tile(tile, st); }
  unsigned int it = 0;
  (void)n_work; (void)stride; (void)g0;
  for (unsigned int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {  // block-uniform trip count (the barrier below)
    const unsigned int gi = (tile << 7) + threadIdx.x;
    {
      const int st = (int)(it % B200RT_SHADE_BULK);
      mbar_wait(smem_u32(&s_bar[st]), (it / B200RT_SHADE_BULK) & 1u);
      if (gi < n_ext) {
        const float4* r = reinterpret_cast<const float4*>(&s_tile[st][0]) + 2 * threadIdx.x;
        const float4* t = reinterpret_cast<const float4*>(&s_tile[st][4096]) + 2 * threadIdx.x;
        n_o = r[0]; n_d = r[1]; n_sa = t[0]; n_sb = t[1]; n_hh = reinterpret_cast<const uint4*>(&s_tile[st][8192])[threadIdx.x];
      }
      __syncthreads();  // every thread holds its record: the stage can be refilled
      const unsigned int next = tile + B200RT_SHADE_BULK * gridDim.x;
      if (threadIdx.x == 0 && next < n_tiles) issue_tile(next, st);
    }
#else
#if B200RT_SHADE_PREFETCH
  if (g0 < n_ext) load_rec(g0);
#endif
  for (unsigned int gi = g0; gi < n_work; gi += stride) {
#endif
#if !B200RT_SHADE_PREFETCH && !B200RT_SHADE_BULK
#if B200RT_SHADE_L2PF
    if (gi + B200RT_SHADE_L2PF * stride < n_ext) {  // pull a later iteration's records into L2: costs no registers
      const unsigned int nx = gi + B200RT_SHADE_L2PF * stride;
      asm volatile("prefetch.global." B200RT_SHADE_PF_LEVEL " [%0];" ::"l"(B.ray + 2 * (size_t)nx));
      asm volatile("prefetch.global." B200RT_SHADE_PF_LEVEL " [%0];" ::"l"(B.st + 2 * (size_t)nx));
      asm volatile("prefetch.global." B200RT_SHADE_PF_LEVEL " [%0];" ::"l"(B.hit + nx));
    }
#endif
    if (gi < n_ext) load_rec(gi);
#endif
    const float4 o = n_o, d = n_d, sa = n_sa;
    const uint4 sb = make_uint4(__float_as_uint(n_sb.x), __float_as_uint(n_sb.y), __float_as_uint(n_sb.z), 0u);
    const uint4 hh = n_hh;
    const int hprim = wf_unpack_prim(hh.z);
    const unsigned int rank = hh.w;
#if B200RT_SHADE_PREFETCH
    if (gi + stride < n_ext) load_rec(gi + stride);
#endif
    const bool active = gi < n_ext && hprim != WF_PRIM_DEAD;
    bool cont = false, shadow = false;
    Ray ray;
    PathState ps;
    ShadowReq sr;
    sr.valid = false;
#if B200RT_SHADE_COOP_PERLIN
    // hit record per lane, then the NOISE textures of the warp one after the other with all 32 lanes on each (perlin_turb_warp),
    // then the rest of the shading per lane
    HitInfo hi;
    float4 mat = make_float4(0.f, 0.f, 0.f, 0.f);
    float turb = -1.f;
    bool surf = false;
    if (active) {
      ray.o = f3(o); ray.time = o.w; ray.d = f3(d); ray.tmax = __int_as_float(0x7f800000); ray.mseed = __float_as_uint(d.w);
      ps.throughput = f3(sa); ps.prev_pdf_b = sa.w; ps.radiance = f3(0, 0, 0);
      ps.pixel = sb.x; ps.sample = sb.y; ps.bounce = (int)sb.z;
      surf = hprim >= 0 && ps.bounce < P.max_depth;  // photon_mapper.rs:336-339: the depth budget is checked after the hit
      if (surf) {
        HitRaw h;
        h.t = __longlong_as_double((long long)(((unsigned long long)hh.y << 32) | hh.x)); h.prim = hprim; h.aux = wf_unpack_aux(hh.z);
        shade_hit_begin(S, ray, h, hi, mat);
      }
    }
    {
      unsigned noisy = __ballot_sync(0xffffffffu, surf && texture_is_noise(S, __float_as_int(mat.y)));
      while (noisy) {
        const int src = __ffs((int)noisy) - 1;
        noisy &= noisy - 1u;
        const float3 q = f3(__shfl_sync(0xffffffffu, hi.p.x, src), __shfl_sync(0xffffffffu, hi.p.y, src), __shfl_sync(0xffffffffu, hi.p.z, src));
        const float tb = perlin_turb_warp(S, q, threadIdx.x & 31u);
        if ((int)(threadIdx.x & 31u) == src) turb = tb;
      }
    }
    if (active) {
      if (hprim < 0) ps.radiance = ps.throughput * background_color(P, ray.d);  // miss
      else if (surf) {
        cont = shade_hit_end<INTEG>(S, P, ray, hi, mat, ps, sr, turb);
        shadow = sr.valid;
      }
      accum_add(accum, ps.pixel, ps.radiance);
    }
#else
    if (active) {
      ray.o = f3(o); ray.time = o.w; ray.d = f3(d); ray.tmax = __int_as_float(0x7f800000); ray.mseed = __float_as_uint(d.w);
      ps.throughput = f3(sa); ps.prev_pdf_b = sa.w; ps.radiance = f3(0, 0, 0);
      ps.pixel = sb.x; ps.sample = sb.y; ps.bounce = (int)sb.z;
      if (hprim < 0) {  // miss
        ps.radiance = ps.throughput * background_color(P, ray.d);
      } else if (ps.bounce < P.max_depth) {  // photon_mapper.rs:336-339: the depth budget is checked after the hit
        HitRaw h;
        h.t = __longlong_as_double((long long)(((unsigned long long)hh.y << 32) | hh.x)); h.prim = hprim; h.aux = wf_unpack_aux(hh.z);
        cont = shade_hit<INTEG>(S, P, ray, h, ps, sr);
        shadow = sr.valid;
      }
      accum_add(accum, ps.pixel, ps.radiance);
    }
#endif
    const bool writes = active && rank != WF_NO_RANK;  // the place extend reserved in the next set: the new ray, or a dead record
    const unsigned int bin = writes ? 1u + __umulhi((unsigned int)hprim, B.bin_mul) : 0u;  // the bin extend ranked it in
    unsigned int slot = writes ? (off_in_smem ? s_off[bin] : __ldg(&B.bin_off[bin])) + rank : 0u;
    if (SPLIT) {
      // Extend's rank is not used in a two-ended bin, only its count: the place is taken from the front or the back of the bin's
      // range by one atomic per (bin, end) present in the warp.
      const bool split = writes && bin >= B.split_bin;
      const bool front = split && cont && (B.split_box ? wf_ray_hits_box(B.split_lo, B.split_hi, ray, P.t_min) : wf_ray_hits_root(S, ray, P.t_min));
      const unsigned lane_id = threadIdx.x & 31u;
      const unsigned int key = split ? 2u * bin + (front ? 0u : 1u) : (0x80000000u | lane_id);
      const unsigned peers = __match_any_sync(0xffffffffu, key);
      const int leader = __ffs((int)peers) - 1;
      unsigned int r2 = 0;
      if (split && (int)lane_id == leader) r2 = atomicAdd(&B.split_cnt[key], (unsigned)__popc(peers));
      r2 = __shfl_sync(0xffffffffu, r2, leader) + __popc(peers & ((1u << lane_id) - 1u));
      if (split) {
        const unsigned int lo = off_in_smem ? s_off[bin] : __ldg(&B.bin_off[bin]), hi = off_in_smem ? s_off[bin + 1] : __ldg(&B.bin_off[bin + 1]);
        slot = front ? lo + r2 : hi - 1u - r2;
      }
    }
    if (writes) {
      const size_t dst = 2 * (size_t)slot;
      float4* out = B.ray_n + dst;
      if (cont) {
        float4* out_s = B.st_n + dst;
        WF_ST(out + 0, make_float4(ray.o.x, ray.o.y, ray.o.z, ray.time));
        WF_ST(out + 1, make_float4(ray.d.x, ray.d.y, ray.d.z, __uint_as_float(ray.mseed)));
        WF_ST(out_s + 0, make_float4(ps.throughput.x, ps.throughput.y, ps.throughput.z, ps.prev_pdf_b));
        WF_ST(out_s + 1, make_float4(__uint_as_float(ps.pixel), __uint_as_float(ps.sample), __uint_as_float((unsigned int)ps.bounce), 0.f));
      } else {
        WF_ST(out + 0, make_float4(0.f, 0.f, 0.f, 0.f));
        WF_ST(out + 1, make_float4(0.f, 0.f, 0.f, 0.f));
      }
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
template <bool OBJF>
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
      trace<true, OBJF>(S, r, t_min, h);
      if (h.prim < 0) accum_add(accum, __float_as_uint(c.w), f3(c));
    }
  }
}

// tail: the paths still alive when a render has no new paths left and its queue has run low are finished here, one thread per
// path, instead of through up to max_depth more near-empty iterations of five launches each (a 400x225x100-spp render is 9 M paths
// in ~5 ms, of which the last 30 iterations carry < 1000 rays and cost ~1 ms of launch latency and host polling).  The records are
// the survivors of the last shade pass (the NEXT set); the draws are keyed (pixel, sample, bounce) as everywhere, so the paths are
// the ones the queue iterations would have traced: the ray statistics and the image do not change.
template <bool OBJF>
__global__ void __launch_bounds__(128) k_wf_tail(DevScene S, WfBuffers B, RenderConst P, float* __restrict__ accum) {
  const unsigned int n = B.cnt->n_ext_next;
  unsigned long long n_ext = 0, n_shd = 0;
  for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float4 o = WF_LD(B.ray_n + 2 * (size_t)i), d = WF_LD(B.ray_n + 2 * (size_t)i + 1);
    if (d.x == 0.f && d.y == 0.f && d.z == 0.f) continue;  // a dead record
    const float4 sa = WF_LD(B.st_n + 2 * (size_t)i), sb = WF_LD(B.st_n + 2 * (size_t)i + 1);
    Ray ray;
    ray.o = f3(o); ray.time = o.w; ray.d = f3(d); ray.tmax = __int_as_float(0x7f800000); ray.mseed = __float_as_uint(d.w);
    PathState ps;
    ps.throughput = f3(sa); ps.prev_pdf_b = sa.w; ps.radiance = f3(0, 0, 0);
    ps.pixel = __float_as_uint(sb.x); ps.sample = __float_as_uint(sb.y); ps.bounce = (int)__float_as_uint(sb.z);
    for (;;) {
      HitRaw h;
      trace<false, OBJF>(S, ray, P.t_min, h);
      n_ext++;
      if (h.prim < 0) { ps.radiance = ps.radiance + ps.throughput * background_color(P, ray.d); break; }
      if (ps.bounce >= P.max_depth) break;  // photon_mapper.rs:336-339: the depth budget is checked after the hit
      ShadowReq sr;
      const bool cont = shade_hit(S, P, ray, h, ps, sr);
      if (sr.valid) {
        HitRaw sh;
        trace<true, OBJF>(S, sr.ray, P.t_min, sh);
        n_shd++;
        if (sh.prim < 0) ps.radiance = ps.radiance + sr.contrib;
      }
      if (!cont) break;
      ray.tmax = __int_as_float(0x7f800000);
    }
    accum_add(accum, ps.pixel, ps.radiance);
  }
  for (int o = 16; o > 0; o >>= 1) { n_ext += __shfl_down_sync(0xffffffffu, n_ext, o); n_shd += __shfl_down_sync(0xffffffffu, n_shd, o); }
  if ((threadIdx.x & 31) == 0) {
    if (n_ext) atomicAdd(&B.cnt->extend_rays, n_ext);
    if (n_shd) atomicAdd(&B.cnt->shadow_rays, n_shd);
  }
}
