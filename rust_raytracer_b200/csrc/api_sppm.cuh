// api_sppm.cuh — SPPMIntegrator::new + capture_image (photon_mapper.rs:139-236, main.rs:51-53) on the device: the
// host loop over the SPPM iterations (photon pass -> grid build for both maps -> eye pass) and the final pass.
// Included by api.cu after api_render.cuh.
#pragma once
#include "sppm.cuh"

namespace {

#define SPPM_MAX_NB 8  // photon buffers per call: passes in flight + 1

struct SppmBuffers {
  float4 *pg[SPPM_MAX_NB] = {}, *pc[SPPM_MAX_NB] = {};  // photons in emission order, one pair per photon pass in flight (+ the one being sorted)
  float4 *sg = nullptr, *sc = nullptr;  // photons in cell order: `sorted_cap` positions, then 2 x `sorted_cap` attribute rows
  size_t sorted_cap = 0;
  uint32_t *tile_sum = nullptr, *cell_count = nullptr, *start_g = nullptr, *start_c = nullptr, *cursor = nullptr;
  float4 *px_g = nullptr, *px_c = nullptr;
  uint2* px_n = nullptr;
  float4 *vp = nullptr, *vb = nullptr;       // eye pass: visible point and bsdf per pixel
  uint32_t *knn_list = nullptr, *knn_n = nullptr;  // eye pass: (pixel, map) pairs for the k-nearest kernel
  SppmCounters* cnt = nullptr;
  float* accum = nullptr;
  cudaError_t alloc_maps(size_t cap, size_t n_pix) {  // the cell-ordered maps and the eye-pass scratch (both entry points)
    sorted_cap = cap;
    cudaError_t e = cudaMalloc((void**)&sg, cap * 48);
    if (e == cudaSuccess) e = cudaMalloc((void**)&sc, cap * 48);
    if (e == cudaSuccess) e = cudaMalloc((void**)&vp, 16 * n_pix);
    if (e == cudaSuccess) e = cudaMalloc((void**)&vb, 16 * n_pix);
    if (e == cudaSuccess) e = cudaMalloc((void**)&knn_list, 4 * 2 * n_pix);
    if (e == cudaSuccess) e = cudaMalloc((void**)&knn_n, 256);
    return e;
  }
  SppmEye eye() const { return SppmEye{vp, vb, knn_list, knn_n}; }
  void release() {
    for (int k = 0; k < SPPM_MAX_NB; ++k) { if (pg[k]) cudaFree(pg[k]); if (pc[k]) cudaFree(pc[k]); }
    void* all[] = {sg, sc, tile_sum, cell_count, start_g, start_c, cursor, px_g, px_c, px_n, vp, vb, knn_list, knn_n, cnt, accum};
    for (void* p : all) if (p) cudaFree(p);
  }
};

const int SPPM_MAX_RES = 128;

// grid resolution and placement for a map of n photons inside the photon bounding box [lo, hi]
SppmGrid sppm_make_grid(const uint32_t* start, uint32_t n, const float lo[3], const float hi[3]) {
  SppmGrid G;
  G.pos = nullptr; G.attr = nullptr; G.cell_start = start; G.n = n;  // pos / attr: set by sppm_build_grid
  int res = (int)std::cbrt((double)n / 2.0) + 1;
  G.res = res < 1 ? 1 : (res > SPPM_MAX_RES ? SPPM_MAX_RES : res);
  float ext = std::fmax(hi[0] - lo[0], std::fmax(hi[1] - lo[1], hi[2] - lo[2]));
  if (!(ext > 0.f)) ext = 1.f;
  G.cell = ext * 1.0001f / (float)G.res;
  G.inv_cell = 1.0f / G.cell;
  G.lo = make_float3(lo[0], lo[1], lo[2]);
  return G;
}

SppmSegs sppm_one_seg(const float4* photons, uint32_t n) {
  SppmSegs S;
  std::memset(&S, 0, sizeof(S));
  S.p[0] = photons; S.start[0] = 0; S.start[1] = n; S.n = 1;
  return S;
}

// `segs`: where the G.n photons of the map live (one local buffer, or the shards of several devices)
int sppm_build_grid(b200rt_ctx* ctx, SppmBuffers& B, SppmGrid& G, const SppmSegs& segs, uint32_t* start, float4* sorted) {
  cudaStream_t q = ctx->stream;
  const uint32_t n_cells = (uint32_t)G.res * G.res * G.res;
  CU_CHECK(ctx, cudaMemsetAsync(B.cell_count, 0, sizeof(uint32_t) * n_cells, q));
  const int blocks = ctx->prop.multiProcessorCount * 4;
  if (G.n) k_sppm_grid_count<<<blocks, 256, 0, q>>>(segs, G.n, G.lo, G.inv_cell, G.res, B.cell_count);
  const uint32_t n_tiles = (n_cells + SPPM_SCAN_TILE - 1) / SPPM_SCAN_TILE;  // <= 512 at 128^3 cells
  k_sppm_scan_tile_sums<<<n_tiles, 1024, 0, q>>>(B.cell_count, n_cells, B.tile_sum);
  k_sppm_scan_tile_offsets<<<1, 1024, 0, q>>>(B.tile_sum, n_tiles, start, n_cells);
  k_sppm_scan_cells<<<n_tiles, 1024, 0, q>>>(B.cell_count, n_cells, B.tile_sum, start, B.cursor);
  float4* attr = sorted + B.sorted_cap;
  if (G.n) k_sppm_grid_fill<<<blocks, 256, 0, q>>>(segs, G.n, G.lo, G.inv_cell, G.res, B.cursor, sorted, attr);
  CU_CHECK(ctx, cudaGetLastError());
  G.pos = sorted; G.attr = attr;  // the eye pass reads the cell-ordered copy
  return B200RT_OK;
}

// SPPMIntegrator::update_sppm for the pixels [pix_first, pix_first + pix_count) on the context stream: the eye rays to their
// visible points, the radius gather for the pixels that have statistics, the k-nearest start for those that do not
int sppm_eye_pass(b200rt_ctx* ctx, const b200rt_scene* s, const RenderConst& Peye, const SppmConst& sp, int it, const SppmGrid& Gg, const SppmGrid& Gc,
                  const SppmPixels& px, SppmBuffers& B, SppmCounters* eye_cnt, uint32_t pix_first, uint32_t pix_count) {
  cudaStream_t q = ctx->stream;
  const int sms = ctx->prop.multiProcessorCount;
  const SppmEye E = B.eye();
  if (pix_count == 0) return B200RT_OK;
  CU_CHECK(ctx, cudaMemsetAsync(B.knn_n, 0, 4, q));
  const int trace_blocks = (int)std::min<uint32_t>((pix_count + 127u) / 128u, (uint32_t)sms * 8u);
  if (s->dev.has_objf) k_sppm_eye_trace<true><<<trace_blocks, 128, 0, q>>>(s->dev, Peye, it, px, E, eye_cnt, pix_first, pix_count);
  else k_sppm_eye_trace<false><<<trace_blocks, 128, 0, q>>>(s->dev, Peye, it, px, E, eye_cnt, pix_first, pix_count);
  const uint32_t gather_threads = pix_count * SPPM_GATHER_LANES;
  const int gather_blocks = (int)std::min<uint32_t>((gather_threads + 255u) / 256u, (uint32_t)sms * 32u);
  k_sppm_gather<<<gather_blocks, 256, 0, q>>>(Gg, Gc, px, E, sp.alpha, pix_first, pix_count);
  // every pixel needs its k-nearest start once (the first iteration it sees a diffuse surface): a full grid then, stragglers later
  const int knn_blocks = it == 0 ? sms * 16 : sms * 2;
  k_sppm_knn<<<knn_blocks, 128, 0, q>>>(Gg, Gc, px, E, sp.k_global, sp.k_caustic);
  CU_CHECK(ctx, cudaGetLastError());
  return B200RT_OK;
}

// ---- the photon pass through the wavefront queues (kernels: sppm.cuh) ---------------------------------------------
struct SppmPhotonWf {
  void* block = nullptr;
  WfBuffers B{};
  PathMap M{};
  WfCounters* host = nullptr;  // pinned: the queue counters after the last launched wavefront iteration
  unsigned int* n_log = nullptr;  // device: rays in the queue per wavefront iteration
  unsigned int* n_log_host = nullptr;  // pinned
  int launched = 0;   // wavefront iterations queued for the current pass
  int plan = 12;      // wavefront iterations per pass before the tail kernel (adapted from the last pass's queue lengths)
  int grid_ext = 0, grid_shade = 0, sms = 0;
  bool objf = false;
  static const int MAX_ITERS = 4096;
  void release() {
    if (block) cudaFree(block);
    if (n_log) cudaFree(n_log);
    if (host) cudaFreeHost(host);
    if (n_log_host) cudaFreeHost(n_log_host);
    block = nullptr; n_log = nullptr; host = nullptr; n_log_host = nullptr;
  }
};

// scene-dependent part of a photon queue set: bins, kernel variants, launch sizes (cheap; every call)
int sppm_photon_wf_configure(b200rt_ctx* ctx, const b200rt_scene* s, SppmPhotonWf& W) {
  wf_bin_layout(s->n_prims, W.B.n_bins, W.B.bin_mul);
  W.B.split_cnt = nullptr; W.B.split_bin = W.B.n_bins; W.B.split_box = 0;  // (the two-ended bins are the path tracer's shade kernel's)
  W.sms = ctx->prop.multiProcessorCount;
  W.objf = s->dev.has_objf != 0;
  int occ_ext = 1, occ_shade = 1;
  CU_CHECK(ctx, W.objf ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_ext, k_wf_extend<true>, 128, 0) : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_ext, k_wf_extend<false>, 128, 0));
  CU_CHECK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_shade, k_sppm_ph_shade, 128, 0));
  W.grid_ext = W.sms * (occ_ext > 0 ? occ_ext : 1);
  W.grid_shade = W.sms * (occ_shade > 0 ? occ_shade : 1);
  W.plan = env_int("B200RT_SPPM_PHOTON_ITERS", 12, 1, SppmPhotonWf::MAX_ITERS);
  return B200RT_OK;
}
int sppm_photon_wf_alloc(b200rt_ctx* ctx, const b200rt_scene* s, SppmPhotonWf& W, unsigned int n_photons) {
  const size_t P = ((size_t)n_photons + 31u) & ~(size_t)31u;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off += align_up(bytes, 256); return o; };
  size_t o_ray[2], o_st[2];
  for (int k = 0; k < 2; ++k) { o_ray[k] = take(P * 32); o_st[k] = take(P * 32); }
  const size_t o_hit = take(P * 16), o_bc = take(4 * WF_MAX_BINS), o_bo = take(4 * (WF_MAX_BINS + 1)), o_cnt = take(sizeof(WfCounters));
  CU_CHECK(ctx, cudaMalloc(&W.block, off));
  CU_CHECK(ctx, cudaMalloc((void**)&W.n_log, 4 * SppmPhotonWf::MAX_ITERS));
  CU_CHECK(ctx, cudaMallocHost((void**)&W.host, sizeof(WfCounters)));
  CU_CHECK(ctx, cudaMallocHost((void**)&W.n_log_host, 4 * SppmPhotonWf::MAX_ITERS));
  char* b = (char*)W.block;
  W.B.ray = (float4*)(b + o_ray[0]); W.B.ray_n = (float4*)(b + o_ray[1]); W.B.st = (float4*)(b + o_st[0]); W.B.st_n = (float4*)(b + o_st[1]);
  W.B.hit = (uint4*)(b + o_hit); W.B.bin_count = (unsigned int*)(b + o_bc); W.B.bin_off = (unsigned int*)(b + o_bo); W.B.cnt = (WfCounters*)(b + o_cnt);
  W.B.sh_o = W.B.sh_d = W.B.sh_c = nullptr; W.B.sh_seed = nullptr;
  W.B.pool = (unsigned int)P;
  W.M.tiles_x = 0; W.M.per_sample = n_photons; W.M.total = n_photons;
  return sppm_photon_wf_configure(ctx, s, W);
}

// One photon pass on `qp`: `W.plan` wavefront iterations carry the bulk of the photons (the queue shrinks by a third per bounce),
// then ONE launch of the per-thread walk finishes whatever the queue still holds.  The tail is what a photon pass cannot
// shorten: a photon caught inside the glass sphere by total internal reflection survives every roulette (max f = 1,
// material.rs:27-45) and bounces hundreds of times, one dependent traversal after the other — as queue iterations that is
// hundreds of near-empty launches (measured: 500 per pass, 17-30 ms), as one thread's loop a few ms that the NEXT passes'
// bulk overlaps (the host keeps several passes in flight on their own streams).  No host round trip inside a pass.
int sppm_photon_wf_begin(b200rt_ctx* ctx, const b200rt_scene* s, SppmPhotonWf& W, const SppmConst& sp, uint2 key, int it, float4* pg, float4* pc, unsigned int cap,
                         SppmCounters* cnt_dev, unsigned int id_first, cudaStream_t qp) {
  CU_CHECK(ctx, cudaMemsetAsync(W.B.cnt, 0, sizeof(WfCounters), qp));
  W.launched = 0;
  for (int j = 0; j < W.plan; ++j, ++W.launched) {
    std::swap(W.B.ray, W.B.ray_n);
    std::swap(W.B.st, W.B.st_n);
    // launch sizes follow the queue: it shrinks geometrically (a bound, not a promise: the kernels are persistent / grid-stride)
    const double est = (double)W.M.total * std::pow(0.8, (double)j);
    const int want = (int)std::min<double>(est / 128.0 + 1.0, 1e9);
    const int g_ext = std::max(std::min(W.grid_ext, want), W.sms), g_shade = std::max(std::min(W.grid_shade, want), W.sms);
    if (j == 0) k_sppm_ph_generate<<<W.sms * 8, 256, 0, qp>>>(s->dev, W.B, W.M, sp, key, it, id_first);
    k_sppm_ph_control<<<1, WF_CTRL_THREADS, 0, qp>>>(W.B, W.M, W.n_log, j);
    if (W.objf) k_wf_extend<true><<<g_ext, 128, 0, qp>>>(s->dev, W.B, 0.0001f, 1);  // photon_mapper.rs:246
    else k_wf_extend<false><<<g_ext, 128, 0, qp>>>(s->dev, W.B, 0.0001f, 1);
    k_wf_control_shade<<<1, WF_CTRL_THREADS, 0, qp>>>(W.B);
    k_sppm_ph_shade<<<g_shade, 128, 0, qp>>>(s->dev, W.B, key, it, pg, pc, cap, cnt_dev);
  }
  // the survivors of the last shade pass sit in the next set
  if (W.objf) k_sppm_photons<true><<<W.sms, 128, 0, qp>>>(s->dev, sp, key, it, pg, pc, cap, cnt_dev, 0u, 0u, W.B.ray_n, W.B.st_n, &W.B.cnt->n_ext_next);
  else k_sppm_photons<false><<<W.sms, 128, 0, qp>>>(s->dev, sp, key, it, pg, pc, cap, cnt_dev, 0u, 0u, W.B.ray_n, W.B.st_n, &W.B.cnt->n_ext_next);
  CU_CHECK(ctx, cudaGetLastError());
  CU_CHECK(ctx, cudaMemcpyAsync(W.host, W.B.cnt, sizeof(WfCounters), cudaMemcpyDeviceToHost, qp));
  CU_CHECK(ctx, cudaMemcpyAsync(W.n_log_host, W.n_log, 4 * (size_t)W.launched, cudaMemcpyDeviceToHost, qp));
  return B200RT_OK;
}
// after cudaStreamSynchronize(qp) and the read-back of *cnt_host: the rays of the pass (queue iterations + tail), and the
// number of queue iterations the next pass gets — up to where the queue fell below SPPM_TAIL_RAYS
#define SPPM_TAIL_RAYS 2048u
void sppm_photon_wf_finish(SppmPhotonWf& W, SppmCounters* cnt_host) {
  cnt_host->photon_rays += W.host->extend_rays - W.host->dead_rays;  // extend counts dead records as well
  if (std::getenv("B200RT_SPPM_PHOTON_ITERS")) return;
  int used = 0;
  while (used < W.launched && W.n_log_host[used] >= SPPM_TAIL_RAYS) ++used;
  W.plan = std::min(std::max(used == W.launched ? used + 4 : used, 2), 64);
}

// What b200rt_render_sppm keeps on its context between calls of the same shape (the cudaMalloc / cudaFree of these GBs was a
// third of the call's wall time: 116 ms on the host clock against 81 ms of device time).  Taken out of the context for the
// duration of a call and put back at its end; a call of another shape, or a failed one, releases it.
struct SppmCache {
  SppmBuffers B;
  SppmPhotonWf PW[SPPM_MAX_NB];
  unsigned int cap = 0, photons = 0;   // photon buffers: records each, photons per pass (one device: all of them; several: this device's shard)
  unsigned int cap_maps = 0;           // records of each cell-ordered map (= cap on one device, the complete map on several)
  uint32_t n_pix = 0;
  int NB = 0;
  bool mega = false;
  void release() { B.release(); for (int k = 0; k < SPPM_MAX_NB; ++k) PW[k].release(); }
  static void destroy(void* p) { SppmCache* c = (SppmCache*)p; c->release(); delete c; }
};

}  // namespace

extern "C" int b200rt_render_sppm(b200rt_scene* s, const b200rt_camera* cam, const b200rt_params* params, const b200rt_sppm_params* sppm,
                                  float* accum_host, float* pixel_stats_host, b200rt_sppm_stats* stats) {
  if (!s || !cam || !params || !sppm || !accum_host) return B200RT_ERR_INVALID;
  SCENE_ALIVE(s);
  b200rt_ctx* ctx = s->ctx;
  int rc = validate_params(ctx, params);
  if (rc) return rc;
  if (sppm->iterations < 1 || sppm->photons_per_iter < 1 || sppm->k_global < 1 || sppm->k_caustic < 1 || sppm->k_global > SPPM_MAX_K ||
      sppm->k_caustic > SPPM_MAX_K || !(sppm->alpha > 0.0) || !(sppm->light_scale > 0.0) || sppm->photons_per_iter > (1 << 26)) {
    ctx->err = "sppm: iterations, photons_per_iter >= 1, 1 <= k <= 128, alpha > 0, light_scale > 0";
    return B200RT_ERR_INVALID;
  }
  if (s->dev.n_lights < 1) { ctx->err = "sppm: the scene has no marked light (World::new takes a light list, world.rs:15-25)"; return B200RT_ERR_INVALID; }
  CU_CHECK(ctx, cudaSetDevice(ctx->device));
  cudaGetLastError();  // a non-sticky error an earlier, unrelated call left in this thread is not this call's (the launch checks below poll the same state)
  cudaStream_t q = ctx->stream;
  const uint32_t n_pix = (uint32_t)params->width * params->height;
  const unsigned int cap = (unsigned int)sppm->photons_per_iter * 16u;  // stored photons per iteration (~1.5 per emitted photon in a Cornell box)
  const uint32_t max_cells = (uint32_t)SPPM_MAX_RES * SPPM_MAX_RES * SPPM_MAX_RES;
  cudaError_t e = cudaSuccess;
  auto alloc = [&](void** p, size_t bytes) { if (e == cudaSuccess) e = cudaMalloc(p, bytes); };
  // Photon passes in flight: a pass ends in a tail of a few photons that bounce hundreds of times (one dependent traversal after
  // the other: ms), and no pass depends on the pixel statistics (photon_mapper.rs:154-160) — so several run ahead of the eye pass,
  // each on its own stream with its own queues and buffers, and the tail of one hides under the bulk of the next.
  int inflight = std::min(env_int("B200RT_SPPM_INFLIGHT", 7, 1, SPPM_MAX_NB - 1), sppm->iterations);
  const bool photons_mega = env_int("B200RT_SPPM_PHOTONS_MEGA", 0, 0, 1) != 0;  // A/B: the one-thread-per-photon kernel of round 1
  // the buffers: the context's cached set when it has this call's shape (and at least the photon buffers this call wants), else a fresh one
  SppmCache* cache = (SppmCache*)ctx->sppm_cache;
  ctx->sppm_cache = nullptr;
  if (cache && !(cache->cap == cap && cache->cap_maps == cap && cache->n_pix == n_pix && cache->NB >= inflight + 1 && cache->photons == (unsigned int)sppm->photons_per_iter && cache->mega == photons_mega)) {
    SppmCache::destroy(cache);
    cache = nullptr;
  }
  const bool fresh = cache == nullptr;
  if (fresh) {  // as many passes in flight as the free memory holds (each: two photon buffers of `cap` records + its queues)
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) {
      const size_t per_pass = 2 * (size_t)cap * 64 + 160 * (size_t)sppm->photons_per_iter + (1u << 20);
      const size_t base = 2 * (size_t)cap * 48 + 72 * (size_t)n_pix + 16 * (size_t)max_cells + (64u << 20);
      const size_t budget = free_b - free_b / 8;  // leave an eighth
      const size_t fit = budget > base + 2 * per_pass ? (budget - base) / per_pass - 1 : 1;
      if ((size_t)inflight > fit) inflight = (int)std::max<size_t>(fit, 1);
    }
  }
  const int NB = fresh ? inflight + 1 : cache->NB;
  if (fresh) {
    cache = new (std::nothrow) SppmCache();
    if (!cache) return B200RT_ERR_NOMEM;
    cache->cap = cap; cache->cap_maps = cap; cache->n_pix = n_pix; cache->NB = NB; cache->photons = (unsigned int)sppm->photons_per_iter; cache->mega = photons_mega;
  }
  SppmBuffers& B = cache->B;
  SppmPhotonWf* PW = cache->PW;
  if (fresh) {
    for (int k = 0; k < NB; ++k) { alloc((void**)&B.pg[k], (size_t)cap * 64); alloc((void**)&B.pc[k], (size_t)cap * 64); }
    if (e == cudaSuccess) e = B.alloc_maps(cap, n_pix);
    alloc((void**)&B.tile_sum, 4 * 1024); alloc((void**)&B.cell_count, 4 * (size_t)max_cells); alloc((void**)&B.cursor, 4 * (size_t)max_cells);
    alloc((void**)&B.start_g, 4 * ((size_t)max_cells + 1)); alloc((void**)&B.start_c, 4 * ((size_t)max_cells + 1));
    alloc((void**)&B.px_g, 16 * (size_t)n_pix); alloc((void**)&B.px_c, 16 * (size_t)n_pix); alloc((void**)&B.px_n, 8 * (size_t)n_pix);
    alloc((void**)&B.accum, 16 * (size_t)n_pix);
    alloc((void**)&B.cnt, (SPPM_MAX_NB + 1) * sizeof(SppmCounters));  // [k < NB]: the photon pass into buffer k; [SPPM_MAX_NB]: eye + final pass
  }
  if (e != cudaSuccess) { SppmCache::destroy(cache); ctx->err = std::string("sppm: ") + cudaGetErrorString(e); return B200RT_ERR_NOMEM; }
  auto fail = [&](int code) { SppmCache::destroy(cache); cache = nullptr; return code; };
#define SPPM_CHECK(call)                                                                          \
  do {                                                                                            \
    cudaError_t e_ = (call);                                                                      \
    if (e_ != cudaSuccess) { ctx->err = std::string(#call) + ": " + cudaGetErrorString(e_); return fail(B200RT_ERR_CUDA); } \
  } while (0)
  SPPM_CHECK(cudaMemsetAsync(B.px_g, 0, 16 * (size_t)n_pix, q));
  SPPM_CHECK(cudaMemsetAsync(B.px_c, 0, 16 * (size_t)n_pix, q));
  SPPM_CHECK(cudaMemsetAsync(B.px_n, 0, 8 * (size_t)n_pix, q));
  SPPM_CHECK(cudaMemsetAsync(B.accum, 0, 16 * (size_t)n_pix, q));

  b200rt_sppm_stats local;
  std::memset(&local, 0, sizeof(local));
  SppmConst sp;
  sp.iterations = sppm->iterations; sp.photons_per_iter = sppm->photons_per_iter; sp.k_global = sppm->k_global; sp.k_caustic = sppm->k_caustic;
  sp.alpha = sppm->alpha; sp.light_scale = (float)sppm->light_scale;
  RenderConst P = make_render_const(cam, params);
  RenderConst Peye = P;
  Peye.key.y ^= 0x5EE05EE0u;  // the eye pass draws its own jitter, independent of the final pass
  const SppmPixels px{B.px_g, B.px_c, B.px_n};
  const int sms = ctx->prop.multiProcessorCount;
  float ms = 0.f;
  // Two streams: the photon pass of iteration i+1 (stream qp, the other photon buffer) runs under the grid build and
  // the eye pass of iteration i (context stream).  Both kernels follow one photon / one pixel per thread through
  // data-dependent loops and leave most issue slots empty on their own (28 % / 14 % issue-active), so they overlap
  // almost for free.  Photon maps do not depend on the pixel statistics, only the eye pass does (photon_mapper.rs:154-160).
  // blocks per SM (150 / 128 registers per thread let 3 / 4 blocks of 128 threads reside).  Measured at the reference
  // configuration, serial 339 ms; overlapped with a photon grid of 1 / 2 / 3 / 4 blocks per SM: 303 / 276 / 264 / 253 ms
  const int photon_grid = env_int("B200RT_SPPM_PHOTON_GRID", 4, 1, 8);
  // photon pass: through the wavefront queues (default), or the one-thread-per-photon kernel of round 1 (photons_mega, A/B)
  const bool trace = env_int("B200RT_SPPM_TRACE", 0, 0, 1) != 0;
  const char* dump_prefix = std::getenv("B200RT_SPPM_DUMP");
  cudaStream_t qp[SPPM_MAX_NB] = {};
  cudaEvent_t ev_ph[SPPM_MAX_NB][2] = {}, ev_eye = nullptr, ev_grid = nullptr;  // ev_ph, ev_eye: timing only
  SppmCounters* hcnt = nullptr;  // pinned: [k < NB] photon counters, [SPPM_MAX_NB] eye counters
  const int EYE = SPPM_MAX_NB;
  auto fail2 = [&](int code) {
    for (int k = 0; k < SPPM_MAX_NB; ++k) if (qp[k]) { cudaStreamSynchronize(qp[k]); cudaStreamDestroy(qp[k]); }
    cudaStreamSynchronize(q);
    for (int a = 0; a < SPPM_MAX_NB; ++a) for (int b = 0; b < 2; ++b) if (ev_ph[a][b]) cudaEventDestroy(ev_ph[a][b]);
    if (ev_eye) cudaEventDestroy(ev_eye);
    if (ev_grid) cudaEventDestroy(ev_grid);
    if (hcnt) cudaFreeHost(hcnt);
    if (code == B200RT_OK) {  // the buffers stay with the context for the next call of this shape
      if (ctx->sppm_cache && ctx->sppm_cache_free) ctx->sppm_cache_free(ctx->sppm_cache);
      ctx->sppm_cache = cache; ctx->sppm_cache_free = &SppmCache::destroy;
      return code;
    }
    return fail(code);
  };
#undef SPPM_CHECK
#define SPPM_CHECK(call)                                                                          \
  do {                                                                                            \
    cudaError_t e_ = (call);                                                                      \
    if (e_ != cudaSuccess) { ctx->err = std::string(#call) + ": " + cudaGetErrorString(e_); return fail2(B200RT_ERR_CUDA); } \
  } while (0)
  for (int k = 0; k < NB; ++k) {
    SPPM_CHECK(cudaStreamCreateWithFlags(&qp[k], cudaStreamNonBlocking));
    for (int b = 0; b < 2; ++b) SPPM_CHECK(cudaEventCreate(&ev_ph[k][b]));
    if (!photons_mega && (rc = fresh ? sppm_photon_wf_alloc(ctx, s, PW[k], (unsigned int)sp.photons_per_iter) : sppm_photon_wf_configure(ctx, s, PW[k])) != 0) return fail2(rc);
  }
  SPPM_CHECK(cudaEventCreate(&ev_eye));
  SPPM_CHECK(cudaEventCreate(&ev_grid));
  SPPM_CHECK(cudaMallocHost(&hcnt, (SPPM_MAX_NB + 1) * sizeof(SppmCounters)));
  std::memset(hcnt, 0, (SPPM_MAX_NB + 1) * sizeof(SppmCounters));
  SPPM_CHECK(cudaMemsetAsync(B.cnt + EYE, 0, sizeof(SppmCounters), q));
  SPPM_CHECK(cudaEventRecord(ctx->ev[0], q));
  for (int k = 0; k < NB; ++k) SPPM_CHECK(cudaStreamWaitEvent(qp[k], ctx->ev[0], 0));
  auto launch_photons = [&](int it) -> cudaError_t {  // on stream / into buffer it % NB
    const int k = it % NB;
    SppmCounters h0;
    std::memset(&h0, 0, sizeof(h0));
    for (int a = 0; a < 3; ++a) { h0.lo[a] = 0xffffffffu; h0.hi[a] = 0u; }
    hcnt[k] = h0;
    cudaError_t e = cudaMemcpyAsync(B.cnt + k, hcnt + k, sizeof(SppmCounters), cudaMemcpyHostToDevice, qp[k]);
    if (e == cudaSuccess) e = cudaEventRecord(ev_ph[k][0], qp[k]);
    if (e != cudaSuccess) return e;
    if (!photons_mega) {
      if (sppm_photon_wf_begin(ctx, s, PW[k], sp, P.key, it, B.pg[k], B.pc[k], cap, B.cnt + k, 0u, qp[k])) return cudaErrorUnknown;
    } else if (s->dev.has_objf) k_sppm_photons<true><<<sms * photon_grid, 128, 0, qp[k]>>>(s->dev, sp, P.key, it, B.pg[k], B.pc[k], cap, B.cnt + k, 0u, (unsigned int)sp.photons_per_iter, nullptr, nullptr, nullptr);
    else k_sppm_photons<false><<<sms * photon_grid, 128, 0, qp[k]>>>(s->dev, sp, P.key, it, B.pg[k], B.pc[k], cap, B.cnt + k, 0u, (unsigned int)sp.photons_per_iter, nullptr, nullptr, nullptr);
    e = cudaEventRecord(ev_ph[k][1], qp[k]);
    if (e == cudaSuccess) e = cudaMemcpyAsync(hcnt + k, B.cnt + k, sizeof(SppmCounters), cudaMemcpyDeviceToHost, qp[k]);
    return e;
  };
  for (int it = 0; it < inflight && it < sp.iterations; ++it) SPPM_CHECK(launch_photons(it));
  for (int it = 0; it < sp.iterations; ++it) {
    const int k = it % NB;
    SPPM_CHECK(cudaStreamSynchronize(qp[k]));  // photon map of this iteration is complete, its counters are on the host
    if (!photons_mega) {
      sppm_photon_wf_finish(PW[k], hcnt + k);
      for (int j = 0; j < NB; ++j) PW[j].plan = PW[k].plan;  // every later pass: queue iterations up to where this one's queue ran low
    }
    const SppmCounters h = hcnt[k];
    if (cudaEventElapsedTime(&ms, ev_ph[k][0], ev_ph[k][1]) == cudaSuccess) local.ms_photons += ms;
    if (h.overflow) { ctx->err = "sppm: photon buffer overflow (more than 16 stored photons per emitted photon)"; return fail2(B200RT_ERR_NOMEM); }
    float lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
    if (h.n_global) for (int a = 0; a < 3; ++a) { lo[a] = sppm_unord(h.lo[a]); hi[a] = sppm_unord(h.hi[a]); }
    SppmGrid Gg = sppm_make_grid(B.start_g, h.n_global, lo, hi);
    SppmGrid Gc = sppm_make_grid(B.start_c, h.n_caustic, lo, hi);
    SPPM_CHECK(cudaEventRecord(ctx->sev[0], q));
    if ((rc = sppm_build_grid(ctx, B, Gg, sppm_one_seg(B.pg[k], h.n_global), B.start_g, B.sg)) != 0) return fail2(rc);
    if ((rc = sppm_build_grid(ctx, B, Gc, sppm_one_seg(B.pc[k], h.n_caustic), B.start_c, B.sc)) != 0) return fail2(rc);
    SPPM_CHECK(cudaEventRecord(ev_grid, q));
    // the eye pass is queued BEFORE the next photon pass: that one is ~60 launches (0.2 ms of host time), which would otherwise
    // leave the context stream idle between the grid build and the eye pass of every iteration
    if ((rc = sppm_eye_pass(ctx, s, Peye, sp, it, Gg, Gc, px, B, B.cnt + EYE, 0u, n_pix)) != 0) return fail2(rc);
    SPPM_CHECK(cudaEventRecord(ev_eye, q));
    if (it + inflight < sp.iterations) {
      // pass it + inflight writes the buffer iteration it - 1 used: its grid build read it and has finished (the host
      // synchronised the context stream at the end of that iteration).  It starts after THIS iteration's grid build:
      // launched earlier, its blocks fill every SM while the dozen small grid kernels are still trickling through their
      // stream, and each of those then waits for photon blocks to retire (grid 0.96 ms instead of 0.10 ms per iteration)
      SPPM_CHECK(cudaStreamWaitEvent(qp[(it + inflight) % NB], ev_grid, 0));
      SPPM_CHECK(launch_photons(it + inflight));
    }
    // the grid tables (cell_count, cursor, start, sg, sc) are single-buffered and the stage events are reused:
    // the host lets the eye pass finish before it builds the next grid — the next photon passes are already running
    SPPM_CHECK(cudaStreamSynchronize(q));
    if (dump_prefix && *dump_prefix && it < 4) {  // test hook: this iteration's cell-ordered maps and visible points, for a brute-force check of the look-ups
      std::vector<float4> buf;
      const std::string path = std::string(dump_prefix) + ".it" + std::to_string(it) + ".bin";
      FILE* f = std::fopen(path.c_str(), "wb");
      if (f) {
        const uint32_t head[4] = {h.n_global, h.n_caustic, n_pix, 0u};
        std::fwrite(head, 4, 4, f);
        auto put = [&](const float4* dev, size_t n) {
          buf.resize(n);
          if (n && cudaMemcpy(buf.data(), dev, 16 * n, cudaMemcpyDeviceToHost) == cudaSuccess) std::fwrite(buf.data(), 16, n, f);
        };
        put(Gg.pos, h.n_global); put(Gg.attr, 2 * (size_t)h.n_global); put(Gc.pos, h.n_caustic); put(Gc.attr, 2 * (size_t)h.n_caustic);
        put(B.vp, n_pix); put(B.vb, n_pix);
        std::fclose(f);
      }
    }
    float ms_g = 0.f, ms_e = 0.f;
    if (cudaEventElapsedTime(&ms_g, ctx->sev[0], ev_grid) == cudaSuccess) local.ms_grid += ms_g;
    if (cudaEventElapsedTime(&ms_e, ev_grid, ev_eye) == cudaSuccess) local.ms_eye += ms_e;
    if (trace) std::fprintf(stderr, "sppm it %d: photons %.3f ms (%u global, %u caustic; %d queue iterations), grid %.3f ms (res %d / %d), eye %.3f ms\n", it, ms, h.n_global, h.n_caustic,
                            photons_mega ? 0 : PW[k].launched, ms_g, Gg.res, Gc.res, ms_e);
    local.global_photons += h.n_global; local.caustic_photons += h.n_caustic;
    local.photon_rays += h.photon_rays;
  }
  // final pass: capture_image over the SPPM integrator (photon_mapper.rs:327-365)
  const PathMap M = make_path_map(params->width, params->height, params->spp_count);
  const float n_emitted = (float)((double)sp.iterations * (double)sp.photons_per_iter);  // photon_mapper.rs:229-230
  SPPM_CHECK(cudaEventRecord(ctx->sev[0], q));
  if (M.total > 0) {
    if (s->dev.has_objf) k_sppm_final<true><<<sms * 8, 128, 0, q>>>(s->dev, P, M, px, n_emitted, B.accum, B.cnt + EYE);
    else k_sppm_final<false><<<sms * 8, 128, 0, q>>>(s->dev, P, M, px, n_emitted, B.accum, B.cnt + EYE);
    k_add_count<<<(n_pix + 255) / 256, 256, 0, q>>>(B.accum, (int)n_pix, (float)params->spp_count);
  }
  SPPM_CHECK(cudaEventRecord(ctx->sev[1], q));
  SPPM_CHECK(cudaEventRecord(ctx->ev[1], q));
  SPPM_CHECK(cudaMemcpyAsync(hcnt + EYE, B.cnt + EYE, sizeof(SppmCounters), cudaMemcpyDeviceToHost, q));
  SPPM_CHECK(cudaMemcpyAsync(accum_host, B.accum, 16 * (size_t)n_pix, cudaMemcpyDeviceToHost, q));
  std::vector<float4> hg, hc;
  std::vector<uint2> hn;
  if (pixel_stats_host) {
    hg.resize(n_pix); hc.resize(n_pix); hn.resize(n_pix);
    SPPM_CHECK(cudaMemcpyAsync(hg.data(), B.px_g, 16 * (size_t)n_pix, cudaMemcpyDeviceToHost, q));
    SPPM_CHECK(cudaMemcpyAsync(hc.data(), B.px_c, 16 * (size_t)n_pix, cudaMemcpyDeviceToHost, q));
    SPPM_CHECK(cudaMemcpyAsync(hn.data(), B.px_n, 8 * (size_t)n_pix, cudaMemcpyDeviceToHost, q));
  }
  SPPM_CHECK(cudaStreamSynchronize(q));
  SPPM_CHECK(cudaGetLastError());
  if (pixel_stats_host)
    for (uint32_t i = 0; i < n_pix; ++i) {
      float* o = pixel_stats_host + 10 * (size_t)i;
      o[0] = hg[i].x; o[1] = hg[i].y; o[2] = hg[i].z; o[3] = hg[i].w; o[4] = (float)hn[i].x;
      o[5] = hc[i].x; o[6] = hc[i].y; o[7] = hc[i].z; o[8] = hc[i].w; o[9] = (float)hn[i].y;
    }
  local.eye_rays = hcnt[EYE].eye_rays;  // eye passes + final pass
  if (cudaEventElapsedTime(&ms, ctx->sev[0], ctx->sev[1]) == cudaSuccess) local.ms_final = ms;
  if (cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]) == cudaSuccess) local.ms_total = ms;
#undef SPPM_CHECK
  fail2(B200RT_OK);  // releases the streams, the events and the pinned counters; the device buffers go back to the context
  if (stats) *stats = local;
  return B200RT_OK;
}
