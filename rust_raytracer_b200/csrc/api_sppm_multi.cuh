// api_sppm_multi.cuh — the SPPM integrator (photon_mapper.rs:139-236, 327-365) over several GPUs of one node, inside the
// C ABI.  SURVEY.md section 8(e): "SPPM would need a photon all-gather per iteration".  Decomposition, per iteration:
//   * photons are sharded by id: device g emits and traces ids [g n/G, (g+1) n/G) into its own emission-ordered buffers
//     (every draw is keyed by the GLOBAL photon id, so the set of stored photons is the one a single device stores);
//   * the all-gather of the two photon maps is FUSED into the grid build: each device's counting sort reads all G shards in
//     place — the peers' over NVLink peer mappings (SppmSegs) — and writes the complete, cell-ordered map into local memory;
//     no staging copy, no collective call;
//   * pixels are sharded by index: device g owns the SPPM statistics of pixels [g P/G, (g+1) P/G) and runs the eye pass
//     (SPPM::update) for them against its local copy of the complete map.
// After the last iteration the per-pixel statistics are gathered once (peer copies, 40 B per pixel), the final pass
// (capture_image over SPPMIntegrator::sample_ray) is sharded by sample index like b200rt_render_multi, and the root adds the
// accumulation buffers with the same peer-memory kernel (k_reduce_peers).
// One host thread per device; the threads meet at a barrier once per iteration (after the photon passes) — the photon pass of
// iteration i+1 still overlaps the eye pass of iteration i on every device, as in b200rt_render_sppm.
// Several contexts on ONE device are accepted too (that is how a one-GPU box tests this entry point).
// Included by api.cu after api_sppm.cuh and api_multi.cuh.
#pragma once
#include <atomic>
#include <condition_variable>

namespace {

struct HostBarrier {  // C++17: no std::barrier
  std::mutex m;
  std::condition_variable cv;
  int n, waiting = 0;
  unsigned long long generation = 0;
  explicit HostBarrier(int n_) : n(n_) {}
  void wait() {
    std::unique_lock<std::mutex> lock(m);
    const unsigned long long gen = generation;
    if (++waiting == n) { waiting = 0; ++generation; cv.notify_all(); }
    else cv.wait(lock, [&] { return generation != gen; });
  }
};

struct SppmShared {  // what the device threads publish to each other
  int n_dev = 0;
  std::atomic<int> failed{0};
  std::mutex err_mutex;
  std::string err;
  int err_code = B200RT_OK;
  // per device: photon buffers of both parities, the per-pixel statistics, the accumulation buffer
  float4* pg[B200RT_MULTI_MAX_DEV][SPPM_MAX_NB];
  float4* pc[B200RT_MULTI_MAX_DEV][SPPM_MAX_NB];
  SppmPixels px[B200RT_MULTI_MAX_DEV];
  float* accum[B200RT_MULTI_MAX_DEV];
  // per device and parity: what the photon pass stored
  SppmCounters counts[B200RT_MULTI_MAX_DEV][SPPM_MAX_NB];
  int inflight = 1;  // photon passes ahead of the eye pass (the same on every device)
  // totals
  std::atomic<unsigned long long> global_photons{0}, caustic_photons{0}, photon_rays{0}, eye_rays{0};
  void fail(int code, const std::string& what) {
    std::lock_guard<std::mutex> lock(err_mutex);
    if (!failed.exchange(1)) { err = what; err_code = code; }
  }
};

void sppm_split(uint32_t total, int rank, int world, uint32_t& first, uint32_t& count) {  // contiguous, remainder to the low ranks
  const uint32_t base = total / (uint32_t)world, rem = total % (uint32_t)world;
  count = base + ((uint32_t)rank < rem ? 1u : 0u);
  first = (uint32_t)rank * base + ((uint32_t)rank < rem ? (uint32_t)rank : rem);
}

// the work of one device
void sppm_multi_worker(int g, SppmShared& sh, HostBarrier& bar, b200rt_scene* const* scenes, const b200rt_camera* cam, const b200rt_params* params,
                       const b200rt_sppm_params* sppm) {
  b200rt_scene* s = scenes[g];
  b200rt_ctx* ctx = s->ctx;
  const int G = sh.n_dev;
  const uint32_t n_pix = (uint32_t)params->width * params->height;
  uint32_t id_first, id_count, pix_first, pix_count;
  sppm_split((uint32_t)sppm->photons_per_iter, g, G, id_first, id_count);
  sppm_split(n_pix, g, G, pix_first, pix_count);
  const unsigned int cap_total = (unsigned int)sppm->photons_per_iter * 16u;  // the complete map (cell order) lives on every device
  const unsigned int cap_shard = (id_count + 1u) * 16u;
  const uint32_t max_cells = (uint32_t)SPPM_MAX_RES * SPPM_MAX_RES * SPPM_MAX_RES;
  cudaStream_t q = ctx->stream, qp[SPPM_MAX_NB] = {};
  const int inflight = sh.inflight, EYE = SPPM_MAX_NB;  // as in b200rt_render_sppm: buffer / stream / queues k = pass % NB
  const bool photons_mega = env_int("B200RT_SPPM_PHOTONS_MEGA", 0, 0, 1) != 0;  // as in b200rt_render_sppm
  // this device's buffers: its context's cached set when it has this call's shape, else a fresh one (b200rt_render_sppm's rule)
  SppmCache* cache = (SppmCache*)ctx->sppm_cache;
  ctx->sppm_cache = nullptr;
  if (cache && !(cache->cap == cap_shard && cache->cap_maps == cap_total && cache->n_pix == n_pix && cache->NB == inflight + 1 && cache->photons == id_count && cache->mega == photons_mega)) {  // NB: the same on every device (peers index each other's buffers by pass % NB)
    SppmCache::destroy(cache);
    cache = nullptr;
  }
  const bool fresh = cache == nullptr;
  SppmCache local_cache;  // stands in when the allocation of the cache object itself fails (reported below)
  if (fresh) {
    cache = new (std::nothrow) SppmCache();
    if (cache) { cache->cap = cap_shard; cache->cap_maps = cap_total; cache->n_pix = n_pix; cache->NB = inflight + 1; cache->photons = id_count; cache->mega = photons_mega; }
  }
  const int NB = cache ? cache->NB : inflight + 1;
  SppmBuffers& B = cache ? cache->B : local_cache.B;
  SppmPhotonWf* PW = cache ? cache->PW : local_cache.PW;
  cudaEvent_t ev_grid = nullptr;
  SppmCounters* hcnt = nullptr;
  auto ok = [&](cudaError_t e, const char* what) {
    if (e == cudaSuccess) return true;
    sh.fail(e == cudaErrorMemoryAllocation ? B200RT_ERR_NOMEM : B200RT_ERR_CUDA, std::string("render_sppm_multi, device ") + std::to_string(ctx->device) + ": " + what + ": " + cudaGetErrorString(e));
    return false;
  };
  // ---- set-up: peer mappings, buffers
  bool good = ok(cudaSetDevice(ctx->device), "cudaSetDevice");
  cudaGetLastError();  // as in b200rt_render_sppm: stale non-sticky errors are not this call's
  if (!cache) { sh.fail(B200RT_ERR_NOMEM, "render_sppm_multi: out of host memory"); good = false; }
  for (int h = 0; good && h < G; ++h) {
    const int peer = scenes[h]->ctx->device;
    if (peer == ctx->device) continue;
    int can = 0;
    cudaDeviceCanAccessPeer(&can, ctx->device, peer);
    if (!can) { sh.fail(B200RT_ERR_CUDA, "render_sppm_multi: no peer access from device " + std::to_string(ctx->device) + " to " + std::to_string(peer) + " (the photon all-gather reads peer memory)"); good = false; break; }
    const cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) good = ok(e, "cudaDeviceEnablePeerAccess");
    cudaGetLastError();
  }
  if (good) {
    cudaError_t e = cudaSuccess;
    auto alloc = [&](void** p, size_t bytes) { if (e == cudaSuccess) e = cudaMalloc(p, bytes); };
    if (fresh) {
      for (int k = 0; k < NB; ++k) { alloc((void**)&B.pg[k], (size_t)cap_shard * 64); alloc((void**)&B.pc[k], (size_t)cap_shard * 64); }
      if (e == cudaSuccess) e = B.alloc_maps(cap_total, n_pix);
      alloc((void**)&B.tile_sum, 4 * 1024); alloc((void**)&B.cell_count, 4 * (size_t)max_cells); alloc((void**)&B.cursor, 4 * (size_t)max_cells);
      alloc((void**)&B.start_g, 4 * ((size_t)max_cells + 1)); alloc((void**)&B.start_c, 4 * ((size_t)max_cells + 1));
      alloc((void**)&B.px_g, 16 * (size_t)n_pix); alloc((void**)&B.px_c, 16 * (size_t)n_pix); alloc((void**)&B.px_n, 8 * (size_t)n_pix);
      alloc((void**)&B.accum, 16 * (size_t)n_pix);
      alloc((void**)&B.cnt, (SPPM_MAX_NB + 1) * sizeof(SppmCounters));
    }
    good = ok(e, "cudaMalloc");
    for (int k = 0; good && k < NB; ++k) good = ok(cudaStreamCreateWithFlags(&qp[k], cudaStreamNonBlocking), "cudaStreamCreate");
    if (good) good = ok(cudaEventCreateWithFlags(&ev_grid, cudaEventDisableTiming), "cudaEventCreate") &&
                     ok(cudaMallocHost(&hcnt, (SPPM_MAX_NB + 1) * sizeof(SppmCounters)), "cudaMallocHost");
    if (good) good = ok(cudaMemsetAsync(B.px_g, 0, 16 * (size_t)n_pix, q), "memset") && ok(cudaMemsetAsync(B.px_c, 0, 16 * (size_t)n_pix, q), "memset") &&
                     ok(cudaMemsetAsync(B.px_n, 0, 8 * (size_t)n_pix, q), "memset") && ok(cudaMemsetAsync(B.accum, 0, 16 * (size_t)n_pix, q), "memset") &&
                     ok(cudaMemsetAsync(B.cnt + EYE, 0, sizeof(SppmCounters), q), "memset") && ok(cudaStreamSynchronize(q), "sync");
  }
  for (int k = 0; k < NB; ++k) { sh.pg[g][k] = B.pg[k]; sh.pc[g][k] = B.pc[k]; }
  sh.px[g] = SppmPixels{B.px_g, B.px_c, B.px_n};
  sh.accum[g] = B.accum;
  bar.wait();  // every device has published its buffers (or failed)

  SppmConst sp;
  sp.iterations = sppm->iterations; sp.photons_per_iter = sppm->photons_per_iter; sp.k_global = sppm->k_global; sp.k_caustic = sppm->k_caustic;
  sp.alpha = sppm->alpha; sp.light_scale = (float)sppm->light_scale;
  const RenderConst P = make_render_const(cam, params);
  RenderConst Peye = P;
  Peye.key.y ^= 0x5EE05EE0u;  // as in b200rt_render_sppm
  const int sms = ctx->prop.multiProcessorCount;
  const int photon_grid = env_int("B200RT_SPPM_PHOTON_GRID", 4, 1, 8);
  for (int k = 0; good && k < NB; ++k)
    if (!photons_mega && id_count && (fresh ? sppm_photon_wf_alloc(ctx, s, PW[k], id_count) : sppm_photon_wf_configure(ctx, s, PW[k]))) { sh.fail(B200RT_ERR_NOMEM, "render_sppm_multi: photon queues: " + ctx->err); good = false; }
  auto launch_photons = [&](int it) -> bool {  // on stream / into the buffers it % NB
    const int k = it % NB;
    cudaStream_t qpk = qp[k];
    SppmCounters h0;
    std::memset(&h0, 0, sizeof(h0));
    for (int a = 0; a < 3; ++a) { h0.lo[a] = 0xffffffffu; h0.hi[a] = 0u; }
    hcnt[k] = h0;
    if (!ok(cudaMemcpyAsync(B.cnt + k, hcnt + k, sizeof(SppmCounters), cudaMemcpyHostToDevice, qpk), "memcpy")) return false;
    if (id_count && !photons_mega) {
      if (sppm_photon_wf_begin(ctx, s, PW[k], sp, P.key, it, B.pg[k], B.pc[k], cap_shard, B.cnt + k, id_first, qpk)) { sh.fail(B200RT_ERR_CUDA, "render_sppm_multi: photon pass: " + ctx->err); return false; }
    } else if (id_count) {
      if (s->dev.has_objf) k_sppm_photons<true><<<sms * photon_grid, 128, 0, qpk>>>(s->dev, sp, P.key, it, B.pg[k], B.pc[k], cap_shard, B.cnt + k, id_first, id_count, nullptr, nullptr, nullptr);
      else k_sppm_photons<false><<<sms * photon_grid, 128, 0, qpk>>>(s->dev, sp, P.key, it, B.pg[k], B.pc[k], cap_shard, B.cnt + k, id_first, id_count, nullptr, nullptr, nullptr);
    }
    return ok(cudaGetLastError(), "k_sppm_photons") && ok(cudaMemcpyAsync(hcnt + k, B.cnt + k, sizeof(SppmCounters), cudaMemcpyDeviceToHost, qpk), "memcpy");
  };
  for (int it = 0; it < inflight && it < sp.iterations; ++it) if (good && !sh.failed) good = launch_photons(it);
  // ---- the SPPM iterations
  for (int it = 0; it < sp.iterations; ++it) {
    const int k = it % NB;
    if (good && !sh.failed) {
      good = ok(cudaStreamSynchronize(qp[k]), "photon pass");  // this device's shard of the map is complete, its counters are on the host
      if (good && id_count && !photons_mega) {
        sppm_photon_wf_finish(PW[k], hcnt + k);
        for (int j = 0; j < NB; ++j) PW[j].plan = PW[k].plan;
      }
      if (good && hcnt[k].overflow) { sh.fail(B200RT_ERR_NOMEM, "render_sppm_multi: photon buffer overflow (more than 16 stored photons per emitted photon)"); good = false; }
      sh.counts[g][k] = hcnt[k];
    }
    bar.wait();  // A: every shard of iteration `it` is complete and its size and bounds are published
    if (!good || sh.failed) continue;
    // the complete map = the concatenation of the shards; bounds = union of the shards' bounds
    SppmSegs seg_g, seg_c;
    std::memset(&seg_g, 0, sizeof(seg_g)); std::memset(&seg_c, 0, sizeof(seg_c));
    unsigned int lo_u[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, hi_u[3] = {0u, 0u, 0u};
    for (int h = 0; h < G; ++h) {
      const SppmCounters& c = sh.counts[h][k];
      seg_g.p[h] = sh.pg[h][k]; seg_g.start[h + 1] = seg_g.start[h] + c.n_global;
      seg_c.p[h] = sh.pc[h][k]; seg_c.start[h + 1] = seg_c.start[h] + c.n_caustic;
      if (c.n_global) for (int a = 0; a < 3; ++a) { lo_u[a] = std::min(lo_u[a], c.lo[a]); hi_u[a] = std::max(hi_u[a], c.hi[a]); }
    }
    seg_g.n = seg_c.n = G;
    const uint32_t n_g = seg_g.start[G], n_c = seg_c.start[G];
    float lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
    if (n_g) for (int a = 0; a < 3; ++a) { lo[a] = sppm_unord(lo_u[a]); hi[a] = sppm_unord(hi_u[a]); }
    SppmGrid Gg = sppm_make_grid(B.start_g, n_g, lo, hi);
    SppmGrid Gc = sppm_make_grid(B.start_c, n_c, lo, hi);
    // counting sort over all shards, peers' read in place over NVLink: the all-gather is this kernel pair's input side
    if (sppm_build_grid(ctx, B, Gg, seg_g, B.start_g, B.sg) || sppm_build_grid(ctx, B, Gc, seg_c, B.start_c, B.sc)) {
      sh.fail(B200RT_ERR_CUDA, "render_sppm_multi: grid build: " + ctx->err); good = false; continue;
    }
    good = ok(cudaEventRecord(ev_grid, q), "event");
    // eye pass first, then the ~60 launches of the next photon pass (as in b200rt_render_sppm)
    if (good && pix_count) {
      const SppmPixels px = sh.px[g];
      if (sppm_eye_pass(ctx, s, Peye, sp, it, Gg, Gc, px, B, B.cnt + EYE, pix_first, pix_count)) { sh.fail(B200RT_ERR_CUDA, "render_sppm_multi: eye pass: " + ctx->err); good = false; }
    }
    if (good && it + inflight < sp.iterations) {
      // pass it + inflight writes the buffers (it - 1) % NB, which every peer's grid build of iteration it - 1 has finished
      // reading: each thread synchronised its stream at the end of that iteration, before it arrived at barrier A of this one.
      // It starts after the local grid build (see b200rt_render_sppm: photon blocks starve the small grid kernels).
      good = ok(cudaStreamWaitEvent(qp[(it + inflight) % NB], ev_grid, 0), "wait") && launch_photons(it + inflight);
    }
    // grid tables are single-buffered, and the peers may not overwrite the shards this grid build reads before it is done
    if (good) good = ok(cudaStreamSynchronize(q), "eye pass");
    if (good) { sh.global_photons += hcnt[k].n_global; sh.caustic_photons += hcnt[k].n_caustic; sh.photon_rays += hcnt[k].photon_rays; }
  }
  bar.wait();  // B: every pixel's statistics are final on the device that owns it
  // ---- gather the per-pixel statistics (once), final pass over this device's sample range
  if (good && !sh.failed) {
    for (int h = 0; good && h < G; ++h) {
      if (h == g) continue;
      uint32_t f, c;
      sppm_split(n_pix, h, G, f, c);
      if (!c) continue;
      good = ok(cudaMemcpyAsync(B.px_g + f, sh.px[h].g + f, 16 * (size_t)c, cudaMemcpyDefault, q), "gather") &&
             ok(cudaMemcpyAsync(B.px_c + f, sh.px[h].c + f, 16 * (size_t)c, cudaMemcpyDefault, q), "gather") &&
             ok(cudaMemcpyAsync(B.px_n + f, sh.px[h].n + f, 8 * (size_t)c, cudaMemcpyDefault, q), "gather");
    }
    int spp_first = 0, spp_cnt = 0;
    multi_spp_range(g, G, params->spp_begin, params->spp_count, spp_first, spp_cnt);
    RenderConst Pf = P;
    Pf.spp_begin = spp_first; Pf.spp_count = spp_cnt;
    const PathMap M = make_path_map(params->width, params->height, spp_cnt);
    const float n_emitted = (float)((double)sp.iterations * (double)sp.photons_per_iter);  // photon_mapper.rs:229-230
    if (good && M.total > 0) {
      if (s->dev.has_objf) k_sppm_final<true><<<sms * 8, 128, 0, q>>>(s->dev, Pf, M, sh.px[g], n_emitted, B.accum, B.cnt + EYE);
      else k_sppm_final<false><<<sms * 8, 128, 0, q>>>(s->dev, Pf, M, sh.px[g], n_emitted, B.accum, B.cnt + EYE);
      good = ok(cudaGetLastError(), "k_sppm_final");
    }
    if (good) good = ok(cudaMemcpyAsync(hcnt + EYE, B.cnt + EYE, sizeof(SppmCounters), cudaMemcpyDeviceToHost, q), "memcpy") && ok(cudaStreamSynchronize(q), "final pass");
    if (good) sh.eye_rays += hcnt[EYE].eye_rays;
  }
  bar.wait();  // C: every accumulation buffer is complete; the root reduces and reads them in b200rt_render_sppm_multi
  bar.wait();  // D: the root is done with the peers' buffers
  for (int k = 0; k < SPPM_MAX_NB; ++k) if (qp[k]) { cudaStreamSynchronize(qp[k]); cudaStreamDestroy(qp[k]); }
  if (ev_grid) cudaEventDestroy(ev_grid);
  if (hcnt) cudaFreeHost(hcnt);
  cudaStreamSynchronize(q);
  if (cache && good && !sh.failed) {  // the buffers stay with this device's context for the next call of this shape
    if (ctx->sppm_cache && ctx->sppm_cache_free) ctx->sppm_cache_free(ctx->sppm_cache);
    ctx->sppm_cache = cache; ctx->sppm_cache_free = &SppmCache::destroy;
  } else if (cache) SppmCache::destroy(cache);
  else local_cache.release();
}

}  // namespace

extern "C" int b200rt_render_sppm_multi(int32_t n_dev, b200rt_scene* const* scenes, const b200rt_camera* cam, const b200rt_params* params,
                                        const b200rt_sppm_params* sppm, float* accum_host, float* pixel_stats_host, b200rt_sppm_stats* stats) {
  if (n_dev < 1 || n_dev > B200RT_MULTI_MAX_DEV || n_dev > SPPM_MAX_SEGS || !scenes || !cam || !params || !sppm || !accum_host) return B200RT_ERR_INVALID;
  for (int g = 0; g < n_dev; ++g) { if (!scenes[g]) return B200RT_ERR_INVALID; SCENE_ALIVE(scenes[g]); }
  b200rt_ctx* root = scenes[0]->ctx;
  for (int g = 0; g < n_dev; ++g)
    for (int k = 0; k < g; ++k)
      if (scenes[k]->ctx == scenes[g]->ctx) { root->err = "render_sppm_multi: every scene needs its own context"; return B200RT_ERR_INVALID; }
  int rc = validate_params(root, params);
  if (rc) return rc;
  if (sppm->iterations < 1 || sppm->photons_per_iter < 1 || sppm->k_global < 1 || sppm->k_caustic < 1 || sppm->k_global > SPPM_MAX_K ||
      sppm->k_caustic > SPPM_MAX_K || !(sppm->alpha > 0.0) || !(sppm->light_scale > 0.0) || sppm->photons_per_iter > (1 << 26)) {
    root->err = "sppm: iterations, photons_per_iter >= 1, 1 <= k <= 128, alpha > 0, light_scale > 0";
    return B200RT_ERR_INVALID;
  }
  for (int g = 0; g < n_dev; ++g)
    if (scenes[g]->dev.n_lights < 1) { root->err = "sppm: the scene has no marked light (World::new takes a light list, world.rs:15-25)"; return B200RT_ERR_INVALID; }
  const uint32_t n_pix = (uint32_t)params->width * params->height;
  SppmShared sh;
  sh.n_dev = n_dev;
  sh.inflight = std::min(env_int("B200RT_SPPM_INFLIGHT", 7, 1, SPPM_MAX_NB - 1), sppm->iterations);
  std::memset(sh.pg, 0, sizeof(sh.pg)); std::memset(sh.pc, 0, sizeof(sh.pc)); std::memset(sh.px, 0, sizeof(sh.px)); std::memset(sh.accum, 0, sizeof(sh.accum));
  std::memset(sh.counts, 0, sizeof(sh.counts));
  HostBarrier bar(n_dev + 1);  // the device threads + this one (which does the root's reduce between barriers C and D)
  const auto t0 = std::chrono::steady_clock::now();
  std::vector<std::thread> threads;
  for (int g = 0; g < n_dev; ++g) threads.emplace_back(sppm_multi_worker, g, std::ref(sh), std::ref(bar), scenes, cam, params, sppm);
  bar.wait();                                                 // buffers published
  const auto t1 = std::chrono::steady_clock::now();
  for (int it = 0; it < sppm->iterations; ++it) bar.wait();   // A of every iteration
  bar.wait();                                                 // B
  const auto t2 = std::chrono::steady_clock::now();
  bar.wait();                                                 // C
  cudaError_t e = cudaSuccess;
  if (!sh.failed) {
    e = cudaSetDevice(root->device);
    if (e == cudaSuccess && n_dev > 1) {
      PeerSrc src;
      src.n = n_dev - 1;
      for (int g = 1; g < n_dev; ++g) src.p[g - 1] = (const float4*)sh.accum[g];
      k_reduce_peers<<<root->prop.multiProcessorCount * 4, 256, 0, root->stream>>>((float4*)sh.accum[0], src, (size_t)n_pix);
    }
    if (e == cudaSuccess) k_add_count<<<(n_pix + 255) / 256, 256, 0, root->stream>>>(sh.accum[0], (int)n_pix, (float)params->spp_count);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(accum_host, sh.accum[0], 16 * (size_t)n_pix, cudaMemcpyDeviceToHost, root->stream);
    std::vector<float4> hg, hc;
    std::vector<uint2> hn;
    if (e == cudaSuccess && pixel_stats_host) {  // the root holds every pixel's statistics after the gather
      hg.resize(n_pix); hc.resize(n_pix); hn.resize(n_pix);
      e = cudaMemcpyAsync(hg.data(), sh.px[0].g, 16 * (size_t)n_pix, cudaMemcpyDeviceToHost, root->stream);
      if (e == cudaSuccess) e = cudaMemcpyAsync(hc.data(), sh.px[0].c, 16 * (size_t)n_pix, cudaMemcpyDeviceToHost, root->stream);
      if (e == cudaSuccess) e = cudaMemcpyAsync(hn.data(), sh.px[0].n, 8 * (size_t)n_pix, cudaMemcpyDeviceToHost, root->stream);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(root->stream);
    if (e == cudaSuccess && pixel_stats_host)
      for (uint32_t i = 0; i < n_pix; ++i) {
        float* o = pixel_stats_host + 10 * (size_t)i;
        o[0] = hg[i].x; o[1] = hg[i].y; o[2] = hg[i].z; o[3] = hg[i].w; o[4] = (float)hn[i].x;
        o[5] = hc[i].x; o[6] = hc[i].y; o[7] = hc[i].z; o[8] = hc[i].w; o[9] = (float)hn[i].y;
      }
  }
  bar.wait();                                                 // D
  for (auto& t : threads) t.join();
  cudaSetDevice(root->device);
  if (sh.failed) { root->err = sh.err; return sh.err_code; }
  if (e != cudaSuccess) { root->err = std::string("render_sppm_multi: ") + cudaGetErrorString(e); return B200RT_ERR_CUDA; }
  if (stats) {
    std::memset(stats, 0, sizeof(*stats));
    stats->global_photons = sh.global_photons; stats->caustic_photons = sh.caustic_photons; stats->photon_rays = sh.photon_rays; stats->eye_rays = sh.eye_rays;
    const auto t3 = std::chrono::steady_clock::now();
    stats->ms_total = (float)(std::chrono::duration<double>(t3 - t0).count() * 1e3);  // host clock around the whole call (buffer set-up included)
    stats->ms_eye = (float)(std::chrono::duration<double>(t2 - t1).count() * 1e3);    // host clock of the SPPM iterations (all devices: photon passes, grid builds, eye passes)
    stats->ms_final = (float)(std::chrono::duration<double>(t3 - t2).count() * 1e3);  // host clock of the statistics gather, the final pass, the reduce, the copies out and the release
  }
  return B200RT_OK;
}
