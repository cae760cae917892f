// api_sppm.cuh — SPPMIntegrator::new + capture_image (photon_mapper.rs:139-236, main.rs:51-53) on the device: the
// host loop over the SPPM iterations (photon pass -> grid build for both maps -> eye pass) and the final pass.
// Included by api.cu after api_render.cuh.
#pragma once
#include "sppm.cuh"

namespace {

struct SppmBuffers {
  float4 *pg[2] = {nullptr, nullptr}, *pc[2] = {nullptr, nullptr};  // photons in emission order, double-buffered over the iterations
  float4 *sg = nullptr, *sc = nullptr;  // photons in cell order
  uint32_t *tile_sum = nullptr, *cell_count = nullptr, *start_g = nullptr, *start_c = nullptr, *cursor = nullptr, *index_g = nullptr, *index_c = nullptr;
  float4 *px_g = nullptr, *px_c = nullptr;
  uint2* px_n = nullptr;
  SppmCounters* cnt = nullptr;
  float* accum = nullptr;
  void release() {
    void* all[] = {pg[0], pg[1], pc[0], pc[1], sg, sc, tile_sum, cell_count, start_g, start_c, cursor, index_g, index_c, px_g, px_c, px_n, cnt, accum};
    for (void* p : all) if (p) cudaFree(p);
  }
};

const int SPPM_MAX_RES = 128;

// grid resolution and placement for a map of n photons inside the photon bounding box [lo, hi]
SppmGrid sppm_make_grid(const float4* photons, const uint32_t* start, const uint32_t* index, uint32_t n, const float lo[3], const float hi[3]) {
  SppmGrid G;
  G.photons = photons; G.cell_start = start; G.index = index; G.n = n;
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
int sppm_build_grid(b200rt_ctx* ctx, SppmBuffers& B, SppmGrid& G, const SppmSegs& segs, uint32_t* start, uint32_t* index, float4* sorted) {
  cudaStream_t q = ctx->stream;
  const uint32_t n_cells = (uint32_t)G.res * G.res * G.res;
  CU_CHECK(ctx, cudaMemsetAsync(B.cell_count, 0, sizeof(uint32_t) * n_cells, q));
  const int blocks = ctx->prop.multiProcessorCount * 4;
  if (G.n) k_sppm_grid_count<<<blocks, 256, 0, q>>>(segs, G.n, G.lo, G.inv_cell, G.res, B.cell_count);
  const uint32_t n_tiles = (n_cells + SPPM_SCAN_TILE - 1) / SPPM_SCAN_TILE;  // <= 512 at 128^3 cells
  k_sppm_scan_tile_sums<<<n_tiles, 1024, 0, q>>>(B.cell_count, n_cells, B.tile_sum);
  k_sppm_scan_tile_offsets<<<1, 1024, 0, q>>>(B.tile_sum, n_tiles, start, n_cells);
  k_sppm_scan_cells<<<n_tiles, 1024, 0, q>>>(B.cell_count, n_cells, B.tile_sum, start, B.cursor);
  if (G.n) k_sppm_grid_fill<<<blocks, 256, 0, q>>>(segs, G.n, G.lo, G.inv_cell, G.res, B.cursor, index, sorted);
  CU_CHECK(ctx, cudaGetLastError());
  G.photons = sorted;  // the eye pass reads the cell-ordered copy
  return B200RT_OK;
}

}  // namespace

extern "C" int b200rt_render_sppm(b200rt_scene* s, const b200rt_camera* cam, const b200rt_params* params, const b200rt_sppm_params* sppm,
                                  float* accum_host, float* pixel_stats_host, b200rt_sppm_stats* stats) {
  if (!s || !cam || !params || !sppm || !accum_host) return B200RT_ERR_INVALID;
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
  cudaStream_t q = ctx->stream;
  const uint32_t n_pix = (uint32_t)params->width * params->height;
  const unsigned int cap = (unsigned int)sppm->photons_per_iter * 16u;  // stored photons per iteration (~1.5 per emitted photon in a Cornell box)
  const uint32_t max_cells = (uint32_t)SPPM_MAX_RES * SPPM_MAX_RES * SPPM_MAX_RES;
  SppmBuffers B;
  cudaError_t e = cudaSuccess;
  auto alloc = [&](void** p, size_t bytes) { if (e == cudaSuccess) e = cudaMalloc(p, bytes); };
  for (int k = 0; k < 2; ++k) { alloc((void**)&B.pg[k], (size_t)cap * 64); alloc((void**)&B.pc[k], (size_t)cap * 64); }
  alloc((void**)&B.sg, (size_t)cap * 64); alloc((void**)&B.sc, (size_t)cap * 64);
  alloc((void**)&B.tile_sum, 4 * 1024); alloc((void**)&B.cell_count, 4 * (size_t)max_cells); alloc((void**)&B.cursor, 4 * (size_t)max_cells);
  alloc((void**)&B.start_g, 4 * ((size_t)max_cells + 1)); alloc((void**)&B.start_c, 4 * ((size_t)max_cells + 1));
  alloc((void**)&B.index_g, 4 * (size_t)cap); alloc((void**)&B.index_c, 4 * (size_t)cap);
  alloc((void**)&B.px_g, 16 * (size_t)n_pix); alloc((void**)&B.px_c, 16 * (size_t)n_pix); alloc((void**)&B.px_n, 8 * (size_t)n_pix);
  alloc((void**)&B.accum, 16 * (size_t)n_pix);
  alloc((void**)&B.cnt, 3 * sizeof(SppmCounters));  // [0], [1]: the photon pass of even / odd iterations; [2]: eye + final pass
  if (e != cudaSuccess) { B.release(); ctx->err = std::string("sppm: ") + cudaGetErrorString(e); return B200RT_ERR_NOMEM; }
  auto fail = [&](int code) { B.release(); return code; };
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
  const int photon_grid = env_int("B200RT_SPPM_PHOTON_GRID", 4, 1, 8), eye_grid = env_int("B200RT_SPPM_EYE_GRID", 4, 1, 8);
  const bool trace = env_int("B200RT_SPPM_TRACE", 0, 0, 1) != 0;
  cudaStream_t qp = nullptr;
  cudaEvent_t ev_ph[2][2] = {{nullptr, nullptr}, {nullptr, nullptr}}, ev_eye[2] = {nullptr, nullptr}, ev_grid = nullptr;  // timing only
  SppmCounters* hcnt = nullptr;  // pinned: [0], [1] photon counters, [2] eye counters
  auto fail2 = [&](int code) {
    if (qp) { cudaStreamSynchronize(qp); cudaStreamDestroy(qp); }
    cudaStreamSynchronize(q);
    for (int a = 0; a < 2; ++a) { for (int b = 0; b < 2; ++b) if (ev_ph[a][b]) cudaEventDestroy(ev_ph[a][b]); if (ev_eye[a]) cudaEventDestroy(ev_eye[a]); }
    if (ev_grid) cudaEventDestroy(ev_grid);
    if (hcnt) cudaFreeHost(hcnt);
    return fail(code);
  };
#undef SPPM_CHECK
#define SPPM_CHECK(call)                                                                          \
  do {                                                                                            \
    cudaError_t e_ = (call);                                                                      \
    if (e_ != cudaSuccess) { ctx->err = std::string(#call) + ": " + cudaGetErrorString(e_); return fail2(B200RT_ERR_CUDA); } \
  } while (0)
  SPPM_CHECK(cudaStreamCreateWithFlags(&qp, cudaStreamNonBlocking));
  for (int a = 0; a < 2; ++a) { for (int b = 0; b < 2; ++b) SPPM_CHECK(cudaEventCreate(&ev_ph[a][b])); SPPM_CHECK(cudaEventCreate(&ev_eye[a])); }
  SPPM_CHECK(cudaEventCreate(&ev_grid));
  SPPM_CHECK(cudaMallocHost(&hcnt, 3 * sizeof(SppmCounters)));
  std::memset(hcnt, 0, 3 * sizeof(SppmCounters));
  SPPM_CHECK(cudaMemsetAsync(B.cnt + 2, 0, sizeof(SppmCounters), q));
  SPPM_CHECK(cudaEventRecord(ctx->ev[0], q));
  SPPM_CHECK(cudaStreamWaitEvent(qp, ctx->ev[0], 0));
  auto launch_photons = [&](int it) -> cudaError_t {  // on qp, into buffer it & 1
    const int k = it & 1;
    SppmCounters h0;
    std::memset(&h0, 0, sizeof(h0));
    for (int a = 0; a < 3; ++a) { h0.lo[a] = 0xffffffffu; h0.hi[a] = 0u; }
    hcnt[k] = h0;
    cudaError_t e = cudaMemcpyAsync(B.cnt + k, hcnt + k, sizeof(SppmCounters), cudaMemcpyHostToDevice, qp);
    if (e == cudaSuccess) e = cudaEventRecord(ev_ph[k][0], qp);
    if (e != cudaSuccess) return e;
    if (s->dev.has_objf) k_sppm_photons<true><<<sms * photon_grid, 128, 0, qp>>>(s->dev, sp, P.key, it, B.pg[k], B.pc[k], cap, B.cnt + k, 0u, (unsigned int)sp.photons_per_iter);
    else k_sppm_photons<false><<<sms * photon_grid, 128, 0, qp>>>(s->dev, sp, P.key, it, B.pg[k], B.pc[k], cap, B.cnt + k, 0u, (unsigned int)sp.photons_per_iter);
    e = cudaEventRecord(ev_ph[k][1], qp);
    if (e == cudaSuccess) e = cudaMemcpyAsync(hcnt + k, B.cnt + k, sizeof(SppmCounters), cudaMemcpyDeviceToHost, qp);
    return e;
  };
  SPPM_CHECK(launch_photons(0));
  for (int it = 0; it < sp.iterations; ++it) {
    const int k = it & 1;
    SPPM_CHECK(cudaStreamSynchronize(qp));  // photon map of this iteration is complete, its counters are on the host
    const SppmCounters h = hcnt[k];
    if (cudaEventElapsedTime(&ms, ev_ph[k][0], ev_ph[k][1]) == cudaSuccess) local.ms_photons += ms;
    if (h.overflow) { ctx->err = "sppm: photon buffer overflow (more than 16 stored photons per emitted photon)"; return fail2(B200RT_ERR_NOMEM); }
    float lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
    if (h.n_global) for (int a = 0; a < 3; ++a) { lo[a] = sppm_unord(h.lo[a]); hi[a] = sppm_unord(h.hi[a]); }
    SppmGrid Gg = sppm_make_grid(B.pg[k], B.start_g, B.index_g, h.n_global, lo, hi);
    SppmGrid Gc = sppm_make_grid(B.pc[k], B.start_c, B.index_c, h.n_caustic, lo, hi);
    SPPM_CHECK(cudaEventRecord(ctx->sev[0], q));
    if ((rc = sppm_build_grid(ctx, B, Gg, sppm_one_seg(B.pg[k], h.n_global), B.start_g, B.index_g, B.sg)) != 0) return fail2(rc);
    if ((rc = sppm_build_grid(ctx, B, Gc, sppm_one_seg(B.pc[k], h.n_caustic), B.start_c, B.index_c, B.sc)) != 0) return fail2(rc);
    SPPM_CHECK(cudaEventRecord(ev_grid, q));
    if (it + 1 < sp.iterations) {
      // the next photon pass writes the buffer iteration it - 1 used: its grid build read it and has finished (the host
      // synchronised the context stream at the end of that iteration).  It starts after THIS iteration's grid build:
      // launched earlier, its blocks fill every SM while the dozen small grid kernels are still trickling through their
      // stream, and each of those then waits for photon blocks to retire (grid 0.96 ms instead of 0.10 ms per iteration)
      SPPM_CHECK(cudaStreamWaitEvent(qp, ev_grid, 0));
      SPPM_CHECK(launch_photons(it + 1));
    }
    if (s->dev.has_objf) k_sppm_eye<true><<<sms * eye_grid, 128, 0, q>>>(s->dev, Peye, sp, it, Gg, Gc, px, B.cnt + 2, 0u, n_pix);
    else k_sppm_eye<false><<<sms * eye_grid, 128, 0, q>>>(s->dev, Peye, sp, it, Gg, Gc, px, B.cnt + 2, 0u, n_pix);
    SPPM_CHECK(cudaEventRecord(ev_eye[k], q));
    // the grid tables (cell_count, cursor, start, index, sg, sc) are single-buffered and the stage events are reused:
    // the host lets the eye pass finish before it builds the next grid — the next photon pass is already running
    SPPM_CHECK(cudaStreamSynchronize(q));
    float ms_g = 0.f, ms_e = 0.f;
    if (cudaEventElapsedTime(&ms_g, ctx->sev[0], ev_grid) == cudaSuccess) local.ms_grid += ms_g;
    if (cudaEventElapsedTime(&ms_e, ev_grid, ev_eye[k]) == cudaSuccess) local.ms_eye += ms_e;
    if (trace) std::fprintf(stderr, "sppm it %d: photons %.3f ms (%u global, %u caustic), grid %.3f ms (res %d / %d), eye %.3f ms\n", it, ms, h.n_global, h.n_caustic, ms_g, Gg.res, Gc.res, ms_e);
    local.global_photons += h.n_global; local.caustic_photons += h.n_caustic;
    local.photon_rays += h.photon_rays;
  }
  // final pass: capture_image over the SPPM integrator (photon_mapper.rs:327-365)
  const PathMap M = make_path_map(params->width, params->height, params->spp_count);
  const float n_emitted = (float)((double)sp.iterations * (double)sp.photons_per_iter);  // photon_mapper.rs:229-230
  SPPM_CHECK(cudaEventRecord(ctx->sev[0], q));
  if (M.total > 0) {
    if (s->dev.has_objf) k_sppm_final<true><<<sms * 8, 128, 0, q>>>(s->dev, P, M, px, n_emitted, B.accum, B.cnt + 2);
    else k_sppm_final<false><<<sms * 8, 128, 0, q>>>(s->dev, P, M, px, n_emitted, B.accum, B.cnt + 2);
    k_add_count<<<(n_pix + 255) / 256, 256, 0, q>>>(B.accum, (int)n_pix, (float)params->spp_count);
  }
  SPPM_CHECK(cudaEventRecord(ctx->sev[1], q));
  SPPM_CHECK(cudaEventRecord(ctx->ev[1], q));
  SPPM_CHECK(cudaMemcpyAsync(hcnt + 2, B.cnt + 2, sizeof(SppmCounters), cudaMemcpyDeviceToHost, q));
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
  local.eye_rays = hcnt[2].eye_rays;  // eye passes + final pass
  if (cudaEventElapsedTime(&ms, ctx->sev[0], ctx->sev[1]) == cudaSuccess) local.ms_final = ms;
  if (cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]) == cudaSuccess) local.ms_total = ms;
#undef SPPM_CHECK
  fail2(B200RT_OK);  // releases the stream, the events, the pinned counters and the device buffers
  if (stats) *stats = local;
  return B200RT_OK;
}
