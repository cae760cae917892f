// api.cu — the C ABI of include/b200rt.h: context, scene upload, World::hit batch operator, capture_image
// (wavefront and megakernel pipelines), tonemap.  Every entry point returns a status; nothing throws across the ABI.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "scene_compile.hpp"
#include "wavefront.cuh"

struct b200rt_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaDeviceProp prop;
  std::string err;
  // wavefront pools, one per concurrent pipeline (allocated lazily, reused across renders)
  WfBuffers wf[WF_MAX_PIPES] = {};
  void* wf_block[WF_MAX_PIPES] = {};
  void* wf_shadow_block[WF_MAX_PIPES] = {};
  size_t wf_bytes = 0;
  cudaStream_t pipe_stream[WF_MAX_PIPES] = {};  // [0] unused: pipeline 0 runs on `stream`
  cudaEvent_t fork_ev = nullptr;
  MegaCounters* mega_cnt = nullptr;
  WfCounters* host_cnt = nullptr;  // pinned, WF_MAX_PIPES entries
  cudaEvent_t ev[2] = {nullptr, nullptr};
  // World::new staging: one grow-only pinned buffer and one recycled device block per context, so that rebuilding
  // a scene (the e2e path of bench.py) pays no cudaMallocHost / cudaMalloc / cudaFree (each of which can stall
  // for hundreds of ms on a busy driver)
  void* stage_host = nullptr;
  size_t stage_bytes = 0;
  void* spare_block = nullptr;
  size_t spare_bytes = 0;
  // b200rt_render_sppm's device buffers (photon buffers of the passes in flight, cell-ordered maps, photon queues: GBs) kept
  // between calls of the same shape, like the wavefront pools above; owned through an opaque pointer (api_sppm.cuh)
  void* sppm_cache = nullptr;
  void (*sppm_cache_free)(void*) = nullptr;
  const WfBuffers* static_wf = nullptr;  // set for the duration of b200rt_render_device_static: render_wavefront takes its queues from there
  // the scenes created on this context and not yet destroyed: b200rt_ctx_destroy orphans them (ctx = nullptr), so that a host
  // that releases a context before its scenes (a garbage collector does) gets B200RT_ERR_INVALID from them, not a dangling pointer
  std::mutex scenes_mutex;
  std::vector<struct b200rt_scene*> scenes;
  cudaEvent_t sev[128 * WF_MAX_PIPES] = {};  // stage events per pipeline: 2 alternating sets x up to 8 iterations in flight x 8 (generate, extend, shade, shadow boundaries)
};

struct b200rt_scene {
  b200rt_ctx* ctx = nullptr;  // nullptr once the context has been destroyed (every entry point then returns B200RT_ERR_INVALID)
  int device = 0;
  DevScene dev{};
  void* block = nullptr;  // one allocation holding every table
  size_t block_bytes = 0;
  uint64_t h2d_bytes = 0;
  uint32_t n_nodes = 0, n_prims = 0;
  bool has_dense_box = false;  // scene_compile.cpp dense_box: the wavefront queues keep the rays that cross it apart from the rest
  float dense_lo[3] = {0, 0, 0}, dense_hi[3] = {0, 0, 0};
};

namespace { void multi_forget_device(int device); }  // api_multi.cuh
static thread_local std::string g_global_err;
// a scene whose context has been destroyed
#define SCENE_ALIVE(s)                                                                                              \
  do {                                                                                                               \
    if (!(s)->ctx) { g_global_err = "the scene's context has been destroyed (b200rt_ctx_destroy before b200rt_scene_destroy)"; return B200RT_ERR_INVALID; } \
  } while (0)

#define CU_CHECK(ctx, call)                                                                                      \
  do {                                                                                                           \
    cudaError_t e_ = (call);                                                                                     \
    if (e_ != cudaSuccess) {                                                                                     \
      (ctx)->err = std::string(#call) + ": " + cudaGetErrorString(e_);                                           \
      return B200RT_ERR_CUDA;                                                                                    \
    }                                                                                                            \
  } while (0)

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

#ifdef B200RT_COUNTERS
extern "C" int b200rt_debug_counters(unsigned long long* out, int reset) {  // debug builds only (tools/dbg)
  cudaDeviceSynchronize();
  if (out) cudaMemcpyFromSymbol(out, g_dbg, sizeof(g_dbg));
  if (reset) { static unsigned long long z[64]; cudaMemcpyToSymbol(g_dbg, z, sizeof(z)); }
  return 0;
}
#endif

extern "C" {

int b200rt_abi_version(void) { return B200RT_ABI_VERSION; }
const char* b200rt_global_error(void) { return g_global_err.c_str(); }
const char* b200rt_last_error(const b200rt_ctx* ctx) { return ctx ? ctx->err.c_str() : g_global_err.c_str(); }

int b200rt_device_count(int* count) {
  if (!count) return B200RT_ERR_INVALID;
  cudaError_t e = cudaGetDeviceCount(count);
  if (e != cudaSuccess) { *count = 0; g_global_err = cudaGetErrorString(e); return B200RT_ERR_CUDA; }
  return B200RT_OK;
}

int b200rt_ctx_create(int device, b200rt_ctx** out) {
  if (!out) return B200RT_ERR_INVALID;
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    g_global_err = std::string("no CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") + " (this library has no CPU fallback)";
    return B200RT_ERR_CUDA;
  }
  if (device < 0 || device >= n) { g_global_err = "device index out of range"; return B200RT_ERR_INVALID; }
  b200rt_ctx* c = new (std::nothrow) b200rt_ctx();
  if (!c) return B200RT_ERR_NOMEM;
  c->device = device;
  if ((e = cudaSetDevice(device)) != cudaSuccess || (e = cudaGetDeviceProperties(&c->prop, device)) != cudaSuccess ||
      (e = cudaMalloc(&c->mega_cnt, sizeof(MegaCounters))) != cudaSuccess || (e = cudaMallocHost(&c->host_cnt, sizeof(WfCounters) * WF_MAX_PIPES)) != cudaSuccess ||
      (e = cudaEventCreate(&c->ev[0])) != cudaSuccess || (e = cudaEventCreate(&c->ev[1])) != cudaSuccess ||
      (e = cudaEventCreate(&c->fork_ev)) != cudaSuccess) {
    g_global_err = std::string("ctx_create: ") + cudaGetErrorString(e);
    delete c;
    return B200RT_ERR_CUDA;
  }
  for (int i = 0; i < 128 * WF_MAX_PIPES && e == cudaSuccess; ++i) e = cudaEventCreate(&c->sev[i]);
  if (e != cudaSuccess) {
    g_global_err = std::string("ctx_create: ") + cudaGetErrorString(e);
    delete c;
    return B200RT_ERR_CUDA;
  }
  *out = c;
  return B200RT_OK;
}

int b200rt_ctx_destroy(b200rt_ctx* c) {
  if (!c) return B200RT_OK;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  multi_forget_device(c->device);
  {
    std::lock_guard<std::mutex> lock(c->scenes_mutex);
    for (b200rt_scene* s : c->scenes) s->ctx = nullptr;  // orphaned: b200rt_scene_destroy still frees their tables
    c->scenes.clear();
  }
  for (int k = 0; k < WF_MAX_PIPES; ++k) {
    if (c->wf_block[k]) cudaFree(c->wf_block[k]);
    if (c->wf_shadow_block[k]) cudaFree(c->wf_shadow_block[k]);
    if (c->pipe_stream[k]) cudaStreamDestroy(c->pipe_stream[k]);
  }
  if (c->sppm_cache && c->sppm_cache_free) c->sppm_cache_free(c->sppm_cache);
  if (c->fork_ev) cudaEventDestroy(c->fork_ev);
  if (c->spare_block) cudaFree(c->spare_block);
  if (c->stage_host) cudaFreeHost(c->stage_host);
  if (c->mega_cnt) cudaFree(c->mega_cnt);
  if (c->host_cnt) cudaFreeHost(c->host_cnt);
  if (c->ev[0]) cudaEventDestroy(c->ev[0]);
  if (c->ev[1]) cudaEventDestroy(c->ev[1]);
  for (int i = 0; i < 128 * WF_MAX_PIPES; ++i) if (c->sev[i]) cudaEventDestroy(c->sev[i]);
  delete c;
  return B200RT_OK;
}

int b200rt_ctx_set_stream(b200rt_ctx* c, void* s) {
  if (!c) return B200RT_ERR_INVALID;
  c->stream = (cudaStream_t)s;
  return B200RT_OK;
}

int b200rt_ctx_device_info(const b200rt_ctx* c, int32_t* sm_count, int32_t* clock_khz, int64_t* l2_bytes, char* name, int name_len) {
  if (!c) return B200RT_ERR_INVALID;
  if (sm_count) *sm_count = c->prop.multiProcessorCount;
  if (clock_khz) { int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, c->device); *clock_khz = khz; }
  if (l2_bytes) *l2_bytes = c->prop.l2CacheSize;
  if (name && name_len > 0) { std::snprintf(name, (size_t)name_len, "%s", c->prop.name); }
  return B200RT_OK;
}

// ------------------------------------------------------------------------------------------------ scene
int b200rt_scene_create(b200rt_ctx* ctx, const b200rt_scene_desc* desc, b200rt_scene** out) {
  if (!ctx || !desc || !out) return B200RT_ERR_INVALID;
  *out = nullptr;
  b200rt::CompiledScene cs;
  int rc = b200rt::compile_scene(desc, cs, ctx->err);
  if (rc) return rc;
  CU_CHECK(ctx, cudaSetDevice(ctx->device));
  struct Part { const void* src; size_t bytes; size_t off; };
  Part parts[15] = {{cs.nodes.data(), cs.nodes.size() * 4, 0},           {cs.prim_data.data(), cs.prim_data.size() * 4, 0},
                    {cs.prim_meta.data(), cs.prim_meta.size() * 4, 0},   {cs.tri_normals.data(), cs.tri_normals.size() * 4, 0},
                    {cs.instances.data(), cs.instances.size() * 4, 0},   {cs.materials.data(), cs.materials.size() * 4, 0},
                    {cs.textures.data(), cs.textures.size() * 4, 0},     {cs.images.data(), cs.images.size() * 4, 0},
                    {cs.image_data.data(), cs.image_data.size(), 0},     {cs.perlin_vec.data(), cs.perlin_vec.size() * 4, 0},
                    {cs.perlin_perm.data(), cs.perlin_perm.size() * 4, 0}, {cs.lights.data(), cs.lights.size() * 4, 0},
                    {cs.prim_obj.data(), cs.prim_obj.size() * 4, 0},     {cs.instances_d.data(), cs.instances_d.size() * 8, 0},
                    {cs.inst_eps.data(), cs.inst_eps.size() * 4, 0}};
  size_t total = 0;
  for (auto& p : parts) { p.off = total; total += align_up(std::max<size_t>(p.bytes, 16), 256); }
  // stage everything in one pinned buffer -> one H2D copy
  if (ctx->stage_bytes < total) {
    if (ctx->stage_host) cudaFreeHost(ctx->stage_host);
    ctx->stage_host = nullptr; ctx->stage_bytes = 0;
    CU_CHECK(ctx, cudaMallocHost(&ctx->stage_host, total));
    ctx->stage_bytes = total;
  }
  void* host = ctx->stage_host;
  std::memset(host, 0, total);
  for (auto& p : parts) if (p.bytes) std::memcpy((char*)host + p.off, p.src, p.bytes);
  b200rt_scene* sc = new (std::nothrow) b200rt_scene();
  if (!sc) return B200RT_ERR_NOMEM;
  cudaError_t e = cudaSuccess;
  if (ctx->spare_block && ctx->spare_bytes >= total && ctx->spare_bytes <= 2 * total + (1u << 20)) {
    sc->block = ctx->spare_block; sc->block_bytes = ctx->spare_bytes;
    ctx->spare_block = nullptr; ctx->spare_bytes = 0;
  } else {
    e = cudaMalloc(&sc->block, total);
    sc->block_bytes = total;
  }
  if (e == cudaSuccess) e = cudaMemcpyAsync(sc->block, host, total, cudaMemcpyHostToDevice, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  if (e != cudaSuccess) {
    ctx->err = std::string("scene upload: ") + cudaGetErrorString(e);
    if (sc->block) cudaFree(sc->block);
    delete sc;
    return e == cudaErrorMemoryAllocation ? B200RT_ERR_NOMEM : B200RT_ERR_CUDA;
  }
  char* base = (char*)sc->block;
  sc->ctx = ctx;
  sc->dev.nodes = (const float4*)(base + parts[0].off);
  sc->dev.prim_data = (const float4*)(base + parts[1].off);
  sc->dev.prim_meta = (const int4*)(base + parts[2].off);
  sc->dev.tri_normals = (const float4*)(base + parts[3].off);
  sc->dev.instances = (const float4*)(base + parts[4].off);
  sc->dev.materials = (const float4*)(base + parts[5].off);
  sc->dev.textures = (const float4*)(base + parts[6].off);
  sc->dev.images = (const int4*)(base + parts[7].off);
  sc->dev.image_data = (const uint8_t*)(base + parts[8].off);
  sc->dev.perlin_vec = (const float4*)(base + parts[9].off);
  sc->dev.perlin_perm = (const int*)(base + parts[10].off);
  sc->dev.lights = (const float4*)(base + parts[11].off);
  sc->dev.prim_obj = (const float4*)(base + parts[12].off);
  sc->dev.instances_d = (const double*)(base + parts[13].off);
  sc->dev.inst_eps = (const float*)(base + parts[14].off);
  sc->has_dense_box = cs.has_dense_box;
  for (int a = 0; a < 3; ++a) { sc->dense_lo[a] = cs.dense_lo[a]; sc->dense_hi[a] = cs.dense_hi[a]; }
  sc->dev.n_prims = (int)cs.n_prims; sc->dev.n_bvh_prims = (int)cs.n_bvh_prims; sc->dev.n_nodes = (int)cs.n_nodes; sc->dev.n_lights = (int)cs.n_lights; sc->dev.flat = cs.flat ? 1 : 0; sc->dev.has_objf = cs.has_objf ? 1 : 0;
  sc->h2d_bytes = total; sc->n_nodes = cs.n_nodes; sc->n_prims = cs.n_prims;
  sc->device = ctx->device;
  {
    std::lock_guard<std::mutex> lock(ctx->scenes_mutex);
    ctx->scenes.push_back(sc);
  }
  *out = sc;
  return B200RT_OK;
}

int b200rt_scene_destroy(b200rt_scene* s) {
  if (!s) return B200RT_OK;
  b200rt_ctx* ctx = s->ctx;
  cudaSetDevice(s->device);
  if (!ctx) {  // its context went first: nothing of it is left to wait for or to recycle into
    if (s->block) { cudaDeviceSynchronize(); cudaFree(s->block); }
    delete s;
    return B200RT_OK;
  }
  {
    std::lock_guard<std::mutex> lock(ctx->scenes_mutex);
    ctx->scenes.erase(std::remove(ctx->scenes.begin(), ctx->scenes.end(), s), ctx->scenes.end());
  }
  if (s->block) {
    cudaStreamSynchronize(ctx->stream);  // nothing in flight may still read the tables
    if (!ctx->spare_block) { ctx->spare_block = s->block; ctx->spare_bytes = s->block_bytes; }
    else cudaFree(s->block);
  }
  delete s;
  return B200RT_OK;
}

int b200rt_scene_info(const b200rt_scene* s, uint64_t* h2d_bytes, uint32_t* bvh_nodes, uint32_t* device_prims) {
  if (!s) return B200RT_ERR_INVALID;
  if (h2d_bytes) *h2d_bytes = s->h2d_bytes;
  if (bvh_nodes) *bvh_nodes = s->n_nodes;
  if (device_prims) *device_prims = s->n_prims;
  return B200RT_OK;
}

// ------------------------------------------------------------------------------------------------ World::hit
int b200rt_intersect_batch_device(b200rt_scene* s, const float* o_time, const float* d_tmax, const uint32_t* medium_seed, uint64_t n, float t_min, b200rt_hit* out) {
  if (!s || (n && (!o_time || !d_tmax || !out))) return B200RT_ERR_INVALID;
  if (n == 0) return B200RT_OK;  // empty batch
  SCENE_ALIVE(s);
  b200rt_ctx* ctx = s->ctx;
  CU_CHECK(ctx, cudaSetDevice(ctx->device));
  cudaGetLastError();  // stale non-sticky errors of earlier, unrelated calls are not this launch's
  const int threads = 128;
  unsigned long long want = (n + threads - 1) / threads;
  unsigned long long cap = (unsigned long long)ctx->prop.multiProcessorCount * 64ull;
  int blocks = (int)(want < cap ? want : cap);
  k_intersect_batch<<<blocks, threads, 0, ctx->stream>>>(s->dev, (const float4*)o_time, (const float4*)d_tmax, medium_seed, n, t_min, out);
  CU_CHECK(ctx, cudaGetLastError());
  return B200RT_OK;
}

int b200rt_intersect_batch(b200rt_scene* s, const float* o_time, const float* d_tmax, const uint32_t* medium_seed, uint64_t n, float t_min, b200rt_hit* out) {
  if (!s || (n && (!o_time || !d_tmax || !out))) return B200RT_ERR_INVALID;
  if (n == 0) return B200RT_OK;
  SCENE_ALIVE(s);
  b200rt_ctx* ctx = s->ctx;
  CU_CHECK(ctx, cudaSetDevice(ctx->device));
  float *d_o = nullptr, *d_d = nullptr;
  uint32_t* d_s = nullptr;
  b200rt_hit* d_h = nullptr;
  int rc = B200RT_OK;
  cudaError_t e;
  auto fail = [&](cudaError_t er) { ctx->err = std::string("intersect_batch: ") + cudaGetErrorString(er); rc = er == cudaErrorMemoryAllocation ? B200RT_ERR_NOMEM : B200RT_ERR_CUDA; };
  if ((e = cudaMalloc(&d_o, n * 16)) != cudaSuccess || (e = cudaMalloc(&d_d, n * 16)) != cudaSuccess || (e = cudaMalloc(&d_h, n * sizeof(b200rt_hit))) != cudaSuccess ||
      (medium_seed && (e = cudaMalloc(&d_s, n * 4)) != cudaSuccess)) fail(e);
  if (!rc && ((e = cudaMemcpyAsync(d_o, o_time, n * 16, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess ||
              (e = cudaMemcpyAsync(d_d, d_tmax, n * 16, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess ||
              (medium_seed && (e = cudaMemcpyAsync(d_s, medium_seed, n * 4, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess))) fail(e);
  if (!rc) rc = b200rt_intersect_batch_device(s, d_o, d_d, d_s, n, t_min, d_h);
  if (!rc && ((e = cudaMemcpyAsync(out, d_h, n * sizeof(b200rt_hit), cudaMemcpyDeviceToHost, ctx->stream)) != cudaSuccess || (e = cudaStreamSynchronize(ctx->stream)) != cudaSuccess)) fail(e);
  cudaFree(d_o); cudaFree(d_d); cudaFree(d_s); cudaFree(d_h);
  return rc;
}

}  // extern "C"

#include "api_render.cuh"
#include "api_sppm.cuh"
#include "api_multi.cuh"
#include "api_sppm_multi.cuh"
#include "api_debug.cuh"
#include "api_static.cuh"
