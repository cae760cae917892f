// api_static.cuh — b200rt_render_device_static: Camera::capture_image through the wavefront pipeline WITHOUT any device
// allocation.  No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// The wavefront kernels need scratch: the two ray / state sets, the hit stream, the shadow-request buffers, the bin tables
// and the counters (wavefront.cuh).  b200rt_render_device takes them from a pool it cudaMallocs on first use; this variant
// takes them from ONE file-scope `static __device__` array of B200RT_STATIC_POOL_BYTES (12.9 MB: 2^16 queue positions),
// which every GPU that loads the library holds until the process exits.  In host code the array's name is only a host-side
// stand-in, not a device address: the address comes from cudaGetSymbolAddress, once per device, with that device current.
//
// Every call shares the array and the array keeps the last call's contents: what must start at zero (the counters; the bin
// sizes, by k_wf_control every iteration) is cleared per call, everything else is written before it is read.  Two calls
// that overlap on the GPU (two contexts / streams on one device) RACE on it: serialise them.
// Included by api.cu after api_render.cuh.
#pragma once
#include <mutex>

namespace {

constexpr unsigned int STATIC_POSITIONS = B200RT_STATIC_POOL_POSITIONS;  // queue positions (paths in flight)
constexpr size_t static_align(size_t v) { return (v + 255) / 256 * 256; }
constexpr size_t ST_RAY = static_align((size_t)STATIC_POSITIONS * 32), ST_HIT = static_align((size_t)STATIC_POSITIONS * 16),
                 ST_SH16 = static_align((size_t)STATIC_POSITIONS * 16), ST_SH4 = static_align((size_t)STATIC_POSITIONS * 4),
                 ST_BINS = static_align(4 * (size_t)(WF_MAX_BINS + 1)), ST_CNT = static_align(sizeof(WfCounters));
constexpr size_t STATIC_BYTES = 4 * ST_RAY + ST_HIT + 3 * ST_SH16 + ST_SH4 + 2 * ST_BINS + ST_CNT;
static_assert(STATIC_BYTES <= B200RT_STATIC_POOL_BYTES, "B200RT_STATIC_POOL_BYTES (include/b200rt.h) is smaller than the queue block");

}  // namespace

static __device__ __align__(256) unsigned char g_static_pool[B200RT_STATIC_POOL_BYTES];

namespace {

std::mutex g_static_mutex;
unsigned char* g_static_addr[64] = {};  // device address of g_static_pool, per device ordinal

// set-up, once per device, with that device current
int static_pool_address(b200rt_ctx* ctx, unsigned char** out) {
  std::lock_guard<std::mutex> lock(g_static_mutex);
  if (ctx->device < 0 || ctx->device >= 64) { ctx->err = "render_device_static: device ordinal out of range"; return B200RT_ERR_INVALID; }
  if (!g_static_addr[ctx->device]) {
    unsigned char* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_static_pool);
    if (e != cudaSuccess || !p) { ctx->err = std::string("cudaGetSymbolAddress(g_static_pool): ") + cudaGetErrorString(e); return B200RT_ERR_CUDA; }
    g_static_addr[ctx->device] = p;
  }
  *out = g_static_addr[ctx->device];
  return B200RT_OK;
}

}  // namespace

extern "C" int b200rt_render_device_static(b200rt_scene* s, const b200rt_camera* cam, const b200rt_params* params, float* accum_dev, b200rt_stats* stats) {
  if (!s || !cam || !params || !accum_dev) return B200RT_ERR_INVALID;
  SCENE_ALIVE(s);
  b200rt_ctx* ctx = s->ctx;
  int rc = validate_params(ctx, params);
  if (rc) return rc;
  if (params->pipeline != B200RT_PIPELINE_WAVEFRONT) { ctx->err = "render_device_static: the static queue block serves the wavefront pipeline only"; return B200RT_ERR_INVALID; }
  const unsigned long long paths = (unsigned long long)params->width * (unsigned long long)params->height * (unsigned long long)params->spp_count;
  if (paths > B200RT_STATIC_MAX_PATHS) {
    ctx->err = "render_device_static: " + std::to_string(paths) + " paths in one call exceed the limit of " + std::to_string((unsigned long long)B200RT_STATIC_MAX_PATHS) +
               " (width * height * spp_count; the static queue block holds " + std::to_string(STATIC_POSITIONS) + " paths in flight): split the sample range or use b200rt_render_device";
    return B200RT_ERR_INVALID;
  }
  CU_CHECK(ctx, cudaSetDevice(ctx->device));
  unsigned char* base = nullptr;
  rc = static_pool_address(ctx, &base);
  if (rc) return rc;
  WfBuffers B;
  std::memset(&B, 0, sizeof(B));
  size_t off = 0;
  auto take = [&](size_t bytes) { unsigned char* p = base + off; off += bytes; return p; };
  B.ray = (float4*)take(ST_RAY); B.st = (float4*)take(ST_RAY); B.ray_n = (float4*)take(ST_RAY); B.st_n = (float4*)take(ST_RAY);
  B.hit = (uint4*)take(ST_HIT);
  B.sh_o = (float4*)take(ST_SH16); B.sh_d = (float4*)take(ST_SH16); B.sh_c = (float4*)take(ST_SH16); B.sh_seed = (unsigned int*)take(ST_SH4);
  B.bin_count = (unsigned int*)take(ST_BINS); B.bin_off = (unsigned int*)take(ST_BINS);
  B.cnt = (WfCounters*)take(ST_CNT);
  B.pool = STATIC_POSITIONS;
  ctx->static_wf = &B;
  rc = b200rt_render_device(s, cam, params, accum_dev, stats);  // no wf_ensure_pool on this path: nothing is allocated or freed
  ctx->static_wf = nullptr;
  return rc;
}
