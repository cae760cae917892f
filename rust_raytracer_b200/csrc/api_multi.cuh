// api_multi.cuh — Camera::capture_image over several GPUs of one node inside the C ABI:
// one host thread and one context per device, every device renders a disjoint range of the global sample index
// (the reference splits the image into row bands over a thread pool, camera.rs:74-111), then ONE sum-reduce of the
// W x H x 4 float accumulation buffers onto the first device:
//   * a peer-memory kernel (default when the root can map its peers): the root GPU loads the other devices' buffers
//     directly over NVLink (P2P) and adds them — 10 MB per peer, 0.04 ms on 2 x B200, nothing to set up;
//   * or NCCL (ncclReduce over NVLink / NVSwitch; B200RT_MULTI_REDUCE=nccl, or when there is no peer access), loaded
//     at run time with dlopen("libnccl.so.2") so that the library has no link-time dependency and shares the NCCL a
//     host process has already loaded (torch's).  Its first call pays ncclCommInitAll and the lazy channel set-up
//     (0.9 s + 60 ms measured on 2 x B200); the clique is kept for the process lifetime.
// There is no host-staged sum: without peer access and without NCCL the call fails.
// Included by api.cu after api_render.cuh.
#pragma once
#include <dlfcn.h>

#include <chrono>
#include <map>
#include <mutex>
#include <thread>
#include <vector>

namespace {

#define B200RT_MULTI_MAX_DEV 16

// ---- NCCL through dlopen (types restated from nccl.h: opaque communicator, int enums)
typedef void* rt_ncclComm_t;
struct NcclApi {
  void* handle = nullptr;
  int (*CommInitAll)(rt_ncclComm_t*, int, const int*) = nullptr;
  int (*CommDestroy)(rt_ncclComm_t) = nullptr;
  int (*Reduce)(const void*, void*, size_t, int, int, int, rt_ncclComm_t, cudaStream_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  bool ok = false;
};
const int RT_NCCL_FLOAT = 7, RT_NCCL_SUM = 0;  // ncclFloat32, ncclSum

std::mutex g_multi_mutex;
NcclApi g_nccl;
bool g_nccl_tried = false;
std::map<std::vector<int>, std::vector<rt_ncclComm_t>> g_nccl_comms;  // one clique per device list, kept for the process lifetime

NcclApi& nccl_api() {  // call with g_multi_mutex held
  if (g_nccl_tried) return g_nccl;
  g_nccl_tried = true;
  // Order: the NCCL the host process already has (RTLD_NOLOAD: a torch process keeps its own, bundled one), then the one the
  // host names (B200RT_NCCL_LIB: the Python wrapper points it at the interpreter's bundled copy), then the system's.  The
  // order matters beyond taste: the loader keeps ONE object per SONAME, so a process that maps the system libnccl.so.2 first
  // and imports torch later hands torch that copy — and a torch built against a newer NCCL then fails to import
  // (libtorch_cuda.so: undefined symbol ncclDevCommCreate; met on a 2-GPU box when the two-device tests ran before a torch import).
  g_nccl.handle = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
  if (!g_nccl.handle) {
    const char* named = std::getenv("B200RT_NCCL_LIB");
    if (named && *named) g_nccl.handle = dlopen(named, RTLD_NOW | RTLD_LOCAL);
  }
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* n : names) {
    if (g_nccl.handle) break;
    g_nccl.handle = dlopen(n, RTLD_NOW | RTLD_LOCAL);
  }
  if (!g_nccl.handle) return g_nccl;
  g_nccl.CommInitAll = (int (*)(rt_ncclComm_t*, int, const int*))dlsym(g_nccl.handle, "ncclCommInitAll");
  g_nccl.CommDestroy = (int (*)(rt_ncclComm_t))dlsym(g_nccl.handle, "ncclCommDestroy");
  g_nccl.Reduce = (int (*)(const void*, void*, size_t, int, int, int, rt_ncclComm_t, cudaStream_t))dlsym(g_nccl.handle, "ncclReduce");
  g_nccl.GroupStart = (int (*)())dlsym(g_nccl.handle, "ncclGroupStart");
  g_nccl.GroupEnd = (int (*)())dlsym(g_nccl.handle, "ncclGroupEnd");
  g_nccl.GetErrorString = (const char* (*)(int))dlsym(g_nccl.handle, "ncclGetErrorString");
  g_nccl.ok = g_nccl.CommInitAll && g_nccl.Reduce && g_nccl.GroupStart && g_nccl.GroupEnd && g_nccl.GetErrorString;
  return g_nccl;
}

// b200rt_ctx_destroy: the NCCL cliques that include the context's device are destroyed with it (a clique is cached for the
// calls to come, not for the process exit: communicators left alive there are torn down by NCCL's own exit handlers, after
// the CUDA context they point into)
void multi_forget_device(int device) {
  std::lock_guard<std::mutex> lock(g_multi_mutex);
  for (auto it = g_nccl_comms.begin(); it != g_nccl_comms.end();) {
    if (std::find(it->first.begin(), it->first.end(), device) == it->first.end()) { ++it; continue; }
    if (g_nccl.CommDestroy) for (rt_ncclComm_t c : it->second) if (c) g_nccl.CommDestroy(c);
    it = g_nccl_comms.erase(it);
  }
}

// ---- peer-memory reduce: dst (on the launching device) += sum of the peers' buffers, read over NVLink
struct PeerSrc { const float4* p[B200RT_MULTI_MAX_DEV]; int n; };
__global__ void __launch_bounds__(256) k_reduce_peers(float4* __restrict__ dst, PeerSrc src, size_t n4) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    float4 a = dst[i];
    for (int k = 0; k < src.n; ++k) {
      const float4 b = __ldcs(src.p[k] + i);  // remote, read once
      a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    }
    dst[i] = a;
  }
}

// the reference's split (camera.rs:85-86 for rows): contiguous, disjoint, exhaustive, remainder to the low ranks
void multi_spp_range(int rank, int world, int spp_begin, int spp_total, int& begin, int& count) {
  const int base = spp_total / world, rem = spp_total % world;
  count = base + (rank < rem ? 1 : 0);
  begin = spp_begin + rank * base + (rank < rem ? rank : rem);
}

}  // namespace

extern "C" int b200rt_render_multi(int32_t n_dev, b200rt_scene* const* scenes, const b200rt_camera* cam, const b200rt_params* params,
                                   float* accum_host, b200rt_stats* stats_per_device, b200rt_multi_stats* multi) {
  if (n_dev < 1 || n_dev > B200RT_MULTI_MAX_DEV || !scenes || !cam || !params || !accum_host) return B200RT_ERR_INVALID;
  for (int g = 0; g < n_dev; ++g) { if (!scenes[g]) return B200RT_ERR_INVALID; SCENE_ALIVE(scenes[g]); }
  b200rt_ctx* root = scenes[0]->ctx;
  // One context per scene, normally one device per context.  Several contexts on ONE device are allowed as well (each has
  // its own streams and queue pools; the "peer" buffers are then plain local memory): that is how a one-GPU box exercises this
  // entry point, sample ranges, host threads and k_reduce_peers included.  NCCL cannot rank one device twice.
  std::vector<int> devs(n_dev);
  bool shared_device = false;
  for (int g = 0; g < n_dev; ++g) {
    devs[g] = scenes[g]->ctx->device;
    for (int k = 0; k < g; ++k) {
      if (scenes[k]->ctx == scenes[g]->ctx) { root->err = "render_multi: every scene needs its own context"; return B200RT_ERR_INVALID; }
      if (devs[k] == devs[g]) shared_device = true;
    }
  }
  int rc = validate_params(root, params);
  if (rc) return rc;
  const auto t_start = std::chrono::steady_clock::now();
  const size_t n4 = (size_t)params->width * params->height, bytes = n4 * 16;

  // 1. every device renders its sample range into its own resident buffer (one host thread per device)
  std::vector<float*> accum(n_dev, nullptr);
  std::vector<int> rcs(n_dev, B200RT_OK);
  std::vector<b200rt_stats> st(n_dev);
  auto worker = [&](int g) {
    b200rt_ctx* ctx = scenes[g]->ctx;
    cudaError_t e = cudaSetDevice(ctx->device);
    if (e == cudaSuccess) e = cudaMalloc(&accum[g], bytes);
    if (e == cudaSuccess) e = cudaMemsetAsync(accum[g], 0, bytes, ctx->stream);
    if (e != cudaSuccess) { ctx->err = std::string("render_multi: ") + cudaGetErrorString(e); rcs[g] = B200RT_ERR_CUDA; return; }
    b200rt_params p = *params;
    multi_spp_range(g, n_dev, params->spp_begin, params->spp_count, p.spp_begin, p.spp_count);
    std::memset(&st[g], 0, sizeof(b200rt_stats));
    rcs[g] = b200rt_render_device(scenes[g], cam, &p, accum[g], &st[g]);  // synchronises the device's stream
  };
  {
    std::vector<std::thread> threads;
    for (int g = 1; g < n_dev; ++g) threads.emplace_back(worker, g);
    worker(0);
    for (auto& t : threads) t.join();
  }
  auto release = [&]() {
    for (int g = 0; g < n_dev; ++g)
      if (accum[g]) { cudaSetDevice(devs[g]); cudaFree(accum[g]); }
    cudaSetDevice(devs[0]);
  };
  for (int g = 0; g < n_dev; ++g)
    if (rcs[g]) {
      if (g) root->err = "render_multi: device " + std::to_string(devs[g]) + ": " + scenes[g]->ctx->err;
      release();
      return rcs[g];
    }

  // 2. one sum-reduce onto the first device
  int path = B200RT_REDUCE_NONE;
  float ms_reduce = 0.f;
  if (n_dev > 1) {
    std::lock_guard<std::mutex> lock(g_multi_mutex);
    const char* want = getenv("B200RT_MULTI_REDUCE");
    const bool want_peer = want && std::strcmp(want, "peer") == 0, want_nccl = want && std::strcmp(want, "nccl") == 0;
    bool can_peer = true;
    for (int g = 1; g < n_dev; ++g) {
      int can = 0;
      if (devs[g] != devs[0] && (cudaDeviceCanAccessPeer(&can, devs[0], devs[g]) != cudaSuccess || !can)) can_peer = false;
    }
    cudaGetLastError();
    const bool use_nccl = want_nccl || (!can_peer && !want_peer);
    if (use_nccl && shared_device) { root->err = "render_multi: NCCL cannot reduce between contexts that share a device (use the peer path)"; release(); return B200RT_ERR_INVALID; }
    NcclApi none;
    NcclApi& N = use_nccl ? nccl_api() : none;
    if (use_nccl && !N.ok) {
      root->err = want_nccl ? "render_multi: B200RT_MULTI_REDUCE=nccl but libnccl.so.2 could not be loaded" : "render_multi: no peer access between the devices and libnccl.so.2 could not be loaded";
      release();
      return B200RT_ERR_CUDA;
    }
    if (use_nccl) {
      auto it = g_nccl_comms.find(devs);
      if (it == g_nccl_comms.end()) {
        std::vector<rt_ncclComm_t> comms(n_dev, nullptr);
        const int r = N.CommInitAll(comms.data(), n_dev, devs.data());
        if (r != 0) { root->err = std::string("ncclCommInitAll: ") + N.GetErrorString(r); release(); return B200RT_ERR_CUDA; }
        it = g_nccl_comms.emplace(devs, comms).first;
      }
      cudaSetDevice(devs[0]);
      cudaEventRecord(root->ev[0], root->stream);
      int r = N.GroupStart();
      for (int g = 0; g < n_dev && r == 0; ++g) {
        cudaSetDevice(devs[g]);
        r = N.Reduce(accum[g], accum[g], n4 * 4, RT_NCCL_FLOAT, RT_NCCL_SUM, 0, it->second[g], scenes[g]->ctx->stream);
      }
      const int r2 = N.GroupEnd();
      if (r == 0) r = r2;
      if (r != 0) { root->err = std::string("ncclReduce: ") + N.GetErrorString(r); release(); return B200RT_ERR_CUDA; }
      for (int g = 1; g < n_dev; ++g) { cudaSetDevice(devs[g]); cudaStreamSynchronize(scenes[g]->ctx->stream); }
      path = B200RT_REDUCE_NCCL;
    } else {
      PeerSrc src;
      src.n = n_dev - 1;
      cudaSetDevice(devs[0]);
      for (int g = 1; g < n_dev; ++g) {
        src.p[g - 1] = (const float4*)accum[g];
        if (devs[g] == devs[0]) continue;  // same device: local memory
        int can = 0;
        cudaDeviceCanAccessPeer(&can, devs[0], devs[g]);
        if (!can) { root->err = "render_multi: B200RT_MULTI_REDUCE=peer but no peer access from device " + std::to_string(devs[0]) + " to " + std::to_string(devs[g]); release(); return B200RT_ERR_CUDA; }
        const cudaError_t e = cudaDeviceEnablePeerAccess(devs[g], 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { root->err = std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e); release(); return B200RT_ERR_CUDA; }
        cudaGetLastError();  // clears "already enabled"
      }
      cudaEventRecord(root->ev[0], root->stream);
      k_reduce_peers<<<root->prop.multiProcessorCount * 4, 256, 0, root->stream>>>((float4*)accum[0], src, n4);
      path = B200RT_REDUCE_PEER;
    }
    cudaSetDevice(devs[0]);
    cudaEventRecord(root->ev[1], root->stream);
    cudaError_t e = cudaEventSynchronize(root->ev[1]);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) { root->err = std::string("render_multi reduce: ") + cudaGetErrorString(e); release(); return B200RT_ERR_CUDA; }
    cudaEventElapsedTime(&ms_reduce, root->ev[0], root->ev[1]);
  }
  // 3. the image leaves the root device
  cudaSetDevice(devs[0]);
  cudaError_t e = cudaMemcpyAsync(accum_host, accum[0], bytes, cudaMemcpyDeviceToHost, root->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(root->stream);
  release();
  if (e != cudaSuccess) { root->err = std::string("render_multi: ") + cudaGetErrorString(e); return B200RT_ERR_CUDA; }
  if (stats_per_device) for (int g = 0; g < n_dev; ++g) stats_per_device[g] = st[g];
  if (multi) {
    std::memset(multi, 0, sizeof(*multi));
    for (int g = 0; g < n_dev; ++g) {
      multi->paths += st[g].paths; multi->extend_rays += st[g].extend_rays; multi->shadow_rays += st[g].shadow_rays;
      if (st[g].ms_total > multi->ms_render_max) multi->ms_render_max = st[g].ms_total;
    }
    multi->ms_reduce = ms_reduce;
    multi->ms_wall = (float)(std::chrono::duration<double>(std::chrono::steady_clock::now() - t_start).count() * 1e3);
    multi->reduce_path = path;
    multi->n_dev = n_dev;
  }
  return B200RT_OK;
}
