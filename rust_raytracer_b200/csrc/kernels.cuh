// kernels.cuh — batch World::hit operator, the persistent megakernel path tracer (bring-up / cross-check
// pipeline), tonemap and the FP32-issue microbenchmark.  The wavefront pipeline lives in wavefront.cuh.
#pragma once
#include "shade.cuh"

// Path id -> (pixel, sample): consecutive ids walk 8x4 pixel tiles so one warp's primary rays are coherent.
struct PathMap {
  int tiles_x;                    // > 0: 8x4 tiling (width % 8 == 0 and height % 4 == 0); 0: row-major
  unsigned long long per_sample;  // width * height
  unsigned long long total;       // per_sample * spp_count
};
__host__ __device__ inline PathMap make_path_map(int w, int h, int spp) {
  PathMap m;
  m.tiles_x = (w % 8 == 0 && h % 4 == 0) ? w / 8 : 0;
  m.per_sample = (unsigned long long)w * (unsigned long long)h;
  m.total = m.per_sample * (unsigned long long)spp;
  return m;
}
DEV void path_to_pixel(const PathMap& m, const RenderConst& P, unsigned long long id, uint32_t& pixel, uint32_t& sample) {
  unsigned long long s = id / m.per_sample, r = id % m.per_sample;
  sample = (uint32_t)(P.spp_begin + (int)s);
  if (m.tiles_x > 0) {
    int tile = (int)(r >> 5), lane = (int)(r & 31);
    int x = (tile % m.tiles_x) * 8 + (lane & 7), y = (tile / m.tiles_x) * 4 + (lane >> 3);
    pixel = (uint32_t)(y * P.width + x);
  } else pixel = (uint32_t)r;
}

DEV void accum_add(float* accum, uint32_t pixel, float3 c) {
  // NaN/inf guard: one bad sample must not poison the pixel (vec3.rs:223-231 maps NaN to 0 only at the very end)
  if (!(c.x == c.x) || !(c.y == c.y) || !(c.z == c.z) || isinf(c.x) || isinf(c.y) || isinf(c.z)) return;
  if (c.x == 0.f && c.y == 0.f && c.z == 0.f) return;
  atomicAdd(reinterpret_cast<float4*>(accum) + pixel, make_float4(c.x, c.y, c.z, 0.f));
}

// ---- World::hit over a batch (world.rs:27-29) ------------------------------------------------------------
__global__ void __launch_bounds__(128) k_intersect_batch(DevScene S, const float4* __restrict__ o_time, const float4* __restrict__ d_tmax,
                                                         const uint32_t* __restrict__ mseed, unsigned long long n, float t_min, b200rt_hit* __restrict__ out) {
  // whole warps walk the batch together (the loop bound is warp-uniform), so the set of lanes that trace is known from a
  // converged ballot and trace() can let them wait for each other before the out-of-BVH tail (the same code path as extend)
  const unsigned lane = threadIdx.x & 31u;
  for (unsigned long long base = blockIdx.x * (unsigned long long)blockDim.x + (threadIdx.x - lane); base < n; base += (unsigned long long)gridDim.x * blockDim.x) {
    const unsigned long long i = base + lane;
    const bool active = i < n;
    const unsigned tracing = __ballot_sync(0xffffffffu, active);
    if (!active) continue;
    const float4 o = o_time[i], d = d_tmax[i];
    Ray r;
    r.o = f3(o); r.time = o.w; r.d = f3(d); r.tmax = d.w; r.mseed = mseed ? mseed[i] : 0u;
    HitRaw h;
    trace<false>(S, r, t_min, h, tracing);
    b200rt_hit res;
    if (h.prim < 0) { res.t = __int_as_float(0x7f800000); res.prim = -1; res.nx = res.ny = res.nz = 0.f; res.front_face = 0; res.u = res.v = 0.f; }
    else {
      HitInfo hi;
      finish_hit(S, r, h, hi, true);
      res.t = hi.t; res.prim = hi.prim_orig; res.nx = hi.n.x; res.ny = hi.n.y; res.nz = hi.n.z; res.front_face = hi.front_face; res.u = hi.u; res.v = hi.v;
    }
    out[i] = res;
  }
}

// ---- megakernel: one persistent thread per path, regenerated in place --------------------------------------
struct MegaCounters { unsigned long long next_path, extend_rays, shadow_rays; };

__global__ void __launch_bounds__(128) k_megakernel(DevScene S, RenderConst P, PathMap M, float* __restrict__ accum, MegaCounters* __restrict__ cnt) {
  unsigned long long n_ext = 0, n_shd = 0;
  const unsigned lane = threadIdx.x & 31;
  for (;;) {
    unsigned long long base = 0;
    if (lane == 0) base = atomicAdd(&cnt->next_path, 32ull);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (base >= M.total) break;
    PathState ps;
    if (base + lane < M.total) {
      path_to_pixel(M, P, base + lane, ps.pixel, ps.sample);
      ps.throughput = f3(1, 1, 1); ps.radiance = f3(0, 0, 0); ps.prev_pdf_b = -1.f; ps.bounce = 0;
      Ray ray = camera_ray(P, ps.pixel, ps.sample);
      int depth = P.max_depth;
      for (;;) {
        HitRaw h;
        trace<false>(S, ray, P.t_min, h);
        n_ext++;
        if (h.prim < 0) { ps.radiance = ps.radiance + ps.throughput * background_color(P, ray.d); break; }
        if (depth <= 0) break;  // photon_mapper.rs:336-338
        depth--;
        ShadowReq sr;
        bool cont = shade_hit(S, P, ray, h, ps, sr);
        if (sr.valid) {
          HitRaw sh;
          trace<true>(S, sr.ray, P.t_min, sh);
          n_shd++;
          if (sh.prim < 0) ps.radiance = ps.radiance + sr.contrib;
        }
        if (!cont) break;
        ray.tmax = __int_as_float(0x7f800000);
      }
      accum_add(accum, ps.pixel, ps.radiance);
    }
  }
  // warp-aggregated statistics
  for (int o = 16; o > 0; o >>= 1) { n_ext += __shfl_down_sync(0xffffffffu, n_ext, o); n_shd += __shfl_down_sync(0xffffffffu, n_shd, o); }
  if (lane == 0) { atomicAdd(&cnt->extend_rays, n_ext); if (n_shd) atomicAdd(&cnt->shadow_rays, n_shd); }
}

// accum.w += spp for every pixel (the sample count channel)
__global__ void k_add_count(float* accum, int n_pixels, float spp) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_pixels) accum[4 * i + 3] += spp;
}

// From<Vec3> for Rgb<u8> (vec3.rs:223-231): floor(clamp(sqrt(mean), 0, 1) * 255); NaN -> 0
__global__ void k_tonemap(const float4* __restrict__ accum, int n_pixels, uint8_t* __restrict__ rgb) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_pixels) return;
  float4 a = accum[i];
  float inv = a.w > 0.f ? 1.0f / a.w : 0.f;
  float c[3] = {a.x * inv, a.y * inv, a.z * inv};
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    float s = sqrtf(c[k]);
    s = s < 0.f ? 0.f : (s > 1.f ? 1.f : s);
    float f = floorf(s * 255.f);
    rgb[3 * i + k] = (f == f) ? (uint8_t)f : (uint8_t)0;
  }
}

// FP32 FMA-issue microbenchmark: 8 independent FMA chains per thread
__global__ void __launch_bounds__(256) k_fma_peak(float* out, int iters, float a, float b) {
  float x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
      x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}
