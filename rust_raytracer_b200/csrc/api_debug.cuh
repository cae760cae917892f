// api_debug.cuh — b200rt_debug_eval: single functions of the hot path evaluated on the device at caller-supplied
// inputs, so that tests can compare Texture::get_color, Material::scatter and Camera::get_ray with the f64 oracle point
// for point (the image tests only compare means).  The kernel calls the very functions the shade / generate kernels
// call (texture_color, shade_surface<PATH>, camera_ray).  Included by api.cu.
#pragma once

__global__ void __launch_bounds__(128) k_debug_eval(DevScene S, RenderConst P, int kind, unsigned long long n, const float* __restrict__ in, float* __restrict__ out) {
  for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < n; i += (unsigned long long)gridDim.x * blockDim.x) {
    if (kind == B200RT_EVAL_TEXTURE) {  // in: 8 floats (texture id bits, p.xyz, u, v, -, -); out: 4 floats (rgb, -)
      const float* a = in + 8 * i;
      HitInfo h;
      h.p = f3(a[1], a[2], a[3]); h.u = a[4]; h.v = a[5];
      h.t = 0.f; h.n = f3(0, 1, 0); h.front_face = 1; h.material = 0; h.prim_orig = 0; h.type = 0;
      const float3 c = texture_color(S, __float_as_int(a[0]), h);
      out[4 * i] = c.x; out[4 * i + 1] = c.y; out[4 * i + 2] = c.z; out[4 * i + 3] = 0.f;
    } else if (kind == B200RT_EVAL_SCATTER) {
      // in: 20 floats (material id bits, front_face bits, p.xyz, n.xyz (unit, facing the ray), u, v, d.xyz, u1, u2, u3, time, -, -, -)
      // out: 16 floats (continues, dir.xyz | attenuation.rgb, - | emitted.rgb, - | sphere sample of (u1, u2).xyz, -)
      const float* a = in + 20 * i;
      HitInfo h;
      h.material = __float_as_int(a[0]); h.front_face = __float_as_int(a[1]);
      h.p = f3(a[2], a[3], a[4]); h.n = f3(a[5], a[6], a[7]); h.u = a[8]; h.v = a[9];
      h.t = 1.f; h.prim_orig = -1; h.type = 0;
      Ray ray;
      ray.o = h.p - f3(a[10], a[11], a[12]); ray.d = f3(a[10], a[11], a[12]); ray.time = a[16]; ray.tmax = __int_as_float(0x7f800000); ray.mseed = 0u;
      PathState ps;
      ps.throughput = f3(1, 1, 1); ps.radiance = f3(0, 0, 0); ps.prev_pdf_b = -1.f; ps.pixel = 0u; ps.sample = 0u; ps.bounce = 1;
      ShadowReq sr;
      sr.valid = false;
      const float4 mat = __ldg(&S.materials[h.material]);
      const bool cont = shade_surface<B200RT_INTEGRATOR_PATH>(S, P, ray, h, mat, a[13], a[14], a[15], 0u, ps, sr);
      const float3 sv = sample_unit_sphere(a[13], a[14]);
      float* o = out + 16 * i;
      o[0] = cont ? 1.f : 0.f; o[1] = cont ? ray.d.x : 0.f; o[2] = cont ? ray.d.y : 0.f; o[3] = cont ? ray.d.z : 0.f;
      o[4] = ps.throughput.x; o[5] = ps.throughput.y; o[6] = ps.throughput.z; o[7] = 0.f;
      o[8] = ps.radiance.x; o[9] = ps.radiance.y; o[10] = ps.radiance.z; o[11] = 0.f;
      o[12] = sv.x; o[13] = sv.y; o[14] = sv.z; o[15] = 0.f;
    } else if (kind == B200RT_EVAL_CAMERA) {
      // in: 2 u32 (pixel, sample); out: 16 floats (o.xyz, time | d.xyz, - | the draws: jitter x, jitter y, lens disk x, lens disk y | time draw, -, -, -)
      const uint32_t pixel = __float_as_uint(in[2 * i]), sample = __float_as_uint(in[2 * i + 1]);
      const Ray r = camera_ray(P, pixel, sample);
      const uint4 r0 = philox4x32(make_uint4(pixel, sample, 0u, 0u), P.key);
      const uint4 r1 = philox4x32(make_uint4(pixel, sample, 0u, 1u), P.key);
      float2 rd = make_float2(0.f, 0.f);
      if (P.cam.lens_radius > 0.f) rd = sample_unit_disk(u01(r0.z), u01(r0.w));
      float* o = out + 16 * i;
      o[0] = r.o.x; o[1] = r.o.y; o[2] = r.o.z; o[3] = r.time;
      o[4] = r.d.x; o[5] = r.d.y; o[6] = r.d.z; o[7] = 0.f;
      o[8] = u01(r0.x); o[9] = u01(r0.y); o[10] = rd.x; o[11] = rd.y;
      o[12] = u01(r1.x); o[13] = 0.f; o[14] = 0.f; o[15] = 0.f;
    } else {  // B200RT_EVAL_SAMPLERS — in: 8 floats (u1, u2, n.xyz, -, -, -); out: 12 floats (on-sphere xyz, - | cosine hemisphere xyz, - | disk xy, -, -)
      const float* a = in + 8 * i;
      const float3 sv = sample_unit_sphere(a[0], a[1]);
      const float3 ch = sample_cosine_hemisphere(f3(a[2], a[3], a[4]), a[0], a[1]);
      const float2 dk = sample_unit_disk(a[0], a[1]);
      float* o = out + 12 * i;
      o[0] = sv.x; o[1] = sv.y; o[2] = sv.z; o[3] = 0.f; o[4] = ch.x; o[5] = ch.y; o[6] = ch.z; o[7] = 0.f; o[8] = dk.x; o[9] = dk.y; o[10] = 0.f; o[11] = 0.f;
    }
  }
}

extern "C" int b200rt_debug_eval(b200rt_scene* s, const b200rt_camera* cam, const b200rt_params* params, int32_t kind, uint64_t n, const float* in_host, float* out_host) {
  if (!s || !cam || !params || kind < B200RT_EVAL_TEXTURE || kind > B200RT_EVAL_SAMPLERS || (n && (!in_host || !out_host))) return B200RT_ERR_INVALID;
  if (n == 0) return B200RT_OK;
  SCENE_ALIVE(s);
  b200rt_ctx* ctx = s->ctx;
  CU_CHECK(ctx, cudaSetDevice(ctx->device));
  static const int in_w[4] = {8, 20, 2, 8}, out_w[4] = {4, 16, 16, 12};
  const size_t in_bytes = (size_t)n * in_w[kind] * 4, out_bytes = (size_t)n * out_w[kind] * 4;
  float *d_in = nullptr, *d_out = nullptr;
  cudaError_t e = cudaMalloc(&d_in, in_bytes);
  if (e == cudaSuccess) e = cudaMalloc(&d_out, out_bytes);
  if (e == cudaSuccess) e = cudaMemcpyAsync(d_in, in_host, in_bytes, cudaMemcpyHostToDevice, ctx->stream);
  if (e == cudaSuccess) {
    const RenderConst P = make_render_const(cam, params);
    const unsigned long long want = (n + 127) / 128, cap = (unsigned long long)ctx->prop.multiProcessorCount * 16ull;
    k_debug_eval<<<(int)(want < cap ? want : cap), 128, 0, ctx->stream>>>(s->dev, P, kind, n, d_in, d_out);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpyAsync(out_host, d_out, out_bytes, cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  cudaFree(d_in); cudaFree(d_out);
  if (e != cudaSuccess) { ctx->err = std::string("debug_eval: ") + cudaGetErrorString(e); return B200RT_ERR_CUDA; }
  return B200RT_OK;
}
