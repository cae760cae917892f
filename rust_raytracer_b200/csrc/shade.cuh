// shade.cuh — Texture::get_color (material.rs:18-84), Material::scatter / emitted (material.rs:86-231), the
// book-3 PDFs (absent upstream) and one step of the per-sample loop (photon_mapper.rs:327-365 skeleton).
#pragma once
#include "intersect.cuh"

struct RenderConst {
  b200rt_camera cam;
  int width, height;
  int spp_begin, spp_count;
  int max_depth;
  int integrator;
  int background_mode;
  uint32_t flags;
  float t_min;
  float3 background;
  uint2 key;  // Philox key = seed
};

// ---- textures ---------------------------------------------------------------------------------------------
DEV float perlin_noise(const DevScene& S, float3 p) {
  float fx = floorf(p.x), fy = floorf(p.y), fz = floorf(p.z);
  float u = p.x - fx, v = p.y - fy, w = p.z - fz;
  int i = (int)fx, j = (int)fy, k = (int)fz;
  float uu = u * u * (3 - 2 * u), vv = v * v * (3 - 2 * v), ww = w * w * (3 - 2 * w);
  float accum = 0;
#pragma unroll
  for (int di = 0; di < 2; ++di)
#pragma unroll
    for (int dj = 0; dj < 2; ++dj)
#pragma unroll
      for (int dk = 0; dk < 2; ++dk) {
        int idx = __ldg(&S.perlin_perm[(i + di) & 255]) ^ __ldg(&S.perlin_perm[256 + ((j + dj) & 255)]) ^ __ldg(&S.perlin_perm[512 + ((k + dk) & 255)]);
        float3 c = f3(__ldg(&S.perlin_vec[idx]));
        float3 wv = f3(u - di, v - dj, w - dk);
        accum += (di ? uu : 1 - uu) * (dj ? vv : 1 - vv) * (dk ? ww : 1 - ww) * dot(c, wv);
      }
  return accum;
}
DEV float perlin_turb(const DevScene& S, float3 p) {
  float accum = 0, weight = 1;
  for (int i = 0; i < 7; ++i) { accum += weight * perlin_noise(S, p); weight *= 0.5f; p = p * 2.0f; }
  return fabsf(accum);
}
// perlin_turb with the whole (converged) warp on ONE point: 7 octaves x 8 lattice corners = 56 terms, a lane per term (4 octaves
// per round, 2 rounds), the 8 corners of an octave summed inside their 8-lane segment, the octaves across the segments.  The
// shade kernel evaluates NOISE textures this way: a warp streams rays in the order of where they STARTED, so the one Perlin
// sphere of book-2 is hit by ~2 of its 32 lanes at a time, and those two spent ~1000 instructions each while the others
// waited (16 % of k_wf_shade's instructions at 2.1 lanes).  Same lattice terms as perlin_noise; the sums associate differently
// (a few ulp).  Every lane returns the value.
DEV float perlin_turb_warp(const DevScene& S, float3 p, unsigned lane) {
  const int corner = (int)(lane & 7u), di = corner >> 2, dj = (corner >> 1) & 1, dk = corner & 1;
  float total = 0.f;
#pragma unroll
  for (int round = 0; round < 2; ++round) {
    const int oct = round * 4 + (int)(lane >> 3);
    float term = 0.f;
    if (oct < 7) {
      const float sc = (float)(1 << oct);
      const float3 q = p * sc;  // exact: the serial version doubles p per octave
      const float fx = floorf(q.x), fy = floorf(q.y), fz = floorf(q.z);
      const float u = q.x - fx, v = q.y - fy, w = q.z - fz;
      const int i = (int)fx, j = (int)fy, k = (int)fz;
      const float uu = u * u * (3 - 2 * u), vv = v * v * (3 - 2 * v), ww = w * w * (3 - 2 * w);
      const int idx = __ldg(&S.perlin_perm[(i + di) & 255]) ^ __ldg(&S.perlin_perm[256 + ((j + dj) & 255)]) ^ __ldg(&S.perlin_perm[512 + ((k + dk) & 255)]);
      const float3 c = f3(__ldg(&S.perlin_vec[idx]));
      const float3 wv = f3(u - di, v - dj, w - dk);
      term = (di ? uu : 1 - uu) * (dj ? vv : 1 - vv) * (dk ? ww : 1 - ww) * dot(c, wv) * (1.0f / sc);
    }
    term += __shfl_xor_sync(0xffffffffu, term, 1); term += __shfl_xor_sync(0xffffffffu, term, 2); term += __shfl_xor_sync(0xffffffffu, term, 4);
    term += __shfl_xor_sync(0xffffffffu, term, 8); term += __shfl_xor_sync(0xffffffffu, term, 16);
    total += term;
  }
  return fabsf(total);
}
DEV bool texture_needs_uv(const DevScene& S, int tex) { return __float_as_int(__ldg(&S.textures[3 * tex]).x) == B200RT_TEX_IMAGE; }

DEV bool texture_is_noise(const DevScene& S, int tex) { return __float_as_int(__ldg(&S.textures[3 * tex]).x) == B200RT_TEX_NOISE; }

// turb >= 0: the turbulence at h.p, already evaluated (perlin_turb_warp)
DEV float3 texture_color(const DevScene& S, int tex, const HitInfo& h, float turb = -1.f) {
  const float4 hd = __ldg(&S.textures[3 * tex]);
  const int type = __float_as_int(hd.x);
  DBG_INC(48 + type);
  if (type == B200RT_TEX_CONSTANT) return f3(__ldg(&S.textures[3 * tex + 1]));
  if (type == B200RT_TEX_CHECKER) {  // material.rs:58-68
    float sines = sinf(10.f * h.p.x) * sinf(10.f * h.p.y) * sinf(10.f * h.p.z);
    return f3(__ldg(&S.textures[3 * tex + (sines < 0.f ? 1 : 2)]));
  }
  if (type == B200RT_TEX_IMAGE) {  // material.rs:70-84 (index clamped; upstream would panic at u == 1)
    const int4 im = __ldg(&S.images[__float_as_int(hd.y)]);
    // the texel index in f64 from the f32 (u, v): w * u is exact there, so the floor agrees with the f64 reference at every
    // texel boundary (two DMULs per image-textured hit)
    const double u = (double)fminf(fmaxf(h.u, 0.f), 1.f), v = 1.0 - (double)fminf(fmaxf(h.v, 0.f), 1.f);
    int x = min((int)floor((double)im.y * u), im.y - 1), y = min((int)floor((double)im.z * v), im.z - 1);
    const uint8_t* px = S.image_data + im.x + 3 * ((size_t)y * im.y + x);
    return f3(px[0] * (1.f / 255.f), px[1] * (1.f / 255.f), px[2] * (1.f / 255.f));
  }
  return f3(1, 1, 1) * (0.5f * (1.f + sinf(hd.z * h.p.z + 10.f * (turb >= 0.f ? turb : perlin_turb(S, h.p)))));  // NOISE
}

// ---- light PDFs (book-3 HittablePDF over the marked lights) -----------------------------------------------
DEV float light_pdf_one(const DevScene& S, int li, float3 o, float3 d, float t_min) {
  const float4 L0 = __ldg(&S.lights[3 * li]), L1 = __ldg(&S.lights[3 * li + 1]);
  if (__float_as_int(L0.x) == 1) {
    const float4 L2 = __ldg(&S.lights[3 * li + 2]);
    const int axis = __float_as_int(L0.z), ia = axis == 0 ? 1 : 0, ib = axis == 2 ? 1 : 2;
    float da = comp(d, axis);
    float t = (L2.x - comp(o, axis)) / da;
    if (!(t >= t_min) || !(t < 3.0e38f)) return 0.f;
    float a = comp(o, ia) + t * comp(d, ia), b = comp(o, ib) + t * comp(d, ib);
    if (a < L1.x || a > L1.z || b < L1.y || b > L1.w) return 0.f;
    float d2 = dot(d, d);
    float cosine = fabsf(da) * rsqrtf(d2);
    return t * t * d2 / (cosine * L2.y);
  }
  float3 oc = o - f3(L1);
  float a = dot(d, d), hb = dot(oc, d), cc = dot(oc, oc) - L1.w * L1.w;
  float disc = hb * hb - a * cc;
  if (disc < 0.f) return 0.f;
  float s = sqrtf(disc);
  float r0 = (-hb - s) / a, r1 = (-hb + s) / a;
  if (!(r0 >= t_min) && !(r1 >= t_min)) return 0.f;
  float cos_max = sqrtf(fmaxf(0.f, 1.f - L1.w * L1.w / dot(oc, oc)));
  return 1.f / (2.f * PI_F * (1.f - cos_max));
}
DEV float light_pdf_all(const DevScene& S, float3 o, float3 d, float t_min) {
  float s = 0;
  for (int i = 0; i < S.n_lights; ++i) s += light_pdf_one(S, i, o, d, t_min);
  return s / (float)S.n_lights;
}
DEV float3 light_sample_dir(const DevScene& S, int li, float3 o, float u1, float u2) {
  DBG_INC(56);
  const float4 L0 = __ldg(&S.lights[3 * li]), L1 = __ldg(&S.lights[3 * li + 1]);
  if (__float_as_int(L0.x) == 1) {
    const float4 L2 = __ldg(&S.lights[3 * li + 2]);
    const int axis = __float_as_int(L0.z), ia = axis == 0 ? 1 : 0, ib = axis == 2 ? 1 : 2;
    float c[3];
    c[ia] = L1.x + u1 * (L1.z - L1.x); c[ib] = L1.y + u2 * (L1.w - L1.y); c[axis] = L2.x;
    return f3(c[0], c[1], c[2]) - o;
  }
  float3 dc = f3(L1) - o;
  float d2 = dot(dc, dc);
  float3 w = dc * rsqrtf(d2), u, v;
  onb_from_w(w, u, v);
  float z = 1.f + u2 * (sqrtf(fmaxf(0.f, 1.f - L1.w * L1.w / d2)) - 1.f);
  float s, c;
  sincospif(2.f * u1, &s, &c);
  float r = sqrtf(fmaxf(0.f, 1.f - z * z));
  return (c * r) * u + (s * r) * v + z * w;
}
DEV int light_index_of_prim(const DevScene& S, int prim_orig) {
  for (int i = 0; i < S.n_lights; ++i) if (__float_as_int(__ldg(&S.lights[3 * i]).y) == prim_orig) return i;
  return -1;
}

// ---- camera (camera.rs:57-64, 97-99) ---------------------------------------------------------------------
DEV float3 ld3(const float* p) { return f3(p[0], p[1], p[2]); }
DEV Ray camera_ray(const RenderConst& P, uint32_t pixel, uint32_t sample) {
  DBG_INC(9);
  const uint4 r0 = philox4x32(make_uint4(pixel, sample, 0u, 0u), P.key);
  const uint4 r1 = philox4x32(make_uint4(pixel, sample, 0u, 1u), P.key);
  const int x = pixel % P.width, y = pixel / P.width;
  const float s = ((float)x + u01(r0.x)) / (float)(P.width - 1);
  const float t = 1.0f - ((float)y + u01(r0.y)) / (float)(P.height - 1);  // get_ray(u, 1 - v)
  float3 offset = f3(0, 0, 0);
  if (P.cam.lens_radius > 0.f) {
    float2 rd = sample_unit_disk(u01(r0.z), u01(r0.w));
    offset = ld3(P.cam.u) * (P.cam.lens_radius * rd.x) + ld3(P.cam.v) * (P.cam.lens_radius * rd.y);
  }
  Ray r;
  r.o = ld3(P.cam.origin) + offset;
  r.d = ld3(P.cam.lower_left_corner) + s * ld3(P.cam.horizontal) + t * ld3(P.cam.vertical) - ld3(P.cam.origin) - offset;
  r.time = P.cam.time1 > P.cam.time0 ? P.cam.time0 + u01(r1.x) * (P.cam.time1 - P.cam.time0) : P.cam.time0;
  r.tmax = __int_as_float(0x7f800000);
  r.mseed = r1.w;
  return r;
}

DEV float3 background_color(const RenderConst& P, float3 d) {
  if (P.background_mode == B200RT_BACKGROUND_SKY) {
    float t = 0.5f * (normalize(d).y + 1.0f);
    return (1.0f - t) * f3(1, 1, 1) + t * P.background;
  }
  return P.background;
}

// ---- one bounce ---------------------------------------------------------------------------------------------
struct PathState {
  float3 throughput;
  float3 radiance;
  float prev_pdf_b;  // NEE: bsdf pdf of the direction that produced the current ray (< 0: camera / specular)
  uint32_t pixel, sample;
  int bounce;  // hits processed so far
};
struct ShadowReq {  // light-sample request produced by the NEE integrator
  bool valid;
  Ray ray;
  float3 contrib;
};

DEV float schlick(float cosine, float ref_idx) {  // material.rs:150-154
  float r0 = (1.f - ref_idx) / (1.f + ref_idx);
  r0 = r0 * r0;
  float m = 1.f - cosine;
  return r0 + (1.f - r0) * (m * m * m * m * m);
}
DEV float3 reflect3(float3 v, float3 n) { return v - 2.f * dot(v, n) * n; }  // vec3.rs:163-165
DEV float3 refract3(float3 uv, float3 n, float eta) {                        // vec3.rs:167-172
  float cos_theta = fminf(dot(-uv, n), 1.0f);
  float3 perp = eta * (uv + cos_theta * n);
  float3 par = -sqrtf(fabsf(1.0f - dot(perp, perp))) * n;
  return perp + par;
}

// Material::scatter / emitted for a finished HitRecord `h` with the three uniforms of this bounce supplied by the caller
// (u1, u2 -> direction sample, u3 -> choice draw): adds emission, samples the next direction.  Returns true when the
// path continues (ray is replaced).  Split from shade_hit so that the function-level parity operator
// (b200rt_debug_eval) runs exactly the code the shade kernel runs, on injected inputs.
// INTEG >= 0 fixes the integrator at compile time (the wavefront shade kernel is instantiated per integrator, so
// the PATH build carries no light-sampling code: fewer registers, a third of the instructions); -1 reads P.integrator.
template <int INTEG = -1>
DEV bool shade_surface(const DevScene& S, const RenderConst& P, Ray& ray, const HitInfo& h, const float4 mat, const float u1, const float u2, const float u3,
                       const uint32_t next_seed, PathState& ps, ShadowReq& sreq, const float turb = -1.f) {
  const int integrator = INTEG >= 0 ? INTEG : P.integrator;
  const int mtype = __float_as_int(mat.x), tex = __float_as_int(mat.y);
  const float3 color = texture_color(S, tex, h, turb);
  DBG_INC(40 + mtype);

  if (mtype == B200RT_MAT_DIFFUSE_LIGHT) {  // emitted = emit.get_color(rec), both faces (material.rs:209-211)
    float w = 1.f;
    if (integrator == B200RT_INTEGRATOR_NEE && ps.prev_pdf_b >= 0.f && S.n_lights > 0) {
      int li = light_index_of_prim(S, h.prim_orig);
      if (li >= 0) {
        float pl = light_pdf_one(S, li, ray.o, ray.d, P.t_min) / (float)S.n_lights;
        w = ps.prev_pdf_b / (ps.prev_pdf_b + pl);
      }
    }
    ps.radiance = ps.radiance + ps.throughput * color * w;
    if (!(P.flags & B200RT_FLAG_LIGHT_SCATTERS)) return false;
    // material.rs:201-208: scatters like Lambertian with bsdf 1/pi
    float3 dir = h.n + sample_unit_sphere(u1, u2);
    if (fabsf(dir.x) < 1e-8f && fabsf(dir.y) < 1e-8f && fabsf(dir.z) < 1e-8f) dir = h.n;
    ps.throughput = ps.throughput * (1.0f / PI_F);
    ps.prev_pdf_b = -1.f;
    ray.o = h.p; ray.d = dir; ray.mseed = next_seed;
    return true;
  }
  const bool diffuse = mtype == B200RT_MAT_LAMBERTIAN || mtype == B200RT_MAT_ISOTROPIC;
  if (diffuse && integrator != B200RT_INTEGRATOR_PATH) {
    const bool iso = mtype == B200RT_MAT_ISOTROPIC;
    const bool use_lights = S.n_lights > 0;
    if (integrator == B200RT_INTEGRATOR_MIXTURE) {  // MixturePDF: 1/2 lights + 1/2 bsdf lobe
      float3 dir;
      if (use_lights && u3 < 0.5f) {
        const uint4 r2 = philox4x32(make_uint4(ps.pixel, ps.sample, (uint32_t)ps.bounce, 1u), P.key);
        int li = min(S.n_lights - 1, (int)(u01(r2.z) * S.n_lights));
        dir = light_sample_dir(S, li, h.p, u01(r2.x), u01(r2.y));
      } else dir = iso ? sample_unit_sphere(u1, u2) : sample_cosine_hemisphere(h.n, u1, u2);
      float pb = iso ? 0.25f / PI_F : fmaxf(0.f, dot(h.n, normalize(dir))) / PI_F;
      float pdf = use_lights ? 0.5f * light_pdf_all(S, h.p, dir, P.t_min) + 0.5f * pb : pb;
      if (!(pdf > 0.f) || !(pb > 0.f)) return false;
      ps.throughput = ps.throughput * color * (pb / pdf);
      ps.prev_pdf_b = -1.f;
      ray.o = h.p; ray.d = dir; ray.mseed = next_seed;
      return true;
    }
    // NEE + MIS (balance heuristic): one light sample, visibility resolved by the shadow kernel
    if (use_lights) {
      const uint4 r2 = philox4x32(make_uint4(ps.pixel, ps.sample, (uint32_t)ps.bounce, 1u), P.key);
      int li = min(S.n_lights - 1, (int)(u01(r2.z) * S.n_lights));
      float3 dl = light_sample_dir(S, li, h.p, u01(r2.x), u01(r2.y));
      float pb = iso ? 0.25f / PI_F : fmaxf(0.f, dot(h.n, normalize(dl))) / PI_F;
      float pl = light_pdf_one(S, li, h.p, dl, P.t_min) / (float)S.n_lights;
      if (pb > 0.f && pl > 0.f) {
        // distance to the light along dl (parametric): rect lights sit at t = 1, sphere lights at their near root
        const float4 L0 = __ldg(&S.lights[3 * li]), L1 = __ldg(&S.lights[3 * li + 1]);
        float tl = 1.0f;
        if (__float_as_int(L0.x) == 0) {
          float3 oc = h.p - f3(L1);
          float a = dot(dl, dl), hb = dot(oc, dl), cc = dot(oc, oc) - L1.w * L1.w;
          float s = sqrtf(fmaxf(0.f, hb * hb - a * cc));
          float r0 = (-hb - s) / a;
          tl = r0 >= P.t_min ? r0 : (-hb + s) / a;
        }
        HitInfo lh = h;  // emission of the light at the sampled point (constant / checker / noise textures use p)
        lh.p = h.p + tl * dl; lh.u = 0.f; lh.v = 0.f;
        const float4 lmat = __ldg(&S.materials[__float_as_int(L0.w)]);
        float3 le = texture_color(S, __float_as_int(lmat.y), lh);
        float w = pl / (pl + pb);
        sreq.valid = true;
        sreq.ray.o = h.p; sreq.ray.d = dl; sreq.ray.time = ray.time; sreq.ray.tmax = tl * (1.0f - 1e-4f); sreq.ray.mseed = r2.w;
        sreq.contrib = ps.throughput * color * le * (pb / pl * w);
      }
    }
    float3 dir = iso ? sample_unit_sphere(u1, u2) : sample_cosine_hemisphere(h.n, u1, u2);
    float pb = iso ? 0.25f / PI_F : fmaxf(0.f, dot(h.n, normalize(dir))) / PI_F;
    if (!(pb > 0.f)) return false;
    ps.throughput = ps.throughput * color;
    ps.prev_pdf_b = pb;
    ray.o = h.p; ray.d = dir; ray.mseed = next_seed;
    return true;
  }
  ps.prev_pdf_b = -1.f;
  if (mtype == B200RT_MAT_LAMBERTIAN) {  // material.rs:92-98,109-112
    float3 dir = h.n + sample_unit_sphere(u1, u2);
    if (fabsf(dir.x) < 1e-8f && fabsf(dir.y) < 1e-8f && fabsf(dir.z) < 1e-8f) dir = h.n;
    ps.throughput = ps.throughput * color;
    ray.o = h.p; ray.d = dir; ray.mseed = next_seed;
    return true;
  }
  if (mtype == B200RT_MAT_ISOTROPIC) {  // material.rs:213-231 (commented out upstream)
    ps.throughput = ps.throughput * color;
    ray.o = h.p; ray.d = sample_unit_sphere(u1, u2); ray.mseed = next_seed;
    return true;
  }
  if (mtype == B200RT_MAT_METAL) {  // material.rs:130-138
    float3 reflected = reflect3(normalize(ray.d), h.n);
    float3 dir = reflected + mat.z * sample_unit_sphere(u1, u2);
    if (!(dot(dir, h.n) > 0.f)) return false;  // Absorb
    ps.throughput = ps.throughput * color;
    ray.o = h.p; ray.d = dir; ray.mseed = next_seed;
    return true;
  }
  // Dielectric, material.rs:161-187
  const float ratio = h.front_face ? 1.0f / mat.z : mat.z;
  const float3 ud = normalize(ray.d);
  const float cos_theta = fminf(dot(-ud, h.n), 1.0f);
  const float sin_theta = sqrtf(fmaxf(0.f, 1.0f - cos_theta * cos_theta));
  const bool cannot_refract = ratio * sin_theta > 1.0f;
  float3 dir = (cannot_refract || schlick(cos_theta, ratio) > u3) ? reflect3(ud, h.n) : refract3(ud, h.n, ratio);
  ps.throughput = ps.throughput * color;
  ray.o = h.p; ray.d = dir; ray.mseed = next_seed;
  return true;
}

// Processes the hit `hr` of `ray`: HitRecord::new, the Philox draws of this bounce, then shade_surface.
// depth bookkeeping (photon_mapper.rs:334-339) is done by the caller.
// In two halves for the shade kernel, which evaluates NOISE textures between them with the warp converged.
DEV void shade_hit_begin(const DevScene& S, const Ray& ray, const HitRaw& hr, HitInfo& h, float4& mat) {
  const int4 meta = __ldg(&S.prim_meta[hr.prim]);
  mat = __ldg(&S.materials[meta.x >> 8]);
  finish_hit(S, ray, hr, h, texture_needs_uv(S, __float_as_int(mat.y)));
}
template <int INTEG = -1>
DEV bool shade_hit_end(const DevScene& S, const RenderConst& P, Ray& ray, const HitInfo& h, const float4 mat, PathState& ps, ShadowReq& sreq, const float turb = -1.f) {
  sreq.valid = false;
  ps.bounce++;
  const uint4 rnd = philox4x32(make_uint4(ps.pixel, ps.sample, (uint32_t)ps.bounce, 0u), P.key);
  return shade_surface<INTEG>(S, P, ray, h, mat, u01(rnd.x), u01(rnd.y), u01(rnd.z), rnd.w, ps, sreq, turb);
}
template <int INTEG = -1>
DEV bool shade_hit(const DevScene& S, const RenderConst& P, Ray& ray, const HitRaw& hr, PathState& ps, ShadowReq& sreq) {
  HitInfo h;
  float4 mat;
  shade_hit_begin(S, ray, hr, h, mat);
  return shade_hit_end<INTEG>(S, P, ray, h, mat, ps, sreq);
}
