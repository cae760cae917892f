/*
 * b200rt.h — C ABI of the B200-native path-tracing backend.
 *
 * This is the drop-in boundary for the per-pixel Monte Carlo loop of
 * BlackCloud37/rust-raytracer.  Every entry point names the reference
 * interface it replaces (paths relative to the reference checkout,
 * `raytracer/src/...`).  A Rust `-sys` crate binds exactly these symbols
 * (see INTEGRATION.md); the C++ mirror of the reference's scene API
 * (`rust_raytracer_b200/csrc/frontend.hpp`) and the ctypes mirror
 * (`rust_raytracer_b200/api.py`) sit on top of the same symbols.
 *
 * Conventions
 *   - every function returns an int status (B200RT_OK == 0); nothing throws or
 *     aborts across the ABI (the reference panics instead: vec3.rs:88,
 *     bvh.rs:43,57, transform.rs:146, mesh.rs:158);
 *   - plain pointers and sizes only; the caller owns host buffers, the library
 *     owns device memory behind opaque handles;
 *   - a context is bound to one CUDA device and is not thread-safe: one context
 *     per host thread (the reference shares an immutable Arc<World> between
 *     pool threads, camera.rs:79-111; a b200rt_scene is likewise immutable);
 *   - there is NO CPU fallback: without a CUDA device every compute entry point
 *     fails with B200RT_ERR_CUDA.
 */
#ifndef B200RT_H
#define B200RT_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200RT_ABI_VERSION 1

/* ---- status codes ------------------------------------------------------ */
#define B200RT_OK 0
#define B200RT_ERR_INVALID 1     /* bad argument / malformed scene table       */
#define B200RT_ERR_CUDA 2        /* CUDA runtime error, or no device           */
#define B200RT_ERR_NOMEM 3
#define B200RT_ERR_UNSUPPORTED 4 /* e.g. non-similarity Transform of a sphere  */
#define B200RT_ERR_IO 5
#define B200RT_ERR_PARSE 6       /* JSON / OBJ syntax                          */

/* ---- scene tables (flat, host side) ------------------------------------ */

/* Primitive kinds.  One row of SURVEY.md §8(a) each. */
enum b200rt_prim_type {
  B200RT_PRIM_SPHERE = 0,        /* objects/sphere.rs:23-62                     */
  B200RT_PRIM_MOVING_SPHERE = 1, /* absent upstream; book-2 semantics           */
  B200RT_PRIM_XY_RECT = 2,       /* objects/rectangle.rs:7-42                   */
  B200RT_PRIM_XZ_RECT = 3,       /* objects/rectangle.rs:44-80                  */
  B200RT_PRIM_YZ_RECT = 4,       /* objects/rectangle.rs:82-117                 */
  B200RT_PRIM_BOX = 5,           /* objects/cube.rs:9-70 (`Cube`)               */
  B200RT_PRIM_TRIANGLE = 6,      /* objects/mesh.rs:8-142                       */
  B200RT_PRIM_MEDIUM_SPHERE = 7, /* objects/medium.rs:9-57, sphere boundary     */
  B200RT_PRIM_MEDIUM_BOX = 8     /* objects/medium.rs:9-57, Cube boundary       */
};

/*
 * One primitive, in OBJECT space (64 bytes).  p[] by type:
 *   SPHERE         cx cy cz r
 *   MOVING_SPHERE  c0x c0y c0z r  c1x c1y c1z  time0 time1
 *   XY_RECT        x0 y0 x1 y1 k(z)     XZ_RECT  x0 z0 x1 z1 k(y)
 *   YZ_RECT        y0 z0 y1 z1 k(x)
 *   BOX            minx miny minz maxx maxy maxz
 *   TRIANGLE       ax ay az bx by bz cx cy cz ; aux = index into tri_normals
 *   MEDIUM_SPHERE  cx cy cz r density    (material = phase function)
 *   MEDIUM_BOX     minx miny minz maxx maxy maxz density
 * instance: index into instances[] (objects/transform.rs) or -1 for none.
 */
typedef struct b200rt_prim {
  int32_t type;
  int32_t material;
  int32_t instance;
  int32_t aux;
  float p[12];
} b200rt_prim;

enum b200rt_material_type {
  B200RT_MAT_LAMBERTIAN = 0,    /* material.rs:89-113  */
  B200RT_MAT_METAL = 1,         /* material.rs:115-139 */
  B200RT_MAT_DIELECTRIC = 2,    /* material.rs:141-188 */
  B200RT_MAT_DIFFUSE_LIGHT = 3, /* material.rs:191-212 */
  B200RT_MAT_ISOTROPIC = 4      /* material.rs:213-231 (commented out upstream) */
};

/* param: Metal fuzz / Dielectric ir; unused otherwise. */
typedef struct b200rt_material {
  int32_t type;
  int32_t texture;
  float param;
  float pad;
} b200rt_material;

enum b200rt_texture_type {
  B200RT_TEX_CONSTANT = 0, /* material.rs:48,52-56 */
  B200RT_TEX_CHECKER = 1,  /* material.rs:49,58-68 : sin(10x)sin(10y)sin(10z) < 0 -> c0 */
  B200RT_TEX_IMAGE = 2,    /* material.rs:50,70-84 */
  B200RT_TEX_NOISE = 3     /* Perlin marble, absent upstream; book-2 semantics */
};

typedef struct b200rt_texture {
  int32_t type;
  int32_t image; /* IMAGE: index into images[]                    */
  float scale;   /* NOISE: frequency                               */
  float pad;
  float c0[4]; /* CONSTANT colour / CHECKER first colour         */
  float c1[4]; /* CHECKER second colour                          */
} b200rt_texture;

typedef struct b200rt_image {
  int32_t offset; /* byte offset of texel (0,0) in image_rgb8      */
  int32_t width;
  int32_t height;
  int32_t pad;
} b200rt_image;

/* Forward matrix M = T*S*Rx*Ry*Rz, row-major 3x4 (transform.rs:28-106). */
typedef struct b200rt_instance {
  double m[12];
} b200rt_instance;

/* Perlin lattice (256 gradient vectors + 3 permutations). */
typedef struct b200rt_perlin {
  float ranvec[256][4];
  int32_t perm_x[256];
  int32_t perm_y[256];
  int32_t perm_z[256];
} b200rt_perlin;

#define B200RT_SCENE_FLAT_LIST 1u /* linear closest-hit scan, hit.rs:56-67 (no BVH) */

/*
 * The scene hand-off: what `World::new(list, cam, lights)` (world.rs:15-25)
 * receives, flattened.  All pointers are host memory, read during
 * b200rt_scene_create only.
 */
typedef struct b200rt_scene_desc {
  uint32_t n_prims;
  uint32_t n_tri_normals; /* entries of 9 floats (na nb nc), object space */
  uint32_t n_materials;
  uint32_t n_textures;
  uint32_t n_images;
  uint32_t n_instances;
  uint32_t n_lights;
  uint32_t flags;
  uint64_t image_bytes;
  const b200rt_prim* prims;
  const float* tri_normals;
  const b200rt_material* materials;
  const b200rt_texture* textures;
  const b200rt_image* images;
  const uint8_t* image_rgb8;
  const b200rt_instance* instances;
  const int32_t* light_prims;  /* prim indices sampled by the light PDF  */
  const b200rt_perlin* perlin; /* may be NULL when no NOISE texture      */
  const float* light_scales;   /* n_lights, or NULL (all 1): the `scale` of XZRectLight::new / SphereDiffuseLight::new
                                  (light.rs:71,131) per light, RELATIVE to b200rt_sppm_params.light_scale (photon power and the
                                  choice between lights; the light PDFs of the path integrators do not use it) */
} b200rt_scene_desc;

/* camera.rs:11-21 plus a shutter interval (MovingSphere). */
typedef struct b200rt_camera {
  float origin[3];
  float lower_left_corner[3];
  float horizontal[3];
  float vertical[3];
  float u[3];
  float v[3];
  float w[3];
  float lens_radius;
  float time0;
  float time1;
} b200rt_camera;

enum b200rt_integrator {
  B200RT_INTEGRATOR_PATH = 0,    /* loop skeleton photon_mapper.rs:327-365, continued after diffuse hits */
  B200RT_INTEGRATOR_MIXTURE = 1, /* book-3 MixturePDF: 1/2 light + 1/2 cosine, one ray per bounce       */
  B200RT_INTEGRATOR_NEE = 2      /* light-sample + shadow ray with MIS (light.rs:107-124,170-183)       */
};

enum b200rt_pipeline {
  B200RT_PIPELINE_WAVEFRONT = 0, /* generate / extend / shade / shadow / compact over SoA ray queues */
  B200RT_PIPELINE_MEGAKERNEL = 1 /* one persistent thread per path; bring-up and cross-check       */
};

#define B200RT_BACKGROUND_CONSTANT 0
#define B200RT_BACKGROUND_SKY 1 /* book-1 gradient: lerp(white, background, 0.5*(unit(d).y+1)) */

#define B200RT_FLAG_LIGHT_SCATTERS 1u /* DiffuseLight also scatters diffusely with 1/pi, material.rs:201-208 */

/* Render parameters: the compile-time constants of main.rs:34-46, camera.rs:73, photon_mapper.rs:334-335. */
typedef struct b200rt_params {
  int32_t width;
  int32_t height;
  int32_t spp_begin; /* global index of the first sample of this call (multi-GPU shard) */
  int32_t spp_count;
  int32_t max_depth; /* 50 upstream */
  int32_t integrator;
  int32_t pipeline;
  int32_t background_mode;
  float t_min; /* 0.001 upstream */
  float background[3];
  uint32_t flags;
  uint32_t pad;
  uint64_t seed;
} b200rt_params;

typedef struct b200rt_stats {
  uint64_t paths;
  uint64_t extend_rays;
  uint64_t shadow_rays;
  uint64_t kernel_launches;
  float ms_total; /* CUDA events on the context stream around the whole render */
  float ms_generate;
  float ms_extend;
  float ms_shade;
  float ms_shadow;
  float ms_other;
  uint32_t wavefront_iterations;
  uint32_t pipelines; /* concurrent wavefront pipelines this call ran (1 or 2) */
  /* With 2 pipelines the stage times above are the length of the UNION of that stage's CUDA-event intervals over the
   * pipelines (time during which at least one pipeline was in the stage; they overlap and add up to more than ms_total).
   * The *_sum fields add up every launch's own event-pair duration instead (queueing behind the other pipeline's blocks
   * included): generate + control | extend | control + shade | shadow. */
  float ms_generate_sum;
  float ms_extend_sum;
  float ms_shade_sum;
  float ms_shadow_sum;
} b200rt_stats;

/* Result of `World::hit` (world.rs:27-29, hit.rs:7-39) for one ray. prim = index into desc.prims or -1. */
typedef struct b200rt_hit {
  float t;
  int32_t prim;
  float nx, ny, nz; /* unit, facing the ray (hit.rs:25-33) */
  int32_t front_face;
  float u, v;
} b200rt_hit;

typedef struct b200rt_ctx b200rt_ctx;
typedef struct b200rt_scene b200rt_scene;
typedef struct b200rt_builder b200rt_builder;

/* ---- context ------------------------------------------------------------ */
int b200rt_abi_version(void);
int b200rt_device_count(int* count);
int b200rt_ctx_create(int device, b200rt_ctx** out);
/* Destroy the context's scenes first.  Scenes that are still alive are orphaned, not left dangling: every entry point then
 * returns B200RT_ERR_INVALID for them (b200rt_global_error says why) and b200rt_scene_destroy still frees their tables. */
int b200rt_ctx_destroy(b200rt_ctx* ctx);
/* Run all work of this context on an existing cudaStream_t (NULL = default stream). */
int b200rt_ctx_set_stream(b200rt_ctx* ctx, void* cuda_stream);
const char* b200rt_last_error(const b200rt_ctx* ctx);
const char* b200rt_global_error(void); /* for calls that have no context */
/* Device facts for roofline arithmetic: SM count, clock kHz, L2 bytes. */
int b200rt_ctx_device_info(const b200rt_ctx* ctx, int32_t* sm_count, int32_t* clock_khz, int64_t* l2_bytes, char* name, int name_len);

/* ---- scene: replaces World::new (world.rs:15-25) ------------------------ */
/* Validates the tables, bakes instance transforms (transform.rs:110-148), builds the wide BVH
 * (replaces BVHNode::new, bvh.rs:60-83) and uploads everything. */
int b200rt_scene_create(b200rt_ctx* ctx, const b200rt_scene_desc* desc, b200rt_scene** out);
int b200rt_scene_destroy(b200rt_scene* scene);
/* Bytes uploaded host->device by scene_create, BVH node count, device prim count. */
int b200rt_scene_info(const b200rt_scene* scene, uint64_t* h2d_bytes, uint32_t* bvh_nodes, uint32_t* device_prims);

/* Host only (no CUDA): the wide BVH b200rt_scene_create would build for `desc` in place of BVHNode::new (bvh.rs:60-83) —
 * node count, depth in BVH4 levels, and the traversal-stack entries the deepest path can need (3 per level) against the
 * per-thread capacity.  scene_create rebuilds an over-deep tree under a depth budget, so needed <= capacity always. */
int b200rt_scene_compile_stats(const b200rt_scene_desc* desc, uint32_t* bvh_nodes, uint32_t* bvh4_depth,
                               uint32_t* stack_needed, uint32_t* stack_capacity);

/* Host only (no CUDA): the surface-area estimate of that BVH — BVH4 node steps and primitive tests a random line through the
 * scene's box pays (sum of node areas, and of leaf area x leaf size, over the root's area) — and the scene's DENSE BOX if it has
 * one: a box of at most half of the root's surface holding >= 90 % of the BVH's primitives (a 300,000-triangle mesh in a
 * room of a dozen walls).  capture_image's wavefront queues then keep the rays that cross it apart from the rest
 * (DESIGN.md section 4).  No counterpart in the reference (bvh.rs:60-83 splits at the median of a random axis).
 * Any out pointer may be NULL; dense_lo / dense_hi: 3 floats each, zero when there is no dense box. */
int b200rt_scene_bvh_quality(const b200rt_scene_desc* desc, float* node_steps, float* prim_tests, int32_t* has_dense_box,
                             float* dense_lo, float* dense_hi);

/* ---- World::hit as a batch operator (world.rs:27-29) -------------------- */
/* o_time: n x (ox,oy,oz,time); d_tmax: n x (dx,dy,dz,t_max); medium_seed: n x u32 or NULL (the draw
 * ConstantMedium::hit takes inside `hit`, medium.rs:37-38, derived as u01(hash(seed, prim))).
 * Host pointers; copies in and out. */
int b200rt_intersect_batch(b200rt_scene* scene, const float* o_time, const float* d_tmax, const uint32_t* medium_seed,
                           uint64_t n, float t_min, b200rt_hit* out);
/* Same with device pointers, asynchronous on the context stream. */
int b200rt_intersect_batch_device(b200rt_scene* scene, const float* o_time, const float* d_tmax,
                                  const uint32_t* medium_seed, uint64_t n, float t_min, b200rt_hit* out);

/* ---- Camera::capture_image (camera.rs:66-128) --------------------------- */
/* Adds spp_count samples per pixel into accum (width*height float4: r,g,b sums and sample count).
 * _device: accum is device memory, left resident (the multi-GPU reduce runs on it);
 * host variant: zero-initialised device buffer, rendered, copied back into accum_host. */
int b200rt_render_device(b200rt_scene* scene, const b200rt_camera* cam, const b200rt_params* params, float* accum_dev,
                         b200rt_stats* stats);
int b200rt_render(b200rt_scene* scene, const b200rt_camera* cam, const b200rt_params* params, float* accum_host,
                  b200rt_stats* stats);
/* ---- capture_image without any device allocation ----
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same contract as b200rt_render_device
 * (wavefront pipeline, any integrator; accum_dev is the caller's device buffer), but the wavefront queues live in ONE
 * file-scope static __device__ array of B200RT_STATIC_POOL_BYTES that every GPU which loads the library holds until exit;
 * the call makes no cudaMalloc / cudaFree.  Limit: width * height * spp_count <= B200RT_STATIC_MAX_PATHS per call, larger
 * requests fail with B200RT_ERR_INVALID and a message.  All calls share the array (it keeps the last call's contents; what
 * must start at zero is cleared per call): calls that overlap on one GPU race on it — serialise them. */
#define B200RT_STATIC_POOL_POSITIONS 65536u        /* paths in flight */
#define B200RT_STATIC_POOL_BYTES (13u << 20)       /* 13 MiB of device memory per GPU */
#define B200RT_STATIC_MAX_PATHS (1ull << 24)       /* per call */
int b200rt_render_device_static(b200rt_scene* scene, const b200rt_camera* cam, const b200rt_params* params, float* accum_dev,
                                b200rt_stats* stats);

/* ---- capture_image over several GPUs of one node ---- */
#define B200RT_REDUCE_NONE 0 /* one device: nothing to sum */
#define B200RT_REDUCE_NCCL 1 /* ncclReduce over NVLink (libnccl.so.2 loaded at run time) */
#define B200RT_REDUCE_PEER 2 /* the root GPU loads the peers' buffers over NVLink (P2P) and adds them */
typedef struct b200rt_multi_stats {
  uint64_t paths;       /* over all devices */
  uint64_t extend_rays;
  uint64_t shadow_rays;
  float ms_render_max;  /* slowest device's CUDA-event render time */
  float ms_reduce;      /* CUDA events on the root stream around the sum-reduce */
  float ms_wall;        /* host clock around the whole call (first call: includes ncclCommInitAll) */
  int32_t reduce_path;  /* B200RT_REDUCE_* */
  int32_t n_dev;
  uint32_t pad;
} b200rt_multi_stats;
/* Camera::capture_image (camera.rs:66-128) sharded over n_dev GPUs: scenes[g] is the same scene created on n_dev
 * different devices (one context each); device g renders the g-th contiguous part of the global sample range
 * [spp_begin, spp_begin + spp_count) of every pixel — the reference's own decomposition is row bands over a thread
 * pool (camera.rs:74-111) — on its own host thread, then the W x H float4 accumulation buffers are summed onto
 * scenes[0]'s device with ONE reduce (peer-memory loads over NVLink; NCCL without peer access or with B200RT_MULTI_REDUCE=nccl)
 * and copied into accum_host.  The union equals the one-device render up to float summation order (the Philox
 * counter carries the global sample index).  stats_per_device (n_dev entries) and multi may be NULL. */
int b200rt_render_multi(int32_t n_dev, b200rt_scene* const* scenes, const b200rt_camera* cam,
                        const b200rt_params* params, float* accum_host, b200rt_stats* stats_per_device,
                        b200rt_multi_stats* multi);
/* ---- SPPMIntegrator (integrator/photon_mapper.rs:17-366) ---- */
typedef struct b200rt_sppm_params {
  int32_t iterations;       /* max_iter_cnt = 50 (photon_mapper.rs:151) */
  int32_t photons_per_iter; /* photon_per_iter = 500000 (photon_mapper.rs:152) */
  int32_t k_global;         /* GLOBAL_INIT_PHOTONS = 100 (photon_mapper.rs:18); <= 128 */
  int32_t k_caustic;        /* CAUSTIC_INIT_PHOTONS = 50 (photon_mapper.rs:19); <= 128 */
  double alpha;             /* ALPHA = 0.7 (photon_mapper.rs:17) */
  double light_scale;       /* `scale` of XZRectLight::new / SphereDiffuseLight::new (1e6 in scene.rs:26-32) */
} b200rt_sppm_params;
typedef struct b200rt_sppm_stats {
  uint64_t global_photons; /* stored over all iterations */
  uint64_t caustic_photons;
  uint64_t photon_rays;
  uint64_t eye_rays; /* eye passes + final pass */
  float ms_total;    /* CUDA events on the context stream */
  float ms_photons;
  float ms_grid;
  float ms_eye;
  float ms_final;
  uint32_t pad;
} b200rt_sppm_stats;
/* SPPMIntegrator::new(world) followed by Camera::capture_image(integrator) (main.rs:51-53): `iterations` rounds of
 * photon map + one eye ray per pixel (SPPM::update), then spp_count samples per pixel of SPPMIntegrator::sample_ray.
 * The lights are the scene's marked lights (World::new's light list); rectangle lights emit towards the negative
 * axis like XZRectLight (light.rs:157-165).  accum_host: width*height float4 (r,g,b sums and sample count);
 * pixel_stats_host (may be NULL): width*height x 10 floats: global flux rgb, radius2, photons, caustic ditto. */
int b200rt_render_sppm(b200rt_scene* scene, const b200rt_camera* cam, const b200rt_params* params,
                       const b200rt_sppm_params* sppm, float* accum_host, float* pixel_stats_host,
                       b200rt_sppm_stats* stats);

/* The same integrator over n_dev GPUs of one node ("SPPM would need a photon all-gather per iteration";
 * the reference runs photon_mapper.rs:158-233 on one thread pool).  scenes[g]: the same scene on n_dev contexts (normally one
 * device each; several contexts on one device are accepted).  Per iteration the photons are sharded by id, each device's grid
 * build reads all shards in place over NVLink peer mappings (the all-gather is fused into the counting sort), and the pixels'
 * SPPM statistics are sharded by pixel index; the final pass is sharded by sample index and summed on scenes[0]'s device.
 * The stored photon counts and the ray counts equal the one-device call's exactly (every draw is keyed by the global photon /
 * pixel id); the image up to float summation order.  Needs peer access between the devices (no NCCL, no host staging).
 * Here the stats times are HOST clocks: ms_total around the whole call (buffer set-up and release included), ms_eye over the
 * SPPM iterations (photon passes, grid builds and eye passes of all devices), ms_final over everything after them (statistics
 * gather, final pass, reduce, copies out, release); ms_photons and ms_grid are not filled. */
int b200rt_render_sppm_multi(int32_t n_dev, b200rt_scene* const* scenes, const b200rt_camera* cam,
                             const b200rt_params* params, const b200rt_sppm_params* sppm, float* accum_host,
                             float* pixel_stats_host, b200rt_sppm_stats* stats);

/* From<Vec3> for Rgb<u8> (vec3.rs:223-231): floor(clamp(sqrt(sum/count),0,1)*255). Device kernel; host buffers. */
int b200rt_tonemap_rgb8(b200rt_ctx* ctx, const float* accum_host, int32_t width, int32_t height, uint8_t* rgb_out);
int b200rt_tonemap_rgb8_device(b200rt_ctx* ctx, const float* accum_dev, int32_t width, int32_t height, uint8_t* rgb_dev);

/* RgbImage::save (main.rs:54: `result.save("output/test.png")`): 8-bit RGB PNG, host only, no dependency. */
int b200rt_write_png(const char* path, const uint8_t* rgb8, int32_t width, int32_t height);

/* Camera::new (camera.rs:24-55), computed in f64 then rounded. */
int b200rt_camera_make(const double look_from[3], const double look_at[3], const double vup[3], double vfov_deg,
                       double aspect, double aperture, double focus_dist, double time0, double time1,
                       b200rt_camera* out);

/* ---- function-level parity operator (tests) ---- */
/* Evaluates single functions of the hot path on the device at caller-supplied inputs, through the same device functions
 * the kernels call, so that they can be compared with the reference's point for point:
 *   TEXTURE  Texture::get_color (material.rs:52-84)       in 8 floats: texture id (i32 bits), p.xyz, u, v, -, -
 *                                                         out 4: rgb, -
 *   SCATTER  Material::scatter + emitted (material.rs:86-212) with the bounce's three uniforms injected
 *                                                         in 20: material id (bits), front_face (bits), p.xyz, n.xyz (unit,
 *                                                         facing the ray), u, v, incoming d.xyz, u1, u2, u3, time, -, -, -
 *                                                         out 16: continues, dir.xyz | attenuation rgb, - | emitted rgb, - |
 *                                                         the on-sphere sample made of (u1, u2) xyz, -
 *   CAMERA   Camera::get_ray + jitter (camera.rs:57-64, 97-99)  in 2 u32: pixel, sample
 *                                                         out 16: o.xyz, time | d.xyz, - | jitter x, y, lens disk x, y | time draw
 *   SAMPLERS the sampling transforms standing in for vec3.rs:111-162   in 8: u1, u2, n.xyz, -, -, -
 *                                                         out 12: on-sphere xyz, - | cosine hemisphere about n xyz, - | disk xy
 * Host pointers; cam / params supply the camera, image size, seed and flags. */
enum b200rt_eval_kind { B200RT_EVAL_TEXTURE = 0, B200RT_EVAL_SCATTER = 1, B200RT_EVAL_CAMERA = 2, B200RT_EVAL_SAMPLERS = 3 };
int b200rt_debug_eval(b200rt_scene* scene, const b200rt_camera* cam, const b200rt_params* params, int32_t kind, uint64_t n,
                      const float* in_host, float* out_host);

/* FP32 FMA-issue microbenchmark (roofline denominator). Returns TFLOP/s (FMA = 2). */
int b200rt_measure_fp32_peak(b200rt_ctx* ctx, double* tflops, double* sm_mhz_effective);

/* ---- scene builder: the `flatten` target of the front end --------------- */
/* Host only (no CUDA needed). Every add_* returns the new id (>= 0) or a negative status. */
int b200rt_builder_create(b200rt_builder** out);
int b200rt_builder_destroy(b200rt_builder* b);
const char* b200rt_builder_error(const b200rt_builder* b);
int b200rt_builder_texture_constant(b200rt_builder* b, double r, double g, double bl);
int b200rt_builder_texture_checker(b200rt_builder* b, const double c0[3], const double c1[3]);
int b200rt_builder_texture_image(b200rt_builder* b, const uint8_t* rgb8, int32_t width, int32_t height);
int b200rt_builder_texture_noise(b200rt_builder* b, double scale);
int b200rt_builder_material(b200rt_builder* b, int32_t type, int32_t texture, double param);
/* Transform::new(rotate_in_degree, scale, translate, obj) (transform.rs:17-148): pushes M = T*S*Rx*Ry*Rz;
 * primitives added until the matching pop carry it.  Nested pushes compose (outer * inner). */
int b200rt_builder_push_transform(b200rt_builder* b, const double rotate_deg[3], const double scale[3],
                                  const double translate[3]);
int b200rt_builder_pop_transform(b200rt_builder* b);
int b200rt_builder_sphere(b200rt_builder* b, const double center[3], double radius, int32_t material);
int b200rt_builder_moving_sphere(b200rt_builder* b, const double center0[3], const double center1[3], double time0,
                                 double time1, double radius, int32_t material);
/* axis: 0 = YZ (k = x), 1 = XZ (k = y), 2 = XY (k = z); a/b are the in-plane bounds in reference order. */
int b200rt_builder_rect(b200rt_builder* b, int32_t axis, double a0, double b0, double a1, double b1, double k,
                        int32_t material);
int b200rt_builder_cube(b200rt_builder* b, const double box_min[3], const double box_max[3], int32_t material);
int b200rt_builder_triangle(b200rt_builder* b, const double pa[3], const double pb[3], const double pc[3],
                            const double na[3], const double nb[3], const double nc[3], int32_t material);
/* Mesh::load_obj (mesh.rs:148-199): first model only, triangulated, single index; vertex normals are required
 * by Triangle::hit (mesh.rs:62-64) — with make_normals != 0 missing normals are generated (area-weighted). */
int b200rt_builder_mesh_obj(b200rt_builder* b, const char* path, int32_t material, int32_t make_normals,
                            int32_t subdivide);
/* ConstantMedium::new(density, boundary, phase) with a Sphere / Cube boundary (medium.rs:15-23). */
int b200rt_builder_medium_sphere(b200rt_builder* b, const double center[3], double radius, double density,
                                 int32_t phase_material);
int b200rt_builder_medium_cube(b200rt_builder* b, const double box_min[3], const double box_max[3], double density,
                               int32_t phase_material);
int b200rt_builder_mark_light(b200rt_builder* b, int32_t prim);
/* The same with the light's own `scale` (light.rs:71,131), relative to b200rt_sppm_params.light_scale: a scene whose lights
 * carry different scales upstream marks each with its scale and renders with light_scale = 1. */
int b200rt_builder_mark_light_scaled(b200rt_builder* b, int32_t prim, double scale);
int b200rt_builder_set_flags(b200rt_builder* b, uint32_t flags);
int b200rt_builder_seed_perlin(b200rt_builder* b, uint64_t seed);
/* Appendix-B JSON scene (data/scene_*.json): fills the builder and the camera. */
int b200rt_builder_load_json(b200rt_builder* b, const char* text, uint64_t len, b200rt_camera* cam_out);
/* The YAML twins (data/scene_*.yaml, README.md:86-89: same objects as the JSON files; block-style subset). */
int b200rt_builder_load_yaml(b200rt_builder* b, const char* text, uint64_t len, b200rt_camera* cam_out);
/* Canonical workloads C1..C5: "book1", "cornell_ref", "cornell_book3", "book2", "mesh".
 * aux_path: OBJ path for "mesh"/"cornell_ref"; earth_*: decoded earthmap for "book2" (may be NULL). */
int b200rt_builder_make_scene(b200rt_builder* b, const char* name, uint64_t seed, const char* aux_path,
                              const uint8_t* earth_rgb8, int32_t earth_w, int32_t earth_h, double aspect,
                              b200rt_camera* cam_out, b200rt_params* params_out);
/* The finished tables; pointers stay valid until the builder is destroyed or modified. */
int b200rt_builder_finish(b200rt_builder* b, b200rt_scene_desc* out);

#ifdef __cplusplus
}
#endif
#endif /* B200RT_H */
