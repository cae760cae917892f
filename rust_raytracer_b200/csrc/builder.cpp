// builder.cpp — the `flatten` target of the scene front end: turns the reference's construction calls
// (Sphere{..}, XYRectangle{..}, Cube::new, Transform::new, Mesh::load_obj, ConstantMedium::new, materials,
// textures; raytracer/src/scene.rs:16-112 is the model usage) into the flat tables of b200rt_scene_desc.
// Host only; no CUDA.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <map>

#include "host_common.hpp"

using namespace b200rt;

namespace {

const double PI = 3.14159265358979323846;

int fail(b200rt_builder* b, int code, const std::string& msg) {
  if (b) b->err = msg;
  return -code;
}

int32_t current_instance(const b200rt_builder* b) { return b->xf_instance.empty() ? -1 : b->xf_instance.back(); }

void set3(float* dst, D3 v) { dst[0] = (float)v.x; dst[1] = (float)v.y; dst[2] = (float)v.z; }

int push_prim(b200rt_builder* b, const b200rt_prim& p) {
  if (p.material < 0 || (size_t)p.material >= b->materials.size()) return fail(b, B200RT_ERR_INVALID, "material id out of range");
  b->prims.push_back(p);
  return (int)b->prims.size() - 1;
}

// Every shape keeps its own type under any Transform (transform.rs:17-165): a rotated Cube stays ONE box primitive with
// an instance matrix (the device filters and decides it in object space), so rectangle uv (rectangle.rs:32,70,107) and
// the six-rect tie rule of cube.rs:63-66 survive the rotation.
int add_triangle(b200rt_builder* b, D3 pa, D3 pb, D3 pc, D3 na, D3 nb, D3 nc, int32_t material) {
  b200rt_prim p;
  std::memset(&p, 0, sizeof(p));
  p.type = B200RT_PRIM_TRIANGLE; p.material = material; p.instance = current_instance(b);
  p.aux = (int32_t)(b->tri_normals.size() / 9);
  set3(p.p, pa); set3(p.p + 3, pb); set3(p.p + 6, pc);
  float n[9];
  set3(n, na); set3(n + 3, nb); set3(n + 6, nc);
  int id = push_prim(b, p);
  if (id >= 0) b->tri_normals.insert(b->tri_normals.end(), n, n + 9);
  return id;
}

}  // namespace

extern "C" {

int b200rt_builder_create(b200rt_builder** out) {
  if (!out) return B200RT_ERR_INVALID;
  *out = new (std::nothrow) b200rt_builder();
  return *out ? B200RT_OK : B200RT_ERR_NOMEM;
}
int b200rt_builder_destroy(b200rt_builder* b) { delete b; return B200RT_OK; }
const char* b200rt_builder_error(const b200rt_builder* b) { return b ? b->err.c_str() : "null builder"; }

int b200rt_builder_texture_constant(b200rt_builder* b, double r, double g, double bl) {
  if (!b) return -B200RT_ERR_INVALID;
  b200rt_texture t;
  std::memset(&t, 0, sizeof(t));
  t.type = B200RT_TEX_CONSTANT; t.image = -1;
  t.c0[0] = (float)r; t.c0[1] = (float)g; t.c0[2] = (float)bl;
  b->textures.push_back(t);
  return (int)b->textures.size() - 1;
}
int b200rt_builder_texture_checker(b200rt_builder* b, const double c0[3], const double c1[3]) {
  if (!b || !c0 || !c1) return -B200RT_ERR_INVALID;
  b200rt_texture t;
  std::memset(&t, 0, sizeof(t));
  t.type = B200RT_TEX_CHECKER; t.image = -1;
  for (int i = 0; i < 3; ++i) { t.c0[i] = (float)c0[i]; t.c1[i] = (float)c1[i]; }
  b->textures.push_back(t);
  return (int)b->textures.size() - 1;
}
int b200rt_builder_texture_image(b200rt_builder* b, const uint8_t* rgb8, int32_t width, int32_t height) {
  if (!b || !rgb8 || width <= 0 || height <= 0) return fail(b, B200RT_ERR_INVALID, "texture_image: bad arguments");
  b200rt_image im;
  im.offset = (int32_t)b->image_data.size(); im.width = width; im.height = height; im.pad = 0;
  b->image_data.insert(b->image_data.end(), rgb8, rgb8 + (size_t)3 * width * height);
  while (b->image_data.size() % 4) b->image_data.push_back(0);
  b->images.push_back(im);
  b200rt_texture t;
  std::memset(&t, 0, sizeof(t));
  t.type = B200RT_TEX_IMAGE; t.image = (int32_t)b->images.size() - 1;
  b->textures.push_back(t);
  return (int)b->textures.size() - 1;
}
int b200rt_builder_seed_perlin(b200rt_builder* b, uint64_t seed) {
  if (!b) return B200RT_ERR_INVALID;
  Pcg32 rng(seed, 7);
  for (int i = 0; i < 256; ++i) {
    const double x = rng.range(-1, 1), y = rng.range(-1, 1), z = rng.range(-1, 1);  // sequenced: x, y, z (argument evaluation order is unspecified)
    D3 v = unit(D3(x, y, z));
    b->perlin.ranvec[i][0] = (float)v.x; b->perlin.ranvec[i][1] = (float)v.y; b->perlin.ranvec[i][2] = (float)v.z; b->perlin.ranvec[i][3] = 0;
  }
  int32_t* perms[3] = {b->perlin.perm_x, b->perlin.perm_y, b->perlin.perm_z};
  for (int a = 0; a < 3; ++a) {
    for (int i = 0; i < 256; ++i) perms[a][i] = i;
    for (int i = 255; i > 0; --i) {
      int t = (int)(rng.next() * (i + 1));
      if (t > i) t = i;
      std::swap(perms[a][i], perms[a][t]);
    }
  }
  b->has_perlin = true;
  return B200RT_OK;
}
int b200rt_builder_texture_noise(b200rt_builder* b, double scale) {
  if (!b) return -B200RT_ERR_INVALID;
  if (!b->has_perlin) b200rt_builder_seed_perlin(b, 0x9E3779B97F4A7C15ull);
  b200rt_texture t;
  std::memset(&t, 0, sizeof(t));
  t.type = B200RT_TEX_NOISE; t.image = -1; t.scale = (float)scale;
  b->textures.push_back(t);
  return (int)b->textures.size() - 1;
}
int b200rt_builder_material(b200rt_builder* b, int32_t type, int32_t texture, double param) {
  if (!b) return -B200RT_ERR_INVALID;
  if (type < 0 || type > B200RT_MAT_ISOTROPIC) return fail(b, B200RT_ERR_INVALID, "unknown material type");
  if (texture < 0 || (size_t)texture >= b->textures.size()) return fail(b, B200RT_ERR_INVALID, "texture id out of range");
  b200rt_material m;
  m.type = type; m.texture = texture; m.param = (float)param; m.pad = 0;
  b->materials.push_back(m);
  return (int)b->materials.size() - 1;
}

// transform.rs:17-106 — M = T * S * Rx * Ry * Rz
int b200rt_builder_push_transform(b200rt_builder* b, const double rotate_deg[3], const double scale[3], const double translate[3]) {
  if (!b || !rotate_deg || !scale || !translate) return B200RT_ERR_INVALID;
  double rx = rotate_deg[0] * PI / 180., ry = rotate_deg[1] * PI / 180., rz = rotate_deg[2] * PI / 180.;
  Affine T = Affine::identity(), S = Affine::identity(), RX = Affine::identity(), RY = Affine::identity(), RZ = Affine::identity();
  T.m[3] = translate[0]; T.m[7] = translate[1]; T.m[11] = translate[2];
  S.m[0] = scale[0]; S.m[5] = scale[1]; S.m[10] = scale[2];
  RX.m[5] = std::cos(rx); RX.m[6] = -std::sin(rx); RX.m[9] = std::sin(rx); RX.m[10] = std::cos(rx);
  RY.m[0] = std::cos(ry); RY.m[2] = std::sin(ry); RY.m[8] = -std::sin(ry); RY.m[10] = std::cos(ry);
  RZ.m[0] = std::cos(rz); RZ.m[1] = -std::sin(rz); RZ.m[4] = std::sin(rz); RZ.m[5] = std::cos(rz);
  Affine M = T * S * RX * RY * RZ;
  Affine inv;
  if (!M.inverse(inv)) { b->err = "Invalid transform matrix"; return B200RT_ERR_INVALID; }  // transform.rs:146
  if (!b->xf_stack.empty()) M = b->xf_stack.back() * M;
  b200rt_instance inst;
  std::memcpy(inst.m, M.m, sizeof(inst.m));
  b->instances.push_back(inst);
  b->xf_stack.push_back(M);
  b->xf_instance.push_back((int32_t)b->instances.size() - 1);
  return B200RT_OK;
}
int b200rt_builder_pop_transform(b200rt_builder* b) {
  if (!b || b->xf_stack.empty()) return B200RT_ERR_INVALID;
  b->xf_stack.pop_back();
  b->xf_instance.pop_back();
  return B200RT_OK;
}

int b200rt_builder_sphere(b200rt_builder* b, const double c[3], double radius, int32_t material) {
  if (!b || !c) return -B200RT_ERR_INVALID;
  b200rt_prim p;
  std::memset(&p, 0, sizeof(p));
  p.type = B200RT_PRIM_SPHERE; p.material = material; p.instance = current_instance(b); p.aux = -1;
  set3(p.p, D3(c)); p.p[3] = (float)radius;
  return push_prim(b, p);
}
int b200rt_builder_moving_sphere(b200rt_builder* b, const double c0[3], const double c1[3], double time0, double time1, double radius, int32_t material) {
  if (!b || !c0 || !c1) return -B200RT_ERR_INVALID;
  if (!(time1 > time0)) return fail(b, B200RT_ERR_INVALID, "moving sphere needs time1 > time0");
  b200rt_prim p;
  std::memset(&p, 0, sizeof(p));
  p.type = B200RT_PRIM_MOVING_SPHERE; p.material = material; p.instance = current_instance(b); p.aux = -1;
  set3(p.p, D3(c0)); p.p[3] = (float)radius; set3(p.p + 4, D3(c1)); p.p[7] = (float)time0; p.p[8] = (float)time1;
  return push_prim(b, p);
}
int b200rt_builder_rect(b200rt_builder* b, int32_t axis, double a0, double b0, double a1, double b1, double k, int32_t material) {
  if (!b || axis < 0 || axis > 2) return -B200RT_ERR_INVALID;
  b200rt_prim p;
  std::memset(&p, 0, sizeof(p));
  p.type = axis == 2 ? B200RT_PRIM_XY_RECT : (axis == 1 ? B200RT_PRIM_XZ_RECT : B200RT_PRIM_YZ_RECT);
  p.material = material; p.instance = current_instance(b); p.aux = -1;
  p.p[0] = (float)a0; p.p[1] = (float)b0; p.p[2] = (float)a1; p.p[3] = (float)b1; p.p[4] = (float)k;
  return push_prim(b, p);
}
int b200rt_builder_cube(b200rt_builder* b, const double mn[3], const double mx[3], int32_t material) {
  if (!b || !mn || !mx) return -B200RT_ERR_INVALID;
  b200rt_prim p;
  std::memset(&p, 0, sizeof(p));
  p.type = B200RT_PRIM_BOX; p.material = material; p.instance = current_instance(b); p.aux = -1;
  set3(p.p, D3(mn)); set3(p.p + 3, D3(mx));
  return push_prim(b, p);
}
int b200rt_builder_triangle(b200rt_builder* b, const double pa[3], const double pb[3], const double pc[3], const double na[3], const double nb[3], const double nc[3], int32_t material) {
  if (!b || !pa || !pb || !pc || !na || !nb || !nc) return -B200RT_ERR_INVALID;
  return add_triangle(b, D3(pa), D3(pb), D3(pc), D3(na), D3(nb), D3(nc), material);
}
int b200rt_builder_medium_sphere(b200rt_builder* b, const double c[3], double radius, double density, int32_t phase) {
  if (!b || !c || !(density > 0)) return fail(b, B200RT_ERR_INVALID, "medium_sphere: bad arguments");
  b200rt_prim p;
  std::memset(&p, 0, sizeof(p));
  p.type = B200RT_PRIM_MEDIUM_SPHERE; p.material = phase; p.instance = current_instance(b); p.aux = -1;
  set3(p.p, D3(c)); p.p[3] = (float)radius; p.p[4] = (float)density;
  return push_prim(b, p);
}
int b200rt_builder_medium_cube(b200rt_builder* b, const double mn[3], const double mx[3], double density, int32_t phase) {
  if (!b || !mn || !mx || !(density > 0)) return fail(b, B200RT_ERR_INVALID, "medium_cube: bad arguments");
  b200rt_prim p;
  std::memset(&p, 0, sizeof(p));
  p.type = B200RT_PRIM_MEDIUM_BOX; p.material = phase; p.instance = current_instance(b); p.aux = -1;
  set3(p.p, D3(mn)); set3(p.p + 3, D3(mx)); p.p[6] = (float)density;
  return push_prim(b, p);
}
int b200rt_builder_mark_light(b200rt_builder* b, int32_t prim) { return b200rt_builder_mark_light_scaled(b, prim, 1.0); }
int b200rt_builder_mark_light_scaled(b200rt_builder* b, int32_t prim, double scale) {
  if (!b) return B200RT_ERR_INVALID;
  if (!(scale > 0.0) || !std::isfinite(scale)) { b->err = "mark_light: the light's scale must be positive and finite"; return B200RT_ERR_INVALID; }
  if (prim < 0 || (size_t)prim >= b->prims.size()) { b->err = "mark_light: prim out of range"; return B200RT_ERR_INVALID; }
  const b200rt_prim& p = b->prims[prim];
  bool ok = p.instance < 0 && (p.type == B200RT_PRIM_SPHERE || p.type == B200RT_PRIM_XY_RECT || p.type == B200RT_PRIM_XZ_RECT || p.type == B200RT_PRIM_YZ_RECT);
  if (!ok) { b->err = "only untransformed spheres and axis rectangles can be sampled as lights"; return B200RT_ERR_UNSUPPORTED; }
  b->lights.push_back(prim);
  b->light_scales.push_back((float)scale);
  return B200RT_OK;
}
int b200rt_builder_set_flags(b200rt_builder* b, uint32_t flags) {
  if (!b) return B200RT_ERR_INVALID;
  b->flags = flags;
  return B200RT_OK;
}
int b200rt_builder_finish(b200rt_builder* b, b200rt_scene_desc* out) {
  if (!b || !out) return B200RT_ERR_INVALID;
  if (!b->xf_stack.empty()) { b->err = "finish with an open transform"; return B200RT_ERR_INVALID; }
  std::memset(out, 0, sizeof(*out));
  out->n_prims = (uint32_t)b->prims.size();
  out->n_tri_normals = (uint32_t)(b->tri_normals.size() / 9);
  out->n_materials = (uint32_t)b->materials.size();
  out->n_textures = (uint32_t)b->textures.size();
  out->n_images = (uint32_t)b->images.size();
  out->n_instances = (uint32_t)b->instances.size();
  out->n_lights = (uint32_t)b->lights.size();
  out->flags = b->flags;
  out->image_bytes = b->image_data.size();
  out->prims = b->prims.data();
  out->tri_normals = b->tri_normals.data();
  out->materials = b->materials.data();
  out->textures = b->textures.data();
  out->images = b->images.data();
  out->image_rgb8 = b->image_data.data();
  out->instances = b->instances.data();
  out->light_prims = b->lights.data();
  out->light_scales = b->light_scales.data();
  out->perlin = b->has_perlin ? &b->perlin : nullptr;
  return B200RT_OK;
}

// Camera::new (camera.rs:24-55)
int b200rt_camera_make(const double look_from[3], const double look_at[3], const double vup[3], double vfov_deg, double aspect, double aperture, double focus_dist, double time0, double time1, b200rt_camera* out) {
  if (!look_from || !look_at || !vup || !out) return B200RT_ERR_INVALID;
  double theta = vfov_deg * PI / 180.;
  double h = std::tan(theta / 2.);
  double vh = 2.0 * h, vw = aspect * vh;
  D3 from(look_from), at(look_at);
  D3 w = from - at;
  if (length(w) == 0) return B200RT_ERR_INVALID;  // vec3.rs:88 would panic
  w = unit(w);
  D3 u = cross(D3(vup), w);
  if (length(u) == 0) return B200RT_ERR_INVALID;
  u = unit(u);
  D3 v = cross(w, u);
  D3 horizontal = (focus_dist * vw) * u, vertical = (focus_dist * vh) * v;
  D3 llc = from - horizontal / 2. - vertical / 2. - focus_dist * w;
  auto s3 = [](float* d, D3 a) { d[0] = (float)a.x; d[1] = (float)a.y; d[2] = (float)a.z; };
  s3(out->origin, from); s3(out->lower_left_corner, llc); s3(out->horizontal, horizontal); s3(out->vertical, vertical);
  s3(out->u, u); s3(out->v, v); s3(out->w, w);
  out->lens_radius = (float)(aperture / 2.); out->time0 = (float)time0; out->time1 = (float)time1;
  return B200RT_OK;
}

}  // extern "C"
