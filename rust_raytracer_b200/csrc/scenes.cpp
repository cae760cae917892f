// scenes.cpp — the canonical workloads C1..C5 of SURVEY.md §8(d), written against the C++ mirror of the
// reference's construction API (frontend.hpp) the way raytracer/src/scene.rs:16-112 writes its Cornell box.
// Only "cornell_ref" exists upstream; the book scenes follow the public book series the reference's README
// cites and are OUR fixtures (deterministic in `seed`).
#include <cstdio>

#include "frontend.hpp"
#include "host_common.hpp"

using namespace b200rt::api;
using b200rt::Pcg32;
using std::make_shared;
using std::make_pair;

namespace {

void default_params(b200rt_params* p, int w, int h, int spp, int integrator, int bg_mode, double r, double g, double b) {
  if (!p) return;
  std::memset(p, 0, sizeof(*p));
  p->width = w; p->height = h; p->spp_begin = 0; p->spp_count = spp;
  p->max_depth = 50;  // photon_mapper.rs:334
  p->t_min = 0.001f;  // photon_mapper.rs:335
  p->integrator = integrator; p->pipeline = B200RT_PIPELINE_WAVEFRONT; p->background_mode = bg_mode;
  p->background[0] = (float)r; p->background[1] = (float)g; p->background[2] = (float)b;
  p->seed = 1;
}

// C1: book-1 final random-spheres scene.
void book1(SceneBuilder& sb, uint64_t seed, double aspect, b200rt_camera* cam, b200rt_params* params) {
  Pcg32 rng(seed);
  HitableList world;
  MaterialPtr ground = make_shared<Lambertian>(ConstantTexture(Vec3(0.5, 0.5, 0.5)));
  world.push_back(make_shared<Sphere>(Vec3(0, -1000, 0), 1000, ground));
  for (int a = -11; a < 11; ++a)
    for (int b = -11; b < 11; ++b) {
      // every draw is sequenced explicitly (the evaluation order of function arguments is unspecified in C++): the fixture
      // must be the same whatever the compiler, and the oracle's own generator (oracle/oracle_scenes.inc) draws in this order
      const double choose = rng.next();
      const double cx = a + 0.9 * rng.next(), cz = b + 0.9 * rng.next();
      Vec3 c(cx, 0.2, cz);
      double dx = c.x - 4, dy = c.y - 0.2, dz = c.z;
      if (std::sqrt(dx * dx + dy * dy + dz * dz) <= 0.9) continue;
      MaterialPtr m;
      if (choose < 0.8) { const double r0 = rng.next() * rng.next(), g0 = rng.next() * rng.next(), b0 = rng.next() * rng.next(); m = make_shared<Lambertian>(ConstantTexture(Vec3(r0, g0, b0))); }
      else if (choose < 0.95) { const double r0 = rng.range(0.5, 1), g0 = rng.range(0.5, 1), b0 = rng.range(0.5, 1); const double fuzz = rng.range(0, 0.5); m = make_shared<Metal>(ConstantTexture(Vec3(r0, g0, b0)), fuzz); }
      else m = make_shared<Dielectric>(1.5, ConstantTexture(Vec3::ones()));
      world.push_back(make_shared<Sphere>(c, 0.2, m));
    }
  world.push_back(make_shared<Sphere>(Vec3(0, 1, 0), 1.0, make_shared<Dielectric>(1.5, ConstantTexture(Vec3::ones()))));
  world.push_back(make_shared<Sphere>(Vec3(-4, 1, 0), 1.0, make_shared<Lambertian>(ConstantTexture(Vec3(0.4, 0.2, 0.1)))));
  world.push_back(make_shared<Sphere>(Vec3(4, 1, 0), 1.0, make_shared<Metal>(ConstantTexture(Vec3(0.7, 0.6, 0.5)), 0.0)));
  World w(world, Camera(make_pair(Vec3(13, 2, 3), Vec3(0, 0, 0)), Vec3(0, 1, 0), 20, aspect > 0 ? aspect : 16.0 / 9.0, 0.1, 10.0), {}, sb);
  *cam = w.cam.cam;
  default_params(params, 400, 225, 100, B200RT_INTEGRATOR_PATH, B200RT_BACKGROUND_SKY, 0.5, 0.7, 1.0);
}

// The five walls of scene.rs:33-66.
void cornell_walls(HitableList& list, MaterialPtr white) {
  MaterialPtr red = make_shared<Lambertian>(ConstantTexture(Vec3(0.75, 0.25, 0.25)));
  MaterialPtr blue = make_shared<Lambertian>(ConstantTexture(Vec3(0.25, 0.25, 0.75)));
  list.push_back(make_shared<YZRectangle>(make_pair(0.0, 0.0), make_pair(555.0, 555.0), 555., red));
  list.push_back(make_shared<YZRectangle>(make_pair(0., 0.), make_pair(555., 555.), 0., blue));
  list.push_back(make_shared<XZRectangle>(make_pair(0., 0.), make_pair(555., 555.), 0., white));
  list.push_back(make_shared<XZRectangle>(make_pair(0., 0.), make_pair(555., 555.), 555., white));
  list.push_back(make_shared<XYRectangle>(make_pair(0., 0.), make_pair(555., 555.), 555., white));
}
// XZRectLight::new((213,227),(343,332),554, flux, scale) (light.rs:134-146, scene.rs:26-32): an XZRectangle with a DiffuseLight
HitablePtr cornell_light(double emit) {
  return make_shared<XZRectangle>(make_pair(213., 227.), make_pair(343., 332.), 554., make_shared<DiffuseLight>(ConstantTexture(Vec3(emit, emit, emit))));
}

// The reference's own scene (scene.rs:16-112): glass + mirror spheres, Transform(Mesh(cube.obj)), Cube.
// Emission: the reference's light carries flux (1,1,1) x scale 1e6 for its photon pass; a path tracer uses
// the DiffuseLight colour directly, so the fixture uses (15,15,15).
void cornell_ref(SceneBuilder& sb, const char* obj_path, double aspect, b200rt_camera* cam, b200rt_params* params) {
  MaterialPtr white = make_shared<Lambertian>(ConstantTexture(Vec3(0.75, 0.75, 0.75)));
  HitableList list;
  cornell_walls(list, white);                                                                                                                     // scene.rs:33-66
  list.push_back(make_shared<Sphere>(Vec3(140., 100., 240.), 100., make_shared<Dielectric>(1.5, ConstantTexture(Vec3(0.999, 0.999, 0.999)))));  // scene.rs:67-74
  list.push_back(make_shared<Sphere>(Vec3(400., 100., 360.), 100., make_shared<Metal>(ConstantTexture(Vec3(0.999, 0.999, 0.999)), 0.)));        // scene.rs:75-82
  HitablePtr light = cornell_light(15.0);                                                                                                         // scene.rs:83
  list.push_back(light);
  list.push_back(make_shared<Transform>(Vec3::zero(), Vec3::ones() * 50., Vec3(100., 50., 100.), Mesh::load_obj(obj_path ? obj_path : "data/mesh/cube.obj", white)));
  list.push_back(make_shared<Cube>(Vec3(300., 0., 100.), Vec3(380., 100., 180.), white));
  World w(list, Camera(make_pair(Vec3(278., 278., -800.), Vec3(278., 278., 278.)), Vec3(0., 1., 0.), 50., aspect > 0 ? aspect : 1.0, 0.0, 10.0), {light}, sb);
  *cam = w.cam.cam;
  default_params(params, 800, 800, 256, B200RT_INTEGRATOR_PATH, B200RT_BACKGROUND_CONSTANT, 0, 0, 0);  // main.rs:34-35, camera.rs:73
}

// C2: Cornell box with the book-3 contents (two rotated boxes) and the MixturePDF estimator.
void cornell_book3(SceneBuilder& sb, double aspect, b200rt_camera* cam, b200rt_params* params) {
  MaterialPtr white = make_shared<Lambertian>(ConstantTexture(Vec3(0.73, 0.73, 0.73)));
  HitableList list;
  cornell_walls(list, white);
  HitablePtr light = cornell_light(15.0);
  list.push_back(light);
  list.push_back(make_shared<Transform>(Vec3(0, 15, 0), Vec3::ones(), Vec3(265, 0, 295), make_shared<Cube>(Vec3(0, 0, 0), Vec3(165, 330, 165), white)));
  list.push_back(make_shared<Transform>(Vec3(0, -18, 0), Vec3::ones(), Vec3(130, 0, 65), make_shared<Cube>(Vec3(0, 0, 0), Vec3(165, 165, 165), white)));
  World w(list, Camera(make_pair(Vec3(278., 278., -800.), Vec3(278., 278., 0.)), Vec3(0., 1., 0.), 40., aspect > 0 ? aspect : 1.0, 0.0, 10.0), {light}, sb);
  *cam = w.cam.cam;
  default_params(params, 600, 600, 1000, B200RT_INTEGRATOR_MIXTURE, B200RT_BACKGROUND_CONSTANT, 0, 0, 0);
}

// C4: book-2 final scene.
void book2(SceneBuilder& sb, uint64_t seed, const uint8_t* earth, int ew, int eh, double aspect, b200rt_camera* cam, b200rt_params* params) {
  Pcg32 rng(seed);
  sb.ok(b200rt_builder_seed_perlin(sb.raw(), seed ^ 0xA5A5A5A5ull));
  HitableList list;
  MaterialPtr ground = make_shared<Lambertian>(ConstantTexture(Vec3(0.48, 0.83, 0.53)));
  HitableList boxes1;
  for (int i = 0; i < 20; ++i)
    for (int j = 0; j < 20; ++j) {
      double w = 100.0, x0 = -1000.0 + i * w, z0 = -1000.0 + j * w, y1 = rng.range(1, 101);
      boxes1.push_back(make_shared<Cube>(Vec3(x0, 0.0, z0), Vec3(x0 + w, y1, z0 + w), ground));
    }
  list.push_back(make_shared<BVHNode>(boxes1));
  HitablePtr light = make_shared<XZRectangle>(make_pair(123., 147.), make_pair(423., 412.), 554., make_shared<DiffuseLight>(ConstantTexture(Vec3(7, 7, 7))));
  list.push_back(light);
  list.push_back(make_shared<MovingSphere>(Vec3(400, 400, 200), Vec3(430, 400, 200), 0.0, 1.0, 50.0, make_shared<Lambertian>(ConstantTexture(Vec3(0.7, 0.3, 0.1)))));
  list.push_back(make_shared<Sphere>(Vec3(260, 150, 45), 50, make_shared<Dielectric>(1.5, ConstantTexture(Vec3::ones()))));
  list.push_back(make_shared<Sphere>(Vec3(0, 150, 145), 50, make_shared<Metal>(ConstantTexture(Vec3(0.8, 0.8, 0.9)), 1.0)));
  auto boundary = make_shared<Sphere>(Vec3(360, 150, 145), 70, make_shared<Dielectric>(1.5, ConstantTexture(Vec3::ones())));
  list.push_back(boundary);
  list.push_back(make_shared<ConstantMedium>(0.2, boundary, make_shared<Isotropic>(ConstantTexture(Vec3(0.2, 0.4, 0.9)))));
  auto fog = make_shared<Sphere>(Vec3(0, 0, 0), 5000, make_shared<Dielectric>(1.5, ConstantTexture(Vec3::ones())));
  list.push_back(make_shared<ConstantMedium>(0.0001, fog, make_shared<Isotropic>(ConstantTexture(Vec3(1, 1, 1)))));
  MaterialPtr emat;
  if (earth && ew > 0 && eh > 0) emat = make_shared<Lambertian>(ImageTexture(earth, ew, eh));
  else emat = make_shared<Lambertian>(CheckerTexture(ConstantTexture(Vec3(0.1, 0.2, 0.6)), ConstantTexture(Vec3(0.2, 0.6, 0.2))));
  list.push_back(make_shared<Sphere>(Vec3(400, 200, 400), 100, emat));
  list.push_back(make_shared<Sphere>(Vec3(220, 280, 300), 80, make_shared<Lambertian>(NoiseTexture(0.1))));
  HitableList boxes2;
  MaterialPtr white = make_shared<Lambertian>(ConstantTexture(Vec3(.73, .73, .73)));
  for (int j = 0; j < 1000; ++j) { const double x = rng.range(0, 165), y = rng.range(0, 165), z = rng.range(0, 165); boxes2.push_back(make_shared<Sphere>(Vec3(x, y, z), 10, white)); }
  list.push_back(make_shared<Transform>(Vec3(0, 15, 0), Vec3::ones(), Vec3(-100, 270, 395), make_shared<BVHNode>(boxes2)));
  World w(list, Camera(make_pair(Vec3(478, 278, -600), Vec3(278, 278, 0)), Vec3(0, 1, 0), 40.0, aspect > 0 ? aspect : 1.0, 0.0, 10.0, 0.0, 1.0), {light}, sb);
  *cam = w.cam.cam;
  default_params(params, 800, 800, 10000, B200RT_INTEGRATOR_PATH, B200RT_BACKGROUND_CONSTANT, 0, 0, 0);
}

// C5: triangle mesh in the reference Cornell box. The reference ships no >=100k mesh: bun315.obj (4,968
// triangles, no vn lines) is midpoint-subdivided 3x -> 317,952 triangles with generated vertex normals.
void mesh_scene(SceneBuilder& sb, const char* obj_path, int subdivide, double aspect, b200rt_camera* cam, b200rt_params* params) {
  MaterialPtr white = make_shared<Lambertian>(ConstantTexture(Vec3(0.75, 0.75, 0.75)));
  HitableList list;
  cornell_walls(list, white);
  HitablePtr light = cornell_light(15.0);
  list.push_back(light);
  list.push_back(make_shared<Sphere>(Vec3(430., 80., 150.), 80., make_shared<Metal>(ConstantTexture(Vec3(0.999, 0.999, 0.999)), 0.)));
  // bun315 spans roughly [-0.095,0.061] x [0.033,0.187] x [-0.062,0.059]: scale 2200 -> ~340 units tall
  list.push_back(make_shared<Transform>(Vec3(0, 180, 0), Vec3::ones() * 2200., Vec3(240., -72., 300.), Mesh::load_obj(obj_path ? obj_path : "data/mesh/bun315.obj", white, true, subdivide)));
  World w(list, Camera(make_pair(Vec3(278., 278., -800.), Vec3(278., 278., 278.)), Vec3(0., 1., 0.), 50., aspect > 0 ? aspect : 16.0 / 9.0, 0.0, 10.0), {light}, sb);
  *cam = w.cam.cam;
  default_params(params, 1920, 1080, 4096, B200RT_INTEGRATOR_MIXTURE, B200RT_BACKGROUND_CONSTANT, 0, 0, 0);
}

}  // namespace

extern "C" int b200rt_builder_make_scene(b200rt_builder* b, const char* name, uint64_t seed, const char* aux_path, const uint8_t* earth_rgb8, int32_t earth_w, int32_t earth_h, double aspect, b200rt_camera* cam_out, b200rt_params* params_out) {
  if (!b || !name || !cam_out) return B200RT_ERR_INVALID;
  b200rt_params dummy;
  if (!params_out) params_out = &dummy;
  try {
    SceneBuilder sb(b);
    std::string n(name);
    if (n == "book1") book1(sb, seed, aspect, cam_out, params_out);
    else if (n == "cornell_ref") cornell_ref(sb, aux_path, aspect, cam_out, params_out);
    else if (n == "cornell_book3") cornell_book3(sb, aspect, cam_out, params_out);
    else if (n == "book2") book2(sb, seed, earth_rgb8, earth_w, earth_h, aspect, cam_out, params_out);
    else if (n == "mesh") mesh_scene(sb, aux_path, 3, aspect, cam_out, params_out);
    else if (n == "mesh_small") mesh_scene(sb, aux_path, 0, aspect, cam_out, params_out);
    else { b->err = "unknown scene '" + n + "'"; return B200RT_ERR_INVALID; }
  } catch (const std::exception& e) {
    if (b->err.empty()) b->err = e.what();
    return B200RT_ERR_INVALID;
  }
  return B200RT_OK;
}
