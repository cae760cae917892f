// frontend.hpp — C++ mirror of the reference's scene-construction API over the C ABI builder.
//
// Same type names, constructor arguments and nesting as the Rust front end (Rust is not available in this
// image; INTEGRATION.md shows the Rust binding of the same builder calls):
//   Hitable / Material / Texture  (objects/hit.rs:51-54, material.rs:18-46)   -> one extra method: flatten()
//   Sphere, XYRectangle, XZRectangle, YZRectangle, Cube, Mesh::load_obj, Transform::new, ConstantMedium::new,
//   Vec<Arc<dyn Hitable>> (HitableList), BVHNode::new, World::new, Camera::new   (scene.rs:16-112 is the model)
// `flatten` is the hook SURVEY.md §8(b) asks for: trait objects expose only hit/bounding_box, so each type
// lowers itself into the flat tables.  BVHNode::new keeps its signature but only forwards its children: the
// device BVH is rebuilt wide by b200rt_scene_create (topology does not change closest-hit results).
#pragma once
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/b200rt.h"

namespace b200rt {
namespace api {

struct Vec3 {
  double x, y, z;
  Vec3(double x_ = 0, double y_ = 0, double z_ = 0) : x(x_), y(y_), z(z_) {}
  static Vec3 ones() { return Vec3(1, 1, 1); }
  static Vec3 zero() { return Vec3(0, 0, 0); }
  Vec3 operator*(double s) const { return Vec3(x * s, y * s, z * s); }
};

class SceneBuilder {
 public:
  SceneBuilder() { if (b200rt_builder_create(&b_)) throw std::runtime_error("builder_create failed"); }
  explicit SceneBuilder(b200rt_builder* borrowed) : b_(borrowed), owns_(false) {}
  ~SceneBuilder() { if (owns_) b200rt_builder_destroy(b_); }
  SceneBuilder(const SceneBuilder&) = delete;
  b200rt_builder* raw() { return b_; }
  int check(int id) { if (id < 0) throw std::runtime_error(b200rt_builder_error(b_)); return id; }
  void ok(int rc) { if (rc) throw std::runtime_error(b200rt_builder_error(b_)); }
  std::map<const void*, int> cache;  // Arc identity -> table id
 private:
  b200rt_builder* b_ = nullptr;
  bool owns_ = true;
};

struct Texture {
  virtual ~Texture() {}
  virtual int flatten(SceneBuilder& sb) const = 0;
};
struct ConstantTexture : Texture {
  Vec3 color;
  explicit ConstantTexture(Vec3 c) : color(c) {}
  int flatten(SceneBuilder& sb) const override { return sb.check(b200rt_builder_texture_constant(sb.raw(), color.x, color.y, color.z)); }
};
struct CheckerTexture : Texture {
  ConstantTexture t0, t1;
  CheckerTexture(ConstantTexture a, ConstantTexture b) : t0(a), t1(b) {}
  int flatten(SceneBuilder& sb) const override {
    double c0[3] = {t0.color.x, t0.color.y, t0.color.z}, c1[3] = {t1.color.x, t1.color.y, t1.color.z};
    return sb.check(b200rt_builder_texture_checker(sb.raw(), c0, c1));
  }
};
struct ImageTexture : Texture {  // ImageTexture(DynamicImage): decoded RGB8, decode stays on the caller's side
  std::vector<uint8_t> rgb; int width, height;
  ImageTexture(const uint8_t* rgb8, int w, int h) : rgb(rgb8, rgb8 + (size_t)3 * w * h), width(w), height(h) {}
  int flatten(SceneBuilder& sb) const override { return sb.check(b200rt_builder_texture_image(sb.raw(), rgb.data(), width, height)); }
};
struct NoiseTexture : Texture {  // absent upstream
  double scale;
  explicit NoiseTexture(double s) : scale(s) {}
  int flatten(SceneBuilder& sb) const override { return sb.check(b200rt_builder_texture_noise(sb.raw(), scale)); }
};

struct Material {
  virtual ~Material() {}
  virtual int flatten_new(SceneBuilder& sb) const = 0;
  int flatten(SceneBuilder& sb) const {  // one table row per Arc<dyn Material>
    auto it = sb.cache.find(this);
    if (it != sb.cache.end()) return it->second;
    int id = flatten_new(sb);
    sb.cache[this] = id;
    return id;
  }
};
typedef std::shared_ptr<Material> MaterialPtr;
typedef std::shared_ptr<Texture> TexturePtr;
template <int TYPE>
struct TexturedMaterial : Material {
  TexturePtr tex; double param;
  TexturedMaterial(TexturePtr t, double p) : tex(t), param(p) {}
  int flatten_new(SceneBuilder& sb) const override { return sb.check(b200rt_builder_material(sb.raw(), TYPE, tex->flatten(sb), param)); }
};
struct Lambertian : TexturedMaterial<B200RT_MAT_LAMBERTIAN> {
  template <class T> explicit Lambertian(T albedo) : TexturedMaterial(std::make_shared<T>(albedo), 0) {}
};
struct Metal : TexturedMaterial<B200RT_MAT_METAL> {
  template <class T> Metal(T albedo, double fuzz) : TexturedMaterial(std::make_shared<T>(albedo), fuzz) {}
};
struct Dielectric : TexturedMaterial<B200RT_MAT_DIELECTRIC> {
  template <class T> Dielectric(double ir, T albedo) : TexturedMaterial(std::make_shared<T>(albedo), ir) {}
};
struct DiffuseLight : TexturedMaterial<B200RT_MAT_DIFFUSE_LIGHT> {
  template <class T> explicit DiffuseLight(T emit) : TexturedMaterial(std::make_shared<T>(emit), 0) {}
};
struct Isotropic : TexturedMaterial<B200RT_MAT_ISOTROPIC> {
  template <class T> explicit Isotropic(T albedo) : TexturedMaterial(std::make_shared<T>(albedo), 0) {}
};

struct Hitable {
  virtual ~Hitable() {}
  // returns the id of the (first) primitive emitted
  virtual int flatten(SceneBuilder& sb) const = 0;
};
typedef std::shared_ptr<Hitable> HitablePtr;
typedef std::vector<HitablePtr> HitableList;  // impl Hitable for Vec<Arc<dyn Hitable>>

struct Sphere : Hitable {
  Vec3 center; double radius; MaterialPtr material;
  Sphere(Vec3 c, double r, MaterialPtr m) : center(c), radius(r), material(m) {}
  int flatten(SceneBuilder& sb) const override {
    double c[3] = {center.x, center.y, center.z};
    return sb.check(b200rt_builder_sphere(sb.raw(), c, radius, material->flatten(sb)));
  }
};
struct MovingSphere : Hitable {  // absent upstream
  Vec3 center0, center1; double time0, time1, radius; MaterialPtr material;
  MovingSphere(Vec3 c0, Vec3 c1, double t0, double t1, double r, MaterialPtr m) : center0(c0), center1(c1), time0(t0), time1(t1), radius(r), material(m) {}
  int flatten(SceneBuilder& sb) const override {
    double c0[3] = {center0.x, center0.y, center0.z}, c1[3] = {center1.x, center1.y, center1.z};
    return sb.check(b200rt_builder_moving_sphere(sb.raw(), c0, c1, time0, time1, radius, material->flatten(sb)));
  }
};
template <int AXIS>
struct AxisRectangle : Hitable {
  double a0, b0, a1, b1, k; MaterialPtr material;
  AxisRectangle(std::pair<double, double> p0, std::pair<double, double> p1, double k_, MaterialPtr m) : a0(p0.first), b0(p0.second), a1(p1.first), b1(p1.second), k(k_), material(m) {}
  int flatten(SceneBuilder& sb) const override { return sb.check(b200rt_builder_rect(sb.raw(), AXIS, a0, b0, a1, b1, k, material->flatten(sb))); }
};
typedef AxisRectangle<2> XYRectangle;  // {xy0, xy1, z, material}
typedef AxisRectangle<1> XZRectangle;  // {xz0, xz1, y, material}
typedef AxisRectangle<0> YZRectangle;  // {yz0, yz1, x, material}

struct Cube : Hitable {
  Vec3 box_min, box_max; MaterialPtr material;
  Cube(Vec3 mn, Vec3 mx, MaterialPtr m) : box_min(mn), box_max(mx), material(m) {}
  int flatten(SceneBuilder& sb) const override {
    double mn[3] = {box_min.x, box_min.y, box_min.z}, mx[3] = {box_max.x, box_max.y, box_max.z};
    return sb.check(b200rt_builder_cube(sb.raw(), mn, mx, material->flatten(sb)));
  }
};
struct Mesh : Hitable {
  std::string obj_file; MaterialPtr material; bool make_normals; int subdivide;
  static std::shared_ptr<Mesh> load_obj(const std::string& f, MaterialPtr m, bool make_normals = false, int subdivide = 0) {
    auto p = std::make_shared<Mesh>();
    p->obj_file = f; p->material = m; p->make_normals = make_normals; p->subdivide = subdivide;
    return p;
  }
  int flatten(SceneBuilder& sb) const override { return sb.check(b200rt_builder_mesh_obj(sb.raw(), obj_file.c_str(), material->flatten(sb), make_normals, subdivide)); }
};
struct ListNode : Hitable {  // a nested Vec<Arc<dyn Hitable>> used as one Hitable
  HitableList items;
  explicit ListNode(HitableList l) : items(std::move(l)) {}
  int flatten(SceneBuilder& sb) const override {
    int first = -1;
    for (auto& h : items) { int id = h->flatten(sb); if (first < 0) first = id; }
    return first;
  }
};
struct BVHNode : ListNode {  // BVHNode::new(src_objects), bvh.rs:60-83
  explicit BVHNode(HitableList l) : ListNode(std::move(l)) {}
};
struct Transform : Hitable {  // Transform::new(rotate_in_degree, scale, translate, obj), transform.rs:17-22
  Vec3 rotate, scale, translate; HitablePtr obj;
  Transform(Vec3 r, Vec3 s, Vec3 t, HitablePtr o) : rotate(r), scale(s), translate(t), obj(o) {}
  int flatten(SceneBuilder& sb) const override {
    double r[3] = {rotate.x, rotate.y, rotate.z}, s[3] = {scale.x, scale.y, scale.z}, t[3] = {translate.x, translate.y, translate.z};
    sb.ok(b200rt_builder_push_transform(sb.raw(), r, s, t));
    int id = obj->flatten(sb);
    sb.ok(b200rt_builder_pop_transform(sb.raw()));
    return id;
  }
};
struct ConstantMedium : Hitable {  // ConstantMedium::new(d, boundary, phase_function), medium.rs:15-23
  double density; HitablePtr boundary; MaterialPtr phase;
  ConstantMedium(double d, HitablePtr b, MaterialPtr p) : density(d), boundary(b), phase(p) {}
  int flatten(SceneBuilder& sb) const override {
    if (auto s = dynamic_cast<const Sphere*>(boundary.get())) {
      double c[3] = {s->center.x, s->center.y, s->center.z};
      return sb.check(b200rt_builder_medium_sphere(sb.raw(), c, s->radius, density, phase->flatten(sb)));
    }
    if (auto c = dynamic_cast<const Cube*>(boundary.get())) {
      double mn[3] = {c->box_min.x, c->box_min.y, c->box_min.z}, mx[3] = {c->box_max.x, c->box_max.y, c->box_max.z};
      return sb.check(b200rt_builder_medium_cube(sb.raw(), mn, mx, density, phase->flatten(sb)));
    }
    throw std::runtime_error("ConstantMedium boundary must be a Sphere or a Cube");
  }
};

struct Camera {  // Camera::new((look_from, look_at), vup, vfov, aspect_ratio, aperture, focus_dist), camera.rs:24-31
  b200rt_camera cam;
  Camera(std::pair<Vec3, Vec3> from_to, Vec3 vup, double vfov, double aspect, double aperture, double focus_dist, double time0 = 0, double time1 = 0) {
    double f[3] = {from_to.first.x, from_to.first.y, from_to.first.z}, a[3] = {from_to.second.x, from_to.second.y, from_to.second.z}, u[3] = {vup.x, vup.y, vup.z};
    if (b200rt_camera_make(f, a, u, vfov, aspect, aperture, focus_dist, time0, time1, &cam)) throw std::runtime_error("Camera::new: degenerate look_from/look_at/vup");
  }
};

// World::new(hitable_list, cam, lights) (world.rs:15-25): flattens the list; `lights` are Hitables that are
// ALSO in the list (scene.rs:86,110) and are marked for light sampling.
struct World {
  Camera cam;
  World(const HitableList& list, Camera c, const HitableList& lights, SceneBuilder& sb) : cam(c) {
    std::map<const Hitable*, int> ids;
    for (auto& h : list) ids[h.get()] = h->flatten(sb);
    for (auto& l : lights) {
      auto it = ids.find(l.get());
      if (it == ids.end()) throw std::runtime_error("World::new: a light must also be in the hitable list");
      sb.ok(b200rt_builder_mark_light(sb.raw(), it->second));
    }
  }
};

}  // namespace api
}  // namespace b200rt
