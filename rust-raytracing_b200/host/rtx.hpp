// rtx.hpp — C++ host-side mirror of the `rtx` crate's public API over the C ABI (include/rtx_cuda.h).
//
// The reference is compiled code (Rust) and its toolchain is absent from this image, so the host side
// above the C ABI is written in C++ with the reference's names, argument meaning and error behaviour:
//   rtx::Vector3            math/vector.rs:12-20
//   rtx::Camera             raytracing/camera.rs:7-68
//   rtx::Config             raytracing/scene.rs:16-65   (with_* builders clone, like reassign!)
//   rtx::object::Material   raytracing/object.rs:79-135
//   rtx::object::{Sphere,Plane,Triangle,Object}   raytracing/object/*.rs, object.rs:10-28
//   rtx::Scene              raytracing/scene.rs:78-190  (new / add_object / render / render_to_image)
// Errors: the reference panics; here every non-zero rtx_status throws rtx::Error.
#pragma once
#include <array>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <variant>
#include <vector>

#include "rtx_cuda.h"

namespace rtx {

struct Error : std::runtime_error {
  rtx_status status;
  Error(rtx_status s, const std::string& m) : std::runtime_error(m), status(s) {}
};
inline void check(rtx_status s, const rtx_scene* scene = nullptr) {
  if (s != RTX_OK) throw Error(s, std::string("rtx: ") + rtx_last_error(scene));
}

struct Vector3 {
  double x = 0, y = 0, z = 0;
  static Vector3 zeros() { return {0, 0, 0}; }
  static Vector3 ones() { return {1, 1, 1}; }
  static Vector3 X() { return {1, 0, 0}; }
  static Vector3 Y() { return {0, 1, 0}; }
  static Vector3 Z() { return {0, 0, 1}; }
  bool operator==(const Vector3& o) const { return x == o.x && y == o.y && z == o.z; }
  Vector3 operator-() const { return {-x, -y, -z}; }
  rtx_vec3 c() const { return {x, y, z}; }
};
inline Vector3 from_c(const rtx_vec3& v) { return {v.x, v.y, v.z}; }

class Camera {
 public:
  Camera(Vector3 position, Vector3 direction, double fov) {  // camera.rs:19-28
    rtx_vec3 p = position.c(), d = direction.c();
    check(rtx_camera_new(&p, &d, fov, &c_));
  }
  double fov() const { return c_.fov; }
  Vector3 position() const { return from_c(c_.position); }
  void set_position(Vector3 p) { c_.position = p.c(); }
  Vector3 get_direction() const { return from_c(c_.direction); }
  void set_direction(Vector3 d) {  // camera.rs:35-40 (matrices lag one call, as upstream)
    rtx_vec3 v = d.c();
    check(rtx_camera_set_direction(&c_, &v));
  }
  Vector3 to_cam_space(Vector3 v) const { return xf(rtx_camera_to_cam_space, v); }
  Vector3 to_world_space(Vector3 v) const { return xf(rtx_camera_to_world_space, v); }
  Vector3 rotate_to_world_space(Vector3 v) const { return xf(rtx_camera_rotate_to_world_space, v); }
  const rtx_camera& c() const { return c_; }

 private:
  template <typename F>
  Vector3 xf(F fn, Vector3 v) const {
    rtx_vec3 in = v.c(), out;
    check(fn(&c_, &in, &out));
    return from_c(out);
  }
  rtx_camera c_{};
};

struct Config {  // scene.rs:16-65
  size_t rays_per_pixel = 16, max_bounces = 10;
  double focal_length = 10.0, focal_offset = 1e-4, non_focal_offset = 1e-1;
  Config with_rays_per_pixel(size_t v) const { Config c = *this; c.rays_per_pixel = v; return c; }
  Config with_max_bounces(size_t v) const { Config c = *this; c.max_bounces = v; return c; }
  Config with_focal_length(double v) const { Config c = *this; c.focal_length = v; return c; }
  Config with_focal_offset(double v) const { Config c = *this; c.focal_offset = v; return c; }
  Config with_non_focal_offset(double v) const { Config c = *this; c.non_focal_offset = v; return c; }
  rtx_config c() const { return {rays_per_pixel, max_bounces, focal_length, focal_offset, non_focal_offset}; }
};

namespace object {
struct Material {  // object.rs:79-135
  Vector3 base_color, emission_color;
  double roughness = 1.0;
  static Material make(Vector3 b, Vector3 e, double r) { return {b, e, r}; }
  static Material colored(Vector3 color) { return {color, Vector3::zeros(), 1.0}; }
  static Material light(Vector3 light_color) { return {Vector3::zeros(), light_color, 1.0}; }
  static Material mirror() { return {Vector3::ones(), Vector3::zeros(), 1.0}; }  // roughness 1.0 on the CPU path
  rtx_material c() const { return {base_color.c(), emission_color.c(), roughness}; }
};
struct Sphere { Vector3 position; double radius; };
struct Plane { Vector3 position, normal; };
struct Triangle { std::array<Vector3, 3> vertices; };
// Only the three built-in shapes can cross the C ABI (CustomShape is open-ended, object.rs:53-76).
struct Object {
  std::variant<Sphere, Plane, Triangle> shape;
  Material material;
};
}  // namespace object

class Scene {
 public:
  Scene(Config config, Camera camera, int device = 0) : config(config), camera(camera) {  // scene.rs:112-118
    rtx_config cc = config.c();
    check(rtx_scene_create(&camera.c(), &cc, device, &h_));
  }
  ~Scene() { rtx_scene_destroy(h_); }
  Scene(const Scene&) = delete;
  Scene& operator=(const Scene&) = delete;

  void add_object(const object::Object& o) {  // scene.rs:126-128
    rtx_material m = o.material.c();
    if (auto* s = std::get_if<object::Sphere>(&o.shape)) {
      rtx_sphere c{s->position.c(), s->radius};
      check(rtx_scene_add_spheres(h_, &c, &m, 1), h_);
    } else if (auto* p = std::get_if<object::Plane>(&o.shape)) {
      rtx_plane c{p->position.c(), p->normal.c()};
      check(rtx_scene_add_planes(h_, &c, &m, 1), h_);
    } else {
      const auto& t = std::get<object::Triangle>(o.shape);
      rtx_triangle c{{t.vertices[0].c(), t.vertices[1].c(), t.vertices[2].c()}};
      check(rtx_scene_add_triangles(h_, &c, &m, 1), h_);
    }
  }
  // Replaces object `index` (global insertion order) in place by an object of the SAME shape kind — what
  // mutating a shape behind its Arc<Mutex<…>> (object.rs:12) does in the reference.  Only that record is
  // re-uploaded and the LBVH is refitted instead of rebuilt (rtx_scene_update_*).
  void update_object(uint64_t index, const object::Object& o) {
    rtx_material m = o.material.c();
    if (auto* s = std::get_if<object::Sphere>(&o.shape)) {
      rtx_sphere c{s->position.c(), s->radius};
      check(rtx_scene_update_spheres(h_, &index, &c, &m, 1), h_);
    } else if (auto* p = std::get_if<object::Plane>(&o.shape)) {
      rtx_plane c{p->position.c(), p->normal.c()};
      check(rtx_scene_update_planes(h_, &index, &c, &m, 1), h_);
    } else {
      const auto& t = std::get<object::Triangle>(o.shape);
      rtx_triangle c{{t.vertices[0].c(), t.vertices[1].c(), t.vertices[2].c()}};
      check(rtx_scene_update_triangles(h_, &index, &c, &m, 1), h_);
    }
  }
  // scene.rs:144-170 → img[y][x], row 0 = bottom.  `seed` is the one added parameter.
  std::vector<std::vector<Vector3>> render(size_t width, size_t height, uint64_t seed = 0, rtx_stats* stats = nullptr) {
    sync();
    std::vector<double> buf(width * height * 3);
    check(rtx_scene_render(h_, (uint32_t)width, (uint32_t)height, seed, RTX_ACCEL_AUTO, buf.data(), nullptr, stats), h_);
    std::vector<std::vector<Vector3>> img(height, std::vector<Vector3>(width));
    for (size_t y = 0; y < height; ++y)
      for (size_t x = 0; x < width; ++x) {
        const double* p = &buf[(y * width + x) * 3];
        img[y][x] = {p[0], p[1], p[2]};
      }
    return img;
  }
  // scene.rs:173-178 → RGB8 [height][width][3], row 0 = top.
  std::vector<uint8_t> render_to_image(size_t width, size_t height, uint64_t seed = 0) {
    sync();
    std::vector<uint8_t> out(width * height * 3);
    check(rtx_scene_render_to_rgb8(h_, (uint32_t)width, (uint32_t)height, seed, RTX_ACCEL_AUTO, out.data(), nullptr), h_);
    return out;
  }
  // ≙ gpu::Scene::render (raytracing/gpu/scene.rs:27-29): the wgpu backend's f32 image (base_shader.wgsl), row 0 = bottom
  std::vector<float> render_preview(size_t width, size_t height) {
    sync();
    std::vector<float> out(width * height * 3);
    check(rtx_scene_render_preview_f32(h_, (uint32_t)width, (uint32_t)height, out.data(), nullptr, nullptr), h_);
    return out;
  }
  Config config;  // pub fields (scene.rs:80-84): edits take effect at the next render
  Camera camera;
  rtx_scene* handle() { return h_; }

 private:
  void sync() {
    rtx_config cc = config.c();
    check(rtx_scene_set_config(h_, &cc), h_);
    check(rtx_scene_set_camera(h_, &camera.c()), h_);
  }
  rtx_scene* h_ = nullptr;
};

// Scene over every listed GPU of this process (rtx_multi_*): same verbs, render() blocks and returns the whole image;
// byte-identical to the one-GPU image.  ≙ raytracing/scene.rs:112-170.
class MultiScene {
 public:
  MultiScene(Config config, Camera camera, const std::vector<int32_t>& devices) : config(config), camera(camera) {
    rtx_config cc = config.c();
    check(rtx_multi_create(&camera.c(), &cc, devices.data(), (uint32_t)devices.size(), &h_));
  }
  ~MultiScene() { rtx_multi_destroy(h_); }
  MultiScene(const MultiScene&) = delete;
  MultiScene& operator=(const MultiScene&) = delete;
  void add_object(const object::Object& o) {
    rtx_material m = o.material.c();
    if (auto* s = std::get_if<object::Sphere>(&o.shape)) {
      rtx_sphere c{s->position.c(), s->radius};
      check(rtx_multi_add_spheres(h_, &c, &m, 1));
    } else if (auto* p = std::get_if<object::Plane>(&o.shape)) {
      rtx_plane c{p->position.c(), p->normal.c()};
      check(rtx_multi_add_planes(h_, &c, &m, 1));
    } else {
      const auto& t = std::get<object::Triangle>(o.shape);
      rtx_triangle c{{t.vertices[0].c(), t.vertices[1].c(), t.vertices[2].c()}};
      check(rtx_multi_add_triangles(h_, &c, &m, 1));
    }
  }
  // rgb: double[height][width][3], row 0 = bottom (pinned memory from rtx_host_alloc makes the per-GPU copies DMA)
  void render(size_t width, size_t height, uint64_t seed, double* rgb, rtx_stats* stats = nullptr) {
    rtx_config cc = config.c();
    check(rtx_multi_set_camera(h_, &camera.c()));
    check(rtx_multi_set_config(h_, &cc));
    check(rtx_multi_render(h_, (uint32_t)width, (uint32_t)height, seed, RTX_ACCEL_AUTO, rgb, nullptr, stats));
  }
  Config config;
  Camera camera;

 private:
  rtx_multi* h_ = nullptr;
};

}  // namespace rtx
