// rtx_host_test.cpp — drives librtx_cuda.so through the C++ mirror (rtx.hpp) exactly as a user of the
// reference would drive `rtx`: the 4 camera tests of raytracing/camera.rs:79-110, then the demo scene.
#include <cmath>
#include <cstdio>
#include <cstdlib>

#include "rtx.hpp"

using namespace rtx;
using namespace rtx::object;

#define EXPECT(c)                                               \
  do {                                                          \
    if (!(c)) {                                                 \
      std::printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #c); \
      std::exit(1);                                             \
    }                                                           \
  } while (0)

int main() {
  const double deg90 = 90.0 * (3.14159265358979323846 / 180.0);  // 90f64.to_radians()
  {  // camera.rs:82-109
    Camera cam(Vector3::zeros(), Vector3::X(), deg90);
    EXPECT(cam.to_cam_space(Vector3::X()) == Vector3::Z());
    EXPECT(cam.to_cam_space(Vector3::Y()) == Vector3::X());
    EXPECT(cam.to_cam_space(Vector3::Z()) == Vector3::Y());
    EXPECT(cam.to_world_space(Vector3::X()) == Vector3::Y());
    EXPECT(cam.to_world_space(Vector3::Y()) == Vector3::Z());
    EXPECT(cam.to_world_space(Vector3::Z()) == Vector3::X());
    Camera cam2(Vector3::zeros(), Vector3::Y(), deg90);
    EXPECT(cam2.to_world_space(Vector3::X()) == -Vector3::X());
    EXPECT(cam2.to_world_space(Vector3::Y()) == Vector3::Z());
    EXPECT(cam2.to_world_space(Vector3::Z()) == Vector3::Y());
    EXPECT(cam2.to_cam_space(Vector3::X()) == -Vector3::X());
    EXPECT(cam2.to_cam_space(Vector3::Y()) == Vector3::Z());
    EXPECT(cam2.to_cam_space(Vector3::Z()) == Vector3::Y());
    std::printf("camera tests ok\n");
  }
  if (rtx_device_count() < 1) {
    // no CPU fallback: creating a scene must fail loudly
    try {
      Scene s(Config{}, Camera({0, 0, 1}, {1, 0, 0}, deg90));
      std::printf("FAILED: scene creation succeeded without a GPU\n");
      return 1;
    } catch (const Error& e) {
      EXPECT(e.status == RTX_ERR_CUDA);
      std::printf("no GPU: %s\nALL OK (cpu-only checks)\n", e.what());
      return 0;
    }
  }
  // S1 sphere demo
  Scene scene(Config{}.with_rays_per_pixel(4).with_max_bounces(8), Camera({0, 0, 1}, {1, 0, 0}, deg90));
  scene.add_object({Plane{{0, 0, 0}, {0, 0, 1}}, Material::make({0.8, 0.8, 0.8}, {0, 0, 0}, 1.0)});
  scene.add_object({Sphere{{6, 0, 1}, 1}, Material::make({0.9, 0.2, 0.2}, {0, 0, 0}, 1.0)});
  scene.add_object({Sphere{{6, -2.5, 1}, 1}, Material::make({0.9, 0.9, 0.9}, {0, 0, 0}, 0.05)});
  scene.add_object({Sphere{{6, 2.5, 1}, 1}, Material::make({0.2, 0.9, 0.2}, {0, 0, 0}, 0.5)});
  scene.add_object({Sphere{{4, 1.2, 0.4}, 0.4}, Material::make({0.2, 0.3, 0.9}, {0, 0, 0}, 1.0)});
  scene.add_object({Sphere{{4, -1.2, 0.4}, 0.4}, Material::light({2, 1.6, 1})});
  scene.add_object({Sphere{{5, 0, 12}, 6}, Material::light({4, 4, 4})});
  rtx_stats st{};
  auto img = scene.render(320, 200, 1, &st);
  EXPECT(img.size() == 200 && img[0].size() == 320);
  double sum = 0;
  for (auto& row : img)
    for (auto& p : row) sum += p.x + p.y + p.z;
  EXPECT(std::isfinite(sum) && sum > 0);
  EXPECT(st.rays > 320 * 200 * 4 && st.paths == 320ull * 200 * 4 && st.kernel_launches >= 1 && st.kernel_launches <= 2);
  // top row (y = h-1) looks at the sky: black; bottom row (y = 0) looks at the lit ground
  EXPECT(img[199][160] == Vector3::zeros());
  EXPECT(img[0][160].x > 0);
  auto again = scene.render(320, 200, 1);
  for (size_t y = 0; y < 200; ++y)
    for (size_t x = 0; x < 320; ++x) EXPECT(again[y][x] == img[y][x]);  // deterministic in (scene, size, seed)
  auto rgb8 = scene.render_to_image(320, 200, 1);
  EXPECT(rgb8.size() == 320u * 200 * 3);
  EXPECT(rgb8[0] == 0 && rgb8[(199 * 320 + 160) * 3] > 0);  // flipped: row 0 = top = sky
  // pub-field edits take effect at the next render (scene.rs:80-84)
  scene.config = scene.config.with_max_bounces(0).with_rays_per_pixel(1);
  rtx_stats st2{};
  scene.render(64, 64, 1, &st2);
  EXPECT(st2.rays == 64 * 64 && st2.paths == 64 * 64);
  // in-place update: moving the red sphere away changes the picture; moving it back restores every bit
  scene.config = scene.config.with_max_bounces(8).with_rays_per_pixel(4);
  scene.update_object(1, {Sphere{{6, 0, 3}, 1}, Material::make({0.9, 0.2, 0.2}, {0, 0, 0}, 1.0)});
  auto moved = scene.render(320, 200, 1);
  bool differs = false;
  for (size_t y = 0; y < 200 && !differs; ++y)
    for (size_t x = 0; x < 320; ++x)
      if (!(moved[y][x] == img[y][x])) { differs = true; break; }
  EXPECT(differs);
  scene.update_object(1, {Sphere{{6, 0, 1}, 1}, Material::make({0.9, 0.2, 0.2}, {0, 0, 0}, 1.0)});
  auto back = scene.render(320, 200, 1);
  for (size_t y = 0; y < 200; ++y)
    for (size_t x = 0; x < 320; ++x) EXPECT(back[y][x] == img[y][x]);
  bool threw = false;
  try {
    scene.update_object(0, {Sphere{{0, 0, 0}, 1}, Material::mirror()});  // object 0 is the plane
  } catch (const rtx::Error&) {
    threw = true;
  }
  EXPECT(threw);
  {  // the same scene over every GPU of the process behind one handle (rtx_multi_*): the one-GPU image, bit for bit
    std::vector<int32_t> devices;
    for (int32_t d = 0; d < rtx_device_count() && d < 8; ++d) devices.push_back(d);
    MultiScene multi(Config{}.with_rays_per_pixel(4).with_max_bounces(8), Camera({0, 0, 1}, {1, 0, 0}, deg90), devices);
    multi.add_object({Plane{{0, 0, 0}, {0, 0, 1}}, Material::make({0.8, 0.8, 0.8}, {0, 0, 0}, 1.0)});
    multi.add_object({Sphere{{6, 0, 1}, 1}, Material::make({0.9, 0.2, 0.2}, {0, 0, 0}, 1.0)});
    multi.add_object({Sphere{{6, -2.5, 1}, 1}, Material::make({0.9, 0.9, 0.9}, {0, 0, 0}, 0.05)});
    multi.add_object({Sphere{{6, 2.5, 1}, 1}, Material::make({0.2, 0.9, 0.2}, {0, 0, 0}, 0.5)});
    multi.add_object({Sphere{{4, 1.2, 0.4}, 0.4}, Material::make({0.2, 0.3, 0.9}, {0, 0, 0}, 1.0)});
    multi.add_object({Sphere{{4, -1.2, 0.4}, 0.4}, Material::light({2, 1.6, 1})});
    multi.add_object({Sphere{{5, 0, 12}, 6}, Material::light({4, 4, 4})});
    std::vector<double> buf(320 * 200 * 3);
    rtx_stats ms{};
    multi.render(320, 200, 1, buf.data(), &ms);
    for (size_t y = 0; y < 200; ++y)
      for (size_t x = 0; x < 320; ++x) {
        const double* p = &buf[(y * 320 + x) * 3];
        EXPECT(p[0] == img[y][x].x && p[1] == img[y][x].y && p[2] == img[y][x].z);
      }
    EXPECT(ms.rays == st.rays && ms.pixels == 320u * 200);
    std::printf("multi ok: %zu device(s)\n", devices.size());
  }
  {  // the wgpu backend's f32 image of the same scene (rtx_scene_render_preview_f32)
    auto pv = scene.render_preview(160, 100);
    double s32 = 0;
    for (float v : pv) s32 += v;
    EXPECT(pv.size() == 160u * 100 * 3 && std::isfinite(s32) && s32 > 0);
  }
  std::printf("render ok: %llu rays, %.3f ms device, mean radiance %.5f\nALL OK\n", (unsigned long long)st.rays,
              st.device_ms, sum / (320.0 * 200 * 3));
  return 0;
}
