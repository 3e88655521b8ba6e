// rtx_wire.cu — ingest of the reference's f32 "GpuSerialize" wire format.
//
// The reference's wgpu backend describes a scene to the device as little-endian f32 records:
//   camera   32 B  position(12) pad(4) direction(12) fov(4)                          raytracing/camera.rs:70-78
//   config   20 B  rays_per_pixel u32, max_bounces u32, focal_length, focal_offset,
//                  non_focal_offset f32                                              raytracing/scene.rs:67-76
//   object   48 B  base_color(12) roughness(4) emission(12) shape_id u32 index u32 pad(12)   gpu/object.rs:34-42
//   sphere   16 B  position(12) radius(4)                                            object/sphere.rs:36-42
//   plane    32 B  position(12) pad(4) normal(12) pad(4)                             object/plane.rs:38-46
//   triangle 48 B  3 x { vertex(12) pad(4) }                                         object/triangle.rs:130-136
// gpu_state.rs:53-77 keeps one shape buffer per shape type (keyed by GpuShape::object_type(), e.g.
// "rtx::raytracing::object::sphere::sphere"), numbers the types in order of first appearance (shape_id) and
// appends one object record per object, in insertion order, pointing at (shape_id, index within that buffer).
// These entry points take exactly those buffers, so a `gpu::Scene` user can hand its serialized scene to the
// CUDA backend.  Every f32 widens to f64 exactly; what is rendered is the f64 CPU-path semantics
// (Scene::render, scene.rs:144-291) on the f32-rounded scene.  The WGSL path's own arithmetic (f32, PCG RNG,
// Moeller-Trumbore) is a different image and is not reproduced (out of scope).
#include <cstring>
#include <string>
#include <vector>

#include "../../include/rtx_cuda.h"
#include "rtx_internal.h"

namespace {
inline float rd_f32(const unsigned char* p) {
  uint32_t u = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
  float f;
  std::memcpy(&f, &u, 4);
  return f;
}
inline uint32_t rd_u32(const unsigned char* p) {
  return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
}
inline rtx_vec3 rd_vec3(const unsigned char* p) {
  rtx_vec3 v;
  v.x = (double)rd_f32(p), v.y = (double)rd_f32(p + 4), v.z = (double)rd_f32(p + 8);
  return v;
}
// GpuShape::object_type() is format!("{}::sphere", module_path!()): the last path segment names the shape.
int kind_of(const char* tag) {
  if (!tag) return -1;
  const std::string t(tag);
  const size_t pos = t.rfind("::");
  const std::string last = pos == std::string::npos ? t : t.substr(pos + 2);
  if (last == "sphere") return RTX_WIRE_SPHERE;
  if (last == "plane") return RTX_WIRE_PLANE;
  if (last == "triangle") return RTX_WIRE_TRIANGLE;
  return -1;
}
const uint64_t SHAPE_BYTES[3] = {RTX_WIRE_SPHERE_BYTES, RTX_WIRE_PLANE_BYTES, RTX_WIRE_TRIANGLE_BYTES};
}  // namespace

extern "C" {

rtx_status rtx_wire_decode_camera(const void* bytes, rtx_camera* out) {
  if (!bytes || !out) return rtx::fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "null argument");
  const unsigned char* p = static_cast<const unsigned char*>(bytes);
  const rtx_vec3 position = rd_vec3(p), direction = rd_vec3(p + 16);
  return rtx_camera_new(&position, &direction, (double)rd_f32(p + 28), out);  // Camera::new, camera.rs:19-28
}

rtx_status rtx_wire_decode_config(const void* bytes, rtx_config* out) {
  if (!bytes || !out) return rtx::fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "null argument");
  const unsigned char* p = static_cast<const unsigned char*>(bytes);
  out->rays_per_pixel = rd_u32(p);
  out->max_bounces = rd_u32(p + 4);
  out->focal_length = (double)rd_f32(p + 8);
  out->focal_offset = (double)rd_f32(p + 12);
  out->non_focal_offset = (double)rd_f32(p + 16);
  return RTX_OK;
}

rtx_status rtx_wire_decode_objects(const void* object_records, uint64_t n_objects,
                                   const rtx_wire_shape_buffer* shape_buffers, uint32_t n_shape_buffers,
                                   uint32_t* out_kinds, rtx_triangle* out_shapes, rtx_material* out_materials) {
  if ((n_objects && !object_records) || (n_shape_buffers && !shape_buffers) || !out_kinds || !out_shapes || !out_materials)
    return rtx::fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "null argument");
  std::vector<int> kinds(n_shape_buffers);
  for (uint32_t b = 0; b < n_shape_buffers; ++b) {
    kinds[b] = kind_of(shape_buffers[b].object_type);
    if (kinds[b] < 0)  // a user-defined GpuShape: its distance function is WGSL source, not data
      return rtx::fail(nullptr, RTX_ERR_UNSUPPORTED_SHAPE,
                       std::string("unsupported shape type '") + (shape_buffers[b].object_type ? shape_buffers[b].object_type : "") + "'");
    if (shape_buffers[b].shape_count && !shape_buffers[b].shape_bytes)
      return rtx::fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "shape buffer without bytes");
  }
  const unsigned char* rec = static_cast<const unsigned char*>(object_records);
  for (uint64_t i = 0; i < n_objects; ++i, rec += RTX_WIRE_OBJECT_BYTES) {
    rtx_material& m = out_materials[i];
    m.base_color = rd_vec3(rec);
    m.roughness = (double)rd_f32(rec + 12);
    m.emission_color = rd_vec3(rec + 16);
    const uint32_t shape_id = rd_u32(rec + 28), index = rd_u32(rec + 32);
    if (shape_id >= n_shape_buffers || index >= shape_buffers[shape_id].shape_count)
      return rtx::fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "object record points outside its shape buffer");
    const int kind = kinds[shape_id];
    const unsigned char* sp = static_cast<const unsigned char*>(shape_buffers[shape_id].shape_bytes) + (uint64_t)index * SHAPE_BYTES[kind];
    out_kinds[i] = (uint32_t)kind;
    rtx_triangle& t = out_shapes[i];
    std::memset(&t, 0, sizeof(t));
    if (kind == RTX_WIRE_SPHERE) {
      t.vertices[0] = rd_vec3(sp);
      t.vertices[1].x = (double)rd_f32(sp + 12);  // radius
    } else if (kind == RTX_WIRE_PLANE) {
      t.vertices[0] = rd_vec3(sp);
      t.vertices[1] = rd_vec3(sp + 16);
    } else {
      t.vertices[0] = rd_vec3(sp), t.vertices[1] = rd_vec3(sp + 16), t.vertices[2] = rd_vec3(sp + 32);
    }
  }
  return RTX_OK;
}

rtx_status rtx_scene_add_wire_objects(rtx_scene* scene, const void* object_records, uint64_t n_objects,
                                      const rtx_wire_shape_buffer* shape_buffers, uint32_t n_shape_buffers) {
  if (!scene) return rtx::fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "null scene");
  if (n_objects == 0) return RTX_OK;
  std::vector<uint32_t> kinds(n_objects);
  std::vector<rtx_triangle> shapes(n_objects);
  std::vector<rtx_material> mats(n_objects);
  rtx_status st = rtx_wire_decode_objects(object_records, n_objects, shape_buffers, n_shape_buffers, kinds.data(),
                                          shapes.data(), mats.data());
  if (st != RTX_OK) return rtx::fail(scene, st, rtx_last_error(nullptr));
  // insertion order = record order (the object index is the nearest-hit tie-break, scene.rs:250); runs of one
  // kind go through the batched add entry points
  uint64_t i = 0;
  while (i < n_objects) {
    uint64_t j = i;
    while (j < n_objects && kinds[j] == kinds[i]) ++j;
    const uint64_t n = j - i;
    if (kinds[i] == RTX_WIRE_SPHERE) {
      std::vector<rtx_sphere> v(n);
      for (uint64_t k = 0; k < n; ++k) v[k].position = shapes[i + k].vertices[0], v[k].radius = shapes[i + k].vertices[1].x;
      st = rtx_scene_add_spheres(scene, v.data(), &mats[i], n);
    } else if (kinds[i] == RTX_WIRE_PLANE) {
      std::vector<rtx_plane> v(n);
      for (uint64_t k = 0; k < n; ++k) v[k].position = shapes[i + k].vertices[0], v[k].normal = shapes[i + k].vertices[1];
      st = rtx_scene_add_planes(scene, v.data(), &mats[i], n);
    } else {
      st = rtx_scene_add_triangles(scene, &shapes[i], &mats[i], n);
    }
    if (st != RTX_OK) return st;
    i = j;
  }
  return RTX_OK;
}

rtx_status rtx_scene_set_wire_camera(rtx_scene* scene, const void* bytes) {
  rtx_camera c;
  rtx_status st = rtx_wire_decode_camera(bytes, &c);
  if (st != RTX_OK) return st;
  return rtx_scene_set_camera(scene, &c);
}
rtx_status rtx_scene_set_wire_config(rtx_scene* scene, const void* bytes) {
  rtx_config c;
  rtx_status st = rtx_wire_decode_config(bytes, &c);
  if (st != RTX_OK) return st;
  return rtx_scene_set_config(scene, &c);
}

}  // extern "C"
