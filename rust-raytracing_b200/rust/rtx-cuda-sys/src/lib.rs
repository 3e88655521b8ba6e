//! rtx-cuda-sys — `extern "C"` declarations for include/rtx_cuda.h with `#[repr(C)]` mirrors of the
//! `rtx` crate's types.  UNBUILT IN THIS ENVIRONMENT (no cargo/rustc); layout is asserted at compile time.
#![allow(non_camel_case_types)]
use std::os::raw::{c_char, c_void};

#[repr(C)] #[derive(Clone, Copy, Debug, Default, PartialEq)]
pub struct rtx_vec3 { pub x: f64, pub y: f64, pub z: f64 }            // math/vector.rs:12-20
#[repr(C)] #[derive(Clone, Copy, Debug, Default)]
pub struct rtx_mat3 { pub x: rtx_vec3, pub y: rtx_vec3, pub z: rtx_vec3 } // math/mat.rs:10-18
#[repr(C)] #[derive(Clone, Copy, Debug, Default)]
pub struct rtx_camera {                                                // raytracing/camera.rs:7-15
    pub fov: f64, pub position: rtx_vec3, pub direction: rtx_vec3,
    pub to_cam_space: rtx_mat3, pub to_world_space: rtx_mat3,
}
#[repr(C)] #[derive(Clone, Copy, Debug)]
pub struct rtx_config {                                                // raytracing/scene.rs:16-28
    pub rays_per_pixel: u64, pub max_bounces: u64,
    pub focal_length: f64, pub focal_offset: f64, pub non_focal_offset: f64,
}
#[repr(C)] #[derive(Clone, Copy, Debug)]
pub struct rtx_material { pub base_color: rtx_vec3, pub emission_color: rtx_vec3, pub roughness: f64 }
#[repr(C)] #[derive(Clone, Copy, Debug)] pub struct rtx_sphere { pub position: rtx_vec3, pub radius: f64 }
#[repr(C)] #[derive(Clone, Copy, Debug)] pub struct rtx_plane { pub position: rtx_vec3, pub normal: rtx_vec3 }
#[repr(C)] #[derive(Clone, Copy, Debug)] pub struct rtx_triangle { pub vertices: [rtx_vec3; 3] }
#[repr(C)] #[derive(Clone, Copy, Debug, Default)]
pub struct rtx_stats {
    pub rays: u64, pub paths: u64, pub pixels: u64, pub triangle_panics: u64,
    pub device_ms: f32, pub total_ms: f32, pub kernel_launches: u32, pub accel_used: u32,
    pub nodes_visited: u64, pub prims_tested: u64, pub hits: u64,
}
#[repr(C)] #[derive(Clone, Copy, Debug)] pub struct rtx_shard { pub index: u32, pub count: u32 }
#[repr(C)] pub struct rtx_scene { _private: [u8; 0] }
/// One per-type shape buffer of gpu_state.rs:53-77 (GpuSerialize wire format).
#[repr(C)] #[derive(Clone, Copy, Debug)]
pub struct rtx_wire_shape_buffer { pub object_type: *const c_char, pub shape_bytes: *const c_void, pub shape_count: u64 }

const _: () = assert!(std::mem::size_of::<rtx_vec3>() == 24);
const _: () = assert!(std::mem::size_of::<rtx_mat3>() == 72);
const _: () = assert!(std::mem::size_of::<rtx_camera>() == 200);
const _: () = assert!(std::mem::size_of::<rtx_config>() == 40);
const _: () = assert!(std::mem::size_of::<rtx_material>() == 56);
const _: () = assert!(std::mem::size_of::<rtx_sphere>() == 32);
const _: () = assert!(std::mem::size_of::<rtx_plane>() == 48);
const _: () = assert!(std::mem::size_of::<rtx_triangle>() == 72);

pub type rtx_status = i32;
pub const RTX_OK: rtx_status = 0;
pub const RTX_ACCEL_AUTO: u32 = 0;
pub const RTX_FLAG_NO_RADIANCE: u32 = 4;
pub const RTX_FLAG_LANE_KERNEL: u32 = 8;
pub const RTX_FLAG_SHARD_IN_PLACE: u32 = 16;
pub const RTX_FLAG_PIXEL_QUEUE: u32 = 32;
pub const RTX_FLAG_SAMPLE_QUEUE: u32 = 64;
pub const RTX_PEER_HANDLE_BYTES: usize = 64;

extern "C" {
    pub fn rtx_wire_decode_camera(bytes: *const c_void, out: *mut rtx_camera) -> rtx_status;
    pub fn rtx_wire_decode_config(bytes: *const c_void, out: *mut rtx_config) -> rtx_status;
    pub fn rtx_wire_decode_objects(object_records: *const c_void, n_objects: u64, shape_buffers: *const rtx_wire_shape_buffer,
                                   n_shape_buffers: u32, out_kinds: *mut u32, out_shapes: *mut rtx_triangle,
                                   out_materials: *mut rtx_material) -> rtx_status;
    pub fn rtx_scene_add_wire_objects(scene: *mut rtx_scene, object_records: *const c_void, n_objects: u64,
                                      shape_buffers: *const rtx_wire_shape_buffer, n_shape_buffers: u32) -> rtx_status;
    pub fn rtx_scene_set_wire_camera(scene: *mut rtx_scene, bytes: *const c_void) -> rtx_status;
    pub fn rtx_scene_set_wire_config(scene: *mut rtx_scene, bytes: *const c_void) -> rtx_status;
    pub fn rtx_scene_update_spheres(scene: *mut rtx_scene, object_indices: *const u64, shapes: *const rtx_sphere, materials: *const rtx_material, n: u64) -> rtx_status;
    pub fn rtx_scene_update_planes(scene: *mut rtx_scene, object_indices: *const u64, shapes: *const rtx_plane, materials: *const rtx_material, n: u64) -> rtx_status;
    pub fn rtx_scene_update_triangles(scene: *mut rtx_scene, object_indices: *const u64, shapes: *const rtx_triangle, materials: *const rtx_material, n: u64) -> rtx_status;
    pub fn rtx_scene_rebuild_accel(scene: *mut rtx_scene) -> rtx_status;
    pub fn rtx_scene_accel_refits(scene: *const rtx_scene) -> u32;
    pub fn rtx_scene_accel_builds(scene: *const rtx_scene) -> u32;
    pub fn rtx_scene_last_launches(scene: *const rtx_scene) -> u32;
    pub fn rtx_peer_alloc(device: i32, bytes: u64, out_d_ptr: *mut *mut c_void) -> rtx_status;
    pub fn rtx_peer_free(device: i32, d_ptr: *mut c_void) -> rtx_status;
    pub fn rtx_peer_export(device: i32, d_ptr: *mut c_void, out_handle: *mut u8) -> rtx_status;
    pub fn rtx_peer_open(device: i32, handle: *const u8, out_d_ptr: *mut *mut c_void) -> rtx_status;
    pub fn rtx_peer_close(device: i32, d_ptr: *mut c_void) -> rtx_status;
    pub fn rtx_camera_new(position: *const rtx_vec3, direction: *const rtx_vec3, fov: f64, out: *mut rtx_camera) -> rtx_status;
    pub fn rtx_camera_set_direction(camera: *mut rtx_camera, direction: *const rtx_vec3) -> rtx_status;
    pub fn rtx_config_default(out: *mut rtx_config) -> rtx_status;
    pub fn rtx_scene_create(camera: *const rtx_camera, config: *const rtx_config, device: i32, out: *mut *mut rtx_scene) -> rtx_status;
    pub fn rtx_scene_destroy(scene: *mut rtx_scene) -> rtx_status;
    pub fn rtx_scene_set_camera(scene: *mut rtx_scene, camera: *const rtx_camera) -> rtx_status;
    pub fn rtx_scene_set_config(scene: *mut rtx_scene, config: *const rtx_config) -> rtx_status;
    pub fn rtx_scene_add_spheres(scene: *mut rtx_scene, shapes: *const rtx_sphere, materials: *const rtx_material, n: u64) -> rtx_status;
    pub fn rtx_scene_add_planes(scene: *mut rtx_scene, shapes: *const rtx_plane, materials: *const rtx_material, n: u64) -> rtx_status;
    pub fn rtx_scene_add_triangles(scene: *mut rtx_scene, shapes: *const rtx_triangle, materials: *const rtx_material, n: u64) -> rtx_status;
    pub fn rtx_scene_clear_objects(scene: *mut rtx_scene) -> rtx_status;
    pub fn rtx_scene_object_count(scene: *const rtx_scene) -> u64;
    pub fn rtx_scene_render(scene: *mut rtx_scene, width: u32, height: u32, seed: u64, flags: u32,
                            out_rgb: *mut f64, out_primary_hit_id: *mut u32, out_stats: *mut rtx_stats) -> rtx_status;
    pub fn rtx_scene_render_device(scene: *mut rtx_scene, width: u32, height: u32, seed: u64, flags: u32, shard: *const rtx_shard,
                                   d_out_rgb: *mut c_void, d_out_ids: *mut c_void, d_out_counters: *mut c_void, stream: *mut c_void) -> rtx_status;
    pub fn rtx_scene_render_device_band(scene: *mut rtx_scene, width: u32, height: u32, seed: u64, flags: u32, shard: *const rtx_shard,
                                        tile_row_first: u32, tile_row_count: u32, d_out_rgb: *mut c_void, d_out_ids: *mut c_void,
                                        d_out_counters: *mut c_void, stream: *mut c_void) -> rtx_status;
    pub fn rtx_shard_band_tiles(width: u32, height: u32, index: u32, count: u32, tile_row_first: u32, tile_row_count: u32,
                                out_first_local_tile: *mut u32, out_tile_count: *mut u32);
    pub fn rtx_scene_render_to_rgb8(scene: *mut rtx_scene, width: u32, height: u32, seed: u64, flags: u32,
                                    out_rgb8: *mut u8, out_stats: *mut rtx_stats) -> rtx_status;
    pub fn rtx_last_error(scene: *const rtx_scene) -> *const c_char;
    pub fn rtx_device_count() -> i32;
}
