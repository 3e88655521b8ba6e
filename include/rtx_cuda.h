/* rtx_cuda.h — C ABI of the B200-native render hot path for the `rtx` (rust-raytracing) crate.
 *
 * This is the drop-in boundary: plain pointers and sizes, no C++/torch types.  Every struct is a
 * layout-compatible mirror of a reference type (the Rust side declares them #[repr(C)]); every entry
 * point names the reference interface it replaces (paths relative to the reference's src/).
 *
 * The path behind it — camera-ray generation, nearest-hit ray/sphere + ray/plane + ray/triangle,
 * shade/bounce/accumulate — runs ONLY as hand-written CUDA (sm_100a).  There is no CPU fallback: every
 * render entry point returns RTX_ERR_CUDA when no usable device exists.
 *
 * Numerics contract: all path arithmetic is IEEE-754 binary64, round-to-nearest, NEVER fused
 * (the reference is Rust f64, which does not contract a*b+c).  Results are a pure function of
 * (scene, width, height, seed): independent of GPU count, tile shape and launch shape.
 */
#ifndef RTX_CUDA_H_
#define RTX_CUDA_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTX_CUDA_ABI_VERSION 1

/* ---- mirrors ------------------------------------------------------------------------------- */
typedef struct { double x, y, z; } rtx_vec3;                 /* math/vector.rs:12-20           24 B */
typedef struct { rtx_vec3 x, y, z; } rtx_mat3;               /* math/mat.rs:10-18 (ROW vectors) 72 B */
typedef struct {                                             /* raytracing/camera.rs:7-15     200 B */
  double   fov;                                              /*   horizontal, radians               */
  rtx_vec3 position;
  rtx_vec3 direction;
  rtx_mat3 to_cam_space;
  rtx_mat3 to_world_space;
} rtx_camera;
typedef struct {                                             /* raytracing/scene.rs:16-28      40 B */
  uint64_t rays_per_pixel;                                   /*   usize                             */
  uint64_t max_bounces;                                      /*   usize; loop runs max_bounces + 1  */
  double   focal_length;
  double   focal_offset;
  double   non_focal_offset;
} rtx_config;
typedef struct {                                             /* raytracing/object.rs:79-89     56 B */
  rtx_vec3 base_color;
  rtx_vec3 emission_color;
  double   roughness;
} rtx_material;
typedef struct { rtx_vec3 position; double radius; } rtx_sphere;   /* object/sphere.rs:9-12    32 B */
typedef struct { rtx_vec3 position, normal; } rtx_plane;           /* object/plane.rs:9-12     48 B */
typedef struct { rtx_vec3 vertices[3]; } rtx_triangle;             /* object/triangle.rs:9-11  72 B */

typedef struct rtx_scene rtx_scene;                          /* opaque; ≙ raytracing/scene.rs:78-85 */

/* ---- status -------------------------------------------------------------------------------- */
typedef int32_t rtx_status;
enum {
  RTX_OK                    = 0,
  RTX_ERR_INVALID_ARGUMENT  = 1,
  RTX_ERR_CUDA              = 2,  /* no device / CUDA runtime error (message in rtx_last_error)     */
  RTX_ERR_OUT_OF_MEMORY     = 3,
  RTX_ERR_UNSUPPORTED_SHAPE = 4,  /* user-defined CustomShape (object.rs:53-76) cannot cross the ABI */
  RTX_ERR_INTERNAL          = 5,
  RTX_ERR_REFERENCE_PANIC   = 6   /* rtx_scene_render / rtx_scene_render_to_rgb8: some ray met a triangle for which the
                                     reference's assert_eq! (object/triangle.rs:76-94) fails — there the row thread
                                     panics and Scene::render aborts through join().unwrap() (scene.rs:167-169).  The
                                     outputs and statistics ARE complete (such hits count as misses;
                                     rtx_stats.triangle_panics says how many), but the reference would have produced
                                     no image.  The device-buffer entry points report the count in d_out_counters[2] */
};

/* ---- render flags -------------------------------------------------------------------------- */
enum {
  RTX_ACCEL_AUTO      = 0u,       /* brute force for small scenes, LBVH above a primitive threshold  */
  RTX_ACCEL_BRUTE     = 1u,       /* scene.rs:243-251 as written: O(N) scan                          */
  RTX_ACCEL_LBVH      = 2u,       /* GPU-built Morton LBVH; identical results to RTX_ACCEL_BRUTE     */
  RTX_ACCEL_MASK      = 3u,
  RTX_FLAG_NO_RADIANCE = 4u,      /* hit-ID-only render (out_rgb may be NULL)                        */
  RTX_FLAG_LANE_KERNEL = 8u,      /* force the one-pixel-per-lane kernel where the context-pool kernel
                                     would run (same image; for A/B tests and measurements)            */
  RTX_FLAG_PIXEL_QUEUE = 32u,     /* context-pool kernel: force whole-pixel queue entries (a context runs all
                                     samples of its pixel)                                             */
  RTX_FLAG_SAMPLE_QUEUE = 64u,    /* context-pool kernel: force (pixel, sample) queue entries + ordered sum
                                     pass (short tail: small images, multi-GPU shards).  Default: chosen
                                     from pixels per resident context.  Same image either way.         */
  RTX_FLAG_RNG_ROW    = 256u,     /* the reference's literal RNG consumption: one stream per image ROW (≙ the row thread's
                                     thread-local fastrand, scene.rs:149-161, math/vector.rs:29-45), every draw in the
                                     reference's order, one CUDA thread per row.  A different image from the default
                                     per-(pixel, sample) streams (only the seeding differs); slow by construction; whole
                                     image only.  Bit-exact against the oracle's RNG_ROW mode.                        */
  RTX_FLAG_SHARD_ROWS = 128u,     /* shard by full-width STRIPS of 4 pixel rows (strip s belongs to shard s % count)
                                     instead of 16x16 tiles, written in place like RTX_FLAG_SHARD_IN_PLACE: a
                                     shard's pixels are whole contiguous rows of the frame, which each GPU's copy
                                     engine can deliver to a shared host frame over its own PCIe link
                                     (rtx_scene_render_rows_to_host).  The image does not depend on ownership.  */
  RTX_FLAG_SHARD_IN_PLACE = 16u   /* with shard->count > 1: write this shard's pixels at their final
                                     position of the FULL image layout instead of a compact buffer.
                                     d_out_* may then be another GPU's memory (rtx_peer_open): every
                                     rank stores straight into rank 0's frame over NVLink while it
                                     renders — the gather of scene.rs:167-169 fused into the kernel   */
};

/* Per-render statistics (all ranks report their own shard). */
typedef struct {
  uint64_t rays;            /* closest_object queries executed (scene.rs:231) — the metric's "ray"   */
  uint64_t paths;           /* camera samples started                                                */
  uint64_t pixels;          /* pixels this call rendered                                             */
  uint64_t triangle_panics; /* hits rejected where the reference's assert_eq! (triangle.rs:76-94) would abort */
  float    device_ms;       /* CUDA-event time of the render kernel(s) on the launching stream       */
  float    total_ms;        /* host-side wall time of the whole call (copies included)               */
  uint32_t kernel_launches; /* kernels launched by this call                                         */
  uint32_t accel_used;      /* RTX_ACCEL_BRUTE or RTX_ACCEL_LBVH                                     */
  uint64_t nodes_visited;   /* LBVH: internal nodes popped (0 for brute force)                       */
  uint64_t prims_tested;    /* primitive distance evaluations                                        */
  uint64_t hits;            /* queries that hit an object (rays - hits = misses)                     */
} rtx_stats;

/* Image shard for multi-GPU rendering: the image is cut into RTX_TILE×RTX_TILE pixel tiles numbered
 * row-major; this call renders tiles t with t % count == index (interleaved for load balance). */
#define RTX_TILE 16u
typedef struct { uint32_t index, count; } rtx_shard;

/* ---- host-side helpers (plain f64, same evaluation order as the reference) ---------------------- */
/* ≙ Camera::new                       raytracing/camera.rs:19-28 (+ derive_to_world_space_mat :42-49) */
rtx_status rtx_camera_new(const rtx_vec3* position, const rtx_vec3* direction, double fov, rtx_camera* out);
/* ≙ Camera::set_direction             raytracing/camera.rs:35-40 — matrices lag one call, as upstream */
rtx_status rtx_camera_set_direction(rtx_camera* camera, const rtx_vec3* direction);
/* ≙ Camera::to_cam_space / to_world_space / rotate_to_world_space   raytracing/camera.rs:51-67       */
rtx_status rtx_camera_to_cam_space(const rtx_camera* camera, const rtx_vec3* v, rtx_vec3* out);
rtx_status rtx_camera_to_world_space(const rtx_camera* camera, const rtx_vec3* v, rtx_vec3* out);
rtx_status rtx_camera_rotate_to_world_space(const rtx_camera* camera, const rtx_vec3* v, rtx_vec3* out);
/* ≙ Config::default                   raytracing/scene.rs:55-65                                     */
rtx_status rtx_config_default(rtx_config* out);

/* ---- scene lifetime ------------------------------------------------------------------------ */
/* ≙ Scene::new                        raytracing/scene.rs:112-118.  device = CUDA ordinal.          */
rtx_status rtx_scene_create(const rtx_camera* camera, const rtx_config* config, int32_t device, rtx_scene** out);
rtx_status rtx_scene_destroy(rtx_scene* scene);
/* ≙ pub fields Scene.camera / Scene.config   raytracing/scene.rs:80-84                              */
rtx_status rtx_scene_set_camera(rtx_scene* scene, const rtx_camera* camera);
rtx_status rtx_scene_set_config(rtx_scene* scene, const rtx_config* config);
rtx_status rtx_scene_get_camera(const rtx_scene* scene, rtx_camera* out);
rtx_status rtx_scene_get_config(const rtx_scene* scene, rtx_config* out);
/* ≙ Scene::add_object ×n              raytracing/scene.rs:126-128.  Object index = global insertion
 * order across all three calls (it is the nearest-hit tie-break, scene.rs:250).                    */
rtx_status rtx_scene_add_spheres(rtx_scene* scene, const rtx_sphere* shapes, const rtx_material* materials, uint64_t n);
rtx_status rtx_scene_add_planes(rtx_scene* scene, const rtx_plane* shapes, const rtx_material* materials, uint64_t n);
rtx_status rtx_scene_add_triangles(rtx_scene* scene, const rtx_triangle* shapes, const rtx_material* materials, uint64_t n);
rtx_status rtx_scene_clear_objects(rtx_scene* scene);
/* In-place replacement of existing objects between renders.  ≙ mutating a shape behind its Arc<Mutex<…>>
 * (raytracing/object.rs:12) / overwriting a record of the wgpu backend's FrequentlyChangedBuffer
 * (raytracing/gpu/gpu_state/buffer.rs:69-115).  object_indices[i] = global insertion index of an object of the
 * SAME shape kind; materials may be NULL (keep the old ones).  Only the touched records are re-uploaded, and the
 * LBVH is REFITTED (same hierarchy, new boxes) instead of rebuilt — results stay identical to the scan because
 * the tree only culls.  A change that alters tree membership (a primitive becoming non-finite or a triangle
 * changing its elimination pivots) falls back to a rebuild.  rtx_scene_rebuild_accel forces a rebuild, e.g.
 * after primitives moved far and the refitted tree has become slow.                                   */
rtx_status rtx_scene_update_spheres(rtx_scene* scene, const uint64_t* object_indices, const rtx_sphere* shapes,
                                    const rtx_material* materials, uint64_t n);
rtx_status rtx_scene_update_planes(rtx_scene* scene, const uint64_t* object_indices, const rtx_plane* shapes,
                                   const rtx_material* materials, uint64_t n);
rtx_status rtx_scene_update_triangles(rtx_scene* scene, const uint64_t* object_indices, const rtx_triangle* shapes,
                                      const rtx_material* materials, uint64_t n);
rtx_status rtx_scene_rebuild_accel(rtx_scene* scene);
/* Number of LBVH refits / full builds this scene has performed (statistics; tests). */
uint32_t   rtx_scene_accel_refits(const rtx_scene* scene);
uint32_t   rtx_scene_accel_builds(const rtx_scene* scene);
/* Kernels the most recent rtx_scene_render_device call on this scene launched (statistics; bench.py). */
uint32_t   rtx_scene_last_launches(const rtx_scene* scene);
uint64_t   rtx_scene_object_count(const rtx_scene* scene);

/* ---- render -------------------------------------------------------------------------------- */
/* ≙ Scene::render                     raytracing/scene.rs:144-170.  HOST buffers; blocks until done.
 * out_rgb: double[height][width][3], row 0 = bottom of the picture, linear unclamped radiance
 * (≙ Vec<Vec<Vector3>> img[y][x]).  out_primary_hit_id: uint32[height][width] (object index hit by the
 * first ray of sample 0; 0xFFFFFFFF = miss) or NULL.  seed: selects the RNG streams (the reference
 * has no seed — its thread-local fastrand is time-seeded).                           */
rtx_status rtx_scene_render(rtx_scene* scene, uint32_t width, uint32_t height, uint64_t seed, uint32_t flags,
                            double* out_rgb, uint32_t* out_primary_hit_id, rtx_stats* out_stats);

/* Device-resident variant, asynchronous on `cuda_stream` (a cudaStream_t; NULL = default stream).
 * Renders the tiles of `shard` (NULL = whole image).  With shard->count == 1 (or NULL) d_out_rgb /
 * d_out_primary_hit_id use the full-image layout above.  With count > 1 they are COMPACT: local tile
 * j (global tile j*count + index) occupies [j][RTX_TILE][RTX_TILE][3] doubles / [j][RTX_TILE][RTX_TILE]
 * ids — equal-sized per-rank buffers that a gather concatenates (see rtx_deinterleave_tiles).
 * d_out_counters: uint64[8] device counters {rays, paths, triangle_panics, hits, nodes_visited, prims_tested, -, -} to accumulate into, or
 * NULL.  Time it with events on your stream.                                                       */
rtx_status rtx_scene_render_device(rtx_scene* scene, uint32_t width, uint32_t height, uint64_t seed, uint32_t flags,
                                   const rtx_shard* shard, void* d_out_rgb, void* d_out_primary_hit_id,
                                   void* d_out_counters, void* cuda_stream);
/* The same for a BAND of the image only: the rows of RTX_TILE x RTX_TILE tiles [tile_row_first, tile_row_first +
 * tile_row_count), i.e. pixel rows [16 first, min(height, 16 (first + count))), of the shard's tiles.  Buffers are
 * those of the whole frame (or of the whole compact shard); pixels outside the band are not touched.  Rendering
 * all bands of a frame gives the bytes of one rtx_scene_render_device call.  A caller pipelines with it: while
 * band b+1 renders, band b — complete on every rank after a barrier — is already being copied out.
 * (≙ the row-wise hand-over of Scene::render's join, raytracing/scene.rs:167-169)                          */
rtx_status rtx_scene_render_device_band(rtx_scene* scene, uint32_t width, uint32_t height, uint64_t seed, uint32_t flags,
                                        const rtx_shard* shard, uint32_t tile_row_first, uint32_t tile_row_count,
                                        void* d_out_rgb, void* d_out_primary_hit_id, void* d_out_counters,
                                        void* cuda_stream);
/* The whole-image rtx_scene_render_device for callers that cannot allocate device memory (CUDA graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * It runs the sample-queue kernel pair, whose scratch (work counter + 25 bytes per camera
 * path: pixels rounded up to whole 16x16 tiles x rays_per_pixel) lives in a static device array of
 * RTX_NOALLOC_SCRATCH_BYTES that every GPU which loads the library holds until exit; the call performs no
 * cudaMalloc / cudaFree.  Preconditions, else RTX_ERR_INVALID_ARGUMENT with a message: the scene is already
 * resident and has camera tables for this width/height/fov (one earlier render at this size through
 * rtx_scene_render_device during set-up — that one allocates), it is a scanned scene (RTX_ACCEL_BRUTE applies;
 * packed scene <= 16 KB), 1 <= rays_per_pixel, and the frame's scratch fits.  flags: RTX_ACCEL_AUTO/BRUTE,
 * RTX_FLAG_SAMPLE_QUEUE.  Output bytes equal rtx_scene_render_device's.  All calls on one GPU share the static
 * array: calls that overlap on the GPU (other streams, other scenes) race — the caller orders them.           */
#define RTX_NOALLOC_SCRATCH_BYTES 33554432u   /* 32 MiB */
rtx_status rtx_scene_render_device_noalloc(rtx_scene* scene, uint32_t width, uint32_t height, uint64_t seed,
                                           uint32_t flags, void* d_out_rgb, void* d_out_primary_hit_id,
                                           void* cuda_stream);
/* Multi-GPU frames bound for HOST memory without a rank-0 funnel: every GPU delivers its own pixels to their final
 * place in ONE host frame over its own PCIe link.  The shard is cut by rows (RTX_FLAG_SHARD_ROWS is implied),
 * rendered in bands into the scene's own device frame, and after each band the GPU's copy engine moves the band's
 * strips to host_rgb ([height][width][3] doubles, the WHOLE frame; pinned: rtx_host_alloc, or rtx_host_frame_* when
 * the GPUs belong to different processes) while the next band renders.  Asynchronous on `cuda_stream`, which is done
 * when this shard's rows are in host memory; the frame is complete when every shard's call is (≙ the join of
 * scene.rs:167-169).  host_ids: uint32[height][width] or NULL.  d_out_counters as in rtx_scene_render_device.   */
rtx_status rtx_scene_render_rows_to_host(rtx_scene* scene, uint32_t width, uint32_t height, uint64_t seed, uint32_t flags,
                                         const rtx_shard* shard, double* host_rgb, uint32_t* host_ids,
                                         void* d_out_counters, void* cuda_stream);
/* One pinned host frame shared by the processes of a node: POSIX shared memory `name` ("/rtx-frame-...") mapped and
 * registered with CUDA in each process.  create: exclusive, sized `bytes`; open: an existing one; close: unregister +
 * unmap, and unlink the name when unlink_name != NULL (the creator).                                             */
rtx_status rtx_host_frame_create(const char* name, uint64_t bytes, void** out);
rtx_status rtx_host_frame_open(const char* name, uint64_t bytes, void** out);
rtx_status rtx_host_frame_close(void* ptr, uint64_t bytes, const char* unlink_name);

/* ---- N GPUs of one process behind one handle ---------------------------------------------------------------
 * ≙ Scene::{new, add_object, render} (raytracing/scene.rs:112-170) for a caller that wants every GPU of the box
 * without processes, NCCL or Python: one scene per device, one worker thread per device inside the library.
 * rtx_multi_render blocks like Scene::render and returns the whole image in HOST memory (pinned for speed); each
 * GPU renders its rows and delivers them itself (rtx_scene_render_rows_to_host).  The image equals the one-GPU
 * image byte for byte.  rtx_multi_scene(i) is device i's scene handle, for the calls that have no rtx_multi_ form
 * (updates, wire ingest) — apply them to every device's scene.                                                  */
typedef struct rtx_multi rtx_multi;
rtx_status rtx_multi_create(const rtx_camera* camera, const rtx_config* config, const int32_t* devices, uint32_t n_devices,
                            rtx_multi** out);
rtx_status rtx_multi_destroy(rtx_multi* multi);
uint32_t   rtx_multi_device_count(const rtx_multi* multi);
rtx_scene* rtx_multi_scene(rtx_multi* multi, uint32_t i);
rtx_status rtx_multi_set_camera(rtx_multi* multi, const rtx_camera* camera);
rtx_status rtx_multi_set_config(rtx_multi* multi, const rtx_config* config);
rtx_status rtx_multi_add_spheres(rtx_multi* multi, const rtx_sphere* shapes, const rtx_material* materials, uint64_t n);
rtx_status rtx_multi_add_planes(rtx_multi* multi, const rtx_plane* shapes, const rtx_material* materials, uint64_t n);
rtx_status rtx_multi_add_triangles(rtx_multi* multi, const rtx_triangle* shapes, const rtx_material* materials, uint64_t n);
rtx_status rtx_multi_clear_objects(rtx_multi* multi);
rtx_status rtx_multi_render(rtx_multi* multi, uint32_t width, uint32_t height, uint64_t seed, uint32_t flags,
                            double* out_rgb, uint32_t* out_primary_hit_id, rtx_stats* out_stats);
/* RTX_FLAG_SHARD_ROWS arithmetic (pure host code): the strips of 4 pixel rows shard `index` of `count` owns (strip s
 * belongs to shard s % count and is its local strip s / count), and the local strips of the band of tile rows
 * [tile_row_first, tile_row_first + tile_row_count) = pixel rows [16 first, 16 (first + count)).               */
uint64_t   rtx_shard_strip_count(uint32_t height, uint32_t index, uint32_t count);
void       rtx_shard_band_strips(uint32_t height, uint32_t index, uint32_t count, uint32_t tile_row_first,
                                 uint32_t tile_row_count, uint32_t* out_first_local_strip, uint32_t* out_strip_count);
/* Number of tiles (and therefore compact-buffer slots) shard `index` of `count` owns. */
uint64_t   rtx_shard_tile_count(uint32_t width, uint32_t height, uint32_t index, uint32_t count);
/* The local tiles (first, count — indices into the shard's tile list, which is also the order of its compact
 * buffer) that rtx_scene_render_device_band renders for the tile rows [tile_row_first, tile_row_first +
 * tile_row_count) of shard `index` of `count`.  Pure arithmetic; no device needed.                        */
void       rtx_shard_band_tiles(uint32_t width, uint32_t height, uint32_t index, uint32_t count,
                                uint32_t tile_row_first, uint32_t tile_row_count,
                                uint32_t* out_first_local_tile, uint32_t* out_tile_count);
/* Scatter gathered compact shard buffers (rank-major: [count][slots][RTX_TILE][RTX_TILE][...]) into
 * the full image layout on the device.  ≙ the row-ordered join of scene.rs:167-169.  Either pair of
 * pointers may be NULL.  slots = rtx_shard_tile_count(width, height, 0, count).                     */
rtx_status rtx_deinterleave_tiles(uint32_t width, uint32_t height, uint32_t count,
                                  const void* d_gathered_rgb, void* d_out_rgb,
                                  const void* d_gathered_ids, void* d_out_ids, void* cuda_stream);

/* ---- the reference's f32 wire format (GpuSerialize) ---------------------------------------------
 * ≙ gpu::Scene::new / add_object (raytracing/gpu/scene.rs:11-25, gpu/gpu_state.rs:31-77): the serialized
 * camera (camera.rs:70-78), config (scene.rs:67-76), object records (gpu/object.rs:34-42) and per-type shape
 * buffers (object/sphere.rs:36-42, plane.rs:38-46, triangle.rs:130-136) of a `gpu::Scene` are accepted as they
 * are; each f32 widens to f64 exactly and the scene is rendered with the CPU path's f64 semantics.
 * object_type = GpuShape::object_type() of the buffer ("…::sphere", "…::plane", "…::triangle"); a buffer's
 * position in the array is the shape_id the object records refer to.  The decode functions are pure host
 * code (no device needed).                                                                          */
#define RTX_WIRE_CAMERA_BYTES   32u
#define RTX_WIRE_CONFIG_BYTES   20u
#define RTX_WIRE_OBJECT_BYTES   48u
#define RTX_WIRE_SPHERE_BYTES   16u
#define RTX_WIRE_PLANE_BYTES    32u
#define RTX_WIRE_TRIANGLE_BYTES 48u
enum { RTX_WIRE_SPHERE = 0, RTX_WIRE_PLANE = 1, RTX_WIRE_TRIANGLE = 2 };
typedef struct { const char* object_type; const void* shape_bytes; uint64_t shape_count; } rtx_wire_shape_buffer;
rtx_status rtx_wire_decode_camera(const void* bytes /* [32] */, rtx_camera* out);
rtx_status rtx_wire_decode_config(const void* bytes /* [20] */, rtx_config* out);
/* out_kinds[i] = RTX_WIRE_*; out_shapes[i]: sphere = {vertices[0] = position, vertices[1].x = radius},
 * plane = {vertices[0] = position, vertices[1] = normal}, triangle = its three vertices.             */
rtx_status rtx_wire_decode_objects(const void* object_records, uint64_t n_objects,
                                   const rtx_wire_shape_buffer* shape_buffers, uint32_t n_shape_buffers,
                                   uint32_t* out_kinds, rtx_triangle* out_shapes, rtx_material* out_materials);
rtx_status rtx_scene_add_wire_objects(rtx_scene* scene, const void* object_records, uint64_t n_objects,
                                      const rtx_wire_shape_buffer* shape_buffers, uint32_t n_shape_buffers);
rtx_status rtx_scene_set_wire_camera(rtx_scene* scene, const void* bytes /* [32] */);
rtx_status rtx_scene_set_wire_config(rtx_scene* scene, const void* bytes /* [20] */);

/* ---- peer frame buffers (multi-GPU, one process per GPU) ------------------------------------------
 * The owner (rank 0) allocates the frame with rtx_peer_alloc and exports a 64-byte handle; the other
 * processes of the node open it and pass the returned pointer as d_out_rgb / d_out_primary_hit_id of
 * rtx_scene_render_device together with RTX_FLAG_SHARD_IN_PLACE.  Their render kernels then write
 * pixels directly into the owner's HBM through NVLink/NVSwitch peer stores; no gather, no
 * de-interleave.  (cudaIpc*; the handle is plain bytes and may be sent over any channel.)            */
#define RTX_PEER_HANDLE_BYTES 64u
rtx_status rtx_peer_alloc(int32_t device, uint64_t bytes, void** out_d_ptr);
rtx_status rtx_peer_free(int32_t device, void* d_ptr);
rtx_status rtx_peer_export(int32_t device, void* d_ptr, uint8_t* out_handle /* [64] */);
rtx_status rtx_peer_open(int32_t device, const uint8_t* handle /* [64] */, void** out_d_ptr);
rtx_status rtx_peer_close(int32_t device, void* d_ptr);

/* ≙ Scene::render_to_image            raytracing/scene.rs:173-178: vertical flip, ×256, saturating
 * `as u8`, no gamma.  out: uint8[height][width][3] (HOST), row 0 = top of the picture.             */
rtx_status rtx_scene_render_to_rgb8(rtx_scene* scene, uint32_t width, uint32_t height, uint64_t seed, uint32_t flags,
                                    uint8_t* out_rgb8, rtx_stats* out_stats);
/* ---- the f32 "preview" mode: the arithmetic of the reference's wgpu backend ----------------------------------
 * ≙ gpu::Scene::render (raytracing/gpu/scene.rs:27-29) → GpuState::render (gpu/gpu_state.rs:78-107) → the fragment
 * shader raytracing/gpu/base_shader.wgsl:64-267 with the shape snippets of object/sphere.rs:50-86, plane.rs:48-85,
 * triangle.rs:138-194.  Same scene handle, other semantics: f32 throughout, PCG-hash RNG seeded from the pixel (no
 * `seed`), max_bounces segments (not + 1), AABB pre-test, Moeller-Trumbore triangles, raw plane normals.  The scene,
 * camera and config are narrowed to f32 as GpuSerialize does (`as f32`).  Running the shader itself needs wgpu, which
 * is unavailable on this box; this is its arithmetic as a CUDA kernel (csrc/rtx_preview.cu), bit-exact against
 * oracle/wgsl_oracle.cpp.  out_rgb: float[height][width][3], row 0 = bottom, linear (what fs_main returns, before the
 * surface format); out_first_hit_id: uint32[height][width] object index hit by sample 0's first ray (0xFFFFFFFF =
 * miss) or NULL.  The host variant blocks; the device variant is asynchronous on `cuda_stream`
 * (d_out_counters: uint64[8], [0] += closest_object calls, or NULL).                                         */
rtx_status rtx_scene_render_preview_f32(rtx_scene* scene, uint32_t width, uint32_t height, float* out_rgb,
                                        uint32_t* out_first_hit_id, rtx_stats* out_stats);
rtx_status rtx_scene_render_preview_f32_device(rtx_scene* scene, uint32_t width, uint32_t height, void* d_out_rgb,
                                               void* d_out_first_hit_id, void* d_out_counters, void* cuda_stream);
/* Device-side quantiser used by the call above (d_rgb: full-image doubles → d_out: uint8 image). */
rtx_status rtx_quantize_rgb8_device(const void* d_rgb, uint32_t width, uint32_t height, void* d_out_rgb8, void* cuda_stream);

/* ---- diagnostics --------------------------------------------------------------------------- */
/* Message for the most recent failure on `scene` (or the last scene-less failure when NULL). */
const char* rtx_last_error(const rtx_scene* scene);
uint32_t    rtx_abi_version(void);
/* Number of usable CUDA devices (0 when none: every render call then fails with RTX_ERR_CUDA). */
int32_t     rtx_device_count(void);
/* Pinned host memory for out_rgb / out_primary_hit_id (optional; makes the D2H copy a single DMA). */
rtx_status  rtx_host_alloc(size_t bytes, void** out);
rtx_status  rtx_host_free(void* p);

/* Measured FP64 issue-rate peaks of the device, the roofline denominators for the brute-force regime
 * (non-fused DADD+DMUL, as the parity contract forbids FMA; and DFMA for context).  Tflop/s.       */
rtx_status rtx_probe_fp64_peak(int32_t device, double* out_nonfused_tflops, double* out_fma_tflops);

/* Per-function device evaluation for parity tests (HOST buffers, n items each).
 * rays[n][6] = position.xyz, direction.xyz.                                                        */
/* ≙ Scene::closest_object             raytracing/scene.rs:243-251 */
rtx_status rtx_debug_closest_object(rtx_scene* scene, uint32_t flags, const double* rays, uint64_t n,
                                    double* out_t, uint32_t* out_id);
/* ≙ Object::normal_at                 raytracing/object.rs:37-39 — ids[n], positions[n][3] → out[n][3] */
rtx_status rtx_debug_normal_at(rtx_scene* scene, const uint32_t* ids, const double* positions, uint64_t n, double* out);
/* ≙ random_bounce_dir                 raytracing/scene.rs:279-291 — in[n][7] = dir.xyz, normal.xyz,
 * roughness; states[n] = WyRand stream states (updated) → out[n][3]                                */
rtx_status rtx_debug_random_bounce_dir(int32_t device, uint64_t* states, const double* in, uint64_t n, double* out);
/* ≙ fastrand::f64()                   call sites math/vector.rs:31-33,37-38 — n draws from *state   */
rtx_status rtx_debug_wyrand_f64(int32_t device, uint64_t* state, uint64_t n, double* out);
/* ≙ Vector3 / f64                      math/vector/div.rs:11-21 — the device's shared-reciprocal form
 * (out) next to three plain IEEE divisions (ref); a[n][3], s[n] → out[n][3], ref[n][3]              */
rtx_status rtx_debug_vdiv3(int32_t device, const double* a, const double* s, uint64_t n, double* out, double* ref);
/* Returns fl(fl(a*b)+c) as the library's device code computes `a*b+c`: proves no FMA contraction.  */
rtx_status rtx_debug_contraction_probe(int32_t device, double a, double b, double c, double* out);
/* The seeding convention: WyRand state of the stream with the given key. */
uint64_t   rtx_stream_state(uint64_t seed, uint64_t key);

#ifdef __cplusplus
}
#endif
#endif /* RTX_CUDA_H_ */
