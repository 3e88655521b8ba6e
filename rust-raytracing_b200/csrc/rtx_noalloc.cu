// rtx_noalloc.cu — rtx_scene_render_device_noalloc: the sample-queue render (render_pool_kernel + tail_sum_kernel,
// rtx_render_pool.cu) for callers that cannot allocate device memory: code under CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc.  It has no counterpart in the
// reference.
//
// The kernel pair needs scratch memory — the work-queue counter and one radiance slot + flag byte per camera path
// (pool_plan_tail) — which every other entry point takes from cudaMalloc'ed buffers the scene owns.  Here the
// scratch is a file-scope __device__ array of RTX_NOALLOC_SCRATCH_BYTES: every GPU that loads the library holds it
// until the process exits, nothing is allocated or freed per call, and a frame that needs more is refused.
//
// The array's NAME is a host-side stand-in in host code, not a device address, so the address is taken once per
// device with cudaGetSymbolAddress (that device current) and that pointer is what the kernels get.
//
// All calls on a device share the array.  What has to start at zero — the work counter — is cleared on the caller's
// stream before the launch; the radiance slots and flags are written before they are read within one call.  Two
// calls that overlap on the GPU (different streams, or different scenes) race on it: the caller orders them.
#include <cstring>
#include <string>

#include "rtx_internal.h"

namespace rtx {

constexpr size_t NOALLOC_COUNTER_BYTES = 256;  // work counter, padded so the radiance slots stay 256-aligned
static_assert(RTX_NOALLOC_SCRATCH_BYTES > NOALLOC_COUNTER_BYTES, "scratch too small");

static __device__ __align__(256) unsigned char g_noalloc_scratch[RTX_NOALLOC_SCRATCH_BYTES];

constexpr int NOALLOC_MAX_DEVICES = 64;
static std::mutex g_noalloc_mu;
static unsigned char* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};

// Device address of this device's copy of g_noalloc_scratch.  `device` is current.
static cudaError_t noalloc_scratch(int device, unsigned char** out) {
  *out = nullptr;
  if (device < 0 || device >= NOALLOC_MAX_DEVICES) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_ptr[device]) {
    void* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
    if (e != cudaSuccess) return e;
    if (!ptr) return cudaErrorInvalidSymbol;
    g_noalloc_ptr[device] = static_cast<unsigned char*>(ptr);
  }
  *out = g_noalloc_ptr[device];
  return cudaSuccess;
}

}  // namespace rtx

using namespace rtx;

extern "C" rtx_status rtx_scene_render_device_noalloc(rtx_scene* s, uint32_t width, uint32_t height, uint64_t seed,
                                                      uint32_t flags, void* d_out_rgb, void* d_out_primary_hit_id,
                                                      void* cuda_stream) {
  if (!s) return fail(s, RTX_ERR_INVALID_ARGUMENT, "null scene");
  if (!d_out_rgb) return fail(s, RTX_ERR_INVALID_ARGUMENT, "rtx_scene_render_device_noalloc: d_out_rgb is null");
  if (flags & ~(uint32_t)(RTX_ACCEL_MASK | RTX_FLAG_SAMPLE_QUEUE))
    return fail(s, RTX_ERR_INVALID_ARGUMENT,
                "rtx_scene_render_device_noalloc: only RTX_ACCEL_AUTO / RTX_ACCEL_BRUTE and RTX_FLAG_SAMPLE_QUEUE are accepted");
  if (width == 0 || height == 0) return fail(s, RTX_ERR_INVALID_ARGUMENT, "width and height must be non-zero");
  if ((uint64_t)width * height > 0xFFFFFFFFull) return fail(s, RTX_ERR_INVALID_ARGUMENT, "image too large");
  std::lock_guard<std::mutex> lk(s->mu);
  // Uploading the scene or the per-size sin/cos tables allocates: that belongs to the caller's set-up phase.
  const bool pending = !s->upd_sph.empty() || !s->upd_pl.empty() || !s->upd_tri.empty() || !s->upd_mat.empty();
  if (s->dirty || !s->d_blob || pending)
    return fail(s, RTX_ERR_INVALID_ARGUMENT,
                "rtx_scene_render_device_noalloc: the scene is not resident on the device (objects were added or "
                "updated since the last render); render once through rtx_scene_render_device during set-up");
  if (!s->tab_valid || s->tab_w != width || s->tab_h != height || std::memcmp(&s->tab_fov, &s->camera.fov, 8) != 0)
    return fail(s, RTX_ERR_INVALID_ARGUMENT,
                "rtx_scene_render_device_noalloc: no camera tables for this width/height/fov on the device; render "
                "once at this size through rtx_scene_render_device during set-up");
  if (choose_accel(s, flags) != RTX_ACCEL_BRUTE)
    return fail(s, RTX_ERR_INVALID_ARGUMENT,
                "rtx_scene_render_device_noalloc: only scanned scenes (RTX_ACCEL_BRUTE) use the scratch-backed sample queue");
  RTX_CUDA_TRY(s, cudaSetDevice(s->device));
  unsigned char* scratch = nullptr;
  RTX_CUDA_TRY(s, noalloc_scratch(s->device, &scratch));

  RenderParams p{};
  p.cam_pos[0] = s->camera.position.x, p.cam_pos[1] = s->camera.position.y, p.cam_pos[2] = s->camera.position.z;
  const rtx_mat3& m = s->camera.to_world_space;
  const double rows[9] = {m.x.x, m.x.y, m.x.z, m.y.x, m.y.y, m.y.z, m.z.x, m.z.y, m.z.z};
  std::memcpy(p.to_world, rows, sizeof(rows));
  p.focal_length = s->config.focal_length;
  p.focal_offset = s->config.focal_offset;
  p.non_focal_offset = s->config.non_focal_offset;
  p.rays_per_pixel = s->config.rays_per_pixel;
  p.max_bounces = s->config.max_bounces;
  p.seed = seed;
  p.width = width, p.height = height;
  p.sin_x = s->d_tables;
  p.cos_x = p.sin_x + width;
  p.sin_y = p.cos_x + width;
  p.cos_y = p.sin_y + height;
  p.scene_blob = s->d_blob;
  p.layout = s->layout;
  cull_constants(&p.layout, s->camera);
  p.queue_mode = 2;  // the sample queue: the kernel pair that uses scratch
  p.shard_index = 0, p.shard_count = 1;
  p.tiles_x = (width + TILE - 1) / TILE, p.tiles_y = (height + TILE - 1) / TILE;
  p.my_tiles = p.tiles_x * p.tiles_y;
  p.grans_per_unit = GRANS_PER_TILE;
  p.work_counter = reinterpret_cast<uint32_t*>(scratch);
  p.out_rgb = static_cast<double*>(d_out_rgb);
  p.out_ids = static_cast<uint32_t*>(d_out_primary_hit_id);
  if (!pool_kernel_eligible(p))
    return fail(s, RTX_ERR_INVALID_ARGUMENT,
                "rtx_scene_render_device_noalloc: only the context-pool kernel uses scratch — it takes scanned scenes whose "
                "packed size is <= 16 KB, rendered with max_bounces >= 1");
  const size_t need = pool_plan_tail(p, s->sm_count);
  if (need == 0)
    return fail(s, RTX_ERR_INVALID_ARGUMENT,
                "rtx_scene_render_device_noalloc: the sample queue needs 1 <= rays_per_pixel <= 65536 and < 2^32 camera paths");
  if (need > RTX_NOALLOC_SCRATCH_BYTES - NOALLOC_COUNTER_BYTES)
    return fail(s, RTX_ERR_INVALID_ARGUMENT,
                "rtx_scene_render_device_noalloc: this frame needs " + std::to_string(need) +
                    " bytes of scratch (25 per camera path, paths rounded up to whole 16x16 tiles), the static scratch holds " +
                    std::to_string(RTX_NOALLOC_SCRATCH_BYTES - NOALLOC_COUNTER_BYTES) +
                    "; use rtx_scene_render_device or render fewer pixels / samples per call");
  p.tail_scratch = reinterpret_cast<double*>(scratch + NOALLOC_COUNTER_BYTES);
  uint32_t launches = 0;
  // the work queue's two words must start at zero (a launch leaves them zero, but the array is shared with every other
  // caller on this GPU: clear them); then render_pool_kernel + tail_sum_kernel: no allocation
  RTX_CUDA_TRY(s, cudaMemsetAsync(scratch, 0, 2 * sizeof(uint32_t), (cudaStream_t)cuda_stream));
  const cudaError_t e = launch_render_pool(p, s->sm_count, (cudaStream_t)cuda_stream, &launches);
  if (e != cudaSuccess) return cuda_fail(s, e, "rtx_scene_render_device_noalloc launch");
  s->last_launches = launches;
  return RTX_OK;
}
