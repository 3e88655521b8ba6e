// rtx_internal.h — the opaque scene handle and helpers shared by the C-ABI translation units.
#pragma once
#include <cuda_runtime.h>

#include <mutex>
#include <string>
#include <vector>

#include "../../include/rtx_cuda.h"
#include "rtx_lbvh.cuh"
#include "rtx_render.cuh"

struct rtx_scene {
  int device = 0;
  int sm_count = 148;
  rtx_camera camera{};
  rtx_config config{};
  std::mutex mu;                 // serialises render / upload on this handle (Scene::render is &self)
  mutable std::string last_error;

  // host-side packed records (see rtx_scene.cuh); object index = insertion order across all three
  std::vector<double> sph, pl, tri, mat;
  std::vector<uint32_t> sph_id, pl_id, tri_id;
  std::vector<float> sph32;     // f32 copy of `sph` for the scan's conservative cull (rebuilt at upload)
  std::vector<double> tri_raw;  // 9 doubles per triangle: the LBVH builder needs the vertices for boxes
  std::vector<double> sph_radius;  // per sphere: the radius as given (`sph` keeps radius*radius); the f32 preview needs it
  std::vector<uint32_t> obj_ref;  // per object: REF_* | index within its shape array
  uint64_t n_obj = 0;
  bool dirty = true;              // layout changed: rebuild and upload the whole blob
  // in-place updates since the last upload (indices within the shape arrays / object indices for materials)
  std::vector<uint32_t> upd_sph, upd_pl, upd_tri, upd_mat;

  // device copies
  unsigned char* d_blob = nullptr;
  size_t d_blob_cap = 0;
  rtx::SceneLayout layout{};

  // LBVH: [0] 3-D tree (spheres + triangles), [1..3] prism trees per dropped axis (rtx_lbvh.cuh)
  rtx::WideNode* d_bvh[4] = {nullptr, nullptr, nullptr, nullptr};
  float4* d_bvh_aux[4] = {nullptr, nullptr, nullptr, nullptr};  // prism trees: per-child plane-gradient bounds (rtx_lbvh.cuh)
  uint32_t bvh_root[4] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu};
  uint32_t bvh_count[4] = {0, 0, 0, 0};
  uint32_t bvh_build_launches = 0;
  bool bvh_valid = false;
  bool bvh_refit = false;   // primitives moved since the trees were built/refitted: same hierarchy, new boxes
  uint32_t bvh_refits = 0, bvh_builds = 0;  // refits / full builds performed (statistics)
  uint32_t last_launches = 0;               // kernels launched by the most recent rtx_scene_render_device
  // what a refit needs of each tree: its primitive list, validity, sorted order and hierarchy
  struct TreeTopo {
    uint32_t n = 0, m = 0;
    int open_axis = 3;
    uint32_t *d_prims = nullptr, *d_vals = nullptr, *d_left = nullptr, *d_right = nullptr, *d_parent = nullptr;
    uint8_t* d_valid = nullptr;
  } topo[4];
  double* d_tri_raw = nullptr;  // device copy of tri_raw (boxes are computed from the vertices)
  size_t d_tri_raw_cap = 0;

  // separable sin/cos tables of get_ray_dir, cached per (width, height, fov)
  double* d_tables = nullptr;  // [sin_x W][cos_x W][sin_y H][cos_y H]
  size_t d_tables_cap = 0;
  uint32_t tab_w = 0, tab_h = 0;
  double tab_fov = 0.0;
  bool tab_valid = false;

  double* d_tail = nullptr;  // per-sample radiance of the sample-granular end of the pool kernel's queue
  size_t d_tail_cap = 0;
  uint32_t* d_work_counter = nullptr;
  unsigned long long* d_counters = nullptr;  // [0] rays [1] paths [2] panics

  // framebuffers owned by the host-buffer entry points
  double* d_rgb = nullptr;
  size_t d_rgb_cap = 0;
  uint32_t* d_ids = nullptr;
  size_t d_ids_cap = 0;
  uint8_t* d_rgb8 = nullptr;
  size_t d_rgb8_cap = 0;

  // f32 preview (rtx_preview.cu): nine float4 arrays of the object list; pv_valid drops whenever an object changes
  float* d_pv = nullptr;
  size_t d_pv_cap = 0;
  uint32_t pv_stride = 0, pv_bytes = 0;
  bool pv_valid = false;

  cudaStream_t stream = nullptr;  // used by the blocking host-buffer entry points
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  // device-side ordering of renders of this scene that were enqueued on different streams (rtx_api.cu, render_common)
  cudaEvent_t ev_last = nullptr;
  cudaStream_t last_stream = nullptr;
  bool last_valid = false;
  uint32_t wc_slot = 0;  // next work-counter slot for the asynchronous entry points
  // banded host output (rtx_scene_render into pinned memory): a copy stream and one event per band + one for the end
  cudaStream_t copy_stream = nullptr;
  static constexpr uint32_t MAX_BANDS = 8;
  cudaEvent_t ev_band[MAX_BANDS + 1] = {};
};

namespace rtx {
rtx_status fail(const rtx_scene* s, rtx_status code, const std::string& msg);
rtx_status cuda_fail(const rtx_scene* s, cudaError_t e, const char* what);
// Uploads the packed scene if it changed; fills s->layout.  Asynchronous on `stream`.
rtx_status ensure_uploaded(rtx_scene* s, cudaStream_t stream);
// Builds (or reuses) the sin/cos tables for this image size.  Asynchronous on `stream`.
// Per-launch margins of the scan's sphere cull (SceneLayout::cull_*)
void cull_constants(SceneLayout* L, const rtx_camera& cam);
rtx_status ensure_tables(rtx_scene* s, uint32_t width, uint32_t height, cudaStream_t stream);
cudaError_t launch_render(const RenderParams& p, int sm_count, cudaStream_t stream, uint32_t* out_launches);
// RTX_FLAG_RNG_ROW: one thread per image row, one RNG stream per row (rtx_render.cu)
cudaError_t launch_render_rows(const RenderParams& p, cudaStream_t stream, uint32_t* out_launches);
// rtx_render_pool.cu: the context-pool kernel for scanned, shared-memory-resident scenes
bool pool_kernel_eligible(const RenderParams& p);
// Fills p.tail_first / p.tail_granules; returns the bytes p.tail_scratch must hold (0 = no tail region).
size_t pool_plan_tail(RenderParams& p, int sm_count);
// Bands (1..8) a large frame bound for pinned host memory is rendered in, sample queue each, so that a band's
// device-to-host copy overlaps the next band's render; 1 = not worth it / not possible.
uint32_t pool_host_bands(const RenderParams& p, int sm_count);
cudaError_t launch_render_pool(const RenderParams& p, int sm_count, cudaStream_t stream, uint32_t* out_launches);
// Builds the LBVH trees if the object list changed since the last build.  Synchronous.
rtx_status ensure_lbvh(rtx_scene* s, cudaStream_t stream);
void free_lbvh(rtx_scene* s);
void fill_bvh_view(const rtx_scene* s, bool use_bvh, BvhView* bv);
// Which structure a render with these flags will use (RTX_ACCEL_BRUTE or RTX_ACCEL_LBVH).
uint32_t choose_accel(const rtx_scene* s, uint32_t flags);
}  // namespace rtx

#define RTX_CUDA_TRY(scene, expr)                                      \
  do {                                                                 \
    cudaError_t _e = (expr);                                           \
    if (_e != cudaSuccess) return rtx::cuda_fail((scene), _e, #expr);  \
  } while (0)
