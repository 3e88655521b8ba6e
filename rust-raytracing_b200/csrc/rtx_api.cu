// rtx_api.cu — the C ABI declared in include/rtx_cuda.h: scene lifetime, upload, render entry points.
//
// Host-side mirror of raytracing/scene.rs (Scene::new / add_object / render / render_to_image) and
// raytracing/camera.rs (Camera::new / set_direction / to_*_space).  The hot path itself lives in
// rtx_render.cu; nothing here computes a pixel on the CPU.
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <cerrno>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>

#include "rtx_internal.h"
#include "rtx_trace.cuh"

namespace rtx {

static thread_local std::string g_last_error;

rtx_status fail(const rtx_scene* s, rtx_status code, const std::string& msg) {
  g_last_error = msg;
  if (s) s->last_error = msg;
  return code;
}
rtx_status cuda_fail(const rtx_scene* s, cudaError_t e, const char* what) {
  std::string msg = std::string("CUDA error: ") + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ") in " + what;
  cudaGetLastError();  // clear the sticky-less error state
  return fail(s, e == cudaErrorMemoryAllocation ? RTX_ERR_OUT_OF_MEMORY : RTX_ERR_CUDA, msg);
}

static inline V3 to_v3(const rtx_vec3& v) { return mk3(v.x, v.y, v.z); }
static inline rtx_vec3 from_v3(V3 v) {
  rtx_vec3 r;
  r.x = v.x, r.y = v.y, r.z = v.z;
  return r;
}
static inline M3 to_m3(const rtx_mat3& m) {
  M3 r;
  r.x = to_v3(m.x), r.y = to_v3(m.y), r.z = to_v3(m.z);
  return r;
}
static inline rtx_mat3 from_m3(const M3& m) {
  rtx_mat3 r;
  r.x = from_v3(m.x), r.y = from_v3(m.y), r.z = from_v3(m.z);
  return r;
}

// raytracing/camera.rs:42-49
static M3 derive_to_world_space_mat(V3 direction) {
  const V3 cam_forward = vnorm(direction);
  const V3 cam_right = cross(cam_forward, mk3(0., 0., -1.));
  const V3 cam_up = cross(cam_forward, cam_right);
  M3 m;
  m.x = cam_right, m.y = cam_up, m.z = cam_forward;
  return transpose(m);
}

template <typename T>
static cudaError_t grow(T** p, size_t* cap, size_t need_bytes) {
  if (*cap >= need_bytes && *p) return cudaSuccess;
  if (*p) cudaFree(*p);
  *p = nullptr;
  *cap = 0;
  cudaError_t e = cudaMalloc((void**)p, need_bytes);
  if (e == cudaSuccess) *cap = need_bytes;
  return e;
}

static inline uint32_t align16(size_t v) { return (uint32_t)((v + 15) & ~(size_t)15); }

// Re-uploads only the records touched by rtx_scene_update_* (the layout is unchanged): element by element for a
// few, the whole array of that kind for many.
template <typename T>
static cudaError_t upload_touched(unsigned char* d_base, const std::vector<T>& host, size_t stride_elems,
                                  std::vector<uint32_t>& touched, cudaStream_t stream) {
  if (touched.empty()) return cudaSuccess;
  cudaError_t e = cudaSuccess;
  if (touched.size() > 256) {
    e = cudaMemcpyAsync(d_base, host.data(), host.size() * sizeof(T), cudaMemcpyHostToDevice, stream);
  } else {
    for (uint32_t i : touched) {
      const size_t o = (size_t)i * stride_elems;
      e = cudaMemcpyAsync(d_base + o * sizeof(T), host.data() + o, stride_elems * sizeof(T), cudaMemcpyHostToDevice, stream);
      if (e != cudaSuccess) break;
    }
  }
  touched.clear();
  return e;
}

// f32 copy of the sphere records for the scan's conservative cull (rtx_trace.cuh, SphereCull) and the scene-wide
// bounds its error margin is built from.  Rounded to nearest; the bounds are rounded up.
static void build_sphere_cull(const std::vector<double>& sph, std::vector<float>& out, float* cmax, float* r2max) {
  out.resize(sph.size());
  double c = 0.0, r2 = 0.0;
  bool finite = true;
  for (size_t i = 0; i < sph.size(); ++i) {
    out[i] = (float)sph[i];
    finite = finite && std::isfinite(sph[i]);
    const double a = std::fabs(sph[i]);
    if ((i % SPH_STRIDE) == 3) r2 = a > r2 ? a : r2;
    else c = a > c ? a : c;
  }
  const float inf = std::numeric_limits<float>::infinity();
  *cmax = finite ? std::nextafterf((float)c, inf) : inf;
  *r2max = finite ? std::nextafterf((float)r2, inf) : inf;
}

// The sphere cull's margins for a whole launch (SceneLayout::cull_*): the formulas of sphere_cull_setup (rtx_trace.cuh)
// with R = pmax + sph_cmax, D1 = sqrt(3) and a = 1 + 1e-6, every step rounded up (computed in double, then to the next
// float above).  pmax covers the camera and four times the spheres' extent: bounce rays start on objects, and the few
// that start far out on an unbounded plane fall back to "keep every sphere".
void cull_constants(SceneLayout* L, const rtx_camera& cam) {
  L->cull_pmax = 0.0f, L->cull_neg_mq = 0.0f, L->cull_mh = 0.0f;
  const double cmax = L->sph_cmax, r2max = L->sph_r2max;
  const double camax = std::fmax(std::fabs(cam.position.x), std::fmax(std::fabs(cam.position.y), std::fabs(cam.position.z)));
  const double pmax = std::fmax(4.0 * cmax + 16.0, 2.0 * camax + 1.0);
  if (!(pmax < 1e15) || !(r2max < 1e30) || L->n_sph == 0) return;  // non-finite or huge: per-ray margins
  const double up = 1.0 + 1e-6;
  const double R = (pmax * up + cmax) * up;
  const double rd = R * 1.7320508075688774 * up;
  const double B = (rd * rd + (1.0 + 1e-6) * (3.0 * R * R + r2max)) * up;
  if (!(B < 1e37)) return;
  const float inf = std::numeric_limits<float>::infinity();
  L->cull_pmax = std::nextafterf((float)pmax, 0.0f);  // (rounded DOWN: the test is pmax_ray <= cull_pmax)
  L->cull_neg_mq = -std::nextafterf((float)(B * 1.9073486328125e-6 * up + 1e-30), inf);
  L->cull_mh = std::nextafterf((float)(rd * 9.5367431640625e-7 * up + 1e-30), inf);
}

rtx_status ensure_uploaded(rtx_scene* s, cudaStream_t stream) {
  if (!s->dirty && s->d_blob) {
    const bool any = !s->upd_sph.empty() || !s->upd_pl.empty() || !s->upd_tri.empty() || !s->upd_mat.empty();
    if (!any) return RTX_OK;
    const SceneLayout& L = s->layout;
    if (s->d_tri_raw && !s->upd_tri.empty()) {  // the LBVH's vertex copy follows the same updates
      std::vector<uint32_t> t = s->upd_tri;
      RTX_CUDA_TRY(s, upload_touched(reinterpret_cast<unsigned char*>(s->d_tri_raw), s->tri_raw, 9, t, stream));
    }
    if (!s->upd_sph.empty()) {
      std::vector<uint32_t> t = s->upd_sph;
      build_sphere_cull(s->sph, s->sph32, &s->layout.sph_cmax, &s->layout.sph_r2max);
      RTX_CUDA_TRY(s, upload_touched(s->d_blob + L.off_sph32, s->sph32, SPH_STRIDE, t, stream));
    }
    RTX_CUDA_TRY(s, upload_touched(s->d_blob + L.off_sph, s->sph, SPH_STRIDE, s->upd_sph, stream));
    RTX_CUDA_TRY(s, upload_touched(s->d_blob + L.off_pl, s->pl, PL_STRIDE, s->upd_pl, stream));
    RTX_CUDA_TRY(s, upload_touched(s->d_blob + L.off_tri, s->tri, TRI_STRIDE, s->upd_tri, stream));
    RTX_CUDA_TRY(s, upload_touched(s->d_blob + L.off_mat, s->mat, MAT_STRIDE, s->upd_mat, stream));
    RTX_CUDA_TRY(s, cudaStreamSynchronize(stream));  // pageable sources
    return RTX_OK;
  }
  s->upd_sph.clear(), s->upd_pl.clear(), s->upd_tri.clear(), s->upd_mat.clear();
  SceneLayout L{};
  L.n_obj = (uint32_t)s->n_obj;
  L.n_sph = (uint32_t)s->sph_id.size();
  L.n_pl = (uint32_t)s->pl_id.size();
  L.n_tri = (uint32_t)s->tri_id.size();
  size_t off = 0;
  L.off_sph = align16(off), off = L.off_sph + s->sph.size() * 8;
  L.off_pl = align16(off), off = L.off_pl + s->pl.size() * 8;
  L.off_tri = align16(off), off = L.off_tri + s->tri.size() * 8;
  L.off_mat = align16(off), off = L.off_mat + s->mat.size() * 8;
  L.off_sph_id = align16(off), off = L.off_sph_id + s->sph_id.size() * 4;
  L.off_pl_id = align16(off), off = L.off_pl_id + s->pl_id.size() * 4;
  L.off_tri_id = align16(off), off = L.off_tri_id + s->tri_id.size() * 4;
  build_sphere_cull(s->sph, s->sph32, &L.sph_cmax, &L.sph_r2max);
  L.off_sph32 = align16(off), off = L.off_sph32 + s->sph32.size() * 4;
  if (off > 0xFFFFFFF0ull) return fail(s, RTX_ERR_INVALID_ARGUMENT, "scene larger than 4 GiB is not supported");
  L.bytes = align16(off);
  if (L.bytes == 0) L.bytes = 16;
  std::vector<unsigned char> blob(L.bytes, 0);
  auto put = [&](uint32_t o, const void* src, size_t n) {
    if (n) std::memcpy(blob.data() + o, src, n);
  };
  put(L.off_sph, s->sph.data(), s->sph.size() * 8);
  put(L.off_pl, s->pl.data(), s->pl.size() * 8);
  put(L.off_tri, s->tri.data(), s->tri.size() * 8);
  put(L.off_mat, s->mat.data(), s->mat.size() * 8);
  put(L.off_sph_id, s->sph_id.data(), s->sph_id.size() * 4);
  put(L.off_pl_id, s->pl_id.data(), s->pl_id.size() * 4);
  put(L.off_tri_id, s->tri_id.data(), s->tri_id.size() * 4);
  put(L.off_sph32, s->sph32.data(), s->sph32.size() * 4);
  RTX_CUDA_TRY(s, grow(&s->d_blob, &s->d_blob_cap, L.bytes));
  // pageable source: the runtime stages it before returning, so `blob` may die at scope exit
  RTX_CUDA_TRY(s, cudaMemcpyAsync(s->d_blob, blob.data(), L.bytes, cudaMemcpyHostToDevice, stream));
  RTX_CUDA_TRY(s, cudaStreamSynchronize(stream));
  s->layout = L;
  s->dirty = false;
  return RTX_OK;
}

rtx_status ensure_tables(rtx_scene* s, uint32_t width, uint32_t height, cudaStream_t stream) {
  if (s->tab_valid && s->tab_w == width && s->tab_h == height &&
      std::memcmp(&s->tab_fov, &s->camera.fov, 8) == 0)
    return RTX_OK;
  const size_t n = 2 * ((size_t)width + height);
  std::vector<double> tab(n);
  const double fov = s->camera.fov;
  const double vertical_fov = (double)height / (double)width * fov;  // scene.rs:145
  double* sin_x = tab.data();
  double* cos_x = sin_x + width;
  double* sin_y = cos_x + width;
  double* cos_y = sin_y + height;
  for (uint32_t x = 0; x < width; ++x) {
    const double u = (double)x / (double)width;  // scene.rs:157
    const double angle_x = fov * (u - 0.5);      // scene.rs:214
    sin_x[x] = std::sin(angle_x);
    cos_x[x] = std::cos(angle_x);
  }
  for (uint32_t y = 0; y < height; ++y) {
    const double v = (double)y / (double)height;   // scene.rs:153
    const double angle_y = vertical_fov * (v - 0.5);  // scene.rs:215
    sin_y[y] = std::sin(angle_y);
    cos_y[y] = std::cos(angle_y);
  }
  RTX_CUDA_TRY(s, grow(&s->d_tables, &s->d_tables_cap, n * 8));
  RTX_CUDA_TRY(s, cudaMemcpyAsync(s->d_tables, tab.data(), n * 8, cudaMemcpyHostToDevice, stream));
  RTX_CUDA_TRY(s, cudaStreamSynchronize(stream));
  s->tab_w = width, s->tab_h = height, s->tab_fov = fov, s->tab_valid = true;
  return RTX_OK;
}

static uint64_t shard_tiles(uint32_t width, uint32_t height, uint32_t index, uint32_t count) {
  if (count == 0 || index >= count) return 0;
  const uint64_t tiles = (uint64_t)((width + TILE - 1) / TILE) * ((height + TILE - 1) / TILE);
  if (index >= tiles) return 0;
  return (tiles - index + count - 1) / count;
}

// The band of tile rows [row_first, row_first + row_count) as a range of a shard's LOCAL tiles: tile t of the image
// belongs to shard t mod count as its local tile t div count, so a contiguous range of image tiles is a contiguous
// range of local tiles.  Rows beyond the image are clipped.
static void shard_band_tiles(uint32_t width, uint32_t height, uint32_t index, uint32_t count, uint32_t row_first,
                             uint32_t row_count, uint32_t* first, uint32_t* n) {
  *first = 0, *n = 0;
  if (count == 0 || index >= count) return;
  const uint64_t tiles_x = (width + TILE - 1) / TILE, tiles_y = (height + TILE - 1) / TILE;
  const uint64_t lo = std::min<uint64_t>(row_first, tiles_y) * tiles_x;
  const uint64_t hi = std::min<uint64_t>((uint64_t)row_first + row_count, tiles_y) * tiles_x;
  auto local_below = [&](uint64_t g) { return g > index ? (g - index + count - 1) / count : 0ull; };  // local tiles with image index < g
  *first = (uint32_t)local_below(lo);
  *n = (uint32_t)(local_below(hi) - local_below(lo));
}
// RTX_FLAG_SHARD_ROWS: the ownership unit is a full-width strip of GRAN_H (4) pixel rows, strip s -> shard s mod count.
static uint64_t strip_count(uint32_t height) { return ((uint64_t)height + GRAN_H - 1) / GRAN_H; }
static uint64_t shard_strips(uint32_t height, uint32_t index, uint32_t count) {
  if (count == 0 || index >= count) return 0;
  const uint64_t strips = strip_count(height);
  return index >= strips ? 0 : (strips - index + count - 1) / count;
}
// The same band of tile rows — pixel rows [16 row_first, 16 (row_first + row_count)) — as a range of LOCAL strips.
static void shard_band_strips(uint32_t height, uint32_t index, uint32_t count, uint32_t row_first, uint32_t row_count,
                              uint32_t* first, uint32_t* n) {
  *first = 0, *n = 0;
  if (count == 0 || index >= count) return;
  const uint64_t strips = strip_count(height), per_row = TILE / GRAN_H;
  const uint64_t lo = std::min<uint64_t>((uint64_t)row_first * per_row, strips);
  const uint64_t hi = std::min<uint64_t>(((uint64_t)row_first + row_count) * per_row, strips);
  auto local_below = [&](uint64_t g) { return g > index ? (g - index + count - 1) / count : 0ull; };
  *first = (uint32_t)local_below(lo);
  *n = (uint32_t)(local_below(hi) - local_below(lo));
}

// ---- small device kernels owned by the API layer ------------------------------------------------
// ≙ the row-ordered join of scene.rs:167-169: scatter gathered per-rank tile buffers into the image.
__global__ void deinterleave_kernel(uint32_t width, uint32_t height, uint32_t tiles_x, uint32_t count,
                                    uint64_t slots, const double* __restrict__ g_rgb, double* __restrict__ out_rgb,
                                    const uint32_t* __restrict__ g_ids, uint32_t* __restrict__ out_ids) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (uint64_t)width * height) return;
  const uint32_t x = (uint32_t)(i % width), y = (uint32_t)(i / width);
  const uint32_t gtile = (y / TILE) * tiles_x + (x / TILE);
  const uint64_t src = (((uint64_t)(gtile % count) * slots + gtile / count) * TILE + (y % TILE)) * TILE + (x % TILE);
  if (g_rgb) {
    out_rgb[i * 3 + 0] = g_rgb[src * 3 + 0];
    out_rgb[i * 3 + 1] = g_rgb[src * 3 + 1];
    out_rgb[i * 3 + 2] = g_rgb[src * 3 + 2];
  }
  if (g_ids) out_ids[i] = g_ids[src];
}

// ≙ render_to_image, scene.rs:173-178: flip y, `* 256`, Rust's saturating `as u8` (NaN → 0).
__device__ __forceinline__ uint8_t sat_u8(double c) {
  const double s = fmul(c, 256.0);
  if (!(s > 0.0)) return 0;
  if (s >= 255.0) return 255;
  return (uint8_t)(int)s;
}
__global__ void quantize_kernel(const double* __restrict__ rgb, uint32_t width, uint32_t height,
                                uint8_t* __restrict__ out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (uint64_t)width * height) return;
  const uint32_t x = (uint32_t)(i % width), y = (uint32_t)(i / width);
  const double* src = rgb + ((uint64_t)(height - y - 1) * width + x) * 3;
  out[i * 3 + 0] = sat_u8(src[0]);
  out[i * 3 + 1] = sat_u8(src[1]);
  out[i * 3 + 2] = sat_u8(src[2]);
}

// RTX_ACCEL_AUTO: the O(N) scan while the whole scene fits the kernel's shared-memory stage and is small;
// the LBVH above that.  Planes never enter a tree, so a planes-only scene is always scanned.
uint32_t choose_accel(const rtx_scene* s, uint32_t flags) {
  const uint32_t a = flags & RTX_ACCEL_MASK;
  const size_t boxed = s->sph_id.size() + s->tri_id.size();
  if (boxed == 0) return RTX_ACCEL_BRUTE;
  if (a == RTX_ACCEL_BRUTE) return RTX_ACCEL_BRUTE;
  if (a == RTX_ACCEL_LBVH) return RTX_ACCEL_LBVH;
  return boxed > 48 ? RTX_ACCEL_LBVH : RTX_ACCEL_BRUTE;
}

// The device-side alias of a host pointer when the allocation is pinned and mapped (zero-copy), else null.
template <typename T>
static T* mapped_device_pointer(T* host) {
#ifdef RTX_MEASUREMENT_KNOBS  // measurement builds only: a stray environment variable must not change the product's path
  static const bool enabled = [] { const char* v = getenv("RTX_ZERO_COPY"); return !v || atoi(v) != 0; }();
  if (!enabled) return nullptr;
#endif
  if (!host) return nullptr;
  cudaPointerAttributes a{};
  if (cudaPointerGetAttributes(&a, host) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  if (a.type != cudaMemoryTypeHost || !a.devicePointer) return nullptr;
  return static_cast<T*>(a.devicePointer);
}

// Fills RenderParams and launches.  Counters / outputs are device pointers.
static rtx_status render_common_on_stream(rtx_scene* s, uint32_t width, uint32_t height, uint64_t seed, uint32_t flags,
                                const rtx_shard* shard, double* d_rgb, uint32_t* d_ids,
                                unsigned long long* d_counters, uint32_t* d_work_counter, cudaStream_t stream,
                                uint32_t* launches, double* host_rgb = nullptr, const uint32_t* tile_rows = nullptr) {
  // tile_rows != null: {first, count} — only that band of tile rows (rtx_scene_render_device_band)
  // host_rgb != null: d_rgb is the device alias of that pinned caller buffer (rtx_scene_render)
  const bool rgb_is_host = host_rgb != nullptr;
  if (width == 0 || height == 0) return fail(s, RTX_ERR_INVALID_ARGUMENT, "width and height must be non-zero");
  if ((uint64_t)width * height > 0xFFFFFFFFull) return fail(s, RTX_ERR_INVALID_ARGUMENT, "image too large");
  const uint32_t accel = flags & RTX_ACCEL_MASK;
  if (accel == RTX_ACCEL_MASK) return fail(s, RTX_ERR_INVALID_ARGUMENT, "invalid accel flags");
  rtx_shard sh{0, 1};
  if (shard) sh = *shard;
  if (sh.count == 0 || sh.index >= sh.count) return fail(s, RTX_ERR_INVALID_ARGUMENT, "invalid shard");
  rtx_status st = ensure_uploaded(s, stream);
  if (st != RTX_OK) return st;
  st = ensure_tables(s, width, height, stream);
  if (st != RTX_OK) return st;
  const bool use_bvh = choose_accel(s, flags) == RTX_ACCEL_LBVH;
  if (use_bvh) {
    st = ensure_lbvh(s, stream);
    if (st != RTX_OK) return st;
  }

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
  p.use_bvh = use_bvh ? 1u : 0u;
  p.lane_kernel = (flags & RTX_FLAG_LANE_KERNEL) ? 1u : 0u;
  p.queue_mode = (flags & RTX_FLAG_PIXEL_QUEUE) ? 1u : ((flags & RTX_FLAG_SAMPLE_QUEUE) ? 2u : 0u);
  fill_bvh_view(s, use_bvh, &p.bvh);
  p.shard_index = sh.index, p.shard_count = sh.count;
  p.tiles_x = (width + TILE - 1) / TILE, p.tiles_y = (height + TILE - 1) / TILE;
  const bool strips = (flags & RTX_FLAG_SHARD_ROWS) != 0;
  p.strip_mode = strips ? 1u : 0u;
  p.grans_per_unit = strips ? (width + GRAN_W - 1) / GRAN_W : GRANS_PER_TILE;
  p.my_tiles = strips ? (uint32_t)shard_strips(height, sh.index, sh.count) : (uint32_t)shard_tiles(width, height, sh.index, sh.count);
  p.compact = (sh.count > 1 && !(flags & (RTX_FLAG_SHARD_IN_PLACE | RTX_FLAG_SHARD_ROWS))) ? 1u : 0u;
  p.work_counter = d_work_counter;
  p.out_rgb = (flags & RTX_FLAG_NO_RADIANCE) ? nullptr : d_rgb;
  p.out_ids = d_ids;
  p.out_is_host = rgb_is_host ? 1u : 0u;
  p.ray_counter = d_counters;
  if (tile_rows) {
    // tile t of the image belongs to shard t mod count as its local tile t div count: a band of tile rows is a
    // contiguous range of local tiles
    if (strips) shard_band_strips(height, sh.index, sh.count, tile_rows[0], tile_rows[1], &p.tile_first, &p.my_tiles);
    else shard_band_tiles(width, height, sh.index, sh.count, tile_rows[0], tile_rows[1], &p.tile_first, &p.my_tiles);
  }
  if (flags & RTX_FLAG_RNG_ROW) {
    if (sh.count != 1 || tile_rows) return fail(s, RTX_ERR_INVALID_ARGUMENT, "RTX_FLAG_RNG_ROW renders the whole image: no shard, no band");
    cudaError_t e = launch_render_rows(p, stream, launches);
    if (e != cudaSuccess) return cuda_fail(s, e, "render_rows_kernel launch");
    return RTX_OK;
  }
  if (p.my_tiles == 0) return RTX_OK;
  p.tail_first = p.my_tiles * p.grans_per_unit, p.tail_granules = 0, p.tail_scratch = nullptr;
  const bool pool = !p.lane_kernel && pool_kernel_eligible(p);
  // A large frame bound for pinned host memory: bands of tile rows, each rendered with the sample queue into the
  // scene's device frame and copied out by the copy engine while the next band renders (C2 e2e 12.2 -> 11.5 ms
  // against the pixel queue storing every pixel over PCIe itself).
  const uint32_t bands = (pool && rgb_is_host && p.out_rgb) ? pool_host_bands(p, s->sm_count) : 1u;
  if (bands > 1u) {
    const size_t npix = (size_t)width * height;
    RTX_CUDA_TRY(s, grow(&s->d_rgb, &s->d_rgb_cap, npix * 3 * sizeof(double)));
    if (!s->copy_stream) RTX_CUDA_TRY(s, cudaStreamCreateWithFlags(&s->copy_stream, cudaStreamNonBlocking));
    for (cudaEvent_t& ev : s->ev_band)
      if (!ev) RTX_CUDA_TRY(s, cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    p.out_rgb = s->d_rgb;
    p.out_is_host = 0;
    for (uint32_t b = 0; b < bands; ++b) {
      // the last band is half as tall as the others: its copy is the one nothing overlaps
      const uint64_t den = 2ull * bands - 1ull;
      const uint32_t r0 = (uint32_t)((uint64_t)p.tiles_y * (2ull * b) / den);
      const uint32_t r1 = b + 1 == bands ? p.tiles_y : (uint32_t)((uint64_t)p.tiles_y * (2ull * (b + 1)) / den);
      if (r1 == r0) continue;
      if (strips) shard_band_strips(height, 0, 1, r0, r1 - r0, &p.tile_first, &p.my_tiles);
      else p.tile_first = r0 * p.tiles_x, p.my_tiles = (r1 - r0) * p.tiles_x;
      const size_t need = pool_plan_tail(p, s->sm_count);
      if (need) {
        RTX_CUDA_TRY(s, grow(&s->d_tail, &s->d_tail_cap, need));
        p.tail_scratch = s->d_tail;
      }
      cudaError_t e = launch_render(p, s->sm_count, stream, launches);
      if (e != cudaSuccess) return cuda_fail(s, e, "render kernel launch");
      RTX_CUDA_TRY(s, cudaEventRecord(s->ev_band[b], stream));
      RTX_CUDA_TRY(s, cudaStreamWaitEvent(s->copy_stream, s->ev_band[b], 0));
      const size_t row0 = (size_t)r0 * TILE, row1 = std::min((size_t)r1 * TILE, (size_t)height);
      RTX_CUDA_TRY(s, cudaMemcpyAsync(host_rgb + row0 * width * 3, s->d_rgb + row0 * width * 3,
                                      (row1 - row0) * width * 3 * sizeof(double), cudaMemcpyDeviceToHost, s->copy_stream));
    }
    // the caller's stream is done when the last copy is
    RTX_CUDA_TRY(s, cudaEventRecord(s->ev_band[rtx_scene::MAX_BANDS], s->copy_stream));
    RTX_CUDA_TRY(s, cudaStreamWaitEvent(stream, s->ev_band[rtx_scene::MAX_BANDS], 0));
    return RTX_OK;
  }
  if (pool) {
    const size_t need = pool_plan_tail(p, s->sm_count);
    if (need) {
      RTX_CUDA_TRY(s, grow(&s->d_tail, &s->d_tail_cap, need));
      p.tail_scratch = s->d_tail;
    }
  }
  cudaError_t e = launch_render(p, s->sm_count, stream, launches);
  if (e != cudaSuccess) return cuda_fail(s, e, "render kernel launch");
  return RTX_OK;
}

// Renders of one scene share device state the scene owns — the sample queue's scratch, the band frame, the sin/cos
// tables, the work counters — so they must not overlap on the GPU.  The host side is serialised by s->mu; the device
// side by an event: every render records s->ev_last behind its last piece of work, and a render enqueued on a
// DIFFERENT stream first makes that stream wait for it.  Two threads (or two streams) rendering one scene therefore
// get correct frames, one after the other (≙ Scene::render(&self) being safe to call concurrently, scene.rs:144).
static rtx_status render_common(rtx_scene* s, uint32_t width, uint32_t height, uint64_t seed, uint32_t flags,
                                const rtx_shard* shard, double* d_rgb, uint32_t* d_ids, unsigned long long* d_counters,
                                uint32_t* d_work_counter, cudaStream_t stream, uint32_t* launches, double* host_rgb = nullptr,
                                const uint32_t* tile_rows = nullptr) {
  // A stream under CUDA graph capture takes no part in this: an event recorded into a capture cannot be waited for
  // outside it.  The caller who captures renders orders them against its other work itself.
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(stream, &cap) != cudaSuccess) cudaGetLastError(), cap = cudaStreamCaptureStatusNone;
  if (cap != cudaStreamCaptureStatusNone)
    return render_common_on_stream(s, width, height, seed, flags, shard, d_rgb, d_ids, d_counters, d_work_counter, stream,
                                   launches, host_rgb, tile_rows);
  if (!s->ev_last) RTX_CUDA_TRY(s, cudaEventCreateWithFlags(&s->ev_last, cudaEventDisableTiming));
  if (s->last_valid && s->last_stream != stream) RTX_CUDA_TRY(s, cudaStreamWaitEvent(stream, s->ev_last, 0));
  const rtx_status st = render_common_on_stream(s, width, height, seed, flags, shard, d_rgb, d_ids, d_counters, d_work_counter,
                                                stream, launches, host_rgb, tile_rows);
  // (also after a failure: whatever was enqueued before it is still in flight)
  if (cudaEventRecord(s->ev_last, stream) == cudaSuccess) {
    s->last_stream = stream;
    s->last_valid = true;
  } else {
    cudaGetLastError();
  }
  return st;
}

}  // namespace rtx

using namespace rtx;

extern "C" {

uint32_t rtx_abi_version(void) { return RTX_CUDA_ABI_VERSION; }

const char* rtx_last_error(const rtx_scene* scene) {
  if (scene) return scene->last_error.c_str();
  return g_last_error.c_str();
}

int32_t rtx_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

rtx_status rtx_host_alloc(size_t bytes, void** out) {
  if (!out) return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "null out");
  RTX_CUDA_TRY(nullptr, cudaMallocHost(out, bytes ? bytes : 1));
  return RTX_OK;
}
// ---- peer frame buffers (cudaIpc) ----------------------------------------------------------------
rtx_status rtx_peer_alloc(int32_t device, uint64_t bytes, void** out_d_ptr) {
  if (!out_d_ptr) return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "null out");
  RTX_CUDA_TRY(nullptr, cudaSetDevice(device));
  RTX_CUDA_TRY(nullptr, cudaMalloc(out_d_ptr, bytes ? bytes : 1));
  return RTX_OK;
}
rtx_status rtx_peer_free(int32_t device, void* d_ptr) {
  RTX_CUDA_TRY(nullptr, cudaSetDevice(device));
  if (d_ptr) RTX_CUDA_TRY(nullptr, cudaFree(d_ptr));
  return RTX_OK;
}
rtx_status rtx_peer_export(int32_t device, void* d_ptr, uint8_t* out_handle) {
  static_assert(sizeof(cudaIpcMemHandle_t) == RTX_PEER_HANDLE_BYTES, "handle size");
  if (!d_ptr || !out_handle) return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "null argument");
  RTX_CUDA_TRY(nullptr, cudaSetDevice(device));
  cudaIpcMemHandle_t h;
  RTX_CUDA_TRY(nullptr, cudaIpcGetMemHandle(&h, d_ptr));
  std::memcpy(out_handle, &h, sizeof(h));
  return RTX_OK;
}
rtx_status rtx_peer_open(int32_t device, const uint8_t* handle, void** out_d_ptr) {
  if (!handle || !out_d_ptr) return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "null argument");
  RTX_CUDA_TRY(nullptr, cudaSetDevice(device));
  cudaIpcMemHandle_t h;
  std::memcpy(&h, handle, sizeof(h));
  RTX_CUDA_TRY(nullptr, cudaIpcOpenMemHandle(out_d_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return RTX_OK;
}
rtx_status rtx_peer_close(int32_t device, void* d_ptr) {
  RTX_CUDA_TRY(nullptr, cudaSetDevice(device));
  if (d_ptr) RTX_CUDA_TRY(nullptr, cudaIpcCloseMemHandle(d_ptr));
  return RTX_OK;
}

rtx_status rtx_host_free(void* p) {
  if (p) RTX_CUDA_TRY(nullptr, cudaFreeHost(p));
  return RTX_OK;
}

// ---- camera / config helpers ---------------------------------------------------------------------
rtx_status rtx_camera_new(const rtx_vec3* position, const rtx_vec3* direction, double fov, rtx_camera* out) {
  if (!position || !direction || !out) return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "null argument");
  const M3 to_world = derive_to_world_space_mat(to_v3(*direction));  // camera.rs:20
  out->fov = fov;
  out->position = *position;
  out->direction = *direction;
  out->to_cam_space = from_m3(inverse(to_world));  // camera.rs:25
  out->to_world_space = from_m3(to_world);
  return RTX_OK;
}
rtx_status rtx_camera_set_direction(rtx_camera* camera, const rtx_vec3* direction) {
  if (!camera || !direction) return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "null argument");
  // camera.rs:35-40 — the matrices are derived from the direction stored BEFORE this call.
  const M3 to_world = derive_to_world_space_mat(to_v3(camera->direction));
  camera->to_cam_space = from_m3(inverse(to_world));
  camera->to_world_space = from_m3(to_world);
  camera->direction = *direction;
  return RTX_OK;
}
rtx_status rtx_camera_to_cam_space(const rtx_camera* c, const rtx_vec3* v, rtx_vec3* out) {
  if (!c || !v || !out) return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "null argument");
  *out = from_v3(matvec(to_m3(c->to_cam_space), to_v3(*v) - to_v3(c->position)));  // camera.rs:51-53
  return RTX_OK;
}
rtx_status rtx_camera_to_world_space(const rtx_camera* c, const rtx_vec3* v, rtx_vec3* out) {
  if (!c || !v || !out) return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "null argument");
  *out = from_v3(matvec(to_m3(c->to_world_space), to_v3(*v)) + to_v3(c->position));  // camera.rs:55-57
  return RTX_OK;
}
rtx_status rtx_camera_rotate_to_world_space(const rtx_camera* c, const rtx_vec3* v, rtx_vec3* out) {
  if (!c || !v || !out) return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "null argument");
  *out = from_v3(matvec(to_m3(c->to_world_space), to_v3(*v)));  // camera.rs:65-67
  return RTX_OK;
}
rtx_status rtx_config_default(rtx_config* out) {
  if (!out) return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "null argument");
  out->rays_per_pixel = 16;  // scene.rs:55-65
  out->max_bounces = 10;
  out->focal_length = 10.0;
  out->focal_offset = 1e-4;
  out->non_focal_offset = 1e-1;
  return RTX_OK;
}

uint64_t rtx_stream_state(uint64_t seed, uint64_t key) { return stream_state(seed, key); }

// ---- scene lifetime ------------------------------------------------------------------------------
rtx_status rtx_scene_create(const rtx_camera* camera, const rtx_config* config, int32_t device, rtx_scene** out) {
  if (!camera || !config || !out) return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "null argument");
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    cudaGetLastError();
    return fail(nullptr, RTX_ERR_CUDA,
                "no CUDA device available: the rtx CUDA path has no CPU fallback (cudaGetDeviceCount failed or returned 0)");
  }
  if (device < 0 || device >= n) return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "device ordinal out of range");
  RTX_CUDA_TRY(nullptr, cudaSetDevice(device));
  rtx_scene* s = new (std::nothrow) rtx_scene();
  if (!s) return fail(nullptr, RTX_ERR_OUT_OF_MEMORY, "host allocation failed");
  s->device = device;
  s->camera = *camera;
  s->config = *config;
  cudaDeviceProp prop{};
  e = cudaGetDeviceProperties(&prop, device);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaEventCreate(&s->ev0);
  if (e == cudaSuccess) e = cudaEventCreate(&s->ev1);
  // 64 work queues of {next entry, CTAs out}: zero here, and every render kernel leaves its queue zero (queue_release)
  if (e == cudaSuccess) e = cudaMalloc((void**)&s->d_work_counter, 64 * 2 * sizeof(uint32_t));
  if (e == cudaSuccess) e = cudaMemset(s->d_work_counter, 0, 64 * 2 * sizeof(uint32_t));
  if (e == cudaSuccess) e = cudaMalloc((void**)&s->d_counters, 8 * sizeof(unsigned long long));
  if (e != cudaSuccess) {
    rtx_status st = cuda_fail(nullptr, e, "rtx_scene_create");
    rtx_scene_destroy(s);
    return st;
  }
  s->sm_count = prop.multiProcessorCount > 0 ? prop.multiProcessorCount : 148;
  *out = s;
  return RTX_OK;
}

rtx_status rtx_scene_destroy(rtx_scene* s) {
  if (!s) return RTX_OK;
  cudaSetDevice(s->device);
  if (s->stream) cudaStreamSynchronize(s->stream);
  free_lbvh(s);
  cudaFree(s->d_blob);
  cudaFree(s->d_tables);
  cudaFree(s->d_tail);
  cudaFree(s->d_tri_raw);
  cudaFree(s->d_work_counter);
  cudaFree(s->d_counters);
  cudaFree(s->d_rgb);
  cudaFree(s->d_ids);
  cudaFree(s->d_rgb8);
  cudaFree(s->d_pv);
  if (s->ev0) cudaEventDestroy(s->ev0);
  if (s->ev1) cudaEventDestroy(s->ev1);
  if (s->ev_last) cudaEventDestroy(s->ev_last);
  if (s->stream) cudaStreamDestroy(s->stream);
  if (s->copy_stream) cudaStreamDestroy(s->copy_stream);
  for (cudaEvent_t ev : s->ev_band)
    if (ev) cudaEventDestroy(ev);
  cudaGetLastError();
  delete s;
  return RTX_OK;
}

rtx_status rtx_scene_set_camera(rtx_scene* s, const rtx_camera* camera) {
  if (!s || !camera) return fail(s, RTX_ERR_INVALID_ARGUMENT, "null argument");
  std::lock_guard<std::mutex> lk(s->mu);
  s->camera = *camera;
  return RTX_OK;
}
rtx_status rtx_scene_set_config(rtx_scene* s, const rtx_config* config) {
  if (!s || !config) return fail(s, RTX_ERR_INVALID_ARGUMENT, "null argument");
  std::lock_guard<std::mutex> lk(s->mu);
  s->config = *config;
  return RTX_OK;
}
rtx_status rtx_scene_get_camera(const rtx_scene* s, rtx_camera* out) {
  if (!s || !out) return fail(s, RTX_ERR_INVALID_ARGUMENT, "null argument");
  *out = s->camera;
  return RTX_OK;
}
rtx_status rtx_scene_get_config(const rtx_scene* s, rtx_config* out) {
  if (!s || !out) return fail(s, RTX_ERR_INVALID_ARGUMENT, "null argument");
  *out = s->config;
  return RTX_OK;
}

// (called with s->mu held: n_obj is read here)
static rtx_status check_add(rtx_scene* s, const void* shapes, const void* materials, uint64_t n) {
  if (n && (!shapes || !materials)) return fail(s, RTX_ERR_INVALID_ARGUMENT, "null shapes/materials");
  if (s->n_obj + n >= (1ull << 30)) return fail(s, RTX_ERR_INVALID_ARGUMENT, "too many objects (limit 2^30)");
  return RTX_OK;
}
static void push_material(rtx_scene* s, const rtx_material& m) {
  const size_t o = s->mat.size();
  s->mat.resize(o + MAT_STRIDE);
  pack_material(&m.base_color.x, &m.emission_color.x, m.roughness, &s->mat[o]);
}

rtx_status rtx_scene_add_spheres(rtx_scene* s, const rtx_sphere* shapes, const rtx_material* materials, uint64_t n) {
  if (!s) return fail(s, RTX_ERR_INVALID_ARGUMENT, "null scene");
  std::lock_guard<std::mutex> lk(s->mu);
  rtx_status st = check_add(s, shapes, materials, n);
  if (st != RTX_OK) return st;
  s->sph.reserve(s->sph.size() + n * SPH_STRIDE);
  for (uint64_t i = 0; i < n; ++i) {
    const size_t o = s->sph.size();
    s->sph.resize(o + SPH_STRIDE);
    pack_sphere(&shapes[i].position.x, shapes[i].radius, &s->sph[o]);
    s->sph_radius.push_back(shapes[i].radius);
    s->obj_ref.push_back(REF_SPHERE | (uint32_t)s->sph_id.size());
    s->sph_id.push_back((uint32_t)s->n_obj++);
    push_material(s, materials[i]);
  }
  s->dirty = true;
  s->pv_valid = false;
  s->bvh_valid = false;
  return RTX_OK;
}
rtx_status rtx_scene_add_planes(rtx_scene* s, const rtx_plane* shapes, const rtx_material* materials, uint64_t n) {
  if (!s) return fail(s, RTX_ERR_INVALID_ARGUMENT, "null scene");
  std::lock_guard<std::mutex> lk(s->mu);
  rtx_status st = check_add(s, shapes, materials, n);
  if (st != RTX_OK) return st;
  for (uint64_t i = 0; i < n; ++i) {
    const size_t o = s->pl.size();
    s->pl.resize(o + PL_STRIDE);
    pack_plane(&shapes[i].position.x, &shapes[i].normal.x, &s->pl[o]);
    s->obj_ref.push_back(REF_PLANE | (uint32_t)s->pl_id.size());
    s->pl_id.push_back((uint32_t)s->n_obj++);
    push_material(s, materials[i]);
  }
  s->dirty = true;
  s->pv_valid = false;
  s->bvh_valid = false;
  return RTX_OK;
}
rtx_status rtx_scene_add_triangles(rtx_scene* s, const rtx_triangle* shapes, const rtx_material* materials, uint64_t n) {
  if (!s) return fail(s, RTX_ERR_INVALID_ARGUMENT, "null scene");
  std::lock_guard<std::mutex> lk(s->mu);
  rtx_status st = check_add(s, shapes, materials, n);
  if (st != RTX_OK) return st;
  s->tri.reserve(s->tri.size() + n * TRI_STRIDE);
  for (uint64_t i = 0; i < n; ++i) {
    const size_t o = s->tri.size();
    s->tri.resize(o + TRI_STRIDE);
    pack_triangle(&shapes[i].vertices[0].x, &s->tri[o]);
    s->tri_raw.insert(s->tri_raw.end(), &shapes[i].vertices[0].x, &shapes[i].vertices[0].x + 9);
    s->obj_ref.push_back(REF_TRIANGLE | (uint32_t)s->tri_id.size());
    s->tri_id.push_back((uint32_t)s->n_obj++);
    push_material(s, materials[i]);
  }
  s->dirty = true;
  s->pv_valid = false;
  s->bvh_valid = false;
  return RTX_OK;
}
// ---- in-place updates (N2) -------------------------------------------------------------------------
static rtx_status check_update(rtx_scene* s, const uint64_t* idx, const void* shapes, uint64_t n, uint32_t kind) {
  if (!s) return fail(s, RTX_ERR_INVALID_ARGUMENT, "null scene");
  if (n && (!idx || !shapes)) return fail(s, RTX_ERR_INVALID_ARGUMENT, "null indices/shapes");
  for (uint64_t i = 0; i < n; ++i) {
    if (idx[i] >= s->n_obj) return fail(s, RTX_ERR_INVALID_ARGUMENT, "object index out of range");
    if ((s->obj_ref[idx[i]] & ~REF_INDEX_MASK) != kind)
      return fail(s, RTX_ERR_INVALID_ARGUMENT, "object is of a different shape kind");
  }
  return RTX_OK;
}
static void update_material(rtx_scene* s, uint64_t obj, const rtx_material& m) {
  pack_material(&m.base_color.x, &m.emission_color.x, m.roughness, &s->mat[(size_t)obj * MAT_STRIDE]);
  s->upd_mat.push_back((uint32_t)obj);
}
rtx_status rtx_scene_update_spheres(rtx_scene* s, const uint64_t* idx, const rtx_sphere* shapes, const rtx_material* materials,
                                    uint64_t n) {
  if (s == nullptr) return fail(s, RTX_ERR_INVALID_ARGUMENT, "null scene");
  std::lock_guard<std::mutex> lk(s->mu);
  rtx_status st = check_update(s, idx, shapes, n, REF_SPHERE);
  if (st != RTX_OK) return st;
  for (uint64_t i = 0; i < n; ++i) {
    const uint32_t k = s->obj_ref[idx[i]] & REF_INDEX_MASK;
    pack_sphere(&shapes[i].position.x, shapes[i].radius, &s->sph[(size_t)k * SPH_STRIDE]);
    s->sph_radius[k] = shapes[i].radius;
    s->upd_sph.push_back(k);
    if (materials) update_material(s, idx[i], materials[i]);
  }
  if (n) s->bvh_refit = true;
  s->pv_valid = false;
  return RTX_OK;
}
rtx_status rtx_scene_update_planes(rtx_scene* s, const uint64_t* idx, const rtx_plane* shapes, const rtx_material* materials,
                                   uint64_t n) {
  if (s == nullptr) return fail(s, RTX_ERR_INVALID_ARGUMENT, "null scene");
  std::lock_guard<std::mutex> lk(s->mu);
  rtx_status st = check_update(s, idx, shapes, n, REF_PLANE);
  if (st != RTX_OK) return st;
  for (uint64_t i = 0; i < n; ++i) {  // planes are outside the trees: nothing to refit
    const uint32_t k = s->obj_ref[idx[i]] & REF_INDEX_MASK;
    pack_plane(&shapes[i].position.x, &shapes[i].normal.x, &s->pl[(size_t)k * PL_STRIDE]);
    s->upd_pl.push_back(k);
    if (materials) update_material(s, idx[i], materials[i]);
  }
  s->pv_valid = false;
  return RTX_OK;
}
rtx_status rtx_scene_update_triangles(rtx_scene* s, const uint64_t* idx, const rtx_triangle* shapes,
                                      const rtx_material* materials, uint64_t n) {
  if (s == nullptr) return fail(s, RTX_ERR_INVALID_ARGUMENT, "null scene");
  std::lock_guard<std::mutex> lk(s->mu);
  rtx_status st = check_update(s, idx, shapes, n, REF_TRIANGLE);
  if (st != RTX_OK) return st;
  for (uint64_t i = 0; i < n; ++i) {
    const uint32_t k = s->obj_ref[idx[i]] & REF_INDEX_MASK;
    double* rec = &s->tri[(size_t)k * TRI_STRIDE];
    uint64_t old_code, new_code;
    std::memcpy(&old_code, &rec[13], 8);
    pack_triangle(&shapes[i].vertices[0].x, rec);
    std::memcpy(&new_code, &rec[13], 8);
    std::memcpy(&s->tri_raw[(size_t)k * 9], &shapes[i].vertices[0].x, 9 * sizeof(double));
    s->upd_tri.push_back(k);
    // the pivots decide which prism tree holds the triangle, the flags whether any tree does
    if (old_code != new_code) s->bvh_valid = false;
    if (materials) update_material(s, idx[i], materials[i]);
  }
  if (n) s->bvh_refit = true;
  s->pv_valid = false;
  return RTX_OK;
}
rtx_status rtx_scene_rebuild_accel(rtx_scene* s) {
  if (!s) return fail(s, RTX_ERR_INVALID_ARGUMENT, "null scene");
  std::lock_guard<std::mutex> lk(s->mu);
  s->bvh_valid = false;
  return RTX_OK;
}

uint32_t rtx_scene_accel_refits(const rtx_scene* s) { return s ? s->bvh_refits : 0; }
uint32_t rtx_scene_accel_builds(const rtx_scene* s) { return s ? s->bvh_builds : 0; }
uint32_t rtx_scene_last_launches(const rtx_scene* s) { return s ? s->last_launches : 0; }

rtx_status rtx_scene_clear_objects(rtx_scene* s) {
  if (!s) return fail(s, RTX_ERR_INVALID_ARGUMENT, "null scene");
  std::lock_guard<std::mutex> lk(s->mu);
  s->sph.clear(), s->pl.clear(), s->tri.clear(), s->mat.clear(), s->tri_raw.clear(), s->sph_radius.clear();
  s->sph_id.clear(), s->pl_id.clear(), s->tri_id.clear(), s->obj_ref.clear();
  s->upd_sph.clear(), s->upd_pl.clear(), s->upd_tri.clear(), s->upd_mat.clear();
  s->n_obj = 0;
  s->dirty = true;
  s->pv_valid = false;
  s->bvh_valid = false;
  return RTX_OK;
}
uint64_t rtx_scene_object_count(const rtx_scene* s) { return s ? s->n_obj : 0; }

void rtx_shard_band_tiles(uint32_t width, uint32_t height, uint32_t index, uint32_t count, uint32_t tile_row_first,
                          uint32_t tile_row_count, uint32_t* out_first_local_tile, uint32_t* out_tile_count) {
  uint32_t first = 0, n = 0;
  shard_band_tiles(width, height, index, count, tile_row_first, tile_row_count, &first, &n);
  if (out_first_local_tile) *out_first_local_tile = first;
  if (out_tile_count) *out_tile_count = n;
}

uint64_t rtx_shard_tile_count(uint32_t width, uint32_t height, uint32_t index, uint32_t count) {
  return shard_tiles(width, height, index, count);
}
uint64_t rtx_shard_strip_count(uint32_t height, uint32_t index, uint32_t count) { return shard_strips(height, index, count); }
void rtx_shard_band_strips(uint32_t height, uint32_t index, uint32_t count, uint32_t tile_row_first, uint32_t tile_row_count,
                           uint32_t* out_first_local_strip, uint32_t* out_strip_count) {
  uint32_t first = 0, n = 0;
  shard_band_strips(height, index, count, tile_row_first, tile_row_count, &first, &n);
  if (out_first_local_strip) *out_first_local_strip = first;
  if (out_strip_count) *out_strip_count = n;
}

// ---- render --------------------------------------------------------------------------------------
rtx_status rtx_scene_render_device(rtx_scene* s, uint32_t width, uint32_t height, uint64_t seed, uint32_t flags,
                                   const rtx_shard* shard, void* d_out_rgb, void* d_out_primary_hit_id,
                                   void* d_out_counters, void* cuda_stream) {
  if (!s) return fail(s, RTX_ERR_INVALID_ARGUMENT, "null scene");
  if (!d_out_rgb && !(flags & RTX_FLAG_NO_RADIANCE) && !d_out_primary_hit_id)
    return fail(s, RTX_ERR_INVALID_ARGUMENT, "no output buffer");
  std::lock_guard<std::mutex> lk(s->mu);
  RTX_CUDA_TRY(s, cudaSetDevice(s->device));
  uint32_t* wc = s->d_work_counter + 2u * (s->wc_slot++ & 63u);  // under s->mu
  uint32_t launches = 0;
  const rtx_status st = render_common(s, width, height, seed, flags, shard, (double*)d_out_rgb, (uint32_t*)d_out_primary_hit_id,
                                      (unsigned long long*)d_out_counters, wc, (cudaStream_t)cuda_stream, &launches);
  s->last_launches = launches;
  return st;
}

rtx_status rtx_scene_render_device_band(rtx_scene* s, uint32_t width, uint32_t height, uint64_t seed, uint32_t flags,
                                        const rtx_shard* shard, uint32_t tile_row_first, uint32_t tile_row_count,
                                        void* d_out_rgb, void* d_out_primary_hit_id, void* d_out_counters,
                                        void* cuda_stream) {
  if (!s) return fail(s, RTX_ERR_INVALID_ARGUMENT, "null scene");
  if (!d_out_rgb && !(flags & RTX_FLAG_NO_RADIANCE) && !d_out_primary_hit_id)
    return fail(s, RTX_ERR_INVALID_ARGUMENT, "no output buffer");
  std::lock_guard<std::mutex> lk(s->mu);
  RTX_CUDA_TRY(s, cudaSetDevice(s->device));
  uint32_t* wc = s->d_work_counter + 2u * (s->wc_slot++ & 63u);  // under s->mu
  uint32_t launches = 0;
  const uint32_t rows[2] = {tile_row_first, tile_row_count};
  const rtx_status st = render_common(s, width, height, seed, flags, shard, (double*)d_out_rgb, (uint32_t*)d_out_primary_hit_id,
                                      (unsigned long long*)d_out_counters, wc, (cudaStream_t)cuda_stream, &launches, nullptr, rows);
  s->last_launches = launches;
  return st;
}

// Every GPU delivers ITS OWN pixels to their final place in one host frame, over its own PCIe link: the shard is cut
// by rows (RTX_FLAG_SHARD_ROWS), rendered band by band into the scene's device frame, and after each band the copy
// engine moves the band's strips — whole rows of the frame — to host memory with one 2-D copy (row pitch = count
// strips) while the next band renders.  No rank-0 funnel, no collective on the data path.
rtx_status rtx_scene_render_rows_to_host(rtx_scene* s, uint32_t width, uint32_t height, uint64_t seed, uint32_t flags,
                                         const rtx_shard* shard, double* host_rgb, uint32_t* host_ids, void* d_out_counters,
                                         void* cuda_stream) {
  if (!s) return fail(s, RTX_ERR_INVALID_ARGUMENT, "null scene");
  if (!host_rgb || (flags & RTX_FLAG_NO_RADIANCE)) return fail(s, RTX_ERR_INVALID_ARGUMENT, "host_rgb is required");
  if (width == 0 || height == 0 || (uint64_t)width * height > 0xFFFFFFFFull)
    return fail(s, RTX_ERR_INVALID_ARGUMENT, "bad image size");
  rtx_shard sh{0, 1};
  if (shard) sh = *shard;
  if (sh.count == 0 || sh.index >= sh.count) return fail(s, RTX_ERR_INVALID_ARGUMENT, "invalid shard");
  std::lock_guard<std::mutex> lk(s->mu);
  RTX_CUDA_TRY(s, cudaSetDevice(s->device));
  cudaStream_t stream = (cudaStream_t)cuda_stream;
  const size_t npix = (size_t)width * height;
  RTX_CUDA_TRY(s, grow(&s->d_rgb, &s->d_rgb_cap, npix * 3 * sizeof(double)));
  if (host_ids) RTX_CUDA_TRY(s, grow(&s->d_ids, &s->d_ids_cap, npix * sizeof(uint32_t)));
  if (!s->copy_stream) RTX_CUDA_TRY(s, cudaStreamCreateWithFlags(&s->copy_stream, cudaStreamNonBlocking));
  for (cudaEvent_t& ev : s->ev_band)
    if (!ev) RTX_CUDA_TRY(s, cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  const uint32_t tiles_y = (height + TILE - 1) / TILE;
  // bands of tile rows, the last one half as tall (its copy is the one nothing overlaps); small shards: fewer bands
  const uint64_t my_pixels = shard_strips(height, sh.index, sh.count) * GRAN_H * (uint64_t)width;
  uint32_t bands = my_pixels >= (1u << 19) ? 4u : (my_pixels >= (1u << 17) ? 3u : (my_pixels >= (1u << 15) ? 2u : 1u));
  bands = std::min(bands, tiles_y);
  const uint64_t strips = strip_count(height);
  const size_t strip_elems = (size_t)GRAN_H * width;  // pixels of a full strip
  uint32_t launches = 0;
  const uint32_t fl = (flags & ~(uint32_t)RTX_FLAG_SHARD_IN_PLACE) | RTX_FLAG_SHARD_ROWS;
  for (uint32_t b = 0; b < bands; ++b) {
    const uint64_t den = 2ull * bands - 1ull;
    const uint32_t r0 = (uint32_t)((uint64_t)tiles_y * (2ull * b) / den);
    const uint32_t r1 = b + 1 == bands ? tiles_y : (uint32_t)((uint64_t)tiles_y * (2ull * (b + 1)) / den);
    uint32_t first = 0, n = 0;
    shard_band_strips(height, sh.index, sh.count, r0, r1 - r0, &first, &n);
    if (n == 0) continue;
    const uint32_t rows[2] = {r0, r1 - r0};
    const rtx_status st = render_common(s, width, height, seed, fl, &sh, s->d_rgb, host_ids ? s->d_ids : nullptr,
                                        (unsigned long long*)d_out_counters, s->d_work_counter + 2u * (s->wc_slot++ & 63u), stream,
                                        &launches, nullptr, rows);
    if (st != RTX_OK) return st;
    RTX_CUDA_TRY(s, cudaEventRecord(s->ev_band[b], stream));
    RTX_CUDA_TRY(s, cudaStreamWaitEvent(s->copy_stream, s->ev_band[b], 0));
    // local strips [first, first + n) are the image's strips (first + j) * count + index: a 2-D copy whose "rows" are
    // strips and whose pitch is `count` strips; a ragged last strip of the image goes separately
    const uint64_t g0 = (uint64_t)first * sh.count + sh.index, g_last = (uint64_t)(first + n - 1) * sh.count + sh.index;
    const bool ragged = g_last == strips - 1 && (height % GRAN_H) != 0;
    const uint32_t full = ragged ? n - 1 : n;
    if (full) {
      RTX_CUDA_TRY(s, cudaMemcpy2DAsync(host_rgb + g0 * strip_elems * 3, (size_t)sh.count * strip_elems * 24,
                                        s->d_rgb + g0 * strip_elems * 3, (size_t)sh.count * strip_elems * 24, strip_elems * 24, full,
                                        cudaMemcpyDeviceToHost, s->copy_stream));
      if (host_ids)
        RTX_CUDA_TRY(s, cudaMemcpy2DAsync(host_ids + g0 * strip_elems, (size_t)sh.count * strip_elems * 4, s->d_ids + g0 * strip_elems,
                                          (size_t)sh.count * strip_elems * 4, strip_elems * 4, full, cudaMemcpyDeviceToHost,
                                          s->copy_stream));
    }
    if (ragged) {
      const size_t px0 = g_last * strip_elems, px = npix - px0;
      RTX_CUDA_TRY(s, cudaMemcpyAsync(host_rgb + px0 * 3, s->d_rgb + px0 * 3, px * 24, cudaMemcpyDeviceToHost, s->copy_stream));
      if (host_ids)
        RTX_CUDA_TRY(s, cudaMemcpyAsync(host_ids + px0, s->d_ids + px0, px * 4, cudaMemcpyDeviceToHost, s->copy_stream));
    }
  }
  // the caller's stream is done when the last copy is (and so is the scene's device frame, for the next render)
  RTX_CUDA_TRY(s, cudaEventRecord(s->ev_band[rtx_scene::MAX_BANDS], s->copy_stream));
  RTX_CUDA_TRY(s, cudaStreamWaitEvent(stream, s->ev_band[rtx_scene::MAX_BANDS], 0));
  {  // (not under graph capture: see render_common)
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(stream, &cap) != cudaSuccess) cudaGetLastError(), cap = cudaStreamCaptureStatusNone;
    if (cap == cudaStreamCaptureStatusNone && s->ev_last && cudaEventRecord(s->ev_last, stream) == cudaSuccess)
      s->last_stream = stream, s->last_valid = true;
  }
  s->last_launches = launches;
  return RTX_OK;
}

// ---- one pinned host frame shared by the processes of a node (POSIX shared memory + cudaHostRegister) -------------
static rtx_status host_frame_map(const char* name, uint64_t bytes, bool create, void** out) {
  if (!name || !out || bytes == 0) return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "bad host frame arguments");
  *out = nullptr;
  const int fd = shm_open(name, create ? (O_CREAT | O_EXCL | O_RDWR) : O_RDWR, 0600);
  if (fd < 0) return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, std::string("shm_open(") + name + ") failed: " + std::strerror(errno));
  if (create && ftruncate(fd, (off_t)bytes) != 0) {
    const std::string why = std::strerror(errno);
    close(fd);
    shm_unlink(name);
    return fail(nullptr, RTX_ERR_OUT_OF_MEMORY, "ftruncate of the host frame failed: " + why);
  }
  void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_POPULATE, fd, 0);
  close(fd);
  if (p == MAP_FAILED) {
    if (create) shm_unlink(name);
    return fail(nullptr, RTX_ERR_OUT_OF_MEMORY, std::string("mmap of the host frame failed: ") + std::strerror(errno));
  }
  const cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterPortable);
  if (e != cudaSuccess) {
    munmap(p, bytes);
    if (create) shm_unlink(name);
    return cuda_fail(nullptr, e, "cudaHostRegister of the host frame");
  }
  *out = p;
  return RTX_OK;
}
rtx_status rtx_host_frame_create(const char* name, uint64_t bytes, void** out) { return host_frame_map(name, bytes, true, out); }
rtx_status rtx_host_frame_open(const char* name, uint64_t bytes, void** out) { return host_frame_map(name, bytes, false, out); }
rtx_status rtx_host_frame_close(void* ptr, uint64_t bytes, const char* unlink_name) {
  if (ptr) {
    if (cudaHostUnregister(ptr) != cudaSuccess) cudaGetLastError();
    munmap(ptr, bytes);
  }
  if (unlink_name) shm_unlink(unlink_name);
  return RTX_OK;
}

rtx_status rtx_scene_render(rtx_scene* s, uint32_t width, uint32_t height, uint64_t seed, uint32_t flags,
                            double* out_rgb, uint32_t* out_primary_hit_id, rtx_stats* out_stats) {
  if (!s) return fail(s, RTX_ERR_INVALID_ARGUMENT, "null scene");
  const bool want_rgb = !(flags & RTX_FLAG_NO_RADIANCE);
  if (want_rgb && !out_rgb) return fail(s, RTX_ERR_INVALID_ARGUMENT, "out_rgb is null");
  if (!want_rgb && !out_primary_hit_id) return fail(s, RTX_ERR_INVALID_ARGUMENT, "no output buffer");
  const auto t0 = std::chrono::steady_clock::now();
  std::lock_guard<std::mutex> lk(s->mu);
  RTX_CUDA_TRY(s, cudaSetDevice(s->device));
  const size_t npix = (size_t)width * height;
  // A caller buffer in pinned, device-mapped host memory (rtx_host_alloc, cudaHostAlloc, cudaHostRegister) is
  // written by the render kernel itself: every finished pixel is a posted store over PCIe while the rest of the
  // frame is still rendering, so there is no device frame and no copy after the kernel (a measurement build can turn it
  // off).  Pageable buffers go through the scene's device frame + cudaMemcpyAsync.
  double* k_rgb = want_rgb ? mapped_device_pointer(out_rgb) : nullptr;
  uint32_t* k_ids = out_primary_hit_id ? mapped_device_pointer(out_primary_hit_id) : nullptr;
  const bool copy_rgb = want_rgb && !k_rgb, copy_ids = out_primary_hit_id && !k_ids;
  if (copy_rgb) {
    RTX_CUDA_TRY(s, grow(&s->d_rgb, &s->d_rgb_cap, npix * 3 * sizeof(double)));
    k_rgb = s->d_rgb;
  }
  if (copy_ids) {
    RTX_CUDA_TRY(s, grow(&s->d_ids, &s->d_ids_cap, npix * sizeof(uint32_t)));
    k_ids = s->d_ids;
  }
  RTX_CUDA_TRY(s, cudaMemsetAsync(s->d_counters, 0, 8 * sizeof(unsigned long long), s->stream));
  uint32_t launches = 0;
  const bool had_bvh = s->bvh_valid;
  RTX_CUDA_TRY(s, cudaEventRecord(s->ev0, s->stream));
  rtx_status st = render_common(s, width, height, seed, flags, nullptr, k_rgb, k_ids, s->d_counters, s->d_work_counter,
                                s->stream, &launches, (want_rgb && !copy_rgb) ? out_rgb : nullptr);
  if (st != RTX_OK) return st;
  RTX_CUDA_TRY(s, cudaEventRecord(s->ev1, s->stream));
  if (copy_rgb)
    RTX_CUDA_TRY(s, cudaMemcpyAsync(out_rgb, s->d_rgb, npix * 3 * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
  if (copy_ids)
    RTX_CUDA_TRY(s, cudaMemcpyAsync(out_primary_hit_id, s->d_ids, npix * sizeof(uint32_t), cudaMemcpyDeviceToHost,
                                    s->stream));
  unsigned long long counters[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  RTX_CUDA_TRY(s, cudaMemcpyAsync(counters, s->d_counters, sizeof(counters), cudaMemcpyDeviceToHost, s->stream));
  RTX_CUDA_TRY(s, cudaStreamSynchronize(s->stream));
  if (out_stats) {
    const uint32_t build_launches = (!had_bvh && s->bvh_valid) ? s->bvh_build_launches : 0u;
    std::memset(out_stats, 0, sizeof(*out_stats));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, s->ev0, s->ev1);
    out_stats->rays = counters[0];
    out_stats->paths = counters[1];
    out_stats->triangle_panics = counters[2];
    out_stats->hits = counters[3];
    out_stats->pixels = npix;
    out_stats->device_ms = ms;
    out_stats->kernel_launches = launches;
    out_stats->accel_used = choose_accel(s, flags);
    out_stats->nodes_visited = counters[4];
    out_stats->prims_tested = counters[5];
    if (out_stats->accel_used == RTX_ACCEL_LBVH) out_stats->kernel_launches += build_launches;
    out_stats->total_ms =
        std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
  }
  if (counters[2] != 0)
    return fail(s, RTX_ERR_REFERENCE_PANIC,
                std::to_string(counters[2]) + " ray/triangle tests hit an assert_eq! of the reference (triangle.rs:76-94): "
                "the reference's render would have panicked; the outputs hold the image with those hits rejected");
  return RTX_OK;
}

rtx_status rtx_deinterleave_tiles(uint32_t width, uint32_t height, uint32_t count, const void* d_gathered_rgb,
                                  void* d_out_rgb, const void* d_gathered_ids, void* d_out_ids, void* cuda_stream) {
  if (width == 0 || height == 0 || count == 0) return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "bad geometry");
  if ((d_gathered_rgb && !d_out_rgb) || (d_gathered_ids && !d_out_ids))
    return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "missing output buffer");
  const uint64_t slots = shard_tiles(width, height, 0, count);
  const uint64_t n = (uint64_t)width * height;
  const unsigned block = 256;
  const unsigned grid = (unsigned)((n + block - 1) / block);
  deinterleave_kernel<<<grid, block, 0, (cudaStream_t)cuda_stream>>>(
      width, height, (width + TILE - 1) / TILE, count, slots, (const double*)d_gathered_rgb, (double*)d_out_rgb,
      (const uint32_t*)d_gathered_ids, (uint32_t*)d_out_ids);
  RTX_CUDA_TRY(nullptr, cudaGetLastError());
  return RTX_OK;
}

rtx_status rtx_quantize_rgb8_device(const void* d_rgb, uint32_t width, uint32_t height, void* d_out_rgb8,
                                    void* cuda_stream) {
  if (!d_rgb || !d_out_rgb8 || width == 0 || height == 0)
    return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "bad argument");
  const uint64_t n = (uint64_t)width * height;
  const unsigned block = 256;
  quantize_kernel<<<(unsigned)((n + block - 1) / block), block, 0, (cudaStream_t)cuda_stream>>>(
      (const double*)d_rgb, width, height, (uint8_t*)d_out_rgb8);
  RTX_CUDA_TRY(nullptr, cudaGetLastError());
  return RTX_OK;
}

rtx_status rtx_scene_render_to_rgb8(rtx_scene* s, uint32_t width, uint32_t height, uint64_t seed, uint32_t flags,
                                    uint8_t* out_rgb8, rtx_stats* out_stats) {
  if (!s || !out_rgb8) return fail(s, RTX_ERR_INVALID_ARGUMENT, "null argument");
  if (flags & RTX_FLAG_NO_RADIANCE) return fail(s, RTX_ERR_INVALID_ARGUMENT, "rgb8 output needs radiance");
  const auto t0 = std::chrono::steady_clock::now();
  std::lock_guard<std::mutex> lk(s->mu);
  RTX_CUDA_TRY(s, cudaSetDevice(s->device));
  const size_t npix = (size_t)width * height;
  RTX_CUDA_TRY(s, grow(&s->d_rgb, &s->d_rgb_cap, npix * 3 * sizeof(double)));
  RTX_CUDA_TRY(s, grow(&s->d_rgb8, &s->d_rgb8_cap, npix * 3));
  RTX_CUDA_TRY(s, cudaMemsetAsync(s->d_counters, 0, 8 * sizeof(unsigned long long), s->stream));
  uint32_t launches = 0;
  const bool had_bvh = s->bvh_valid;
  RTX_CUDA_TRY(s, cudaEventRecord(s->ev0, s->stream));
  rtx_status st = render_common(s, width, height, seed, flags, nullptr, s->d_rgb, nullptr, s->d_counters,
                                s->d_work_counter, s->stream, &launches);
  if (st != RTX_OK) return st;
  st = rtx_quantize_rgb8_device(s->d_rgb, width, height, s->d_rgb8, s->stream);
  if (st != RTX_OK) return st;
  ++launches;
  RTX_CUDA_TRY(s, cudaEventRecord(s->ev1, s->stream));
  RTX_CUDA_TRY(s, cudaMemcpyAsync(out_rgb8, s->d_rgb8, npix * 3, cudaMemcpyDeviceToHost, s->stream));
  unsigned long long counters[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  RTX_CUDA_TRY(s, cudaMemcpyAsync(counters, s->d_counters, sizeof(counters), cudaMemcpyDeviceToHost, s->stream));
  RTX_CUDA_TRY(s, cudaStreamSynchronize(s->stream));
  if (out_stats) {
    const uint32_t build_launches = (!had_bvh && s->bvh_valid) ? s->bvh_build_launches : 0u;
    std::memset(out_stats, 0, sizeof(*out_stats));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, s->ev0, s->ev1);
    out_stats->rays = counters[0];
    out_stats->paths = counters[1];
    out_stats->triangle_panics = counters[2];
    out_stats->hits = counters[3];
    out_stats->pixels = npix;
    out_stats->device_ms = ms;
    out_stats->kernel_launches = launches;
    out_stats->accel_used = choose_accel(s, flags);
    out_stats->nodes_visited = counters[4];
    out_stats->prims_tested = counters[5];
    if (out_stats->accel_used == RTX_ACCEL_LBVH) out_stats->kernel_launches += build_launches;
    out_stats->total_ms =
        std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
  }
  if (counters[2] != 0)
    return fail(s, RTX_ERR_REFERENCE_PANIC,
                std::to_string(counters[2]) + " ray/triangle tests hit an assert_eq! of the reference (triangle.rs:76-94): "
                "the reference's render would have panicked; the outputs hold the image with those hits rejected");
  return RTX_OK;
}

}  // extern "C"
