// rtx_render.cuh — launch parameters shared between the API layer and the render kernels.
#pragma once
#include "rtx_lbvh.cuh"
#include "rtx_scene.cuh"

namespace rtx {

// Work granule pulled from the atomic queue: an 8×4-pixel block inside an RTX_TILE×RTX_TILE tile.
// Tiles (16×16) are the multi-GPU ownership unit; granules keep the queue fine enough for a short tail.
constexpr uint32_t TILE = 16;
constexpr uint32_t GRAN_W = 8, GRAN_H = 4;
constexpr uint32_t GRANS_PER_TILE = (TILE / GRAN_W) * (TILE / GRAN_H);  // 8
constexpr uint32_t GRAN_PIXELS = GRAN_W * GRAN_H;                       // 32

struct RenderParams {
  // camera (camera.rs:7-15) and config (scene.rs:16-28)
  double cam_pos[3];
  double to_world[9];  // rows of Camera.to_world_space
  double focal_length, focal_offset, non_focal_offset;
  uint64_t rays_per_pixel;
  uint64_t max_bounces;
  uint64_t seed;
  uint32_t width, height;
  // get_ray_dir (scene.rs:213-222) is separable: sin/cos of the column angle and of the row angle,
  // computed on the host with the same libm the reference uses.
  const double* sin_x;  // [width]
  const double* cos_x;  // [width]
  const double* sin_y;  // [height]
  const double* cos_y;  // [height]
  // scene
  const unsigned char* scene_blob;
  SceneLayout layout;
  BvhView bvh;       // used when use_bvh != 0
  uint32_t use_bvh;
  uint32_t lane_kernel;  // RTX_FLAG_LANE_KERNEL: never pick the context-pool kernel
  uint32_t queue_mode;   // 0 auto, 1 RTX_FLAG_PIXEL_QUEUE, 2 RTX_FLAG_SAMPLE_QUEUE
  // image shard
  uint32_t shard_index, shard_count;
  uint32_t tiles_x, tiles_y;
  uint32_t my_tiles;    // ownership units (tiles, or strips) of this shard this launch renders ...
  uint32_t tile_first;  // ... starting at this one (a band of the shard's unit list; 0 = from the start)
  uint32_t compact;     // 1: outputs use the compact per-shard tile layout
  // Ownership unit.  0: RTX_TILE x RTX_TILE tiles, row-major, 8 granules each (the default: the best load balance).
  // 1 (RTX_FLAG_SHARD_ROWS): full-width STRIPS of GRAN_H = 4 pixel rows, strip s -> shard s mod count — a shard's
  // pixels are whole contiguous rows of the frame, which a copy engine moves to host memory with one 2-D copy per band.
  uint32_t strip_mode;
  uint32_t grans_per_unit;  // GRANS_PER_TILE, or ceil(width / GRAN_W) in strip mode
  // The atomic work queue: [0] next entry, [1] CTAs that have left the queue.  Both are zero when a launch starts and
  // the launch leaves them zero (queue_release: the last CTA out resets them) — no memset between frames, which in
  // the launch-bound regime (C5 at 256x256) was one of the three stream operations of a frame.
  uint32_t* work_counter;
  // outputs
  double* out_rgb;                  // may be null (RTX_FLAG_NO_RADIANCE)
  uint32_t* out_ids;                // may be null
  uint32_t out_is_host;             // out_rgb is pinned host memory written over PCIe (queue plan, rtx_render_pool.cu)
  // Sample queue (rtx_render_pool.cu, pool_plan_tail): when tail_granules != 0 (then tail_first == 0 and
  // tail_granules == all granules of the shard) the work queue hands out (granule, sample) pairs — one camera path
  // per context — and per-sample radiance goes to tail_scratch[((g * rpp + s) * 32 + k) * 3 ..], which
  // tail_sum_kernel adds up per pixel in sample order (the reference's left fold).
  uint32_t tail_first, tail_granules;
  double* tail_scratch;
  unsigned long long* ray_counter;  // [0] rays, [1] paths, [2] triangle panics, [3] hits, [4] nodes visited, [5] prims tested; may be null
};

#if defined(__CUDACC__)
// ---- work-queue granule → pixels ---------------------------------------------------------------
struct Granule {
  uint32_t x0, y0, w, npix, base_out, tx0, ty0;
};
// gidx = index of the granule within this shard's unit list (local unit = gidx / grans_per_unit).
__device__ __forceinline__ Granule decode_granule(const RenderParams& p, uint32_t gidx) {
  Granule g;
  if (p.strip_mode) {  // (warp-uniform)
    const uint32_t lstrip = p.tile_first + gidx / p.grans_per_unit, col = gidx % p.grans_per_unit;
    g.tx0 = g.ty0 = g.base_out = 0;
    g.x0 = col * GRAN_W;
    g.y0 = (lstrip * p.shard_count + p.shard_index) * GRAN_H;
    const uint32_t gw = g.x0 < p.width ? min(GRAN_W, p.width - g.x0) : 0u;
    const uint32_t gh = g.y0 < p.height ? min(GRAN_H, p.height - g.y0) : 0u;
    g.w = gw;
    g.npix = gw * gh;
    return g;
  }
  const uint32_t ltile = p.tile_first + gidx / GRANS_PER_TILE, sub = gidx % GRANS_PER_TILE;
  const uint32_t gtile = ltile * p.shard_count + p.shard_index;
  const uint32_t ty = gtile / p.tiles_x, tx = gtile % p.tiles_x;
  g.tx0 = tx * TILE;
  g.ty0 = ty * TILE;
  g.x0 = g.tx0 + (sub % (TILE / GRAN_W)) * GRAN_W;
  g.y0 = g.ty0 + (sub / (TILE / GRAN_W)) * GRAN_H;
  const uint32_t gw = g.x0 < p.width ? min(GRAN_W, p.width - g.x0) : 0u;
  const uint32_t gh = g.y0 < p.height ? min(GRAN_H, p.height - g.y0) : 0u;
  g.w = gw;
  g.npix = gw * gh;
  g.base_out = ltile * (TILE * TILE);
  return g;
}
// The same out of line, for the context-pool kernel: once per queue entry, two runtime divisions — not worth a place
// in a loop that is larger than the instruction cache (the lane kernel and the per-thread kernels keep it inline:
// primary-only frames decode a granule per ~150 warp instructions, and a call there costs 4 %).
static __device__ __noinline__ Granule decode_granule_cold(const RenderParams& p, uint32_t gidx) { return decode_granule(p, gidx); }
static __device__ __noinline__ uint32_t ragged_div(uint32_t k, uint32_t w) { return k / w; }
// pixel k (row-major inside the granule) → image pixel index and output index
__device__ __forceinline__ void granule_pixel(const RenderParams& p, const Granule& g, uint32_t k, uint32_t* x, uint32_t* y,
                                              uint32_t* out_index) {
  if (g.w == GRAN_W) {  // (warp-uniform) a full-width granule: no division
    *x = g.x0 + (k & (GRAN_W - 1u));
    *y = g.y0 + k / GRAN_W;
  } else {  // the image's ragged right edge
    const uint32_t q = ragged_div(k, g.w);
    *x = g.x0 + (k - q * g.w);
    *y = g.y0 + q;
  }
  *out_index = p.compact ? g.base_out + (*y - g.ty0) * TILE + (*x - g.tx0) : *y * p.width + *x;
}

// Last thing a render kernel does (every thread of the CTA must get here): the last CTA out zeroes the queue.
__device__ __forceinline__ void queue_release(const RenderParams& p) {
  __syncthreads();  // every warp of this CTA has seen the queue run dry: no more atomics on [0] from here
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(p.work_counter + 1, 1u) + 1u == gridDim.x) {
      p.work_counter[0] = 0u;
      p.work_counter[1] = 0u;
    }
  }
}

// ---- TMA bulk copy (global → shared) with an mbarrier ------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void stage_blob(unsigned char* dst, const unsigned char* src, uint32_t bytes,
                                           uint64_t* mbar) {
  const uint32_t bar = smem_u32(mbar);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(dst)),
        "l"(src), "r"(bytes), "r"(bar)
        : "memory");
  }
  // every thread waits for phase 0 of the barrier
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], 0;\n"
      "@P1 bra DONE;\n"
      "bra WAIT_LOOP;\n"
      "DONE:\n"
      "}\n" ::"r"(bar)
      : "memory");
}

#endif  // __CUDACC__

}  // namespace rtx
