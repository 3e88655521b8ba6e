// rtx_render.cu — the render mega-kernel: camera-ray generation → nearest hit → shade/bounce →
// per-pixel accumulation, fused, for sm_100a.
//
// Replaces the per-pixel closure of Scene::render (raytracing/scene.rs:144-170) and everything it
// calls: render_pixel :194-212, get_ray_dir :213-222, render_ray :223-242, closest_object :243-251,
// avg :253-259, ray_hit :260-278, random_bounce_dir :279-291.
//
// Shape of the kernel:
//  * persistent CTAs (a multiple of the SM count); each WARP pulls 8×4-pixel granules from one global
//    atomic counter, and each LANE pulls the next pixel of its warp's granule with a ballot/popc rank —
//    so a lane whose pixel is finished never waits for its neighbours (path lengths differ per sample);
//  * one loop iteration = one nearest-hit query per lane (a "ray" of the Mrays/s metric); lanes whose
//    path ended regenerate a camera sample in the same iteration instead of idling;
//  * the scene blob is staged once per CTA into shared memory with a TMA bulk copy
//    (cp.async.bulk + mbarrier) when it fits, and read with 128-bit broadcast loads;
//  * a pixel's samples are accumulated sequentially by the lane that owns it, in sample order, so
//    the f64 sum is the reference's left fold (iter_ops.rs:4-8) bit for bit.
#include <cuda_runtime.h>
#include <math_constants.h>

#include "rtx_internal.h"
#include "rtx_lbvh.cuh"
#include "rtx_render.cuh"
#include "rtx_trace.cuh"

namespace rtx {

constexpr int RENDER_BLOCK = 256;
#ifndef RTX_BVH_BLOCKS_PER_SM
#define RTX_BVH_BLOCKS_PER_SM 3
#endif
// LBVH walks wait on L2 most of the time (ncu: long_scoreboard is the top stall), so the LBVH instantiations
// trade registers for resident warps.
constexpr int BVH_BLOCKS_PER_SM = RTX_BVH_BLOCKS_PER_SM;
constexpr uint32_t FULL = 0xFFFFFFFFu;

// K2 for scanned scenes whose blob does not fit the shared-memory stage (> 64 KB): the sphere arrays — the bulk of such
// a scene — go through shared memory CHUNK by chunk.  One CTA = one chunk pipeline: thread 0 issues three TMA bulk copies
// per chunk (f32 cull records, f64 records, object ids) into one of two buffers while every thread scans the other for
// its own ray; an mbarrier per buffer, a __syncthreads() per chunk.  That makes the nearest-hit query a CTA-wide step
// (every thread takes part in the barriers, with or without a ray), so the CHUNKED instantiation of the kernel leaves
// its loop on a CTA-wide vote.  Planes, triangles and materials stay where they are (L1/L2).
constexpr uint32_t CHUNK_SPHERES = 1024;
constexpr uint32_t CHUNK_BYTES = CHUNK_SPHERES * (SPH_STRIDE * 8 + SPH_STRIDE * 4 + 4);  // 52 KB per buffer
struct ChunkPipe {
  unsigned char* buf;  // [2][CHUNK_BYTES]: f64 records | f32 records | ids
  uint64_t* mbar;      // [2]
  uint32_t uses;       // chunks loaded so far (the same number in every thread)
};
__device__ __forceinline__ void chunk_issue(const ChunkPipe& cp, const SceneView& sc, uint32_t c, uint32_t use) {
  if (threadIdx.x != 0) return;
  const uint32_t b = use & 1u, first = c * CHUNK_SPHERES, n = min(CHUNK_SPHERES, sc.n_sph - first);
  unsigned char* dst = cp.buf + (size_t)b * CHUNK_BYTES;
  const uint32_t bytes64 = n * SPH_STRIDE * 8u, bytes32 = n * SPH_STRIDE * 4u, bytes_id = (n * 4u + 15u) & ~15u;
  const uint32_t bar = smem_u32(cp.mbar + b);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // the buffer was read through the generic proxy
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes64 + bytes32 + bytes_id) : "memory");
  auto copy = [&](unsigned char* d, const void* s, uint32_t bytes) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(d)),
                 "l"(s), "r"(bytes), "r"(bar)
                 : "memory");
  };
  copy(dst, sc.sph + (size_t)first * SPH_STRIDE, bytes64);
  copy(dst + CHUNK_SPHERES * SPH_STRIDE * 8u, sc.sph32 + (size_t)first * SPH_STRIDE, bytes32);
  copy(dst + CHUNK_SPHERES * SPH_STRIDE * 12u, sc.sph_id + first, bytes_id);
}
__device__ __forceinline__ void chunk_wait(const ChunkPipe& cp, uint32_t use) {
  const uint32_t bar = smem_u32(cp.mbar + (use & 1u)), parity = (use >> 1) & 1u;
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "CHUNK_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra CHUNK_DONE;\n"
      "bra CHUNK_WAIT;\n"
      "CHUNK_DONE:\n"
      "}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}
// scene.rs:243-251 for every thread of the CTA at once (active = this thread has a ray).  Same Hit as closest_brute.
template <bool HAS_TRI>
__device__ __forceinline__ Hit closest_chunked(ChunkPipe& cp, const SceneView& sc, bool active, V3 pos, V3 dir, uint32_t* panics) {
  Hit h;
  h.t = CUDART_INF;
  h.id = MISS_ID;
  h.ref = 0;
  const RayInv ri = ray_invariants(dir);
  const SphereCull cull = sphere_cull_setup(sc, pos, ri);
  const uint32_t n_chunks = (sc.n_sph + CHUNK_SPHERES - 1u) / CHUNK_SPHERES;
  if (n_chunks) chunk_issue(cp, sc, 0u, cp.uses);
  for (uint32_t c = 0; c < n_chunks; ++c) {
    if (c + 1u < n_chunks) chunk_issue(cp, sc, c + 1u, cp.uses + 1u);  // its buffer was released by the barrier of chunk c - 1
    chunk_wait(cp, cp.uses);
    if (active) {
      const unsigned char* b = cp.buf + (size_t)(cp.uses & 1u) * CHUNK_BYTES;
      const uint32_t first = c * CHUNK_SPHERES;
      scan_spheres(reinterpret_cast<const float*>(b + CHUNK_SPHERES * SPH_STRIDE * 8u), reinterpret_cast<const double*>(b),
                   reinterpret_cast<const uint32_t*>(b + CHUNK_SPHERES * SPH_STRIDE * 12u), first,
                   min(CHUNK_SPHERES, sc.n_sph - first), pos, ri, cull, h);
    }
    __syncthreads();
    ++cp.uses;
  }
  if (active) scan_planes_triangles<HAS_TRI>(sc, pos, dir, ri, h, panics);
  return h;
}

template <bool SMEM_SCENE, bool USE_BVH, bool HAS_TRI = true, bool CHUNKED = false>
__global__ void __launch_bounds__(RENDER_BLOCK, USE_BVH ? BVH_BLOCKS_PER_SM : 2) render_megakernel(const __grid_constant__ RenderParams p) {
  static_assert(!CHUNKED || (!SMEM_SCENE && !USE_BVH), "chunked staging is for scanned scenes too large to stage whole");
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const unsigned char* base = p.scene_blob;
  ChunkPipe cp;
  cp.buf = smem_raw + 128, cp.mbar = reinterpret_cast<uint64_t*>(smem_raw), cp.uses = 0;
  if (CHUNKED) {
    if (threadIdx.x == 0) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(cp.mbar)));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(cp.mbar + 1)));
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
  }
  if (SMEM_SCENE) {
    uint64_t* mbar = reinterpret_cast<uint64_t*>(smem_raw);
    unsigned char* dst = smem_raw + 128;
    stage_blob(dst, p.scene_blob, p.layout.bytes, mbar);
    base = dst;
  }
  const SceneView sc = make_view(base, p.layout);

  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t lt_mask = (1u << lane) - 1u;

  const V3 cam_pos = mk3(p.cam_pos[0], p.cam_pos[1], p.cam_pos[2]);
  M3 to_world;
  to_world.x = mk3(p.to_world[0], p.to_world[1], p.to_world[2]);
  to_world.y = mk3(p.to_world[3], p.to_world[4], p.to_world[5]);
  to_world.z = mk3(p.to_world[6], p.to_world[7], p.to_world[8]);
  const uint64_t rpp = p.rays_per_pixel;
  const uint64_t max_segments = p.max_bounces + 1ull;  // scene.rs:227
  const uint32_t total_work = p.my_tiles * p.grans_per_unit;

  // warp-uniform queue state
  uint32_t g_x0 = 0, g_y0 = 0, g_w = 0, g_npix = 0, g_next = 0, g_base_out = 0, g_tx0 = 0, g_ty0 = 0;
  bool exhausted = false;

  // per-lane pixel state
  bool have_pixel = false;
  uint32_t out_index = 0;
  uint64_t key_base = 0, sample = 0;
  V3 pix_dir = mk3(0, 0, 0), focal = mk3(0, 0, 0), acc = mk3(0, 0, 0);
  // per-lane path state (ray.rs:4-21)
  bool alive = false;
  V3 pos = mk3(0, 0, 0), dir = mk3(0, 0, 0), light = mk3(0, 0, 0), res = mk3(0, 0, 0);
  uint64_t segment = 0;
  WyRand rng;
  rng.s = 0;
  // counters
  uint32_t n_rays = 0, n_paths = 0, n_panics = 0, n_hits = 0;
  TraversalCounters tc;
  tc.nodes = 0, tc.prims = 0;

  for (;;) {
    // ---- 1. pixel acquisition: lanes without a pixel take the next slots of the warp's granule ----
    uint32_t need = __ballot_sync(FULL, !have_pixel);
    while (need != 0u && !exhausted) {
      if (g_next >= g_npix) {
        uint32_t w = 0;
        if (lane == 0) w = atomicAdd(p.work_counter, 1u);
        w = __shfl_sync(FULL, w, 0);
        if (w >= total_work) {
          exhausted = true;
          break;
        }
        const Granule g = decode_granule(p, w);
        g_tx0 = g.tx0, g_ty0 = g.ty0, g_x0 = g.x0, g_y0 = g.y0, g_w = g.w, g_npix = g.npix, g_base_out = g.base_out;
        g_next = 0;
        continue;  // re-check (granule may be empty at the image border)
      }
      const uint32_t slot = g_next + __popc(need & lt_mask);
      if (!have_pixel && slot < g_npix) {
        const uint32_t x = g_x0 + slot % g_w, y = g_y0 + slot / g_w;
        have_pixel = true;
        const uint32_t pixel_index = y * p.width + x;
        out_index = p.compact ? g_base_out + (y - g_ty0) * TILE + (x - g_tx0) : pixel_index;
        key_base = (uint64_t)pixel_index * rpp;
        sample = 0;
        acc = mk3(0.0, 0.0, 0.0);  // iter_ops.rs:4-8: fold starts from zeros()
        // scene.rs:213-222 with the libm calls hoisted to the host tables
        const V3 cam_dir = mk3(p.sin_x[x], p.sin_y[y], fmul(p.cos_x[x], p.cos_y[y]));
        pix_dir = matvec(to_world, cam_dir);                  // camera.rs:65-67
        focal = cam_pos + pix_dir * p.focal_length;            // scene.rs:203
        if (p.out_ids) p.out_ids[out_index] = MISS_ID;
        if (rpp == 0) {  // avg of an empty iterator: zeros / 0.0 = NaN (scene.rs:253-259)
          if (p.out_rgb) {
            const V3 px = acc / (double)rpp;
            p.out_rgb[(size_t)out_index * 3 + 0] = px.x;
            p.out_rgb[(size_t)out_index * 3 + 1] = px.y;
            p.out_rgb[(size_t)out_index * 3 + 2] = px.z;
          }
          have_pixel = false;
        }
      }
      g_next = min(g_next + (uint32_t)__popc(need), g_npix);
      need = __ballot_sync(FULL, !have_pixel);
    }
    if (CHUNKED) {
      if (!__syncthreads_or(have_pixel ? 1 : 0)) break;  // the nearest-hit query below is a CTA-wide step
    } else if (!__any_sync(FULL, have_pixel)) {
      break;  // queue exhausted and every lane is done
    }

    // ---- 2. start the next camera sample (scene.rs:199-210) ----
    if (have_pixel && !alive) {
      rng.s = stream_state(p.seed, key_base + sample);
      const double u0 = wy_f64(rng), u1 = wy_f64(rng), u2 = wy_f64(rng);
      pos = cam_pos + mk3(u0, u1, u2) * p.non_focal_offset;
      const double u3 = wy_f64(rng), u4 = wy_f64(rng), u5 = wy_f64(rng);
      const V3 target = focal + mk3(u3, u4, u5) * p.focal_offset;
      dir = vnorm(target - pos);
      light = mk3(1.0, 1.0, 1.0);
      res = mk3(0.0, 0.0, 0.0);
      segment = 0;
      alive = true;
      ++n_paths;
    }

    // ---- 3. one segment of render_ray (scene.rs:227-240) ----
    Hit hc;
    hc.t = CUDART_INF, hc.id = MISS_ID, hc.ref = 0;
    if (CHUNKED) hc = closest_chunked<HAS_TRI>(cp, sc, alive && sc.n_obj != 0, pos, dir, &n_panics);
    if (alive) {
      bool end = false;
      if (sc.n_obj == 0) {
        end = true;  // scene.rs:224-226
      } else {
        const Hit h = CHUNKED ? hc
                              : (USE_BVH ? closest_lbvh<HAS_TRI>(sc, p.bvh, pos, dir, &n_panics, tc)
                                         : closest_brute<HAS_TRI>(sc, pos, dir, &n_panics));
        ++n_rays;
        if (segment == 0 && sample == 0 && p.out_ids) p.out_ids[out_index] = h.id;
        if (h.id == MISS_ID) {
          end = true;
        } else {
          pos = pos + dir * h.t;  // scene.rs:234
          ++n_hits;
          const double* m = sc.mat + (size_t)h.id * MAT_STRIDE;
          const V3 base_color = mk3(m[0], m[1], m[2]);
          const V3 emission = mk3(m[3], m[4], m[5]);
          const V3 new_light = light * base_color;
          ++segment;
          // The bounce direction (scene.rs:275) is only consumed by a later segment.  When this was the
          // last allowed segment (:227) or light_color is now zero (:228 breaks before the next query)
          // nothing reads it, and the stream is private to this (pixel, sample): skip it.
          const bool last = (segment == max_segments) ||
                            (new_light.x == 0.0 && new_light.y == 0.0 && new_light.z == 0.0);
          if (!last) {
            const V3 n = normal_at(sc, h.ref, pos);
            dir = random_bounce_dir(rng, dir, n, m[6]);
          }
          res = res + light * emission;  // scene.rs:276
          light = new_light;             // scene.rs:277
          end = last;
        }
      }
      if (end) {
        acc = acc + res;  // Sum: a + b, left to right
        alive = false;
        ++sample;
        if (sample == rpp) {
          if (p.out_rgb) {
            const V3 px = acc / (double)rpp;  // avg: `/ length` (scene.rs:258)
            p.out_rgb[(size_t)out_index * 3 + 0] = px.x;
            p.out_rgb[(size_t)out_index * 3 + 1] = px.y;
            p.out_rgb[(size_t)out_index * 3 + 2] = px.z;
          }
          have_pixel = false;
        }
      }
    }
  }

  // ---- counters: one atomic per warp ----
  if (p.ray_counter) {
    unsigned long long r = n_rays, q = n_paths, e = n_panics, hh = n_hits;
    unsigned long long nv = tc.nodes, pt = USE_BVH ? (unsigned long long)tc.prims : (unsigned long long)n_rays * sc.n_obj;
    for (int o = 16; o > 0; o >>= 1) {
      r += __shfl_down_sync(FULL, r, o);
      q += __shfl_down_sync(FULL, q, o);
      e += __shfl_down_sync(FULL, e, o);
      hh += __shfl_down_sync(FULL, hh, o);
      nv += __shfl_down_sync(FULL, nv, o);
      pt += __shfl_down_sync(FULL, pt, o);
    }
    if (lane == 0) {
      atomicAdd(&p.ray_counter[0], r);
      atomicAdd(&p.ray_counter[1], q);
      if (e) atomicAdd(&p.ray_counter[2], e);
      atomicAdd(&p.ray_counter[3], hh);
      if (nv) atomicAdd(&p.ray_counter[4], nv);
      atomicAdd(&p.ray_counter[5], pt);
    }
  }
  queue_release(p);
}

// RTX_FLAG_RNG_ROW: the reference's own consumption of its random source, literally.  In Scene::render every row is
// one thread (scene.rs:149-151) and every draw of that thread comes from its one thread-local generator, pixel after
// pixel, sample after sample, bounce after bounce (math/vector.rs:29-45) — including the bounce directions whose ray is
// never traced (ray_hit always draws, scene.rs:275; the other kernels skip those because their streams are private to
// a camera path).  Here one CUDA thread is one row thread with one WyRand stream, state0 = stream_state(seed ^
// "ROW_STRE", y) (the reference's seed is the clock).  It exists to show that the per-(pixel, sample) streams of the
// production kernels differ from the literal reference in the SEEDING only: the oracle has the same mode
// (oracle_render, RNG_ROW) and the two agree bit for bit.  A frame is `height` threads: a demonstration, not a fast path.
constexpr uint64_t ROW_STREAM_TAG = 0x524F575F53545245ULL;
template <bool USE_BVH>
__global__ void __launch_bounds__(32) render_rows_kernel(const __grid_constant__ RenderParams p) {
  const uint32_t y = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t n_rays = 0, n_paths = 0, n_panics = 0, n_hits = 0;
  TraversalCounters tc;
  tc.nodes = 0, tc.prims = 0;
  const SceneView sc = make_view(p.scene_blob, p.layout);
  if (y < p.height) {
    const V3 cam_pos = mk3(p.cam_pos[0], p.cam_pos[1], p.cam_pos[2]);
    M3 to_world;
    to_world.x = mk3(p.to_world[0], p.to_world[1], p.to_world[2]);
    to_world.y = mk3(p.to_world[3], p.to_world[4], p.to_world[5]);
    to_world.z = mk3(p.to_world[6], p.to_world[7], p.to_world[8]);
    WyRand rng;
    rng.s = stream_state(p.seed ^ ROW_STREAM_TAG, y);
    for (uint32_t x = 0; x < p.width; ++x) {
      const V3 cam_dir = mk3(p.sin_x[x], p.sin_y[y], fmul(p.cos_x[x], p.cos_y[y]));  // scene.rs:213-222
      const V3 pix_dir = matvec(to_world, cam_dir);
      const size_t out_index = (size_t)y * p.width + x;
      V3 acc = mk3(0.0, 0.0, 0.0);
      uint32_t first = MISS_ID;
      for (uint64_t s = 0; s < p.rays_per_pixel; ++s) {  // scene.rs:199-210
        const double u0 = wy_f64(rng), u1 = wy_f64(rng), u2 = wy_f64(rng);
        V3 pos = cam_pos + mk3(u0, u1, u2) * p.non_focal_offset;
        const V3 focal = cam_pos + pix_dir * p.focal_length;
        const double u3 = wy_f64(rng), u4 = wy_f64(rng), u5 = wy_f64(rng);
        V3 dir = vnorm((focal + mk3(u3, u4, u5) * p.focal_offset) - pos);
        V3 light = mk3(1.0, 1.0, 1.0), res = mk3(0.0, 0.0, 0.0);
        ++n_paths;
        if (sc.n_obj != 0) {  // scene.rs:223-242
          for (uint64_t seg = 0; seg < p.max_bounces + 1ull; ++seg) {
            if (light.x == 0.0 && light.y == 0.0 && light.z == 0.0) break;
            const Hit h = USE_BVH ? closest_lbvh(sc, p.bvh, pos, dir, &n_panics, tc) : closest_brute(sc, pos, dir, &n_panics);
            ++n_rays;
            if (seg == 0 && s == 0) first = h.id;
            if (h.id == MISS_ID) break;
            ++n_hits;
            pos = pos + dir * h.t;
            const double* m = sc.mat + (size_t)h.id * MAT_STRIDE;
            dir = random_bounce_dir(rng, dir, normal_at(sc, h.ref, pos), m[6]);  // scene.rs:275 — always drawn
            res = res + light * mk3(m[3], m[4], m[5]);
            light = light * mk3(m[0], m[1], m[2]);
          }
        }
        acc = acc + res;
      }
      if (p.out_rgb) {
        const V3 px = acc / (double)p.rays_per_pixel;
        p.out_rgb[out_index * 3 + 0] = px.x, p.out_rgb[out_index * 3 + 1] = px.y, p.out_rgb[out_index * 3 + 2] = px.z;
      }
      if (p.out_ids) p.out_ids[out_index] = first;
    }
  }
  if (p.ray_counter) {
    unsigned long long r = n_rays, q = n_paths, e = n_panics, hh = n_hits;
    unsigned long long nv = tc.nodes, pt = USE_BVH ? (unsigned long long)tc.prims : (unsigned long long)n_rays * sc.n_obj;
    for (int o = 16; o > 0; o >>= 1) {
      r += __shfl_down_sync(FULL, r, o), q += __shfl_down_sync(FULL, q, o), e += __shfl_down_sync(FULL, e, o);
      hh += __shfl_down_sync(FULL, hh, o), nv += __shfl_down_sync(FULL, nv, o), pt += __shfl_down_sync(FULL, pt, o);
    }
    if ((threadIdx.x & 31u) == 0u) {
      atomicAdd(&p.ray_counter[0], r), atomicAdd(&p.ray_counter[1], q), atomicAdd(&p.ray_counter[3], hh);
      if (e) atomicAdd(&p.ray_counter[2], e);
      if (nv) atomicAdd(&p.ray_counter[4], nv);
      atomicAdd(&p.ray_counter[5], pt);
    }
  }
}
cudaError_t launch_render_rows(const RenderParams& p, cudaStream_t stream, uint32_t* out_launches) {
  const unsigned grid = (p.height + 31u) / 32u;
  if (p.use_bvh) render_rows_kernel<true><<<grid, 32, 0, stream>>>(p);
  else render_rows_kernel<false><<<grid, 32, 0, stream>>>(p);
  if (out_launches) *out_launches += 1;
  return cudaGetLastError();
}

// Launch helper used by the API layer.  Returns the cudaError of the launch.
cudaError_t launch_render(const RenderParams& p, int sm_count, cudaStream_t stream, uint32_t* out_launches) {
  if (!p.lane_kernel && pool_kernel_eligible(p)) return launch_render_pool(p, sm_count, stream, out_launches);
  constexpr uint32_t SMEM_LIMIT = 64u * 1024u;
  const bool use_smem = p.layout.bytes > 0 && p.layout.bytes <= SMEM_LIMIT;
  // a scanned scene beyond the stage with enough spheres to matter: chunked staging (two 52 KB buffers per CTA)
  const bool chunked = !use_smem && !p.use_bvh && p.layout.n_sph >= 256;
  const size_t smem = use_smem ? (size_t)p.layout.bytes + 128 : (chunked ? 128 + 2 * (size_t)CHUNK_BYTES : 0);
  cudaError_t e = cudaSuccess;
  const int blocks_per_sm = p.use_bvh ? BVH_BLOCKS_PER_SM : 2;
  const uint64_t total_work = (uint64_t)p.my_tiles * p.grans_per_unit;
  // persistent grid: a multiple of the SM count, never more warps than granules
  uint64_t want_blocks = (total_work + (RENDER_BLOCK / 32) - 1) / (RENDER_BLOCK / 32);
  uint64_t grid = (uint64_t)sm_count * blocks_per_sm;
  if (want_blocks < grid) grid = want_blocks ? want_blocks : 1;
  auto launch = [&](auto kernel) -> cudaError_t {
    if (smem) {
      cudaError_t a = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (a != cudaSuccess) return a;
    }
    kernel<<<(unsigned)grid, RENDER_BLOCK, smem, stream>>>(p);
    return cudaSuccess;
  };
  // specialised on triangle-free scenes (C3): the triangle code and its registers are compiled out
  const bool tri = p.layout.n_tri != 0;
  if (chunked)
    e = tri ? launch(render_megakernel<false, false, true, true>) : launch(render_megakernel<false, false, false, true>);
  else if (use_smem)
    e = p.use_bvh ? (tri ? launch(render_megakernel<true, true, true>) : launch(render_megakernel<true, true, false>))
                  : (tri ? launch(render_megakernel<true, false, true>) : launch(render_megakernel<true, false, false>));
  else
    e = p.use_bvh ? (tri ? launch(render_megakernel<false, true, true>) : launch(render_megakernel<false, true, false>))
                  : (tri ? launch(render_megakernel<false, false, true>) : launch(render_megakernel<false, false, false>));
  if (e != cudaSuccess) return e;
  if (out_launches) *out_launches += 1;
  return cudaGetLastError();
}

}  // namespace rtx
