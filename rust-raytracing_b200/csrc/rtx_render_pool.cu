// rtx_render_pool.cu — render kernel v2 for scenes that are scanned (no LBVH) from shared memory:
// the same arithmetic as render_megakernel (rtx_render.cu), scheduled so that a warp instruction almost
// always has 32 live lanes.
//
// Replaces the per-pixel closure of Scene::render (raytracing/scene.rs:144-170) and what it calls:
// render_pixel :194-212, get_ray_dir :213-222, render_ray :223-242, closest_object :243-251, avg :253-259,
// ray_hit :260-278, random_bounce_dir :279-291.
//
// Why: in the one-pixel-per-lane kernel a loop trip runs "start a camera sample" (for the lanes whose path
// just ended), "nearest hit" (all lanes) and "shade + bounce" (the lanes that hit) one after the other, each
// under its own lane mask — ncu: 18.7 of 32 lanes per instruction (profiles/r01_SUMMARY.md).  Here a warp
// owns POOL_SLOTS (64) pixel contexts in shared memory, twice as many as it has lanes.  Every context is in
// exactly one of three states — REGEN (needs a new camera sample, or a new pixel), TRACE (needs a nearest-hit
// query), SHADE (hit something, needs a bounce direction) — kept as three warp-uniform 64-bit masks.  A loop
// trip picks the fullest state, hands its first 32 contexts to the 32 lanes, runs that one phase for all of
// them, and moves each context to its next state.  With 64 contexts over 3 states some state nearly always
// holds >= 32.  A context still owns its pixel for all its samples and adds them up in sample order, so the
// f64 sum is the reference's left fold (iter_ops.rs:4-8) bit for bit, and every (pixel, sample) still owns
// its RNG stream: the image is identical to render_megakernel's, whatever the schedule.
#include <cuda_runtime.h>
#include <math_constants.h>

#include <cstdlib>

#include "rtx_render.cuh"
#include "rtx_trace.cuh"

namespace rtx {

constexpr int POOL_WARPS = 8;    // warps per CTA
constexpr int POOL_BLOCK = POOL_WARPS * 32;
constexpr int POOL_SLOTS_MAX = 64;  // pixel contexts per warp (one bit each in a 64-bit state mask)
constexpr uint32_t FULLMASK = 0xFFFFFFFFu;
constexpr uint32_t NO_PIXEL = 0xFFFFFFFFu;

// Context fields, SoA per warp: word(f, slot) = pool[f * POOL_SLOTS + slot], 8 bytes each.
enum : int {
  F_POS = 0,     // 3  ray.position            (ray.rs:5)
  F_DIR = 3,     // 3  ray.direction           (ray.rs:6)
  F_LIGHT = 6,   // 3  ray.light_color         (ray.rs:8)
  F_RES = 9,     // 3  ray.resulting_color     (ray.rs:7)
  F_ACC = 12,    // 3  running sum of the pixel's samples (scene.rs:253-259)
  F_FOCAL = 15,  // 3  camera.position + pixel_dir * focal_length (scene.rs:203), per pixel
  F_RNG = 18,    // 1  WyRand state of the current (pixel, sample) stream
  F_M0 = 19,     // 1  sample index | segment << 32
  F_M1 = 20,     // 1  out_index | hit ref << 32
  F_M2 = 21,     // 1  pixel index (NO_PIXEL = none) | hit object id << 32
  F_M3 = 22,     // 1  SAMPLE_QUEUE: tail_scratch slot of this context's (pixel, sample)
  F_COUNT = 23
};
template <int POOL_SLOTS>
struct Pool {
  static constexpr size_t BYTES_PER_WARP = (size_t)F_COUNT * POOL_SLOTS * 8;
  unsigned long long* w;  // this warp's fields
  __device__ __forceinline__ double ld(int f, uint32_t s) const { return __longlong_as_double((long long)w[f * POOL_SLOTS + s]); }
  __device__ __forceinline__ void st(int f, uint32_t s, double v) { w[f * POOL_SLOTS + s] = (unsigned long long)__double_as_longlong(v); }
  __device__ __forceinline__ V3 ld3(int f, uint32_t s) const { return mk3(ld(f, s), ld(f + 1, s), ld(f + 2, s)); }
  __device__ __forceinline__ void st3(int f, uint32_t s, V3 v) { st(f, s, v.x), st(f + 1, s, v.y), st(f + 2, s, v.z); }
  __device__ __forceinline__ unsigned long long ldu(int f, uint32_t s) const { return w[f * POOL_SLOTS + s]; }
  __device__ __forceinline__ void stu(int f, uint32_t s, unsigned long long v) { w[f * POOL_SLOTS + s] = v; }
};
__device__ __forceinline__ unsigned long long pack2(uint32_t lo, uint32_t hi) { return (unsigned long long)lo | ((unsigned long long)hi << 32); }

// 64-bit OR over the warp of one bit per lane (lanes without a context contribute nothing).
__device__ __forceinline__ unsigned long long warp_or_bit(bool on, uint32_t slot) {
  const uint32_t lo = (on && slot < 32u) ? (1u << slot) : 0u;
  const uint32_t hi = (on && slot >= 32u) ? (1u << (slot - 32u)) : 0u;
  return pack2(__reduce_or_sync(FULLMASK, lo), __reduce_or_sync(FULLMASK, hi));
}

// Ordered sum after a SAMPLE_QUEUE render: one thread per pixel adds its rpp per-sample radiances in sample
// order — the reference's left fold from zeros (iter_ops.rs:4-8) — then avg's division (scene.rs:253-259).
__global__ void __launch_bounds__(256) tail_sum_kernel(const __grid_constant__ RenderParams p) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t tg = t / GRAN_PIXELS, k = t % GRAN_PIXELS;
  if (tg >= p.tail_granules) return;
  const Granule g = decode_granule(p, p.tail_first + tg);
  if (k >= g.npix) return;
  uint32_t x, y, out_index;
  granule_pixel(p, g, k, &x, &y, &out_index);
  const uint32_t rpp = (uint32_t)p.rays_per_pixel;
  V3 acc = mk3(0.0, 0.0, 0.0);
  const double* src = p.tail_scratch + ((size_t)tg * rpp * GRAN_PIXELS + k) * 3;
#pragma unroll 4
  for (uint32_t s = 0; s < rpp; ++s, src += GRAN_PIXELS * 3) acc = acc + mk3(__ldcs(src), __ldcs(src + 1), __ldcs(src + 2));
  const V3 px = acc / (double)rpp;
  p.out_rgb[(size_t)out_index * 3 + 0] = px.x;
  p.out_rgb[(size_t)out_index * 3 + 1] = px.y;
  p.out_rgb[(size_t)out_index * 3 + 2] = px.z;
}

template <int POOL_SLOTS, int MIN_BLOCKS, bool SAMPLE_QUEUE>
__global__ void __launch_bounds__(POOL_BLOCK, MIN_BLOCKS) render_pool_kernel(const __grid_constant__ RenderParams p) {
  constexpr size_t POOL_BYTES_PER_WARP = Pool<POOL_SLOTS>::BYTES_PER_WARP;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  // [0,128) mbarrier | scene blob | (128-aligned) POOL_WARPS pools | POOL_WARPS x 64-byte index scratch |
  // glibc sin/cos table (3520 B)
  uint64_t* mbar = reinterpret_cast<uint64_t*>(smem_raw);
  unsigned char* blob = smem_raw + 128;
  stage_blob(blob, p.scene_blob, p.layout.bytes, mbar);
  const SceneView sc = make_view(blob, p.layout);
  const uint32_t pools_off = (128u + p.layout.bytes + 127u) & ~127u;

  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  const uint32_t lt_mask = (1u << lane) - 1u;
  Pool<POOL_SLOTS> P;
  P.w = reinterpret_cast<unsigned long long*>(smem_raw + pools_off + warp * POOL_BYTES_PER_WARP);
  unsigned char* idx = smem_raw + pools_off + POOL_WARPS * POOL_BYTES_PER_WARP + warp * POOL_SLOTS_MAX;
  double* sctab = reinterpret_cast<double*>(smem_raw + pools_off + POOL_WARPS * POOL_BYTES_PER_WARP + POOL_WARPS * POOL_SLOTS_MAX);
  for (uint32_t i = threadIdx.x; i < 440u; i += POOL_BLOCK) sctab[i] = SINCOS_TAB_DEV[i];
  __syncthreads();

  const V3 cam_pos = mk3(p.cam_pos[0], p.cam_pos[1], p.cam_pos[2]);
  const uint32_t rpp = (uint32_t)p.rays_per_pixel;
  const uint32_t max_segments = (uint32_t)p.max_bounces + 1u;  // scene.rs:227
  const uint32_t total_work = p.tail_first + p.tail_granules * rpp;

  // every context starts without a pixel, in REGEN
  P.stu(F_M2, lane, pack2(NO_PIXEL, 0));
  P.stu(F_M0, lane, 0);
  if (lane + 32u < (uint32_t)POOL_SLOTS) {
    P.stu(F_M2, lane + 32u, pack2(NO_PIXEL, 0));
    P.stu(F_M0, lane + 32u, 0);
  }
  __syncwarp();
  unsigned long long m_regen = POOL_SLOTS == 64 ? ~0ull : ((1ull << (POOL_SLOTS & 63)) - 1ull), m_trace = 0ull, m_shade = 0ull;

  // warp-uniform work-queue state (granule = 8x4 pixels of a 16x16 tile, as in render_megakernel)
  Granule g;
  g.x0 = g.y0 = g.w = g.npix = g.base_out = g.tx0 = g.ty0 = 0;
  uint32_t g_next = 0, g_sample = 0, g_slot_base = 0;
  bool g_single = false;
  bool exhausted = false;
  uint32_t n_rays = 0, n_paths = 0, n_panics = 0, n_hits = 0;

  for (;;) {
    const int c_t = __popcll(m_trace), c_s = __popcll(m_shade), c_r = __popcll(m_regen);
    if ((c_t | c_s | c_r) == 0) break;
    // the fullest state runs; ties: TRACE, then SHADE, then REGEN
    int phase = 0;  // 0 TRACE, 1 SHADE, 2 REGEN
    unsigned long long m = m_trace;
    int c = c_t;
    if (c_s > c) phase = 1, m = m_shade, c = c_s;
    if (c_r > c) phase = 2, m = m_regen, c = c_r;
    // hand the first min(32, c) contexts of `m` to lanes 0..: lane j looks at bits j and j+32
    {
      const uint32_t lo = (uint32_t)m, hi = (uint32_t)(m >> 32);
      const uint32_t c_lo = __popc(lo);
      if ((lo >> lane) & 1u) idx[__popc(lo & lt_mask)] = (unsigned char)lane;
      if ((hi >> lane) & 1u) idx[c_lo + __popc(hi & lt_mask)] = (unsigned char)(lane + 32u);
      __syncwarp();
    }
    const bool has = (int)lane < c;  // c may exceed 32: then every lane has one
    const uint32_t slot = has ? idx[lane] : 0u;
    __syncwarp();
    const unsigned long long sel = warp_or_bit(has, slot);

    if (phase == 0) {
      // ---- TRACE: one segment of render_ray (scene.rs:227-240) up to, not including, the bounce direction ----
      bool to_shade = false;
      if (has) {
        V3 pos = P.ld3(F_POS, slot);
        const V3 dir = P.ld3(F_DIR, slot);
        const unsigned long long m0 = P.ldu(F_M0, slot);
        uint32_t sample = (uint32_t)m0, segment = (uint32_t)(m0 >> 32);
        bool end = false;
        V3 res;
        if (sc.n_obj == 0) {
          end = true;  // scene.rs:224-226
          res = P.ld3(F_RES, slot);
        } else {
          const Hit h = closest_brute(sc, pos, dir, &n_panics);
          ++n_rays;
          if (segment == 0 && sample == 0 && p.out_ids) p.out_ids[(uint32_t)P.ldu(F_M1, slot)] = h.id;
          res = P.ld3(F_RES, slot);
          if (h.id == MISS_ID) {
            end = true;
          } else {
            pos = pos + dir * h.t;  // scene.rs:234
            ++n_hits;
            const double* mt = sc.mat + (size_t)h.id * MAT_STRIDE;
            const V3 base_color = mk3(mt[0], mt[1], mt[2]);
            const V3 emission = mk3(mt[3], mt[4], mt[5]);
            const V3 light = P.ld3(F_LIGHT, slot);
            const V3 new_light = light * base_color;
            ++segment;
            // The bounce direction (scene.rs:275) is only consumed by a later segment: when this was the last
            // allowed segment (:227) or light_color is now zero (:228 breaks before the next query) it is dead,
            // and the stream is private to this (pixel, sample) — skip SHADE.
            const bool last = (segment == max_segments) ||
                              (new_light.x == 0.0 && new_light.y == 0.0 && new_light.z == 0.0);
            res = res + light * emission;  // scene.rs:276
            end = last;
            if (!last) {
              P.st3(F_POS, slot, pos);
              P.st3(F_LIGHT, slot, new_light);  // scene.rs:277
              P.st3(F_RES, slot, res);
              P.stu(F_M0, slot, pack2(sample, segment));
              P.stu(F_M1, slot, pack2((uint32_t)P.ldu(F_M1, slot), h.ref));
              P.stu(F_M2, slot, pack2((uint32_t)P.ldu(F_M2, slot), h.id));
              to_shade = true;
            }
          }
        }
        if (end) {
          if (SAMPLE_QUEUE) {
            // single-sample context: this sample's radiance goes to the scratch, the context is free again
            double* dst = p.tail_scratch + (size_t)(uint32_t)P.ldu(F_M3, slot) * 3;
            dst[0] = res.x, dst[1] = res.y, dst[2] = res.z;
            P.stu(F_M2, slot, pack2(NO_PIXEL, 0));
            P.stu(F_M0, slot, 0);
          } else {
            P.st3(F_ACC, slot, P.ld3(F_ACC, slot) + res);  // Sum: a + b, left to right
            P.stu(F_M0, slot, pack2(sample + 1u, 0u));
          }
        }
      }
      const unsigned long long to_s = warp_or_bit(to_shade, slot);
      m_trace &= ~sel;
      m_shade |= to_s;
      m_regen |= sel & ~to_s;
    } else if (phase == 1) {
      // ---- SHADE: normal_at + random_bounce_dir (scene.rs:274-275, 279-291) ----
      if (has) {
        const V3 pos = P.ld3(F_POS, slot);
        const V3 dir = P.ld3(F_DIR, slot);
        const uint32_t ref = (uint32_t)(P.ldu(F_M1, slot) >> 32);
        const uint32_t id = (uint32_t)(P.ldu(F_M2, slot) >> 32);
        WyRand rng;
        rng.s = P.ldu(F_RNG, slot);
        const V3 n = normal_at(sc, ref, pos);
        const V3 nd = random_bounce_dir(rng, dir, n, sc.mat[(size_t)id * MAT_STRIDE + 6], sctab);
        P.st3(F_DIR, slot, nd);
        P.stu(F_RNG, slot, rng.s);
      }
      m_shade &= ~sel;
      m_trace |= sel;
    } else {
      // ---- REGEN: finish the pixel (avg, scene.rs:253-259), take the next one from the queue, start the next
      // camera sample (scene.rs:199-210) ----
      uint32_t pix = NO_PIXEL, sample = 0, out_index = 0;
      if (has) {
        pix = (uint32_t)P.ldu(F_M2, slot);
        sample = (uint32_t)P.ldu(F_M0, slot);
        out_index = (uint32_t)P.ldu(F_M1, slot);
        if (pix != NO_PIXEL && sample == rpp) {
          if (p.out_rgb) {
            const V3 px = P.ld3(F_ACC, slot) / (double)rpp;  // avg: `/ length` (scene.rs:258)
            p.out_rgb[(size_t)out_index * 3 + 0] = px.x;
            p.out_rgb[(size_t)out_index * 3 + 1] = px.y;
            p.out_rgb[(size_t)out_index * 3 + 2] = px.z;
          }
          pix = NO_PIXEL;
        }
      }
      bool need = has && pix == NO_PIXEL;
      uint32_t need_m = __ballot_sync(FULLMASK, need);
      while (need_m != 0u && !exhausted) {
        if (g_next >= g.npix) {
          uint32_t w = 0;
          if (lane == 0) w = atomicAdd(p.work_counter, 1u);
          w = __shfl_sync(FULLMASK, w, 0);
          if (w >= total_work) {
            exhausted = true;
            break;
          }
          uint32_t gidx = w;
          g_single = false, g_sample = 0, g_slot_base = 0;
          if (SAMPLE_QUEUE) {  // a (granule, sample) entry: one sample per context
            const uint32_t v = w - p.tail_first, tg = v / rpp;
            g_sample = v - tg * rpp;
            gidx = p.tail_first + tg;
            g_single = true;
            g_slot_base = (tg * rpp + g_sample) * GRAN_PIXELS;
          }
          g = decode_granule(p, gidx);
          g_next = 0;
          continue;  // re-check (the granule may be empty at the image border)
        }
        const uint32_t k = g_next + __popc(need_m & lt_mask);
        if (need && k < g.npix) {
          uint32_t x, y;
          granule_pixel(p, g, k, &x, &y, &out_index);
          pix = y * p.width + x;
          sample = g_sample;
          if (SAMPLE_QUEUE) P.stu(F_M3, slot, pack2(g_slot_base + k, 0u));
          // scene.rs:213-222 with the libm calls hoisted to the host tables; camera.rs:65-67; scene.rs:203
          M3 to_world;
          to_world.x = mk3(p.to_world[0], p.to_world[1], p.to_world[2]);
          to_world.y = mk3(p.to_world[3], p.to_world[4], p.to_world[5]);
          to_world.z = mk3(p.to_world[6], p.to_world[7], p.to_world[8]);
          const V3 cam_dir = mk3(p.sin_x[x], p.sin_y[y], fmul(p.cos_x[x], p.cos_y[y]));
          const V3 pix_dir = matvec(to_world, cam_dir);
          P.st3(F_FOCAL, slot, cam_pos + pix_dir * p.focal_length);
          P.st3(F_ACC, slot, mk3(0.0, 0.0, 0.0));  // iter_ops.rs:4-8: the fold starts from zeros()
          if (p.out_ids && g_sample == 0) p.out_ids[out_index] = MISS_ID;
          need = false;
          if (rpp == 0) {  // avg of an empty iterator: zeros / 0.0 = NaN (scene.rs:253-259)
            if (p.out_rgb) {
              const V3 px = mk3(0.0, 0.0, 0.0) / (double)rpp;
              p.out_rgb[(size_t)out_index * 3 + 0] = px.x;
              p.out_rgb[(size_t)out_index * 3 + 1] = px.y;
              p.out_rgb[(size_t)out_index * 3 + 2] = px.z;
            }
            pix = NO_PIXEL;
            need = true;
          }
        }
        g_next = min(g_next + (uint32_t)__popc(need_m), g.npix);
        need_m = __ballot_sync(FULLMASK, need);
      }
      const bool go = has && pix != NO_PIXEL;
      if (go) {
        WyRand rng;
        rng.s = stream_state(p.seed, (uint64_t)pix * rpp + sample);
        const double u0 = wy_f64(rng), u1 = wy_f64(rng), u2 = wy_f64(rng);
        const V3 pos = cam_pos + mk3(u0, u1, u2) * p.non_focal_offset;
        const double u3 = wy_f64(rng), u4 = wy_f64(rng), u5 = wy_f64(rng);
        const V3 target = P.ld3(F_FOCAL, slot) + mk3(u3, u4, u5) * p.focal_offset;
        P.st3(F_POS, slot, pos);
        P.st3(F_DIR, slot, vnorm(target - pos));
        P.st3(F_LIGHT, slot, mk3(1.0, 1.0, 1.0));
        P.st3(F_RES, slot, mk3(0.0, 0.0, 0.0));
        P.stu(F_RNG, slot, rng.s);
        P.stu(F_M0, slot, pack2(sample, 0u));
        P.stu(F_M1, slot, pack2(out_index, 0u));
        P.stu(F_M2, slot, pack2(pix, 0u));
        ++n_paths;
      }
      m_regen &= ~sel;
      m_trace |= warp_or_bit(go, slot);  // contexts left without a pixel (queue exhausted) leave every mask
    }
    __syncwarp();
  }

  // ---- counters: one atomic per warp ----
  if (p.ray_counter) {
    unsigned long long r = n_rays, q = n_paths, e = n_panics, hh = n_hits;
    unsigned long long pt = (unsigned long long)n_rays * sc.n_obj;
    for (int o = 16; o > 0; o >>= 1) {
      r += __shfl_down_sync(FULLMASK, r, o);
      q += __shfl_down_sync(FULLMASK, q, o);
      e += __shfl_down_sync(FULLMASK, e, o);
      hh += __shfl_down_sync(FULLMASK, hh, o);
      pt += __shfl_down_sync(FULLMASK, pt, o);
    }
    if (lane == 0) {
      atomicAdd(&p.ray_counter[0], r);
      atomicAdd(&p.ray_counter[1], q);
      if (e) atomicAdd(&p.ray_counter[2], e);
      atomicAdd(&p.ray_counter[3], hh);
      atomicAdd(&p.ray_counter[5], pt);
    }
  }
}

bool pool_kernel_eligible(const RenderParams& p) {
  return !p.use_bvh && p.layout.bytes > 0 && p.layout.bytes <= 16u * 1024u && p.rays_per_pixel < 0x7FFFFFFFull &&
         p.max_bounces < 0x7FFFFFFEull;
}

// Which queue this launch uses.  With whole-pixel entries a context runs its pixel's rpp samples one after
// the other, so the kernel cannot end before the most expensive pixel handed out last has finished:
// rpp x (up to max_bounces + 1) segments of latency — on S1 at 64 spp more than a millisecond of under-filled
// machine per launch whatever the shard size, which is what limits strong scaling.  The SAMPLE_QUEUE hands
// out (granule, sample) entries instead: the tail is one path, at the price of 24 B of scratch per path and
// the ordered sum afterwards.  Chosen when rpp >= 4 and the scratch fits the budget
// (RTX_FLAG_PIXEL_QUEUE / RTX_FLAG_SAMPLE_QUEUE force one or the other).
// Returns the scratch bytes needed (0 = whole-pixel queue): [granules * rpp * 32 * 3] f64.
size_t pool_plan_tail(RenderParams& p, int sm_count) {
  const uint64_t granules = (uint64_t)p.my_tiles * GRANS_PER_TILE;
  p.tail_first = (uint32_t)granules;
  p.tail_granules = 0;
  p.tail_scratch = nullptr;
  p.tail_done = nullptr;
  const int force = p.queue_mode == 1 ? 0 : (p.queue_mode == 2 ? 1 : -1);
  const uint64_t rpp = p.rays_per_pixel;
  if (force == 0 || !p.out_rgb || rpp < 1 || rpp > 65536) return 0;
  if (force != 1 && rpp < 4) return 0;
  // Whole-pixel entries are ~4 % faster in the steady state (no scratch, no sum pass): keep them when every
  // resident context gets many pixels, so the tail is a small share (measured cross-over: N = 1 at 1080p,
  // 13.7 pixels per context, 12.32 vs 12.77 ms; N = 2, 6.8 per context, 6.72 vs 6.41 ms).
  const uint64_t contexts = (uint64_t)sm_count * 2 * POOL_WARPS * POOL_SLOTS_MAX;
  if (force != 1 && granules * GRAN_PIXELS >= 10 * contexts) return 0;
  const uint64_t cap_bytes = 4ull << 30;  // scratch budget
  const uint64_t bytes = granules * rpp * GRAN_PIXELS * 24;
  if (bytes > cap_bytes || granules * rpp * GRAN_PIXELS > 0xFFFFFFFFull) return 0;
  p.tail_granules = (uint32_t)granules;
  p.tail_first = 0;
  return (size_t)bytes;
}

template <int POOL_SLOTS, int MIN_BLOCKS, bool SAMPLE_QUEUE>
static cudaError_t launch_pool_variant(const RenderParams& p, int sm_count, cudaStream_t stream, uint32_t* out_launches) {
  const size_t pools_off = (128u + (size_t)p.layout.bytes + 127u) & ~(size_t)127u;
  const size_t smem = pools_off + POOL_WARPS * Pool<POOL_SLOTS>::BYTES_PER_WARP + POOL_WARPS * POOL_SLOTS_MAX + 440 * sizeof(double);
  cudaError_t e = cudaMemsetAsync(p.work_counter, 0, sizeof(uint32_t), stream);
  if (e != cudaSuccess) return e;
  const uint64_t total_work = (uint64_t)p.tail_first + (uint64_t)p.tail_granules * p.rays_per_pixel;
  // persistent grid: MIN_BLOCKS CTAs per SM, never more contexts than queue entries can feed
  const uint64_t per_block = (uint64_t)POOL_WARPS * POOL_SLOTS / GRAN_PIXELS;  // queue entries one CTA can hold at once
  uint64_t want_blocks = (total_work + per_block - 1) / per_block;
  uint64_t grid = (uint64_t)sm_count * MIN_BLOCKS;
  if (want_blocks < grid) grid = want_blocks ? want_blocks : 1;
  e = cudaFuncSetAttribute(render_pool_kernel<POOL_SLOTS, MIN_BLOCKS, SAMPLE_QUEUE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  render_pool_kernel<POOL_SLOTS, MIN_BLOCKS, SAMPLE_QUEUE><<<(unsigned)grid, POOL_BLOCK, smem, stream>>>(p);
  if (out_launches) *out_launches += 1;
  e = cudaGetLastError();
  if (e != cudaSuccess || !SAMPLE_QUEUE) return e;
  const unsigned threads = p.tail_granules * GRAN_PIXELS;
  tail_sum_kernel<<<(threads + 255) / 256, 256, 0, stream>>>(p);
  if (out_launches) *out_launches += 1;
  return cudaGetLastError();
}

cudaError_t launch_render_pool(const RenderParams& p, int sm_count, cudaStream_t stream, uint32_t* out_launches) {
  // RTX_POOL_VARIANT (measurement knob): 0 = 64 contexts/warp, 2 CTAs/SM; 1 = 48 contexts/warp, 3 CTAs/SM
  static const int variant = [] { const char* v = getenv("RTX_POOL_VARIANT"); return v ? atoi(v) : 0; }();
  if (p.tail_granules != 0) return launch_pool_variant<64, 2, true>(p, sm_count, stream, out_launches);
  if (variant == 1) return launch_pool_variant<48, 3, false>(p, sm_count, stream, out_launches);
  return launch_pool_variant<64, 2, false>(p, sm_count, stream, out_launches);
}

}  // namespace rtx
