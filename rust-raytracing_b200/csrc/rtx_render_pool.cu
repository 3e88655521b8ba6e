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

constexpr int POOL_WARPS_PIXEL_QUEUE = 20;  // resident warps per SM of the pixel-queue variant (pool_plan_tail)
constexpr int POOL_SLOTS_MAX = 64;  // pixel contexts per warp (one bit each in a 64-bit state mask)
constexpr int POOL_WARP_SCRATCH = POOL_SLOTS_MAX;  // per warp: compaction indices
constexpr uint32_t FULLMASK = 0xFFFFFFFFu;
constexpr uint32_t NO_PIXEL = 0xFFFFFFFFu;
// Sample queue: consecutive samples of a granule one queue entry covers (one atomic + one granule decode per entry)
constexpr uint32_t SAMPLES_PER_ENTRY = 8;

// Context fields, SoA per warp: word(f, slot) = pool[f * POOL_SLOTS + slot], 8 bytes each.  A pixel-queue
// context keeps its pixel for all rpp samples (running sum + the pixel's focal point: 22 words); a sample-queue
// context lives for one camera path (17 words).  Fewer words = more resident warps per SM.
template <int POOL_SLOTS, bool SAMPLE_QUEUE>
struct Pool {
  enum : int {
    F_POS = 0,     // 3  ray.position            (ray.rs:5)
    F_DIR = 3,     // 3  ray.direction           (ray.rs:6)
    F_LIGHT = 6,   // 3  ray.light_color         (ray.rs:8)
    F_RES = 9,     // 3  ray.resulting_color     (ray.rs:7)
    F_RNG = 12,    // 1  WyRand state of the current (pixel, sample) stream
    F_M0 = 13,     // 1  sample index | segment << 32
    F_M1 = 14,     // 1  out_index | hit ref << 32
    F_M2 = 15,     // 1  pixel index (NO_PIXEL = none) | hit object id << 32
    F_M3 = 16,     // 1  SAMPLE_QUEUE only: tail_scratch slot of this context's (pixel, sample)
    F_ACC = 16,    // 3  pixel queue only: running sum of the pixel's samples (scene.rs:253-259)
    F_FOCAL = 19,  // 3  pixel queue only: camera.position + pixel_dir * focal_length (scene.rs:203), per pixel
    F_COUNT = SAMPLE_QUEUE ? 17 : 22
  };
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
  const size_t slot0 = (size_t)tg * rpp * GRAN_PIXELS + k;
  const double* src = p.tail_scratch + slot0 * 3;
  // one flag byte per camera path, behind the values: 0 = the path's radiance was (+-)0 and its values were never
  // written.  Skipping `acc + 0` is exact: acc starts at +0 and x + (+-0) == x for every x that can arise (a sum
  // that starts at +0 is never -0).
  const unsigned char* flag = reinterpret_cast<const unsigned char*>(p.tail_scratch + (size_t)p.tail_granules * rpp * GRAN_PIXELS * 3) + slot0;
#pragma unroll 4
  for (uint32_t s = 0; s < rpp; ++s, src += GRAN_PIXELS * 3, flag += GRAN_PIXELS)
    if (__ldcs(flag)) acc = acc + mk3(__ldcs(src), __ldcs(src + 1), __ldcs(src + 2));
  const V3 px = acc / (double)rpp;
  p.out_rgb[(size_t)out_index * 3 + 0] = px.x;
  p.out_rgb[(size_t)out_index * 3 + 1] = px.y;
  p.out_rgb[(size_t)out_index * 3 + 2] = px.z;
}

// Registers per thread that still let `warps` warps share an SM's 64 K registers (allocated 8 per thread at a time).
constexpr int pool_max_regs(int warps) { return (65536 / (warps * 32)) / 8 * 8 > 128 ? 128 : (65536 / (warps * 32)) / 8 * 8; }

template <int POOL_SLOTS, int MIN_BLOCKS, bool SAMPLE_QUEUE, int POOL_WARPS = 8, bool HAS_TRI = true>
__global__ void __launch_bounds__(POOL_WARPS * 32, MIN_BLOCKS) __maxnreg__(pool_max_regs(POOL_WARPS * MIN_BLOCKS)) render_pool_kernel(const __grid_constant__ RenderParams p) {
  constexpr int POOL_BLOCK = POOL_WARPS * 32;
  constexpr size_t POOL_BYTES_PER_WARP = Pool<POOL_SLOTS, SAMPLE_QUEUE>::BYTES_PER_WARP;
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
  using PoolT = Pool<POOL_SLOTS, SAMPLE_QUEUE>;
  constexpr int F_POS = PoolT::F_POS, F_DIR = PoolT::F_DIR, F_LIGHT = PoolT::F_LIGHT, F_RES = PoolT::F_RES,
                F_RNG = PoolT::F_RNG, F_M0 = PoolT::F_M0, F_M1 = PoolT::F_M1, F_M2 = PoolT::F_M2, F_M3 = PoolT::F_M3,
                F_ACC = PoolT::F_ACC, F_FOCAL = PoolT::F_FOCAL;
  PoolT P;
  P.w = reinterpret_cast<unsigned long long*>(smem_raw + pools_off + warp * POOL_BYTES_PER_WARP);
  unsigned char* idx = smem_raw + pools_off + POOL_WARPS * POOL_BYTES_PER_WARP + warp * POOL_WARP_SCRATCH;
  double* sctab = reinterpret_cast<double*>(smem_raw + pools_off + POOL_WARPS * POOL_BYTES_PER_WARP + POOL_WARPS * POOL_WARP_SCRATCH);
  // (a depth-0 render never bounces: nothing reads the table)
  if (p.max_bounces != 0)
    for (uint32_t i = threadIdx.x; i < 440u; i += POOL_BLOCK) sctab[i] = SINCOS_TAB_DEV[i];
  __syncthreads();

  const V3 cam_pos = mk3(p.cam_pos[0], p.cam_pos[1], p.cam_pos[2]);
  const uint32_t rpp = (uint32_t)p.rays_per_pixel;
  const uint32_t max_segments = (uint32_t)p.max_bounces + 1u;  // scene.rs:227
  const uint32_t entries_per_granule = (rpp + SAMPLES_PER_ENTRY - 1u) / SAMPLES_PER_ENTRY;
  const uint32_t total_work = p.tail_first + p.tail_granules * entries_per_granule;

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
  uint32_t g_next = 0, g_sample = 0, g_sample_end = 0, g_slot_base = 0;
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
          const Hit h = closest_brute<HAS_TRI>(sc, pos, dir, &n_panics);
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
            const size_t sslot = (size_t)(uint32_t)P.ldu(F_M3, slot);
            const bool nonzero = !(res.x == 0.0 && res.y == 0.0 && res.z == 0.0);  // (NaN counts as non-zero)
            reinterpret_cast<unsigned char*>(p.tail_scratch + (size_t)p.tail_granules * rpp * GRAN_PIXELS * 3)[sslot] = nonzero ? 1 : 0;
            if (nonzero) {  // (plain stores: __stcg's STG.STRONG.GPU cost 11 % of the kernel)
              double* dst = p.tail_scratch + sslot * 3;
              dst[0] = res.x, dst[1] = res.y, dst[2] = res.z;
            }
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
        if (!SAMPLE_QUEUE && pix != NO_PIXEL && sample == rpp) {
          if (p.out_rgb) {
            const V3 px = P.ld3(F_ACC, slot) / (double)rpp;  // avg: `/ length` (scene.rs:258)
            p.out_rgb[(size_t)out_index * 3 + 0] = px.x;
            p.out_rgb[(size_t)out_index * 3 + 1] = px.y;
            p.out_rgb[(size_t)out_index * 3 + 2] = px.z;
          }
          pix = NO_PIXEL;
        }
      }
      V3 focal = mk3(0.0, 0.0, 0.0);  // SAMPLE_QUEUE: a context's focal point lives only inside this trip
      bool need = has && pix == NO_PIXEL;
      uint32_t need_m = __ballot_sync(FULLMASK, need);
      while (need_m != 0u && !exhausted) {
        if (g_next >= g.npix) {
          if (SAMPLE_QUEUE && g_sample + 1u < g_sample_end) {  // next sample of the entry
            ++g_sample;
            g_slot_base += GRAN_PIXELS;
            g_next = 0;
            continue;
          }
          uint32_t w = 0;
          if (lane == 0) w = atomicAdd(p.work_counter, 1u);
          w = __shfl_sync(FULLMASK, w, 0);
          if (w >= total_work) {
            exhausted = true;
            break;
          }
          uint32_t gidx = w;
          g_sample = 0, g_slot_base = 0;
          if (SAMPLE_QUEUE) {  // a (granule, samples [g_sample, g_sample_end)) entry: one sample per context
            const uint32_t tg = w / entries_per_granule;
            g_sample = (w - tg * entries_per_granule) * SAMPLES_PER_ENTRY;
            g_sample_end = min(rpp, g_sample + SAMPLES_PER_ENTRY);
            gidx = tg;
            g_slot_base = (tg * rpp + g_sample) * GRAN_PIXELS;
          }
          g = decode_granule_cold(p, gidx);
          g_next = 0;
          continue;  // re-check (the granule may be empty at the image border)
        }
        const uint32_t k = g_next + __popc(need_m & lt_mask);
        if (need && k < g.npix) {
          uint32_t x, y;
          granule_pixel(p, g, k, &x, &y, &out_index);
          pix = y * p.width + x;
          sample = g_sample;
          // scene.rs:213-222 with the libm calls hoisted to the host tables; camera.rs:65-67; scene.rs:203
          M3 to_world;
          to_world.x = mk3(p.to_world[0], p.to_world[1], p.to_world[2]);
          to_world.y = mk3(p.to_world[3], p.to_world[4], p.to_world[5]);
          to_world.z = mk3(p.to_world[6], p.to_world[7], p.to_world[8]);
          const V3 cam_dir = mk3(p.sin_x[x], p.sin_y[y], fmul(p.cos_x[x], p.cos_y[y]));
          const V3 pix_dir = matvec(to_world, cam_dir);
          focal = cam_pos + pix_dir * p.focal_length;
          if (!SAMPLE_QUEUE) P.st3(F_FOCAL, slot, focal);
          if (SAMPLE_QUEUE)
            P.stu(F_M3, slot, pack2(g_slot_base + k, 0u));
          else
            P.st3(F_ACC, slot, mk3(0.0, 0.0, 0.0));  // iter_ops.rs:4-8: the fold starts from zeros()
          if (p.out_ids && g_sample == 0) p.out_ids[out_index] = MISS_ID;
          need = false;
          if (!SAMPLE_QUEUE && rpp == 0) {  // avg of an empty iterator: zeros / 0.0 = NaN (scene.rs:253-259); the sample queue needs rpp >= 1
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
        // sample queue: every context that starts a path here was handed its pixel in this trip
        const V3 target = (SAMPLE_QUEUE ? focal : P.ld3(F_FOCAL, slot)) + mk3(u3, u4, u5) * p.focal_offset;
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
  queue_release(p);
}

// Scanned scenes whose blob fits next to the pools.  Not for primary-only renders (max_bounces == 0, config C5): every
// path is one ray, there is no divergence for the pool to repair, and the lane kernel's shorter prologue and loop win
// at every size (CUDA-graph replay, tools/bench_c5_small.py: 256x256 7.2 vs 12.7 us per frame, 7680x4320 1.309 vs
// 1.362 ms).
bool pool_kernel_eligible(const RenderParams& p) {
  return !p.use_bvh && p.layout.bytes > 0 && p.layout.bytes <= 16u * 1024u && p.rays_per_pixel < 0x7FFFFFFFull &&
         p.max_bounces != 0 && p.max_bounces < 0x7FFFFFFEull;
}

// Which queue this launch uses.
//   pixel queue   an entry is a whole granule and a context keeps its pixel for all rpp samples (running sum in the
//                 context, 22 words, 20 warps per SM); no scratch, and every finished pixel is stored at once — a frame
//                 in pinned host memory streams out over PCIe while the launch is still rendering.  But a warp's 64
//                 contexts drift onto pixels of unrelated granules at unrelated depths (25.7 of 32 lanes per
//                 instruction), and the launch cannot end before the last pixel handed out has run its rpp samples
//                 one after the other.
//   sample queue  an entry is a (granule, sample): one camera path per context (17 words, 24 warps per SM), per-sample
//                 radiance to a 24 B scratch slot, tail_sum_kernel adds up afterwards.  A warp works on one or two
//                 granules at a time (28.1 lanes per instruction) and the launch ends within one path of the last
//                 entry: C2 11.35 ms against 12.2 ms, a 1/8 shard 1.65 against 2.8 ms.  The frame only appears in
//                 the sum pass, which is an exposed PCIe write when it goes straight to host memory (C2 e2e 14.1 ms).
// Auto: the sample queue when rpp >= 4 and the scratch fits the budget — except for a large frame rendered straight
// into pinned host memory, where the pixel queue's overlapped stores win (C2 e2e 12.3 ms).
// RTX_FLAG_PIXEL_QUEUE / RTX_FLAG_SAMPLE_QUEUE force one.  The image is the same either way.
// Returns the scratch bytes needed (0 = pixel queue): [granules * rpp * 32 * 3] f64 + [granules * rpp * 32] flag bytes.
size_t pool_plan_tail(RenderParams& p, int sm_count) {
  const uint64_t granules = (uint64_t)p.my_tiles * p.grans_per_unit;
  p.tail_first = (uint32_t)granules;
  p.tail_granules = 0;
  p.tail_scratch = nullptr;
  // RTX_QUEUE_MODE (measurement builds, -DRTX_MEASUREMENT_KNOBS, and only when the caller left the choice open):
  // 1 pixel queue, 2 sample queue
#ifdef RTX_MEASUREMENT_KNOBS
  static const int env_mode = [] { const char* v = getenv("RTX_QUEUE_MODE"); return v ? atoi(v) : 0; }();
#else
  constexpr int env_mode = 0;
#endif
  const uint32_t mode = p.queue_mode ? p.queue_mode : (uint32_t)env_mode;
  const uint64_t rpp = p.rays_per_pixel;
  if (mode == 1 || !p.out_rgb || rpp < 1 || rpp > 65536) return 0;
  if (mode != 2 && rpp < 4) return 0;
  // a frame is "large" when every resident context gets many pixels, so that draining is a small share of the launch
  const uint64_t contexts = (uint64_t)sm_count * POOL_WARPS_PIXEL_QUEUE * POOL_SLOTS_MAX;
  const bool large = granules * GRAN_PIXELS >= 10 * contexts;
  if (mode != 2 && large && p.out_is_host) return 0;
  const uint64_t cap_bytes = 4ull << 30;  // scratch budget
  const uint64_t bytes = granules * rpp * GRAN_PIXELS * 25;  // 24 B of radiance + 1 flag byte per camera path
  if (bytes > cap_bytes || granules * rpp * GRAN_PIXELS > 0xFFFFFFFFull) return 0;
  p.tail_granules = (uint32_t)granules;
  p.tail_first = 0;
  return (size_t)bytes;
}

uint32_t pool_host_bands(const RenderParams& p, int sm_count) {
  if (p.queue_mode != 0 || !p.out_rgb || p.shard_count != 1 || p.rays_per_pixel < 4 || p.rays_per_pixel > 65536) return 1;
#ifdef RTX_MEASUREMENT_KNOBS
  static const int env_bands = [] { const char* v = getenv("RTX_HOST_BANDS"); return v ? atoi(v) : 0; }();
#else
  constexpr int env_bands = 0;
#endif
  const uint64_t contexts = (uint64_t)sm_count * 24 * POOL_SLOTS_MAX;
  const uint64_t pixels = (uint64_t)p.width * p.height;
  // measured on C2 (9.1 pixels per context): 1 band (pixel queue, zero-copy) 12.21 ms, 2: 11.57, 3: 11.46, 4: 11.39, 6: 11.42, 8: 11.43
  uint32_t bands = pixels >= 8 * contexts ? 4u : (pixels >= 4 * contexts ? 2u : 1u);
  if (env_bands > 0) bands = (uint32_t)env_bands;
  if (bands > 8u) bands = 8u;
  if (bands > p.tiles_y) bands = p.tiles_y;
  // every band's scratch has to fit the budget of pool_plan_tail
  const uint64_t band_pixels = ((uint64_t)(p.tiles_y + bands - 1) / bands) * p.tiles_x * TILE * TILE;
  if (band_pixels * p.rays_per_pixel * 25 > (4ull << 30)) return 1;
  return bands < 1u ? 1u : bands;
}

// Dynamic shared memory of one CTA of a variant: mbarrier | blob | pools | index scratch | sin/cos table.
template <int POOL_SLOTS, bool SAMPLE_QUEUE, int POOL_WARPS>
static size_t pool_smem_bytes(uint32_t blob_bytes) {
  const size_t pools_off = (128u + (size_t)blob_bytes + 127u) & ~(size_t)127u;
  return pools_off + POOL_WARPS * Pool<POOL_SLOTS, SAMPLE_QUEUE>::BYTES_PER_WARP + POOL_WARPS * POOL_WARP_SCRATCH + 440 * sizeof(double);
}

template <int POOL_SLOTS, int MIN_BLOCKS, bool SAMPLE_QUEUE, int POOL_WARPS = 8, bool HAS_TRI = true>
static cudaError_t launch_pool_variant(const RenderParams& p, int sm_count, cudaStream_t stream, uint32_t* out_launches) {
  constexpr int POOL_BLOCK = POOL_WARPS * 32;
  const size_t smem = pool_smem_bytes<POOL_SLOTS, SAMPLE_QUEUE, POOL_WARPS>(p.layout.bytes);
  cudaError_t e = cudaSuccess;
  const uint64_t total_work =
      (uint64_t)p.tail_first + (uint64_t)p.tail_granules * ((p.rays_per_pixel + SAMPLES_PER_ENTRY - 1) / SAMPLES_PER_ENTRY);
  // persistent grid: MIN_BLOCKS CTAs per SM, never more contexts than queue entries can feed
  // queue entries one CTA can hold at once (a sample-queue entry is up to SAMPLES_PER_ENTRY x 32 camera paths)
  const uint64_t entry_paths = GRAN_PIXELS * (SAMPLE_QUEUE ? (p.rays_per_pixel < SAMPLES_PER_ENTRY ? p.rays_per_pixel : SAMPLES_PER_ENTRY) : 1);
  const uint64_t per_block = ((uint64_t)POOL_WARPS * POOL_SLOTS + entry_paths - 1) / entry_paths;
  uint64_t want_blocks = (total_work + per_block - 1) / per_block;
  uint64_t grid = (uint64_t)sm_count * MIN_BLOCKS;
  if (want_blocks < grid) grid = want_blocks ? want_blocks : 1;
  e = cudaFuncSetAttribute(render_pool_kernel<POOL_SLOTS, MIN_BLOCKS, SAMPLE_QUEUE, POOL_WARPS, HAS_TRI>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  render_pool_kernel<POOL_SLOTS, MIN_BLOCKS, SAMPLE_QUEUE, POOL_WARPS, HAS_TRI><<<(unsigned)grid, POOL_BLOCK, smem, stream>>>(p);
  if (out_launches) *out_launches += 1;
  e = cudaGetLastError();
  if (e != cudaSuccess || p.tail_granules == 0) return e;
  const unsigned threads = p.tail_granules * GRAN_PIXELS;
  tail_sum_kernel<<<(threads + 255) / 256, 256, 0, stream>>>(p);
  if (out_launches) *out_launches += 1;
  return cudaGetLastError();
}

// Largest dynamic shared memory one CTA may ask for on the current device (227 KB on sm_100).
static size_t pool_smem_optin() {
  int dev = 0, v = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev) != cudaSuccess)
    return 0;
  return (size_t)v;
}

// Variants.  The kernel is issue/latency bound (ncu: 62 % of issue slots used with 4 warps per scheduler), so
// the wide variants put ONE CTA of 20 (pixel queue: 22-word contexts, 92 registers) or 24 (sample queue: 17-word
// contexts, 80 registers) warps on an SM, all with 64 contexts per warp; they need almost all of the SM's shared
// memory, so scenes whose blob does not fit next to the pools use the 2 x 8-warp variant.  Measured on C2
// (profiles/r01_SUMMARY.md): 2 x 8 warps 13.00 ms, 1 x 20 warps 12.20 ms.
// The kernel is specialised on what the scene holds: without triangles the triangle loop, its divisions and — what
// matters — its registers are compiled out; at the 80-register cap of the 24-warp variant that alone is 6 % (C2 10.82
// -> 10.20 ms: fewer rematerialised values, 2784 -> 2480 instructions, a smaller loop in the instruction cache).
template <int POOL_SLOTS, int MIN_BLOCKS, bool SAMPLE_QUEUE, int POOL_WARPS>
static cudaError_t launch_pool_for_scene(const RenderParams& p, int sm_count, cudaStream_t stream, uint32_t* out_launches) {
  return p.layout.n_tri == 0 ? launch_pool_variant<POOL_SLOTS, MIN_BLOCKS, SAMPLE_QUEUE, POOL_WARPS, false>(p, sm_count, stream, out_launches)
                             : launch_pool_variant<POOL_SLOTS, MIN_BLOCKS, SAMPLE_QUEUE, POOL_WARPS, true>(p, sm_count, stream, out_launches);
}

cudaError_t launch_render_pool(const RenderParams& p, int sm_count, cudaStream_t stream, uint32_t* out_launches) {
  // RTX_POOL_VARIANT (measurement builds only): 0 = auto; 1 = 2 CTAs x 8 warps; 2 = 1 x 20 warps; 3 = 1 x 24 warps (sample queue only)
#ifdef RTX_MEASUREMENT_KNOBS
  static const int variant = [] { const char* v = getenv("RTX_POOL_VARIANT"); return v ? atoi(v) : 0; }();
#else
  constexpr int variant = 0;
#endif
  static const size_t optin = pool_smem_optin();
  const uint32_t blob = p.layout.bytes;
  if (p.tail_granules != 0) {  // sample queue
    if ((variant == 0 || variant == 3) && pool_smem_bytes<64, true, 24>(blob) <= optin)
      return launch_pool_for_scene<64, 1, true, 24>(p, sm_count, stream, out_launches);
    if ((variant == 0 || variant == 2) && pool_smem_bytes<64, true, 20>(blob) <= optin)
      return launch_pool_for_scene<64, 1, true, 20>(p, sm_count, stream, out_launches);
    return launch_pool_for_scene<64, 2, true, 8>(p, sm_count, stream, out_launches);
  }
  if ((variant == 0 || variant == 2) && pool_smem_bytes<64, false, 20>(blob) <= optin)
    return launch_pool_for_scene<64, 1, false, 20>(p, sm_count, stream, out_launches);
  return launch_pool_for_scene<64, 2, false, 8>(p, sm_count, stream, out_launches);
}

}  // namespace rtx
