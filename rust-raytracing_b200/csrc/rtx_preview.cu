// rtx_preview.cu — the FP32 "preview" mode: the arithmetic of the reference's wgpu backend as a CUDA kernel.
//
// The reference has a second, real-time path: a WGSL fragment shader (raytracing/gpu/base_shader.wgsl, completed by
// the per-shape snippets of object/sphere.rs:50-86, plane.rs:48-85, triangle.rs:138-194 through
// gpu/gpu_state.rs:208-312) whose semantics differ from the CPU path's in every respect that matters to a pixel:
// f32 throughout, a PCG-hash RNG seeded from the pixel's uv (:99-142), target jitter before origin jitter and
// un-normalised random directions (:64-91, :133-142), max_bounces segments instead of max_bounces + 1 (:151-169), a
// slab AABB pre-test per object (:190-224), Moeller-Trumbore triangles, the plane's RAW normal as its shading normal.
// Executing that shader needs wgpu / Vulkan — unavailable on this box — but its arithmetic needs neither: this file
// is that arithmetic, one thread per pixel, checked bit for bit against oracle/wgsl_oracle.cpp (whose header states
// what is fixed where WGSL leaves a choice: IEEE f32 never fused, sin/cos = f64 libm rounded to f32, fminf/fmaxf).
//
// Layout: this is where the north star's "SoA float4" applies.  The scene is nine float4 arrays of n objects each
// (PV_*), everything ray-independent precomputed at upload with the same f32 operations the shader would run per ray:
// edge vectors, cross(r, s), both unit normals, the bounding boxes.  The whole block is staged into shared memory with
// one TMA bulk copy when it fits (<= PV_SMEM_BYTES), and read with 128-bit warp-uniform loads either way.
#include <cuda_runtime.h>
#include <math_constants.h>

#include <chrono>
#include <cmath>
#include <cstring>
#include <vector>

#include "rtx_internal.h"
#include "rtx_sincos.cuh"

namespace rtx {

enum : int { PV_A = 0, PV_B, PV_C, PV_D, PV_E, PV_LO, PV_HI, PV_MAT0, PV_MAT1, PV_FIELDS };
// PV_A   sphere (centre, radius)        plane (position, -)          triangle (v0, -)
// PV_B   -                              (raw normal, -)              (r = v1 - v0, -)
// PV_C   -                              (normalize(normal), -)       (s = v2 - v0, -)
// PV_D   -                              -                            (cross(r, s), -)
// PV_E   -                              -                            (normalize(cross(r, s)), -)
// PV_LO  (box min, type as bits)   PV_HI (box max, -)       — planes have no box (bounding_box_code, plane.rs:83-85)
// PV_MAT0 (max(base_color, 0), roughness)   PV_MAT1 (emission_color, -)
constexpr uint32_t PV_SMEM_BYTES = 96u * 1024u;
constexpr int PV_BLOCK = 128;  // 16 x 8 pixels

struct PreviewParams {
  float cam_pos[3], cam_dir[3], fov;
  uint32_t rays_per_pixel, max_bounces;
  float focal_length, focal_offset, non_focal_offset;
  uint32_t width, height, n_obj, stride;  // stride = objects per field array (n_obj rounded up: 16-byte fields anyway)
  const float4* soa;                      // [PV_FIELDS][stride]
  uint32_t soa_bytes;
  float* out_rgb;
  uint32_t* out_ids;
  unsigned long long* ray_counter;
};

// ---- f32 vector arithmetic, never fused, host + device -----------------------------------------------------------
struct F3 {
  float x, y, z;
};
RTX_HD float pf_add(float a, float b) {
#if defined(__CUDA_ARCH__)
  return __fadd_rn(a, b);
#else
  return a + b;
#endif
}
RTX_HD float pf_sub(float a, float b) {
#if defined(__CUDA_ARCH__)
  return __fsub_rn(a, b);
#else
  return a - b;
#endif
}
RTX_HD float pf_mul(float a, float b) {
#if defined(__CUDA_ARCH__)
  return __fmul_rn(a, b);
#else
  return a * b;
#endif
}
RTX_HD float pf_div(float a, float b) {
#if defined(__CUDA_ARCH__)
  return __fdiv_rn(a, b);
#else
  return a / b;
#endif
}
RTX_HD float pf_sqrt(float a) {
#if defined(__CUDA_ARCH__)
  return __fsqrt_rn(a);
#else
  return std::sqrt(a);
#endif
}
RTX_HD F3 f3(float x, float y, float z) {
  F3 v;
  v.x = x, v.y = y, v.z = z;
  return v;
}
RTX_HD F3 operator+(F3 a, F3 b) { return f3(pf_add(a.x, b.x), pf_add(a.y, b.y), pf_add(a.z, b.z)); }
RTX_HD F3 operator-(F3 a, F3 b) { return f3(pf_sub(a.x, b.x), pf_sub(a.y, b.y), pf_sub(a.z, b.z)); }
RTX_HD F3 operator*(F3 a, F3 b) { return f3(pf_mul(a.x, b.x), pf_mul(a.y, b.y), pf_mul(a.z, b.z)); }
RTX_HD F3 operator*(F3 a, float s) { return f3(pf_mul(a.x, s), pf_mul(a.y, s), pf_mul(a.z, s)); }
RTX_HD F3 operator/(F3 a, float s) { return f3(pf_div(a.x, s), pf_div(a.y, s), pf_div(a.z, s)); }
RTX_HD F3 operator-(F3 a) { return f3(-a.x, -a.y, -a.z); }
RTX_HD float pdot(F3 a, F3 b) { return pf_add(pf_add(pf_mul(a.x, b.x), pf_mul(a.y, b.y)), pf_mul(a.z, b.z)); }
RTX_HD F3 pcross(F3 a, F3 b) {
  return f3(pf_sub(pf_mul(a.y, b.z), pf_mul(a.z, b.y)), pf_sub(pf_mul(a.z, b.x), pf_mul(a.x, b.z)),
            pf_sub(pf_mul(a.x, b.y), pf_mul(a.y, b.x)));
}
RTX_HD float plength(F3 a) { return pf_sqrt(pdot(a, a)); }
RTX_HD F3 pnormalize(F3 a) { return a / plength(a); }

#if defined(__CUDACC__)
__device__ __forceinline__ F3 xyz(float4 v) { return f3(v.x, v.y, v.z); }
// sin / cos: the f64 libm value (glibc's, restated bit for bit in rtx_sincos.cuh), rounded to f32
__device__ __forceinline__ void sincos32(float x, const double* tab, float* s, float* c) {
  double ds, dc;
  libm_sincos((double)x, tab, &ds, &dc);
  *s = __double2float_rn(ds);
  *c = __double2float_rn(dc);
}

// base_shader.wgsl:99-142
struct PcgRng {
  uint32_t seed;
  __device__ __forceinline__ void init(float ux, float uy) {
    const float ax = fminf(fmaxf(pf_add(ux, 0.5f), 0.0f), 0.999999f), ay = fminf(fmaxf(pf_add(uy, 0.5f), 0.0f), 0.999999f);
    const uint32_t ai = __float2uint_rz(pf_mul(ax, 4294967296.0f)), bi = __float2uint_rz(pf_mul(ay, 4294967296.0f));
    uint32_t h = ai ^ (bi * 0x85ebca6bu);
    h = (h ^ (h >> 16)) * 0x85ebca6bu;
    uint32_t x = h ^ (h >> 13);  // jenkinsHash, :115-123
    x += x << 10;
    x ^= x >> 6;
    x += x << 3;
    x ^= x >> 11;
    x += x << 15;
    seed = x;
  }
  __device__ __forceinline__ float next_f32() {
    const uint32_t s = seed * 747796405u + 2891336453u;
    seed = s;
    const uint32_t word = ((s >> ((s >> 28) + 4u)) ^ s) * 277803737u;
    return pf_div(__uint2float_rn((word >> 22) ^ word), 4294967296.0f);
  }
  __device__ __forceinline__ F3 direction(const double* tab) {
    const float pi = pf_mul(1.57079637f, 2.0f);
    const float z = pf_sub(pf_mul(next_f32(), 2.0f), 1.0f);
    const float theta = pf_mul(pf_mul(next_f32(), 2.0f), pi);
    const float r = pf_sqrt(pf_sub(1.0f, pf_mul(z, z)));
    float s, c;
    sincos32(theta, tab, &s, &c);
    return f3(pf_mul(r, c), pf_mul(r, s), z);
  }
};

struct PvScene {
  const float4* f[PV_FIELDS];
  uint32_t n;
};

// base_shader.wgsl:190-224 with the per-shape distance snippets; returns the object index or MISS_ID
__device__ __forceinline__ uint32_t pv_closest(const PvScene& sc, F3 pos, F3 dir, float* out_d) {
  uint32_t best_i = MISS_ID;
  float best = -1.0f;
  // hoisted out of the object loop (the shader recomputes them per object, same operations on the same values)
  const F3 inv = f3(pf_div(1.0f, dir.x), pf_div(1.0f, dir.y), pf_div(1.0f, dir.z));  // :216
  const F3 dn = pnormalize(dir);  // sphere.rs:64, plane.rs:68, triangle.rs:156
  const float a = pdot(dn, dn);   // sphere.rs:65
  for (uint32_t i = 0; i < sc.n; ++i) {
    const float4 lo = sc.f[PV_LO][i];
    const uint32_t type = __float_as_uint(lo.w);
    if (type != 1u) {  // :212-224
      const float4 hi = sc.f[PV_HI][i];
      const F3 t0 = (xyz(lo) - pos) * inv, t1 = (xyz(hi) - pos) * inv;
      const float tmin = fmaxf(fmaxf(fminf(t0.x, t1.x), fminf(t0.y, t1.y)), fminf(t0.z, t1.z));
      const float tmax = fminf(fminf(fmaxf(t0.x, t1.x), fmaxf(t0.y, t1.y)), fmaxf(t0.z, t1.z));
      if (!(tmax >= fmaxf(tmin, 0.0f))) continue;
    }
    const float4 A = sc.f[PV_A][i];
    float d;
    if (type == 0u) {  // sphere.rs:52-75
      const F3 offset = pos - xyz(A);
      const float b = pf_mul(2.0f, pdot(offset, dn));
      const float c = pf_sub(pdot(offset, offset), pf_mul(A.w, A.w));
      const float disc = pf_sub(pf_mul(b, b), pf_mul(pf_mul(4.0f, a), c));
      if (disc < 0.0f) continue;
      d = pf_div(pf_sub(-b, pf_sqrt(disc)), pf_mul(2.0f, a));
    } else if (type == 1u) {  // plane.rs:55-76
      const F3 offset = pos - xyz(A);
      const F3 norm = xyz(sc.f[PV_C][i]);
      if (pdot(dn, norm) >= 0.0f || pdot(offset, norm) <= 0.0f) continue;
      const float t = pf_div(pdot(offset, norm), pdot(norm, dn));
      const F3 ip = offset + dn * t;
      d = plength(offset - ip);
    } else {  // triangle.rs:149-186
      const F3 r = xyz(sc.f[PV_B][i]), s = xyz(sc.f[PV_C][i]);
      const F3 p = pos - xyz(A);
      if (pdot(dn, xyz(sc.f[PV_D][i])) == 0.0f) continue;
      const F3 pvec = pcross(dir, s);
      const float det = pdot(r, pvec);
      if (fabsf(det) < 1e-8f) continue;
      const float inv_det = pf_div(1.0f, det);
      const float u = pf_mul(pdot(p, pvec), inv_det);
      if (u < 0.0f || u > 1.0f) continue;
      const F3 qvec = pcross(p, r);
      const float v = pf_mul(pdot(dir, qvec), inv_det);
      if (v < 0.0f || pf_add(u, v) > 1.0f) continue;
      d = pf_mul(pdot(s, qvec), inv_det);
      if (!(d > 1e-8f)) continue;
    }
    if (d <= 0.0f) continue;  // (a NaN distance passes, as in the shader: `NaN <= 0` is false)
    if (best_i == MISS_ID || d < best) best = d, best_i = i;
  }
  *out_d = best;
  return best_i;
}

__global__ void __launch_bounds__(PV_BLOCK) preview_kernel(const __grid_constant__ PreviewParams p, int smem_scene) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const float4* base = p.soa;
  if (smem_scene) {  // one TMA bulk copy of the nine arrays
    uint64_t* mbar = reinterpret_cast<uint64_t*>(smem_raw);
    stage_blob(smem_raw + 128, reinterpret_cast<const unsigned char*>(p.soa), p.soa_bytes, mbar);
    base = reinterpret_cast<const float4*>(smem_raw + 128);
  }
  PvScene sc;
  for (int k = 0; k < PV_FIELDS; ++k) sc.f[k] = base + (size_t)k * p.stride;
  sc.n = p.n_obj;
  const double* tab = sincos_table();

  const uint32_t tiles_x = (p.width + 15u) / 16u;
  const uint32_t tx = blockIdx.x % tiles_x, ty = blockIdx.x / tiles_x;
  const uint32_t x = tx * 16u + (threadIdx.x & 15u), row = ty * 8u + (threadIdx.x >> 4);
  uint32_t n_rays = 0;
  if (x < p.width && row < p.height) {
    // fs_main, :64-91 (uv as the rasteriser interpolates it at the pixel centre; row 0 = bottom of the picture)
    const float ux = pf_sub(pf_div(pf_add((float)x, 0.5f), (float)p.width), 0.5f);
    const float uy = pf_sub(0.5f, pf_div(pf_add((float)row, 0.5f), (float)p.height));
    PcgRng g;
    g.init(ux, uy);
    const float aspect = pf_div((float)p.width, (float)p.height);
    const float ax = pf_mul(ux, p.fov), ay = pf_mul(uy, pf_div(p.fov, aspect));
    float sx, cx, sy, cy;
    sincos32(ax, tab, &sx, &cx);
    sincos32(ay, tab, &sy, &cy);
    const F3 csd = f3(sx, sy, pf_mul(cx, cy));
    const F3 cam_pos = f3(p.cam_pos[0], p.cam_pos[1], p.cam_pos[2]);
    const F3 fwd = pnormalize(f3(p.cam_dir[0], p.cam_dir[1], p.cam_dir[2]));
    const F3 right = pcross(fwd, f3(0.0f, 0.0f, 1.0f));
    const F3 up = pcross(fwd, right);
    // inverse3x3 (:228-267) of the matrix whose COLUMNS are right, up, forward
    const float ma = right.x, md = right.y, mg = right.z, mb = up.x, me = up.y, mh = up.z, mc = fwd.x, mf = fwd.y, mi = fwd.z;
    const float ei_fh = pf_sub(pf_mul(me, mi), pf_mul(mf, mh)), di_fg = pf_sub(pf_mul(md, mi), pf_mul(mf, mg)),
                dh_eg = pf_sub(pf_mul(md, mh), pf_mul(me, mg));
    const float det = pf_add(pf_sub(pf_mul(ma, ei_fh), pf_mul(mb, di_fg)), pf_mul(mc, dh_eg));
    const float A = ei_fh, B = pf_sub(pf_mul(mf, mg), pf_mul(md, mi)), C = dh_eg;
    const float D = pf_sub(pf_mul(mc, mh), pf_mul(mb, mi)), E = pf_sub(pf_mul(ma, mi), pf_mul(mc, mg)),
                F = pf_sub(pf_mul(mb, mg), pf_mul(ma, mh));
    const float G = pf_sub(pf_mul(mb, mf), pf_mul(mc, me)), H = pf_sub(pf_mul(mc, md), pf_mul(ma, mf)),
                I = pf_sub(pf_mul(ma, me), pf_mul(mb, md));
    const float s = pf_div(1.0f, det);
    const F3 i0 = f3(pf_mul(s, A), pf_mul(s, D), pf_mul(s, G)), i1 = f3(pf_mul(s, B), pf_mul(s, E), pf_mul(s, H)),
             i2 = f3(pf_mul(s, C), pf_mul(s, F), pf_mul(s, I));
    const F3 ray_dir = (i0 * csd.x + i1 * csd.y) + i2 * csd.z;
    const F3 target_point = cam_pos + ray_dir * p.focal_length;
    F3 color = f3(0.0f, 0.0f, 0.0f);
    uint32_t first = MISS_ID;
    // fs_main's sample loop (:79-89) around trace_ray's bounce loop (:151-169), flattened into ONE loop whose trip is one
    // closest_object call: a lane whose path has ended starts its pixel's next sample in the next trip instead of
    // waiting at the inner loop's exit for the longest path of the warp (nested: 16.3 of 32 lanes per instruction).
    // Per lane the operations and the RNG draws are the same, in the same order.
    uint32_t smp = 0, b = 0;
    bool fresh = true;
    F3 pos = cam_pos, dir = ray_dir, light = f3(1.0f, 1.0f, 1.0f), acc = f3(0.0f, 0.0f, 0.0f);
    while (smp < p.rays_per_pixel) {
      if (fresh) {
        const F3 ray_target = target_point + g.direction(tab) * p.focal_offset;
        pos = cam_pos + g.direction(tab) * p.non_focal_offset;
        dir = pnormalize(ray_target - pos);
        light = f3(1.0f, 1.0f, 1.0f), acc = f3(0.0f, 0.0f, 0.0f);
        b = 0;
        fresh = false;
      }
      bool end = b >= p.max_bounces;
      if (!end) {
        float dist;
        const uint32_t hit = pv_closest(sc, pos, dir, &dist);
        ++n_rays;
        if (smp == 0 && b == 0) first = hit;
        if (hit == MISS_ID) {
          end = true;
        } else {
          const float4 m0 = sc.f[PV_MAT0][hit], m1 = sc.f[PV_MAT1][hit];
          pos = pos + dir * dist;
          acc = acc + xyz(m1) * light;
          light = light * xyz(m0);
          if (light.x == 0.0f && light.y == 0.0f && light.z == 0.0f) {
            end = true;
          } else {
            // calculate_normal: sphere.rs:76-78, plane.rs:77-79 (the RAW normal), triangle.rs:142-148
            const uint32_t type = __float_as_uint(sc.f[PV_LO][hit].w);
            F3 n;
            if (type == 0u) n = pnormalize(pos - xyz(sc.f[PV_A][hit]));
            else if (type == 1u) n = xyz(sc.f[PV_B][hit]);
            else n = xyz(sc.f[PV_E][hit]);
            // random_bounce, :170-183 (drawn even when this was the last segment: the stream goes on into the next sample)
            const F3 rd = g.direction(tab);
            const F3 refl = dir - (n * 2.0f) * pdot(dir, n);
            const F3 fin = pnormalize(rd + (refl - rd) * pf_sub(1.0f, m0.w));
            dir = pdot(fin, n) < 0.0f ? -fin : fin;
            ++b;
          }
        }
      }
      if (end) {
        color = color + acc;
        ++smp;
        fresh = true;
      }
    }
    color = color / (float)p.rays_per_pixel;
    const size_t o = ((size_t)row * p.width + x) * 3;
    p.out_rgb[o] = color.x, p.out_rgb[o + 1] = color.y, p.out_rgb[o + 2] = color.z;
    if (p.out_ids) p.out_ids[(size_t)row * p.width + x] = first;
  }
  if (p.ray_counter) {
    unsigned long long r = n_rays;
    for (int o = 16; o > 0; o >>= 1) r += __shfl_down_sync(0xFFFFFFFFu, r, o);
    if ((threadIdx.x & 31u) == 0u && r) atomicAdd(&p.ray_counter[0], r);
  }
}
#endif  // __CUDACC__

// ---- upload: f64 host records -> f32 (Rust `as f32`, GpuSerialize) -> the nine float4 arrays ----------------------
static void build_preview_soa(const rtx_scene* s, std::vector<float>& out, uint32_t* stride) {
  const uint32_t n = (uint32_t)s->n_obj;
  const uint32_t st = n ? n : 1;
  *stride = st;
  out.assign((size_t)PV_FIELDS * st * 4, 0.0f);
  auto put = [&](int field, uint32_t i, float x, float y, float z, float w) {
    float* d = &out[((size_t)field * st + i) * 4];
    d[0] = x, d[1] = y, d[2] = z, d[3] = w;
  };
  auto bits = [](uint32_t u) {
    float f;
    std::memcpy(&f, &u, 4);
    return f;
  };
  for (uint32_t i = 0; i < n; ++i) {
    const uint32_t ref = s->obj_ref[i], k = ref & REF_INDEX_MASK, kind = ref & ~REF_INDEX_MASK;
    const double* m = &s->mat[(size_t)i * MAT_STRIDE];
    put(PV_MAT0, i, std::fmax((float)m[0], 0.0f), std::fmax((float)m[1], 0.0f), std::fmax((float)m[2], 0.0f), (float)m[6]);
    put(PV_MAT1, i, (float)m[3], (float)m[4], (float)m[5], 0.0f);
    if (kind == REF_SPHERE) {
      const double* r = &s->sph[(size_t)k * SPH_STRIDE];
      const float cx = (float)r[0], cy = (float)r[1], cz = (float)r[2], rad = (float)s->sph_radius[k];
      put(PV_A, i, cx, cy, cz, rad);
      put(PV_LO, i, cx - rad, cy - rad, cz - rad, bits(0u));  // sphere.rs:82-86
      put(PV_HI, i, cx + rad, cy + rad, cz + rad, 0.0f);
    } else if (kind == REF_PLANE) {
      const double* r = &s->pl[(size_t)k * PL_STRIDE];
      const F3 nraw = f3((float)r[3], (float)r[4], (float)r[5]);
      const F3 nh = pnormalize(nraw);  // plane.rs:67
      put(PV_A, i, (float)r[0], (float)r[1], (float)r[2], 0.0f);
      put(PV_B, i, nraw.x, nraw.y, nraw.z, 0.0f);
      put(PV_C, i, nh.x, nh.y, nh.z, 0.0f);
      put(PV_LO, i, 0.0f, 0.0f, 0.0f, bits(1u));
    } else {
      const double* v = &s->tri_raw[(size_t)k * 9];
      const F3 v0 = f3((float)v[0], (float)v[1], (float)v[2]), v1 = f3((float)v[3], (float)v[4], (float)v[5]),
               v2 = f3((float)v[6], (float)v[7], (float)v[8]);
      const F3 r = v1 - v0, sv = v2 - v0, nrm = pcross(r, sv), nh = pnormalize(nrm);
      put(PV_A, i, v0.x, v0.y, v0.z, 0.0f);
      put(PV_B, i, r.x, r.y, r.z, 0.0f);
      put(PV_C, i, sv.x, sv.y, sv.z, 0.0f);
      put(PV_D, i, nrm.x, nrm.y, nrm.z, 0.0f);
      put(PV_E, i, nh.x, nh.y, nh.z, 0.0f);
      put(PV_LO, i, std::fmin(std::fmin(v0.x, v1.x), v2.x), std::fmin(std::fmin(v0.y, v1.y), v2.y),
          std::fmin(std::fmin(v0.z, v1.z), v2.z), bits(2u));  // triangle.rs:190-194
      put(PV_HI, i, std::fmax(std::fmax(v0.x, v1.x), v2.x), std::fmax(std::fmax(v0.y, v1.y), v2.y),
          std::fmax(std::fmax(v0.z, v1.z), v2.z), 0.0f);
    }
  }
}

static rtx_status preview_launch(rtx_scene* s, uint32_t width, uint32_t height, float* d_rgb, uint32_t* d_ids,
                                 unsigned long long* d_counters, cudaStream_t stream) {
  if (width == 0 || height == 0) return fail(s, RTX_ERR_INVALID_ARGUMENT, "width and height must be non-zero");
  if ((uint64_t)width * height > 0xFFFFFFFFull) return fail(s, RTX_ERR_INVALID_ARGUMENT, "image too large");
  if (s->config.rays_per_pixel > 0xFFFFFFFFull || s->config.max_bounces > 0xFFFFFFFFull)
    return fail(s, RTX_ERR_INVALID_ARGUMENT, "the f32 preview takes rays_per_pixel and max_bounces as u32 (scene.rs:67-76)");
  if (s->n_obj > 0x7FFFFFFFull) return fail(s, RTX_ERR_INVALID_ARGUMENT, "too many objects");
  // the preview's copy of the scene follows the object list (rebuilt whenever it changed since the last preview)
  if (!s->pv_valid || !s->d_pv) {
    std::vector<float> soa;
    uint32_t stride = 1;
    build_preview_soa(s, soa, &stride);
    const size_t bytes = soa.size() * sizeof(float);
    if (s->d_pv_cap < bytes || !s->d_pv) {
      if (s->d_pv) cudaFree(s->d_pv);
      s->d_pv = nullptr, s->d_pv_cap = 0;
      RTX_CUDA_TRY(s, cudaMalloc((void**)&s->d_pv, bytes));
      s->d_pv_cap = bytes;
    }
    RTX_CUDA_TRY(s, cudaMemcpyAsync(s->d_pv, soa.data(), bytes, cudaMemcpyHostToDevice, stream));
    RTX_CUDA_TRY(s, cudaStreamSynchronize(stream));  // pageable source
    s->pv_stride = stride, s->pv_bytes = (uint32_t)bytes, s->pv_valid = true;
  }
  PreviewParams p{};
  p.cam_pos[0] = (float)s->camera.position.x, p.cam_pos[1] = (float)s->camera.position.y, p.cam_pos[2] = (float)s->camera.position.z;
  p.cam_dir[0] = (float)s->camera.direction.x, p.cam_dir[1] = (float)s->camera.direction.y, p.cam_dir[2] = (float)s->camera.direction.z;
  p.fov = (float)s->camera.fov;  // camera.rs:70-78
  p.rays_per_pixel = (uint32_t)s->config.rays_per_pixel, p.max_bounces = (uint32_t)s->config.max_bounces;
  p.focal_length = (float)s->config.focal_length, p.focal_offset = (float)s->config.focal_offset;
  p.non_focal_offset = (float)s->config.non_focal_offset;
  p.width = width, p.height = height, p.n_obj = (uint32_t)s->n_obj, p.stride = s->pv_stride;
  p.soa = reinterpret_cast<const float4*>(s->d_pv);
  p.soa_bytes = s->pv_bytes;
  p.out_rgb = d_rgb, p.out_ids = d_ids, p.ray_counter = d_counters;
  const int smem_scene = p.soa_bytes <= PV_SMEM_BYTES ? 1 : 0;
  const size_t smem = smem_scene ? 128 + (size_t)p.soa_bytes : 0;
  if (smem > 48 * 1024)
    RTX_CUDA_TRY(s, cudaFuncSetAttribute(preview_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const unsigned grid = ((width + 15u) / 16u) * ((height + 7u) / 8u);
  preview_kernel<<<grid, PV_BLOCK, smem, stream>>>(p, smem_scene);
  RTX_CUDA_TRY(s, cudaGetLastError());
  return RTX_OK;
}

}  // namespace rtx

using namespace rtx;

extern "C" {

rtx_status rtx_scene_render_preview_f32_device(rtx_scene* s, uint32_t width, uint32_t height, void* d_out_rgb,
                                               void* d_out_first_hit_id, void* d_out_counters, void* cuda_stream) {
  if (!s) return fail(s, RTX_ERR_INVALID_ARGUMENT, "null scene");
  if (!d_out_rgb) return fail(s, RTX_ERR_INVALID_ARGUMENT, "d_out_rgb is null");
  std::lock_guard<std::mutex> lk(s->mu);
  RTX_CUDA_TRY(s, cudaSetDevice(s->device));
  return preview_launch(s, width, height, (float*)d_out_rgb, (uint32_t*)d_out_first_hit_id, (unsigned long long*)d_out_counters,
                        (cudaStream_t)cuda_stream);
}

rtx_status rtx_scene_render_preview_f32(rtx_scene* s, uint32_t width, uint32_t height, float* out_rgb,
                                        uint32_t* out_first_hit_id, rtx_stats* out_stats) {
  if (!s) return fail(s, RTX_ERR_INVALID_ARGUMENT, "null scene");
  if (!out_rgb) return fail(s, RTX_ERR_INVALID_ARGUMENT, "out_rgb is null");
  const auto t0 = std::chrono::steady_clock::now();
  std::lock_guard<std::mutex> lk(s->mu);
  RTX_CUDA_TRY(s, cudaSetDevice(s->device));
  const size_t npix = (size_t)width * height;
  // the f64 path's device frame is large enough for an f32 frame of the same size
  if (s->d_rgb_cap < npix * 3 * sizeof(float) || !s->d_rgb) {
    if (s->d_rgb) cudaFree(s->d_rgb);
    s->d_rgb = nullptr, s->d_rgb_cap = 0;
    RTX_CUDA_TRY(s, cudaMalloc((void**)&s->d_rgb, npix * 3 * sizeof(double)));
    s->d_rgb_cap = npix * 3 * sizeof(double);
  }
  if (out_first_hit_id && (s->d_ids_cap < npix * 4 || !s->d_ids)) {
    if (s->d_ids) cudaFree(s->d_ids);
    s->d_ids = nullptr, s->d_ids_cap = 0;
    RTX_CUDA_TRY(s, cudaMalloc((void**)&s->d_ids, npix * 4));
    s->d_ids_cap = npix * 4;
  }
  // the scene's device frame is shared with the f64 entry points: wait for whatever they last enqueued elsewhere
  if (s->last_valid && s->last_stream != s->stream && s->ev_last) RTX_CUDA_TRY(s, cudaStreamWaitEvent(s->stream, s->ev_last, 0));
  RTX_CUDA_TRY(s, cudaMemsetAsync(s->d_counters, 0, 8 * sizeof(unsigned long long), s->stream));
  RTX_CUDA_TRY(s, cudaEventRecord(s->ev0, s->stream));
  const rtx_status st = preview_launch(s, width, height, reinterpret_cast<float*>(s->d_rgb), out_first_hit_id ? s->d_ids : nullptr,
                                       s->d_counters, s->stream);
  if (st != RTX_OK) return st;
  RTX_CUDA_TRY(s, cudaEventRecord(s->ev1, s->stream));
  RTX_CUDA_TRY(s, cudaMemcpyAsync(out_rgb, s->d_rgb, npix * 3 * sizeof(float), cudaMemcpyDeviceToHost, s->stream));
  if (out_first_hit_id)
    RTX_CUDA_TRY(s, cudaMemcpyAsync(out_first_hit_id, s->d_ids, npix * 4, cudaMemcpyDeviceToHost, s->stream));
  unsigned long long counters[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  RTX_CUDA_TRY(s, cudaMemcpyAsync(counters, s->d_counters, sizeof(counters), cudaMemcpyDeviceToHost, s->stream));
  RTX_CUDA_TRY(s, cudaStreamSynchronize(s->stream));
  if (out_stats) {
    std::memset(out_stats, 0, sizeof(*out_stats));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, s->ev0, s->ev1);
    out_stats->rays = counters[0];
    out_stats->paths = (uint64_t)npix * s->config.rays_per_pixel;
    out_stats->pixels = npix;
    out_stats->device_ms = ms;
    out_stats->kernel_launches = 1;
    out_stats->accel_used = RTX_ACCEL_BRUTE;
    out_stats->prims_tested = counters[0] * s->n_obj;
    out_stats->total_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
  }
  return RTX_OK;
}

}  // extern "C"
