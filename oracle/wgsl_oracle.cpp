// wgsl_oracle.cpp — TEST INFRASTRUCTURE ONLY: a CPU restatement, in f32, of the arithmetic of the reference's wgpu
// backend — raytracing/gpu/base_shader.wgsl plus the per-shape WGSL snippets (object/sphere.rs:50-86,
// object/plane.rs:48-85, object/triangle.rs:138-194) that gpu_state.rs:208-312 pastes into it.  It is the checker of
// the product's "preview" kernel (rust-raytracing_b200/csrc/rtx_preview.cu); only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline may call it, never the product path.
//
// PARITY UNPINNED: executing the shader itself (wgpu / Vulkan / GL) is unavailable on this box, and the reference
// holds no golden image of it, so nothing reference-held pins this file — it pins the CUDA kernel to one reading of
// the shader, written down here statement by statement with the WGSL line beside it.
//
// Where WGSL leaves the arithmetic to the implementation, this is what both this file and the kernel do:
//  * every +, -, *, / and sqrt is IEEE-754 binary32, round to nearest, never fused (-ffp-contract=off here);
//    dot(a,b) = (a.x*b.x + a.y*b.y) + a.z*b.z;  length(v) = sqrt(dot(v,v));  normalize(v) = v / length(v);
//    cross as written in the WGSL spec;  M * v = (col0*v.x + col1*v.y) + col2*v.z;
//  * sin, cos: the f64 libm value rounded to f32;  min / max: fminf / fmaxf (the non-NaN operand wins);
//  * literals: 1e-100 (sphere.rs:69) is not representable in f32 and becomes 0.0; EPSILON = f32(1e-8);
//    PI = acos(0.0) * 2.0 evaluated in f32 = 3.14159274;
//  * the fragment's interpolated uv at the pixel centre: vs_main (:48-62) sets uv = clip position * (0.5, -0.5), so
//    uv.x = (x + 0.5) / width - 0.5 and uv.y = -(clip y) / 2.  The image this file returns uses the row order of the
//    f64 path, row 0 = BOTTOM of the picture (clip y = -1 + (2 row + 1) / height): uv.y = 0.5 - (row + 0.5) / height;
//  * f32(u32) in random_float rounds to nearest; u32(f32) truncates.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

namespace {

struct V {
  float x, y, z;
};
inline V mk(float x, float y, float z) { return V{x, y, z}; }
inline V operator+(V a, V b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V operator-(V a, V b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V operator*(V a, V b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
inline V operator*(V a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
inline V operator/(V a, float s) { return mk(a.x / s, a.y / s, a.z / s); }
inline V operator-(V a) { return mk(-a.x, -a.y, -a.z); }
inline float dot(V a, V b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline V cross(V a, V b) { return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
inline float length(V a) { return std::sqrt(dot(a, a)); }
inline V normalize(V a) { return a / length(a); }
inline float sin32(float x) { return (float)std::sin((double)x); }
inline float cos32(float x) { return (float)std::cos((double)x); }

// object records as the host hands them over (f32 already): type 0 sphere {pos, radius}, 1 plane {pos, normal},
// 2 triangle {v0, v1, v2}
struct WObject {
  float base[3], roughness, emission[3];
  uint32_t type;
  float shape[9];
};

struct Scene {
  V cam_pos, cam_dir;
  float fov;
  uint32_t rays_per_pixel, max_bounces;
  float focal_length, focal_offset, non_focal_offset;
  const WObject* obj;
  uint64_t n;
};

// base_shader.wgsl:99-142
struct Rng {
  uint32_t seed;
  static uint32_t jenkins(uint32_t x) {  // :115-123
    x += x << 10;
    x ^= x >> 6;
    x += x << 3;
    x ^= x >> 11;
    x += x << 15;
    return x;
  }
  void init(float ux, float uy) {  // :99-114
    const float ax = std::fmin(std::fmax(ux + 0.5f, 0.0f), 0.999999f), ay = std::fmin(std::fmax(uy + 0.5f, 0.0f), 0.999999f);
    const uint32_t ai = (uint32_t)(ax * 4294967296.0f), bi = (uint32_t)(ay * 4294967296.0f);
    uint32_t h = ai ^ (bi * 0x85ebca6bu);
    h = (h ^ (h >> 16)) * 0x85ebca6bu;
    seed = jenkins(h ^ (h >> 13));
  }
  uint32_t next_u32() {  // :124-129
    const uint32_t s = seed * 747796405u + 2891336453u;
    seed = s;
    const uint32_t word = ((s >> ((s >> 28) + 4u)) ^ s) * 277803737u;
    return (word >> 22) ^ word;
  }
  float next_f32() { return (float)next_u32() / 4294967296.0f; }  // :130-132  f32(0xffffffffu) = 2^32
  V direction() {                                                  // :133-142
    const float pi = 1.57079637f * 2.0f;                           // :1  acos(0.0) * 2.0
    const float z = next_f32() * 2.0f - 1.0f;
    const float theta = next_f32() * 2.0f * pi;
    const float r = std::sqrt(1.0f - z * z);
    return mk(r * cos32(theta), r * sin32(theta), z);
  }
};

struct Dist {
  bool hit;
  float d;
};

// object/sphere.rs:52-75
Dist sphere_distance(const float* s, V pos, V dir) {
  const V offset = pos - mk(s[0], s[1], s[2]);
  const V rd = normalize(dir);
  const float a = dot(rd, rd);
  const float b = 2.0f * dot(offset, rd);
  const float c = dot(offset, offset) - s[3] * s[3];
  const float disc = b * b - 4.0f * a * c;
  if (disc < 0.0f) return Dist{false, 0.0f};  // `< 1e-100`: 0 in f32
  return Dist{true, (-b - std::sqrt(disc)) / (2.0f * a)};
}
// object/plane.rs:55-76
Dist plane_distance(const float* s, V pos, V dir) {
  const V offset = pos - mk(s[0], s[1], s[2]);
  const V norm = normalize(mk(s[3], s[4], s[5]));
  const V d = normalize(dir);
  if (dot(d, norm) >= 0.0f || dot(offset, norm) <= 0.0f) return Dist{false, 0.0f};
  const float t = dot(offset, norm) / dot(norm, d);
  const V ip = offset + d * t;
  return Dist{true, length(offset - ip)};
}
// object/triangle.rs:149-186 (Moeller-Trumbore)
Dist triangle_distance(const float* s, V pos, V dir) {
  const float EPSILON = 1e-8f;
  const V v0 = mk(s[0], s[1], s[2]);
  const V r = mk(s[3], s[4], s[5]) - v0, sv = mk(s[6], s[7], s[8]) - v0;
  const V p = pos - v0;
  const V normal = cross(r, sv);
  if (dot(normalize(dir), normal) == 0.0f) return Dist{false, 0.0f};
  // (:159-161 flips `normal`, which nothing reads afterwards)
  const V pvec = cross(dir, sv);
  const float det = dot(r, pvec);
  if (std::fabs(det) < EPSILON) return Dist{false, 0.0f};
  const float inv_det = 1.0f / det;
  const float u = dot(p, pvec) * inv_det;
  if (u < 0.0f || u > 1.0f) return Dist{false, 0.0f};
  const V qvec = cross(p, r);
  const float v = dot(dir, qvec) * inv_det;
  if (v < 0.0f || (u + v) > 1.0f) return Dist{false, 0.0f};
  const float t = dot(sv, qvec) * inv_det;
  if (t > EPSILON) return Dist{true, t};
  return Dist{false, 0.0f};
}
Dist calculate_distance(const WObject& o, V pos, V dir) {
  return o.type == 0 ? sphere_distance(o.shape, pos, dir) : (o.type == 1 ? plane_distance(o.shape, pos, dir) : triangle_distance(o.shape, pos, dir));
}
// normal_calculation_code: sphere.rs:76-78, plane.rs:77-79 (NOT normalised), triangle.rs:142-148
V calculate_normal(const WObject& o, V world) {
  const float* s = o.shape;
  if (o.type == 0) return normalize(world - mk(s[0], s[1], s[2]));
  if (o.type == 1) return mk(s[3], s[4], s[5]);
  const V v0 = mk(s[0], s[1], s[2]);
  return normalize(cross(mk(s[3], s[4], s[5]) - v0, mk(s[6], s[7], s[8]) - v0));
}
// bounding_box_code (sphere.rs:82-86, plane.rs:83-85, triangle.rs:190-194) + base_shader.wgsl:212-224
bool bounding_box_hit(const WObject& o, V pos, V dir) {
  const float* s = o.shape;
  V lo, hi;
  if (o.type == 0) {
    lo = mk(s[0] - s[3], s[1] - s[3], s[2] - s[3]);
    hi = mk(s[0] + s[3], s[1] + s[3], s[2] + s[3]);
  } else if (o.type == 1) {
    return true;
  } else {
    lo = mk(std::fmin(std::fmin(s[0], s[3]), s[6]), std::fmin(std::fmin(s[1], s[4]), s[7]), std::fmin(std::fmin(s[2], s[5]), s[8]));
    hi = mk(std::fmax(std::fmax(s[0], s[3]), s[6]), std::fmax(std::fmax(s[1], s[4]), s[7]), std::fmax(std::fmax(s[2], s[5]), s[8]));
  }
  const V inv = mk(1.0f / dir.x, 1.0f / dir.y, 1.0f / dir.z);
  const V t0 = (lo - pos) * inv, t1 = (hi - pos) * inv;
  const float tmin = std::fmax(std::fmax(std::fmin(t0.x, t1.x), std::fmin(t0.y, t1.y)), std::fmin(t0.z, t1.z));
  const float tmax = std::fmin(std::fmin(std::fmax(t0.x, t1.x), std::fmax(t0.y, t1.y)), std::fmax(t0.z, t1.z));
  return tmax >= std::fmax(tmin, 0.0f);
}

// base_shader.wgsl:190-211
bool closest_object(const Scene& sc, V pos, V dir, uint64_t* index, float* distance) {
  bool did = false;
  float best = -1.0f;
  for (uint64_t i = 0; i < sc.n; ++i) {
    const WObject& cur = sc.obj[i];
    if (!bounding_box_hit(cur, pos, dir)) continue;
    const Dist d = calculate_distance(cur, pos, dir);
    if (!d.hit) continue;
    if (d.d <= 0.0f) continue;
    if (!did || d.d < best) did = true, best = d.d, *index = i;
  }
  *distance = best;
  return did;
}
// :170-183
V random_bounce(Rng& g, V ray_dir, V n, float roughness) {
  const V rd = g.direction();
  const V refl = ray_dir - (n * 2.0f) * dot(ray_dir, n);
  const V to_refl = refl - rd;
  const float mult = 1.0f - roughness;
  V fin = rd + to_refl * mult;
  fin = normalize(fin);
  return dot(fin, n) < 0.0f ? -fin : fin;
}
// :151-169
V trace_ray(const Scene& sc, Rng& g, V pos, V dir, uint64_t* rays, uint32_t* first_hit) {
  V light = mk(1.0f, 1.0f, 1.0f), color = mk(0.0f, 0.0f, 0.0f);
  for (uint32_t i = 0; i < sc.max_bounces; ++i) {
    uint64_t idx = 0;
    float dist = 0.0f;
    const bool hit = closest_object(sc, pos, dir, &idx, &dist);
    ++*rays;
    if (i == 0 && first_hit) *first_hit = hit ? (uint32_t)idx : 0xFFFFFFFFu;
    if (!hit) break;
    const WObject& o = sc.obj[idx];
    pos = pos + dir * dist;
    color = color + mk(o.emission[0], o.emission[1], o.emission[2]) * light;
    light = light * mk(std::fmax(o.base[0], 0.0f), std::fmax(o.base[1], 0.0f), std::fmax(o.base[2], 0.0f));
    if (light.x == 0.0f && light.y == 0.0f && light.z == 0.0f) break;
    const V n = calculate_normal(o, pos);
    dir = random_bounce(g, dir, n, o.roughness);
  }
  return color;
}

// :228-267: 1/det * adjugate, columns (A,D,G) (B,E,H) (C,F,I); m = columns (right, up, forward)
void inverse3x3(const V c0, const V c1, const V c2, V* o0, V* o1, V* o2) {
  const float a = c0.x, d = c0.y, g = c0.z, b = c1.x, e = c1.y, h = c1.z, c = c2.x, f = c2.y, i = c2.z;
  const float det = a * (e * i - f * h) - b * (d * i - f * g) + c * (d * h - e * g);
  const float A = e * i - f * h, B = f * g - d * i, C = d * h - e * g;
  const float D = c * h - b * i, E = a * i - c * g, F = b * g - a * h;
  const float G = b * f - c * e, H = c * d - a * f, I = a * e - b * d;
  const float s = 1.0f / det;
  *o0 = mk(s * A, s * D, s * G);
  *o1 = mk(s * B, s * E, s * H);
  *o2 = mk(s * C, s * F, s * I);
}

// fs_main, :64-91
void render_pixel(const Scene& sc, uint32_t x, uint32_t row, uint32_t width, uint32_t height, float* out, uint32_t* id,
                  uint64_t* rays) {
  const float ux = ((float)x + 0.5f) / (float)width - 0.5f;
  const float uy = 0.5f - ((float)row + 0.5f) / (float)height;
  Rng g;
  g.init(ux, uy);
  const float aspect = (float)width / (float)height;
  const float ax = ux * sc.fov, ay = uy * (sc.fov / aspect);
  const V cam_space_dir = mk(sin32(ax), sin32(ay), cos32(ax) * cos32(ay));
  const V fwd = normalize(sc.cam_dir);
  const V right = cross(fwd, mk(0.0f, 0.0f, 1.0f));
  const V up = cross(fwd, right);
  V i0, i1, i2;
  inverse3x3(right, up, fwd, &i0, &i1, &i2);
  const V ray_dir = (i0 * cam_space_dir.x + i1 * cam_space_dir.y) + i2 * cam_space_dir.z;
  V color = mk(0.0f, 0.0f, 0.0f);
  const V target_point = sc.cam_pos + ray_dir * sc.focal_length;
  uint32_t first = 0xFFFFFFFFu;
  for (uint32_t i = 0; i < sc.rays_per_pixel; ++i) {
    const V ray_target = target_point + g.direction() * sc.focal_offset;
    const V ray_origin = sc.cam_pos + g.direction() * sc.non_focal_offset;
    const V dir = normalize(ray_target - ray_origin);
    color = color + trace_ray(sc, g, ray_origin, dir, rays, i == 0 ? &first : nullptr);
  }
  color = color / (float)sc.rays_per_pixel;
  out[0] = color.x, out[1] = color.y, out[2] = color.z;
  *id = first;
}

}  // namespace

extern "C" {

uint64_t wgsl_sizeof_object() { return sizeof(WObject); }

// The RNG alone (base_shader.wgsl:99-132): the seed initRngSeed derives from a pixel's uv, then n draws of
// random_int() and — from a second copy of the stream — of random_float().
void wgsl_rng_probe(float ux, float uy, uint32_t n, uint32_t* out_seed, uint32_t* out_u32, float* out_f32) {
  Rng a, b;
  a.init(ux, uy);
  b = a;
  *out_seed = a.seed;
  for (uint32_t i = 0; i < n; ++i) out_u32[i] = a.next_u32(), out_f32[i] = b.next_f32();
}

// cam[7] = pos.xyz, dir.xyz, fov; cfg_u[2] = rays_per_pixel, max_bounces; cfg_f[3] = focal_length, focal_offset,
// non_focal_offset.  out_rgb: float[height][width][3], row 0 = bottom; out_ids: first object hit by sample 0 (or
// 0xFFFFFFFF), may be null; out_rays: closest_object calls.  nthreads <= 0: all hardware threads.
int wgsl_render(const float* cam, const uint32_t* cfg_u, const float* cfg_f, const void* objects, uint64_t n_objects,
                uint32_t width, uint32_t height, int32_t nthreads, float* out_rgb, uint32_t* out_ids, uint64_t* out_rays) {
  Scene sc;
  sc.cam_pos = mk(cam[0], cam[1], cam[2]);
  sc.cam_dir = mk(cam[3], cam[4], cam[5]);
  sc.fov = cam[6];
  sc.rays_per_pixel = cfg_u[0], sc.max_bounces = cfg_u[1];
  sc.focal_length = cfg_f[0], sc.focal_offset = cfg_f[1], sc.non_focal_offset = cfg_f[2];
  sc.obj = static_cast<const WObject*>(objects);
  sc.n = n_objects;
  unsigned nt = nthreads > 0 ? (unsigned)nthreads : std::max(1u, std::thread::hardware_concurrency());
  nt = std::min<unsigned>(nt, height ? height : 1);
  std::atomic<uint32_t> next{0};
  std::atomic<uint64_t> total{0};
  auto work = [&] {
    uint64_t rays = 0;
    for (;;) {
      const uint32_t row = next.fetch_add(1);
      if (row >= height) break;
      for (uint32_t x = 0; x < width; ++x) {
        uint32_t id;
        render_pixel(sc, x, row, width, height, out_rgb + ((size_t)row * width + x) * 3, &id, &rays);
        if (out_ids) out_ids[(size_t)row * width + x] = id;
      }
    }
    total += rays;
  };
  std::vector<std::thread> th;
  for (unsigned t = 1; t < nt; ++t) th.emplace_back(work);
  work();
  for (auto& t : th) t.join();
  if (out_rays) *out_rays = total.load();
  return 0;
}

}  // extern "C"
