// rtx_oracle.cpp — CPU restatement of the rust-raytracing (`rtx` crate) CPU render path.
//
// *** TEST INFRASTRUCTURE ONLY. ***  Nothing in the product path (rust-raytracing_b200/, include/) may
// import, link or execute this file.  Allowed users: tests/, __graft_entry__.smoke(), and bench.py's
// cpu_baseline / --impl reference legs.
//
// PARITY STATUS: "parity unpinned" for every intersection/shading routine and for the f64 mantissa fill of the
// RNG — the reference ships no golden vectors for them.  Pinned: the 12 exact assertions of the
// 4 camera tests (src/raytracing/camera.rs:79-110), which tests/test_oracle.py replays against this file, and — since
// round 2 — the RNG's u64 stream, against the machine code of fastrand 2.3.0 (the reference's Cargo.lock pin) as
// compiled into wheels of this image (tests/test_fastrand_pin.py).  The reference (Rust) cannot be compiled in this
// image (no cargo/rustc), so fidelity is argued line by line: every function cites the reference lines it
// restates, and keeps the reference's exact IEEE-754 f64 evaluation order.
//
// Build: g++ -O2 -std=c++17 -ffp-contract=off -fno-fast-math -shared -fPIC (see oracle/Makefile).
// -ffp-contract=off matters: Rust never fuses a*b+c, so neither may we.
//
// Third-party arithmetic not vendored in the reference: fastrand 2.3.0 (Cargo.lock:338-341).  Restated
// from the published crate: WyRand with the wyhash v4.2 constants, f64() = 52-bit mantissa fill.  The
// reference never seeds it (thread-local, time-seeded), so a seeding convention is ADDED here
// (stream_state below); it is the only thing on the path that is not dictated by the reference.

#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <thread>
#include <utility>
#include <vector>

namespace {

// ---------------------------------------------------------------------------------------------
// math/vector.rs:12-20  — Vector3 {x,y,z: f64}
// ---------------------------------------------------------------------------------------------
struct V3 {
  double x, y, z;
};
// vector/add.rs:17-26, sub.rs:17-26: component-wise
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
// vector/mul.rs:11-21 (scalar broadcast), :22-31 (component-wise)
inline V3 operator*(V3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }
inline V3 operator*(V3 a, V3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
// vector/div.rs:11-21: three divisions, never a reciprocal multiply
inline V3 operator/(V3 a, double s) { return {a.x / s, a.y / s, a.z / s}; }
// math/vector.rs:115-120
inline V3 operator-(V3 a) { return {-a.x, -a.y, -a.z}; }
// math/vector.rs:85-87: (x*x' + y*y') + z*z'
inline double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// math/vector.rs:89-95
inline V3 cross(V3 a, V3 b) {
  return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
// math/vector.rs:97-103: len = sqrt(((x*x)+(y*y))+(z*z))
inline double vsum(V3 a) { return a.x + a.y + a.z; }
inline double vlen(V3 a) { return std::sqrt(vsum(a * a)); }
// math/vector.rs:105-107
inline V3 vnorm(V3 a) { return a / vlen(a); }
// derive(PartialEq): -0.0 == 0.0, NaN != NaN
inline bool veq(V3 a, V3 b) { return a.x == b.x && a.y == b.y && a.z == b.z; }

// ---------------------------------------------------------------------------------------------
// math/mat.rs:10-18 — three ROW vectors
// ---------------------------------------------------------------------------------------------
struct M3 {
  V3 x, y, z;
};
// mat/specific_math.rs:16-21
inline M3 transpose(const M3& m) {
  return {{m.x.x, m.y.x, m.z.x}, {m.x.y, m.y.y, m.z.y}, {m.x.z, m.y.z, m.z.z}};
}
// mat/specific_math.rs:23-43 (Sarrus; each triple product left to right)
inline double determinant(const M3& m) {
  double sum1 = m.x.x * m.y.y * m.z.z;
  double sum2 = m.x.y * m.y.z * m.z.x;
  double sum3 = m.x.z * m.y.x * m.z.y;
  double sub1 = m.z.x * m.y.y * m.x.z;
  double sub2 = m.z.y * m.y.z * m.x.x;
  double sub3 = m.z.z * m.y.x * m.x.y;
  return (sum1 + sum2 + sum3) - (sub1 + sub2 + sub3);
}
// mat/specific_math.rs:45-71
inline M3 adjugate(const M3& m) {
  double a = m.x.x, b = m.x.y, c = m.x.z;
  double d = m.y.x, e = m.y.y, f = m.y.z;
  double g = m.z.x, h = m.z.y, i = m.z.z;
  return {{e * i - f * h, c * h - b * i, b * f - c * e},
          {f * g - d * i, a * i - c * g, c * d - a * f},
          {d * h - e * g, b * g - a * h, a * e - b * d}};
}
// mat/specific_math.rs:10-14 + mat/div.rs:11-21: adjugate / det (9 divisions)
inline M3 inverse(const M3& m) {
  M3 adj = adjugate(m);
  double det = determinant(m);
  return {adj.x / det, adj.y / det, adj.z / det};
}
// mat/mul.rs:42-51: M*v = (v.dot(row_x), v.dot(row_y), v.dot(row_z))
inline V3 matvec(const M3& m, V3 v) { return {dot(v, m.x), dot(v, m.y), dot(v, m.z)}; }

// ---------------------------------------------------------------------------------------------
// fastrand 2.3.0 (third-party; restated) — call sites math/vector.rs:31-33,37-38
// gen_u64 is pinned: tests/test_fastrand_pin.py executes the machine code of this step as compiled into hf_xet 1.5.0
// (whose SBOM lists pkg:cargo/fastrand@2.3.0, the reference's Cargo.lock pin) and gets this stream bit for bit.
// f64()'s mantissa fill is restated from the published source; nothing in this image holds a compiled copy of it.
// ---------------------------------------------------------------------------------------------
struct WyRand {
  uint64_t s;
  inline uint64_t gen_u64() {
    s += 0x2d358dccaa6c78a5ULL;
    unsigned __int128 t = (unsigned __int128)s * (unsigned __int128)(s ^ 0x8bb84b93962eacc9ULL);
    return (uint64_t)t ^ (uint64_t)(t >> 64);
  }
  // f64::from_bits((1<<62) - (1<<52) + (u64 >> 12)) - 1.0   ∈ [0,1)
  inline double f64() {
    uint64_t bits = 0x3FF0000000000000ULL + (gen_u64() >> 12);
    double d;
    std::memcpy(&d, &bits, 8);
    return d - 1.0;
  }
};
// splitmix64 finaliser — the ADDED seeding convention: one stream per key.
inline uint64_t mix64(uint64_t z) {
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}
inline uint64_t stream_state(uint64_t seed, uint64_t key) {
  return mix64(seed + 0x9E3779B97F4A7C15ULL * (key + 1ULL));
}

// math/vector.rs:29-35 — struct-literal field order x, y, z = draw order
inline V3 v_random(WyRand& g) {
  double x = g.f64();
  double y = g.f64();
  double z = g.f64();
  return {x, y, z};
}
// math/vector.rs:36-45
inline V3 v_random_direction(WyRand& g) {
  double z = g.f64() * 2.0 - 1.0;
  double theta = g.f64() * 2.0 * 3.14159265358979323846264338327950288;  // std::f64::consts::PI
  double r = std::sqrt(1.0 - z * z);
  V3 v{r * std::cos(theta), r * std::sin(theta), z};
  return vnorm(v);
}

// ---------------------------------------------------------------------------------------------
// Plain-data mirrors handed over the test ABI
// ---------------------------------------------------------------------------------------------
struct OCamera {  // raytracing/camera.rs:7-15
  double fov;
  V3 position, direction;
  M3 to_cam_space, to_world_space;
};
struct OConfig {  // raytracing/scene.rs:16-28
  uint64_t rays_per_pixel, max_bounces;
  double focal_length, focal_offset, non_focal_offset;
};
enum : int32_t { SHAPE_SPHERE = 0, SHAPE_PLANE = 1, SHAPE_TRIANGLE = 2 };
struct OObject {  // raytracing/object.rs:10-15 (+ Material :79-89), shape flattened
  int32_t type;
  int32_t pad;
  double shape[9];  // sphere: c.xyz, r | plane: pos.xyz, normal.xyz | triangle: v0, v1, v2
  V3 base_color, emission_color;
  double roughness;
};

inline V3 ld3(const double* p) { return {p[0], p[1], p[2]}; }

// raytracing/camera.rs:42-49
M3 derive_to_world_space_mat(V3 direction) {
  V3 cam_forward = vnorm(direction);
  V3 cam_right = cross(cam_forward, V3{0., 0., -1.});
  V3 cam_up = cross(cam_forward, cam_right);
  return transpose(M3{cam_right, cam_up, cam_forward});
}
// raytracing/camera.rs:19-28
OCamera camera_new(V3 position, V3 direction, double fov) {
  M3 to_world = derive_to_world_space_mat(direction);
  OCamera c;
  c.fov = fov;
  c.position = position;
  c.direction = direction;
  c.to_cam_space = inverse(to_world);
  c.to_world_space = to_world;
  return c;
}
// raytracing/camera.rs:35-40 — matrices are derived from the OLD direction (lag quirk)
void camera_set_direction(OCamera& c, V3 direction) {
  M3 to_world = derive_to_world_space_mat(c.direction);
  c.to_cam_space = inverse(to_world);
  c.to_world_space = to_world;
  c.direction = direction;
}

// ---------------------------------------------------------------------------------------------
// Shapes
// ---------------------------------------------------------------------------------------------
// object/sphere.rs:19-30
bool sphere_distance(const double* s, V3 ray_position, V3 ray_direction, double* out) {
  V3 position = ld3(s);
  double radius = s[3];
  V3 offset = ray_position - position;
  V3 rd = vnorm(ray_direction);
  double a = dot(rd, rd);
  double b = 2.0 * dot(offset, rd);
  double c = dot(offset, offset) - radius * radius;
  double discriminant = b * b - 4.0 * a * c;
  if (discriminant <= 1e-100) return false;
  *out = (-b - std::sqrt(discriminant)) / (2.0 * a);
  return true;
}
// object/sphere.rs:31-33
V3 sphere_normal(const double* s, V3 world_position) { return vnorm(world_position - ld3(s)); }

// object/plane.rs:20-31
bool plane_distance(const double* s, V3 ray_pos, V3 ray_dir, double* out) {
  V3 position = ld3(s), normal = ld3(s + 3);
  V3 offset = ray_pos - position;
  V3 norm = vnorm(normal);
  V3 dir = vnorm(ray_dir);
  if (dot(dir, normal) >= 0. || dot(offset, normal) <= 0.) return false;
  double t = dot(offset, norm) / dot(dir, norm);
  V3 intersection_point = offset + dir * t;
  *out = vlen(offset - intersection_point);
  return true;
}
// object/plane.rs:33-35
V3 plane_normal(const double* s, V3) { return ld3(s + 3); }

// object/triangle.rs:104-107 (ignores its argument)
V3 triangle_normal(const double* s) {
  V3 basis = ld3(s);
  V3 a = ld3(s + 3) - basis;  // plane_vectors :20-25
  V3 b = ld3(s + 6) - basis;
  return vnorm(cross(a, b));
}
// object/triangle.rs:26-35
double triangle_plane_distance(const double* s, V3 ray_pos, V3 dir_in) {
  V3 dir = vnorm(dir_in);
  V3 normal = triangle_normal(s);
  V3 self_pos = ld3(s);
  if (dot(dir, normal) == 0.0) return std::numeric_limits<double>::infinity();
  return dot(normal, self_pos - ray_pos) / dot(dir, normal);
}
// object/triangle.rs:37-101.  *panicked is set where the reference's assert_eq! (:76-78, :92-94) would
// abort the render (non-finite rows); the oracle then rejects the hit.
bool triangle_contains(const double* sh, V3 point, bool* panicked) {
  V3 pos = ld3(sh);
  V3 r = ld3(sh + 3) - pos;
  V3 s = ld3(sh + 6) - pos;
  V3 p = point - pos;
  V3 lgs1{r.x, s.x, p.x};
  V3 lgs2{r.y, s.y, p.y};
  V3 lgs3{r.z, s.z, p.z};
  if (lgs1.x == 0.0) {
    if (lgs2.x == 0.0) {
      if (lgs3.x == 0.0) return false;  // "can't handle LGS"
      std::swap(lgs3, lgs1);
    } else {
      std::swap(lgs1, lgs2);
    }
  }
  lgs1 = lgs1 / lgs1.x;
  lgs2 = lgs2 - lgs1 * (lgs2.x / lgs1.x);
  lgs3 = lgs3 - lgs1 * (lgs3.x / lgs1.x);
  if (!(lgs1.x == 1.0) || !(lgs2.x == 0.0) || !(lgs3.x == 0.0)) {
    if (panicked) *panicked = true;
    return false;
  }
  if (lgs2.y == 0.0) {
    if (lgs3.y == 0.0) return false;  // "can't handle LGS"
    std::swap(lgs2, lgs3);
  }
  lgs2 = lgs2 / lgs2.y;
  lgs1 = lgs1 - lgs2 * (lgs1.y / lgs2.y);
  lgs3 = lgs3 - lgs2 * (lgs3.y / lgs2.y);
  if (!(lgs1.y == 0.0) || !(lgs2.y == 1.0) || !(lgs3.y == 0.0)) {
    if (panicked) *panicked = true;
    return false;
  }
  double a = lgs1.z, b = lgs2.z;
  return 0. <= a && a <= 1. && 0. <= b && b <= 1. && (a + b) <= 1.;
}
// object/triangle.rs:108-127
bool triangle_distance(const double* s, V3 pos, V3 dir, double* out, bool* panicked) {
  if (dot(triangle_normal(s), ld3(s) - dir) < 0.0) return false;
  double distance = std::fabs(triangle_plane_distance(s, pos, dir));
  if (distance == std::numeric_limits<double>::infinity()) return false;
  V3 hit_point = pos + dir * distance;
  if (!triangle_contains(s, hit_point, panicked)) return false;
  *out = distance;
  return true;
}

// raytracing/object.rs:49-51 (dispatch; the Mutex + dyn call are not restated)
bool object_distance(const OObject& o, V3 pos, V3 dir, double* out, bool* panicked) {
  switch (o.type) {
    case SHAPE_SPHERE: return sphere_distance(o.shape, pos, dir, out);
    case SHAPE_PLANE: return plane_distance(o.shape, pos, dir, out);
    default: return triangle_distance(o.shape, pos, dir, out, panicked);
  }
}
// raytracing/object.rs:37-39: shape.normal(p).norm()  (second normalisation)
V3 object_normal_at(const OObject& o, V3 world_pos) {
  V3 n;
  switch (o.type) {
    case SHAPE_SPHERE: n = sphere_normal(o.shape, world_pos); break;
    case SHAPE_PLANE: n = plane_normal(o.shape, world_pos); break;
    default: n = triangle_normal(o.shape); break;
  }
  return vnorm(n);
}

// ---------------------------------------------------------------------------------------------
// Scene hot path — raytracing/scene.rs
// ---------------------------------------------------------------------------------------------
struct Ray {  // raytracing/ray.rs:4-21
  V3 position, direction, resulting_color, light_color;
};

// f64::is_normal && f64::is_sign_positive  (scene.rs:249)
inline bool normal_positive(double d) {
  return std::fpclassify(d) == FP_NORMAL && !std::signbit(d);
}

// scene.rs:243-251: filter_map → filter → min_by(total_cmp): the FIRST minimum wins.
// For the values that survive the filter (positive normals) total_cmp is the ordinary order.
int64_t closest_object(const OObject* objs, uint64_t n, const Ray& ray, double* out_dst,
                       uint64_t* panics) {
  int64_t best = -1;
  double best_d = 0.0;
  for (uint64_t i = 0; i < n; ++i) {
    double d;
    bool panicked = false;
    bool some = object_distance(objs[i], ray.position, ray.direction, &d, &panicked);
    if (panicked && panics) ++*panics;
    if (!some) continue;
    if (!normal_positive(d)) continue;
    if (best < 0 || d < best_d) {
      best = (int64_t)i;
      best_d = d;
    }
  }
  *out_dst = best_d;
  return best;
}

// scene.rs:279-291
V3 random_bounce_dir(WyRand& g, V3 ray_dir, V3 surface_normal, double surface_roughness) {
  V3 random_dir = v_random_direction(g);
  V3 reflection_dir = ray_dir - surface_normal * 2.0 * dot(ray_dir, surface_normal);
  V3 random_to_reflection_dir = reflection_dir - random_dir;
  double reflection_mult = 1.0 - surface_roughness;
  V3 final_direction = random_dir + random_to_reflection_dir * reflection_mult;
  final_direction = vnorm(final_direction);
  if (dot(final_direction, surface_normal) > 0.0) return final_direction;
  return -final_direction;
}
// scene.rs:260-278
void ray_hit(WyRand& g, Ray& ray, const OObject& object) {
  ray.direction =
      random_bounce_dir(g, ray.direction, object_normal_at(object, ray.position), object.roughness);
  ray.resulting_color = ray.resulting_color + ray.light_color * object.emission_color;
  ray.light_color = ray.light_color * object.base_color;
}

struct PixelOut {
  V3 color;
  uint32_t primary_id;  // hit of sample 0's first ray; 0xFFFFFFFF = miss / no ray
  uint32_t rays;        // closest_object calls made for this pixel
};

// scene.rs:223-242
V3 render_ray(const OObject* objs, uint64_t n, const OConfig& cfg, WyRand& g, Ray ray,
              uint32_t* first_id, uint32_t* rays, uint64_t* panics) {
  if (n == 0) return ray.resulting_color;
  for (uint64_t i = 0; i < cfg.max_bounces + 1; ++i) {
    if (veq(ray.light_color, V3{0., 0., 0.})) break;
    double dst;
    int64_t hit = closest_object(objs, n, ray, &dst, panics);
    ++*rays;
    if (i == 0 && first_id) *first_id = hit < 0 ? 0xFFFFFFFFu : (uint32_t)hit;
    if (hit < 0) break;
    ray.position = ray.position + ray.direction * dst;
    ray_hit(g, ray, objs[hit]);
  }
  return ray.resulting_color;
}

// scene.rs:213-222
V3 get_ray_dir(const OCamera& cam, double x, double y, double vertical_fov) {
  double angle_x = cam.fov * (x - 0.5);
  double angle_y = vertical_fov * (y - 0.5);
  V3 cam_space_dir{std::sin(angle_x), std::sin(angle_y), std::cos(angle_x) * std::cos(angle_y)};
  return matvec(cam.to_world_space, cam_space_dir);  // camera.rs:65-67
}

enum : int32_t { RNG_PIXEL_SAMPLE = 0, RNG_ROW = 1 };

// scene.rs:194-212 + avg :253-259.  `g` is the row stream in RNG_ROW mode; in RNG_PIXEL_SAMPLE mode a
// fresh stream is started per (pixel, sample) from key = (py*width + px)*rpp + s.
PixelOut render_pixel(const OCamera& cam, const OConfig& cfg, const OObject* objs, uint64_t n,
                      double u, double v, double vertical_fov, int32_t rng_mode, uint64_t seed,
                      uint64_t pixel_index, WyRand& g, uint64_t* panics) {
  PixelOut out{{0., 0., 0.}, 0xFFFFFFFFu, 0};
  V3 ray_dir = get_ray_dir(cam, u, v, vertical_fov);
  Ray base{cam.position, ray_dir, {0., 0., 0.}, {1., 1., 1.}};
  V3 sum{0., 0., 0.};  // iter_ops.rs:4-8: fold from zeros(), left to right
  for (uint64_t s = 0; s < cfg.rays_per_pixel; ++s) {
    WyRand local{stream_state(seed, pixel_index * cfg.rays_per_pixel + s)};
    WyRand& rng = (rng_mode == RNG_ROW) ? g : local;
    Ray ray = base;
    V3 ray_position = ray.position + v_random(rng) * cfg.non_focal_offset;
    V3 focal_point = ray.position + ray.direction * cfg.focal_length;
    V3 target_point = focal_point + v_random(rng) * cfg.focal_offset;
    V3 ray_direction = target_point - ray_position;
    ray.position = ray_position;
    ray.direction = vnorm(ray_direction);
    V3 c = render_ray(objs, n, cfg, rng, ray, s == 0 ? &out.primary_id : nullptr, &out.rays, panics);
    sum = sum + c;
  }
  out.color = sum / (double)cfg.rays_per_pixel;  // `/ length` → `length as f64` (general.rs:9-20)
  return out;
}

}  // namespace

// =============================================================================================
// Test ABI (ctypes)
// =============================================================================================
extern "C" {

// scene.rs:144-170.  nthreads > 0: pool of nthreads pulling rows; 0: hardware_concurrency();
// -1: reference-faithful one std::thread per row (scene.rs:149-151).
// xs/ys: render only pixels with x % xs == 0 and y % ys == 0 of the full w×h image (outputs are the
// compact ceil(w/xs) × ceil(h/ys) grid) — used to brute-force-check large scenes on a sub-grid.
int oracle_render(const OCamera* cam, const OConfig* cfg, const OObject* objs, uint64_t n_objs,
                  uint32_t width, uint32_t height, uint64_t seed, int32_t rng_mode, int32_t nthreads,
                  uint32_t xs, uint32_t ys, double* out_rgb, uint32_t* out_primary_id,
                  uint32_t* out_rays_per_pixel, uint64_t* out_total_rays, uint64_t* out_panics) {
  if (!cam || !cfg || width == 0 || height == 0 || xs == 0 || ys == 0) return 1;
  const double vertical_fov = (double)height / (double)width * cam->fov;  // scene.rs:145
  const uint32_t ow = (width + xs - 1) / xs, oh = (height + ys - 1) / ys;
  std::atomic<uint64_t> total_rays{0}, total_panics{0};
  std::atomic<uint32_t> next_row{0};

  auto do_row = [&](uint32_t oy) {
    const uint32_t py = oy * ys;
    const double v = (double)py / (double)height;  // scene.rs:153
    WyRand row_rng{stream_state(seed ^ 0x524F575F53545245ULL /* "ROW_STRE" */, py)};
    uint64_t rays = 0, panics = 0;
    for (uint32_t ox = 0; ox < ow; ++ox) {
      const uint32_t px = ox * xs;
      const double u = (double)px / (double)width;  // scene.rs:157
      PixelOut o = render_pixel(*cam, *cfg, objs, n_objs, u, v, vertical_fov, rng_mode, seed,
                                (uint64_t)py * width + px, row_rng, &panics);
      const size_t idx = (size_t)oy * ow + ox;
      if (out_rgb) {
        out_rgb[idx * 3 + 0] = o.color.x;
        out_rgb[idx * 3 + 1] = o.color.y;
        out_rgb[idx * 3 + 2] = o.color.z;
      }
      if (out_primary_id) out_primary_id[idx] = o.primary_id;
      if (out_rays_per_pixel) out_rays_per_pixel[idx] = o.rays;
      rays += o.rays;
    }
    total_rays += rays;
    total_panics += panics;
  };

  if (nthreads < 0) {
    std::vector<std::thread> ts;
    ts.reserve(oh);
    for (uint32_t oy = 0; oy < oh; ++oy) ts.emplace_back(do_row, oy);
    for (auto& t : ts) t.join();  // scene.rs:167-169: joined in row order
  } else {
    unsigned nt = nthreads > 0 ? (unsigned)nthreads : std::thread::hardware_concurrency();
    if (nt == 0) nt = 1;
    auto worker = [&]() {
      for (;;) {
        uint32_t oy = next_row.fetch_add(1);
        if (oy >= oh) break;
        do_row(oy);
      }
    };
    std::vector<std::thread> ts;
    for (unsigned i = 1; i < nt; ++i) ts.emplace_back(worker);
    worker();
    for (auto& t : ts) t.join();
  }
  if (out_total_rays) *out_total_rays = total_rays.load();
  if (out_panics) *out_panics = total_panics.load();
  return 0;
}

// scene.rs:173-178: flip y, `* 256` (integer → 256.0), saturating `as u8` (NaN → 0).
void oracle_to_rgb8(const double* rgb, uint32_t width, uint32_t height, uint8_t* out) {
  auto sat = [](double c) -> uint8_t {
    double s = c * 256.0;
    if (!(s > 0.0)) return 0;  // negatives, -0, NaN
    if (s >= 255.0) return 255;
    return (uint8_t)s;  // truncation toward zero
  };
  for (uint32_t y = 0; y < height; ++y)
    for (uint32_t x = 0; x < width; ++x) {
      const double* src = rgb + ((size_t)(height - y - 1) * width + x) * 3;
      uint8_t* dst = out + ((size_t)y * width + x) * 3;
      dst[0] = sat(src[0]);
      dst[1] = sat(src[1]);
      dst[2] = sat(src[2]);
    }
}

unsigned oracle_hardware_threads() { return std::thread::hardware_concurrency(); }

// ---- per-function entry points (unit parity tests) ------------------------------------------
void oracle_wyrand_u64(uint64_t* state, uint64_t n, uint64_t* out) {
  WyRand g{*state};
  for (uint64_t i = 0; i < n; ++i) out[i] = g.gen_u64();
  *state = g.s;
}
void oracle_wyrand_f64(uint64_t* state, uint64_t n, double* out) {
  WyRand g{*state};
  for (uint64_t i = 0; i < n; ++i) out[i] = g.f64();
  *state = g.s;
}
uint64_t oracle_stream_state(uint64_t seed, uint64_t key) { return stream_state(seed, key); }

void oracle_camera_new(const double* position, const double* direction, double fov, OCamera* out) {
  *out = camera_new(ld3(position), ld3(direction), fov);
}
void oracle_camera_set_direction(OCamera* cam, const double* direction) {
  camera_set_direction(*cam, ld3(direction));
}
static void st3(double* o, V3 v) {
  o[0] = v.x;
  o[1] = v.y;
  o[2] = v.z;
}
// camera.rs:51-67
void oracle_camera_to_cam_space(const OCamera* c, const double* v, double* out) {
  st3(out, matvec(c->to_cam_space, ld3(v) - c->position));
}
void oracle_camera_to_world_space(const OCamera* c, const double* v, double* out) {
  st3(out, matvec(c->to_world_space, ld3(v)) + c->position);
}
void oracle_camera_rotate_to_world_space(const OCamera* c, const double* v, double* out) {
  st3(out, matvec(c->to_world_space, ld3(v)));
}
void oracle_get_ray_dir(const OCamera* c, double x, double y, double vertical_fov, double* out) {
  st3(out, get_ray_dir(*c, x, y, vertical_fov));
}
void oracle_mat_inverse(const double* m, double* out) {
  M3 a{ld3(m), ld3(m + 3), ld3(m + 6)};
  M3 r = inverse(a);
  st3(out, r.x);
  st3(out + 3, r.y);
  st3(out + 6, r.z);
}
double oracle_mat_determinant(const double* m) {
  return determinant(M3{ld3(m), ld3(m + 3), ld3(m + 6)});
}

// Batched Object::distance: rays[n][6] = pos.xyz, dir.xyz → out_t[n] (NaN-free: valid flag separately)
void oracle_object_distance(const OObject* obj, const double* rays, uint64_t n, double* out_t,
                            uint8_t* out_some) {
  for (uint64_t i = 0; i < n; ++i) {
    double d = 0.0;
    bool panicked = false;
    bool some = object_distance(*obj, ld3(rays + 6 * i), ld3(rays + 6 * i + 3), &d, &panicked);
    out_some[i] = some ? 1 : 0;
    out_t[i] = some ? d : 0.0;
  }
}
void oracle_object_normal_at(const OObject* obj, const double* pos, uint64_t n, double* out) {
  for (uint64_t i = 0; i < n; ++i) st3(out + 3 * i, object_normal_at(*obj, ld3(pos + 3 * i)));
}
// Batched closest_object over a whole scene.
void oracle_closest_object(const OObject* objs, uint64_t n_objs, const double* rays, uint64_t n,
                           double* out_t, uint32_t* out_id) {
  for (uint64_t i = 0; i < n; ++i) {
    Ray r{ld3(rays + 6 * i), ld3(rays + 6 * i + 3), {0, 0, 0}, {1, 1, 1}};
    double d;
    int64_t id = closest_object(objs, n_objs, r, &d, nullptr);
    out_id[i] = id < 0 ? 0xFFFFFFFFu : (uint32_t)id;
    out_t[i] = id < 0 ? 0.0 : d;
  }
}
// random_bounce_dir with stream states: in[n][7] = dir.xyz, normal.xyz, roughness; states[n] in/out
void oracle_random_bounce_dir(uint64_t* states, const double* in, uint64_t n, double* out) {
  for (uint64_t i = 0; i < n; ++i) {
    WyRand g{states[i]};
    st3(out + 3 * i, random_bounce_dir(g, ld3(in + 7 * i), ld3(in + 7 * i + 3), in[7 * i + 6]));
    states[i] = g.s;
  }
}
void oracle_random_direction(uint64_t* state, uint64_t n, double* out) {
  WyRand g{*state};
  for (uint64_t i = 0; i < n; ++i) st3(out + 3 * i, v_random_direction(g));
  *state = g.s;
}

uint64_t oracle_sizeof_object() { return sizeof(OObject); }
uint64_t oracle_sizeof_camera() { return sizeof(OCamera); }
uint64_t oracle_sizeof_config() { return sizeof(OConfig); }

}  // extern "C"
