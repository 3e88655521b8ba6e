// rtx_trace.cuh — nearest-hit and shading device functions (and the host-side shape precompute).
//
// Every function is the hoisted, bit-identical form of a reference routine; the derivations (why each
// rewrite cannot change a single output bit) are in DESIGN.md §"Exact rewrites" and are enforced by the
// per-function parity tests (tests/test_gpu_functions.py) against the as-written CPU oracle.
#pragma once
#if defined(__CUDACC__)
#include <math_constants.h>
#endif
#include "rtx_rng.cuh"
#include "rtx_scene.cuh"
#include "rtx_sincos.cuh"

namespace rtx {

// ---------------------------------------------------------------------------------------------
// Upload-time precompute (host).  Same non-fused f64 operations the reference performs per ray.
// ---------------------------------------------------------------------------------------------
inline void pack_sphere(const double* c_xyz, double radius, double* rec) {
  rec[0] = c_xyz[0];
  rec[1] = c_xyz[1];
  rec[2] = c_xyz[2];
  rec[3] = fmul(radius, radius);  // sphere.rs:24: self.radius * self.radius
}
inline void pack_plane(const double* pos, const double* normal, double* rec) {
  V3 n = mk3(normal[0], normal[1], normal[2]);
  V3 nh = vnorm(n);  // plane.rs:22 (and object.rs:38 for normal_at)
  rec[0] = pos[0], rec[1] = pos[1], rec[2] = pos[2];
  rec[3] = n.x, rec[4] = n.y, rec[5] = n.z;
  rec[6] = nh.x, rec[7] = nh.y, rec[8] = nh.z;
  rec[9] = 0.0;
}
inline double comp(V3 v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : v.z); }
// triangle.rs:37-94 with the p-column left symbolic.  Rows are L_i = (r[i], s[i], p[i]).
inline void pack_triangle(const double* v, double* rec) {
  V3 v0 = mk3(v[0], v[1], v[2]);
  V3 r = mk3(v[3], v[4], v[5]) - v0;  // plane_vectors :20-25
  V3 s = mk3(v[6], v[7], v[8]) - v0;
  V3 n = vnorm(cross(r, s));  // normal :104-107
  V3 nn = vnorm(n);           // Object::normal_at, object.rs:38
  int i1 = 0, i2 = 1, i3 = 2;
  uint64_t code = 0;
  double r1 = 0, l1y = 0, r2 = 0, d2 = 0;
  bool degenerate = false, panic = false;
  if (comp(r, i1) == 0.0) {  // :60-71
    if (comp(r, i2) == 0.0) {
      if (comp(r, i3) == 0.0) {
        degenerate = true;
      } else {
        int t = i1; i1 = i3; i3 = t;
      }
    } else {
      int t = i1; i1 = i2; i2 = t;
    }
  }
  if (!degenerate) {
    r1 = comp(r, i1);
    const double l1x = fdiv(r1, r1);   // :72  lgs1 /= lgs1.x
    l1y = fdiv(comp(s, i1), r1);
    // :73-74  lgs_k -= lgs1 * (lgs_k.x / lgs1.x)
    const double f2 = fdiv(comp(r, i2), l1x), f3 = fdiv(comp(r, i3), l1x);
    const double l2x = fsub(comp(r, i2), fmul(l1x, f2));
    const double l3x = fsub(comp(r, i3), fmul(l1x, f3));
    double l2y = fsub(comp(s, i2), fmul(l1y, f2));
    double l3y = fsub(comp(s, i3), fmul(l1y, f3));
    if (!(l1x == 1.0) || !(l2x == 0.0) || !(l3x == 0.0)) panic = true;  // :76-78
    if (!panic) {
      if (l2y == 0.0) {  // :81-87
        if (l3y == 0.0) {
          degenerate = true;
        } else {
          int t = i2; i2 = i3; i3 = t;
          double ty = l2y; l2y = l3y; l3y = ty;
        }
      }
      if (!degenerate) {
        d2 = l2y;
        r2 = comp(r, i2);  // == f2 of the (possibly swapped) row: r/1.0
        const double n2x = fdiv(0.0, d2), n2y = fdiv(d2, d2);  // :88 lgs2 /= lgs2.y (x column is ±0)
        const double g1 = fdiv(l1y, n2y), g3 = fdiv(l3y, n2y);
        const double n1x = fsub(l1x, fmul(n2x, g1)), n1y = fsub(l1y, fmul(n2y, g1));
        const double n3y = fsub(l3y, fmul(n2y, g3));
        if (!(n1y == 0.0) || !(n2y == 1.0) || !(n3y == 0.0)) panic = true;  // :92-94
        (void)n1x;
      }
    }
  }
  code = ((uint64_t)i1 << TRI_I1_SHIFT) | ((uint64_t)i2 << TRI_I2_SHIFT);
  if (degenerate) code |= TRI_DEGENERATE;
  if (panic) code |= TRI_PANIC;
  rec[0] = v0.x, rec[1] = v0.y, rec[2] = v0.z;
  rec[3] = n.x, rec[4] = n.y, rec[5] = n.z;
  rec[6] = nn.x, rec[7] = nn.y, rec[8] = nn.z;
  rec[9] = r1, rec[10] = l1y, rec[11] = r2, rec[12] = d2;
  __builtin_memcpy(&rec[13], &code, 8);
}
inline void pack_material(const double* base, const double* emission, double roughness, double* rec) {
  rec[0] = base[0], rec[1] = base[1], rec[2] = base[2];
  rec[3] = emission[0], rec[4] = emission[1], rec[5] = emission[2];
  rec[6] = roughness;
  // scene.rs:228: `light_color == zeros` after `light_color *= base_color` (:277).  Only an exactly-zero
  // base colour is flagged here; the kernel still performs the generic comparison every bounce.
  uint64_t flags = (base[0] == 0.0 && base[1] == 0.0 && base[2] == 0.0) ? MAT_TERMINAL : 0;
  __builtin_memcpy(&rec[7], &flags, 8);
}

#if defined(__CUDACC__)
// ---------------------------------------------------------------------------------------------
// Device: nearest hit
// ---------------------------------------------------------------------------------------------
struct Hit {
  double t;      // +inf when nothing was hit
  uint32_t id;   // object index (insertion order) or MISS_ID
  uint32_t ref;  // REF_* | index within the shape array
};

// Ray-invariant terms every Shape::distance recomputes per object in the reference
// (sphere.rs:21-22, plane.rs:23, triangle.rs:28): dn = dir.norm(), a = dn·dn.
struct RayInv {
  V3 dn;
  double a;
};
RTX_D RayInv ray_invariants(V3 dir) {
  RayInv ri;
  ri.dn = vnorm(dir);
  ri.a = dot(ri.dn, ri.dn);
  return ri;
}

// scene.rs:249-250: keep positive normal distances; min_by(total_cmp) returns the FIRST minimum, i.e.
// the lowest object index among equal distances.
RTX_D void consider(Hit& h, double t, uint32_t id, uint32_t ref) {
  if (normal_positive(t) && (t < h.t || (t == h.t && id < h.id))) {
    h.t = t;
    h.id = id;
    h.ref = ref;
  }
}

// sphere.rs:19-30.  With b = 2h (h = offset·dn): disc = b*b - (4a)c = 4(h*h - a*c) exactly (scaling by
// powers of two commutes with rounding), sqrt(disc) = 2 sqrt(q), and (-b - sqrt(disc)) / (2a) =
// (-h - sqrt(q)) / a — every intermediate differs from the reference's by an exact power of two, the
// result by nothing.  `disc <= 1e-100` becomes `q <= 1e-100/4`.  h >= 0 gives a negative t, which
// scene.rs:249 drops; testing it first skips the sqrt + division.
RTX_D bool sphere_candidate(const double* __restrict__ rec, V3 pos, const RayInv& ri, double* out_h, double* out_q) {
  const double2 cxy = reinterpret_cast<const double2*>(rec)[0], czr = reinterpret_cast<const double2*>(rec)[1];
  const V3 o = pos - mk3(cxy.x, cxy.y, czr.x);
  const double h = dot(o, ri.dn);
  const double c = fsub(dot(o, o), czr.y);
  const double q = fsub(fmul(h, h), fmul(ri.a, c));
  *out_h = h;
  *out_q = q;
  return q > (1e-100 * 0.25) && h < 0.0;
}
RTX_D double sphere_root(double h, double q, const RayInv& ri) { return fdiv(fsub(-h, fsqrt(q)), ri.a); }
RTX_D bool sphere_distance(const double* __restrict__ rec, V3 pos, const RayInv& ri, double* out_t) {
  double h, q;
  if (sphere_candidate(rec, pos, ri, &h, &q)) {
    *out_t = sphere_root(h, q, ri);
    return true;
  }
  return false;
}

// plane.rs:20-31
RTX_D bool plane_distance(const double* __restrict__ rec, V3 pos, const RayInv& ri, double* out_t) {
  const V3 o = pos - mk3(rec[0], rec[1], rec[2]);
  const V3 nraw = mk3(rec[3], rec[4], rec[5]);
  if (dot(ri.dn, nraw) >= 0.0 || dot(o, nraw) <= 0.0) return false;
  const V3 nh = mk3(rec[6], rec[7], rec[8]);
  const double t = fdiv(dot(o, nh), dot(ri.dn, nh));
  const V3 ip = o + ri.dn * t;
  *out_t = vlen(o - ip);
  return true;
}

RTX_D double pick(V3 v, uint32_t i) { return i == 0 ? v.x : (i == 1 ? v.y : v.z); }

// triangle.rs:108-127 (+ plane_distance :26-35, contains :37-101).  Returns true for Some(distance).
// *panic is set when the reference's assert_eq! would have aborted the render.
RTX_D bool triangle_distance(const double* __restrict__ rec, V3 pos, V3 dir, const RayInv& ri, double* out_t,
                             bool* panic) {
  const V3 v0 = mk3(rec[0], rec[1], rec[2]);
  const V3 n = mk3(rec[3], rec[4], rec[5]);
  if (dot(n, v0 - dir) < 0.0) return false;  // :115 — uses dir, not pos (reference quirk)
  const double dnn = dot(ri.dn, n);
  if (dnn == 0.0) return false;  // plane_distance → +INF → None (:31-33, :119)
  const double distance = fabs(fdiv(dot(n, v0 - pos), dnn));  // :34, :118
  if (distance == CUDART_INF) return false;                   // :119
  const V3 p = (pos + dir * distance) - v0;                   // :122, :39
  const uint64_t code = (uint64_t)__double_as_longlong(rec[13]);
  if (code & TRI_DEGENERATE) return false;
  if (code & TRI_PANIC) {
    *panic = true;
    return false;
  }
  const uint32_t i1 = (uint32_t)(code >> TRI_I1_SHIFT) & 3u, i2 = (uint32_t)(code >> TRI_I2_SHIFT) & 3u;
  const double l1z = fdiv(pick(p, i1), rec[9]);             // :72   p1 / r1
  const double l2z = fsub(pick(p, i2), fmul(l1z, rec[11]));  // :73   p2 - l1z * r2
  const double b = fdiv(l2z, rec[12]);                      // :88   / d2
  const double a = fsub(l1z, fmul(b, rec[10]));             // :89   l1z - b * l1y
  if (0.0 <= a && a <= 1.0 && 0.0 <= b && b <= 1.0 && fadd(a, b) <= 1.0) {  // :100
    *out_t = distance;
    return true;
  }
  return false;
}

// Conservative f32 cull for the sphere scan.  It never decides a hit: a sphere it lets through is evaluated by
// sphere_candidate/sphere_root in f64 exactly as before; a sphere it rejects is one that sphere_candidate is
// *proven* to reject (q <= 0 or h >= 0), so the scan's result is unchanged bit for bit.
//   f32 values: o32 = p32 - c32, h32 = o32·d32, c32 = o32·o32 - r2_32, q32 = h32² - a32·c32 (FMA, round to nearest).
//   With u = 2^-24, R >= |p_i| + |c_i| for every component and sphere, D1 = |d_x|+|d_y|+|d_z|:
//     |o32_i - o_i| <= 2uR, |h32 - h| <= 6u·R·D1, |c32 - c| <= 25uR² + 2u·r², and
//     |q32 - q| <= 14u·B with B = (R·D1)² + a(3R² + r²max)   (conversion + one rounding per operation;
//   the f64 roundings of the exact path, ~2^-53·B, disappear in the slack).  Margins used: 32u·B for q and 16u·R·D1
//   for h, plus 1e-30 absolute for f32 underflow.  NaN anywhere makes a comparison false ⇒ "keep"; B >= 1e37
//   (overflow range) switches the cull off for the ray.
struct SphereCull {
  float px, py, pz, dx, dy, dz, a, neg_mq, mh;
};
RTX_D SphereCull sphere_cull_setup(const SceneView& sc, V3 pos, const RayInv& ri) {
  SphereCull k;
  k.px = __double2float_rn(pos.x), k.py = __double2float_rn(pos.y), k.pz = __double2float_rn(pos.z);
  k.dx = __double2float_rn(ri.dn.x), k.dy = __double2float_rn(ri.dn.y), k.dz = __double2float_rn(ri.dn.z);
  k.a = __double2float_ru(ri.a);
  const float pmax = fmaxf(fmaxf(fabsf(k.px), fabsf(k.py)), fabsf(k.pz));
  if (sc.cull_pmax > 0.0f) {  // (warp-uniform) margins precomputed per launch for origins within cull_pmax
    const bool in = pmax <= sc.cull_pmax && k.a <= 1.000001f;
    k.neg_mq = in ? sc.cull_neg_mq : CUDART_NAN_F;
    k.mh = in ? sc.cull_mh : CUDART_NAN_F;
    return k;
  }
  const float R = __fadd_ru(pmax, sc.sph_cmax);
  const float rd = __fmul_ru(R, __fadd_ru(__fadd_ru(fabsf(k.dx), fabsf(k.dy)), fabsf(k.dz)));
  const float B = __fmaf_ru(rd, rd, __fmul_ru(fabsf(k.a), __fmaf_ru(__fmul_ru(3.0f, R), R, sc.sph_r2max)));
  // B far from the f32 range: no intermediate of sphere_cull_keep can overflow to an infinity that would
  // compare as a certain miss.  Otherwise (or NaN) the margins are NaN and every sphere is kept.
  const bool ok = B < 1e37f;
  k.neg_mq = ok ? -__fmaf_ru(B, 1.9073486328125e-6f /* 2^-19 */, 1e-30f) : CUDART_NAN_F;
  k.mh = ok ? __fmaf_ru(rd, 9.5367431640625e-7f /* 2^-20 */, 1e-30f) : CUDART_NAN_F;
  return k;
}
// false only when sphere_candidate is certain to return false for this sphere
RTX_D bool sphere_cull_keep(const float* __restrict__ rec32, const SphereCull& k) {
  const float4 s = *reinterpret_cast<const float4*>(rec32);
  const float ox = k.px - s.x, oy = k.py - s.y, oz = k.pz - s.z;
  const float h = fmaf(oz, k.dz, fmaf(oy, k.dy, ox * k.dx));
  const float c = fmaf(oz, oz, fmaf(oy, oy, fmaf(ox, ox, -s.w)));
  const float q = fmaf(h, h, -(k.a * c));
  return !(q <= k.neg_mq) && !(h >= k.mh);
}

// The sphere part of the scan over `n` spheres whose records start at sph32 / sph / ids (the scene's arrays, or a chunk
// of them staged in shared memory); first = scene index of the first one.  In batches of 32: a branch-free f32 pass
// (SphereCull) records as one bit each the spheres that cannot be ruled out; those few (usually 0-2 per ray) then go
// through the exact f64 test and, if it passes, the sqrt + division of the root.  `consider` is a lexicographic
// (t, id) minimum, so the visiting order is irrelevant.
RTX_D void scan_spheres(const float* __restrict__ sph32, const double* __restrict__ sph, const uint32_t* __restrict__ ids,
                        uint32_t first, uint32_t n, V3 pos, const RayInv& ri, const SphereCull& cull, Hit& h) {
  for (uint32_t base = 0; base < n; base += 32u) {
    const uint32_t cnt = min(32u, n - base);
    const double* __restrict__ chunk = sph + (size_t)base * SPH_STRIDE;
    const float* __restrict__ chunk32 = sph32 + (size_t)base * SPH_STRIDE;
    uint32_t cand = 0u;
    // (not unrolled: the pool kernel's hot loop is larger than the instruction cache, and every copy of this body costs
    // more in fetch stalls than it saves in loop overhead — C2 10.43 ms at unroll 2, 10.35 at 1, 10.66 at 3, 11.01 at 6)
#pragma unroll 1
    for (uint32_t j = 0; j < cnt; ++j) cand |= (sphere_cull_keep(chunk32 + (size_t)j * SPH_STRIDE, cull) ? 1u : 0u) << j;
    while (cand != 0u) {
      const uint32_t j = (uint32_t)__ffs((int)cand) - 1u;
      cand &= cand - 1u;
      double hh, qq;
      if (sphere_candidate(chunk + (size_t)j * SPH_STRIDE, pos, ri, &hh, &qq))
        consider(h, sphere_root(hh, qq, ri), ids[base + j], REF_SPHERE | (first + base + j));
    }
  }
}
// ... and the planes and triangles.
template <bool HAS_TRI = true>
RTX_D void scan_planes_triangles(const SceneView& sc, V3 pos, V3 dir, const RayInv& ri, Hit& h, uint32_t* panics) {
  for (uint32_t i = 0; i < sc.n_pl; ++i) {
    double t;
    if (plane_distance(sc.pl + (size_t)i * PL_STRIDE, pos, ri, &t)) consider(h, t, sc.pl_id[i], REF_PLANE | i);
  }
  for (uint32_t i = 0; HAS_TRI && i < sc.n_tri; ++i) {
    double t;
    bool panic = false;
    if (triangle_distance(sc.tri + (size_t)i * TRI_STRIDE, pos, dir, ri, &t, &panic))
      consider(h, t, sc.tri_id[i], REF_TRIANGLE | i);
    if (panic) ++*panics;
  }
}

// scene.rs:243-251 — brute-force scan over all three shape arrays.
// HAS_TRI = false: the caller knows the scene has no triangles (the loop and its registers are compiled out)
template <bool HAS_TRI = true>
RTX_D Hit closest_brute(const SceneView& sc, V3 pos, V3 dir, uint32_t* panics) {
  Hit h;
  h.t = CUDART_INF;
  h.id = MISS_ID;
  h.ref = 0;
  const RayInv ri = ray_invariants(dir);
  const SphereCull cull = sphere_cull_setup(sc, pos, ri);
  scan_spheres(sc.sph32, sc.sph, sc.sph_id, 0u, sc.n_sph, pos, ri, cull, h);
  scan_planes_triangles<HAS_TRI>(sc, pos, dir, ri, h, panics);
  return h;
}

// ---------------------------------------------------------------------------------------------
// Device: shading
// ---------------------------------------------------------------------------------------------
// object.rs:37-39: shape.normal(p).norm()
RTX_D V3 normal_at(const SceneView& sc, uint32_t ref, V3 pos) {
  const uint32_t idx = ref & REF_INDEX_MASK;
  const uint32_t kind = ref & ~REF_INDEX_MASK;
  if (kind == REF_SPHERE) {
    const double* rec = sc.sph + (size_t)idx * SPH_STRIDE;
    return vnorm(vnorm(pos - mk3(rec[0], rec[1], rec[2])));  // sphere.rs:31-33, then object.rs:38
  } else if (kind == REF_PLANE) {
    const double* rec = sc.pl + (size_t)idx * PL_STRIDE;
    return mk3(rec[6], rec[7], rec[8]);  // plane.rs:33-35 → norm()
  } else {
    const double* rec = sc.tri + (size_t)idx * TRI_STRIDE;
    return mk3(rec[6], rec[7], rec[8]);  // triangle.rs:104-107 → norm() again
  }
}

// math/vector.rs:36-45
RTX_D V3 random_direction(WyRand& g, const double* sctab) {
  const double z = fsub(fmul(wy_f64(g), 2.0), 1.0);
  const double theta = fmul(fmul(wy_f64(g), 2.0), 3.14159265358979323846264338327950288);
  const double r = fsqrt(fsub(1.0, fmul(z, z)));
  // glibc's sin/cos restated bit for bit (rtx_sincos.cuh): CUDA's sincos() differs from it by an ulp in
  // ~15 % of the draws, which is enough to flip later hits on triangle scenes.
  double s, c;
  libm_sincos(theta, sctab, &s, &c);
  return vnorm(mk3(fmul(r, c), fmul(r, s), z));
}

// scene.rs:279-291
// sctab: glibc's sin/cos table — sincos_table() (global memory) or a copy staged in shared memory
RTX_D V3 random_bounce_dir(WyRand& g, V3 ray_dir, V3 n, double roughness, const double* sctab = sincos_table()) {
  const V3 rd = random_direction(g, sctab);
  const V3 refl = ray_dir - (n * 2.0) * dot(ray_dir, n);
  const V3 to_refl = refl - rd;
  const double mult = fsub(1.0, roughness);
  const V3 fin = vnorm(rd + to_refl * mult);
  return dot(fin, n) > 0.0 ? fin : -fin;
}
#endif  // __CUDACC__

}  // namespace rtx
