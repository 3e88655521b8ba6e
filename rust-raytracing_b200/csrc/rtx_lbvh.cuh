// rtx_lbvh.cuh — LBVH node layout and the traversal (K5): near-first, with a per-thread stack (not stackless: the
// escape-pointer walk of the first build visited far more nodes; see DESIGN.md §6 for what was measured).
//
// The reference has no acceleration structure: closest_object is an O(N) scan (raytracing/scene.rs:243-251).
// The LBVH must return EXACTLY what that scan returns: the same distance bits and, among equal distances,
// the lowest object index.  So the tree only culls; every surviving primitive goes through the same exact
// f64 distance routine as the brute-force kernel (rtx_trace.cuh), and culling is conservative:
//  * boxes are built in f64, padded outward by 2^-20 x scene scale (~1e9 x the rounding error of a hit
//    point), then rounded OUTWARD to f32;
//  * the slab test is f32 interval arithmetic with directed rounding (__fsub_rd/_ru, __fmul_rd/_ru) on
//    rigorous f32 enclosures of the f64 ray, so the computed [entry, exit] always contains the true one —
//    it runs on the FP32 pipe, leaving the FP64 pipe to the exact leaf tests;
//  * a child is skipped only if its interval is provably empty, provably behind the origin, or provably
//    beyond the current best distance (strictly: equal distances must still be seen for the tie rule).
// Triangles need a second kind of tree.  A "phantom" hit (plane BEHIND the ray, distance = |t|,
// triangle.rs:118) reports the mirrored point pos + dir*|t|, which is off the plane and only has to lie in
// the infinite prism over the triangle's projection along the coordinate contains() drops
// (triangle.rs:96-100).  One "prism tree" per dropped axis a is traversed after the 3-D tree.  Its nodes keep
// the triangles' true 3-D boxes, but a node is entered on the evidence of the two KEPT axes only (the reported
// point may be anywhere along a), pruned by the best distance so far and by two bounds that hold for every
// triangle T below the node (unit normal n, |n_perp| <= S, |n_a| >= Z; v0 inside the node's box):
//  (A) "not a phantom": if the node lies wholly ahead of the origin along a (distance >= g > 0 in the direction
//      the ray travels along a), g*Z > S*dperp_max and |d_a|*Z > |d_perp|*S, then for every T both n.(v0 - pos)
//      and d.n take the sign of n_a (their a-terms dominate): the plane is in FRONT, the hit is an ordinary
//      one and the 3-D tree finds it;
//  (B) |t| = |n.(v0 - pos)| / |d.n| <= (S*dperp_max + da_max) / (|d_a|*Z - |d_perp|*S) =: U when the denominator
//      is positive: the reported point lies at parameter <= U, so a node whose kept-axes entry parameter exceeds
//      U cannot contain it.
// Without them a ray that leaves a height field upwards walks the footprint of every triangle under its path.
#pragma once
#include "rtx_trace.cuh"

namespace rtx {

constexpr uint32_t BVH_END = 0xFFFFFFFFu;
constexpr uint32_t BVH_NO_PRIM = 0xFFFFFFFFu;

struct __align__(16) BvhNode {  // 64 B = four 128-bit loads
  double lo[3];
  double hi[3];
  uint32_t left;    // internal: first child (its sibling is reached through the child's escape)
  uint32_t escape;  // next node once this subtree is done (BVH_END at the end of the walk)
  uint32_t prim;    // leaf: REF_* | index ; internal: BVH_NO_PRIM
  uint32_t pad;
};

// Traversal node: both children's boxes in f32, rounded OUTWARD from the padded f64 boxes, + child refs.
// Child ref: bit 31 = leaf; leaf payload: bit 30 = triangle (else sphere), bits 0..29 = index in its array.
constexpr uint32_t WIDE_LEAF = 0x80000000u, WIDE_TRI = 0x40000000u, WIDE_IDX = 0x3FFFFFFFu;
struct __align__(16) WideNode {  // 64 B = four 128-bit loads
  float lo0[3], hi0[3];
  float lo1[3], hi1[3];
  uint32_t c0, c1;
  uint32_t pad[2];
};

struct BvhView {
  const WideNode* nodes[4];  // [0] 3-D tree over spheres + triangles; [1..3] prism trees (dropped axis x,y,z)
  uint32_t root[4];          // BVH_END when the tree is empty, else 0
  // Per prism tree (index 1 + a, a = dropped axis), over its triangles' unit normals n:
  //   prism_s >= max |n without its a-component|, prism_z <= min |n_a|  (rounded outward),
  //   prism_sign = +1 / -1 when every n_a has that sign, 0 when mixed.
  float prism_s[4], prism_z[4];
  int prism_sign[4];
};

#if defined(__CUDACC__)
struct TraversalCounters {
  unsigned long long nodes, prims;
};

// Per-ray f32 data for the conservative slab test: one FFMA per slab side.
//   exact      t  = (x - o) / d                      x = a box coordinate (f32), o, d = the f64 ray
//   computed   u  = rn(x * inv + c)                  inv = rn32(1/d),  c = -(o/d) -+ e  rounded away from o/d
// With e >= 2^-21 |o/d| the errors of inv (2^-24 relative, hitting x*inv = t + o/d), of o/d and of c's rounding are
// all inside e, and what is left is the FFMA's own rounding: t >= u_near - 2^-22 |u_near|, t <= u_far + 2^-22 |u_far|.
// Since the entry parameter is the max of three such t and the exit the min, the same relative widening applied
// once to the max / min gives rigorous bounds.  An axis the ray does not move along (1/d or o/d not
// representable, or NaN) gets inv = 0, c = -+inf: its slab never excludes anything.
struct RayF32 {
  float inv[3], cn[3], cf[3];
  bool neg[3];
};
constexpr float SLAB_REL = 2.384185791015625e-07f;  // 2^-22
RTX_D RayF32 make_ray_f32(V3 pos, V3 dn) {
  RayF32 r;
  const double p[3] = {pos.x, pos.y, pos.z}, d[3] = {dn.x, dn.y, dn.z};
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    r.neg[k] = d[k] < 0.0;
    const double inv = 1.0 / d[k];
    const double oi = p[k] * inv;
    if (fabs(inv) < 1.0e37 && fabs(oi) < 1.0e37) {
      const double e = fabs(oi) * 4.76837158203125e-07 + 1.0e-30;  // 2^-21 |o/d|
      r.inv[k] = __double2float_rn(inv);
      r.cn[k] = __double2float_rd(-oi - e);
      r.cf[k] = __double2float_ru(-oi + e);
    } else {
      r.inv[k] = 0.0f;
      r.cn[k] = -CUDART_INF_F;
      r.cf[k] = CUDART_INF_F;
    }
  }
  return r;
}
// Lower bound of the entry parameter (tn) and upper bound of the exit parameter (tf) of the slab of axis K.
template <int K>
RTX_D void slab(const float* lo, const float* hi, const RayF32& r, float& tn, float& tf) {
  const float xn = r.neg[K] ? hi[K] : lo[K], xf = r.neg[K] ? lo[K] : hi[K];
  tn = __fmaf_rn(xn, r.inv[K], r.cn[K]);
  tf = __fmaf_rn(xf, r.inv[K], r.cf[K]);
}
// visit unless provably empty, provably behind the origin, or provably beyond the best distance
RTX_D bool slab_verdict(float tn, float tf, float best_up, float& tnear) {
  tnear = __fmaf_rn(-fabsf(tn), SLAB_REL, tn);
  const float tfar = __fmaf_rn(fabsf(tf), SLAB_REL, tf);
  return !(tfar < fmaxf(tnear, 0.0f)) && !(tnear > best_up);
}
RTX_D bool box_test(const float* lo, const float* hi, const RayF32& r, float best_up, float& tnear) {
  float tn0, tf0, tn1, tf1, tn2, tf2;
  slab<0>(lo, hi, r, tn0, tf0);
  slab<1>(lo, hi, r, tn1, tf1);
  slab<2>(lo, hi, r, tn2, tf2);
  return slab_verdict(fmaxf(fmaxf(tn0, tn1), tn2), fminf(fminf(tf0, tf1), tf2), best_up, tnear);
}

// Per-ray constants of the prism tree that drops axis A: rigorous f32 enclosures of the origin, and the lower
// bound `den` of |d.n| over the tree's triangles: |d_a|*Z - |d_perp|*S (<= 0: no bound).
struct PrismRay {
  float o_dn[3], o_up[3];
  float den, s, z;
};
template <int A>
RTX_D PrismRay make_prism_ray(const BvhView& bv, V3 pos, V3 dn) {
  constexpr int B = (A + 1) % 3, C = (A + 2) % 3;
  PrismRay pr;
  const double p[3] = {pos.x, pos.y, pos.z}, d[3] = {dn.x, dn.y, dn.z};
#pragma unroll
  for (int k = 0; k < 3; ++k) pr.o_dn[k] = __double2float_rd(p[k]), pr.o_up[k] = __double2float_ru(p[k]);
  pr.s = bv.prism_s[1 + A];
  pr.z = bv.prism_z[1 + A];
  const float da = __double2float_rd(fabs(d[A]) * 0.999999);
  const float dp = __double2float_ru(sqrt(d[B] * d[B] + d[C] * d[C]) * 1.000001);
  pr.den = __fsub_rd(__fmul_rd(da, pr.z), __fmul_ru(dp, pr.s));
  return pr;
}
// Entry test of a node of the prism tree that drops axis A: slabs on the kept axes, then bounds (A) and (B)
// above.  All roundings are directed so that a node is only skipped when the exact-arithmetic statement holds
// with margin; the f64 quantities they bound (|t| as computed by triangle_distance, the reported point) differ
// from the exact ones by ~1e-16 relative, far inside the 1e-6 margins and the 2^-20 box padding.
template <int A>
RTX_D bool box_test_prism(const float* lo, const float* hi, const RayF32& r, const PrismRay& pr, float best_up,
                          float& tnear) {
  constexpr int B = (A + 1) % 3, C = (A + 2) % 3;
  float tnb, tfb, tnc, tfc;
  slab<B>(lo, hi, r, tnb, tfb);
  slab<C>(lo, hi, r, tnc, tfc);
  if (!slab_verdict(fmaxf(tnb, tnc), fminf(tfb, tfc), best_up, tnear)) return false;
  if (!(pr.den > 0.0f)) return true;
  const float db = fmaxf(__fsub_ru(hi[B], pr.o_dn[B]), __fsub_ru(pr.o_up[B], lo[B]));  // >= max |v_b - p_b|
  const float dc = fmaxf(__fsub_ru(hi[C], pr.o_dn[C]), __fsub_ru(pr.o_up[C], lo[C]));
  const float dperp = __fmul_ru(__fsqrt_ru(__fadd_ru(__fmul_ru(db, db), __fmul_ru(dc, dc))), 1.000001f);
  const float sd = __fmul_ru(pr.s, dperp);
  // (A): wholly ahead along a (in the direction the ray travels), and far enough that the a-term decides the
  // sign of n.(v0 - pos)
  const float gap = r.neg[A] ? __fsub_rd(pr.o_dn[A], hi[A]) : __fsub_rd(lo[A], pr.o_up[A]);
  if (gap > 0.0f && __fmul_rd(gap, pr.z) > sd) return false;
  // (B): upper bound of |t|
  const float da = fmaxf(__fsub_ru(hi[A], pr.o_dn[A]), __fsub_ru(pr.o_up[A], lo[A]));
  const float u = __fmul_ru(__fdiv_ru(__fadd_ru(sd, da), pr.den), 1.00001f);
  return !(tnear > u);
}

// HAS_TRI = false: the caller knows the scene has no triangles, so every leaf is a sphere
template <bool PRISM, bool HAS_TRI = true>
RTX_D void bvh_leaf_test(const SceneView& sc, uint32_t child, V3 pos, V3 dir, const RayInv& ri, Hit& h,
                         uint32_t* panics) {
  const uint32_t idx = child & WIDE_IDX;
  double t;
  if (!HAS_TRI || !(child & WIDE_TRI)) {
    if (sphere_distance(sc.sph + (size_t)idx * SPH_STRIDE, pos, ri, &t))
      consider(h, t, sc.sph_id[idx], REF_SPHERE | idx);
  } else {
    const double* rec = sc.tri + (size_t)idx * TRI_STRIDE;
    if (PRISM) {
      // The prism trees only have to find PHANTOM hits (plane behind the ray: the quotient of
      // triangle.rs:34 is negative); a plane in front puts the hit point on the triangle, inside its 3-D
      // box, where the 3-D tree finds it.  Same two dot products triangle_distance starts with.
      const V3 n = mk3(rec[3], rec[4], rec[5]);
      const double num = dot(n, mk3(rec[0], rec[1], rec[2]) - pos), den = dot(ri.dn, n);
      if ((num > 0.0 && den > 0.0) || (num < 0.0 && den < 0.0)) return;
    }
    bool panic = false;
    if (triangle_distance(rec, pos, dir, ri, &t, &panic)) consider(h, t, sc.tri_id[idx], REF_TRIANGLE | idx);
    if (panic) ++*panics;
  }
}

// Near-first walk with a per-thread stack of (child ref, entry bound), in "while-while" form for triangles: an
// inner loop descends until the lane holds a TRIANGLE leaf (or has finished), so the lanes of a warp run the
// FP32 box tests together and then meet for the expensive FP64 triangle tests instead of interleaving the two
// (measured on C4: 10.9 ms single loop, 10.3 ms when every leaf is postponed, 10.8 ms in this form).  Sphere
// leaves are cheap and are tested inside the inner loop: making lanes wait for them costs more than it saves
// (C3: 261 ms inside the loop, 305 ms postponed).
// Pops the next stack entry that can still beat the best distance into `cur`; false when the stack ran dry.
RTX_D bool bvh_pop(const uint32_t* stack_ref, const float* stack_t, int& sp, double best, uint32_t& cur) {
  const float best_up = __double2float_ru(best);
  for (;;) {
    if (sp == 0) return false;
    --sp;
    if (!(stack_t[sp] > best_up)) break;
  }
  cur = stack_ref[sp];
  return true;
}

// AXIS < 0: the 3-D tree; AXIS = 0..2: the prism tree that drops that axis.
template <int AXIS, bool HAS_TRI = true>
RTX_D void bvh_walk(const WideNode* __restrict__ nodes, const SceneView& sc, V3 pos, V3 dir, const RayInv& ri,
                    const RayF32& rf, const PrismRay& pr, Hit& h, uint32_t* panics, TraversalCounters& tc) {
  constexpr bool PRISM = AXIS >= 0;
  constexpr int PA = PRISM ? AXIS : 0;
  uint32_t stack_ref[128];
  float stack_t[128];
  int sp = 0;
  uint32_t cur = 0;  // root
  for (;;) {
    for (;;) {
      if (cur & WIDE_LEAF) {
        if (HAS_TRI && (cur & WIDE_TRI)) break;
        ++tc.prims;
        bvh_leaf_test<PRISM, HAS_TRI>(sc, cur, pos, dir, ri, h, panics);
        if (!bvh_pop(stack_ref, stack_t, sp, h.t, cur)) return;
        continue;
      }
      const WideNode* nd = nodes + cur;
      const float4 q0 = __ldg(reinterpret_cast<const float4*>(nd));
      const float4 q1 = __ldg(reinterpret_cast<const float4*>(nd) + 1);
      const float4 q2 = __ldg(reinterpret_cast<const float4*>(nd) + 2);
      const uint4 q3 = __ldg(reinterpret_cast<const uint4*>(nd) + 3);
      ++tc.nodes;
      const float lo0[3] = {q0.x, q0.y, q0.z}, hi0[3] = {q0.w, q1.x, q1.y};
      const float lo1[3] = {q1.z, q1.w, q2.x}, hi1[3] = {q2.y, q2.z, q2.w};
      const float best_up = __double2float_ru(h.t);
      float t0, t1;
      const bool v0 = PRISM ? box_test_prism<PA>(lo0, hi0, rf, pr, best_up, t0) : box_test(lo0, hi0, rf, best_up, t0);
      const bool v1 = PRISM ? box_test_prism<PA>(lo1, hi1, rf, pr, best_up, t1) : box_test(lo1, hi1, rf, best_up, t1);
      if (v0 && v1) {
        const bool first0 = t0 <= t1;
        // depth <= 64 key bits + 32 index bits (Karras: every level lengthens the common prefix), so 128 holds
        stack_ref[sp] = first0 ? q3.y : q3.x;
        stack_t[sp] = first0 ? t1 : t0;
        ++sp;
        cur = first0 ? q3.x : q3.y;
      } else if (v0) {
        cur = q3.x;
      } else if (v1) {
        cur = q3.y;
      } else if (!bvh_pop(stack_ref, stack_t, sp, h.t, cur)) {
        return;
      }
    }
    ++tc.prims;
    bvh_leaf_test<PRISM, HAS_TRI>(sc, cur, pos, dir, ri, h, panics);
    if (!bvh_pop(stack_ref, stack_t, sp, h.t, cur)) return;
  }
}

// scene.rs:243-251 through the trees: planes (unbounded, kept outside) by scan, then the 3-D tree, then the
// prism trees.  Same Hit as closest_brute for every ray.
// HAS_TRI = false (sphere-only scenes, e.g. C3): the prism trees, the triangle leaf test and their registers are
// compiled out.
template <bool HAS_TRI = true>
RTX_D Hit closest_lbvh(const SceneView& sc, const BvhView& bv, V3 pos, V3 dir, uint32_t* panics,
                       TraversalCounters& tc) {
  Hit h;
  h.t = CUDART_INF;
  h.id = MISS_ID;
  h.ref = 0;
  const RayInv ri = ray_invariants(dir);
  for (uint32_t i = 0; i < sc.n_pl; ++i) {
    double t;
    if (plane_distance(sc.pl + (size_t)i * PL_STRIDE, pos, ri, &t)) consider(h, t, sc.pl_id[i], REF_PLANE | i);
  }
  tc.prims += sc.n_pl;
  // slab parameters are in units of dn, like every distance the shapes return
  const RayF32 rf = make_ray_f32(pos, ri.dn);
  PrismRay pr;
  pr.den = 0.0f, pr.s = 1.0f, pr.z = 0.0f;
  if (bv.root[0] != BVH_END) bvh_walk<-1, HAS_TRI>(bv.nodes[0], sc, pos, dir, ri, rf, pr, h, panics, tc);
  if (!HAS_TRI) return h;
  if (bv.root[1] != BVH_END) {
    pr = make_prism_ray<0>(bv, pos, ri.dn);
    bvh_walk<0>(bv.nodes[1], sc, pos, dir, ri, rf, pr, h, panics, tc);
  }
  if (bv.root[2] != BVH_END) {
    pr = make_prism_ray<1>(bv, pos, ri.dn);
    bvh_walk<1>(bv.nodes[2], sc, pos, dir, ri, rf, pr, h, panics, tc);
  }
  if (bv.root[3] != BVH_END) {
    pr = make_prism_ray<2>(bv, pos, ri.dn);
    bvh_walk<2>(bv.nodes[3], sc, pos, dir, ri, rf, pr, h, panics, tc);
  }
  return h;
}
#endif

}  // namespace rtx
