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
// point may be anywhere along a), pruned by the best distance so far and by the gradient bound of
// prism_gradient_test below: per child, bounds of the plane gradients of the triangles under it turn "is there a
// triangle here whose plane, cut BEHIND the origin, reports a point over this box?" into interval arithmetic.
// (Round 1 had two tree-wide bounds instead — (A) "the node lies wholly ahead, so every plane is in front" and (B) an
// upper bound of |t| from tree-wide normal bounds S, Z; the per-child gradient bound implies both and costs neither a
// square root nor a division per child: C4 8.35 -> 4.47 ms.)  Without any such bound a ray that leaves a height field
// upwards walks the footprint of every triangle under its path.
#pragma once
#include "rtx_trace.cuh"

namespace rtx {

constexpr uint32_t BVH_END = 0xFFFFFFFFu;
constexpr uint32_t BVH_NO_PRIM = 0xFFFFFFFFu;

struct __align__(16) BvhNode {  // f64 build node (build-time only)
  double lo[3];
  double hi[3];
  double glo[2], ghi[2];  // prism trees: bounds of the plane gradients (d a / d b, d a / d c) of the triangles below
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
  // prism trees: per node two float4, one per child: (gb_lo, gb_hi, gc_lo, gc_hi) — outward-rounded bounds of the plane
  // gradients g = -n_perp / n_a of the triangles below that child (bound (C) of box_test_prism); +-inf = no bound
  const float4* aux[4];
  uint32_t root[4];          // BVH_END when the tree is empty, else 0
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
RTX_D bool slab_verdict(float tn, float tf, float best_up, float& tnear, float& tfar) {
  tnear = __fmaf_rn(-fabsf(tn), SLAB_REL, tn);
  tfar = __fmaf_rn(fabsf(tf), SLAB_REL, tf);
  return !(tfar < fmaxf(tnear, 0.0f)) && !(tnear > best_up);
}
RTX_D bool slab_verdict(float tn, float tf, float best_up, float& tnear) {
  float tfar;
  return slab_verdict(tn, tf, best_up, tnear, tfar);
}
// [a_lo, a_hi] * [b_lo, b_hi] with outward rounding (finite operands)
RTX_D void imul(float a_lo, float a_hi, float b_lo, float b_hi, float& lo, float& hi) {
  lo = fminf(fminf(__fmul_rd(a_lo, b_lo), __fmul_rd(a_lo, b_hi)), fminf(__fmul_rd(a_hi, b_lo), __fmul_rd(a_hi, b_hi)));
  hi = fmaxf(fmaxf(__fmul_ru(a_lo, b_lo), __fmul_ru(a_lo, b_hi)), fmaxf(__fmul_ru(a_hi, b_lo), __fmul_ru(a_hi, b_hi)));
}
RTX_D bool box_test(const float* lo, const float* hi, const RayF32& r, float best_up, float& tnear) {
  float tn0, tf0, tn1, tf1, tn2, tf2;
  slab<0>(lo, hi, r, tn0, tf0);
  slab<1>(lo, hi, r, tn1, tf1);
  slab<2>(lo, hi, r, tn2, tf2);
  return slab_verdict(fmaxf(fmaxf(tn0, tn1), tn2), fminf(fminf(tf0, tf1), tf2), best_up, tnear);
}

// Per-ray constants of the prism walks: rigorous f32 enclosures of the origin and of the unit direction.
struct PrismRay {
  float o_dn[3], o_up[3];  // enclosures of the origin
  float d_dn[3], d_up[3];  // enclosures of the unit direction
};
RTX_D PrismRay make_prism_ray(V3 pos, V3 dn) {
  PrismRay pr;
  const double p[3] = {pos.x, pos.y, pos.z}, d[3] = {dn.x, dn.y, dn.z};
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    pr.o_dn[k] = __double2float_rd(p[k]), pr.o_up[k] = __double2float_ru(p[k]);
    pr.d_dn[k] = __double2float_rd(d[k]), pr.d_up[k] = __double2float_ru(d[k]);
  }
  return pr;
}
// Entry test of a child of a node of the prism tree that drops axis A: slabs on the kept axes, then the gradient
// bound.  All roundings are directed so that a child is only skipped when the exact-arithmetic statement holds with
// margin; the f64 quantities they bound (|t| as computed by triangle_distance, the reported point) differ from the
// exact ones by ~1e-16 relative, far inside the margins and the 2^-20 box padding.
// The gradient bound.  Write a triangle's plane over the kept axes as a height function along
// a with gradient g = -(n_b, n_c) / n_a.  Then n.(v0 - pos) = n_a * gap and d.n = n_a * slope, with
//   gap   = (v0 - pos)_a - g.(v0 - pos)_perp     (how far the plane, extended to pos, lies above pos along a)
//   slope = d_a - g.d_perp
// so the quotient of triangle.rs:34 is t = gap / slope — n_a cancels — and a phantom hit needs t < 0 with the reported
// point at parameter |t| over the node: |t| inside the kept-axes window [tnear, tfar] (and <= the best distance).
// With v0 inside the child's box and g inside the child's gradient box, gap and slope are intervals; the child is
// skipped when no (gap, slope) in them gives such a t.  All interval arithmetic is f32 with directed rounding on enclosures of
// pos, d and the padded boxes; the f64 evaluation it brackets (triangle_distance) differs from exact arithmetic by
// ~1e-14 absolute, the box padding alone is 2^-20 of the scene scale.
template <int A>
RTX_D bool prism_gradient_test(const float* lo, const float* hi, const PrismRay& pr, float4 g, float tnear, float tfar,
                               float best_up) {
  constexpr int B = (A + 1) % 3, C = (A + 2) % 3;
  if (!(fabsf(g.x) < 1e30f && fabsf(g.y) < 1e30f && fabsf(g.z) < 1e30f && fabsf(g.w) < 1e30f)) return true;  // no bound
  float pb_lo, pb_hi, pc_lo, pc_hi, qb_lo, qb_hi, qc_lo, qc_hi;
  imul(g.x, g.y, __fsub_rd(lo[B], pr.o_up[B]), __fsub_ru(hi[B], pr.o_dn[B]), pb_lo, pb_hi);
  imul(g.z, g.w, __fsub_rd(lo[C], pr.o_up[C]), __fsub_ru(hi[C], pr.o_dn[C]), pc_lo, pc_hi);
  const float gap_lo = __fsub_rd(__fsub_rd(lo[A], pr.o_up[A]), __fadd_ru(pb_hi, pc_hi));
  const float gap_hi = __fsub_ru(__fsub_ru(hi[A], pr.o_dn[A]), __fadd_rd(pb_lo, pc_lo));
  imul(g.x, g.y, pr.d_dn[B], pr.d_up[B], qb_lo, qb_hi);
  imul(g.z, g.w, pr.d_dn[C], pr.d_up[C], qc_lo, qc_hi);
  const float sl_lo = __fsub_rd(pr.d_dn[A], __fadd_ru(qb_hi, qc_hi));
  const float sl_hi = __fsub_ru(pr.d_up[A], __fadd_rd(qb_lo, qc_lo));
  // |t| = G / D with G = -+gap > 0 and D = +-slope > 0
  float G_lo, G_hi, D_lo, D_hi;
  if (sl_lo > 0.0f) {
    G_lo = -gap_hi, G_hi = -gap_lo, D_lo = sl_lo, D_hi = sl_hi;
  } else if (sl_hi < 0.0f) {
    G_lo = gap_lo, G_hi = gap_hi, D_lo = -sl_hi, D_hi = -sl_lo;
  } else {
    return true;  // the slope interval holds 0 (or is NaN): |t| is unbounded
  }
  if (G_hi <= 0.0f) return false;  // t >= 0 for every triangle below: ordinary hits only, the 3-D tree finds them
  // (a NaN anywhere — inf - inf after an overflow — makes every comparison here false: the child is kept)
  const float w_lo = fmaxf(tnear, 0.0f), w_hi = fminf(tfar, best_up);
  // |t| <= G_hi / D_lo < w_lo,  or  |t| >= max(G_lo, 0) / D_hi > w_hi
  if (__fmul_ru(G_hi, 1.00001f) < __fmul_rd(w_lo, D_lo)) return false;
  if (__fmul_rd(fmaxf(G_lo, 0.0f), 0.99999f) > __fmul_ru(w_hi, D_hi)) return false;
  return true;
}

template <int A>
RTX_D bool box_test_prism(const float* lo, const float* hi, const RayF32& r, const PrismRay& pr, float best_up,
                          float& tnear, float4 g) {
  constexpr int B = (A + 1) % 3, C = (A + 2) % 3;
  float tnb, tfb, tnc, tfc, tfar;
  slab<B>(lo, hi, r, tnb, tfb);
  slab<C>(lo, hi, r, tnc, tfc);
  if (!slab_verdict(fmaxf(tnb, tnc), fminf(tfb, tfc), best_up, tnear, tfar)) return false;
  return prism_gradient_test<A>(lo, hi, pr, g, tnear, tfar, best_up);
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
                    const RayF32& rf, const PrismRay& pr, Hit& h, uint32_t* panics, TraversalCounters& tc,
                    const float4* __restrict__ aux = nullptr) {
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
      float4 g0 = make_float4(0.f, 0.f, 0.f, 0.f), g1 = g0;
      if (PRISM) g0 = __ldg(aux + 2u * cur), g1 = __ldg(aux + 2u * cur + 1u);
      const bool v0 = PRISM ? box_test_prism<PA>(lo0, hi0, rf, pr, best_up, t0, g0) : box_test(lo0, hi0, rf, best_up, t0);
      const bool v1 = PRISM ? box_test_prism<PA>(lo1, hi1, rf, pr, best_up, t1, g1) : box_test(lo1, hi1, rf, best_up, t1);
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
  PrismRay pr = make_prism_ray(pos, ri.dn);
  if (bv.root[0] != BVH_END) bvh_walk<-1, HAS_TRI>(bv.nodes[0], sc, pos, dir, ri, rf, pr, h, panics, tc);
  if (!HAS_TRI) return h;
  if (bv.root[1] != BVH_END) {
    bvh_walk<0>(bv.nodes[1], sc, pos, dir, ri, rf, pr, h, panics, tc, bv.aux[1]);
  }
  if (bv.root[2] != BVH_END) {
    bvh_walk<1>(bv.nodes[2], sc, pos, dir, ri, rf, pr, h, panics, tc, bv.aux[2]);
  }
  if (bv.root[3] != BVH_END) {
    bvh_walk<2>(bv.nodes[3], sc, pos, dir, ri, rf, pr, h, panics, tc, bv.aux[3]);
  }
  return h;
}
#endif

}  // namespace rtx
