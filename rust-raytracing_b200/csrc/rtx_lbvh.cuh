// rtx_lbvh.cuh — LBVH node layout and the stackless traversal (K5).
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
constexpr uint32_t WIDE_NONE = 0x7FFFFFFFu;  // empty child slot
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

// Per-ray f32 interval data for the conservative slab test.  Axes with a negative direction are mirrored
// (x -> -x), so every component behaves as d >= 0.  o_dn <= o <= o_up and inv_dn <= 1/|d| <= inv_up hold
// rigorously (directed rounding), inv_dn is finite.
struct RayF32 {
  float o_dn[3], o_up[3], inv_dn[3], inv_up[3];
  bool neg[3];
};
RTX_D RayF32 make_ray_f32(V3 pos, V3 dn) {
  RayF32 r;
  const double p[3] = {pos.x, pos.y, pos.z}, d[3] = {dn.x, dn.y, dn.z};
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    r.neg[k] = d[k] < 0.0;
    const double o = r.neg[k] ? -p[k] : p[k];
    r.o_dn[k] = __double2float_rd(o);
    r.o_up[k] = __double2float_ru(o);
    const double inv = 1.0 / fabs(d[k]);  // +inf for d == 0; correctly rounded, so widen by one f32 step each way
    r.inv_dn[k] = fminf(__double2float_rd(inv * 0.99999999), 3.0e38f);
    r.inv_up[k] = __double2float_ru(inv * 1.00000001);
  }
  return r;
}
// Lower bound of the entry parameter and upper bound of the exit parameter of one (mirrored) slab.
RTX_D void slab(float lo, float hi, bool neg, float o_dn, float o_up, float inv_dn, float inv_up, float& tn, float& tf) {
  const float l = neg ? -hi : lo, h = neg ? -lo : hi;
  const float a = __fsub_rd(l, o_up);  // <= l - o
  const float b = __fsub_ru(h, o_dn);  // >= h - o
  tn = a < 0.0f ? __fmul_rd(a, inv_up) : __fmul_rd(a, inv_dn);
  tf = b > 0.0f ? __fmul_ru(b, inv_up) : __fmul_ru(b, inv_dn);
}
RTX_D bool box_test(const float* lo, const float* hi, const RayF32& r, float best_up, float& tnear) {
  float tn0, tf0, tn1, tf1, tn2, tf2;
  slab(lo[0], hi[0], r.neg[0], r.o_dn[0], r.o_up[0], r.inv_dn[0], r.inv_up[0], tn0, tf0);
  slab(lo[1], hi[1], r.neg[1], r.o_dn[1], r.o_up[1], r.inv_dn[1], r.inv_up[1], tn1, tf1);
  slab(lo[2], hi[2], r.neg[2], r.o_dn[2], r.o_up[2], r.inv_dn[2], r.inv_up[2], tn2, tf2);
  tnear = fmaxf(fmaxf(tn0, tn1), tn2);
  const float tfar = fminf(fminf(tf0, tf1), tf2);
  // visit unless provably empty, provably behind the origin, or provably beyond the best distance
  return !(tfar < fmaxf(tnear, 0.0f)) && !(tnear > best_up);
}

// Per-ray constants of one prism tree (dropped axis a).
struct PrismRay {
  int a, b, c;      // dropped axis, kept axes
  float den;        // lower bound of |d.n| over the tree's triangles: |d_a|*Z - |d_perp|*S (<= 0: no bound)
  float s, z;       // the tree's S and Z
  bool sign_known;  // all n_a of one sign
};
RTX_D PrismRay make_prism_ray(const BvhView& bv, int tree, V3 dn) {
  PrismRay pr;
  pr.a = tree - 1;
  pr.b = (pr.a + 1) % 3;
  pr.c = (pr.a + 2) % 3;
  const double d[3] = {dn.x, dn.y, dn.z};
  pr.s = bv.prism_s[tree];
  pr.z = bv.prism_z[tree];
  pr.sign_known = bv.prism_sign[tree] != 0;
  const float da = __double2float_rd(fabs(d[pr.a]) * 0.999999);
  const float dp = __double2float_ru(sqrt(d[pr.b] * d[pr.b] + d[pr.c] * d[pr.c]) * 1.000001);
  pr.den = __fsub_rd(__fmul_rd(da, pr.z), __fmul_ru(dp, pr.s));
  return pr;
}
// Entry test of a prism-tree node: slabs on the kept axes, then bounds (A) and (B) above.  All roundings are
// directed so that a node is only skipped when the exact-arithmetic statement holds with margin; the f64
// quantities they bound (|t| as computed by triangle_distance, the reported point) differ from the exact ones
// by ~1e-16 relative, far inside the 1e-6 margins and the 2^-20 box padding.
RTX_D bool box_test_prism(const float* lo, const float* hi, const RayF32& r, const PrismRay& pr, float best_up, float& tnear) {
  float tnb, tfb, tnc, tfc;
  slab(lo[pr.b], hi[pr.b], r.neg[pr.b], r.o_dn[pr.b], r.o_up[pr.b], r.inv_dn[pr.b], r.inv_up[pr.b], tnb, tfb);
  slab(lo[pr.c], hi[pr.c], r.neg[pr.c], r.o_dn[pr.c], r.o_up[pr.c], r.inv_dn[pr.c], r.inv_up[pr.c], tnc, tfc);
  tnear = fmaxf(tnb, tnc);
  const float tfar = fminf(tfb, tfc);
  if ((tfar < fmaxf(tnear, 0.0f)) || (tnear > best_up)) return false;
  if (!(pr.den > 0.0f)) return true;
  // mirrored coordinates (the ray travels towards + on every axis): box [l, h], origin in [o_dn, o_up]
  const float lb = r.neg[pr.b] ? -hi[pr.b] : lo[pr.b], hb = r.neg[pr.b] ? -lo[pr.b] : hi[pr.b];
  const float lc = r.neg[pr.c] ? -hi[pr.c] : lo[pr.c], hc = r.neg[pr.c] ? -lo[pr.c] : hi[pr.c];
  const float la = r.neg[pr.a] ? -hi[pr.a] : lo[pr.a], ha = r.neg[pr.a] ? -lo[pr.a] : hi[pr.a];
  const float db = fmaxf(__fsub_ru(hb, r.o_dn[pr.b]), __fsub_ru(r.o_up[pr.b], lb));  // >= max |v_b - p_b|
  const float dc = fmaxf(__fsub_ru(hc, r.o_dn[pr.c]), __fsub_ru(r.o_up[pr.c], lc));
  const float dperp = __fmul_ru(__fsqrt_ru(__fadd_ru(__fmul_ru(db, db), __fmul_ru(dc, dc))), 1.000001f);
  const float sd = __fmul_ru(pr.s, dperp);
  // (A): wholly ahead along a, and far enough that the a-term decides the sign of n.(v0 - pos)
  const float gap = __fsub_rd(la, r.o_up[pr.a]);
  if (gap > 0.0f && __fmul_rd(gap, pr.z) > sd) return false;
  // (B): upper bound of |t|
  const float da = fmaxf(__fsub_ru(ha, r.o_dn[pr.a]), __fsub_ru(r.o_up[pr.a], la));
  const float u = __fmul_ru(__fdiv_ru(__fadd_ru(sd, da), pr.den), 1.00001f);
  return !(tnear > u);
}

template <bool PRISM>
RTX_D void bvh_leaf_test(const SceneView& sc, uint32_t child, V3 pos, V3 dir, const RayInv& ri, Hit& h,
                         uint32_t* panics) {
  const uint32_t idx = child & WIDE_IDX;
  double t;
  if (!(child & WIDE_TRI)) {
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

template <bool PRISM>
RTX_D void bvh_walk(const WideNode* __restrict__ nodes, const SceneView& sc, V3 pos, V3 dir, const RayInv& ri,
                    const RayF32& rf, const PrismRay& pr, Hit& h, uint32_t* panics, TraversalCounters& tc) {
  uint32_t stack_ref[128];
  float stack_t[128];
  int sp = 0;
  uint32_t cur = 0;  // root
  for (;;) {
    for (;;) {
      if (cur & WIDE_LEAF) {
        if (cur & WIDE_TRI) break;
        ++tc.prims;
        bvh_leaf_test<PRISM>(sc, cur, pos, dir, ri, h, panics);
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
      const bool v0 = PRISM ? box_test_prism(lo0, hi0, rf, pr, best_up, t0) : box_test(lo0, hi0, rf, best_up, t0);
      const bool v1 = PRISM ? box_test_prism(lo1, hi1, rf, pr, best_up, t1) : box_test(lo1, hi1, rf, best_up, t1);
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
    bvh_leaf_test<PRISM>(sc, cur, pos, dir, ri, h, panics);
    if (!bvh_pop(stack_ref, stack_t, sp, h.t, cur)) return;
  }
}

// scene.rs:243-251 through the trees: planes (unbounded, kept outside) by scan, then the 3-D tree, then the
// prism trees.  Same Hit as closest_brute for every ray.
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
  pr.a = 0, pr.b = 1, pr.c = 2, pr.den = 0.0f, pr.s = 1.0f, pr.z = 0.0f, pr.sign_known = false;
  if (bv.root[0] != BVH_END) bvh_walk<false>(bv.nodes[0], sc, pos, dir, ri, rf, pr, h, panics, tc);
#pragma unroll 1
  for (int k = 1; k < 4; ++k)
    if (bv.root[k] != BVH_END) {
      pr = make_prism_ray(bv, k, ri.dn);
      bvh_walk<true>(bv.nodes[k], sc, pos, dir, ri, rf, pr, h, panics, tc);
    }
  return h;
}
#endif

}  // namespace rtx
