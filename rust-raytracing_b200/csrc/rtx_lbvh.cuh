// rtx_lbvh.cuh — LBVH node layout and the stackless traversal (K5).
//
// The reference has no acceleration structure: closest_object is an O(N) scan (raytracing/scene.rs:243-251).
// The LBVH must return EXACTLY what that scan returns: the same distance bits and, among equal distances,
// the lowest object index.  So the tree only culls; every surviving primitive goes through the same exact
// f64 distance routine as the brute-force kernel (rtx_trace.cuh), and culling is conservative:
//  * boxes are f64, padded outward by `pad` (1e-9 x scene scale, ~1e7 x the rounding error of a hit point);
//  * a node is skipped only if the padded slab interval is empty, ends behind the origin, or starts
//    beyond the current best distance (strictly: equal distances must still be seen for the tie rule);
//  * NaN slabs (0 * inf) are ignored by fmin/fmax, i.e. treated as "no constraint".
// Triangles need a second kind of tree.  A "phantom" hit (plane BEHIND the ray, distance = |t|,
// triangle.rs:118) reports the mirrored point pos + dir*|t|, which is off the plane and only has to lie in
// the infinite prism over the triangle's projection along the coordinate contains() drops
// (triangle.rs:96-100).  It is bounded by the triangle's box with that axis opened to (-inf, +inf).  One
// such "prism tree" per dropped axis is traversed after the 3-D tree, pruned by the best distance so far.
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

struct BvhView {
  const BvhNode* nodes[4];  // [0] 3-D tree over spheres + triangles; [1..3] prism trees (dropped axis x,y,z)
  uint32_t root[4];         // BVH_END when the tree is empty
};

#if defined(__CUDACC__)
struct TraversalCounters {
  unsigned long long nodes, prims;
};

RTX_D void bvh_leaf_test(const SceneView& sc, uint32_t ref, V3 pos, V3 dir, const RayInv& ri, Hit& h,
                         uint32_t* panics) {
  const uint32_t idx = ref & REF_INDEX_MASK;
  double t;
  if ((ref & ~REF_INDEX_MASK) == REF_SPHERE) {
    if (sphere_distance(sc.sph + (size_t)idx * SPH_STRIDE, pos, ri, &t)) consider(h, t, sc.sph_id[idx], ref);
  } else {
    bool panic = false;
    if (triangle_distance(sc.tri + (size_t)idx * TRI_STRIDE, pos, dir, ri, &t, &panic))
      consider(h, t, sc.tri_id[idx], ref);
    if (panic) ++*panics;
  }
}

// Skip-pointer ("escape index") walk: no stack, fixed depth-first order.
RTX_D void bvh_walk(const BvhNode* __restrict__ nodes, uint32_t root, const SceneView& sc, V3 pos, V3 dir,
                    const RayInv& ri, V3 inv, Hit& h, uint32_t* panics, TraversalCounters& tc) {
  uint32_t cur = root;
  while (cur != BVH_END) {
    const BvhNode* nd = nodes + cur;
    const double2 a0 = __ldg(reinterpret_cast<const double2*>(nd));      // four 128-bit loads per node
    const double2 a1 = __ldg(reinterpret_cast<const double2*>(nd) + 1);
    const double2 b = __ldg(reinterpret_cast<const double2*>(nd) + 2);
    const uint4 c = __ldg(reinterpret_cast<const uint4*>(nd) + 3);       // left escape prim pad
    struct { double x, y, z, w; } a = {a0.x, a0.y, a1.x, a1.y};          // lo.xyz hi.x ; b = hi.yz
    ++tc.nodes;
    const double t1x = (a.x - pos.x) * inv.x, t2x = (a.w - pos.x) * inv.x;
    const double t1y = (a.y - pos.y) * inv.y, t2y = (b.x - pos.y) * inv.y;
    const double t1z = (a.z - pos.z) * inv.z, t2z = (b.y - pos.z) * inv.z;
    const double tmin = fmax(fmax(fmin(t1x, t2x), fmin(t1y, t2y)), fmin(t1z, t2z));
    const double tmax = fmin(fmin(fmax(t1x, t2x), fmax(t1y, t2y)), fmax(t1z, t2z));
    // !(miss) form so that a NaN tmin/tmax (all three slabs NaN) falls through to "visit"
    const bool miss = (tmax < fmax(tmin, 0.0)) || (tmin > h.t);
    if (miss) {
      cur = c.y;
    } else if (c.z != BVH_NO_PRIM) {
      ++tc.prims;
      bvh_leaf_test(sc, c.z, pos, dir, ri, h, panics);
      cur = c.y;
    } else {
      cur = c.x;
    }
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
  const V3 inv = mk3(1.0 / ri.dn.x, 1.0 / ri.dn.y, 1.0 / ri.dn.z);
#pragma unroll 1
  for (int k = 0; k < 4; ++k)
    if (bv.root[k] != BVH_END) bvh_walk(bv.nodes[k], bv.root[k], sc, pos, dir, ri, inv, h, panics, tc);
  return h;
}
#endif

}  // namespace rtx
