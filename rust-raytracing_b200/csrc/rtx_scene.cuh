// rtx_scene.cuh — device scene layout (one contiguous blob in HBM, SoA per shape type).
//
// The reference keeps `Vec<Object>` with each shape behind Arc<Mutex<dyn CustomShape>>
// (raytracing/object.rs:10-15).  Here the three built-in shapes are split by type into dense records of
// doubles, 16-byte aligned so a record is fetched with 128-bit loads, plus one u32 object-index array
// per type (the object index is the nearest-hit tie-break, scene.rs:250) and one material table indexed
// by object index.  Everything that depends only on the shape is precomputed at upload with the same
// non-fused f64 operations the reference would execute per ray, so the per-ray results stay
// bit-identical:
//   sphere   : radius*radius                                   (sphere.rs:24)
//   plane    : normal.norm()                                   (plane.rs:22, object.rs:38)
//   triangle : normal = norm(cross(v1-v0, v2-v0)), norm(normal) (triangle.rs:104-107, object.rs:38),
//              and the ray-independent part of the Gauss-Jordan elimination in contains()
//              (triangle.rs:54-94): pivot selection, row swaps and elimination factors.
#pragma once
#include "rtx_math.cuh"

namespace rtx {

constexpr int SPH_STRIDE = 4;   // cx, cy, cz, r*r
constexpr int PL_STRIDE = 10;   // pos.xyz, normal.xyz (raw), nhat.xyz, pad
constexpr int TRI_STRIDE = 14;  // v0.xyz, n.xyz, nn.xyz, r1, l1y, r2, d2, code(bits)
constexpr int MAT_STRIDE = 8;   // base.xyz, emission.xyz, roughness, flags(bits)

// Triangle `code` bits (stored in the bit pattern of the 14th double).
constexpr uint64_t TRI_I1_SHIFT = 0;   // 2 bits: coordinate index of elimination row 1
constexpr uint64_t TRI_I2_SHIFT = 2;   // 2 bits: coordinate index of elimination row 2
constexpr uint64_t TRI_DEGENERATE = 1ull << 4;  // "can't handle LGS" → contains() is always false
constexpr uint64_t TRI_PANIC = 1ull << 5;       // reference's assert_eq! would abort (non-finite rows)

// Material flag bits.
constexpr uint64_t MAT_TERMINAL = 1ull << 0;  // base_color == (0,0,0): light_color becomes zero → path ends

// Hit reference: shape type in the top 2 bits, index within the type's array below.
constexpr uint32_t REF_SPHERE = 0u << 30, REF_PLANE = 1u << 30, REF_TRIANGLE = 2u << 30;
constexpr uint32_t REF_INDEX_MASK = (1u << 30) - 1u;
constexpr uint32_t MISS_ID = 0xFFFFFFFFu;

// Byte offsets of the arrays inside the blob (all multiples of 16).
struct SceneLayout {
  uint32_t n_obj, n_sph, n_pl, n_tri;
  uint32_t off_sph, off_pl, off_tri, off_mat;
  uint32_t off_sph_id, off_pl_id, off_tri_id;
  uint32_t off_sph32;  // float4 per sphere (cx, cy, cz, r*r rounded to f32): conservative cull of the scan only
  uint32_t bytes;  // total blob size, multiple of 16
  // Scene-wide bounds for the f32 cull (rtx_trace.cuh, SphereCull): max |centre component| and max r*r over
  // all spheres, rounded up; +inf when a sphere is not finite (then nothing is ever culled).
  float sph_cmax, sph_r2max;
  // Per launch (cull_constants, rtx_api.cu): the cull's two margins for every ray whose origin has no component larger
  // than cull_pmax — computed once on the host from the bounds R = cull_pmax + sph_cmax, D1 <= sqrt(3), a <= 1 + 1e-6
  // instead of per ray from the ray's own |p|, D1 and a.  Larger R, D1, a only widen the margins (SphereCull's error
  // bounds are monotone in them), so the cull stays a proof; an origin beyond cull_pmax (a far hit on an unbounded
  // plane) switches the cull off for that ray.  cull_pmax = 0: per-ray margins (debug kernels).
  float cull_pmax, cull_neg_mq, cull_mh;
};

struct SceneView {
  const double* sph;
  const double* pl;
  const double* tri;
  const double* mat;
  const uint32_t* sph_id;
  const uint32_t* pl_id;
  const uint32_t* tri_id;
  const float* sph32;
  float sph_cmax, sph_r2max;
  float cull_pmax, cull_neg_mq, cull_mh;
  uint32_t n_obj, n_sph, n_pl, n_tri;
};

RTX_HD SceneView make_view(const unsigned char* base, const SceneLayout& L) {
  SceneView v;
  v.sph = reinterpret_cast<const double*>(base + L.off_sph);
  v.pl = reinterpret_cast<const double*>(base + L.off_pl);
  v.tri = reinterpret_cast<const double*>(base + L.off_tri);
  v.mat = reinterpret_cast<const double*>(base + L.off_mat);
  v.sph_id = reinterpret_cast<const uint32_t*>(base + L.off_sph_id);
  v.pl_id = reinterpret_cast<const uint32_t*>(base + L.off_pl_id);
  v.tri_id = reinterpret_cast<const uint32_t*>(base + L.off_tri_id);
  v.sph32 = reinterpret_cast<const float*>(base + L.off_sph32);
  v.sph_cmax = L.sph_cmax;
  v.sph_r2max = L.sph_r2max;
  v.cull_pmax = L.cull_pmax, v.cull_neg_mq = L.cull_neg_mq, v.cull_mh = L.cull_mh;
  v.n_obj = L.n_obj;
  v.n_sph = L.n_sph;
  v.n_pl = L.n_pl;
  v.n_tri = L.n_tri;
  return v;
}

}  // namespace rtx
