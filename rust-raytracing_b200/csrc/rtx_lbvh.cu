// rtx_lbvh.cu — GPU LBVH build (K4): AABBs → Morton codes → radix sort → Karras hierarchy → bottom-up refit
// → two-box f32 traversal nodes.  No reference counterpart (the reference scans all objects, scene.rs:243-251); the gate
// is "hit ids and distances identical to the brute-force kernel".  Everything here is our own kernel code
// (no CUB/Thrust): the sort is an 8-bit LSD radix sort, one warp per 1024-key tile, stable by construction.
#include <cuda_runtime.h>
#include <math_constants.h>

#include <cmath>
#include <cstring>
#include <limits>
#include <vector>

#include "rtx_internal.h"
#include "rtx_lbvh.cuh"

namespace rtx {

constexpr uint32_t FULLMASK = 0xFFFFFFFFu;
constexpr int SORT_TILE = 1024;  // keys per warp

__device__ __forceinline__ void atomic_min_f64(double* addr, double v) {
  unsigned long long* a = reinterpret_cast<unsigned long long*>(addr);
  unsigned long long old = *a;
  while (v < __longlong_as_double((long long)old)) {
    const unsigned long long assumed = old;
    old = atomicCAS(a, assumed, (unsigned long long)__double_as_longlong(v));
    if (old == assumed) break;
  }
}
__device__ __forceinline__ void atomic_max_f64(double* addr, double v) {
  unsigned long long* a = reinterpret_cast<unsigned long long*>(addr);
  unsigned long long old = *a;
  while (v > __longlong_as_double((long long)old)) {
    const unsigned long long assumed = old;
    old = atomicCAS(a, assumed, (unsigned long long)__double_as_longlong(v));
    if (old == assumed) break;
  }
}

// ---- 1. per-primitive boxes + scene bounds ------------------------------------------------------------------
// prims[i] = REF_* | index.  open_axis in {0,1,2}: prism tree over that dropped axis (it contributes nothing to
// the Morton code); 3: the 3-D tree.
// Primitives that can never be hit (non-finite data) get valid = 0: they are left out of the tree.
// bounds[0..2] = min, [3..5] = max over the finite box corners (used for Morton quantisation and padding).
__global__ void k_prim_boxes(const double* __restrict__ sph /* cx cy cz r2 */, const double* __restrict__ tri_raw /* 9/tri */,
                             const uint32_t* __restrict__ prims, uint32_t n, int open_axis, double* __restrict__ lo,
                             double* __restrict__ hi, uint8_t* __restrict__ valid, double* __restrict__ bounds,
                             double* __restrict__ grad /* prism trees: [n][2] plane gradient (d a/d b, d a/d c); else null */) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t ref = prims[i], idx = ref & REF_INDEX_MASK;
  double l[3], h[3];
  if ((ref & ~REF_INDEX_MASK) == REF_SPHERE) {
    const double* r = sph + (size_t)idx * SPH_STRIDE;
    const double rad = sqrt(r[3]) * (1.0 + 1e-15);  // >= |radius| (r2 = fl(radius*radius))
    for (int k = 0; k < 3; ++k) l[k] = r[k] - rad, h[k] = r[k] + rad;
  } else {
    const double* v = tri_raw + (size_t)idx * 9;
    for (int k = 0; k < 3; ++k) {
      l[k] = fmin(v[k], fmin(v[3 + k], v[6 + k]));
      h[k] = fmax(v[k], fmax(v[3 + k], v[6 + k]));
      if (!(v[k] == v[k]) || !(v[3 + k] == v[3 + k]) || !(v[6 + k] == v[6 + k])) l[k] = CUDART_NAN;
    }
  }
  bool ok = true;
  for (int k = 0; k < 3; ++k) ok = ok && (l[k] - l[k] == 0.0) && (h[k] - h[k] == 0.0);  // finite
  valid[i] = ok ? 1 : 0;
  if (!ok) {
    for (int k = 0; k < 3; ++k) l[k] = CUDART_INF, h[k] = -CUDART_INF;
  } else {
    for (int k = 0; k < 3; ++k) {
      atomic_min_f64(&bounds[k], l[k]);
      atomic_max_f64(&bounds[3 + k], h[k]);
    }
    // prism trees keep the true extent along the dropped axis too: the traversal ignores it for the entry
    // test (the reported point of a phantom hit may be anywhere along it) but uses it for its |t| bounds
  }
  for (int k = 0; k < 3; ++k) lo[(size_t)i * 3 + k] = l[k], hi[(size_t)i * 3 + k] = h[k];
  if (grad) {
    // the triangle's plane as a height function along the dropped axis a over the kept axes b, c:
    // gradient = -(n_b, n_c) / n_a with n = cross(v1 - v0, v2 - v0) (triangle.rs:104-107 up to normalisation).
    // A plane parallel to a (n_a = 0) or non-finite data: NaN here = "no bound" (k_refit turns it into +-inf).
    double gb = CUDART_NAN, gc = CUDART_NAN;
    if (ok && open_axis < 3 && (ref & ~REF_INDEX_MASK) == REF_TRIANGLE) {
      const double* v = tri_raw + (size_t)idx * 9;
      const double r[3] = {v[3] - v[0], v[4] - v[1], v[5] - v[2]}, s[3] = {v[6] - v[0], v[7] - v[1], v[8] - v[2]};
      const double nrm[3] = {r[1] * s[2] - r[2] * s[1], r[2] * s[0] - r[0] * s[2], r[0] * s[1] - r[1] * s[0]};
      const int a = open_axis, b = (a + 1) % 3, c = (a + 2) % 3;
      if (nrm[a] != 0.0) gb = -nrm[b] / nrm[a], gc = -nrm[c] / nrm[a];
    }
    grad[(size_t)i * 2 + 0] = gb, grad[(size_t)i * 2 + 1] = gc;
  }
}

// ---- 2. Morton keys -----------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t spread3(uint64_t x) {  // 21 bits → every third bit
  x &= 0x1fffffULL;
  x = (x | x << 32) & 0x1f00000000ffffULL;
  x = (x | x << 16) & 0x1f0000ff0000ffULL;
  x = (x | x << 8) & 0x100f00f00f00f00fULL;
  x = (x | x << 4) & 0x10c30c30c30c30c3ULL;
  x = (x | x << 2) & 0x1249249249249249ULL;
  return x;
}
// 63-bit code of the box centre inside `bounds`; an opened axis contributes 0; invalid prims sort last.
__global__ void k_morton(const double* __restrict__ lo, const double* __restrict__ hi, const uint8_t* __restrict__ valid,
                         uint32_t n, int open_axis, const double* __restrict__ bounds, uint64_t* __restrict__ keys,
                         uint32_t* __restrict__ vals) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  vals[i] = i;
  if (!valid[i]) {
    keys[i] = ~0ULL;
    return;
  }
  uint64_t q[3];
  for (int k = 0; k < 3; ++k) {
    if (k == open_axis) {
      q[k] = 0;
      continue;
    }
    const double ext = bounds[3 + k] - bounds[k];
    const double c = 0.5 * lo[(size_t)i * 3 + k] + 0.5 * hi[(size_t)i * 3 + k];
    double u = ext > 0.0 ? (c - bounds[k]) / ext : 0.0;
    u = fmin(fmax(u, 0.0), 1.0);
    q[k] = (uint64_t)fmin(u * 2097152.0, 2097151.0);
  }
  keys[i] = (spread3(q[0]) << 2) | (spread3(q[1]) << 1) | spread3(q[2]);
}

// ---- 3. LSD radix sort, 8 bits per pass, one warp per SORT_TILE keys ------------------------------------------
__global__ void k_sort_hist(const uint64_t* __restrict__ keys, uint32_t n, int shift, uint32_t ntiles,
                            uint32_t* __restrict__ hist /* [256][ntiles] */) {
  __shared__ uint32_t cnt[8][256];
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t tile = blockIdx.x * 8 + warp;
  for (int d = lane; d < 256; d += 32) cnt[warp][d] = 0;
  __syncwarp();
  if (tile < ntiles) {
    const uint32_t base = tile * SORT_TILE;
    for (int c = 0; c < SORT_TILE; c += 32) {
      const uint32_t i = base + c + lane;
      if (i < n) atomicAdd(&cnt[warp][(keys[i] >> shift) & 255], 1u);
    }
    __syncwarp();
    for (int d = lane; d < 256; d += 32) hist[(size_t)d * ntiles + tile] = cnt[warp][d];
  }
}
// exclusive scan of hist (digit-major, then tile) in place; single block
__global__ void k_sort_scan(uint32_t* __restrict__ hist, uint32_t total) {
  __shared__ uint32_t warp_sums[32];
  __shared__ uint32_t carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (uint32_t base = 0; base < total; base += blockDim.x) {
    const uint32_t i = base + threadIdx.x;
    const uint32_t v = i < total ? hist[i] : 0;
    uint32_t x = v;
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_up_sync(FULLMASK, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) warp_sums[warp] = x;
    __syncthreads();
    if (warp == 0) {
      uint32_t s = lane < (blockDim.x >> 5) ? warp_sums[lane] : 0;
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(FULLMASK, s, o);
        if (lane >= o) s += y;
      }
      warp_sums[lane] = s;  // inclusive
    }
    __syncthreads();
    const uint32_t before = (warp ? warp_sums[warp - 1] : 0) + carry;
    if (i < total) hist[i] = before + x - v;
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) carry = before + x;
    __syncthreads();
  }
}
__global__ void k_sort_scatter(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ vals, uint32_t n, int shift,
                               uint32_t ntiles, const uint32_t* __restrict__ hist, uint64_t* __restrict__ keys_out,
                               uint32_t* __restrict__ vals_out) {
  __shared__ uint32_t off[8][256];
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t tile = blockIdx.x * 8 + warp;
  if (tile >= ntiles) return;
  for (int d = lane; d < 256; d += 32) off[warp][d] = hist[(size_t)d * ntiles + tile];
  __syncwarp();
  const uint32_t base = tile * SORT_TILE;
  for (int c = 0; c < SORT_TILE; c += 32) {
    const uint32_t i = base + c + lane;
    const bool act = i < n;
    const uint32_t active = __ballot_sync(FULLMASK, act);
    if (act) {
      const uint64_t k = keys[i];
      const uint32_t d = (uint32_t)(k >> shift) & 255u;
      const uint32_t peers = __match_any_sync(active, d);
      const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
      const uint32_t dst = off[warp][d] + rank;  // all peers read before the leader bumps it
      __syncwarp(active);
      if (rank == 0) off[warp][d] += __popc(peers);
      keys_out[dst] = k;
      vals_out[dst] = vals[i];
    }
    __syncwarp();
  }
}

// ---- 4. Karras 2012 hierarchy over the m sorted valid keys ---------------------------------------------------
// Node numbering: internal nodes 0..m-2 (root = 0), leaf j = node (m-1) + j.  m == 1: the single leaf is root.
__device__ __forceinline__ int delta(const uint64_t* __restrict__ keys, int m, int i, int j) {
  if (j < 0 || j >= m) return -1;
  const uint64_t a = keys[i], b = keys[j];
  if (a == b) return 64 + __clz(i ^ j);  // duplicate codes: fall back to the index
  return __clzll((long long)(a ^ b));
}
__global__ void k_hierarchy(const uint64_t* __restrict__ keys, int m, uint32_t* __restrict__ left,
                            uint32_t* __restrict__ right, uint32_t* __restrict__ parent) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m - 1) return;
  const int d = (delta(keys, m, i, i + 1) - delta(keys, m, i, i - 1)) >= 0 ? 1 : -1;
  const int dmin = delta(keys, m, i, i - d);
  int lmax = 2;
  while (delta(keys, m, i, i + lmax * d) > dmin) lmax <<= 1;
  int l = 0;
  for (int t = lmax >> 1; t >= 1; t >>= 1)
    if (delta(keys, m, i, i + (l + t) * d) > dmin) l += t;
  const int j = i + l * d;
  const int dnode = delta(keys, m, i, j);
  int s = 0, t = l;
  do {
    t = (t + 1) >> 1;
    if (delta(keys, m, i, i + (s + t) * d) > dnode) s += t;
  } while (t > 1);
  const int gamma = i + s * d + min(d, 0);
  const uint32_t lc = (min(i, j) == gamma) ? (uint32_t)(m - 1 + gamma) : (uint32_t)gamma;
  const uint32_t rc = (max(i, j) == gamma + 1) ? (uint32_t)(m - 1 + gamma + 1) : (uint32_t)(gamma + 1);
  left[i] = lc;
  right[i] = rc;
  parent[lc] = (uint32_t)i;
  parent[rc] = (uint32_t)i;
  if (i == 0) parent[0] = BVH_END;
}

// ---- 5. leaves + bottom-up refit (second arrival at a node merges its children) --------------------------------
// pad = 2^-20 x scene scale (largest |coordinate| of the finite box corners; see rtx_lbvh.cuh)
__device__ __forceinline__ double box_pad(const double* __restrict__ bounds) {
  double scale = 1e-300;
  for (int k = 0; k < 6; ++k) scale = fmax(scale, fabs(bounds[k]));
  return scale * 9.5367431640625e-07;
}
__global__ void k_refit(const uint32_t* __restrict__ sorted_vals, const uint32_t* __restrict__ prims,
                        const double* __restrict__ lo, const double* __restrict__ hi, int m,
                        const double* __restrict__ bounds, const uint32_t* __restrict__ left,
                        const uint32_t* __restrict__ right, const uint32_t* __restrict__ parent,
                        uint32_t* __restrict__ flags, BvhNode* __restrict__ nodes, const double* __restrict__ grad) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= m) return;
  const double pad = box_pad(bounds);
  const uint32_t src = sorted_vals[j];
  const uint32_t self = (uint32_t)(m - 1 + j);
  BvhNode nd;
  for (int k = 0; k < 3; ++k) {
    nd.lo[k] = lo[(size_t)src * 3 + k] - pad;  // -inf - pad = -inf for opened axes
    nd.hi[k] = hi[(size_t)src * 3 + k] + pad;
  }
  for (int k = 0; k < 2; ++k) {
    const double g = grad ? grad[(size_t)src * 2 + k] : 0.0;
    const bool fin = g - g == 0.0;
    nd.glo[k] = fin ? g : -CUDART_INF;
    nd.ghi[k] = fin ? g : CUDART_INF;
  }
  nd.left = BVH_END;
  nd.escape = BVH_END;
  nd.prim = prims[src];
  nd.pad = 0;
  nodes[self] = nd;
  if (m == 1) return;
  __threadfence();
  uint32_t cur = parent[self];
  while (cur != BVH_END) {
    if (atomicAdd(&flags[cur], 1u) == 0u) return;  // first arrival: the sibling subtree is not ready
    __threadfence();
    const volatile BvhNode* a = nodes + left[cur];
    const volatile BvhNode* b = nodes + right[cur];
    BvhNode p;
    for (int k = 0; k < 3; ++k) {
      p.lo[k] = fmin(a->lo[k], b->lo[k]);
      p.hi[k] = fmax(a->hi[k], b->hi[k]);
    }
    for (int k = 0; k < 2; ++k) {
      p.glo[k] = fmin(a->glo[k], b->glo[k]);
      p.ghi[k] = fmax(a->ghi[k], b->ghi[k]);
    }
    p.left = left[cur];
    p.escape = BVH_END;
    p.prim = BVH_NO_PRIM;
    p.pad = 0;
    nodes[cur] = p;
    __threadfence();
    cur = parent[cur];
  }
}

// ---- 7. traversal nodes: both children's boxes (f32, rounded outward) + child refs ----------------------------
__device__ __forceinline__ uint32_t child_ref(const BvhNode* __restrict__ nodes, uint32_t c) {
  const uint32_t prim = nodes[c].prim;
  if (prim == BVH_NO_PRIM) return c;  // internal node index (< m-1 < 2^30)
  return WIDE_LEAF | (((prim & ~REF_INDEX_MASK) == REF_TRIANGLE) ? WIDE_TRI : 0u) | (prim & REF_INDEX_MASK);
}
// gradient bounds of a child, f32, rounded outward and widened by 1e-6 relative (they are compared with quantities
// derived from the runtime's own normal, which agrees with this one to ~1e-16)
__device__ __forceinline__ float4 child_gradient(const BvhNode* __restrict__ nodes, uint32_t c) {
  float v[4];
  for (int k = 0; k < 2; ++k) {
    const double lo = nodes[c].glo[k], hi = nodes[c].ghi[k];
    v[2 * k + 0] = __double2float_rd(lo - fabs(lo) * 1e-6 - 1e-30);
    v[2 * k + 1] = __double2float_ru(hi + fabs(hi) * 1e-6 + 1e-30);
  }
  return make_float4(v[0], v[1], v[2], v[3]);
}
__global__ void k_emit_wide(int m, const BvhNode* __restrict__ nodes, const uint32_t* __restrict__ left,
                            const uint32_t* __restrict__ right, WideNode* __restrict__ out, float4* __restrict__ aux) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (m == 1) {  // a single leaf: root = one wide node with an empty second slot
    if (i != 0) return;
    WideNode w;
    for (int k = 0; k < 3; ++k) {
      w.lo0[k] = __double2float_rd(nodes[0].lo[k]);
      w.hi0[k] = __double2float_ru(nodes[0].hi[k]);
      w.lo1[k] = w.lo0[k];  // a one-primitive tree: the second slot repeats the first (testing a primitive twice
      w.hi1[k] = w.hi0[k];  // changes nothing), so the walk never has to recognise an empty slot
    }
    w.c0 = child_ref(nodes, 0);
    w.c1 = w.c0;
    w.pad[0] = w.pad[1] = 0;
    out[0] = w;
    if (aux) aux[0] = aux[1] = child_gradient(nodes, 0);
    return;
  }
  if (i >= m - 1) return;
  const uint32_t l = left[i], r = right[i];
  WideNode w;
  for (int k = 0; k < 3; ++k) {
    w.lo0[k] = __double2float_rd(nodes[l].lo[k]);
    w.hi0[k] = __double2float_ru(nodes[l].hi[k]);
    w.lo1[k] = __double2float_rd(nodes[r].lo[k]);
    w.hi1[k] = __double2float_ru(nodes[r].hi[k]);
  }
  w.c0 = child_ref(nodes, l);
  w.c1 = child_ref(nodes, r);
  w.pad[0] = w.pad[1] = 0;
  out[i] = w;
  if (aux) aux[2 * i] = child_gradient(nodes, l), aux[2 * i + 1] = child_gradient(nodes, r);
}

__global__ void k_count_valid(const uint64_t* __restrict__ sorted_keys, uint32_t n, uint32_t* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  // keys are sorted: the first invalid key marks m
  if (sorted_keys[i] == ~0ULL && (i == 0 || sorted_keys[i - 1] != ~0ULL)) *out = i;
}

// refit precondition: the set of primitives inside the tree is unchanged
__global__ void k_valid_diff(const uint8_t* __restrict__ a, const uint8_t* __restrict__ b, uint32_t n, uint32_t* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && a[i] != b[i]) atomicAdd(out, 1u);
}

// ---- host orchestration ---------------------------------------------------------------------------------------
template <typename T>
struct Tmp {
  T* p = nullptr;
  ~Tmp() {
    if (p) cudaFree(p);
  }
  cudaError_t alloc(size_t n) { return cudaMalloc((void**)&p, (n ? n : 1) * sizeof(T)); }
  T* release() {
    T* q = p;
    p = nullptr;
    return q;
  }
};

#define LBVH_TRY(expr)                                          \
  do {                                                          \
    cudaError_t _e = (expr);                                    \
    if (_e != cudaSuccess) return cuda_fail(s, _e, #expr);      \
  } while (0)

// Builds one tree over `prims` (host list of REF_*|index).  Synchronous on `stream`.
static void free_topo(rtx_scene::TreeTopo& t) {
  cudaFree(t.d_prims), cudaFree(t.d_vals), cudaFree(t.d_left), cudaFree(t.d_right), cudaFree(t.d_parent), cudaFree(t.d_valid);
  t = rtx_scene::TreeTopo{};
}

static rtx_status build_tree(rtx_scene* s, const std::vector<uint32_t>& prims, int open_axis, const double* d_tri_raw,
                             cudaStream_t stream, WideNode** out_nodes, uint32_t* out_root, uint32_t* out_count,
                             uint32_t* launches, rtx_scene::TreeTopo* topo, float4** out_aux) {
  *out_nodes = nullptr;
  *out_aux = nullptr;
  *out_root = BVH_END;
  *out_count = 0;
  free_topo(*topo);
  topo->open_axis = open_axis;
  const uint32_t n = (uint32_t)prims.size();
  if (n == 0) return RTX_OK;
  const unsigned B = 256, G = (n + B - 1) / B;
  Tmp<uint32_t> d_prims, d_vals[2], d_hist, d_left, d_right, d_parent, d_flags, d_m;
  Tmp<double> d_lo, d_hi, d_bounds, d_grad;
  Tmp<uint8_t> d_valid;
  Tmp<uint64_t> d_keys[2];
  const bool prism = open_axis < 3;
  if (prism) LBVH_TRY(d_grad.alloc((size_t)n * 2));
  LBVH_TRY(d_prims.alloc(n));
  LBVH_TRY(d_lo.alloc((size_t)n * 3));
  LBVH_TRY(d_hi.alloc((size_t)n * 3));
  LBVH_TRY(d_valid.alloc(n));
  LBVH_TRY(d_bounds.alloc(6));
  LBVH_TRY(d_keys[0].alloc(n));
  LBVH_TRY(d_keys[1].alloc(n));
  LBVH_TRY(d_vals[0].alloc(n));
  LBVH_TRY(d_vals[1].alloc(n));
  LBVH_TRY(d_m.alloc(1));
  LBVH_TRY(cudaMemcpyAsync(d_prims.p, prims.data(), (size_t)n * 4, cudaMemcpyHostToDevice, stream));
  const double inf = std::numeric_limits<double>::infinity();
  const double b0[6] = {inf, inf, inf, -inf, -inf, -inf};
  LBVH_TRY(cudaMemcpyAsync(d_bounds.p, b0, sizeof(b0), cudaMemcpyHostToDevice, stream));
  const double* d_sph = reinterpret_cast<const double*>(s->d_blob + s->layout.off_sph);
  k_prim_boxes<<<G, B, 0, stream>>>(d_sph, d_tri_raw, d_prims.p, n, open_axis, d_lo.p, d_hi.p, d_valid.p, d_bounds.p,
                                    prism ? d_grad.p : nullptr);
  k_morton<<<G, B, 0, stream>>>(d_lo.p, d_hi.p, d_valid.p, n, open_axis, d_bounds.p, d_keys[0].p, d_vals[0].p);
  *launches += 2;
  // radix sort: 8 passes x 8 bits
  const uint32_t ntiles = (n + SORT_TILE - 1) / SORT_TILE;
  LBVH_TRY(d_hist.alloc((size_t)256 * ntiles));
  int cur = 0;
  for (int pass = 0; pass < 8; ++pass) {
    const unsigned gb = (ntiles + 7) / 8;
    k_sort_hist<<<gb, 256, 0, stream>>>(d_keys[cur].p, n, pass * 8, ntiles, d_hist.p);
    k_sort_scan<<<1, 1024, 0, stream>>>(d_hist.p, 256u * ntiles);
    k_sort_scatter<<<gb, 256, 0, stream>>>(d_keys[cur].p, d_vals[cur].p, n, pass * 8, ntiles, d_hist.p, d_keys[cur ^ 1].p,
                                           d_vals[cur ^ 1].p);
    cur ^= 1;
    *launches += 3;
  }
  // m = number of valid primitives (invalid keys = ~0 sorted last)
  LBVH_TRY(cudaMemcpyAsync(d_m.p, &n, 4, cudaMemcpyHostToDevice, stream));
  k_count_valid<<<G, B, 0, stream>>>(d_keys[cur].p, n, d_m.p);
  *launches += 1;
  uint32_t m = 0;
  LBVH_TRY(cudaMemcpyAsync(&m, d_m.p, 4, cudaMemcpyDeviceToHost, stream));
  LBVH_TRY(cudaStreamSynchronize(stream));
  if (m == 0) return RTX_OK;
  const uint32_t total = 2 * m - 1;
  BvhNode* nodes = nullptr;  // f64 build nodes (freed below); the traversal gets the f32 two-box nodes
  WideNode* wide = nullptr;
  const uint32_t n_wide = m > 1 ? m - 1 : 1;
  LBVH_TRY(cudaMalloc((void**)&nodes, (size_t)total * sizeof(BvhNode)));
  float4* aux = nullptr;
  {
    cudaError_t ew = cudaMalloc((void**)&wide, (size_t)n_wide * sizeof(WideNode));
    if (ew == cudaSuccess && prism) ew = cudaMalloc((void**)&aux, (size_t)n_wide * 2 * sizeof(float4));
    if (ew != cudaSuccess) {
      cudaFree(nodes);
      cudaFree(wide);
      return cuda_fail(s, ew, "lbvh wide nodes");
    }
  }
  rtx_status st = RTX_OK;
  do {
    cudaError_t e;
    if ((e = d_left.alloc(m)) != cudaSuccess || (e = d_right.alloc(m)) != cudaSuccess ||
        (e = d_parent.alloc(total)) != cudaSuccess || (e = d_flags.alloc(m)) != cudaSuccess ||
        (e = cudaMemsetAsync(d_flags.p, 0, (size_t)m * 4, stream)) != cudaSuccess ||
        (e = cudaMemsetAsync(d_parent.p, 0xFF, (size_t)total * 4, stream)) != cudaSuccess) {
      st = cuda_fail(s, e, "lbvh temporaries");
      break;
    }
    const unsigned Gm = (m + B - 1) / B;
    if (m > 1) k_hierarchy<<<Gm, B, 0, stream>>>(d_keys[cur].p, (int)m, d_left.p, d_right.p, d_parent.p);
    k_refit<<<Gm, B, 0, stream>>>(d_vals[cur].p, d_prims.p, d_lo.p, d_hi.p, (int)m, d_bounds.p, d_left.p, d_right.p,
                                  d_parent.p, d_flags.p, nodes, prism ? d_grad.p : nullptr);
    k_emit_wide<<<Gm, B, 0, stream>>>((int)m, nodes, d_left.p, d_right.p, wide, aux);
    *launches += 3;
    if ((e = cudaGetLastError()) != cudaSuccess || (e = cudaStreamSynchronize(stream)) != cudaSuccess) {
      st = cuda_fail(s, e, "lbvh build kernels");
      break;
    }
  } while (0);
  cudaFree(nodes);
  if (st != RTX_OK) {
    cudaFree(wide);
    cudaFree(aux);
    return st;
  }
  *out_aux = aux;
  *out_nodes = wide;
  *out_root = 0u;  // the root is wide node 0
  *out_count = n_wide;
  // keep what a refit needs: primitive list, validity, sorted order, hierarchy
  topo->n = n, topo->m = m;
  topo->d_prims = d_prims.release();
  topo->d_vals = d_vals[cur].release();
  topo->d_left = d_left.release();
  topo->d_right = d_right.release();
  topo->d_parent = d_parent.release();
  topo->d_valid = d_valid.release();
  return RTX_OK;
}

// Same hierarchy, new boxes: recompute the primitive boxes from the (updated) scene, check that the same
// primitives are inside the tree, then redo the bottom-up pass and the traversal nodes in place.
// *needs_rebuild is set when membership changed.  Synchronous on `stream`.
static rtx_status refit_tree(rtx_scene* s, const rtx_scene::TreeTopo& t, WideNode* wide, float4* aux, cudaStream_t stream,
                             bool* needs_rebuild, uint32_t* launches) {
  if (t.n == 0 || t.m == 0) return RTX_OK;
  const unsigned B = 256, G = (t.n + B - 1) / B, Gm = (t.m + B - 1) / B;
  Tmp<double> d_lo, d_hi, d_bounds, d_grad;
  Tmp<uint8_t> d_valid;
  Tmp<uint32_t> d_flags, d_diff;
  Tmp<BvhNode> d_nodes;
  const bool prism = t.open_axis < 3 && aux;
  if (prism) LBVH_TRY(d_grad.alloc((size_t)t.n * 2));
  LBVH_TRY(d_lo.alloc((size_t)t.n * 3));
  LBVH_TRY(d_hi.alloc((size_t)t.n * 3));
  LBVH_TRY(d_valid.alloc(t.n));
  LBVH_TRY(d_bounds.alloc(6));
  LBVH_TRY(d_flags.alloc(t.m));
  LBVH_TRY(d_diff.alloc(1));
  LBVH_TRY(d_nodes.alloc((size_t)2 * t.m - 1));
  const double inf = std::numeric_limits<double>::infinity();
  const double b0[6] = {inf, inf, inf, -inf, -inf, -inf};
  LBVH_TRY(cudaMemcpyAsync(d_bounds.p, b0, sizeof(b0), cudaMemcpyHostToDevice, stream));
  LBVH_TRY(cudaMemsetAsync(d_diff.p, 0, 4, stream));
  LBVH_TRY(cudaMemsetAsync(d_flags.p, 0, (size_t)t.m * 4, stream));
  const double* d_sph = reinterpret_cast<const double*>(s->d_blob + s->layout.off_sph);
  k_prim_boxes<<<G, B, 0, stream>>>(d_sph, s->d_tri_raw, t.d_prims, t.n, t.open_axis, d_lo.p, d_hi.p, d_valid.p, d_bounds.p,
                                    prism ? d_grad.p : nullptr);
  k_valid_diff<<<G, B, 0, stream>>>(d_valid.p, t.d_valid, t.n, d_diff.p);
  uint32_t diff = 0;
  LBVH_TRY(cudaMemcpyAsync(&diff, d_diff.p, 4, cudaMemcpyDeviceToHost, stream));
  LBVH_TRY(cudaStreamSynchronize(stream));
  *launches += 2;
  if (diff != 0) {
    *needs_rebuild = true;
    return RTX_OK;
  }
  k_refit<<<Gm, B, 0, stream>>>(t.d_vals, t.d_prims, d_lo.p, d_hi.p, (int)t.m, d_bounds.p, t.d_left, t.d_right, t.d_parent,
                                d_flags.p, d_nodes.p, prism ? d_grad.p : nullptr);
  k_emit_wide<<<Gm, B, 0, stream>>>((int)t.m, d_nodes.p, t.d_left, t.d_right, wide, prism ? aux : nullptr);
  *launches += 2;
  LBVH_TRY(cudaGetLastError());
  LBVH_TRY(cudaStreamSynchronize(stream));
  return RTX_OK;
}

void free_lbvh(rtx_scene* s) {
  for (int k = 0; k < 4; ++k) {
    if (s->d_bvh[k]) cudaFree(s->d_bvh[k]);
    if (s->d_bvh_aux[k]) cudaFree(s->d_bvh_aux[k]);
    s->d_bvh[k] = nullptr;
    s->d_bvh_aux[k] = nullptr;
    s->bvh_root[k] = BVH_END;
    s->bvh_count[k] = 0;
    free_topo(s->topo[k]);
  }
  s->bvh_valid = false;
  s->bvh_refit = false;
}

void fill_bvh_view(const rtx_scene* s, bool use_bvh, BvhView* bv) {
  for (int k = 0; k < 4; ++k) {
    bv->nodes[k] = use_bvh ? s->d_bvh[k] : nullptr;
    bv->aux[k] = use_bvh ? s->d_bvh_aux[k] : nullptr;
    bv->root[k] = use_bvh ? s->bvh_root[k] : BVH_END;
  }
}

// Builds the 3-D tree (spheres + triangles) and one prism tree per dropped axis (triangles).  The scene blob
// must be uploaded.  Rebuilt whenever the object list changed.
rtx_status ensure_lbvh(rtx_scene* s, cudaStream_t stream) {
  if (s->bvh_valid && s->bvh_refit) {  // rtx_scene_update_*: primitives moved, hierarchy kept
    bool rebuild = false;
    uint32_t launches = 0;
    for (int k = 0; k < 4 && !rebuild; ++k) {
      if (s->bvh_root[k] == BVH_END) continue;
      rtx_status st = refit_tree(s, s->topo[k], s->d_bvh[k], s->d_bvh_aux[k], stream, &rebuild, &launches);
      if (st != RTX_OK) return st;
    }
    s->bvh_refit = false;
    if (!rebuild) {
      ++s->bvh_refits;
      return RTX_OK;
    }
    s->bvh_valid = false;
  }
  if (s->bvh_valid) return RTX_OK;
  free_lbvh(s);
  const uint32_t n_sph = (uint32_t)s->sph_id.size(), n_tri = (uint32_t)s->tri_id.size();
  std::vector<uint32_t> all, prism[3];
  all.reserve((size_t)n_sph + n_tri);
  for (uint32_t i = 0; i < n_sph; ++i) all.push_back(REF_SPHERE | i);
  for (uint32_t i = 0; i < n_tri; ++i) {
    uint64_t code;
    std::memcpy(&code, &s->tri[(size_t)i * TRI_STRIDE + 13], 8);
    if (code & (TRI_DEGENERATE | TRI_PANIC)) continue;  // contains() is never true: cannot be hit
    all.push_back(REF_TRIANGLE | i);
    const int i1 = (int)((code >> TRI_I1_SHIFT) & 3), i2 = (int)((code >> TRI_I2_SHIFT) & 3);
    prism[3 - i1 - i2].push_back(REF_TRIANGLE | i);  // the coordinate contains() never looks at
  }
  if (n_tri) {  // persistent: refits recompute boxes from it
    if (s->d_tri_raw_cap < s->tri_raw.size() * 8 || !s->d_tri_raw) {
      cudaFree(s->d_tri_raw);
      s->d_tri_raw = nullptr, s->d_tri_raw_cap = 0;
      LBVH_TRY(cudaMalloc((void**)&s->d_tri_raw, s->tri_raw.size() * 8));
      s->d_tri_raw_cap = s->tri_raw.size() * 8;
    }
    LBVH_TRY(cudaMemcpyAsync(s->d_tri_raw, s->tri_raw.data(), s->tri_raw.size() * 8, cudaMemcpyHostToDevice, stream));
  }
  uint32_t launches = 0;
  rtx_status st = build_tree(s, all, 3, s->d_tri_raw, stream, &s->d_bvh[0], &s->bvh_root[0], &s->bvh_count[0], &launches,
                             &s->topo[0], &s->d_bvh_aux[0]);
  for (int a = 0; a < 3 && st == RTX_OK; ++a)
    st = build_tree(s, prism[a], a, s->d_tri_raw, stream, &s->d_bvh[1 + a], &s->bvh_root[1 + a], &s->bvh_count[1 + a],
                    &launches, &s->topo[1 + a], &s->d_bvh_aux[1 + a]);
  if (st != RTX_OK) {
    free_lbvh(s);
    return st;
  }
  s->bvh_build_launches = launches;
  s->bvh_valid = true;
  ++s->bvh_builds;
  return RTX_OK;
}

}  // namespace rtx
