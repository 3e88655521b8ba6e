// rtx_debug.cu — per-function device evaluation (parity tests) and the FP64 issue-rate probe
// (roofline denominators).  Same device functions the render kernel inlines; nothing here runs on the CPU.
#include <math_constants.h>

#include "rtx_internal.h"
#include "rtx_lbvh.cuh"
#include "rtx_trace.cuh"

namespace rtx {

__global__ void closest_kernel(const unsigned char* blob, SceneLayout L, BvhView bv, int use_bvh, const double* __restrict__ rays, uint64_t n,
                               double* __restrict__ out_t, uint32_t* __restrict__ out_id) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const SceneView sc = make_view(blob, L);
  const V3 pos = mk3(rays[i * 6 + 0], rays[i * 6 + 1], rays[i * 6 + 2]);
  const V3 dir = mk3(rays[i * 6 + 3], rays[i * 6 + 4], rays[i * 6 + 5]);
  uint32_t panics = 0;
  TraversalCounters tc;
  tc.nodes = 0, tc.prims = 0;
  const Hit h = use_bvh ? closest_lbvh(sc, bv, pos, dir, &panics, tc) : closest_brute(sc, pos, dir, &panics);
  out_id[i] = h.id;
  out_t[i] = h.id == MISS_ID ? 0.0 : h.t;
}

__global__ void normal_kernel(const unsigned char* blob, SceneLayout L, const uint32_t* __restrict__ refs,
                              const double* __restrict__ pos, uint64_t n, double* __restrict__ out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const SceneView sc = make_view(blob, L);
  const V3 nn = normal_at(sc, refs[i], mk3(pos[i * 3], pos[i * 3 + 1], pos[i * 3 + 2]));
  out[i * 3 + 0] = nn.x, out[i * 3 + 1] = nn.y, out[i * 3 + 2] = nn.z;
}

__global__ void bounce_kernel(uint64_t* __restrict__ states, const double* __restrict__ in, uint64_t n,
                              double* __restrict__ out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  WyRand g;
  g.s = states[i];
  const V3 d = random_bounce_dir(g, mk3(in[i * 7], in[i * 7 + 1], in[i * 7 + 2]),
                                 mk3(in[i * 7 + 3], in[i * 7 + 4], in[i * 7 + 5]), in[i * 7 + 6]);
  out[i * 3 + 0] = d.x, out[i * 3 + 1] = d.y, out[i * 3 + 2] = d.z;
  states[i] = g.s;
}

__global__ void wyrand_kernel(uint64_t* state, uint64_t n, double* __restrict__ out) {
  if (blockIdx.x != 0 || threadIdx.x != 0) return;
  WyRand g;
  g.s = *state;
  for (uint64_t i = 0; i < n; ++i) out[i] = wy_f64(g);
  *state = g.s;
}

// vdiv3 (shared-reciprocal division, rtx_math.cuh) next to three plain __ddiv_rn: must be bit-equal.
__global__ void vdiv3_kernel(const double* __restrict__ a, const double* __restrict__ s, uint64_t n,
                             double* __restrict__ out, double* __restrict__ ref) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const V3 v = mk3(a[i * 3], a[i * 3 + 1], a[i * 3 + 2]);
  const V3 q = vdiv3(v, s[i]);
  out[i * 3 + 0] = q.x, out[i * 3 + 1] = q.y, out[i * 3 + 2] = q.z;
  ref[i * 3 + 0] = __ddiv_rn(v.x, s[i]), ref[i * 3 + 1] = __ddiv_rn(v.y, s[i]), ref[i * 3 + 2] = __ddiv_rn(v.z, s[i]);
}

// Register-resident issue-rate probes.  8 independent chains per thread.
template <bool FUSED>
__global__ void __launch_bounds__(256) fp64_probe_kernel(double* out, int iters, double a, double b) {
  double x0 = threadIdx.x * 1e-3, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6,
         x7 = x0 + 7;
  for (int i = 0; i < iters; ++i) {
    if (FUSED) {
      x0 = __fma_rn(x0, a, b), x1 = __fma_rn(x1, a, b), x2 = __fma_rn(x2, a, b), x3 = __fma_rn(x3, a, b);
      x4 = __fma_rn(x4, a, b), x5 = __fma_rn(x5, a, b), x6 = __fma_rn(x6, a, b), x7 = __fma_rn(x7, a, b);
    } else {
      x0 = __dadd_rn(__dmul_rn(x0, a), b), x1 = __dadd_rn(__dmul_rn(x1, a), b);
      x2 = __dadd_rn(__dmul_rn(x2, a), b), x3 = __dadd_rn(__dmul_rn(x3, a), b);
      x4 = __dadd_rn(__dmul_rn(x4, a), b), x5 = __dadd_rn(__dmul_rn(x5, a), b);
      x6 = __dadd_rn(__dmul_rn(x6, a), b), x7 = __dadd_rn(__dmul_rn(x7, a), b);
    }
  }
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}

// Does this translation unit contract a*b+c?  (It must not: parity contract.)  Returns fl(fl(a*b)+c).
__global__ void contraction_probe_kernel(double a, double b, double c, double* out) { *out = a * b + c; }

template <typename T>
struct DevBuf {
  T* p = nullptr;
  ~DevBuf() {
    if (p) cudaFree(p);
  }
  cudaError_t alloc(size_t n) { return cudaMalloc((void**)&p, (n ? n : 1) * sizeof(T)); }
};

}  // namespace rtx

using namespace rtx;

extern "C" {

rtx_status rtx_debug_closest_object(rtx_scene* s, uint32_t flags, const double* rays, uint64_t n, double* out_t,
                                    uint32_t* out_id) {
  if (!s || !rays || !out_t || !out_id) return fail(s, RTX_ERR_INVALID_ARGUMENT, "null argument");
  std::lock_guard<std::mutex> lk(s->mu);
  RTX_CUDA_TRY(s, cudaSetDevice(s->device));
  rtx_status st = ensure_uploaded(s, s->stream);
  if (st != RTX_OK) return st;
  const bool use_bvh = choose_accel(s, flags) == RTX_ACCEL_LBVH;
  BvhView bv{};
  if (use_bvh) {
    st = ensure_lbvh(s, s->stream);
    if (st != RTX_OK) return st;
  }
  fill_bvh_view(s, use_bvh, &bv);
  if (n == 0) return RTX_OK;
  DevBuf<double> d_rays, d_t;
  DevBuf<uint32_t> d_id;
  RTX_CUDA_TRY(s, d_rays.alloc(n * 6));
  RTX_CUDA_TRY(s, d_t.alloc(n));
  RTX_CUDA_TRY(s, d_id.alloc(n));
  RTX_CUDA_TRY(s, cudaMemcpyAsync(d_rays.p, rays, n * 48, cudaMemcpyHostToDevice, s->stream));
  closest_kernel<<<(unsigned)((n + 127) / 128), 128, 0, s->stream>>>(s->d_blob, s->layout, bv, use_bvh ? 1 : 0, d_rays.p, n, d_t.p, d_id.p);
  RTX_CUDA_TRY(s, cudaGetLastError());
  RTX_CUDA_TRY(s, cudaMemcpyAsync(out_t, d_t.p, n * 8, cudaMemcpyDeviceToHost, s->stream));
  RTX_CUDA_TRY(s, cudaMemcpyAsync(out_id, d_id.p, n * 4, cudaMemcpyDeviceToHost, s->stream));
  RTX_CUDA_TRY(s, cudaStreamSynchronize(s->stream));
  return RTX_OK;
}

rtx_status rtx_debug_normal_at(rtx_scene* s, const uint32_t* ids, const double* positions, uint64_t n, double* out) {
  if (!s || !ids || !positions || !out) return fail(s, RTX_ERR_INVALID_ARGUMENT, "null argument");
  std::lock_guard<std::mutex> lk(s->mu);
  RTX_CUDA_TRY(s, cudaSetDevice(s->device));
  rtx_status st = ensure_uploaded(s, s->stream);
  if (st != RTX_OK) return st;
  if (n == 0) return RTX_OK;
  // object index → (type, index) reference
  std::vector<uint32_t> ref_of(s->n_obj, 0);
  for (size_t i = 0; i < s->sph_id.size(); ++i) ref_of[s->sph_id[i]] = REF_SPHERE | (uint32_t)i;
  for (size_t i = 0; i < s->pl_id.size(); ++i) ref_of[s->pl_id[i]] = REF_PLANE | (uint32_t)i;
  for (size_t i = 0; i < s->tri_id.size(); ++i) ref_of[s->tri_id[i]] = REF_TRIANGLE | (uint32_t)i;
  std::vector<uint32_t> refs(n);
  for (uint64_t i = 0; i < n; ++i) {
    if (ids[i] >= s->n_obj) return fail(s, RTX_ERR_INVALID_ARGUMENT, "object index out of range");
    refs[i] = ref_of[ids[i]];
  }
  DevBuf<uint32_t> d_refs;
  DevBuf<double> d_pos, d_out;
  RTX_CUDA_TRY(s, d_refs.alloc(n));
  RTX_CUDA_TRY(s, d_pos.alloc(n * 3));
  RTX_CUDA_TRY(s, d_out.alloc(n * 3));
  RTX_CUDA_TRY(s, cudaMemcpyAsync(d_refs.p, refs.data(), n * 4, cudaMemcpyHostToDevice, s->stream));
  RTX_CUDA_TRY(s, cudaMemcpyAsync(d_pos.p, positions, n * 24, cudaMemcpyHostToDevice, s->stream));
  normal_kernel<<<(unsigned)((n + 127) / 128), 128, 0, s->stream>>>(s->d_blob, s->layout, d_refs.p, d_pos.p, n, d_out.p);
  RTX_CUDA_TRY(s, cudaGetLastError());
  RTX_CUDA_TRY(s, cudaMemcpyAsync(out, d_out.p, n * 24, cudaMemcpyDeviceToHost, s->stream));
  RTX_CUDA_TRY(s, cudaStreamSynchronize(s->stream));
  return RTX_OK;
}

static rtx_status pick_device(int32_t device) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
    cudaGetLastError();
    return fail(nullptr, RTX_ERR_CUDA, "no CUDA device available: the rtx CUDA path has no CPU fallback");
  }
  if (device < 0 || device >= n) return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "device ordinal out of range");
  RTX_CUDA_TRY(nullptr, cudaSetDevice(device));
  return RTX_OK;
}

rtx_status rtx_debug_random_bounce_dir(int32_t device, uint64_t* states, const double* in, uint64_t n, double* out) {
  if (!states || !in || !out) return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "null argument");
  rtx_status st = pick_device(device);
  if (st != RTX_OK) return st;
  if (n == 0) return RTX_OK;
  DevBuf<uint64_t> d_states;
  DevBuf<double> d_in, d_out;
  RTX_CUDA_TRY(nullptr, d_states.alloc(n));
  RTX_CUDA_TRY(nullptr, d_in.alloc(n * 7));
  RTX_CUDA_TRY(nullptr, d_out.alloc(n * 3));
  RTX_CUDA_TRY(nullptr, cudaMemcpy(d_states.p, states, n * 8, cudaMemcpyHostToDevice));
  RTX_CUDA_TRY(nullptr, cudaMemcpy(d_in.p, in, n * 56, cudaMemcpyHostToDevice));
  bounce_kernel<<<(unsigned)((n + 127) / 128), 128>>>(d_states.p, d_in.p, n, d_out.p);
  RTX_CUDA_TRY(nullptr, cudaGetLastError());
  RTX_CUDA_TRY(nullptr, cudaMemcpy(out, d_out.p, n * 24, cudaMemcpyDeviceToHost));
  RTX_CUDA_TRY(nullptr, cudaMemcpy(states, d_states.p, n * 8, cudaMemcpyDeviceToHost));
  return RTX_OK;
}

rtx_status rtx_debug_wyrand_f64(int32_t device, uint64_t* state, uint64_t n, double* out) {
  if (!state || !out) return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "null argument");
  rtx_status st = pick_device(device);
  if (st != RTX_OK) return st;
  DevBuf<uint64_t> d_state;
  DevBuf<double> d_out;
  RTX_CUDA_TRY(nullptr, d_state.alloc(1));
  RTX_CUDA_TRY(nullptr, d_out.alloc(n));
  RTX_CUDA_TRY(nullptr, cudaMemcpy(d_state.p, state, 8, cudaMemcpyHostToDevice));
  wyrand_kernel<<<1, 32>>>(d_state.p, n, d_out.p);
  RTX_CUDA_TRY(nullptr, cudaGetLastError());
  RTX_CUDA_TRY(nullptr, cudaMemcpy(out, d_out.p, n * 8, cudaMemcpyDeviceToHost));
  RTX_CUDA_TRY(nullptr, cudaMemcpy(state, d_state.p, 8, cudaMemcpyDeviceToHost));
  return RTX_OK;
}

rtx_status rtx_debug_vdiv3(int32_t device, const double* a, const double* s, uint64_t n, double* out, double* ref) {
  if (!a || !s || !out || !ref) return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "null argument");
  rtx_status st = pick_device(device);
  if (st != RTX_OK) return st;
  if (n == 0) return RTX_OK;
  DevBuf<double> d_a, d_s, d_out, d_ref;
  RTX_CUDA_TRY(nullptr, d_a.alloc(n * 3));
  RTX_CUDA_TRY(nullptr, d_s.alloc(n));
  RTX_CUDA_TRY(nullptr, d_out.alloc(n * 3));
  RTX_CUDA_TRY(nullptr, d_ref.alloc(n * 3));
  RTX_CUDA_TRY(nullptr, cudaMemcpy(d_a.p, a, n * 24, cudaMemcpyHostToDevice));
  RTX_CUDA_TRY(nullptr, cudaMemcpy(d_s.p, s, n * 8, cudaMemcpyHostToDevice));
  vdiv3_kernel<<<(unsigned)((n + 127) / 128), 128>>>(d_a.p, d_s.p, n, d_out.p, d_ref.p);
  RTX_CUDA_TRY(nullptr, cudaGetLastError());
  RTX_CUDA_TRY(nullptr, cudaMemcpy(out, d_out.p, n * 24, cudaMemcpyDeviceToHost));
  RTX_CUDA_TRY(nullptr, cudaMemcpy(ref, d_ref.p, n * 24, cudaMemcpyDeviceToHost));
  return RTX_OK;
}

rtx_status rtx_debug_contraction_probe(int32_t device, double a, double b, double c, double* out) {
  if (!out) return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "null argument");
  rtx_status st = pick_device(device);
  if (st != RTX_OK) return st;
  DevBuf<double> d;
  RTX_CUDA_TRY(nullptr, d.alloc(1));
  contraction_probe_kernel<<<1, 1>>>(a, b, c, d.p);
  RTX_CUDA_TRY(nullptr, cudaGetLastError());
  RTX_CUDA_TRY(nullptr, cudaMemcpy(out, d.p, 8, cudaMemcpyDeviceToHost));
  return RTX_OK;
}

rtx_status rtx_probe_fp64_peak(int32_t device, double* out_nonfused_tflops, double* out_fma_tflops) {
  rtx_status st = pick_device(device);
  if (st != RTX_OK) return st;
  cudaDeviceProp prop{};
  RTX_CUDA_TRY(nullptr, cudaGetDeviceProperties(&prop, device));
  const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 4096;
  DevBuf<double> d;
  RTX_CUDA_TRY(nullptr, d.alloc((size_t)blocks * threads));
  cudaEvent_t e0, e1;
  RTX_CUDA_TRY(nullptr, cudaEventCreate(&e0));
  RTX_CUDA_TRY(nullptr, cudaEventCreate(&e1));
  double best[2] = {0, 0};
  for (int fused = 0; fused < 2; ++fused) {
    for (int rep = 0; rep < 6; ++rep) {
      cudaEventRecord(e0);
      if (fused)
        fp64_probe_kernel<true><<<blocks, threads>>>(d.p, iters, 1.0000001, 1e-9);
      else
        fp64_probe_kernel<false><<<blocks, threads>>>(d.p, iters, 1.0000001, 1e-9);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms = 0;
      cudaEventElapsedTime(&ms, e0, e1);
      // 8 chains × 2 flops (mul + add, fused or not) per iteration per thread
      const double flops = (double)blocks * threads * (double)iters * 16.0;
      const double tf = flops / (ms * 1e-3) / 1e12;
      if (rep >= 1 && tf > best[fused]) best[fused] = tf;
    }
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  RTX_CUDA_TRY(nullptr, cudaGetLastError());
  if (out_nonfused_tflops) *out_nonfused_tflops = best[0];
  if (out_fma_tflops) *out_fma_tflops = best[1];
  return RTX_OK;
}

}  // extern "C"
