// rtx_rng.cuh — the reference's random source, restated for host + device.
//
// The reference draws every random number from fastrand::f64() (call sites math/vector.rs:31-33 and
// :37-38; fastrand 2.3.0, Cargo.lock:338-341 — third-party, not vendored): WyRand with the wyhash v4.2
// constants, then a 52-bit mantissa fill.  The generator state is a pure counter, so a stream can be
// started anywhere in O(1) — which is what lets every (pixel, sample) own a stream on the GPU.
//
// The reference never seeds the generator (thread-local, time-seeded), so the seeding convention is
// ours: stream key k = (y*width + x)*rays_per_pixel + sample, state0 = mix64(seed + GOLDEN*(k+1)).
#pragma once
#include "rtx_math.cuh"

namespace rtx {

struct WyRand {
  uint64_t s;
};

RTX_HD uint64_t mulhi64(uint64_t a, uint64_t b) {
#if defined(__CUDA_ARCH__)
  return __umul64hi(a, b);
#else
  return (uint64_t)(((unsigned __int128)a * (unsigned __int128)b) >> 64);
#endif
}

RTX_HD uint64_t wy_u64(WyRand& g) {
  g.s += 0x2d358dccaa6c78a5ULL;
  const uint64_t m = g.s ^ 0x8bb84b93962eacc9ULL;
  return (g.s * m) ^ mulhi64(g.s, m);
}

// f64::from_bits((1<<62) - (1<<52) + (u >> 12)) - 1.0  ∈ [0, 1)
RTX_HD double wy_f64(WyRand& g) {
  const uint64_t bits = 0x3FF0000000000000ULL + (wy_u64(g) >> 12);
#if defined(__CUDA_ARCH__)
  return fsub(__longlong_as_double((long long)bits), 1.0);
#else
  double d;
  __builtin_memcpy(&d, &bits, 8);
  return d - 1.0;
#endif
}

RTX_HD uint64_t mix64(uint64_t z) {
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}
RTX_HD uint64_t stream_state(uint64_t seed, uint64_t key) {
  return mix64(seed + 0x9E3779B97F4A7C15ULL * (key + 1ULL));
}

}  // namespace rtx
