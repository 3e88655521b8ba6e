// sincos_check.cpp — compares rtx::libm_sin / rtx::libm_cos (rust-raytracing_b200/csrc/rtx_sincos.cuh, host
// build) with the sin/cos of the libm this process links: the functions the oracle — and, through Rust's
// f64::sin/cos, the reference (math/vector.rs:39-43) — call.  Prints one line per range; exit code 0 when every
// result is bit-identical.  usage: sincos_check [samples per range]
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "rtx_sincos.cuh"

static uint64_t state = 88172645463325252ull;
static inline uint64_t next() {
  state ^= state << 13, state ^= state >> 7, state ^= state << 17;
  return state;
}

int main(int argc, char** argv) {
  const long n = argc > 1 ? atol(argv[1]) : 2000000;
  // the first range is exactly the argument random_direction forms: (u * 2.0) * PI with u = k * 2^-52
  const double ranges[][2] = {{0, 0},       {0, 0.13},   {0.12, 0.9}, {0.8, 2.5},   {2.4, 6.3},
                              {-6.3, 6.3}, {0, 1e-7}, {6, 1000},   {1000, 1.0e8}};
  long total_bad = 0;
  for (auto& r : ranges) {
    long bad_s = 0, bad_c = 0;
    double first = 0;
    for (long i = 0; i < n; i++) {
      const uint64_t bits = next();
      double x;
      if (r[1] == 0) {
        const double u = (double)(bits >> 12) * (1.0 / 4503599627370496.0);
        x = (u * 2.0) * 3.14159265358979323846264338327950288;
      } else {
        x = r[0] + (r[1] - r[0]) * ((bits >> 11) * (1.0 / 9007199254740992.0));
      }
      volatile double xv = x;  // keep the compiler from folding or vectorising the libm calls
      const double a = rtx::libm_sin(x), b = std::sin(xv), c = rtx::libm_cos(x), d = std::cos(xv);
      if (std::memcmp(&a, &b, 8)) {
        if (!bad_s && !bad_c) first = x;
        bad_s++;
      }
      if (std::memcmp(&c, &d, 8)) {
        if (!bad_s && !bad_c) first = x;
        bad_c++;
      }
    }
    std::printf("range [%g, %g)%s: sin mismatches %ld, cos mismatches %ld of %ld (first at %.17g)\n", r[0], r[1],
                r[1] == 0 ? " = theta of random_direction" : "", bad_s, bad_c, n, first);
    total_bad += bad_s + bad_c;
  }
  return total_bad ? 1 : 0;
}
