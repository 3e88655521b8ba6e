// rtx_sincos.cuh — sin/cos with the bits of the reference's libm, host + device.
//
// Vector3::random_direction (math/vector.rs:36-45) calls f64::sin and f64::cos; Rust lowers them to the
// platform libm, on Linux glibc's __sin/__cos (sysdeps/ieee754/dbl-64/s_sin.c, the IBM Accurate Mathematical
// Library routines: table + short polynomials, < 1 ulp, NOT correctly rounded).  CUDA's sincos() is also
// < 1-2 ulp but rounds differently in ~15 % of the draws, and one ulp in a bounce direction can flip a later
// hit on triangle scenes.  This file restates glibc's algorithm operation by operation so
// that the device produces the same bits as the oracle's (and the reference's) libm.
//
// Pinned against: glibc 2.39 (Ubuntu 2.39-0ubuntu8.5), x86-64, the `__sin_fma`/`__cos_fma` ifunc variants —
// the ones every FMA-capable CPU selects.  Those are s_sin.c compiled with -mfma, so the compiler contracted
// a*b+c into FMAs; which ones is not visible in the C source, so the contraction pattern below was read from
// the disassembly of this image's libm.so.6 (each fma() here is a vfmadd/vfnmadd there, each fmul/fadd a
// vmulsd/vaddsd).  tests/test_sincos.py checks 10^6+ arguments per range against the live libm on the CPU,
// tests/test_gpu_functions.py checks the device against it on the GPU.
//
// Range: |x| < 105414350 (0x419921FB) like glibc's fast paths; beyond that glibc runs its multi-precision
// __branred reduction, which is not restated — callers here only pass theta in [0, 2*pi).
//
// Provenance and licence: the algorithm and the constants restated here, and the table in rtx_sincostab.inc, come from
// the GNU C Library (sysdeps/ieee754/dbl-64/s_sin.c, sincostab.c; Copyright (C) Free Software Foundation, Inc.), which
// is distributed under the GNU Lesser General Public License, version 2.1 or later.  This file and rtx_sincostab.inc
// are therefore offered under LGPL-2.1-or-later (SPDX-License-Identifier: LGPL-2.1-or-later), whatever licence the
// rest of this repository carries.
#pragma once
#include "rtx_math.cuh"

namespace rtx {

// glibc's __sincostab (440 doubles), extracted and cross-checked by tools/gen_sincostab.py.  On the device it
// lives in global memory and is read through the L1/read-only path: the index differs per lane, which
// __constant__ memory would serialise.
static const double SINCOS_TAB[440] = {
#include "rtx_sincostab.inc"
};
#if defined(__CUDACC__)
__device__ __align__(32) const double SINCOS_TAB_DEV[440] = {
#include "rtx_sincostab.inc"
};
#endif

namespace sc_detail {
RTX_HD double sc_fma(double a, double b, double c) {
#if defined(__CUDA_ARCH__)
  return __fma_rn(a, b, c);
#else
  return __builtin_fma(a, b, c);
#endif
}
RTX_HD double sc_abs(double x) { return fabs(x); }
RTX_HD double sc_copysign(double mag, double sgn) {
#if defined(__CUDA_ARCH__)
  return copysign(mag, sgn);
#else
  return __builtin_copysign(mag, sgn);
#endif
}
RTX_HD uint32_t sc_lo32(double x) {
#if defined(__CUDA_ARCH__)
  return (uint32_t)__double2loint(x);
#else
  uint64_t u;
  __builtin_memcpy(&u, &x, 8);
  return (uint32_t)u;
#endif
}

RTX_HD void sc_tab4(const double* tab, uint32_t k, double* sn, double* ssn, double* cs, double* ccs) {
#if defined(__CUDA_ARCH__)
  const double2 a = reinterpret_cast<const double2*>(tab + k)[0];  // k is a multiple of 4: 32-byte aligned
  const double2 b = reinterpret_cast<const double2*>(tab + k)[1];
  *sn = a.x, *ssn = a.y, *cs = b.x, *ccs = b.y;
#else
  *sn = tab[k], *ssn = tab[k + 1], *cs = tab[k + 2], *ccs = tab[k + 3];
#endif
}

// s_sin.c constants (values read back from the binary; they equal the source's)
constexpr double SC_BIG = 52776558133248.0;             // 0x1.8p45
constexpr double SC_TOINT = 6755399441055744.0;         // 0x1.8p52
constexpr double SC_HPINV = 0.6366197723675814;         // 2/pi
constexpr double SC_HP0 = 1.5707963267948966;           // pi/2 head
constexpr double SC_HP1 = 6.123233995736766e-17;        // pi/2 tail
constexpr double SC_MP1 = 1.5707963407039642;           // 0x3FF921FB58000000
constexpr double SC_MP2 = -1.3909067564377153e-08;      // 0xBE4DDE973C000000
constexpr double SC_PP3 = -4.9789962314799099e-17;      // 0xBC8CB3B398000000
constexpr double SC_PP4 = -1.9034889620193266e-25;      // 0xBACD747F23E32ED7
constexpr double SC_SN3 = -1.66666666666664880952546298448555E-01;
constexpr double SC_SN5 = 8.33333214285722277379541354343671E-03;
constexpr double SC_CS2 = 4.99999999999999999999950396842453E-01;
constexpr double SC_CS4 = -4.16666666666664434524222570944589E-02;
constexpr double SC_CS6 = 1.38888874007937613028114285595617E-03;
constexpr double SC_S1 = -0.16666666666666666;          // TAYLOR_SIN coefficients
constexpr double SC_S2 = 0.0083333333333323288;
constexpr double SC_S3 = -0.00019841269834414642;
constexpr double SC_S4 = 2.755729806860771e-06;
constexpr double SC_S5 = -2.5022014848318398e-08;

// TAYLOR_SIN(xx, a, da): ((POLYNOMIAL(xx) * a - 0.5 * da) * xx + da) + a
RTX_HD double taylor_sin(double a, double da) {
  const double xx = fmul(a, a);
  double pl = sc_fma(xx, SC_S5, SC_S4);
  pl = sc_fma(xx, pl, SC_S3);
  pl = sc_fma(xx, pl, SC_S2);
  pl = sc_fma(xx, pl, SC_S1);
  const double t1 = sc_fma(pl, a, -fmul(da, 0.5));  // vfmsub: pl * a - 0.5 * da
  const double t = sc_fma(xx, t1, da);
  return fadd(t, a);
}

// do_sin(x, dx) of s_sin.c: sin(x + dx) for |x| < 0.855469.  Its |x| < 0.126 case (TAYLOR_SIN) is evaluated
// alongside and selected, so all lanes execute the same instructions.
RTX_HD double do_sin(double x, double dx_in, const double* tab) {
  const double ax = sc_abs(x);
  const double dx = !(x > 0.0) ? -dx_in : dx_in;  // `if (x <= 0) dx = -dx`
  const double u = fadd(ax, SC_BIG);
  const double xr = fsub(ax, fsub(u, SC_BIG));
  const uint32_t k = sc_lo32(u) << 2;
  const double xx = fmul(xr, xr);
  const double s = fadd(xr, sc_fma(fmul(xr, xx), sc_fma(xx, SC_SN5, SC_SN3), dx));
  const double c = sc_fma(xr, dx, fmul(xx, sc_fma(xx, sc_fma(xx, SC_CS6, SC_CS4), SC_CS2)));
  double sn, ssn, cs, ccs;
  sc_tab4(tab, k, &sn, &ssn, &cs, &ccs);
  const double cor = sc_fma(s, cs, sc_fma(-c, sn, sc_fma(s, ccs, ssn)));
  const double r = sc_copysign(fadd(sn, cor), x);
  const double ty = taylor_sin(x, dx_in);
  return ax < 0.126 ? ty : r;
}

// do_cos(x, dx) of s_sin.c: cos(x + dx) for |x| < 0.855469
RTX_HD double do_cos(double x, double dx_in, const double* tab) {
  const double ax = sc_abs(x);
  const double dx = x < 0.0 ? -dx_in : dx_in;
  const double u = fadd(ax, SC_BIG);
  const double xr = fadd(fsub(ax, fsub(u, SC_BIG)), dx);
  const uint32_t k = sc_lo32(u) << 2;
  const double xx = fmul(xr, xr);
  const double s = sc_fma(fmul(xr, xx), sc_fma(xx, SC_SN5, SC_SN3), xr);
  const double c = fmul(xx, sc_fma(xx, sc_fma(xx, SC_CS6, SC_CS4), SC_CS2));
  double sn, ssn, cs, ccs;
  sc_tab4(tab, k, &sn, &ssn, &cs, &ccs);
  const double cor = sc_fma(-s, sn, sc_fma(-c, cs, sc_fma(-s, ssn, ccs)));
  return fadd(cs, cor);
}

// reduce_sincos: x = n * pi/2 + (a + da), |a| <= pi/4
RTX_HD uint32_t reduce_sincos(double x, double* a, double* da) {
  const double t = sc_fma(x, SC_HPINV, SC_TOINT);
  const double xn = fsub(t, SC_TOINT);
  const uint32_t n = sc_lo32(t) & 3u;
  const double y = sc_fma(-xn, SC_MP2, sc_fma(-xn, SC_MP1, x));
  const double t2 = sc_fma(-xn, SC_PP3, y);
  double db = sc_fma(-xn, SC_PP3, fsub(y, t2));
  const double b = sc_fma(-xn, SC_PP4, t2);
  db = fadd(db, sc_fma(-xn, SC_PP4, fsub(t2, b)));
  *a = b;
  *da = db;
  return n;
}

RTX_HD uint32_t sc_hi_abs(double x) {
#if defined(__CUDA_ARCH__)
  return (uint32_t)__double2hiint(x) & 0x7fffffffu;
#else
  uint64_t u;
  __builtin_memcpy(&u, &x, 8);
  return (uint32_t)(u >> 32) & 0x7fffffffu;
#endif
}
}  // namespace sc_detail

// __sin and __cos of s_sin.c for |x| < 105414350, both at once.  Range dispatch of the original:
//   |x| < 2^-26 (sin) / 2^-27 (cos)      x                                   / 1.0
//   |x| < 0.855469                       do_sin(x, 0)                        / do_cos(x, 0)
//   |x| < 2.426265                       copysign(do_cos(y, hp1), x)         / do_sin(a, da): y = hp0 - |x|,
//                                                                              a = y + hp1, da = (y - a) + hp1
//   |x| < 105414350                      do_sincos(a, da, n)                 / do_sincos(a, da, n + 1)
// In every range exactly one of the two results comes from a do_sin and the other from a do_cos, so this
// evaluates one of each, branch-free: the range only selects their arguments and which result goes where.
// `tab` = the 440-entry table (sincos_table() in global memory, or a copy staged in shared memory).
RTX_HD void libm_sincos(double x, const double* tab, double* out_sin, double* out_cos) {
  using namespace sc_detail;
  const uint32_t k = sc_hi_abs(x);
  const bool in_b = k < 0x3feb6000u, in_c = !in_b && k < 0x400368fdu, in_d = !in_b && !in_c;
  double ra, rda;
  const uint32_t n = reduce_sincos(x, &ra, &rda);
  const double y = fsub(SC_HP0, sc_abs(x));
  const double ac = fadd(y, SC_HP1);
  const double dac = fadd(fsub(y, ac), SC_HP1);
  // arguments of the do_sin and of the do_cos
  const double s_a = in_b ? x : (in_c ? ac : ra), s_da = in_b ? 0.0 : (in_c ? dac : rda);
  const double c_a = in_b ? x : (in_c ? y : ra), c_da = in_b ? 0.0 : (in_c ? SC_HP1 : rda);
  const double vs = do_sin(s_a, s_da, tab), vc = do_cos(c_a, c_da, tab);
  // the sine comes from the do_cos in range C and for odd n in range D
  const bool swap = in_c || (in_d && (n & 1u));
  double rs = swap ? vc : vs, rc = swap ? vs : vc;
  if (in_c) rs = sc_copysign(rs, x);
  if (in_d && (n & 2u)) rs = -rs;
  if (in_d && ((n + 1u) & 2u)) rc = -rc;
  if (k < 0x3e500000u) rs = x;
  if (k < 0x3e400000u) rc = 1.0;
  *out_sin = rs;
  *out_cos = rc;
}

RTX_HD const double* sincos_table() {
#if defined(__CUDA_ARCH__)
  return SINCOS_TAB_DEV;
#else
  return SINCOS_TAB;
#endif
}
RTX_HD double libm_sin(double x) {
  double s, c;
  libm_sincos(x, sincos_table(), &s, &c);
  return s;
}
RTX_HD double libm_cos(double x) {
  double s, c;
  libm_sincos(x, sincos_table(), &s, &c);
  return c;
}

}  // namespace rtx
