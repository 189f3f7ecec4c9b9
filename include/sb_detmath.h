/*
 * sb_detmath.h -- the numeric contract of the smarts_b200 C-ABI.
 *
 * Every result the library produces is a pure function of its inputs evaluated with IEEE-754
 * binary64 +, -, *, /, sqrt, fmod, rint and comparisons only, in the order written in the
 * sources, with NO fused multiply-add contraction (nvcc -fmad=false, gcc -ffp-contract=off).
 * The transcendental functions below are therefore spelled out as fixed operation sequences
 * (argument reduction + the classic fdlibm minimax polynomials) instead of calling a vendor
 * libm, whose last-bit behaviour differs between glibc and CUDA.  That makes a GPU step and a
 * CPU step bit-identical, which is what lets the parity tests demand exact equality for
 * neighbour sets, waypoint indices and grid occupancy, and lets a rollout be replayed on
 * either side.
 *
 * This header is shared by the CUDA kernels (smarts_b200/csrc) and by the CPU oracle
 * (oracle/); it contains only libm-level primitives, none of the reference's algorithms.
 * Accuracy (checked in tests/test_oracle_golden.py against numpy): <= 2 ulp for sin/cos on
 * |x| <= 1e5 and for atan2.
 */
#ifndef SB_DETMATH_H
#define SB_DETMATH_H

#include <math.h>

#if defined(__CUDACC__)
#define SB_HD __host__ __device__ __forceinline__
#else
#define SB_HD static inline
#endif

#define SB_PI 3.14159265358979323846
#define SB_TWO_PI 6.28318530717958647692
#define SB_HALF_PI 1.57079632679489661923

/* Python's float % for b > 0 (used by Heading.__new__, coordinates.py:175-181, and
 * min_angles_difference_signed, utils/math.py:447-449). fmod is exact in IEEE arithmetic. */
/* fmod(a, b) for finite b > 0, bit-identical to the libm function: for |a| < 4b the quotient is
 * 0..3 and every subtraction below is exact (Sterbenz: y/2 <= x <= 2y), so the result is the
 * exact remainder that fmod returns -- without libm's iterative long division, which dominated
 * the angle-wrapping-heavy controller code on the GPU.  Larger |a|, inf and NaN go to fmod. */
SB_HD double sb_fmod_pos(double a, double b) {
  double r = fabs(a);
  if (r < b) return a;
  if (r < 4.0 * b) {
    if (r >= 2.0 * b) r -= 2.0 * b;
    if (r >= b) r -= b;
    return copysign(r, a);
  }
  return fmod(a, b);
}

SB_HD double sb_pymod(double a, double b) {
  double m = sb_fmod_pos(a, b);
  if (m != 0.0 && m < 0.0) m += b;
  return m;
}

SB_HD double sb_kernel_sin(double x) {
  const double S1 = -1.66666666666666324348e-01, S2 = 8.33333333332248946124e-03,
               S3 = -1.98412698298579493134e-04, S4 = 2.75573137070700676789e-06,
               S5 = -2.50507602534068634195e-08, S6 = 1.58969099521155010221e-10;
  double z = x * x;
  double v = z * x;
  double r = S2 + z * (S3 + z * (S4 + z * (S5 + z * S6)));
  return x + v * (S1 + z * r);
}

SB_HD double sb_kernel_cos(double x) {
  const double C1 = 4.16666666666666019037e-02, C2 = -1.38888888888741095749e-03,
               C3 = 2.48015872894767294178e-05, C4 = -2.75573143513906633035e-07,
               C5 = 2.08757232129817482790e-09, C6 = -1.13596475577881948265e-11;
  double z = x * x;
  double r = z * (C1 + z * (C2 + z * (C3 + z * (C4 + z * (C5 + z * C6)))));
  double hz = 0.5 * z;
  double w = 1.0 - hz;
  return w + (((1.0 - w) - hz) + z * r);
}

/* sin and cos of x (radians), |x| <= ~1e6. */
SB_HD void sb_sincos(double x, double *s, double *c) {
  const double INV_PIO2 = 6.36619772367581382433e-01;
  const double PIO2_1 = 1.57079632673412561417e+00;  /* first 33 bits of pi/2 */
  const double PIO2_1T = 6.07710050650619224932e-11; /* pi/2 - PIO2_1 */
  double k = rint(x * INV_PIO2);
  double r = (x - k * PIO2_1) - k * PIO2_1T;
  double ks = sb_kernel_sin(r), kc = sb_kernel_cos(r);
  long long q = (long long)k;
  switch ((int)(q & 3)) {
    case 0: *s = ks; *c = kc; break;
    case 1: *s = kc; *c = -ks; break;
    case 2: *s = -ks; *c = -kc; break;
    default: *s = -kc; *c = ks; break;
  }
}

SB_HD double sb_sin(double x) { double s, c; sb_sincos(x, &s, &c); return s; }
SB_HD double sb_cos(double x) { double s, c; sb_sincos(x, &s, &c); return c; }

/* atan(x), fdlibm s_atan.c scheme */
SB_HD double sb_atan(double x) {
  const double aT0 = 3.33333333333329318027e-01, aT1 = -1.99999999998764832476e-01,
               aT2 = 1.42857142725034663711e-01, aT3 = -1.11111104054623557880e-01,
               aT4 = 9.09088713343650656196e-02, aT5 = -7.69187620504482999495e-02,
               aT6 = 6.66107313738753120669e-02, aT7 = -5.83357013379057348645e-02,
               aT8 = 4.97687799461593236017e-02, aT9 = -3.65315727442169155270e-02,
               aT10 = 1.62858201153657823623e-02;
  double hi, lo;
  int id;
  double ax = fabs(x);
  if (ax > 1e300) return x > 0 ? SB_HALF_PI : -SB_HALF_PI;
  if (ax < 0.4375) {
    id = -1; hi = 0.0; lo = 0.0;
  } else if (ax < 1.1875) {
    if (ax < 0.6875) {
      id = 0; hi = 4.63647609000806093515e-01; lo = 2.26987774529616870924e-17;
      ax = (2.0 * ax - 1.0) / (2.0 + ax);
    } else {
      id = 1; hi = 7.85398163397448278999e-01; lo = 3.06161699786838301793e-17;
      ax = (ax - 1.0) / (ax + 1.0);
    }
  } else if (ax < 2.4375) {
    id = 2; hi = 9.82793723247329054082e-01; lo = 1.39033110312309984516e-17;
    ax = (ax - 1.5) / (1.0 + 1.5 * ax);
  } else {
    id = 3; hi = 1.57079632679489655800e+00; lo = 6.12323399573676603587e-17;
    ax = -1.0 / ax;
  }
  double z = ax * ax;
  double w = z * z;
  double s1 = z * (aT0 + w * (aT2 + w * (aT4 + w * (aT6 + w * (aT8 + w * aT10)))));
  double s2 = w * (aT1 + w * (aT3 + w * (aT5 + w * (aT7 + w * aT9))));
  double r;
  if (id < 0) {
    r = ax - ax * (s1 + s2);
  } else {
    r = hi - ((ax * (s1 + s2) - lo) - ax);
  }
  return x < 0 ? -r : r;
}

SB_HD double sb_atan2(double y, double x) {
  if (x > 0.0) return sb_atan(y / x);
  if (x < 0.0) {
    double a = sb_atan(y / x);
    return (y >= 0.0) ? a + SB_PI : a - SB_PI;
  }
  if (y > 0.0) return SB_HALF_PI;
  if (y < 0.0) return -SB_HALF_PI;
  return 0.0;
}

/* The simulation clock.  SMARTS advances it as  t <- round(t + dt, d)  with d the number of decimals of dt
 * (smarts.py:262, utils/math.py:553-569): the value after k ticks is the binary64 nearest to the decimal k * dt,
 * which is the correctly rounded quotient below (k * dt_num and dt_den = 10^d are exact).  Time-window tests
 * (DrivenPathSensor.distance_travelled, sensors.py:866-871) compare such clock values after a floating-point
 * subtraction, so they must be reproduced to the bit, not approximated by tick counts. */
SB_HD double sb_clock(long long ticks, long long dt_num, double dt_den) { return (double)(ticks * dt_num) / dt_den; }

/* d of the above for a given dt: the digits after the point in its shortest decimal form (<= 12). */
static inline int sb_clock_decimals(double dt, long long *dt_num, double *dt_den) {
  double den = 1.0;
  for (int d = 0; d <= 12; ++d, den *= 10.0) {
    const double num = rint(dt * den);
    if (num / den == dt) {
      *dt_num = (long long)num;
      *dt_den = den;
      return d;
    }
  }
  *dt_num = (long long)rint(dt * 1e12);
  *dt_den = 1e12;
  return 12;
}

#endif /* SB_DETMATH_H */
