/*
 * fkmath.h -- portable, deterministic double-precision sin / cos / asin / atan2.
 *
 * WHY: tinyfk's FK path calls libm (reference urdf_parser/include/urdf_model/pose.h:148-156,
 * 181-189, 210-213; joint.h:219-236; src/kinematics.cpp:18-21) and three of its Jacobian
 * columns are a forward finite difference with eps = 1e-7 (src/kinematics.cpp:201-232), which
 * amplifies a 1-ulp difference in any libm result to ~1e-9.  The 1e-12 parity contract on those
 * columns is therefore only reachable with bit-identical transcendental functions on the CPU
 * and on the GPU.  CUDA's libdevice and glibc disagree in the last bit on a few percent of the
 * arguments, so this header provides ONE implementation used by
 *   - the sm_100a kernels (nvcc, strict translation unit built with -fmad=false), and
 *   - the host model code and the "strict" oracle (gcc, -ffp-contract=off).
 * It uses only IEEE-754 binary64  + - * / sqrt and fusedMultiplyAdd (TFK_FMA: one correctly rounded
 * operation on both sides -- `__fma_rn` on the device, `fma()` on the host, which glibc evaluates exactly
 * with or without FMA hardware), comparisons and integer bit moves, so the two sides round identically.
 * The explicit fused operations are part of the DEFINITION of these functions (they do not depend on
 * -fmad / -ffp-contract).
 *
 * Algorithms (table-free, straight-line on ordinary arguments; details at each function): sin / cos -- one
 * three-step fused reduction by pi/2 for every argument below 2^20 * pi/2 and Horner kernels on the reduced
 * argument, no tail word; asin -- one polynomial for both ranges, no division, the square root's rounding error
 * restored from its exact residual; atan2 -- one division (numerator and denominator chosen before dividing), one
 * polynomial on |t| <= 1/2, offsets k * pi/4.  tfk_atan (only behind the `atan` symbol of the pinned reference
 * build) keeps the classic four-interval reduction.  The coefficients are our own fits
 * (tools/gen_fkmath_consts.py, mpmath).  Measured accuracy vs mpmath (tests/test_fkmath.py): sin, cos < 1.5 ulp,
 * asin < 1 ulp (0.60), atan2 < 2 ulp (1.52).  DESIGN.md section 3, "Arithmetic building blocks", has the history.
 *
 * Domain note: sin/cos are accurate for |x| < 2^38; beyond that the reduction is crude (but still
 * deterministic and identical on both sides).
 */
#ifndef TFK_FKMATH_H
#define TFK_FKMATH_H

#if defined(__CUDACC_RTC__) /* run-time compilation (jit.cpp): no libc headers, the CUDA built-ins are implicit */
typedef signed int int32_t;
typedef unsigned int uint32_t;
typedef long long int64_t;
typedef unsigned long long uint64_t;
#else
#include <math.h>
#include <stdint.h>
#include <string.h>
#endif

/* rare branches: keep the common path the fall-through one (taken branches cost instruction-fetch bubbles) */
#if defined(__GNUC__) || defined(__CUDACC__)
#define TFK_UNLIKELY(x) __builtin_expect(!!(x), 0)
#define TFK_LIKELY(x) __builtin_expect(!!(x), 1)
#else
#define TFK_UNLIKELY(x) (x)
#define TFK_LIKELY(x) (x)
#endif

#if defined(__CUDACC__)
#define TFK_HD __host__ __device__ inline
/* out-of-line on the device: the kernels call these from a dozen sites and the instruction cache is
   the scarcer resource ("code size") */
#define TFK_HD_CALL static __host__ __device__ __noinline__
#else
#define TFK_HD static inline
#define TFK_HD_CALL static inline
#endif

#include "fkmath_consts.inc"

/* changes whenever a result of this header may change in any bit (keys the caches of values derived from it) */
#define TFK_FKMATH_VERSION "fkmath 4 (fma, one-division atan2, tail-free sincos, division-free asin)"

/* a*b + c with ONE rounding, identical on host and device */
#if defined(__CUDA_ARCH__)
#define TFK_FMA(a, b, c) __fma_rn((a), (b), (c))
#else
#define TFK_FMA(a, b, c) __builtin_fma((a), (b), (c))
#endif

TFK_HD int32_t tfk_hi32(double x) {
#if defined(__CUDA_ARCH__)
  return __double2hiint(x);
#else
  uint64_t u;
  memcpy(&u, &x, sizeof u);
  return (int32_t)(u >> 32);
#endif
}

TFK_HD double tfk_from_hi32(int32_t hi) { /* low word = 0 */
#if defined(__CUDA_ARCH__)
  return __hiloint2double(hi, 0);
#else
  uint64_t u = ((uint64_t)(uint32_t)hi) << 32;
  double x;
  memcpy(&x, &u, sizeof x);
  return x;
#endif
}

TFK_HD int tfk_signbit(double x) { return (int)(((uint32_t)tfk_hi32(x)) >> 31); }

/* ---- sin / cos ------------------------------------------------------------------------------------------------
 * One reduction for every argument below 2^20 * pi/2 (no "small argument" case: n = 0 reduces x to itself), three
 * fused steps x - n*(P1 + P1T + P1TT) with 33 + 53 + 53 bits of pi/2 (the first is exact, the second is the one rounding of
 * the reduced argument, the third only acts right next to a multiple of pi/2, where the second leaves few bits), and
 * plain Horner kernels on the reduced argument.  No tail word, no data-dependent branch: 26 FP64 operations for sin and
 * cos together.  Accuracy (tests/test_fkmath.py, mpmath): below 1.5 ulp. */
#define TFK_RND_MAGIC (6755399441055744.0) /* 1.5 * 2^52: adding it rounds to the nearest integer (ties to even) */

/* sin(x), |x| <= pi/4 (+ rounding): x + x^3 * S(x^2) */
TFK_HD double tfk_ksin(double x) {
  const double z = x * x;
  const double v = z * x;
  const double r = TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_S6, TFK_S5), TFK_S4), TFK_S3), TFK_S2);
  return TFK_FMA(v, TFK_FMA(z, r, TFK_S1), x);
}

/* cos(x), |x| <= pi/4 (+ rounding): 1 + z * (-1/2 + z * C(z)) */
TFK_HD double tfk_kcos(double x) {
  const double z = x * x;
  const double q =
      TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_C6, TFK_C5), TFK_C4), TFK_C3), TFK_C2), TFK_C1), -0.5);
  return TFK_FMA(z, q, 1.0);
}

/* t = n*(pi/2) + a0 for 0 <= t < 2^20 * pi/2; *n gets the quadrant count */
TFK_HD double tfk_reduce(double t, int32_t *n) {
  const double m = TFK_FMA(t, TFK_INVPIO2, TFK_RND_MAGIC);
  const double fn = m - TFK_RND_MAGIC;
#if defined(__CUDA_ARCH__)
  *n = __double2loint(m);
#else
  uint64_t u;
  memcpy(&u, &m, sizeof u);
  *n = (int32_t)(uint32_t)u;
#endif
  const double r = TFK_FMA(-fn, TFK_PIO2_1, t); /* exact: 33-bit piece times an integer below 2^21 */
  const double a = TFK_FMA(-fn, TFK_PIO2_1T, r);  /* the one rounding that matters */
  return TFK_FMA(-fn, TFK_PIO2_1TT, a);           /* changes a only next to a multiple of pi/2, where a is nearly exact */
}

/* (sin, cos) from the kernels' values on the reduced argument of |x|: quadrant n, sign of x (bit 31 of hx) */
TFK_HD void tfk_sincos_place(double ks, double kc, int32_t n, int32_t hx, double *s, double *c) {
  /* n&3: 0 (ks, kc)  1 (kc, -ks)  2 (-ks, -kc)  3 (-kc, ks);  sin is odd in x */
  double ss = (n & 1) ? kc : ks;
  double cc = (n & 1) ? ks : kc;
  const int32_t ms = (int32_t)((uint32_t)n << 30) ^ hx;  /* (n & 2) xor sign(x) -> bit 31 */
  const int32_t mc = (int32_t)((uint32_t)(n + 1) << 30); /* (n + 1) & 2 -> bit 31 */
#if defined(__CUDA_ARCH__)
  *s = __hiloint2double(__double2hiint(ss) ^ (ms & (int32_t)0x80000000), __double2loint(ss));
  *c = __hiloint2double(__double2hiint(cc) ^ (mc & (int32_t)0x80000000), __double2loint(cc));
#else
  uint64_t u;
  memcpy(&u, &ss, sizeof u);
  u ^= ((uint64_t)((uint32_t)ms & 0x80000000u)) << 32;
  memcpy(s, &u, sizeof u);
  memcpy(&u, &cc, sizeof u);
  u ^= ((uint64_t)((uint32_t)mc & 0x80000000u)) << 32;
  memcpy(c, &u, sizeof u);
#endif
}

/* the whole evaluation for |x| < 2^20 * pi/2 (1647099.0): straight-line code */
TFK_HD void tfk_sincos_core(double x, double *s, double *c) {
  int32_t n;
  const double a0 = tfk_reduce(fabs(x), &n);
  tfk_sincos_place(tfk_ksin(a0), tfk_kcos(a0), n, tfk_hi32(x), s, c);
}

/* sin and cos of the same argument.  tfk_sin / tfk_cos are its two halves, so the reference's separate sin() and
   cos() calls and a compiler-fused sincos() give the same bits. */
TFK_HD_CALL void tfk_sincos(double x, double *s, double *c) {
  double t = fabs(x);
  if (TFK_LIKELY(t < 1647099.0)) { /* ~ 2^20 * pi/2 */
    tfk_sincos_core(x, s, c);
    return;
  }
  if (!(t <= 1.7976931348623157e308)) { /* inf, nan */
    *s = *c = x - x;
    return;
  }
  /* big arguments: first strip a quadrant-neutral multiple of 4 quadrants, keep the rounding error as a tail */
  double tl = 0.0;
  if (t >= 274877906944.0) { /* 2^38: crude, documented */
    double k = floor(t / (4.0 * TFK_PIO2_HI));
    t = t - k * (4.0 * TFK_PIO2_HI);
    if (t < 0.0) t = 0.0;
  } else {
    /* remove a multiple of 2^18 quadrants (<= 20 significant bits => products with the
       33-bit pieces are exact); keep the rounding error of the second subtraction */
    double fq = floor(t * TFK_INVPIO2 * (1.0 / 262144.0)) * 262144.0;
    double t1 = t - fq * TFK_PIO2_1; /* exact */
    double w = fq * TFK_PIO2_2;      /* exact */
    t = t1 - w;
    tl = ((t1 - t) - w) - fq * TFK_PIO2_2T;
    if (t < 0.0) {
      t = 0.0;
      tl = 0.0;
    }
  }
  int32_t n;
  const double a0 = tfk_reduce(t, &n) + tl;
  tfk_sincos_place(tfk_ksin(a0), tfk_kcos(a0), n, tfk_hi32(x), s, c);
}

TFK_HD double tfk_sin(double x) {
  double s, c;
  tfk_sincos(x, &s, &c);
  return s;
}

TFK_HD double tfk_cos(double x) {
  double s, c;
  tfk_sincos(x, &s, &c);
  return c;
}

/* ---- asin -----------------------------------------------------------------------------------------------------
 * No division.  |x| < 1/2: x + x * (z R(z)), z = x^2, R a 14-term polynomial.  Otherwise asin|x| = pi/2 - 2 asin(s),
 * s = sqrt(t), t = (1 - |x|)/2 <= 1/4, with the SAME polynomial: pi/2 - 2 (s + c + s t R(t)), where c = e / (2 s) restores
 * what the rounding of the square root lost (e = t - s^2 comes exactly out of one fused operation; c is below an ulp of
 * s, so the reciprocal of s in single precision -- an IEEE operation on both sides -- is more than enough), and
 * pi/2 - 2 s is taken as a rounded value v plus its exact error dv.  Accuracy: below 1 ulp (tests/test_fkmath.py). */
TFK_HD double tfk_asin_poly(double z) { /* z * R(z) */
  double r = TFK_AR13;
  r = TFK_FMA(z, r, TFK_AR12);
  r = TFK_FMA(z, r, TFK_AR11);
  r = TFK_FMA(z, r, TFK_AR10);
  r = TFK_FMA(z, r, TFK_AR9);
  r = TFK_FMA(z, r, TFK_AR8);
  r = TFK_FMA(z, r, TFK_AR7);
  r = TFK_FMA(z, r, TFK_AR6);
  r = TFK_FMA(z, r, TFK_AR5);
  r = TFK_FMA(z, r, TFK_AR4);
  r = TFK_FMA(z, r, TFK_AR3);
  r = TFK_FMA(z, r, TFK_AR2);
  r = TFK_FMA(z, r, TFK_AR1);
  r = TFK_FMA(z, r, TFK_AR0);
  return z * r;
}

/* 1 / (2 s) to single precision, identically on host and device (correctly rounded conversions and reciprocal) */
TFK_HD double tfk_half_rcp_f32(double s) {
#if defined(__CUDA_ARCH__)
  return 0.5 * (double)__frcp_rn(__double2float_rn(s));
#else
  const float sf = (float)s;
  const float rf = 1.0f / sf;
  return 0.5 * (double)rf;
#endif
}

/* asin of 1/2 <= ax < 1 (ax = |x|) */
TFK_HD double tfk_asin_upper(double ax, double zr, double s, double t) { /* zr = t R(t), s = sqrt(t) */
  const double e = TFK_FMA(-s, s, t);
  const double c = e * tfk_half_rcp_f32(s);
  const double u = TFK_FMA(s, zr, c);
  const double v = TFK_FMA(-2.0, s, TFK_PIO2_HI);
  const double dv = TFK_FMA(-2.0, s, TFK_PIO2_HI - v); /* exact */
  return v + (TFK_FMA(-2.0, u, TFK_PIO2_LO) + dv);
}

TFK_HD_CALL double tfk_asin(double x) {
  double ax = fabs(x);
  if (!(ax < 1.0)) {
    if (ax == 1.0) return x * TFK_PIO2_HI + x * TFK_PIO2_LO;
    return (x - x) / (x - x); /* |x| > 1 or nan */
  }
  if (ax < 0.5) {
    if (ax < 7.450580596923828125e-9) return x; /* 2^-27 */
    return TFK_FMA(x, tfk_asin_poly(x * x), x);
  }
  const double t = TFK_FMA(-0.5, ax, 0.5); /* exact */
  const double r = tfk_asin_upper(ax, tfk_asin_poly(t), sqrt(t), t);
  return (x > 0.0) ? r : -r;
}

TFK_HD double tfk_atan(double x) {
  double ax = fabs(x);
  if (!(ax < 7.378697629483821e19)) { /* 2^66, inf, nan */
    if (ax != ax) return x + x;
    return (x > 0.0) ? (TFK_ATANHI3 + TFK_ATANLO3) : -(TFK_ATANHI3 + TFK_ATANLO3);
  }
  int id;
  double t, hi = 0.0, lo = 0.0;
  if (ax < 0.4375) {
    if (ax < 1.862645149230957e-9) return x; /* 2^-29 */
    id = -1;
    t = x;
  } else if (ax < 0.6875) {
    id = 0;
    t = TFK_FMA(2.0, ax, -1.0) / (ax + 2.0);
    hi = TFK_ATANHI0;
    lo = TFK_ATANLO0;
  } else if (ax < 1.1875) {
    id = 1;
    t = (ax - 1.0) / (ax + 1.0);
    hi = TFK_ATANHI1;
    lo = TFK_ATANLO1;
  } else if (ax < 2.4375) {
    id = 2;
    t = (ax - 1.5) / TFK_FMA(1.5, ax, 1.0);
    hi = TFK_ATANHI2;
    lo = TFK_ATANLO2;
  } else {
    id = 3;
    t = -1.0 / ax;
    hi = TFK_ATANHI3;
    lo = TFK_ATANLO3;
  }
  double z = t * t;
  double s = TFK_AT11;
  s = TFK_FMA(z, s, TFK_AT10);
  s = TFK_FMA(z, s, TFK_AT9);
  s = TFK_FMA(z, s, TFK_AT8);
  s = TFK_FMA(z, s, TFK_AT7);
  s = TFK_FMA(z, s, TFK_AT6);
  s = TFK_FMA(z, s, TFK_AT5);
  s = TFK_FMA(z, s, TFK_AT4);
  s = TFK_FMA(z, s, TFK_AT3);
  s = TFK_FMA(z, s, TFK_AT2);
  s = TFK_FMA(z, s, TFK_AT1);
  s = TFK_FMA(z, s, TFK_AT0);
  s = z * s;
  if (id < 0) return t - t * s;
  double r = hi - (TFK_FMA(t, s, -lo) - t);
  return (x < 0.0) ? -r : r;
}

/* ---- atan2 ----------------------------------------------------------------------------------------------------
 * ONE division: with mx = max(|x|, |y|), mn = min(|x|, |y|) the angle of (mx, mn) is atan(mn / mx) in [0, pi/4];
 * from mn / mx = 1/2 on it is pi/4 + atan((mn - mx) / (mn + mx)), so numerator and denominator are chosen BEFORE
 * dividing and the quotient t lies in [-1/3, 1/2], inside the range of the atan polynomial (|t| <= 1/2); with the split
 * at 1/2 the difference mn - mx is exact (Sterbenz), which keeps the error below 2 ulp where pi/4 + atan(t) is smallest.  Undoing the
 * swap (pi/2 - .), the reflection for x < 0 (pi - .) and the pi/4 offset all add a multiple of pi/4 and possibly flip the
 * sign of atan(t):  theta = k * pi/4 + sigma * atan(t), k in 0..4, evaluated as hi_k - ((sigma t * s - lo_k) - sigma t)
 * (the form tfk_atan uses); the result takes the sign of y.  Straight-line code with selects.
 * Preconditions (tfk_atan2 establishes them): x, y finite, max(|x|, |y|) in [2^-900, 2^1022). */

TFK_HD double tfk_atan_poly(double z) { /* atan(t) = t - t * (z * A(z)), z = t^2; returns z * A(z) */
  double s = TFK_AT11;
  s = TFK_FMA(z, s, TFK_AT10);
  s = TFK_FMA(z, s, TFK_AT9);
  s = TFK_FMA(z, s, TFK_AT8);
  s = TFK_FMA(z, s, TFK_AT7);
  s = TFK_FMA(z, s, TFK_AT6);
  s = TFK_FMA(z, s, TFK_AT5);
  s = TFK_FMA(z, s, TFK_AT4);
  s = TFK_FMA(z, s, TFK_AT3);
  s = TFK_FMA(z, s, TFK_AT2);
  s = TFK_FMA(z, s, TFK_AT1);
  s = TFK_FMA(z, s, TFK_AT0);
  return z * s;
}

#if !defined(TFK_ATAN2_TABLE)
/* k * pi/4 as hi + lo, k = 0..4 (fkmath_simt.h replaces the lookup by a constant-bank table on the device) */
#define TFK_ATAN2_TABLE(k, hi, lo)                                                                                    \
  do {                                                                                                                \
    const double tfk_hi_[5] = {0.0, TFK_PIO4_HI, TFK_PIO2_HI, TFK_3PIO4_HI, TFK_PI};                                  \
    const double tfk_lo_[5] = {0.0, TFK_PIO4_LO, TFK_PIO2_LO, TFK_3PIO4_LO, TFK_PI_LO};                               \
    (hi) = tfk_hi_[k];                                                                                                \
    (lo) = tfk_lo_[k];                                                                                                \
  } while (0)
#endif

TFK_HD double tfk_atan2_core(double y, double x) {
  const double ax = fabs(x), ay = fabs(y);
  const int swap = ay > ax;
  const double mx = swap ? ay : ax, mn = swap ? ax : ay;
  const int big = mn >= 0.5 * mx;
  const double num = big ? mn - mx : mn;
  const double den = big ? mn + mx : mx;
  const double t = num / den;
  const double s = tfk_atan_poly(t * t);
  const int xneg = tfk_hi32(x) < 0;
  const int flip = swap ^ xneg; /* sigma = -1 */
  const int k = (swap ? 2 : (xneg ? 4 : 0)) + (flip ? -big : big);
  double hi, lo;
  TFK_ATAN2_TABLE(k, hi, lo);
  const double tv = flip ? -t : t;
  const double th = hi - (TFK_FMA(tv, s, -lo) - tv);
  return (tfk_hi32(y) < 0) ? -th : th;
}

TFK_HD_CALL double tfk_atan2(double y, double x) {
  if (x != x || y != y) return x + y;
  int m = tfk_signbit(y) | (tfk_signbit(x) << 1);
  if (y == 0.0) {
    switch (m) {
    case 0:
    case 1:
      return y; /* atan2(+-0, +anything) = +-0 */
    case 2:
      return TFK_PI;
    default:
      return -TFK_PI;
    }
  }
  if (x == 0.0) return (m & 1) ? -TFK_PIO2_HI : TFK_PIO2_HI;
  double ax = fabs(x), ay = fabs(y);
  const double inf = 1.7976931348623157e308;
  if (ax > inf) {
    if (ay > inf) {
      switch (m) {
      case 0:
        return TFK_PIO4_HI;
      case 1:
        return -TFK_PIO4_HI;
      case 2:
        return 3.0 * TFK_PIO4_HI;
      default:
        return -3.0 * TFK_PIO4_HI;
      }
    }
    switch (m) {
    case 0:
      return 0.0;
    case 1:
      return -0.0;
    case 2:
      return TFK_PI;
    default:
      return -TFK_PI;
    }
  }
  if (ay > inf) return (m & 1) ? -TFK_PIO2_HI : TFK_PIO2_HI;
  /* bring the larger magnitude into the core's range; powers of two, so the quotient does not change (a smaller
     operand that leaves the normal range on the way belongs to a quotient far below one ulp of the result) */
  const double mxv = (ay > ax) ? ay : ax;
  if (mxv >= 4.49423283715579e307) { /* 2^1022 */
    x *= 0.0625;
    y *= 0.0625;
  } else if (mxv < 1.1830521861667747e-271) { /* 2^-900 */
    x *= 1.157920892373162e77; /* 2^256 */
    y *= 1.157920892373162e77;
  }
  return tfk_atan2_core(y, x);
}

#endif /* TFK_FKMATH_H */
