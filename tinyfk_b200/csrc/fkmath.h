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
 * -fmad / -ffp-contract): every Horner step and every "product minus correction" of the kernels is one
 * of them, which halves the FP64 instruction count of the polynomials on the GPU (DESIGN.md section 3,
 * "libm with explicit fused operations") and does not hurt accuracy.
 *
 * Algorithms: classic table-free ones (Cody-Waite three-stage pi/2 reduction; odd / even
 * polynomial kernels on [-pi/4, pi/4]; rational kernel for asin with the sqrt identity above 0.5;
 * four-interval reduction for atan).  This structure follows the published fdlibm design
 * (Sun Microsystems, 1993, freely redistributable); the coefficients are our own fit
 * (tools/gen_fkmath_consts.py, mpmath).  Measured accuracy vs mpmath: see tests/test_fkmath.py
 * (sin, cos, asin < 1 ulp; atan2 < 2 ulp).
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
#define TFK_FKMATH_VERSION "fkmath 2 (explicit fma)"

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

/* sin on [-pi/4, pi/4]; (x + y) is the reduced argument, |y| << |x| */
TFK_HD double tfk_ksin(double x, double y) {
  if (fabs(x) < 7.450580596923828125e-9) return x; /* 2^-27 */
  double z = x * x;
  double v = z * x;
  double r = TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_S6, TFK_S5), TFK_S4), TFK_S3), TFK_S2);
  /* x - ((z*(y/2 - v*r) - y) - v*S1) */
  return x - TFK_FMA(-v, TFK_S1, TFK_FMA(z, TFK_FMA(-v, r, 0.5 * y), -y));
}

/* cos on [-pi/4, pi/4] */
TFK_HD double tfk_kcos(double x, double y) {
  double ax = fabs(x);
  if (ax < 7.450580596923828125e-9) return 1.0;
  double z = x * x;
  double r = z * TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_C6, TFK_C5), TFK_C4), TFK_C3), TFK_C2), TFK_C1);
  double e = TFK_FMA(z, r, -(x * y)); /* z*r - x*y */
  if (ax < 0.3) return 1.0 - (0.5 * z - e);
  /* 1 - z/2 would lose a bit here: peel off qx ~ |x|/4 so that (1 - qx) is exact */
  double qx = (ax > 0.78125) ? 0.28125 : tfk_from_hi32((tfk_hi32(ax)) - 0x00200000);
  double hz = TFK_FMA(0.5, z, -qx);
  double a = 1.0 - qx;
  return a - (hz - e);
}

/* x = n*(pi/2) + (y0 + y1), |y0 + y1| <= pi/4.  Precondition: |x| > pi/4, finite. */
TFK_HD int tfk_rem_pio2(double x, double *y0, double *y1) {
  double t = fabs(x);
  double tl = 0.0; /* tail carried over from the big-argument pre-reduction */
  if (t >= 1647099.0) { /* ~ 2^20 * pi/2: first strip a quadrant-neutral multiple of 4 quadrants */
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
  }
  int n = 0;
  double a0, a1;
  if (t <= TFK_PIO4_HI) {
    a0 = t;
    a1 = 0.0;
  } else {
    n = (int)TFK_FMA(t, TFK_INVPIO2, 0.5);
    double fn = (double)n;
    double r = TFK_FMA(-fn, TFK_PIO2_1, t); /* the product is exact (33-bit piece) */
    double w = fn * TFK_PIO2_1T;
    a0 = r - w;
    int j = (tfk_hi32(t) >> 20) & 0x7ff;
    int i = j - ((tfk_hi32(a0) >> 20) & 0x7ff);
    if (i > 16) { /* cancellation: bring in the second piece */
      double t2 = r;
      w = fn * TFK_PIO2_2;
      r = t2 - w;
      w = fn * TFK_PIO2_2T - ((t2 - r) - w);
      a0 = r - w;
      i = j - ((tfk_hi32(a0) >> 20) & 0x7ff);
      if (i > 49) { /* and the third */
        t2 = r;
        w = fn * TFK_PIO2_3;
        r = t2 - w;
        w = fn * TFK_PIO2_3T - ((t2 - r) - w);
        a0 = r - w;
      }
    }
    a1 = (r - a0) - w;
  }
  if (tl != 0.0) {
    a1 = a1 + tl;
    double h = a0 + a1;
    a1 = (a0 - h) + a1;
    a0 = h;
  }
  if (x < 0.0) {
    *y0 = -a0;
    *y1 = -a1;
    return -n;
  }
  *y0 = a0;
  *y1 = a1;
  return n;
}

TFK_HD double tfk_sin(double x) {
  if (fabs(x) <= TFK_PIO4_HI) return tfk_ksin(x, 0.0);
  if (!(fabs(x) <= 1.7976931348623157e308)) return x - x; /* inf, nan */
  double y0, y1;
  int n = tfk_rem_pio2(x, &y0, &y1);
  switch (n & 3) {
  case 0:
    return tfk_ksin(y0, y1);
  case 1:
    return tfk_kcos(y0, y1);
  case 2:
    return -tfk_ksin(y0, y1);
  default:
    return -tfk_kcos(y0, y1);
  }
}

TFK_HD double tfk_cos(double x) {
  if (fabs(x) <= TFK_PIO4_HI) return tfk_kcos(x, 0.0);
  if (!(fabs(x) <= 1.7976931348623157e308)) return x - x;
  double y0, y1;
  int n = tfk_rem_pio2(x, &y0, &y1);
  switch (n & 3) {
  case 0:
    return tfk_kcos(y0, y1);
  case 1:
    return -tfk_ksin(y0, y1);
  case 2:
    return -tfk_kcos(y0, y1);
  default:
    return tfk_ksin(y0, y1);
  }
}

/* sin and cos of the same argument with one reduction; results are bit-identical to the two
   separate calls (same kernels on the same reduced argument). */
TFK_HD_CALL void tfk_sincos(double x, double *s, double *c) {
  if (fabs(x) <= TFK_PIO4_HI) {
    *s = tfk_ksin(x, 0.0);
    *c = tfk_kcos(x, 0.0);
    return;
  }
  if (!(fabs(x) <= 1.7976931348623157e308)) {
    *s = *c = x - x;
    return;
  }
  double y0, y1;
  int n = tfk_rem_pio2(x, &y0, &y1);
  double ks = tfk_ksin(y0, y1), kc = tfk_kcos(y0, y1);
  switch (n & 3) {
  case 0:
    *s = ks;
    *c = kc;
    break;
  case 1:
    *s = kc;
    *c = -ks;
    break;
  case 2:
    *s = -ks;
    *c = -kc;
    break;
  default:
    *s = -kc;
    *c = ks;
    break;
  }
}

TFK_HD double tfk_asin_pq(double z, double *q) {
  *q = TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_AQ4, TFK_AQ3), TFK_AQ2), TFK_AQ1), 1.0);
  return z * TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_AP5, TFK_AP4), TFK_AP3), TFK_AP2), TFK_AP1), TFK_AP0);
}

TFK_HD_CALL double tfk_asin(double x) {
  double ax = fabs(x);
  if (!(ax < 1.0)) {
    if (ax == 1.0) return x * TFK_PIO2_HI + x * TFK_PIO2_LO;
    return (x - x) / (x - x); /* |x| > 1 or nan */
  }
  double p, q;
  if (ax < 0.5) {
    if (ax < 7.450580596923828125e-9) return x;
    p = tfk_asin_pq(x * x, &q);
    return TFK_FMA(x, p / q, x);
  }
  double w = 1.0 - ax;
  double t = w * 0.5;
  p = tfk_asin_pq(t, &q);
  double s = sqrt(t);
  if (ax >= 0.975) {
    w = p / q;
    t = TFK_PIO2_HI - TFK_FMA(2.0, TFK_FMA(s, w, s), -TFK_PIO2_LO);
  } else {
    /* s = w + c with w the top half of s: keeps pi/4 - 2w exact */
    w = tfk_from_hi32(tfk_hi32(s));
    double c = TFK_FMA(-w, w, t) / (s + w);
    double r = p / q;
    p = TFK_FMA(2.0 * s, r, -TFK_FMA(-2.0, c, TFK_PIO2_LO));
    q = TFK_FMA(-2.0, w, TFK_PIO4_HI);
    t = TFK_PIO4_HI - (p - q);
  }
  return (x > 0.0) ? t : -t;
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
  double s = TFK_AT10;
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

TFK_HD_CALL double tfk_atan2(double y, double x) {
  if (x != x || y != y) return x + y;
  if (x == 1.0) return tfk_atan(y);
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
  int k = ((tfk_hi32(ay) >> 20) & 0x7ff) - ((tfk_hi32(ax) >> 20) & 0x7ff);
  double z;
  if (k > 60) {
    z = TFK_PIO2_HI + 0.5 * TFK_PI_LO;
    m &= 1;
  } else if ((m & 2) && k < -60) {
    z = 0.0;
  } else {
    z = tfk_atan(fabs(y / x));
  }
  switch (m) {
  case 0:
    return z;
  case 1:
    return -z;
  case 2:
    return TFK_PI - (z - TFK_PI_LO);
  default:
    return (z - TFK_PI_LO) - TFK_PI;
  }
}

#endif /* TFK_FKMATH_H */
