/*
 * fkmath_simt.h -- the SAME functions as fkmath.h (tfk_sincos, tfk_asin, tfk_atan2) as the kernels call them.
 * Every result is BIT-IDENTICAL to the fkmath.h function of the same name
 * (tests/test_fkmath.py::test_simt_variants_are_bit_identical compares them on millions of arguments;
 * the GPU parity tests compare the kernels, which use these, with the reference linked against
 * fkmath.h).  fkmath.h's evaluations are straight-line code already (one reduction for every sin/cos argument, one
 * division per atan2, no division in asin); what differs here:
 *
 *   - rare inputs (huge arguments, NaN/Inf, magnitudes outside atan2's working range, |x| >= 1 for asin) are detected
 *     with integer tests on the exponent word and handed to the fkmath.h function itself;
 *   - asin evaluates its polynomial once for both ranges (the argument is selected);
 *   - sign changes are one XOR on the high word;
 *   - on the device the polynomial coefficients and the k * pi/4 offsets come from constant memory (operands of the
 *     FP64 instructions) instead of being rebuilt from immediates at every use.
 *
 * Why identical: each value is produced by the same sequence of IEEE operations on the same operands.
 */
#ifndef TFK_FKMATH_SIMT_H
#define TFK_FKMATH_SIMT_H

#include "fkmath.h"

/* the one quotient of atan2.  Strict mode: the IEEE division (part of the function's definition).  The FP64 fast mode
   (fk_math.h:MathFast64, 1e-12 contract) defines it as a reciprocal-multiply before including this header. */
#ifndef TFK_QUOT
#define TFK_QUOT(a, b) ((a) / (b))
#endif

#if defined(__CUDACC__)
/* constant-bank copies of the coefficients (same literals, so same bits) */
struct TfkKc {
  double s1, s2, s3, s4, s5, s6;
  double c1, c2, c3, c4, c5, c6;
  double ar[14];
  double at[12];
  double invpio2, pio2_1, pio2_1t, pio2_1tt;
  double k_hi[5], k_lo[5]; /* k * pi/4 as hi + lo: the atan2 offsets */
};
static __constant__ TfkKc tfk_kc = {
    TFK_S1, TFK_S2, TFK_S3, TFK_S4, TFK_S5, TFK_S6,
    TFK_C1, TFK_C2, TFK_C3, TFK_C4, TFK_C5, TFK_C6,
    {TFK_AR0, TFK_AR1, TFK_AR2, TFK_AR3, TFK_AR4, TFK_AR5, TFK_AR6, TFK_AR7, TFK_AR8, TFK_AR9, TFK_AR10, TFK_AR11, TFK_AR12, TFK_AR13},
    {TFK_AT0, TFK_AT1, TFK_AT2, TFK_AT3, TFK_AT4, TFK_AT5, TFK_AT6, TFK_AT7, TFK_AT8, TFK_AT9, TFK_AT10, TFK_AT11},
    TFK_INVPIO2, TFK_PIO2_1, TFK_PIO2_1T, TFK_PIO2_1TT,
    {0.0, TFK_PIO4_HI, TFK_PIO2_HI, TFK_3PIO4_HI, TFK_PI},
    {0.0, TFK_PIO4_LO, TFK_PIO2_LO, TFK_3PIO4_LO, TFK_PI_LO}};
#endif

#if defined(__CUDA_ARCH__)
#define TFK_KC(field, lit) (tfk_kc.field)
#else
#define TFK_KC(field, lit) (lit)
#endif

TFK_HD int32_t tfk_lo32(double x) {
#if defined(__CUDA_ARCH__)
  return __double2loint(x);
#else
  uint64_t u;
  memcpy(&u, &x, sizeof u);
  return (int32_t)(uint32_t)u;
#endif
}

/* x with its sign bit XORed with bit 31 of `mask` (a conditional negation as ONE integer instruction on
   the high word; -x is exact, so this equals `cond ? -x : x`) */
TFK_HD double tfk_xor_sign(double x, int32_t mask) {
#if defined(__CUDA_ARCH__)
  return __hiloint2double(__double2hiint(x) ^ (mask & (int32_t)0x80000000), __double2loint(x));
#else
  uint64_t u;
  memcpy(&u, &x, sizeof u);
  u ^= ((uint64_t)((uint32_t)mask & 0x80000000u)) << 32;
  memcpy(&x, &u, sizeof x);
  return x;
#endif
}
/* bits of |x| as an unsigned integer: for non-NaN doubles a < b  <=>  bits(|a|) < bits(|b|), so
   threshold tests run on the integer pipe instead of the FP64 pipe */
TFK_HD uint64_t tfk_abs_bits(double x) {
  return (((uint64_t)((uint32_t)tfk_hi32(x) & 0x7fffffffu)) << 32) | (uint32_t)tfk_lo32(x);
}

/* tfk_ksin / tfk_kcos / tfk_reduce of fkmath.h, operation for operation, with the coefficients as constant-bank operands */
TFK_HD double tfk_ksin_simt(double x) {
  const double z = x * x;
  const double v = z * x;
  const double r = TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_KC(s6, TFK_S6), TFK_KC(s5, TFK_S5)), TFK_KC(s4, TFK_S4)), TFK_KC(s3, TFK_S3)),
                           TFK_KC(s2, TFK_S2));
  return TFK_FMA(v, TFK_FMA(z, r, TFK_KC(s1, TFK_S1)), x);
}

TFK_HD double tfk_kcos_simt(double x) {
  const double z = x * x;
  const double q = TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_KC(c6, TFK_C6), TFK_KC(c5, TFK_C5)), TFK_KC(c4, TFK_C4)), TFK_KC(c3, TFK_C3)),
                                           TFK_KC(c2, TFK_C2)),
                                   TFK_KC(c1, TFK_C1)),
                           -0.5);
  return TFK_FMA(z, q, 1.0);
}

TFK_HD double tfk_reduce_simt(double t, int32_t *n) {
  const double m = TFK_FMA(t, TFK_KC(invpio2, TFK_INVPIO2), TFK_RND_MAGIC);
  const double fn = m - TFK_RND_MAGIC;
  *n = tfk_lo32(m);
  const double r = TFK_FMA(-fn, TFK_KC(pio2_1, TFK_PIO2_1), t);
  const double a = TFK_FMA(-fn, TFK_KC(pio2_1t, TFK_PIO2_1T), r);
  return TFK_FMA(-fn, TFK_KC(pio2_1tt, TFK_PIO2_1TT), a);
}

typedef struct {
  double s, c;
} TfkSinCos;

/* sin and cos of x; bit-identical to tfk_sincos (fkmath.h) */
TFK_HD TfkSinCos tfk_sincos_simt(double x) {
  TfkSinCos out;
  const int32_t hx = tfk_hi32(x);
  if (TFK_UNLIKELY((hx & 0x7fffffff) >= 0x41392000)) { /* |x| >= ~1.64e6 (1647099.0 = 0x413921fb_00000000), inf, nan: rare */
    tfk_sincos(x, &out.s, &out.c);
    return out;
  }
  int32_t n;
  const double a0 = tfk_reduce_simt(fabs(x), &n);
  tfk_sincos_place(tfk_ksin_simt(a0), tfk_kcos_simt(a0), n, hx, &out.s, &out.c);
  return out;
}

/* sin and cos of TWO arguments at once: the two evaluations are independent dependency chains, and written as one
   straight-line body with one shared test for the rare inputs the compiler interleaves them -- twice the
   instruction-level parallelism for a thread whose next step needs both (the walk down a kinematic chain knows every
   joint angle up front).  Same operations per argument as tfk_sincos_simt: bit-identical results. */
typedef struct {
  TfkSinCos a, b;
} TfkSinCos2;

TFK_HD TfkSinCos2 tfk_sincos2_simt(double xa, double xb) {
  TfkSinCos2 out;
  const int32_t hxa = tfk_hi32(xa), hxb = tfk_hi32(xb);
  if (TFK_UNLIKELY(((hxa & 0x7fffffff) >= 0x41392000) | ((hxb & 0x7fffffff) >= 0x41392000))) { /* |x| >= ~1.64e6, inf, nan: rare */
    out.a = tfk_sincos_simt(xa);
    out.b = tfk_sincos_simt(xb);
    return out;
  }
  int32_t na, nb;
  const double a0a = tfk_reduce_simt(fabs(xa), &na), a0b = tfk_reduce_simt(fabs(xb), &nb);
  const double ksa = tfk_ksin_simt(a0a), kca = tfk_kcos_simt(a0a);
  const double ksb = tfk_ksin_simt(a0b), kcb = tfk_kcos_simt(a0b);
  tfk_sincos_place(ksa, kca, na, hxa, &out.a.s, &out.a.c);
  tfk_sincos_place(ksb, kcb, nb, hxb, &out.b.s, &out.b.c);
  return out;
}

/* bit-identical to tfk_asin (fkmath.h): the same operations with the coefficients as constant-bank operands; one
   evaluation of the polynomial serves both ranges (its argument is selected) */
TFK_HD double tfk_asin_simt(double x) {
  const int32_t hx = tfk_hi32(x) & 0x7fffffff;
  if (TFK_UNLIKELY(hx >= 0x3ff00000)) return tfk_asin(x); /* |x| >= 1, nan: rare */
  if (TFK_UNLIKELY(hx < 0x3e400000)) return x; /* |x| < 2^-27 */
  const double ax = fabs(x);
  const int big = hx >= 0x3fe00000; /* |x| >= 0.5 */
  const double t = TFK_FMA(-0.5, ax, 0.5);
  const double z = big ? t : x * x;
  double r = TFK_KC(ar[13], TFK_AR13);
  r = TFK_FMA(z, r, TFK_KC(ar[12], TFK_AR12));
  r = TFK_FMA(z, r, TFK_KC(ar[11], TFK_AR11));
  r = TFK_FMA(z, r, TFK_KC(ar[10], TFK_AR10));
  r = TFK_FMA(z, r, TFK_KC(ar[9], TFK_AR9));
  r = TFK_FMA(z, r, TFK_KC(ar[8], TFK_AR8));
  r = TFK_FMA(z, r, TFK_KC(ar[7], TFK_AR7));
  r = TFK_FMA(z, r, TFK_KC(ar[6], TFK_AR6));
  r = TFK_FMA(z, r, TFK_KC(ar[5], TFK_AR5));
  r = TFK_FMA(z, r, TFK_KC(ar[4], TFK_AR4));
  r = TFK_FMA(z, r, TFK_KC(ar[3], TFK_AR3));
  r = TFK_FMA(z, r, TFK_KC(ar[2], TFK_AR2));
  r = TFK_FMA(z, r, TFK_KC(ar[1], TFK_AR1));
  r = TFK_FMA(z, r, TFK_KC(ar[0], TFK_AR0));
  const double zr = z * r;
  if (!big) return TFK_FMA(x, zr, x);
  const double res = tfk_asin_upper(ax, zr, sqrt(t), t);
  return tfk_xor_sign(res, tfk_hi32(x)); /* x > 0 ? res : -res (x is neither 0 nor NaN here) */
}

/* bit-identical to tfk_atan2 (fkmath.h): the same core (tfk_atan2_core, operation for operation) with the atan
   coefficients and the k * pi/4 offsets as constant-bank operands; everything the core does not take -- NaN / Inf, a
   larger magnitude outside [2^-900, 2^1022) (which includes x = y = 0) -- goes to the fkmath.h function. */
TFK_HD double tfk_atan2_simt(double y, double x) {
  const int32_t hxs = tfk_hi32(x), hys = tfk_hi32(y);
  const uint32_t hx = (uint32_t)hxs & 0x7fffffffu, hy = (uint32_t)hys & 0x7fffffffu;
  const uint32_t hm = hx > hy ? hx : hy;
  if (TFK_UNLIKELY(hm - 0x07b00000u >= 0x7fd00000u - 0x07b00000u)) return tfk_atan2(y, x);
  const double ax = fabs(x), ay = fabs(y);
  const int swap = ay > ax;
  const double mx = swap ? ay : ax, mn = swap ? ax : ay;
  const int big = mn >= 0.5 * mx;
  const double num = big ? mn - mx : mn;
  const double den = big ? mn + mx : mx;
  const double t = TFK_QUOT(num, den);
  const double z = t * t;
  double s = TFK_KC(at[11], TFK_AT11);
  s = TFK_FMA(z, s, TFK_KC(at[10], TFK_AT10));
  s = TFK_FMA(z, s, TFK_KC(at[9], TFK_AT9));
  s = TFK_FMA(z, s, TFK_KC(at[8], TFK_AT8));
  s = TFK_FMA(z, s, TFK_KC(at[7], TFK_AT7));
  s = TFK_FMA(z, s, TFK_KC(at[6], TFK_AT6));
  s = TFK_FMA(z, s, TFK_KC(at[5], TFK_AT5));
  s = TFK_FMA(z, s, TFK_KC(at[4], TFK_AT4));
  s = TFK_FMA(z, s, TFK_KC(at[3], TFK_AT3));
  s = TFK_FMA(z, s, TFK_KC(at[2], TFK_AT2));
  s = TFK_FMA(z, s, TFK_KC(at[1], TFK_AT1));
  s = TFK_FMA(z, s, TFK_KC(at[0], TFK_AT0));
  s = z * s;
  const int xneg = hxs < 0;
  const int flip = swap ^ xneg;
  const int k = (swap ? 2 : (xneg ? 4 : 0)) + (flip ? -big : big);
  double hi, lo;
#if defined(__CUDA_ARCH__)
  hi = tfk_kc.k_hi[k];
  lo = tfk_kc.k_lo[k];
#else
  TFK_ATAN2_TABLE(k, hi, lo);
#endif
  const double tv = tfk_xor_sign(t, flip ? (int32_t)0x80000000 : 0);
  const double th = hi - (TFK_FMA(tv, s, -lo) - tv);
  return tfk_xor_sign(th, hys);
}

#endif /* TFK_FKMATH_SIMT_H */
