/*
 * fkmath_simt.h -- the SAME functions as fkmath.h (tfk_sincos, tfk_asin, tfk_atan2), restructured for
 * a SIMT machine.  Every result is BIT-IDENTICAL to the fkmath.h function of the same name
 * (tests/test_fkmath.py::test_simt_variants_are_bit_identical compares them on millions of arguments;
 * the GPU parity tests compare the kernels, which use these, with the reference linked against
 * fkmath.h).  What changes is only the control flow:
 *
 *   - data-dependent branches that split a warp (|x| <= pi/4 or not, the quadrant switch, the three
 *     cosine kernels, asin below/above 0.5, the five atan intervals) become selects around ONE copy
 *     of the arithmetic, so a warp executes the polynomial and the division once instead of once per
 *     taken path;
 *   - rare inputs (huge arguments, NaN/Inf, exact zeros, |y/x| beyond 2^60) are detected with integer
 *     tests on the exponent word and handed to the fkmath.h function itself;
 *   - FP64 compares against thresholds whose low word is zero are done on the high word (the FP64
 *     pipe is the bottleneck of the strict kernels; integer compares run on the ALU pipe);
 *   - on the device the polynomial coefficients come from constant memory (operands of the FP64
 *     instructions) instead of being rebuilt from immediates at every use.
 *
 * Why identical: each value is produced by the same sequence of IEEE operations on the same
 * operands; where two code paths of fkmath.h are merged the merged expression reduces to each of
 * them exactly (x - 0.0 = x, 1.0 - 0.0 = 1.0, x / 1.0 = x, 1.0 * x = x, 0.0 * x + c = c for finite x,
 * -(a - b) = b - a), as noted at each merge.
 */
#ifndef TFK_FKMATH_SIMT_H
#define TFK_FKMATH_SIMT_H

#include "fkmath.h"

#if defined(__CUDACC__)
/* constant-bank copies of the coefficients (same literals, so same bits) */
struct TfkKc {
  double s1, s2, s3, s4, s5, s6;
  double c1, c2, c3, c4, c5, c6;
  double ap0, ap1, ap2, ap3, ap4, ap5, aq1, aq2, aq3, aq4;
  double at[11];
  double invpio2, pio2_1, pio2_1t;
  /* atan intervals: t = (na*ax + nb) / (da*ax + db), result hi - ((t*s - lo) - t) */
  double iv[5][6];
};
static __constant__ TfkKc tfk_kc = {
    TFK_S1, TFK_S2, TFK_S3, TFK_S4, TFK_S5, TFK_S6,
    TFK_C1, TFK_C2, TFK_C3, TFK_C4, TFK_C5, TFK_C6,
    TFK_AP0, TFK_AP1, TFK_AP2, TFK_AP3, TFK_AP4, TFK_AP5, TFK_AQ1, TFK_AQ2, TFK_AQ3, TFK_AQ4,
    {TFK_AT0, TFK_AT1, TFK_AT2, TFK_AT3, TFK_AT4, TFK_AT5, TFK_AT6, TFK_AT7, TFK_AT8, TFK_AT9, TFK_AT10},
    TFK_INVPIO2, TFK_PIO2_1, TFK_PIO2_1T,
    {{1.0, 0.0, 0.0, 1.0, 0.0, 0.0},
     {2.0, -1.0, 1.0, 2.0, TFK_ATANHI0, TFK_ATANLO0},
     {1.0, -1.0, 1.0, 1.0, TFK_ATANHI1, TFK_ATANLO1},
     {1.0, -1.5, 1.5, 1.0, TFK_ATANHI2, TFK_ATANLO2},
     {0.0, -1.0, 1.0, 0.0, TFK_ATANHI3, TFK_ATANLO3}}};
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

/* tfk_ksin / tfk_kcos with the early exits turned into final selects */
TFK_HD double tfk_ksin_simt(double x, double y) {
  double z = x * x;
  double v = z * x;
  double r = TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_KC(s6, TFK_S6), TFK_KC(s5, TFK_S5)), TFK_KC(s4, TFK_S4)), TFK_KC(s3, TFK_S3)),
                     TFK_KC(s2, TFK_S2));
  double p = x - TFK_FMA(-v, TFK_KC(s1, TFK_S1), TFK_FMA(z, TFK_FMA(-v, r, 0.5 * y), -y));
  /* |x| < 2^-27 <=> high word (sign cleared) < 0x3e400000 (the threshold's low word is zero) */
  return ((tfk_hi32(x) & 0x7fffffff) < 0x3e400000) ? x : p;
}

TFK_HD double tfk_kcos_simt(double x, double y) {
  const int32_t hx = tfk_hi32(x) & 0x7fffffff;
  double z = x * x;
  double r = z * TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_KC(c6, TFK_C6), TFK_KC(c5, TFK_C5)), TFK_KC(c4, TFK_C4)), TFK_KC(c3, TFK_C3)),
                                 TFK_KC(c2, TFK_C2)),
                         TFK_KC(c1, TFK_C1));
  /* ax < 0.3: qx = 0 gives hz = fma(0.5, z, -0) = 0.5*z and a = 1 - 0 = 1, i.e. fkmath.h's first form.
     (0.3 has a non-zero low word: keep the FP64 compare.)  0.78125 = 0x3fe90000_00000000. */
  const uint64_t ux = tfk_abs_bits(x);
  double qx = tfk_from_hi32(hx - 0x00200000);
  qx = (ux > 0x3fe9000000000000ull) ? 0.28125 : qx; /* ax > 0.78125 */
  qx = (ux < 0x3fd3333333333333ull) ? 0.0 : qx;     /* ax < 0.3 = 0x3fd33333_33333333 */
  double hz = TFK_FMA(0.5, z, -qx);
  double a = 1.0 - qx;
  /* fkmath.h returns 1.0 below 2^-27: there qx = 0 and |0.5*z - (z*r - x*y)| < 2^-55, a quarter of the spacing below 1.0, so
     the expression itself rounds to 1.0 -- no select needed */
  return a - (hz - TFK_FMA(z, r, -(x * y)));
}

typedef struct {
  double s, c;
} TfkSinCos;

/* sin and cos of x; bit-identical to tfk_sincos (fkmath.h) */
TFK_HD TfkSinCos tfk_sincos_simt(double x) {
  TfkSinCos out;
  const int32_t hx = tfk_hi32(x) & 0x7fffffff;
  if (TFK_UNLIKELY(hx >= 0x41392000)) { /* |x| >= ~1.64e6 (1647099.0 = 0x413921fb_00000000), inf, nan: rare */
    tfk_sincos(x, &out.s, &out.c);
    return out;
  }
  const double t = fabs(x);
  const int small = tfk_abs_bits(x) <= 0x3fe921fb54442d18ull; /* |x| <= TFK_PIO4_HI */
  /* first reduction stage for every lane (harmless when small) */
  int n = (int)TFK_FMA(t, TFK_KC(invpio2, TFK_INVPIO2), 0.5);
  const double fn = (double)n;
  double r = TFK_FMA(-fn, TFK_KC(pio2_1, TFK_PIO2_1), t);
  double w = fn * TFK_KC(pio2_1t, TFK_PIO2_1T);
  double a0 = r - w;
  if (!small) {
    const int j = (hx >> 20) & 0x7ff;
    int i = j - ((tfk_hi32(a0) >> 20) & 0x7ff);
    if (TFK_UNLIKELY(i > 16)) { /* cancellation near a multiple of pi/2: rare */
      double t2 = r;
      w = fn * TFK_PIO2_2;
      r = t2 - w;
      w = fn * TFK_PIO2_2T - ((t2 - r) - w);
      a0 = r - w;
      i = j - ((tfk_hi32(a0) >> 20) & 0x7ff);
      if (i > 49) {
        t2 = r;
        w = fn * TFK_PIO2_3;
        r = t2 - w;
        w = fn * TFK_PIO2_3T - ((t2 - r) - w);
        a0 = r - w;
      }
    }
  }
  double a1 = (r - a0) - w;
  if (small) {
    n = 0;
    a0 = t;
    a1 = 0.0;
  }
  /* fkmath.h: small -> kernels on (x, 0.0); else on (+-a0, +-a1) with n negated for x < 0 */
  const int32_t sx = tfk_hi32(x); /* bit 31 = sign of x */
  const double y0 = tfk_xor_sign(a0, sx);
  const double y1 = tfk_xor_sign(a1, small ? 0 : sx);
  if (sx < 0) n = -n;
  const double ks = tfk_ksin_simt(y0, y1), kc = tfk_kcos_simt(y0, y1);
  /* n&3: 0 (ks, kc)  1 (kc, -ks)  2 (-ks, -kc)  3 (-kc, ks) */
  double ss = (n & 1) ? kc : ks;
  double cc = (n & 1) ? ks : kc;
  ss = tfk_xor_sign(ss, (int32_t)((uint32_t)n << 30));       /* n & 2 */
  cc = tfk_xor_sign(cc, (int32_t)((uint32_t)(n + 1) << 30)); /* (n + 1) & 2 */
  out.s = ss;
  out.c = cc;
  return out;
}

/* sin and cos of TWO arguments at once: the two evaluations are independent dependency chains, and written as one
   straight-line body (selects instead of the `small` branch, one shared test for the rare inputs) the compiler
   interleaves them -- twice the instruction-level parallelism for a thread whose next step needs both (the walk
   down a kinematic chain knows every joint angle up front).  Same operations per argument as tfk_sincos_simt:
   bit-identical results. */
typedef struct {
  TfkSinCos a, b;
} TfkSinCos2;

TFK_HD TfkSinCos2 tfk_sincos2_simt(double xa, double xb) {
  TfkSinCos2 out;
  const int32_t hxa = tfk_hi32(xa) & 0x7fffffff, hxb = tfk_hi32(xb) & 0x7fffffff;
  if (TFK_UNLIKELY((hxa >= 0x41392000) | (hxb >= 0x41392000))) { /* |x| >= ~1.64e6, inf, nan: rare */
    out.a = tfk_sincos_simt(xa);
    out.b = tfk_sincos_simt(xb);
    return out;
  }
  const double ta = fabs(xa), tb = fabs(xb);
  const int smalla = tfk_abs_bits(xa) <= 0x3fe921fb54442d18ull, smallb = tfk_abs_bits(xb) <= 0x3fe921fb54442d18ull;
  int na = (int)TFK_FMA(ta, TFK_KC(invpio2, TFK_INVPIO2), 0.5), nb = (int)TFK_FMA(tb, TFK_KC(invpio2, TFK_INVPIO2), 0.5);
  const double fna = (double)na, fnb = (double)nb;
  const double ra = TFK_FMA(-fna, TFK_KC(pio2_1, TFK_PIO2_1), ta), rb = TFK_FMA(-fnb, TFK_KC(pio2_1, TFK_PIO2_1), tb);
  const double wa = fna * TFK_KC(pio2_1t, TFK_PIO2_1T), wb = fnb * TFK_KC(pio2_1t, TFK_PIO2_1T);
  double a0a = ra - wa, a0b = rb - wb;
  {
    const int ia = ((hxa >> 20) & 0x7ff) - ((tfk_hi32(a0a) >> 20) & 0x7ff);
    const int ib = ((hxb >> 20) & 0x7ff) - ((tfk_hi32(a0b) >> 20) & 0x7ff);
    if (TFK_UNLIKELY((!smalla & (ia > 16)) | (!smallb & (ib > 16)))) { /* cancellation near a multiple of pi/2: rare */
      out.a = tfk_sincos_simt(xa);
      out.b = tfk_sincos_simt(xb);
      return out;
    }
  }
  double a1a = (ra - a0a) - wa, a1b = (rb - a0b) - wb;
  if (smalla) {
    na = 0;
    a0a = ta;
    a1a = 0.0;
  }
  if (smallb) {
    nb = 0;
    a0b = tb;
    a1b = 0.0;
  }
  const int32_t sxa = tfk_hi32(xa), sxb = tfk_hi32(xb);
  const double y0a = tfk_xor_sign(a0a, sxa), y0b = tfk_xor_sign(a0b, sxb);
  const double y1a = tfk_xor_sign(a1a, smalla ? 0 : sxa), y1b = tfk_xor_sign(a1b, smallb ? 0 : sxb);
  if (sxa < 0) na = -na;
  if (sxb < 0) nb = -nb;
  const double ksa = tfk_ksin_simt(y0a, y1a), kca = tfk_kcos_simt(y0a, y1a);
  const double ksb = tfk_ksin_simt(y0b, y1b), kcb = tfk_kcos_simt(y0b, y1b);
  double ssa = (na & 1) ? kca : ksa, cca = (na & 1) ? ksa : kca;
  double ssb = (nb & 1) ? kcb : ksb, ccb = (nb & 1) ? ksb : kcb;
  out.a.s = tfk_xor_sign(ssa, (int32_t)((uint32_t)na << 30));
  out.a.c = tfk_xor_sign(cca, (int32_t)((uint32_t)(na + 1) << 30));
  out.b.s = tfk_xor_sign(ssb, (int32_t)((uint32_t)nb << 30));
  out.b.c = tfk_xor_sign(ccb, (int32_t)((uint32_t)(nb + 1) << 30));
  return out;
}

TFK_HD double tfk_asin_pq_simt(double z, double *q) {
  *q = TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_KC(aq4, TFK_AQ4), TFK_KC(aq3, TFK_AQ3)), TFK_KC(aq2, TFK_AQ2)), TFK_KC(aq1, TFK_AQ1)), 1.0);
  return z * TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_FMA(z, TFK_KC(ap5, TFK_AP5), TFK_KC(ap4, TFK_AP4)), TFK_KC(ap3, TFK_AP3)), TFK_KC(ap2, TFK_AP2)),
                                TFK_KC(ap1, TFK_AP1)),
                     TFK_KC(ap0, TFK_AP0));
}

/* bit-identical to tfk_asin (fkmath.h): one rational kernel + one division for both ranges */
TFK_HD double tfk_asin_simt(double x) {
  const int32_t hx = tfk_hi32(x) & 0x7fffffff;
  if (TFK_UNLIKELY(hx >= 0x3ff00000)) return tfk_asin(x); /* |x| >= 1, nan: rare */
  if (TFK_UNLIKELY(hx < 0x3e400000)) return x; /* |x| < 2^-27 */
  const double ax = fabs(x);
  const int big = hx >= 0x3fe00000; /* |x| >= 0.5 */
  const double t = (1.0 - ax) * 0.5;
  double q;
  const double p = tfk_asin_pq_simt(big ? t : x * x, &q);
  const double rr = p / q;
  if (!big) return TFK_FMA(x, rr, x);
  const double s = sqrt(t);
  double res;
  if (tfk_abs_bits(x) >= 0x3fef333333333333ull) { /* ax >= 0.975 */
    res = TFK_PIO2_HI - TFK_FMA(2.0, TFK_FMA(s, rr, s), -TFK_PIO2_LO);
  } else {
    const double w = tfk_from_hi32(tfk_hi32(s));
    const double c = TFK_FMA(-w, w, t) / (s + w);
    const double pp = TFK_FMA(2.0 * s, rr, -TFK_FMA(-2.0, c, TFK_PIO2_LO));
    const double qq = TFK_FMA(-2.0, w, TFK_PIO4_HI);
    res = TFK_PIO4_HI - (pp - qq);
  }
  return tfk_xor_sign(res, tfk_hi32(x)); /* x > 0 ? res : -res (x is neither 0 nor NaN here) */
}

/* atan of ax >= 2^-29, finite, < 2^66; returns atan(ax) >= 0.  Bit-identical to tfk_atan(ax):
   interval id from the high word (all four thresholds have a zero low word), its six constants from a
   table, ONE division:  t = (na*ax + nb) / (da*ax + db)
     id -1: (1*ax + 0) / (0*ax + 1) = ax / 1 = ax            fkmath.h: t = x
     id  0: (2*ax - 1) / (1*ax + 2)                          fma(2.0, ax, -1.0) / (ax + 2.0)
     id  1: (1*ax - 1) / (1*ax + 1)                          (ax - 1.0) / (ax + 1.0)
     id  2: (1*ax - 1.5) / (1.5*ax + 1)                      (ax - 1.5) / fma(1.5, ax, 1.0)
     id  3: (0*ax - 1) / (1*ax + 0) = -1 / ax                -1.0 / ax
   (numerator and denominator are ONE fused operation each: exact products with 0, 1 and 2 make them equal to the
   plain forms) and hi - (fma(t, s, -lo) - t) with hi = lo = 0 for id -1 equals -(t*s - t) = t - t*s. */
TFK_HD double tfk_atan_pos_simt(double ax) {
  const int32_t hx = tfk_hi32(ax);
  const int id = (hx >= 0x3fdc0000) + (hx >= 0x3fe60000) + (hx >= 0x3ff30000) + (hx >= 0x40038000);
#if defined(__CUDA_ARCH__)
  const double *iv = tfk_kc.iv[id];
  const double na = iv[0], nb = iv[1], da = iv[2], db = iv[3], hi = iv[4], lo = iv[5];
#else
  static const double tab[5][6] = {{1.0, 0.0, 0.0, 1.0, 0.0, 0.0},
                                   {2.0, -1.0, 1.0, 2.0, TFK_ATANHI0, TFK_ATANLO0},
                                   {1.0, -1.0, 1.0, 1.0, TFK_ATANHI1, TFK_ATANLO1},
                                   {1.0, -1.5, 1.5, 1.0, TFK_ATANHI2, TFK_ATANLO2},
                                   {0.0, -1.0, 1.0, 0.0, TFK_ATANHI3, TFK_ATANLO3}};
  const double na = tab[id][0], nb = tab[id][1], da = tab[id][2], db = tab[id][3], hi = tab[id][4], lo = tab[id][5];
#endif
  const double t = TFK_FMA(na, ax, nb) / TFK_FMA(da, ax, db);
  const double z = t * t;
  double s = TFK_KC(at[10], TFK_AT10);
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
  return hi - (TFK_FMA(t, s, -lo) - t);
}

/* bit-identical to tfk_atan2 (fkmath.h) */
TFK_HD double tfk_atan2_simt(double y, double x) {
  const int32_t hxs = tfk_hi32(x), hys = tfk_hi32(y);
  const int32_t hx = hxs & 0x7fffffff, hy = hys & 0x7fffffff;
  const int k = (hy >> 20) - (hx >> 20);
  /* rare: nan/inf, x == 1.0, y == 0, x == 0, exponents more than 60 apart, |y/x| outside atan's
     plain range -- the fkmath.h function handles them */
  const int special = (hx >= 0x7ff00000) | (hy >= 0x7ff00000) | ((hy | tfk_lo32(y)) == 0) | ((hx | tfk_lo32(x)) == 0) |
                      ((hxs == 0x3ff00000) & (tfk_lo32(x) == 0)) | (k > 60) | (k < -60);
  if (TFK_UNLIKELY(special)) return tfk_atan2(y, x);
  const double a = fabs(y / x);
  const int32_t ha = tfk_hi32(a);
  double z = (ha < 0x3e200000) ? a : tfk_atan_pos_simt(a); /* |a| < 2^-29: atan(a) = a */
  /* m = signbit(y) | signbit(x) << 1:  0: z   1: -z   2: pi - (z - pi_lo)   3: (z - pi_lo) - pi */
  if (hxs < 0) z = TFK_PI - (z - TFK_PI_LO);
  return tfk_xor_sign(z, hys);
}

#endif /* TFK_FKMATH_SIMT_H */
