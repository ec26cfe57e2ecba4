/*
 * strict_libm.c -- TEST INFRASTRUCTURE (oracle).  Pins libm for the "strict" reference build.
 *
 * Linked (with -Wl,-Bsymbolic) into oracle/_ref/libtinyfk_ref_strict.so and
 * oracle/build/libfk_oracle_strict.so so that every sin/cos/sincos/asin/atan2 call made by the
 * UNMODIFIED reference sources (urdf_model/pose.h:148-156,181-189,210-213; joint.h:219-236;
 * src/kinematics.cpp:18-21) resolves to the portable implementation in
 * tinyfk_b200/csrc/fkmath.h -- the same code the sm_100a kernels run.  sqrt stays the IEEE
 * correctly rounded instruction on both sides.
 * Also exports array helpers used by tests/test_fkmath.py.
 */
#include "../tinyfk_b200/csrc/fkmath.h"

double sin(double x) { return tfk_sin(x); }
double cos(double x) { return tfk_cos(x); }
void sincos(double x, double *s, double *c) { tfk_sincos(x, s, c); }
double asin(double x) { return tfk_asin(x); }
double atan2(double y, double x) { return tfk_atan2(y, x); }
double atan(double x) { return tfk_atan(x); }

/* ---- test helpers -------------------------------------------------------- */
void tfkm_sin_v(const double *x, double *o, long n) { for (long i = 0; i < n; ++i) o[i] = tfk_sin(x[i]); }
void tfkm_cos_v(const double *x, double *o, long n) { for (long i = 0; i < n; ++i) o[i] = tfk_cos(x[i]); }
void tfkm_asin_v(const double *x, double *o, long n) { for (long i = 0; i < n; ++i) o[i] = tfk_asin(x[i]); }
void tfkm_atan2_v(const double *y, const double *x, double *o, long n) { for (long i = 0; i < n; ++i) o[i] = tfk_atan2(y[i], x[i]); }
void tfkm_sincos_v(const double *x, double *s, double *c, long n) { for (long i = 0; i < n; ++i) tfk_sincos(x[i], s + i, c + i); }

/* SIMT-restructured variants (tinyfk_b200/csrc/fkmath_simt.h): must be bit-identical to the above */
#include "../tinyfk_b200/csrc/fkmath_simt.h"
void tfkm_sincos_simt_v(const double *x, double *s, double *c, long n) { for (long i = 0; i < n; ++i) { TfkSinCos r = tfk_sincos_simt(x[i]); s[i] = r.s; c[i] = r.c; } }
void tfkm_sincos2_simt_v(const double *xa, const double *xb, double *sa, double *ca, double *sb, double *cb, long n) { for (long i = 0; i < n; ++i) { TfkSinCos2 r = tfk_sincos2_simt(xa[i], xb[i]); sa[i] = r.a.s; ca[i] = r.a.c; sb[i] = r.b.s; cb[i] = r.b.c; } }
void tfkm_asin_simt_v(const double *x, double *o, long n) { for (long i = 0; i < n; ++i) o[i] = tfk_asin_simt(x[i]); }
void tfkm_atan2_simt_v(const double *y, const double *x, double *o, long n) { for (long i = 0; i < n; ++i) o[i] = tfk_atan2_simt(y[i], x[i]); }
