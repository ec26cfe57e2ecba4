// fk_math.h -- pose / quaternion / Jacobian primitives shared by the host model code (g++) and the
// sm_100a kernels (nvcc).  Header only, no CUDA API in here.
//
// Every function states the reference expression it evaluates.  In the FP64 "strict" mode the
// operation ORDER is the reference's (left-to-right as written there) and the translation unit is
// built without FMA contraction (nvcc -fmad=false / g++ -ffp-contract=off), so results are
// bit-identical to the reference built with the same libm (oracle/_ref/libtinyfk_ref_strict.so).
// The only liberty taken: products with a literal 0.0 operand are dropped where the reference
// multiplies a pure-vector quaternion (w == 0) -- that can flip the sign of an exact zero, never a
// non-zero bit (documented in DESIGN.md, "signed zeros").
//
// Reference files: urdf_parser/include/urdf_model/pose.h (Rotation, Pose), urdf_parser/src/pose.cpp
// (pose_transform, cross_product), urdf_parser/include/urdf_model/joint.h:227-253 (Joint::transform),
// src/kinematics.cpp:14-35 (rpy_derivative, q_derivative).
#ifndef TFK_FK_MATH_H
#define TFK_FK_MATH_H

#include "fkmath.h"
#if defined(TFK_FAST_PORTABLE_LIBM) && defined(__CUDA_ARCH__)
// FP64 fast mode only (a translation unit that never holds strict code), device pass: atan2's quotient as one correctly
// rounded reciprocal times the numerator
#define TFK_QUOT(a, b) ((a) * __drcp_rn(b))
#endif
#include "fkmath_simt.h"

#if defined(__CUDACC__)
#define TFK_FN __host__ __device__ __forceinline__
// out-of-line on the device: called from several sites, each holding two or three independent libm
// evaluations that the compiler interleaves (ILP) -- one copy keeps the instruction cache footprint small
#if defined(TFK_SPEC_INLINE)
#define TFK_FN_CALL __host__ __device__ __forceinline__
#else
#define TFK_FN_CALL __host__ __device__ __noinline__
#endif
#define TFK_FN_RARE __host__ __device__ __noinline__
#else
#define TFK_FN inline
#define TFK_FN_CALL inline
#define TFK_FN_RARE inline
#endif
#if defined(__CUDACC__)
#define TFK_MATH_UNROLL _Pragma("unroll")
#else
#define TFK_MATH_UNROLL
#endif

namespace tfk {

template <class T> struct Vec3 {
  T x, y, z;
};
template <class T> struct Quat {
  T x, y, z, w;
};
template <class T> struct PoseT {
  Vec3<T> p;
  Quat<T> q;
};

// ---- math policies --------------------------------------------------------------------------
// Portable libm: identical bits on host and device (fkmath.h).  FP64 only.
struct MathPortable {
  typedef double real;
  // registers to spare for prefetching the next item's 7-real link record in the several-link kernels
  // (fk_kernels.cuh:fk_twophase_body)?  Strict FP64 sits at its register cap.
  enum { SPARE_REGS = 0 };
  static TFK_FN double sin_(double x) { return tfk_sin(x); }
  static TFK_FN double cos_(double x) { return tfk_cos(x); }
  // fkmath_simt.h: bit-identical to fkmath.h's functions, restructured for warps
#if defined(TFK_SPEC_SINCOS_CALL)
  static __host__ __device__ __noinline__ TfkSinCos sincos_call_(double x) { return tfk_sincos_simt(x); }
  static TFK_FN void sincos_(double x, double *s, double *c) {
    const TfkSinCos r = sincos_call_(x);
    *s = r.s;
    *c = r.c;
  }
#else
  static TFK_FN void sincos_(double x, double *s, double *c) {
    const TfkSinCos r = tfk_sincos_simt(x);
    *s = r.s;
    *c = r.c;
  }
#endif
  // two arguments at once (fkmath_simt.h: the two evaluations interleave; same bits as two sincos_ calls)
#if defined(__CUDACC__) // (inlined it costs registers: cfg2 strict 0.265 against 0.198 ms)
  static __host__ __device__ __noinline__ TfkSinCos2 sincos2_call_(double xa, double xb) { return tfk_sincos2_simt(xa, xb); }
#else
  static inline TfkSinCos2 sincos2_call_(double xa, double xb) { return tfk_sincos2_simt(xa, xb); }
#endif
  static TFK_FN void sincos2_(double xa, double xb, double *sa, double *ca, double *sb, double *cb) {
    const TfkSinCos2 r = sincos2_call_(xa, xb);
    *sa = r.a.s, *ca = r.a.c, *sb = r.b.s, *cb = r.b.c;
  }
  static TFK_FN double asin_(double x) { return tfk_asin_simt(x); }
  static TFK_FN double atan2_(double y, double x) { return tfk_atan2_simt(y, x); }
  static TFK_FN double atan2_rare_(double y, double x) { return tfk_atan2(y, x); } // same bits, out of line
  static TFK_FN double sqrt_(double x) { return sqrt(x); }

  // ---- several IEEE divisions by the same denominator ------------------------------------------
  // The reference divides the four components of a quaternion by one norm (pose.h:284-299,
  // 230-250), five rpy_derivative coefficients by one cosine (kinematics.cpp:14-21) and every
  // finite-difference entry by eps (kinematics.cpp:212-231).  den_() takes y = RN(1/b) once
  // (correctly rounded reciprocal: __drcp_rn on the device, 1.0/b on the host); div_() then gets
  // RN(a/b) with two residual corrections (Markstein): q0 = RN(a*y) is within 1.5 ulp of a/b,
  // r0 = a - b*q0 is exact in an FMA, q1 = RN(q0 + r0*y) is a faithful quotient (its unrounded value
  // is a/b*(1 + O(2^-104))), and for a faithful q1 and a correctly rounded y the second step
  // q2 = RN(q1 + (a - b*q1)*y) is the correctly rounded quotient [Markstein 1990; Muller et al.,
  // Handbook of Floating-Point Arithmetic, sec. 4.7].  Exponent ranges that could leave the
  // normal range on the way (and NaN/Inf/0 denominators) use the plain division instead; an exact
  // zero numerator gives a*y = +-0 with the quotient's sign.  tests/test_fkmath.py compares
  // div_() with `/` bit for bit.
  struct Den {
    double b, y;
    bool ok;
  };
  static TFK_FN Den den_(double b) {
    Den d;
    d.b = b;
    const uint32_t t = (uint32_t)tfk_hi32(b) & 0x7fffffffu;
    d.ok = (t - 0x20000000u) < 0x40000000u; // 2^-511 <= |b| < 2^513
#if defined(__CUDA_ARCH__)
    d.y = __drcp_rn(b);
#else
    d.y = 1.0 / b;
#endif
    return d;
  }
  // the compiler's full IEEE division (its slow path is ~100 instructions): one copy, out of line
  static TFK_FN_RARE double div_rare_(double a, double b) { return a / b; }
  static TFK_FN double div_(double a, const Den &d) {
    const uint32_t t = (uint32_t)tfk_hi32(a) & 0x7fffffffu;
    if (TFK_UNLIKELY(!d.ok || (t - 0x27000000u) >= 0x32000000u)) { // |a| outside [2^-399, 2^401): rare, or an exact zero
      if (d.ok && (t | (uint32_t)tfk_lo32(a)) == 0u) return a * d.y;
      return div_rare_(a, d.b);
    }
#if defined(__CUDA_ARCH__)
    double q = __dmul_rn(a, d.y);
    double r = __fma_rn(-d.b, q, a);
    q = __fma_rn(r, d.y, q);
    r = __fma_rn(-d.b, q, a);
    return __fma_rn(r, d.y, q);
#else
    double q = a * d.y;
    double r = fma(-d.b, q, a);
    q = fma(r, d.y, q);
    r = fma(-d.b, q, a);
    return fma(r, d.y, q);
#endif
  }
  // N quotients by one denominator with ONE range test for all numerators (the common case: every
  // numerator is an ordinary number); otherwise each goes through div_ on its own.
  template <int N> static TFK_FN void divn_(double (&a)[N], const Den &d) {
    uint32_t bad = d.ok ? 0u : 1u;
    TFK_MATH_UNROLL
    for (int i = 0; i < N; ++i) bad |= (((uint32_t)tfk_hi32(a[i]) & 0x7fffffffu) - 0x27000000u) >= 0x32000000u ? 1u : 0u;
    if (TFK_UNLIKELY(bad != 0u)) {
      TFK_MATH_UNROLL
      for (int i = 0; i < N; ++i) a[i] = div_(a[i], d);
      return;
    }
    TFK_MATH_UNROLL
    for (int i = 0; i < N; ++i) {
#if defined(__CUDA_ARCH__)
      double q = __dmul_rn(a[i], d.y);
      double r = __fma_rn(-d.b, q, a[i]);
      q = __fma_rn(r, d.y, q);
      r = __fma_rn(-d.b, q, a[i]);
      a[i] = __fma_rn(r, d.y, q);
#else
      double q = a[i] * d.y;
      double r = fma(-d.b, q, a[i]);
      q = fma(r, d.y, q);
      r = fma(-d.b, q, a[i]);
      a[i] = fma(r, d.y, q);
#endif
    }
  }
};

// The platform's own libm (glibc on the host, libdevice on the GPU).
struct MathNative64 {
  typedef double real;
  enum { SPARE_REGS = 0 };
  static TFK_FN double sin_(double x) { return ::sin(x); }
  static TFK_FN double cos_(double x) { return ::cos(x); }
  static TFK_FN void sincos_(double x, double *s, double *c) {
#if defined(__CUDA_ARCH__)
    ::sincos(x, s, c);
#else
    *s = ::sin(x);
    *c = ::cos(x);
#endif
  }
  static TFK_FN void sincos2_(double xa, double xb, double *sa, double *ca, double *sb, double *cb) {
    sincos_(xa, sa, ca);
    sincos_(xb, sb, cb);
  }
  static TFK_FN double asin_(double x) { return ::asin(x); }
  static TFK_FN double atan2_(double y, double x) { return ::atan2(y, x); }
  static TFK_FN double atan2_rare_(double y, double x) { return ::atan2(y, x); }
  static TFK_FN double sqrt_(double x) { return ::sqrt(x); }
  struct Den {
    double b;
  };
  static TFK_FN Den den_(double b) {
    Den d;
    d.b = b;
    return d;
  }
  static TFK_FN double div_(double a, const Den &d) { return a / d.b; }
  template <int N> static TFK_FN void divn_(double (&a)[N], const Den &d) {
    for (int i = 0; i < N; ++i) a[i] = a[i] / d.b;
  }
};

struct MathNative32 {
  typedef float real;
  enum { SPARE_REGS = 1 }; // see MathPortable (cfg5 FP32 0.348 -> 0.340 ms)
  static TFK_FN float sin_(float x) { return ::sinf(x); }
  static TFK_FN float cos_(float x) { return ::cosf(x); }
  static TFK_FN void sincos_(float x, float *s, float *c) {
#if defined(__CUDA_ARCH__)
    ::sincosf(x, s, c);
#else
    *s = ::sinf(x);
    *c = ::cosf(x);
#endif
  }
  static TFK_FN void sincos2_(float xa, float xb, float *sa, float *ca, float *sb, float *cb) {
    sincos_(xa, sa, ca);
    sincos_(xb, sb, cb);
  }
  static TFK_FN float asin_(float x) { return ::asinf(x); }
  static TFK_FN float atan2_(float y, float x) { return ::atan2f(y, x); }
  static TFK_FN float atan2_rare_(float y, float x) { return ::atan2f(y, x); }
  static TFK_FN float sqrt_(float x) { return ::sqrtf(x); }
  // Several quotients by one denominator (quaternion / norm, rpy_derivative coefficients / cos): ONE reciprocal --
  // MUFU.RCP (1 ulp) and one Newton step -- then a multiplication per quotient (relative error < 2^-22, against the 1e-5
  // contract of this mode).  The IEEE division the compiler emits for `a / b` is ~10 instructions and a slow-path branch
  // per quotient: in the round-1 capture (profiles/r2_full_cfg2_fp32_v12.txt) 24 % of the kernel's instructions.
  struct Den {
    float b, y;
  };
  static TFK_FN Den den_(float b) {
    Den d;
    d.b = b;
#if defined(__CUDA_ARCH__)
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(b));
    const float y1 = fmaf(y, fmaf(-b, y, 1.0f), y);
    d.y = (fabsf(y) < 3.0e38f) ? y1 : y; // b == 0 keeps the infinity
#else
    d.y = 1.0f / b;
#endif
    return d;
  }
  static TFK_FN float div_(float a, const Den &d) { return a * d.y; }
  template <int N> static TFK_FN void divn_(float (&a)[N], const Den &d) {
    TFK_MATH_UNROLL
    for (int i = 0; i < N; ++i) a[i] = a[i] * d.y;
  }
};

// Opt-in FP64 "fast" mode (SURVEY.md section 7 hard part 4): double arithmetic with FMA contraction (the translation
// unit is built with the default -fmad=true), the portable libm's device variant on the GPU (TFK_FAST_PORTABLE_LIBM, set by
// fk_kernels_fast.cu and jit.cpp: cfg2 0.1389 -> 0.1356 ms, cfg3 0.2086 -> 0.2039 ms against libdevice, which builds every
// polynomial coefficient from immediates) and the platform libm on the host, quotients through one correctly rounded
// reciprocal and a multiplication (1 ulp), base-rotation columns analytic.  Not bit-identical to the reference: the
// contract is 1e-12 on poses and analytic Jacobian entries.
struct MathFast64 {
  typedef double real;
  enum { SPARE_REGS = 0 }; // see MathPortable (measured: no gain on cfg4 / cfg5)
  static TFK_FN double sin_(double x) { return ::sin(x); }
  static TFK_FN double cos_(double x) { return ::cos(x); }
#if defined(TFK_FAST_PORTABLE_LIBM) && defined(__CUDA_ARCH__)
  // fkmath_simt.h (explicit fused operations, coefficients as constant-bank operands, straight-line evaluations; atan2's
  // quotient as a reciprocal-multiply): libdevice builds every polynomial coefficient from immediates
  static __device__ __noinline__ TfkSinCos sincos_call_(double x) { return tfk_sincos_simt(x); }
  static TFK_FN void sincos_(double x, double *s, double *c) {
    const TfkSinCos r = sincos_call_(x);
    *s = r.s;
    *c = r.c;
  }
  static __device__ __noinline__ TfkSinCos2 sincos2_call_(double xa, double xb) { return tfk_sincos2_simt(xa, xb); }
  static TFK_FN void sincos2_(double xa, double xb, double *sa, double *ca, double *sb, double *cb) {
    const TfkSinCos2 r = sincos2_call_(xa, xb);
    *sa = r.a.s, *ca = r.a.c, *sb = r.b.s, *cb = r.b.c;
  }
  static TFK_FN double asin_(double x) { return tfk_asin_simt(x); }
  static TFK_FN double atan2_(double y, double x) { return tfk_atan2_simt(y, x); }
  static TFK_FN double atan2_rare_(double y, double x) { return tfk_atan2(y, x); }
#else
  static TFK_FN void sincos_(double x, double *s, double *c) {
#if defined(__CUDA_ARCH__)
    ::sincos(x, s, c);
#else
    *s = ::sin(x);
    *c = ::cos(x);
#endif
  }
  static TFK_FN void sincos2_(double xa, double xb, double *sa, double *ca, double *sb, double *cb) {
    sincos_(xa, sa, ca);
    sincos_(xb, sb, cb);
  }
  static TFK_FN double asin_(double x) { return ::asin(x); }
  static TFK_FN double atan2_(double y, double x) { return ::atan2(y, x); }
  static TFK_FN double atan2_rare_(double y, double x) { return ::atan2(y, x); }
#endif
  static TFK_FN double sqrt_(double x) { return ::sqrt(x); }
  struct Den {
    double b, y;
  };
  static TFK_FN Den den_(double b) {
    Den d;
    d.b = b;
#if defined(__CUDA_ARCH__)
    d.y = __drcp_rn(b);
#else
    d.y = 1.0 / b;
#endif
    return d;
  }
  static TFK_FN double div_(double a, const Den &d) { return a * d.y; }
  template <int N> static TFK_FN void divn_(double (&a)[N], const Den &d) {
    TFK_MATH_UNROLL
    for (int i = 0; i < N; ++i) a[i] = a[i] * d.y;
  }
};

template <class M> struct Ops {
  typedef typename M::real T;
  typedef Vec3<T> V;
  typedef Quat<T> Q;
  typedef PoseT<T> P;

  static TFK_FN Q qidentity() {
    Q q;
    q.x = q.y = q.z = T(0);
    q.w = T(1);
    return q;
  }

  // Rotation::operator*(Rotation), pose.h:253-263
  static TFK_FN Q qmul(const Q &a, const Q &b) {
    Q c;
    c.x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
    c.y = a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x;
    c.z = a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w;
    c.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
    return c;
  }

  // Rotation::GetInverse, pose.h:284-299 (returns the default-constructed identity if norm <= 0)
  static TFK_FN Q qget_inverse(const Q &q) {
    T norm = q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z;
    // the quotients first, the reference's test second: the common path is straight-line code and nothing is preset
    // (a zero / negative / NaN norm takes divn_'s rare path and is overwritten here)
    T a[4] = {q.w, -q.x, -q.y, -q.z};
    M::divn_(a, M::den_(norm));
    Q r;
    r.w = a[0];
    r.x = a[1];
    r.y = a[2];
    r.z = a[3];
    if (TFK_UNLIKELY(!(norm > T(0)))) r = qidentity();
    return r;
  }

  // Rotation::inverse, pose.h:301-304 (no norm check)
  static TFK_FN Q qinverse(const Q &q) {
    T norm = q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z;
    Q r;
    T a[4] = {-q.x, -q.y, -q.z, q.w};
    M::divn_(a, M::den_(norm));
    r.x = a[0];
    r.y = a[1];
    r.z = a[2];
    r.w = a[3];
    return r;
  }

  // Rotation::operator*(Vector3), pose.h:265-282:  q * ((v,0) * GetInverse(q)), xyz part.
  // (v,0) * inv with the literal-zero products dropped (see file header).
  static TFK_FN V rotate(const Q &q, const V &v) {
    Q i = qget_inverse(q);
    Q t;
    t.x = v.x * i.w + v.y * i.z - v.z * i.y;
    t.y = -(v.x * i.z) + v.y * i.w + v.z * i.x;
    t.z = v.x * i.y - v.y * i.x + v.z * i.w;
    t.w = -(v.x * i.x) - v.y * i.y - v.z * i.z;
    V r;
    r.x = q.w * t.x + q.x * t.w + q.y * t.z - q.z * t.y;
    r.y = q.w * t.y - q.x * t.z + q.y * t.w + q.z * t.x;
    r.z = q.w * t.z + q.x * t.y - q.y * t.x + q.z * t.w;
    return r;
  }

  // ---- sparsity-aware pieces of Rotation::operator*(Vector3) (pose.h:265-282) -------------------
  // rotate(q, v) = xyz( q * ((v,0) * inv) ) with inv = GetInverse(q).  The two Hamilton products are
  // split so that (a) `inv` can be kept while q does not change (the reference recomputes it per
  // call: same operands, same bits) and (b) products whose constant factor is an exact zero are not
  // evaluated (fk_plan.h "sparsity classes"; only the sign of an exact zero can differ).

  // (v,0) * i for v = (vx,0,0) / (0,vy,0) / (0,0,vz) / dense
  static TFK_FN Q vq_x(T vx, const Q &i) {
    Q t;
    t.x = vx * i.w;
    t.y = -(vx * i.z);
    t.z = vx * i.y;
    t.w = -(vx * i.x);
    return t;
  }
  static TFK_FN Q vq_y(T vy, const Q &i) {
    Q t;
    t.x = vy * i.z;
    t.y = vy * i.w;
    t.z = -(vy * i.x);
    t.w = -(vy * i.y);
    return t;
  }
  static TFK_FN Q vq_z(T vz, const Q &i) {
    Q t;
    t.x = -(vz * i.y);
    t.y = vz * i.x;
    t.z = vz * i.w;
    t.w = -(vz * i.z);
    return t;
  }
  static TFK_FN Q vq_full(const V &v, const Q &i) {
    Q t;
    t.x = v.x * i.w + v.y * i.z - v.z * i.y;
    t.y = -(v.x * i.z) + v.y * i.w + v.z * i.x;
    t.z = v.x * i.y - v.y * i.x + v.z * i.w;
    t.w = -(v.x * i.x) - v.y * i.y - v.z * i.z;
    return t;
  }
  // cls: V_X = 1, V_Y = 2, V_Z = 3, anything else dense (fk_plan.h)
  static TFK_FN Q vq(int cls, const V &v, const Q &i) {
    if (cls == 1) return vq_x(v.x, i);
    if (cls == 2) return vq_y(v.y, i);
    if (cls == 3) return vq_z(v.z, i);
    return vq_full(v, i);
  }
  // xyz part of q * t
  static TFK_FN V qt_xyz(const Q &q, const Q &t) {
    V r;
    r.x = q.w * t.x + q.x * t.w + q.y * t.z - q.z * t.y;
    r.y = q.w * t.y - q.x * t.z + q.y * t.w + q.z * t.x;
    r.z = q.w * t.z + q.x * t.y - q.y * t.x + q.z * t.w;
    return r;
  }

  // Rotation::operator*(Rotation) (pose.h:253-263) with b = (bx,0,0,bw) / (0,by,0,bw) / (0,0,bz,bw)
  static TFK_FN Q qmul_bx(const Q &a, T bx, T bw) {
    Q c;
    c.x = a.w * bx + a.x * bw;
    c.y = a.y * bw + a.z * bx;
    c.z = -(a.y * bx) + a.z * bw;
    c.w = a.w * bw - a.x * bx;
    return c;
  }
  static TFK_FN Q qmul_by(const Q &a, T by, T bw) {
    Q c;
    c.x = a.x * bw - a.z * by;
    c.y = a.w * by + a.y * bw;
    c.z = a.x * by + a.z * bw;
    c.w = a.w * bw - a.y * by;
    return c;
  }
  static TFK_FN Q qmul_bz(const Q &a, T bz, T bw) {
    Q c;
    c.x = a.x * bw + a.y * bz;
    c.y = -(a.x * bz) + a.y * bw;
    c.z = a.w * bz + a.z * bw;
    c.w = a.w * bw - a.z * bz;
    return c;
  }
  // a * (axis*s, c) where axis has class cls
  static TFK_FN Q qmul_axis(const Q &a, int cls, const V &axis, T s, T c) {
    if (cls == 1) return qmul_bx(a, axis.x * s, c);
    if (cls == 2) return qmul_by(a, axis.y * s, c);
    if (cls == 3) return qmul_bz(a, axis.z * s, c);
    Q b;
    b.x = axis.x * s;
    b.y = axis.y * s;
    b.z = axis.z * s;
    b.w = c;
    return qmul(a, b);
  }

  // urdf::pose_transform, pose.cpp:139-146
  static TFK_FN P compose(const P &a, const P &b) {
    P o;
    o.q = qmul(a.q, b.q);
    V r = rotate(a.q, b.p);
    o.p.x = r.x + a.p.x;
    o.p.y = r.y + a.p.y;
    o.p.z = r.z + a.p.z;
    return o;
  }

  // Rotation::normalize, pose.h:230-250
  static TFK_FN void qnormalize(Q &q) {
    T s = M::sqrt_(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
    if (TFK_UNLIKELY(s == T(0))) {
      q = qidentity();
    } else {
      T a[4] = {q.x, q.y, q.z, q.w};
      M::divn_(a, M::den_(s));
      q.x = a[0];
      q.y = a[1];
      q.z = a[2];
      q.w = a[3];
    }
  }

  // Rotation::setFromRPY, pose.h:202-216
  static TFK_FN_CALL Q from_rpy(T roll, T pitch, T yaw) {
    T phi = roll / T(2), the = pitch / T(2), psi = yaw / T(2);
    T sph, cph, sth, cth, sps, cps;
    M::sincos_(phi, &sph, &cph);
    M::sincos_(the, &sth, &cth);
    M::sincos_(psi, &sps, &cps);
    Q q;
    q.x = sph * cth * cps - cph * sth * sps;
    q.y = cph * sth * cps + sph * cth * sps;
    q.z = cph * cth * sps - sph * sth * cps;
    q.w = cph * cth * cps + sph * sth * sps;
    qnormalize(q);
    return q;
  }

  // Rotation::getRPY, pose.h:161-192
  static TFK_FN_CALL V to_rpy(Q q) {
    T sqx = q.x * q.x, sqy = q.y * q.y, sqz = q.z * q.z, sqw = q.w * q.w;
    T sarg = T(-2) * (q.x * q.z - q.w * q.y);
    const T pi_2 = T(1.57079632679489661923);
    V r;
    if (TFK_UNLIKELY(sarg <= T(-0.99999))) {
      r.y = -pi_2;
      r.x = T(0);
      r.z = T(2) * M::atan2_rare_(q.x, -q.y);
    } else if (TFK_UNLIKELY(sarg >= T(0.99999))) {
      r.y = pi_2;
      r.x = T(0);
      r.z = T(2) * M::atan2_rare_(-q.x, q.y);
    } else {
      r.y = M::asin_(sarg);
      r.x = M::atan2_(T(2) * (q.y * q.z + q.w * q.x), sqw - sqx - sqy + sqz);
      r.z = M::atan2_(T(2) * (q.x * q.y + q.w * q.z), sqw + sqx - sqy - sqz);
    }
    return r;
  }

  // Joint::transform for REVOLUTE/CONTINUOUS, joint.h:231-240 (axis is NOT normalised)
  static TFK_FN Q joint_quat(const V &axis, T angle) {
    T s, c;
    M::sincos_(angle * T(0.5), &s, &c);
    Q q;
    q.x = axis.x * s;
    q.y = axis.y * s;
    q.z = axis.z * s;
    q.w = c;
    return q;
  }

  // pose_transform(origin, Joint::transform(angle)) for a revolute joint (src/kinematics.cpp:77-80).
  // Joint::transform leaves position = 0, so rotate(origin.q, 0) = (+-0,+-0,+-0) and the sum is
  // origin.p (up to the sign of an exact zero).
  static TFK_FN P local_revolute(const P &origin, const V &axis, T angle) {
    P o;
    o.q = qmul(origin.q, joint_quat(axis, angle));
    o.p = origin.p;
    return o;
  }

  // same for PRISMATIC (joint.h:241-247): rotation identity, position axis*angle.  Evaluated in full.
  static TFK_FN P local_prismatic(const P &origin, const V &axis, T angle) {
    P j;
    j.q = qidentity();
    j.p.x = axis.x * angle;
    j.p.y = axis.y * angle;
    j.p.z = axis.z * angle;
    return compose(origin, j);
  }

  // Pose::inverse, pose.h:329-339 (Rotation::inverse_inplace :306-312, Vector3::inverse_inplace :95-99)
  static TFK_FN P inverse(const P &a) {
    P o;
    T norm = a.q.w * a.q.w + a.q.x * a.q.x + a.q.y * a.q.y + a.q.z * a.q.z;
    T qa[4] = {a.q.w, -a.q.x, -a.q.y, -a.q.z};
    M::divn_(qa, M::den_(norm));
    o.q.w = qa[0];
    o.q.x = qa[1];
    o.q.y = qa[2];
    o.q.z = qa[3];
    V n;
    n.x = a.p.x * T(-1);
    n.y = a.p.y * T(-1);
    n.z = a.p.z * T(-1);
    o.p = rotate(o.q, n);
    return o;
  }

  // ---- yaw-only rotations (PLANER base) ------------------------------------------------------------
  // With base_type PLANER every base quantity is built from q = (+-0, +-0, z, w): setFromRPY(0, 0, yaw)
  // (sin(0) = 0, cos(0) = 1 exactly), its inverse, and the yaw-tweaked rotation of the finite-difference
  // column.  These are the reference's dense expressions (pose.h) with the products by those two exact
  // zeros dropped -- the sparsity policy of fk_plan.h: only the sign of an exact zero can differ.
  struct QZ {
    T z, w;
  };
  struct PZ {
    V p;
    QZ q;
  };
  static TFK_FN Q qz_dense(const QZ &q) {
    Q r;
    r.x = T(0);
    r.y = T(0);
    r.z = q.z;
    r.w = q.w;
    return r;
  }
  // setFromRPY(0, 0, yaw) incl. normalize (pose.h:202-216, 230-250)
  static TFK_FN QZ qz_from_yaw(T yaw) {
    T s, c;
    M::sincos_(yaw / T(2), &s, &c);
    QZ q;
    q.z = s;
    q.w = c;
    const T n = M::sqrt_(s * s + c * c); // x*x + y*y + z*z + w*w with x = y = 0
    if (TFK_UNLIKELY(n == T(0))) {
      q.z = T(0);
      q.w = T(1);
    } else {
      T a[2] = {s, c};
      M::divn_(a, M::den_(n));
      q.z = a[0];
      q.w = a[1];
    }
    return q;
  }
  // getRPY (pose.h:161-192) of a yaw-only rotation: roll = atan2(+-0, w^2 + z^2) = +-0, pitch = asin(+-0) = +-0
  static TFK_FN T qz_yaw(const QZ &q) { return M::atan2_(T(2) * (q.w * q.z), q.w * q.w - q.z * q.z); }
  // GetInverse (pose.h:284-299)
  static TFK_FN QZ qz_get_inverse(const QZ &q) {
    const T norm = q.w * q.w + q.z * q.z;
    T a[2] = {q.w, -q.z}; // quotients first, the reference's test second (see qget_inverse)
    M::divn_(a, M::den_(norm));
    QZ r;
    r.w = a[0];
    r.z = a[1];
    if (TFK_UNLIKELY(!(norm > T(0)))) {
      r.z = T(0);
      r.w = T(1);
    }
    return r;
  }
  // a * b, a yaw-only (pose.h:253-263)
  static TFK_FN Q qmul_zb(const QZ &a, const Q &b) {
    Q c;
    c.x = a.w * b.x - a.z * b.y;
    c.y = a.w * b.y + a.z * b.x;
    c.z = a.w * b.z + a.z * b.w;
    c.w = a.w * b.w - a.z * b.z;
    return c;
  }
  // Rotation * Vector3 (pose.h:265-282), q yaw-only; i = GetInverse(q) (the reference recomputes it per call: same
  // operands, same bits, so callers that rotate several vectors by one q take it once)
  static TFK_FN V rotate_z_i(const QZ &q, const QZ &i, const V &v) {
    Q t;
    t.x = v.x * i.w + v.y * i.z;
    t.y = -(v.x * i.z) + v.y * i.w;
    t.z = v.z * i.w;
    t.w = -(v.z * i.z);
    V r;
    r.x = q.w * t.x - q.z * t.y;
    r.y = q.w * t.y + q.z * t.x;
    r.z = q.w * t.z + q.z * t.w;
    return r;
  }
  static TFK_FN V rotate_z(const QZ &q, const V &v) { return rotate_z_i(q, qz_get_inverse(q), v); }
  // Pose::inverse (pose.h:329-339) of a yaw-only pose
  static TFK_FN PZ inverse_z(const PZ &a) {
    PZ o;
    const T norm = a.q.w * a.q.w + a.q.z * a.q.z;
    T qa[2] = {a.q.w, -a.q.z};
    M::divn_(qa, M::den_(norm));
    o.q.w = qa[0];
    o.q.z = qa[1];
    V n;
    n.x = a.p.x * T(-1);
    n.y = a.p.y * T(-1);
    n.z = a.p.z * T(-1);
    o.p = rotate_z(o.q, n);
    return o;
  }

  // urdf::cross_product, pose.cpp:90-94
  static TFK_FN V cross(const V &a, const V &b) {
    V o;
    o.x = a.y * b.z - a.z * b.y;
    o.y = a.z * b.x - a.x * b.z;
    o.z = a.x * b.y - a.y * b.x;
    return o;
  }

  // Coefficients of rpy_derivative (src/kinematics.cpp:14-23) that depend only on the link's rpy.
  // The reference recomputes cos/sin per joint; the values are the same, so hoisting is exact.
  struct RpyCoef {
    T k_xx, k_xy; //  cos(a3)/cos(a2),  sin(a3)/cos(a2)
    T s3, c3;     //  sin(a3), cos(a3)
    T k_zx, k_zy; // -cos(a3)*sin(a2)/cos(a2),  sin(a3)*sin(a2)/cos(a2)
  };
  static TFK_FN_CALL RpyCoef rpy_coef(V rpy) {
    T a2 = -rpy.y, a3 = -rpy.z;
    T s2, c2, s3, c3;
    M::sincos_(a2, &s2, &c2);
    M::sincos_(a3, &s3, &c3);
    RpyCoef k;
    T a[4] = {c3, s3, -c3 * s2, s3 * s2};
    M::divn_(a, M::den_(c2));
    k.k_xx = a[0];
    k.k_xy = a[1];
    k.s3 = s3;
    k.c3 = c3;
    k.k_zx = a[2];
    k.k_zy = a[3];
    return k;
  }
  static TFK_FN V rpy_derivative(const RpyCoef &k, const V &axis) {
    V d;
    d.x = k.k_xx * axis.x - k.k_xy * axis.y;
    d.y = k.s3 * axis.x + k.c3 * axis.y;
    d.z = k.k_zx * axis.x + k.k_zy * axis.y + axis.z;
    return d;
  }

  // q_derivative, src/kinematics.cpp:25-35 (literal 0 * q terms dropped); returns (dx,dy,dz,-dw)
  static TFK_FN Q q_derivative(const Q &q, const V &om) {
    Q d;
    d.x = T(0.5) * (om.z * q.y - om.y * q.z + om.x * q.w);
    d.y = T(0.5) * (-(om.z * q.x) + om.x * q.z + om.y * q.w);
    d.z = T(0.5) * (om.y * q.x - om.x * q.y + om.z * q.w);
    d.w = -(T(0.5) * (-(om.x * q.x) - om.y * q.y - om.z * q.z));
    return d;
  }
};

} // namespace tfk
#endif
