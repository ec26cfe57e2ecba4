// TEST INFRASTRUCTURE: applies the kernels' libm (tinyfk_b200/csrc/fkmath_simt.h -- what MathPortable calls on the device) to
// arrays on the GPU, so that tests/test_libm_device.py can compare it bit for bit with the host build of fkmath.h that the pinned
// reference links (oracle/strict_libm.c).  Built at test time with the flags of the strict translation unit
// (nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false).
#include <cuda_runtime.h>

#include "../../tinyfk_b200/csrc/fkmath_simt.h"

namespace {
__global__ void k_sincos(const double *x, double *s, double *c, long n) {
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    const TfkSinCos r = tfk_sincos_simt(x[i]);
    s[i] = r.s;
    c[i] = r.c;
  }
}
__global__ void k_sincos2(const double *xa, const double *xb, double *sa, double *ca, double *sb, double *cb, long n) {
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    const TfkSinCos2 r = tfk_sincos2_simt(xa[i], xb[i]);
    sa[i] = r.a.s, ca[i] = r.a.c, sb[i] = r.b.s, cb[i] = r.b.c;
  }
}
__global__ void k_asin(const double *x, double *o, long n) {
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) o[i] = tfk_asin_simt(x[i]);
}
__global__ void k_atan2(const double *y, const double *x, double *o, long n) {
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) o[i] = tfk_atan2_simt(y[i], x[i]);
}
// the fkmath.h functions themselves as device code (the kernels reach them through the rare-input branches)
__global__ void k_plain(const double *y, const double *x, double *s, double *c, double *as, double *at, long n) {
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    tfk_sincos(x[i], &s[i], &c[i]);
    as[i] = tfk_asin(x[i]);
    at[i] = tfk_atan2(y[i], x[i]);
  }
}

struct Buf {
  double *d = nullptr;
  Buf(const double *h, long n) {
    if (cudaMalloc(&d, n * sizeof(double)) == cudaSuccess && h) cudaMemcpy(d, h, n * sizeof(double), cudaMemcpyHostToDevice);
  }
  ~Buf() { cudaFree(d); }
  int get(double *h, long n) const { return cudaMemcpy(h, d, n * sizeof(double), cudaMemcpyDeviceToHost) == cudaSuccess ? 0 : 1; }
};
int done() { return (cudaDeviceSynchronize() == cudaSuccess && cudaGetLastError() == cudaSuccess) ? 0 : 1; }
} // namespace

extern "C" {
int probe_sincos(const double *x, double *s, double *c, long n) {
  Buf X(x, n), S(nullptr, n), C(nullptr, n);
  k_sincos<<<592, 256>>>(X.d, S.d, C.d, n);
  return done() | S.get(s, n) | C.get(c, n);
}
int probe_sincos2(const double *xa, const double *xb, double *sa, double *ca, double *sb, double *cb, long n) {
  Buf XA(xa, n), XB(xb, n), SA(nullptr, n), CA(nullptr, n), SB(nullptr, n), CB(nullptr, n);
  k_sincos2<<<592, 256>>>(XA.d, XB.d, SA.d, CA.d, SB.d, CB.d, n);
  return done() | SA.get(sa, n) | CA.get(ca, n) | SB.get(sb, n) | CB.get(cb, n);
}
int probe_asin(const double *x, double *o, long n) {
  Buf X(x, n), O(nullptr, n);
  k_asin<<<592, 256>>>(X.d, O.d, n);
  return done() | O.get(o, n);
}
int probe_atan2(const double *y, const double *x, double *o, long n) {
  Buf Y(y, n), X(x, n), O(nullptr, n);
  k_atan2<<<592, 256>>>(Y.d, X.d, O.d, n);
  return done() | O.get(o, n);
}
int probe_plain(const double *y, const double *x, double *s, double *c, double *as, double *at, long n) {
  Buf Y(y, n), X(x, n), S(nullptr, n), C(nullptr, n), AS(nullptr, n), AT(nullptr, n);
  k_plain<<<592, 256>>>(Y.d, X.d, S.d, C.d, AS.d, AT.d, n);
  return done() | S.get(s, n) | C.get(c, n) | AS.get(as, n) | AT.get(at, n);
}
}
