// fk_debug_math.cu -- test hook (include/tinyfk_b200.h:tfk_debug_flag_math): evaluates the branch-free strict arithmetic
// of fkmath_flag.h element by element on the GPU, next to the functions it must equal, so that tests can compare them
// bit for bit on millions of arguments (tests/test_flag_math_gpu.py).  Built with -fmad=false like the strict kernels.
#include <cuda_runtime.h>

#include "fk_debug_math.h"
#include "fk_math.h"

namespace tfk {
namespace {

// op: 0 rcp  1 div  2 sqrt  3 sincos (out[i], out[n + i])  4 asin  5 atan2(x[i], y[i])
// out: the flag variant; ref: the ordinary device function of the strict kernels; flags[i]: 1 where the variant reported
// a rare input (its result is then not to be trusted)
__global__ void flag_math_kernel(int op, long long n, const double *x, const double *y, double *out, double *ref, int *flags) {
#if defined(__CUDA_ARCH__) // fkmath_flag.h exists in the device pass only
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    int bad = 0;
    const double a = x[i], b = y ? y[i] : 0.0;
    if (op == 0) {
      out[i] = tfk_rcp_flag(a, bad);
      ref[i] = __drcp_rn(a);
    } else if (op == 1) {
      out[i] = tfk_div_flag(a, b, bad);
      ref[i] = a / b;
    } else if (op == 2) {
      out[i] = tfk_sqrt_flag(a, bad);
      ref[i] = sqrt(a);
    } else if (op == 3) {
      const TfkSinCos r = tfk_sincos_flag(a, bad), s = tfk_sincos_simt(a);
      out[i] = r.s, out[n + i] = r.c;
      ref[i] = s.s, ref[n + i] = s.c;
    } else if (op == 4) {
      out[i] = tfk_asin_flag(a, bad);
      ref[i] = tfk_asin_simt(a);
    } else {
      out[i] = tfk_atan2_flag(a, b, bad);
      ref[i] = tfk_atan2_simt(a, b);
    }
    flags[i] = bad;
  }
#endif
}

} // namespace

cudaError_t debug_flag_math(int op, long long n, const double *h_x, const double *h_y, double *h_out, double *h_ref, int *h_flags) {
  const size_t nb = (size_t)n * sizeof(double), no = nb * (op == 3 ? 2 : 1);
  double *x = nullptr, *y = nullptr, *out = nullptr, *ref = nullptr;
  int *flags = nullptr;
  cudaError_t e = cudaMalloc(&x, nb);
  if (e == cudaSuccess && h_y) e = cudaMalloc(&y, nb);
  if (e == cudaSuccess) e = cudaMalloc(&out, no);
  if (e == cudaSuccess) e = cudaMalloc(&ref, no);
  if (e == cudaSuccess) e = cudaMalloc(&flags, (size_t)n * sizeof(int));
  if (e == cudaSuccess) e = cudaMemcpy(x, h_x, nb, cudaMemcpyHostToDevice);
  if (e == cudaSuccess && h_y) e = cudaMemcpy(y, h_y, nb, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    flag_math_kernel<<<1184, 256>>>(op, n, x, y, out, ref, flags);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpy(h_out, out, no, cudaMemcpyDeviceToHost);
  if (e == cudaSuccess) e = cudaMemcpy(h_ref, ref, no, cudaMemcpyDeviceToHost);
  if (e == cudaSuccess) e = cudaMemcpy(h_flags, flags, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost);
  cudaFree(x);
  cudaFree(y);
  cudaFree(out);
  cudaFree(ref);
  cudaFree(flags);
  return e;
}

} // namespace tfk
