// Microbenchmark: FP64 pipe latency / throughput on one SM (B200).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3
// usage: fp64_pipe   -> prints cycles per instruction for (warps per SM, independent chains per thread)
#include <cstdio>
#include <cuda_runtime.h>

template <int ILP, int OP>
__global__ void chain(double *out, long long *cyc, int iters, double a, double b) {
  double x[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) x[i] = a + threadIdx.x + i;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 16; ++u) {
#pragma unroll
      for (int i = 0; i < ILP; ++i) {
        if (OP == 0) x[i] = __fma_rn(x[i], b, a);
        if (OP == 1) x[i] = __dmul_rn(x[i], b);
        if (OP == 2) x[i] = __dadd_rn(x[i], b);
      }
    }
  }
  long long t1 = clock64();
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

// mixed: per FP64 instruction, K integer instructions (independent) -- does the ALU issue in the FP64 pipe's shadow?
template <int K>
__global__ void mixed(double *out, long long *cyc, int iters, double a, double b, int m) {
  double x[4];
  int y[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) x[i] = a + threadIdx.x + i, y[i] = threadIdx.x + i;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 16; ++u) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        x[i] = __fma_rn(x[i], b, a);
#pragma unroll
        for (int k = 0; k < K; ++k) y[(i + k) & 3] = (y[(i + k) & 3] ^ m) + k;
      }
    }
  }
  long long t1 = clock64();
  double s = 0;
  int t = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) s += x[i], t += y[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s + t;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <class F> static void run(const char *name, F launch, int warps, int ilp_total_per_iter) {
  double *out;
  long long *cyc, h;
  cudaMalloc(&out, 1024 * 8);
  cudaMalloc(&cyc, 8);
  const int iters = 2000;
  launch(out, cyc, 10, warps);
  launch(out, cyc, iters, warps);
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("%-28s warps/SM %2d  cycles per warp-instr per SMSP-warp: %.3f   (SM-wide instr/cycle %.3f)\n", name, warps,
         (double)h / ((double)iters * ilp_total_per_iter), (double)iters * ilp_total_per_iter * warps / (double)h);
  cudaFree(out);
  cudaFree(cyc);
}

int main() {
  const int W[] = {1, 2, 4, 8, 16, 20, 32};
  for (int w : W) {
    run("DFMA ilp1", [](double *o, long long *c, int it, int wp) { chain<1, 0><<<1, 32 * wp>>>(o, c, it, 1.0, 1.0000001); }, w, 16);
    run("DFMA ilp2", [](double *o, long long *c, int it, int wp) { chain<2, 0><<<1, 32 * wp>>>(o, c, it, 1.0, 1.0000001); }, w, 32);
    run("DFMA ilp4", [](double *o, long long *c, int it, int wp) { chain<4, 0><<<1, 32 * wp>>>(o, c, it, 1.0, 1.0000001); }, w, 64);
    run("DMUL ilp1", [](double *o, long long *c, int it, int wp) { chain<1, 1><<<1, 32 * wp>>>(o, c, it, 1.0, 1.0000001); }, w, 16);
    run("DADD ilp1", [](double *o, long long *c, int it, int wp) { chain<1, 2><<<1, 32 * wp>>>(o, c, it, 1.0, 1.0000001); }, w, 16);
    run("DFMA ilp4 + 1 int each", [](double *o, long long *c, int it, int wp) { mixed<1><<<1, 32 * wp>>>(o, c, it, 1.0, 1.0000001, 5); }, w, 64);
    run("DFMA ilp4 + 2 int each", [](double *o, long long *c, int it, int wp) { mixed<2><<<1, 32 * wp>>>(o, c, it, 1.0, 1.0000001, 5); }, w, 64);
    run("DFMA ilp4 + 4 int each", [](double *o, long long *c, int it, int wp) { mixed<4><<<1, 32 * wp>>>(o, c, it, 1.0, 1.0000001, 5); }, w, 64);
  }
  cudaError_t e = cudaDeviceSynchronize();
  printf("%s\n", cudaGetErrorString(e));
  return 0;
}
