// fk_kernels.cuh -- the sm_100a batch kernel: persistent blocks, one tile of TFK_TILE configurations
// per iteration, phase A (thread per configuration) -> block-private records -> phase B (thread per
// (configuration, link)), see fk_core.h.  Included by fk_kernels_strict.cu (built with -fmad=false:
// bit-faithful FP64) and fk_kernels_fast.cu (FMA contraction on: FP32 mode).
//
// Memory plan (DESIGN.md "HBM layout"):
//   q        [N][n_in]           row-major input, read once
//   P        [pose_dim][N*n_out] column-major, written once      (reference layout, wrapper.cpp:80,123)
//   J        [n_out_cols][N*n_out*pose_dim] column-major, written once, every entry (zeros included)
//   records  [grid][rec_size][TFK_TILE] block-private scratch, written in phase A and re-read in phase B
//            of the same tile; grid*TILE*rec_size*8 B total (L2-resident: 45 MB for the PR2-arm case)
//   plan     a few KB, copied to shared memory once per block, read with warp-uniform loads
#ifndef TFK_FK_KERNELS_CUH
#define TFK_FK_KERNELS_CUH

#include <cuda_runtime.h>

#include "fk_core.h"
#include "fk_launch.h"

namespace tfk {

template <class Core>
__global__ void __launch_bounds__(TFK_TILE) fk_batch_kernel(const KernelArgs a) {
  typedef typename Core::T T;
  extern __shared__ __align__(16) unsigned char plan_smem[];
  for (int i = threadIdx.x * 16; i < a.plan_bytes; i += TFK_TILE * 16)
    *reinterpret_cast<int4 *>(plan_smem + i) = *reinterpret_cast<const int4 *>(a.plan + i);
  __syncthreads();
  const PlanView pv = plan_view(plan_smem);
  const PlanHeader &h = *pv.h;

  T *rec = static_cast<T *>(a.scratch) + (size_t)blockIdx.x * TFK_TILE * h.rec_size;
  const T *q = static_cast<const T *>(a.q);
  T *Pout = static_cast<T *>(a.P);
  T *Jout = static_cast<T *>(a.J);
  const long long ldP = a.ld_cfg * h.n_out;
  const long long ldJ = ldP * h.pose_dim;
  const long long n_tiles = (a.n + TFK_TILE - 1) / TFK_TILE;

  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long i0 = tile * TFK_TILE;
    const int nv = (a.n - i0 < TFK_TILE) ? (int)(a.n - i0) : (int)TFK_TILE;
    if ((int)threadIdx.x < nv) Core::phase_a(pv, q + (i0 + threadIdx.x) * h.n_in, rec + threadIdx.x, TFK_TILE);
    __syncthreads();
    const int items = nv * h.n_out;
    for (int it = threadIdx.x; it < items; it += TFK_TILE) {
      const int il = it / h.n_out;
      const int o = it - il * h.n_out;
      Core::phase_b(pv, rec + il, TFK_TILE, a.row0 + i0 + il, o, Pout, Jout, ldP, ldJ);
    }
    __syncthreads();
  }
}

template <class Core> cudaError_t launch_fk(const KernelArgs &a, int grid, cudaStream_t stream) {
  fk_batch_kernel<Core><<<grid, TFK_TILE, a.plan_bytes, stream>>>(a);
  return cudaGetLastError();
}

template <class Core> cudaError_t occupancy_fk(int plan_bytes, int *blocks_per_sm) {
  cudaError_t e = cudaSuccess;
  if (plan_bytes > 48 * 1024)
    e = cudaFuncSetAttribute(fk_batch_kernel<Core>, cudaFuncAttributeMaxDynamicSharedMemorySize, plan_bytes);
  if (e != cudaSuccess) return e;
  return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, fk_batch_kernel<Core>, TFK_TILE, plan_bytes);
}

} // namespace tfk
#endif
