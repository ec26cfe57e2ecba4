// fk_kernels_fast.cu -- opt-in FP32 mode (FMA contraction on, libdevice libm, analytic base-rotation
// columns: the reference's eps = 1e-7 forward difference is below one FP32 ulp).  Tolerance vs the
// reference: 1e-5 (BASELINE.json north_star).
#include "fk_kernels.cuh"

namespace tfk {
cudaError_t launch_fk_f32(const KernelArgs &a, const PlanParams &pp, int variant, int grid, int smem_bytes,
    cudaStream_t stream) {
  return launch_fk<MathNative32, true>(a, pp, variant, grid, smem_bytes, stream);
}
cudaError_t occupancy_fk_f32(int rot_type, int variant, int smem_bytes, int *blocks_per_sm) {
  return occupancy_fk<MathNative32, true>(rot_type, variant, smem_bytes, blocks_per_sm);
}
} // namespace tfk
