// fk_kernels_fast.cu -- the opt-in modes.  FP32 mode (FMA contraction on, libdevice libm, analytic base-rotation
// columns: the reference's eps = 1e-7 forward difference is below one FP32 ulp).  Tolerance vs the
// reference: 1e-5 (BASELINE.json north_star).
// The FP64 fast mode evaluates sin / cos / asin / atan2 with the portable libm's device variant (fkmath_simt.h; atan2's quotient as
// a reciprocal-multiply) -- measured faster than libdevice's (fk_math.h:MathFast64); no strict code lives in this translation unit.
#define TFK_FAST_PORTABLE_LIBM 1
#include "fk_kernels.cuh"

namespace tfk {
cudaError_t launch_fk_f32(const KernelArgs &a, const PlanParams &pp, int variant, int grid, int smem_bytes,
    cudaStream_t stream) {
  return launch_fk<MathNative32, true>(a, pp, variant, grid, smem_bytes, stream);
}
cudaError_t occupancy_fk_f32(int rot_type, int variant, int smem_bytes, int *blocks_per_sm) {
  return occupancy_fk<MathNative32, true>(rot_type, variant, smem_bytes, blocks_per_sm);
}
// opt-in FP64 fast mode (fk_math.h MathFast64): 1e-12 on poses and analytic Jacobian entries, not bit-identical
cudaError_t launch_fk_f64_fast(const KernelArgs &a, const PlanParams &pp, int variant, int grid, int smem_bytes,
    cudaStream_t stream) {
  return launch_fk<MathFast64, true>(a, pp, variant, grid, smem_bytes, stream);
}
cudaError_t occupancy_fk_f64_fast(int rot_type, int variant, int smem_bytes, int *blocks_per_sm) {
  return occupancy_fk<MathFast64, true>(rot_type, variant, smem_bytes, blocks_per_sm);
}
} // namespace tfk
