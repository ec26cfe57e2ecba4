// fk_kernels_strict.cu -- FP64 "strict" instantiation.  MUST be compiled with -fmad=false: every
// multiply and add rounds separately, as in the reference's x86-64 build (CMakeLists.txt:3-14 has no
// -march/-ffast-math, so no FMA contraction).  libm = fkmath.h, shared bit for bit with the host.
#include "fk_kernels.cuh"

namespace tfk {
cudaError_t launch_fk_f64_strict(const KernelArgs &a, const PlanParams &pp, int variant, int grid, int smem_bytes,
    cudaStream_t stream) {
  return launch_fk<MathPortable, false>(a, pp, variant, grid, smem_bytes, stream);
}
cudaError_t occupancy_fk_f64_strict(int rot_type, int variant, int smem_bytes, int *blocks_per_sm) {
  return occupancy_fk<MathPortable, false>(rot_type, variant, smem_bytes, blocks_per_sm);
}
} // namespace tfk
