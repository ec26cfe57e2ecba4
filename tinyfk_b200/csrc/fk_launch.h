// fk_launch.h -- the narrow interface between the host engine (engine.cpp) and the CUDA translation
// units.  One entry per precision mode; each TU is compiled with the FP contraction setting its mode
// needs (strict: nvcc -fmad=false).
#ifndef TFK_FK_LAUNCH_H
#define TFK_FK_LAUNCH_H

#include <cuda_runtime.h>

namespace tfk {

enum Precision {
  PREC_F64_STRICT = 0, // reference operation order, portable libm, FD base columns, no FMA contraction
  PREC_F32 = 1,        // float arithmetic, libdevice libm, analytic base-rotation columns
};

struct KernelArgs {
  const unsigned char *plan; // device copy of the plan blob (16-byte aligned, padded to 16)
  int plan_bytes;
  const void *q;     // [n][n_in] reals
  long long n;       // configurations in this launch
  long long row0;    // global row of configuration 0 inside P / J
  long long ld_cfg;  // configurations the destination arrays are dimensioned for (ldP = ld_cfg*n_out)
  void *P, *J;       // column-major outputs (J may be null when the plan has no Jacobian)
  void *scratch;     // grid * TFK_TILE * rec_size reals
};

cudaError_t launch_fk_f64_strict(const KernelArgs &a, int grid, cudaStream_t stream);
cudaError_t occupancy_fk_f64_strict(int plan_bytes, int *blocks_per_sm);
cudaError_t launch_fk_f32(const KernelArgs &a, int grid, cudaStream_t stream);
cudaError_t occupancy_fk_f32(int plan_bytes, int *blocks_per_sm);

} // namespace tfk
#endif
