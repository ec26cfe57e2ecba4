// Experiment: the fused kernel compiled against the cfg2 plan as compile-time constants (spec_cfg2.inc from
// plan_to_source).  Exposes spec_launch() so a Python harness can time it and compare with the generic kernel.
#define TFK_SPEC_PLAN 1
#include <stdint.h>
#include "../../tinyfk_b200/csrc/fk_plan.h"
#include "spec_cfg2.inc"
#include "../../tinyfk_b200/csrc/fk_kernels.cuh"

using namespace tfk;
typedef FkCore<MathPortable, false, 1, TFK_TILE> Core;

extern "C" __global__ void __launch_bounds__(TFK_TILE, 512 / TFK_TILE) tfk_spec_kernel(const KernelArgs a) {
  fk_fused_spec_body<Core, true, 512 / TFK_TILE>(a);
}

extern "C" int spec_tile() { return TFK_TILE; }
extern "C" int spec_smem() {
  return smem_layout(0, spec::k_h.n_in, spec::k_h.fz_size, 8).total;
}
extern "C" int spec_launch(const void *q, long long n, void *P, void *J, int grid, void *stream) {
  KernelArgs a = KernelArgs();
  a.q = q;
  a.n = n;
  a.row0 = 0;
  a.ld_cfg = n;
  a.P = P;
  a.J = J;
  const int smem = spec_smem();
  static bool once = false;
  if (!once) {
    cudaFuncSetAttribute(tfk_spec_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    once = true;
  }
  tfk_spec_kernel<<<grid, TFK_TILE, smem, (cudaStream_t)stream>>>(a);
  return (int)cudaGetLastError();
}
