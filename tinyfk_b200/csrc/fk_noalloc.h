// fk_noalloc.h -- interface between the host engine and fk_noalloc.cu (the allocation-free variant of the several-link
// kernel; include/tinyfk_b200.h, tfk_solve_fk_device_noalloc).  No counterpart in the reference.
#ifndef TFK_FK_NOALLOC_H
#define TFK_FK_NOALLOC_H

#include <cuda_runtime.h>

#include "fk_launch.h"

namespace tfk {

// The static device array every GPU that loads the module holds until exit, and its split (fk_noalloc.cu).  Keep in step
// with TFK_NOALLOC_* in include/tinyfk_b200.h (capi.cpp static_asserts it).
enum {
  TFK_NOALLOC_SCRATCH_BYTES = 32 << 20, // 32 MiB in all
  TFK_NOALLOC_TABLE_BYTES = 64 << 10,   // head of the array: the plan's per-output tables
  TFK_NOALLOC_MAX_CONFIGS_ = 1 << 16,   // configurations per call
};

cudaError_t occupancy_fk_noalloc(int rot_type, int smem_bytes, int *blocks_per_sm);
cudaError_t launch_fk_noalloc(int device, KernelArgs a, const PlanParams &pp, const unsigned char *tables_host, int table_bytes,
                              int grid, int smem_bytes, cudaStream_t stream);

} // namespace tfk
#endif
