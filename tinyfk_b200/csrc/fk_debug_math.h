// fk_debug_math.h -- test hook behind tfk_debug_flag_math (include/tinyfk_b200.h); see fk_debug_math.cu
#ifndef TFK_FK_DEBUG_MATH_H
#define TFK_FK_DEBUG_MATH_H

#include <cuda_runtime.h>

namespace tfk {
cudaError_t debug_flag_math(int op, long long n, const double *h_x, const double *h_y, double *h_out, double *h_ref, int *h_flags);
}
#endif
