// fk_launch.h -- the narrow interface between the host engine (engine.cpp) and the CUDA translation
// units.  One entry per precision mode; each TU is compiled with the FP contraction setting its mode
// needs (strict: nvcc -fmad=false).
#ifndef TFK_FK_LAUNCH_H
#define TFK_FK_LAUNCH_H

#if !defined(__CUDACC_RTC__)
#include <cuda_runtime.h>
#endif

#include "fk_plan.h"

namespace tfk {

enum Precision {
  PREC_F64_STRICT = 0, // reference operation order, portable libm, FD base columns, no FMA contraction
  PREC_F32 = 1,        // float arithmetic, libdevice libm, analytic base-rotation columns
  PREC_F64_FAST = 2,   // double arithmetic with FMA contraction, libdevice libm, reciprocal-multiply quotients, analytic
                       // base-rotation columns: 1e-12 on poses and analytic columns, not bit-identical
  PREC_COUNT = 3,
};

enum KernelVariant {
  KV_FUSED_SMEM = 0, // one requested link, joint records in shared memory
  KV_FUSED_GMEM = 1, // one requested link, records too large for shared memory: block-private global scratch
  KV_TWOPHASE = 2,   // several links: phase A -> records in global scratch -> phase B
  KV_SQDIST = 4,      // inter-link squared distances: phase A -> records -> one thread per (configuration, pair)
  KV_COM = 5,         // centre of mass (+ Jacobian): phase A -> records -> mass-weighted sums in link-id order
  KV_INERTIA = 6,
  KV_TWOPHASE_SMEM = 8, // several links, [base | arec | prec] in shared memory, frames and link records in scratch
  KV_SQDIST_SMEM = 9,   // KV_SQDIST with [base | arec | prec] in shared memory
  KV_FUSED_SPEC = 7,  // plan-specialised build of the fused kernel, compiled at run time (jit.cpp)
     // total inertia matrix about the centre of mass: phase A (with rotations) -> records -> 3x3 sums
  KV_TWOPHASE_SPEC = 12, // plan-specialised build of the two-phase kernel (several links), compiled at run time
  KV_FUSED_SMEM5 = 3, // KV_FUSED_SMEM compiled for 5 resident blocks per SM (<= 96 registers), used when the
                      // shared-memory footprint leaves room for the fifth block
};

struct KernelArgs {
  const unsigned char *tables; // two-phase: device copy of the plan's per-output tables (lrec_of_out, masks)
  int table_bytes;             // 16-byte multiple
  const void *q;     // [n][n_in] reals
  long long n;       // configurations in this launch
  long long row0;    // global row of configuration 0 inside P / J
  long long ld_cfg;  // configurations the destination arrays are dimensioned for (ldP = ld_cfg*n_out)
  void *P, *J;       // column-major outputs (J may be null when the plan has no Jacobian)
  void *scratch;     // grid * TFK_TILE * (rec_size | fz_size) reals (unused by KV_FUSED_SMEM)
  // KV_SQDIST only: P = sqdists [ld_cfg*n_pairs], J = grads column-major (ld_cfg*n_pairs) x n_dof (or null)
  const int *pairs;  // device: o1[n_pairs] then o2[n_pairs], indices into the plan's elink list
  int n_pairs;
  // KV_INERTIA: P = out (9, ld_cfg); weights = 7 reals per com link (mass, ixx, ixy, ixz, iyy, iyz, izz); n_com, mass_total
  // KV_COM only: P = com (3, ld_cfg), J = column-major (3*ld_cfg) x n_dof (or null)
  int sub_cfg;         // fk_twophase_kernel: configurations per tile (0 = TFK_TILE), twophase_tile_cfgs()
  int stage = 0;       // fk_twophase_kernel without SMEM_HOT: 1 = the launch reserved the per-warp staging area of
                       // twophase_stage_reals() and phase B reads [base | axes | positions] through it
  const void *weights; // device: n_com masses then n_act forces/total_mass, in the kernel's real type
  int n_com, n_act;
  double mass_total, vfs, com_diag;
};

// Dynamic shared memory of a block: [per-output tables | q tile (doubles as the warps' staging strips) | records]
struct SmemLayout {
  int plan_bytes, q_pitch, q_off, rec_off, total;
};
#if defined(__CUDACC__)
#define TFK_HDI __host__ __device__ inline
#else
#define TFK_HDI inline
#endif
// q tile pitch: odd (conflict-free 64-bit reads at one row per thread) and >= 7 so that a warp's own 32
// rows can double as its staging strip (32 x pose_dim reals) once the angles have been consumed
TFK_HDI int q_pitch_of(int n_in) { return (n_in < 7 ? 7 : n_in) | 1; }
// fused kernels (one requested link): [one staging strip of 32 x 7 reals per warp | records], no q tile
TFK_HDI int fused_strip_reals(int tile) { return (tile / 32) * (32 * 7); }
TFK_HDI int fused_smem_bytes(int tile, int rec_reals_in_smem, int real_bytes) {
  return (fused_strip_reals(tile) + rec_reals_in_smem * tile) * real_bytes;
}
TFK_HDI SmemLayout smem_layout_tile(int tile, int table_bytes, int n_in, int rec_reals_in_smem, int real_bytes) {
  SmemLayout L;
  L.plan_bytes = (table_bytes + 15) & ~15;
  L.q_pitch = q_pitch_of(n_in);
  L.q_off = L.plan_bytes;
  L.rec_off = L.q_off + ((tile * L.q_pitch * real_bytes + 15) & ~15);
  L.total = L.rec_off + rec_reals_in_smem * tile * real_bytes;
  return L;
}
TFK_HDI SmemLayout smem_layout(int table_bytes, int n_in, int rec_reals_in_smem, int real_bytes) {
  return smem_layout_tile(TFK_TILE, table_bytes, n_in, rec_reals_in_smem, real_bytes);
}

// Two-phase kernel whose joint records do not fit in shared memory for the whole tile (cfg5: 107 reals x 128 configurations):
// a warp of phase B works on 32 consecutive (configuration, link) items, i.e. on at most TWO configurations when at least
// 32 links are requested.  Their [base | axes | positions] records are copied from the block-private scratch into a
// warp-private shared-memory area once per 32 items -- all loads in flight together, one L2 round trip -- instead of one
// dependent round trip per Jacobian column and item.  Returns the area's size expressed in reals per thread of the block
// (what smem_layout() takes), 0 when the request has fewer than 32 links.
TFK_HDI int twophase_stage_reals(int n_out, int off_base, int off_frame) {
  if (n_out < 32) return 0;
  return ((TFK_TILE / 32) * 2 * (off_frame - off_base) + TFK_TILE - 1) / TFK_TILE;
}

// fk_twophase_kernel: configurations per tile.  When the batch needs more than one wave of resident blocks, the last
// wave is partly empty (cfg5 FP32: 512 tiles of 128 on 444 resident blocks = two rounds for 68 blocks, one for the
// rest); half tiles even that out at the price of idle threads in phase A, so they are used only when the busiest
// block gets at least 8 % fewer configurations.  (With a single wave -- cfg5 FP64: 512 tiles, 592 slots -- half tiles
// measured 3 % slower: the SMs are latency-bound and three or four co-resident blocks take about the same time.)
TFK_HDI int twophase_tile_cfgs(long long n, long long resident_blocks) {
  const long long full = (((n + TFK_TILE - 1) / TFK_TILE + resident_blocks - 1) / resident_blocks) * TFK_TILE;
  const int tc = TFK_TILE / 2;
  const long long half = (((n + tc - 1) / tc + resident_blocks - 1) / resident_blocks) * tc;
  return half * 108 < full * 100 ? tc : (int)TFK_TILE;
}

#if !defined(__CUDACC_RTC__)
cudaError_t launch_fk_f64_strict(const KernelArgs &a, const PlanParams &pp, int variant, int grid, int smem_bytes,
                                 cudaStream_t stream);
cudaError_t occupancy_fk_f64_strict(int rot_type, int variant, int smem_bytes, int *blocks_per_sm);
cudaError_t launch_fk_f32(const KernelArgs &a, const PlanParams &pp, int variant, int grid, int smem_bytes,
                          cudaStream_t stream);
cudaError_t occupancy_fk_f32(int rot_type, int variant, int smem_bytes, int *blocks_per_sm);
cudaError_t launch_fk_f64_fast(const KernelArgs &a, const PlanParams &pp, int variant, int grid, int smem_bytes,
                               cudaStream_t stream);
cudaError_t occupancy_fk_f64_fast(int rot_type, int variant, int smem_bytes, int *blocks_per_sm);
#endif

} // namespace tfk
#endif
