// fk_plan.h -- the flattened link tree ("plan") exactly as the kernels read it.  POD only; shared by
// the host flattening (plan.cpp, g++) and the device code (fk_kernels.cuh, nvcc).
//
// One plan is built per call signature (elink_ids, joint_ids, rot_type, base_type, with_jacobian) and
// model state version.  It replaces, for the batch path, the reference's per-config machinery:
//   - get_link_pose_inner's leaf->root walk + transform_stack_/transform_cache_
//     (reference src/kinematics.cpp:47-103, include/data_structure.hpp:5-61) becomes a PRUNED,
//     TOPOLOGICALLY SORTED slot array (only ancestors of requested links, DFS pre-order);
//   - RelevancePredicateTable lookups (include/tinyfk.hpp:31-46) become one bitmask per output link;
//   - joints that are not in joint_ids keep the model's stored angle (src/tinyfk.cpp:122-128 leaves
//     them untouched), so their local transform is a host-evaluated constant (kind SLOT_FIXED).
//
// SPARSITY CLASSES.  URDF origins are mostly "rpy = 0 0 0", translations mostly along one axis and
// joint axes mostly +-e_k.  The host classifies every slot once (rot / trn / axis class below) and the
// kernel evaluates only the terms of the reference's expressions whose constant operand is not an
// exact zero.  Dropping `x * 0.0` products and `+ 0.0` terms never changes a non-zero bit of an IEEE
// result (it can only flip the sign of an exact zero; DESIGN.md "signed zeros"), so the strict FP64
// mode stays bit-identical to the reference while doing less than half of its arithmetic.
//
// The plan is a single byte blob: PlanHeader, then PlanSlot[n_slots], int lrec_of_out[n_out],
// uint32 mask[n_out][mask_words], PlanCol[n_cols].  The kernel copies it to shared memory once per
// block and reads it with warp-uniform (broadcast) loads.
#ifndef TFK_FK_PLAN_H
#define TFK_FK_PLAN_H

#if defined(__CUDACC_RTC__)
typedef signed int int32_t;
typedef unsigned int uint32_t;
typedef long long int64_t;
typedef unsigned long long uint64_t;
#else
#include <stdint.h>
#endif

namespace tfk {

enum { SLOT_FIXED = 0, SLOT_REVOLUTE = 1, SLOT_PRISMATIC = 2 };
enum { FRAME_RUNNING = -1, FRAME_BASE = -2 };
enum { TFK_MAX_FRAMES = 64 };     // live branch-point frames per config, kept in the per-config records
#ifndef TFK_TILE_CFG
#define TFK_TILE_CFG 128
#endif
enum { TFK_TILE = TFK_TILE_CFG }; // configs per block tile == threads per block
enum { TFK_BASE_REC = 26 };       // base pose 7 + inverse 7 + three tweaked rotations 3*4

// how the slot's local rotation (parent link -> this link) arises
enum {
  ROT_NONE = 0,    // exactly the identity quaternion: q = parent.q
  ROT_CONST = 1,   // a constant dense quaternion (oquat)
  ROT_REV = 2,     // revolute, origin rotation is the identity: local.q = (axis*sin, cos)
  ROT_REV_OQ = 3,  // revolute behind a rotated origin: local.q = oquat * (axis*sin, cos)
};
// class of a 3-vector constant: which components are not exact zeros
enum { V_NONE = 0, V_X = 1, V_Y = 2, V_Z = 3, V_FULL = 4 };
// slot flags
enum {
  SF_INV_BEFORE = 1,  // GetInverse(parent.q) is not in the running `inv` yet and the translation needs it
  SF_INV_AFTER = 2,   // GetInverse(cur.q) is needed after the slot (joint record) and is not valid yet
  SF_PRISM_OQ = 4,    // prismatic behind a rotated origin: local.p = oquat * (axis*d) * oinv + opos
  SF_LEAF = 8,        // fixed link without needed children (sphere centre, tool frame, "_com" link) of a several-link
                      // plan: its pose is computed on the side and the running frame stays its parent's, so siblings
                      // need no branch-point frame (same arithmetic per link; only the bookkeeping differs)
};

struct PlanSlot {
  double opos[3];  // parent link -> this link when the joint is at its (constant) angle, or the joint origin
  double oquat[4]; // x y z w
  double axis[3];  // joint axis in the joint frame (NOT normalised, as the reference)
  double oinv[4];  // GetInverse(oquat) evaluated on the host (only read with SF_PRISM_OQ)
  int32_t kind;    // SLOT_*
  int32_t qcol;    // column of the input row that drives this joint (selected joints only)
  int32_t frame_ld; // where the parent frame comes from: FRAME_RUNNING, FRAME_BASE or a frame-file index
  int32_t frame_st; // frame-file index to keep this link's frame in (it has >1 needed child), or -1
  int32_t arec;    // world-axis record to emit after this slot (3 reals), or -1
  int32_t prec;    // position record to emit after this slot (3 reals, shared by co-located joints), or -1
  int32_t lrec;    // link record to emit after this slot (position [+ quaternion]), or -1
  int32_t rot;     // ROT_*
  int32_t trn;     // V_* class of the local translation (opos [+ axis*d for a prismatic joint])
  int32_t axc;     // V_* class of the joint axis
  int32_t flags;   // SF_*
  int32_t pad_;
};

struct PlanCol {
  int32_t arec;      // axis record feeding Jacobian column c, -1 if the joint moves none of the links
  int32_t prec;      // position record of that joint
  int32_t prismatic; // 1: dpos = axis, rotation rows zero (src/kinematics.cpp:149-151,161-162)
  int32_t pad_;
};

struct PlanHeader {
  int32_t n_slots, n_out, n_cols, n_arec, n_prec, n_lrec;
  int32_t rot_type, base_type, pose_dim, with_jacobian;
  int32_t n_in;       // reals per input row: n_cols (+3 PLANER, +6 FLOATING)
  int32_t n_out_cols; // Jacobian columns: same count (PLANER keeps x, y, yaw: __init__.py:154-158)
  int32_t lrec_size;  // 3 (IGNORE) or 7
  // per-config record layout, in reals.  Two-phase kernel: [base 7|26 | arec | prec | frames | lrec];
  // fused single-link kernel: [base pose 7 | arec | prec | frames] (fz_*).  has_base_ctx: the base record
  // carries the inverse pose and tweaked rotations (Jacobian with a moving base).
  int32_t off_base, off_arec, off_prec, off_lrec, off_frame, rec_size;
  int32_t fz_arec, fz_prec, fz_frame, fz_size;
  int32_t root_lrec;  // link record that receives the base pose (root link requested), or -1
  int32_t mask_words;
  int32_t off_slots, off_outs, off_masks, off_cols; // byte offsets into the blob
  int32_t total_bytes;
  int32_t n_frames;
  int32_t pad_;
  double base_pose[7]; // stored base pose used when base_type == FIXED (include/tinyfk.hpp:72)
};

// What the kernels read on their hot path, passed BY VALUE as a __grid_constant__ kernel parameter: it
// lives in the constant bank, so slot fields arrive through the uniform datapath (no shared-memory
// loads, branches on them are uniform branches) and FP64 instructions take the constants as operands.
// The per-output tables (lrec_of_out, masks) of the two-phase kernel are indexed per lane and stay in
// shared memory; mask0 is the relevance mask of the fused kernel's only output.
enum { TFK_PSLOTS = 176, TFK_PCOLS = 64 };
struct PlanParams {
  PlanHeader h;
  PlanSlot slots[TFK_PSLOTS];
  PlanCol cols[TFK_PCOLS];
  uint32_t mask0[2];
  uint32_t pad_[2];
};

} // namespace tfk
#endif
