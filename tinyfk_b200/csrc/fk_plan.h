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
// The plan is a single byte blob: PlanHeader, then PlanSlot[n_slots], int lrec_of_out[n_out],
// uint32 mask[n_out][mask_words], PlanCol[n_cols].  The kernel copies it to shared memory once per
// block and reads it with warp-uniform (broadcast) loads.
#ifndef TFK_FK_PLAN_H
#define TFK_FK_PLAN_H

#include <stdint.h>

namespace tfk {

enum { SLOT_FIXED = 0, SLOT_REVOLUTE = 1, SLOT_PRISMATIC = 2 };
enum { FRAME_RUNNING = -1, FRAME_BASE = -2 };
enum { TFK_MAX_FRAMES = 12 };     // live branch-point frames per config (PR2 both arms + head + grippers: 4)
enum { TFK_TILE = 128 };          // configs per block tile == threads per block
enum { TFK_BASE_REC = 26 };       // base pose 7 + inverse 7 + three tweaked rotations 3*4

struct PlanSlot {
  double opos[3];  // parent link -> this link when the joint is at its (constant) angle, or the joint origin
  double oquat[4]; // x y z w
  double axis[3];  // joint axis in the joint frame (NOT normalised, as the reference)
  int32_t kind;    // SLOT_*
  int32_t qcol;    // column of the input row that drives this joint (selected joints only)
  int32_t frame_ld; // where the parent frame comes from: FRAME_RUNNING, FRAME_BASE or a frame-file index
  int32_t frame_st; // frame-file index to keep this link's frame in (it has >1 needed child), or -1
  int32_t jrec;    // joint record to emit after this slot (world axis + position), or -1
  int32_t lrec;    // link record to emit after this slot (position [+ quaternion]), or -1
};

struct PlanCol {
  int32_t jrec;      // joint record feeding Jacobian column c, -1 if the joint moves none of the links
  int32_t prismatic; // 1: dpos = axis, rotation rows zero (src/kinematics.cpp:149-151,161-162)
};

struct PlanHeader {
  int32_t n_slots, n_out, n_cols, n_jrec, n_lrec;
  int32_t rot_type, base_type, pose_dim, with_jacobian;
  int32_t n_in;       // reals per input row: n_cols (+3 PLANER, +6 FLOATING)
  int32_t n_out_cols; // Jacobian columns: same count (PLANER keeps x, y, yaw: __init__.py:154-158)
  int32_t lrec_size;  // 3 (IGNORE) or 7
  int32_t off_base, off_jrec, off_lrec, rec_size; // per-config scratch record layout, in reals
  int32_t root_lrec;  // link record that receives the base pose (root link requested), or -1
  int32_t mask_words;
  int32_t off_slots, off_outs, off_masks, off_cols; // byte offsets into the blob
  int32_t total_bytes;
  int32_t n_frames;
  double base_pose[7]; // stored base pose used when base_type == FIXED (include/tinyfk.hpp:72)
};

} // namespace tfk
#endif
