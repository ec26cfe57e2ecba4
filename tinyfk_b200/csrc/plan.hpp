// plan.hpp -- host flattening: model + call signature -> plan blob (fk_plan.h).
#ifndef TFK_PLAN_HPP
#define TFK_PLAN_HPP

#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#include "fk_plan.h"
#include "model.hpp"

namespace tfk {

struct PlanKey {
  std::vector<int> elink_ids, joint_ids;
  int rot_type, base_type, with_jacobian;
  uint64_t model_version;
  bool operator==(const PlanKey &o) const {
    return rot_type == o.rot_type && base_type == o.base_type && with_jacobian == o.with_jacobian &&
           model_version == o.model_version && elink_ids == o.elink_ids && joint_ids == o.joint_ids;
  }
};

// Builds the plan blob.  Throws std::invalid_argument on out-of-range ids or too deep branching.
std::vector<unsigned char> build_plan(const KinematicModel &model, const PlanKey &key);

// Kernel-parameter form of a plan blob.  Throws std::invalid_argument if the pruned tree or the column
// list exceeds what the parameter block holds (TFK_PSLOTS links, TFK_PCOLS joints).
std::unique_ptr<PlanParams> make_plan_params(const std::vector<unsigned char> &blob);

// Constants of solve_com_forward_kinematics (src/wrapper.cpp:126-179, src/kinematics.cpp:237-270), evaluated in the
// reference's own order: links = "_com" links in id order then the action links; w = masses then forces/total_mass.
struct ComTables {
  std::vector<int> links;
  std::vector<double> w;
  int n_com, n_act;
  double mass_total; // sum of the com links' masses (kinematics.cpp:242, 260)
  double vfs;        // 1 + sum of forces/total_mass (wrapper.cpp:153-157)
  double diag;       // d com / d base translation: ((sum m*1)/mass_total [+ sum f*1] [/ vfs])
};
// get_total_inertia_matrix (src/kinematics.cpp:272-317): per "_com" link (mass, ixx, ixy, ixz, iyy, iyz, izz)
struct InertiaTables {
  std::vector<int> links;
  std::vector<double> w; // 7 per link
  double mass_total;
};
InertiaTables build_inertia_tables(const KinematicModel &model);
ComTables build_com_tables(const KinematicModel &model, const int *action_link_ids, const double *forces, int n_action);

// CUDA source of a plan as compile-time constants (namespace tfk::spec: k_h, k_slots[], k_cols[], k_mask0[]):
// what the plan-specialised build of the fused kernel is compiled against (fk_kernels.cuh, TFK_SPEC_PLAN).
// Reals are printed as hexadecimal floating literals: the constants are the blob's, bit for bit.
std::string plan_to_source(const std::vector<unsigned char> &blob);

inline int pose_dim_of(int rot_type) { return 3 + (rot_type == ROT_RPY) * 3 + (rot_type == ROT_XYZW) * 4; }
inline int base_dof_of(int base_type) { return base_type == BASE_PLANER ? 3 : (base_type == BASE_FLOATING ? 6 : 0); }

} // namespace tfk
#endif
