// tinyfk_b200.hpp -- header-only C++ view of the C ABI (tinyfk_b200.h) with the shape of the reference's C++ class, for
// callers that use tinyfk from C++ rather than from Python (reference bench/bench_tinyfk.cpp:17-60 drives
// tinyfk::KinematicModel directly).  Replaces, for the FK path, include/tinyfk.hpp:55-161:
//
//   reference (tinyfk::KinematicModel)                         here (tinyfk_b200::KinematicModel)
//   KinematicModel(xml_string)               tinyfk.hpp:78     KinematicModel(xml_string)
//   set_joint_angles(ids, angles)            :82               set_joint_angles(ids, angles)
//   set_base_pose(Transform)                 :90               set_base_pose(xyz, rpy)  (wrapper.cpp:37-47 builds the Transform so)
//   get_joint_ids / get_link_ids             :103 / :122       get_joint_ids / get_link_ids
//   get_link_pose(link_id, Transform &)      :132              get_link_pose(link_id, Pose &)
//   get_jacobian(elink, ids, rot, with_base) :134              get_jacobian(elink, ids, rot, with_base) -> Matrix (column-major)
//   add_new_link(name, parent, pos, rpy)     :149              add_new_link(name, parent, pos, rpy)
// plus what the reference has no C++ entry for: solve_fk over N configurations (the wrapper's batch loop,
// src/wrapper.cpp:60-124, on the GPU).
//
// No Eigen, no urdfdom: poses are plain structs (position xyz, rotation as quaternion xyzw -- urdf::Pose's members), the
// Jacobian is a small column-major matrix.  The single-configuration calls evaluate the model's STORED state (as the
// reference does) by sending a one-row batch through tfk_solve_fk_host: one kernel launch per call -- they exist so that
// existing C++ call sites keep working; throughput comes from solve_fk().  get_jacobian(.., with_base = true) hands the
// stored base pose to the batch entry as (xyz, rpy) read back from the stored quaternion: equal to the stored pose to
// within rounding (1e-16), which the reference's eps = 1e-7 finite-difference base columns amplify to ~1e-9.
// Errors become exceptions (std::invalid_argument / std::runtime_error), as in the reference.  Link with -ltinyfk_b200.
#ifndef TINYFK_B200_HPP
#define TINYFK_B200_HPP

#include <array>
#include <stdexcept>
#include <string>
#include <vector>

#include "tinyfk_b200.h"

namespace tinyfk_b200 {

enum class RotationType { IGNORE = TFK_ROT_IGNORE, RPY = TFK_ROT_RPY, XYZW = TFK_ROT_XYZW }; // data_structure.hpp
enum class BaseType { FIXED = TFK_BASE_FIXED, PLANER = TFK_BASE_PLANER, FLOATING = TFK_BASE_FLOATING };

struct Pose { // urdf::Pose (urdf_model/pose.h): position + rotation quaternion
  std::array<double, 3> position;
  std::array<double, 4> rotation; // x, y, z, w
};

struct Matrix { // column-major, as Eigen::MatrixXd of the reference
  size_t rows = 0, cols = 0;
  std::vector<double> data;
  double operator()(size_t r, size_t c) const { return data[c * rows + r]; }
};

class KinematicModel {
public:
  explicit KinematicModel(const std::string &xml_string) {
    check(tfk_model_create(xml_string.data(), xml_string.size(), TFK_LIBM_PORTABLE, &h_));
  }
  ~KinematicModel() { tfk_model_destroy(h_); }
  KinematicModel(const KinematicModel &) = delete;
  KinematicModel &operator=(const KinematicModel &) = delete;

  std::vector<size_t> get_joint_ids(const std::vector<std::string> &names) const { return ids_of(names, tfk_get_joint_ids); }
  std::vector<size_t> get_link_ids(const std::vector<std::string> &names) const { return ids_of(names, tfk_get_link_ids); }

  void set_joint_angles(const std::vector<size_t> &joint_ids, const std::vector<double> &angles) {
    if (joint_ids.size() != angles.size()) throw std::invalid_argument("set_joint_angles: size mismatch");
    const std::vector<int> ids = ints(joint_ids);
    check(tfk_set_joint_angles(h_, ids.data(), (int)ids.size(), angles.data()));
  }
  std::vector<double> get_joint_angles(const std::vector<size_t> &joint_ids) const {
    const std::vector<int> ids = ints(joint_ids);
    std::vector<double> a(ids.size());
    check(tfk_get_joint_angles(h_, ids.data(), (int)ids.size(), a.data()));
    return a;
  }
  void set_base_pose(const std::array<double, 3> &xyz, const std::array<double, 3> &rpy) {
    const double b[6] = {xyz[0], xyz[1], xyz[2], rpy[0], rpy[1], rpy[2]};
    check(tfk_set_base_pose(h_, b));
  }

  size_t add_new_link(const std::string &name, size_t parent_id, const std::array<double, 3> &position,
                      const std::array<double, 3> &rpy) {
    int id = -1;
    check(tfk_add_new_link(h_, name.c_str(), (int)parent_id, position.data(), rpy.data(), &id));
    return (size_t)id;
  }

  // pose of a link in the root frame at the stored joint angles and base pose
  void get_link_pose(size_t link_id, Pose &out) const {
    const int lid = (int)link_id;
    double P[7];
    double q_dummy = 0.0; // a 1 x 0 batch: every joint takes its stored angle, the base its stored pose (BaseType FIXED)
    check(tfk_solve_fk_host(h_, &q_dummy, 1, &lid, 1, nullptr, 0, TFK_ROT_XYZW, TFK_BASE_FIXED, 0, TFK_F64_STRICT, nullptr, 0, P,
                            nullptr));
    out.position = {P[0], P[1], P[2]};
    out.rotation = {P[3], P[4], P[5], P[6]};
  }

  // (3 | 6 | 7) x (n_joint [+ 6]) Jacobian of a link at the stored state, src/kinematics.cpp:105-235
  Matrix get_jacobian(size_t elink_id, const std::vector<size_t> &joint_ids, RotationType rot_type = RotationType::IGNORE,
                      bool with_base = false) {
    const int lid = (int)elink_id;
    const std::vector<int> ids = ints(joint_ids);
    std::vector<double> q = get_joint_angles(joint_ids);
    if (with_base) {
      double b[6];
      check(tfk_get_base_pose(h_, b));
      q.insert(q.end(), b, b + 6);
    }
    if (q.empty()) q.push_back(0.0);
    const size_t pd = rot_type == RotationType::IGNORE ? 3 : (rot_type == RotationType::RPY ? 6 : 7);
    Matrix J;
    J.rows = pd;
    J.cols = ids.size() + (with_base ? 6 : 0);
    J.data.assign(J.rows * J.cols, 0.0);
    double P[7];
    double unused = 0.0;
    check(tfk_solve_fk_host(h_, q.data(), 1, &lid, 1, ids.data(), (int)ids.size(), (int)rot_type,
                            with_base ? TFK_BASE_FLOATING : TFK_BASE_FIXED, 1, TFK_F64_STRICT, nullptr, 0, P,
                            J.data.empty() ? &unused : J.data.data()));
    return J;
  }

  // N configurations at once: qs is N x (n_joint + 0 | 3 | 6) row-major; P is (N * n_link) x pose_dim and J is
  // (N * n_link * pose_dim) x n_dof, both column-major (the layout the reference's wrapper returns, src/wrapper.cpp:80-123)
  void solve_fk(const std::vector<double> &qs, size_t n, const std::vector<size_t> &elink_ids, const std::vector<size_t> &joint_ids,
                RotationType rot_type, BaseType base_type, bool with_jacobian, Matrix &P, Matrix &J,
                const std::vector<int> &devices = std::vector<int>()) {
    const std::vector<int> lids = ints(elink_ids), jids = ints(joint_ids);
    const size_t n_dof = jids.size() + (base_type == BaseType::FIXED ? 0 : (base_type == BaseType::PLANER ? 3 : 6));
    if (qs.size() != n * n_dof) throw std::invalid_argument("solve_fk: qs must hold n x n_dof values");
    const size_t pd = rot_type == RotationType::IGNORE ? 3 : (rot_type == RotationType::RPY ? 6 : 7);
    P.rows = n * lids.size(), P.cols = pd;
    P.data.assign(P.rows * P.cols, 0.0);
    J.rows = n * lids.size() * pd, J.cols = n_dof;
    J.data.assign(with_jacobian ? J.rows * J.cols : 0, 0.0);
    double unused = 0.0;
    check(tfk_solve_fk_host(h_, qs.empty() ? &unused : qs.data(), (int64_t)n, lids.data(), (int)lids.size(), jids.data(),
                            (int)jids.size(), (int)rot_type, (int)base_type, with_jacobian ? 1 : 0, TFK_F64_STRICT,
                            devices.empty() ? nullptr : devices.data(), (int)devices.size(), P.data.empty() ? &unused : P.data.data(),
                            with_jacobian && !J.data.empty() ? J.data.data() : nullptr));
  }

  tfk_model *handle() const { return h_; }

private:
  tfk_model *h_ = nullptr;

  static void check(int rc) {
    if (rc == TFK_OK) return;
    const std::string msg = tfk_last_error();
    if (rc == TFK_ERR_INVALID) throw std::invalid_argument(msg);
    throw std::runtime_error(msg);
  }
  static std::vector<int> ints(const std::vector<size_t> &v) { return std::vector<int>(v.begin(), v.end()); }
  template <class F> std::vector<size_t> ids_of(const std::vector<std::string> &names, F f) const {
    std::vector<const char *> p;
    for (const std::string &s : names) p.push_back(s.c_str());
    std::vector<int> ids(names.size());
    check(f(h_, p.data(), (int)p.size(), ids.data()));
    return std::vector<size_t>(ids.begin(), ids.end());
  }
};

} // namespace tinyfk_b200
#endif
