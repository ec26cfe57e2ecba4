// model.hpp -- host-side kinematic model: URDF -> id tables, joint/base state, flat link table.
//
// Mirrors the public surface of tinyfk::KinematicModel (reference include/tinyfk.hpp:55-161,
// src/tinyfk.cpp:16-297) for everything the FK path needs; the numerics of the batch path live in
// plan.hpp (flattening) and fk_kernels_*.cu (CUDA).  Link and joint NUMBERING reproduces the
// reference exactly (src/tinyfk.cpp:23-64, 89-111):
//   link ids  = URDF link names in std::map order (bytewise lexicographic), then "<name>_com" for every
//               link with an <inertial> (in id order), then user links in add_new_link order;
//   joint ids = names of revolute / continuous / prismatic joints in std::map order.
// Invalid input raises std::invalid_argument / std::runtime_error where the reference would raise
// or run into undefined behaviour ("Errors").
#ifndef TFK_MODEL_HPP
#define TFK_MODEL_HPP

#include <cstdint>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

#include "fk_math.h"

namespace tfk {

// numeric values = urdf::Joint's enum (urdf_parser/include/urdf_model/joint.h:174-177)
enum JointType { JT_UNKNOWN = 0, JT_REVOLUTE = 1, JT_CONTINUOUS = 2, JT_PRISMATIC = 3, JT_FLOATING = 4, JT_PLANAR = 5, JT_FIXED = 6 };

enum RotationType { ROT_IGNORE = 0, ROT_RPY = 1, ROT_XYZW = 2 }; // include/tinyfk.hpp:53
enum BaseType { BASE_FIXED = 0, BASE_PLANER = 1, BASE_FLOATING = 2 }; // python/tinyfk/__init__.py:29-32

// Which libm evaluates the HOST-side constants (origin quaternions from rpy, stored base pose, local
// transforms of non-selected movable joints).  PORTABLE = fkmath.h, the same code the GPU runs, so
// that CPU constants + GPU arithmetic together equal the libm-pinned reference bit for bit.
// SYSTEM = the platform libm, as stock tinyfk would on this host.
enum HostLibm { LIBM_PORTABLE = 0, LIBM_SYSTEM = 1 };

typedef PoseT<double> Pose64;

struct LinkRec {
  std::string name;
  int parent;     // link id, -1 for the root
  int jtype;      // JointType of the parent joint (JT_FIXED for user/com links, 0 for the root)
  int jid;        // movable-joint id of the parent joint, -1 if fixed / root
  Vec3<double> axis;
  Pose64 origin;  // parent link -> joint frame (parent_to_joint_origin_transform)
  std::vector<int> children;
  bool has_inertial;
  double mass;
  double inertia[6]; // ixx ixy ixz iyy iyz izz (urdf::Inertial, link.h:60-75)
};

struct JointRec {
  std::string name;
  int type;
  int child_link;
  bool has_limits;
  double lower, upper, velocity, effort;
};

class KinematicModel {
public:
  explicit KinematicModel(const std::string &xml_text, HostLibm libm = LIBM_PORTABLE);

  // ---- ids / names (reference src/tinyfk.cpp:150-162, 203-215; include/tinyfk.hpp:114-130)
  std::vector<size_t> get_joint_ids(const std::vector<std::string> &names) const;
  std::vector<size_t> get_link_ids(const std::vector<std::string> &names) const;
  std::vector<std::string> get_joint_names() const;
  std::vector<std::string> get_link_names() const;
  const std::string &root_link_name() const { return links_[root_].name; }
  int root_link_id() const { return root_; }
  size_t n_links() const { return links_.size(); }
  size_t n_joints() const { return joints_.size(); }
  double total_mass() const { return total_mass_; }
  // ids of the "<name>_com" links in creation (= id) order: com_dummy_links_ of src/tinyfk.cpp:89-111
  const std::vector<int> &com_links() const { return com_links_; }

  // ---- limits (src/tinyfk.cpp:164-201): continuous joints -> (-inf, +inf)
  std::vector<std::pair<double, double>> get_joint_position_limits(const std::vector<size_t> &ids) const;
  std::vector<double> get_joint_velocity_limits(const std::vector<size_t> &ids) const;
  std::vector<double> get_joint_effort_limits(const std::vector<size_t> &ids) const;

  // ---- state (src/tinyfk.cpp:116-148; src/wrapper.cpp:25-55)
  void set_joint_angles(const std::vector<size_t> &ids, const std::vector<double> &angles);
  std::vector<double> get_joint_angles(const std::vector<size_t> &ids) const;
  void set_base_pose_xyzrpy(const double xyzrpy[6]);
  void get_base_pose_xyzrpy(double xyzrpy[6]) const;
  const Pose64 &base_pose() const { return base_pose_; }
  void set_init_angles();

  // ---- structure edits (src/tinyfk.cpp:217-258).  Returns the new link id.
  int add_new_link(const std::string &name, size_t parent_id, const double pos[3], const double rpy[3]);

  // joint's child link is an ancestor-or-self of link (include/tinyfk.hpp:31-46, src/tinyfk.cpp:260-283)
  bool is_relevant(size_t link_id, size_t joint_id) const;

  // ---- read access for the flattening
  const std::vector<LinkRec> &links() const { return links_; }
  const std::vector<JointRec> &joints() const { return joints_; }
  const std::vector<double> &joint_angles() const { return joint_angles_; }
  HostLibm libm() const { return libm_; }
  // bumps whenever structure or stored state changes; plan caches key on it
  uint64_t version() const { return version_; }
  // true when the constructor took the link / joint tables from the disk cache instead of reading the XML
  bool tables_from_cache() const { return tables_from_cache_; }

  // host-libm dispatch used by everything that builds constants
  Quat<double> quat_from_rpy(double r, double p, double y) const;
  Vec3<double> rpy_from_quat(const Quat<double> &q) const;
  // pose_transform(origin, Joint::transform(angle)) for a movable link at a given angle
  Pose64 local_transform(const LinkRec &l, double angle) const;

private:
  int add_link_pose(const std::string &name, size_t parent_id, const Pose64 &pose);
  size_t build_from_xml(const std::string &xml_text); // returns the number of URDF links (the "_com" links follow)
  std::vector<char> save_tables(size_t n_urdf_links) const;
  bool load_tables(const std::vector<char> &blob);

  std::vector<LinkRec> links_;
  std::vector<JointRec> joints_;
  std::unordered_map<std::string, int> link_ids_, joint_ids_;
  std::vector<double> joint_angles_;
  std::vector<int> com_links_;
  Pose64 base_pose_;
  int root_;
  double total_mass_;
  HostLibm libm_;
  uint64_t version_;
  bool tables_from_cache_;
};

std::string load_urdf(const std::string &path); // src/tinyfk.cpp:285-295

} // namespace tfk
#endif
