/*
 * ref_driver.cpp -- TEST INFRASTRUCTURE (oracle/_ref).  NOT part of the product path.
 *
 * A plain-C handle API around the UNMODIFIED reference core (src/tinyfk.cpp, src/kinematics.cpp,
 * urdf_parser/, compiled where they lie under /root/reference by oracle/Makefile into
 * oracle/_ref/).  Only the batch loop and output layout of the pybind11 wrapper and the PLANER
 * glue of the Python facade are restated here, because those two files need real Eigen/pybind11:
 *   - loop + layout ............ reference src/wrapper.cpp:60-124
 *   - set_base_pose(xyzrpy) .... reference src/wrapper.cpp:25-43
 *   - PLANER pad / column pick . reference python/tinyfk/__init__.py:98-103, 154-158
 * Everything arithmetic (get_link_pose, get_jacobian, pose math, libm calls) is the reference's
 * own object code.
 *
 * Two shared objects are produced from this file:
 *   libtinyfk_ref.so         stock: libm = the host's glibc            (cpu_baseline "reference")
 *   libtinyfk_ref_strict.so  libm pinned to tinyfk_b200/csrc/fkmath.h  (bit-level parity oracle)
 */
#include "tinyfk.hpp"

#include <cmath>
#include <cstring>
#include <string>
#include <vector>

using tinyfk::KinematicModel;
using tinyfk::RotationType;
using tinyfk::Transform;

namespace {
struct Handle {
  KinematicModel *kin;
  std::string err;
  std::vector<std::string> link_names, joint_names;
};
thread_local std::string g_err;
} // namespace

extern "C" {

const char *tfkref_last_error() { return g_err.c_str(); }

void *tfkref_create(const char *xml) {
  try {
    // reference would dereference a null model on a parse failure (src/tinyfk.cpp:27): probe first
    urdf::ModelInterfaceSharedPtr probe = urdf::parseURDF(std::string(xml));
    if (!probe) {
      g_err = "urdf parse failed";
      return nullptr;
    }
    Handle *h = new Handle;
    h->kin = new KinematicModel(std::string(xml));
    return h;
  } catch (const std::exception &e) {
    g_err = e.what();
    return nullptr;
  }
}

void tfkref_destroy(void *p) {
  Handle *h = static_cast<Handle *>(p);
  if (!h) return;
  delete h->kin;
  delete h;
}

int tfkref_num_links(void *p) { return (int)static_cast<Handle *>(p)->kin->links_.size(); }
int tfkref_num_joints(void *p) { return (int)static_cast<Handle *>(p)->kin->joints_.size(); }
int tfkref_root_link_id(void *p) { return (int)static_cast<Handle *>(p)->kin->root_link_id_; }
double tfkref_total_mass(void *p) { return static_cast<Handle *>(p)->kin->total_mass_; }

const char *tfkref_link_name(void *p, int id) {
  return static_cast<Handle *>(p)->kin->links_[id]->name.c_str();
}
const char *tfkref_joint_name(void *p, int id) {
  return static_cast<Handle *>(p)->kin->joints_[id]->name.c_str();
}

int tfkref_link_id(void *p, const char *name) {
  try {
    return (int)static_cast<Handle *>(p)->kin->get_link_ids({std::string(name)})[0];
  } catch (const std::exception &e) {
    g_err = e.what();
    return -1;
  }
}
int tfkref_joint_id(void *p, const char *name) {
  try {
    return (int)static_cast<Handle *>(p)->kin->get_joint_ids({std::string(name)})[0];
  } catch (const std::exception &e) {
    g_err = e.what();
    return -1;
  }
}

/* Per-link view of the parsed tree ("urdf_parser output"), used to validate the product's own
 * parser/flattening: parent link id (-1 root), joint type code of the parent joint
 * (urdf::Joint enum: 1 REVOLUTE 2 CONTINUOUS 3 PRISMATIC 6 FIXED), movable-joint id (-1 if fixed),
 * axis[3], origin position[3] + quaternion xyzw[4]. */
void tfkref_link_info(void *p, int id, int *parent, int *jtype, int *jid, double *axis, double *origin7) {
  KinematicModel *kin = static_cast<Handle *>(p)->kin;
  const urdf::LinkSharedPtr &l = kin->links_[id];
  urdf::LinkSharedPtr par = l->getParent();
  *parent = par ? (int)par->id : -1;
  *jtype = 0;
  *jid = -1;
  for (int k = 0; k < 3; ++k) axis[k] = 0.0;
  for (int k = 0; k < 7; ++k) origin7[k] = 0.0;
  origin7[6] = 1.0;
  if (par && l->parent_joint) {
    const urdf::JointSharedPtr &j = l->parent_joint;
    *jtype = (int)j->type;
    if (j->type != urdf::Joint::FIXED) *jid = (int)j->id;
    axis[0] = j->axis.x;
    axis[1] = j->axis.y;
    axis[2] = j->axis.z;
    const Transform &o = j->parent_to_joint_origin_transform;
    origin7[0] = o.position.x;
    origin7[1] = o.position.y;
    origin7[2] = o.position.z;
    origin7[3] = o.rotation.x;
    origin7[4] = o.rotation.y;
    origin7[5] = o.rotation.z;
    origin7[6] = o.rotation.w;
  }
}

void tfkref_joint_limits(void *p, int jid, double *lo, double *hi) {
  KinematicModel *kin = static_cast<Handle *>(p)->kin;
  auto b = kin->get_joint_position_limits({(size_t)jid});
  *lo = b[0].first;
  *hi = b[0].second;
}

int tfkref_is_relevant(void *p, int link_id, int joint_id) {
  return static_cast<Handle *>(p)->kin->rptable_.isRelevant(link_id, joint_id) ? 1 : 0;
}

int tfkref_add_link(void *p, const char *name, int parent_id, const double *pos, const double *rpy) {
  try {
    std::array<double, 3> a = {pos[0], pos[1], pos[2]}, b = {rpy[0], rpy[1], rpy[2]};
    static_cast<Handle *>(p)->kin->add_new_link(std::string(name), (size_t)parent_id, a, b);
    return 0;
  } catch (const std::exception &e) {
    g_err = e.what();
    return 1;
  }
}

void tfkref_set_joint_angles(void *p, const int *ids, int n, const double *angles) {
  std::vector<size_t> v(ids, ids + n);
  std::vector<double> a(angles, angles + n);
  static_cast<Handle *>(p)->kin->set_joint_angles(v, a);
}

void tfkref_get_joint_angles(void *p, const int *ids, int n, double *angles) {
  std::vector<size_t> v(ids, ids + n);
  auto a = static_cast<Handle *>(p)->kin->get_joint_angles(v);
  for (int i = 0; i < n; ++i) angles[i] = a[i];
}

static void set_base_xyzrpy(KinematicModel *kin, const double *b) { // wrapper.cpp:25-35
  Transform pose;
  pose.position.x = b[0];
  pose.position.y = b[1];
  pose.position.z = b[2];
  pose.rotation.setFromRPY(b[3], b[4], b[5]);
  kin->set_base_pose(pose);
}

void tfkref_set_base_pose(void *p, const double *xyzrpy) {
  set_base_xyzrpy(static_cast<Handle *>(p)->kin, xyzrpy);
}

void tfkref_get_base_pose(void *p, double *xyzrpy) { // wrapper.cpp:45-55
  KinematicModel *kin = static_cast<Handle *>(p)->kin;
  xyzrpy[0] = kin->base_pose_.position.x;
  xyzrpy[1] = kin->base_pose_.position.y;
  xyzrpy[2] = kin->base_pose_.position.z;
  auto rpy = kin->base_pose_.rotation.getRPY();
  xyzrpy[3] = rpy.x;
  xyzrpy[4] = rpy.y;
  xyzrpy[5] = rpy.z;
}

/*
 * solve_fk as the Python user sees it (facade + wrapper).
 *   qs        row-major n_wp x n_dof_in, n_dof_in = n_joint (+3 PLANER, +6 FLOATING)
 *   rot_type  0 IGNORE, 1 RPY, 2 XYZW      base_type 0 FIXED, 1 PLANER, 2 FLOATING
 *   P         column-major (n_wp*n_link) x pose_dim           (what numpy receives: F-order)
 *   J         column-major (n_wp*n_link*pose_dim) x n_dof_out, n_dof_out = n_joint (+3 / +6);
 *             all zeros when !with_jacobian (wrapper.cpp:81-82).  May be NULL if !with_jacobian.
 */
int tfkref_solve_fk(void *p, const double *qs, long n_wp, const int *elink_ids, int n_link,
                    const int *joint_ids, int n_joint, int rot_type, int base_type, int with_jacobian,
                    int use_cache, double *P, double *J) {
  try {
    KinematicModel *kin = static_cast<Handle *>(p)->kin;
    const RotationType rt = rot_type == 0 ? RotationType::IGNORE
                                          : (rot_type == 1 ? RotationType::RPY : RotationType::XYZW);
    const int pose_dim = 3 + (rot_type == 1) * 3 + (rot_type == 2) * 4;
    const bool with_base = base_type != 0;
    const int n_in = n_joint + (base_type == 1 ? 3 : (base_type == 2 ? 6 : 0));
    const int n_out = n_in;
    const int n_dof = n_joint + (with_base ? 6 : 0);
    std::vector<size_t> lids(elink_ids, elink_ids + n_link), jids(joint_ids, joint_ids + n_joint);
    const size_t ldP = (size_t)n_wp * n_link;
    const size_t ldJ = ldP * pose_dim;
    if (J) std::memset(J, 0, sizeof(double) * ldJ * n_out);
    std::vector<double> q(n_joint), b(6);
    for (long i = 0; i < n_wp; ++i) {
      const double *row = qs + (size_t)i * n_in;
      for (int k = 0; k < n_joint; ++k) q[k] = row[k];
      kin->_set_joint_angles(jids, q); // wrapper.cpp:85
      if (with_base) {                 // wrapper.cpp:86-89 (+ __init__.py:98-103 for PLANER)
        if (base_type == 1) {
          b[0] = row[n_joint];
          b[1] = row[n_joint + 1];
          b[2] = 0.0;
          b[3] = 0.0;
          b[4] = 0.0;
          b[5] = row[n_joint + 2];
        } else {
          for (int k = 0; k < 6; ++k) b[k] = row[n_joint + k];
        }
        set_base_xyzrpy(kin, b.data());
      }
      if (!use_cache) kin->clear_cache(); // wrapper.cpp:90-92
      for (int j = 0; j < n_link; ++j) {  // wrapper.cpp:94-113
        const size_t head = (size_t)i * n_link + j;
        Transform pose;
        kin->get_link_pose(lids[j], pose);
        P[0 * ldP + head] = pose.position.x;
        P[1 * ldP + head] = pose.position.y;
        P[2 * ldP + head] = pose.position.z;
        if (rot_type == 1) {
          urdf::Vector3 rpy = pose.rotation.getRPY();
          P[3 * ldP + head] = rpy.x;
          P[4 * ldP + head] = rpy.y;
          P[5 * ldP + head] = rpy.z;
        } else if (rot_type == 2) {
          P[3 * ldP + head] = pose.rotation.x;
          P[4 * ldP + head] = pose.rotation.y;
          P[5 * ldP + head] = pose.rotation.z;
          P[6 * ldP + head] = pose.rotation.w;
        }
      }
      if (with_jacobian && J) { // wrapper.cpp:114-121
        for (int j = 0; j < n_link; ++j) {
          const size_t head = ((size_t)i * n_link + j) * pose_dim;
          Eigen::MatrixXd jac = kin->get_jacobian(lids[j], jids, rt, with_base);
          for (int c = 0; c < n_out; ++c) {
            // __init__.py:154-158: PLANER keeps columns [joints, n, n+1, n+5]
            int src = c;
            if (base_type == 1 && c == n_joint + 2) src = n_joint + 5;
            for (int r = 0; r < pose_dim; ++r) J[(size_t)c * ldJ + head + r] = jac(r, src);
          }
        }
      }
    }
    (void)n_dof;
    return 0;
  } catch (const std::exception &e) {
    g_err = e.what();
    return 1;
  }
}

/*
 * compute_inter_link_sqdists as the Python user sees it (python/tinyfk/__init__.py:233-266 +
 * src/wrapper.cpp:189-237).  The link poses and Jacobians come from the reference's own get_link_pose /
 * get_jacobian; the last three lines of arithmetic are Eigen expressions in the reference
 * (diff_vec.squaredNorm(), 2 * (jac1 - jac2).transpose() * diff_vec) and Eigen is not available here, so
 * they are restated in the order written below -- Eigen's internal summation order is NOT pinned (parity
 * for this entry is a 1e-12 statement, not a bit statement).
 *   sq     n_wp*n_check values, head = i*n_check + j
 *   G      column-major (n_wp*n_check) x n_dof_out (n_joint (+3 PLANER: columns n, n+1, n+5 | +6)); may be NULL
 */
int tfkref_sqdists(void *p, const double *qs, long n_wp, const int *link_ids1, const int *link_ids2, int n_check,
                   const int *joint_ids, int n_joint, int base_type, int with_grads, int use_cache, double *sq,
                   double *G) {
  try {
    KinematicModel *kin = static_cast<Handle *>(p)->kin;
    const bool with_base = base_type != 0;
    const int n_in = n_joint + (base_type == 1 ? 3 : (base_type == 2 ? 6 : 0));
    const int n_out = n_in;
    std::vector<size_t> jids(joint_ids, joint_ids + n_joint);
    const size_t ld = (size_t)n_wp * n_check;
    std::vector<double> q(n_joint), b(6);
    for (long i = 0; i < n_wp; ++i) {
      const double *row = qs + (size_t)i * n_in;
      for (int k = 0; k < n_joint; ++k) q[k] = row[k];
      kin->_set_joint_angles(jids, q); // wrapper.cpp:205
      if (with_base) {                 // wrapper.cpp:206-209 (+ __init__.py:243-246 for PLANER)
        if (base_type == 1) {
          b[0] = row[n_joint];
          b[1] = row[n_joint + 1];
          b[2] = 0.0;
          b[3] = 0.0;
          b[4] = 0.0;
          b[5] = row[n_joint + 2];
        } else {
          for (int k = 0; k < 6; ++k) b[k] = row[n_joint + k];
        }
        set_base_xyzrpy(kin, b.data());
      }
      if (!use_cache) kin->clear_cache(); // wrapper.cpp:210-212
      for (int j = 0; j < n_check; ++j) {
        const size_t head = (size_t)i * n_check + j;
        Transform pose1, pose2;
        kin->get_link_pose(link_ids1[j], pose1);
        kin->get_link_pose(link_ids2[j], pose2);
        const urdf::Vector3 diff = pose1.position - pose2.position; // wrapper.cpp:219
        const double d[3] = {diff.x, diff.y, diff.z};
        sq[head] = (d[0] * d[0] + d[1] * d[1]) + d[2] * d[2]; // squaredNorm(), wrapper.cpp:222
        if (with_grads && G) {
          Eigen::MatrixXd jac1 = kin->get_jacobian(link_ids1[j], jids, RotationType::IGNORE, with_base);
          Eigen::MatrixXd jac2 = kin->get_jacobian(link_ids2[j], jids, RotationType::IGNORE, with_base);
          for (int c = 0; c < n_out; ++c) {
            int src = c; // __init__.py:259-264: PLANER keeps columns [joints, n, n+1, n+5]
            if (base_type == 1 && c == n_joint + 2) src = n_joint + 5;
            const double a0 = jac1(0, src) - jac2(0, src), a1 = jac1(1, src) - jac2(1, src),
                         a2 = jac1(2, src) - jac2(2, src);
            G[(size_t)c * ld + head] = 2.0 * ((a0 * d[0] + a1 * d[1]) + a2 * d[2]); // wrapper.cpp:230-232
          }
        }
      }
    }
    return 0;
  } catch (const std::exception &e) {
    g_err = e.what();
    return 1;
  }
}

/*
 * solve_com_fk as the Python user sees it (python/tinyfk/__init__.py:160-195 + src/wrapper.cpp:126-179).  The
 * reference's own get_com / get_com_jacobian / get_link_pose / get_jacobian do the work; this function restates the
 * loop of wrapper.cpp:136-178, whose Eigen expressions are element-wise only (block += force * jac, block /= s: no
 * reductions, so no summation order to pin).
 * ONE DELIBERATE DIFFERENCE: wrapper.cpp:137-141 calls _set_joint_angles (no cache clear) and only set_base_pose
 * clears the transform cache, so with a FIXED base the reference answers every row from the cache left by the first
 * one.  `clear_each` != 0 clears the cache per row (what a caller gets from one call per configuration after
 * clear_cache()); `clear_each` == 0 reproduces the reference literally.
 *   q rows [joints | x y yaw (PLANER, padded as __init__.py:98-103) | xyzrpy (FLOATING)]
 *   com (n_wp, 3) row-major;  J column-major (3*n_wp) x (n_joint + 0|6), may be NULL
 */
int tfkref_com(void *p, const double *qs, long n_wp, const int *joint_ids, int n_joint, const int *action_link_ids,
               const double *forces, int n_action, int base_type, int with_jacobian, int clear_each, double *com,
               double *J) {
  try {
    KinematicModel *kin = static_cast<Handle *>(p)->kin;
    const bool with_base = base_type != 0;
    const int n_in = n_joint + (base_type == 1 ? 3 : (base_type == 2 ? 6 : 0));
    const int n_dof = n_joint + (with_base ? 6 : 0);
    std::vector<size_t> jids(joint_ids, joint_ids + n_joint);
    const size_t ldJ = (size_t)3 * n_wp;
    std::vector<double> q(n_joint), b(6);
    for (long i = 0; i < n_wp; ++i) {
      const double *row = qs + (size_t)i * n_in;
      for (int k = 0; k < n_joint; ++k) q[k] = row[k];
      kin->_set_joint_angles(jids, q); // wrapper.cpp:137
      if (with_base) {                 // wrapper.cpp:138-141
        if (base_type == 1) {
          b[0] = row[n_joint];
          b[1] = row[n_joint + 1];
          b[2] = 0.0;
          b[3] = 0.0;
          b[4] = 0.0;
          b[5] = row[n_joint + 2];
        } else {
          for (int k = 0; k < 6; ++k) b[k] = row[n_joint + k];
        }
        set_base_xyzrpy(kin, b.data());
      }
      if (clear_each) kin->clear_cache();
      const urdf::Vector3 c0 = kin->get_com(); // wrapper.cpp:143
      double c[3] = {c0.x, c0.y, c0.z};
      Eigen::MatrixXd jblock = Eigen::MatrixXd::Zero(3, n_dof);
      if (with_jacobian) jblock = kin->get_com_jacobian(jids, with_base); // wrapper.cpp:147-150
      if (n_action > 0) {                                                  // wrapper.cpp:152-177
        double vertical_force_sum = 1.0;
        Transform pose;
        for (int j = 0; j < n_action; ++j) {
          const double force = forces[j] / kin->total_mass_;
          vertical_force_sum += force;
          kin->get_link_pose(action_link_ids[j], pose);
          c[0] += force * pose.position.x;
          c[1] += force * pose.position.y;
          c[2] += force * pose.position.z;
          if (with_jacobian) {
            Eigen::MatrixXd jac = kin->get_jacobian(action_link_ids[j], jids, RotationType::IGNORE, with_base);
            for (int cc = 0; cc < n_dof; ++cc)
              for (int r = 0; r < 3; ++r) jblock(r, cc) += force * jac(r, cc);
          }
        }
        c[0] /= vertical_force_sum;
        c[1] /= vertical_force_sum;
        c[2] /= vertical_force_sum;
        if (with_jacobian)
          for (int cc = 0; cc < n_dof; ++cc)
            for (int r = 0; r < 3; ++r) jblock(r, cc) /= vertical_force_sum;
      }
      for (int k = 0; k < 3; ++k) com[(size_t)i * 3 + k] = c[k];
      if (with_jacobian && J)
        for (int cc = 0; cc < n_dof; ++cc)
          for (int r = 0; r < 3; ++r) J[(size_t)cc * ldJ + (size_t)3 * i + r] = jblock(r, cc);
    }
    return 0;
  } catch (const std::exception &e) {
    g_err = e.what();
    return 1;
  }
}

/*
 * compute_total_inertia_matrix (src/wrapper.cpp:181-187): _set_joint_angles + the reference's own
 * get_total_inertia_matrix (src/kinematics.cpp:272-317), whose Eigen expressions run on oracle/eigen_stub (3x3 products
 * summed k = 0, 1, 2 from 0.0; Eigen's own order is NOT pinned).  clear != 0 clears the transform cache first (the
 * reference does not: see tfkref_com).  out: 9 doubles row-major per configuration.
 */
int tfkref_inertia(void *p, const double *qs, long n_wp, const int *joint_ids, int n_joint, int clear, double *out) {
  try {
    KinematicModel *kin = static_cast<Handle *>(p)->kin;
    std::vector<size_t> jids(joint_ids, joint_ids + n_joint);
    std::vector<double> q(n_joint);
    for (long i = 0; i < n_wp; ++i) {
      for (int k = 0; k < n_joint; ++k) q[k] = qs[(size_t)i * n_joint + k];
      kin->_set_joint_angles(jids, q);
      if (clear) kin->clear_cache();
      const Eigen::Matrix3d m = kin->get_total_inertia_matrix();
      for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) out[(size_t)i * 9 + 3 * r + c] = m(r, c);
    }
    return 0;
  } catch (const std::exception &e) {
    g_err = e.what();
    return 1;
  }
}

/* which libm is this object bound to?  (tests check the strict build really is pinned) */
double tfkref_probe_sin(double x) {
  volatile double v = x;
  return sin(v);
}
double tfkref_probe_atan2(double y, double x) {
  volatile double a = y, b = x;
  return atan2(a, b);
}

} // extern "C"
