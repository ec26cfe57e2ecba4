// wrapper.cpp -- pybind11 module `_tinyfk`: the reference's Python-visible class RobotModel
// (reference src/wrapper.cpp:246-283), re-exported on top of the C ABI (include/tinyfk_b200.h).
// numpy in / numpy out, no PyTorch.  Method names, argument order and error types follow the
// reference; everything numerical happens behind tfk_solve_fk_host (CUDA, no CPU fallback).
//
// SURVEY.md section 8(f) rows built on the same path: compute_inter_link_squared_dists (rank 1),
// solve_com_forward_kinematics (rank 2), compute_total_inertia_matrix (rank 4).
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <array>
#include <cstring>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "../../include/tinyfk_b200.h"

namespace py = pybind11;

namespace {

enum class RotationType { IGNORE = 0, RPY = 1, XYZW = 2 };

[[noreturn]] void raise_for(int rc) {
  const std::string msg = tfk_last_error();
  if (rc == TFK_ERR_INVALID) throw std::invalid_argument(msg); // -> ValueError, as the reference
  throw std::runtime_error(msg);                               // -> RuntimeError
}
inline void check(int rc) {
  if (rc != TFK_OK) raise_for(rc);
}

std::vector<int> to_int(const std::vector<size_t> &v) { return std::vector<int>(v.begin(), v.end()); }

// numpy array backed by pooled pinned memory (device->host copies land in it at full PCIe speed)
template <class T> py::array_t<T, py::array::f_style> pinned_array(size_t rows, size_t cols, bool pinned) {
  const size_t bytes = rows * cols * sizeof(T);
  if (!pinned || bytes == 0) return py::array_t<T, py::array::f_style>({rows, cols});
  void *p = nullptr;
  check(tfk_host_alloc(bytes, &p));
  py::capsule owner(p, [](void *q) { tfk_host_free(q); });
  return py::array_t<T, py::array::f_style>({rows, cols}, {sizeof(T), sizeof(T) * rows}, static_cast<T *>(p), owner);
}

class RobotModel {
public:
  explicit RobotModel(const std::string &xml, int host_libm = TFK_LIBM_PORTABLE) : m_(nullptr) {
    check(tfk_model_create(xml.data(), xml.size(), host_libm, &m_));
  }
  ~RobotModel() { tfk_model_destroy(m_); }
  RobotModel(const RobotModel &) = delete;
  RobotModel &operator=(const RobotModel &) = delete;

  double total_mass() const { return tfk_total_mass(m_); }
  std::string get_root_link_name() const { return tfk_link_name(m_, tfk_root_link_id(m_)); }

  std::vector<std::string> get_joint_names() const {
    std::vector<std::string> v;
    for (int i = 0; i < tfk_num_joints(m_); ++i) v.push_back(tfk_joint_name(m_, i));
    return v;
  }
  std::vector<std::string> get_link_names() const {
    std::vector<std::string> v;
    for (int i = 0; i < tfk_num_links(m_); ++i) v.push_back(tfk_link_name(m_, i));
    return v;
  }
  std::vector<size_t> get_joint_ids(const std::vector<std::string> &names) const {
    std::vector<const char *> c;
    for (auto &s : names) c.push_back(s.c_str());
    std::vector<int> ids(names.size());
    check(tfk_get_joint_ids(m_, c.data(), (int)c.size(), ids.data()));
    return std::vector<size_t>(ids.begin(), ids.end());
  }
  std::vector<size_t> get_link_ids(const std::vector<std::string> &names) const {
    std::vector<const char *> c;
    for (auto &s : names) c.push_back(s.c_str());
    std::vector<int> ids(names.size());
    check(tfk_get_link_ids(m_, c.data(), (int)c.size(), ids.data()));
    return std::vector<size_t>(ids.begin(), ids.end());
  }
  std::vector<std::pair<double, double>> get_joint_limits(const std::vector<size_t> &ids) const {
    std::vector<int> i = to_int(ids);
    std::vector<double> lo(i.size()), hi(i.size());
    check(tfk_get_joint_limits(m_, i.data(), (int)i.size(), lo.data(), hi.data()));
    std::vector<std::pair<double, double>> out;
    for (size_t k = 0; k < i.size(); ++k) out.emplace_back(lo[k], hi[k]);
    return out;
  }
  std::vector<double> get_joint_angles(const std::vector<size_t> &ids) const {
    std::vector<int> i = to_int(ids);
    std::vector<double> a(i.size());
    check(tfk_get_joint_angles(m_, i.data(), (int)i.size(), a.data()));
    return a;
  }
  void set_joint_angles(const std::vector<size_t> &ids, const std::vector<double> &angles) {
    if (ids.size() != angles.size()) throw std::invalid_argument("joint id / angle count mismatch");
    std::vector<int> i = to_int(ids);
    check(tfk_set_joint_angles(m_, i.data(), (int)i.size(), angles.data()));
  }
  std::vector<double> get_base_pose() const {
    std::vector<double> b(6);
    check(tfk_get_base_pose(m_, b.data()));
    return b;
  }
  void set_base_pose(const std::vector<double> &xyzrpy) {
    if (xyzrpy.size() != 6) throw std::invalid_argument("argument must have size 6"); // src/wrapper.cpp:38-40
    check(tfk_set_base_pose(m_, xyzrpy.data()));
  }
  void add_new_link(const std::string &name, size_t parent_id, std::array<double, 3> pos, std::array<double, 3> rpy) {
    check(tfk_add_new_link(m_, name.c_str(), (int)parent_id, pos.data(), rpy.data(), nullptr));
  }
  bool is_relevant(size_t link_id, size_t joint_id) const {
    int r = 0;
    check(tfk_is_relevant(m_, (int)link_id, (int)joint_id, &r));
    return r != 0;
  }
  void clear_cache() {} // there is no transform cache to clear: every batch is computed from its inputs

  // The new batched overload: qs is an (N, n_dof) ndarray; base_type 0/1/2; returns (P, J) F-ordered
  // exactly like the reference.  J is all zeros (lazily allocated) when with_jacobian is false, as
  // src/wrapper.cpp:81-82 returns it.
  py::tuple solve_batch(py::object qs_in, const std::vector<size_t> &elink_ids, const std::vector<size_t> &joint_ids,
                        int rot_type, int base_type, bool with_jacobian, int precision, const std::vector<int> &devices,
                        bool pinned_output) {
    if (rot_type < 0 || rot_type > 2) throw std::invalid_argument("bad rot_type");
    if (base_type < 0 || base_type > 2) throw std::invalid_argument("bad base_type");
    const int pose_dim = 3 + (rot_type == 1) * 3 + (rot_type == 2) * 4;
    const size_t n_in = joint_ids.size() + (base_type == 1 ? 3 : (base_type == 2 ? 6 : 0));
    std::vector<int> l = to_int(elink_ids), j = to_int(joint_ids);
    if (precision == TFK_F32)
      return run<float>(py::array_t<float, py::array::c_style | py::array::forcecast>::ensure(qs_in), l, j, rot_type,
                        base_type, with_jacobian, precision, devices, pinned_output, pose_dim, n_in);
    return run<double>(py::array_t<double, py::array::c_style | py::array::forcecast>::ensure(qs_in), l, j, rot_type,
                       base_type, with_jacobian, precision, devices, pinned_output, pose_dim, n_in);
  }

  // The reference's 7-argument entry (src/wrapper.cpp:60-124): with_base means the 6-D floating base.
  py::tuple solve_forward_kinematics(py::object qs, const std::vector<size_t> &elink_ids,
                                     const std::vector<size_t> &joint_ids, RotationType rot_type, bool with_base,
                                     bool with_jacobian, bool /*use_cache: no cache exists*/) {
    py::array_t<double, py::array::c_style | py::array::forcecast> a =
        py::array_t<double, py::array::c_style | py::array::forcecast>::ensure(qs);
    if (!a) throw std::invalid_argument("joint_angles_sequence must be a sequence of float sequences");
    if (with_base && a.ndim() == 2 && (size_t)a.shape(1) != joint_ids.size() + 6)
      throw std::invalid_argument("dof mismatch!! Probably you forget base's (x, y, theta)"); // src/wrapper.cpp:74-78
    return solve_batch(a, elink_ids, joint_ids, (int)rot_type, with_base ? TFK_BASE_FLOATING : TFK_BASE_FIXED,
                       with_jacobian, TFK_F64_STRICT, {}, true);
  }

  // compute_inter_link_squared_dists (src/wrapper.cpp:189-237), batched: returns (sqdists [n*n_check],
  // grads (n*n_check, n_dof) F-ordered; a 1x1 zero matrix when !with_grads, as the reference)
  py::tuple sqdists_batch(py::object qs_in, const std::vector<size_t> &link_ids1, const std::vector<size_t> &link_ids2,
                          const std::vector<size_t> &joint_ids, int base_type, bool with_grads, int precision) {
    if (base_type < 0 || base_type > 2) throw std::invalid_argument("bad base_type");
    if (link_ids1.size() != link_ids2.size()) throw std::invalid_argument("link_ids1 and link_ids2 differ in length");
    const size_t n_in = joint_ids.size() + (base_type == 1 ? 3 : (base_type == 2 ? 6 : 0));
    std::vector<int> l1 = to_int(link_ids1), l2 = to_int(link_ids2), j = to_int(joint_ids);
    if (precision == TFK_F32)
      return run_sq<float>(py::array_t<float, py::array::c_style | py::array::forcecast>::ensure(qs_in), l1, l2, j, base_type,
                           with_grads, precision, n_in);
    return run_sq<double>(py::array_t<double, py::array::c_style | py::array::forcecast>::ensure(qs_in), l1, l2, j, base_type,
                          with_grads, precision, n_in);
  }
  // the reference's 7-argument entry: with_base means the 6-D floating base; use_cache has no meaning here
  py::tuple compute_inter_link_squared_dists(py::object qs, const std::vector<size_t> &link_ids1,
                                             const std::vector<size_t> &link_ids2, const std::vector<size_t> &joint_ids,
                                             bool with_base, bool with_grads, bool /*use_cache*/) {
    return sqdists_batch(qs, link_ids1, link_ids2, joint_ids, with_base ? TFK_BASE_FLOATING : TFK_BASE_FIXED, with_grads,
                         TFK_F64_STRICT);
  }

  // solve_com_forward_kinematics (src/wrapper.cpp:126-179), batched: returns (coms (n, 3), J (3n, n_dof) F-ordered,
  // all zeros when !with_jacobian as the reference).  base_type PLANER takes (x, y, yaw) rows and returns the six
  // base columns of the padded row, as KinematicModel.solve_com_fk does (python/tinyfk/__init__.py:180-195).
  py::tuple com_batch(py::object qs_in, const std::vector<size_t> &joint_ids, const std::vector<size_t> &action_link_ids,
                      const std::vector<double> &forces, int base_type, bool with_jacobian, int precision) {
    if (base_type < 0 || base_type > 2) throw std::invalid_argument("bad base_type");
    if (action_link_ids.size() != forces.size()) throw std::invalid_argument("action_link_ids and forces differ in length");
    const size_t n_in = joint_ids.size() + (base_type == 1 ? 3 : (base_type == 2 ? 6 : 0));
    std::vector<int> j = to_int(joint_ids), al = to_int(action_link_ids);
    if (precision == TFK_F32)
      return run_com<float>(py::array_t<float, py::array::c_style | py::array::forcecast>::ensure(qs_in), j, al, forces, base_type,
                            with_jacobian, precision, n_in);
    return run_com<double>(py::array_t<double, py::array::c_style | py::array::forcecast>::ensure(qs_in), j, al, forces, base_type,
                           with_jacobian, precision, n_in);
  }
  // the reference's 6-argument entry: with_base means the 6-D floating base
  py::tuple solve_com_forward_kinematics(py::object qs, const std::vector<size_t> &joint_ids,
                                         const std::vector<size_t> &action_link_ids, const std::vector<double> &forces,
                                         bool with_base, bool with_jacobian) {
    return com_batch(qs, joint_ids, action_link_ids, forces, with_base ? TFK_BASE_FLOATING : TFK_BASE_FIXED, with_jacobian,
                     TFK_F64_STRICT);
  }

  // compute_total_inertia_matrix (src/wrapper.cpp:181-187): one configuration -> 3x3; the batched form returns (n, 3, 3)
  py::array inertia_batch(py::object qs_in, const std::vector<size_t> &joint_ids, int precision) {
    std::vector<int> j = to_int(joint_ids);
    if (precision == TFK_F32)
      return run_inertia<float>(py::array_t<float, py::array::c_style | py::array::forcecast>::ensure(qs_in), j, precision);
    return run_inertia<double>(py::array_t<double, py::array::c_style | py::array::forcecast>::ensure(qs_in), j, precision);
  }
  py::array compute_total_inertia_matrix(const std::vector<double> &q, const std::vector<size_t> &joint_ids) {
    if (q.size() != joint_ids.size()) throw std::invalid_argument("q and joint_ids differ in length");
    py::array_t<double> qa(std::vector<py::ssize_t>{1, (py::ssize_t)q.size()});
    std::memcpy(qa.mutable_data(), q.data(), q.size() * sizeof(double));
    py::array r = inertia_batch(qa, joint_ids, TFK_F64_STRICT);
    return r.attr("reshape")(3, 3).cast<py::array>();
  }

  // tfk_specialize: compile the single-link kernel against this query's plan (include/tinyfk_b200.h)
  py::dict specialize(const std::vector<size_t> &elink_ids, const std::vector<size_t> &joint_ids, int rot_type, int base_type,
                      bool with_jacobian, int precision, bool wait, bool compile_only) {
    std::vector<int> l = to_int(elink_ids), j = to_int(joint_ids);
    int64_t info[6] = {0, 0, 0, 0, 0, 0};
    int rc;
    {
      py::gil_scoped_release nogil;
      rc = tfk_specialize(m_, l.data(), (int)l.size(), j.data(), (int)j.size(), rot_type, base_type, with_jacobian ? 1 : 0, precision,
                          (wait ? TFK_SPEC_WAIT : 0) | (compile_only ? TFK_SPEC_COMPILE_ONLY : 0), info);
    }
    check(rc);
    py::dict d;
    d["state"] = info[0];
    d["block"] = info[1];
    d["blocks_per_sm"] = info[2];
    d["smem_bytes"] = info[3];
    d["cubin_bytes"] = info[4];
    d["from_cache"] = info[5];
    return d;
  }
  void set_jit_mode(int mode) { check(tfk_set_jit_mode(m_, mode)); }
  int get_jit_mode() const { return tfk_get_jit_mode(m_); }

  int64_t kernel_launches() const { return tfk_kernel_launch_count(); }

private:
  template <class T>
  py::tuple run(py::array_t<T, py::array::c_style | py::array::forcecast> qs, const std::vector<int> &l,
                const std::vector<int> &j, int rot_type, int base_type, bool with_jacobian, int precision,
                const std::vector<int> &devices, bool pinned_output, int pose_dim, size_t n_in) {
    if (!qs) throw std::invalid_argument("qs must be convertible to a 2-D float array");
    if (qs.ndim() != 2) throw std::invalid_argument("qs must be 2-D (n_configs, n_dof)");
    if ((size_t)qs.shape(1) != n_in)
      throw std::invalid_argument("dof mismatch: expected " + std::to_string(n_in) + " columns, got " +
                                  std::to_string(qs.shape(1)));
    const size_t n = (size_t)qs.shape(0);
    const size_t rowsP = n * l.size();
    auto P = pinned_array<T>(rowsP, (size_t)pose_dim, pinned_output);
    py::array_t<T, py::array::f_style> J;
    void *Jp = nullptr;
    if (with_jacobian) {
      J = pinned_array<T>(rowsP * pose_dim, n_in, pinned_output);
      Jp = J.mutable_data();
    }
    int rc;
    {
      py::gil_scoped_release nogil;
      rc = tfk_solve_fk_host(m_, qs.data(), (int64_t)n, l.data(), (int)l.size(), j.data(), (int)j.size(), rot_type,
                             base_type, with_jacobian ? 1 : 0, precision, devices.empty() ? nullptr : devices.data(),
                             (int)devices.size(), P.mutable_data(), Jp);
    }
    check(rc);
    if (!with_jacobian) {
      // zeros of the full shape (calloc-backed: no pages are touched until someone reads them)
      py::object zeros = py::module_::import("numpy").attr("zeros")(
          py::make_tuple(rowsP * pose_dim, n_in), py::arg("dtype") = py::dtype::of<T>(), py::arg("order") = "F");
      return py::make_tuple(P, zeros);
    }
    return py::make_tuple(P, J);
  }

  template <class T>
  py::tuple run_sq(py::array_t<T, py::array::c_style | py::array::forcecast> qs, const std::vector<int> &l1,
                   const std::vector<int> &l2, const std::vector<int> &j, int base_type, bool with_grads, int precision,
                   size_t n_in) {
    if (!qs) throw std::invalid_argument("qs must be convertible to a 2-D float array");
    if (qs.ndim() != 2) throw std::invalid_argument("qs must be 2-D (n_configs, n_dof)");
    if ((size_t)qs.shape(1) != n_in)
      throw std::invalid_argument("dof mismatch: expected " + std::to_string(n_in) + " columns, got " +
                                  std::to_string(qs.shape(1)));
    const size_t n = (size_t)qs.shape(0), rows = n * l1.size();
    py::array_t<T> sq((py::ssize_t)rows);
    py::array_t<T, py::array::f_style> G(with_grads ? std::vector<py::ssize_t>{(py::ssize_t)rows, (py::ssize_t)n_in}
                                                    : std::vector<py::ssize_t>{1, 1});
    if (!with_grads) *G.mutable_data() = T(0);
    int rc;
    {
      py::gil_scoped_release nogil;
      rc = tfk_inter_link_sqdists_host(m_, qs.data(), (int64_t)n, l1.data(), l2.data(), (int)l1.size(), j.data(),
                                       (int)j.size(), base_type, with_grads ? 1 : 0, precision, sq.mutable_data(),
                                       with_grads ? G.mutable_data() : nullptr);
    }
    check(rc);
    return py::make_tuple(sq, G);
  }

  template <class T>
  py::tuple run_com(py::array_t<T, py::array::c_style | py::array::forcecast> qs, const std::vector<int> &j,
                    const std::vector<int> &al, const std::vector<double> &forces, int base_type, bool with_jacobian,
                    int precision, size_t n_in) {
    if (!qs) throw std::invalid_argument("qs must be convertible to a 2-D float array");
    if (qs.ndim() != 2) throw std::invalid_argument("qs must be 2-D (n_configs, n_dof)");
    if ((size_t)qs.shape(1) != n_in)
      throw std::invalid_argument("dof mismatch: expected " + std::to_string(n_in) + " columns, got " +
                                  std::to_string(qs.shape(1)));
    const size_t n = (size_t)qs.shape(0), n_dof = j.size() + (base_type == 0 ? 0 : 6);
    py::array_t<T> com(std::vector<py::ssize_t>{(py::ssize_t)n, 3});
    if (!with_jacobian) {
      int rc;
      {
        py::gil_scoped_release nogil;
        rc = tfk_solve_com_fk_host(m_, qs.data(), (int64_t)n, j.data(), (int)j.size(), al.data(), forces.data(), (int)al.size(),
                                   base_type, 0, precision, com.mutable_data(), nullptr);
      }
      check(rc);
      py::object zeros = py::module_::import("numpy").attr("zeros")(
          py::make_tuple(3 * n, n_dof), py::arg("dtype") = py::dtype::of<T>(), py::arg("order") = "F");
      return py::make_tuple(com, zeros);
    }
    py::array_t<T, py::array::f_style> J(std::vector<py::ssize_t>{(py::ssize_t)(3 * n), (py::ssize_t)n_dof});
    int rc;
    {
      py::gil_scoped_release nogil;
      rc = tfk_solve_com_fk_host(m_, qs.data(), (int64_t)n, j.data(), (int)j.size(), al.data(), forces.data(), (int)al.size(),
                                 base_type, 1, precision, com.mutable_data(), J.mutable_data());
    }
    check(rc);
    return py::make_tuple(com, J);
  }

  template <class T> py::array run_inertia(py::array_t<T, py::array::c_style | py::array::forcecast> qs, const std::vector<int> &j, int precision) {
    if (!qs) throw std::invalid_argument("qs must be convertible to a 2-D float array");
    if (qs.ndim() != 2) throw std::invalid_argument("qs must be 2-D (n_configs, n_joint)");
    if ((size_t)qs.shape(1) != j.size())
      throw std::invalid_argument("dof mismatch: expected " + std::to_string(j.size()) + " columns, got " + std::to_string(qs.shape(1)));
    const size_t n = (size_t)qs.shape(0);
    py::array_t<T> out(std::vector<py::ssize_t>{(py::ssize_t)n, 3, 3});
    int rc;
    {
      py::gil_scoped_release nogil;
      rc = tfk_total_inertia_matrix_host(m_, qs.data(), (int64_t)n, j.data(), (int)j.size(), precision, out.mutable_data());
    }
    check(rc);
    return out;
  }

  tfk_model *m_;
};

} // namespace

PYBIND11_MODULE(_tinyfk, m) {
  m.doc() = "B200-native batched forward kinematics behind tinyfk's RobotModel interface";

  py::enum_<RotationType>(m, "RotationType")
      .value("IGNORE", RotationType::IGNORE)
      .value("RPY", RotationType::RPY)
      .value("XYZW", RotationType::XYZW);

  py::class_<RobotModel>(m, "RobotModel")
      .def(py::init<const std::string &, int>(), py::arg("xml"), py::arg("host_libm") = (int)TFK_LIBM_PORTABLE)
      .def_property_readonly("total_mass_", &RobotModel::total_mass)
      .def("get_root_link_name", &RobotModel::get_root_link_name)
      .def("solve_forward_kinematics", &RobotModel::solve_forward_kinematics)
      .def("solve_forward_kinematics_batch", &RobotModel::solve_batch, py::arg("qs"), py::arg("elink_ids"),
           py::arg("joint_ids"), py::arg("rot_type"), py::arg("base_type"), py::arg("with_jacobian"),
           py::arg("precision") = (int)TFK_F64_STRICT, py::arg("devices") = std::vector<int>(),
           py::arg("pinned_output") = true)
      .def("get_joint_angles", &RobotModel::get_joint_angles)
      .def("set_joint_angles", &RobotModel::set_joint_angles)
      .def("solve_com_forward_kinematics", &RobotModel::solve_com_forward_kinematics)
      .def("solve_com_forward_kinematics_batch", &RobotModel::com_batch, py::arg("qs"), py::arg("joint_ids"),
           py::arg("action_link_ids"), py::arg("forces"), py::arg("base_type"), py::arg("with_jacobian"),
           py::arg("precision") = (int)TFK_F64_STRICT)
      .def("compute_total_inertia_matrix", &RobotModel::compute_total_inertia_matrix)
      .def("compute_total_inertia_matrix_batch", &RobotModel::inertia_batch, py::arg("qs"), py::arg("joint_ids"),
           py::arg("precision") = (int)TFK_F64_STRICT)
      .def("get_joint_names", &RobotModel::get_joint_names)
      .def("get_joint_ids", &RobotModel::get_joint_ids)
      .def("get_joint_limits", &RobotModel::get_joint_limits)
      .def("get_base_pose", &RobotModel::get_base_pose)
      .def("set_base_pose", &RobotModel::set_base_pose)
      .def("get_link_ids", &RobotModel::get_link_ids)
      .def("get_link_names", &RobotModel::get_link_names)
      .def("add_new_link", &RobotModel::add_new_link)
      .def("compute_inter_link_squared_dists", &RobotModel::compute_inter_link_squared_dists)
      .def("compute_inter_link_squared_dists_batch", &RobotModel::sqdists_batch, py::arg("qs"), py::arg("link_ids1"),
           py::arg("link_ids2"), py::arg("joint_ids"), py::arg("base_type"), py::arg("with_grads"),
           py::arg("precision") = (int)TFK_F64_STRICT)
      .def("specialize", &RobotModel::specialize, py::arg("elink_ids"), py::arg("joint_ids"), py::arg("rot_type"),
           py::arg("base_type"), py::arg("with_jacobian"), py::arg("precision") = (int)TFK_F64_STRICT, py::arg("wait") = true,
           py::arg("compile_only") = false)
      .def("set_jit_mode", &RobotModel::set_jit_mode)
      .def("get_jit_mode", &RobotModel::get_jit_mode)
      .def("is_relevant", &RobotModel::is_relevant)
      .def("kernel_launches", &RobotModel::kernel_launches)
      .def("clear_cache", &RobotModel::clear_cache);
}
