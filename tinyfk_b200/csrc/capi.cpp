// capi.cpp -- extern "C" surface declared in include/tinyfk_b200.h.  Translates C++ exceptions into
// error codes; holds the per-handle lock.
#include "../../include/tinyfk_b200.h"

#include <cstring>
#include <string>
#include <vector>

#include "engine.hpp"
#include "fk_noalloc.h"

using namespace tfk;

struct tfk_model {
  Engine engine;
  tfk_model(const std::string &xml, HostLibm libm) : engine(xml, libm) {}
};

namespace {
thread_local std::string g_err;

template <class F> int guarded(F &&f) {
  try {
    f();
    return TFK_OK;
  } catch (const CudaError &e) {
    g_err = e.what();
    return TFK_ERR_CUDA;
  } catch (const std::invalid_argument &e) {
    g_err = e.what();
    return TFK_ERR_INVALID;
  } catch (const std::exception &e) {
    g_err = e.what();
    return TFK_ERR_RUNTIME;
  } catch (...) {
    g_err = "unknown error";
    return TFK_ERR_RUNTIME;
  }
}

std::vector<size_t> ids_of(const int *ids, int n, size_t limit, const char *what) {
  if (n < 0 || (n > 0 && !ids)) throw std::invalid_argument(std::string("bad ") + what + " id array");
  std::vector<size_t> v((size_t)n);
  for (int i = 0; i < n; ++i) {
    if (ids[i] < 0 || (size_t)ids[i] >= limit)
      throw std::invalid_argument(std::string(what) + " id " + std::to_string(ids[i]) + " out of range");
    v[i] = (size_t)ids[i];
  }
  return v;
}

void need(const void *p, const char *what) {
  if (!p) throw std::invalid_argument(std::string(what) + " is null");
}
} // namespace

extern "C" {

const char *tfk_last_error(void) { return g_err.c_str(); }
const char *tfk_version(void) { return "tinyfk_b200 0.1 (sm_100a)"; }

int tfk_model_create(const char *xml, size_t xml_len, int host_libm, tfk_model **out) {
  return guarded([&] {
    need(out, "out");
    *out = nullptr;
    if (!xml) xml_len = 0;
    if (host_libm != TFK_LIBM_PORTABLE && host_libm != TFK_LIBM_SYSTEM) throw std::invalid_argument("bad host_libm");
    *out = new tfk_model(std::string(xml ? xml : "", xml_len), host_libm == TFK_LIBM_SYSTEM ? LIBM_SYSTEM : LIBM_PORTABLE);
  });
}

void tfk_model_destroy(tfk_model *m) { delete m; }

int tfk_num_links(const tfk_model *m) { return m ? (int)m->engine.model.n_links() : 0; }
int tfk_num_joints(const tfk_model *m) { return m ? (int)m->engine.model.n_joints() : 0; }
int tfk_root_link_id(const tfk_model *m) { return m ? m->engine.model.root_link_id() : -1; }
double tfk_total_mass(const tfk_model *m) { return m ? m->engine.model.total_mass() : 0.0; }

const char *tfk_link_name(const tfk_model *m, int id) {
  if (!m || id < 0 || (size_t)id >= m->engine.model.n_links()) return nullptr;
  return m->engine.model.links()[id].name.c_str();
}
const char *tfk_joint_name(const tfk_model *m, int id) {
  if (!m || id < 0 || (size_t)id >= m->engine.model.n_joints()) return nullptr;
  return m->engine.model.joints()[id].name.c_str();
}

int tfk_get_link_ids(const tfk_model *m, const char *const *names, int n, int *out) {
  return guarded([&] {
    need(m, "model");
    std::vector<std::string> v;
    for (int i = 0; i < n; ++i) v.push_back(names[i] ? names[i] : "");
    std::vector<size_t> ids = m->engine.model.get_link_ids(v);
    for (int i = 0; i < n; ++i) out[i] = (int)ids[i];
  });
}
int tfk_get_joint_ids(const tfk_model *m, const char *const *names, int n, int *out) {
  return guarded([&] {
    need(m, "model");
    std::vector<std::string> v;
    for (int i = 0; i < n; ++i) v.push_back(names[i] ? names[i] : "");
    std::vector<size_t> ids = m->engine.model.get_joint_ids(v);
    for (int i = 0; i < n; ++i) out[i] = (int)ids[i];
  });
}

int tfk_get_joint_limits(const tfk_model *m, const int *joint_ids, int n, double *lower, double *upper) {
  return guarded([&] {
    need(m, "model");
    auto lim = m->engine.model.get_joint_position_limits(ids_of(joint_ids, n, m->engine.model.n_joints(), "joint"));
    for (int i = 0; i < n; ++i) {
      lower[i] = lim[i].first;
      upper[i] = lim[i].second;
    }
  });
}
int tfk_get_joint_velocity_limits(const tfk_model *m, const int *joint_ids, int n, double *out) {
  return guarded([&] {
    need(m, "model");
    auto v = m->engine.model.get_joint_velocity_limits(ids_of(joint_ids, n, m->engine.model.n_joints(), "joint"));
    for (int i = 0; i < n; ++i) out[i] = v[i];
  });
}
int tfk_get_joint_effort_limits(const tfk_model *m, const int *joint_ids, int n, double *out) {
  return guarded([&] {
    need(m, "model");
    auto v = m->engine.model.get_joint_effort_limits(ids_of(joint_ids, n, m->engine.model.n_joints(), "joint"));
    for (int i = 0; i < n; ++i) out[i] = v[i];
  });
}
int tfk_link_info(const tfk_model *m, int link_id, int *parent, int *jtype, int *jid, double axis[3], double origin[7]) {
  return guarded([&] {
    need(m, "model");
    if (link_id < 0 || (size_t)link_id >= m->engine.model.n_links()) throw std::invalid_argument("link id out of range");
    const LinkRec &l = m->engine.model.links()[link_id];
    if (parent) *parent = l.parent;
    if (jtype) *jtype = l.jtype;
    if (jid) *jid = l.jid;
    if (axis) axis[0] = l.axis.x, axis[1] = l.axis.y, axis[2] = l.axis.z;
    if (origin) {
      origin[0] = l.origin.p.x, origin[1] = l.origin.p.y, origin[2] = l.origin.p.z;
      origin[3] = l.origin.q.x, origin[4] = l.origin.q.y, origin[5] = l.origin.q.z, origin[6] = l.origin.q.w;
    }
  });
}
int tfk_is_relevant(const tfk_model *m, int link_id, int joint_id, int *out) {
  return guarded([&] {
    need(m, "model");
    if (link_id < 0 || joint_id < 0) throw std::invalid_argument("id out of range");
    *out = m->engine.model.is_relevant((size_t)link_id, (size_t)joint_id) ? 1 : 0;
  });
}

int tfk_set_joint_angles(tfk_model *m, const int *joint_ids, int n, const double *angles) {
  return guarded([&] {
    need(m, "model");
    std::lock_guard<std::mutex> lk(m->engine.mu);
    m->engine.model.set_joint_angles(ids_of(joint_ids, n, m->engine.model.n_joints(), "joint"),
                                     std::vector<double>(angles, angles + n));
  });
}
int tfk_get_joint_angles(const tfk_model *m, const int *joint_ids, int n, double *angles) {
  return guarded([&] {
    need(m, "model");
    auto a = m->engine.model.get_joint_angles(ids_of(joint_ids, n, m->engine.model.n_joints(), "joint"));
    for (int i = 0; i < n; ++i) angles[i] = a[i];
  });
}
int tfk_set_base_pose(tfk_model *m, const double xyzrpy[6]) {
  return guarded([&] {
    need(m, "model");
    need(xyzrpy, "xyzrpy");
    std::lock_guard<std::mutex> lk(m->engine.mu);
    m->engine.model.set_base_pose_xyzrpy(xyzrpy);
  });
}
int tfk_get_base_pose(const tfk_model *m, double xyzrpy[6]) {
  return guarded([&] {
    need(m, "model");
    m->engine.model.get_base_pose_xyzrpy(xyzrpy);
  });
}
int tfk_add_new_link(tfk_model *m, const char *name, int parent_id, const double pos[3], const double rpy[3],
                     int *new_id_out) {
  return guarded([&] {
    need(m, "model");
    need(name, "name");
    if (parent_id < 0) throw std::invalid_argument("parent link id out of range");
    std::lock_guard<std::mutex> lk(m->engine.mu);
    int id = m->engine.model.add_new_link(name, (size_t)parent_id, pos, rpy);
    if (new_id_out) *new_id_out = id;
  });
}

static SolveSpec make_spec(const int *elink_ids, int n_link, const int *joint_ids, int n_joint, int rot_type,
                           int base_type, int with_jacobian, int precision) {
  if (n_link < 0 || n_joint < 0) throw std::invalid_argument("negative id count");
  if (n_link > 0) need(elink_ids, "elink_ids");
  if (n_joint > 0) need(joint_ids, "joint_ids");
  SolveSpec s;
  s.elink_ids = elink_ids;
  s.n_link = n_link;
  s.joint_ids = joint_ids;
  s.n_joint = n_joint;
  s.rot_type = rot_type;
  s.base_type = base_type;
  s.with_jacobian = with_jacobian;
  s.precision = precision;
  return s;
}

int tfk_solve_fk_host(tfk_model *m, const void *q, int64_t n, const int *elink_ids, int n_link, const int *joint_ids,
                      int n_joint, int rot_type, int base_type, int with_jacobian, int precision, const int *devices,
                      int n_devices, void *P, void *J) {
  return guarded([&] {
    need(m, "model");
    SolveSpec s = make_spec(elink_ids, n_link, joint_ids, n_joint, rot_type, base_type, with_jacobian, precision);
    std::lock_guard<std::mutex> lk(m->engine.mu);
    m->engine.solve_host(q, n, s, devices, n_devices, P, J);
  });
}

int tfk_solve_fk_device(tfk_model *m, int device, void *stream, const void *d_q, int64_t n, const int *elink_ids,
                        int n_link, const int *joint_ids, int n_joint, int rot_type, int base_type, int with_jacobian,
                        int precision, void *d_P, void *d_J, int64_t ld_cfg, int64_t row0) {
  return guarded([&] {
    need(m, "model");
    SolveSpec s = make_spec(elink_ids, n_link, joint_ids, n_joint, rot_type, base_type, with_jacobian, precision);
    std::lock_guard<std::mutex> lk(m->engine.mu);
    m->engine.solve_device(device, static_cast<cudaStream_t>(stream), d_q, n, s, d_P, d_J, ld_cfg, row0);
  });
}

// No counterpart in the reference (include/tinyfk_b200.h).
static_assert(TFK_NOALLOC_MAX_CONFIGS == (int)tfk::TFK_NOALLOC_MAX_CONFIGS_ && TFK_NOALLOC_STATIC_BYTES == (int)tfk::TFK_NOALLOC_SCRATCH_BYTES,
              "include/tinyfk_b200.h and csrc/fk_noalloc.h disagree on the limits of the allocation-free entry");
int tfk_solve_fk_device_noalloc(tfk_model *m, int device, void *stream, const void *d_q, int64_t n, const int *elink_ids,
                                int n_link, const int *joint_ids, int n_joint, int rot_type, int base_type, int with_jacobian,
                                void *d_P, void *d_J, int64_t ld_cfg, int64_t row0) {
  return guarded([&] {
    need(m, "model");
    SolveSpec s = make_spec(elink_ids, n_link, joint_ids, n_joint, rot_type, base_type, with_jacobian, PREC_F64_STRICT);
    std::lock_guard<std::mutex> lk(m->engine.mu);
    m->engine.solve_device_noalloc(device, static_cast<cudaStream_t>(stream), d_q, n, s, d_P, d_J, ld_cfg, row0);
  });
}

static PairSpec make_pair_spec(const tfk_model *m, const int *l1, const int *l2, int n_check, const int *joint_ids, int n_joint,
                               int base_type, int with_grads, int precision) {
  if (n_check < 0 || (n_check > 0 && (!l1 || !l2))) throw std::invalid_argument("bad link pair arrays");
  if (n_joint < 0 || (n_joint > 0 && !joint_ids)) throw std::invalid_argument("bad joint id array");
  ids_of(l1, n_check, m->engine.model.n_links(), "link");
  ids_of(l2, n_check, m->engine.model.n_links(), "link");
  ids_of(joint_ids, n_joint, m->engine.model.n_joints(), "joint");
  if (base_type < 0 || base_type > 2) throw std::invalid_argument("base_type must be 0 (FIXED), 1 (PLANER) or 2 (FLOATING)");
  PairSpec ps;
  ps.link_ids1 = l1;
  ps.link_ids2 = l2;
  ps.n_check = n_check;
  ps.joint_ids = joint_ids;
  ps.n_joint = n_joint;
  ps.base_type = base_type;
  ps.with_grads = with_grads ? 1 : 0;
  ps.precision = precision;
  return ps;
}

int tfk_inter_link_sqdists_host(tfk_model *m, const void *q, int64_t n, const int *link_ids1, const int *link_ids2,
                                int n_check, const int *joint_ids, int n_joint, int base_type, int with_grads,
                                int precision, void *sqdists, void *grads) {
  return guarded([&] {
    need(m, "model");
    PairSpec ps = make_pair_spec(m, link_ids1, link_ids2, n_check, joint_ids, n_joint, base_type, with_grads, precision);
    std::lock_guard<std::mutex> lk(m->engine.mu);
    m->engine.sqdists_host(q, n, ps, sqdists, grads);
  });
}

int tfk_inter_link_sqdists_device(tfk_model *m, int device, void *stream, const void *d_q, int64_t n,
                                  const int *link_ids1, const int *link_ids2, int n_check, const int *joint_ids,
                                  int n_joint, int base_type, int with_grads, int precision, void *d_sqdists,
                                  void *d_grads, int64_t ld_cfg, int64_t row0) {
  return guarded([&] {
    need(m, "model");
    PairSpec ps = make_pair_spec(m, link_ids1, link_ids2, n_check, joint_ids, n_joint, base_type, with_grads, precision);
    std::lock_guard<std::mutex> lk(m->engine.mu);
    m->engine.sqdists_device(device, static_cast<cudaStream_t>(stream), d_q, n, ps, d_sqdists, d_grads, ld_cfg, row0);
  });
}

static ComSpec make_com_spec(const tfk_model *m, const int *joint_ids, int n_joint, const int *action_link_ids,
                             const double *forces, int n_action, int base_type, int with_jacobian, int precision) {
  if (n_joint < 0 || (n_joint > 0 && !joint_ids)) throw std::invalid_argument("bad joint id array");
  if (n_action < 0 || (n_action > 0 && (!action_link_ids || !forces))) throw std::invalid_argument("bad action link / force arrays");
  ids_of(joint_ids, n_joint, m->engine.model.n_joints(), "joint");
  ids_of(action_link_ids, n_action, m->engine.model.n_links(), "link");
  if (base_type < 0 || base_type > 2) throw std::invalid_argument("base_type must be 0 (FIXED), 1 (PLANER) or 2 (FLOATING)");
  ComSpec cs;
  cs.joint_ids = joint_ids;
  cs.n_joint = n_joint;
  cs.action_link_ids = action_link_ids;
  cs.forces = forces;
  cs.n_action = n_action;
  cs.base_type = base_type;
  cs.with_jacobian = with_jacobian ? 1 : 0;
  cs.precision = precision;
  return cs;
}

int tfk_solve_com_fk_host(tfk_model *m, const void *q, int64_t n, const int *joint_ids, int n_joint,
                          const int *action_link_ids, const double *forces, int n_action, int base_type,
                          int with_jacobian, int precision, void *com, void *J) {
  return guarded([&] {
    need(m, "model");
    ComSpec cs = make_com_spec(m, joint_ids, n_joint, action_link_ids, forces, n_action, base_type, with_jacobian, precision);
    std::lock_guard<std::mutex> lk(m->engine.mu);
    m->engine.com_host(q, n, cs, com, J);
  });
}

int tfk_solve_com_fk_device(tfk_model *m, int device, void *stream, const void *d_q, int64_t n, const int *joint_ids,
                            int n_joint, const int *action_link_ids, const double *forces, int n_action,
                            int base_type, int with_jacobian, int precision, void *d_com, void *d_J, int64_t ld_cfg,
                            int64_t row0) {
  return guarded([&] {
    need(m, "model");
    ComSpec cs = make_com_spec(m, joint_ids, n_joint, action_link_ids, forces, n_action, base_type, with_jacobian, precision);
    std::lock_guard<std::mutex> lk(m->engine.mu);
    m->engine.com_device(device, static_cast<cudaStream_t>(stream), d_q, n, cs, d_com, d_J, ld_cfg, row0);
  });
}

int tfk_total_inertia_matrix_host(tfk_model *m, const void *q, int64_t n, const int *joint_ids, int n_joint, int precision,
                                  void *out) {
  return guarded([&] {
    need(m, "model");
    if (n_joint < 0 || (n_joint > 0 && !joint_ids)) throw std::invalid_argument("bad joint id array");
    ids_of(joint_ids, n_joint, m->engine.model.n_joints(), "joint");
    std::lock_guard<std::mutex> lk(m->engine.mu);
    m->engine.inertia_host(q, n, joint_ids, n_joint, precision, out);
  });
}

int tfk_total_inertia_matrix_device(tfk_model *m, int device, void *stream, const void *d_q, int64_t n, const int *joint_ids,
                                    int n_joint, int precision, void *d_out, int64_t ld_cfg, int64_t row0) {
  return guarded([&] {
    need(m, "model");
    if (n_joint < 0 || (n_joint > 0 && !joint_ids)) throw std::invalid_argument("bad joint id array");
    ids_of(joint_ids, n_joint, m->engine.model.n_joints(), "joint");
    std::lock_guard<std::mutex> lk(m->engine.mu);
    m->engine.inertia_device(device, static_cast<cudaStream_t>(stream), d_q, n, joint_ids, n_joint, precision, d_out, ld_cfg, row0);
  });
}

int tfk_device_count(int *count) {
  return guarded([&] {
    need(count, "count");
    *count = 0;
    cudaError_t e = cudaGetDeviceCount(count);
    if (e != cudaSuccess) throw CudaError(std::string("cudaGetDeviceCount: ") + cudaGetErrorString(e));
  });
}

int tfk_host_alloc(size_t bytes, void **ptr) {
  return guarded([&] {
    need(ptr, "ptr");
    *ptr = pinned_alloc(bytes);
  });
}
int tfk_host_free(void *ptr) {
  return guarded([&] { pinned_free(ptr); });
}

int64_t tfk_kernel_launch_count(void) { return kernel_launch_count(); }

int tfk_last_launch_info(const tfk_model *m, int *grid, int *block, int *smem_bytes, int *blocks_per_sm, int *rec_reals) {
  return guarded([&] {
    need(m, "model");
    const LaunchInfo &l = m->engine.last_launch;
    if (grid) *grid = l.grid;
    if (block) *block = l.block;
    if (smem_bytes) *smem_bytes = l.smem_bytes;
    if (blocks_per_sm) *blocks_per_sm = l.blocks_per_sm;
    if (rec_reals) *rec_reals = l.rec_reals;
  });
}

int tfk_last_launch_variant(const tfk_model *m) { return m ? m->engine.last_launch.variant : -1; }

int tfk_plan_cache_size(const tfk_model *m) { return m ? (int)m->engine.plan_cache_size() : -1; }

int tfk_model_tables_from_cache(const tfk_model *m) { return m ? (m->engine.model.tables_from_cache() ? 1 : 0) : -1; }

int tfk_specialize(tfk_model *m, const int *elink_ids, int n_link, const int *joint_ids, int n_joint, int rot_type,
                   int base_type, int with_jacobian, int precision, int flags, int64_t *info) {
  return guarded([&] {
    need(m, "model");
    SolveSpec s = make_spec(elink_ids, n_link, joint_ids, n_joint, rot_type, base_type, with_jacobian, precision);
    std::lock_guard<std::mutex> lk(m->engine.mu);
    const SpecInfo si = m->engine.specialize(s, (flags & TFK_SPEC_WAIT) != 0, (flags & TFK_SPEC_COMPILE_ONLY) != 0);
    if (info) {
      info[0] = si.state;
      info[1] = si.tile;
      info[2] = si.blocks_per_sm;
      info[3] = si.smem_bytes;
      info[4] = si.cubin_bytes;
      info[5] = si.from_cache;
    }
  });
}

int tfk_set_jit_mode(tfk_model *m, int mode) {
  return guarded([&] {
    need(m, "model");
    if (mode < 0 || mode > 2) throw std::invalid_argument("jit mode must be 0 (off), 1 (auto) or 2 (on)");
    std::lock_guard<std::mutex> lk(m->engine.mu);
    m->engine.jit_mode = mode;
  });
}
int tfk_get_jit_mode(const tfk_model *m) { return m ? m->engine.jit_mode : -1; }

} // extern "C"
