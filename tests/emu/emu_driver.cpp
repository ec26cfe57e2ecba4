// emu_driver.cpp -- TEST INFRASTRUCTURE, not a product path and never loaded by tinyfk_b200.
//
// Runs the SAME per-thread code the CUDA kernels run (tinyfk_b200/csrc/fk_core.h, phase A / phase B)
// serially on the CPU, driven by the product's own host model + plan flattening.  Purpose: this
// container has no GPU, so the kernel logic is debugged here against oracle/_ref before GPU time is
// spent; `-m "not gpu"` tests also use it to pin the host logic (plan pruning, masks, layouts).
// Built by tests/emu/build.sh with g++ -ffp-contract=off (the CPU equivalent of nvcc -fmad=false).
#include <cstring>
#include <string>
#include <vector>

#include "../../tinyfk_b200/csrc/fk_core.h"
#include "../../tinyfk_b200/csrc/model.hpp"
#include "../../tinyfk_b200/csrc/plan.hpp"

#include <stdexcept>

using namespace tfk;

static thread_local std::string g_err;

template <class T> struct DirectFactory { // the fused driver builds its sink after the walk
  typedef DirectSink<T> Sink;
  DirectSink<T> proto;
  Sink make() const { return proto; }
};

template <class Core, class T>
static void run(const std::vector<unsigned char> &blob, const double *qs, long n, double *Pd, double *Jd, bool two_phase) {
  PlanView pv = plan_view(blob.data());
  const PlanHeader &h = *pv.h;
  const long long ldP = (long long)n * h.n_out, ldJ = ldP * h.pose_dim;
  const bool fused = h.n_out == 1 && !two_phase; // the kernels' rule (engine.cpp)
  std::vector<T> rec((size_t)((fused ? h.fz_size : h.rec_size) > 0 ? (fused ? h.fz_size : h.rec_size) : 1));
  std::vector<T> qrow((size_t)(h.n_in > 0 ? h.n_in : 1));
  std::vector<T> P((size_t)ldP * h.pose_dim), J(h.with_jacobian ? (size_t)ldJ * h.n_out_cols : 0);
  DirectSink<T> out;
  out.P = P.data();
  out.J = J.data();
  out.ldP = ldP;
  out.ldJ = ldJ;
  out.D = h.pose_dim;
  for (long i = 0; i < n; ++i) {
    for (int c = 0; c < h.n_in; ++c) qrow[c] = (T)qs[(size_t)i * h.n_in + c];
    if (fused) {
      out.head = i;
      DirectFactory<T> f;
      f.proto = out;
      Core::run_fused(pv, qrow.data(), rec.data(), 1, f);
    } else {
      Core::phase_a(pv, qrow.data(), rec.data(), 1);
      {
        for (int o = 0; o < h.n_out; ++o) {
          out.head = (long long)i * h.n_out + o;
          Core::phase_b(pv, rec.data(), 1, o, out);
        }
      }
    }
  }
  for (size_t k = 0; k < P.size(); ++k) Pd[k] = (double)P[k];
  if (Jd)
    for (size_t k = 0; k < J.size(); ++k) Jd[k] = (double)J[k];
}

// the pair kernel's per-thread code (FkCore::phase_a + pair_sqdist) on the CPU
template <class Core, class T>
static void run_pairs(const std::vector<unsigned char> &blob, const std::vector<int> &o12, int n_check, const double *qs, long n,
                      double *sq, double *G) {
  PlanView pv = plan_view(blob.data());
  const PlanHeader &h = *pv.h;
  std::vector<T> rec((size_t)(h.rec_size > 0 ? h.rec_size : 1)), qrow((size_t)(h.n_in > 0 ? h.n_in : 1));
  const long long ld = (long long)n * n_check;
  std::vector<T> s((size_t)ld), g(G ? (size_t)ld * h.n_out_cols : 1);
  for (long i = 0; i < n; ++i) {
    for (int c = 0; c < h.n_in; ++c) qrow[c] = (T)qs[(size_t)i * h.n_in + c];
    Core::phase_a(pv, qrow.data(), rec.data(), 1);
    for (int j = 0; j < n_check; ++j) {
      const long long head = (long long)i * n_check + j;
      Core::pair_sqdist(pv, rec.data(), o12[j], o12[n_check + j], true, &s[head], G ? &g[head] : nullptr, ld);
    }
  }
  for (long long k = 0; k < ld; ++k) sq[k] = (double)s[k];
  if (G)
    for (size_t k = 0; k < (size_t)ld * h.n_out_cols; ++k) G[k] = (double)g[k];
}

// the centre-of-mass kernel's per-thread code (FkCore::phase_a + com_item, item loop of fk_com_kernel) on the CPU
template <class Core, class T>
static void run_com(const std::vector<unsigned char> &blob, const ComTables &ct, const double *qs, long n, double *com, double *J) {
  PlanView pv = plan_view(blob.data());
  const PlanHeader &h = *pv.h;
  std::vector<T> rec((size_t)(h.rec_size > 0 ? h.rec_size : 1)), qrow((size_t)(h.n_in > 0 ? h.n_in : 1));
  std::vector<T> w(ct.w.begin(), ct.w.end());
  typename Core::ComArgs ca;
  ca.w = w.data();
  ca.n_com = ct.n_com;
  ca.n_act = ct.n_act;
  ca.mass_total = (T)ct.mass_total;
  ca.vfs = (T)ct.vfs;
  const int n_base = h.base_type == 0 ? 0 : 6;
  const int n_items = 1 + (J ? h.n_cols + (n_base ? 3 : 0) : 0);
  const size_t ldJ = (size_t)3 * n;
  for (long i = 0; i < n; ++i) {
    for (int c = 0; c < h.n_in; ++c) qrow[c] = (T)qs[(size_t)i * h.n_in + c];
    Core::phase_a(pv, qrow.data(), rec.data(), 1);
    for (int item = 0; item < n_items; ++item) {
      const typename Core::V v = Core::com_item(pv, rec.data(), item, ca);
      const double d[3] = {(double)v.x, (double)v.y, (double)v.z};
      if (item == 0) {
        for (int k = 0; k < 3; ++k) com[(size_t)i * 3 + k] = d[k];
      } else {
        const int col = (item <= h.n_cols) ? item - 1 : item + 2;
        for (int k = 0; k < 3; ++k) J[(size_t)col * ldJ + 3 * i + k] = d[k];
      }
    }
    if (J && n_base)
      for (int k = 0; k < 3; ++k)
        for (int r = 0; r < 3; ++r) J[(size_t)(h.n_cols + k) * ldJ + 3 * i + r] = (r == k) ? (double)(T)ct.diag : 0.0;
  }
}

// the inertia kernel's per-thread code on the CPU: out 9 doubles row-major per configuration
template <class Core, class T>
static void run_inertia(const std::vector<unsigned char> &blob, const InertiaTables &it, const double *qs, long n, double *out) {
  PlanView pv = plan_view(blob.data());
  const PlanHeader &h = *pv.h;
  std::vector<T> rec((size_t)(h.rec_size > 0 ? h.rec_size : 1)), qrow((size_t)(h.n_in > 0 ? h.n_in : 1));
  std::vector<T> w(it.w.begin(), it.w.end());
  for (long i = 0; i < n; ++i) {
    for (int c = 0; c < h.n_in; ++c) qrow[c] = (T)qs[(size_t)i * h.n_in + c];
    Core::phase_a(pv, qrow.data(), rec.data(), 1);
    T m[9];
    Core::inertia_total(pv, rec.data(), w.data(), (int)it.links.size(), (T)it.mass_total, m);
    for (int k = 0; k < 9; ++k) out[(size_t)i * 9 + k] = (double)m[k];
  }
}

extern "C" {
int emu_sqdists(void *m, const double *qs, long n, const int *l1, const int *l2, int n_check, const int *joints, int n_joint,
                int base_type, int with_grads, int mode, double *sq, double *G) {
  try {
    std::vector<int> links, o12((size_t)2 * n_check);
    for (int side = 0; side < 2; ++side)
      for (int j = 0; j < n_check; ++j) {
        const int id = side == 0 ? l1[j] : l2[j];
        size_t k = 0;
        while (k < links.size() && links[k] != id) ++k;
        if (k == links.size()) links.push_back(id);
        o12[(size_t)side * n_check + j] = (int)k;
      }
    PlanKey key;
    key.elink_ids = links;
    key.joint_ids.assign(joints, joints + n_joint);
    key.rot_type = 0;
    key.base_type = base_type;
    key.with_jacobian = with_grads;
    key.model_version = static_cast<KinematicModel *>(m)->version();
    std::vector<unsigned char> blob = build_plan(*static_cast<KinematicModel *>(m), key);
    if (mode == 1)
      run_pairs<FkCore<MathNative32, true, 0, 1>, float>(blob, o12, n_check, qs, n, sq, with_grads ? G : nullptr);
    else
      run_pairs<FkCore<MathPortable, false, 0, 1>, double>(blob, o12, n_check, qs, n, sq, with_grads ? G : nullptr);
    return 0;
  } catch (const std::exception &e) {
    g_err = e.what();
    return 1;
  }
}
// base_type 0 (FIXED) or 2 (FLOATING); com (n, 3) row-major, J column-major (3n) x (n_joint + 0|6)
int emu_com(void *m, const double *qs, long n, const int *joints, int n_joint, const int *act, const double *forces, int n_act,
            int base_type, int with_jacobian, int mode, double *com, double *J) {
  try {
    const KinematicModel &km = *static_cast<KinematicModel *>(m);
    const ComTables ct = build_com_tables(km, act, forces, n_act);
    PlanKey key;
    key.elink_ids = ct.links;
    key.joint_ids.assign(joints, joints + n_joint);
    key.rot_type = 0;
    key.base_type = base_type;
    key.with_jacobian = with_jacobian;
    key.model_version = km.version();
    std::vector<unsigned char> blob = build_plan(km, key);
    make_plan_params(blob); // the kernel-parameter limits apply (throws when exceeded)
    if (mode == 1)
      run_com<FkCore<MathNative32, true, 0, 1>, float>(blob, ct, qs, n, com, with_jacobian ? J : nullptr);
    else
      run_com<FkCore<MathPortable, false, 0, 1>, double>(blob, ct, qs, n, com, with_jacobian ? J : nullptr);
    return 0;
  } catch (const std::exception &e) {
    g_err = e.what();
    return 1;
  }
}
// plan.cpp:plan_to_source for a single-link query; returns the length written (0 on error)
long emu_plan_source(void *m, const int *elinks, int n_link, const int *joints, int n_joint, int rot_type, int base_type,
                     int with_jacobian, char *buf, long cap) {
  try {
    PlanKey key;
    key.elink_ids.assign(elinks, elinks + n_link);
    key.joint_ids.assign(joints, joints + n_joint);
    key.rot_type = rot_type;
    key.base_type = base_type;
    key.with_jacobian = with_jacobian;
    key.model_version = static_cast<KinematicModel *>(m)->version();
    const std::string src = plan_to_source(build_plan(*static_cast<KinematicModel *>(m), key));
    if ((long)src.size() + 1 > cap) throw std::invalid_argument("buffer too small");
    std::memcpy(buf, src.c_str(), src.size() + 1);
    return (long)src.size();
  } catch (const std::exception &e) {
    g_err = e.what();
    return 0;
  }
}
int emu_inertia(void *m, const double *qs, long n, const int *joints, int n_joint, int mode, double *out) {
  try {
    const KinematicModel &km = *static_cast<KinematicModel *>(m);
    const InertiaTables it = build_inertia_tables(km);
    PlanKey key;
    key.elink_ids = it.links;
    key.joint_ids.assign(joints, joints + n_joint);
    key.rot_type = 2;
    key.base_type = 0;
    key.with_jacobian = 0;
    key.model_version = km.version();
    std::vector<unsigned char> blob = build_plan(km, key);
    make_plan_params(blob);
    if (mode == 1)
      run_inertia<FkCore<MathNative32, true, 2, 1>, float>(blob, it, qs, n, out);
    else
      run_inertia<FkCore<MathPortable, false, 2, 1>, double>(blob, it, qs, n, out);
    return 0;
  } catch (const std::exception &e) {
    g_err = e.what();
    return 1;
  }
}
// MathPortable::div_ (shared-reciprocal IEEE division, fk_math.h) on arrays: must equal a[i] / b[i]
void emu_div(const double *a, const double *b, double *o, long n) {
  for (long i = 0; i < n; ++i) o[i] = MathPortable::div_(a[i], MathPortable::den_(b[i]));
}
const char *emu_last_error() { return g_err.c_str(); }

void *emu_create(const char *xml, int system_libm) {
  try {
    return new KinematicModel(std::string(xml), system_libm ? LIBM_SYSTEM : LIBM_PORTABLE);
  } catch (const std::exception &e) {
    g_err = e.what();
    return nullptr;
  }
}
void emu_destroy(void *m) { delete static_cast<KinematicModel *>(m); }
int emu_add_link(void *m, const char *name, int parent, const double *pos, const double *rpy) {
  try {
    return static_cast<KinematicModel *>(m)->add_new_link(name, (size_t)parent, pos, rpy);
  } catch (const std::exception &e) {
    g_err = e.what();
    return -1;
  }
}
void emu_set_joint_angles(void *m, const int *ids, int n, const double *a) {
  std::vector<size_t> v(ids, ids + n);
  std::vector<double> x(a, a + n);
  static_cast<KinematicModel *>(m)->set_joint_angles(v, x);
}
void emu_set_base_pose(void *m, const double *b) { static_cast<KinematicModel *>(m)->set_base_pose_xyzrpy(b); }
int emu_num_links(void *m) { return (int)static_cast<KinematicModel *>(m)->n_links(); }
int emu_num_joints(void *m) { return (int)static_cast<KinematicModel *>(m)->n_joints(); }
int emu_root(void *m) { return static_cast<KinematicModel *>(m)->root_link_id(); }
const char *emu_link_name(void *m, int i) { return static_cast<KinematicModel *>(m)->links()[i].name.c_str(); }
const char *emu_joint_name(void *m, int i) { return static_cast<KinematicModel *>(m)->joints()[i].name.c_str(); }
void emu_link_info(void *m, int id, int *parent, int *jtype, int *jid, double *axis, double *origin7) {
  const LinkRec &l = static_cast<KinematicModel *>(m)->links()[id];
  *parent = l.parent;
  *jtype = l.jtype;
  *jid = l.jid;
  axis[0] = l.axis.x, axis[1] = l.axis.y, axis[2] = l.axis.z;
  origin7[0] = l.origin.p.x, origin7[1] = l.origin.p.y, origin7[2] = l.origin.p.z;
  origin7[3] = l.origin.q.x, origin7[4] = l.origin.q.y, origin7[5] = l.origin.q.z, origin7[6] = l.origin.q.w;
}
int emu_is_relevant(void *m, int l, int j) { return static_cast<KinematicModel *>(m)->is_relevant(l, j) ? 1 : 0; }

// mode 0: FP64 strict (portable libm, FD base columns); 1: FP32 (native libm, analytic base columns);
// 2: FP64 with the platform libm (FD base columns); 3: FP64 fast; +16: force the two-phase driver for a single link;
// +32 (modes 0 and 1): the core over the pair layout of the records
int emu_solve(void *m, const double *qs, long n, const int *elinks, int n_link, const int *joints, int n_joint,
              int rot_type, int base_type, int with_jacobian, int mode, double *P, double *J) {
  try {
    PlanKey key;
    key.elink_ids.assign(elinks, elinks + n_link);
    key.joint_ids.assign(joints, joints + n_joint);
    key.rot_type = rot_type;
    key.base_type = base_type;
    key.with_jacobian = with_jacobian;
    key.model_version = static_cast<KinematicModel *>(m)->version();
    std::vector<unsigned char> blob = build_plan(*static_cast<KinematicModel *>(m), key);
    const bool two_phase = (mode & 16) != 0;
    const bool paired = (mode & 32) != 0; // the record accessors of the pair layout (fk_core.h "record layout"; stride 1 here)
    mode &= 15;
    double *Jp = with_jacobian ? J : nullptr;
#define TFK_EMU_RUN(R)                                                                      \
  if (mode == 0 && paired)                                                                  \
    run<FkCore<MathPortable, false, R, 1, true>, double>(blob, qs, n, P, Jp, two_phase);       \
  else if (mode == 1 && paired)                                                             \
    run<FkCore<MathNative32, true, R, 1, true>, float>(blob, qs, n, P, Jp, two_phase);         \
  else if (mode == 0)                                                                       \
    run<FkCore<MathPortable, false, R, 1>, double>(blob, qs, n, P, Jp, two_phase);             \
  else if (mode == 1)                                                                       \
    run<FkCore<MathNative32, true, R, 1>, float>(blob, qs, n, P, Jp, two_phase);               \
  else if (mode == 3)                                                                       \
    run<FkCore<MathFast64, true, R, 1>, double>(blob, qs, n, P, Jp, two_phase);                \
  else                                                                                      \
    run<FkCore<MathNative64, false, R, 1>, double>(blob, qs, n, P, Jp, two_phase);
    if (rot_type == 0) {
      TFK_EMU_RUN(0)
    } else if (rot_type == 1) {
      TFK_EMU_RUN(1)
    } else {
      TFK_EMU_RUN(2)
    }
#undef TFK_EMU_RUN
    return 0;
  } catch (const std::exception &e) {
    g_err = e.what();
    return 1;
  }
}
}
