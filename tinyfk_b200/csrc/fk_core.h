// fk_core.h -- the per-config work of the batched FK + Jacobian path, as two phases.
//
//   phase A (one thread per configuration): base pose from the row tail, then ONE pass down the
//     pruned, topologically sorted link tree composing (quaternion, translation) frames in registers.
//     Replaces the reference's set_base_pose + clear_cache + get_link_pose/get_link_pose_inner walk
//     (src/wrapper.cpp:84-97, src/kinematics.cpp:37-103).  Emits small records: the base record,
//     one (world axis, position) record per driven joint, one (position[, quaternion]) record per
//     requested link.
//   phase B (one thread per (configuration, requested link)): pose row of P and all Jacobian
//     columns for that link from the records.  Replaces src/wrapper.cpp:98-121 + get_jacobian
//     (src/kinematics.cpp:105-235) and fuses the PLANER column pick of python/tinyfk/__init__.py:154-158.
//
// Header only, TFK_FN = __host__ __device__: the kernels (fk_kernels.cuh) call these per thread; the
// CPU test harness tests/emu (test infrastructure, not a product path) can run the very same code.
#ifndef TFK_FK_CORE_H
#define TFK_FK_CORE_H

#include "fk_math.h"
#include "fk_plan.h"

namespace tfk {

// Views into a plan blob (host memory on the CPU, shared memory on the GPU).
struct PlanView {
  const PlanHeader *h;
  const PlanSlot *slots;
  const int32_t *lrec_of_out;
  const uint32_t *masks;
  const PlanCol *cols;
};

TFK_FN PlanView plan_view(const unsigned char *blob) {
  PlanView v;
  v.h = reinterpret_cast<const PlanHeader *>(blob);
  v.slots = reinterpret_cast<const PlanSlot *>(blob + v.h->off_slots);
  v.lrec_of_out = reinterpret_cast<const int32_t *>(blob + v.h->off_outs);
  v.masks = reinterpret_cast<const uint32_t *>(blob + v.h->off_masks);
  v.cols = reinterpret_cast<const PlanCol *>(blob + v.h->off_cols);
  return v;
}

// M: math policy (fk_math.h).  ANALYTIC_BASE: base-rotation Jacobian columns from the closed form
// instead of the reference's forward difference (FP32 mode: eps = 1e-7 is below one FP32 ulp).
template <class M, bool ANALYTIC_BASE> struct FkCore {
  typedef typename M::real T;
  typedef Ops<M> O;
  typedef Vec3<T> V;
  typedef Quat<T> Q;
  typedef PoseT<T> P;

  static TFK_FN void put_pose(T *rec, int stride, int at, const P &a) {
    rec[(at + 0) * stride] = a.p.x;
    rec[(at + 1) * stride] = a.p.y;
    rec[(at + 2) * stride] = a.p.z;
    rec[(at + 3) * stride] = a.q.x;
    rec[(at + 4) * stride] = a.q.y;
    rec[(at + 5) * stride] = a.q.z;
    rec[(at + 6) * stride] = a.q.w;
  }
  static TFK_FN P get_pose(const T *rec, int stride, int at) {
    P a;
    a.p.x = rec[(at + 0) * stride];
    a.p.y = rec[(at + 1) * stride];
    a.p.z = rec[(at + 2) * stride];
    a.q.x = rec[(at + 3) * stride];
    a.q.y = rec[(at + 4) * stride];
    a.q.z = rec[(at + 5) * stride];
    a.q.w = rec[(at + 6) * stride];
    return a;
  }
  static TFK_FN V get_vec(const T *rec, int stride, int at) {
    V a;
    a.x = rec[(at + 0) * stride];
    a.y = rec[(at + 1) * stride];
    a.z = rec[(at + 2) * stride];
    return a;
  }
  static TFK_FN void put_vec(T *rec, int stride, int at, const V &a) {
    rec[(at + 0) * stride] = a.x;
    rec[(at + 1) * stride] = a.y;
    rec[(at + 2) * stride] = a.z;
  }

  // ---------------------------------------------------------------------------------------------
  // qrow: this configuration's input row (n_in reals).  rec: this configuration's record, element k
  // at rec[k * stride].
  static TFK_FN void phase_a(const PlanView &pv, const T *qrow, T *rec, int stride) {
    const PlanHeader &h = *pv.h;
    P base;
    if (h.base_type == 0) { // FIXED: the model's stored base pose (include/tinyfk.hpp:72, kinematics.cpp:51)
      base.p.x = T(h.base_pose[0]);
      base.p.y = T(h.base_pose[1]);
      base.p.z = T(h.base_pose[2]);
      base.q.x = T(h.base_pose[3]);
      base.q.y = T(h.base_pose[4]);
      base.q.z = T(h.base_pose[5]);
      base.q.w = T(h.base_pose[6]);
    } else {
      // src/wrapper.cpp:25-35 on the last 6 entries; PLANER rows are (x, y, yaw) padded to
      // (x, y, 0, 0, 0, yaw) by python/tinyfk/__init__.py:98-103
      const T *b = qrow + h.n_cols;
      T roll = T(0), pitch = T(0), yaw;
      base.p.x = b[0];
      base.p.y = b[1];
      if (h.base_type == 1) {
        base.p.z = T(0);
        yaw = b[2];
      } else {
        base.p.z = b[2];
        roll = b[3];
        pitch = b[4];
        yaw = b[5];
      }
      base.q = O::from_rpy(roll, pitch, yaw);
    }

    if (h.off_jrec > h.off_base) { // base record wanted (Jacobian with a moving base)
      put_pose(rec, stride, h.off_base, base);
      const V rpy_b = O::to_rpy(base.q); // src/kinematics.cpp:187
      if (ANALYTIC_BASE) {
        // world angular-velocity axes of roll, pitch, yaw for R = Rz(yaw) Ry(pitch) Rx(roll)
        T sp, cp, sy, cy;
        M::sincos_(rpy_b.y, &sp, &cp);
        M::sincos_(rpy_b.z, &sy, &cy);
        V wr, wp, wy;
        wr.x = cy * cp;
        wr.y = sy * cp;
        wr.z = -sp;
        wp.x = -sy;
        wp.y = cy;
        wp.z = T(0);
        wy.x = T(0);
        wy.y = T(0);
        wy.z = T(1);
        put_vec(rec, stride, h.off_base + 7, wr);
        put_vec(rec, stride, h.off_base + 10, wp);
        put_vec(rec, stride, h.off_base + 13, wy);
      } else {
        put_pose(rec, stride, h.off_base + 7, O::inverse(base)); // src/kinematics.cpp:186
        const T eps = T(1e-7);                                   // src/kinematics.cpp:201
        for (int k = (h.base_type == 1 ? 2 : 0); k < 3; ++k) {
          V t = rpy_b;
          if (k == 0) t.x += eps;
          if (k == 1) t.y += eps;
          if (k == 2) t.z += eps;
          const Q qk = O::from_rpy(t.x, t.y, t.z); // src/kinematics.cpp:205-210
          const int at = h.off_base + 14 + 4 * k;
          rec[(at + 0) * stride] = qk.x;
          rec[(at + 1) * stride] = qk.y;
          rec[(at + 2) * stride] = qk.z;
          rec[(at + 3) * stride] = qk.w;
        }
      }
    }
    if (h.root_lrec >= 0) { // the root link's pose is the base pose (src/kinematics.cpp:51-58,102)
      const int at = h.off_lrec + h.root_lrec * h.lrec_size;
      put_vec(rec, stride, at, base.p);
      if (h.lrec_size == 7) {
        rec[(at + 3) * stride] = base.q.x;
        rec[(at + 4) * stride] = base.q.y;
        rec[(at + 5) * stride] = base.q.z;
        rec[(at + 6) * stride] = base.q.w;
      }
    }

    T frames[TFK_MAX_FRAMES][7]; // branch-point frames; touched only by trees, never by chains
    P cur = base;
    for (int s = 0; s < h.n_slots; ++s) {
      const PlanSlot &sl = pv.slots[s];
      P parent;
      if (sl.frame_ld == FRAME_RUNNING) {
        parent = cur;
      } else if (sl.frame_ld == FRAME_BASE) {
        parent = base;
      } else {
        const T *f = frames[sl.frame_ld];
        parent.p.x = f[0];
        parent.p.y = f[1];
        parent.p.z = f[2];
        parent.q.x = f[3];
        parent.q.y = f[4];
        parent.q.z = f[5];
        parent.q.w = f[6];
      }
      P origin;
      origin.p.x = T(sl.opos[0]);
      origin.p.y = T(sl.opos[1]);
      origin.p.z = T(sl.opos[2]);
      origin.q.x = T(sl.oquat[0]);
      origin.q.y = T(sl.oquat[1]);
      origin.q.z = T(sl.oquat[2]);
      origin.q.w = T(sl.oquat[3]);
      V axis;
      axis.x = T(sl.axis[0]);
      axis.y = T(sl.axis[1]);
      axis.z = T(sl.axis[2]);
      P local;
      if (sl.kind == SLOT_FIXED) {
        local = origin; // src/kinematics.cpp:74-75
      } else if (sl.kind == SLOT_REVOLUTE) {
        local = O::local_revolute(origin, axis, qrow[sl.qcol]); // src/kinematics.cpp:77-80
      } else {
        local = O::local_prismatic(origin, axis, qrow[sl.qcol]);
      }
      cur = O::compose(parent, local); // src/kinematics.cpp:96-97
      if (sl.frame_st >= 0) {
        T *f = frames[sl.frame_st];
        f[0] = cur.p.x;
        f[1] = cur.p.y;
        f[2] = cur.p.z;
        f[3] = cur.q.x;
        f[4] = cur.q.y;
        f[5] = cur.q.z;
        f[6] = cur.q.w;
      }
      if (sl.jrec >= 0) { // src/kinematics.cpp:143-147: axis of the joint in the root frame
        const int at = h.off_jrec + 6 * sl.jrec;
        put_vec(rec, stride, at, O::rotate(cur.q, axis));
        put_vec(rec, stride, at + 3, cur.p);
      }
      if (sl.lrec >= 0) {
        const int at = h.off_lrec + sl.lrec * h.lrec_size;
        put_vec(rec, stride, at, cur.p);
        if (h.lrec_size == 7) {
          rec[(at + 3) * stride] = cur.q.x;
          rec[(at + 4) * stride] = cur.q.y;
          rec[(at + 5) * stride] = cur.q.z;
          rec[(at + 6) * stride] = cur.q.w;
        }
      }
    }
  }

  // ---------------------------------------------------------------------------------------------
  // One (configuration, requested link) pair.  row = global configuration index (row0 + i) used for
  // addressing; o = index into elink_ids.  Outputs are column-major exactly as the reference hands
  // them to numpy (src/wrapper.cpp:80-82,123):  P[d * ldP + row*n_out + o],
  // J[c * ldJ + (row*n_out + o)*pose_dim + r]  with ldP = N*n_out, ldJ = ldP*pose_dim.
  static TFK_FN void phase_b(const PlanView &pv, const T *rec, int stride, long long row, int o, T *Pout,
                             T *Jout, long long ldP, long long ldJ) {
    const PlanHeader &h = *pv.h;
    const int lat = h.off_lrec + pv.lrec_of_out[o] * h.lrec_size;
    P e;
    e.p = get_vec(rec, stride, lat);
    e.q = O::qidentity();
    if (h.lrec_size == 7) {
      e.q.x = rec[(lat + 3) * stride];
      e.q.y = rec[(lat + 4) * stride];
      e.q.z = rec[(lat + 5) * stride];
      e.q.w = rec[(lat + 6) * stride];
    }
    const long long head = row * h.n_out + o;
    Pout[0 * ldP + head] = e.p.x; // src/wrapper.cpp:98-100
    Pout[1 * ldP + head] = e.p.y;
    Pout[2 * ldP + head] = e.p.z;
    V erpy;
    erpy.x = erpy.y = erpy.z = T(0);
    if (h.rot_type == 1) { // src/wrapper.cpp:101-106
      erpy = O::to_rpy(e.q);
      Pout[3 * ldP + head] = erpy.x;
      Pout[4 * ldP + head] = erpy.y;
      Pout[5 * ldP + head] = erpy.z;
    } else if (h.rot_type == 2) { // src/wrapper.cpp:107-112
      Pout[3 * ldP + head] = e.q.x;
      Pout[4 * ldP + head] = e.q.y;
      Pout[5 * ldP + head] = e.q.z;
      Pout[6 * ldP + head] = e.q.w;
    }
    if (!h.with_jacobian) return;

    const int D = h.pose_dim;
    T *Jrow = Jout + head * D;
    typename O::RpyCoef coef;
    Q einv = O::qidentity();
    if (h.rot_type == 1) coef = O::rpy_coef(erpy);   // src/kinematics.cpp:121-123 + :14-21
    if (h.rot_type == 2) einv = O::qinverse(e.q);    // src/kinematics.cpp:124-126

    // joint columns, src/kinematics.cpp:131-180
    const uint32_t *mask = pv.masks + (size_t)o * h.mask_words;
    for (int c = 0; c < h.n_cols; ++c) {
      T *Jc = Jrow + (long long)c * ldJ;
      const bool relevant = (mask[c >> 5] >> (c & 31)) & 1u;
      if (!relevant) {
        for (int r = 0; r < D; ++r) Jc[r] = T(0);
        continue;
      }
      const int jat = h.off_jrec + 6 * pv.cols[c].jrec;
      const V aw = get_vec(rec, stride, jat);
      if (pv.cols[c].prismatic) {
        Jc[0] = aw.x;
        Jc[1] = aw.y;
        Jc[2] = aw.z;
        for (int r = 3; r < D; ++r) Jc[r] = T(0);
        continue;
      }
      const V cp = get_vec(rec, stride, jat + 3);
      V d;
      d.x = e.p.x - cp.x;
      d.y = e.p.y - cp.y;
      d.z = e.p.z - cp.z;
      const V dpos = O::cross(aw, d);
      Jc[0] = dpos.x;
      Jc[1] = dpos.y;
      Jc[2] = dpos.z;
      if (h.rot_type == 1) {
        const V dr = O::rpy_derivative(coef, aw);
        Jc[3] = dr.x;
        Jc[4] = dr.y;
        Jc[5] = dr.z;
      } else if (h.rot_type == 2) {
        const Q dq = O::q_derivative(einv, aw);
        Jc[3] = dq.x;
        Jc[4] = dq.y;
        Jc[5] = dq.z;
        Jc[6] = dq.w;
      }
    }

    if (h.base_type == 0) return;
    // base translation columns, src/kinematics.cpp:194-196
    const int n_trans = (h.base_type == 1) ? 2 : 3;
    for (int k = 0; k < n_trans; ++k) {
      T *Jc = Jrow + (long long)(h.n_cols + k) * ldJ;
      for (int r = 0; r < D; ++r) Jc[r] = (r == k) ? T(1) : T(0);
    }
    // base rotation columns: FLOATING keeps roll, pitch, yaw; PLANER keeps yaw only
    const P base = get_pose(rec, stride, h.off_base);
    const int k0 = (h.base_type == 1) ? 2 : 0;
    if (ANALYTIC_BASE) {
      V d;
      d.x = e.p.x - base.p.x;
      d.y = e.p.y - base.p.y;
      d.z = e.p.z - base.p.z;
      for (int k = k0; k < 3; ++k) {
        T *Jc = Jrow + (long long)(h.n_cols + n_trans + (k - k0)) * ldJ;
        const V w = get_vec(rec, stride, h.off_base + 7 + 3 * k);
        const V dpos = O::cross(w, d);
        Jc[0] = dpos.x;
        Jc[1] = dpos.y;
        Jc[2] = dpos.z;
        if (h.rot_type == 1) {
          const V dr = O::rpy_derivative(coef, w);
          Jc[3] = dr.x;
          Jc[4] = dr.y;
          Jc[5] = dr.z;
        } else if (h.rot_type == 2) {
          const Q dq = O::q_derivative(einv, w);
          Jc[3] = dq.x;
          Jc[4] = dq.y;
          Jc[5] = dq.z;
          Jc[6] = dq.w;
        }
      }
      return;
    }
    // forward finite difference replica, src/kinematics.cpp:182-232
    const T eps = T(1e-7);
    const P binv = get_pose(rec, stride, h.off_base + 7);
    P be; // tf_blink_to_elink = pose_transform(tf_blink_to_rlink, tf_rlink_to_elink)
    if (h.lrec_size == 7) {
      be = O::compose(binv, e);
    } else { // IGNORE: only the position rows exist and they never read the link rotation
      const V r = O::rotate(binv.q, e.p);
      be.p.x = r.x + binv.p.x;
      be.p.y = r.y + binv.p.y;
      be.p.z = r.z + binv.p.z;
      be.q = O::qidentity();
    }
    for (int k = k0; k < 3; ++k) {
      T *Jc = Jrow + (long long)(h.n_cols + n_trans + (k - k0)) * ldJ;
      const int at = h.off_base + 14 + 4 * k;
      P bk;
      bk.p = base.p;
      bk.q.x = rec[(at + 0) * stride];
      bk.q.y = rec[(at + 1) * stride];
      bk.q.z = rec[(at + 2) * stride];
      bk.q.w = rec[(at + 3) * stride];
      V pk;
      Q qk = O::qidentity();
      if (h.lrec_size == 7) {
        const P t = O::compose(bk, be);
        pk = t.p;
        qk = t.q;
      } else {
        const V r = O::rotate(bk.q, be.p);
        pk.x = r.x + bk.p.x;
        pk.y = r.y + bk.p.y;
        pk.z = r.z + bk.p.z;
      }
      Jc[0] = (pk.x - e.p.x) / eps;
      Jc[1] = (pk.y - e.p.y) / eps;
      Jc[2] = (pk.z - e.p.z) / eps;
      if (h.rot_type == 1) {
        const V rk = O::to_rpy(qk);
        Jc[3] = (rk.x - erpy.x) / eps;
        Jc[4] = (rk.y - erpy.y) / eps;
        Jc[5] = (rk.z - erpy.z) / eps;
      } else if (h.rot_type == 2) {
        Jc[3] = (qk.x - e.q.x) / eps;
        Jc[4] = (qk.y - e.q.y) / eps;
        Jc[5] = (qk.z - e.q.z) / eps;
        Jc[6] = (qk.w - e.q.w) / eps;
      }
    }
  }
};

} // namespace tfk
#endif
