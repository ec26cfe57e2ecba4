// fk_core.h -- the per-configuration work of the batched FK + Jacobian path.
//
//   walk():      base pose -> ONE pass down the pruned, topologically sorted link tree, composing
//                (quaternion, translation) frames in registers.  Replaces the reference's
//                set_base_pose + clear_cache + get_link_pose/get_link_pose_inner walk
//                (src/wrapper.cpp:84-97, src/kinematics.cpp:37-103).  Emits small records (world axis
//                per driven joint, position per distinct joint position, branch-point frames) and
//                hands every requested link's pose to a sink.
//   emit_link(): pose row of P and all Jacobian columns of one (configuration, link) pair from the
//                records.  Replaces src/wrapper.cpp:98-121 + get_jacobian (src/kinematics.cpp:105-235)
//                and fuses the PLANER column pick of python/tinyfk/__init__.py:154-158.
//
// Two drivers use them (fk_kernels.cuh):
//   fused     (one requested link): run_fused() = walk + emit_link in the same thread, records are
//             thread-private (shared memory), nothing but q, P and J touches HBM;
//   two-phase (several links): phase_a() = walk per configuration storing link records, phase_b() =
//             emit_link per (configuration, link) so that consecutive threads write consecutive rows.
//
// Header only, TFK_FN = __host__ __device__: the kernels call these per thread; the CPU test harness
// tests/emu (test infrastructure, not a product path) runs the very same code.
#ifndef TFK_FK_CORE_H
#define TFK_FK_CORE_H

#include "fk_math.h"
#include "fk_plan.h"

#if defined(__CUDACC__)
#define TFK_UNROLL _Pragma("unroll")
#define TFK_NOUNROLL _Pragma("unroll 1")
// loops over the plan (slots, Jacobian columns, base columns): one copy of the body in the generic kernels; fully
// unrolled in a plan-specialised build, where the plan is a compile-time constant and every test on it folds away
#if defined(TFK_SPEC_PLAN)
#define TFK_PLAN_LOOP _Pragma("unroll")
#else
#define TFK_PLAN_LOOP _Pragma("unroll 1")
#endif
// the walk down the tree: unrolled with the plan for a single chain (fused kernel); a several-link plan keeps ONE copy
// of the slot body (TFK_SPEC_ROLLED_WALK: 50-130 slots of straight-line code do not fit the instruction cache of an SM
// whose 12-16 warps are spread over it -- measured slower than the generic kernel) and unrolls only the column loops
#if defined(TFK_SPEC_PLAN) && !defined(TFK_SPEC_ROLLED_WALK)
#define TFK_WALK_LOOP _Pragma("unroll")
#else
#define TFK_WALK_LOOP _Pragma("unroll 1")
#endif
#else
#define TFK_UNROLL
#define TFK_NOUNROLL
#define TFK_PLAN_LOOP
#define TFK_WALK_LOOP
#endif

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

// Output sink used on the CPU (tests/emu) and wherever no staging is wanted: plain stores into the
// column-major arrays of the reference (src/wrapper.cpp:80-82,123):
//   P[d*ldP + head],  J[c*ldJ + head*D + r],  head = row*n_out + o.
template <class T> struct DirectSink {
  T *P, *J;
  long long ldP, ldJ, head;
  int D;
  TFK_FN void begin() const {}
  TFK_FN void put_P(int d, T v) const { P[d * ldP + head] = v; }
  TFK_FN void put_J(int c, const T (&v)[7]) const {
    T *Jc = J + c * ldJ + head * D;
    for (int r = 0; r < D; ++r) Jc[r] = v[r];
  }
  TFK_FN void put_J2(int c, const T (&v0)[7], const T (&v1)[7]) const { // columns c and c + 1
    put_J(c, v0);
    put_J(c + 1, v1);
  }
  TFK_FN void put_J_unit(int c, int k) const { // column c = unit vector e_k on every link (base translation columns)
    T *Jc = J + c * ldJ + head * D;
    for (int r = 0; r < D; ++r) Jc[r] = (r == k) ? T(1) : T(0);
  }
};

// two adjacent record elements moved as one (pair layout of FkCore, below)
template <class T> struct RecPair {
  struct alignas(2 * sizeof(T)) type {
    T x, y;
  };
};

// M: math policy (fk_math.h).  ANALYTIC_BASE: base-rotation Jacobian columns from the closed form
// instead of the reference's forward difference (FP32 mode: eps = 1e-7 is below one FP32 ulp).
// ROT: the call's RotationType (0 IGNORE, 1 RPY, 2 XYZW) as a compile-time constant -- it fixes the
// pose dimension, so the kernels are instantiated once per value instead of branching per column.
// STRIDE: distance in reals between consecutive record elements of one configuration ([k][STRIDE] layout:
// TFK_TILE in the kernels, 1 in the CPU harness) -- compile-time, so record addresses are immediates; the
// `stride` arguments below must equal it.
template <class M, bool ANALYTIC_BASE, int ROT, int STRIDE, bool PAIRS = false> struct FkCore {
  enum { D = ROT == 0 ? 3 : (ROT == 1 ? 6 : 7), LREC = ROT == 0 ? 3 : 7 };
  enum { SPARE = M::SPARE_REGS }; // the mode's arithmetic leaves registers for software prefetch (fk_twophase_kernel)
  // the same core over records laid out with another stride (the warp-private staged copy of fk_twophase_kernel)
  template <int STRIDE2> struct Restride {
    typedef FkCore<M, ANALYTIC_BASE, ROT, STRIDE2, PAIRS> type;
  };
  // the same core over the pair layout / the plain layout of the records (below)
  typedef FkCore<M, ANALYTIC_BASE, ROT, STRIDE, true> Paired;
  typedef FkCore<M, ANALYTIC_BASE, ROT, STRIDE, false> Plain;
  typedef typename M::real T;
  typedef Ops<M> O;
  typedef Vec3<T> V;
  typedef Quat<T> Q;
  typedef PoseT<T> P;

  // ---- record layout ------------------------------------------------------------------------------
  // Element k of a configuration's records lives at rec[ix(k)], `rec` being the configuration's base pointer.
  //   PAIRS = false  rec = array + i,          ix(k) = k * STRIDE                     ([k][STRIDE])
  //   PAIRS = true   rec = array + 2 * i,      ix(k) = (k / 2) * 2 * STRIDE + k % 2   ([k / 2][STRIDE][2])
  // The pair layout (the several-link device kernels) keeps elements 2m and 2m + 1 of one configuration adjacent and
  // 16-byte aligned (FP64), so a run of N elements is moved as whole pairs -- a 3-vector is one 128-bit and one 64-bit
  // access instead of three 64-bit ones, a pose four instead of seven: phase B of the several-link kernels waits on
  // exactly those shared-memory loads (profiles/r2_full_cfg4_fp64.txt before: short scoreboard 3.1 + MIO throttle 1.6
  // cycles per issue).  The plan keeps off_frame, rec_size and fz_size even (plan.cpp), so the split of the records
  // between shared memory and the scratch falls on a pair boundary.
  enum { TB = PAIRS ? 2 : 1 }; // distance between the base pointers of consecutive configurations
  static TFK_FN int ix(int k) { return PAIRS ? (k >> 1) * (2 * STRIDE) + (k & 1) : k * STRIDE; }

  // N consecutive elements at .. at + N - 1.  Pair layout: whole pairs as one 128-bit (FP64) access, the odd element at
  // either end alone; `at & 1` is a constant of the plan in unrolled code and a uniform value in the rolled walk.
  template <int N> static TFK_FN void ldn(const T *rec, int at, T (&v)[N]) {
    if (PAIRS) {
    typedef typename RecPair<T>::type T2;
    const T *r_ = rec + ix(at & ~1); // the pair that holds element `at`
    if (at & 1) {
      v[0] = r_[1];
      TFK_UNROLL
      for (int m = 1; 2 * m <= N - 1; ++m) {
        const T2 p = *reinterpret_cast<const T2 *>(r_ + m * (2 * STRIDE));
        v[2 * m - 1] = p.x, v[2 * m] = p.y;
      }
      if ((N & 1) == 0) v[N - 1] = r_[(N / 2) * (2 * STRIDE)];
    } else {
      TFK_UNROLL
      for (int m = 0; 2 * m + 1 <= N - 1; ++m) {
        const T2 p = *reinterpret_cast<const T2 *>(r_ + m * (2 * STRIDE));
        v[2 * m] = p.x, v[2 * m + 1] = p.y;
      }
      if (N & 1) v[N - 1] = r_[((N - 1) / 2) * (2 * STRIDE)];
    }
    return;
    }
    const T *r_ = rec + at * STRIDE; // one address, constant offsets
    TFK_UNROLL
    for (int k = 0; k < N; ++k) v[k] = r_[k * STRIDE];
  }
  template <int N> static TFK_FN void stn(T *rec, int at, const T (&v)[N]) {
    if (PAIRS) {
    typedef typename RecPair<T>::type T2;
    T *r_ = rec + ix(at & ~1);
    if (at & 1) {
      r_[1] = v[0];
      TFK_UNROLL
      for (int m = 1; 2 * m <= N - 1; ++m) {
        T2 p;
        p.x = v[2 * m - 1], p.y = v[2 * m];
        *reinterpret_cast<T2 *>(r_ + m * (2 * STRIDE)) = p;
      }
      if ((N & 1) == 0) r_[(N / 2) * (2 * STRIDE)] = v[N - 1];
    } else {
      TFK_UNROLL
      for (int m = 0; 2 * m + 1 <= N - 1; ++m) {
        T2 p;
        p.x = v[2 * m], p.y = v[2 * m + 1];
        *reinterpret_cast<T2 *>(r_ + m * (2 * STRIDE)) = p;
      }
      if (N & 1) r_[((N - 1) / 2) * (2 * STRIDE)] = v[N - 1];
    }
    return;
    }
    T *r_ = rec + at * STRIDE; // one address, constant offsets
    TFK_UNROLL
    for (int k = 0; k < N; ++k) r_[k * STRIDE] = v[k];
  }

  static TFK_FN void put_pose(T *rec, int stride, int at, const P &a) {
    const T v[7] = {a.p.x, a.p.y, a.p.z, a.q.x, a.q.y, a.q.z, a.q.w};
    stn(rec, at, v);
  }
  static TFK_FN P get_pose(const T *rec, int stride, int at) {
    T v[7];
    ldn(rec, at, v);
    P a;
    a.p.x = v[0], a.p.y = v[1], a.p.z = v[2];
    a.q.x = v[3], a.q.y = v[4], a.q.z = v[5], a.q.w = v[6];
    return a;
  }
  static TFK_FN V get_vec(const T *rec, int stride, int at) {
    T v[3];
    ldn(rec, at, v);
    V a;
    a.x = v[0], a.y = v[1], a.z = v[2];
    return a;
  }
  static TFK_FN void put_vec(T *rec, int stride, int at, const V &a) {
    const T v[3] = {a.x, a.y, a.z};
    stn(rec, at, v);
  }
  static TFK_FN Q get_quat(const T *rec, int stride, int at) {
    T v[4];
    ldn(rec, at, v);
    Q a;
    a.x = v[0], a.y = v[1], a.z = v[2], a.w = v[3];
    return a;
  }
  static TFK_FN void put_quat(T *rec, int stride, int at, const Q &a) {
    const T v[4] = {a.x, a.y, a.z, a.w};
    stn(rec, at, v);
  }
  // per-lane record offsets (phase B reads the link record of ITS output: `at` differs from lane to lane): element-wise
  static TFK_FN P get_pose_lane(const T *rec, int at, bool with_q) {
    P a;
    a.p.x = rec[ix(at)], a.p.y = rec[ix(at + 1)], a.p.z = rec[ix(at + 2)];
    a.q = O::qidentity();
    if (with_q) a.q.x = rec[ix(at + 3)], a.q.y = rec[ix(at + 4)], a.q.z = rec[ix(at + 5)], a.q.w = rec[ix(at + 6)];
    return a;
  }

  // ---------------------------------------------------------------------------------------------
  // Base pose of this configuration.  FIXED: the model's stored pose (include/tinyfk.hpp:72,
  // kinematics.cpp:51); otherwise src/wrapper.cpp:25-35 on the row tail (PLANER rows are (x, y, yaw),
  // padded to (x, y, 0, 0, 0, yaw) by python/tinyfk/__init__.py:98-103).
  static TFK_FN P base_pose_of(const PlanHeader &h, const T *qrow) { return base_pose_from_tail(h, qrow + h.n_cols); }
  // b: the row tail (x, y, yaw | x, y, z, roll, pitch, yaw); not read for a FIXED base
  static TFK_FN P base_pose_from_tail(const PlanHeader &h, const T *b) {
    P base;
    if (h.base_type == 0) {
      base.p.x = T(h.base_pose[0]);
      base.p.y = T(h.base_pose[1]);
      base.p.z = T(h.base_pose[2]);
      base.q.x = T(h.base_pose[3]);
      base.q.y = T(h.base_pose[4]);
      base.q.z = T(h.base_pose[5]);
      base.q.w = T(h.base_pose[6]);
      return base;
    }
    base.p.x = b[0];
    base.p.y = b[1];
    if (h.base_type == 1) {
      // setFromRPY(0, 0, yaw), pose.h:202-216: sin(0) = 0, cos(0) = 1 exactly, so the products with
      // them are exact and x = y = 0, z = sin(yaw/2), w = cos(yaw/2) before normalize
      base.p.z = T(0);
      base.q = O::qz_dense(O::qz_from_yaw(b[2]));
    } else {
      base.p.z = b[2];
      base.q = O::from_rpy(b[3], b[4], b[5]);
    }
    return base;
  }

  // ---------------------------------------------------------------------------------------------
  // What the base columns of the Jacobian need (src/kinematics.cpp:182-232), derived from the base
  // pose.  The reference recomputes all of it per link; the operands are the same for every link of a
  // configuration, so hoisting is exact.
  //   strict:   rpy_b = getRPY(base) (:187), binv = Pose::inverse(base) (:186),
  //             bk[k] = setFromRPY(rpy_b + eps*e_k) (:201-210)
  //   analytic: world angular-velocity axes of roll, pitch, yaw for R = Rz(yaw) Ry(pitch) Rx(roll)
  static TFK_FN Q tweaked_rotation(const V &rpy_b, int k) {
    const T eps = T(1e-7); // src/kinematics.cpp:201
    V t = rpy_b;
    if (k == 0) t.x += eps;
    if (k == 1) t.y += eps;
    if (k == 2) t.z += eps;
    return O::from_rpy(t.x, t.y, t.z);
  }
  static TFK_FN void analytic_axes(const V &rpy_b, V (&w)[3]) {
    T sp, cp, sy, cy;
    M::sincos_(rpy_b.y, &sp, &cp);
    M::sincos_(rpy_b.z, &sy, &cy);
    w[0].x = cy * cp;
    w[0].y = sy * cp;
    w[0].z = -sp;
    w[1].x = -sy;
    w[1].y = cy;
    w[1].z = T(0);
    w[2].x = T(0);
    w[2].y = T(0);
    w[2].z = T(1);
  }

  // two-phase: phase A stores the context once per configuration, phase B reads what it needs, late
  // PLANER: everything is yaw-only (fk_math.h "yaw-only rotations"); same record layout, x = y = 0
  struct PlanerCtx {
    V base_p;
    typename O::PZ binv;
    typename O::QZ bk; // setFromRPY(0, 0, yaw_b + eps), the only tweaked rotation PLANER keeps
  };
  static TFK_FN PlanerCtx planer_ctx(const P &base) {
    PlanerCtx c;
    c.base_p = base.p;
    typename O::PZ b;
    b.p = base.p;
    b.q.z = base.q.z;
    b.q.w = base.q.w;
    c.binv = O::inverse_z(b);
    c.bk = O::qz_from_yaw(O::qz_yaw(b.q) + T(1e-7)); // src/kinematics.cpp:187, 201-210
    return c;
  }
  static TFK_FN void store_base_ctx(const PlanHeader &h, const P &base, T *rec, int stride) {
    if (!ANALYTIC_BASE && h.base_type == 1) {
      const PlanerCtx c = planer_ctx(base);
      P bi;
      bi.p = c.binv.p;
      bi.q = O::qz_dense(c.binv.q);
      put_pose(rec, stride, h.off_base + 7, bi);
      put_quat(rec, stride, h.off_base + 14 + 4 * 2, O::qz_dense(c.bk));
      return;
    }
    const V rpy_b = O::to_rpy(base.q);
    if (ANALYTIC_BASE) {
      V w[3];
      analytic_axes(rpy_b, w);
      TFK_UNROLL
      for (int k = 0; k < 3; ++k) put_vec(rec, stride, h.off_base + 7 + 3 * k, w[k]);
      return;
    }
    put_pose(rec, stride, h.off_base + 7, O::inverse(base));
    TFK_NOUNROLL
    for (int k = (h.base_type == 1 ? 2 : 0); k < 3; ++k)
      put_quat(rec, stride, h.off_base + 14 + 4 * k, tweaked_rotation(rpy_b, k));
  }
  struct BaseFromRec {
    const T *rec;
    int stride, at;
    TFK_FN void prepare() {}
    TFK_FN V base_p() const { return get_vec(rec, stride, at); }
    TFK_FN P binv() const { return get_pose(rec, stride, at + 7); }
    TFK_FN Q bk(int k) const { return get_quat(rec, stride, at + 14 + 4 * k); }
    TFK_FN V w(int k) const { return get_vec(rec, stride, at + 7 + 3 * k); }
    TFK_FN PlanerCtx planer() const {
      PlanerCtx c;
      c.base_p = get_vec(rec, stride, at);
      c.binv.p = get_vec(rec, stride, at + 7);
      c.binv.q.z = rec[ix(at + 12)];
      c.binv.q.w = rec[ix(at + 13)];
      c.bk.z = rec[ix(at + 24)];
      c.bk.w = rec[ix(at + 25)];
      return c;
    }
  };
  // fused: one link per configuration, so everything is computed where it is used; only the base pose
  // itself is re-read from the record (it exists only for a Jacobian with a moving base)
  // the base record of the fused driver: a PLANER pose is (x, y, 0) and (0, 0, z, w) -- four reals
  static TFK_FN void put_base_fused(const PlanHeader &h, T *rec, const P &base) {
    if (h.base_type == 1) {
      const T v[4] = {base.p.x, base.p.y, base.q.z, base.q.w};
      stn(rec, 0, v);
    } else {
      put_pose(rec, STRIDE, 0, base);
    }
  }
  static TFK_FN P get_base_fused(const PlanHeader &h, const T *rec) {
    if (h.base_type != 1) return get_pose(rec, STRIDE, 0);
    T v[4];
    ldn(rec, 0, v);
    P b;
    b.p.x = v[0], b.p.y = v[1], b.p.z = T(0);
    b.q.x = T(0), b.q.y = T(0), b.q.z = v[2], b.q.w = v[3];
    return b;
  }
  struct BaseComputed {
    const T *rec;
    int stride;
    const PlanHeader *h;
    P b;
    V rpy_b;
    V wk[3];
    TFK_FN void prepare() {
      b = get_base_fused(*h, rec);
      rpy_b = O::to_rpy(b.q);
      if (ANALYTIC_BASE) analytic_axes(rpy_b, wk);
    }
    TFK_FN V base_p() const { return b.p; }
    TFK_FN P binv() const { return O::inverse(b); }
    TFK_FN Q bk(int k) const { return tweaked_rotation(rpy_b, k); }
    TFK_FN V w(int k) const { return (k == 0) ? wk[0] : ((k == 1) ? wk[1] : wk[2]); }
    TFK_FN PlanerCtx planer() const { return planer_ctx(get_base_fused(*h, rec)); }
  };

  // ---------------------------------------------------------------------------------------------
  // One pass down the tree.  qrow: this configuration's input row.  rec: this configuration's records
  // (element k at rec[k*stride]); o_base / o_arec / o_prec / o_frame: where the base pose (already
  // stored by the caller), axis, position and frame records are.  sink.link(lrec, pose) is called for
  // every requested link, in tree order.  Returns the pose of the last slot.
  // BASE_REC = false (fused driver): no base record; the only child of the root that is walked is slot 0,
  // whose parent frame is the initial `cur`.
  template <bool BASE_REC, class LinkSink>
  static TFK_FN P walk(const PlanView &pv, const T *qrow, T *rec, T *rec_f, int stride, int o_base, int o_arec, int o_prec,
                       int o_frame, P cur, const LinkSink &sink) {
    const PlanHeader &h = *pv.h;
    Q inv = O::qidentity(); // GetInverse(q of the frame the next rotation uses), valid where the plan says so
#if defined(TFK_SPEC_PLAN) && defined(TFK_SPEC_SINCOS_PAIR) && !defined(TFK_SPEC_ROLLED_WALK)
    // Plan-specialised build: the slot loop is unrolled over a compile-time plan, so "is there a second revolute
    // joint ahead" folds.  Revolute joints are taken in pairs: the first one's slot evaluates sin/cos of BOTH half
    // angles in one call (M::sincos2_: two independent chains that interleave), the second one's slot uses the
    // kept pair.  Same operations per angle, same bits.
    T sn_kept = T(0), cs_kept = T(0);
    bool kept = false;
#endif
    TFK_WALK_LOOP
    for (int s = 0; s < h.n_slots; ++s) {
      const PlanSlot &sl = pv.slots[s];
      // SF_LEAF (several-link plans only): a fixed link without needed children is computed on the side -- the
      // running frame is put back afterwards, so its siblings continue from their parent without a frame reload
      const bool leaf = BASE_REC && (sl.flags & SF_LEAF);
      const P parent = cur;
      if (sl.frame_ld != FRAME_RUNNING && (BASE_REC || sl.frame_ld != FRAME_BASE)) // a child of the root, or the
        cur = (sl.frame_ld == FRAME_BASE) ? (h.base_type == 0 ? base_pose_from_tail(h, qrow) : get_pose(rec, stride, o_base)) // (FIXED: plan constant)
                                          : get_pose(rec_f, stride, o_frame + 7 * sl.frame_ld); // 2nd.. child of a branch point
      // pose_transform(parent, local), pose.cpp:139-146: position first (it uses the parent's rotation)
      T ang = T(0);
      if (sl.kind != SLOT_FIXED) ang = qrow[sl.qcol];
      if (sl.trn != V_NONE) {
        if (sl.flags & SF_INV_BEFORE) inv = O::qget_inverse(cur.q);
        V lp; // local translation
        lp.x = T(sl.opos[0]);
        lp.y = T(sl.opos[1]);
        lp.z = T(sl.opos[2]);
        if (sl.kind == SLOT_PRISMATIC) {
          // Joint::transform (joint.h:241-247): position axis*d, rotation identity; then
          // pose_transform(origin, .) (src/kinematics.cpp:77-80)
          V d;
          d.x = T(sl.axis[0]) * ang;
          d.y = T(sl.axis[1]) * ang;
          d.z = T(sl.axis[2]) * ang;
          if (sl.flags & SF_PRISM_OQ) {
            Q oq, oi;
            oq.x = T(sl.oquat[0]), oq.y = T(sl.oquat[1]), oq.z = T(sl.oquat[2]), oq.w = T(sl.oquat[3]);
            oi.x = T(sl.oinv[0]), oi.y = T(sl.oinv[1]), oi.z = T(sl.oinv[2]), oi.w = T(sl.oinv[3]);
            d = O::qt_xyz(oq, O::vq_full(d, oi));
          }
          // components whose axis and origin entries are both exact zeros stay exact zeros
          lp.x = d.x + lp.x;
          lp.y = d.y + lp.y;
          lp.z = d.z + lp.z;
        }
        const V r = O::qt_xyz(cur.q, O::vq(sl.trn, lp, inv));
        cur.p.x = r.x + cur.p.x;
        cur.p.y = r.y + cur.p.y;
        cur.p.z = r.z + cur.p.z;
      }
      if (sl.rot != ROT_NONE) {
        Q oq;
        oq.x = T(sl.oquat[0]), oq.y = T(sl.oquat[1]), oq.z = T(sl.oquat[2]), oq.w = T(sl.oquat[3]);
        if (sl.rot == ROT_CONST) {
          cur.q = O::qmul(cur.q, oq);
        } else {
          // Joint::transform for REVOLUTE/CONTINUOUS, joint.h:231-240 (axis is NOT normalised)
          T sn, cs;
#if defined(TFK_SPEC_PLAN) && defined(TFK_SPEC_SINCOS_PAIR) && !defined(TFK_SPEC_ROLLED_WALK)
          if (kept) {
            sn = sn_kept, cs = cs_kept;
            kept = false;
          } else {
            int nx = -1;
            TFK_UNROLL
            for (int t = h.n_slots - 1; t > s; --t)
              if (pv.slots[t].rot >= ROT_REV) nx = t;
            if (nx >= 0) {
              M::sincos2_(ang * T(0.5), qrow[pv.slots[nx].qcol] * T(0.5), &sn, &cs, &sn_kept, &cs_kept);
              kept = true;
            } else {
              M::sincos_(ang * T(0.5), &sn, &cs);
            }
          }
#else
          M::sincos_(ang * T(0.5), &sn, &cs);
#endif
          V ax;
          ax.x = T(sl.axis[0]), ax.y = T(sl.axis[1]), ax.z = T(sl.axis[2]);
          if (sl.rot == ROT_REV)
            cur.q = O::qmul_axis(cur.q, sl.axc, ax, sn, cs);
          else
            cur.q = O::qmul(cur.q, O::qmul_axis(oq, sl.axc, ax, sn, cs));
        }
      }
      if (sl.frame_st >= 0) put_pose(rec_f, stride, o_frame + 7 * sl.frame_st, cur);
      if (sl.arec >= 0) { // src/kinematics.cpp:143-147: axis of the joint in the root frame
        if (sl.flags & SF_INV_AFTER) inv = O::qget_inverse(cur.q);
        V ax;
        ax.x = T(sl.axis[0]), ax.y = T(sl.axis[1]), ax.z = T(sl.axis[2]);
        put_vec(rec, stride, o_arec + 3 * sl.arec, O::qt_xyz(cur.q, O::vq(sl.axc, ax, inv)));
        if (sl.prec >= 0) put_vec(rec, stride, o_prec + 3 * sl.prec, cur.p);
      }
      if (sl.lrec >= 0) sink.link(sl.lrec, cur);
      if (leaf) cur = parent;
    }
    return cur;
  }

  // ---------------------------------------------------------------------------------------------
  // One (configuration, requested link) pair: pose row (src/wrapper.cpp:98-112) and Jacobian
  // (src/kinematics.cpp:105-235).  e: the link's pose; o: index into elink_ids; rec/o_arec/o_prec: the
  // configuration's joint records; bs: where the base context comes from (only touched with a moving
  // base and a Jacobian, and only after the joint columns, to keep few values live).
  // `out` gets put_P(d, value) and put_J(column, values[pose_dim]); columns arrive in increasing
  // order and every entry of J is written.
  // RotationType IGNORE: column c of the position Jacobian (src/kinematics.cpp:131-162, rows 0..2) without a divergent
  // region -- the joint's records are read whether or not it moves this link and a select discards them -- so that the
  // loads and cross products of neighbouring columns overlap.  Same expressions as the general loop of emit_link.
  static TFK_FN void pos_col(const PlanView &pv, const P &e, const uint32_t *mask, int c, const T *rec, int o_arec, int o_prec,
                             T (&v)[7]) {
    v[0] = v[1] = v[2] = v[3] = v[4] = v[5] = v[6] = T(0);
    const PlanCol &pc = pv.cols[c];
    if (pc.arec < 0) return; // the joint moves none of the requested links (uniform)
    const bool relevant = (mask[c >> 5] >> (c & 31)) & 1u;
    const V aw = get_vec(rec, STRIDE, o_arec + 3 * pc.arec);
    V r = aw; // prismatic: the axis itself (src/kinematics.cpp:149-151)
    if (!pc.prismatic) {
      const V cp = get_vec(rec, STRIDE, o_prec + 3 * pc.prec);
      V d;
      d.x = e.p.x - cp.x;
      d.y = e.p.y - cp.y;
      d.z = e.p.z - cp.z;
      r = O::cross(aw, d);
    }
    v[0] = relevant ? r.x : T(0);
    v[1] = relevant ? r.y : T(0);
    v[2] = relevant ? r.z : T(0);
  }

  template <class BaseSrc, class OutSink>
  static TFK_FN void emit_link(const PlanView &pv, const P &e, int o, const T *rec, int stride, int o_arec, int o_prec,
                               BaseSrc &bs, OutSink &out) {
    const PlanHeader &h = *pv.h;
    out.begin();
    out.put_P(0, e.p.x);
    out.put_P(1, e.p.y);
    out.put_P(2, e.p.z);
    V erpy;
    erpy.x = erpy.y = erpy.z = T(0);
    if (ROT == 1) { // src/wrapper.cpp:101-106
      erpy = O::to_rpy(e.q);
      out.put_P(3, erpy.x);
      out.put_P(4, erpy.y);
      out.put_P(5, erpy.z);
    } else if (ROT == 2) { // src/wrapper.cpp:107-112
      out.put_P(3, e.q.x);
      out.put_P(4, e.q.y);
      out.put_P(5, e.q.z);
      out.put_P(6, e.q.w);
    }
    if (!h.with_jacobian) return;

    typename O::RpyCoef coef;
    coef.k_xx = coef.k_xy = coef.s3 = coef.c3 = coef.k_zx = coef.k_zy = T(0);
    Q einv = O::qidentity();
    if (ROT == 1) coef = O::rpy_coef(erpy);  // src/kinematics.cpp:121-123 + :14-21
    if (ROT == 2) einv = O::qinverse(e.q);   // src/kinematics.cpp:124-126

    // joint columns, src/kinematics.cpp:131-180
    const uint32_t *mask = pv.masks + (size_t)o * h.mask_words;
    if (ROT == 0) { // position rows only: two columns per round of the output sink (pos_col: no divergent region)
      TFK_PLAN_LOOP
      for (int c = 0; c + 1 < h.n_cols; c += 2) {
        T v0[7], v1[7];
        pos_col(pv, e, mask, c, rec, o_arec, o_prec, v0);
        pos_col(pv, e, mask, c + 1, rec, o_arec, o_prec, v1);
        out.put_J2(c, v0, v1);
      }
      if (h.n_cols & 1) {
        T v[7];
        pos_col(pv, e, mask, h.n_cols - 1, rec, o_arec, o_prec, v);
        out.put_J(h.n_cols - 1, v);
      }
    }
    TFK_PLAN_LOOP
    for (int c = 0; ROT != 0 && c < h.n_cols; ++c) {
      T v[7];
      v[0] = v[1] = v[2] = v[3] = v[4] = v[5] = v[6] = T(0);
      const bool relevant = (mask[c >> 5] >> (c & 31)) & 1u;
      if (relevant) {
        const PlanCol &pc = pv.cols[c];
        const V aw = get_vec(rec, stride, o_arec + 3 * pc.arec);
        if (pc.prismatic) {
          v[0] = aw.x;
          v[1] = aw.y;
          v[2] = aw.z;
        } else {
          const V cp = get_vec(rec, stride, o_prec + 3 * pc.prec);
          V d;
          d.x = e.p.x - cp.x;
          d.y = e.p.y - cp.y;
          d.z = e.p.z - cp.z;
          const V dpos = O::cross(aw, d);
          v[0] = dpos.x;
          v[1] = dpos.y;
          v[2] = dpos.z;
          if (ROT == 1) {
            const V dr = O::rpy_derivative(coef, aw);
            v[3] = dr.x;
            v[4] = dr.y;
            v[5] = dr.z;
          } else if (ROT == 2) {
            const Q dq = O::q_derivative(einv, aw);
            v[3] = dq.x;
            v[4] = dq.y;
            v[5] = dq.z;
            v[6] = dq.w;
          }
        }
      }
      out.put_J(c, v);
    }

    if (h.base_type == 0) return;
    // base translation columns, src/kinematics.cpp:194-196
    const int n_trans = (h.base_type == 1) ? 2 : 3;
    TFK_PLAN_LOOP
    for (int k = 0; k < n_trans; ++k) out.put_J_unit(h.n_cols + k, k); // a constant pattern: no staging needed
    // base rotation columns: FLOATING keeps roll, pitch, yaw; PLANER keeps yaw only
    if (!ANALYTIC_BASE && h.base_type == 1) {
      emit_planer_yaw(h, e, erpy, bs, out);
      return;
    }
    bs.prepare();
    const V base_p = bs.base_p();
    const int k0 = (h.base_type == 1) ? 2 : 0;
    if (ANALYTIC_BASE) {
      V d;
      d.x = e.p.x - base_p.x;
      d.y = e.p.y - base_p.y;
      d.z = e.p.z - base_p.z;
      TFK_PLAN_LOOP
      for (int k = k0; k < 3; ++k) {
        T v[7];
        v[6] = T(0);
        const V w = bs.w(k);
        const V dpos = O::cross(w, d);
        v[0] = dpos.x;
        v[1] = dpos.y;
        v[2] = dpos.z;
        v[3] = v[4] = v[5] = T(0);
        if (ROT == 1) {
          const V dr = O::rpy_derivative(coef, w);
          v[3] = dr.x;
          v[4] = dr.y;
          v[5] = dr.z;
        } else if (ROT == 2) {
          const Q dq = O::q_derivative(einv, w);
          v[3] = dq.x;
          v[4] = dq.y;
          v[5] = dq.z;
          v[6] = dq.w;
        }
        out.put_J(h.n_cols + n_trans + (k - k0), v);
      }
      return;
    }
    // forward finite difference replica, src/kinematics.cpp:182-232
    const typename M::Den deps = M::den_(T(1e-7));
    P be; // tf_blink_to_elink = pose_transform(tf_blink_to_rlink, tf_rlink_to_elink)
    {
      const P binv = bs.binv();
      if (LREC == 7) {
        be = O::compose(binv, e);
      } else { // IGNORE: only the position rows exist and they never read the link rotation
        const V r = O::rotate(binv.q, e.p);
        be.p.x = r.x + binv.p.x;
        be.p.y = r.y + binv.p.y;
        be.p.z = r.z + binv.p.z;
        be.q = O::qidentity();
      }
    }
    TFK_PLAN_LOOP
    for (int k = k0; k < 3; ++k) { // one copy of the code
      P bk;
      bk.p = base_p;
      bk.q = bs.bk(k);
      V pk;
      Q qk = O::qidentity();
      if (LREC == 7) {
        const P t = O::compose(bk, be);
        pk = t.p;
        qk = t.q;
      } else {
        const V r = O::rotate(bk.q, be.p);
        pk.x = r.x + bk.p.x;
        pk.y = r.y + bk.p.y;
        pk.z = r.z + bk.p.z;
      }
      T v[7];
      v[0] = pk.x - e.p.x;
      v[1] = pk.y - e.p.y;
      v[2] = pk.z - e.p.z;
      v[3] = v[4] = v[5] = v[6] = T(0);
      if (ROT == 1) {
        const V rk = O::to_rpy(qk);
        v[3] = rk.x - erpy.x;
        v[4] = rk.y - erpy.y;
        v[5] = rk.z - erpy.z;
      } else if (ROT == 2) {
        v[3] = qk.x - e.q.x;
        v[4] = qk.y - e.q.y;
        v[5] = qk.z - e.q.z;
        v[6] = qk.w - e.q.w;
      }
      { // (.. - ..) / eps for the D entries of the column, src/kinematics.cpp:212-231
        T num[D];
        TFK_UNROLL
        for (int r = 0; r < D; ++r) num[r] = v[r];
        M::divn_(num, deps);
        TFK_UNROLL
        for (int r = 0; r < D; ++r) v[r] = num[r];
      }
      out.put_J(h.n_cols + n_trans + (k - k0), v);
    }
  }

  // The yaw column of a PLANER base: the forward finite difference of src/kinematics.cpp:182-232 with every base
  // quantity yaw-only (fk_math.h), i.e. half of the products of the two pose_transforms, one atan2 instead of
  // getRPY(base) and one sincos instead of three in setFromRPY(rpy_b + eps*e_z).
  template <class BaseSrc, class OutSink>
  static TFK_FN void emit_planer_yaw(const PlanHeader &h, const P &e, const V &erpy, BaseSrc &bs, OutSink &out) {
    const PlanerCtx c = bs.planer();
    V bep; // tf_blink_to_elink = pose_transform(tf_blink_to_rlink, tf_rlink_to_elink)
    {
      const V r = O::rotate_z(c.binv.q, e.p);
      bep.x = r.x + c.binv.p.x;
      bep.y = r.y + c.binv.p.y;
      bep.z = r.z + c.binv.p.z;
    }
    V pk; // pose_out = pose_transform(tf_rlink_to_blink_tweaked, tf_blink_to_elink)
    {
      const V r = O::rotate_z(c.bk, bep);
      pk.x = r.x + c.base_p.x;
      pk.y = r.y + c.base_p.y;
      pk.z = r.z + c.base_p.z;
    }
    T v[7];
    v[0] = pk.x - e.p.x;
    v[1] = pk.y - e.p.y;
    v[2] = pk.z - e.p.z;
    v[3] = v[4] = v[5] = v[6] = T(0);
    if (LREC == 7) {
      const Q qk = O::qmul_zb(c.bk, O::qmul_zb(c.binv.q, e.q));
      if (ROT == 1) {
        const V rk = O::to_rpy(qk);
        v[3] = rk.x - erpy.x;
        v[4] = rk.y - erpy.y;
        v[5] = rk.z - erpy.z;
      } else {
        v[3] = qk.x - e.q.x;
        v[4] = qk.y - e.q.y;
        v[5] = qk.z - e.q.z;
        v[6] = qk.w - e.q.w;
      }
    }
    {
      T num[D];
      TFK_UNROLL
      for (int r = 0; r < D; ++r) num[r] = v[r];
      M::divn_(num, M::den_(T(1e-7)));
      TFK_UNROLL
      for (int r = 0; r < D; ++r) v[r] = num[r];
    }
    out.put_J(h.n_cols + 2, v);
  }

  // ---------------------------------------------------------------------------------------------
  // two-phase driver, phase A: one thread per configuration
  struct StoreLinks {
    const PlanHeader *h;
    T *rec;
    int stride;
    TFK_FN void link(int lrec, const P &cur) const {
      const int at = h->off_lrec + lrec * LREC;
      put_vec(rec, stride, at, cur.p);
      if (LREC == 7) put_quat(rec, stride, at + 3, cur.q);
    }
  };
  // rec: [base | arec | prec] of this configuration; rec_link: where its frames and link records go (the same array,
  // or a second one biased so that the absolute offsets h.off_lrec + .. still apply: fk_twophase_kernel)
  static TFK_FN void phase_a(const PlanView &pv, const T *qrow, T *rec, int stride) { phase_a(pv, qrow, rec, rec, stride); }
  static TFK_FN void phase_a(const PlanView &pv, const T *qrow, T *rec, T *rec_link, int stride) {
    StoreLinks sink;
    sink.h = pv.h;
    sink.rec = rec_link;
    sink.stride = stride;
    phase_a_sink(pv, qrow, rec, rec_link, stride, sink);
  }
  // the same with the caller's link sink
  template <class LinkSink>
  static TFK_FN void phase_a_sink(const PlanView &pv, const T *qrow, T *rec, T *rec_f, int stride, const LinkSink &sink) {
    const PlanHeader &h = *pv.h;
    const P base = base_pose_of(h, qrow);
    if (h.base_type != 0) put_pose(rec, stride, h.off_base, base); // a FIXED base is a constant of the plan: no record
    if (h.off_arec - h.off_base > 7) store_base_ctx(h, base, rec, stride); // Jacobian with a moving base
    if (h.root_lrec >= 0) sink.link(h.root_lrec, base); // the root link's pose is the base pose (kinematics.cpp:51-58)
    walk<true>(pv, qrow, rec, rec_f, stride, h.off_base, h.off_arec, h.off_prec, h.off_frame, base, sink);
  }

  // two-phase driver, phase B: one thread per (configuration, requested link)
  template <class OutSink> static TFK_FN void phase_b(const PlanView &pv, const T *rec, int stride, int o, OutSink &out) {
    phase_b(pv, rec, rec, stride, o, out);
  }
  template <class OutSink>
  static TFK_FN void phase_b(const PlanView &pv, const T *rec, const T *rec_link, int stride, int o, OutSink &out) {
    phase_b(pv, rec, load_link(pv, rec_link, o), o, out);
  }
  // the pose phase A recorded for output o (the kernel loads the NEXT item's while it works on the current one)
  static TFK_FN P load_link(const PlanView &pv, const T *rec_link, int o) {
    const int lat = pv.h->off_lrec + pv.lrec_of_out[o] * LREC;
    return get_pose_lane(rec_link, lat, LREC == 7);
  }
  template <class OutSink> static TFK_FN void phase_b(const PlanView &pv, const T *rec, const P &e, int o, OutSink &out) {
    const PlanHeader &h = *pv.h;
    BaseFromRec bs;
    bs.rec = rec;
    bs.stride = STRIDE;
    bs.at = h.off_base;
    emit_link(pv, e, o, rec, STRIDE, h.off_arec, h.off_prec, bs, out);
  }

  // ---------------------------------------------------------------------------------------------
  // compute_inter_link_squared_dists (src/wrapper.cpp:189-237) for one (configuration, link pair), on the
  // records of phase A (plan built for the union of the pairs' links, RotationType IGNORE):
  //   sqdist = |p1 - p2|^2,   grad_c = 2 * (J1[:,c] - J2[:,c]) . (p1 - p2)
  // J1/J2 are the position Jacobians get_jacobian(.., IGNORE, with_base) would return -- their columns
  // are produced one at a time (same expressions as emit_link) and never stored.  o1, o2: indices into
  // the plan's elink list.  G points at this pair's row of column 0, columns are ldG apart.
  static TFK_FN V pos_joint_col(const PlanView &pv, const T *rec, const V &ep, int o, int c) {
    const PlanHeader &h = *pv.h;
    V v;
    v.x = v.y = v.z = T(0);
    const uint32_t *mask = pv.masks + (size_t)o * h.mask_words;
    if (!((mask[c >> 5] >> (c & 31)) & 1u)) return v;
    const PlanCol &pc = pv.cols[c];
    const V aw = get_vec(rec, STRIDE, h.off_arec + 3 * pc.arec);
    if (pc.prismatic) return aw;
    const V cp = get_vec(rec, STRIDE, h.off_prec + 3 * pc.prec);
    V d;
    d.x = ep.x - cp.x;
    d.y = ep.y - cp.y;
    d.z = ep.z - cp.z;
    return O::cross(aw, d);
  }
  static TFK_FN T grad_entry(const V &j1, const V &j2, const V &d) {
    const T a0 = j1.x - j2.x, a1 = j1.y - j2.y, a2 = j1.z - j2.z;
    return T(2) * ((a0 * d.x + a1 * d.y) + a2 * d.z);
  }
  static TFK_FN void pair_sqdist(const PlanView &pv, const T *rec, int o1, int o2, bool valid, T *sq, T *G, long long ldG) {
    pair_sqdist(pv, rec, rec, o1, o2, valid, sq, G, ldG);
  }
  // rec: [base | axes | positions] of the configuration; rec_link: its link records (see phase_a)
  static TFK_FN void pair_sqdist(const PlanView &pv, const T *rec, const T *rec_link, int o1, int o2, bool valid, T *sq, T *G,
                                 long long ldG) {
    const PlanHeader &h = *pv.h;
    const V p1 = get_vec(rec_link, STRIDE, h.off_lrec + pv.lrec_of_out[o1] * LREC);
    const V p2 = get_vec(rec_link, STRIDE, h.off_lrec + pv.lrec_of_out[o2] * LREC);
    V d;
    d.x = p1.x - p2.x;
    d.y = p1.y - p2.y;
    d.z = p1.z - p2.z;
    if (valid) *sq = (d.x * d.x + d.y * d.y) + d.z * d.z;
    if (!h.with_jacobian) return;
    TFK_NOUNROLL
    for (int c = 0; c < h.n_cols; ++c) {
      const T g = grad_entry(pos_joint_col(pv, rec, p1, o1, c), pos_joint_col(pv, rec, p2, o2, c), d);
      if (valid) G[c * ldG] = g;
    }
    if (h.base_type == 0) return;
    // base translation columns are the identity for both links (src/kinematics.cpp:194-196): difference 0
    const int n_trans = (h.base_type == 1) ? 2 : 3;
    for (int k = 0; k < n_trans; ++k) {
      V e1;
      e1.x = (k == 0) ? T(1) : T(0);
      e1.y = (k == 1) ? T(1) : T(0);
      e1.z = (k == 2) ? T(1) : T(0);
      const T g = grad_entry(e1, e1, d);
      if (valid) G[(h.n_cols + k) * ldG] = g;
    }
    BaseFromRec bs;
    bs.rec = rec;
    bs.stride = STRIDE;
    bs.at = h.off_base;
    const V base_p = bs.base_p();
    const int k0 = (h.base_type == 1) ? 2 : 0;
    if (ANALYTIC_BASE) {
      V d1, d2;
      d1.x = p1.x - base_p.x, d1.y = p1.y - base_p.y, d1.z = p1.z - base_p.z;
      d2.x = p2.x - base_p.x, d2.y = p2.y - base_p.y, d2.z = p2.z - base_p.z;
      for (int k = k0; k < 3; ++k) {
        const V w = bs.w(k);
        const T g = grad_entry(O::cross(w, d1), O::cross(w, d2), d);
        if (valid) G[(h.n_cols + n_trans + (k - k0)) * ldG] = g;
      }
      return;
    }
    // forward finite difference columns, src/kinematics.cpp:182-232 (position rows only)
    const typename M::Den deps = M::den_(T(1e-7));
    if (h.base_type == 1) { // PLANER: yaw column only, every base quantity yaw-only (fk_math.h, emit_planer_yaw)
      const PlanerCtx c = bs.planer();
      const typename O::QZ ii = O::qz_get_inverse(c.binv.q), ki = O::qz_get_inverse(c.bk);
      V n1, n2;
      {
        const V a1 = O::rotate_z_i(c.binv.q, ii, p1), a2 = O::rotate_z_i(c.binv.q, ii, p2);
        V b1, b2;
        b1.x = a1.x + c.binv.p.x, b1.y = a1.y + c.binv.p.y, b1.z = a1.z + c.binv.p.z;
        b2.x = a2.x + c.binv.p.x, b2.y = a2.y + c.binv.p.y, b2.z = a2.z + c.binv.p.z;
        const V r1 = O::rotate_z_i(c.bk, ki, b1), r2 = O::rotate_z_i(c.bk, ki, b2);
        n1.x = (r1.x + base_p.x) - p1.x, n1.y = (r1.y + base_p.y) - p1.y, n1.z = (r1.z + base_p.z) - p1.z;
        n2.x = (r2.x + base_p.x) - p2.x, n2.y = (r2.y + base_p.y) - p2.y, n2.z = (r2.z + base_p.z) - p2.z;
      }
      T m1[3] = {n1.x, n1.y, n1.z}, m2[3] = {n2.x, n2.y, n2.z};
      M::divn_(m1, deps);
      M::divn_(m2, deps);
      V j1, j2;
      j1.x = m1[0], j1.y = m1[1], j1.z = m1[2];
      j2.x = m2[0], j2.y = m2[1], j2.z = m2[2];
      const T g = grad_entry(j1, j2, d);
      if (valid) G[(h.n_cols + n_trans) * ldG] = g;
      return;
    }
    V be1, be2;
    {
      const P binv = bs.binv();
      const Q bi = O::qget_inverse(binv.q);
      const V r1 = O::qt_xyz(binv.q, O::vq_full(p1, bi)), r2 = O::qt_xyz(binv.q, O::vq_full(p2, bi));
      be1.x = r1.x + binv.p.x, be1.y = r1.y + binv.p.y, be1.z = r1.z + binv.p.z;
      be2.x = r2.x + binv.p.x, be2.y = r2.y + binv.p.y, be2.z = r2.z + binv.p.z;
    }
    TFK_NOUNROLL
    for (int k = k0; k < 3; ++k) {
      const Q bq = bs.bk(k);
      const Q bi = O::qget_inverse(bq);
      const V r1 = O::qt_xyz(bq, O::vq_full(be1, bi)), r2 = O::qt_xyz(bq, O::vq_full(be2, bi));
      T n1[3] = {(r1.x + base_p.x) - p1.x, (r1.y + base_p.y) - p1.y, (r1.z + base_p.z) - p1.z};
      T n2[3] = {(r2.x + base_p.x) - p2.x, (r2.y + base_p.y) - p2.y, (r2.z + base_p.z) - p2.z};
      M::divn_(n1, deps);
      M::divn_(n2, deps);
      V j1, j2;
      j1.x = n1[0], j1.y = n1[1], j1.z = n1[2];
      j2.x = n2[0], j2.y = n2[1], j2.z = n2[2];
      const T g = grad_entry(j1, j2, d);
      if (valid) G[(h.n_cols + n_trans + (k - k0)) * ldG] = g;
    }
  }

  // ---------------------------------------------------------------------------------------------
  // solve_com_forward_kinematics (src/wrapper.cpp:126-179; get_com / get_com_jacobian,
  // src/kinematics.cpp:237-270) on the records of phase A.  The plan's links are the model's "<name>_com"
  // links in id order (weights w[o] = link mass) followed by the caller's action links (w[o] = force /
  // total mass).  One work item produces one 3-vector of the result for one configuration:
  //   item 0                    the centre of mass itself
  //   item 1 + c                column c of the COM Jacobian (joint c)
  //   item 1 + n_cols + k       base rotation column k (roll, pitch, yaw), finite difference as get_jacobian
  // and the reference's order of accumulation is kept per element:
  //   acc = 0; for com links: acc += m * v;  acc /= mass_total;  for action links: acc += f * v;  acc /= vfs
  // (the base translation columns are the constant `diag` times the identity: evaluated on the host).
  struct ComArgs {
    const T *w;     // n_com + n_act weights
    int n_com, n_act;
    T mass_total, vfs;
  };
  static TFK_FN V com_item(const PlanView &pv, const T *rec, int item, const ComArgs &ca) {
    const PlanHeader &h = *pv.h;
    const int kind = (item == 0) ? 0 : ((item <= h.n_cols) ? 1 : 2); // warp-uniform
    const int c = item - 1;
    const int k = item - 1 - h.n_cols;
    BaseFromRec bs;
    bs.rec = rec;
    bs.stride = STRIDE;
    bs.at = h.off_base;
    V base_p, wk;
    P binv;
    Q bi, bq, bqi;
    base_p.x = base_p.y = base_p.z = T(0);
    wk = base_p;
    binv.p = base_p;
    binv.q = bi = bq = bqi = O::qidentity();
    if (kind == 2) {
      base_p = bs.base_p();
      if (ANALYTIC_BASE) {
        wk = bs.w(k);
      } else {
        binv = bs.binv();
        bi = O::qget_inverse(binv.q);
        bq = bs.bk(k);
        bqi = O::qget_inverse(bq);
      }
    }
    const typename M::Den deps = M::den_(T(1e-7));
    V acc;
    acc.x = acc.y = acc.z = T(0);
    const int n_tot = ca.n_com + ca.n_act;
    TFK_NOUNROLL
    for (int o = 0; o < n_tot; ++o) {
      if (o == ca.n_com) {
        T a[3] = {acc.x, acc.y, acc.z};
        M::divn_(a, M::den_(ca.mass_total));
        acc.x = a[0], acc.y = a[1], acc.z = a[2];
      }
      const V p = get_vec(rec, STRIDE, h.off_lrec + pv.lrec_of_out[o] * LREC);
      V v = p;
      if (kind == 1) {
        v = pos_joint_col(pv, rec, p, o, c);
      } else if (kind == 2) {
        if (ANALYTIC_BASE) {
          V d;
          d.x = p.x - base_p.x, d.y = p.y - base_p.y, d.z = p.z - base_p.z;
          v = O::cross(wk, d);
        } else { // src/kinematics.cpp:182-232, position rows
          const V r0 = O::qt_xyz(binv.q, O::vq_full(p, bi));
          V be;
          be.x = r0.x + binv.p.x, be.y = r0.y + binv.p.y, be.z = r0.z + binv.p.z;
          const V r1 = O::qt_xyz(bq, O::vq_full(be, bqi));
          T nn[3] = {(r1.x + base_p.x) - p.x, (r1.y + base_p.y) - p.y, (r1.z + base_p.z) - p.z};
          M::divn_(nn, deps);
          v.x = nn[0], v.y = nn[1], v.z = nn[2];
        }
      }
      const T wo = ca.w[o];
      acc.x = acc.x + wo * v.x;
      acc.y = acc.y + wo * v.y;
      acc.z = acc.z + wo * v.z;
    }
    {
      T a[3] = {acc.x, acc.y, acc.z};
      M::divn_(a, M::den_(ca.n_act > 0 ? ca.vfs : ca.mass_total));
      acc.x = a[0], acc.y = a[1], acc.z = a[2];
    }
    return acc;
  }

  // ---------------------------------------------------------------------------------------------
  // get_total_inertia_matrix (src/kinematics.cpp:272-317) on the records of phase A (plan: the "_com" links in id
  // order, RotationType XYZW so that the link rotations are recorded).  w: 7 reals per link (mass, ixx, ixy, ixz,
  // iyy, iyz, izz).  out[3*i + j] = sum over links of (R I R^T + m (|v|^2 Id - v v^T))_ij, v = p_link - com.
  // The Eigen 3x3 products of the reference are evaluated as (a0*b0 + a1*b1) + a2*b2 (Eigen's order is not
  // pinned in this repo: oracle/eigen_stub).
  static TFK_FN T dot3_(T a0, T b0, T a1, T b1, T a2, T b2) { return (a0 * b0 + a1 * b1) + a2 * b2; }
  static TFK_FN void inertia_total(const PlanView &pv, const T *rec, const T *w, int n_com, T mass_total, T (&out)[9]) {
    const PlanHeader &h = *pv.h;
    V com; // get_com, src/kinematics.cpp:237-253
    com.x = com.y = com.z = T(0);
    TFK_NOUNROLL
    for (int o = 0; o < n_com; ++o) {
      const V p = get_vec(rec, STRIDE, h.off_lrec + pv.lrec_of_out[o] * LREC);
      const T m = w[7 * o];
      com.x = com.x + m * p.x;
      com.y = com.y + m * p.y;
      com.z = com.z + m * p.z;
    }
    {
      T a[3] = {com.x, com.y, com.z};
      M::divn_(a, M::den_(mass_total));
      com.x = a[0], com.y = a[1], com.z = a[2];
    }
    TFK_UNROLL
    for (int k = 0; k < 9; ++k) out[k] = T(0);
    TFK_NOUNROLL
    for (int o = 0; o < n_com; ++o) {
      const int at = h.off_lrec + pv.lrec_of_out[o] * LREC;
      const V p = get_vec(rec, STRIDE, at);
      Q q = O::qidentity();
      if (LREC == 7) q = get_quat(rec, STRIDE, at + 3);
      const T *wl = w + 7 * o;
      const T mass = wl[0];
      const T I[3][3] = {{wl[1], wl[2], wl[3]}, {wl[2], wl[4], wl[5]}, {wl[3], wl[5], wl[6]}};
      const T vx = p.x - com.x, vy = p.y - com.y, vz = p.z - com.z;
      const T xy2 = T(2) * (q.x * q.y), xz2 = T(2) * (q.x * q.z), xw2 = T(2) * (q.x * q.w);
      const T yz2 = T(2) * (q.y * q.z), yw2 = T(2) * (q.y * q.w), zw2 = T(2) * (q.z * q.w);
      const T xx2 = T(2) * (q.x * q.x), yy2 = T(2) * (q.y * q.y), zz2 = T(2) * (q.z * q.z);
      const T R[3][3] = {{T(1) - yy2 - zz2, xy2 - zw2, xz2 + yw2},
                         {xy2 + zw2, T(1) - xx2 - zz2, yz2 - xw2},
                         {xz2 - yw2, yz2 + xw2, T(1) - xx2 - yy2}};
      T RI[3][3];
      TFK_UNROLL
      for (int i = 0; i < 3; ++i)
        TFK_UNROLL
        for (int j = 0; j < 3; ++j) RI[i][j] = dot3_(R[i][0], I[0][j], R[i][1], I[1][j], R[i][2], I[2][j]);
      const T nrm = M::sqrt_((vx * vx + vy * vy) + vz * vz);
      const T n2 = nrm * nrm;
      const T v[3] = {vx, vy, vz};
      TFK_UNROLL
      for (int i = 0; i < 3; ++i)
        TFK_UNROLL
        for (int j = 0; j < 3; ++j) {
          const T rirt = dot3_(RI[i][0], R[j][0], RI[i][1], R[j][1], RI[i][2], R[j][2]);
          const T tt = mass * ((i == j ? n2 : T(0)) - v[i] * v[j]);
          out[3 * i + j] = out[3 * i + j] + (rirt + tt);
        }
    }
  }

  // ---------------------------------------------------------------------------------------------
  // fused driver (n_out == 1): the requested link is the last slot of the pruned chain (or the root),
  // so walk() returns its pose and the same thread emits the outputs; records use the fz_* layout.
  // OutFactory::make() builds the output sink only once the walk is done (nothing of it stays live
  // across the walk).
  struct NoLinks {
    TFK_FN void link(int, const P &) const {}
  };
  template <class OutFactory>
  static TFK_FN void run_fused(const PlanView &pv, const T *qrow, T *rec, int stride, const OutFactory &make_out) {
    const PlanHeader &h = *pv.h;
    const P base = base_pose_of(h, qrow);
    if (h.fz_arec > 0) put_base_fused(h, rec, base); // Jacobian with a moving base: re-read for the base columns
    const P e = walk<false>(pv, qrow, rec, rec, stride, 0, h.fz_arec, h.fz_prec, h.fz_frame, base, NoLinks());
    BaseComputed bs;
    bs.rec = rec;
    bs.stride = stride;
    bs.h = &h;
    typename OutFactory::Sink out = make_out.make();
    emit_link(pv, e, 0, rec, stride, h.fz_arec, h.fz_prec, bs, out);
  }
};

} // namespace tfk
#endif
