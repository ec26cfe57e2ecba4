// plan.cpp -- see plan.hpp / fk_plan.h.
#include "plan.hpp"

#include <cstdio>
#include <initializer_list>

#include <cstring>
#include <stdexcept>
#include <string>

namespace tfk {

namespace {

struct Builder {
  const KinematicModel &m;
  const PlanKey &key;
  std::vector<char> needed;       // per link
  std::vector<int> qcol_of_joint; // per joint id: last column that names it, or -1
  std::vector<int> slot_of_link;  // per link
  std::vector<PlanSlot> slots;
  std::vector<int> parent_slot;   // per slot: slot of the parent link, -1 when the parent is the root
  int frames_live, frames_max;
  bool leaf_mode; // several requested links: fixed leaves are computed on the side (SF_LEAF); the fused single-link
                  // kernel returns the running frame as the link's pose and keeps the plain order

  Builder(const KinematicModel &model, const PlanKey &k)
      : m(model), key(k), needed(model.n_links(), 0), qcol_of_joint(model.n_joints(), -1),
        slot_of_link(model.n_links(), -1), frames_live(0), frames_max(0), leaf_mode(k.elink_ids.size() > 1) {}

  static int vclass(double x, double y, double z) {
    const int nz = (x != 0.0) + (y != 0.0) + (z != 0.0);
    if (nz == 0) return V_NONE;
    if (nz > 1) return V_FULL;
    return x != 0.0 ? V_X : (y != 0.0 ? V_Y : V_Z);
  }

  // sparsity classes of the slot's constants (fk_plan.h); every test is an exact comparison
  static void classify(PlanSlot &s) {
    const bool oq_ident = s.oquat[0] == 0.0 && s.oquat[1] == 0.0 && s.oquat[2] == 0.0 && s.oquat[3] == 1.0;
    s.axc = vclass(s.axis[0], s.axis[1], s.axis[2]);
    s.flags = 0;
    {
      // GetInverse(oquat), pose.h:284-299, with the host's own IEEE arithmetic (same bits as the device)
      Quat<double> oq;
      oq.x = s.oquat[0], oq.y = s.oquat[1], oq.z = s.oquat[2], oq.w = s.oquat[3];
      const Quat<double> oi = Ops<MathPortable>::qget_inverse(oq);
      s.oinv[0] = oi.x, s.oinv[1] = oi.y, s.oinv[2] = oi.z, s.oinv[3] = oi.w;
    }
    if (s.kind == SLOT_FIXED) {
      s.rot = oq_ident ? ROT_NONE : ROT_CONST;
      s.trn = vclass(s.opos[0], s.opos[1], s.opos[2]);
    } else if (s.kind == SLOT_REVOLUTE) {
      s.rot = oq_ident ? ROT_REV : ROT_REV_OQ;
      s.trn = vclass(s.opos[0], s.opos[1], s.opos[2]);
    } else { // prismatic: local = pose_transform(origin, (identity, axis*d))
      s.rot = oq_ident ? ROT_NONE : ROT_CONST;
      if (oq_ident) {
        // rotate(identity, v) == v exactly, so local.p[k] = axis[k]*d + opos[k]
        s.trn = vclass((s.opos[0] != 0.0) || (s.axis[0] != 0.0), (s.opos[1] != 0.0) || (s.axis[1] != 0.0),
                       (s.opos[2] != 0.0) || (s.axis[2] != 0.0));
      } else {
        s.trn = V_FULL;
        s.flags |= SF_PRISM_OQ;
      }
    }
  }

  void visit(int u, int frame_ld) {
    const LinkRec &l = m.links()[u];
    PlanSlot s;
    std::memset(&s, 0, sizeof s);
    s.qcol = -1;
    s.frame_ld = frame_ld;
    s.frame_st = -1;
    s.arec = -1;
    s.prec = -1;
    s.lrec = -1;
    const bool selected = l.jid >= 0 && qcol_of_joint[l.jid] >= 0;
    Pose64 local;
    if (selected) {
      s.kind = (l.jtype == JT_PRISMATIC) ? SLOT_PRISMATIC : SLOT_REVOLUTE;
      s.qcol = qcol_of_joint[l.jid];
      local = l.origin;
    } else {
      // not driven by the input: the joint sits at the model's stored angle (src/tinyfk.cpp:122-128)
      s.kind = SLOT_FIXED;
      local = (l.jid >= 0) ? m.local_transform(l, m.joint_angles()[l.jid]) : l.origin;
    }
    s.opos[0] = local.p.x;
    s.opos[1] = local.p.y;
    s.opos[2] = local.p.z;
    s.oquat[0] = local.q.x;
    s.oquat[1] = local.q.y;
    s.oquat[2] = local.q.z;
    s.oquat[3] = local.q.w;
    s.axis[0] = l.axis.x;
    s.axis[1] = l.axis.y;
    s.axis[2] = l.axis.z;
    classify(s);

    // needed children: fixed leaves first (computed on the side, SF_LEAF: they leave the running frame alone), then
    // the subtrees; only a link with two or more SUBTREES has to keep its frame
    std::vector<int> leaves, inner;
    for (int c : l.children) {
      if (!needed[c]) continue;
      bool has_needed_child = false;
      for (int g : m.links()[c].children)
        if (needed[g]) has_needed_child = true;
      const LinkRec &lc = m.links()[c];
      const bool driven = lc.jid >= 0 && qcol_of_joint[lc.jid] >= 0;
      (leaf_mode && !has_needed_child && !driven ? leaves : inner).push_back(c);
    }
    if (inner.size() >= 2) {
      s.frame_st = frames_live++;
      if (frames_live > frames_max) frames_max = frames_live;
    }
    slot_of_link[u] = (int)slots.size();
    slots.push_back(s);
    parent_slot.push_back(l.parent >= 0 ? slot_of_link[l.parent] : -1);
    const int my_frame = s.frame_st;
    for (int c : leaves) {
      visit(c, FRAME_RUNNING);
      slots[slot_of_link[c]].flags |= SF_LEAF;
    }
    bool first = true;
    for (int c : inner) {
      visit(c, first ? (int)FRAME_RUNNING : my_frame);
      first = false;
    }
    if (my_frame >= 0) --frames_live;
  }
};

} // namespace

std::vector<unsigned char> build_plan(const KinematicModel &model, const PlanKey &key) {
  const int n_out = (int)key.elink_ids.size();
  const int n_cols = (int)key.joint_ids.size();
  if (key.rot_type < 0 || key.rot_type > 2) throw std::invalid_argument("rot_type must be 0 (IGNORE), 1 (RPY) or 2 (XYZW)");
  if (key.base_type < 0 || key.base_type > 2) throw std::invalid_argument("base_type must be 0 (FIXED), 1 (PLANER) or 2 (FLOATING)");
  for (int id : key.elink_ids)
    if (id < 0 || (size_t)id >= model.n_links()) throw std::invalid_argument("link id " + std::to_string(id) + " out of range");
  for (int id : key.joint_ids)
    if (id < 0 || (size_t)id >= model.n_joints()) throw std::invalid_argument("joint id " + std::to_string(id) + " out of range");

  Builder b(model, key);
  const std::vector<LinkRec> &links = model.links();
  const int root = model.root_link_id();
  for (int id : key.elink_ids)
    for (int l = id; l >= 0 && l != root; l = links[l].parent) {
      if (b.needed[l]) break;
      b.needed[l] = 1;
    }
  for (int c = 0; c < n_cols; ++c) b.qcol_of_joint[key.joint_ids[c]] = c; // last one wins

  for (int c : links[root].children)
    if (b.needed[c]) b.visit(c, FRAME_BASE);
  if (b.frames_max > TFK_MAX_FRAMES)
    throw std::invalid_argument("kinematic tree branches too deeply for the kernel's frame file (" +
                                std::to_string(b.frames_max) + " > " + std::to_string((int)TFK_MAX_FRAMES) + ")");

  const bool jac = key.with_jacobian != 0;
  // joint records: a world-axis record per selected joint whose child link is on a requested chain and
  // a position record per distinct joint position (a joint whose local translation is exactly zero sits
  // where its parent link sits: o.p = 0 + a.p, so co-located joints share one record)
  int n_arec = 0, n_prec = 0;
  std::vector<int> arec_of_joint(model.n_joints(), -1), prec_of_joint(model.n_joints(), -1);
  std::vector<PlanCol> cols(n_cols);
  if (jac) {
    const int ns = (int)b.slots.size();
    std::vector<int> cls(ns), prec_of_cls(ns, -1);
    for (int s = 0; s < ns; ++s)
      cls[s] = (b.slots[s].trn == V_NONE && b.parent_slot[s] >= 0) ? cls[b.parent_slot[s]] : s;
    for (int s = 0; s < ns; ++s) {
      PlanSlot &sl = b.slots[s];
      if (sl.kind == SLOT_FIXED) continue;
      sl.arec = n_arec++;
      if (sl.kind == SLOT_REVOLUTE) {
        if (prec_of_cls[cls[s]] < 0) {
          prec_of_cls[cls[s]] = n_prec++;
          sl.prec = prec_of_cls[cls[s]]; // the first joint of the class writes the record
        }
      }
    }
    for (size_t l = 0; l < links.size(); ++l) {
      const int s = b.slot_of_link[l];
      if (s < 0 || b.slots[s].arec < 0) continue;
      arec_of_joint[links[l].jid] = b.slots[s].arec;
      prec_of_joint[links[l].jid] = b.slots[s].kind == SLOT_REVOLUTE ? prec_of_cls[cls[s]] : -1;
    }
  }
  for (int c = 0; c < n_cols; ++c) {
    cols[c].arec = jac ? arec_of_joint[key.joint_ids[c]] : -1;
    cols[c].prec = jac ? prec_of_joint[key.joint_ids[c]] : -1;
    cols[c].prismatic = model.joints()[key.joint_ids[c]].type == JT_PRISMATIC ? 1 : 0;
    cols[c].pad_ = 0;
  }

  // running-inverse bookkeeping: `inv` = GetInverse(q of the running frame) is computed only where a
  // vector rotation needs it and it is not already there (the reference recomputes it in every
  // Rotation::operator*(Vector3), pose.h:265-282; same operands, same bits)
  {
    bool valid = false;
    for (PlanSlot &sl : b.slots) {
      if (sl.frame_ld != FRAME_RUNNING) valid = false;
      if (sl.trn != V_NONE) {
        if (!valid) sl.flags |= SF_INV_BEFORE;
        valid = true;
      }
      if (sl.flags & SF_LEAF) continue; // computed on the side: the running frame (and its inverse) stay the parent's
      if (sl.rot != ROT_NONE) valid = false;
      if (sl.arec >= 0) {
        if (!valid) sl.flags |= SF_INV_AFTER;
        valid = true;
      }
    }
  }

  // link records: one per distinct requested link
  int n_lrec = 0, root_lrec = -1;
  std::vector<int> lrec_of_out(n_out, -1);
  for (int o = 0; o < n_out; ++o) {
    const int id = key.elink_ids[o];
    if (id == root) {
      if (root_lrec < 0) root_lrec = n_lrec++;
      lrec_of_out[o] = root_lrec;
    } else {
      PlanSlot &s = b.slots[b.slot_of_link[id]];
      if (s.lrec < 0) s.lrec = n_lrec++;
      lrec_of_out[o] = s.lrec;
    }
  }

  const int mask_words = (n_cols + 31) / 32 > 0 ? (n_cols + 31) / 32 : 1;
  std::vector<uint32_t> masks((size_t)n_out * mask_words, 0u);
  if (jac)
    for (int o = 0; o < n_out; ++o)
      for (int c = 0; c < n_cols; ++c)
        if (model.is_relevant(key.elink_ids[o], key.joint_ids[c]))
          masks[(size_t)o * mask_words + c / 32] |= (1u << (c % 32));

  PlanHeader h;
  std::memset(&h, 0, sizeof h);
  h.n_slots = (int)b.slots.size();
  h.n_out = n_out;
  h.n_cols = n_cols;
  h.n_arec = n_arec;
  h.n_prec = n_prec;
  h.n_lrec = n_lrec;
  h.rot_type = key.rot_type;
  h.base_type = key.base_type;
  h.pose_dim = pose_dim_of(key.rot_type);
  h.with_jacobian = jac ? 1 : 0;
  h.n_in = n_cols + base_dof_of(key.base_type);
  h.n_out_cols = h.n_in;
  // IGNORE needs positions only -- also for the base columns, whose position rows never read the
  // link's rotation (src/kinematics.cpp:189-219 with pose.cpp:139-146)
  h.lrec_size = key.rot_type == ROT_IGNORE ? 3 : 7;
  h.off_base = 0;
  // the base pose always has a record (7 reals; children of the root re-read it); a Jacobian with a moving
  // base adds the inverse pose and the tweaked rotations of the finite-difference columns
  // (a FIXED base is a constant of the plan, h.base_pose: no record at all)
  const int base_rec = key.base_type == BASE_FIXED ? 0 : (jac ? (int)TFK_BASE_REC : 7);
  h.off_arec = h.off_base + base_rec;
  h.off_prec = h.off_arec + 3 * n_arec;
  // [base | arec | prec] are what every (configuration, link) item re-reads (small: shared memory when it fits);
  // branch-point frames and link records come last (block-private scratch in L2)
  // (off_frame, rec_size and fz_size are even: the pair layout of the plan-specialised kernels, fk_core.h "record layout",
  // splits the records between shared memory and the scratch at a pair boundary)
  h.off_frame = (h.off_prec + 3 * n_prec + 1) & ~1;
  h.off_lrec = h.off_frame + 7 * b.frames_max;
  h.rec_size = (h.off_lrec + h.lrec_size * n_lrec + 1) & ~1;
  // fused single-link kernel: the link pose and the base context stay in registers
  // base pose, re-read for the base columns: 7 reals, or the 4 that a PLANER pose has (x, y, sin, cos of yaw / 2)
  h.fz_arec = (jac && key.base_type != BASE_FIXED) ? (key.base_type == BASE_PLANER ? 4 : 7) : 0;
  h.fz_prec = h.fz_arec + 3 * n_arec;
  h.fz_frame = h.fz_prec + 3 * n_prec;
  h.fz_size = (h.fz_frame + 7 * b.frames_max + 1) & ~1;
  h.root_lrec = root_lrec;
  h.mask_words = mask_words;
  h.n_frames = b.frames_max;
  const Pose64 &bp = model.base_pose();
  h.base_pose[0] = bp.p.x;
  h.base_pose[1] = bp.p.y;
  h.base_pose[2] = bp.p.z;
  h.base_pose[3] = bp.q.x;
  h.base_pose[4] = bp.q.y;
  h.base_pose[5] = bp.q.z;
  h.base_pose[6] = bp.q.w;

  size_t off = sizeof(PlanHeader);
  h.off_slots = (int)off;
  off += sizeof(PlanSlot) * b.slots.size();
  off = (off + 15) & ~(size_t)15; // the two-phase kernel copies [outs | masks] to shared memory with 16-byte loads
  h.off_outs = (int)off;
  off += sizeof(int32_t) * n_out;
  h.off_masks = (int)off;
  off += sizeof(uint32_t) * masks.size();
  h.off_cols = (int)off;
  off += sizeof(PlanCol) * n_cols;
  off = (off + 15) & ~(size_t)15;
  h.total_bytes = (int)off;

  std::vector<unsigned char> blob(off, 0);
  std::memcpy(blob.data(), &h, sizeof h);
  if (!b.slots.empty()) std::memcpy(blob.data() + h.off_slots, b.slots.data(), sizeof(PlanSlot) * b.slots.size());
  if (n_out) std::memcpy(blob.data() + h.off_outs, lrec_of_out.data(), sizeof(int32_t) * n_out);
  if (!masks.empty()) std::memcpy(blob.data() + h.off_masks, masks.data(), sizeof(uint32_t) * masks.size());
  if (n_cols) std::memcpy(blob.data() + h.off_cols, cols.data(), sizeof(PlanCol) * n_cols);
  return blob;
}

std::unique_ptr<PlanParams> make_plan_params(const std::vector<unsigned char> &blob) {
  const PlanHeader *h = reinterpret_cast<const PlanHeader *>(blob.data());
  if (h->n_slots > (int)TFK_PSLOTS)
    throw std::invalid_argument("request needs " + std::to_string(h->n_slots) + " links of the tree; the kernels take at most " +
                                std::to_string((int)TFK_PSLOTS));
  if (h->n_cols > (int)TFK_PCOLS)
    throw std::invalid_argument("request names " + std::to_string(h->n_cols) + " joints; the kernels take at most " +
                                std::to_string((int)TFK_PCOLS));
  std::unique_ptr<PlanParams> p(new PlanParams);
  std::memset(p.get(), 0, sizeof(PlanParams));
  p->h = *h;
  if (h->n_slots) std::memcpy(p->slots, blob.data() + h->off_slots, sizeof(PlanSlot) * h->n_slots);
  if (h->n_cols) std::memcpy(p->cols, blob.data() + h->off_cols, sizeof(PlanCol) * h->n_cols);
  if (h->n_out >= 1) std::memcpy(p->mask0, blob.data() + h->off_masks, sizeof(uint32_t) * (h->mask_words < 2 ? h->mask_words : 2));
  return p;
}

namespace {
void put_real(std::string &o, double v) {
  char b[64];
  std::snprintf(b, sizeof b, "%a", v);
  o += b;
}
template <int N> void put_reals(std::string &o, const double (&v)[N]) {
  o += "{";
  for (int i = 0; i < N; ++i) {
    if (i) o += ", ";
    put_real(o, v[i]);
  }
  o += "}";
}
void put_ints(std::string &o, std::initializer_list<int> v) {
  for (int x : v) {
    o += ", ";
    o += std::to_string(x);
  }
}
} // namespace

std::string plan_to_source(const std::vector<unsigned char> &blob) {
  const PlanHeader &h = *reinterpret_cast<const PlanHeader *>(blob.data());
  const PlanSlot *slots = reinterpret_cast<const PlanSlot *>(blob.data() + h.off_slots);
  const PlanCol *cols = reinterpret_cast<const PlanCol *>(blob.data() + h.off_cols);
  const uint32_t *masks = reinterpret_cast<const uint32_t *>(blob.data() + h.off_masks);
  // several links: the per-output tables (link record of every output, relevance masks) are indexed per lane and stay
  // a run-time table (KernelArgs::tables); k_mask0 is the first output's mask (all the fused kernel needs)
  if (h.mask_words > 2) throw std::invalid_argument("plan specialisation: more than 64 joints");
  std::string o = "namespace tfk { namespace spec {\n";
  // PlanHeader, in declaration order (fk_plan.h)
  o += "static __device__ constexpr PlanHeader k_h = {" + std::to_string(h.n_slots);
  put_ints(o, {h.n_out, h.n_cols, h.n_arec, h.n_prec, h.n_lrec, h.rot_type, h.base_type, h.pose_dim, h.with_jacobian, h.n_in,
               h.n_out_cols, h.lrec_size, h.off_base, h.off_arec, h.off_prec, h.off_lrec, h.off_frame, h.rec_size, h.fz_arec,
               h.fz_prec, h.fz_frame, h.fz_size, h.root_lrec, h.mask_words, h.off_slots, h.off_outs, h.off_masks, h.off_cols,
               h.total_bytes, h.n_frames, h.pad_});
  o += ", ";
  put_reals(o, h.base_pose);
  o += "};\n";
  o += "static __constant__ constexpr PlanSlot k_slots[" + std::to_string(std::max(h.n_slots, 1)) + "] = {\n";
  for (int s = 0; s < h.n_slots; ++s) {
    const PlanSlot &sl = slots[s];
    o += "  {";
    put_reals(o, sl.opos);
    o += ", ";
    put_reals(o, sl.oquat);
    o += ", ";
    put_reals(o, sl.axis);
    o += ", ";
    put_reals(o, sl.oinv);
    put_ints(o, {sl.kind, sl.qcol, sl.frame_ld, sl.frame_st, sl.arec, sl.prec, sl.lrec, sl.rot, sl.trn, sl.axc, sl.flags, sl.pad_});
    o += "},\n";
  }
  if (h.n_slots == 0) o += "  {}\n";
  o += "};\n";
  o += "static __constant__ constexpr PlanCol k_cols[" + std::to_string(std::max(h.n_cols, 1)) + "] = {\n";
  for (int c = 0; c < h.n_cols; ++c)
    o += "  {" + std::to_string(cols[c].arec) + ", " + std::to_string(cols[c].prec) + ", " + std::to_string(cols[c].prismatic) + ", 0},\n";
  if (h.n_cols == 0) o += "  {}\n";
  o += "};\n";
  o += "static __device__ constexpr uint32_t k_mask0[2] = {" + std::to_string(h.mask_words > 0 ? masks[0] : 0u) + "u, " +
       std::to_string(h.mask_words > 1 ? masks[1] : 0u) + "u};\n";
  o += "}} // namespace tfk::spec\n";
  return o;
}

InertiaTables build_inertia_tables(const KinematicModel &model) {
  InertiaTables t;
  const std::vector<int> &cl = model.com_links();
  if (cl.empty()) throw std::invalid_argument("the model has no link with an <inertial>: inertia matrix undefined");
  t.links.assign(cl.begin(), cl.end());
  t.mass_total = 0.0;
  for (int id : cl) {
    const LinkRec &l = model.links()[id];
    t.w.push_back(l.mass);
    for (int k = 0; k < 6; ++k) t.w.push_back(l.inertia[k]);
    t.mass_total += l.mass;
  }
  return t;
}

ComTables build_com_tables(const KinematicModel &model, const int *action_link_ids, const double *forces, int n_action) {
  ComTables t;
  const std::vector<int> &cl = model.com_links();
  if (cl.empty()) throw std::invalid_argument("the model has no link with an <inertial>: centre of mass undefined");
  if (n_action < 0 || (n_action > 0 && (!action_link_ids || !forces))) throw std::invalid_argument("bad action link / force arrays");
  t.links.assign(cl.begin(), cl.end());
  t.n_com = (int)cl.size();
  t.n_act = n_action;
  t.mass_total = 0.0;
  t.diag = 0.0;
  for (int id : cl) {
    const double m = model.links()[id].mass;
    t.w.push_back(m);
    t.mass_total += m;
    t.diag = t.diag + m * 1.0; // jac_average(k, n_joint + k) += mass * 1.0 (kinematics.cpp:194-196, 263)
  }
  t.diag = t.diag / t.mass_total;
  t.vfs = 1.0;
  for (int j = 0; j < n_action; ++j) {
    const double f = forces[j] / model.total_mass(); // wrapper.cpp:156
    t.links.push_back(action_link_ids[j]);
    t.w.push_back(f);
    t.vfs += f;
    t.diag = t.diag + f * 1.0;
  }
  if (n_action > 0) t.diag = t.diag / t.vfs;
  return t;
}

} // namespace tfk
