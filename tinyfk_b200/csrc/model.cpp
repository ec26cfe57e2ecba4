// model.cpp -- see model.hpp.  Host-side, one-time work per robot; no device code here.
#include "model.hpp"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <limits>
#include <map>
#include <sstream>
#include <stdexcept>

#include <gnu/libc-version.h>

#include "disk_cache.hpp"
#include "xml_lite.hpp"

namespace tfk {

namespace {

typedef Ops<MathPortable> OpsP;
typedef Ops<MathNative64> OpsS;

// urdf::strToDouble (urdf_parser/include/urdf_model/utils.h:73-88): classic-locale stream extraction
// of the whole token.  strtod is what the stream extraction bottoms out in; we only add the
// "nothing but a decimal literal" filter the stream imposes.
double parse_double(const std::string &tok, const char *what) {
  if (tok.empty()) throw std::invalid_argument(std::string("urdf: empty number in ") + what);
  for (char c : tok) {
    bool ok = (c >= '0' && c <= '9') || c == '+' || c == '-' || c == '.' || c == 'e' || c == 'E';
    if (!ok) throw std::invalid_argument(std::string("urdf: '") + tok + "' is not a number (" + what + ")");
  }
  char *end = nullptr;
  double v = std::strtod(tok.c_str(), &end);
  if (end == tok.c_str() || *end != '\0')
    throw std::invalid_argument(std::string("urdf: '") + tok + "' is not a number (" + what + ")");
  return v;
}

// urdf::Vector3::init (pose.h:69-91): split on ' ' only, exactly three non-empty pieces.
Vec3<double> parse_vec3(const std::string &s, const char *what) {
  std::vector<double> v;
  size_t i = 0;
  while (i <= s.size()) {
    size_t j = s.find(' ', i);
    if (j == std::string::npos) j = s.size();
    if (j > i) v.push_back(parse_double(s.substr(i, j - i), what));
    i = j + 1;
  }
  if (v.size() != 3)
    throw std::invalid_argument(std::string("urdf: expected 3 numbers in ") + what + " [" + s + "]");
  Vec3<double> o;
  o.x = v[0];
  o.y = v[1];
  o.z = v[2];
  return o;
}

struct RawJoint {
  std::string name, parent, child;
  int type;
  Vec3<double> axis;
  Vec3<double> oxyz, orpy;
  bool has_limits;
  double lower, upper, velocity, effort;
};

struct RawLink {
  std::string name;
  bool has_inertial;
  Vec3<double> com;
  double mass;
  double inertia[6]; // ixx ixy ixz iyy iyz izz
};

// urdf::parsePose (urdf_parser/src/pose.cpp:96-128): missing attributes leave zeros
void parse_origin(const xml::Element *o, Vec3<double> &xyz, Vec3<double> &rpy, const char *what) {
  xyz.x = xyz.y = xyz.z = 0.0;
  rpy.x = rpy.y = rpy.z = 0.0;
  if (!o) return;
  if (const std::string *a = o->attr("xyz")) xyz = parse_vec3(*a, what);
  if (const std::string *a = o->attr("rpy")) rpy = parse_vec3(*a, what);
}

// urdf::parseLink / parseInertial (urdf_parser/src/link.cpp:269-340, 415-440): name, optional inertial
RawLink parse_link(const xml::Element &e) {
  RawLink l;
  const std::string *name = e.attr("name");
  if (!name) throw std::invalid_argument("urdf: link without a name");
  l.name = *name;
  l.has_inertial = false;
  l.com.x = l.com.y = l.com.z = 0.0;
  l.mass = 0.0;
  for (int i = 0; i < 6; ++i) l.inertia[i] = 0.0;
  if (const xml::Element *in = e.child("inertial")) {
    Vec3<double> rpy;
    parse_origin(in->child("origin"), l.com, rpy, "inertial origin");
    const xml::Element *m = in->child("mass");
    if (!m || !m->attr("value")) throw std::invalid_argument("urdf: inertial of link [" + l.name + "] needs <mass value=>");
    l.mass = parse_double(*m->attr("value"), "mass");
    const xml::Element *it = in->child("inertia");
    if (!it) throw std::invalid_argument("urdf: inertial of link [" + l.name + "] needs <inertia>");
    static const char *k[6] = {"ixx", "ixy", "ixz", "iyy", "iyz", "izz"};
    for (int i = 0; i < 6; ++i) {
      const std::string *a = it->attr(k[i]);
      if (!a) throw std::invalid_argument("urdf: inertia of link [" + l.name + "] misses " + k[i]);
      l.inertia[i] = parse_double(*a, "inertia");
    }
    l.has_inertial = true;
  }
  return l;
}

// urdf::parseJoint (urdf_parser/src/joint.cpp:337-520), the fields the FK path and the limit getters use
RawJoint parse_joint(const xml::Element &e) {
  RawJoint j;
  const std::string *name = e.attr("name");
  if (!name) throw std::invalid_argument("urdf: joint without a name");
  j.name = *name;
  parse_origin(e.child("origin"), j.oxyz, j.orpy, "joint origin");
  if (const xml::Element *p = e.child("parent"))
    if (const std::string *a = p->attr("link")) j.parent = *a;
  if (const xml::Element *c = e.child("child"))
    if (const std::string *a = c->attr("link")) j.child = *a;
  const std::string *type = e.attr("type");
  if (!type) throw std::invalid_argument("urdf: joint [" + j.name + "] has no type");
  if (*type == "planar")
    j.type = JT_PLANAR;
  else if (*type == "floating")
    j.type = JT_FLOATING;
  else if (*type == "revolute")
    j.type = JT_REVOLUTE;
  else if (*type == "continuous")
    j.type = JT_CONTINUOUS;
  else if (*type == "prismatic")
    j.type = JT_PRISMATIC;
  else if (*type == "fixed")
    j.type = JT_FIXED;
  else
    throw std::invalid_argument("urdf: joint [" + j.name + "] has unknown type [" + *type + "]");

  j.axis.x = j.axis.y = j.axis.z = 0.0;
  if (j.type != JT_FLOATING && j.type != JT_FIXED) {
    const xml::Element *ax = e.child("axis");
    if (!ax) {
      j.axis.x = 1.0; // default (1,0,0), joint.cpp:433-436
    } else if (const std::string *a = ax->attr("xyz")) {
      j.axis = parse_vec3(*a, "joint axis");
    } // <axis> without xyz leaves (0,0,0), joint.cpp:438-449
  }

  j.has_limits = false;
  j.lower = j.upper = j.velocity = j.effort = 0.0;
  if (const xml::Element *lim = e.child("limit")) {
    // parseJointLimits, joint.cpp:97-160: lower/upper default 0, effort and velocity required
    if (const std::string *a = lim->attr("lower")) j.lower = parse_double(*a, "limit lower");
    if (const std::string *a = lim->attr("upper")) j.upper = parse_double(*a, "limit upper");
    const std::string *ef = lim->attr("effort"), *ve = lim->attr("velocity");
    if (!ef || !ve) throw std::invalid_argument("urdf: limit of joint [" + j.name + "] needs effort and velocity");
    j.effort = parse_double(*ef, "limit effort");
    j.velocity = parse_double(*ve, "limit velocity");
    j.has_limits = true;
  } else if (j.type == JT_REVOLUTE || j.type == JT_PRISMATIC) {
    throw std::invalid_argument("urdf: joint [" + j.name + "] is revolute/prismatic but has no <limit>");
  }
  return j;
}

} // namespace

// ---- host libm dispatch -----------------------------------------------------------------------
Quat<double> KinematicModel::quat_from_rpy(double r, double p, double y) const {
  return libm_ == LIBM_PORTABLE ? OpsP::from_rpy(r, p, y) : OpsS::from_rpy(r, p, y);
}

Vec3<double> KinematicModel::rpy_from_quat(const Quat<double> &q) const {
  return libm_ == LIBM_PORTABLE ? OpsP::to_rpy(q) : OpsS::to_rpy(q);
}

Pose64 KinematicModel::local_transform(const LinkRec &l, double angle) const {
  if (l.jtype == JT_REVOLUTE || l.jtype == JT_CONTINUOUS)
    return libm_ == LIBM_PORTABLE ? OpsP::local_revolute(l.origin, l.axis, angle)
                                  : OpsS::local_revolute(l.origin, l.axis, angle);
  if (l.jtype == JT_PRISMATIC) return OpsP::local_prismatic(l.origin, l.axis, angle); // no libm inside
  return l.origin;
}

// ---- construction (reference src/tinyfk.cpp:16-114) ---------------------------------------------
// ---- flat tables on disk ("URDF -> flat-table cache on disk") -------------------
// What the constructor derives from the XML -- link / joint records with their ids, origins as quaternions, the
// "<name>_com" links, root and total mass -- is a pure function of (URDF text, host libm).  It is stored under the
// library's disk cache (disk_cache.hpp: private directory, header with 128 bits of key, payload length and hash) and
// a later process with the same text skips the XML reader and the table building.  The blob is still treated as
// input: every count, index and string length is bounds-checked and the tree checks of the parsing path are repeated;
// anything odd means "parse the XML".  TFK_TABLE_CACHE=0 turns it off.
namespace {
const char k_table_magic[8] = {'T', 'F', 'K', 'T', 'A', 'B', 'L', '1'};

struct BlobWriter {
  std::vector<char> b;
  void raw(const void *p, size_t n) { b.insert(b.end(), static_cast<const char *>(p), static_cast<const char *>(p) + n); }
  void i32(int32_t v) { raw(&v, sizeof v); }
  void f64(double v) { raw(&v, sizeof v); }
  void str(const std::string &v) {
    i32((int32_t)v.size());
    raw(v.data(), v.size());
  }
};
struct BlobReader {
  const char *p, *end;
  bool ok = true;
  bool raw(void *out, size_t n) {
    if (!ok || (size_t)(end - p) < n) return ok = false;
    std::memcpy(out, p, n);
    p += n;
    return true;
  }
  int32_t i32() {
    int32_t v = 0;
    raw(&v, sizeof v);
    return v;
  }
  double f64() {
    double v = 0;
    raw(&v, sizeof v);
    return v;
  }
  std::string str() {
    const int32_t n = i32();
    if (!ok || n < 0 || (size_t)(end - p) < (size_t)n) {
      ok = false;
      return std::string();
    }
    std::string v(p, (size_t)n);
    p += n;
    return v;
  }
};

bool table_cache_enabled() {
  const char *e = std::getenv("TFK_TABLE_CACHE");
  return !(e && e[0] == '0');
}
} // namespace

std::vector<char> KinematicModel::save_tables(size_t n_urdf_links) const {
  BlobWriter w;
  w.i32((int32_t)n_urdf_links);
  w.i32((int32_t)links_.size());
  w.i32((int32_t)joints_.size());
  w.i32(root_);
  w.f64(total_mass_);
  for (const LinkRec &l : links_) {
    w.str(l.name);
    w.i32(l.parent);
    w.i32(l.jtype);
    w.i32(l.jid);
    w.raw(&l.axis, sizeof l.axis);
    w.raw(&l.origin, sizeof l.origin);
    w.i32(l.has_inertial ? 1 : 0);
    w.f64(l.mass);
    w.raw(l.inertia, sizeof l.inertia);
  }
  for (const JointRec &j : joints_) {
    w.str(j.name);
    w.i32(j.type);
    w.i32(j.child_link);
    w.i32(j.has_limits ? 1 : 0);
    w.f64(j.lower);
    w.f64(j.upper);
    w.f64(j.velocity);
    w.f64(j.effort);
  }
  return w.b;
}

bool KinematicModel::load_tables(const std::vector<char> &blob) {
  BlobReader r{blob.data(), blob.data() + blob.size()};
  const int32_t n_urdf = r.i32(), n_links = r.i32(), n_joints = r.i32(), root = r.i32();
  const double mass = r.f64();
  if (!r.ok || n_urdf <= 0 || n_links < n_urdf || n_links > (1 << 20) || n_joints < 0 || n_joints > n_urdf ||
      root < 0 || root >= n_urdf)
    return false;
  std::vector<LinkRec> links((size_t)n_links);
  std::vector<JointRec> joints((size_t)n_joints);
  for (LinkRec &l : links) {
    l.name = r.str();
    l.parent = r.i32();
    l.jtype = r.i32();
    l.jid = r.i32();
    r.raw(&l.axis, sizeof l.axis);
    r.raw(&l.origin, sizeof l.origin);
    l.has_inertial = r.i32() != 0;
    l.mass = r.f64();
    r.raw(l.inertia, sizeof l.inertia);
    if (!r.ok || l.parent < -1 || l.parent >= n_links || l.jid < -1 || l.jid >= n_joints || l.jtype < 0 || l.jtype > JT_FIXED)
      return false;
  }
  for (JointRec &j : joints) {
    j.name = r.str();
    j.type = r.i32();
    j.child_link = r.i32();
    j.has_limits = r.i32() != 0;
    j.lower = r.f64();
    j.upper = r.f64();
    j.velocity = r.f64();
    j.effort = r.f64();
    if (!r.ok || j.child_link < 0 || j.child_link >= n_urdf || links[(size_t)j.child_link].jid != (int)(&j - joints.data()))
      return false;
  }
  if (r.p != r.end) return false;
  // the same tree checks as the parsing path: one root, every link reaches it
  for (int32_t i = 0; i < n_links; ++i) {
    if ((links[(size_t)i].parent < 0) != (i == root)) return false;
    int32_t hops = 0;
    for (int l = i; l >= 0; l = links[(size_t)l].parent)
      if (++hops > n_links) return false;
  }
  std::unordered_map<std::string, int> link_ids, joint_ids;
  for (int32_t i = 0; i < n_links; ++i) {
    if (!link_ids.emplace(links[(size_t)i].name, i).second) return false;
    if (links[(size_t)i].parent >= 0) links[(size_t)links[(size_t)i].parent].children.push_back(i); // id order, as built
  }
  for (int32_t i = 0; i < n_joints; ++i)
    if (!joint_ids.emplace(joints[(size_t)i].name, i).second) return false;
  links_.swap(links);
  joints_.swap(joints);
  link_ids_.swap(link_ids);
  joint_ids_.swap(joint_ids);
  com_links_.clear();
  for (int32_t i = n_urdf; i < n_links; ++i) com_links_.push_back(i);
  root_ = root;
  total_mass_ = mass;
  joint_angles_.assign(joints_.size(), 0.0);
  base_pose_.p.x = base_pose_.p.y = base_pose_.p.z = 0.0;
  base_pose_.q = OpsP::qidentity();
  return true;
}

KinematicModel::KinematicModel(const std::string &xml_text, HostLibm libm)
    : root_(-1), total_mass_(0.0), libm_(libm), version_(1), tables_from_cache_(false) {
  if (xml_text.empty()) throw std::runtime_error("xml string is empty"); // src/tinyfk.cpp:17-19
  const bool cache = table_cache_enabled();
  uint64_t key_a = 0, key_b = 0;
  char file[64] = "";
  if (cache) {
    // the origin quaternions come from the host libm: its identity is part of the key
    // ... and so is this translation unit's build (a rebuilt library never trusts tables written by another build)
    const std::string what = std::string("tables v1 built " __DATE__ " " __TIME__ " libm ") +
                             (libm == LIBM_PORTABLE ? std::string("portable " TFK_FKMATH_VERSION) : std::string("system glibc ") + gnu_get_libc_version());
    key_a = diskcache::fnv1a(xml_text, diskcache::fnv1a(what));
    key_b = diskcache::fnv1a(xml_text, diskcache::fnv1a(what, diskcache::k_second_offset));
    std::snprintf(file, sizeof file, "tables_%016llx.tfkt", (unsigned long long)key_a);
    std::vector<char> blob;
    if (diskcache::read(file, k_table_magic, key_a, key_b, blob, 32) && load_tables(blob)) {
      tables_from_cache_ = true;
      return;
    }
  }
  const size_t n_urdf_links = build_from_xml(xml_text);
  if (cache) {
    const std::vector<char> blob = save_tables(n_urdf_links);
    diskcache::write(file, k_table_magic, key_a, key_b, blob.data(), blob.size()); // best effort
  }
}

size_t KinematicModel::build_from_xml(const std::string &xml_text) {
  std::unique_ptr<xml::Element> doc = xml::parse(xml_text);
  if (doc->tag != "robot") throw std::invalid_argument("urdf: could not find the 'robot' element");
  if (!doc->attr("name")) throw std::invalid_argument("urdf: no name given for the robot");
  if (const std::string *v = doc->attr("version")) {
    // urdf_parser/src/model.cpp:124-138: only version 1.0 ("1" or "1.0")
    if (!(*v == "1.0" || *v == "1")) throw std::invalid_argument("urdf: only version 1.0 is supported");
  }

  // std::map => bytewise lexicographic order, duplicates are errors (model.cpp:170-230)
  std::map<std::string, RawLink> raw_links;
  for (const xml::Element *e : doc->children_named("link")) {
    RawLink l = parse_link(*e);
    if (!raw_links.emplace(l.name, l).second) throw std::invalid_argument("urdf: link '" + l.name + "' is not unique");
  }
  if (raw_links.empty()) throw std::invalid_argument("urdf: no link elements found");
  std::map<std::string, RawJoint> raw_joints;
  for (const xml::Element *e : doc->children_named("joint")) {
    RawJoint j = parse_joint(*e);
    if (!raw_joints.emplace(j.name, j).second) throw std::invalid_argument("urdf: joint '" + j.name + "' is not unique");
  }

  // link ids (src/tinyfk.cpp:23-35)
  for (const auto &kv : raw_links) {
    LinkRec l;
    l.name = kv.first;
    l.parent = -1;
    l.jtype = JT_UNKNOWN;
    l.jid = -1;
    l.axis.x = l.axis.y = l.axis.z = 0.0;
    l.origin.p.x = l.origin.p.y = l.origin.p.z = 0.0;
    l.origin.q = OpsP::qidentity();
    l.has_inertial = kv.second.has_inertial;
    l.mass = kv.second.mass;
    for (int i = 0; i < 6; ++i) l.inertia[i] = kv.second.inertia[i];
    link_ids_[l.name] = (int)links_.size();
    links_.push_back(l);
    if (l.has_inertial) total_mass_ += l.mass; // src/tinyfk.cpp:38-44 (id order)
  }

  // joint ids (src/tinyfk.cpp:46-64): movable joints in map order; floating/planar are rejected
  for (const auto &kv : raw_joints) {
    const RawJoint &j = kv.second;
    if (j.type == JT_REVOLUTE || j.type == JT_CONTINUOUS || j.type == JT_PRISMATIC) {
      JointRec r;
      r.name = j.name;
      r.type = j.type;
      r.child_link = -1;
      r.has_limits = j.has_limits;
      r.lower = j.lower;
      r.upper = j.upper;
      r.velocity = j.velocity;
      r.effort = j.effort;
      joint_ids_[r.name] = (int)joints_.size();
      joints_.push_back(r);
    } else if (j.type != JT_FIXED) {
      throw std::invalid_argument("unsuported joint type is detected");
    }
  }

  // tree wiring (urdf_model/model.h:113-156 initTree): joints in map order; a later joint
  // overrides the parent of a child claimed twice
  for (const auto &kv : raw_joints) {
    const RawJoint &j = kv.second;
    if (j.parent.empty() || j.child.empty())
      throw std::invalid_argument("urdf: joint [" + j.name + "] is missing a parent and/or child link");
    auto ci = link_ids_.find(j.child), pi = link_ids_.find(j.parent);
    if (ci == link_ids_.end()) throw std::invalid_argument("urdf: child link [" + j.child + "] of joint [" + j.name + "] not found");
    if (pi == link_ids_.end()) throw std::invalid_argument("urdf: parent link [" + j.parent + "] of joint [" + j.name + "] not found");
    LinkRec &c = links_[ci->second];
    c.parent = pi->second;
    c.jtype = j.type;
    c.axis = j.axis;
    c.origin.p = j.oxyz;
    c.origin.q = quat_from_rpy(j.orpy.x, j.orpy.y, j.orpy.z); // Rotation::init, pose.h:220-226
    auto ji = joint_ids_.find(j.name);
    c.jid = (ji == joint_ids_.end()) ? -1 : ji->second;
    if (c.jid >= 0) joints_[c.jid].child_link = ci->second;
  }
  for (size_t i = 0; i < links_.size(); ++i)
    if (links_[i].parent >= 0) links_[links_[i].parent].children.push_back((int)i);

  // root (model.h:158-185 initRoot): exactly one link without a parent
  for (size_t i = 0; i < links_.size(); ++i) {
    if (links_[i].parent < 0) {
      if (root_ >= 0)
        throw std::invalid_argument("urdf: two root links found: [" + links_[root_].name + "] and [" + links_[i].name + "]");
      root_ = (int)i;
    }
  }
  if (root_ < 0) throw std::invalid_argument("urdf: no root link found, the robot is not a tree");
  // every link must reach the root (a cycle would hang the chain walk)
  for (size_t i = 0; i < links_.size(); ++i) {
    size_t hops = 0;
    for (int l = (int)i; l >= 0; l = links_[l].parent)
      if (++hops > links_.size()) throw std::invalid_argument("urdf: the link graph has a cycle");
  }

  joint_angles_.assign(joints_.size(), 0.0);
  base_pose_.p.x = base_pose_.p.y = base_pose_.p.z = 0.0;
  base_pose_.q = OpsP::qidentity();

  // "<name>_com" dummy links (src/tinyfk.cpp:89-111): position = inertial origin xyz, identity rotation
  const size_t n_urdf_links = links_.size();
  for (size_t i = 0; i < n_urdf_links; ++i) {
    if (!links_[i].has_inertial) continue;
    Pose64 p;
    p.p = raw_links.at(links_[i].name).com;
    p.q = OpsP::qidentity();
    int id = add_link_pose(links_[i].name + "_com", i, p);
    links_[id].has_inertial = true;
    links_[id].mass = links_[i].mass;
    for (int k = 0; k < 6; ++k) links_[id].inertia[k] = links_[i].inertia[k]; // the dummy link shares the parent's inertial
    com_links_.push_back(id);
  }
  return n_urdf_links;
}

// ---- ids / names --------------------------------------------------------------------------------
std::vector<size_t> KinematicModel::get_joint_ids(const std::vector<std::string> &names) const {
  std::vector<size_t> out(names.size());
  for (size_t i = 0; i < names.size(); ++i) {
    auto it = joint_ids_.find(names[i]);
    if (it == joint_ids_.end()) throw std::invalid_argument("no joint named " + names[i]); // src/tinyfk.cpp:157
    out[i] = (size_t)it->second;
  }
  return out;
}

std::vector<size_t> KinematicModel::get_link_ids(const std::vector<std::string> &names) const {
  std::vector<size_t> out(names.size());
  for (size_t i = 0; i < names.size(); ++i) {
    auto it = link_ids_.find(names[i]);
    if (it == link_ids_.end()) throw std::invalid_argument("no link named " + names[i]); // src/tinyfk.cpp:210
    out[i] = (size_t)it->second;
  }
  return out;
}

std::vector<std::string> KinematicModel::get_joint_names() const {
  std::vector<std::string> out;
  for (const JointRec &j : joints_) out.push_back(j.name);
  return out;
}

std::vector<std::string> KinematicModel::get_link_names() const {
  std::vector<std::string> out;
  for (const LinkRec &l : links_) out.push_back(l.name);
  return out;
}

// ---- limits -------------------------------------------------------------------------------------
static void check_joint_ids(const std::vector<size_t> &ids, size_t n) {
  for (size_t id : ids)
    if (id >= n) throw std::invalid_argument("joint id " + std::to_string(id) + " out of range");
}

std::vector<std::pair<double, double>>
KinematicModel::get_joint_position_limits(const std::vector<size_t> &ids) const {
  check_joint_ids(ids, joints_.size());
  std::vector<std::pair<double, double>> out(ids.size());
  const double inf = std::numeric_limits<double>::infinity();
  for (size_t i = 0; i < ids.size(); ++i) {
    const JointRec &j = joints_[ids[i]];
    if (j.type == JT_CONTINUOUS)
      out[i] = std::make_pair(-inf, inf);
    else
      out[i] = std::make_pair(j.lower, j.upper);
  }
  return out;
}

std::vector<double> KinematicModel::get_joint_velocity_limits(const std::vector<size_t> &ids) const {
  check_joint_ids(ids, joints_.size());
  std::vector<double> out(ids.size());
  for (size_t i = 0; i < ids.size(); ++i) {
    const JointRec &j = joints_[ids[i]];
    if (!j.has_limits) throw std::invalid_argument("joint " + j.name + " has no <limit>"); // reference: null deref
    out[i] = j.velocity;
  }
  return out;
}

std::vector<double> KinematicModel::get_joint_effort_limits(const std::vector<size_t> &ids) const {
  check_joint_ids(ids, joints_.size());
  std::vector<double> out(ids.size());
  for (size_t i = 0; i < ids.size(); ++i) {
    const JointRec &j = joints_[ids[i]];
    if (!j.has_limits) throw std::invalid_argument("joint " + j.name + " has no <limit>");
    out[i] = j.effort;
  }
  return out;
}

// ---- state --------------------------------------------------------------------------------------
void KinematicModel::set_joint_angles(const std::vector<size_t> &ids, const std::vector<double> &angles) {
  if (ids.size() != angles.size()) throw std::invalid_argument("joint id / angle count mismatch");
  check_joint_ids(ids, joints_.size());
  bool changed = false; // an unchanged state keeps the version: plans built from it stay valid (Engine::plan_for)
  for (size_t i = 0; i < ids.size(); ++i) {
    changed |= std::memcmp(&joint_angles_[ids[i]], &angles[i], sizeof(double)) != 0;
    joint_angles_[ids[i]] = angles[i];
  }
  if (changed) ++version_;
}

std::vector<double> KinematicModel::get_joint_angles(const std::vector<size_t> &ids) const {
  check_joint_ids(ids, joints_.size());
  std::vector<double> out(ids.size());
  for (size_t i = 0; i < ids.size(); ++i) out[i] = joint_angles_[ids[i]];
  return out;
}

void KinematicModel::set_init_angles() {
  joint_angles_.assign(joints_.size(), 0.0);
  ++version_;
}

void KinematicModel::set_base_pose_xyzrpy(const double b[6]) { // src/wrapper.cpp:25-35
  Pose64 nb;
  nb.p.x = b[0];
  nb.p.y = b[1];
  nb.p.z = b[2];
  nb.q = quat_from_rpy(b[3], b[4], b[5]);
  if (std::memcmp(&nb, &base_pose_, sizeof nb) != 0) ++version_; // see set_joint_angles
  base_pose_ = nb;
}

void KinematicModel::get_base_pose_xyzrpy(double b[6]) const { // src/wrapper.cpp:45-55
  b[0] = base_pose_.p.x;
  b[1] = base_pose_.p.y;
  b[2] = base_pose_.p.z;
  Vec3<double> rpy = rpy_from_quat(base_pose_.q);
  b[3] = rpy.x;
  b[4] = rpy.y;
  b[5] = rpy.z;
}

// ---- structure edits ------------------------------------------------------------------------------
int KinematicModel::add_link_pose(const std::string &name, size_t parent_id, const Pose64 &pose) {
  if (link_ids_.find(name) != link_ids_.end())
    throw std::invalid_argument("link name : " + name + " already exists"); // src/tinyfk.cpp:233-236
  if (parent_id >= links_.size()) throw std::invalid_argument("parent link id out of range");
  LinkRec l;
  l.name = name;
  l.parent = (int)parent_id;
  l.jtype = JT_FIXED;
  l.jid = -1;
  l.axis.x = l.axis.y = l.axis.z = 0.0;
  l.origin = pose;
  l.has_inertial = false;
  l.mass = 0.0;
  for (int k = 0; k < 6; ++k) l.inertia[k] = 0.0;
  int id = (int)links_.size();
  link_ids_[name] = id;
  links_.push_back(l);
  links_[parent_id].children.push_back(id);
  ++version_;
  return id;
}

int KinematicModel::add_new_link(const std::string &name, size_t parent_id, const double pos[3],
                                 const double rpy[3]) { // src/tinyfk.cpp:217-227
  Pose64 p;
  p.p.x = pos[0];
  p.p.y = pos[1];
  p.p.z = pos[2];
  p.q = quat_from_rpy(rpy[0], rpy[1], rpy[2]);
  return add_link_pose(name, parent_id, p);
}

bool KinematicModel::is_relevant(size_t link_id, size_t joint_id) const {
  if (link_id >= links_.size() || joint_id >= joints_.size()) throw std::invalid_argument("id out of range");
  const int target = joints_[joint_id].child_link;
  for (int l = (int)link_id; l >= 0; l = links_[l].parent)
    if (l == target) return true;
  return false;
}

std::string load_urdf(const std::string &path) {
  std::ifstream f(path.c_str(), std::ios::in | std::ios::binary);
  if (!f) throw std::runtime_error("cannot open " + path);
  std::ostringstream ss;
  ss << f.rdbuf();
  return ss.str();
}

} // namespace tfk
