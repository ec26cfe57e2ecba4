/*
 * fk_oracle.c -- TEST INFRASTRUCTURE (see fk_oracle.h).  CPU restatement, in plain C, of the
 * reference algorithm for the batched FK + Jacobian path.  Each function names the reference code it
 * follows (paths relative to the tinyfk tree).  Written independently of the product's fk_math.h /
 * fk_core.h: it keeps the reference's own structure (leaf->root walk with a stack and a per-config
 * transform cache, full Hamilton products including the multiplications by zero) so that it checks,
 * rather than shares, the product's shortcuts.
 * libm: the platform's, or tinyfk_b200/csrc/fkmath.h when linked with strict_libm.c (-Bsymbolic).
 */
#include "fk_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef struct { double x, y, z; } vec3;
typedef struct { double x, y, z, w; } quat;
typedef struct { vec3 p; quat q; } pose;

enum { J_REVOLUTE = 1, J_CONTINUOUS = 2, J_PRISMATIC = 3, J_FIXED = 6 };

struct fko_model {
  int n_links, n_joints, root;
  int *parent, *jtype, *jid;
  vec3 *axis;
  pose *origin;
  int *child_of_joint;   /* joint id -> child link id */
  double *angles;        /* joint_angles_, include/tinyfk.hpp:67 */
  pose base;             /* base_pose_,    include/tinyfk.hpp:68 */
  pose *cache;           /* transform_cache_ (include/data_structure.hpp:5-38) */
  char *cached;
  int *stack_id;         /* transform_stack_ (include/data_structure.hpp:40-61) */
  pose *stack_tf;
};

/* ---- urdf_model/pose.h ------------------------------------------------------------------- */
static quat q_mul(quat a, quat b) { /* Rotation::operator*(Rotation), pose.h:253-263 */
  quat c;
  c.x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
  c.y = a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x;
  c.z = a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w;
  c.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
  return c;
}

static quat q_get_inverse(quat q) { /* Rotation::GetInverse, pose.h:284-299 */
  quat r = {0.0, 0.0, 0.0, 1.0};
  double norm = q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z;
  if (norm > 0.0) {
    r.w = q.w / norm;
    r.x = -q.x / norm;
    r.y = -q.y / norm;
    r.z = -q.z / norm;
  }
  return r;
}

static vec3 q_rotate(quat q, vec3 v) { /* Rotation::operator*(Vector3), pose.h:265-282 */
  quat tmp = {v.x, v.y, v.z, 0.0};
  vec3 r;
  tmp = q_mul(q, q_mul(tmp, q_get_inverse(q)));
  r.x = tmp.x;
  r.y = tmp.y;
  r.z = tmp.z;
  return r;
}

static void q_normalize(quat *q) { /* Rotation::normalize, pose.h:230-250 */
  double s = sqrt(q->x * q->x + q->y * q->y + q->z * q->z + q->w * q->w);
  if (s == 0.0) {
    q->x = q->y = q->z = 0.0;
    q->w = 1.0;
  } else {
    q->x /= s;
    q->y /= s;
    q->z /= s;
    q->w /= s;
  }
}

static quat q_from_rpy(double roll, double pitch, double yaw) { /* Rotation::setFromRPY, pose.h:202-216 */
  double phi = roll / 2.0, the = pitch / 2.0, psi = yaw / 2.0;
  quat q;
  q.x = sin(phi) * cos(the) * cos(psi) - cos(phi) * sin(the) * sin(psi);
  q.y = cos(phi) * sin(the) * cos(psi) + sin(phi) * cos(the) * sin(psi);
  q.z = cos(phi) * cos(the) * sin(psi) - sin(phi) * sin(the) * cos(psi);
  q.w = cos(phi) * cos(the) * cos(psi) + sin(phi) * sin(the) * sin(psi);
  q_normalize(&q);
  return q;
}

static vec3 q_to_rpy(quat q) { /* Rotation::getRPY, pose.h:161-192 */
  double sqx = q.x * q.x, sqy = q.y * q.y, sqz = q.z * q.z, sqw = q.w * q.w;
  double sarg = -2 * (q.x * q.z - q.w * q.y);
  const double pi_2 = 1.57079632679489661923;
  vec3 r;
  if (sarg <= -0.99999) {
    r.y = -pi_2;
    r.x = 0;
    r.z = 2 * atan2(q.x, -q.y);
  } else if (sarg >= 0.99999) {
    r.y = pi_2;
    r.x = 0;
    r.z = 2 * atan2(-q.x, q.y);
  } else {
    r.y = asin(sarg);
    r.x = atan2(2 * (q.y * q.z + q.w * q.x), sqw - sqx - sqy + sqz);
    r.z = atan2(2 * (q.x * q.y + q.w * q.z), sqw + sqx - sqy - sqz);
  }
  return r;
}

static pose pose_identity(void) {
  pose p = {{0.0, 0.0, 0.0}, {0.0, 0.0, 0.0, 1.0}};
  return p;
}

static pose pose_transform(pose a, pose b) { /* urdf_parser/src/pose.cpp:139-146 */
  pose o;
  vec3 r = q_rotate(a.q, b.p);
  o.q = q_mul(a.q, b.q);
  o.p.x = r.x + a.p.x;
  o.p.y = r.y + a.p.y;
  o.p.z = r.z + a.p.z;
  return o;
}

static pose pose_inverse(pose a) { /* Pose::inverse, pose.h:329-339 */
  pose o;
  double norm = a.q.w * a.q.w + a.q.x * a.q.x + a.q.y * a.q.y + a.q.z * a.q.z;
  vec3 n;
  o.q.w = a.q.w / norm;
  o.q.x = -a.q.x / norm;
  o.q.y = -a.q.y / norm;
  o.q.z = -a.q.z / norm;
  n.x = a.p.x * -1;
  n.y = a.p.y * -1;
  n.z = a.p.z * -1;
  o.p = q_rotate(o.q, n);
  return o;
}

static pose joint_transform(int type, vec3 axis, double angle) { /* Joint::transform, joint.h:227-253 */
  pose tf = pose_identity();
  if (type == J_REVOLUTE || type == J_CONTINUOUS) {
    tf.q.x = axis.x * sin(angle * 0.5);
    tf.q.y = axis.y * sin(angle * 0.5);
    tf.q.z = axis.z * sin(angle * 0.5);
    tf.q.w = cos(angle * 0.5);
  } else if (type == J_PRISMATIC) {
    tf.p.x = axis.x * angle;
    tf.p.y = axis.y * angle;
    tf.p.z = axis.z * angle;
  }
  return tf;
}

/* ---- model ----------------------------------------------------------------------------------- */
fko_model *fko_create(int n_links, int n_joints, const int *parent, const int *jtype, const int *jid,
                      const double *axis3, const double *origin7, int root) {
  fko_model *m = (fko_model *)calloc(1, sizeof *m);
  int i;
  m->n_links = n_links;
  m->n_joints = n_joints;
  m->root = root;
  m->parent = (int *)malloc(sizeof(int) * n_links);
  m->jtype = (int *)malloc(sizeof(int) * n_links);
  m->jid = (int *)malloc(sizeof(int) * n_links);
  m->axis = (vec3 *)malloc(sizeof(vec3) * n_links);
  m->origin = (pose *)malloc(sizeof(pose) * n_links);
  m->child_of_joint = (int *)malloc(sizeof(int) * (n_joints > 0 ? n_joints : 1));
  m->angles = (double *)calloc(n_joints > 0 ? n_joints : 1, sizeof(double));
  m->cache = (pose *)malloc(sizeof(pose) * n_links);
  m->cached = (char *)calloc(n_links, 1);
  m->stack_id = (int *)malloc(sizeof(int) * n_links);
  m->stack_tf = (pose *)malloc(sizeof(pose) * n_links);
  for (i = 0; i < n_joints; ++i) m->child_of_joint[i] = -1;
  for (i = 0; i < n_links; ++i) {
    m->parent[i] = parent[i];
    m->jtype[i] = jtype[i];
    m->jid[i] = jid[i];
    m->axis[i].x = axis3[3 * i], m->axis[i].y = axis3[3 * i + 1], m->axis[i].z = axis3[3 * i + 2];
    m->origin[i].p.x = origin7[7 * i], m->origin[i].p.y = origin7[7 * i + 1], m->origin[i].p.z = origin7[7 * i + 2];
    m->origin[i].q.x = origin7[7 * i + 3], m->origin[i].q.y = origin7[7 * i + 4];
    m->origin[i].q.z = origin7[7 * i + 5], m->origin[i].q.w = origin7[7 * i + 6];
    if (jid[i] >= 0 && jid[i] < n_joints) m->child_of_joint[jid[i]] = i;
  }
  m->base = pose_identity(); /* src/tinyfk.cpp:113 */
  return m;
}

void fko_destroy(fko_model *m) {
  if (!m) return;
  free(m->parent), free(m->jtype), free(m->jid), free(m->axis), free(m->origin), free(m->child_of_joint);
  free(m->angles), free(m->cache), free(m->cached), free(m->stack_id), free(m->stack_tf);
  free(m);
}

static void clear_cache(fko_model *m) { memset(m->cached, 0, (size_t)m->n_links); }

void fko_set_joint_angles(fko_model *m, const int *ids, int n, const double *angles) { /* src/tinyfk.cpp:116-128 */
  int i;
  for (i = 0; i < n; ++i) m->angles[ids[i]] = angles[i];
  clear_cache(m);
}

void fko_set_base_pose(fko_model *m, const double *b) { /* src/wrapper.cpp:25-35 */
  m->base.p.x = b[0];
  m->base.p.y = b[1];
  m->base.p.z = b[2];
  m->base.q = q_from_rpy(b[3], b[4], b[5]);
  clear_cache(m);
}

/* KinematicModel::get_link_pose / get_link_pose_inner, src/kinematics.cpp:37-103 */
static pose get_link_pose(fko_model *m, int link_id) {
  int h = link_id, top = 0;
  pose tf_root_to_b = m->base, cur;
  if (m->cached[link_id]) return m->cache[link_id];
  for (;;) {
    pose local;
    if (m->parent[h] < 0) break; /* hit the root link */
    if (m->cached[h]) {
      tf_root_to_b = m->cache[h];
      break;
    }
    if (m->jtype[h] == J_FIXED) {
      local = m->origin[h];
    } else {
      double angle = m->angles[m->jid[h]];
      local = pose_transform(m->origin[h], joint_transform(m->jtype[h], m->axis[h], angle));
    }
    m->stack_id[top] = h;
    m->stack_tf[top] = local;
    ++top;
    h = m->parent[h];
  }
  cur = tf_root_to_b;
  while (top > 0) {
    --top;
    cur = pose_transform(cur, m->stack_tf[top]);
    m->cache[m->stack_id[top]] = cur;
    m->cached[m->stack_id[top]] = 1;
  }
  return cur;
}

/* RelevancePredicateTable::isRelevant as built by update_rptable (src/tinyfk.cpp:260-283):
 * the joint's child link is an ancestor-or-self of the link */
static int is_relevant(const fko_model *m, int link_id, int joint_id) {
  int target = m->child_of_joint[joint_id], l;
  for (l = link_id; l >= 0; l = m->parent[l])
    if (l == target) return 1;
  return 0;
}

static vec3 rpy_derivative(vec3 rpy, vec3 axis) { /* src/kinematics.cpp:14-23 */
  vec3 d;
  double a2 = -rpy.y, a3 = -rpy.z;
  d.x = cos(a3) / cos(a2) * axis.x - sin(a3) / cos(a2) * axis.y;
  d.y = sin(a3) * axis.x + cos(a3) * axis.y;
  d.z = -cos(a3) * sin(a2) / cos(a2) * axis.x + sin(a3) * sin(a2) / cos(a2) * axis.y + axis.z;
  return d;
}

static quat q_derivative(quat q, vec3 omega) { /* src/kinematics.cpp:25-35 */
  quat d;
  double dxdt = 0.5 * (0 * q.x + omega.z * q.y - omega.y * q.z + omega.x * q.w);
  double dydt = 0.5 * (-omega.z * q.x + 0 * q.y + omega.x * q.z + omega.y * q.w);
  double dzdt = 0.5 * (omega.y * q.x - omega.x * q.y + 0 * q.z + omega.z * q.w);
  double dwdt = 0.5 * (-omega.x * q.x - omega.y * q.y - omega.z * q.z + 0 * q.w);
  d.x = dxdt, d.y = dydt, d.z = dzdt, d.w = -dwdt;
  return d;
}

/* KinematicModel::get_jacobian, src/kinematics.cpp:105-235.  jac: pose_dim x n_dof, column-major, zeroed here */
static void get_jacobian(fko_model *m, int elink, const int *joints, int n_joint, int rot_type, int with_base,
                         double *jac, int pose_dim) {
  const int n_dof = n_joint + (with_base ? 6 : 0);
  pose te = get_link_pose(m, elink);
  vec3 epos = te.p, erpy = {0, 0, 0};
  quat erot = te.q, erot_inv = {0, 0, 0, 1};
  int i, k;
  memset(jac, 0, sizeof(double) * (size_t)pose_dim * n_dof);
  if (rot_type == 1) erpy = q_to_rpy(erot);
  if (rot_type == 2) { /* Rotation::inverse, pose.h:301-304 */
    double norm = erot.w * erot.w + erot.x * erot.x + erot.y * erot.y + erot.z * erot.z;
    erot_inv.x = -erot.x / norm, erot_inv.y = -erot.y / norm, erot_inv.z = -erot.z / norm, erot_inv.w = erot.w / norm;
  }
  for (i = 0; i < n_joint; ++i) {
    int jid = joints[i], clink, type;
    pose tc;
    vec3 waxis, dpos;
    double *col = jac + (size_t)i * pose_dim;
    if (!is_relevant(m, elink, jid)) continue;
    clink = m->child_of_joint[jid];
    type = m->jtype[clink];
    tc = get_link_pose(m, clink);
    waxis = q_rotate(tc.q, m->axis[clink]);
    if (type == J_PRISMATIC) {
      dpos = waxis;
    } else {
      vec3 v = {epos.x - tc.p.x, epos.y - tc.p.y, epos.z - tc.p.z};
      dpos.x = waxis.y * v.z - waxis.z * v.y; /* cross_product, pose.cpp:90-94 */
      dpos.y = waxis.z * v.x - waxis.x * v.z;
      dpos.z = waxis.x * v.y - waxis.y * v.x;
    }
    col[0] = dpos.x, col[1] = dpos.y, col[2] = dpos.z;
    if (type != J_PRISMATIC) {
      if (rot_type == 1) {
        vec3 d = rpy_derivative(erpy, waxis);
        col[3] = d.x, col[4] = d.y, col[5] = d.z;
      }
      if (rot_type == 2) {
        quat d = q_derivative(erot_inv, waxis);
        col[3] = d.x, col[4] = d.y, col[5] = d.z, col[6] = d.w;
      }
    }
  }
  if (with_base) { /* src/kinematics.cpp:182-233 */
    const double eps = 1e-7;
    pose tb = get_link_pose(m, m->root);
    pose tb_inv = pose_inverse(tb);
    vec3 rpy_b = q_to_rpy(tb.q);
    pose t_be = pose_transform(tb_inv, te);
    jac[(size_t)(n_joint + 0) * pose_dim + 0] = 1.0;
    jac[(size_t)(n_joint + 1) * pose_dim + 1] = 1.0;
    jac[(size_t)(n_joint + 2) * pose_dim + 2] = 1.0;
    for (k = 0; k < 3; ++k) {
      double *col = jac + (size_t)(n_joint + 3 + k) * pose_dim;
      vec3 t = rpy_b;
      pose tbk = tb, out;
      if (k == 0) t.x += eps;
      if (k == 1) t.y += eps;
      if (k == 2) t.z += eps;
      tbk.q = q_from_rpy(t.x, t.y, t.z);
      out = pose_transform(tbk, t_be);
      col[0] = (out.p.x - te.p.x) / eps;
      col[1] = (out.p.y - te.p.y) / eps;
      col[2] = (out.p.z - te.p.z) / eps;
      if (rot_type == 1) {
        vec3 r = q_to_rpy(out.q);
        col[3] = (r.x - erpy.x) / eps, col[4] = (r.y - erpy.y) / eps, col[5] = (r.z - erpy.z) / eps;
      }
      if (rot_type == 2) {
        col[3] = (out.q.x - erot.x) / eps, col[4] = (out.q.y - erot.y) / eps;
        col[5] = (out.q.z - erot.z) / eps, col[6] = (out.q.w - erot.w) / eps;
      }
    }
  }
}

/* batch loop + layout: src/wrapper.cpp:60-124; PLANER glue: python/tinyfk/__init__.py:98-103,154-158 */
int fko_solve_fk(fko_model *m, const double *qs, long n_wp, const int *elinks, int n_link, const int *joints,
                 int n_joint, int rot_type, int base_type, int with_jacobian, double *P, double *J) {
  const int pose_dim = 3 + (rot_type == 1) * 3 + (rot_type == 2) * 4;
  const int with_base = base_type != 0;
  const int n_in = n_joint + (base_type == 1 ? 3 : (base_type == 2 ? 6 : 0));
  const int n_dof = n_joint + (with_base ? 6 : 0);
  const size_t ldP = (size_t)n_wp * n_link, ldJ = ldP * pose_dim;
  double *jac = (double *)malloc(sizeof(double) * (size_t)pose_dim * (n_dof > 0 ? n_dof : 1));
  long i;
  int j, c, r;
  if (J) memset(J, 0, sizeof(double) * ldJ * n_in);
  for (i = 0; i < n_wp; ++i) {
    const double *row = qs + (size_t)i * n_in;
    for (j = 0; j < n_joint; ++j) m->angles[joints[j]] = row[j]; /* _set_joint_angles, no cache clear */
    if (with_base) {
      double b[6];
      if (base_type == 1) {
        b[0] = row[n_joint], b[1] = row[n_joint + 1], b[2] = 0.0, b[3] = 0.0, b[4] = 0.0, b[5] = row[n_joint + 2];
      } else {
        for (j = 0; j < 6; ++j) b[j] = row[n_joint + j];
      }
      fko_set_base_pose(m, b);
    }
    clear_cache(m); /* use_cache = false */
    for (j = 0; j < n_link; ++j) {
      const size_t head = (size_t)i * n_link + j;
      pose t = get_link_pose(m, elinks[j]);
      P[0 * ldP + head] = t.p.x, P[1 * ldP + head] = t.p.y, P[2 * ldP + head] = t.p.z;
      if (rot_type == 1) {
        vec3 rpy = q_to_rpy(t.q);
        P[3 * ldP + head] = rpy.x, P[4 * ldP + head] = rpy.y, P[5 * ldP + head] = rpy.z;
      } else if (rot_type == 2) {
        P[3 * ldP + head] = t.q.x, P[4 * ldP + head] = t.q.y, P[5 * ldP + head] = t.q.z, P[6 * ldP + head] = t.q.w;
      }
    }
    if (with_jacobian && J) {
      for (j = 0; j < n_link; ++j) {
        const size_t head = ((size_t)i * n_link + j) * pose_dim;
        get_jacobian(m, elinks[j], joints, n_joint, rot_type, with_base, jac, pose_dim);
        for (c = 0; c < n_in; ++c) {
          int src = c;
          if (base_type == 1 && c == n_joint + 2) src = n_joint + 5;
          for (r = 0; r < pose_dim; ++r) J[(size_t)c * ldJ + head + r] = jac[(size_t)src * pose_dim + r];
        }
      }
    }
  }
  free(jac);
  return 0;
}
