/*
 * fk_oracle.h -- TEST INFRASTRUCTURE.  Plain-C CPU restatement of tinyfk's batched solve_fk path.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use it; the product never
 * links or loads it.
 *
 * Input is the parsed link table ("urdf_parser output"): per link its parent link id (-1 for the
 * root), the type code of its parent joint (urdf::Joint enum: 1 REVOLUTE, 2 CONTINUOUS, 3 PRISMATIC,
 * 6 FIXED), the movable-joint id (-1 if fixed), the joint axis and the parent->joint origin
 * (x y z  qx qy qz qw).  Pinning: checked against oracle/_ref (the reference's own object code) and
 * the reference's golden vectors in tests/test_oracle.py.
 */
#ifndef FK_ORACLE_H
#define FK_ORACLE_H
#ifdef __cplusplus
extern "C" {
#endif

typedef struct fko_model fko_model;

fko_model *fko_create(int n_links, int n_joints, const int *parent, const int *jtype, const int *jid,
                      const double *axis3, const double *origin7, int root);
void fko_destroy(fko_model *m);
void fko_set_joint_angles(fko_model *m, const int *ids, int n, const double *angles);
void fko_set_base_pose(fko_model *m, const double *xyzrpy);
/* rot_type 0 IGNORE / 1 RPY / 2 XYZW; base_type 0 FIXED / 1 PLANER / 2 FLOATING.
 * qs row-major n_wp x (n_joint + 0|3|6).  P column-major (n_wp*n_link) x pose_dim.
 * J column-major (n_wp*n_link*pose_dim) x (n_joint + 0|3|6), zero-filled here; may be NULL if !with_jacobian. */
int fko_solve_fk(fko_model *m, const double *qs, long n_wp, const int *elinks, int n_link, const int *joints,
                 int n_joint, int rot_type, int base_type, int with_jacobian, double *P, double *J);

#ifdef __cplusplus
}
#endif
#endif
