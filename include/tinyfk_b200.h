/*
 * tinyfk_b200.h -- C ABI of the B200-native batched forward-kinematics engine.
 *
 * This is the drop-in boundary for tinyfk's FK path: plain pointers and sizes, no C++ types, no
 * exceptions, no torch types.  Each entry names the reference interface it replaces (paths relative
 * to the tinyfk source tree, HiroIshida/tinyfk @ f18ea42).  The pybind11 module _tinyfk
 * (tinyfk_b200/csrc/wrapper.cpp) re-exports it to Python as numpy-in / numpy-out under the reference's
 * own class and method names (src/wrapper.cpp:246-283).
 *
 * Conventions
 *   - Every function returns 0 on success or a TFK_ERR_* code; tfk_last_error() gives the message
 *     for the calling thread.  TFK_ERR_INVALID mirrors std::invalid_argument (-> Python ValueError),
 *     TFK_ERR_RUNTIME mirrors std::runtime_error (-> RuntimeError), TFK_ERR_CUDA is a CUDA failure
 *     (including "no CUDA device": there is NO CPU fallback).
 *   - ids are the reference's ids: links in lexicographic name order, then "<link>_com" links, then
 *     user links; joints = movable joints in lexicographic order (src/tinyfk.cpp:23-64, 89-111).
 *   - rot_type: 0 IGNORE, 1 RPY, 2 XYZW (include/tinyfk.hpp:53).  base_type: 0 FIXED, 1 PLANER,
 *     2 FLOATING (python/tinyfk/__init__.py:29-32).  pose_dim = 3 / 6 / 7.
 *   - Input rows: [q(joint_ids[0]) ... q(joint_ids[n-1]), base...] with base = (x, y, yaw) for PLANER,
 *     (x, y, z, roll, pitch, yaw) for FLOATING (python/tinyfk/__init__.py:98-103, 133-142).
 *   - Outputs are COLUMN-MAJOR exactly as the reference hands them to numpy (src/wrapper.cpp:80-82,123):
 *       P[d * (N*n_link) + i*n_link + j]                         shape (N*n_link, pose_dim)
 *       J[c * (N*n_link*pose_dim) + (i*n_link + j)*pose_dim + r] shape (N*n_link*pose_dim, n_dof)
 *     with n_dof = n_joint (+3 PLANER: x, y, yaw columns, python/tinyfk/__init__.py:154-158; +6 FLOATING).
 *   - precision: TFK_F64_STRICT computes in double in the reference's operation order (matches the
 *     reference to 1e-12 and the libm-pinned reference bit for bit); TFK_F32 is the opt-in float mode
 *     (arrays are float, tolerance 1e-5); TFK_F64_FAST is the opt-in double mode with FMA contraction,
 *     reciprocal-multiply quotients and ANALYTIC base-rotation Jacobian columns: poses and analytic Jacobian entries
 *     within 1e-12 of the reference; its base-rotation columns are the exact derivative, which the reference's
 *     eps = 1e-7 forward difference (src/kinematics.cpp:201-232) only approximates to ~1e-7 * |f''|.
 */
#ifndef TINYFK_B200_H
#define TINYFK_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define TFK_API __attribute__((visibility("default")))
#else
#define TFK_API
#endif

typedef struct tfk_model tfk_model; /* opaque: tinyfk::KinematicModel (include/tinyfk.hpp:55-161) */

enum { TFK_OK = 0, TFK_ERR_INVALID = 1, TFK_ERR_RUNTIME = 2, TFK_ERR_CUDA = 3 };
enum { TFK_ROT_IGNORE = 0, TFK_ROT_RPY = 1, TFK_ROT_XYZW = 2 };
enum { TFK_BASE_FIXED = 0, TFK_BASE_PLANER = 1, TFK_BASE_FLOATING = 2 };
enum { TFK_F64_STRICT = 0, TFK_F32 = 1, TFK_F64_FAST = 2 };
enum { TFK_LIBM_PORTABLE = 0, TFK_LIBM_SYSTEM = 1 };

/* message of the last failing call on this thread */
TFK_API const char *tfk_last_error(void);

/* ---- model life cycle ------------------------------------------------------------------------ */
/* KinematicModel::KinematicModel(xml_string), src/tinyfk.cpp:16-114 (RobotModel(xml), src/wrapper.cpp:255).
 * host_libm selects the libm that evaluates host-side constants (TFK_LIBM_PORTABLE: bit-compatible with
 * the kernels; TFK_LIBM_SYSTEM: the platform libm, as stock tinyfk on this host). */
TFK_API int tfk_model_create(const char *xml, size_t xml_len, int host_libm, tfk_model **out);
TFK_API void tfk_model_destroy(tfk_model *m);

/* ---- ids, names, limits ------------------------------------------------------------------------ */
TFK_API int tfk_num_links(const tfk_model *m);   /* links_.size(), include/tinyfk.hpp:61 */
TFK_API int tfk_num_joints(const tfk_model *m);  /* joints_.size(), include/tinyfk.hpp:65 */
TFK_API int tfk_root_link_id(const tfk_model *m); /* root_link_id_, include/tinyfk.hpp:60 */
TFK_API double tfk_total_mass(const tfk_model *m); /* total_mass_, src/wrapper.cpp:256 */
/* get_link_names / get_joint_names, include/tinyfk.hpp:114-130; NULL if id is out of range */
TFK_API const char *tfk_link_name(const tfk_model *m, int link_id);
TFK_API const char *tfk_joint_name(const tfk_model *m, int joint_id);
/* get_link_ids / get_joint_ids, src/tinyfk.cpp:203-215 / 150-162; unknown name -> TFK_ERR_INVALID */
TFK_API int tfk_get_link_ids(const tfk_model *m, const char *const *names, int n, int *ids_out);
TFK_API int tfk_get_joint_ids(const tfk_model *m, const char *const *names, int n, int *ids_out);
/* get_joint_position_limits, src/tinyfk.cpp:164-179 (continuous joints: -inf / +inf) */
TFK_API int tfk_get_joint_limits(const tfk_model *m, const int *joint_ids, int n, double *lower, double *upper);
/* get_joint_velocity_limits / get_joint_effort_limits, src/tinyfk.cpp:181-201 */
TFK_API int tfk_get_joint_velocity_limits(const tfk_model *m, const int *joint_ids, int n, double *out);
TFK_API int tfk_get_joint_effort_limits(const tfk_model *m, const int *joint_ids, int n, double *out);
/* Parsed tree as the reference holds it per link (urdf::Link / urdf::Joint, link.h:195-220, joint.h:174-199):
 * parent link id (-1 root), type code of the parent joint (urdf::Joint enum: 1 REVOLUTE 2 CONTINUOUS
 * 3 PRISMATIC 6 FIXED, 0 for the root), movable-joint id (-1 if fixed), axis[3], origin x y z qx qy qz qw. */
TFK_API int tfk_link_info(const tfk_model *m, int link_id, int *parent, int *jtype, int *jid, double axis[3],
                          double origin[7]);
/* rptable_.isRelevant(link, joint), include/tinyfk.hpp:43; writes 0/1 */
TFK_API int tfk_is_relevant(const tfk_model *m, int link_id, int joint_id, int *out);

/* ---- stored state ---------------------------------------------------------------------------- */
/* set_joint_angles / get_joint_angles, src/tinyfk.cpp:116-148 */
TFK_API int tfk_set_joint_angles(tfk_model *m, const int *joint_ids, int n, const double *angles);
TFK_API int tfk_get_joint_angles(const tfk_model *m, const int *joint_ids, int n, double *angles);
/* set_base_pose(xyzrpy) / get_base_pose(), src/wrapper.cpp:37-55 */
TFK_API int tfk_set_base_pose(tfk_model *m, const double xyzrpy[6]);
TFK_API int tfk_get_base_pose(const tfk_model *m, double xyzrpy[6]);
/* add_new_link(name, parent_id, position, rpy), src/tinyfk.cpp:217-258; duplicate name -> TFK_ERR_INVALID */
TFK_API int tfk_add_new_link(tfk_model *m, const char *name, int parent_id, const double pos[3],
                             const double rpy[3], int *new_id_out);

/* ---- the hot path ---------------------------------------------------------------------------- */
/* Batched solve_fk on HOST buffers: KinematicsModelPyWrapper::solve_forward_kinematics
 * (src/wrapper.cpp:60-124) + the facade glue of KinematicModel.solve_fk (python/tinyfk/__init__.py:105-159).
 * q: n x n_in row-major (double, or float for TFK_F32).  P, J: column-major outputs as described above
 * (J may be NULL when with_jacobian == 0).  devices/n_devices: CUDA ordinals to shard the batch over
 * (contiguous slices, no collective); NULL/0 = the current device.  Copies are chunked and overlapped
 * with the kernel; pass pinned memory (tfk_host_alloc) for full PCIe speed.  Like the reference, the
 * model's stored joint angles / base pose afterwards are those of the LAST configuration. */
TFK_API int tfk_solve_fk_host(tfk_model *m, const void *q, int64_t n, const int *elink_ids, int n_link,
                              const int *joint_ids, int n_joint, int rot_type, int base_type,
                              int with_jacobian, int precision, const int *devices, int n_devices, void *P,
                              void *J);

/* Same computation on DEVICE-resident buffers of `device`, enqueued on `stream` (a cudaStream_t; NULL =
 * the legacy default stream) and not synchronised.  The destination arrays are dimensioned for ld_cfg
 * configurations and this call fills configurations [row0, row0 + n).  Does not touch the model's stored
 * state.  This is the entry the throughput metric is timed on.  After the first call with a signature on a stream the
 * call only enqueues one kernel: it may be captured into a CUDA graph on that stream.  Capture rules: (1) run the
 * same call once on the stream before capturing (the record scratch of a stream only ever grows, and growing it
 * inside a capture fails with TFK_ERR_INVALID); (2) buffers a captured kernel names stay allocated until
 * tfk_model_destroy, so replaying the graph after larger launches is safe; (3) while the stream is capturing the
 * engine never loads or compiles a plan-specialised kernel: the capture records the specialised kernel only if it has
 * already run on this device, otherwise the generic one; (4) keep the model's stored state that feeds the signature's
 * plan (angles of non-selected joints, the FIXED base pose) unchanged between capture and replay. */
TFK_API int tfk_solve_fk_device(tfk_model *m, int device, void *stream, const void *d_q, int64_t n,
                                const int *elink_ids, int n_link, const int *joint_ids, int n_joint,
                                int rot_type, int base_type, int with_jacobian, int precision, void *d_P,
                                void *d_J, int64_t ld_cfg, int64_t row0);

/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA-graph capture, tight serving loops, harnesses that forbid allocation).  tfk_solve_fk_device in
 * TFK_F64_STRICT (same arguments minus `precision`, same outputs, same bits) computed by the several-link kernel
 * (fk_twophase_kernel, any n_link >= 1) whose scratch -- the per-configuration records and the plan's per-link tables --
 * is a file-scope static device array of TFK_NOALLOC_STATIC_BYTES (32 MiB) instead of cudaMalloc'ed memory
 * (tinyfk_b200/csrc/fk_noalloc.cu).  The call allocates NO device memory: no cudaMalloc, cudaFree or allocating library
 * routine; it enqueues one small host->device copy (the tables) and one kernel on `stream`.  (The plan of a signature is
 * built in host memory on its first call and cached.)
 *   - Limits: n <= TFK_NOALLOC_MAX_CONFIGS (65 536) configurations per call, per-link tables <= 64 KiB, one tile's records
 *     <= the array; beyond them the call fails with TFK_ERR_INVALID and a message that names the limit.  Requests with
 *     large records (many links) run with fewer resident blocks, never with more memory.
 *   - Every GPU on which the entry has been used holds the array until the process exits.  (The first launch of a kernel
 *     instance on a device lets the driver load its code, as for any kernel; callers that must see no change in free
 *     memory at all make one call per RotationType at set-up.)
 *   - Every call shares the array, which keeps the previous call's contents (nothing in it has to start at zero: every
 *     record is written before it is read, the tables are re-written per call).  Calls that overlap on the GPU -- two
 *     streams, two host threads, two models on one device -- RACE on it: serialise them (one stream is enough). */
enum { TFK_NOALLOC_MAX_CONFIGS = 1 << 16, TFK_NOALLOC_STATIC_BYTES = 32 << 20 };
TFK_API int tfk_solve_fk_device_noalloc(tfk_model *m, int device, void *stream, const void *d_q, int64_t n,
                                        const int *elink_ids, int n_link, const int *joint_ids, int n_joint,
                                        int rot_type, int base_type, int with_jacobian, void *d_P, void *d_J,
                                        int64_t ld_cfg, int64_t row0);

/* ---- inter-link squared distances ------------------------------- */
/* KinematicsModelPyWrapper::compute_inter_link_squared_dists (src/wrapper.cpp:189-237) + the facade glue of
 * KinematicModel.compute_inter_link_sqdists (python/tinyfk/__init__.py:233-266), batched on the GPU: for every
 * configuration i and pair j = (link_ids1[j], link_ids2[j])
 *     sqdists[i*n_check + j] = |p1 - p2|^2
 *     grads[c * (N*n_check) + i*n_check + j] = 2 * (J1[:,c] - J2[:,c]) . (p1 - p2)       (column-major, n_dof columns)
 * with J1, J2 the position Jacobians of the two links (never materialised).  q, base_type, n_dof, precision
 * and the stored-state rule are those of tfk_solve_fk_host; grads may be NULL when with_grads == 0.  Runs on the
 * current device. */
TFK_API int tfk_inter_link_sqdists_host(tfk_model *m, const void *q, int64_t n, const int *link_ids1,
                                        const int *link_ids2, int n_check, const int *joint_ids, int n_joint,
                                        int base_type, int with_grads, int precision, void *sqdists, void *grads);
/* same on device-resident buffers (see tfk_solve_fk_device for stream / ld_cfg / row0) */
TFK_API int tfk_inter_link_sqdists_device(tfk_model *m, int device, void *stream, const void *d_q, int64_t n,
                                          const int *link_ids1, const int *link_ids2, int n_check,
                                          const int *joint_ids, int n_joint, int base_type, int with_grads,
                                          int precision, void *d_sqdists, void *d_grads, int64_t ld_cfg, int64_t row0);

/* ---- centre of mass ---------------------------------------------- */
/* KinematicsModelPyWrapper::solve_com_forward_kinematics (src/wrapper.cpp:126-179) over get_com / get_com_jacobian
 * (src/kinematics.cpp:237-270) + the facade glue of KinematicModel.solve_com_fk (python/tinyfk/__init__.py:160-195),
 * batched on the GPU.  For every configuration i:
 *     com[i*3 + k]                           mass-weighted mean of the "<link>_com" link positions, in link-id order;
 *                                            with action links: (com + sum_j (forces[j]/total_mass) * p_j) / (1 + sum_j ...)
 *     J[c * (3*N) + 3*i + k]                 its Jacobian, column-major (3N) x n_dof, n_dof = n_joint (+6 with a moving base)
 * q rows are [joints | base] as for tfk_solve_fk_host; a PLANER row (x, y, yaw) is padded to (x, y, 0, 0, 0, yaw) and
 * J KEEPS ALL SIX base columns (the reference's facade does not pick columns on this entry).  J may be NULL when
 * with_jacobian == 0.  Every configuration is evaluated from its own row (the reference skips clear_cache() on this
 * entry with a FIXED base, src/wrapper.cpp:137-141, and then repeats the first row's result: not reproduced).
 * Runs on the current device; the model's stored state follows the last configuration. */
TFK_API int tfk_solve_com_fk_host(tfk_model *m, const void *q, int64_t n, const int *joint_ids, int n_joint,
                                  const int *action_link_ids, const double *forces, int n_action, int base_type,
                                  int with_jacobian, int precision, void *com, void *J);
/* same on device-resident buffers (see tfk_solve_fk_device for stream / ld_cfg / row0); base_type FIXED or FLOATING;
 *     d_com[k * ld_cfg + row]   (3, ld_cfg) component-major,   d_J[c * (3*ld_cfg) + 3*row + k] */
TFK_API int tfk_solve_com_fk_device(tfk_model *m, int device, void *stream, const void *d_q, int64_t n,
                                    const int *joint_ids, int n_joint, const int *action_link_ids,
                                    const double *forces, int n_action, int base_type, int with_jacobian,
                                    int precision, void *d_com, void *d_J, int64_t ld_cfg, int64_t row0);

/* ---- total inertia matrix ----------------------------------------- */
/* KinematicsModelPyWrapper::compute_total_inertia_matrix (src/wrapper.cpp:181-187) over get_total_inertia_matrix
 * (src/kinematics.cpp:272-317), batched: q is [n][n_joint] (the stored base pose applies), out is [n][3][3]:
 *     sum over the links with an <inertial> of  R I R^T + m (|v|^2 Id - v v^T),   v = p_link - centre of mass.
 * The reference evaluates the 3x3 products with Eigen, whose summation order is not pinned in this repository
 * (no Eigen here): the contract for this entry is 1e-12 relative, not bit equality.  Every row is evaluated from its
 * own configuration (the reference skips clear_cache() on this entry).  Current device; stored angles follow the
 * last configuration. */
TFK_API int tfk_total_inertia_matrix_host(tfk_model *m, const void *q, int64_t n, const int *joint_ids, int n_joint,
                                          int precision, void *out);
/* device-resident buffers: d_out[(3*i + j) * ld_cfg + row] (9, ld_cfg) component-major */
TFK_API int tfk_total_inertia_matrix_device(tfk_model *m, int device, void *stream, const void *d_q, int64_t n,
                                            const int *joint_ids, int n_joint, int precision, void *d_out,
                                            int64_t ld_cfg, int64_t row0);

/* ---- plan specialisation ------------------------------------------------------------------------ */
/* No reference equivalent.  A query signature (links, joints, rot_type, base_type, with_jacobian) that is evaluated
 * over and over is a constant of the application: the single-link kernel can be compiled once more against THAT plan
 * (NVRTC at run time, same source, same flags, same bits; cubins cached under TFK_CACHE_DIR, default
 * ~/.cache/tinyfk_b200) and then runs ~1.25x faster.  Policy (tfk_set_jit_mode; env TFK_JIT=off|auto|on):
 *   TFK_JIT_OFF   generic kernels only
 *   TFK_JIT_AUTO  (default) a plan that has served TFK_JIT_THRESHOLD (default 2^22) configurations is compiled in a
 *                 background thread; calls keep using the generic kernel until it is ready
 *   TFK_JIT_ON    compile synchronously on a plan's first use
 * tfk_specialize does it explicitly: flags TFK_SPEC_WAIT blocks until the kernel is compiled (and, unless
 * TFK_SPEC_COMPILE_ONLY, loaded on the current device); TFK_SPEC_COMPILE_ONLY needs no GPU.  info (may be NULL)
 * receives {state, threads per block, blocks per SM, shared memory bytes, cubin bytes, came from the disk cache};
 * state: 1 compiling, 2 compiled, 3 ready.  TFK_ERR_INVALID: the plan is not eligible (several links, very long
 * chain); TFK_ERR_RUNTIME: NVRTC not found or the compile failed (the generic kernel keeps serving the plan). */
enum { TFK_JIT_OFF = 0, TFK_JIT_AUTO = 1, TFK_JIT_ON = 2 };
enum { TFK_SPEC_WAIT = 1, TFK_SPEC_COMPILE_ONLY = 2 };
TFK_API int tfk_specialize(tfk_model *m, const int *elink_ids, int n_link, const int *joint_ids, int n_joint,
                           int rot_type, int base_type, int with_jacobian, int precision, int flags, int64_t *info);
TFK_API int tfk_set_jit_mode(tfk_model *m, int mode);
TFK_API int tfk_get_jit_mode(const tfk_model *m);

/* ---- helpers ---------------------------------------------------------------------------------- */
TFK_API int tfk_device_count(int *count);
/* pinned host memory for the host path (pooled; tfk_host_free returns the block to the pool) */
TFK_API int tfk_host_alloc(size_t bytes, void **ptr);
TFK_API int tfk_host_free(void *ptr);
/* kernels launched by this library since load (bench.py's gpu_launches) */
TFK_API int64_t tfk_kernel_launch_count(void);
/* name of the kernel last launched for (m, precision) and its launch shape, for logs */
TFK_API int tfk_last_launch_info(const tfk_model *m, int *grid, int *block, int *smem_bytes, int *blocks_per_sm,
                                 int *rec_reals);
/* which kernel the last launch used: 0 fused single-link (records in shared memory), 1 fused single-link
   (records in global scratch), 2 two-phase multi-link, 3 = 0 built for five resident blocks per SM -- see tinyfk_b200/csrc/fk_kernels.cuh */
TFK_API int tfk_last_launch_variant(const tfk_model *m);
/* call signatures the engine currently holds a plan for (bounded at 64, least recently used evicted; a signature whose
 * plan constants change with the stored model state is replaced in place).  Diagnostics / tests. */
TFK_API int tfk_plan_cache_size(const tfk_model *m);
/* 1 when tfk_model_create took the link / joint tables of this URDF text from the disk cache (TFK_CACHE_DIR, default
 * ~/.cache/tinyfk_b200; TFK_TABLE_CACHE=0 disables it) instead of reading the XML, else 0.  The cache is SURVEY.md section
 * 8(f) rank 4 ("URDF -> flat-table cache on disk"); it replaces the per-process work of src/tinyfk.cpp:16-114 +
 * urdf_parser/src/model.cpp:91-270.  Diagnostics / tests. */
TFK_API int tfk_model_tables_from_cache(const tfk_model *m);
TFK_API const char *tfk_version(void);

#ifdef __cplusplus
}
#endif
#endif /* TINYFK_B200_H */
