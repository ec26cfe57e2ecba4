// engine.hpp -- host runtime around the batch kernel: plan cache, per-device workspaces, the chunked
// host<->device pipeline and the contiguous multi-GPU split.  Replaces the sequential, stateful batch
// loop of the reference (src/wrapper.cpp:84-122).  No CPU compute path exists here: without a CUDA
// device every solve fails with TFK_ERR_CUDA.
#ifndef TFK_ENGINE_HPP
#define TFK_ENGINE_HPP

#include <cuda_runtime.h>

#include <cstdint>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <stdexcept>
#include <thread>
#include <string>
#include <tuple>
#include <utility>
#include <vector>

#include "fk_launch.h"
#include "model.hpp"
#include "plan.hpp"

namespace tfk {

struct CudaError : std::runtime_error {
  explicit CudaError(const std::string &w) : std::runtime_error(w) {}
};

struct SolveSpec {
  const int *elink_ids;
  int n_link;
  const int *joint_ids;
  int n_joint;
  int rot_type, base_type, with_jacobian, precision;
};

struct PairSpec { // compute_inter_link_squared_dists (reference src/wrapper.cpp:189-237)
  const int *link_ids1, *link_ids2;
  int n_check;
  const int *joint_ids;
  int n_joint;
  int base_type, with_grads, precision;
};

struct ComSpec { // solve_com_forward_kinematics (reference src/wrapper.cpp:126-179)
  const int *joint_ids;
  int n_joint;
  const int *action_link_ids;
  const double *forces;
  int n_action;
  int base_type, with_jacobian, precision; // base_type: FIXED or FLOATING on the device entry (PLANER: host entry pads)
};

enum JitMode {
  JIT_OFF = 0,  // generic kernels only
  JIT_AUTO = 1, // a single-link plan that has served TFK_JIT_THRESHOLD configurations is specialised in the background
  JIT_ON = 2,   // specialise synchronously on first use (first call pays the compile, a few seconds)
};

struct SpecInfo { // what Engine::specialize reports
  int state;      // SPEC_*
  int tile, blocks_per_sm, smem_bytes;
  int64_t cubin_bytes;
  int from_cache;
};
enum { SPEC_NONE = 0, SPEC_COMPILING = 1, SPEC_COMPILED = 2, SPEC_READY = 3, SPEC_FAILED = 4 };

struct LaunchInfo {
  int grid, block, smem_bytes, blocks_per_sm, rec_reals, variant;
};

class Engine {
public:
  Engine(const std::string &xml, HostLibm libm);
  ~Engine();

  KinematicModel model;
  std::mutex mu; // the C ABI serialises calls per model handle

  void solve_device(int device, cudaStream_t stream, const void *d_q, int64_t n, const SolveSpec &spec, void *d_P,
                    void *d_J, int64_t ld_cfg, int64_t row0);
  void solve_host(const void *q, int64_t n, const SolveSpec &spec, const int *devices, int n_devices, void *P, void *J);
  // solve_device without any device-memory allocation (fk_noalloc.cu): the several-link kernel with its record scratch and
  // the plan's per-output tables in a static device array; strict FP64 (spec.precision is ignored), n <= 65 536
  void solve_device_noalloc(int device, cudaStream_t stream, const void *d_q, int64_t n, const SolveSpec &spec, void *d_P,
                            void *d_J, int64_t ld_cfg, int64_t row0);
  void sqdists_device(int device, cudaStream_t stream, const void *d_q, int64_t n, const PairSpec &ps, void *d_sq,
                      void *d_G, int64_t ld_cfg, int64_t row0);
  void sqdists_host(const void *q, int64_t n, const PairSpec &ps, void *sq, void *G);
  void com_device(int device, cudaStream_t stream, const void *d_q, int64_t n, const ComSpec &cs, void *d_com, void *d_J,
                  int64_t ld_cfg, int64_t row0);
  void com_host(const void *q, int64_t n, const ComSpec &cs, void *com, void *J);
  // total inertia matrix about the centre of mass (reference src/wrapper.cpp:181-187, src/kinematics.cpp:272-317);
  // q rows are the selected joints (stored base pose); device out (9, ld_cfg) component-major, host out (n, 3, 3)
  void inertia_device(int device, cudaStream_t stream, const void *d_q, int64_t n, const int *joint_ids, int n_joint,
                      int precision, void *d_out, int64_t ld_cfg, int64_t row0);
  // Plan specialisation (jit.hpp).  wait: block until the kernel is compiled; compile_only: never touch a device.
  // Throws std::invalid_argument when the plan is not eligible (several links, records beyond shared memory, very
  // long chains) and std::runtime_error when NVRTC is missing or the compile fails.
  SpecInfo specialize(const SolveSpec &spec, bool wait, bool compile_only);
  int jit_mode;
  int64_t jit_threshold;
  void inertia_host(const void *q, int64_t n, const int *joint_ids, int n_joint, int precision, void *out);
  LaunchInfo last_launch;
  size_t plan_cache_size() const; // signatures held (diagnostics, tests)

private:
  struct SpecKernel { // one per (plan, precision)
    std::mutex mu;     // guards state / cubin / error against the compile thread
    int state = SPEC_NONE;
    std::string error;
    std::vector<char> cubin;
    bool from_cache = false;
    std::thread worker;
    int tile = 0, bps = 0, smem = 0;
    int kind = 0;      // jit.hpp SpecOptions::kind: 0 fused, 1 two-phase, 2 two-phase with the joint records in shared memory
    cudaLibrary_t lib = nullptr;
    cudaKernel_t kernel = nullptr;
    std::map<int, int> blocks_per_sm; // device -> occupancy (also marks the shared-memory attribute as set there)
    ~SpecKernel();
  };
  struct PlanEntry {
    PlanKey key;                     // the call signature; key.model_version = the model state the blob was built from
    uint64_t last_used = 0;          // LRU tick
    std::set<int> launched_on;       // devices that may still run a kernel of this entry (synchronised before it is freed)
    std::vector<unsigned char> blob;
    std::unique_ptr<PlanParams> params; // the kernel-parameter form of the blob (fk_plan.h)
    std::map<int, unsigned char *> on_device; // device ordinal -> uploaded copy
    SpecKernel spec[PREC_COUNT];              // by precision mode
    int64_t configs_served = 0;
  };
  bool spec_eligible(const PlanEntry &e, int precision, int *tile, int *bps, int *smem, int *kind, std::string *why) const;
  void spec_start(PlanEntry &e, int precision, bool wait);
  bool spec_launch(int device, cudaStream_t stream, PlanEntry &e, int precision, const KernelArgs &a);
  struct Workspace { // one per (device, stream): concurrent launches never share records
    void *scratch;
    size_t scratch_bytes;
    std::vector<void *> outgrown; // grow-only: a captured CUDA graph may still name an older buffer (freed with the engine)
  };
  struct HostPipe { // per device: two slots of staging buffers, one stream each
    cudaStream_t stream[2];
    void *dq[2], *dP[2], *dJ[2];
    size_t cap_q[2], cap_P[2], cap_J[2];
    bool ready;
  };

  struct LaunchExtra { // what the pair / centre-of-mass kernels need beyond the plan
    int variant;
    const int *d_pairs;
    int n_pairs;
    const void *d_weights;
    int n_com, n_act;
    double mass_total, vfs, com_diag;
  };
  PlanEntry &plan_for(const SolveSpec &spec);
  void retire(std::unique_ptr<PlanEntry> &e); // waits for the devices it ran on, frees its tables and its specialised kernels
  uint64_t plan_tick_ = 0;
  std::mutex launch_mu_; // Engine::solve_host drives several devices from one thread each; launch() and its caches are shared
  size_t max_pitch(int device);
  std::map<int, size_t> max_pitch_;
  // device -> host copy of `ncols` columns of `width` bytes (column-major source and destination): one strided copy, or one
  // plain copy per column when a pitch exceeds what cudaMemcpy2D accepts (cudaDevAttrMaxPitch, 2 GiB - 1 on B200)
  void copy_columns_d2h(int device, char *dst, size_t dst_pitch, const char *src, size_t src_pitch, size_t width, size_t ncols,
                        cudaStream_t st, const char *what);
  const unsigned char *plan_on_device(PlanEntry &e, int device, cudaStream_t stream);
  Workspace &workspace(int device, cudaStream_t stream, size_t bytes);
  HostPipe &pipe(int device);
  int sm_count(int device);
  void launch(int device, cudaStream_t stream, PlanEntry &e, int precision, const void *d_q, int64_t n, void *d_P,
              void *d_J, int64_t ld_cfg, int64_t row0, const LaunchExtra *ex = nullptr);
  PlanEntry &plan_for_pairs(const PairSpec &ps, std::vector<int> &o12);
  struct PairBuf {
    int *d = nullptr;
    size_t bytes = 0;
    std::vector<int> host;
  };
  std::map<std::pair<int, cudaStream_t>, PairBuf> pair_bufs_;
  struct WeightBuf {
    void *d = nullptr;
    size_t bytes = 0;
    std::vector<double> host;
    int precision = -1;
  };
  std::map<std::pair<int, cudaStream_t>, WeightBuf> weight_bufs_;
  const void *upload_weights(int device, cudaStream_t stream, const std::vector<double> &w, int precision);
  void release_device_memory();

  std::vector<std::unique_ptr<PlanEntry>> plans_;
  std::map<std::pair<int, cudaStream_t>, Workspace> workspaces_;
  std::map<int, HostPipe> pipes_;
  std::map<int, int> sm_counts_;
  std::map<std::tuple<int, int, int, int>, int> occupancy_; // (device, precision, variant, smem bytes) -> blocks per SM
  std::map<std::tuple<int, int, int>, int> occupancy_noalloc_;   // (device, rot_type, smem bytes) -> blocks per SM
};

int64_t kernel_launch_count();

// pooled pinned host memory
void *pinned_alloc(size_t bytes);
void pinned_free(void *p);

} // namespace tfk
#endif
