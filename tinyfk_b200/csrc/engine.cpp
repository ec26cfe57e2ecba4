// engine.cpp -- see engine.hpp.
#include "engine.hpp"
#include "fk_noalloc.h"
#include "jit.hpp"

#include <nvtx3/nvToolsExt.h> // header-only NVTX 3: ranges are no-ops unless a profiler is attached

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstring>
#include <exception>

namespace tfk {

namespace {

std::atomic<int64_t> g_launches(0);

inline void ck(cudaError_t e, const char *what) {
  if (e != cudaSuccess) throw CudaError(std::string(what) + ": " + cudaGetErrorString(e));
}

struct DeviceGuard { // restore the caller's current device
  int prev;
  bool ok;
  explicit DeviceGuard(int dev) : prev(0), ok(false) {
    ck(cudaGetDevice(&prev), "cudaGetDevice (is a CUDA device present? there is no CPU fallback)");
    ok = true;
    if (dev != prev) ck(cudaSetDevice(dev), "cudaSetDevice");
  }
  ~DeviceGuard() {
    if (ok) cudaSetDevice(prev);
  }
};

struct NvtxRange { // NVTX range around a host-side section (H2D / kernel / D2H enqueue of a chunk)
  explicit NvtxRange(const char *name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
};

inline size_t real_size(int precision) { return precision == PREC_F32 ? sizeof(float) : sizeof(double); }

// ---- pinned pool --------------------------------------------------------------------------------
std::mutex g_pin_mu;
std::map<void *, size_t> g_pin_live;            // handed out
std::multimap<size_t, void *> g_pin_free;       // available, by size

} // namespace

static bool stream_is_capturing(cudaStream_t stream) {
  cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(stream, &st) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return st != cudaStreamCaptureStatusNone;
}

int64_t kernel_launch_count() { return g_launches.load(); }

void *pinned_alloc(size_t bytes) {
  if (bytes == 0) bytes = 1;
  std::lock_guard<std::mutex> lk(g_pin_mu);
  auto it = g_pin_free.lower_bound(bytes);
  if (it != g_pin_free.end() && it->first <= bytes + bytes / 4 + 4096) {
    void *p = it->second;
    g_pin_live[p] = it->first;
    g_pin_free.erase(it);
    return p;
  }
  void *p = nullptr;
  ck(cudaHostAlloc(&p, bytes, cudaHostAllocPortable), "cudaHostAlloc");
  g_pin_live[p] = bytes;
  return p;
}

void pinned_free(void *p) {
  if (!p) return;
  std::lock_guard<std::mutex> lk(g_pin_mu);
  auto it = g_pin_live.find(p);
  if (it == g_pin_live.end()) throw std::invalid_argument("tfk_host_free: pointer was not returned by tfk_host_alloc");
  g_pin_free.emplace(it->second, p);
  g_pin_live.erase(it);
  // keep the pool bounded: drop the largest free blocks beyond 8 GiB
  size_t total = 0;
  for (auto &kv : g_pin_free) total += kv.first;
  while (total > (size_t(8) << 30) && !g_pin_free.empty()) {
    auto last = std::prev(g_pin_free.end());
    total -= last->first;
    cudaFreeHost(last->second);
    g_pin_free.erase(last);
  }
}

// ---- engine -------------------------------------------------------------------------------------
Engine::Engine(const std::string &xml, HostLibm libm) : model(xml, libm) {
  std::memset(&last_launch, 0, sizeof last_launch);
  jit_mode = JIT_AUTO;
  jit_threshold = int64_t(1) << 22;
  if (const char *m = std::getenv("TFK_JIT")) {
    const std::string v(m);
    if (v == "off" || v == "0") jit_mode = JIT_OFF;
    if (v == "on" || v == "2") jit_mode = JIT_ON;
  }
  if (const char *t = std::getenv("TFK_JIT_THRESHOLD")) jit_threshold = std::atoll(t);
}

Engine::SpecKernel::~SpecKernel() {
  if (worker.joinable()) worker.join();
  if (lib) cudaLibraryUnload(lib);
}

// configurations per tile of the two-phase kernels: the rule of fk_launch.h, or TFK_TWOPHASE_TILE_CFGS (32 | 64 | 128:
// A/B knob for measurements -- smaller tiles touch less of the block-private record scratch)
// small_tiles: the plan-specialised FP32 build.  Its phase A (the straight-line walk of THIS tree in FP32) is short, so
// idle threads there cost little, while a 32-configuration tile touches a quarter of the block-private record scratch
// (L2 hits instead of HBM round trips) and evens out the last wave: measured on B200 (profiles/README.md round 2) cfg5
// FP32 0.437 -> 0.416 ms, cfg4 FP32 0.432 -> 0.421 ms.  FP64 builds keep 128 or 64 by the rule of fk_launch.h: a long
// FP64 phase A with three quarters of the threads idle costs more than the scratch saves (cfg4 strict at 32: 1.27 ms
// against 0.80 ms, cfg4 FP64-fast 1.32 ms; cfg5 FP64-fast gains 2 % at 64 or 32, within the rule's reach).
static int tile_cfgs_for(long long n, long long resident_blocks, bool small_tiles) {
  if (const char *t = std::getenv("TFK_TWOPHASE_TILE_CFGS")) {
    const int v = std::atoi(t);
    if (v == 32 || v == 64 || v == (int)TFK_TILE) return v;
  }
  if (small_tiles) return 32;
  return twophase_tile_cfgs(n, resident_blocks);
}

// shared memory (reals per thread) of the per-warp staging area of the two-phase kernel whose joint records stay in the
// scratch: fk_launch.h:twophase_stage_reals; TFK_NO_STAGE=1 turns it off (A/B knob for measurements)
static int twophase_stage_for(const PlanHeader &h) {
  if (std::getenv("TFK_NO_STAGE")) return 0;
  return twophase_stage_reals(h.n_out, h.off_base, h.off_frame);
}

// ---- plan specialisation (jit.hpp) -------------------------------------------------------------------------------
bool Engine::spec_eligible(const PlanEntry &e, int precision, int *tile, int *bps, int *smem, int *kind, std::string *why) const {
  const PlanHeader *h = reinterpret_cast<const PlanHeader *>(e.blob.data());
  *kind = 0;
  if (h->n_out != 1) {
    // several links: the two-phase kernel against this plan (same block shape and shared-memory rule as the generic
    // launch: [base | arec | prec] in shared memory when three blocks fit)
    if (h->n_slots > 128 || h->n_cols > 32) {
      if (why) *why = "tree too large to unroll (more than 128 links or 32 joints)";
      return false;
    }
    // measured (B200, profiles/README.md v10): FP32 gains 28 % (cfg4) with everything unrolled; strict FP64 gains 5 % on
    // position-only requests (cfg4, walk kept rolled) and LOSES 12 % with orientation rows (cfg5: the unrolled RPY /
    // finite-difference columns of 21 columns outgrow the instruction cache) -- those stay with the generic kernel
    if (precision == PREC_F64_STRICT && h->rot_type != 0 && !std::getenv("TFK_SPEC_STRICT_ORIENT")) { // (env: A/B knob)
      if (why) *why = "several links with orientation rows in strict FP64: the generic two-phase kernel is faster";
      return false;
    }
    const int rb = (int)real_size(precision);
    const int table_bytes = (h->off_cols - h->off_outs + 15) & ~15;
    const SmemLayout Lh = smem_layout(table_bytes, h->n_in, h->off_frame, rb);
    const bool hot = 3 * (Lh.total + 1024) <= 227 * 1024 && !std::getenv("TFK_NO_SMEM_HOT");
    const SmemLayout L = hot ? Lh : smem_layout(table_bytes, h->n_in, twophase_stage_for(*h), rb);
    if (L.total > 200 * 1024) {
      if (why) *why = "plan too large for shared memory";
      return false;
    }
    *tile = TFK_TILE;
    *bps = hot ? std::min(4, (227 * 1024) / (L.total + 1024)) : 4;
    *smem = L.total;
    *kind = hot ? 2 : 1;
    return true;
  }
  if (h->n_slots > 48 || h->n_cols > 32) {
    if (why) *why = "chain too long to unroll (more than 48 links or 32 joints)";
    return false;
  }
  const int rb = (int)real_size(precision);
  const int limit = 227 * 1024; // registers allow 512 (FP64) / 768 (FP32) resident threads
  // Shared memory of the specialised kernel: one staging strip per warp + the thread-private records (no q tile: every
  // thread reads its own input row from global memory at constant offsets).
  // 256-thread blocks, two per SM: the blocks of an SM then walk the long straight-line code close together, which the
  // instruction cache rewards (measured on cfg2: 2 x 256 threads 0.245 ms, 4 x 128 0.258, 3 x 192 0.266, 1 x 512 0.256)
  if (const char *t = std::getenv("TFK_SPEC_SHAPE")) { // A/B knob for measurements: "threads,blocks" (e.g. "128,5")
    int tt = 0, bb = 0;
    if (std::sscanf(t, "%d,%d", &tt, &bb) == 2 && tt >= 32 && tt <= 512 && tt % 32 == 0 && bb >= 1) {
      const int sm = fused_smem_bytes(tt, h->fz_size, rb);
      if (bb * (sm + 1024) <= limit || std::getenv("TFK_SPEC_SHAPE_FORCE")) { // (FORCE: compile-only register studies)
        *tile = tt;
        *bps = bb;
        *smem = sm;
        return true;
      }
    }
  }
  // FP64 fast (since it runs the portable libm: 2.2 k instead of 3.3 k warp instructions per 32 cfg2 configurations, long
  // scoreboard 1.8 and barrier 0.5 cycles per issue at 2 x 256): four blocks of 128 -- cfg2 0.1386 -> 0.1348 ms, cfg3 0.2036 ->
  // 0.1949 ms (83 % of HBM); 2 x 320: cfg2 0.1319 ms but no room for cfg3's records.  Strict keeps 2 x 256 (cfg2 0.1909 ms against
  // 0.2107 at 4 x 128, 0.1957 at 2 x 320, 0.2238 at 3 x 192; cfg3 0.2606 against 0.2551 / - / 0.2640)
  if (precision == PREC_F64_FAST) {
    const int sm = fused_smem_bytes(128, h->fz_size, rb);
    if (4 * (sm + 1024) <= limit) {
      *tile = 128;
      *bps = 4;
      *smem = sm;
      return true;
    }
  }
  {
    const int sm = fused_smem_bytes(256, h->fz_size, rb);
    int b = precision == PREC_F32 ? 3 : 2;
    while (b > 2 && b * (sm + 1024) > limit) --b;
    if (b * (sm + 1024) <= limit) {
      *tile = 256;
      *bps = b;
      *smem = sm;
      return true;
    }
  }
  // larger records: 128-thread blocks, as many as fit (at least two)
  const int sm = fused_smem_bytes(128, h->fz_size, rb);
  int b = precision == PREC_F32 ? 6 : 4;
  while (b > 2 && b * (sm + 1024) > limit) --b;
  if (b * (sm + 1024) > limit) {
    if (why) *why = "joint records do not fit in shared memory twice";
    return false;
  }
  *tile = 128;
  *bps = b;
  *smem = sm;
  return true;
}

void Engine::spec_start(PlanEntry &e, int precision, bool wait) {
  SpecKernel &sk = e.spec[precision];
  {
    std::lock_guard<std::mutex> lk(sk.mu);
    if (sk.state != SPEC_NONE) {
      if (!(wait && sk.state == SPEC_COMPILING)) return;
    } else {
      std::string why;
      if (!spec_eligible(e, precision, &sk.tile, &sk.bps, &sk.smem, &sk.kind, &why)) throw std::invalid_argument(why);
      if (!nvrtc_available(&why)) {
        sk.state = SPEC_FAILED;
        sk.error = why;
        throw std::runtime_error(why);
      }
      sk.state = SPEC_COMPILING;
      const PlanHeader *h = reinterpret_cast<const PlanHeader *>(e.blob.data());
      SpecOptions o;
      o.precision = precision;
      o.rot_type = h->rot_type;
      o.tile = sk.tile;
      o.blocks_per_sm = sk.bps;
      o.kind = sk.kind;
      const std::string src = plan_to_source(e.blob);
      SpecKernel *p = &sk; // heap-allocated with the plan entry; its destructor joins this thread
      sk.worker = std::thread([p, src, o] {
        std::vector<char> cubin;
        std::string err;
        bool cached = false;
        try {
          cubin = spec_compile(src, o, &cached);
        } catch (const std::exception &ex) {
          err = ex.what();
        }
        std::lock_guard<std::mutex> lk2(p->mu);
        if (err.empty()) {
          p->cubin.swap(cubin);
          p->from_cache = cached;
          p->state = SPEC_COMPILED;
        } else {
          p->error = err;
          p->state = SPEC_FAILED;
        }
      });
    }
  }
  if (wait && sk.worker.joinable()) sk.worker.join();
}

SpecInfo Engine::specialize(const SolveSpec &spec, bool wait, bool compile_only) {
  if (spec.precision < 0 || spec.precision >= PREC_COUNT) throw std::invalid_argument("unknown precision mode");
  PlanEntry &e = plan_for(spec);
  SpecKernel &sk = e.spec[spec.precision];
  spec_start(e, spec.precision, wait);
  SpecInfo info;
  {
    std::lock_guard<std::mutex> lk(sk.mu);
    if (sk.state == SPEC_FAILED) throw std::runtime_error(sk.error);
    info.state = sk.state;
    info.tile = sk.tile;
    info.blocks_per_sm = sk.bps;
    info.smem_bytes = sk.smem;
    info.cubin_bytes = (int64_t)sk.cubin.size();
    info.from_cache = sk.from_cache ? 1 : 0;
  }
  if (!compile_only && info.state == SPEC_COMPILED) { // load now so that the first launch does not pay for it
    int cur = 0;
    ck(cudaGetDevice(&cur), "cudaGetDevice (is a CUDA device present?)");
    std::lock_guard<std::mutex> lk(sk.mu);
    if (sk.state == SPEC_COMPILED) {
      ck(cudaLibraryLoadData(&sk.lib, sk.cubin.data(), nullptr, nullptr, 0, nullptr, nullptr, 0), "cudaLibraryLoadData(specialised kernel)");
      ck(cudaLibraryGetKernel(&sk.kernel, sk.lib, "tfk_spec_kernel"), "cudaLibraryGetKernel");
      sk.state = SPEC_READY;
    }
    info.state = sk.state;
  }
  return info;
}

// Launches the specialised kernel if there is one; otherwise advances the JIT policy and returns false (the caller
// launches the generic kernel).
bool Engine::spec_launch(int device, cudaStream_t stream, PlanEntry &e, int precision, const KernelArgs &a) {
  SpecKernel &sk = e.spec[precision];
  int state;
  {
    std::lock_guard<std::mutex> lk(sk.mu);
    state = sk.state;
  }
  // A stream capture forbids what the JIT does behind the scenes (cudaLibraryLoadData, cudaFuncSetAttribute, cudaMalloc of
  // a differently sized scratch): while the stream is capturing, only a kernel that is already loaded AND has been
  // launched on this device is used; otherwise the capture records the generic kernel and the policy does not advance.
  const bool capturing = stream_is_capturing(stream);
  if (capturing && (state != SPEC_READY || sk.blocks_per_sm.find(device) == sk.blocks_per_sm.end())) return false;
  if (state == SPEC_NONE) {
    if (jit_mode == JIT_AUTO) {
      e.configs_served += a.n;
      if (e.configs_served < jit_threshold) return false;
    }
    try {
      spec_start(e, precision, jit_mode == JIT_ON);
    } catch (const std::exception &ex) { // not eligible / no NVRTC: remember, keep serving with the generic kernel
      std::lock_guard<std::mutex> lk(sk.mu);
      sk.state = SPEC_FAILED;
      sk.error = ex.what();
      return false;
    }
    std::lock_guard<std::mutex> lk(sk.mu);
    state = sk.state;
  }
  if (state == SPEC_COMPILED) {
    std::lock_guard<std::mutex> lk(sk.mu);
    if (cudaLibraryLoadData(&sk.lib, sk.cubin.data(), nullptr, nullptr, 0, nullptr, nullptr, 0) != cudaSuccess ||
        cudaLibraryGetKernel(&sk.kernel, sk.lib, "tfk_spec_kernel") != cudaSuccess) {
      cudaGetLastError();
      sk.state = SPEC_FAILED;
      sk.error = "loading the specialised kernel failed";
      return false;
    }
    sk.state = state = SPEC_READY;
  }
  if (state != SPEC_READY) return false;
  auto it = sk.blocks_per_sm.find(device);
  if (it == sk.blocks_per_sm.end()) {
    const void *fn = reinterpret_cast<const void *>(sk.kernel);
    if (sk.smem > 48 * 1024) ck(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, sk.smem), "cudaFuncSetAttribute(specialised kernel)");
    int b = 0;
    ck(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, fn, sk.tile, sk.smem), "occupancy query (specialised kernel)");
    if (b < 1) throw CudaError("specialised kernel does not fit on an SM");
    it = sk.blocks_per_sm.insert(std::make_pair(device, b)).first;
  }
  // two-phase: configurations per tile (fk_launch.h:twophase_tile_cfgs; the two-phase builds use TFK_TILE threads)
  const int tile_cfgs = sk.kind != 0 ? tile_cfgs_for(a.n, (int64_t)sm_count(device) * it->second, precision == PREC_F32) : sk.tile;
  const int64_t n_tiles = (a.n + tile_cfgs - 1) / tile_cfgs;
  const int grid = (int)std::min<int64_t>(n_tiles, (int64_t)sm_count(device) * it->second);
  KernelArgs args = a;
  if (sk.kind != 0) args.sub_cfg = tile_cfgs;
  args.stage = (sk.kind == 1 && twophase_stage_for(*reinterpret_cast<const PlanHeader *>(e.blob.data())) > 0) ? 1 : 0;
  const PlanHeader *ph = reinterpret_cast<const PlanHeader *>(e.blob.data());
  int rec_reals = ph->fz_size;
  if (sk.kind != 0) { // two-phase: per-output tables and the block-private record scratch, sized for THIS grid
    const int table_bytes = (ph->off_cols - ph->off_outs + 15) & ~15;
    rec_reals = sk.kind == 2 ? ph->rec_size - ph->off_frame : ph->rec_size;
    args.tables = plan_on_device(e, device, stream) + ph->off_outs;
    args.table_bytes = table_bytes;
    args.scratch = workspace(device, stream, (size_t)grid * sk.tile * (size_t)std::max(rec_reals, 1) * real_size(precision)).scratch;
  }
  void *params[] = {&args};
  ck(cudaLaunchKernel(reinterpret_cast<const void *>(sk.kernel), dim3(grid), dim3(sk.tile), params, (size_t)sk.smem, stream), "launch (specialised kernel)");
  e.launched_on.insert(device);
  g_launches.fetch_add(1);
  last_launch.grid = grid;
  last_launch.block = sk.tile;
  last_launch.smem_bytes = sk.smem;
  last_launch.blocks_per_sm = it->second;
  last_launch.rec_reals = rec_reals;
  last_launch.variant = sk.kind == 0 ? KV_FUSED_SPEC : KV_TWOPHASE_SPEC;
  return true;
}

Engine::~Engine() { release_device_memory(); }

void Engine::release_device_memory() {
  int prev = 0;
  if (cudaGetDevice(&prev) != cudaSuccess) return;
  for (auto &e : plans_) retire(e);
  plans_.clear();
  for (auto &kv : workspaces_) {
    cudaSetDevice(kv.first.first);
    cudaFree(kv.second.scratch);
    for (void *p : kv.second.outgrown) cudaFree(p);
  }
  workspaces_.clear();
  for (auto &kv : pipes_) {
    cudaSetDevice(kv.first);
    for (int s = 0; s < 2; ++s) {
      cudaStreamSynchronize(kv.second.stream[s]);
      cudaFree(kv.second.dq[s]);
      cudaFree(kv.second.dP[s]);
      cudaFree(kv.second.dJ[s]);
      cudaStreamDestroy(kv.second.stream[s]);
    }
  }
  pipes_.clear();
  for (auto &kv : pair_bufs_) {
    cudaSetDevice(kv.first.first);
    cudaFree(kv.second.d);
  }
  pair_bufs_.clear();
  for (auto &kv : weight_bufs_) {
    cudaSetDevice(kv.first.first);
    cudaFree(kv.second.d);
  }
  weight_bufs_.clear();
  cudaSetDevice(prev);
}

int Engine::sm_count(int device) {
  auto it = sm_counts_.find(device);
  if (it != sm_counts_.end()) return it->second;
  int n = 0;
  ck(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device), "cudaDeviceGetAttribute");
  sm_counts_[device] = n;
  return n;
}

// Frees everything a plan entry owns on the devices.  Kernels of the entry may still be in flight (the device entry is
// asynchronous, and a fused single-link plan never uploads tables, so `on_device` says nothing about where it ran):
// every device the entry was LAUNCHED on is synchronised first; only then are the tables freed and the specialised
// kernels unloaded (~SpecKernel joins the compile thread and calls cudaLibraryUnload).
void Engine::retire(std::unique_ptr<PlanEntry> &e) {
  if (!e) return;
  int prev = 0;
  const bool have_dev = cudaGetDevice(&prev) == cudaSuccess;
  std::set<int> devs = e->launched_on;
  for (auto &kv : e->on_device) devs.insert(kv.first);
  for (int d : devs)
    if (cudaSetDevice(d) == cudaSuccess) cudaDeviceSynchronize();
  for (auto &kv : e->on_device)
    if (cudaSetDevice(kv.first) == cudaSuccess) cudaFree(kv.second);
  if (have_dev) cudaSetDevice(prev);
  e.reset();
}

// One entry per call SIGNATURE (links, joints, rotation type, base type, Jacobian flag).  The stored model state feeds
// the plan too (angles of joints that are not selected, the base pose of a FIXED base), and every *_host call writes the
// last configuration back into the model: when the state has moved on, the blob is rebuilt; identical bytes (the usual
// case: the state that changed is not read by this plan) keep the entry, its uploaded tables and its specialised
// kernels; different bytes REPLACE the entry in place (IK loops that alternate joint sets do not grow the cache).
// Bounded at 64 signatures, least recently used evicted.
Engine::PlanEntry &Engine::plan_for(const SolveSpec &spec) {
  if (spec.n_link < 0 || spec.n_joint < 0) throw std::invalid_argument("negative id count");
  PlanKey key;
  key.elink_ids.assign(spec.elink_ids, spec.elink_ids + spec.n_link);
  key.joint_ids.assign(spec.joint_ids, spec.joint_ids + spec.n_joint);
  key.rot_type = spec.rot_type;
  key.base_type = spec.base_type;
  key.with_jacobian = spec.with_jacobian ? 1 : 0;
  key.model_version = model.version();
  const uint64_t tick = ++plan_tick_;
  size_t slot = plans_.size();
  for (size_t i = 0; i < plans_.size(); ++i) {
    PlanKey k2 = plans_[i]->key;
    const uint64_t built_for = k2.model_version;
    k2.model_version = key.model_version;
    if (!(k2 == key)) continue;
    if (built_for == key.model_version) {
      plans_[i]->last_used = tick;
      return *plans_[i];
    }
    slot = i;
    break;
  }
  std::vector<unsigned char> blob = build_plan(model, key);
  if (slot < plans_.size()) {
    PlanEntry &e = *plans_[slot];
    if (e.blob == blob) {
      e.key.model_version = key.model_version;
      e.last_used = tick;
      return e;
    }
    retire(plans_[slot]); // same signature, different constants: replaced below
  } else if (plans_.size() >= 64) {
    slot = 0;
    for (size_t i = 1; i < plans_.size(); ++i)
      if (plans_[i]->last_used < plans_[slot]->last_used) slot = i;
    retire(plans_[slot]);
  } else {
    plans_.emplace_back();
  }
  std::unique_ptr<PlanEntry> e(new PlanEntry);
  e->key = key;
  e->last_used = tick;
  e->blob.swap(blob);
  e->params = make_plan_params(e->blob);
  plans_[slot] = std::move(e);
  return *plans_[slot];
}

size_t Engine::plan_cache_size() const { return plans_.size(); }

const unsigned char *Engine::plan_on_device(PlanEntry &e, int device, cudaStream_t stream) {
  auto it = e.on_device.find(device);
  if (it != e.on_device.end()) return it->second;
  unsigned char *d = nullptr;
  ck(cudaMalloc(&d, e.blob.size()), "cudaMalloc(plan)");
  // pageable source: the runtime stages it before returning, and the copy is ordered before the kernel
  cudaError_t err = cudaMemcpyAsync(d, e.blob.data(), e.blob.size(), cudaMemcpyHostToDevice, stream);
  if (err != cudaSuccess) {
    cudaFree(d);
    ck(err, "cudaMemcpyAsync(plan)");
  }
  // other streams may use this copy right away: make it visible device-wide first
  ck(cudaStreamSynchronize(stream), "cudaStreamSynchronize(plan upload)");
  e.on_device[device] = d;
  return d;
}

// Record scratch of one (device, stream).  GROW-ONLY: a buffer that became too small is kept until the engine is
// destroyed, because a CUDA graph captured earlier may still name it in a kernel node.  Growing needs cudaMalloc, which
// a stream capture forbids: a launch that would grow the scratch during a capture fails with a clear message (run the
// signature once on the stream before capturing, include/tinyfk_b200.h).
Engine::Workspace &Engine::workspace(int device, cudaStream_t stream, size_t bytes) {
  Workspace &w = workspaces_[std::make_pair(device, stream)];
  if (w.scratch_bytes < bytes) {
    if (stream_is_capturing(stream))
      throw std::invalid_argument("tfk_*_device inside a stream capture needs more record scratch than this stream has: run the "
                                  "same call once on the stream before capturing");
    void *fresh = nullptr;
    ck(cudaMalloc(&fresh, bytes), "cudaMalloc(records)");
    if (w.scratch) w.outgrown.push_back(w.scratch);
    w.scratch = fresh;
    w.scratch_bytes = bytes;
  }
  return w;
}

size_t Engine::max_pitch(int device) {
  auto it = max_pitch_.find(device);
  if (it != max_pitch_.end()) return it->second;
  int v = 0;
  ck(cudaDeviceGetAttribute(&v, cudaDevAttrMaxPitch, device), "cudaDeviceGetAttribute(max pitch)");
  size_t mp = v > 0 ? (size_t)v : (size_t)0x7fffffff;
  if (const char *e = std::getenv("TFK_MAX_PITCH")) mp = std::min<size_t>(mp, (size_t)std::atoll(e)); // test knob
  max_pitch_[device] = mp;
  return mp;
}

void Engine::copy_columns_d2h(int device, char *dst, size_t dst_pitch, const char *src, size_t src_pitch, size_t width,
                              size_t ncols, cudaStream_t st, const char *what) {
  if (ncols == 0 || width == 0) return;
  const size_t mp = max_pitch(device);
  if (ncols == 1 || (dst_pitch <= mp && src_pitch <= mp)) {
    if (ncols == 1)
      ck(cudaMemcpyAsync(dst, src, width, cudaMemcpyDeviceToHost, st), what);
    else
      ck(cudaMemcpy2DAsync(dst, dst_pitch, src, src_pitch, width, ncols, cudaMemcpyDeviceToHost, st), what);
    return;
  }
  // a 2-D copy with such a pitch fails with cudaErrorInvalidPitchValue (cfg4 at 2^22 configurations: 3.0 GB per column)
  for (size_t c = 0; c < ncols; ++c)
    ck(cudaMemcpyAsync(dst + c * dst_pitch, src + c * src_pitch, width, cudaMemcpyDeviceToHost, st), what);
}

void Engine::launch(int device, cudaStream_t stream, PlanEntry &e, int precision, const void *d_q, int64_t n,
                    void *d_P, void *d_J, int64_t ld_cfg, int64_t row0, const LaunchExtra *ex) {
  const PlanHeader *h = reinterpret_cast<const PlanHeader *>(e.blob.data());
  if (n <= 0 || h->n_out == 0) return;
  const bool sqdist = ex && ex->variant == KV_SQDIST, com = ex && (ex->variant == KV_COM || ex->variant == KV_INERTIA);
  const int n_pairs = sqdist ? ex->n_pairs : 0;
  if (precision < 0 || precision >= PREC_COUNT) throw std::invalid_argument("unknown precision mode");
  const int rb = (int)real_size(precision);
  // per-output tables of the two-phase kernel: [lrec_of_out | masks], contiguous inside the blob
  const int table_bytes = (h->n_out == 1 && !ex) ? 0 : ((h->off_cols - h->off_outs + 15) & ~15);
  const int plan_bytes = table_bytes;
  // kernel variant (fk_kernels.cuh): one requested link -> fused walk + emit with thread-private records
  // in shared memory; several links -> two-phase through block-private scratch
  int variant, rec_reals, stage_reals = 0;
  SmemLayout L;
  if (sqdist) {
    rec_reals = h->rec_size;
    L = smem_layout(table_bytes + ((2 * n_pairs * 4 + 15) & ~15), h->n_in, 0, rb);
    variant = KV_SQDIST;
    const SmemLayout Lh = smem_layout(table_bytes + ((2 * n_pairs * 4 + 15) & ~15), h->n_in, h->off_frame, rb);
    if (3 * (Lh.total + 1024) <= 227 * 1024 && !std::getenv("TFK_NO_SMEM_HOT")) { // (env: A/B knob for measurements)
      L = Lh;
      variant = KV_SQDIST_SMEM;
      rec_reals = h->rec_size - h->off_frame;
    }
  } else if (com) {
    rec_reals = h->rec_size;
    const int n_w = ex->variant == KV_INERTIA ? 7 * ex->n_com : ex->n_com + ex->n_act;
    L = smem_layout(table_bytes + ((n_w * rb + 15) & ~15), h->n_in, 0, rb);
    variant = ex->variant;
  } else if (h->n_out == 1) {
    rec_reals = h->fz_size;
    L = SmemLayout();
    L.total = fused_smem_bytes(TFK_TILE, rec_reals, rb);
    variant = KV_FUSED_SMEM;
    if (L.total > 100 * 1024) { // keep at least two blocks per SM; beyond that records go to L2
      L.total = fused_smem_bytes(TFK_TILE, 0, rb);
      variant = KV_FUSED_GMEM;
    } else if (5 * (L.total + 1024) <= 228 * 1024) {
      variant = KV_FUSED_SMEM5; // a fifth block fits next to four: take the build that leaves registers for it
    }
  } else {
    rec_reals = h->rec_size;
    L = smem_layout(plan_bytes, h->n_in, 0, rb);
    variant = KV_TWOPHASE;
    // the records every (configuration, link) item re-reads go to shared memory when three blocks still fit on an SM
    const SmemLayout Lh = smem_layout(plan_bytes, h->n_in, h->off_frame, rb);
    if (3 * (Lh.total + 1024) <= 227 * 1024 && !std::getenv("TFK_NO_SMEM_HOT")) {
      L = Lh;
      variant = KV_TWOPHASE_SMEM;
      rec_reals = h->rec_size - h->off_frame;
    } else if ((stage_reals = twophase_stage_for(*h)) > 0) {
      L = smem_layout(plan_bytes, h->n_in, stage_reals, rb); // per-warp staging of the joint records (fk_launch.h)
    }
  }
  if ((variant == KV_FUSED_SMEM || variant == KV_FUSED_SMEM5 || variant == KV_TWOPHASE || variant == KV_TWOPHASE_SMEM) && !ex &&
      jit_mode != JIT_OFF) {
    KernelArgs sa = KernelArgs();
    sa.q = d_q;
    sa.n = n;
    sa.row0 = row0;
    sa.ld_cfg = ld_cfg;
    sa.P = d_P;
    sa.J = d_J;
    if (spec_launch(device, stream, e, precision, sa)) return;
  }
  const int smem = L.total;
  if (smem > 200 * 1024) throw std::invalid_argument("plan too large for shared memory");
  auto okey = std::make_tuple(device, precision * 4 + h->rot_type, variant, smem);
  auto oit = occupancy_.find(okey);
  int bps = 0;
  if (oit == occupancy_.end()) {
    // also raises the kernel's dynamic shared memory limit on this device when smem > 48 KB
    ck(precision == PREC_F32        ? occupancy_fk_f32(h->rot_type, variant, smem, &bps)
       : precision == PREC_F64_FAST ? occupancy_fk_f64_fast(h->rot_type, variant, smem, &bps)
                                    : occupancy_fk_f64_strict(h->rot_type, variant, smem, &bps),
       "occupancy query");
    if (bps < 1) throw CudaError("kernel does not fit on an SM");
    occupancy_[okey] = bps;
  } else {
    bps = oit->second;
  }
  const bool twophase = variant == KV_TWOPHASE || variant == KV_TWOPHASE_SMEM;
  const int64_t resident = (int64_t)sm_count(device) * bps;
  const int tile_cfgs = twophase ? tile_cfgs_for(n, resident, false) : (int)TFK_TILE;
  const int64_t n_tiles = (n + tile_cfgs - 1) / tile_cfgs;
  const int grid = (int)std::min<int64_t>(n_tiles, resident);
  KernelArgs a;
  a.scratch = nullptr;
  if (variant != KV_FUSED_SMEM && variant != KV_FUSED_SMEM5) {
    const size_t rec_bytes = (size_t)grid * TFK_TILE * (size_t)std::max(rec_reals, 1) * real_size(precision);
    a.scratch = workspace(device, stream, rec_bytes).scratch;
  }
  a.tables = table_bytes ? plan_on_device(e, device, stream) + h->off_outs : nullptr;
  a.table_bytes = table_bytes;
  a.q = d_q;
  a.n = n;
  a.row0 = row0;
  a.ld_cfg = ld_cfg;
  a.P = d_P;
  a.J = d_J;
  a.pairs = sqdist ? ex->d_pairs : nullptr;
  a.n_pairs = n_pairs;
  a.sub_cfg = twophase ? tile_cfgs : 0;
  a.stage = (variant == KV_TWOPHASE && stage_reals > 0) ? 1 : 0;
  a.weights = com ? ex->d_weights : nullptr;
  a.n_com = com ? ex->n_com : 0;
  a.n_act = com ? ex->n_act : 0;
  a.mass_total = com ? ex->mass_total : 0.0;
  a.vfs = com ? ex->vfs : 0.0;
  a.com_diag = com ? ex->com_diag : 0.0;
  ck(precision == PREC_F32        ? launch_fk_f32(a, *e.params, variant, grid, smem, stream)
     : precision == PREC_F64_FAST ? launch_fk_f64_fast(a, *e.params, variant, grid, smem, stream)
                                  : launch_fk_f64_strict(a, *e.params, variant, grid, smem, stream),
     "kernel launch");
  e.launched_on.insert(device);
  g_launches.fetch_add(1);
  last_launch.grid = grid;
  last_launch.block = TFK_TILE;
  last_launch.smem_bytes = smem;
  last_launch.blocks_per_sm = bps;
  last_launch.rec_reals = rec_reals;
  last_launch.variant = variant;
}

void Engine::solve_device(int device, cudaStream_t stream, const void *d_q, int64_t n, const SolveSpec &spec,
                          void *d_P, void *d_J, int64_t ld_cfg, int64_t row0) {
  if (n < 0) throw std::invalid_argument("negative batch size");
  if (row0 < 0 || row0 + n > ld_cfg) throw std::invalid_argument("rows [row0, row0+n) exceed ld_cfg");
  PlanEntry &e = plan_for(spec);
  if (n == 0) return;
  if (!d_q && reinterpret_cast<const PlanHeader *>(e.blob.data())->n_in > 0) throw std::invalid_argument("q is null");
  if (!d_P) throw std::invalid_argument("P is null");
  if (spec.with_jacobian && !d_J) throw std::invalid_argument("J is null but with_jacobian is set");
  DeviceGuard g(device);
  launch(device, stream, e, spec.precision, d_q, n, d_P, d_J, ld_cfg, row0);
}

// The allocation-free device entry (tfk_solve_fk_device_noalloc; no counterpart in the reference):
// fk_twophase_kernel with its record scratch and the plan's per-output tables in the static device
// array of fk_noalloc.cu.  Calls no cudaMalloc / cudaFree (nor Engine::workspace / plan_on_device, which do); the plan
// itself is host memory (built on the first call with a signature, then cached).  Strict FP64.  The grid is what the
// array holds: a request whose records are large runs with fewer resident blocks, not with more memory.
void Engine::solve_device_noalloc(int device, cudaStream_t stream, const void *d_q, int64_t n, const SolveSpec &spec_in,
                                  void *d_P, void *d_J, int64_t ld_cfg, int64_t row0) {
  if (n < 0) throw std::invalid_argument("negative batch size");
  if (n > (int64_t)TFK_NOALLOC_MAX_CONFIGS_)
    throw std::invalid_argument("tfk_solve_fk_device_noalloc: " + std::to_string(n) + " configurations exceed the maximum of " +
                                std::to_string((int)TFK_NOALLOC_MAX_CONFIGS_) + " per call (TFK_NOALLOC_MAX_CONFIGS); split the batch "
                                "or use tfk_solve_fk_device");
  if (row0 < 0 || row0 + n > ld_cfg) throw std::invalid_argument("rows [row0, row0+n) exceed ld_cfg");
  SolveSpec spec = spec_in;
  spec.precision = PREC_F64_STRICT;
  PlanEntry &e = plan_for(spec);
  const PlanHeader *h = reinterpret_cast<const PlanHeader *>(e.blob.data());
  if (n == 0 || h->n_out == 0) return;
  if (!d_q && h->n_in > 0) throw std::invalid_argument("q is null");
  if (!d_P) throw std::invalid_argument("P is null");
  if (spec.with_jacobian && !d_J) throw std::invalid_argument("J is null but with_jacobian is set");
  const int table_bytes = (h->off_cols - h->off_outs + 15) & ~15;
  if (table_bytes > (int)TFK_NOALLOC_TABLE_BYTES)
    throw std::invalid_argument("tfk_solve_fk_device_noalloc: the per-link tables of this request (" + std::to_string(table_bytes) +
                                " bytes) exceed the " + std::to_string((int)TFK_NOALLOC_TABLE_BYTES) + " bytes reserved for them");
  const size_t per_block = (size_t)TFK_TILE * (size_t)std::max(h->rec_size, 1) * sizeof(double);
  const size_t room = (size_t)TFK_NOALLOC_SCRATCH_BYTES - TFK_NOALLOC_TABLE_BYTES;
  if (per_block > room)
    throw std::invalid_argument("tfk_solve_fk_device_noalloc: one tile of this request needs " + std::to_string(per_block) +
                                " bytes of records, more than the static scratch holds (" + std::to_string(room) + ")");
  const SmemLayout L = smem_layout(table_bytes, h->n_in, 0, (int)sizeof(double));
  if (L.total > 200 * 1024) throw std::invalid_argument("plan too large for shared memory");
  DeviceGuard g(device);
  const auto okey = std::make_tuple(device, (int)h->rot_type, L.total);
  auto oit = occupancy_noalloc_.find(okey);
  if (oit == occupancy_noalloc_.end()) {
    int bps = 0;
    ck(occupancy_fk_noalloc(h->rot_type, L.total, &bps), "occupancy query (allocation-free kernel)");
    if (bps < 1) throw CudaError("kernel does not fit on an SM");
    oit = occupancy_noalloc_.insert(std::make_pair(okey, bps)).first;
  }
  const int64_t n_tiles = (n + TFK_TILE - 1) / TFK_TILE;
  const int64_t resident = (int64_t)sm_count(device) * oit->second;
  const int grid = (int)std::min<int64_t>(std::min<int64_t>(n_tiles, resident), (int64_t)(room / per_block));
  KernelArgs a = KernelArgs();
  a.q = d_q;
  a.n = n;
  a.row0 = row0;
  a.ld_cfg = ld_cfg;
  a.P = d_P;
  a.J = d_J;
  a.sub_cfg = 0; // whole tiles of TFK_TILE configurations
  ck(launch_fk_noalloc(device, a, *e.params, e.blob.data() + h->off_outs, table_bytes, grid, L.total, stream),
     "launch (allocation-free kernel)");
  e.launched_on.insert(device);
  g_launches.fetch_add(1);
  last_launch.grid = grid;
  last_launch.block = TFK_TILE;
  last_launch.smem_bytes = L.total;
  last_launch.blocks_per_sm = oit->second;
  last_launch.rec_reals = h->rec_size;
  last_launch.variant = KV_TWOPHASE;
}

Engine::HostPipe &Engine::pipe(int device) {
  HostPipe &p = pipes_[device];
  if (!p.ready) {
    for (int s = 0; s < 2; ++s) ck(cudaStreamCreateWithFlags(&p.stream[s], cudaStreamNonBlocking), "cudaStreamCreate");
    p.ready = true;
  }
  return p;
}

static void grow(void **buf, size_t *cap, size_t need, cudaStream_t st) {
  if (*cap >= need) return;
  if (*buf) {
    cudaStreamSynchronize(st); // the slot's buffers are only ever used on the slot's stream
    cudaFree(*buf);
    *buf = nullptr;
    *cap = 0;
  }
  ck(cudaMalloc(buf, need), "cudaMalloc(staging)");
  *cap = need;
}

// ---- inter-link squared distances (reference src/wrapper.cpp:189-237) ------------------------------------
// The plan is built over the distinct links of the pairs (RotationType IGNORE); o1/o2 index into it.
Engine::PlanEntry &Engine::plan_for_pairs(const PairSpec &ps, std::vector<int> &o12) {
  if (ps.n_check < 0 || ps.n_joint < 0) throw std::invalid_argument("negative id count");
  std::vector<int> links;
  o12.assign((size_t)2 * ps.n_check, 0);
  for (int side = 0; side < 2; ++side)
    for (int j = 0; j < ps.n_check; ++j) {
      const int id = side == 0 ? ps.link_ids1[j] : ps.link_ids2[j];
      size_t k = 0;
      while (k < links.size() && links[k] != id) ++k;
      if (k == links.size()) links.push_back(id);
      o12[(size_t)side * ps.n_check + j] = (int)k;
    }
  SolveSpec s;
  s.elink_ids = links.data();
  s.n_link = (int)links.size();
  s.joint_ids = ps.joint_ids;
  s.n_joint = ps.n_joint;
  s.rot_type = ROT_IGNORE;
  s.base_type = ps.base_type;
  s.with_jacobian = ps.with_grads;
  s.precision = ps.precision;
  return plan_for(s);
}

void Engine::sqdists_device(int device, cudaStream_t stream, const void *d_q, int64_t n, const PairSpec &ps, void *d_sq,
                            void *d_G, int64_t ld_cfg, int64_t row0) {
  if (n < 0) throw std::invalid_argument("negative batch size");
  if (row0 < 0 || row0 + n > ld_cfg) throw std::invalid_argument("rows [row0, row0+n) exceed ld_cfg");
  std::vector<int> o12;
  PlanEntry &e = plan_for_pairs(ps, o12);
  if (n == 0 || ps.n_check == 0) return;
  if (!d_q && reinterpret_cast<const PlanHeader *>(e.blob.data())->n_in > 0) throw std::invalid_argument("q is null");
  if (!d_sq) throw std::invalid_argument("sqdists is null");
  if (ps.with_grads && !d_G) throw std::invalid_argument("grads is null but with_grads is set");
  DeviceGuard g(device);
  // the pair table is tiny and changes per call: staged through a per-(device, stream) buffer
  PairBuf &pb = pair_bufs_[std::make_pair(device, stream)];
  const size_t need = o12.size() * sizeof(int);
  if (pb.bytes < need) {
    if (pb.d) {
      ck(cudaStreamSynchronize(stream), "cudaStreamSynchronize(pairs)");
      cudaFree(pb.d);
      pb.d = nullptr;
      pb.bytes = 0;
    }
    ck(cudaMalloc(reinterpret_cast<void **>(&pb.d), need), "cudaMalloc(pairs)");
    pb.bytes = need;
  }
  if (pb.host != o12) { // unchanged table: nothing to copy
    ck(cudaStreamSynchronize(stream), "cudaStreamSynchronize(pairs)"); // earlier launches still read the old table
    ck(cudaMemcpyAsync(pb.d, o12.data(), need, cudaMemcpyHostToDevice, stream), "cudaMemcpyAsync(pairs)");
    ck(cudaStreamSynchronize(stream), "cudaStreamSynchronize(pairs)"); // o12 is a pageable temporary
    pb.host = o12;
  }
  LaunchExtra ex = LaunchExtra();
  ex.variant = KV_SQDIST;
  ex.d_pairs = pb.d;
  ex.n_pairs = ps.n_check;
  launch(device, stream, e, ps.precision, d_q, n, d_sq, ps.with_grads ? d_G : nullptr, ld_cfg, row0, &ex);
}

// Host buffers in, host buffers out, current device.  Outputs are small next to the FK arrays
// (n*n_check*(1+n_dof) reals), so the slice is processed in plain chunks on one stream.
void Engine::sqdists_host(const void *q, int64_t n, const PairSpec &ps, void *sq, void *G) {
  if (n < 0) throw std::invalid_argument("negative batch size");
  std::vector<int> o12;
  PlanEntry &e = plan_for_pairs(ps, o12);
  const PlanHeader h = *reinterpret_cast<const PlanHeader *>(e.blob.data());
  if (n == 0 || ps.n_check == 0) return;
  if (!q && h.n_in > 0) throw std::invalid_argument("q is null");
  if (!sq) throw std::invalid_argument("sqdists is null");
  if (ps.with_grads && !G) throw std::invalid_argument("grads is null but with_grads is set");
  int cur = 0;
  ck(cudaGetDevice(&cur), "cudaGetDevice (is a CUDA device present? there is no CPU fallback)");
  HostPipe &pp = pipe(cur);
  cudaStream_t st = pp.stream[0];
  const size_t sz = real_size(ps.precision);
  const size_t in_row = (size_t)h.n_in * sz, row_out = (size_t)ps.n_check * sz;
  const int64_t chunk = std::max<int64_t>(8192, (int64_t)((size_t(64) << 20) / std::max<size_t>(row_out * (1 + h.n_out_cols), 1)));
  const int64_t capn = std::min<int64_t>(chunk, n);
  grow(&pp.dq[0], &pp.cap_q[0], std::max<size_t>((size_t)capn * in_row, 16), st);
  grow(&pp.dP[0], &pp.cap_P[0], (size_t)capn * row_out, st);
  if (ps.with_grads) grow(&pp.dJ[0], &pp.cap_J[0], (size_t)capn * row_out * h.n_out_cols, st);
  const char *qb = static_cast<const char *>(q);
  for (int64_t r0 = 0; r0 < n; r0 += chunk) {
    const int64_t nc = std::min<int64_t>(chunk, n - r0);
    if (in_row) ck(cudaMemcpyAsync(pp.dq[0], qb + (size_t)r0 * in_row, (size_t)nc * in_row, cudaMemcpyHostToDevice, st), "cudaMemcpyAsync(q)");
    sqdists_device(cur, st, pp.dq[0], nc, ps, pp.dP[0], ps.with_grads ? pp.dJ[0] : nullptr, nc, 0);
    ck(cudaMemcpyAsync(static_cast<char *>(sq) + (size_t)r0 * row_out, pp.dP[0], (size_t)nc * row_out, cudaMemcpyDeviceToHost, st), "cudaMemcpyAsync(sq)");
    if (ps.with_grads && h.n_out_cols > 0)
      copy_columns_d2h(cur, static_cast<char *>(G) + (size_t)r0 * row_out, (size_t)n * row_out, static_cast<const char *>(pp.dJ[0]),
                       (size_t)nc * row_out, (size_t)nc * row_out, (size_t)h.n_out_cols, st, "device->host copy of the gradients");
    ck(cudaStreamSynchronize(st), "cudaStreamSynchronize");
  }
  // like the reference, the model is left at the LAST configuration (src/wrapper.cpp:205-209)
  std::vector<size_t> ids(ps.joint_ids, ps.joint_ids + ps.n_joint);
  std::vector<double> last((size_t)h.n_in);
  for (int c = 0; c < h.n_in; ++c)
    last[c] = ps.precision == PREC_F32 ? (double)reinterpret_cast<const float *>(qb)[(size_t)(n - 1) * h.n_in + c]
                                       : reinterpret_cast<const double *>(qb)[(size_t)(n - 1) * h.n_in + c];
  if (ps.n_joint > 0) model.set_joint_angles(ids, std::vector<double>(last.begin(), last.begin() + ps.n_joint));
  if (h.base_type == BASE_PLANER) {
    const double b[6] = {last[ps.n_joint], last[ps.n_joint + 1], 0.0, 0.0, 0.0, last[ps.n_joint + 2]};
    model.set_base_pose_xyzrpy(b);
  } else if (h.base_type == BASE_FLOATING) {
    model.set_base_pose_xyzrpy(&last[ps.n_joint]);
  }
}

// The weight tables of the centre-of-mass / inertia kernels are tiny and rarely change: one buffer per (device, stream),
// re-uploaded (in the kernel's real type) only when the table or the precision differs from what is there.
const void *Engine::upload_weights(int device, cudaStream_t stream, const std::vector<double> &w, int precision) {
  WeightBuf &wb = weight_bufs_[std::make_pair(device, stream)];
  const size_t need = std::max<size_t>(w.size() * real_size(precision), 16);
  if (wb.bytes < need) {
    if (wb.d) {
      ck(cudaStreamSynchronize(stream), "cudaStreamSynchronize(weights)");
      cudaFree(wb.d);
      wb.d = nullptr;
      wb.bytes = 0;
    }
    ck(cudaMalloc(&wb.d, need), "cudaMalloc(weights)");
    wb.bytes = need;
    wb.host.clear();
  }
  if (wb.host != w || wb.precision != precision) {
    ck(cudaStreamSynchronize(stream), "cudaStreamSynchronize(weights)"); // earlier launches still read the old table
    std::vector<float> wf(w.begin(), w.end());
    const void *src = precision == PREC_F32 ? static_cast<const void *>(wf.data()) : static_cast<const void *>(w.data());
    ck(cudaMemcpyAsync(wb.d, src, w.size() * real_size(precision), cudaMemcpyHostToDevice, stream), "cudaMemcpyAsync(weights)");
    ck(cudaStreamSynchronize(stream), "cudaStreamSynchronize(weights)"); // pageable temporaries
    wb.host = w;
    wb.precision = precision;
  }
  return wb.d;
}

// ---- centre of mass (reference src/wrapper.cpp:126-179) --------------------------------------------------------
// Plan links = the model's "_com" links in id order, then the action links (RotationType IGNORE).
void Engine::com_device(int device, cudaStream_t stream, const void *d_q, int64_t n, const ComSpec &cs, void *d_com,
                        void *d_J, int64_t ld_cfg, int64_t row0) {
  if (n < 0) throw std::invalid_argument("negative batch size");
  if (row0 < 0 || row0 + n > ld_cfg) throw std::invalid_argument("rows [row0, row0+n) exceed ld_cfg");
  if (cs.n_action < 0 || cs.n_joint < 0) throw std::invalid_argument("negative id count");
  if (cs.base_type != BASE_FIXED && cs.base_type != BASE_FLOATING)
    throw std::invalid_argument("centre of mass on device buffers takes a FIXED or FLOATING base (pad a PLANER row to x y 0 0 0 yaw as "
                                "python/tinyfk/__init__.py:98-103 does; the host entry does it)");
  const ComTables ct = build_com_tables(model, cs.action_link_ids, cs.forces, cs.n_action);
  const std::vector<int> &links = ct.links;
  SolveSpec s;
  s.elink_ids = links.data();
  s.n_link = (int)links.size();
  s.joint_ids = cs.joint_ids;
  s.n_joint = cs.n_joint;
  s.rot_type = ROT_IGNORE;
  s.base_type = cs.base_type;
  s.with_jacobian = cs.with_jacobian;
  s.precision = cs.precision;
  PlanEntry &e = plan_for(s);
  if (n == 0) return;
  if (!d_q && reinterpret_cast<const PlanHeader *>(e.blob.data())->n_in > 0) throw std::invalid_argument("q is null");
  if (!d_com) throw std::invalid_argument("com is null");
  if (cs.with_jacobian && !d_J) throw std::invalid_argument("J is null but with_jacobian is set");
  DeviceGuard g(device);
  const void *d_w = upload_weights(device, stream, ct.w, cs.precision);
  LaunchExtra ex = LaunchExtra();
  ex.variant = KV_COM;
  ex.d_weights = d_w;
  ex.n_com = ct.n_com;
  ex.n_act = ct.n_act;
  ex.mass_total = ct.mass_total;
  ex.vfs = ct.vfs;
  ex.com_diag = ct.diag;
  launch(device, stream, e, cs.precision, d_q, n, d_com, cs.with_jacobian ? d_J : nullptr, ld_cfg, row0, &ex);
}

// ---- total inertia matrix (reference src/wrapper.cpp:181-187, src/kinematics.cpp:272-317) ---------------------
void Engine::inertia_device(int device, cudaStream_t stream, const void *d_q, int64_t n, const int *joint_ids, int n_joint,
                            int precision, void *d_out, int64_t ld_cfg, int64_t row0) {
  if (n < 0) throw std::invalid_argument("negative batch size");
  if (row0 < 0 || row0 + n > ld_cfg) throw std::invalid_argument("rows [row0, row0+n) exceed ld_cfg");
  if (n_joint < 0) throw std::invalid_argument("negative id count");
  const InertiaTables it = build_inertia_tables(model);
  SolveSpec s;
  s.elink_ids = it.links.data();
  s.n_link = (int)it.links.size();
  s.joint_ids = joint_ids;
  s.n_joint = n_joint;
  s.rot_type = ROT_XYZW; // the link rotations are needed (R I R^T)
  s.base_type = BASE_FIXED;
  s.with_jacobian = 0;
  s.precision = precision;
  PlanEntry &e = plan_for(s);
  if (n == 0) return;
  if (!d_q && n_joint > 0) throw std::invalid_argument("q is null");
  if (!d_out) throw std::invalid_argument("out is null");
  DeviceGuard g(device);
  const void *d_w = upload_weights(device, stream, it.w, precision);
  LaunchExtra ex = LaunchExtra();
  ex.variant = KV_INERTIA;
  ex.d_weights = d_w;
  ex.n_com = (int)it.links.size();
  ex.mass_total = it.mass_total;
  launch(device, stream, e, precision, d_q, n, d_out, nullptr, ld_cfg, row0, &ex);
}

void Engine::inertia_host(const void *q, int64_t n, const int *joint_ids, int n_joint, int precision, void *out) {
  if (n < 0) throw std::invalid_argument("negative batch size");
  if (n == 0) {
    inertia_device(0, nullptr, nullptr, 0, joint_ids, n_joint, precision, nullptr, 0, 0);
    return;
  }
  if (!q && n_joint > 0) throw std::invalid_argument("q is null");
  if (!out) throw std::invalid_argument("out is null");
  int cur = 0;
  ck(cudaGetDevice(&cur), "cudaGetDevice (is a CUDA device present? there is no CPU fallback)");
  HostPipe &pp = pipe(cur);
  cudaStream_t st = pp.stream[0];
  const size_t sz = real_size(precision), in_row = (size_t)n_joint * sz;
  const int64_t chunk = 1 << 16;
  const int64_t capn = std::min<int64_t>(chunk, n);
  grow(&pp.dq[0], &pp.cap_q[0], std::max<size_t>((size_t)capn * in_row, 16), st);
  grow(&pp.dP[0], &pp.cap_P[0], (size_t)capn * 9 * sz, st);
  const char *qb = static_cast<const char *>(q);
  std::vector<char> tmp;
  for (int64_t r0 = 0; r0 < n; r0 += chunk) {
    const int64_t nc = std::min<int64_t>(chunk, n - r0);
    if (in_row) ck(cudaMemcpyAsync(pp.dq[0], qb + (size_t)r0 * in_row, (size_t)nc * in_row, cudaMemcpyHostToDevice, st), "cudaMemcpyAsync(q)");
    inertia_device(cur, st, pp.dq[0], nc, joint_ids, n_joint, precision, pp.dP[0], nc, 0);
    tmp.resize((size_t)nc * 9 * sz);
    ck(cudaMemcpyAsync(tmp.data(), pp.dP[0], tmp.size(), cudaMemcpyDeviceToHost, st), "cudaMemcpyAsync(inertia)");
    ck(cudaStreamSynchronize(st), "cudaStreamSynchronize");
    for (int64_t i = 0; i < nc; ++i)
      for (int k = 0; k < 9; ++k)
        std::memcpy(static_cast<char *>(out) + ((size_t)(r0 + i) * 9 + k) * sz, tmp.data() + ((size_t)k * nc + i) * sz, sz);
  }
  // the model is left at the LAST configuration (src/wrapper.cpp:185)
  if (n_joint > 0) {
    std::vector<size_t> ids(joint_ids, joint_ids + n_joint);
    std::vector<double> last((size_t)n_joint);
    for (int c = 0; c < n_joint; ++c)
      last[c] = precision == PREC_F32 ? (double)reinterpret_cast<const float *>(qb)[(size_t)(n - 1) * n_joint + c]
                                      : reinterpret_cast<const double *>(qb)[(size_t)(n - 1) * n_joint + c];
    model.set_joint_angles(ids, last);
  }
}

// Host buffers in, host buffers out, current device.  q rows are [joints | base]; a PLANER row (x, y, yaw) is padded
// to the FLOATING row the reference's facade builds (python/tinyfk/__init__.py:98-103, 180-182) and the Jacobian keeps
// all six base columns, exactly as solve_com_fk returns it.
//   com  (n, 3) row-major;  J  column-major (3n) x (n_joint + 0|6)
void Engine::com_host(const void *q, int64_t n, const ComSpec &cs_in, void *com, void *J) {
  if (n < 0) throw std::invalid_argument("negative batch size");
  if (cs_in.base_type < 0 || cs_in.base_type > 2) throw std::invalid_argument("unknown base type");
  ComSpec cs = cs_in;
  const int nj = cs.n_joint;
  const int n_in_user = nj + (cs.base_type == BASE_PLANER ? 3 : (cs.base_type == BASE_FLOATING ? 6 : 0));
  if (cs.base_type == BASE_PLANER) cs.base_type = BASE_FLOATING;
  const int n_in = nj + (cs.base_type == BASE_FLOATING ? 6 : 0);
  const int n_dof = n_in;
  if (n == 0) {
    com_device(0, nullptr, nullptr, 0, cs, nullptr, nullptr, 0, 0); // validates ids
    return;
  }
  if (!q && n_in_user > 0) throw std::invalid_argument("q is null");
  if (!com) throw std::invalid_argument("com is null");
  if (cs.with_jacobian && !J) throw std::invalid_argument("J is null but with_jacobian is set");
  int cur = 0;
  ck(cudaGetDevice(&cur), "cudaGetDevice (is a CUDA device present? there is no CPU fallback)");
  HostPipe &pp = pipe(cur);
  cudaStream_t st = pp.stream[0];
  const size_t sz = real_size(cs.precision);
  const size_t in_row = (size_t)n_in * sz;
  const int64_t chunk = std::max<int64_t>(8192, (int64_t)((size_t(64) << 20) / std::max<size_t>(3 * sz * (1 + n_dof), 1)));
  const int64_t capn = std::min<int64_t>(chunk, n);
  grow(&pp.dq[0], &pp.cap_q[0], std::max<size_t>((size_t)capn * in_row, 16), st);
  grow(&pp.dP[0], &pp.cap_P[0], (size_t)capn * 3 * sz, st);
  if (cs.with_jacobian) grow(&pp.dJ[0], &pp.cap_J[0], std::max<size_t>((size_t)capn * 3 * sz * n_dof, 16), st);
  const char *qb = static_cast<const char *>(q);
  std::vector<char> padded, comT;
  for (int64_t r0 = 0; r0 < n; r0 += chunk) {
    const int64_t nc = std::min<int64_t>(chunk, n - r0);
    const char *src = qb + (size_t)r0 * n_in_user * sz;
    if (cs_in.base_type == BASE_PLANER) {
      padded.assign((size_t)nc * in_row, 0);
      for (int64_t i = 0; i < nc; ++i) {
        const char *r = src + (size_t)i * n_in_user * sz;
        char *d = padded.data() + (size_t)i * in_row;
        std::memcpy(d, r, (size_t)(nj + 2) * sz);                          // joints, x, y
        std::memcpy(d + (size_t)(nj + 5) * sz, r + (size_t)(nj + 2) * sz, sz); // yaw
      }
      src = padded.data();
    }
    if (in_row) ck(cudaMemcpyAsync(pp.dq[0], src, (size_t)nc * in_row, cudaMemcpyHostToDevice, st), "cudaMemcpyAsync(q)");
    com_device(cur, st, pp.dq[0], nc, cs, pp.dP[0], cs.with_jacobian ? pp.dJ[0] : nullptr, nc, 0);
    comT.resize((size_t)nc * 3 * sz);
    ck(cudaMemcpyAsync(comT.data(), pp.dP[0], (size_t)nc * 3 * sz, cudaMemcpyDeviceToHost, st), "cudaMemcpyAsync(com)");
    if (cs.with_jacobian && n_dof > 0)
      copy_columns_d2h(cur, static_cast<char *>(J) + (size_t)r0 * 3 * sz, (size_t)n * 3 * sz, static_cast<const char *>(pp.dJ[0]),
                       (size_t)nc * 3 * sz, (size_t)nc * 3 * sz, (size_t)n_dof, st, "device->host copy of the COM Jacobian");
    ck(cudaStreamSynchronize(st), "cudaStreamSynchronize");
    for (int64_t i = 0; i < nc; ++i) // (3, nc) component-major -> (n, 3) rows
      for (int k = 0; k < 3; ++k)
        std::memcpy(static_cast<char *>(com) + ((size_t)(r0 + i) * 3 + k) * sz, comT.data() + ((size_t)k * nc + i) * sz, sz);
  }
  // like the reference, the model is left at the LAST configuration (src/wrapper.cpp:137-141)
  std::vector<size_t> ids(cs.joint_ids, cs.joint_ids + nj);
  std::vector<double> last((size_t)n_in_user);
  for (int c = 0; c < n_in_user; ++c)
    last[c] = cs.precision == PREC_F32 ? (double)reinterpret_cast<const float *>(qb)[(size_t)(n - 1) * n_in_user + c]
                                       : reinterpret_cast<const double *>(qb)[(size_t)(n - 1) * n_in_user + c];
  if (nj > 0) model.set_joint_angles(ids, std::vector<double>(last.begin(), last.begin() + nj));
  if (cs_in.base_type == BASE_PLANER) {
    const double b[6] = {last[nj], last[nj + 1], 0.0, 0.0, 0.0, last[nj + 2]};
    model.set_base_pose_xyzrpy(b);
  } else if (cs_in.base_type == BASE_FLOATING) {
    model.set_base_pose_xyzrpy(&last[nj]);
  }
}


// Columns of J that are constants of the plan never cross PCIe on the host path: a joint that moves none of the
// requested links gives an all-zero column (PlanCol::arec < 0, include/tinyfk.hpp:31-46) and the base-translation
// columns are the unit vectors of src/kinematics.cpp:194-196.  The kernel still writes them into the device staging
// buffer (the device entry's contract is "every entry written"); the host entry fills them with a few CPU threads while
// the copies of the other columns run (cfg2: 96 of 528 output bytes per configuration).
namespace {
template <class T> void fill_const_rows(T *col, size_t r0, size_t r1, int pd, int unit_k) {
  const size_t block = (size_t)pd * 4096;
  for (size_t b = r0; b < r1; b += block) {
    const size_t e = std::min(r1, b + block);
    std::memset(col + b, 0, (e - b) * sizeof(T));
    if (unit_k >= 0) {
      size_t r = b + ((size_t)unit_k + pd - b % pd) % pd;
      for (; r < e; r += pd) col[r] = T(1);
    }
  }
}
} // namespace

// Host buffers in, host buffers out.  The batch is cut into contiguous per-device slices (no
// collective: configurations are independent) and each slice into chunks that ping-pong between two
// stream slots, so chunk k's device->host copies overlap chunk k+1's host->device copy and kernel.
// With several devices every device is driven by its own host thread (a pageable source or destination makes the
// copies synchronous for the issuing thread; one thread per device keeps the devices independent).
void Engine::solve_host(const void *q, int64_t n, const SolveSpec &spec, const int *devices, int n_devices, void *P,
                        void *J) {
  if (n < 0) throw std::invalid_argument("negative batch size");
  PlanEntry &e = plan_for(spec);
  const PlanHeader h = *reinterpret_cast<const PlanHeader *>(e.blob.data());
  if (n == 0 || h.n_out == 0) return;
  if (!q && h.n_in > 0) throw std::invalid_argument("q is null");
  if (!P) throw std::invalid_argument("P is null");
  const bool jac = h.with_jacobian != 0;
  if (jac && !J) throw std::invalid_argument("J is null but with_jacobian is set");

  int cur = 0;
  ck(cudaGetDevice(&cur), "cudaGetDevice (is a CUDA device present? there is no CPU fallback)");
  std::vector<int> devs;
  if (devices && n_devices > 0)
    devs.assign(devices, devices + n_devices);
  else
    devs.push_back(cur);
  int count = 0;
  ck(cudaGetDeviceCount(&count), "cudaGetDeviceCount");
  for (size_t i = 0; i < devs.size(); ++i) {
    if (devs[i] < 0 || devs[i] >= count) throw std::invalid_argument("device ordinal " + std::to_string(devs[i]) + " out of range");
    for (size_t k = 0; k < i; ++k)
      if (devs[k] == devs[i]) throw std::invalid_argument("device ordinal " + std::to_string(devs[i]) + " listed twice");
  }
  const int G = (int)devs.size();

  const size_t sz = real_size(spec.precision);
  const size_t in_row = (size_t)h.n_in * sz;
  const size_t p_cfg = (size_t)h.n_out * sz;                  // bytes per configuration in one P column
  const size_t j_cfg = (size_t)h.n_out * h.pose_dim * sz;     // bytes per configuration in one J column
  const size_t out_cfg = p_cfg * h.pose_dim + (jac ? j_cfg * h.n_out_cols : 0);
  // chunk: ~48 MB of output per slot, a whole number of tiles, at least one full wave of tiles
  int64_t chunk = (int64_t)((size_t(48) << 20) / std::max<size_t>(out_cfg, 1));
  chunk = std::max<int64_t>(chunk, 8192);
  chunk = (chunk + TFK_TILE - 1) / TFK_TILE * TFK_TILE;

  // which columns of J are constants of the plan: -2 = copied from the device, -1 = zeros, k >= 0 = unit vector e_k
  std::vector<int> col_kind(jac ? h.n_out_cols : 0, -2);
  if (jac && !std::getenv("TFK_D2H_ALL")) { // (env: A/B knob for measurements)
    const PlanCol *cols = reinterpret_cast<const PlanCol *>(e.blob.data() + h.off_cols);
    for (int c = 0; c < h.n_cols; ++c)
      if (cols[c].arec < 0) col_kind[c] = -1;
    const int n_trans = h.base_type == BASE_FIXED ? 0 : (h.base_type == BASE_PLANER ? 2 : 3);
    for (int k = 0; k < n_trans; ++k) col_kind[h.n_cols + k] = k;
  }
  std::vector<std::pair<int, int>> copy_runs; // [first, last) runs of columns that do come from the device
  for (int c = 0; c < (int)col_kind.size();) {
    if (col_kind[c] != -2) {
      ++c;
      continue;
    }
    int c1 = c;
    while (c1 < (int)col_kind.size() && col_kind[c1] == -2) ++c1;
    copy_runs.emplace_back(c, c1);
    c = c1;
  }

  const int64_t per_dev = (n + G - 1) / G;
  const char *qb = static_cast<const char *>(q);
  char *Pb = static_cast<char *>(P);
  char *Jb = static_cast<char *>(J);

  // host threads that fill the constant columns (whole batch, split by rows) while the devices work
  std::vector<std::thread> fillers;
  {
    std::vector<int> cc;
    for (int c = 0; c < (int)col_kind.size(); ++c)
      if (col_kind[c] != -2) cc.push_back(c);
    if (!cc.empty()) {
      const size_t rows = (size_t)n * h.n_out * h.pose_dim;
      int nt = 4;
      if (const char *t = std::getenv("TFK_FILL_THREADS")) nt = std::max(1, std::atoi(t));
      nt = (int)std::min<size_t>((size_t)nt, std::max<size_t>(1, rows * cc.size() * sz >> 20)); // one thread per MB at least
      const int pd = h.pose_dim, prec = spec.precision;
      const size_t col_bytes = (size_t)n * j_cfg;
      for (int t = 0; t < nt; ++t) {
        const size_t r0 = rows * t / nt / pd * pd, r1 = (t + 1 == nt) ? rows : rows * (t + 1) / nt / pd * pd;
        fillers.emplace_back([=]() {
          for (int c : cc) {
            char *col = Jb + (size_t)c * col_bytes;
            if (prec == PREC_F32)
              fill_const_rows(reinterpret_cast<float *>(col), r0, r1, pd, col_kind[c]);
            else
              fill_const_rows(reinterpret_cast<double *>(col), r0, r1, pd, col_kind[c]);
          }
        });
      }
    }
  }
  struct JoinAll {
    std::vector<std::thread> &t;
    ~JoinAll() {
      for (auto &x : t)
        if (x.joinable()) x.join();
    }
  } join_fillers{fillers};

  // stream slots and staging buffers of every device, created before the worker threads start (shared maps)
  for (int g = 0; g < G; ++g) {
    const int64_t begin = std::min<int64_t>(n, (int64_t)g * per_dev), end = std::min<int64_t>(n, (int64_t)(g + 1) * per_dev);
    if (begin >= end) continue;
    DeviceGuard dg(devs[g]);
    HostPipe &pp = pipe(devs[g]);
    const int64_t capn = std::min<int64_t>(chunk, per_dev); // size once for a full chunk
    for (int s = 0; s < ((end - begin > chunk) ? 2 : 1); ++s) {
      grow(&pp.dq[s], &pp.cap_q[s], std::max<size_t>((size_t)capn * in_row, 16), pp.stream[s]);
      grow(&pp.dP[s], &pp.cap_P[s], (size_t)capn * p_cfg * h.pose_dim, pp.stream[s]);
      if (jac) grow(&pp.dJ[s], &pp.cap_J[s], std::max<size_t>((size_t)capn * j_cfg * h.n_out_cols, 16), pp.stream[s]);
    }
  }

  auto run_slice = [&](int g) {
    const int64_t begin = std::min<int64_t>(n, (int64_t)g * per_dev), end = std::min<int64_t>(n, (int64_t)(g + 1) * per_dev);
    if (begin >= end) return;
    const int dev = devs[g];
    ck(cudaSetDevice(dev), "cudaSetDevice");
    HostPipe &pp = pipes_.find(dev)->second;
    int64_t k = 0;
    for (int64_t r0 = begin; r0 < end; ++k) {
      const int64_t nc = std::min<int64_t>(chunk, end - r0);
      const int s = (int)(k & 1);
      cudaStream_t st = pp.stream[s];
      if (in_row) {
        NvtxRange r("tfk_solve_fk_host: H2D q");
        ck(cudaMemcpyAsync(pp.dq[s], qb + (size_t)r0 * in_row, (size_t)nc * in_row, cudaMemcpyHostToDevice, st), "cudaMemcpyAsync(q)");
      }
      {
        NvtxRange r("tfk_solve_fk_host: kernel");
        std::lock_guard<std::mutex> lk(launch_mu_);
        launch(dev, st, e, spec.precision, pp.dq[s], nc, pp.dP[s], jac ? pp.dJ[s] : nullptr, nc, 0);
      }
      // one contiguous run per output column (column-major destination dimensioned for all n): a single strided
      // copy per array keeps the copy engine busy back to back (16 separate copies per chunk cost ~5-10 % of PCIe time;
      // letting the kernel store straight into the pinned outputs was measured too and is slower)
      NvtxRange r("tfk_solve_fk_host: D2H P, J");
      copy_columns_d2h(dev, Pb + (size_t)r0 * p_cfg, (size_t)n * p_cfg, static_cast<const char *>(pp.dP[s]), (size_t)nc * p_cfg,
                       (size_t)nc * p_cfg, (size_t)h.pose_dim, st, "device->host copy of P");
      for (const auto &run : copy_runs)
        copy_columns_d2h(dev, Jb + (size_t)run.first * n * j_cfg + (size_t)r0 * j_cfg, (size_t)n * j_cfg,
                         static_cast<const char *>(pp.dJ[s]) + (size_t)run.first * nc * j_cfg, (size_t)nc * j_cfg, (size_t)nc * j_cfg,
                         (size_t)(run.second - run.first), st, "device->host copy of J");
      r0 += nc;
    }
    ck(cudaStreamSynchronize(pp.stream[0]), "cudaStreamSynchronize");
    ck(cudaStreamSynchronize(pp.stream[1]), "cudaStreamSynchronize");
  };

  std::vector<std::exception_ptr> errs(G);
  if (G == 1) {
    try {
      run_slice(0);
    } catch (...) {
      errs[0] = std::current_exception();
    }
  } else {
    std::vector<std::thread> workers;
    for (int g = 0; g < G; ++g)
      workers.emplace_back([&, g]() {
        try {
          run_slice(g);
        } catch (...) {
          errs[g] = std::current_exception();
        }
      });
    for (auto &w : workers) w.join();
  }
  cudaSetDevice(cur);
  for (auto &x : fillers) x.join();
  fillers.clear();
  for (int g = 0; g < G; ++g)
    if (errs[g]) std::rethrow_exception(errs[g]);

  // like the reference, the model is left at the LAST configuration (src/wrapper.cpp:85-89)
  std::vector<size_t> ids(spec.joint_ids, spec.joint_ids + spec.n_joint);
  std::vector<double> last((size_t)h.n_in);
  for (int c = 0; c < h.n_in; ++c)
    last[c] = spec.precision == PREC_F32 ? (double)reinterpret_cast<const float *>(qb)[(size_t)(n - 1) * h.n_in + c]
                                         : reinterpret_cast<const double *>(qb)[(size_t)(n - 1) * h.n_in + c];
  if (spec.n_joint > 0) model.set_joint_angles(ids, std::vector<double>(last.begin(), last.begin() + spec.n_joint));
  if (h.base_type == BASE_PLANER) {
    const double b[6] = {last[spec.n_joint], last[spec.n_joint + 1], 0.0, 0.0, 0.0, last[spec.n_joint + 2]};
    model.set_base_pose_xyzrpy(b);
  } else if (h.base_type == BASE_FLOATING) {
    model.set_base_pose_xyzrpy(&last[spec.n_joint]);
  }
}

} // namespace tfk
