// jit.hpp -- run-time specialisation of the fused kernel on one plan.
//
// The generic kernels interpret the plan: every slot costs uniform loads, tests and branches on its fields.  For a
// query that is evaluated millions of times (one robot, one end effector, one joint set) the plan is a constant, so
// the SAME kernel source (fk_kernels.cuh, fk_core.h) is compiled once more with the plan as compile-time constants
// (plan.cpp:plan_to_source): the plan loops unroll, every test on a plan field folds, URDF constants become
// immediates.  Same arithmetic, same compiler flags, same bits -- fewer instructions.
//
// NVRTC is loaded with dlopen (libnvrtc.so.12 from the CUDA toolkit or the one PyTorch ships); the headers are
// embedded in this library (embed_headers.py).  When NVRTC is not available nothing is lost but the speed-up: the
// generic kernel keeps serving the plan.  Compiled cubins are cached on disk (TFK_CACHE_DIR, default
// ~/.cache/tinyfk_b200).
#ifndef TFK_JIT_HPP
#define TFK_JIT_HPP

#include <string>
#include <vector>

namespace tfk {

struct SpecOptions {
  int precision;  // PREC_F64_STRICT | PREC_F32 | PREC_F64_FAST
  int rot_type;   // compile-time RotationType of FkCore
  int tile;       // threads per block == configurations per tile (a multiple of 32)
  int blocks_per_sm;
  int kind = 0;   // 0: fused kernel (one requested link); 1: two-phase kernel; 2: two-phase, joint records in shared memory
};

// The translation unit handed to NVRTC for this plan (exposed for tests and for the cache key).
std::string spec_source(const std::string &plan_source, const SpecOptions &o);

// Compiles to an sm_100a cubin.  Throws std::runtime_error with the compiler log on failure, or when NVRTC cannot
// be loaded.  Needs no GPU.  `from_cache` (optional) tells whether the cubin came from the disk cache.
std::vector<char> spec_compile(const std::string &plan_source, const SpecOptions &o, bool *from_cache = nullptr);

bool nvrtc_available(std::string *why = nullptr);

} // namespace tfk
#endif
