// jit.cpp -- see jit.hpp
#include "jit.hpp"

#include <dlfcn.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <mutex>
#include <stdexcept>

#include "disk_cache.hpp"
#include "fk_launch.h"
#include "jit_headers.inc" // build/: k_jit_header_names / k_jit_header_sources / k_jit_n_headers

namespace tfk {
namespace {

// ---- NVRTC through dlopen -----------------------------------------------------------------------------------
typedef struct _nvrtcProgram *nvrtcProgram;
struct Nvrtc {
  void *so = nullptr;
  std::string error;
  int (*Version)(int *, int *) = nullptr;
  int (*CreateProgram)(nvrtcProgram *, const char *, const char *, int, const char *const *, const char *const *) = nullptr;
  int (*DestroyProgram)(nvrtcProgram *) = nullptr;
  int (*CompileProgram)(nvrtcProgram, int, const char *const *) = nullptr;
  int (*GetCUBINSize)(nvrtcProgram, size_t *) = nullptr;
  int (*GetCUBIN)(nvrtcProgram, char *) = nullptr;
  int (*GetProgramLogSize)(nvrtcProgram, size_t *) = nullptr;
  int (*GetProgramLog)(nvrtcProgram, char *) = nullptr;
  const char *(*GetErrorString)(int) = nullptr;
  int major = 0, minor = 0;
};

template <class F> bool sym(void *so, const char *name, F &f) {
  f = reinterpret_cast<F>(dlsym(so, name));
  return f != nullptr;
}

Nvrtc &nvrtc() {
  static Nvrtc n;
  static std::once_flag once;
  std::call_once(once, [] {
    std::vector<std::string> names;
    if (const char *p = std::getenv("TFK_NVRTC")) names.push_back(p);
    names.push_back("libnvrtc.so.12");
    names.push_back("/usr/local/cuda/lib64/libnvrtc.so.12");
    if (const char *h = std::getenv("CUDA_HOME")) names.push_back(std::string(h) + "/lib64/libnvrtc.so.12");
    names.push_back("libnvrtc.so");
    names.push_back("libnvrtc.so.13");
    for (const std::string &s : names) {
      n.so = dlopen(s.c_str(), RTLD_NOW | RTLD_LOCAL);
      if (n.so) break;
    }
    if (!n.so) {
      n.error = "NVRTC not found (libnvrtc.so.12; set TFK_NVRTC to its path)";
      return;
    }
    const bool ok = sym(n.so, "nvrtcVersion", n.Version) && sym(n.so, "nvrtcCreateProgram", n.CreateProgram) &&
                    sym(n.so, "nvrtcDestroyProgram", n.DestroyProgram) && sym(n.so, "nvrtcCompileProgram", n.CompileProgram) &&
                    sym(n.so, "nvrtcGetCUBINSize", n.GetCUBINSize) && sym(n.so, "nvrtcGetCUBIN", n.GetCUBIN) &&
                    sym(n.so, "nvrtcGetProgramLogSize", n.GetProgramLogSize) && sym(n.so, "nvrtcGetProgramLog", n.GetProgramLog) &&
                    sym(n.so, "nvrtcGetErrorString", n.GetErrorString);
    if (!ok) {
      n.error = "NVRTC library lacks a required entry point";
      dlclose(n.so);
      n.so = nullptr;
      return;
    }
    n.Version(&n.major, &n.minor);
  });
  return n;
}

// ---- disk cache: disk_cache.hpp (header with 128 bits of key, payload length and payload hash; private directory) ------
const char k_cache_magic[8] = {'T', 'F', 'K', 'C', 'U', 'B', 'N', '1'};

std::vector<std::string> compile_options(const SpecOptions &o) {
  std::vector<std::string> v;
  v.push_back("--gpu-architecture=sm_100a");
  v.push_back("--std=c++17");
  v.push_back("--generate-line-info");
  if (o.precision == PREC_F64_STRICT) v.push_back("--fmad=false"); // as fk_kernels_strict.cu: no FMA contraction
  return v;
}

} // namespace

bool nvrtc_available(std::string *why) {
  Nvrtc &n = nvrtc();
  if (!n.so && why) *why = n.error;
  return n.so != nullptr;
}

std::string spec_source(const std::string &plan_source, const SpecOptions &o) {
  if (o.tile < 32 || o.tile > 512 || o.tile % 32) throw std::invalid_argument("tile must be a multiple of 32 up to 512");
  if (o.rot_type < 0 || o.rot_type > 2) throw std::invalid_argument("bad rot_type");
  const bool f32 = o.precision == PREC_F32;
  std::string s;
  s += "#define TFK_SPEC_PLAN 1\n";
  // which helpers are inlined into the straight-line code (tuning knob for experiments: TFK_JIT_DEFINES="A=1;B=1")
  if (const char *d = std::getenv("TFK_JIT_DEFINES")) {
    std::string x(d);
    size_t i = 0;
    while (i < x.size()) {
      size_t j = x.find(';', i);
      if (j == std::string::npos) j = x.size();
      std::string kv = x.substr(i, j - i);
      const size_t eq = kv.find('=');
      if (!kv.empty()) s += "#define " + (eq == std::string::npos ? kv : kv.substr(0, eq) + " " + kv.substr(eq + 1)) + "\n";
      i = j + 1;
    }
  } else {
    s += "#define TFK_SPEC_INLINE 1\n";
    // sin/cos out of line and two joints per call: pays wherever the libm is libdevice's (FP32 cfg4 0.340 against 0.357 ms
    // without, FP64-fast cfg3 0.212 against 0.219 ms).  The strict single-link kernels inline the portable sin/cos (26 FP64
    // operations, fkmath.h) one joint at a time: cfg2 0.1987 -> 0.1961 ms, cfg3 0.267 -> 0.256 ms
    if (!(o.kind == 0 && o.precision == PREC_F64_STRICT)) s += "#define TFK_SPEC_SINCOS_CALL 1\n#define TFK_SPEC_SINCOS_PAIR 1\n";
  }
  // several links, FP64 (strict and fast): one rolled copy of the walk, fk_core.h TFK_WALK_LOOP (FP64-fast cfg4: 0.873 ->
  // 0.811 ms with the walk rolled, cfg5 0.880 -> 0.872 ms; FP32 code is half the size and keeps the unrolled walk)
  if (o.kind != 0 && o.precision != PREC_F32) s += "#define TFK_SPEC_ROLLED_WALK 1\n";
  // FP64 fast: the portable libm (explicit fused operations, constant-bank coefficients, one division per atan2 taken as a
  // reciprocal-multiply, none in asin) instead of libdevice: cfg2 0.1389 -> 0.1356 ms, cfg3 0.2086 -> 0.2039 ms (fk_math.h)
  if (o.precision == PREC_F64_FAST) s += "#define TFK_FAST_PORTABLE_LIBM 1\n";
  s += "#define TFK_TILE_CFG " + std::to_string(o.tile) + "\n";
  s += "#include \"fk_plan.h\"\n";
  s += plan_source;
  s += "#include \"fk_kernels.cuh\"\n";
  s += "extern \"C\" __global__ void __launch_bounds__(tfk::TFK_TILE, " + std::to_string(o.blocks_per_sm) +
       ") tfk_spec_kernel(const tfk::KernelArgs a) {\n";
  const std::string core = std::string("tfk::FkCore<") +
                           (f32 ? "tfk::MathNative32, true, " : (o.precision == PREC_F64_FAST ? "tfk::MathFast64, true, " : "tfk::MathPortable, false, ")) +
                           std::to_string(o.rot_type) + ", tfk::TFK_TILE" +
                           // several links: records in the pair layout of fk_core.h ("record layout"): every record access moves
                           // whole pairs (cfg4 strict 0.808 -> 0.742 ms, cfg4 FP32 0.421 -> 0.340 ms, cfg5 FP32 0.411 -> 0.348 ms,
                           // cfg5 FP64-fast 0.862 -> 0.788 ms).  Single-link kernels: FP32 gains 1 % (cfg2 0.0681 -> 0.0674 ms, cfg3
                           // 0.1111 -> 0.1097 ms) and takes it, FP64 loses 1 % and keeps the plain layout.  TFK_NO_REC_PAIRS=1: A/B knob
                           (((o.kind != 0 || f32) && !std::getenv("TFK_NO_REC_PAIRS")) ? ", true>" : ">");
  if (o.kind == 0)
    s += "  tfk::fk_fused_spec_body<" + core + ", true, " + std::to_string(o.blocks_per_sm) + ">(a);\n}\n";
  else
    s += "  tfk::fk_twophase_spec_body<" + core + ", " + (o.kind == 2 ? "true" : "false") + ">(a);\n}\n";
  return s;
}

std::vector<char> spec_compile(const std::string &plan_source, const SpecOptions &o, bool *from_cache) {
  if (from_cache) *from_cache = false;
  Nvrtc &n = nvrtc();
  if (!n.so) throw std::runtime_error(n.error);
  const std::string src = spec_source(plan_source, o);
  const std::vector<std::string> opts = compile_options(o);
  // cache key: everything that determines the cubin (two FNV-1a streams with different offsets: 128 bits)
  using diskcache::fnv1a;
  uint64_t key = fnv1a(src), key_b = fnv1a(src, diskcache::k_second_offset);
  for (const std::string &x : opts) key = fnv1a(x, key), key_b = fnv1a(x, key_b);
  for (int i = 0; i < k_jit_n_headers; ++i) key = fnv1a(k_jit_header_sources[i], key), key_b = fnv1a(k_jit_header_sources[i], key_b);
  const std::string ver = "nvrtc " + std::to_string(n.major) + "." + std::to_string(n.minor) + " sm_100a";
  key = fnv1a(ver, key), key_b = fnv1a(ver, key_b);
  char name[64];
  std::snprintf(name, sizeof name, "spec_%016llx.cubin", (unsigned long long)key);
  {
    std::vector<char> cached;
    if (diskcache::read(name, k_cache_magic, key, key_b, cached, 65)) {
      if (from_cache) *from_cache = true;
      return cached;
    }
  }
  nvrtcProgram prog = nullptr;
  int rc = n.CreateProgram(&prog, src.c_str(), "tfk_spec.cu", k_jit_n_headers, k_jit_header_sources, k_jit_header_names);
  if (rc != 0) throw std::runtime_error(std::string("nvrtcCreateProgram: ") + n.GetErrorString(rc));
  std::vector<const char *> copts;
  for (const std::string &x : opts) copts.push_back(x.c_str());
  rc = n.CompileProgram(prog, (int)copts.size(), copts.data());
  if (rc != 0) {
    size_t ls = 0;
    n.GetProgramLogSize(prog, &ls);
    std::string log(ls, '\0');
    if (ls) n.GetProgramLog(prog, &log[0]);
    n.DestroyProgram(&prog);
    throw std::runtime_error(std::string("plan specialisation failed to compile (") + n.GetErrorString(rc) + "):\n" + log);
  }
  size_t cs = 0;
  rc = n.GetCUBINSize(prog, &cs);
  std::vector<char> cubin(cs);
  if (rc == 0 && cs) rc = n.GetCUBIN(prog, cubin.data());
  n.DestroyProgram(&prog);
  if (rc != 0 || cs == 0) throw std::runtime_error("NVRTC produced no cubin");
  diskcache::write(name, k_cache_magic, key, key_b, cubin.data(), cubin.size()); // best effort
  return cubin;
}

} // namespace tfk
