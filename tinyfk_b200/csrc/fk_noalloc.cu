// fk_noalloc.cu -- the allocation-free variant of the several-link kernel (tfk_solve_fk_device_noalloc,
// include/tinyfk_b200.h).  No counterpart in the reference: added for callers
// that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
//
// fk_twophase_kernel (fk_kernels.cuh) keeps its per-configuration records in a block-private global SCRATCH and reads
// the plan's per-output tables from a device copy; the ordinary entry allocates both with cudaMalloc (Engine::workspace,
// Engine::plan_on_device).  Here both live in ONE file-scope static device array of TFK_NOALLOC_SCRATCH_BYTES:
//
//     [ 0, TFK_NOALLOC_TABLE_BYTES )                            the plan's per-output tables (re-written by every call)
//     [ TFK_NOALLOC_TABLE_BYTES, TFK_NOALLOC_SCRATCH_BYTES )    records, grid x TFK_TILE x rec_size reals
//
// Every GPU that loads this module holds the array until the process exits.  In HOST code the array's name is only a
// host-side stand-in, not a device address (a kernel handed it would fault or, where the GPU can reach host memory,
// silently use the wrong memory): the device address is taken with cudaGetSymbolAddress, once per device, with that
// device current (noalloc_scratch_address below), and that pointer is what the kernel gets.
//
// Every call shares the array and finds the previous call's contents in it.  Nothing has to start at zero: phase A of a
// tile writes every record that phase B of the same tile reads (records of the valid configurations only; tail lanes are
// clamped to a valid item), and the tables are copied in front of every launch.  Calls that OVERLAP ON THE GPU (two
// streams, two host threads, two models on the same device) race on the array: serialise them.
//
// Strict FP64 (this file is compiled with -fmad=false, like fk_kernels_strict.cu): same arithmetic, same bits as
// tfk_solve_fk_device in TFK_F64_STRICT.
#include <map>
#include <mutex>

#include "fk_kernels.cuh"
#include "fk_noalloc.h"

namespace tfk {

static __device__ __align__(256) unsigned char g_noalloc_scratch[TFK_NOALLOC_SCRATCH_BYTES];

namespace {
std::mutex g_noalloc_mu;
std::map<int, unsigned char *> g_noalloc_addr; // device ordinal -> device address of g_noalloc_scratch on that device

// `device` must be the current device (the engine's DeviceGuard sees to it): the symbol is resolved in the current context
cudaError_t noalloc_scratch_address(int device, unsigned char **out) {
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  auto it = g_noalloc_addr.find(device);
  if (it != g_noalloc_addr.end()) {
    *out = it->second;
    return cudaSuccess;
  }
  int cur = -1;
  cudaError_t e = cudaGetDevice(&cur);
  if (e != cudaSuccess) return e;
  if (cur != device) return cudaErrorInvalidDevice;
  void *p = nullptr;
  e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
  if (e != cudaSuccess) return e;
  if (p == nullptr) return cudaErrorInvalidSymbol;
  g_noalloc_addr[device] = static_cast<unsigned char *>(p);
  *out = static_cast<unsigned char *>(p);
  return cudaSuccess;
}

template <int ROT> const void *noalloc_kernel() {
  return reinterpret_cast<const void *>(&fk_twophase_kernel<FkCore<MathPortable, false, ROT, TFK_TILE>, false>);
}
const void *noalloc_kernel_of(int rot_type) {
  return rot_type == 0 ? noalloc_kernel<0>() : (rot_type == 1 ? noalloc_kernel<1>() : noalloc_kernel<2>());
}
} // namespace

cudaError_t occupancy_fk_noalloc(int rot_type, int smem_bytes, int *blocks_per_sm) {
  const void *fn = noalloc_kernel_of(rot_type);
  if (smem_bytes > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
    if (e != cudaSuccess) return e;
  }
  return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, fn, TFK_TILE, smem_bytes);
}

// a: every field but `tables` and `scratch` filled in by the caller.  tables_host / table_bytes: the plan's per-output
// tables (inside the plan blob, which outlives the call).  Enqueues one small host->device copy and one kernel on `stream`.
cudaError_t launch_fk_noalloc(int device, KernelArgs a, const PlanParams &pp, const unsigned char *tables_host, int table_bytes,
                              int grid, int smem_bytes, cudaStream_t stream) {
  if (table_bytes < 0 || table_bytes > TFK_NOALLOC_TABLE_BYTES) return cudaErrorInvalidValue;
  const size_t rec_bytes = (size_t)grid * TFK_TILE * (size_t)(pp.h.rec_size > 0 ? pp.h.rec_size : 1) * sizeof(double);
  if (grid < 1 || rec_bytes > (size_t)TFK_NOALLOC_SCRATCH_BYTES - TFK_NOALLOC_TABLE_BYTES) return cudaErrorInvalidValue;
  unsigned char *base = nullptr;
  cudaError_t e = noalloc_scratch_address(device, &base);
  if (e != cudaSuccess) return e;
  if (table_bytes > 0) {
    // stream-ordered in front of the kernel, and behind the previous call's kernel on this stream (which reads the same bytes)
    e = cudaMemcpyAsync(base, tables_host, (size_t)table_bytes, cudaMemcpyHostToDevice, stream);
    if (e != cudaSuccess) return e;
  }
  a.tables = base;
  a.table_bytes = table_bytes;
  a.scratch = base + TFK_NOALLOC_TABLE_BYTES;
  if (pp.h.rot_type == 0)
    fk_twophase_kernel<FkCore<MathPortable, false, 0, TFK_TILE>, false><<<grid, TFK_TILE, smem_bytes, stream>>>(a, pp);
  else if (pp.h.rot_type == 1)
    fk_twophase_kernel<FkCore<MathPortable, false, 1, TFK_TILE>, false><<<grid, TFK_TILE, smem_bytes, stream>>>(a, pp);
  else
    fk_twophase_kernel<FkCore<MathPortable, false, 2, TFK_TILE>, false><<<grid, TFK_TILE, smem_bytes, stream>>>(a, pp);
  return cudaGetLastError();
}

} // namespace tfk
