// fk_kernels.cuh -- the sm_100a batch kernels.  Persistent blocks of TFK_TILE threads, one tile of
// TFK_TILE configurations per iteration.  Included by fk_kernels_strict.cu (built with -fmad=false:
// bit-faithful FP64) and fk_kernels_fast.cu (FMA contraction on: FP32 mode).
//
//   fk_fused_kernel     one requested link (the usual end-effector query): one thread per configuration
//                       walks the chain and writes the pose row and every Jacobian column itself.  The
//                       joint records are thread-private and live in shared memory, so HBM sees only
//                       q (read once), P and J (written once).
//   fk_twophase_kernel  several links: phase A (thread per configuration) -> records -> phase B (thread
//                       per (configuration, link)), so that consecutive threads own consecutive rows.
//
// Memory plan (DESIGN.md "Data layout in HBM"):
//   q        [N][n_in] row-major input.  A tile's rows are one contiguous run: staged into shared memory
//            with coalesced loads (odd row pitch -> conflict-free per-thread reads).
//   P        [pose_dim][N*n_out] column-major (reference layout, wrapper.cpp:80,123): consecutive
//            threads write consecutive addresses.
//   J        [n_out_cols][N*n_out*pose_dim] column-major, every entry written (zeros included).  A
//            warp's share of one column is ONE contiguous run of 32*pose_dim reals: it is transposed
//            through a warp-private shared-memory strip and written with full-sector, fully coalesced
//            streaming stores (st.global.cs).
//   records  fused: [fz_size][TILE] in shared memory (global scratch only if they do not fit);
//            two-phase: [grid][rec_size][TILE] block-private scratch, written in phase A and re-read in
//            phase B of the same tile (L2-resident; the output stores are streaming so they do not evict it).
//   plan     a by-value __grid_constant__ kernel parameter (PlanParams, fk_plan.h): constant bank, uniform
//            datapath; only the two-phase kernel's per-output tables (indexed per lane) go to shared memory.
#ifndef TFK_FK_KERNELS_CUH
#define TFK_FK_KERNELS_CUH

#if !defined(__CUDACC_RTC__)
#include <cuda_runtime.h>

#include <type_traits>
#endif

#include "fk_core.h"
#include "fk_launch.h"

namespace tfk {

enum { TFK_STG = 32 * 7 }; // reals in a warp's staging strip: 32 rows x pose_dim <= 7

template <class T> struct Pair2;
template <> struct Pair2<double> { typedef double2 type; };
template <> struct Pair2<float> { typedef float2 type; };

// Output sink of the kernels.  put_J must be called by all 32 lanes of the warp together, once per
// Jacobian column, in increasing column order (it advances a running column pointer).
template <class T, int D> struct StagedSink {
  T *pP;   // P + head: this lane's row in column 0 of P (head = row*n_out + o)
  T *pJ;   // J + head0*D + lane: this lane's first element of the warp's run in the current column of J
  long long ldP, ldJ;
  T *stg_w; // warp-private strip of TFK_STG reals, + lane*D (where this lane's column entries go)
  T *stg_r; // the strip + lane (what this lane copies out)
  int lane, cnt; // cnt = valid lanes * D
  bool valid;
  bool wide; // every column of the warp's run and the strip are aligned for pair (2 * sizeof(T)) accesses

  __device__ __forceinline__ void begin() const { __syncwarp(); }
  __device__ __forceinline__ void put_P(int d, T v) const {
    if (valid) __stcs(pP + d * ldP, v);
  }
  __device__ __forceinline__ void put_J(int, const T (&v)[7]) {
#pragma unroll
    for (int r = 0; r < D; ++r) stg_w[r] = v[r];
    __syncwarp();
    if (cnt == 32 * D) { // full warp: D coalesced 256-byte (FP64) stores, no predicates
#pragma unroll
      for (int k = 0; k < D; ++k) __stcs(pJ + 32 * k, stg_r[32 * k]);
    } else {
#pragma unroll
      for (int k = 0; k < D; ++k)
        if (lane + 32 * k < cnt) __stcs(pJ + 32 * k, stg_r[32 * k]);
    }
    __syncwarp();
    pJ += ldJ;
  }
  // a column that is the unit vector e_k on every row group (base translation columns, src/kinematics.cpp:194-196):
  // element j of the warp's run is row j % D -- written straight from registers, no strip, no barrier
  __device__ __forceinline__ void put_J_unit(int, int k) {
    const int row0 = lane % D;
#pragma unroll
    for (int i = 0; i < D; ++i) {
      const int row = (row0 + (32 * i) % D) % D;
      if (lane + 32 * i < cnt) __stcs(pJ + 32 * i, row == k ? T(1) : T(0));
    }
    pJ += ldJ;
  }
  // two consecutive columns in one round (needs 2 * 32 * D <= TFK_STG reals: pose_dim 3): half the warp barriers, and
  // the second column's strip reads and stores overlap the first one's
  __device__ __forceinline__ void put_J2(int, const T (&v0)[7], const T (&v1)[7]) {
    static_assert(D != 3 || 2 * 32 * D <= TFK_STG, "strip too small for two columns");
#pragma unroll
    for (int r = 0; r < D; ++r) {
      stg_w[r] = v0[r];
      stg_w[32 * D + r] = v1[r];
    }
    __syncwarp();
    T *pJ1 = pJ + ldJ;
    if (cnt == 32 * D && wide) {
      // the two strips are contiguous (2 * 32 * D reals = 32 * D pairs): D rounds of 16-byte (FP64) strip reads and
      // stores instead of 2 * D rounds of 8-byte ones; pair j < 16 * D belongs to the first column
      typedef typename Pair2<T>::type T2;
      const T2 *s2 = reinterpret_cast<const T2 *>(stg_r - lane);
#pragma unroll
      for (int k = 0; k < D; ++k) {
        const int j = lane + 32 * k;
        T *col = j < 16 * D ? pJ - lane + 2 * j : pJ1 - lane + 2 * (j - 16 * D);
        __stcs(reinterpret_cast<T2 *>(col), s2[j]);
      }
    } else if (cnt == 32 * D) {
#pragma unroll
      for (int k = 0; k < D; ++k) __stcs(pJ + 32 * k, stg_r[32 * k]);
#pragma unroll
      for (int k = 0; k < D; ++k) __stcs(pJ1 + 32 * k, stg_r[32 * D + 32 * k]);
    } else {
#pragma unroll
      for (int k = 0; k < D; ++k)
        if (lane + 32 * k < cnt) {
          __stcs(pJ + 32 * k, stg_r[32 * k]);
          __stcs(pJ1 + 32 * k, stg_r[32 * D + 32 * k]);
        }
    }
    __syncwarp();
    pJ += 2 * ldJ;
  }
};

// builds the sink of one warp-load of consecutive rows: head0 = P row of lane 0, nvalid = rows present
template <class T, int D> struct StagedFactory {
  typedef StagedSink<T, D> Sink;
  T *P, *J;
  long long ldP, head0;
  int nvalid;
  T *stg;
  __device__ __forceinline__ Sink make() const {
    Sink s;
    const int lane = threadIdx.x & 31;
    s.lane = lane;
    s.ldP = ldP;
    s.ldJ = ldP * D;
    s.pP = P + head0 + lane;
    s.pJ = J + head0 * D + lane;
    s.stg_w = stg + lane * D;
    s.stg_r = stg + lane;
    s.cnt = nvalid * D;
    s.valid = lane < nvalid;
    s.wide = ((reinterpret_cast<unsigned long long>(J + head0 * D) | reinterpret_cast<unsigned long long>(stg)) % (2 * sizeof(T)) == 0) &&
             s.ldJ % 2 == 0;
    return s;
  }
};

template <class T>
__device__ __forceinline__ void stage_q_tile(T *qs, int q_pitch, const T *q, long long i0, int nv, int n_in) {
  // the tile's rows are contiguous in HBM: coalesced loads, scattered (row, col) stores
  if (n_in == 0) return;
  const T *src = q + i0 * n_in;
  const int total = nv * n_in;
  int r = (int)threadIdx.x / n_in, c = (int)threadIdx.x - r * n_in;
  const int dr = TFK_TILE / n_in, dc = TFK_TILE - dr * n_in;
  for (int k = threadIdx.x; k < total; k += TFK_TILE) {
    qs[r * q_pitch + c] = __ldcs(src + k);
    r += dr;
    c += dc;
    if (c >= n_in) {
      c -= n_in;
      ++r;
    }
  }
}

// pull the block's NEXT tile of q into L2 while this one is being worked on (no registers, no smem)
template <class T> __device__ __forceinline__ void prefetch_q_tile(const T *q, long long i0, long long n, int n_in) {
  if (i0 >= n || n_in == 0) return;
  const long long rows = (n - i0 < TFK_TILE) ? (n - i0) : (long long)TFK_TILE;
  const char *p = reinterpret_cast<const char *>(q + i0 * n_in) + (size_t)threadIdx.x * 128;
  if ((long long)threadIdx.x * 128 < rows * n_in * (long long)sizeof(T)) asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}

// BPS: resident blocks per SM the register allocation aims at (5 when the shared-memory footprint allows it)
template <class Core, bool SMEM_REC, int BPS>
__global__ void __launch_bounds__(TFK_TILE, BPS) fk_fused_kernel(const KernelArgs a, const __grid_constant__ PlanParams pp) {
  typedef typename Core::T T;
  extern __shared__ __align__(16) unsigned char smem[];
  PlanView pv;
  pv.h = &pp.h;
  pv.slots = pp.slots;
  pv.cols = pp.cols;
  pv.masks = pp.mask0;
  pv.lrec_of_out = nullptr; // one output: never read
  const PlanHeader &h = pp.h;
  // shared memory: [one staging strip per warp | thread-private records].  There is no q tile: every thread reads its
  // own input row from global memory (a warp's rows are one contiguous run; L1 serves the repeated sectors), which
  // measured faster than staging the tile (no staging pass, one barrier less).
  T *stg0 = reinterpret_cast<T *>(smem);
  T *rec = SMEM_REC ? stg0 + fused_strip_reals(TFK_TILE)
                    : static_cast<T *>(a.scratch) + (size_t)blockIdx.x * TFK_TILE * h.fz_size;
  rec += threadIdx.x;
  const T *q = static_cast<const T *>(a.q);
  const int warp = threadIdx.x >> 5;

  StagedFactory<T, Core::D> fac;
  fac.P = static_cast<T *>(a.P);
  fac.J = static_cast<T *>(a.J);
  fac.ldP = a.ld_cfg; // n_out == 1
  fac.stg = stg0 + warp * TFK_STG;

  const long long n_tiles = (a.n + TFK_TILE - 1) / TFK_TILE;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long i0 = tile * TFK_TILE;
    const int nv = (a.n - i0 < TFK_TILE) ? (int)(a.n - i0) : (int)TFK_TILE;
    prefetch_q_tile(q, i0 + (long long)gridDim.x * TFK_TILE, a.n, h.n_in);
    int nvalid = nv - warp * 32;
    nvalid = nvalid < 0 ? 0 : (nvalid > 32 ? 32 : nvalid);
    if (nvalid > 0) {
      const int row = (int)threadIdx.x < nv ? (int)threadIdx.x : nv - 1; // tail lanes redo the last row, store nothing
      fac.nvalid = nvalid;
      fac.head0 = a.row0 + i0 + warp * 32;
      Core::run_fused(pv, q + (i0 + row) * h.n_in, rec, TFK_TILE, fac);
    }
    __syncthreads(); // keeps the block's warps together in the code (instruction cache)
  }
}

#if defined(TFK_SPEC_PLAN)
// Plan-specialised build of the fused kernel: the plan is the compile-time constant tfk::spec::k_* (generated by
// plan.cpp:plan_to_source and compiled at run time, jit.cpp), the plan loops are fully unrolled and every test on a
// plan field is folded, so what is left is the straight-line arithmetic of THIS robot/query.  Same source, same
// compiler flags, same bits as the generic kernel.
template <class Core, bool SMEM_REC, int BPS>
__device__ __forceinline__ void fk_fused_spec_body(const KernelArgs &a) {
  typedef typename Core::T T;
  extern __shared__ __align__(16) unsigned char smem[];
  PlanView pv;
  pv.h = &spec::k_h;
  pv.slots = spec::k_slots;
  pv.cols = spec::k_cols;
  pv.masks = spec::k_mask0;
  pv.lrec_of_out = nullptr;
  const PlanHeader &h = spec::k_h;
  // no q tile: every thread reads its own input row straight from global memory (constant offsets; the rows of a warp
  // are one contiguous run, L1 serves the repeats) -- shared memory holds only the warps' staging strips and the
  // records, which lets more warps stay resident
  T *stg0 = reinterpret_cast<T *>(smem);
  T *rec = stg0 + fused_strip_reals(TFK_TILE);
  rec += Core::TB * threadIdx.x; // fk_core.h "record layout"
  const T *q = static_cast<const T *>(a.q);
  const int warp = threadIdx.x >> 5;

  StagedFactory<T, Core::D> fac;
  fac.P = static_cast<T *>(a.P);
  fac.J = static_cast<T *>(a.J);
  fac.ldP = a.ld_cfg;
  fac.stg = stg0 + warp * TFK_STG;

  const long long n_tiles = (a.n + TFK_TILE - 1) / TFK_TILE;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long i0 = tile * TFK_TILE;
    const int nv = (a.n - i0 < TFK_TILE) ? (int)(a.n - i0) : (int)TFK_TILE;
    prefetch_q_tile(q, i0 + (long long)gridDim.x * TFK_TILE, a.n, h.n_in);
    int nvalid = nv - warp * 32;
    nvalid = nvalid < 0 ? 0 : (nvalid > 32 ? 32 : nvalid);
    if (nvalid > 0) {
      const int row = (int)threadIdx.x < nv ? (int)threadIdx.x : nv - 1;
      fac.nvalid = nvalid;
      fac.head0 = a.row0 + i0 + warp * 32;
      Core::run_fused(pv, q + (i0 + row) * h.n_in, rec, TFK_TILE, fac);
    }
    __syncthreads(); // keeps the block's warps together in the straight-line code (instruction cache)
  }
}
#endif

// SMEM_HOT: [base | arec | prec] (what every (configuration, link) item re-reads) live in shared memory and only the
// branch-point frames and the link records go through the block-private scratch; otherwise all records do.
// pv0: where the plan's header, slots and columns are (the kernel parameter, or the compile-time tables of a
// plan-specialised build); the per-output tables (indexed per lane in phase B) are copied to shared memory either way.
template <class Core, bool SMEM_HOT>
__device__ __forceinline__ void fk_twophase_body(const KernelArgs &a, const PlanView &pv0) {
  typedef typename Core::T T;
  extern __shared__ __align__(16) unsigned char smem[];
  // per-output tables (indexed per lane in phase B) -> shared memory; everything else stays in the constant bank
  for (int i = threadIdx.x * 16; i < a.table_bytes; i += TFK_TILE * 16)
    *reinterpret_cast<int4 *>(smem + i) = *reinterpret_cast<const int4 *>(a.tables + i);
  __syncthreads();
  const PlanHeader &h = *pv0.h;
  PlanView pv = pv0;
  pv.lrec_of_out = reinterpret_cast<const int32_t *>(smem);
  pv.masks = reinterpret_cast<const uint32_t *>(smem + (h.off_masks - h.off_outs));
  // without SMEM_HOT the record area of the layout is the warps' staging area for [base | axes | positions] (fk_launch.h:
  // twophase_stage_reals), when the launch reserved it
  const bool stage = !SMEM_HOT && a.stage != 0;
  const int hot = h.off_frame - h.off_base;
  const SmemLayout L = smem_layout(a.table_bytes, h.n_in,
                                   SMEM_HOT ? h.off_frame : (stage ? twophase_stage_reals(h.n_out, h.off_base, h.off_frame) : 0), (int)sizeof(T));
  T *qs = reinterpret_cast<T *>(smem + L.q_off);
  // frames and link records: element k of configuration i at rec_link[(h.off_frame + k) * TILE + i] (biased when the hot part is
  // not in the scratch, so that the plan's absolute offsets apply to both arrays)
  T *rec_link = SMEM_HOT ? static_cast<T *>(a.scratch) + ((long long)blockIdx.x * (h.rec_size - h.off_frame) - h.off_frame) * TFK_TILE
                         : static_cast<T *>(a.scratch) + (size_t)blockIdx.x * TFK_TILE * h.rec_size;
  T *rec = SMEM_HOT ? reinterpret_cast<T *>(smem + L.rec_off) : rec_link;
  const T *q = static_cast<const T *>(a.q);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

  StagedFactory<T, Core::D> fac;
  fac.P = static_cast<T *>(a.P);
  fac.J = static_cast<T *>(a.J);
  fac.ldP = a.ld_cfg * h.n_out;
  fac.stg = qs + warp * TFK_STG; // the q tile is dead during phase B (TILE*q_pitch >= 4*TFK_STG)

  // configurations per tile: TFK_TILE, or less when the batch has too few full tiles to load the SMs evenly
  // (twophase_tile_cfgs, fk_launch.h) -- phase A then leaves threads idle, phase B spreads the tile's items over all
  const int tc = a.sub_cfg > 0 ? a.sub_cfg : (int)TFK_TILE;
  const long long n_tiles = (a.n + tc - 1) / tc;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long i0 = tile * tc;
    const int nv = (a.n - i0 < tc) ? (int)(a.n - i0) : tc;
    stage_q_tile(qs, L.q_pitch, q, i0, nv, h.n_in);
    prefetch_q_tile(q, i0 + (long long)gridDim.x * tc, a.n, h.n_in);
    __syncthreads();
    if ((int)threadIdx.x < nv)
      Core::phase_a(pv, qs + threadIdx.x * L.q_pitch, rec + Core::TB * threadIdx.x, rec_link + Core::TB * threadIdx.x, TFK_TILE);
    __syncthreads();
    const int items = nv * h.n_out;
    const long long hbase = (a.row0 + i0) * h.n_out;
    // one (configuration, link) item per lane.  The NEXT item's link record is requested from the scratch (L2) before the
    // current item is worked on, so its latency hides behind a whole item instead of stalling the item's first arithmetic:
    // always for position-only records (3 reals), for 7-real records where the arithmetic leaves registers (Core::SPARE:
    // the FP32 mode, cfg5 0.348 -> 0.340 ms; the FP64 kernels sit at or near their register cap and do not gain).
    enum { PREFETCH = Core::LREC == 3 || Core::SPARE };
    int il = 0, o = 0;
    typename Core::P e;
    e.p.x = e.p.y = e.p.z = T(0);
    e.q.x = e.q.y = e.q.z = e.q.w = T(0);
    if (PREFETCH && warp * 32 < items) {
      const int itc = warp * 32 + lane < items ? warp * 32 + lane : items - 1;
      il = itc / h.n_out;
      o = itc - il * h.n_out;
      e = Core::load_link(pv, rec_link + Core::TB * il, o);
    }
    for (int it0 = warp * 32; it0 < items; it0 += TFK_TILE) {
      if (!PREFETCH) {
        const int itc = it0 + lane < items ? it0 + lane : items - 1;
        il = itc / h.n_out;
        o = itc - il * h.n_out;
        e = Core::load_link(pv, rec_link + Core::TB * il, o);
      }
      const int il_c = il, o_c = o;
      const typename Core::P e_c = e;
      if (PREFETCH && it0 + TFK_TILE < items) {
        const int itn = it0 + TFK_TILE + lane < items ? it0 + TFK_TILE + lane : items - 1;
        il = itn / h.n_out;
        o = itn - il * h.n_out;
        e = Core::load_link(pv, rec_link + Core::TB * il, o);
      }
      fac.nvalid = items - it0 < 32 ? items - it0 : 32;
      fac.head0 = hbase + it0;
      StagedSink<T, Core::D> out = fac.make();
      if (!SMEM_HOT && stage) {
        // the (at most two) configurations of this warp's 32 items: their joint records into the warp's staging area,
        // element k of configuration c0 + s at wst[CoreS::ix(k) + TB * s]; every load is in flight before the first one
        // is used.  (Issuing the NEXT batch's loads before this batch is worked on -- eight more registers per lane --
        // measured slower in the one mode that would need it: FP64-fast cfg5 0.840 vs 0.788 ms.)
        typedef typename Core::template Restride<2>::type CoreS;
        T *wst = reinterpret_cast<T *>(smem + L.rec_off) + warp * 2 * hot;
        const int last = it0 + 31 < items ? it0 + 31 : items - 1;
        const int c0 = it0 / h.n_out, nc = last / h.n_out - c0 + 1;
        __syncwarp();
        for (int idx = lane; idx < 2 * hot; idx += 32) {
          const int k = idx >> 1, s = idx & 1;
          if (s < nc) wst[CoreS::ix(k) + CoreS::TB * s] = rec[Core::ix(h.off_base + k) + Core::TB * (c0 + s)];
        }
        __syncwarp();
        CoreS::phase_b(pv, wst - CoreS::ix(h.off_base) + CoreS::TB * (il_c - c0), e_c, o_c, out);
      } else {
        Core::phase_b(pv, rec + Core::TB * il_c, e_c, o_c, out);
      }
    }
    __syncthreads();
  }
}

template <class Core, bool SMEM_HOT>
__global__ void __launch_bounds__(TFK_TILE, 4) fk_twophase_kernel(const KernelArgs a, const __grid_constant__ PlanParams pp) {
  PlanView pv;
  pv.h = &pp.h;
  pv.slots = pp.slots;
  pv.cols = pp.cols;
  pv.lrec_of_out = nullptr; // set by the body (shared memory)
  pv.masks = nullptr;
#if defined(TFK_GENERIC_PLAIN_RECORDS) // A/B build: the generic kernel on the plain record layout
  fk_twophase_body<Core, SMEM_HOT>(a, pv);
#else
  fk_twophase_body<typename Core::Paired, SMEM_HOT>(a, pv); // records in the pair layout (fk_core.h "record layout")
#endif
}

#if defined(TFK_SPEC_PLAN)
// Plan-specialised build of the two-phase kernel (several requested links): header, slots and columns are the
// compile-time tables tfk::spec::k_*, so the walk of phase A and the column loops of phase B unroll into the
// straight-line arithmetic of THIS robot/query; the per-output tables still come from a.tables.  Same source, same
// compiler flags, same bits as the generic kernel.
template <class Core, bool SMEM_HOT> __device__ __forceinline__ void fk_twophase_spec_body(const KernelArgs &a) {
  PlanView pv;
  pv.h = &spec::k_h;
  pv.slots = spec::k_slots;
  pv.cols = spec::k_cols;
  pv.lrec_of_out = nullptr;
  pv.masks = nullptr;
  fk_twophase_body<Core, SMEM_HOT>(a, pv);
}
#endif

// compute_inter_link_squared_dists (src/wrapper.cpp:189-237): phase A as in the two-phase kernel (plan over the
// union of the pairs' links, positions only), then one thread per (configuration, pair): consecutive threads
// write consecutive entries of sqdists and of every gradient column -- coalesced without staging.
template <class Core, bool SMEM_HOT>
__global__ void __launch_bounds__(TFK_TILE, 4) fk_sqdist_kernel(const KernelArgs a, const __grid_constant__ PlanParams pp) {
  typedef typename Core::T T;
  extern __shared__ __align__(16) unsigned char smem[];
  for (int i = threadIdx.x * 16; i < a.table_bytes; i += TFK_TILE * 16)
    *reinterpret_cast<int4 *>(smem + i) = *reinterpret_cast<const int4 *>(a.tables + i);
  const PlanHeader &h = pp.h;
  const int tab = (a.table_bytes + 15) & ~15;
  int *pairs = reinterpret_cast<int *>(smem + tab);
  for (int i = threadIdx.x; i < 2 * a.n_pairs; i += TFK_TILE) pairs[i] = a.pairs[i];
  __syncthreads();
  PlanView pv;
  pv.h = &pp.h;
  pv.slots = pp.slots;
  pv.cols = pp.cols;
  pv.lrec_of_out = reinterpret_cast<const int32_t *>(smem);
  pv.masks = reinterpret_cast<const uint32_t *>(smem + (h.off_masks - h.off_outs));
  const SmemLayout L = smem_layout(tab + ((2 * a.n_pairs * 4 + 15) & ~15), h.n_in, SMEM_HOT ? h.off_frame : 0, (int)sizeof(T));
  T *qs = reinterpret_cast<T *>(smem + L.q_off);
  // as fk_twophase_kernel: base/axis/position records in shared memory when they fit, frames and link records in scratch
  T *rec_link = SMEM_HOT ? static_cast<T *>(a.scratch) + ((long long)blockIdx.x * (h.rec_size - h.off_frame) - h.off_frame) * TFK_TILE
                         : static_cast<T *>(a.scratch) + (size_t)blockIdx.x * TFK_TILE * h.rec_size;
  T *rec = SMEM_HOT ? reinterpret_cast<T *>(smem + L.rec_off) : rec_link;
  const T *q = static_cast<const T *>(a.q);
  T *sq = static_cast<T *>(a.P);
  T *G = static_cast<T *>(a.J);
  const long long ldG = a.ld_cfg * a.n_pairs;

  const long long n_tiles = (a.n + TFK_TILE - 1) / TFK_TILE;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long i0 = tile * TFK_TILE;
    const int nv = (a.n - i0 < TFK_TILE) ? (int)(a.n - i0) : (int)TFK_TILE;
    stage_q_tile(qs, L.q_pitch, q, i0, nv, h.n_in);
    prefetch_q_tile(q, i0 + (long long)gridDim.x * TFK_TILE, a.n, h.n_in);
    __syncthreads();
    if ((int)threadIdx.x < nv) Core::phase_a(pv, qs + threadIdx.x * L.q_pitch, rec + threadIdx.x, rec_link + threadIdx.x, TFK_TILE);
    __syncthreads();
    const int items = nv * a.n_pairs;
    const long long hbase = (a.row0 + i0) * a.n_pairs;
    for (int it = threadIdx.x; it < items; it += TFK_TILE) {
      const int il = it / a.n_pairs;
      const int j = it - il * a.n_pairs;
      Core::pair_sqdist(pv, rec + il, rec_link + il, pairs[j], pairs[a.n_pairs + j], true, sq + hbase + it, G ? G + hbase + it : nullptr, ldG);
    }
    __syncthreads();
  }
}

// ---- centre of mass (src/wrapper.cpp:126-179) ----------------------------------------------------------------
// Phase A as above over the model's "_com" links (+ action links).  Phase B: the same thread walks the work
// items of its configuration (Core::com_item: COM, then one Jacobian column at a time), so a warp works on ONE
// column for 32 consecutive configurations: the relevance test and the column's plan entries are warp-uniform,
// record reads are coalesced (element k of configuration i at rec[k*TILE + i]) and the three stores of a warp
// cover 768 contiguous bytes.
//   com  (3, ld_cfg) component-major: com[k*ld + row]
//   J    column-major (3*ld_cfg) x n_dof: J[c*3*ld + 3*row + k]   (Eigen's layout of wrapper.cpp:134)
template <class Core>
__global__ void __launch_bounds__(TFK_TILE, 4) fk_com_kernel(const KernelArgs a, const __grid_constant__ PlanParams pp) {
  typedef typename Core::T T;
  extern __shared__ __align__(16) unsigned char smem[];
  for (int i = threadIdx.x * 16; i < a.table_bytes; i += TFK_TILE * 16)
    *reinterpret_cast<int4 *>(smem + i) = *reinterpret_cast<const int4 *>(a.tables + i);
  const PlanHeader &h = pp.h;
  const int tab = (a.table_bytes + 15) & ~15;
  const int n_w = a.n_com + a.n_act;
  T *w = reinterpret_cast<T *>(smem + tab);
  for (int i = threadIdx.x; i < n_w; i += TFK_TILE) w[i] = static_cast<const T *>(a.weights)[i];
  __syncthreads();
  PlanView pv;
  pv.h = &pp.h;
  pv.slots = pp.slots;
  pv.cols = pp.cols;
  pv.lrec_of_out = reinterpret_cast<const int32_t *>(smem);
  pv.masks = reinterpret_cast<const uint32_t *>(smem + (h.off_masks - h.off_outs));
  const SmemLayout L = smem_layout(tab + ((n_w * (int)sizeof(T) + 15) & ~15), h.n_in, 0, (int)sizeof(T));
  T *qs = reinterpret_cast<T *>(smem + L.q_off);
  T *rec = static_cast<T *>(a.scratch) + (size_t)blockIdx.x * TFK_TILE * h.rec_size;
  const T *q = static_cast<const T *>(a.q);
  T *com = static_cast<T *>(a.P);
  T *J = static_cast<T *>(a.J);
  typename Core::ComArgs ca;
  ca.w = w;
  ca.n_com = a.n_com;
  ca.n_act = a.n_act;
  ca.mass_total = (T)a.mass_total;
  ca.vfs = (T)a.vfs;
  const int n_base = (h.base_type == 0) ? 0 : 6;
  // work items per configuration: COM, joint columns, three base rotation columns (translation columns are constant)
  const int n_items = 1 + (J ? h.n_cols + (n_base ? 3 : 0) : 0);

  const long long n_tiles = (a.n + TFK_TILE - 1) / TFK_TILE;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long i0 = tile * TFK_TILE;
    const int nv = (a.n - i0 < TFK_TILE) ? (int)(a.n - i0) : (int)TFK_TILE;
    stage_q_tile(qs, L.q_pitch, q, i0, nv, h.n_in);
    prefetch_q_tile(q, i0 + (long long)gridDim.x * TFK_TILE, a.n, h.n_in);
    __syncthreads();
    if ((int)threadIdx.x < nv) Core::phase_a(pv, qs + threadIdx.x * L.q_pitch, rec + threadIdx.x, TFK_TILE);
    __syncthreads();
    const long long row = a.row0 + i0 + threadIdx.x;
    if ((int)threadIdx.x < nv) {
      TFK_NOUNROLL
      for (int item = 0; item < n_items; ++item) {
        const typename Core::V v = Core::com_item(pv, rec + threadIdx.x, item, ca);
        if (item == 0) {
          __stcs(com + row, v.x);
          __stcs(com + a.ld_cfg + row, v.y);
          __stcs(com + 2 * a.ld_cfg + row, v.z);
        } else {
          const int col = (item <= h.n_cols) ? item - 1 : item + 2; // rotation columns follow the three translation columns
          T *o = J + (size_t)col * 3 * a.ld_cfg + 3 * row;
          __stcs(o, v.x);
          __stcs(o + 1, v.y);
          __stcs(o + 2, v.z);
        }
      }
      if (J && n_base) { // d com / d base translation = diag * I (host-evaluated constant, engine.cpp)
        const T dg = (T)a.com_diag;
        for (int k = 0; k < 3; ++k) {
          T *o = J + (size_t)(h.n_cols + k) * 3 * a.ld_cfg + 3 * row;
          __stcs(o, k == 0 ? dg : T(0));
          __stcs(o + 1, k == 1 ? dg : T(0));
          __stcs(o + 2, k == 2 ? dg : T(0));
        }
      }
    }
    __syncthreads();
  }
}

// ---- total inertia matrix (src/kinematics.cpp:272-317) ---------------------------------------------------------
// Phase A over the "_com" links with their rotations recorded (RotationType XYZW plan), then the same thread
// reduces its configuration: out (9, ld_cfg) component-major, out[(3*i + j)*ld + row].
template <class Core>
__global__ void __launch_bounds__(TFK_TILE, 4) fk_inertia_kernel(const KernelArgs a, const __grid_constant__ PlanParams pp) {
  typedef typename Core::T T;
  extern __shared__ __align__(16) unsigned char smem[];
  for (int i = threadIdx.x * 16; i < a.table_bytes; i += TFK_TILE * 16)
    *reinterpret_cast<int4 *>(smem + i) = *reinterpret_cast<const int4 *>(a.tables + i);
  const PlanHeader &h = pp.h;
  const int tab = (a.table_bytes + 15) & ~15;
  const int n_w = 7 * a.n_com;
  T *w = reinterpret_cast<T *>(smem + tab);
  for (int i = threadIdx.x; i < n_w; i += TFK_TILE) w[i] = static_cast<const T *>(a.weights)[i];
  __syncthreads();
  PlanView pv;
  pv.h = &pp.h;
  pv.slots = pp.slots;
  pv.cols = pp.cols;
  pv.lrec_of_out = reinterpret_cast<const int32_t *>(smem);
  pv.masks = reinterpret_cast<const uint32_t *>(smem + (h.off_masks - h.off_outs));
  const SmemLayout L = smem_layout(tab + ((n_w * (int)sizeof(T) + 15) & ~15), h.n_in, 0, (int)sizeof(T));
  T *qs = reinterpret_cast<T *>(smem + L.q_off);
  T *rec = static_cast<T *>(a.scratch) + (size_t)blockIdx.x * TFK_TILE * h.rec_size;
  const T *q = static_cast<const T *>(a.q);
  T *out = static_cast<T *>(a.P);
  const long long n_tiles = (a.n + TFK_TILE - 1) / TFK_TILE;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long i0 = tile * TFK_TILE;
    const int nv = (a.n - i0 < TFK_TILE) ? (int)(a.n - i0) : (int)TFK_TILE;
    stage_q_tile(qs, L.q_pitch, q, i0, nv, h.n_in);
    __syncthreads();
    if ((int)threadIdx.x < nv) {
      Core::phase_a(pv, qs + threadIdx.x * L.q_pitch, rec + threadIdx.x, TFK_TILE);
      T m[9];
      Core::inertia_total(pv, rec + threadIdx.x, w, a.n_com, (T)a.mass_total, m);
      const long long row = a.row0 + i0 + threadIdx.x;
#pragma unroll
      for (int k = 0; k < 9; ++k) __stcs(out + k * a.ld_cfg + row, m[k]);
    }
    __syncthreads();
  }
}

#if !defined(__CUDACC_RTC__) // host-side launch helpers (not part of a run-time compiled translation unit)
// the pair kernel exists for RotationType IGNORE only (positions); other instantiations never reach it
template <class Core> typename std::enable_if<Core::D == 3, void>::type launch_sqdist(const KernelArgs &a, const PlanParams &pp, bool hot, int grid, int smem_bytes, cudaStream_t stream) {
  if (hot)
    fk_sqdist_kernel<Core, true><<<grid, TFK_TILE, smem_bytes, stream>>>(a, pp);
  else
    fk_sqdist_kernel<Core, false><<<grid, TFK_TILE, smem_bytes, stream>>>(a, pp);
}
template <class Core> typename std::enable_if<Core::D != 3, void>::type launch_sqdist(const KernelArgs &, const PlanParams &, bool, int, int, cudaStream_t) {}
template <class Core> typename std::enable_if<Core::D == 3, cudaError_t>::type occupancy_sqdist(bool hot, int smem_bytes, int *blocks_per_sm) {
  const void *fn = hot ? (const void *)fk_sqdist_kernel<Core, true> : (const void *)fk_sqdist_kernel<Core, false>;
  if (smem_bytes > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
    if (e != cudaSuccess) return e;
  }
  return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, fn, TFK_TILE, smem_bytes);
}
template <class Core> typename std::enable_if<Core::D != 3, cudaError_t>::type occupancy_sqdist(bool, int, int *blocks_per_sm) {
  *blocks_per_sm = 0;
  return cudaSuccess;
}
template <class Core> typename std::enable_if<Core::D == 7, void>::type launch_inertia(const KernelArgs &a, const PlanParams &pp, int grid, int smem_bytes, cudaStream_t stream) {
  fk_inertia_kernel<Core><<<grid, TFK_TILE, smem_bytes, stream>>>(a, pp);
}
template <class Core> typename std::enable_if<Core::D != 7, void>::type launch_inertia(const KernelArgs &, const PlanParams &, int, int, cudaStream_t) {}
template <class Core> typename std::enable_if<Core::D == 7, cudaError_t>::type occupancy_inertia(int smem_bytes, int *blocks_per_sm) {
  return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, fk_inertia_kernel<Core>, TFK_TILE, smem_bytes);
}
template <class Core> typename std::enable_if<Core::D != 7, cudaError_t>::type occupancy_inertia(int, int *blocks_per_sm) {
  *blocks_per_sm = 0;
  return cudaSuccess;
}
template <class Core> typename std::enable_if<Core::D == 3, void>::type launch_com(const KernelArgs &a, const PlanParams &pp, int grid, int smem_bytes, cudaStream_t stream) {
  fk_com_kernel<Core><<<grid, TFK_TILE, smem_bytes, stream>>>(a, pp);
}
template <class Core> typename std::enable_if<Core::D != 3, void>::type launch_com(const KernelArgs &, const PlanParams &, int, int, cudaStream_t) {}
template <class Core> typename std::enable_if<Core::D == 3, cudaError_t>::type occupancy_com(int smem_bytes, int *blocks_per_sm) {
  return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, fk_com_kernel<Core>, TFK_TILE, smem_bytes);
}
template <class Core> typename std::enable_if<Core::D != 3, cudaError_t>::type occupancy_com(int, int *blocks_per_sm) {
  *blocks_per_sm = 0;
  return cudaSuccess;
}

template <class Core>
cudaError_t launch_fk_core(const KernelArgs &a, const PlanParams &pp, int variant, int grid, int smem_bytes, cudaStream_t stream) {
  if (variant == KV_FUSED_SMEM5)
    fk_fused_kernel<Core, true, 5><<<grid, TFK_TILE, smem_bytes, stream>>>(a, pp);
  else if (variant == KV_FUSED_SMEM)
    fk_fused_kernel<Core, true, 4><<<grid, TFK_TILE, smem_bytes, stream>>>(a, pp);
  else if (variant == KV_FUSED_GMEM)
    fk_fused_kernel<Core, false, 4><<<grid, TFK_TILE, smem_bytes, stream>>>(a, pp);
  else if (variant == KV_TWOPHASE_SMEM)
    fk_twophase_kernel<Core, true><<<grid, TFK_TILE, smem_bytes, stream>>>(a, pp);
  else if (variant == KV_SQDIST || variant == KV_SQDIST_SMEM)
    launch_sqdist<Core>(a, pp, variant == KV_SQDIST_SMEM, grid, smem_bytes, stream);
  else if (variant == KV_COM)
    launch_com<Core>(a, pp, grid, smem_bytes, stream);
  else if (variant == KV_INERTIA)
    launch_inertia<Core>(a, pp, grid, smem_bytes, stream);
  else
    fk_twophase_kernel<Core, false><<<grid, TFK_TILE, smem_bytes, stream>>>(a, pp);
  return cudaGetLastError();
}

template <class Core> cudaError_t occupancy_fk_core(int variant, int smem_bytes, int *blocks_per_sm) {
  if (variant == KV_SQDIST || variant == KV_SQDIST_SMEM) return occupancy_sqdist<Core>(variant == KV_SQDIST_SMEM, smem_bytes, blocks_per_sm);
  if (variant == KV_COM) return occupancy_com<Core>(smem_bytes, blocks_per_sm);
  if (variant == KV_INERTIA) return occupancy_inertia<Core>(smem_bytes, blocks_per_sm);
  const void *fn = variant == KV_FUSED_SMEM5  ? (const void *)fk_fused_kernel<Core, true, 5>
                   : variant == KV_FUSED_SMEM ? (const void *)fk_fused_kernel<Core, true, 4>
                   : variant == KV_FUSED_GMEM ? (const void *)fk_fused_kernel<Core, false, 4>
                   : variant == KV_TWOPHASE_SMEM ? (const void *)fk_twophase_kernel<Core, true>
                                              : (const void *)fk_twophase_kernel<Core, false>;
  cudaError_t e = cudaSuccess;
  if (smem_bytes > 48 * 1024) e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
  if (e != cudaSuccess) return e;
  return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, fn, TFK_TILE, smem_bytes);
}

// one instantiation per RotationType (the pose dimension is a compile-time constant of the kernels)
template <class M, bool ANALYTIC_BASE>
cudaError_t launch_fk(const KernelArgs &a, const PlanParams &pp, int variant, int grid, int smem_bytes, cudaStream_t stream) {
  if (pp.h.rot_type == 0) return launch_fk_core<FkCore<M, ANALYTIC_BASE, 0, TFK_TILE>>(a, pp, variant, grid, smem_bytes, stream);
  if (pp.h.rot_type == 1) return launch_fk_core<FkCore<M, ANALYTIC_BASE, 1, TFK_TILE>>(a, pp, variant, grid, smem_bytes, stream);
  return launch_fk_core<FkCore<M, ANALYTIC_BASE, 2, TFK_TILE>>(a, pp, variant, grid, smem_bytes, stream);
}
template <class M, bool ANALYTIC_BASE> cudaError_t occupancy_fk(int rot_type, int variant, int smem_bytes, int *blocks_per_sm) {
  if (rot_type == 0) return occupancy_fk_core<FkCore<M, ANALYTIC_BASE, 0, TFK_TILE>>(variant, smem_bytes, blocks_per_sm);
  if (rot_type == 1) return occupancy_fk_core<FkCore<M, ANALYTIC_BASE, 1, TFK_TILE>>(variant, smem_bytes, blocks_per_sm);
  return occupancy_fk_core<FkCore<M, ANALYTIC_BASE, 2, TFK_TILE>>(variant, smem_bytes, blocks_per_sm);
}

#endif // !__CUDACC_RTC__

} // namespace tfk
#endif
