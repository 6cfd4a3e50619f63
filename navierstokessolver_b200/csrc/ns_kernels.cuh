// ns_kernels.cuh -- sm_100a kernels of the time-step loop (fp64, HBM-bound stencils; no tensor cores).
//
//   k_smooth_rb        one full red-black Gauss-Seidel sweep in ONE pass over HBM: shared-memory tile with a
//                      2-cell halo, red update on tile+1 ring, black update on the tile, out of place.
//                      Algorithmic traffic 24 B per cell-update (read x, read b, write x).
//   k_residual         r = b - A x and/or ||r||^2            (16-24 B/cell)
//   k_restrict         coarse b = average of the children's residuals (fused residual + restriction)
//   k_prolong_add      x_fine += x_coarse(parent)            (piecewise-constant prolongation)
//   k_apply            y = A x (+ optional dots)
//   k_rhs_build        ConstructRHS_V                        (64 B/cell)
//   k_div_rhs          ConstructRHS_phi + block sums for MatNullSpaceRemove   (24 B/cell)
//   k_correct          CorrectVelocities + GradP + VecMin/VecMax              (40+16 B/cell)
// Reductions are two-stage and deterministic: warp shuffles -> one partial per block -> the last block to
// finish adds the partials in a fixed order.
#pragma once
#include <cuda_runtime.h>
#include "ns_cell.cuh"

namespace nsgpu {

constexpr int RED_MAX_BLOCKS = 1 << 16;

struct Reducer {
  double* partials;   // [NV * RED_MAX_BLOCKS]
  unsigned* counter;  // zero between launches
  double* tmp;        // 8 doubles: where a reduction lands when its epilogue distributes the values itself
};

// reduction operators; the index k is the position in the reduced vector (OpMax4Sum: four maxima and one sum in one reduction)
struct OpSum { __device__ static double f(int, double a, double b) { return a + b; } __device__ static double id(int = 0) { return 0.0; } };
struct OpMax { __device__ static double f(int, double a, double b) { return fmax(a, b); } __device__ static double id(int = 0) { return -1.79769313486231570e308; } };
struct OpMax4Sum {
  __device__ static double f(int k, double a, double b) { return k < 4 ? fmax(a, b) : a + b; }
  __device__ static double id(int k = 0) { return k < 4 ? -1.79769313486231570e308 : 0.0; }
};

// blockIdx.y owns a contiguous run of columns, so the x-neighbour columns a block re-reads are still in its L1
__device__ __forceinline__ int chunk_lo(int nx) { return (int)(((long long)nx * blockIdx.y) / gridDim.y); }
__device__ __forceinline__ int chunk_hi(int nx) { return (int)(((long long)nx * (blockIdx.y + 1)) / gridDim.y); }

template <class OP>
__device__ __forceinline__ double warp_red(int k, double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = OP::f(k, v, __shfl_down_sync(0xffffffffu, v, o));
  return v;
}

// NSGPU_XRED_DBG=1: clock stamps of the reduction tails (cycles, summed; printed by nsgpu_destroy): [0] calls, [1] last block's sum of the
// partials, [2] whole epilogue, [3] xreduce value stores, [4] system fence, [5] flag stores, [6] waiting for the other ranks, [7] xreduce calls
// Compiled in only with -DNSGPU_STAMPS=1 (NSGPU_BUILD_STAMPS=1 python -c "import __graft_entry__ as g; g.build()"): even switched off at run time the
// stamps change the register allocation of the fused passes (4096^2: last pass of a cycle 96 -> 117 us), so the product is built without them.
#ifndef NSGPU_STAMPS
#define NSGPU_STAMPS 0
#endif
constexpr bool NS_STAMPS = NSGPU_STAMPS != 0;
static __device__ int g_ns_xdbg = 0;
static __device__ long long g_ns_xdbg_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
constexpr unsigned NS_DBG_CTAS = 2048;
static __device__ unsigned long long g_ns_cta_t[4 * NS_DBG_CTAS];
static __device__ unsigned g_ns_cta_grid[2];
__device__ __forceinline__ unsigned long long globaltimer_ns() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }

// Block-level then grid-level reduction of NV values held by every thread.  All threads of the block must call.
struct NoEpilogue { __device__ void operator()(const double*) const {} };

// `epi(out)` runs once, in the thread that wrote the results, after they are in place.
template <int NV, class OP, class EPI = NoEpilogue>
__device__ void grid_reduce(double (&v)[NV], Reducer red, double* out, EPI epi = EPI()) {
  __shared__ double sm[NV][32];
  __shared__ bool is_last;
  const int tid = threadIdx.x + blockDim.x * (threadIdx.y + blockDim.y * threadIdx.z);
  const int nth = blockDim.x * blockDim.y * blockDim.z;
  const int lane = tid & 31, wid = tid >> 5, nw = (nth + 31) >> 5;
#pragma unroll
  for (int k = 0; k < NV; k++) {
    double w = warp_red<OP>(k, v[k]);
    if (lane == 0) sm[k][wid] = w;
  }
  __syncthreads();
  const unsigned bid = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
  const unsigned nb = gridDim.x * gridDim.y * gridDim.z;
  if (wid == 0) {
#pragma unroll
    for (int k = 0; k < NV; k++) {
      double w = (lane < nw) ? sm[k][lane] : OP::id(k);
      w = warp_red<OP>(k, w);
      if (lane == 0) red.partials[(size_t)k * RED_MAX_BLOCKS + bid] = w;
    }
    if (lane == 0) {
      __threadfence();
      unsigned t = atomicAdd(red.counter, 1u);
      is_last = (t == nb - 1);
    }
  }
  __syncthreads();
  if (!is_last) return;
  const long long dbg_t0 = (NS_STAMPS && g_ns_xdbg) ? clock64() : 0;
  __threadfence();
#pragma unroll
  for (int k = 0; k < NV; k++) {
    double w = OP::id(k);
    for (unsigned b = tid; b < nb; b += nth) w = OP::f(k, w, __ldcg(&red.partials[(size_t)k * RED_MAX_BLOCKS + b]));
    w = warp_red<OP>(k, w);
    __syncthreads();
    if (lane == 0) sm[k][wid] = w;
    __syncthreads();
    if (wid == 0) {
      double z = (lane < nw) ? sm[k][lane] : OP::id(k);
      z = warp_red<OP>(k, z);
      if (lane == 0) out[k] = z;
    }
  }
  if (tid == 0) {
    *red.counter = 0u;
    if (NS_STAMPS && g_ns_xdbg) {
      const long long t1 = clock64();
      epi(out);
      const long long t2 = clock64();
      g_ns_xdbg_acc[0] += 1; g_ns_xdbg_acc[1] += t1 - dbg_t0; g_ns_xdbg_acc[2] += t2 - t1;
    } else epi(out);
  }
}

// Convergence control block of one linear solve, resident on the device (one per solve slot).
struct SolveCtl {
  double bn2;        // |b|^2
  double rn2;        // |b - A x|^2 after the last cycle
  double tol2;       // rtol^2
  int cycles, max_cycles;
  int go;            // loop condition computed by the last residual reduction
  int zero_rhs;      // b == 0: the solution is x = 0 (k_zero_if)
  int total_cycles;  // cumulative over the life of the context (never reset on the device)
  int fails;         // cumulative number of solves that stopped at max_cycles above the tolerance
  unsigned long long handle;   // cudaGraphConditionalHandle of the WHILE node (0: none)
  double xsum;       // phi slot: sum of phi over all cells, reduced by the projection kernel that reads phi anyway (lazy mean removal)
};
static_assert(sizeof(SolveCtl) == 64, "SolveCtl is one 64-byte block");

// ---------------------------------------------------------------------------------------------------------
// Several ranks (one process per GPU): every rank's arena is mapped into every other rank's address space (CUDA IPC).
// The first words of an arena are its control header; off[r] is the distance, in 8-byte words, from this rank's
// arena to rank r's, so `local_pointer + off[r]` is the same object on rank r.
constexpr int MAX_RANKS = 8;
enum : int { HDR_ARR_L = 0, HDR_ARR_R = 1, HDR_TAIL_ARR = 2, HDR_SENT_L = 3, HDR_SENT_R = 4, HDR_TICKET = 5, HDR_TAIL_NEED = 6,
             HDR_XSEQ = 7, /* 8: HDR_XVSEQ */ HDR_HX_ARR_L = 9, HDR_HX_ARR_R = 10, HDR_HX_NEED = 11, HDR_HX_TICKET = 12,
             HDR_XSLOTS = 16 /* 4 buffers x MAX_RANKS x 3 words */, HDR_WORDS = 512 };
struct PeerTable {
  int world, rank;
  long long off[MAX_RANKS];
  unsigned long long* hdr;   // this rank's header
  int* err;                  // set when a wait times out (instead of hanging the GPU)
  int nowait;                // TIMING EXPERIMENT ONLY (NSGPU_XRED_NOWAIT=1): cross-rank reductions return world x the local value without waiting
};

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
// TIMING EXPERIMENTS ONLY (NSGPU_NOWAIT=1, set at nsgpu_create): no device-side wait waits -- results are wrong, kernel times are the pure execution times
static __device__ int g_ns_nowait = 0;
// spin until *p >= need (exact == false) or *p == need (exact == true); gives up after ~3 s and raises the error flag
__device__ __forceinline__ void wait_word(const unsigned long long* p, unsigned long long need, bool exact, int* err) {
  if (g_ns_nowait) return;
  const long long t0 = clock64();
  for (;;) {
    const unsigned long long v = ld_acquire_sys(p);
    if (exact ? v == need : v >= need) break;
    if (clock64() - t0 > 6000000000LL) { *err = 1; break; }
  }
}

// Sum of two doubles over all ranks, called by ONE thread per rank; every rank adds the contributions in rank order,
// so all ranks get the same bits.  Four slot buffers: a rank can be at most one reduction ahead of the slowest one.
// Slot of (buffer, rank): three words {value a, value b, sequence}.
__device__ inline void xreduce_sum2(const PeerTable* pt, double a, double b, double* ra, double* rb) {
  if (pt->nowait) { *ra = a * pt->world; *rb = b * pt->world; return; }
  const bool dbg = NS_STAMPS && g_ns_xdbg != 0;
  const long long t0 = dbg ? clock64() : 0;
  unsigned long long* hdr = pt->hdr;
  const unsigned long long s = hdr[HDR_XSEQ] + 1;
  hdr[HDR_XSEQ] = s;
  const int buf = (int)(s & 3);
  unsigned long long* my_slot = hdr + HDR_XSLOTS + 3 * (buf * MAX_RANKS + pt->rank);
  for (int r = 0; r < pt->world; r++) {
    reinterpret_cast<double*>(my_slot + pt->off[r])[0] = a;
    reinterpret_cast<double*>(my_slot + pt->off[r])[1] = b;
  }
  const long long t1 = dbg ? clock64() : 0;
  __threadfence_system();
  const long long t2 = dbg ? clock64() : 0;
  for (int r = 0; r < pt->world; r++) {
    unsigned long long* q = my_slot + pt->off[r] + 2;
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" :: "l"(q), "l"(s) : "memory");   // ordered after the values by the fence
  }
  const long long t3 = dbg ? clock64() : 0;
  double sa = 0.0, sb = 0.0;
  for (int r = 0; r < pt->world; r++) {
    const unsigned long long* q = hdr + HDR_XSLOTS + 3 * (buf * MAX_RANKS + r);
    wait_word(q + 2, s, true, pt->err);
    sa += reinterpret_cast<const volatile double*>(q)[0];
    sb += reinterpret_cast<const volatile double*>(q)[1];
  }
  if (dbg) {
    const long long t4 = clock64();
    g_ns_xdbg_acc[3] += t1 - t0; g_ns_xdbg_acc[4] += t2 - t1; g_ns_xdbg_acc[5] += t3 - t2; g_ns_xdbg_acc[6] += t4 - t3; g_ns_xdbg_acc[7] += 1;
  }
  *ra = sa; *rb = sb;
}
// Exchange of a small vector (N <= XV_MAX doubles) between all ranks, called by ONE thread per rank: all[r * N + k] = value k of rank r,
// the same bits on every rank; the caller combines them in rank order with whatever operator each entry needs.  Own slot region
// behind the sum slots: four buffers x MAX_RANKS x (XV_MAX + 1) words.
constexpr int XV_MAX = 6;
constexpr int HDR_XVSLOTS = HDR_XSLOTS + 4 * MAX_RANKS * 3;   // = 112
constexpr int HDR_XVSEQ = 8;
static_assert(HDR_XVSLOTS + 4 * MAX_RANKS * (XV_MAX + 1) <= HDR_WORDS, "control header too small");
template <int N>
__device__ inline void xexchange(const PeerTable* pt, const double (&mine)[N], double* all) {
  static_assert(N <= XV_MAX, "vector too long");
  unsigned long long* hdr = pt->hdr;
  const unsigned long long s = hdr[HDR_XVSEQ] + 1;
  hdr[HDR_XVSEQ] = s;
  const int buf = (int)(s & 3);
  unsigned long long* my_slot = hdr + HDR_XVSLOTS + (XV_MAX + 1) * (buf * MAX_RANKS + pt->rank);
  for (int r = 0; r < pt->world; r++)
#pragma unroll
    for (int k = 0; k < N; k++) reinterpret_cast<double*>(my_slot + pt->off[r])[k] = mine[k];
  __threadfence_system();
  for (int r = 0; r < pt->world; r++) {
    unsigned long long* q = my_slot + pt->off[r] + XV_MAX;
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" :: "l"(q), "l"(s) : "memory");
  }
  for (int r = 0; r < pt->world; r++) {
    const unsigned long long* q = hdr + HDR_XVSLOTS + (XV_MAX + 1) * (buf * MAX_RANKS + r);
    wait_word(q + XV_MAX, s, true, pt->err);
#pragma unroll
    for (int k = 0; k < N; k++) all[r * N + k] = reinterpret_cast<const volatile double*>(q)[k];
  }
}

__device__ inline double xreduce_sum(const PeerTable* pt, double mine) {
  double a, b;
  xreduce_sum2(pt, mine, 0.0, &a, &b);
  return a;
}

// Several ranks: the slabs of a replicated right-hand side are stored by every rank's restriction pass (peer memory);
// this one-thread kernel holds the stream until all of them have arrived.
__global__ void k_wait_gather(const PeerTable* pt, unsigned gather_ctas) {
  const unsigned long long need = pt->hdr[HDR_TAIL_NEED] + (unsigned long long)pt->world * gather_ctas;
  wait_word(pt->hdr + HDR_TAIL_ARR, need, false, pt->err);
  pt->hdr[HDR_TAIL_NEED] = need;
}

struct SumEpilogue {   // out[0] <- sum over ranks
  const PeerTable* pt;
  double* out_w;
  __device__ void operator()(const double* out) const { if (pt) out_w[0] = xreduce_sum(pt, out[0]); }
};

// out[0..N) <- sums over the ranks (rank order: the same bits everywhere); no-op on one rank
template <int N>
struct XSumEpilogue {
  const PeerTable* pt;
  __device__ void operator()(const double* out_c) const {
    if (!pt) return;
    double* out = const_cast<double*>(out_c);
    double mine[N], all[MAX_RANKS * N];
#pragma unroll
    for (int k = 0; k < N; k++) mine[k] = out[k];
    xexchange<N>(pt, mine, all);
#pragma unroll
    for (int k = 0; k < N; k++) {
      double t = 0.0;
      for (int r = 0; r < pt->world; r++) t += all[r * N + k];
      out[k] = t;
    }
  }
};

// Halo exchange of up to HXP_MAX fields outside the solves, without a library call: every rank stores the `width` first / last columns of
// each field straight into its neighbours' halo columns (peer memory over NVLink), the last block to finish tells both neighbours
// (system-scope fence, one arrival counter per side) and then waits for THEIR arrival -- all ranks run the same sequence of exchanges,
// so "the k-th exchange has arrived" is a count, kept on the device (graph-safe).  When the kernel ends, this rank's halo columns are in
// place.  Why a neighbour's early push cannot overwrite halo columns still being read: ns_solver.inl, halo_exchange_push.
constexpr int HXP_MAX = 4;
struct HaloPushArgs {
  double* f[HXP_MAX];
  int width[HXP_MAX];
  int nf, ld, nx;
  long long offL, offR;        // arena offsets (doubles) of the neighbours; 0 = none
  unsigned long long* hdr;
  int* err;
};
__global__ void __launch_bounds__(256) k_halo_push(HaloPushArgs a) {
  const int k = blockIdx.y;
  if (k < a.nf) {
    const size_t n2 = (size_t)a.width[k] * a.ld / 2;                 // double2 per side (ld is a multiple of 4)
    const double2* sL = reinterpret_cast<const double2*>(a.f[k] + (size_t)HX * a.ld);
    const double2* sR = reinterpret_cast<const double2*>(a.f[k] + (size_t)(HX + a.nx - a.width[k]) * a.ld);
    // my column c (0 .. w) is the left neighbour's halo column nx + c; my column nx - w + c the right neighbour's column -w + c
    double2* dL = reinterpret_cast<double2*>(a.f[k] + (size_t)HX * a.ld + a.offL + (long long)a.nx * a.ld);
    double2* dR = reinterpret_cast<double2*>(a.f[k] + (size_t)(HX + a.nx - a.width[k]) * a.ld + a.offR - (long long)a.nx * a.ld);
    for (size_t q = blockIdx.x * (size_t)blockDim.x + threadIdx.x; q < n2; q += (size_t)gridDim.x * blockDim.x) {
      if (a.offL) dL[q] = sL[q];
      if (a.offR) dR[q] = sR[q];
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();                                           // this block's peer stores are visible before its ticket
    const unsigned total = gridDim.x * gridDim.y;
    if (atomicAdd(a.hdr + HDR_HX_TICKET, 1ull) == total - 1) {
      a.hdr[HDR_HX_TICKET] = 0;
      __threadfence_system();
      if (a.offL) atomicAdd_system(a.hdr + a.offL + HDR_HX_ARR_R, 1ull);
      if (a.offR) atomicAdd_system(a.hdr + a.offR + HDR_HX_ARR_L, 1ull);
      const unsigned long long need = a.hdr[HDR_HX_NEED] + 1;
      a.hdr[HDR_HX_NEED] = need;
      if (a.offL) wait_word(a.hdr + HDR_HX_ARR_L, need, false, a.err);
      if (a.offR) wait_word(a.hdr + HDR_HX_ARR_R, need, false, a.err);
    }
  }
}

// projection kernel: out[0..3] = the four maxima, out[4] = sum(phi) of this rank; over all ranks when pt != nullptr
struct CorrectEpilogue {
  const PeerTable* pt;
  double* xsum;    // SolveCtl::xsum of the phi slot (may be nullptr)
  __device__ void operator()(const double* out_c) const {
    double* out = const_cast<double*>(out_c);
    double s = out[4];
    if (pt) {
      double all[MAX_RANKS * 5];
      const double mine[5] = {out[0], out[1], out[2], out[3], out[4]};
      xexchange<5>(pt, mine, all);
      s = 0.0;
      for (int r = 0; r < pt->world; r++) {
        for (int k = 0; k < 4; k++) out[k] = r == 0 ? all[k] : fmax(out[k], all[r * 5 + k]);
        s += all[r * 5 + 4];
      }
    }
    if (xsum) *xsum = s;
  }
};

struct CtlEpilogue {
  SolveCtl* c;
  const PeerTable* pt;
  double* out_w;
  __device__ void operator()(const double* out) const {
    if (!c && !pt) return;
    double rn2 = out[0];
    if (pt) { rn2 = xreduce_sum(pt, rn2); if (out_w) out_w[0] = rn2; }
    if (!c) return;
    c->rn2 = rn2;
    c->cycles += 1;
    if (c->cycles > 0) c->total_cycles += 1;
    const bool zero = !(c->bn2 > 0.0);
    c->zero_rhs = zero ? 1 : 0;
    if (zero) c->cycles = 0;   // x = 0 is the answer, whatever ran before
    const bool go = !zero && (rn2 > c->tol2 * c->bn2) && (c->cycles < c->max_cycles) && (rn2 == rn2);
    c->go = go ? 1 : 0;
    if (!go && !zero && !(rn2 <= c->tol2 * c->bn2)) c->fails += 1;
    if (c->handle) cudaGraphSetConditional((cudaGraphConditionalHandle)c->handle, go ? 1u : 0u);
  }
};

// ctl <- start of a solve: cycles0 = -1 when the first residual reduction is the initial guess' (it brings it to 0),
// 0 when it already closes the first iteration
// skipped: iterations the head runs without a stopping test before the first tested one (they count as cycles)
__global__ void k_ctl_init(SolveCtl* c, double tol2, int max_cycles, unsigned long long handle, int cycles0, const double* bn2_src, int skipped) {
  if (bn2_src) c->bn2 = *bn2_src;   // |b|^2 came out of the kernel that wrote b
  c->tol2 = tol2; c->max_cycles = max_cycles; c->cycles = cycles0; c->go = 0; c->zero_rhs = 0; c->handle = handle; c->rn2 = 0.0;
  c->total_cycles += skipped;
}
// x <- 0 when the right-hand side was identically zero (KSPSolve returns x = 0 for b = 0)
// Also publishes the control block to the host: `pub` is page-locked host memory mapped into the device's address space, written by
// one thread -- no copy-engine transfer, so the readback never queues behind the field copies of nsgpu_step_host / nsgpu_export_begin.
__global__ void __launch_bounds__(256) k_zero_if(size_t n, double* __restrict__ x, const SolveCtl* __restrict__ c, SolveCtl* pub) {
  if (pub && blockIdx.x == 0 && threadIdx.x == 0) *pub = *c;
  if (!c->zero_rhs) return;
  for (size_t k = blockIdx.x * (size_t)blockDim.x + threadIdx.x; k < n; k += (size_t)gridDim.x * blockDim.x) x[k] = 0.0;
}

// ---------------------------------------------------------------------------------------------------------
// Fused red-black Gauss-Seidel sweep.  Colour of a cell = (global i + j) & 1, red = 0 first.
// The outflow rows of L_phi use the previous sweep's value for their second inward cell (same colour).
constexpr int SM_TI = 16, SM_TJ = 128, SM_THREADS = 256;

template <int OP, bool RECT, bool XZERO>
__global__ void __launch_bounds__(SM_THREADS)
k_smooth_rb(LevelGeom g, const double* __restrict__ xin, const double* __restrict__ b, double* __restrict__ xout,
            double alpha, double beta) {
  constexpr int SI = SM_TI + 4, SJ = SM_TJ + 4;
  __shared__ double s[SI][SJ];
  const int tid = threadIdx.x;
  const int i0 = blockIdx.y * SM_TI, j0 = blockIdx.x * SM_TJ;

  // phase 1: tile + 2-cell halo -> shared memory (double2, 16-byte aligned since j0 - 2 + PJ is even)
  if (XZERO) {
    for (int q = tid; q < SI * SJ; q += SM_THREADS) (&s[0][0])[q] = 0.0;
  } else {
    for (int q = tid; q < SI * (SJ / 2); q += SM_THREADS) {
      const int a = q / (SJ / 2), c = 2 * (q - a * (SJ / 2));
      const int i = i0 - 2 + a, j = j0 - 2 + c;
      double2 val = make_double2(0.0, 0.0);
      if (i < g.nx + HX && j < g.ny + PJ) val = *reinterpret_cast<const double2*>(xin + cidx(g, i, j));
      *reinterpret_cast<double2*>(&s[a][c]) = val;
    }
  }
  __syncthreads();

  // phase 2: red cells of the tile and of the one-cell ring around it
  {
    constexpr int PER = SM_TJ / 2 + 1;
    for (int q = tid; q < (SM_TI + 2) * PER; q += SM_THREADS) {
      const int a = 1 + q / PER, t = q % PER;
      const int i = i0 - 2 + a;
      const int par = (g.gi0 + i + j0) & 1;          // parity of the cell at c == 2
      const int c = (par ? 1 : 2) + 2 * t;           // red: (gi + j) even
      const int j = j0 - 2 + c;
      if (i >= g.nx + 1 || j >= g.ny + 1) continue;
      const unsigned f = cell_flags<RECT>(g, i, j);
      if (f == F_OUTSIDE) continue;
      const size_t id = cidx(g, i, j);
      s[a][c] = relax_cell<OP>(g, f, i, j, s[a - 1][c], s[a + 1][c], s[a][c - 1], s[a][c + 1], xin + id, b[id], alpha, beta);
    }
  }
  __syncthreads();

  // phase 3: black cells of the tile
  {
    constexpr int PER = SM_TJ / 2;
    for (int q = tid; q < SM_TI * PER; q += SM_THREADS) {
      const int a = 2 + q / PER, t = q % PER;
      const int i = i0 - 2 + a;
      const int par = (g.gi0 + i + j0) & 1;
      const int c = (par ? 2 : 3) + 2 * t;           // black: (gi + j) odd
      const int j = j0 - 2 + c;
      if (i >= g.nx || j >= g.ny) continue;
      const unsigned f = cell_flags<RECT>(g, i, j);
      if (f == F_OUTSIDE) continue;
      const size_t id = cidx(g, i, j);
      s[a][c] = relax_cell<OP>(g, f, i, j, s[a - 1][c], s[a + 1][c], s[a][c - 1], s[a][c + 1], xin + id, b[id], alpha, beta);
    }
  }
  __syncthreads();

  // phase 4: tile -> xout
  for (int q = tid; q < SM_TI * (SM_TJ / 2); q += SM_THREADS) {
    const int a = 2 + q / (SM_TJ / 2), c = 2 + 2 * (q % (SM_TJ / 2));
    const int i = i0 - 2 + a, j = j0 - 2 + c;
    if (i >= g.nx || j >= g.ny) continue;
    if (j + 1 < g.ny) *reinterpret_cast<double2*>(xout + cidx(g, i, j)) = *reinterpret_cast<const double2*>(&s[a][c]);
    else xout[cidx(g, i, j)] = s[a][c];
  }
}

// ---------------------------------------------------------------------------------------------------------
// r = b - A x  (WRITE) and / or sum r^2 (NORM)
template <int OP, bool RECT, bool WRITE, bool NORM>
__global__ void __launch_bounds__(256)
k_residual(LevelGeom g, const double* __restrict__ x, const double* __restrict__ b, double* __restrict__ r,
           double alpha, double beta, Reducer red, double* out, SolveCtl* ctl, const PeerTable* pt) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  double acc[1] = {0.0};
  for (int i = chunk_lo(g.nx), i_hi = chunk_hi(g.nx); i < i_hi; i++) {
    if (j < g.ny) {
      const unsigned f = cell_flags<RECT>(g, i, j);
      if (f != F_OUTSIDE) {
        const size_t id = cidx(g, i, j);
        const double rv = b[id] - apply_cell<OP>(g, f, i, j, x + id, alpha, beta);
        if (WRITE) r[id] = rv;
        if (NORM) acc[0] += rv * rv;
      }
    }
  }
  if (NORM) grid_reduce<1, OpSum>(acc, red, out, CtlEpilogue{ctl, pt, out});
}

// y = A x, optionally with the dots (w, y) and (y, y)
template <int OP, bool RECT, bool DOTS>
__global__ void __launch_bounds__(256)
k_apply(LevelGeom g, const double* __restrict__ x, double* __restrict__ y, const double* __restrict__ w,
        double alpha, double beta, Reducer red, double* out) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  double acc[2] = {0.0, 0.0};
  for (int i = chunk_lo(g.nx), i_hi = chunk_hi(g.nx); i < i_hi; i++) {
    if (j < g.ny) {
      const unsigned f = cell_flags<RECT>(g, i, j);
      if (f != F_OUTSIDE) {
        const size_t id = cidx(g, i, j);
        const double yv = apply_cell<OP>(g, f, i, j, x + id, alpha, beta);
        y[id] = yv;
        if (DOTS) { acc[0] += w[id] * yv; acc[1] += yv * yv; }
      }
    }
  }
  if (DOTS) grid_reduce<2, OpSum>(acc, red, out);
}

// coarse b = (1/(cx*cy)) * sum over children of (b - A x); one thread per coarse cell
template <int OP, bool RECT>
__global__ void __launch_bounds__(256)
k_restrict(LevelGeom gf, LevelGeom gc, const double* __restrict__ x, const double* __restrict__ b,
           double* __restrict__ bc, double alpha, double beta, int cx, int cy) {
  const int J = blockIdx.x * blockDim.x + threadIdx.x;
  const int I = blockIdx.y;
  if (J >= gc.ny || I >= gc.nx) return;
  double acc = 0.0;
  for (int di = 0; di < cx; di++)
    for (int dj = 0; dj < cy; dj++) {
      const int i = I * cx + di, j = J * cy + dj;
      const unsigned f = cell_flags<RECT>(gf, i, j);
      if (f == F_OUTSIDE) continue;
      const size_t id = cidx(gf, i, j);
      acc += b[id] - apply_cell<OP>(gf, f, i, j, x + id, alpha, beta);
    }
  const unsigned fc = cell_flags<RECT>(gc, I, J);
  bc[cidx(gc, I, J)] = (fc == F_OUTSIDE) ? 0.0 : acc / (double)(cx * cy);
}

// x_fine += x_coarse(parent); one thread per fine cell
template <bool RECT>
__global__ void __launch_bounds__(256)
k_prolong_add(LevelGeom gf, LevelGeom gc, double* __restrict__ xf, const double* __restrict__ xc, int cx, int cy) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= gf.ny) return;
  for (int i = chunk_lo(gf.nx), i_hi = chunk_hi(gf.nx); i < i_hi; i++) {
    if (cell_flags<RECT>(gf, i, j) == F_OUTSIDE) continue;
    xf[cidx(gf, i, j)] += xc[cidx(gc, i / cx, j / cy)];
  }
}

// ---------------------------------------------------------------------------------------------------------
// Time-step physics

template <bool RECT>
__global__ void __launch_bounds__(128)
k_rhs_build(LevelGeom g, const EdgeDev* __restrict__ edges, StepConst sc, const double* __restrict__ u,
            const double* __restrict__ v, double* __restrict__ nu0, double* __restrict__ nv0,
            const double* __restrict__ gpx, const double* __restrict__ gpy, double* __restrict__ rhs_u,
            double* __restrict__ rhs_v) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  const int i = blockIdx.y;
  if (j >= g.ny || i >= g.nx) return;
  if (cell_flags<RECT>(g, i, j) == F_OUTSIDE) return;
  const size_t id = cidx(g, i, j);
  RhsOut o = rhs_cell<RECT>(g, edges, sc, i, j, u + id, v + id, nu0[id], nv0[id], gpx + id, gpy + id);
  rhs_u[id] = o.rhs_u;
  rhs_v[id] = o.rhs_v;
  nu0[id] = o.nu;
  nv0[id] = o.nv;
}

template <bool RECT>
__global__ void __launch_bounds__(256)
k_div_rhs(LevelGeom g, const EdgeDev* __restrict__ edges, StepConst sc, const double* __restrict__ u,
          const double* __restrict__ v, double* __restrict__ b, Reducer red, double* out) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  double acc[2] = {0.0, 0.0};   // sum and sum of squares: the mean to remove and |b - mean|^2 = S2 - S1^2 / n
  if (RECT) {
    // Rectangles: interior cells (all four neighbours present) take a straight-line path with the row's geometry in
    // registers and running pointers; the expressions are div_cell's.  Boundary cells go through div_cell.
    const bool row = j < g.ny, fS = j > 0, fN = j < g.ny - 1;
    const int ja = min(j, g.ny - 1) + PJ;
    const double hyj = g.hy[ja], rS = hyj * g.rsy[ja - 1], rN = hyj * g.rsy[ja], rhy = g.rhy[ja];
    const int i_lo = chunk_lo(g.nx), i_hi = chunk_hi(g.nx);
    size_t id = cidx(g, i_lo, min(j, g.ny - 1));
    for (int i = i_lo; i < i_hi; i++, id += g.ld) {
      if (!row) continue;
      const int ig = g.gi0 + i;
      double d;
      if (ig > 0 && ig < g.gnx - 1 && fS && fN) {
        const double hxi = g.hx[i + HX], rW = hxi * g.rsx[i - 1 + HX], rE = hxi * g.rsx[i + HX];
        const double uc = u[id], vc = v[id];
        const double VW = u[id - g.ld] * rW + uc * (1.0 - rW), VE = u[id + g.ld] * rE + uc * (1.0 - rE);
        const double VS = v[id - 1] * rS + vc * (1.0 - rS), VN = v[id + 1] * rN + vc * (1.0 - rN);
        d = ((VE - VW) * g.rhx[i + HX] + (VN - VS) * rhy) / sc.dt;
      } else {
        d = div_cell<RECT>(g, edges, sc, i, j, u + id, v + id);
      }
      b[id] = d;
      acc[0] += d;
      acc[1] = fma(d, d, acc[1]);
    }
  } else {
    for (int i = chunk_lo(g.nx), i_hi = chunk_hi(g.nx); i < i_hi; i++) {
      if (j < g.ny && cell_flags<RECT>(g, i, j) != F_OUTSIDE) {
        const size_t id = cidx(g, i, j);
        const double d = div_cell<RECT>(g, edges, sc, i, j, u + id, v + id);
        b[id] = d;
        acc[0] += d;
        acc[1] = fma(d, d, acc[1]);
      }
    }
  }
  grid_reduce<2, OpSum>(acc, red, out);
}

// b -= *sum / n over fluid cells (MatNullSpaceRemove: plain arithmetic mean over ids, FS.cpp:550)
template <bool RECT>
__global__ void __launch_bounds__(256)
k_sub_mean(LevelGeom g, double* __restrict__ b, const double* __restrict__ sum, double inv_n) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= g.ny) return;
  const double m = *sum * inv_n;
  if constexpr (RECT) {   // four columns in flight per thread: one load per iteration leaves the pass latency-bound (3.7 TB/s)
    const int i_lo = chunk_lo(g.nx), i_hi = chunk_hi(g.nx);
    const size_t ld = g.ld;
    double* p = b + cidx(g, i_lo, j);
    int i = i_lo;
    for (; i + 4 <= i_hi; i += 4, p += 4 * ld) {
      const double a0 = p[0], a1 = p[ld], a2 = p[2 * ld], a3 = p[3 * ld];
      p[0] = a0 - m; p[ld] = a1 - m; p[2 * ld] = a2 - m; p[3 * ld] = a3 - m;
    }
    for (; i < i_hi; i++, p += ld) p[0] -= m;
  } else {
    for (int i = chunk_lo(g.nx), i_hi = chunk_hi(g.nx); i < i_hi; i++)
      if (cell_flags<RECT>(g, i, j) != F_OUTSIDE) b[cidx(g, i, j)] -= m;
  }
}

// out[0..3] = max(-u), max(u), max(-v), max(v)
template <bool RECT>
__global__ void __launch_bounds__(256)
k_correct(LevelGeom g, StepConst sc, const double* __restrict__ phi, const double* __restrict__ us,
          const double* __restrict__ vs, double* __restrict__ u, double* __restrict__ v, double* __restrict__ gpx,
          double* __restrict__ gpy, Reducer red, double* out) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  double acc[4] = {OpMax::id(), OpMax::id(), OpMax::id(), OpMax::id()};
  // (unrolling this loop by 2 or 4 was measured: 48 / 54 registers, fewer resident warps, 0.256 -> 0.29 ms)
#pragma unroll 1
  for (int i = chunk_lo(g.nx), i_hi = chunk_hi(g.nx); i < i_hi; i++) {
    if (j < g.ny) {
      const unsigned f = cell_flags<RECT>(g, i, j);
      if (f != F_OUTSIDE) {
        const size_t id = cidx(g, i, j);
        const UV gr = grad_cell(g, f, i, j, phi + id);   // (an interior fast path was measured: no gain, the pass is latency-bound)
        gpx[id] = gr.u;
        gpy[id] = gr.v;
        const double un = us[id] - sc.dt * gr.u, vn = vs[id] - sc.dt * gr.v;
        u[id] = un;
        v[id] = vn;
        // NaN-propagating: fmax drops NaNs, so map a non-finite value to +inf on all four
        const bool bad = !(isfinite(un) && isfinite(vn));
        const double inf = 1.0 / 0.0;
        acc[0] = fmax(acc[0], bad ? inf : -un);
        acc[1] = fmax(acc[1], bad ? inf : un);
        acc[2] = fmax(acc[2], bad ? inf : -vn);
        acc[3] = fmax(acc[3], bad ? inf : vn);
      }
    }
  }
  grid_reduce<4, OpMax>(acc, red, out);
}

// out[0] = sum over the fluid cells of hx_i hy_j b_ij: the product of b with the left null vector of LHS_phi on domains without an
// outflow edge (ns_hostgeom.h, plan_is_uniform) -- the inconsistency the reference's arithmetic mean removal leaves on stretched grids
template <bool RECT>
__global__ void __launch_bounds__(256)
k_vol_dot(LevelGeom g, const double* __restrict__ b, Reducer red, double* out) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  double acc[1] = {0.0};
  if (j < g.ny) {
    const double hyj = g.hy[j + PJ];
    for (int i = chunk_lo(g.nx), i_hi = chunk_hi(g.nx); i < i_hi; i++)
      if (RECT || cell_flags<RECT>(g, i, j) != F_OUTSIDE) acc[0] = fma(g.hx[i + HX] * hyj, b[cidx(g, i, j)], acc[0]);
  }
  grid_reduce<1, OpSum>(acc, red, out);
}

// (gpx, gpy) = GradP(phi) alone (FS.cpp:420-456): divPhi is a pure function of phi, so a caller that keeps its state on the host
// need not ship it (nsgpu_step_host with gradphi_x = gradphi_y = NULL)
template <bool RECT>
__global__ void __launch_bounds__(256)
k_gradphi(LevelGeom g, const double* __restrict__ phi, double* __restrict__ gpx, double* __restrict__ gpy) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= g.ny) return;
  for (int i = chunk_lo(g.nx), i_hi = chunk_hi(g.nx); i < i_hi; i++) {
    const unsigned f = cell_flags<RECT>(g, i, j);
    if (f == F_OUTSIDE) continue;
    const size_t id = cidx(g, i, j);
    const UV gr = grad_cell(g, f, i, j, phi + id);
    gpx[id] = gr.u;
    gpy[id] = gr.v;
  }
}

// Snapshot for ExportData (FS.cpp:569-615): copies of u, v, phi plus lapPhi = LHS_phi phi (the reference's MatMult, FS.cpp:577) and the
// physical pressure Pr = phi - dt/(2Re) lapPhi (FS.cpp:579-581) in one pass, into five staging arrays (streaming stores: nothing reads
// them but the copy engine), so the time loop can go on while the copies travel home.
template <bool RECT>
__global__ void __launch_bounds__(256)
k_export_snapshot(LevelGeom g, double half, const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ phi,
                  double* __restrict__ su, double* __restrict__ sv, double* __restrict__ sphi, double* __restrict__ spr,
                  double* __restrict__ slap) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= g.ny) return;
  for (int i = chunk_lo(g.nx), i_hi = chunk_hi(g.nx); i < i_hi; i++) {
    const unsigned f = cell_flags<RECT>(g, i, j);
    if (f == F_OUTSIDE) continue;
    const size_t id = cidx(g, i, j);
    const double lap = apply_cell<1>(g, f, i, j, phi + id, 0.0, -1.0);
    const double pc = phi[id];
    __stcs(su + id, u[id]);
    __stcs(sv + id, v[id]);
    __stcs(sphi + id, pc);
    __stcs(slap + id, lap);
    __stcs(spr + id, pc - half * lap);
  }
}

// ---------------------------------------------------------------------------------------------------------
// Contiguous vector kernels over the owned columns [HX*ld, (HX+nx)*ld): padding rows and masked cells hold 0
// in every vector, so plain linear loops are exact.

__global__ void __launch_bounds__(256) k_axpby(size_t n, double a, const double* __restrict__ x, double bb, double* __restrict__ y) {
  for (size_t k = blockIdx.x * (size_t)blockDim.x + threadIdx.x; k < n; k += (size_t)gridDim.x * blockDim.x)
    y[k] = a * x[k] + bb * y[k];
}

// scalars live on the device (sc[]) so the Krylov loop needs no host round trip between kernels
// p = r + beta*(p - omega*v),  beta = (rho_new/rho_old)*(alpha/omega)
__global__ void __launch_bounds__(256) k_bicg_p(size_t n, const double* __restrict__ r, double* __restrict__ p,
                                                const double* __restrict__ v, const double* __restrict__ scal) {
  const double beta = (scal[0] / scal[1]) * (scal[2] / scal[3]), omega = scal[3];
  for (size_t k = blockIdx.x * (size_t)blockDim.x + threadIdx.x; k < n; k += (size_t)gridDim.x * blockDim.x)
    p[k] = r[k] + beta * (p[k] - omega * v[k]);
}
// s = r - alpha*v ; out[0] = (s,s)
__global__ void __launch_bounds__(256) k_bicg_s(size_t n, const double* __restrict__ r, const double* __restrict__ v,
                                                double* __restrict__ s, const double* __restrict__ alpha_p, Reducer red, double* out) {
  const double alpha = *alpha_p;
  double acc[1] = {0.0};
  for (size_t k = blockIdx.x * (size_t)blockDim.x + threadIdx.x; k < n; k += (size_t)gridDim.x * blockDim.x) {
    const double sv = r[k] - alpha * v[k];
    s[k] = sv;
    acc[0] += sv * sv;
  }
  grid_reduce<1, OpSum>(acc, red, out);
}
// x += alpha*ph + omega*sh ; r = s - omega*t ; out = {(r,r), (r0,r)}
__global__ void __launch_bounds__(256) k_bicg_xr(size_t n, double* __restrict__ x, const double* __restrict__ ph,
                                                 const double* __restrict__ sh, double* __restrict__ r, const double* __restrict__ s,
                                                 const double* __restrict__ t, const double* __restrict__ r0,
                                                 const double* __restrict__ alpha_p, const double* __restrict__ omega_p,
                                                 Reducer red, double* out) {
  const double alpha = *alpha_p, omega = *omega_p;
  double acc[2] = {0.0, 0.0};
  for (size_t k = blockIdx.x * (size_t)blockDim.x + threadIdx.x; k < n; k += (size_t)gridDim.x * blockDim.x) {
    x[k] += alpha * ph[k] + omega * sh[k];
    const double rv = s[k] - omega * t[k];
    r[k] = rv;
    acc[0] += rv * rv;
    acc[1] += r0[k] * rv;
  }
  grid_reduce<2, OpSum>(acc, red, out);
}
// out = {(a,b)}
__global__ void __launch_bounds__(256) k_dot(size_t n, const double* __restrict__ a, const double* __restrict__ b, Reducer red, double* out,
                                             const PeerTable* pt) {
  double acc[1] = {0.0};
  for (size_t k = blockIdx.x * (size_t)blockDim.x + threadIdx.x; k < n; k += (size_t)gridDim.x * blockDim.x) acc[0] += a[k] * b[k];
  grid_reduce<1, OpSum>(acc, red, out, SumEpilogue{pt, out});
}
// out = {sum a}
__global__ void __launch_bounds__(256) k_sum(size_t n, const double* __restrict__ a, Reducer red, double* out) {
  double acc[1] = {0.0};
  for (size_t k = blockIdx.x * (size_t)blockDim.x + threadIdx.x; k < n; k += (size_t)gridDim.x * blockDim.x) acc[0] += a[k];
  grid_reduce<1, OpSum>(acc, red, out);
}
// y += scale * (*s) * w ; out = {(d, y), (y, y)}  (d may be null: out[0] = 0) -- the rank-one term of the augmented
// operator of outflow domains, A^ x = A x + mean(x) w, added to y = A x with the two dot products BiCGStab needs next
__global__ void __launch_bounds__(256) k_aug_axpy(size_t n, double* __restrict__ y, const double* __restrict__ w, const double* __restrict__ s,
                                                  double scale, const double* __restrict__ d, Reducer red, double* out) {
  const double a = scale * *s;
  double acc[2] = {0.0, 0.0};
  for (size_t k = blockIdx.x * (size_t)blockDim.x + threadIdx.x; k < n; k += (size_t)gridDim.x * blockDim.x) {
    const double yv = fma(a, w[k], y[k]);
    y[k] = yv;
    if (d) acc[0] = fma(d[k], yv, acc[0]);
    acc[1] = fma(yv, yv, acc[1]);
  }
  grid_reduce<2, OpSum>(acc, red, out);
}
// n doubles (n <= 32) from device memory to mapped page-locked host memory, in place of a cudaMemcpyAsync: a copy-engine transfer would wait
// behind every field copy already queued in that direction
__global__ void k_publish(double* __restrict__ host_mapped, const double* __restrict__ src, int n) {
  if ((int)threadIdx.x < n) host_mapped[threadIdx.x] = src[threadIdx.x];
}
__global__ void k_publish_ctl(SolveCtl* host_mapped, const SolveCtl* src) { *host_mapped = *src; }
// tiny scalar kernels of the Krylov recurrences: dst = num/den
__global__ void k_scalar_div(double* dst, const double* num, const double* den) { *dst = *num / *den; }
__global__ void k_scalar_copy(double* dst, const double* src) { *dst = *src; }
__global__ void k_scalar_set(double* dst, double v) { *dst = v; }
// Device-side control of the Krylov solves (BiCGStab, solve_bicgstab): the recurrence scalars live in d_scal (S), the loop condition in
// the solve's control block, and the iteration is the body of a CUDA-graph WHILE node -- the host looks at a residual once per solve.
//   k_kry_first: start of a solve (|b|^2 = S[11], iteration counter);  k_kry_begin: start of a (re)start from the TRUE residual in S[9];
//   k_kry_ctl: end of an iteration (S[9] = (r, r) of the recurrence);  k_kry_end: the true residual after the loop.
__global__ void k_kry_first(const double* S, SolveCtl* c, double tol2, int max_cycles) {
  c->bn2 = S[11]; c->tol2 = tol2; c->max_cycles = max_cycles; c->cycles = 0; c->zero_rhs = !(S[11] > 0.0) ? 1 : 0;
}
__global__ void k_kry_begin(double* S, SolveCtl* c, unsigned long long handle) {
  const double rn2 = S[9];
  S[0] = rn2; S[1] = 1.0; S[2] = 1.0; S[3] = 1.0;   // rho_new = (r0, r) = (r, r); rho_old = alpha = omega = 1
  c->rn2 = rn2; c->handle = handle;
  const bool go = !c->zero_rhs && (rn2 > c->tol2 * c->bn2) && (c->cycles < c->max_cycles) && (rn2 == rn2);
  c->go = go ? 1 : 0;
  if (handle) cudaGraphSetConditional((cudaGraphConditionalHandle)handle, go ? 1u : 0u);
}
__global__ void k_kry_ctl(const double* S, SolveCtl* c) {
  const double rn2 = S[9];
  c->rn2 = rn2; c->cycles += 1;
  const bool go = (rn2 > c->tol2 * c->bn2) && (c->cycles < c->max_cycles) && (rn2 == rn2);
  c->go = go ? 1 : 0;
  if (c->handle) cudaGraphSetConditional((cudaGraphConditionalHandle)c->handle, go ? 1u : 0u);
}
__global__ void k_kry_end(const double* S, SolveCtl* c, SolveCtl* pub) { c->rn2 = S[9]; c->go = 0; *pub = *c; }
// |b - mean|^2 from s = {sum b, sum b^2} over n cells (inv_n = 1 / n; 0: the mean stays in)
__global__ void k_bn2_from_sums(double* dst, const double* s, double inv_n) { *dst = fmax(s[1] - s[0] * s[0] * inv_n, 0.0); }

}  // namespace nsgpu
