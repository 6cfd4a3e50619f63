// ns_march.cuh -- the hot kernel of both linear solves: NSW fused red-black Gauss-Seidel sweeps in ONE pass over
// HBM, register-marching along x (the outer index), with the neighbouring multigrid transfers fused in:
//   PROLONG   the coarse-grid correction is added while the column is loaded      (x <- x + P xc)
//   RES == 1  the residual of the smoothed iterate is restricted to the coarse rhs (bc <- R (b - A x))
//   RES == 2  the squared residual norm of the smoothed iterate is reduced         (out <- |b - A x|^2)
// so that one V(2,2) cycle touches every level exactly twice.  Rectangle fast path (1-D folded coefficients).
//
// Replaces the inner iterations of KSPSolve(phiSolver) / KSPSolve(uSolver) (reference FS.cpp:547-551) on the
// operators ConstructLHS assembles (FS.cpp:105-145); same arithmetic order as k_smooth_rb (red = (gi + j) even
// first, then black, out of place), with the division replaced by a reciprocal good to 1 ulp.
//
// Mapping.  Memory is y-fastest (id = i*ny + j, Grid.cpp:149-161).  A warp owns a strip of 64 consecutive j
// (lane l holds the pair j0 = strip - HJ + 2l, j0 + 1 as one double2: one red and one black cell of every
// column) and marches over the columns i of its chunk.  Per column it issues one 16-byte load of x and of b
// per lane (512 contiguous bytes per warp), keeps the most recent columns in registers, and applies the 2*NSW
// half-sweeps as a software pipeline: when column n arrives, half-sweep s updates column n - 1 - s; with RES
// the residual of column n - 2*NSW - 1 follows.  y-neighbours come from the partner cell of the pair or from
// the adjacent lane (one shuffle per half-sweep), x-neighbours from the register ring.  The halo cells on each
// side of the strip (2*NSW, two more with RES) are recomputed, chunks overlap by twice that many columns.
// No shared-memory tile and no block barrier inside the march: warps are independent, which is what hides the
// HBM latency.  Algorithmic traffic: 24 B per cell per launch (read x, read b, write x) for NSW cell-updates.
#pragma once
#include <cuda_runtime.h>
#include "ns_cell.cuh"

namespace nsgpu {

constexpr int MARCH_WARPS = 4;                 // warps per CTA (adjacent strips)
constexpr int MARCH_THREADS = 32 * MARCH_WARPS;
constexpr int MARCH_PF = 3;                    // default number of columns prefetched ahead in registers
#ifndef NS_MARCH_PFD
#define NS_MARCH_PFD 4
#endif
constexpr int MARCH_PFD = NS_MARCH_PFD;   // L2 prefetch distance beyond the register ring (fused variants): lead + PF + MARCH_PFD + 2 <= HX + SLACK_COLS

__host__ __device__ constexpr int march_halo(int nsw, int res) { return 2 * nsw + (res ? 2 : 0); }      // cells, y
__host__ __device__ constexpr int march_lead(int nsw, int res) { return 2 * nsw + (res ? 1 : 0); }      // columns, x
__host__ __device__ constexpr int march_uj(int nsw, int res) { return 64 - 2 * march_halo(nsw, res); }  // useful j per warp

// reciprocal to ~1 ulp: MUFU.RCP64H seed (>= 20 bits) + two Newton steps
__device__ __forceinline__ double fast_rcp(double d) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
  double e = fma(-d, r, 1.0);
  r = fma(r, e, r);
  e = fma(-d, r, 1.0);
  return fma(r, e, r);
}

struct ColConst { double csa, cna, dya, csb, cnb, dyb; };   // y-coefficients of the lane's two cells

// one half-sweep on column `c` (current state), neighbours cm (i-1) and cp (i+1); TA: the updated cell is .x
// SOR: over-relaxed update x <- x + omega (x_gs - x) (red-black SOR, used for the diagonally dominant LHS_V)
template <bool TA, bool SOR>
__device__ __forceinline__ void half_update(double2& c, const double2& cm, const double2& cp, const double2& bb,
                                            double cw, double ce, double dx, const ColConst& k, double omega) {
  if (TA) {
    const double S = __shfl_up_sync(0xffffffffu, c.y, 1), N = c.y;
    double v = fma(cw, cm.x, bb.x);
    v = fma(ce, cp.x, v);
    v = fma(k.csa, S, v);
    v = fma(k.cna, N, v);
    if (SOR) c.x = fma(omega, fma(v, fast_rcp(dx + k.dya), -c.x), c.x);
    else c.x = v * fast_rcp(dx + k.dya);
  } else {
    const double S = c.x, N = __shfl_down_sync(0xffffffffu, c.x, 1);
    double v = fma(cw, cm.y, bb.y);
    v = fma(ce, cp.y, v);
    v = fma(k.csb, S, v);
    v = fma(k.cnb, N, v);
    if (SOR) c.y = fma(omega, fma(v, fast_rcp(dx + k.dyb), -c.y), c.y);
    else c.y = v * fast_rcp(dx + k.dyb);
  }
}

// residual b - A x of both cells of the pair:  A x = (dx + dy) x_c - (cw xW + ce xE + cs xS + cn xN)
__device__ __forceinline__ double2 pair_residual(const double2& c, const double2& cm, const double2& cp, const double2& bb,
                                                 double cw, double ce, double dx, const ColConst& k) {
  const double Sa = __shfl_up_sync(0xffffffffu, c.y, 1), Nb = __shfl_down_sync(0xffffffffu, c.x, 1);
  double2 r;
  double v = fma(cw, cm.x, bb.x);
  v = fma(ce, cp.x, v);
  v = fma(k.csa, Sa, v);
  v = fma(k.cna, c.y, v);
  r.x = fma(-(dx + k.dya), c.x, v);
  v = fma(cw, cm.y, bb.y);
  v = fma(ce, cp.y, v);
  v = fma(k.csb, c.x, v);
  v = fma(k.cnb, Nb, v);
  r.y = fma(-(dx + k.dyb), c.y, v);
  return r;
}

struct MarchXfer {
  const double* xc;   // PROLONG: coarse correction (level l+1 layout)
  int ldc, pcx, pcy;
  double* bc;         // RES == 1: coarse right-hand side (level l+1 layout)
  int ldbc, rcx, rcy;
  double rscale;      // 1 / (rcx * rcy)
  double omega;       // SOR relaxation factor
  Reducer red;        // RES == 2
  double* out;
  SolveCtl* ctl;      // RES == 2: optional convergence epilogue
  // Several ranks (x-slabs, one process per GPU): the CTAs of the first / last chunk also store the outermost `pushw`
  // columns they produce straight into the neighbour's halo columns (peer memory over NVLink) and then bump the
  // neighbour's arrival counter; before reading their own halo columns they wait until the neighbours' earlier
  // pushes have arrived.  Offsets are in doubles: local cell -> the same cell seen from the neighbour's halo.
  long long dL, dR;                          // xout  (0: no neighbour on that side / no push)
  long long dbL, dbR;                        // coarse right-hand side (RES == 1)
  int pushw, nxc;                            // columns pushed per side; coarse columns of this slab
  int zone;                                  // chunks starting within `zone` columns of a slab end push, signal and wait
  const PeerTable* pt;                       // peer map (gather, reductions)
  unsigned long long* hdr;                   // this rank's control header (arrival / sent counters)
  long long offL, offR;                      // arena offsets (8-byte words) of the neighbours, 0 = none
  unsigned incL, incR;                       // CTAs of this launch that signal leftwards / rightwards (the same on every rank)
  int rb;                                    // > 0: the first two chunks are the short boundary chunks [0, rb) and [nx - rb, nx)
  int gather;                                // RES == 1 onto a replicated coarse level: store the slab on every rank
  const double* shift;                       // != nullptr: xin - *shift * shift_scale is what the pass smooths (phi's lazily removed mean)
  double shift_scale;
  int stamp;                                 // -DNSGPU_STAMPS=1 builds: this launch writes its per-CTA time stamps (NSGPU_STAMP_KEY)
};

// Per-thread state of the march: one ring of columns (PF prefetched ones ahead of the K being worked on) and
// running pointers (one add per column instead of a 64-bit multiply per access).
template <int NSW, bool PROLONG, int RES, int PF, bool SOR, bool PUSH>
struct Marcher {
  static constexpr int L = 2 * NSW;                 // half-sweep stages
  static constexpr int K = L + 2 + (RES ? 1 : 0);   // columns being worked on
  static constexpr int RL = K + PF;                 // ring length
  double2 X[RL], Bq[RL];                            // X[PF + q]: column n - q;  X[PF - 1 - p]: column n + 1 + p
  ColConst k;
  const double* cc;                                 // shared memory: (cw, ce, dx) of column n
  const double* __restrict__ px;                    // next column to prefetch (nullptr: zero initial guess)
  const double* __restrict__ pb;
  double* __restrict__ po;                          // column n - L of xout
  double* __restrict__ pbc;                         // RES == 1: next coarse column to emit
  const double* __restrict__ pxc;                   // PROLONG: coarse column base + row offset
  int ld, ny, j0, ib, ie, nxl;
  size_t pfo;                                       // MARCH_PFD columns, in doubles
  bool useful, jstore2, jstore1, pushL, pushR;
  MarchXfer xf;
  double2 racc;
  double acc;                                       // RES == 2: |r|^2
  double xshift;                                    // subtracted from every loaded x (0 unless the pass removes phi's mean)

  // Unconditional loads (no branch, so they can be scheduled early).  Lanes past the column end read the zero
  // padding (their row offset is clamped); columns past the chunk exist (SLACK_COLS extra columns are allocated
  // behind every field) and only feed pipeline stages whose results are never stored.
  __device__ __forceinline__ void load_col(int n, double2& vx, double2& vb) {
    vx = make_double2(0.0, 0.0);
    // The register ring prefetches PF columns ahead, which is shorter than the DRAM latency on the fused variants; a
    // register-free L2 prefetch MARCH_PFD columns further ahead turns those loads into L2 hits.  Unconditional: the
    // SLACK_COLS columns behind every field cover it.  Not compiled into the plain sweep (PF = 6 is enough there).
    if (MARCH_PFD > 0 && (NSW > 1 || PROLONG || RES)) {
      if (px) asm volatile("prefetch.global.L2 [%0];" :: "l"(px + pfo));
      asm volatile("prefetch.global.L2 [%0];" :: "l"(pb + pfo));
    }
    if (px) { vx = *reinterpret_cast<const double2*>(px); px += ld; }
    vb = *reinterpret_cast<const double2*>(pb);
    pb += ld;
    if (PROLONG) {
      // (measured: adding the correction when the column enters the working set instead, with the coarse value loaded one step
      // ahead, is slower -- 126 us against 96 us at 4096^2: the fine level's 400 MB per pass push the coarse array out of L2 and one
      // step does not cover a DRAM access; a coarse register ring as deep as the fine one costs the third resident CTA)
      const int ic = (xf.pcx == 2 ? ((n + 2 * HX) >> 1) : n + HX);
      const double* pc = pxc + (size_t)ic * xf.ldc;
      if (xf.pcy == 2) { const double c0 = *pc; vx.x += c0; vx.y += c0; }
      else { const double2 c2 = *reinterpret_cast<const double2*>(pc); vx.x += c2.x; vx.y += c2.y; }
    }
  }

  // column n (loaded PF steps ago, now X[PF]) enters the working set.  phi's lazily removed mean is subtracted HERE, not behind the load:
  // an instruction that consumes a load's result right away waits for it and the register prefetch is gone (every pass 20-50 % slower).
  // Only the restricting passes carry the code: the first pass of a phi solve is one (solve_head), everywhere else the shift is 0.
  __device__ __forceinline__ void arrive(int n) {
    if (RES == 1 && px) { X[PF].x -= xshift; X[PF].y -= xshift; }
  }

  // column n enters the pipeline; TA: this step's half-sweeps all update the .x cell (they do: the colour of
  // half-sweep s on column n - 1 - s depends on n only)
  template <bool TA>
  __device__ __forceinline__ void step(int n) {
#pragma unroll
    for (int q = RL - 1; q > 0; q--) { X[q] = X[q - 1]; Bq[q] = Bq[q - 1]; }
    load_col(n + PF, X[0], Bq[0]);
    arrive(n);
#pragma unroll
    for (int s = 0; s < L; s++) {
      const double* c3 = cc - 3 * (1 + s);
      half_update<TA, SOR>(X[PF + s + 1], X[PF + s + 2], X[PF + s], Bq[PF + s + 1], c3[0], c3[1], c3[2], k, xf.omega);
    }
    const int rs = n - L;                              // this column is final
    if ((unsigned)(rs - ib) < (unsigned)(ie - ib)) {
      if (jstore2) *reinterpret_cast<double2*>(po) = X[PF + L];
      else if (jstore1) *po = X[PF + L].x;
      if (PUSH && pushL && rs < xf.pushw) {                    // first columns of the slab -> left neighbour's right halo
        if (jstore2) *reinterpret_cast<double2*>(po + xf.dL) = X[PF + L];
        else if (jstore1) po[xf.dL] = X[PF + L].x;
      }
      if (PUSH && pushR && rs >= nxl - xf.pushw) {             // last columns -> right neighbour's left halo
        if (jstore2) *reinterpret_cast<double2*>(po + xf.dR) = X[PF + L];
        else if (jstore1) po[xf.dR] = X[PF + L].x;
      }
    }
    po += ld;
    if (RES) {
      const int rr = n - L - 1;                        // its x-neighbours n - L - 2 and n - L are final
      const double* c3 = cc - 3 * (L + 1);
      const double2 res = pair_residual(X[PF + L + 1], X[PF + L + 2], X[PF + L], Bq[PF + L + 1], c3[0], c3[1], c3[2], k);
      if ((unsigned)(rr - ib) < (unsigned)(ie - ib)) {
        if (RES == 2) {
          acc += (jstore2 || jstore1 ? res.x * res.x : 0.0) + (jstore2 ? res.y * res.y : 0.0);   // branch-free
        } else {
          // chunks start on even columns, so a coarse pair never straddles two chunks
          double2 t = res;
          bool emit = true;
          if (xf.rcx == 2) {
            if (rr & 1) { t.x += racc.x; t.y += racc.y; }
            else { racc = res; emit = false; }
          }
          if (emit) {
            const int I = xf.rcx == 2 ? (rr >> 1) : rr;   // coarse column of this slab
            const bool pl = PUSH && pushL && xf.dbL != 0 && I < xf.pushw, pr = PUSH && pushR && xf.dbR != 0 && I >= xf.nxc - xf.pushw;
            if (xf.rcy == 2) {
              if (useful && j0 < ny) {
                const double v = (t.x + t.y) * xf.rscale;
                *pbc = v;
                if (pl) pbc[xf.dbL] = v;
                if (pr) pbc[xf.dbR] = v;
                if (PUSH && xf.gather) for (int r = 0; r < xf.pt->world; r++) if (r != xf.pt->rank) pbc[xf.pt->off[r]] = v;
              }
            } else {
              const double2 v = make_double2(t.x * xf.rscale, t.y * xf.rscale);
              if (jstore2) {
                *reinterpret_cast<double2*>(pbc) = v;
                if (pl) *reinterpret_cast<double2*>(pbc + xf.dbL) = v;
                if (pr) *reinterpret_cast<double2*>(pbc + xf.dbR) = v;
                if (PUSH && xf.gather) for (int r = 0; r < xf.pt->world; r++) if (r != xf.pt->rank) *reinterpret_cast<double2*>(pbc + xf.pt->off[r]) = v;
              } else if (jstore1) {
                *pbc = v.x;
                if (pl) pbc[xf.dbL] = v.x;
                if (pr) pbc[xf.dbR] = v.x;
                if (PUSH && xf.gather) for (int r = 0; r < xf.pt->world; r++) if (r != xf.pt->rank) pbc[xf.pt->off[r]] = v.x;
              }
            }
            pbc += xf.ldbc;
          }
        }
      }
    }
    cc += 3;
  }
};

// The march of one warp over its chunk: sets the Marcher up, runs the pipelined column loop, returns the warp lane's share of |r|^2
// (RES == 2).  P: this CTA stores boundary columns into a neighbour's halo (peer memory).  A kernel of the pushing family holds BOTH
// instances and picks one per CTA (uniform branch): the conditional peer stores in the loop body cost every CTA 50 % (the loads of the
// register prefetch are no longer hoisted over them: 148 us against 100 us per fine-level pass at 4096^2, long-scoreboard stalls
// doubled -- profiles/r2_t_up_push_ncu_raw.csv), so only the two short boundary chunks of a slab run the pushing loop.
template <int NSW, bool PROLONG, int RES, int PF, bool SOR, bool P>
__device__ __forceinline__ double march_run(const LevelGeom& g, const double* __restrict__ mcy, const double* __restrict__ xin,
                                            const double* __restrict__ b, double* __restrict__ xout, const MarchXfer& xf, const double* s_c,
                                            int ib, int ie, int n0, int nend, int r0, int lane, int wj) {
#include "ns_march_body.inl"
  return m.acc;
}

// NOTE: no minBlocksPerSM here -- naming one (even 1) makes ptxas spend 15-40 more registers on every variant.
template <int NSW, bool PROLONG, int RES, int PF, bool SOR = false, bool PUSH = false>
__global__ void __launch_bounds__(MARCH_THREADS)
k_march(LevelGeom g, const double* __restrict__ mcx, const double* __restrict__ mcy, const double* __restrict__ xin,
        const double* __restrict__ b, double* __restrict__ xout, int RI, MarchXfer xf) {
  using M = Marcher<NSW, PROLONG, RES, PF, SOR, PUSH>;
  constexpr int L = M::L;
  constexpr int HJ = march_halo(NSW, RES);   // halo cells per side of the strip
  constexpr int LL = march_lead(NSW, RES);   // halo columns per side of the chunk
  extern __shared__ double s_c[];            // (cw, ce, dx) of columns [r0, r0 + nrows)
  const int lane = threadIdx.x & 31;
  const int wj = blockIdx.x * MARCH_WARPS + (threadIdx.x >> 5);
  // NSGPU_XRED_DBG=2|3: per-CTA time stamps of the last k_march launch (start, after the halo wait, after the march, end)
  const unsigned dbg_cta = blockIdx.x + gridDim.x * blockIdx.y;
  const bool dbg = NS_STAMPS && xf.stamp && threadIdx.x == 0 && dbg_cta < NS_DBG_CTAS;
  if (dbg) { g_ns_cta_t[4 * dbg_cta] = globaltimer_ns(); if (dbg_cta == 0) { g_ns_cta_grid[0] = gridDim.x; g_ns_cta_grid[1] = gridDim.y; } }
  // Several ranks: the columns a neighbour needs come out of two SHORT chunks that are scheduled first, so that they
  // are pushed and signalled long before the neighbour's next pass asks for them (and a late start costs little).
  int ib = blockIdx.y * RI, ie = min(ib + RI, g.nx);
  if (PUSH && xf.rb) {
    const int y = blockIdx.y;
    if (y == 0) { ib = 0; ie = xf.rb; }
    else if (y == 1) { ib = g.nx - xf.rb; ie = g.nx; }
    else { ib = xf.rb + (y - 2) * RI; ie = min(ib + RI, g.nx - xf.rb); }
  }
  const int nbeg = ib - LL, nend = ie + LL;  // columns needed: [nbeg, nend) inside [-HX + 1, nx + HX)
  // start on a column whose half-sweeps update the .x cells: (gi0 + n0 - 1) even; n0 >= -HX
  const int n0 = nbeg - ((g.gi0 + nbeg - 1) & 1);
  const int r0 = n0 - 1 - L - (RES ? 1 : 0), nrows = nend + 1 - r0;
  for (int q = threadIdx.x; q < nrows; q += MARCH_THREADS) {
    const int a = min(max(r0 + q + HX, 0), g.nx + 2 * HX - 1);
    s_c[3 * q] = mcx[3 * a];
    s_c[3 * q + 1] = mcx[3 * a + 1];
    s_c[3 * q + 2] = mcx[3 * a + 2];
  }
  if (PUSH && xf.zone && threadIdx.x == 0) {   // halo columns of this chunk come from a neighbour's push
    // what has arrived from the left is what every rank has sent rightwards so far, and vice versa; the counts are
    // kept on the device (updated by the last CTA of every launch) so that a replayed graph stays consistent
    if (ib < xf.zone && xf.offL) {
      const unsigned long long need = xf.hdr[HDR_SENT_R];
      if (need) wait_word(xf.hdr + HDR_ARR_L, need, false, xf.pt->err);
    }
    if (ie > g.nx - xf.zone && xf.offR) {
      const unsigned long long need = xf.hdr[HDR_SENT_L];
      if (need) wait_word(xf.hdr + HDR_ARR_R, need, false, xf.pt->err);
    }
  }
  __syncthreads();
  if (dbg) g_ns_cta_t[4 * dbg_cta + 1] = globaltimer_ns();

  // warps past the last strip run with every predicate false (no divergent exit: the shuffles stay convergent)
  // Kernels without peer stores hold the march DIRECTLY (included, not called): the same statements reached through march_run compile to a
  // plain sweep of 81 us instead of 70 us at 4096^2 (100 against 96 registers, another schedule).  The pushing family calls march_run
  // for both of its loops (one per CTA, uniform branch); the plain loop as a __noinline__ function of its own, or the pushing one, were
  // measured too and are slower (per-kind times of all eight arrangements: DESIGN section 7).
  double acc_r2;
  if (PUSH && (xf.gather || ib < xf.zone || ie > g.nx - xf.zone))   // uniform over the CTA
    acc_r2 = march_run<NSW, PROLONG, RES, PF, SOR, PUSH>(g, mcy, xin, b, xout, xf, s_c, ib, ie, n0, nend, r0, lane, wj);
  else if (PUSH)
    acc_r2 = march_run<NSW, PROLONG, RES, PF, SOR, false>(g, mcy, xin, b, xout, xf, s_c, ib, ie, n0, nend, r0, lane, wj);
  else {
    constexpr bool P = false;
#include "ns_march_body.inl"
    acc_r2 = m.acc;
  }
  if (dbg) g_ns_cta_t[4 * dbg_cta + 2] = globaltimer_ns();
  if (PUSH) {
    __syncthreads();          // every store of this CTA happens before thread 0's fence (cumulativity)
    if (threadIdx.x == 0) {
      const PeerTable* pt = xf.pt;
      unsigned long long* hdr = xf.hdr;
      const bool sl = xf.zone && ib < xf.zone && xf.offL, sr = xf.zone && ie > g.nx - xf.zone && xf.offR;
      const bool ga = RES == 1 && xf.gather;
      // pushed columns are visible system-wide before the neighbour is told; CTAs that pushed nothing skip the
      // system-scope fence (it is expensive, and thousands of them at the end of a pass serialise)
      if (sl || sr || ga) __threadfence_system();
      if (sl) atomicAdd_system(hdr + xf.offL + HDR_ARR_R, 1ull);
      if (sr) atomicAdd_system(hdr + xf.offR + HDR_ARR_L, 1ull);
      if (ga)
        for (int r = 0; r < pt->world; r++) atomicAdd_system(hdr + pt->off[r] + HDR_TAIL_ARR, 1ull);
      const unsigned long long total = (unsigned long long)gridDim.x * gridDim.y;
      if (atomicAdd(hdr + HDR_TICKET, 1ull) == total - 1) {   // every CTA has read the sent counters: advance them
        hdr[HDR_SENT_L] += xf.incL;
        hdr[HDR_SENT_R] += xf.incR;
        hdr[HDR_TICKET] = 0;
        __threadfence();
      }
    }
  }
  if (dbg) g_ns_cta_t[4 * dbg_cta + 3] = globaltimer_ns();
  if (RES == 2) {
    double acc[1] = {acc_r2};
    grid_reduce<1, OpSum>(acc, xf.red, xf.out, CtlEpilogue{xf.ctl, xf.pt, xf.out});
  }
}

}  // namespace nsgpu
