// ns_tail.cuh -- the coarse tail of the V-cycle in ONE single-CTA kernel.
//
// Every multigrid level of at most TAIL_MAX_CELLS cells lives in shared memory; the kernel receives the
// right-hand side of the first tail level (written by the fused restriction of the level above), runs the rest
// of the V(nu_pre, nu_post) cycle -- smoothing, residual restriction, coarsest-level sweeps, prolongation --
// with block barriers instead of kernel launches, and writes the correction back.  Same algorithm and same
// folded coefficient tables as the marching passes (ns_march.cuh), so swapping one for the other changes the
// result by rounding only.  Part of the replacement of KSPSolve (reference FS.cpp:547-551).
#pragma once
#include <cuda_runtime.h>
#include "ns_cell.cuh"
#include "ns_march.cuh"

namespace nsgpu {

constexpr int TAIL_MAX_LEVELS = 10;
constexpr int TAIL_MAX_CELLS = 4096;     // first tail level: at most 64 x 64 cells
constexpr int TAIL_THREADS = 1024;

struct TailLevel {
  int nx, ny, cx, cy;        // cells; coarsening factors towards the next level
  int ld;                    // pitch of the global arrays of this level
  const double* mcx;         // folded tables (global): 3 per column (index i + HX), 3 per row (index j + PJ)
  const double* mcy;
  int xoff, boff, sxoff, syoff;   // offsets (doubles) into shared memory
};

struct TailDesc {
  int nlev, nu_pre, nu_post, coarse_sweeps, smem_doubles, gi0;
  double coarse_omega;   // over-relaxation of the sweeps on the coarsest level (1: plain Gauss-Seidel)
  TailLevel lv[TAIL_MAX_LEVELS];
};

// shared layout of a level: (nx + 2) x (ny + 2) with a zero ring, value(i, j) at (i + 1) * (ny + 2) + j + 1
__device__ __forceinline__ int tidx(const TailLevel& L, int i, int j) { return (i + 1) * (L.ny + 2) + j + 1; }

__device__ void tail_smooth(const TailDesc& d, const TailLevel& L, double* sm, int sweeps, double omega = 1.0) {
  double* x = sm + L.xoff;
  const double* b = sm + L.boff;
  const double* sx = sm + L.sxoff;
  const double* sy = sm + L.syoff;
  const int hny = (L.ny + 1) >> 1, n = L.nx * hny, p = L.ny + 2;
  for (int h = 0; h < 2 * sweeps; h++) {
    for (int c = threadIdx.x; c < n; c += TAIL_THREADS) {   // only the cells of this half-sweep's colour
      const int i = c / hny, j = 2 * (c - i * hny) + ((d.gi0 + i + h) & 1);
      if (j < L.ny) {
        const int q = tidx(L, i, j);
        double v = fma(sx[3 * i], x[q - p], b[q]);
        v = fma(sx[3 * i + 1], x[q + p], v);
        v = fma(sy[3 * j], x[q - 1], v);
        v = fma(sy[3 * j + 1], x[q + 1], v);
        const double gs = v * fast_rcp(sx[3 * i + 2] + sy[3 * j + 2]);
        x[q] = omega == 1.0 ? gs : fma(omega, gs - x[q], x[q]);
      }
    }
    __syncthreads();
  }
}

// b_coarse <- average over the children of (b - A x)
__device__ void tail_restrict(const TailLevel& F, const TailLevel& C, double* sm) {
  const double* x = sm + F.xoff;
  const double* b = sm + F.boff;
  const double* sx = sm + F.sxoff;
  const double* sy = sm + F.syoff;
  double* bc = sm + C.boff;
  const int n = C.nx * C.ny, p = F.ny + 2;
  const double scale = 1.0 / (double)(F.cx * F.cy);
  for (int c = threadIdx.x; c < n; c += TAIL_THREADS) {
    const int I = c / C.ny, J = c - I * C.ny;
    double acc = 0.0;
    for (int di = 0; di < F.cx; di++)
      for (int dj = 0; dj < F.cy; dj++) {
        const int i = I * F.cx + di, j = J * F.cy + dj, q = tidx(F, i, j);
        double v = fma(sx[3 * i], x[q - p], b[q]);
        v = fma(sx[3 * i + 1], x[q + p], v);
        v = fma(sy[3 * j], x[q - 1], v);
        v = fma(sy[3 * j + 1], x[q + 1], v);
        acc += fma(-(sx[3 * i + 2] + sy[3 * j + 2]), x[q], v);
      }
    bc[tidx(C, I, J)] = acc * scale;
  }
  __syncthreads();
}

__device__ void tail_prolong(const TailLevel& F, const TailLevel& C, double* sm) {
  double* x = sm + F.xoff;
  const double* xc = sm + C.xoff;
  const int n = F.nx * F.ny;
  for (int c = threadIdx.x; c < n; c += TAIL_THREADS) {
    const int i = c / F.ny, j = c - i * F.ny;
    x[tidx(F, i, j)] += xc[tidx(C, i / F.cx, j / F.cy)];
  }
  __syncthreads();
}

// b0: right-hand side of the first tail level (global, padded layout); x0: its solution array (global)
__global__ void __launch_bounds__(TAIL_THREADS)
k_coarse_tail(TailDesc d, const double* __restrict__ b0, double* __restrict__ x0, const PeerTable* pt, unsigned gather_ctas) {
  extern __shared__ double sm[];
  for (int c = threadIdx.x; c < d.smem_doubles; c += TAIL_THREADS) sm[c] = 0.0;
  if (pt && threadIdx.x == 0) {   // several ranks: every rank's restriction pass stores its slab of b0 here (peer memory)
    const unsigned long long need = pt->hdr[HDR_TAIL_NEED] + (unsigned long long)pt->world * gather_ctas;
    wait_word(pt->hdr + HDR_TAIL_ARR, need, false, pt->err);
    pt->hdr[HDR_TAIL_NEED] = need;
  }
  __syncthreads();
  for (int l = 0; l < d.nlev; l++) {
    const TailLevel& L = d.lv[l];
    for (int c = threadIdx.x; c < 3 * L.nx; c += TAIL_THREADS) sm[L.sxoff + c] = L.mcx[3 * HX + c];
    for (int c = threadIdx.x; c < 3 * L.ny; c += TAIL_THREADS) sm[L.syoff + c] = L.mcy[3 * PJ + c];
  }
  {
    const TailLevel& L = d.lv[0];
    const int n = L.nx * L.ny;
    for (int c = threadIdx.x; c < n; c += TAIL_THREADS) {
      const int i = c / L.ny, j = c - i * L.ny;
      sm[L.boff + tidx(L, i, j)] = b0[(size_t)(i + HX) * L.ld + j + PJ];
    }
  }
  __syncthreads();
  for (int l = 0; l + 1 < d.nlev; l++) {          // down: x = 0, pre-smooth, restrict
    tail_smooth(d, d.lv[l], sm, d.nu_pre);
    tail_restrict(d.lv[l], d.lv[l + 1], sm);
  }
  tail_smooth(d, d.lv[d.nlev - 1], sm, d.coarse_sweeps, d.coarse_omega);   // the last tail level is the coarsest level of the hierarchy
  for (int l = d.nlev - 2; l >= 0; l--) {         // up: prolong, post-smooth
    tail_prolong(d.lv[l], d.lv[l + 1], sm);
    tail_smooth(d, d.lv[l], sm, d.nu_post);
  }
  {
    const TailLevel& L = d.lv[0];
    const int n = L.nx * L.ny;
    for (int c = threadIdx.x; c < n; c += TAIL_THREADS) {
      const int i = c / L.ny, j = c - i * L.ny;
      x0[(size_t)(i + HX) * L.ld + j + PJ] = sm[L.xoff + tidx(L, i, j)];
    }
  }
}

}  // namespace nsgpu
