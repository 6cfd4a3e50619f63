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
  int xoff, boff, rdoff, sxoff, syoff;   // offsets (doubles) into shared memory; rd = 1 / diagonal per cell
};

struct TailDesc {
  int nlev, nu_pre, nu_post, coarse_sweeps, smem_doubles, gi0;
  double coarse_omega;   // over-relaxation of the sweeps on the coarsest level (1: plain Gauss-Seidel)
  TailLevel lv[TAIL_MAX_LEVELS];
};

// shared layout of a level: (nx + 2) x (ny + 2) with a zero ring, value(i, j) at (i + 1) * (ny + 2) + j + 1
__device__ __forceinline__ int tidx(const TailLevel& L, int i, int j) { return (i + 1) * (L.ny + 2) + j + 1; }

// Levels whose colour fits one warp are swept by warp 0 alone with warp barriers; the other warps wait at one block
// barrier behind the sweeps.  The block barrier of a 1024-thread CTA is what a small level's half-sweep costs otherwise.
__device__ void tail_smooth(const TailDesc& d, const TailLevel& L, double* sm, int sweeps, double omega = 1.0) {
  double* x = sm + L.xoff;
  const double* b = sm + L.boff;
  const double* rd = sm + L.rdoff;
  const double* sx = sm + L.sxoff;
  const double* sy = sm + L.syoff;
  const int hny = (L.ny + 1) >> 1, n = L.nx * hny, p = L.ny + 2;
  const bool one_warp = n <= 32;
  if (one_warp && threadIdx.x >= 32) { __syncthreads(); return; }
  // the cells of a colour this thread owns (n <= 2112 for 4096 cells: at most three), fixed over the half-sweeps
  constexpr int KC = 3;
  static_assert(TAIL_MAX_CELLS / 2 + 64 <= KC * TAIL_THREADS, "cells of one colour per thread");
  int ci[KC], cjj[KC];
#pragma unroll
  for (int k = 0; k < KC; k++) {
    const int c = threadIdx.x + k * TAIL_THREADS;
    ci[k] = c < n ? c / hny : -1;
    cjj[k] = c < n ? c - ci[k] * hny : 0;
  }
  for (int h = 0; h < 2 * sweeps; h++) {
#pragma unroll
    for (int k = 0; k < KC; k++) {   // only the cells of this half-sweep's colour
      const int i = ci[k], j = 2 * cjj[k] + ((d.gi0 + i + h) & 1);
      if (i >= 0 && j < L.ny) {
        const int q = (i + 1) * p + j + 1;
        double v = fma(sx[3 * i], x[q - p], b[q]);
        v = fma(sx[3 * i + 1], x[q + p], v);
        v = fma(sy[3 * j], x[q - 1], v);
        v = fma(sy[3 * j + 1], x[q + 1], v);
        const double gs = v * rd[q];
        x[q] = omega == 1.0 ? gs : fma(omega, gs - x[q], x[q]);
      }
    }
    if (one_warp) __syncwarp(); else __syncthreads();
  }
  if (one_warp) __syncthreads();
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

// Everything in the tail's shared memory that does not change from launch to launch -- the folded coefficient tables,
// the reciprocal diagonals, the zero rings and the zero initial guess -- is laid out once, at nsgpu_create, as an image in
// global memory; a launch starts by copying the image (one coalesced pass, L2 hits) instead of rebuilding it.
__global__ void __launch_bounds__(TAIL_THREADS) k_tail_image(TailDesc d, double* __restrict__ img) {
  for (int c = threadIdx.x; c < d.smem_doubles; c += TAIL_THREADS) img[c] = 0.0;
  __syncthreads();
  for (int l = 0; l < d.nlev; l++) {
    const TailLevel& L = d.lv[l];
    for (int c = threadIdx.x; c < 3 * L.nx; c += TAIL_THREADS) img[L.sxoff + c] = L.mcx[3 * HX + c];
    for (int c = threadIdx.x; c < 3 * L.ny; c += TAIL_THREADS) img[L.syoff + c] = L.mcy[3 * PJ + c];
    const int n = L.nx * L.ny;
    for (int c = threadIdx.x; c < n; c += TAIL_THREADS) {
      const int i = c / L.ny, j = c - i * L.ny;
      img[L.rdoff + tidx(L, i, j)] = fast_rcp(L.mcx[3 * (HX + i) + 2] + L.mcy[3 * (PJ + j) + 2]);
    }
  }
}

// b0: right-hand side of the first tail level (global, padded layout); x0: its solution array (global)
__global__ void __launch_bounds__(TAIL_THREADS)
k_coarse_tail(TailDesc d, const double* __restrict__ img, const double* __restrict__ b0, double* __restrict__ x0,
              const PeerTable* pt, unsigned gather_ctas, long long* dbg) {
  extern __shared__ double sm[];
  int ts = 0;
#define TAIL_TS() do { if (dbg && threadIdx.x == 0) dbg[ts++] = clock64(); } while (0)
  TAIL_TS();
  if (pt) {   // several ranks: every rank's restriction pass stores its slab of b0 here (peer memory)
    if (threadIdx.x == 0) {
      const unsigned long long need = pt->hdr[HDR_TAIL_NEED] + (unsigned long long)pt->world * gather_ctas;
      wait_word(pt->hdr + HDR_TAIL_ARR, need, false, pt->err);
      pt->hdr[HDR_TAIL_NEED] = need;
    }
    __syncthreads();
  }
  constexpr int KB = TAIL_MAX_CELLS / TAIL_THREADS;
  double bv[KB];
  {
    const TailLevel& L = d.lv[0];
    const int n = L.nx * L.ny;
#pragma unroll
    for (int k = 0; k < KB; k++) {   // issued before the image copy, stored behind it
      const int c = threadIdx.x + k * TAIL_THREADS;
      const int i = c / L.ny, j = c - i * L.ny;
      bv[k] = c < n ? b0[(size_t)(i + HX) * L.ld + j + PJ] : 0.0;
    }
  }
  {
    const double2* src = reinterpret_cast<const double2*>(img);
    double2* dst = reinterpret_cast<double2*>(sm);
    for (int c = threadIdx.x; c < d.smem_doubles / 2; c += TAIL_THREADS) dst[c] = src[c];   // smem_doubles is even
  }
  __syncthreads();
  {
    const TailLevel& L = d.lv[0];
    const int n = L.nx * L.ny;
#pragma unroll
    for (int k = 0; k < KB; k++) {
      const int c = threadIdx.x + k * TAIL_THREADS;
      const int i = c / L.ny, j = c - i * L.ny;
      if (c < n) sm[L.boff + tidx(L, i, j)] = bv[k];
    }
  }
  __syncthreads();
  TAIL_TS();
  for (int l = 0; l + 1 < d.nlev; l++) {          // down: x = 0, pre-smooth, restrict
    tail_smooth(d, d.lv[l], sm, d.nu_pre);
    TAIL_TS();
    tail_restrict(d.lv[l], d.lv[l + 1], sm);
    TAIL_TS();
  }
  tail_smooth(d, d.lv[d.nlev - 1], sm, d.coarse_sweeps, d.coarse_omega);   // the last tail level is the coarsest level of the hierarchy
  TAIL_TS();
  for (int l = d.nlev - 2; l >= 0; l--) {         // up: prolong, post-smooth
    tail_prolong(d.lv[l], d.lv[l + 1], sm);
    TAIL_TS();
    tail_smooth(d, d.lv[l], sm, d.nu_post);
    TAIL_TS();
  }
  {
    const TailLevel& L = d.lv[0];
    const int n = L.nx * L.ny;
    for (int c = threadIdx.x; c < n; c += TAIL_THREADS) {
      const int i = c / L.ny, j = c - i * L.ny;
      x0[(size_t)(i + HX) * L.ld + j + PJ] = sm[L.xoff + tidx(L, i, j)];
    }
  }
  TAIL_TS();
#undef TAIL_TS
}

}  // namespace nsgpu
