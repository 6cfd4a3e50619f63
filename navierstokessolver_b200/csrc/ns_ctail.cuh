// ns_ctail.cuh -- the coarse tail of the V-cycle on a thread-block CLUSTER (sm_90+ hardware feature, 8 CTAs on 8 SMs of one GPC).
//
// ns_tail.cuh runs every level of at most 64 x 64 cells inside ONE CTA: 43-50 us per cycle at 4096^2, of which a 64 x 64 half-sweep alone
// takes 1.6 us (1024 threads, one block barrier per colour), and the two levels above it (128^2, 256^2) still cost two marching launches
// of 11-13 us each -- launch and pipeline-fill latency, not bandwidth.  Here the levels from 128 x 128 (CT_FIRST_CELLS) down are
// distributed over the CTAs of one cluster: CTA r keeps the columns [r nx / C, (r + 1) nx / C) of every distributed level in its shared
// memory (iterate with one halo column per side, right-hand side, reciprocal diagonal, folded coefficient tables), a half-sweep
// updates one colour and STORES the two edge columns straight into the neighbours' halo columns through distributed shared memory
// (cluster.map_shared_rank), and cluster.sync() replaces both the block barrier of the one-CTA tail and the kernel boundaries of the
// marching passes.  Levels whose column count no longer splits evenly (8 x 8 and below) live on CTA 0 alone; the restriction onto the
// first of them writes into CTA 0's shared memory, the prolongation from it reads from there.
// Same algorithm and the same folded tables as ns_march.cuh / ns_tail.cuh (red = (i + j) even first, V(nu_pre, nu_post), averaged
// residual restriction, piecewise-constant prolongation, over-relaxed sweeps on the coarsest level): swapping one for the other changes
// the result by rounding only.  Part of the replacement of KSPSolve (reference FS.cpp:547-551).
#pragma once
#include <cuda_runtime.h>
#include <cooperative_groups.h>
#include "ns_cell.cuh"
#include "ns_march.cuh"

namespace nsgpu {

namespace cg = cooperative_groups;

constexpr int CT_C = 8;                  // CTAs per cluster: the portable maximum (NSGPU_CTAIL_C=16: non-portable size, with NSGPU_CTAIL_CELLS=65536 the tail starts at 256 x 256)
constexpr int CT_THREADS = 512;
constexpr int CT_MAX_LEVELS = 12;
constexpr int CT_KC = 4;                 // cells of one colour per thread, at most
constexpr int CT_FIRST_CELLS = 128 * 128;   // first tail level: at most this many cells (NSGPU_CTAIL_CELLS)

struct CtLevel {
  int nx, ny, cx, cy, ld;      // cells of the whole level; coarsening factors towards the next level; pitch of the global arrays
  int dist;                    // 1: columns distributed over the cluster (nc = nx / C per CTA), 0: the whole level on CTA 0
  int nc;                      // columns per CTA (distributed) or nx (single)
  const double* mcx;           // folded tables (global): 3 per column (index i + HX), 3 per row (index j + PJ)
  const double* mcy;
  int xoff, boff, rdoff, sxoff, syoff;   // offsets (doubles) into a CTA's shared memory, the same in every CTA
};

struct CtDesc {
  int nlev, nu_pre, nu_post, coarse_sweeps, smem_doubles;
  double coarse_omega;
  CtLevel lv[CT_MAX_LEVELS];
};

// shared layout of a level in one CTA: x is (nc + 2) x (ny + 2) with a halo column per side and a zero ring row per end,
// value(i, j) at (i + 1) * (ny + 2) + j + 1 (i local); b and rd are nc x ny, value(i, j) at i * ny + j
__device__ __forceinline__ int ctx_(const CtLevel& L, int i, int j) { return (i + 1) * (L.ny + 2) + j + 1; }

// columns of level L owned by cluster rank r: [c0, c0 + n)
__device__ __forceinline__ void ct_cols(const CtLevel& L, int r, int& c0, int& n) {
  if (L.dist) { n = L.nc; c0 = r * L.nc; }
  else { n = r == 0 ? L.nx : 0; c0 = 0; }
}

// `sweeps` red-black sweeps on level L.  Distributed levels: every CTA of the cluster calls, one cluster barrier per half-sweep.
__device__ void ct_smooth(cg::cluster_group& cl, const CtLevel& L, double* sm, int sweeps, double omega) {
  const int r = (int)cl.block_rank();
  int c0, nc;
  ct_cols(L, r, c0, nc);
  if (!L.dist && r != 0) return;
  double* x = sm + L.xoff;
  const double* b = sm + L.boff;
  const double* rd = sm + L.rdoff;
  const double* sx = sm + L.sxoff;
  const double* sy = sm + L.syoff;
  const int hny = (L.ny + 1) >> 1, n = nc * hny, p = L.ny + 2;
  // neighbours' copies of x (distributed shared memory): my first column is the left neighbour's right halo and vice versa
  double* xl = (L.dist && r > 0) ? cl.map_shared_rank(x, r - 1) : nullptr;
  double* xr = (L.dist && r < (int)cl.num_blocks() - 1) ? cl.map_shared_rank(x, r + 1) : nullptr;
  int ci[CT_KC], cjj[CT_KC];
#pragma unroll
  for (int k = 0; k < CT_KC; k++) {
    const int c = threadIdx.x + k * CT_THREADS;
    ci[k] = c < n ? c / hny : -1;
    cjj[k] = c < n ? c - ci[k] * hny : 0;
  }
  for (int h = 0; h < 2 * sweeps; h++) {
#pragma unroll
    for (int k = 0; k < CT_KC; k++) {
      const int i = ci[k], j = 2 * cjj[k] + ((c0 + i + h) & 1);
      if (i >= 0 && j < L.ny) {
        const int q = (i + 1) * p + j + 1;
        double v = fma(sx[3 * i], x[q - p], b[i * L.ny + j]);
        v = fma(sx[3 * i + 1], x[q + p], v);
        v = fma(sy[3 * j], x[q - 1], v);
        v = fma(sy[3 * j + 1], x[q + 1], v);
        const double gs = v * rd[i * L.ny + j];
        const double xn = omega == 1.0 ? gs : fma(omega, gs - x[q], x[q]);
        x[q] = xn;
        if (i == 0 && xl) xl[(nc + 1) * p + j + 1] = xn;          // left neighbour's right halo column (local index nc)
        if (i == nc - 1 && xr) xr[j + 1] = xn;                     // right neighbour's left halo column (local index -1)
      }
    }
    if (L.dist) cl.sync(); else __syncthreads();
  }
}

// b_coarse <- average over the children of (b - A x); x of the fine level (with halo columns) is final
__device__ void ct_restrict(cg::cluster_group& cl, const CtLevel& F, const CtLevel& C, double* sm) {
  const int r = (int)cl.block_rank();
  int f0, nf, k0, nk;
  ct_cols(F, r, f0, nf);
  ct_cols(C, r, k0, nk);
  if (nf > 0) {
    const double* x = sm + F.xoff;
    const double* b = sm + F.boff;
    const double* sx = sm + F.sxoff;
    const double* sy = sm + F.syoff;
    // coarse cells whose children this CTA owns: global coarse columns [f0 / cx, (f0 + nf) / cx)
    const int I0 = f0 / F.cx, nI = nf / F.cx;
    double* bc = (C.dist || r == 0) ? sm + C.boff : cl.map_shared_rank(sm + C.boff, 0);   // single coarse level: CTA 0's memory
    const int n = nI * C.ny, p = F.ny + 2;
    const double scale = 1.0 / (double)(F.cx * F.cy);
    for (int c = threadIdx.x; c < n; c += CT_THREADS) {
      const int I = c / C.ny, J = c - I * C.ny;
      double acc = 0.0;
      for (int di = 0; di < F.cx; di++)
        for (int dj = 0; dj < F.cy; dj++) {
          const int i = I * F.cx + di, j = J * F.cy + dj, q = ctx_(F, i, j);
          double v = fma(sx[3 * i], x[q - p], b[i * F.ny + j]);
          v = fma(sx[3 * i + 1], x[q + p], v);
          v = fma(sy[3 * j], x[q - 1], v);
          v = fma(sy[3 * j + 1], x[q + 1], v);
          acc += fma(-(sx[3 * i + 2] + sy[3 * j + 2]), x[q], v);
        }
      bc[(C.dist ? I0 + I - k0 : I0 + I) * C.ny + J] = acc * scale;   // local coarse column, or the global one in CTA 0's memory
    }
  }
  if (F.dist && !C.dist) cl.sync(); else __syncthreads();
}

// x_fine += x_coarse(parent), halo columns included (their parents are the coarse halo columns, or CTA 0's cells)
__device__ void ct_prolong(cg::cluster_group& cl, const CtLevel& F, const CtLevel& C, double* sm) {
  const int r = (int)cl.block_rank();
  int f0, nf, k0, nk;
  ct_cols(F, r, f0, nf);
  ct_cols(C, r, k0, nk);
  if (F.dist && !C.dist) cl.sync();          // CTA 0 has finished the single levels; everybody reads its x
  if (nf > 0) {
    double* x = sm + F.xoff;
    const bool remote = F.dist && !C.dist && r != 0;
    const double* xc = remote ? cl.map_shared_rank(sm + C.xoff, 0) : sm + C.xoff;
    const int ilo = (F.dist && f0 > 0) ? -1 : 0, ihi = (F.dist && f0 + nf < F.nx) ? nf + 1 : nf;   // halo columns inside the domain
    const int n = (ihi - ilo) * F.ny;
    for (int c = threadIdx.x; c < n; c += CT_THREADS) {
      const int a = c / F.ny, j = c - a * F.ny, i = ilo + a;
      const int Ig = (f0 + i) / F.cx;                            // global coarse column (f0 + i >= 0)
      x[ctx_(F, i, j)] += xc[ctx_(C, Ig - k0, j / F.cy)];
    }
  }
  // the first half-sweep that follows stores into the neighbours' halo columns, which they have just updated here themselves
  if (F.dist) cl.sync(); else __syncthreads();
}

// Everything in a CTA's shared memory that does not change from launch to launch (tables, reciprocal diagonals, zeros): built once per
// context as C images in global memory, copied in at the start of every launch.
__global__ void __launch_bounds__(CT_THREADS) k_ctail_image(CtDesc d, double* __restrict__ img) {
  const int r = blockIdx.x;
  double* im = img + (size_t)r * d.smem_doubles;
  for (int c = threadIdx.x; c < d.smem_doubles; c += CT_THREADS) im[c] = 0.0;
  __syncthreads();
  for (int l = 0; l < d.nlev; l++) {
    const CtLevel& L = d.lv[l];
    const int c0 = L.dist ? r * L.nc : 0, nc = L.dist ? L.nc : (r == 0 ? L.nx : 0);
    for (int c = threadIdx.x; c < 3 * nc; c += CT_THREADS) im[L.sxoff + c] = L.mcx[3 * (HX + c0) + c];
    for (int c = threadIdx.x; c < 3 * L.ny; c += CT_THREADS) im[L.syoff + c] = L.mcy[3 * PJ + c];
    for (int c = threadIdx.x; c < nc * L.ny; c += CT_THREADS) {
      const int i = c / L.ny, j = c - i * L.ny;
      im[L.rdoff + c] = fast_rcp(L.mcx[3 * (HX + c0 + i) + 2] + L.mcy[3 * (PJ + j) + 2]);
    }
  }
}

// b0: right-hand side of the first tail level (global, padded layout); x0: its solution array (global).  Launched as ONE cluster.
__global__ void __launch_bounds__(CT_THREADS)
k_cluster_tail(CtDesc d, const double* __restrict__ img, const double* __restrict__ b0, double* __restrict__ x0, long long* dbg) {
  extern __shared__ double sm[];
  cg::cluster_group cl = cg::this_cluster();
  const int r = (int)cl.block_rank();
  int ts = 0;
#define CT_TS() do { if (dbg && r == 0 && threadIdx.x == 0) dbg[ts++] = clock64(); } while (0)
  CT_TS();
  {
    const CtLevel& L = d.lv[0];
    const double2* src = reinterpret_cast<const double2*>(img + (size_t)r * d.smem_doubles);
    double2* dst = reinterpret_cast<double2*>(sm);
    for (int c = threadIdx.x; c < d.smem_doubles / 2; c += CT_THREADS) dst[c] = src[c];   // smem_doubles is even
    __syncthreads();
    const int c0 = r * L.nc, n = L.nc * L.ny;
    for (int c = threadIdx.x; c < n; c += CT_THREADS) {
      const int i = c / L.ny, j = c - i * L.ny;
      sm[L.boff + c] = b0[(size_t)(c0 + i + HX) * L.ld + j + PJ];
    }
  }
  cl.sync();                                                        // every CTA's memory is initialised before anybody stores into it
  CT_TS();
  for (int l = 0; l + 1 < d.nlev; l++) {                            // down: x = 0, pre-smooth, restrict
    ct_smooth(cl, d.lv[l], sm, d.nu_pre, 1.0);
    CT_TS();
    ct_restrict(cl, d.lv[l], d.lv[l + 1], sm);
    CT_TS();
  }
  ct_smooth(cl, d.lv[d.nlev - 1], sm, d.coarse_sweeps, d.coarse_omega);
  CT_TS();
  for (int l = d.nlev - 2; l >= 0; l--) {                           // up: prolong, post-smooth
    ct_prolong(cl, d.lv[l], d.lv[l + 1], sm);
    CT_TS();
    ct_smooth(cl, d.lv[l], sm, d.nu_post, 1.0);
    CT_TS();
  }
  {
    const CtLevel& L = d.lv[0];
    const int c0 = r * L.nc, n = L.nc * L.ny;
    for (int c = threadIdx.x; c < n; c += CT_THREADS) {
      const int i = c / L.ny, j = c - i * L.ny;
      x0[(size_t)(c0 + i + HX) * L.ld + j + PJ] = sm[L.xoff + ctx_(L, i, j)];
    }
  }
  cl.sync();                                                        // nobody exits while a neighbour may still read its shared memory
  CT_TS();
#undef CT_TS
}

}  // namespace nsgpu
