// ns_march.cuh -- the fine-level hot kernel: NSW fused red-black Gauss-Seidel sweeps in ONE pass over HBM,
// register-marching along x (the outer index), rectangle fast path.
//
// Replaces the inner iterations of KSPSolve(phiSolver) / KSPSolve(uSolver) (reference FS.cpp:547-551) on the
// operators ConstructLHS assembles (FS.cpp:105-145); identical arithmetic order to k_smooth_rb (red = (gi + j)
// even first, then black, out of place), with the division replaced by a reciprocal good to 1 ulp.
//
// Mapping.  Memory is y-fastest (id = i*ny + j, Grid.cpp:149-161).  A warp owns a strip of 64 consecutive j
// (lane l holds the pair j0 = strip - L + 2l, j0 + 1 as one double2: one red and one black cell of every
// column) and marches over the columns i of its chunk.  Per column it issues one 16-byte load of x and of b
// per lane (512 contiguous bytes per warp), keeps the 2*NSW + 2 most recent columns in registers, and applies
// the 2*NSW half-sweeps as a software pipeline: when column n arrives, half-sweep s updates column n - 1 - s.
// y-neighbours come from the partner cell of the pair or from the adjacent lane (one shuffle per half-sweep),
// x-neighbours from the register ring.  L = 2*NSW halo cells on each side of the strip are recomputed
// (60 of 64 lanes useful for NSW = 1, 56 for NSW = 2); chunks overlap by 2L columns.  No shared-memory tile, no
// block barrier inside the march: warps are independent, which is what hides the HBM latency.
// Algorithmic traffic: 24 B per cell per launch (read x, read b, write x) for NSW cell-updates.
#pragma once
#include <cuda_runtime.h>
#include "ns_cell.cuh"

namespace nsgpu {

constexpr int MARCH_WARPS = 4;                 // warps per CTA (adjacent strips)
constexpr int MARCH_THREADS = 32 * MARCH_WARPS;
constexpr int MARCH_PF = 2;                    // columns prefetched ahead in registers

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

// one half-sweep on column `c` (current state), neighbours cm (i-1) and cp (i+1); ta: the updated cell is .x
__device__ __forceinline__ void half_update(double2& c, const double2& cm, const double2& cp, const double2& bb,
                                            bool ta, double cw, double ce, double dx, const ColConst& k) {
  if (ta) {
    const double S = __shfl_up_sync(0xffffffffu, c.y, 1), N = c.y;
    double v = fma(cw, cm.x, bb.x);
    v = fma(ce, cp.x, v);
    v = fma(k.csa, S, v);
    v = fma(k.cna, N, v);
    c.x = v * fast_rcp(dx + k.dya);
  } else {
    const double S = c.x, N = __shfl_down_sync(0xffffffffu, c.x, 1);
    double v = fma(cw, cm.y, bb.y);
    v = fma(ce, cp.y, v);
    v = fma(k.csb, S, v);
    v = fma(k.cnb, N, v);
    c.y = v * fast_rcp(dx + k.dyb);
  }
}

template <int NSW>
__global__ void __launch_bounds__(MARCH_THREADS)
k_sweep_march(LevelGeom g, const double* __restrict__ mcx, const double* __restrict__ mcy,
              const double* __restrict__ xin, const double* __restrict__ b, double* __restrict__ xout, int RI) {
  constexpr int L = 2 * NSW;        // halo cells per side == pipeline lead in columns
  constexpr int K = 2 * NSW + 2;    // columns held in registers
  constexpr int UJ = 64 - 2 * L;    // useful rows per warp strip
  extern __shared__ double s_c[];   // (cw, ce, dx) of columns [r0, r0 + nrows)
  const int lane = threadIdx.x & 31;
  const int wj = blockIdx.x * MARCH_WARPS + (threadIdx.x >> 5);
  const int ib = blockIdx.y * RI, ie = min(ib + RI, g.nx);
  const int r0 = ib - 2 * L, nrows = ie + L - r0;
  for (int q = threadIdx.x; q < nrows; q += MARCH_THREADS) {
    const int a = min(max(r0 + q + HX, 0), g.nx + 2 * HX - 1);
    s_c[3 * q] = mcx[3 * a];
    s_c[3 * q + 1] = mcx[3 * a + 1];
    s_c[3 * q + 2] = mcx[3 * a + 2];
  }
  __syncthreads();
  if (wj * UJ >= g.ny) return;

  const int j0 = wj * UJ - L + 2 * lane;                     // even
  const bool jload = (j0 + 1 < g.ny + PJ);                   // pair inside the padded column (j0 >= -L >= -PJ)
  ColConst k;
  {
    const int a = min(j0 + PJ, g.ny + 2 * PJ - 2);           // clamped entries are outside the domain: cs = cn = 0
    k.csa = mcy[3 * a]; k.cna = mcy[3 * a + 1]; k.dya = mcy[3 * a + 2];
    k.csb = mcy[3 * a + 3]; k.cnb = mcy[3 * a + 4]; k.dyb = mcy[3 * a + 5];
  }
  const size_t joff = (size_t)(j0 + PJ);
  const int ld = g.ld;
  const bool jstore2 = lane >= NSW && lane < 32 - NSW && (j0 + 1 < g.ny);
  const bool jstore1 = lane >= NSW && lane < 32 - NSW && (j0 < g.ny) && !jstore2;

  double2 R[K], B[K - 1];           // R[q]: column n - q; B[q]: right-hand side of column n - 1 - q ... see below
#pragma unroll
  for (int q = 0; q < K; q++) R[q] = make_double2(0.0, 0.0);
#pragma unroll
  for (int q = 0; q < K - 1; q++) B[q] = make_double2(0.0, 0.0);

  const int nbeg = ib - L, nend = ie + L;                    // columns loaded: [nbeg, nend) within [-HX, nx + HX)
  double2 px[MARCH_PF], pb[MARCH_PF];
#pragma unroll
  for (int q = 0; q < MARCH_PF; q++) {
    px[q] = make_double2(0.0, 0.0);
    pb[q] = make_double2(0.0, 0.0);
    const int n = nbeg + q;
    if (jload && n < nend) {
      const size_t o = (size_t)(n + HX) * ld + joff;
      px[q] = *reinterpret_cast<const double2*>(xin + o);
      pb[q] = *reinterpret_cast<const double2*>(b + o);
    }
  }

  const int par0 = g.gi0 & 1;
#pragma unroll(K)
  for (int n = nbeg; n < nend; n++) {
    // rotate the rings and take the prefetched column
#pragma unroll
    for (int q = K - 1; q > 0; q--) R[q] = R[q - 1];
#pragma unroll
    for (int q = K - 2; q > 0; q--) B[q] = B[q - 1];
    R[0] = px[0];
    B[0] = pb[0];                    // B[q] = b of column n - q
#pragma unroll
    for (int q = 0; q < MARCH_PF - 1; q++) { px[q] = px[q + 1]; pb[q] = pb[q + 1]; }
    {
      const int m = n + MARCH_PF;
      double2 vx = make_double2(0.0, 0.0), vb = vx;
      if (jload && m < nend) {
        const size_t o = (size_t)(m + HX) * ld + joff;
        vx = *reinterpret_cast<const double2*>(xin + o);
        vb = *reinterpret_cast<const double2*>(b + o);
      }
      px[MARCH_PF - 1] = vx;
      pb[MARCH_PF - 1] = vb;
    }
    // half-sweep s (colour s & 1) on column r = n - 1 - s
#pragma unroll
    for (int s = 0; s < 2 * NSW; s++) {
      const int r = n - 1 - s;
      const double* cc = s_c + 3 * (r - r0);
      const bool ta = (((par0 + r + s) & 1) == 0);           // colour s&1 sits on .x when (gi + j0 + colour) is even
      half_update(R[s + 1], R[s + 2], R[s], B[s + 1], ta, cc[0], cc[1], cc[2], k);
    }
    // column n - L is final
    const int rs = n - L;
    if (rs >= ib && rs < ie) {
      double* o = xout + (size_t)(rs + HX) * ld + joff;
      if (jstore2) *reinterpret_cast<double2*>(o) = R[L];
      else if (jstore1) *o = R[L].x;
    }
  }
}

}  // namespace nsgpu
