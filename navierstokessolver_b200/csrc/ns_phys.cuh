// ns_phys.cuh -- the two projection kernels of the time step on rectangles as register-marching passes:
//   k_div_march      ConstructRHS_phi = Div_V(u*, v*) / dt + the sums MatNullSpaceRemove needs   (FS.cpp:365-418, 550)   24 B / cell
//   k_correct_march  CorrectVelocities + GradP + VecMin / VecMax x 4                            (FS.cpp:420-456, 512-534, 554-557)   56 B / cell
// The one-thread-per-row versions (k_div_rhs, k_correct in ns_kernels.cuh; kept for general domains) have one column in flight per
// thread and run at 55-61 % of the HBM peak, latency-bound.  Here, as in ns_march.cuh, a lane owns the pair of rows (j0, j0 + 1) as one
// double2 -- every load and store is 16 bytes per lane, 512 contiguous bytes per warp -- and marches over the columns of its chunk with
// the next PF columns of every input already on their way (register prefetch, statically indexed: the column loop is unrolled by PF);
// x neighbours come from a three-column register ring, y neighbours from the pair partner or the adjacent lane (one halo pair per side
// of the warp: 60 useful rows of 64).  Interior cells take a straight-line path whose expressions are div_cell's / grad_cell's
// (ns_cell.cuh); cells that touch an edge go through div_cell / grad_cell themselves (first / last column: the whole warp, first / last
// row: one lane), so every boundary-condition type keeps its one implementation.
#pragma once
#include <cuda_runtime.h>
#include "ns_cell.cuh"
#include "ns_kernels.cuh"

namespace nsgpu {

constexpr int PH_WARPS = 4;
constexpr int PH_THREADS = 32 * PH_WARPS;
constexpr int PH_UJ = 60;      // useful rows per warp
constexpr int PH_PF = 4;       // columns in flight per input array (PF <= SLACK_COLS)
#ifndef NS_PH_MIN_CTAS
#define NS_PH_MIN_CTAS 5
#endif
constexpr int PH_MIN_CTAS = NS_PH_MIN_CTAS;   // resident CTAs per SM the register allocation is held to (5 x 4 warps)

struct PhysLane {
  int j0, ld;
  size_t row;                  // offset of the pair inside a column (clamped into the zero padding for lanes past the last row)
  bool st2, st1;               // the lane stores both cells / only the first (odd row count)
  bool inA, inB;               // cell a = (i, j0), b = (i, j0 + 1): interior in y (both y neighbours are fluid cells)
  double rSa, rNa, rhya, rSb, rNb, rhyb;
};

__device__ __forceinline__ PhysLane phys_lane(const LevelGeom& g, int lane, int wj) {
  PhysLane p;
  p.j0 = wj * PH_UJ - 2 + 2 * lane;                               // even, >= -2
  p.ld = g.ld;
  const int jl = min(p.j0, (g.ny + PJ - 2) & ~1);
  p.row = (size_t)(jl + PJ);
  const bool useful = lane >= 1 && lane < 31;
  p.st2 = useful && p.j0 + 1 < g.ny;
  p.st1 = useful && p.j0 < g.ny && !p.st2;
  p.inA = p.j0 > 0 && p.j0 < g.ny - 1;
  p.inB = p.j0 + 1 > 0 && p.j0 + 1 < g.ny - 1;
  const int ja = min(max(p.j0, 0), g.ny - 1) + PJ, jb = min(max(p.j0 + 1, 0), g.ny - 1) + PJ;
  const double hya = g.hy[ja], hyb = g.hy[jb];
  p.rSa = hya * g.rsy[ja - 1]; p.rNa = hya * g.rsy[ja]; p.rhya = g.rhy[ja];
  p.rSb = hyb * g.rsy[jb - 1]; p.rNb = hyb * g.rsy[jb]; p.rhyb = g.rhy[jb];
  return p;
}

// (hx_i rsx_{i-1}, hx_i rsx_i, 1 / hx_i) of the columns [ib, ie) of a chunk -> shared memory
__device__ __forceinline__ void phys_stage_x(const LevelGeom& g, int ib, int ie, double* s_x) {
  for (int q = threadIdx.x; q < ie - ib; q += PH_THREADS) {
    const int a = ib + q + HX;
    const double h = g.hx[a];
    s_x[3 * q] = h * g.rsx[a - 1];
    s_x[3 * q + 1] = h * g.rsx[a];
    s_x[3 * q + 2] = g.rhx[a];
  }
  __syncthreads();
}

__device__ __forceinline__ double2 ld2(const double* p) { return *reinterpret_cast<const double2*>(p); }
__device__ __forceinline__ void st_pair(double* p, const double2& v, bool st2, bool st1) {
  if (st2) *reinterpret_cast<double2*>(p) = v;
  else if (st1) *p = v.x;
}

// ---------------------------------------------------------------------------------------------------------
// b = Div_V(u*, v*) / dt ; out = {sum b, sum b^2}
__global__ void __launch_bounds__(PH_THREADS, PH_MIN_CTAS)
k_div_march(LevelGeom g, const EdgeDev* __restrict__ edges, StepConst sc, const double* __restrict__ u, const double* __restrict__ v,
            double* __restrict__ b, int RI, Reducer red, double* __restrict__ out) {
  extern __shared__ double s_x[];
  const int lane = threadIdx.x & 31, wj = blockIdx.x * PH_WARPS + (threadIdx.x >> 5);
  const int ib = blockIdx.y * RI, ie = min(ib + RI, g.nx);
  phys_stage_x(g, ib, ie, s_x);
  const PhysLane L = phys_lane(g, lane, wj);
  const int ld = L.ld;
  const double rdt = sc.dt;
  const double* pu = u + (size_t)(ib + HX) * ld + L.row;          // column ib
  const double* pv = v + (size_t)(ib + HX) * ld + L.row;
  double* pb = b + (size_t)(ib + HX) * ld + L.row;
  double2 um = ld2(pu - ld), uc = ld2(pu);
  double2 Fu[PH_PF], Fv[PH_PF];                                    // Fu[p]: column i + 1 + p of u; Fv[p]: column i + p of v
#pragma unroll
  for (int p = 0; p < PH_PF; p++) { Fu[p] = ld2(pu + (size_t)(1 + p) * ld); Fv[p] = ld2(pv + (size_t)p * ld); }
  double acc[2] = {0.0, 0.0};
  for (int i0 = ib; i0 < ie; i0 += PH_PF) {
#pragma unroll
    for (int p = 0; p < PH_PF; p++) {
      const int i = i0 + p;
      if (i < ie) {                                                // uniform
        const double2 up = Fu[p], vc = Fv[p];
        Fu[p] = ld2(pu + (size_t)(1 + PH_PF) * ld);               // columns i + 1 + PF, i + PF: SLACK_COLS columns exist behind every field
        Fv[p] = ld2(pv + (size_t)PH_PF * ld);
        const double vS = __shfl_up_sync(0xffffffffu, vc.y, 1), vN = __shfl_down_sync(0xffffffffu, vc.x, 1);
        const int ig = g.gi0 + i;
        const bool xin = ig > 0 && ig < g.gnx - 1;
        const double* c3 = s_x + 3 * (i - ib);
        const double rW = c3[0], rE = c3[1], rhx = c3[2];
        double2 d;
        {
          const double VW = um.x * rW + uc.x * (1.0 - rW), VE = up.x * rE + uc.x * (1.0 - rE);
          const double VS = vS * L.rSa + vc.x * (1.0 - L.rSa), VN = vc.y * L.rNa + vc.x * (1.0 - L.rNa);
          d.x = ((VE - VW) * rhx + (VN - VS) * L.rhya) / rdt;
        }
        {
          const double VW = um.y * rW + uc.y * (1.0 - rW), VE = up.y * rE + uc.y * (1.0 - rE);
          const double VS = vc.x * L.rSb + vc.y * (1.0 - L.rSb), VN = vN * L.rNb + vc.y * (1.0 - L.rNb);
          d.y = ((VE - VW) * rhx + (VN - VS) * L.rhyb) / rdt;
        }
        if (L.st2 || L.st1) {                                      // cells on an edge: the general cell function (reads its neighbours from memory)
          const size_t off = (size_t)(i + HX) * ld + L.row;
          if (!(xin && L.inA)) d.x = div_cell<true>(g, edges, sc, i, L.j0, u + off, v + off);
          if (L.st2 && !(xin && L.inB)) d.y = div_cell<true>(g, edges, sc, i, L.j0 + 1, u + off + 1, v + off + 1);
        }
        st_pair(pb, d, L.st2, L.st1);
        if (L.st2 || L.st1) { acc[0] += d.x; acc[1] = fma(d.x, d.x, acc[1]); }
        if (L.st2) { acc[0] += d.y; acc[1] = fma(d.y, d.y, acc[1]); }
        um = uc; uc = up;
        pu += ld; pv += ld; pb += ld;
      }
    }
  }
  grid_reduce<2, OpSum>(acc, red, out);
}

// ---------------------------------------------------------------------------------------------------------
// divPhi = GradP(phi); u = u* - dt divPhi_x; v = v* - dt divPhi_y; out[0..3] = max(-u), max(u), max(-v), max(v) (NaN-propagating)
// phi may carry a constant offset (the mean is removed lazily, ns_solver.inl): only differences of phi enter.
__global__ void __launch_bounds__(PH_THREADS, PH_MIN_CTAS)
k_correct_march(LevelGeom g, StepConst sc, const double* __restrict__ phi, const double* __restrict__ us, const double* __restrict__ vs,
                double* __restrict__ u, double* __restrict__ v, double* __restrict__ gpx, double* __restrict__ gpy, int RI, Reducer red,
                double* __restrict__ out) {
  extern __shared__ double s_x[];
  const int lane = threadIdx.x & 31, wj = blockIdx.x * PH_WARPS + (threadIdx.x >> 5);
  const int ib = blockIdx.y * RI, ie = min(ib + RI, g.nx);
  phys_stage_x(g, ib, ie, s_x);
  const PhysLane L = phys_lane(g, lane, wj);
  const int ld = L.ld;
  const size_t base = (size_t)(ib + HX) * ld + L.row;
  const double* pp = phi + base;
  const double* pus = us + base;
  const double* pvs = vs + base;
  size_t o = base;                                                 // running offset of the output pair
  double2 pm = ld2(pp - ld), pc = ld2(pp);
  double2 Fp[PH_PF], Fu[PH_PF], Fv[PH_PF];                         // Fp[p]: column i + 1 + p of phi; Fu / Fv[p]: column i + p of u*, v*
#pragma unroll
  for (int p = 0; p < PH_PF; p++) { Fp[p] = ld2(pp + (size_t)(1 + p) * ld); Fu[p] = ld2(pus + (size_t)p * ld); Fv[p] = ld2(pvs + (size_t)p * ld); }
  double acc[4] = {OpMax::id(), OpMax::id(), OpMax::id(), OpMax::id()};
  const double inf = 1.0 / 0.0;
  for (int i0 = ib; i0 < ie; i0 += PH_PF) {
#pragma unroll
    for (int p = 0; p < PH_PF; p++) {
      const int i = i0 + p;
      if (i < ie) {                                                // uniform
        const double2 pn = Fp[p], usc = Fu[p], vsc = Fv[p];
        Fp[p] = ld2(pp + (size_t)(1 + PH_PF) * ld);
        Fu[p] = ld2(pus + (size_t)PH_PF * ld);
        Fv[p] = ld2(pvs + (size_t)PH_PF * ld);
        const double pS = __shfl_up_sync(0xffffffffu, pc.y, 1), pN = __shfl_down_sync(0xffffffffu, pc.x, 1);
        const int ig = g.gi0 + i;
        const bool xin = ig > 0 && ig < g.gnx - 1;
        const double* c3 = s_x + 3 * (i - ib);
        const double rW = c3[0], rE = c3[1], rhx = c3[2];
        double2 gx, gy;
        {
          const double VW = pm.x * rW + pc.x * (1.0 - rW), VE = pn.x * rE + pc.x * (1.0 - rE);
          const double VS = pS * L.rSa + pc.x * (1.0 - L.rSa), VN = pc.y * L.rNa + pc.x * (1.0 - L.rNa);
          gx.x = (VE - VW) * rhx; gy.x = (VN - VS) * L.rhya;
        }
        {
          const double VW = pm.y * rW + pc.y * (1.0 - rW), VE = pn.y * rE + pc.y * (1.0 - rE);
          const double VS = pc.x * L.rSb + pc.y * (1.0 - L.rSb), VN = pN * L.rNb + pc.y * (1.0 - L.rNb);
          gx.y = (VE - VW) * rhx; gy.y = (VN - VS) * L.rhyb;
        }
        if (L.st2 || L.st1) {                                      // cells on an edge: grad_cell itself (wall and outflow closures)
          if (!(xin && L.inA)) { const UV gr = grad_cell(g, cell_flags<true>(g, i, L.j0), i, L.j0, phi + o); gx.x = gr.u; gy.x = gr.v; }
          if (L.st2 && !(xin && L.inB)) { const UV gr = grad_cell(g, cell_flags<true>(g, i, L.j0 + 1), i, L.j0 + 1, phi + o + 1); gx.y = gr.u; gy.y = gr.v; }
        }
        const double2 un = make_double2(usc.x - sc.dt * gx.x, usc.y - sc.dt * gx.y);
        const double2 vn = make_double2(vsc.x - sc.dt * gy.x, vsc.y - sc.dt * gy.y);
        st_pair(gpx + o, gx, L.st2, L.st1);
        st_pair(gpy + o, gy, L.st2, L.st1);
        st_pair(u + o, un, L.st2, L.st1);
        st_pair(v + o, vn, L.st2, L.st1);
        if (L.st2 || L.st1) {
          const bool bad = !(isfinite(un.x) && isfinite(vn.x));
          acc[0] = fmax(acc[0], bad ? inf : -un.x); acc[1] = fmax(acc[1], bad ? inf : un.x);
          acc[2] = fmax(acc[2], bad ? inf : -vn.x); acc[3] = fmax(acc[3], bad ? inf : vn.x);
        }
        if (L.st2) {
          const bool bad = !(isfinite(un.y) && isfinite(vn.y));
          acc[0] = fmax(acc[0], bad ? inf : -un.y); acc[1] = fmax(acc[1], bad ? inf : un.y);
          acc[2] = fmax(acc[2], bad ? inf : -vn.y); acc[3] = fmax(acc[3], bad ? inf : vn.y);
        }
        pm = pc; pc = pn;
        pp += ld; pus += ld; pvs += ld; o += ld;
      }
    }
  }
  grid_reduce<4, OpMax>(acc, red, out);
}

}  // namespace nsgpu
