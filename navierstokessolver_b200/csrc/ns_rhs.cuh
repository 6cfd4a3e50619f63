// ns_rhs.cuh -- ConstructRHS_V (FS.cpp:183-363, 458-510) on rectangles as a register-marching kernel.
//
// k_rhs_build (ns_kernels.cuh) follows the reference cell by cell: every cell evaluates its own MUSCL slopes AND
// those of its four neighbours, and every face flux is evaluated by both cells that share the face: 6 slope pairs
// and 8 Roe fluxes per cell, ~650 us at 4096^2 (64 B/cell would take 165 us).  Here a lane owns one row and marches
// over the columns of its chunk:
//   * x direction: u, v of columns i-1 .. i+2 and the slopes of columns i, i+1 stay in registers; the east-face
//     fluxes of cell i are carried to cell i+1 as its west-face fluxes;
//   * y direction: the neighbours' values, slopes and the south-face fluxes come from the adjacent lanes by shuffle
//     (2 halo lanes per side of the warp, 28 useful rows);
// so a cell costs 2 slope pairs and 4 fluxes, and u, v are read once per column.  The arithmetic of every quantity is
// the expression rhs_cell uses (ns_cell.cuh), evaluated once instead of twice.
#pragma once
#include <cuda_runtime.h>
#include "ns_cell.cuh"
#include "ns_kernels.cuh"

namespace nsgpu {

constexpr int RHS_WARPS = 4;
constexpr int RHS_THREADS = 32 * RHS_WARPS;
constexpr int RHS_UJ = 28;   // useful rows per warp

__device__ __forceinline__ double shfl_up1(double v) { return __shfl_up_sync(0xffffffffu, v, 1); }
__device__ __forceinline__ double shfl_dn1(double v) { return __shfl_down_sync(0xffffffffu, v, 1); }

// Ghost value behind side k (W, E, S, N) of a cell with value c: w * c + constant (EvaluateGhostStencil_V, FS.cpp:166-173).
// Only boundary cells come here, so the edge record is read where it is needed instead of living in registers.
struct GhostUV { double u, v; };
__device__ __forceinline__ GhostUV ghost_uv(const LevelGeom& g, const EdgeDev* __restrict__ edges, int k, double uc, double vc) {
  const EdgeDev e = edges[g.side_edge[k]];
  const double w = e.outflow ? 1.0 : -1.0;
  return GhostUV{w * uc + e.cu, w * vc + e.cv};
}

// minmod slopes of (u, v) along one direction; k_hi / k_lo: the sides towards the higher / lower index
__device__ __forceinline__ GhostUV slope_uv(const LevelGeom& g, const EdgeDev* __restrict__ edges, bool has_hi, bool has_lo, int k_hi,
                                            int k_lo, double uc, double vc, double uhi, double vhi, double ulo, double vlo,
                                            double s_hi, double s_lo, double rh) {
  double au, av, bu, bv;
  if (has_hi) { au = (uhi - uc) * s_hi; av = (vhi - vc) * s_hi; }
  else { const GhostUV gh = ghost_uv(g, edges, k_hi, uc, vc); au = (gh.u - uc) * rh; av = (gh.v - vc) * rh; }
  if (has_lo) { bu = (uc - ulo) * s_lo; bv = (vc - vlo) * s_lo; }
  else { const GhostUV gl = ghost_uv(g, edges, k_lo, uc, vc); bu = (uc - gl.u) * rh; bv = (vc - gl.v) * rh; }
  return GhostUV{minmod(au, bu), minmod(av, bv)};
}

__global__ void __launch_bounds__(RHS_THREADS)
k_rhs_march(LevelGeom g, const EdgeDev* __restrict__ edges, StepConst sc, const double* __restrict__ u,
            const double* __restrict__ v, double* __restrict__ nu0, double* __restrict__ nv0,
            const double* __restrict__ gpx, const double* __restrict__ gpy, double* __restrict__ rhs_u,
            double* __restrict__ rhs_v, int RI, Reducer red, double* __restrict__ norms, int i_begin, int i_end) {
  const int lane = threadIdx.x & 31;
  const int wj = blockIdx.x * RHS_WARPS + (threadIdx.x >> 5);
  const int j = wj * RHS_UJ + lane - 2;                  // rows -2, -1, ny, ny+1 are zero padding (PJ = 8)
  // columns [i_begin, i_end) of the slab (the whole slab, or one upload chunk of nsgpu_step_host), RI columns per CTA row
  const int ib = i_begin + blockIdx.y * RI, ie = min(ib + RI, i_end);
  const int jc = min(j, g.ny + 1);                       // rows past the padding: clamp (never stored)
  const bool in_dom = j >= 0 && j < g.ny;
  const bool store = in_dom && lane >= 2 && lane < 30;
  const int ld = g.ld;
  const double ire = 1.0 / sc.re;

  // y geometry and y-side ghost constants of this lane's row
  const int ja = max(min(jc, g.ny - 1), 0) + PJ;
  const double hyj = g.hy[ja], rhy = g.rhy[ja], rsym = g.rsy[ja - 1], rsyp = g.rsy[ja];
  const double hyN = 0.5 * g.hy[ja + 1];
  const bool fS = j > 0, fN = j < g.ny - 1;

  const size_t row = (size_t)(jc + PJ);
  const double* pu = u + (size_t)(ib - 2 + HX) * ld + row;   // next column to load: ib - 2 (then ib - 1, ...)
  const double* pv = v + (size_t)(ib - 2 + HX) * ld + row;
  auto col_ok = [&](int n) { const int ng = g.gi0 + n; return ng >= 0 && ng < g.gnx && n >= -HX && n < g.nx + HX; };
  double uF = col_ok(ib - 2) ? *pu : 0.0, vF = col_ok(ib - 2) ? *pv : 0.0;   // prefetched column

  // ring: columns i-1, i, i+1, i+2 while cell i is finished
  double uM = 0, uC = 0, uP = 0, uQ = 0, vM = 0, vC = 0, vP = 0, vQ = 0;
  double sxCu = 0, sxCv = 0, sxPu = 0, sxPv = 0;   // x slopes of columns i and i+1
  double cWsq = 0, cWuv = 0;                       // east-face fluxes of cell i-1 = west-face fluxes of cell i
  double nrm[2] = {0.0, 0.0};                      // |rhs_u|^2, |rhs_v|^2 of the cells this thread stores

  // n: column entering the ring; cell i = n - 2 is finished; the first finished cell is ib - 1 (its east face only)
  for (int n = ib - 2; n < ie + 2; n++) {
    uM = uC; uC = uP; uP = uQ; vM = vC; vC = vP; vP = vQ;
    uQ = uF; vQ = vF;
    pu += ld; pv += ld;
    { const bool ok = col_ok(n + 1) && n + 1 < ie + 2; uF = ok ? *pu : 0.0; vF = ok ? *pv : 0.0; }   // next column, one step ahead
    const int ic = n - 2;   // history terms of the cell that is finished in this step
    const bool ld_hist = store && ic >= ib;
    const size_t idc = (size_t)(ic + HX) * ld + row;
    const double hu = ld_hist ? nu0[idc] : 0.0, hv = ld_hist ? nv0[idc] : 0.0;
    sxCu = sxPu; sxCv = sxPv;
    {  // x slope of column n - 1 (needs columns n - 2, n - 1, n)
      const int i1 = n - 1, ig = g.gi0 + i1;
      if (ig >= 0 && ig < g.gnx && i1 >= -HX + 1) {
        const bool hW = ig > 0, hE = ig < g.gnx - 1;
        const int a = i1 + HX;
        const double sE = 2.0 * g.rsx[a], sW = 2.0 * g.rsx[a - 1], rh = g.rhx[a];
        const GhostUV sl = slope_uv(g, edges, hE, hW, 1, 0, uP, vP, uQ, vQ, uC, vC, sE, sW, rh);
        sxPu = sl.u; sxPv = sl.v;
      } else { sxPu = 0.0; sxPv = 0.0; }
    }
    const int i = n - 2, ig = g.gi0 + i;
    if (i < ib - 1) continue;
    if (ig < 0 || ig >= g.gnx) { cWsq = 0.0; cWuv = 0.0; continue; }   // uniform over the block
    const bool fW = ig > 0, fE = ig < g.gnx - 1;
    const int a = i + HX;
    const double hxi = g.hx[a], rhx = g.rhx[a];
    // east face (FS.cpp:232-237)
    double cEsq, cEuv;
    {
      const double u1 = uC + 0.5 * hxi * sxCu, v1 = vC + 0.5 * hxi * sxCv;
      double u2, v2;
      if (fE) { const double h = 0.5 * g.hx[a + 1]; u2 = uP - h * sxPu; v2 = vP - h * sxPv; }
      else { const GhostUV gh = ghost_uv(g, edges, 1, uC, vC); u2 = 0.5 * (uC + gh.u); v2 = 0.5 * (vC + gh.v); }
      cEsq = flux_sq(u1, u2); cEuv = flux_uv(u1, v1, u2, v2);
    }
    if (i < ib) { cWsq = cEsq; cWuv = cEuv; continue; }   // warm-up column: only its east face is needed

    // y neighbours of (i, j) from the adjacent lanes
    const double uS = shfl_up1(uC), uN = shfl_dn1(uC), vS = shfl_up1(vC), vN = shfl_dn1(vC);
    double syu = 0.0, syv = 0.0;
    if (in_dom) { const GhostUV sl = slope_uv(g, edges, fN, fS, 3, 2, uC, vC, uN, vN, uS, vS, 2.0 * rsyp, 2.0 * rsym, rhy); syu = sl.u; syv = sl.v; }
    const double syNu = shfl_dn1(syu), syNv = shfl_dn1(syv);
    // north face (FS.cpp:246-251); the south face is the north face of the lane below
    double cNsq, cNuv;
    {
      const double u1 = uC + 0.5 * hyj * syu, v1 = vC + 0.5 * hyj * syv;
      double u2, v2;
      if (fN) { u2 = uN - hyN * syNu; v2 = vN - hyN * syNv; }
      else { const GhostUV gh = ghost_uv(g, edges, 3, uC, vC); u2 = 0.5 * (uC + gh.u); v2 = 0.5 * (vC + gh.v); }
      cNsq = flux_sq(v1, v2); cNuv = flux_uv(u1, v1, u2, v2);
    }
    double cSsq = shfl_up1(cNsq), cSuv = shfl_up1(cNuv);
    if (!fS) {   // south side is an edge (FS.cpp:239-244)
      const double u2 = uC - 0.5 * hyj * syu, v2 = vC - 0.5 * hyj * syv;
      const GhostUV gh = ghost_uv(g, edges, 2, uC, vC);
      const double u1 = 0.5 * (uC + gh.u), v1 = 0.5 * (vC + gh.v);
      cSsq = flux_sq(v1, v2); cSuv = flux_uv(u1, v1, u2, v2);
    }
    if (!fW) {   // west side is an edge (FS.cpp:225-230)
      const double u2 = uC - 0.5 * hxi * sxCu, v2 = vC - 0.5 * hxi * sxCv;
      const GhostUV gh = ghost_uv(g, edges, 0, uC, vC);
      const double u1 = 0.5 * (uC + gh.u), v1 = 0.5 * (vC + gh.v);
      cWsq = flux_sq(u1, u2); cWuv = flux_uv(u1, v1, u2, v2);
    }
    if (store) {
      // DiffusiveFlux (FS.cpp:183-203)
      double DuW, DuE, DuS, DuN, DvW, DvE, DvS, DvN;
      if (fW) { const double s = ire * g.rsx[a - 1]; DuW = s * (uC - uM); DvW = s * (vC - vM); }
      else { const double s = 0.5 * ire * rhx; const GhostUV gh = ghost_uv(g, edges, 0, uC, vC); DuW = s * (uC - gh.u); DvW = s * (vC - gh.v); }
      if (fE) { const double s = ire * g.rsx[a]; DuE = s * (uP - uC); DvE = s * (vP - vC); }
      else { const double s = -0.5 * ire * rhx; const GhostUV gh = ghost_uv(g, edges, 1, uC, vC); DuE = s * (uC - gh.u); DvE = s * (vC - gh.v); }
      if (fS) { const double s = ire * rsym; DuS = s * (uC - uS); DvS = s * (vC - vS); }
      else { const double s = 0.5 * ire * rhy; const GhostUV gh = ghost_uv(g, edges, 2, uC, vC); DuS = s * (uC - gh.u); DvS = s * (vC - gh.v); }
      if (fN) { const double s = ire * rsyp; DuN = s * (uN - uC); DvN = s * (vN - vC); }
      else { const double s = -0.5 * ire * rhy; const GhostUV gh = ghost_uv(g, edges, 3, uC, vC); DuN = s * (uC - gh.u); DvN = s * (vC - gh.v); }
      const double diff_u = sc.dt * ((DuE - DuW) * rhx + (DuN - DuS) * rhy);
      const double diff_v = sc.dt * ((DvE - DvW) * rhx + (DvN - DvS) * rhy);
      const size_t id = (size_t)a * ld + row;
      const double nu = (cEsq - cWsq) * rhx + (cNuv - cSuv) * rhy;   // FS.cpp:352
      const double nv = (cEuv - cWuv) * rhx + (cNsq - cSsq) * rhy;   // FS.cpp:356
      double ru = uC + 0.5 * sc.dt * hu + diff_u - 1.5 * sc.dt * nu;   // FS.cpp:335-359
      double rv = vC + 0.5 * sc.dt * hv + diff_v - 1.5 * sc.dt * nv;
      if (!(fW && fE && fS && fN)) {
        const unsigned f = (fW ? F_W : (g.side_out[0] ? F_OW : 0u)) | (fE ? F_E : (g.side_out[1] ? F_OE : 0u)) |
                           (fS ? F_S : (g.side_out[2] ? F_OS : 0u)) | (fN ? F_N : (g.side_out[3] ? F_ON : 0u));
        rhs_boundary_terms<true>(g, edges, sc, f, i, j, gpx + id, gpy + id, ru, rv);
      }
      rhs_u[id] = ru;
      rhs_v[id] = rv;
      nrm[0] = fma(ru, ru, nrm[0]);
      nrm[1] = fma(rv, rv, nrm[1]);
      nu0[id] = nu;
      nv0[id] = nv;
    }
    cWsq = cEsq; cWuv = cEuv;
  }
  if (norms) grid_reduce<2, OpSum>(nrm, red, norms);   // the velocity solves' |b|^2 (uniform branch)
}

}  // namespace nsgpu
