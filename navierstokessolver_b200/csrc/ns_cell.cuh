// ns_cell.cuh -- per-cell arithmetic of the time-step loop, shared by every kernel.
//
// Device data layout (all fields, every multigrid level):
//   value(i, j) lives at  (i + HX) * ld + (j + PJ),   i in [-HX, nx + HX),  j in [-PJ, ny + PJ)
//   i = local x index inside this rank's slab (outer, like the reference's cells[i][j]), j = y (inner,
//   unit stride: the reference's id = i*ny + j, Grid.cpp:149-161).  HX = 6 halo columns on each side hold
//   the neighbour slab's cells (or zeros at a domain boundary): wide enough for two fused sweeps plus the
//   fused residual (5 columns, ns_march.cuh); PJ = 8 padding rows keep every strip-halo read in bounds and
//   every column 32-byte aligned (ld is a multiple of 4).
//
// The functions here are the matrix-free restatement of the reference's operators and per-cell physics:
//   row_eval      rows of LHS_V / LHS_phi            FS.cpp:105-164 (ConstructLHS, AddGhostStencils)
//   rhs_cell      ConstructRHS_V for one cell        FS.cpp:183-363, 458-510
//   div_cell      Div_V / ConstructRHS_phi           FS.cpp:365-418
//   grad_cell     GradP                              FS.cpp:420-456
// They are NS_HD so that tests/emul can compile the very same arithmetic for the host and check it against
// the oracle without a GPU; libnsgpu.so itself contains no host execution path for them.
#pragma once
#include <stdint.h>
#include <math.h>

#if defined(__CUDACC__)
#define NS_HD __host__ __device__ __forceinline__
#else
#define NS_HD inline
#endif

namespace nsgpu {

constexpr int HX = 6;
constexpr int PJ = 8;
constexpr int SLACK_COLS = 8;   // allocated (zero) columns behind the last halo column: the marching kernel prefetches past its chunk

// per-cell flag byte: bits 0-3 = neighbour W,E,S,N is a fluid cell; bits 4-7 = that side is an outflow
// (bcType 4) edge, meaningful only where the neighbour is absent; 0xFF = cell outside the domain.
enum : unsigned { F_W = 1, F_E = 2, F_S = 4, F_N = 8, F_OW = 16, F_OE = 32, F_OS = 64, F_ON = 128, F_OUTSIDE = 0xFF };

// Edge::ghost as ConstructGhostStencils builds it (FS.cpp:84-103): ghost_V = w*V_c + c[comp],
// w = -1 (types 0, 2) or +1 (type 4); ghost_P = +phi_c, or the 2.5/-2/0.5 extrapolation on type 4.
struct EdgeDev {
  double cu, cv;     // ghost[0].constant[0], [1]
  int32_t outflow;   // bcType == NEUMANN
  int32_t nx, ny;    // outward normal
  int32_t pad;
};

struct alignas(16) Pair { double m, p; };  // coefficient towards the lower / upper neighbour

struct LevelGeom {
  int32_t nx, ny;     // cells owned by this rank on this level
  int32_t ld;         // column pitch in doubles
  int32_t gi0, gnx;   // global index of local column 0, global column count (for colours and RECT flags)
  const double* hx;   // [nx + 2HX] (index i + HX), [ny + 2PJ] (index j + PJ); replicated past the domain
  const double* hy;
  const double* rhx;  // 1 / hx
  const double* rhy;
  const double* rsx;  // rsx[i] = 1 / (hx[i] + hx[i+1])
  const double* rsy;
  const Pair* ax;     // ax[i] = { 2/(hx_i (hx_i + hx_{i-1})), 2/(hx_i (hx_i + hx_{i+1})) }   FS.cpp:119
  const Pair* ay;
  const double* bx;   // 1 / hx_i^2                                                           FS.cpp:125
  const double* by;
  const uint8_t* flags;   // [ncols * ld] or nullptr (RECT: flags follow from the indices)
  const uint8_t* cedges;  // fine level, general domains: 4 edge ids per cell; nullptr for RECT
  uint8_t side_out[4];    // RECT: side k (W,E,S,N) is an outflow edge
  uint8_t side_edge[4];   // RECT: edge id of side k
};

NS_HD size_t cidx(const LevelGeom& g, int i, int j) { return (size_t)(i + HX) * g.ld + (j + PJ); }

template <bool RECT>
NS_HD unsigned cell_flags(const LevelGeom& g, int i, int j) {
  if (RECT) {
    int ig = g.gi0 + i;
    if (ig < 0 || ig >= g.gnx || j < 0 || j >= g.ny) return F_OUTSIDE;
    unsigned f = 0;
    if (ig > 0) f |= F_W; else if (g.side_out[0]) f |= F_OW;
    if (ig < g.gnx - 1) f |= F_E; else if (g.side_out[1]) f |= F_OE;
    if (j > 0) f |= F_S; else if (g.side_out[2]) f |= F_OS;
    if (j < g.ny - 1) f |= F_N; else if (g.side_out[3]) f |= F_ON;
    return f;
  } else {
    return g.flags[cidx(g, i, j)];
  }
}

template <bool RECT>
NS_HD int side_edge_id(const LevelGeom& g, int i, int j, int k) {
  return RECT ? (int)g.side_edge[k] : (int)g.cedges[4 * cidx(g, i, j) + k];
}

// ---------------------------------------------------------------------------------------------------------
// Operators.  OP 0: L_V rows (velocity ghosts, ghost[0]); OP 1: L_phi rows (ghost[1]); OP 2: L_phi with every
// boundary side treated as a wall (ghost = phi_c) -- never the system that is solved, only the symmetric operator whose
// V-cycle preconditions the Krylov iteration on outflow domains (ns_solver.inl, solve_bicgstab).
//   (L x)_c = diagL * x_c + offL
// x points at the cell itself (so x[+-1] are the y neighbours, x[+-ld] the x neighbours).
// The outflow rows of L_phi reach two cells inward (FS.cpp:100).
template <int OP>
NS_HD void row_eval(const LevelGeom& g, unsigned f, int i, int j, double xW, double xE, double xS, double xN,
                    const double* xfar, double& diagL, double& offL) {
  const Pair ax = g.ax[i + HX];
  const Pair ay = g.ay[j + PJ];
  double d = 0.0, o = 0.0;
  if (f & F_W) { d -= ax.m; o += ax.m * xW; }
  else if (OP == 0) { if (!(f & F_OW)) d -= 2.0 * g.bx[i + HX]; }
  else if (OP == 1 && (f & F_OW)) { double w = g.bx[i + HX]; d += 1.5 * w; o += w * (-2.0 * xE + 0.5 * xfar[2 * g.ld]); }
  if (f & F_E) { d -= ax.p; o += ax.p * xE; }
  else if (OP == 0) { if (!(f & F_OE)) d -= 2.0 * g.bx[i + HX]; }
  else if (OP == 1 && (f & F_OE)) { double w = g.bx[i + HX]; d += 1.5 * w; o += w * (-2.0 * xW + 0.5 * xfar[-2 * g.ld]); }
  if (f & F_S) { d -= ay.m; o += ay.m * xS; }
  else if (OP == 0) { if (!(f & F_OS)) d -= 2.0 * g.by[j + PJ]; }
  else if (OP == 1 && (f & F_OS)) { double w = g.by[j + PJ]; d += 1.5 * w; o += w * (-2.0 * xN + 0.5 * xfar[2]); }
  if (f & F_N) { d -= ay.p; o += ay.p * xN; }
  else if (OP == 0) { if (!(f & F_ON)) d -= 2.0 * g.by[j + PJ]; }
  else if (OP == 1 && (f & F_ON)) { double w = g.by[j + PJ]; d += 1.5 * w; o += w * (-2.0 * xS + 0.5 * xfar[-2]); }
  diagL = d;
  offL = o;
}

// A = alpha*I - beta*L.  LHS_V: alpha = 1, beta = dt/(2Re) (FS.cpp:140-141); LHS_phi: alpha = 0, beta = -1.
template <int OP>
NS_HD double apply_cell(const LevelGeom& g, unsigned f, int i, int j, const double* x, double alpha, double beta) {
  double d, o;
  row_eval<OP>(g, f, i, j, x[-g.ld], x[g.ld], x[-1], x[1], x, d, o);
  return alpha * x[0] - beta * (d * x[0] + o);
}

// Gauss-Seidel value of one cell from its current neighbours.
template <int OP>
NS_HD double relax_cell(const LevelGeom& g, unsigned f, int i, int j, double xW, double xE, double xS, double xN,
                        const double* xfar, double rhs, double alpha, double beta) {
  double d, o;
  row_eval<OP>(g, f, i, j, xW, xE, xS, xN, xfar, d, o);
  const double den = alpha - beta * d;
  return (den != 0.0) ? (rhs + beta * o) / den : 0.0;   // den == 0: isolated all-Neumann cell of a coarse level
}

// ---------------------------------------------------------------------------------------------------------
// Explicit part of the momentum step.

NS_HD double minmod(double a, double b) {  // minmode(), FS.cpp:671-674:  a*min(1,|b/a|) if ab > 0
  if (a * b > 0.0) return (fabs(b) < fabs(a)) ? b : a;
  return 0.0;
}

struct UV { double u, v; };

// MUSCL slope of (u, v) in cell (i, j) along x (SlopeLimiter d == 0, FS.cpp:288-305)
template <bool RECT>
NS_HD UV slope_x(const LevelGeom& g, const EdgeDev* edges, unsigned f, int i, int j, const double* u, const double* v) {
  const double uc = u[0], vc = v[0];
  double au, av, bu, bv;
  if (f & F_E) { double s = 2.0 * g.rsx[i + HX]; au = (u[g.ld] - uc) * s; av = (v[g.ld] - vc) * s; }
  else { const EdgeDev e = edges[side_edge_id<RECT>(g, i, j, 1)]; double w = e.outflow ? 1.0 : -1.0, r = g.rhx[i + HX];
         au = ((w * uc + e.cu) - uc) * r; av = ((w * vc + e.cv) - vc) * r; }
  if (f & F_W) { double s = 2.0 * g.rsx[i - 1 + HX]; bu = (uc - u[-g.ld]) * s; bv = (vc - v[-g.ld]) * s; }
  else { const EdgeDev e = edges[side_edge_id<RECT>(g, i, j, 0)]; double w = e.outflow ? 1.0 : -1.0, r = g.rhx[i + HX];
         bu = (uc - (w * uc + e.cu)) * r; bv = (vc - (w * vc + e.cv)) * r; }
  return UV{minmod(au, bu), minmod(av, bv)};
}

template <bool RECT>
NS_HD UV slope_y(const LevelGeom& g, const EdgeDev* edges, unsigned f, int i, int j, const double* u, const double* v) {
  const double uc = u[0], vc = v[0];
  double au, av, bu, bv;
  if (f & F_N) { double s = 2.0 * g.rsy[j + PJ]; au = (u[1] - uc) * s; av = (v[1] - vc) * s; }
  else { const EdgeDev e = edges[side_edge_id<RECT>(g, i, j, 3)]; double w = e.outflow ? 1.0 : -1.0, r = g.rhy[j + PJ];
         au = ((w * uc + e.cu) - uc) * r; av = ((w * vc + e.cv) - vc) * r; }
  if (f & F_S) { double s = 2.0 * g.rsy[j - 1 + PJ]; bu = (uc - u[-1]) * s; bv = (vc - v[-1]) * s; }
  else { const EdgeDev e = edges[side_edge_id<RECT>(g, i, j, 2)]; double w = e.outflow ? 1.0 : -1.0, r = g.rhy[j + PJ];
         bu = (uc - (w * uc + e.cu)) * r; bv = (vc - (w * vc + e.cv)) * r; }
  return UV{minmod(au, bu), minmod(av, bv)};
}

// Roe-type fluxes of ConvectiveFlux (FS.cpp:227-228, 260-261); state 1 = low side, state 2 = high side
NS_HD double flux_sq(double a1, double a2) { return 0.5 * (a1 * a1 + a2 * a2 - fabs(a1 + a2) * (a2 - a1)); }
NS_HD double flux_uv(double u1, double v1, double u2, double v2) {
  return 0.5 * (u1 * v1 + u2 * v2 - 0.5 * fabs(v1 + v2) * (u2 - u1) - 0.5 * fabs(u2 + u1) * (v2 - v1));
}

struct StepConst { double dt, re; };

struct RhsOut { double rhs_u, rhs_v, nu, nv; };

// ApplyBoundaryConditions (FS.cpp:458-510) for a cell that touches an edge: adds the boundary terms to (ru, rv).
// gpx, gpy point at the cell.
template <bool RECT>
NS_HD void rhs_boundary_terms(const LevelGeom& g, const EdgeDev* edges, StepConst sc, unsigned f, int i, int j,
                              const double* gpx, const double* gpy, double& ru, double& rv) {
  const int ld = g.ld;
  const double hxi = g.hx[i + HX], hyj = g.hy[j + PJ];
  const double ire = 1.0 / sc.re;
  int first = 0;
  while (f & (1u << first)) first++;
  double D;
  if (first < 2) {  // first adjacent edge is vertical: d/dy of dphi/dx
    const double r0 = g.rsy[j - 1 + PJ], r1 = g.rsy[j + PJ];
    if (!(f & F_N)) D = 2.0 * (gpx[0] - gpx[-1]) * r0;
    else if (!(f & F_S)) D = 2.0 * (gpx[1] - gpx[0]) * r1;
    else D = gpx[1] * r1 - gpx[-1] * r0 - gpx[0] * (r1 - r0);
  } else {          // horizontal: d/dx of dphi/dy
    const double r0 = g.rsx[i - 1 + HX], r1 = g.rsx[i + HX];
    if (!(f & F_E)) D = 2.0 * (gpy[0] - gpy[-ld]) * r0;
    else if (!(f & F_W)) D = 2.0 * (gpy[ld] - gpy[0]) * r1;
    else D = gpy[ld] * r1 - gpy[-ld] * r0 - gpy[0] * (r1 - r0);
  }
  for (int k = 0; k < 4; k++) {
    if (f & (1u << k)) continue;
    const EdgeDev e = edges[side_edge_id<RECT>(g, i, j, k)];
    const double n = (k & 1) ? 1.0 : -1.0;
    if (k < 2) {
      const double cf = sc.dt * (0.5 * ire * g.bx[i + HX]);
      if (e.outflow) { rv += cf * (sc.dt * n * hxi * D); }
      else { ru += cf * e.cu; rv += cf * (e.cv + 2.0 * sc.dt * (gpy[0] + n * hxi * D * 0.5)); }
    } else {
      const double cf = sc.dt * (0.5 * ire * g.by[j + PJ]);
      if (e.outflow) { ru += cf * (sc.dt * n * hyj * D); }
      else { rv += cf * e.cv; ru += cf * (e.cu + 2.0 * sc.dt * (gpx[0] + n * hyj * D * 0.5)); }
    }
  }
}

// ConstructRHS_V for cell (i, j).  u, v, gpx, gpy point at the cell; nu0/nv0 = convectiveDer_{u,v}0[id].
template <bool RECT>
NS_HD RhsOut rhs_cell(const LevelGeom& g, const EdgeDev* edges, StepConst sc, int i, int j, const double* u,
                      const double* v, double nu0, double nv0, const double* gpx, const double* gpy) {
  const int ld = g.ld;
  const unsigned f = cell_flags<RECT>(g, i, j);
  const double uc = u[0], vc = v[0];
  const double hxi = g.hx[i + HX], hyj = g.hy[j + PJ], rhx = g.rhx[i + HX], rhy = g.rhy[j + PJ];
  const double ire = 1.0 / sc.re;

  // ghost values on boundary sides (EvaluateGhostStencil_V, FS.cpp:166-173)
  double gu[4], gv[4];
  for (int k = 0; k < 4; k++) { gu[k] = 0.0; gv[k] = 0.0; }
  if ((f & 15u) != 15u) {
    for (int k = 0; k < 4; k++)
      if (!(f & (1u << k))) {
        const EdgeDev e = edges[side_edge_id<RECT>(g, i, j, k)];
        double w = e.outflow ? 1.0 : -1.0;
        gu[k] = w * uc + e.cu;
        gv[k] = w * vc + e.cv;
      }
  }

  // --- DiffusiveFlux (FS.cpp:183-203) for u and v
  double DuW, DuE, DuS, DuN, DvW, DvE, DvS, DvN;
  if (f & F_W) { double s = ire * g.rsx[i - 1 + HX]; DuW = s * (uc - u[-ld]); DvW = s * (vc - v[-ld]); }
  else { double s = 0.5 * ire * rhx; DuW = s * (uc - gu[0]); DvW = s * (vc - gv[0]); }
  if (f & F_E) { double s = ire * g.rsx[i + HX]; DuE = s * (u[ld] - uc); DvE = s * (v[ld] - vc); }
  else { double s = -0.5 * ire * rhx; DuE = s * (uc - gu[1]); DvE = s * (vc - gv[1]); }
  if (f & F_S) { double s = ire * g.rsy[j - 1 + PJ]; DuS = s * (uc - u[-1]); DvS = s * (vc - v[-1]); }
  else { double s = 0.5 * ire * rhy; DuS = s * (uc - gu[2]); DvS = s * (vc - gv[2]); }
  if (f & F_N) { double s = ire * g.rsy[j + PJ]; DuN = s * (u[1] - uc); DvN = s * (v[1] - vc); }
  else { double s = -0.5 * ire * rhy; DuN = s * (uc - gu[3]); DvN = s * (vc - gv[3]); }
  const double diff_u = sc.dt * ((DuE - DuW) * rhx + (DuN - DuS) * rhy);
  const double diff_v = sc.dt * ((DvE - DvW) * rhx + (DvN - DvS) * rhy);

  // --- ConvectiveFlux (FS.cpp:205-281)
  const UV sx = slope_x<RECT>(g, edges, f, i, j, u, v);
  const UV sy = slope_y<RECT>(g, edges, f, i, j, u, v);
  double u1, v1, u2, v2;
  // west face
  u2 = uc - 0.5 * hxi * sx.u; v2 = vc - 0.5 * hxi * sx.v;
  if (f & F_W) {
    UV s = slope_x<RECT>(g, edges, cell_flags<RECT>(g, i - 1, j), i - 1, j, u - ld, v - ld);
    double h = 0.5 * g.hx[i - 1 + HX]; u1 = u[-ld] + h * s.u; v1 = v[-ld] + h * s.v;
  } else { u1 = 0.5 * (uc + gu[0]); v1 = 0.5 * (vc + gv[0]); }
  const double C0 = flux_sq(u1, u2), C1 = flux_uv(u1, v1, u2, v2);
  // east face
  u1 = uc + 0.5 * hxi * sx.u; v1 = vc + 0.5 * hxi * sx.v;
  if (f & F_E) {
    UV s = slope_x<RECT>(g, edges, cell_flags<RECT>(g, i + 1, j), i + 1, j, u + ld, v + ld);
    double h = 0.5 * g.hx[i + 1 + HX]; u2 = u[ld] - h * s.u; v2 = v[ld] - h * s.v;
  } else { u2 = 0.5 * (uc + gu[1]); v2 = 0.5 * (vc + gv[1]); }
  const double C2 = flux_sq(u1, u2), C3 = flux_uv(u1, v1, u2, v2);
  // south face
  u2 = uc - 0.5 * hyj * sy.u; v2 = vc - 0.5 * hyj * sy.v;
  if (f & F_S) {
    UV s = slope_y<RECT>(g, edges, cell_flags<RECT>(g, i, j - 1), i, j - 1, u - 1, v - 1);
    double h = 0.5 * g.hy[j - 1 + PJ]; u1 = u[-1] + h * s.u; v1 = v[-1] + h * s.v;
  } else { u1 = 0.5 * (uc + gu[2]); v1 = 0.5 * (vc + gv[2]); }
  const double C5 = flux_sq(v1, v2), C4 = flux_uv(u1, v1, u2, v2);
  // north face
  u1 = uc + 0.5 * hyj * sy.u; v1 = vc + 0.5 * hyj * sy.v;
  if (f & F_N) {
    UV s = slope_y<RECT>(g, edges, cell_flags<RECT>(g, i, j + 1), i, j + 1, u + 1, v + 1);
    double h = 0.5 * g.hy[j + 1 + PJ]; u2 = u[1] - h * s.u; v2 = v[1] - h * s.v;
  } else { u2 = 0.5 * (uc + gu[3]); v2 = 0.5 * (vc + gv[3]); }
  const double C7 = flux_sq(v1, v2), C6 = flux_uv(u1, v1, u2, v2);

  RhsOut out;
  out.nu = (C2 - C0) * rhx + (C6 - C4) * rhy;   // FS.cpp:352
  out.nv = (C3 - C1) * rhx + (C7 - C5) * rhy;   // FS.cpp:356
  double ru = uc + 0.5 * sc.dt * nu0 + diff_u - 1.5 * sc.dt * out.nu;   // FS.cpp:335-359
  double rv = vc + 0.5 * sc.dt * nv0 + diff_v - 1.5 * sc.dt * out.nv;

  // --- ApplyBoundaryConditions (FS.cpp:458-510): only cells that touch an edge
  if ((f & 15u) != 15u) rhs_boundary_terms<RECT>(g, edges, sc, f, i, j, gpx, gpy, ru, rv);
  out.rhs_u = ru;
  out.rhs_v = rv;
  return out;
}

// Div_V(u*, v*)/dt (FS.cpp:374, 380-418)
template <bool RECT>
NS_HD double div_cell(const LevelGeom& g, const EdgeDev* edges, StepConst sc, int i, int j, const double* u, const double* v) {
  const int ld = g.ld;
  const unsigned f = cell_flags<RECT>(g, i, j);
  const double uc = u[0], vc = v[0], hxi = g.hx[i + HX], hyj = g.hy[j + PJ];
  double VW, VE, VS, VN;
  if (f & F_W) { double r = hxi * g.rsx[i - 1 + HX]; VW = u[-ld] * r + uc * (1.0 - r); }
  else { const EdgeDev e = edges[side_edge_id<RECT>(g, i, j, 0)]; VW = 0.5 * (uc + ((e.outflow ? 1.0 : -1.0) * uc + e.cu)); }
  if (f & F_E) { double r = hxi * g.rsx[i + HX]; VE = u[ld] * r + uc * (1.0 - r); }
  else { const EdgeDev e = edges[side_edge_id<RECT>(g, i, j, 1)]; VE = 0.5 * (uc + ((e.outflow ? 1.0 : -1.0) * uc + e.cu)); }
  if (f & F_S) { double r = hyj * g.rsy[j - 1 + PJ]; VS = v[-1] * r + vc * (1.0 - r); }
  else { const EdgeDev e = edges[side_edge_id<RECT>(g, i, j, 2)]; VS = 0.5 * (vc + ((e.outflow ? 1.0 : -1.0) * vc + e.cv)); }
  if (f & F_N) { double r = hyj * g.rsy[j + PJ]; VN = v[1] * r + vc * (1.0 - r); }
  else { const EdgeDev e = edges[side_edge_id<RECT>(g, i, j, 3)]; VN = 0.5 * (vc + ((e.outflow ? 1.0 : -1.0) * vc + e.cv)); }
  return ((VE - VW) * g.rhx[i + HX] + (VN - VS) * g.rhy[j + PJ]) / sc.dt;
}

// GradP (FS.cpp:420-456); p points at the cell.  ghost_P = phi_c (types 0, 2) or 2.5/-2/0.5 (type 4).
NS_HD UV grad_cell(const LevelGeom& g, unsigned f, int i, int j, const double* p) {
  const int ld = g.ld;
  const double pc = p[0], hxi = g.hx[i + HX], hyj = g.hy[j + PJ];
  double VW, VE, VS, VN;
  if (f & F_W) { double r = hxi * g.rsx[i - 1 + HX]; VW = p[-ld] * r + pc * (1.0 - r); }
  else if (f & F_OW) VW = 0.5 * (pc + (2.5 * pc - 2.0 * p[ld] + 0.5 * p[2 * ld]));
  else VW = 0.5 * (pc + pc);
  if (f & F_E) { double r = hxi * g.rsx[i + HX]; VE = p[ld] * r + pc * (1.0 - r); }
  else if (f & F_OE) VE = 0.5 * (pc + (2.5 * pc - 2.0 * p[-ld] + 0.5 * p[-2 * ld]));
  else VE = 0.5 * (pc + pc);
  if (f & F_S) { double r = hyj * g.rsy[j - 1 + PJ]; VS = p[-1] * r + pc * (1.0 - r); }
  else if (f & F_OS) VS = 0.5 * (pc + (2.5 * pc - 2.0 * p[1] + 0.5 * p[2]));
  else VS = 0.5 * (pc + pc);
  if (f & F_N) { double r = hyj * g.rsy[j + PJ]; VN = p[1] * r + pc * (1.0 - r); }
  else if (f & F_ON) VN = 0.5 * (pc + (2.5 * pc - 2.0 * p[-1] + 0.5 * p[-2]));
  else VN = 0.5 * (pc + pc);
  return UV{(VE - VW) * g.rhx[i + HX], (VN - VS) * g.rhy[j + PJ]};
}

}  // namespace nsgpu
