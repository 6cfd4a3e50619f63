// ns_hostgeom.h -- pure host C++ (no CUDA): turns the flattened Grid descriptor (include/nsgpu.h) into the per-level
// coefficient tables, flag bytes and the coarsening plan the kernels consume.  Restates, matrix-free, what
// ConstructGhostStencils / ConstructLHS set up in the reference (FS.cpp:84-164).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>
#include "../../include/nsgpu.h"
#include "ns_cell.cuh"

namespace nsgpu {

// One level over the GLOBAL domain.
struct HostLevel {
  int gnx = 0, ny = 0;
  std::vector<double> hx, hy;
  std::vector<uint8_t> flags;  // [gnx*ny], empty on the rectangle fast path
};

struct HostPlan {
  bool rect = true;
  uint8_t side_out[4] = {0, 0, 0, 0}, side_edge[4] = {0, 0, 0, 0};
  std::vector<EdgeDev> edges;
  std::vector<uint8_t> cedges0;       // [4*gnx*ny] edge ids, general domains only
  std::vector<HostLevel> levels;
  std::vector<int> lnx, lgi0, lcx, lcy;  // per level: local columns, global index of local column 0, coarsening factors
  int rep_start = 0;                     // several ranks: levels >= rep_start are REPLICATED (every rank holds the whole level); 0 = none
  int64_t n_cells_global = 0;
  int i0 = 0, i1 = 0;
  int64_t id0 = 0, n_local = 0;
  std::vector<int32_t> local_id;      // general domains: compact local id per (i - i0, j), -1 outside
};

// 1-D tables of one level for the local slab [gi0, gi0 + nx), indexed with the halo/padding offsets.
struct GeomArrays {
  int nx = 0, ny = 0, ld = 0, gi0 = 0, gnx = 0;
  size_t stride = 0;
  std::vector<double> hx, hy, rhx, rhy, rsx, rsy, bx, by;
  std::vector<Pair> ax, ay;
  std::vector<uint8_t> flags, cedges;  // padded layout; empty on the rectangle fast path
  // rectangle fast path (k_sweep_march): per operator, 3 doubles per column / row with the boundary rows of
  // LHS_V / LHS_phi and the scaling alpha, beta folded in:  x_new = (b + cw xW + ce xE + cs xS + cn xN) / (dx + dy)
  std::vector<double> mcx[2], mcy[2];
  bool march_ok[2] = {false, false};
};

static inline double clamp_get(const std::vector<double>& a, int k) {
  if (k < 0) k = 0;
  if (k >= (int)a.size()) k = (int)a.size() - 1;
  return a[k];
}

static inline void make_geom_arrays(const HostLevel& H, int gi0, int nx, const std::vector<uint8_t>* cedges_global, GeomArrays& A) {
  const int ny = H.ny;
  A.nx = nx; A.ny = ny; A.gi0 = gi0; A.gnx = H.gnx;
  A.ld = ((ny + 3) / 4) * 4 + 2 * PJ;
  A.stride = (size_t)(nx + 2 * HX + SLACK_COLS) * A.ld;
  const int nxa = nx + 2 * HX, nya = ny + 2 * PJ;
  A.hx.resize(nxa); A.rhx.resize(nxa); A.rsx.resize(nxa); A.bx.resize(nxa); A.ax.resize(nxa);
  A.hy.resize(nya); A.rhy.resize(nya); A.rsy.resize(nya); A.by.resize(nya); A.ay.resize(nya);
  for (int a = 0; a < nxa; a++) {
    const int ig = gi0 + a - HX;
    const double h = clamp_get(H.hx, ig), hm = clamp_get(H.hx, ig - 1), hp = clamp_get(H.hx, ig + 1);
    A.hx[a] = h; A.rhx[a] = 1.0 / h; A.rsx[a] = 1.0 / (h + hp);
    A.ax[a].m = 2.0 / (h * (h + hm)); A.ax[a].p = 2.0 / (h * (h + hp));   // FS.cpp:119
    A.bx[a] = 1.0 / (h * h);                                              // FS.cpp:125
  }
  for (int a = 0; a < nya; a++) {
    const int jg = a - PJ;
    const double h = clamp_get(H.hy, jg), hm = clamp_get(H.hy, jg - 1), hp = clamp_get(H.hy, jg + 1);
    A.hy[a] = h; A.rhy[a] = 1.0 / h; A.rsy[a] = 1.0 / (h + hp);
    A.ay[a].m = 2.0 / (h * (h + hm)); A.ay[a].p = 2.0 / (h * (h + hp));   // FS.cpp:120
    A.by[a] = 1.0 / (h * h);                                              // FS.cpp:126
  }
  A.flags.clear(); A.cedges.clear();
  if (!H.flags.empty()) {
    A.flags.assign(A.stride, (uint8_t)F_OUTSIDE);
    for (int a = 0; a < nxa; a++) {
      const int ig = gi0 + a - HX;
      if (ig < 0 || ig >= H.gnx) continue;
      for (int j = 0; j < ny; j++) A.flags[(size_t)a * A.ld + j + PJ] = H.flags[(size_t)ig * ny + j];
    }
    if (cedges_global) {
      A.cedges.assign(4 * A.stride, (uint8_t)0);
      for (int a = 0; a < nxa; a++) {
        const int ig = gi0 + a - HX;
        if (ig < 0 || ig >= H.gnx) continue;
        for (int j = 0; j < ny; j++)
          for (int k = 0; k < 4; k++) A.cedges[4 * ((size_t)a * A.ld + j + PJ) + k] = (*cedges_global)[4 * ((size_t)ig * ny + j) + k];
      }
    }
  }
}

// Fill a LevelGeom with pointers (host or device copies of the GeomArrays tables).
static inline void geom_scalars(const GeomArrays& A, const HostPlan& P, LevelGeom& g) {
  g.nx = A.nx; g.ny = A.ny; g.ld = A.ld; g.gi0 = A.gi0; g.gnx = A.gnx;
  for (int k = 0; k < 4; k++) { g.side_out[k] = P.side_out[k]; g.side_edge[k] = P.side_edge[k]; }
}

// Folded 1-D coefficient tables of the marching sweep (rectangles only).  op 0: A = I - beta L_V (velocity ghosts:
// Dirichlet-type sides add 2/h^2 to the diagonal, outflow sides drop out, FS.cpp:88,99,147-164); op 1: A = L_phi
// (Neumann sides drop out; the outflow rows of L_phi reach two cells inward and are not separable -> march_ok false).
// Indices outside the global domain get cw = ce = 0 and a dummy diagonal part so that every lane computes finite values.
static inline void make_march_coefs(const HostLevel& H, const HostPlan& P, double dt, double re, GeomArrays& A) {
  if (!P.rect) return;
  const int nxa = A.nx + 2 * HX, nya = A.ny + 2 * PJ;
  for (int op = 0; op < 2; op++) {
    const double alpha = op == 0 ? 1.0 : 0.0, beta = op == 0 ? dt / (2.0 * re) : -1.0;
    const double dummy = op == 0 ? 1.0 : -1.0;
    A.march_ok[op] = !(op == 1 && (P.side_out[0] || P.side_out[1] || P.side_out[2] || P.side_out[3]));
    A.mcx[op].assign((size_t)3 * nxa, 0.0);
    A.mcy[op].assign((size_t)3 * nya, 0.0);
    for (int a = 0; a < nxa; a++) {
      const int ig = A.gi0 + a - HX;
      double cw = 0, ce = 0, d = 0;
      if (ig < 0 || ig >= H.gnx) { A.mcx[op][3 * a + 2] = dummy; continue; }
      if (ig > 0) { cw = beta * A.ax[a].m; d -= A.ax[a].m; } else if (op == 0 && !P.side_out[0]) d -= 2.0 * A.bx[a];
      if (ig < H.gnx - 1) { ce = beta * A.ax[a].p; d -= A.ax[a].p; } else if (op == 0 && !P.side_out[1]) d -= 2.0 * A.bx[a];
      A.mcx[op][3 * a] = cw; A.mcx[op][3 * a + 1] = ce; A.mcx[op][3 * a + 2] = 0.5 * alpha - beta * d;
    }
    for (int a = 0; a < nya; a++) {
      const int jg = a - PJ;
      double cs = 0, cn = 0, d = 0;
      if (jg < 0 || jg >= H.ny) { A.mcy[op][3 * a + 2] = dummy; continue; }
      if (jg > 0) { cs = beta * A.ay[a].m; d -= A.ay[a].m; } else if (op == 0 && !P.side_out[2]) d -= 2.0 * A.by[a];
      if (jg < H.ny - 1) { cn = beta * A.ay[a].p; d -= A.ay[a].p; } else if (op == 0 && !P.side_out[3]) d -= 2.0 * A.by[a];
      A.mcy[op][3 * a] = cs; A.mcy[op][3 * a + 1] = cn; A.mcy[op][3 * a + 2] = 0.5 * alpha - beta * d;
    }
  }
}

static inline double mean_of(const std::vector<double>& a) {
  double s = 0;
  for (double x : a) s += x;
  return s / (double)a.size();
}

// Validates the descriptor and builds edges, flags, the slab of `rank` and the multigrid coarsening plan.
// Returns 0 or a negative nsgpu_status with `err` filled.
// rep_cells: several ranks replicate every level of at most this many GLOBAL cells (never fewer than the coarse tail's 4096)
static inline int make_host_plan(const nsgpu_grid* G, int rank, int world, int max_levels_in, HostPlan& P, std::string& err,
                                 long long rep_cells = 4096) {
  if (!G) { err = "null grid"; return NSGPU_ERR_ARG; }
  if (G->nx < 1 || G->ny < 1 || !G->hx || !G->hy) { err = "bad grid extents"; return NSGPU_ERR_ARG; }
  if (G->n_edges < 1 || G->n_edges > 255 || !G->edge_nx || !G->edge_ny || !G->edge_bc_type || !G->edge_bc_info) {
    err = "edge list missing (1..255 edges)"; return NSGPU_ERR_ARG;
  }
  if ((G->cell_id == nullptr) != (G->cell_edges == nullptr)) { err = "cell_id and cell_edges must both be given or both be NULL"; return NSGPU_ERR_ARG; }
  if (world < 1 || rank < 0 || rank >= world) { err = "rank outside [0, world)"; return NSGPU_ERR_ARG; }
  const int gnx = G->nx, ny = G->ny;

  // ---- edges (ConstructGhostStencils, FS.cpp:84-103)
  P.edges.assign(G->n_edges, EdgeDev{});
  for (int e = 0; e < G->n_edges; e++) {
    const int t = G->edge_bc_type[e];
    EdgeDev& E = P.edges[e];
    E.nx = G->edge_nx[e]; E.ny = G->edge_ny[e]; E.cu = 0; E.cv = 0; E.outflow = 0; E.pad = 0;
    if ((E.nx != 0) == (E.ny != 0)) { err = "edge normal must be axis aligned"; return NSGPU_ERR_ARG; }
    const double c = 2.0 * G->edge_bc_info[e];
    if (t == 0) { if (E.nx == 0) E.cv = c; else E.cu = c; }          // uniform inlet: normal component (FS.cpp:89-92)
    else if (t == 2) { if (E.nx != 0) E.cv = c; else E.cu = c; }     // wall: tangential component (FS.cpp:93-96)
    else if (t == 4) { E.outflow = 1; }                              // zero-gradient outflow (FS.cpp:98-101)
    else {
      char buf[160];
      snprintf(buf, sizeof buf, "edge %d: bcType %d is not implemented by the reference either (FS.cpp:86-101)", e, t);
      err = buf; return NSGPU_ERR_BC;
    }
  }

  // ---- domain shape
  HostLevel H0;
  H0.gnx = gnx; H0.ny = ny;
  H0.hx.assign(G->hx, G->hx + gnx);
  H0.hy.assign(G->hy, G->hy + ny);
  for (double h : H0.hx) if (!(h > 0)) { err = "non-positive hx"; return NSGPU_ERR_ARG; }
  for (double h : H0.hy) if (!(h > 0)) { err = "non-positive hy"; return NSGPU_ERR_ARG; }
  bool full = true;
  if (G->cell_id) for (size_t c = 0; c < (size_t)gnx * ny; c++) if (G->cell_id[c] < 0) { full = false; break; }
  bool rect = full && G->n_edges == 4;
  if (rect) {
    int seen[4] = {0, 0, 0, 0};
    for (int e = 0; e < 4; e++) {
      int k = P.edges[e].nx == -1 ? 0 : P.edges[e].nx == 1 ? 1 : P.edges[e].ny == -1 ? 2 : 3;
      seen[k]++; P.side_edge[k] = (uint8_t)e; P.side_out[k] = (uint8_t)P.edges[e].outflow;
    }
    for (int k = 0; k < 4; k++) if (seen[k] != 1) rect = false;
  }
  if (!rect && !G->cell_id) { err = "cell_id/cell_edges are required unless the domain is a rectangle with one edge per side"; return NSGPU_ERR_ARG; }
  P.rect = rect;
  if (!rect) { for (int k = 0; k < 4; k++) { P.side_out[k] = 0; P.side_edge[k] = 0; } }

  const int dx[4] = {-1, 1, 0, 0}, dy[4] = {0, 0, -1, 1};
  if (G->cell_id) {   // validate the tags even when the fast path will ignore them
    std::vector<uint8_t> fl((size_t)gnx * ny, (uint8_t)F_OUTSIDE);
    P.cedges0.assign((size_t)4 * gnx * ny, 0);
    int64_t count = 0;
    for (int i = 0; i < gnx; i++)
      for (int j = 0; j < ny; j++) {
        const size_t c = (size_t)i * ny + j;
        if (G->cell_id[c] < 0) continue;
        if (G->cell_id[c] != count) { err = "cell ids must be compact and ordered i-major (Grid.cpp:149-161)"; return NSGPU_ERR_ARG; }
        count++;
        unsigned f = 0;
        for (int k = 0; k < 4; k++) {
          const int in = i + dx[k], jn = j + dy[k];
          const bool present = in >= 0 && in < gnx && jn >= 0 && jn < ny && G->cell_id[(size_t)in * ny + jn] >= 0;
          const int e = G->cell_edges[4 * c + k];
          if (present != (e < 0) || e >= G->n_edges) {
            char buf[220];
            snprintf(buf, sizeof buf, "cell (%d,%d) side %d: edge tag %d inconsistent with the neighbour cell (the reference would index out of bounds)", i, j, k, e);
            err = buf; return NSGPU_ERR_ARG;
          }
          if (present) f |= (1u << k);
          else { P.cedges0[4 * c + k] = (uint8_t)e; if (P.edges[e].outflow) f |= (16u << k); }
        }
        fl[c] = (uint8_t)f;
      }
    if (count != G->n_cells) { err = "n_cells does not match cell_id"; return NSGPU_ERR_ARG; }
    P.n_cells_global = count;
    if (!rect) H0.flags.swap(fl); else P.cedges0.clear();
  } else {
    if (G->n_cells != gnx * ny) { err = "n_cells does not match nx*ny"; return NSGPU_ERR_ARG; }
    P.n_cells_global = (int64_t)gnx * ny;
  }

  // ---- slab of this rank
  if (gnx % world != 0) { err = "nx must be divisible by the number of ranks"; return NSGPU_ERR_ARG; }
  const int nxl = gnx / world;
  if (world > 1 && nxl < 2) { err = "slab thinner than the 2-column halo"; return NSGPU_ERR_ARG; }
  P.i0 = rank * nxl; P.i1 = P.i0 + nxl;
  if (rect) { P.id0 = (int64_t)P.i0 * ny; P.n_local = (int64_t)nxl * ny; }
  else {
    P.local_id.assign((size_t)nxl * ny, -1);
    int64_t before = 0, mine = 0;
    for (int i = 0; i < gnx; i++)
      for (int j = 0; j < ny; j++) {
        if (G->cell_id[(size_t)i * ny + j] < 0) continue;
        if (i < P.i0) before++;
        else if (i < P.i1) P.local_id[(size_t)(i - P.i0) * ny + j] = (int32_t)mine++;
      }
    P.id0 = before; P.n_local = mine;
  }

  // ---- level hierarchy (rediscretisation: coarse h = sum of the children's h)
  P.levels.clear();
  P.levels.push_back(std::move(H0));
  P.lnx.assign(1, nxl); P.lgi0.assign(1, P.i0); P.lcx.clear(); P.lcy.clear();
  const int max_levels = max_levels_in > 0 ? max_levels_in : 64;
  while ((int)P.levels.size() < max_levels) {
    const HostLevel& F = P.levels.back();
    const int nx_l = P.lnx.back();
    const double mx = mean_of(F.hx), my = mean_of(F.hy);
    // Several ranks on a plain rectangle: slabs keep at least HX columns (the fused marching passes take their whole
    // halo from ONE neighbour); as soon as a level is small (rep_cells; at the latest when it fits the single-CTA coarse
    // tail) it is replicated on every rank (all-gather of its right-hand side): on small levels a pass costs launch and
    // neighbour latency, not bandwidth, so redundant work is cheaper than one more synchronisation per pass.
    const bool fast = rect && world > 1 && !(P.side_out[0] || P.side_out[1] || P.side_out[2] || P.side_out[3]);
    const bool replicated = P.rep_start > 0;
    const int min_nx = world == 1 || replicated ? 1 : (fast ? HX : 2);
    const bool can_x = (nx_l % 2 == 0) && (nx_l / 2 >= min_nx) && (F.gnx / 2 >= 2);
    const bool can_y = (F.ny % 2 == 0) && (F.ny / 2 >= 2);
    // semi-coarsen only the strongly coupled (finer) direction while the cells are stretched
    const bool want_x = mx <= 1.6 * my, want_y = my <= 1.6 * mx;
    const int cx = (can_x && want_x) ? 2 : 1, cy = (can_y && want_y) ? 2 : 1;
    if (cx == 1 && cy == 1) break;
    HostLevel C;
    C.gnx = F.gnx / cx; C.ny = F.ny / cy;
    C.hx.resize(C.gnx); C.hy.resize(C.ny);
    for (int I = 0; I < C.gnx; I++) { double s = 0; for (int d = 0; d < cx; d++) s += F.hx[I * cx + d]; C.hx[I] = s; }
    for (int J = 0; J < C.ny; J++) { double s = 0; for (int d = 0; d < cy; d++) s += F.hy[J * cy + d]; C.hy[J] = s; }
    if (!F.flags.empty()) {
      std::vector<uint8_t> in((size_t)C.gnx * C.ny, 0), oc((size_t)C.gnx * C.ny, 0);
      for (int I = 0; I < C.gnx; I++)
        for (int J = 0; J < C.ny; J++)
          for (int di = 0; di < cx; di++)
            for (int dj = 0; dj < cy; dj++) {
              const uint8_t f = F.flags[(size_t)(I * cx + di) * F.ny + J * cy + dj];
              if (f != F_OUTSIDE) { in[(size_t)I * C.ny + J] = 1; oc[(size_t)I * C.ny + J] |= (f & 0xF0); }
            }
      C.flags.assign((size_t)C.gnx * C.ny, (uint8_t)F_OUTSIDE);
      for (int I = 0; I < C.gnx; I++)
        for (int J = 0; J < C.ny; J++) {
          if (!in[(size_t)I * C.ny + J]) continue;
          unsigned f = 0;
          for (int k = 0; k < 4; k++) {
            const int In = I + dx[k], Jn = J + dy[k];
            const bool present = In >= 0 && In < C.gnx && Jn >= 0 && Jn < C.ny && in[(size_t)In * C.ny + Jn];
            if (present) f |= (1u << k);
            else if (oc[(size_t)I * C.ny + J] & (16u << k)) f |= (16u << k);
          }
          C.flags[(size_t)I * C.ny + J] = (uint8_t)f;
        }
    }
    P.lcx.push_back(cx); P.lcy.push_back(cy);
    if (fast && !replicated && (long long)C.gnx * C.ny <= std::max(rep_cells, 4096LL)) P.rep_start = (int)P.levels.size();   // 4096: TAIL_MAX_CELLS
    if (P.rep_start > 0) { P.lnx.push_back(C.gnx); P.lgi0.push_back(0); }
    else { P.lnx.push_back(nx_l / cx); P.lgi0.push_back(P.lgi0.back() / cx); }
    P.levels.push_back(std::move(C));
  }
  P.lcx.push_back(1); P.lcy.push_back(1);
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// Outflow (bcType 4) domains: what the reference's Poisson solve converges to.
//
// With a type-4 edge LHS_phi keeps the constants as right null space (zero row sums) but its left null vector is no
// longer constant (the 2.5/-2/0.5 ghost rows, FS.cpp:100), and the right-hand side the reference builds is not orthogonal to it:
// LHS_phi x = b has NO solution.  KSPSolve still "converges" (FS.cpp:551): with MatSetNullSpace the left-preconditioned
// iteration works on P M^-1 (b - A x), P = removal of the mean, M = the ILU(0) factors of A in natural ordering
// (PCBJACOBI with one block, FS.cpp:78), and its fixed point is
//        A x + c (M 1) = b,   mean(x) = 0        (x, c unknown; measured against the oracle: 8e-11)
// So the reference's answer is defined by w = M 1 = L (U 1).  This function restates ILU(0) of LHS_phi on the host, once per
// context, O(N): rows in compact-id order (Grid.cpp:149-161), columns sorted, the pattern of ConstructLHS + AddGhostStencils
// (FS.cpp:105-164: cell, present neighbours, and the second inward cell of an outflow side).  Returns w in global id order.
static inline bool plan_has_outflow(const HostPlan& P) {
  for (const EdgeDev& e : P.edges) if (e.outflow) return true;
  return false;
}

// When is LHS_phi x = b consistent?  Without an outflow edge the left null vector of LHS_phi is the cell volumes (hx_i hy_j times row c is
// symmetric), and the reference removes the ARITHMETIC mean from b (MatNullSpaceRemove, FS.cpp:550).  On a uniform grid that is the
// compatibility condition, on a stretched grid it is not: a wall-refined lid-driven cavity has |vol.b| / (|vol| |b|) ~ 1e-4, and the
// reference's KSP converges to the same bordered system as on outflow domains, A x + c w = b with c = (vol, b) / (vol, w) known in closed
// form.  So: outflow edge -> Krylov on the augmented operator; no outflow, stretched grid -> b - c w, then the ordinary multigrid;
// no outflow, uniform grid -> c = 0, nothing to do.
static inline bool plan_is_uniform(const HostPlan& P) {
  const HostLevel& H = P.levels[0];
  for (double h : H.hx) if (std::fabs(h - H.hx[0]) > 1e-13 * H.hx[0]) return false;
  for (double h : H.hy) if (std::fabs(h - H.hy[0]) > 1e-13 * H.hy[0]) return false;
  return true;
}

// max(h) / min(h), the larger of the two directions.  Point-smoothed multigrid with rediscretised coarse operators degrades quickly with it
// (host emulation, wall-refined 64 x 64 cavity: 7 V(2,2) cycles at 1.8, 30 at 3.4, divergence at 6), BiCGStab preconditioned by the same
// V-cycle does not (10 iterations at 3.4, 17 at 6, 37 at 19).
static inline double plan_stretch(const HostPlan& P) {
  const HostLevel& H = P.levels[0];
  const auto mmx = std::minmax_element(H.hx.begin(), H.hx.end());
  const auto mmy = std::minmax_element(H.hy.begin(), H.hy.end());
  return std::max(*mmx.second / *mmx.first, *mmy.second / *mmy.first);
}

// (vol, w) over the global domain, w in global compact-id order
static inline double volume_dot(const nsgpu_grid* G, const HostPlan& P, const std::vector<double>& w) {
  const HostLevel& H = P.levels[0];
  double s = 0.0;
  for (int i = 0; i < H.gnx; i++)
    for (int j = 0; j < H.ny; j++) {
      const int64_t id = P.rect ? (int64_t)i * H.ny + j : (int64_t)G->cell_id[(size_t)i * H.ny + j];
      if (id >= 0) s += H.hx[i] * H.hy[j] * w[(size_t)id];
    }
  return s;
}

static inline int phi_ilu0_image_of_ones(const nsgpu_grid* G, const HostPlan& P, std::vector<double>& w, std::string& err) {
  const HostLevel& H = P.levels[0];
  const int gnx = H.gnx, ny = H.ny;
  const int64_t N = P.n_cells_global;
  auto id_of = [&](int i, int j) -> int64_t {
    if (i < 0 || i >= gnx || j < 0 || j >= ny) return -1;
    return P.rect ? (int64_t)i * ny + j : (int64_t)G->cell_id[(size_t)i * ny + j];
  };
  auto flags_of = [&](int i, int j) -> unsigned {
    if (!P.rect) return H.flags[(size_t)i * ny + j];
    unsigned f = 0;
    if (i > 0) f |= F_W; else if (P.side_out[0]) f |= F_OW;
    if (i < gnx - 1) f |= F_E; else if (P.side_out[1]) f |= F_OE;
    if (j > 0) f |= F_S; else if (P.side_out[2]) f |= F_OS;
    if (j < ny - 1) f |= F_N; else if (P.side_out[3]) f |= F_ON;
    return f;
  };
  std::vector<int64_t> rp((size_t)N + 1, 0), diag((size_t)N, 0);
  std::vector<int64_t> col; std::vector<double> val;
  col.reserve((size_t)N * 6); val.reserve((size_t)N * 6);
  const int di[4] = {-1, 1, 0, 0}, dj[4] = {0, 0, -1, 1};
  int64_t row = 0;
  for (int i = 0; i < gnx; i++)
    for (int j = 0; j < ny; j++) {
      const int64_t c = id_of(i, j);
      if (c < 0) continue;
      if (c != row) { err = "internal: compact ids out of order"; return NSGPU_ERR_ARG; }
      const unsigned f = flags_of(i, j);
      const double hx = H.hx[i], hy = H.hy[j];
      int64_t ec[8]; double ev[8]; int ne = 0;
      auto add = [&](int64_t cc, double v) {
        if (cc < 0) return;
        for (int k = 0; k < ne; k++) if (ec[k] == cc) { ev[k] += v; return; }
        ec[ne] = cc; ev[ne] = v; ne++;
      };
      double d = 0.0;
      for (int k = 0; k < 4; k++) {
        const bool xdir = k < 2;
        const double h = xdir ? hx : hy;
        if (f & (1u << k)) {
          const double hn = xdir ? H.hx[i + di[k]] : H.hy[j + dj[k]];
          const double a = 2.0 / (h * (h + hn));                      // FS.cpp:119-120
          add(id_of(i + di[k], j + dj[k]), a); d -= a;
        } else if (f & (16u << k)) {                                  // ghost = 2.5 c - 2 (c - n) + 0.5 (c - 2n), FS.cpp:100
          const double b = 1.0 / (h * h);                             // FS.cpp:125-126
          d += 1.5 * b;
          add(id_of(i - di[k], j - dj[k]), -2.0 * b);
          add(id_of(i - 2 * di[k], j - 2 * dj[k]), 0.5 * b);
        }
      }
      add(c, d);
      // sort the (at most 7) entries by column
      for (int a = 1; a < ne; a++) { int64_t kc = ec[a]; double kv = ev[a]; int b2 = a - 1; while (b2 >= 0 && ec[b2] > kc) { ec[b2 + 1] = ec[b2]; ev[b2 + 1] = ev[b2]; b2--; } ec[b2 + 1] = kc; ev[b2 + 1] = kv; }
      for (int a = 0; a < ne; a++) { if (ec[a] == c) diag[(size_t)row] = (int64_t)col.size(); col.push_back(ec[a]); val.push_back(ev[a]); }
      row++;
      rp[(size_t)row] = (int64_t)col.size();
    }
  if (row != N) { err = "internal: cell count mismatch in the ILU(0) setup"; return NSGPU_ERR_ARG; }
  // ILU(0), IKJ order
  for (int64_t i = 0; i < N; i++) {
    for (int64_t kk = rp[(size_t)i]; kk < diag[(size_t)i]; kk++) {
      const int64_t k = col[(size_t)kk];
      const double piv = val[(size_t)diag[(size_t)k]];
      const double lik = val[(size_t)kk] / piv;
      val[(size_t)kk] = lik;
      for (int64_t jj = diag[(size_t)k] + 1; jj < rp[(size_t)k + 1]; jj++) {
        const int64_t jc = col[(size_t)jj];
        for (int64_t q = kk + 1; q < rp[(size_t)i + 1]; q++)
          if (col[(size_t)q] == jc) { val[(size_t)q] -= lik * val[(size_t)jj]; break; }
      }
    }
    const double piv = val[(size_t)diag[(size_t)i]];
    if (!(piv == piv) || piv == 0.0 || std::isinf(piv)) { err = "ILU(0) of LHS_phi broke down (zero or non-finite pivot)"; return NSGPU_ERR_ARG; }
  }
  // w = L (U 1)
  std::vector<double> t((size_t)N);
  for (int64_t i = 0; i < N; i++) { double s = 0; for (int64_t q = diag[(size_t)i]; q < rp[(size_t)i + 1]; q++) s += val[(size_t)q]; t[(size_t)i] = s; }
  w.assign((size_t)N, 0.0);
  for (int64_t i = 0; i < N; i++) { double s = t[(size_t)i]; for (int64_t q = rp[(size_t)i]; q < diag[(size_t)i]; q++) s += val[(size_t)q] * t[(size_t)col[(size_t)q]]; w[(size_t)i] = s; }
  return 0;
}

static inline int coarsest_sweeps(int gnx, int ny) {
  int n = std::min(400, std::max(8, 2 * std::max(gnx, ny)));
  return n + (n & 1);   // even: the ping-pong ends in the buffer it started from
}

}  // namespace nsgpu
