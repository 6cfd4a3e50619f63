// ns_phys.cuh -- the two projection kernels of the time step on rectangles as register-marching passes:
//   k_div_march      ConstructRHS_phi = Div_V(u*, v*) / dt + the sums MatNullSpaceRemove needs   (FS.cpp:365-418, 550)   24 B / cell
//   k_correct_march  CorrectVelocities + GradP + VecMin / VecMax x 4                            (FS.cpp:420-456, 512-534, 554-557)   56 B / cell
// The one-thread-per-row versions (k_div_rhs, k_correct in ns_kernels.cuh; kept for general domains) have one column in flight per
// thread and run at 55-61 % of the HBM peak, latency-bound.  Here, as in ns_march.cuh, a lane owns the pair of rows (j0, j0 + 1) as one
// double2 -- every load and store is 16 bytes per lane, 512 contiguous bytes per warp -- and marches over the columns of its chunk with
// the next PH_D columns of every input already on their way: one lane issues a 512-byte bulk copy (cp.async.bulk, TMA engine) per field
// column into the warp's own ring of column slots in shared memory, completion is signalled through one mbarrier per slot (ns_tma.cuh),
// so the prefetch costs no registers (the register-ring version needed 130 and, held to 96, spilled the prefetched columns -- which
// waits for the loads and is no prefetch at all: 1.3-1.9 us per column step, long-scoreboard bound);
// x neighbours come from a three-column register ring, y neighbours from the pair partner or the adjacent lane (one halo pair per side
// of the warp: 60 useful rows of 64).  Interior cells take a straight-line path whose expressions are div_cell's / grad_cell's
// (ns_cell.cuh); cells that touch an edge go through div_cell / grad_cell themselves after the march (for_edge_cells), so every
// boundary-condition type keeps its one implementation.
#pragma once
#include <cuda_runtime.h>
#include "ns_cell.cuh"
#include "ns_kernels.cuh"
#include "ns_tma.cuh"

namespace nsgpu {

constexpr int PH_WARPS = 4;
constexpr int PH_THREADS = 32 * PH_WARPS;
constexpr int PH_UJ = 60;      // useful rows per warp
#ifndef NS_PH_D
#define NS_PH_D 6
#endif
constexpr int PH_D = NS_PH_D;  // ring depth: columns in flight per field and warp (PH_D + 1 <= SLACK_COLS)
#ifndef NS_PH_MIN_CTAS
#define NS_PH_MIN_CTAS 5
#endif
constexpr int PH_MIN_CTAS = NS_PH_MIN_CTAS;   // resident CTAs per SM the register allocation is held to (5 x 4 warps)
// dynamic shared memory of a CTA: x coefficients of the chunk (3 doubles per column) | PH_WARPS rings of PH_D slots x A fields x 512 B | barriers
__host__ __device__ constexpr size_t phys_smem(int ri, int fields) {
  return (size_t)(((3 * ri * 8 + 127) / 128) * 128) + (size_t)PH_WARPS * PH_D * fields * 512 + (size_t)PH_WARPS * PH_D * 8;
}

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

// Cells that touch an edge are not computed in the marching loop at all: inlined there, the general cell functions (ghost closures of
// every boundary type) were copied into every unrolled step and the loop body outgrew the instruction caches (32 KB: slower than the
// row kernels).  The loop stores interior cells only; afterwards every warp visits the edge cells of its (strip, chunk) -- first / last
// row: one lane per column, first / last column: one lane per row -- through div_cell / grad_cell, once.
struct EdgeWalk {
  int jlo, jhi;   // useful rows of this warp: [jlo, jhi)
};
__device__ __forceinline__ EdgeWalk edge_walk(const LevelGeom& g, int wj) {
  EdgeWalk w;
  w.jlo = wj * PH_UJ;
  w.jhi = min(w.jlo + PH_UJ, g.ny);
  return w;
}
// calls f(i, j) for every edge cell of the chunk [ib, ie) x rows [jlo, jhi), each exactly once, lanes in parallel
template <class F>
__device__ __forceinline__ void for_edge_cells(const LevelGeom& g, int ib, int ie, const EdgeWalk& w, int lane, F f) {
  if (w.jlo >= w.jhi) return;
#pragma unroll 1
  for (int side = 0; side < 2; side++) {                         // first / last row
    const int j = side ? g.ny - 1 : 0;
    if (side && g.ny == 1) break;
    if (j < w.jlo || j >= w.jhi) continue;
#pragma unroll 1
    for (int i = ib + lane; i < ie; i += 32) f(i, j);
  }
#pragma unroll 1
  for (int side = 0; side < 2; side++) {                         // first / last column of the domain, rows 1 .. ny - 2
    const int ig = side ? g.gnx - 1 : 0;
    if (side && g.gnx == 1) break;
    const int i = ig - g.gi0;
    if (i < ib || i >= ie) continue;
#pragma unroll 1
    for (int j = max(w.jlo, 1) + lane; j < min(w.jhi, g.ny - 1); j += 32) f(i, j);
  }
}

// x / d for a divisor that is the same for every cell (dt): q = x * (1/d) corrected by one fused step, q + (x - q d) * (1/d).  With
// 1/d correctly rounded this is the correctly rounded quotient (Markstein), i.e. the bits of the reference's `/ dt` (FS.cpp:374), in
// 3 instructions instead of the ~30 of a general double division (two per column step were a third of the kernel's instructions).
__device__ __forceinline__ double div_invariant(double x, double d, double rd) {
  const double q = x * rd;
  return fma(fma(-q, d, x), rd, q);
}

// A ring slot may be refilled (async-proxy write by the TMA engine) only after EVERY lane's shared-memory loads from it have
// completed.  Issuing the bulk copy right behind the loads is a write-after-read race: LDS results are scoreboarded, not awaited, and
// a bulk copy whose source sits in L2 can land before a queued LDS has read the slot (seen as wrong phi in a few cells per million,
// run-to-run different -- tests/test_gpu_round2.py::test_projection_kernels_random_fields_repeatable).  So the refill sits at the END
// of the step, behind (1) an instruction that consumes the loaded registers -- `sink`, which the kernel's epilogue keeps alive --
// and (2) a warp barrier.
__device__ __forceinline__ void ring_refill_point(int& sink) {
  asm volatile("" : "+r"(sink) :: "memory");
  __syncwarp();
}

// stores the interior cells of a pair: both, only a = (i, j0), or only b = (i, j0 + 1)
__device__ __forceinline__ void st_interior(double* p, const double2& v, bool sa, bool sb) {
  if (sa && sb) *reinterpret_cast<double2*>(p) = v;
  else if (sa) p[0] = v.x;
  else if (sb) p[1] = v.y;
}

// ---------------------------------------------------------------------------------------------------------
// b = Div_V(u*, v*) / dt ; out = {sum b, sum b^2}
__global__ void __launch_bounds__(PH_THREADS, PH_MIN_CTAS)
k_div_march(LevelGeom g, const EdgeDev* __restrict__ edges, StepConst sc, const double* __restrict__ u, const double* __restrict__ v,
            double* __restrict__ b, int RI, Reducer red, double* __restrict__ out, const PeerTable* pt) {
  extern __shared__ __align__(128) unsigned char s_raw[];
  double* s_x = reinterpret_cast<double*>(s_raw);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, wj = blockIdx.x * PH_WARPS + wid;
  const int ib = blockIdx.y * RI, ie = min(ib + RI, g.nx);
  constexpr int A = 2;                                              // fields in a slot: u (column i + 1), v (column i)
  double2* ring = reinterpret_cast<double2*>(s_raw + ((3 * RI * 8 + 127) / 128) * 128) + (size_t)wid * PH_D * A * 32;
  uint64_t* bar = reinterpret_cast<uint64_t*>(s_raw + ((3 * RI * 8 + 127) / 128) * 128 + (size_t)PH_WARPS * PH_D * A * 512) + wid * PH_D;
  phys_stage_x(g, ib, ie, s_x);
  const PhysLane L = phys_lane(g, lane, wj);
  const int ld = L.ld;
  const bool active = wj * PH_UJ < g.ny;                            // warps past the last strip only take part in the reduction
  const bool sa = (L.st2 || L.st1) && L.inA, sb = L.st2 && L.inB;   // interior cells this lane stores
  const double rdt = 1.0 / sc.dt;
  double acc[2] = {0.0, 0.0};
  if (active) {
    // the warp's 64 rows start at row wj * 60 - 2 (+ PJ padding rows): 512 contiguous bytes of every column
    const size_t wrow = (size_t)(wj * PH_UJ - 2 + PJ);
    const double* gu = u + (size_t)(ib + 1 + HX) * ld + wrow;       // column ib + 1 of u, column ib of v: slot of step ib
    const double* gv = v + (size_t)(ib + HX) * ld + wrow;
    if (lane == 0) {
#pragma unroll
      for (int p = 0; p < PH_D; p++) mbar_init(bar + p, 1);
      mbar_init_fence();
#pragma unroll
      for (int p = 0; p < PH_D; p++)
        if (ib + p < ie) {
          mbar_expect_tx(bar + p, A * 512);
          bulk_g2s(ring + (p * A + 0) * 32, gu + (size_t)p * ld, 512, bar + p);
          bulk_g2s(ring + (p * A + 1) * 32, gv + (size_t)p * ld, 512, bar + p);
        }
    }
    __syncwarp();
    int sink = 0;                                                   // see ring_refill_point
    double* pb = b + (size_t)(ib + HX) * ld + L.row;
    double2 um = ld2(u + (size_t)(ib - 1 + HX) * ld + L.row), uc = ld2(u + (size_t)(ib + HX) * ld + L.row);
    uint32_t parity = 0;
#pragma unroll 1
    for (int i0 = ib; i0 < ie; i0 += PH_D, parity ^= 1) {
#pragma unroll
      for (int p = 0; p < PH_D; p++) {
        const int i = i0 + p;
        if (i < ie) {                                                // uniform
          mbar_wait(bar + p, parity);
          const double2 up = ring[(p * A + 0) * 32 + lane], vc = ring[(p * A + 1) * 32 + lane];
          sink ^= __double2hiint(up.x) ^ __double2hiint(vc.x);
          const double vS = __shfl_up_sync(0xffffffffu, vc.y, 1), vN = __shfl_down_sync(0xffffffffu, vc.x, 1);
          const int ig = g.gi0 + i;
          if (ig > 0 && ig < g.gnx - 1) {                            // uniform: first / last column belong to the edge walk
            const double* c3 = s_x + 3 * (i - ib);
            const double rW = c3[0], rE = c3[1], rhx = c3[2];
            double2 d;
            {
              const double VW = um.x * rW + uc.x * (1.0 - rW), VE = up.x * rE + uc.x * (1.0 - rE);
              const double VS = vS * L.rSa + vc.x * (1.0 - L.rSa), VN = vc.y * L.rNa + vc.x * (1.0 - L.rNa);
              d.x = div_invariant((VE - VW) * rhx + (VN - VS) * L.rhya, sc.dt, rdt);
            }
            {
              const double VW = um.y * rW + uc.y * (1.0 - rW), VE = up.y * rE + uc.y * (1.0 - rE);
              const double VS = vc.x * L.rSb + vc.y * (1.0 - L.rSb), VN = vN * L.rNb + vc.y * (1.0 - L.rNb);
              d.y = div_invariant((VE - VW) * rhx + (VN - VS) * L.rhyb, sc.dt, rdt);
            }
            st_interior(pb, d, sa, sb);
            if (sa) { acc[0] += d.x; acc[1] = fma(d.x, d.x, acc[1]); }
            if (sb) { acc[0] += d.y; acc[1] = fma(d.y, d.y, acc[1]); }
          }
          um = uc; uc = up;
          pb += ld;
          ring_refill_point(sink);
          if (lane == 0 && i + PH_D < ie) {
            mbar_expect_tx(bar + p, A * 512);
            bulk_g2s(ring + (p * A + 0) * 32, gu + (size_t)(i - ib + PH_D) * ld, 512, bar + p);
            bulk_g2s(ring + (p * A + 1) * 32, gv + (size_t)(i - ib + PH_D) * ld, 512, bar + p);
          }
        }
      }
    }
    for_edge_cells(g, ib, ie, edge_walk(g, wj), lane, [&](int i, int j) {
      const size_t id = cidx(g, i, j);
      const double d = div_cell<true>(g, edges, sc, i, j, u + id, v + id);
      b[id] = d;
      acc[0] += d; acc[1] = fma(d, d, acc[1]);
    });
    if (sink == 0x7fc0ffee && RI < 0) acc[0] += 1.0;               // never true (RI > 0): keeps `sink` and with it the consumption of every slot load
  }
  grid_reduce<2, OpSum>(acc, red, out, XSumEpilogue<2>{pt});   // summed over the ranks in the epilogue (peer map)
}

// ---------------------------------------------------------------------------------------------------------
// divPhi = GradP(phi); u = u* - dt divPhi_x; v = v* - dt divPhi_y; out[0..3] = max(-u), max(u), max(-v), max(v) (NaN-propagating)
// phi may carry a constant offset (the mean is removed lazily, ns_solver.inl): only differences of phi enter.
__device__ __forceinline__ void minmax_acc(double (&acc)[5], double un, double vn) {
  const bool bad = !(isfinite(un) && isfinite(vn));   // fmax drops NaNs: map a non-finite value to +inf on all four
  const double inf = 1.0 / 0.0;
  acc[0] = fmax(acc[0], bad ? inf : -un); acc[1] = fmax(acc[1], bad ? inf : un);
  acc[2] = fmax(acc[2], bad ? inf : -vn); acc[3] = fmax(acc[3], bad ? inf : vn);
}

__global__ void __launch_bounds__(PH_THREADS, PH_MIN_CTAS)
k_correct_march(LevelGeom g, StepConst sc, const double* __restrict__ phi, const double* __restrict__ us, const double* __restrict__ vs,
                double* __restrict__ u, double* __restrict__ v, double* __restrict__ gpx, double* __restrict__ gpy, int RI, Reducer red,
                double* __restrict__ out, const PeerTable* pt, double* xsum) {
  extern __shared__ __align__(128) unsigned char s_raw[];
  double* s_x = reinterpret_cast<double*>(s_raw);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, wj = blockIdx.x * PH_WARPS + wid;
  const int ib = blockIdx.y * RI, ie = min(ib + RI, g.nx);
  constexpr int A = 3;                                              // fields in a slot: phi (column i + 1), u*, v* (column i)
  double2* ring = reinterpret_cast<double2*>(s_raw + ((3 * RI * 8 + 127) / 128) * 128) + (size_t)wid * PH_D * A * 32;
  uint64_t* bar = reinterpret_cast<uint64_t*>(s_raw + ((3 * RI * 8 + 127) / 128) * 128 + (size_t)PH_WARPS * PH_D * A * 512) + wid * PH_D;
  phys_stage_x(g, ib, ie, s_x);
  const PhysLane L = phys_lane(g, lane, wj);
  const int ld = L.ld;
  const bool active = wj * PH_UJ < g.ny;
  const bool sa = (L.st2 || L.st1) && L.inA, sb = L.st2 && L.inB;
  double acc[5] = {OpMax::id(), OpMax::id(), OpMax::id(), OpMax::id(), 0.0};   // four maxima and sum(phi) (nsgpu_ctx::lazy_mean)
  if (active) {
    const size_t wrow = (size_t)(wj * PH_UJ - 2 + PJ);
    const double* gp = phi + (size_t)(ib + 1 + HX) * ld + wrow;
    const double* gus = us + (size_t)(ib + HX) * ld + wrow;
    const double* gvs = vs + (size_t)(ib + HX) * ld + wrow;
    if (lane == 0) {
#pragma unroll
      for (int p = 0; p < PH_D; p++) mbar_init(bar + p, 1);
      mbar_init_fence();
#pragma unroll
      for (int p = 0; p < PH_D; p++)
        if (ib + p < ie) {
          mbar_expect_tx(bar + p, A * 512);
          bulk_g2s(ring + (p * A + 0) * 32, gp + (size_t)p * ld, 512, bar + p);
          bulk_g2s(ring + (p * A + 1) * 32, gus + (size_t)p * ld, 512, bar + p);
          bulk_g2s(ring + (p * A + 2) * 32, gvs + (size_t)p * ld, 512, bar + p);
        }
    }
    __syncwarp();
    int sink = 0;                                                   // see ring_refill_point
    size_t o = (size_t)(ib + HX) * ld + L.row;                       // running offset of the output pair
    double2 pm = ld2(phi + o - ld), pc = ld2(phi + o);
    uint32_t parity = 0;
#pragma unroll 1
    for (int i0 = ib; i0 < ie; i0 += PH_D, parity ^= 1) {
#pragma unroll
      for (int p = 0; p < PH_D; p++) {
        const int i = i0 + p;
        if (i < ie) {                                                // uniform
          mbar_wait(bar + p, parity);
          const double2 pn = ring[(p * A + 0) * 32 + lane], usc = ring[(p * A + 1) * 32 + lane], vsc = ring[(p * A + 2) * 32 + lane];
          sink ^= __double2hiint(pn.x) ^ __double2hiint(usc.x) ^ __double2hiint(vsc.x);
          const double pS = __shfl_up_sync(0xffffffffu, pc.y, 1), pN = __shfl_down_sync(0xffffffffu, pc.x, 1);
          const int ig = g.gi0 + i;
          if (ig > 0 && ig < g.gnx - 1) {                            // uniform
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
            const double2 un = make_double2(usc.x - sc.dt * gx.x, usc.y - sc.dt * gx.y);
            const double2 vn = make_double2(vsc.x - sc.dt * gy.x, vsc.y - sc.dt * gy.y);
            st_interior(gpx + o, gx, sa, sb);
            st_interior(gpy + o, gy, sa, sb);
            st_interior(u + o, un, sa, sb);
            st_interior(v + o, vn, sa, sb);
            if (sa) { minmax_acc(acc, un.x, vn.x); acc[4] += pc.x; }
            if (sb) { minmax_acc(acc, un.y, vn.y); acc[4] += pc.y; }
          }
          pm = pc; pc = pn;
          o += ld;
          ring_refill_point(sink);
          if (lane == 0 && i + PH_D < ie) {
            mbar_expect_tx(bar + p, A * 512);
            bulk_g2s(ring + (p * A + 0) * 32, gp + (size_t)(i - ib + PH_D) * ld, 512, bar + p);
            bulk_g2s(ring + (p * A + 1) * 32, gus + (size_t)(i - ib + PH_D) * ld, 512, bar + p);
            bulk_g2s(ring + (p * A + 2) * 32, gvs + (size_t)(i - ib + PH_D) * ld, 512, bar + p);
          }
        }
      }
    }
    for_edge_cells(g, ib, ie, edge_walk(g, wj), lane, [&](int i, int j) {
      const size_t id = cidx(g, i, j);
      const UV gr = grad_cell(g, cell_flags<true>(g, i, j), i, j, phi + id);
      const double un = us[id] - sc.dt * gr.u, vn = vs[id] - sc.dt * gr.v;
      gpx[id] = gr.u; gpy[id] = gr.v; u[id] = un; v[id] = vn;
      minmax_acc(acc, un, vn);
      acc[4] += phi[id];
    });
    if (sink == 0x7fc0ffee && RI < 0) acc[4] += 1.0;               // never true (RI > 0): keeps `sink` alive
  }
  grid_reduce<5, OpMax4Sum>(acc, red, out, CorrectEpilogue{pt, xsum});
}

}  // namespace nsgpu
