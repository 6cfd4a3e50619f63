// ns_march_ring.cuh -- the fused marching passes of ns_march.cuh with the columns of x and b arriving through a ring of column slots in
// SHARED memory (one ring per CTA, shared by its four warps), filled by bulk asynchronous copies (TMA engine, cp.async.bulk + mbarrier:
// ns_tma.cuh) instead of a register prefetch ring.  AN EXPERIMENT, off by default (NSGPU_MARCH_RING=1): it gives the same bits as k_march
// and is slower -- 126 us against 91 us for the first pass of a cycle at 4096^2, because it executes 83 M warp instructions instead of
// 49 M (slot arithmetic, six LDS per step, barriers, the producer's turn) and the pass is issue-bound, not bandwidth-bound.
//
// Why.  k_march<2, *, 1|2> carries x AND b of ten columns in registers (three of them only to cover the load latency, b through all five
// stages that use it): 80 of its 156-168 registers, three CTAs of four warps per SM, 12 warps -- ncu: issue slots 50 % busy, 2.9 warps per
// scheduler, stalls `wait` 1.95 and `long_scoreboard` 1.17 per issue, 71 % of the HBM peak (profiles/r2_zz_down_ncu_key.txt).  Here a column of b
// is read from its slot by each of the five stages that need it (one LDS.128 each), x is read once when the column enters the working
// set; a slot is free again once all four warps have left it (an `empty` mbarrier per slot, one arrival per warp) and is refilled, two steps
// later, with the column AHEAD steps in front, so the prefetch distance is a shared-memory budget (MR_D slots of 3.5 KB per CTA), not a
// register budget.  One bulk copy per field and column serves the whole CTA: with one 512-byte copy per warp the copy engine's issue
// rate is the limit (measured: 3.1 TB/s for the chip, 137 us per pass, the warps spinning on their barriers -- 99 M warp instructions).
//
// Same arithmetic, same order of operations as Marcher::step (ns_march.cuh): the two kernels produce the same bits
// (tests/test_gpu_round2.py::test_ring_march_equals_register_march).  No peer stores here: several ranks keep k_march.
#pragma once
#include <cuda_runtime.h>
#include "ns_march.cuh"
#include "ns_tma.cuh"

namespace nsgpu {

constexpr int MR_D = 12;        // column slots per CTA
constexpr int MR_SLACK = 2;     // a slot is refilled this many steps after its last reader may have left it (the warps of a CTA drift)
__host__ __device__ constexpr int mr_rows(int nsw, int res) { return (MARCH_WARPS - 1) * march_uj(nsw, res) + 64; }   // rows of a CTA's window
__host__ __device__ constexpr int mr_field(int nsw, int res) { return (mr_rows(nsw, res) * 8 + 127) / 128 * 128; }     // bytes of one field's column
__host__ __device__ constexpr int mr_slot(int nsw, int res) { return 2 * mr_field(nsw, res); }                         // x column + b column

__host__ __device__ constexpr int march_sc_doubles(int ri, int nsw, int res) { return 3 * (ri + 2 * march_lead(nsw, res) + 2 * nsw + 8); }
__host__ __device__ constexpr size_t march_ring_smem(int ri, int nsw, int res) {
  return ((size_t)march_sc_doubles(ri, nsw, res) * 8 + 127) / 128 * 128 + (size_t)MR_D * mr_slot(nsw, res) + (size_t)2 * MR_D * 8;
}

template <int NSW, bool PROLONG, int RES, bool SOR>
__global__ void __launch_bounds__(MARCH_THREADS)
k_march_ring(LevelGeom g, const double* __restrict__ mcx, const double* __restrict__ mcy, const double* __restrict__ xin,
             const double* __restrict__ b, double* __restrict__ xout, int RI, MarchXfer xf) {
  static_assert(!PROLONG, "the ring variant has no prolongation yet");
  constexpr int L = 2 * NSW;                    // half-sweep stages
  constexpr int K = L + 2 + (RES ? 1 : 0);      // columns in the working set: X[q] = column n - q
  constexpr int AHEAD = MR_D - K + 1 - MR_SLACK;   // at the end of step n the slot of column n - K + 1 - MR_SLACK takes column n + AHEAD
  constexpr int FIELD = mr_field(NSW, RES), SLOT = mr_slot(NSW, RES), WROWS = mr_rows(NSW, RES);
  constexpr int HJ = march_halo(NSW, RES);
  constexpr int LL = march_lead(NSW, RES);
  constexpr int UJ = 64 - 2 * HJ;
  constexpr int HL = HJ / 2;
  static_assert(AHEAD >= 3 && LL + AHEAD + 1 <= HX + SLACK_COLS, "prefetch distance against the slack columns behind a field");
  extern __shared__ __align__(128) unsigned char s_raw[];
  double* s_c = reinterpret_cast<double*>(s_raw);            // (cw, ce, dx) of columns [r0, r0 + nrows)
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int wj = blockIdx.x * MARCH_WARPS + wid;
  const int ib = blockIdx.y * RI, ie = min(ib + RI, g.nx);
  const int nbeg = ib - LL, nend = ie + LL;
  const int n0 = nbeg - ((g.gi0 + nbeg - 1) & 1);
  const int r0 = n0 - 1 - L - (RES ? 1 : 0), nrows = nend + 1 - r0;
  const size_t ring_off = ((size_t)march_sc_doubles(RI, NSW, RES) * 8 + 127) / 128 * 128;
  unsigned char* ring = s_raw + ring_off;                                                        // shared by the four warps of the CTA
  uint64_t* full = reinterpret_cast<uint64_t*>(s_raw + ring_off + (size_t)MR_D * SLOT);          // slot filled (transaction bytes)
  uint64_t* empty = full + MR_D;                                                                 // slot left by all four warps
  for (int q = threadIdx.x; q < nrows; q += MARCH_THREADS) {
    const int a = min(max(r0 + q + HX, 0), g.nx + 2 * HX - 1);
    s_c[3 * q] = mcx[3 * a];
    s_c[3 * q + 1] = mcx[3 * a + 1];
    s_c[3 * q + 2] = mcx[3 * a + 2];
  }
  // the slots start as zeros: the stages of the columns before n0 work on them (results never stored), as the register ring's did
  for (int q = threadIdx.x; q < MR_D * SLOT / 16; q += MARCH_THREADS) reinterpret_cast<double2*>(ring)[q] = make_double2(0.0, 0.0);
  if (threadIdx.x == 0) {
    for (int p = 0; p < MR_D; p++) { mbar_init(full + p, 1); mbar_init(empty + p, MARCH_WARPS); }
    mbar_init_fence();
  }
  __syncthreads();

  // the CTA's window of a column: WROWS rows from jc0 (one bulk copy per field and column for all four warps: 512-byte copies per warp
  // are limited by the copy engine's issue rate, about one per 50-80 cycles and SM = 3 TB/s for the whole chip); the warp's 64 rows start
  // UJ * wid rows into it (16-byte aligned: UJ and jc0 are even).  Windows of the last strips reach past the column's padding into the
  // next column (allocated: SLACK_COLS): those lanes' coefficients are zero, their values are never stored.
  const int jc0 = blockIdx.x * MARCH_WARPS * UJ - HJ;
  const int jw0 = wj * UJ - HJ;
  const int j0 = jw0 + 2 * lane;
  const int jl = min(j0, (g.ny + PJ - 2) & ~1);
  ColConst k;
  {
    const int a = jl + PJ;
    k.csa = mcy[3 * a]; k.cna = mcy[3 * a + 1]; k.dya = mcy[3 * a + 2];
    k.csb = mcy[3 * a + 3]; k.cnb = mcy[3 * a + 4]; k.dyb = mcy[3 * a + 5];
  }
  // CTAs wholly past the domain copy nothing (their window could leave the allocation); their lanes work on the zero slots
  const bool live = jc0 < g.ny + PJ;
  const size_t joff = (size_t)(jl + PJ);
  const bool useful = lane >= HL && lane < 32 - HL;
  const bool jstore2 = useful && (j0 + 1 < g.ny);
  const bool jstore1 = useful && (j0 < g.ny) && !jstore2;
  const double* cc = s_c + 3 * (n0 - r0);
  const size_t wrow = (size_t)(jc0 + PJ);                                    // window start inside a column
  const double* gx = xin ? xin + (size_t)(n0 + HX) * g.ld + wrow : nullptr;   // column n0 of the warp's window
  const double* gb = b + (size_t)(n0 + HX) * g.ld + wrow;
  double* po = xout + (size_t)(n0 - L + HX) * g.ld + joff;
  double* pbc = nullptr;
  if (RES == 1) pbc = xf.bc + (size_t)((xf.rcx == 2 ? (ib >> 1) : ib) + HX) * xf.ldbc + (xf.rcy == 2 ? (size_t)((j0 >> 1) + PJ) : joff);
  double2 racc = make_double2(0.0, 0.0);
  double acc = 0.0;
  const double xshift = (RES == 1 && xf.shift) ? *xf.shift * xf.shift_scale : 0.0;
  const int ld = g.ld, ny = g.ny;
  const uint32_t ring_s = smem_u32(ring) + wid * (UJ * 8) + lane * 16;       // this lane's pair inside slot 0 (x; b is FIELD bytes further)
  double2 X[K];
#pragma unroll
  for (int q = 0; q < K; q++) X[q] = make_double2(0.0, 0.0);

  // Only columns that WILL be waited for are requested: a bulk copy still in flight when the CTA exits would land in the shared memory
  // of whichever CTA gets it next.  The loop below runs over nn in [n0, nlast).
  const int nlast = n0 + ((nend - n0 + 1) & ~1);
  // first fills: columns n0 .. n0 + AHEAD - 1 into slots 0 .. AHEAD - 1 (column n lives in slot (n - n0) mod MR_D)
  if (threadIdx.x == 0 && live) {
    for (int p = 0; p < AHEAD && n0 + p < nlast; p++) {
      mbar_expect_tx(full + p, (gx ? 2 : 1) * WROWS * 8);
      if (gx) bulk_g2s(ring + p * SLOT, gx + (size_t)p * ld, WROWS * 8, full + p);
      bulk_g2s(ring + p * SLOT + FIELD, gb + (size_t)p * ld, WROWS * 8, full + p);
    }
  }
  __syncwarp();
  auto lds2 = [](uint32_t a) {
    double2 v;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(a));
    return v;
  };
  // slot(c) for the column c = n - d, given sl = slot(n): (sl - d) mod MR_D
  auto slot_back = [](int sl, int d) { const int t = sl - d; return t < 0 ? t + MR_D : t; };

  int sl = 0;                 // slot of column n
  uint32_t parity = 0;        // phase of the barrier of slot sl for column n
  int sink = 0;
#pragma unroll 1
  for (int n = n0; n < nend; n += 2) {
#pragma unroll
    for (int h = 0; h < 2; h++) {
      const bool TA = h == 0;                  // step n0 + even: the half-sweeps update the .x cells
      const int nn = n + h;
      // ---- column nn enters the working set
#pragma unroll
      for (int q = K - 1; q > 0; q--) X[q] = X[q - 1];
      if (live) mbar_wait(full + sl, parity);
      X[0] = xin ? lds2(ring_s + sl * SLOT) : make_double2(0.0, 0.0);
      if (RES == 1 && xin) { X[0].x -= xshift; X[0].y -= xshift; }
      // ---- the 2 NSW half-sweeps: stage s updates column nn - 1 - s
#pragma unroll
      for (int s = 0; s < L; s++) {
        const double* c3 = cc - 3 * (1 + s);
        const double2 bb = lds2(ring_s + slot_back(sl, 1 + s) * SLOT + FIELD);
        if (TA) half_update<true, SOR>(X[s + 1], X[s + 2], X[s], bb, c3[0], c3[1], c3[2], k, xf.omega);
        else half_update<false, SOR>(X[s + 1], X[s + 2], X[s], bb, c3[0], c3[1], c3[2], k, xf.omega);
      }
      const int rs = nn - L;                   // this column is final
      if ((unsigned)(rs - ib) < (unsigned)(ie - ib)) {
        if (jstore2) *reinterpret_cast<double2*>(po) = X[L];
        else if (jstore1) *po = X[L].x;
      }
      po += ld;
      if (RES) {
        const int rr = nn - L - 1;
        const double* c3 = cc - 3 * (L + 1);
        const double2 bb = lds2(ring_s + slot_back(sl, L + 1) * SLOT + FIELD);
        const double2 res = pair_residual(X[L + 1], X[L + 2], X[L], bb, c3[0], c3[1], c3[2], k);
        if ((unsigned)(rr - ib) < (unsigned)(ie - ib)) {
          if (RES == 2) {
            acc += (jstore2 || jstore1 ? res.x * res.x : 0.0) + (jstore2 ? res.y * res.y : 0.0);
          } else {
            double2 t = res;
            bool emit = true;
            if (xf.rcx == 2) {
              if (rr & 1) { t.x += racc.x; t.y += racc.y; }
              else { racc = res; emit = false; }
            }
            if (emit) {
              if (xf.rcy == 2) {
                if (useful && j0 < ny) *pbc = (t.x + t.y) * xf.rscale;
              } else {
                const double2 v = make_double2(t.x * xf.rscale, t.y * xf.rscale);
                if (jstore2) *reinterpret_cast<double2*>(pbc) = v;
                else if (jstore1) *pbc = v.x;
              }
              pbc += xf.ldbc;
            }
          }
        }
        sink ^= __double2hiint(bb.x);
      }
      cc += 3;
      // ---- the slot of column nn - K + 1 was read for the last time one step ago (by the last stage that needs b): every lane has
      // consumed that load (`sink`), the warp meets and leaves the slot.  MR_SLACK steps later one warp (they take turns) makes sure all
      // four have left it and lets the copy engine overwrite it with the column AHEAD steps in front of its own.
      asm volatile("" : "+r"(sink) :: "memory");
      __syncwarp();
      if (lane == 0 && live) {
        const int c_left = nn - (K - 1);                                   // column whose slot this warp leaves now
        if (c_left >= n0) {
          uint64_t* e = empty + slot_back(sl, K - 1);
          asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(e)) : "memory");
        }
        const int c_new = nn + AHEAD;                                      // lives in the slot of column c_new - MR_D
        if ((c_new & (MARCH_WARPS - 1)) == wid && c_new < nlast) {
          const int fs = slot_back(sl, K - 1 + MR_SLACK);
          const int c_old = c_new - MR_D;
          if (c_old >= n0) mbar_wait(empty + fs, (uint32_t)(((c_old - n0) / MR_D) & 1));
          const size_t o = (size_t)(c_new - n0) * ld;
          mbar_expect_tx(full + fs, (gx ? 2 : 1) * WROWS * 8);
          if (gx) bulk_g2s(ring + fs * SLOT, gx + o, WROWS * 8, full + fs);
          bulk_g2s(ring + fs * SLOT + FIELD, gb + o, WROWS * 8, full + fs);
        }
      }
      if (++sl == MR_D) { sl = 0; parity ^= 1; }
    }
  }
  if (sink == 0x7fc0ffee && RI < 0) acc += 1.0;   // never true: keeps `sink`, and with it the consumption of the slot loads, alive
  if (RES == 2) {
    double a1[1] = {acc};
    grid_reduce<1, OpSum>(a1, xf.red, xf.out, CtlEpilogue{xf.ctl, xf.pt, xf.out});
  }
}

}  // namespace nsgpu
