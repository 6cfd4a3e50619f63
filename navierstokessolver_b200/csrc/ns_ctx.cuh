// ns_ctx.cuh -- device-side state of one FluidSolver (host bookkeeping of libnsgpu.so).
#pragma once
#include <cuda_runtime.h>
#include <nccl.h>
#include <nvtx3/nvToolsExt.h>   // header-only: ranges per phase of the time step (visible in nsys / ncu --nvtx; no-ops without a tool attached)
#include <string>
#include <vector>
#include <unordered_map>
#include <cstdio>
#include <cstring>
#include <cmath>
#include <unistd.h>
#include "../../include/nsgpu.h"
#include "ns_kernels.cuh"
#include "ns_phys.cuh"
#include "ns_rhs.cuh"
#include "ns_march.cuh"
#include "ns_march_ring.cuh"
#include "ns_tail.cuh"
#include "ns_ctail.cuh"
#include "ns_hostgeom.h"

namespace nsgpu {

struct Level {
  LevelGeom g{};
  int cx = 1, cy = 1;        // coarsening factors towards the next (coarser) level
  int gny = 0;
  size_t stride = 0;         // doubles per field on this level: (nx + 2HX) * ld
  double* x = nullptr;       // levels >= 1: correction, ping-pong partner, right-hand side
  double* xt = nullptr;
  double* b = nullptr;
  std::vector<void*> owned;  // device allocations of this level (geometry arrays)
  int coarse_sweeps = 0;     // only used on the coarsest level
  bool replicated = false;   // several ranks: every rank holds the whole level (coarse tail), no halo exchange
  const double* mcx[2] = {nullptr, nullptr};   // folded coefficient tables of k_sweep_march, per operator
  const double* mcy[2] = {nullptr, nullptr};
  bool march_ok[2] = {false, false};
};

struct SolveResult { int cycles = 0; double rel = 0.0; bool converged = false; bool pending = false; };

// One linear solve as a CUDA graph: head (|b|^2, first residual) -> WHILE(not converged){ V-cycle } -> tail.
struct SolveGraph {
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t exec = nullptr;
  const double *x = nullptr, *xt = nullptr, *b = nullptr, *x0 = nullptr;   // buffers the graph was built for
  bool bn2_pre = false;                                      // the head takes |b|^2 from d_scal instead of reducing b
  double rtol = 0;
  int launches_fixed = 0, launches_body = 0;                 // kernels outside / inside the loop body
  int head_iterations = 0;                                   // iterations the head runs itself (counted in launches_fixed)
  int k0 = 1;                                                // ... of which the first k0 - 1 carry no stopping test
};

}  // namespace nsgpu

struct nsgpu_ctx {
  std::string err;
  int device = 0, rank = 0, world = 1;
  bool rect = true;
  double dt = 0, re = 0;
  double rtol_u = 1e-8, rtol_phi = 1e-8;
  int max_cycles = 200, nu_pre = 2, nu_post = 2, solver = 0, use_graph = 0;
  int march = 1, march_ri = 0, march_pf = 6, march_wps = 24;   // NSGPU_MARCH=0 forces the generic tile kernel; NSGPU_RI overrides the chunk length
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  cudaEvent_t evp[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // phase boundaries of the last step of a call
  float phase_ms[5] = {0, 0, 0, 0, 0};
  ncclComm_t comm = nullptr;

  // global description
  int gnx = 0, ny = 0;
  int i0 = 0, i1 = 0;           // owned columns
  int64_t id0 = 0, n_local = 0; // compact ids owned by this rank
  int64_t n_cells_global = 0;
  std::vector<int32_t> local_id;  // general domains: compact local id per (i - i0, j) or -1

  std::vector<nsgpu::Level> levels;
  nsgpu::EdgeDev* d_edges = nullptr;

  // fine-level fields (each `stride` doubles)
  double *u = nullptr, *v = nullptr, *phi = nullptr;
  double *conv_u = nullptr, *conv_v = nullptr, *gpx = nullptr, *gpy = nullptr;
  double *rhs_u = nullptr, *rhs_v = nullptr, *rhs_phi = nullptr, *ustar = nullptr, *vstar = nullptr;
  double* tmp = nullptr;        // ping-pong partner of the fine-level smoother
  double* kry[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};  // BiCGStab vectors (solver == 1, or aug)
  // outflow (bcType 4) domains: the phi solve reproduces the fixed point of the reference's ILU(0)-preconditioned, null-space-
  // projected KSP:  A phi + c w = b, mean(phi) = 0, w = ILU(0)(LHS_phi) 1 (ns_hostgeom.h, phi_ilu0_image_of_ones)
  bool aug = false;
  double* aug_w = nullptr;      // w in the field layout (zeros in padding and masked cells)
  double* aug_t = nullptr;      // scratch vector of the preconditioner
  double aug_ww = 0.0;          // (w, w) over the global domain
  // no outflow edge but a stretched grid: LHS_phi x = b is inconsistent too (the reference removes the arithmetic mean, the left null
  // vector is the cell volumes) and the reference returns A x + c w = b, c = (vol, b) / (vol, w): phase_phi solves with b - c w
  bool aug_rhs = false;
  double aug_yw = 0.0;          // (vol, w) over the global domain
  bool krylov_phi = false;      // the phi solve runs BiCGStab: aug, or a grid that cannot be coarsened (one level: no multigrid to keep)

  nsgpu::Reducer red{};
  double* d_scal = nullptr;     // device scalars (64 doubles)
  double* h_scal = nullptr;     // pinned mirror
  double* h_stage = nullptr;    // host staging for general-domain field transfers
  // device-resident convergence control, one block per solve slot (0 = u, 1 = v, 2 = phi)
  nsgpu::SolveCtl* d_ctl = nullptr;
  nsgpu::SolveCtl* h_ctl = nullptr;       // pinned mirror
  int seen_cycles[3] = {0, 0, 0}, seen_fails[3] = {0, 0, 0};
  nsgpu::SolveGraph sgraph[3];
  nsgpu::SolveGraph kgraph[3];            // Krylov solves (BiCGStab): begin -> WHILE { iteration } -> true residual; x0 holds the operator / aug key
  int graph_cycles_pending[3] = {0, 0, 0};
  // Stopping tests that cannot succeed are not run: a velocity solve that took n iterations in the last time step runs its first n - 1
  // without the residual reduction (the iterate does not depend on the test; NSGPU_SKIP_CHECKS=0: test after every iteration)
  int skip_checks = 1, solve_k0[3] = {1, 1, 1}, solve_hist[3][4] = {}, solve_hist_n[3] = {0, 0, 0};
  // several ranks: every buffer a marching pass writes lives in one arena, mapped by the two neighbours through CUDA IPC,
  // so that the pass can store its boundary columns straight into their halo columns (ns_march.cuh)
  char* arena = nullptr;                 // device allocation (owned)
  size_t arena_bytes = 0, arena_used = 0;
  char* peers[nsgpu::MAX_RANKS] = {};                   // every other rank's arena in this process' address space (CUDA IPC)
  bool peer_inproc[nsgpu::MAX_RANKS] = {};              // that rank lives in this process: its arena is reached by peer access, not IPC
  char *peerL = nullptr, *peerR = nullptr;       // the neighbours' (aliases into peers[])
  nsgpu::PeerTable* d_pt = nullptr;                     // device copy of the peer map (nullptr: one rank, or NCCL-only exchange)
  bool capturing = false, capture_nccl = false;  // a solve graph is being captured / an NCCL call was issued meanwhile
  // NSGPU_TRACE=1 (direct launches): CUDA events around every marching pass / coarse tail, table on stderr at destroy
  struct TraceEv { cudaEvent_t a, b; int key; };
  int trace = 0;
  std::vector<TraceEv> trace_ev;
  std::unordered_map<int, std::pair<double, long>> trace_acc;
  cudaStream_t s_in = nullptr, s_out = nullptr;  // copy streams of nsgpu_step_host (created on first use)
  cudaEvent_t ev_io[6] = {};
  std::vector<cudaEvent_t> ev_chunk;            // nsgpu_step_host, pipelined form: upload / ConstructRHS_V events per column chunk
  int io_chunks = 8;                             // NSGPU_IO_CHUNKS (1: no pipelining)
  int fuse_bn2 = 1;                              // NSGPU_FUSE_BN2=0: |b|^2 by a separate reduction pass over b
  bool bn2_ready[3] = {false, false, false};     // |b|^2 of the next solve of slot k is already in d_scal[30 + k] (written with b)
  double* tail_img[2] = {nullptr, nullptr};      // per operator: the launch-invariant part of the coarse tail's shared memory (k_tail_image)
  long long* tail_dbg = nullptr;                 // NSGPU_TAIL_DBG=1: clock64 stamps of the coarse tail's phases (printed at destroy)
  int phys_march = 1;                            // ConstructRHS_phi / CorrectVelocities on rectangles through k_div_march / k_correct_march (NSGPU_PHYS_MARCH=0: row kernels)
  int rhs_march = 1;                             // ConstructRHS_V on rectangles through k_rhs_march (NSGPU_RHS_MARCH=0: cell kernel)
  int vel_guess_rhs = 0;                         // initial guess of the u*, v* solves: 0 = u^n, v^n; 1 = the right-hand side (LHS_V = I - dt/(2Re) L is close to I)
  int march_rb_min_nx = 0;                       // levels with fewer columns keep uniform chunks (NSGPU_RB_MINNX)
  int ctail_c = 8;                               // CTAs of the cluster that runs the coarse tail (NSGPU_CTAIL_C)
  int march_ring = 0;                            // fused passes without peer stores through the shared-memory ring kernels (NSGPU_MARCH_RING)
  int stamp_key = -1;                            // stamps builds: trace key of the marching pass whose CTA stamps are kept (NSGPU_STAMP_KEY; -2: every pass)
  int field_skew = 0, alloc_k = 0;               // dev_zeros: the k-th large array starts k x field_skew bytes later (NSGPU_FIELD_SKEW)
  double rb_cost = 1.5, rb_cost_up = 1.5;        // pushing loop / plain loop, per column (NSGPU_RB_COST, NSGPU_RB_COST_UP: passes that prolong and reduce)
  int march_rb = 0;                              // width of the short boundary chunks of pushing passes (NSGPU_RB, 0 = off)
  unsigned last_gather_ctas = 0;                 // CTAs of the last restriction pass that gathered onto the replicated level
  unsigned long long* selfpush_hdr = nullptr;
  int selfpush = 0;                              // PROFILING ONLY (NSGPU_SELFPUSH=1, one rank): run the pushing kernel variants with this rank as its own neighbour
  int* d_errflag = nullptr;
  int use_p2p = 1;
  int hx_push = 0;                               // halo exchanges outside the solves through peer stores (k_halo_push) instead of NCCL send / recv
  std::unordered_map<const double*, int> halo_w; // valid halo width of a buffer since its last write
  // coarse tail (ns_tail.cuh): levels >= tail_start run in one single-CTA kernel; 0 = no tail
  int tail_start = 0, use_tail = 1;
  double vel_rho = 0.0, vel_omega = 0.0;   // Jacobi radius bound of LHS_V; SOR factor (0: velocity solves use multigrid)
  nsgpu::TailDesc tail[2];
  // cluster tail (ns_ctail.cuh): levels >= ctail_start run in one 8-CTA cluster kernel (takes precedence over the one-CTA tail); 0 = off
  int ctail_start = 0, use_ctail = 1;
  nsgpu::CtDesc ctail[2];
  double* ctail_img[2] = {nullptr, nullptr};
  // non-blocking export (nsgpu_export_begin / _wait): a device snapshot of u, v, phi, Pr, L_phi phi taken on the solver's stream, brought
  // home by a copy stream into one of two pinned buffers while the time loop goes on
  double* exp_dev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};   // device staging: u, v, phi, Pr, lapPhi
  double* exp_host[2] = {nullptr, nullptr};      // 5 * n_host doubles each (n_host = n_local, or the padded stride on general domains)
  std::vector<double> exp_compact[2];            // general domains: compacted copies handed to the caller
  cudaStream_t s_exp = nullptr;
  cudaEvent_t exp_snap = nullptr, exp_home[2] = {nullptr, nullptr};
  cudaEvent_t exp_t0[2] = {nullptr, nullptr}, exp_t1[2] = {nullptr, nullptr};   // around the snapshot kernel (nsgpu_export_last_ms)
  float exp_last_ms = 0.0f;
  int64_t exp_tag[2] = {0, 0};
  int exp_head = 0, exp_pending = 0;             // next slot to fill; snapshots begun and not yet waited for (<= 2)
  // MatSetNullSpace semantics (mean(phi) = 0 after the solve, FS.cpp:143) without a pass of its own: CorrectVelocities, which reads phi anyway, reduces sum(phi)
  // with its four maxima (SolveCtl::xsum of the phi slot); phi stays as it is in memory (only its differences enter CorrectVelocities) and
  // the FIRST pass of the next phi solve subtracts the mean while it loads the warm start.  Whoever needs phi's values (downloads, export,
  // checkpoints, nsgpu_step_host, the host-driven solvers) calls phi_normalize() first.  NSGPU_LAZY_MEAN=0: explicit k_sum + k_sub_mean.
  int lazy_mean = 1;
  bool phi_shift = false;                        // phi in memory still contains its mean
  bool phi_sum_known = false;                    // ... and SolveCtl::xsum of the phi slot holds sum(phi) (written by k_correct_march)
  const double* next_pass_shift = nullptr;       // consumed by the next march_pass (set by the head of the phi solve)
  bool kry_graph_off = false;                    // a Krylov solve could not be captured (several ranks over NCCL, or the driver refused): direct launches from then on
  bool graph_fallback = false;                   // graphs were asked for but a solve had to fall back to direct launches
  int krylov_restarts = 0;                       // BiCGStab restarts from the true residual during the last call
  int64_t launches = 0;
  double fine_sweeps_phi = 0;
  std::vector<void*> owned;
};
