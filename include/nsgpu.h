/* nsgpu.h -- C ABI of the B200-native time-step loop (libnsgpu.so).
 *
 * This is the drop-in boundary for the hot path of shivams15/navierstokessolver: everything
 * FluidSolver::Solve() does per time step (reference SRC/FluidSolver.cpp:536-567) runs behind
 * these entry points as hand-written sm_100a kernels.  The reference has no FFI of its own; the
 * seam replaced here is the set of private member functions Solve() calls, so each entry point
 * cites the member function(s) it stands in for.  Plain C types only: caller-owned host arrays
 * are copied at the call and never retained.
 *
 * Conventions
 *   - every function returns 0 on success, a negative nsgpu_status otherwise; the message of the
 *     last failure is available from nsgpu_last_error() (the reference reports errors only by
 *     printing and by its `setup` flag, FS.cpp:17-32 -- the host FluidSolver maps codes to that).
 *   - field arrays cross the boundary in the reference's Vec order: compact cell id,
 *     id = i*ny + j on rectangles (Grid.cpp:149-161), `n_local` doubles per field.  With more
 *     than one rank every rank passes / receives only the ids of its own x-slab.
 *   - a context is not thread-safe (the reference is single-threaded, FS.h:11-13).
 */
#ifndef NSGPU_H
#define NSGPU_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct nsgpu_ctx nsgpu_ctx;

typedef enum {
  NSGPU_OK = 0,
  NSGPU_ERR_ARG = -1,       /* bad descriptor / parameter                                   */
  NSGPU_ERR_BC = -2,        /* bcType 1 (parabolic inlet) or 3 (pressure): unimplemented in  */
                            /* the reference as well (FS.cpp:86-101, UB there; refused here) */
  NSGPU_ERR_CUDA = -3,      /* CUDA runtime failure, no device, or kernel image missing     */
  NSGPU_ERR_NCCL = -4,
  NSGPU_ERR_NOT_CONVERGED = -5, /* a linear solve hit max_cycles (fields are still updated)  */
  NSGPU_ERR_NAN = -6        /* non-finite velocity after the projection                     */
} nsgpu_status;

/* Flattened copy of the reference's Grid public data (Grid.h:58-67). */
typedef struct {
  int32_t nx, ny;            /* bounding box: hx.size(), hy.size()                              */
  int32_t n_cells;           /* Grid::N (Grid.h:58)                                             */
  const double* hx;          /* [nx]  Grid::hx (Grid.h:62)                                      */
  const double* hy;          /* [ny]  Grid::hy                                                  */
  const int32_t* cell_id;    /* [nx*ny] cells[i][j].id at i*ny+j, -1 outside (Grid.h:33);       */
                             /*   NULL = full rectangle, id = i*ny+j                            */
  const int32_t* cell_edges; /* [nx*ny*4] cells[i][j].edges, order -x,+x,-y,+y, -1 = none       */
                             /*   (Grid.h:37); NULL = rectangle whose 4 sides are one edge each */
  int32_t n_edges;
  const int32_t* edge_nx;    /* [n_edges] Edge::nx, Edge::ny (Grid.h:23)                        */
  const int32_t* edge_ny;
  const int32_t* edge_bc_type; /* [n_edges] Edge::bcType: 0 inlet, 2 wall, 4 outflow (Grid.h:10,25) */
  const double* edge_bc_info;  /* [n_edges] Edge::bcInfo (Grid.h:26)                            */
} nsgpu_grid;

typedef struct {
  double dt, re;             /* FluidSolver::dt, re (FS.h:24,26)                                */
  int32_t device;            /* CUDA device ordinal                                             */
  int32_t rank, world;       /* x-slab decomposition: this process owns slab `rank` of `world`  */
  const void* nccl_id;       /* world > 1: the 128 bytes of nsgpu_nccl_unique_id(), same on all ranks */
  double rtol_u, rtol_phi;   /* true-residual relative tolerances; <= 0 selects 1e-8 (FS.cpp:68,80) */
  int32_t max_cycles;        /* per linear solve; <= 0 selects 200                              */
  int32_t nu_pre, nu_post;   /* multigrid smoothing sweeps; <= 0 selects 2                      */
  int32_t solver;            /* 0 = multigrid iteration, 1 = BiCGStab preconditioned by one V-cycle */
  int32_t use_graph;         /* >= 0: each linear solve is one CUDA graph (head -> WHILE(not converged){cycle}) with the */
                             /* stopping test on the device; < 0: direct launches, one host read per cycle             */
  int32_t max_levels;        /* <= 0: coarsen as far as the grid allows                          */
  int32_t vel_solver;        /* LHS_V solves: 0 = auto, 1 = multigrid, 2 = red-black SOR (optimal factor)            */
  int32_t reserved[6];
} nsgpu_params;

typedef struct {
  double umin, umax, vmin, vmax;   /* VecMin/VecMax of the new velocities (FS.cpp:554-557)      */
  int32_t it_u, it_v, it_phi;      /* cycles / iterations used by the three solves              */
  double res_u, res_v, res_phi;    /* final relative true residuals                             */
  double fine_sweeps;              /* fine-level Poisson smoother sweeps executed in this call  */
  float ms_total;                  /* device time of the call, CUDA events on the ctx stream    */
  int32_t kernel_launches;         /* kernels launched by this call                             */
  int32_t status;
  int32_t flags;                   /* NSGPU_FLAG_* bits                                         */
  int32_t krylov_restarts;         /* BiCGStab restarts from the recomputed true residual       */
} nsgpu_step_info;
enum {
  NSGPU_FLAG_GRAPH_FALLBACK = 1,   /* CUDA graphs were requested but the solves ran as direct launches (capture failed, or a  */
                                   /* solve needs NCCL inside its loop): same results, about twice the launch latency        */
  NSGPU_FLAG_KRYLOV_PHI = 2,       /* the phi solve took the Krylov route (outflow / stretched / uncoarsenable grid)         */
  NSGPU_FLAG_P2P = 4               /* several ranks: halo columns travel through NVLink peer memory from inside the kernels  */
};

/* ids for nsgpu_get_array / nsgpu_set_array (FluidSolver private state, FS.h:22-33) */
enum {
  NSGPU_U = 0, NSGPU_V = 1, NSGPU_PHI = 2,           /* prevField->u, v, phi                    */
  NSGPU_RHS_U = 3, NSGPU_RHS_V = 4, NSGPU_RHS_PHI = 5,
  NSGPU_CONV_U0 = 6, NSGPU_CONV_V0 = 7,              /* convectiveDer_u0, convectiveDer_v0      */
  NSGPU_GRADPHI_X = 8, NSGPU_GRADPHI_Y = 9,          /* divPhi[id][0], divPhi[id][1]            */
  NSGPU_USTAR = 10, NSGPU_VSTAR = 11                 /* the local Vec u, v of Solve() (FS.cpp:538) */
};

int nsgpu_nccl_unique_id(void* out128);

/* FluidSolver::SolverInitialize + SolverSetup (FS.cpp:16-58, 84-164): allocates the state, builds the
 * matrix-free form of LHS_V = I - dt/(2Re) L_V and LHS_phi = L_phi and the multigrid hierarchy. */
int nsgpu_create(const nsgpu_grid* grid, const nsgpu_params* params, nsgpu_ctx** out);
void nsgpu_destroy(nsgpu_ctx* ctx);
const char* nsgpu_last_error(const nsgpu_ctx* ctx); /* ctx may be NULL: error of the last failed create */

/* slab owned by this rank: columns [i0, i1) and the compact-id range [id0, id0 + n_local) */
int nsgpu_local_range(const nsgpu_ctx* ctx, int32_t* i0, int32_t* i1, int64_t* id0, int64_t* n_local);

/* n passes of the Solve() loop body, FS.cpp:546-557 (state stays resident on the device) */
int nsgpu_step(nsgpu_ctx* ctx, int32_t n_steps, nsgpu_step_info* last);

/* One time step with the whole state in HOST memory (what a caller that keeps the reference's Vecs on the host has:
 * u, v, phi, convectiveDer_u0/v0, gradphi_x/y, FS.h:22-33; compact-id order, this rank's range; page-locked buffers
 * make the copies asynchronous).  The seven arrays are uploaded, the step runs, the seven arrays are overwritten with
 * the new state.  Copies overlap the phases: phi travels up while ConstructRHS_V and the velocity solves run, the
 * convective terms travel down during the solves, phi during CorrectVelocities.  Rectangular domains.
 * gradphi_x = gradphi_y = NULL: divPhi is not shipped at all -- it is GradP(phi) (FS.cpp:529), so it is rebuilt on the device from
 * the uploaded phi (same arithmetic as CorrectVelocities) and stays there: five arrays up, five down. */
typedef struct nsgpu_host_state {
  double *u, *v, *phi, *conv_u0, *conv_v0, *gradphi_x, *gradphi_y;
} nsgpu_host_state;
int nsgpu_step_host(nsgpu_ctx* ctx, const nsgpu_host_state* io, nsgpu_step_info* info);

/* the phases one at a time (used by the parity tests) */
int nsgpu_rhs_build(nsgpu_ctx* ctx);                           /* ConstructRHS_V            FS.cpp:327-363 (+183-325, 458-510) */
int nsgpu_solve_velocity(nsgpu_ctx* ctx, nsgpu_step_info* info); /* KSPSolve(uSolver) x2      FS.cpp:547-548 */
int nsgpu_div_rhs(nsgpu_ctx* ctx, int32_t remove_mean);        /* ConstructRHS_phi + MatNullSpaceRemove  FS.cpp:365-418, 550 */
int nsgpu_solve_phi(nsgpu_ctx* ctx, nsgpu_step_info* info);    /* KSPSolve(phiSolver)       FS.cpp:551 */
int nsgpu_correct(nsgpu_ctx* ctx, nsgpu_step_info* info);      /* CorrectVelocities + VecMin/VecMax  FS.cpp:512-534, 420-456, 554-557 */

/* n red-black sweeps of the fine-level Poisson smoother on (phi, RHS_phi); the benchmarked kernel.
 * residual_l2 (may be NULL) receives ||RHS_phi - L_phi phi||_2 afterwards, ms (may be NULL) the device
 * time of the sweeps alone (CUDA events on the ctx stream). */
int nsgpu_poisson_sweeps(nsgpu_ctx* ctx, int32_t n_sweeps, double* residual_l2, float* ms);

/* The fine-level passes the time step really runs, timed alone (n back-to-back launches, CUDA events on the ctx stream): kind 0 = plain
 * sweep (as nsgpu_poisson_sweeps), 1 = first pass of a phi V-cycle (two fused sweeps + residual restricted to the next level), 2 = its
 * last pass (coarse correction added while loading, two fused sweeps, |r|^2 reduced), 3 = red-black SOR double sweep of LHS_V on
 * (u*, RHS_u).  Smooths phi / u* in place: for benchmark contexts, not inside a time loop.  Rectangles with >= 2 levels. */
int nsgpu_time_passes(nsgpu_ctx* ctx, int32_t kind, int32_t n, float* ms);

/* y = LHS x for op 0 = LHS_V, 1 = LHS_phi (the operators ConstructLHS assembles, FS.cpp:105-145);
 * x, y are host arrays of n_local doubles. */
int nsgpu_apply_operator(nsgpu_ctx* ctx, int32_t op, const double* x, double* y);

/* physical pressure Pr = phi - dt/(2Re) LHS_phi phi (ExportData, FS.cpp:577-581) and lapPhi */
int nsgpu_pressure(nsgpu_ctx* ctx, double* pr, double* lap_phi);

/* ExportData without stalling the time loop (FS.cpp:561, 569-615).  nsgpu_export_begin takes a snapshot of u, v, phi, Pr and
 * lapPhi = LHS_phi phi on the device (one kernel on the solver's stream) and starts their copy into page-locked host memory on a
 * separate stream; it returns at once, further nsgpu_step calls overlap the copy.  nsgpu_export_wait blocks until the OLDEST
 * pending snapshot is on the host and returns pointers to its five arrays (n_local doubles each, compact-id order; owned by the
 * context, valid until the second nsgpu_export_begin after this snapshot's) and the tag given to begin.  At most two snapshots
 * may be pending.  Any output pointer may be NULL. */
int nsgpu_export_begin(nsgpu_ctx* ctx, int64_t tag);
int nsgpu_export_wait(nsgpu_ctx* ctx, const double** u, const double** v, const double** phi, const double** pr,
                      const double** lap_phi, int64_t* tag);

int nsgpu_export_last_ms(const nsgpu_ctx* ctx, float* ms);   /* device time of the snapshot kernel of the snapshot last waited for */

int nsgpu_get_array(nsgpu_ctx* ctx, int32_t which, double* out);       /* n_local doubles, compact-id order */
int nsgpu_set_array(nsgpu_ctx* ctx, int32_t which, const double* in);
int nsgpu_get_fields(nsgpu_ctx* ctx, double* u, double* v, double* phi);  /* any pointer may be NULL */
int nsgpu_set_fields(nsgpu_ctx* ctx, const double* u, const double* v, const double* phi);

/* Checkpoint / restart.  The reference has none (its state lives and dies with the process, FS.h:22-33); SURVEY.md section 8(f) lists
 * it for the long runs to steady state.  The file holds the complete state of the time loop for this rank's id range -- the seven
 * arrays of nsgpu_host_state -- plus the caller's step counter, dt, re and the warm-start shift of the augmented phi solve of outflow /
 * stretched domains, so a restarted run continues bit for bit on every domain.  Little-endian binary: 88-byte header
 * ("NSGPUCK2", nx, ny, n_cells, id0, n_local, step, dt, re, shift, reserved) then 7 x n_local doubles.  Several ranks: each rank writes
 * "<path>.<rank>".  load refuses a file whose grid extents, id range, dt or re differ from the context's (NSGPU_ERR_ARG). */
int nsgpu_checkpoint_save(nsgpu_ctx* ctx, const char* path, int64_t step);
int nsgpu_checkpoint_load(nsgpu_ctx* ctx, const char* path, int64_t* step);

/* No counterpart in the reference: added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  VecMin / VecMax of two arrays as the time step computes them
 * (FS.cpp:554-557: out4 = umin, umax, vmin, vmax; a non-finite value anywhere gives -inf, +inf, -inf, +inf), with NO allocation of
 * any kind: the staging copies of u and v and the reduction's scratch live in one static device array of fixed size that every GPU
 * holds from the first use of libnsgpu.so until exit (18 MiB).  u, v: host arrays of n doubles, n <= nsgpu_noalloc_max_cells()
 * (2^20; larger n is refused with NSGPU_ERR_ARG, message from nsgpu_last_error(NULL)).  Synchronous, on the legacy default stream of
 * `device`.  All calls share the static array: calls that overlap on one GPU race -- serialise them. */
int nsgpu_minmax_noalloc(int32_t device, const double* u, const double* v, int64_t n, double* out4);
long long nsgpu_noalloc_max_cells(void);

/* introspection for benchmarks and tests */
int nsgpu_num_levels(const nsgpu_ctx* ctx);
/* device milliseconds of the five phases of the last time step run by nsgpu_step: ConstructRHS_V, the two velocity
 * solves, ConstructRHS_phi + mean removal, the phi solve, CorrectVelocities + min/max (FS.cpp:546-557) */
int nsgpu_phase_ms(const nsgpu_ctx* ctx, float* out5);
int64_t nsgpu_kernel_launches(const nsgpu_ctx* ctx);   /* total kernels launched by this ctx so far */
int nsgpu_synchronize(nsgpu_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif
