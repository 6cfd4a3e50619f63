/* TEST INFRASTRUCTURE ONLY -- per-function access to the UNMODIFIED reference classes.
 *
 * This translation unit is compiled with -Dprivate=public together with the reference's own
 * Grid.cpp / FluidSolver.cpp (taken from /root/reference/SRC where they lie, never copied into
 * this repository) and the PETSc stand-in, into oracle/_ref/libnsref.so.  It lets the parity
 * tests drive the reference's private phases one at a time (ConstructRHS_V, ConstructRHS_phi,
 * CorrectVelocities, the three KSPSolve calls) and read every Vec at full precision.
 * nsref_step() replays the body of FluidSolver::Solve() (FS.cpp:546-557) phase by phase.
 */
#include "FluidSolver.h"
#include "petscksp.h"
#include <iostream>
#include <chrono>

extern "C" {
int shim_vec_size(Vec v);
double* shim_vec_data(Vec v);
int shim_mat_csr(Mat A, int* n, int* nnz, const int** rowptr, const int** col, const double** val);
void shim_ksp_stats(KSP ksp, int* its, int* matvecs, double* rnorm0, double* rnorm, int* converged,
                    double* seconds_total, long* solves, long* iterations_total);
void shim_ksp_set_rtol(KSP ksp, double rtol);
}

struct RefHandle {
  Grid* grid = nullptr;
  FluidSolver* fs = nullptr;
  Vec ustar = nullptr, vstar = nullptr;
  double t_rhs = 0, t_vel = 0, t_div = 0, t_phi = 0, t_corr = 0;
};

static double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

static Vec pick(RefHandle* h, int which) {
  FluidSolver* f = h->fs;
  switch (which) {
    case 0: return f->prevField->u;
    case 1: return f->prevField->v;
    case 2: return f->prevField->phi;
    case 3: return f->RHS_u;
    case 4: return f->RHS_v;
    case 5: return f->RHS_phi;
    case 6: return f->convectiveDer_u0;
    case 7: return f->convectiveDer_v0;
    case 10: return h->ustar;
    case 11: return h->vstar;
  }
  return nullptr;
}

extern "C" {

void* nsref_create(const char* gridfile, const char* simfile, int quiet) {
  std::streambuf* old = nullptr;
  std::ostringstream sink;
  if (quiet) old = std::cout.rdbuf(sink.rdbuf());
  RefHandle* h = new RefHandle;
  h->grid = new Grid(const_cast<char*>(gridfile));
  if (h->grid->setup) h->fs = new FluidSolver(const_cast<char*>(simfile), h->grid);
  if (quiet) std::cout.rdbuf(old);
  if (!h->grid->setup || !h->fs || !h->fs->setup) return nullptr;
  VecCreateSeq(PETSC_COMM_SELF, h->grid->N, &h->ustar);
  VecCreateSeq(PETSC_COMM_SELF, h->grid->N, &h->vstar);
  return h;
}

/* Grid only (the reference Grid has no PETSc dependency); for host-Grid parity tests. */
void* nsref_grid_create(const char* gridfile, int quiet) {
  std::streambuf* old = nullptr;
  std::ostringstream sink;
  if (quiet) old = std::cout.rdbuf(sink.rdbuf());
  RefHandle* h = new RefHandle;
  h->grid = new Grid(const_cast<char*>(gridfile));
  if (quiet) std::cout.rdbuf(old);
  return h;
}
int nsref_grid_setup(void* p) { return ((RefHandle*)p)->grid->setup ? 1 : 0; }

int nsref_n(void* p) { return ((RefHandle*)p)->grid->N; }
int nsref_nx(void* p) { return (int)((RefHandle*)p)->grid->hx.size(); }
int nsref_ny(void* p) { return (int)((RefHandle*)p)->grid->hy.size(); }
int nsref_n_edges(void* p) { return (int)((RefHandle*)p)->grid->edges.size(); }

void nsref_grid_arrays(void* p, double* hx, double* hy, int* cell_id, int* cell_edges, double* xc, double* yc) {
  Grid* g = ((RefHandle*)p)->grid;
  int nx = (int)g->hx.size(), ny = (int)g->hy.size();
  for (int i = 0; i < nx; i++) hx[i] = g->hx[i];
  for (int j = 0; j < ny; j++) hy[j] = g->hy[j];
  for (int i = 0; i < nx; i++)
    for (int j = 0; j < ny; j++) {
      const Cell& c = g->cells[i][j];
      cell_id[i * ny + j] = c.id;
      for (int k = 0; k < 4; k++) cell_edges[(i * ny + j) * 4 + k] = c.edges[k];
      if (xc) xc[i * ny + j] = c.x;
      if (yc) yc[i * ny + j] = c.y;
    }
}
void nsref_edges(void* p, int* enx, int* eny, int* type, double* info, double* loc) {
  Grid* g = ((RefHandle*)p)->grid;
  for (size_t k = 0; k < g->edges.size(); k++) {
    enx[k] = g->edges[k].nx; eny[k] = g->edges[k].ny; type[k] = g->edges[k].bcType; info[k] = g->edges[k].bcInfo;
    if (loc) for (int q = 0; q < 3; q++) loc[3 * k + q] = g->edges[k].loc[q];
  }
}
double nsref_param(void* p, int which) {
  FluidSolver* f = ((RefHandle*)p)->fs;
  switch (which) { case 0: return f->dt; case 1: return f->finalTime; case 2: return f->re; case 3: return f->saveIter; }
  return 0.0;
}

int nsref_get(void* p, int which, double* out) {
  RefHandle* h = (RefHandle*)p;
  int n = h->grid->N;
  if (which == 8 || which == 9) { for (int i = 0; i < n; i++) out[i] = h->fs->divPhi[i][which - 8]; return 0; }
  Vec v = pick(h, which);
  if (!v) return 1;
  std::memcpy(out, shim_vec_data(v), sizeof(double) * n);
  return 0;
}
int nsref_set(void* p, int which, const double* in) {
  RefHandle* h = (RefHandle*)p;
  int n = h->grid->N;
  if (which == 8 || which == 9) { for (int i = 0; i < n; i++) h->fs->divPhi[i][which - 8] = in[i]; return 0; }
  Vec v = pick(h, which);
  if (!v) return 1;
  std::memcpy(shim_vec_data(v), in, sizeof(double) * n);
  return 0;
}

void nsref_construct_rhs_v(void* p) { ((RefHandle*)p)->fs->ConstructRHS_V(); }                       /* FS.cpp:546 */
void nsref_solve_velocity(void* p) {                                                                 /* FS.cpp:547-548 */
  RefHandle* h = (RefHandle*)p;
  KSPSolve(h->fs->uSolver, h->fs->RHS_u, h->ustar);
  KSPSolve(h->fs->uSolver, h->fs->RHS_v, h->vstar);
}
void nsref_construct_rhs_phi(void* p) { RefHandle* h = (RefHandle*)p; h->fs->ConstructRHS_phi(h->ustar, h->vstar); }  /* FS.cpp:549 */
void nsref_nullspace_remove(void* p) { RefHandle* h = (RefHandle*)p; MatNullSpaceRemove(h->fs->NSP, h->fs->RHS_phi); }   /* FS.cpp:550 */
void nsref_solve_phi(void* p) { RefHandle* h = (RefHandle*)p; KSPSolve(h->fs->phiSolver, h->fs->RHS_phi, h->fs->prevField->phi); } /* FS.cpp:551 */
void nsref_correct(void* p) { RefHandle* h = (RefHandle*)p; h->fs->CorrectVelocities(h->ustar, h->vstar); }              /* FS.cpp:552 */

/* n_steps passes of the Solve() loop body; minmax = {umin, umax, vmin, vmax} of the last step (FS.cpp:554-557). */
void nsref_step(void* p, int n_steps, double* minmax) {
  RefHandle* h = (RefHandle*)p;
  for (int s = 0; s < n_steps; s++) {
    double t0 = now_s();
    nsref_construct_rhs_v(p);
    double t1 = now_s();
    nsref_solve_velocity(p);
    double t2 = now_s();
    nsref_construct_rhs_phi(p);
    nsref_nullspace_remove(p);
    double t3 = now_s();
    nsref_solve_phi(p);
    double t4 = now_s();
    nsref_correct(p);
    double t5 = now_s();
    h->t_rhs += t1 - t0; h->t_vel += t2 - t1; h->t_div += t3 - t2; h->t_phi += t4 - t3; h->t_corr += t5 - t4;
  }
  if (minmax) {
    VecMin(h->fs->prevField->u, NULL, &minmax[0]);
    VecMax(h->fs->prevField->u, NULL, &minmax[1]);
    VecMin(h->fs->prevField->v, NULL, &minmax[2]);
    VecMax(h->fs->prevField->v, NULL, &minmax[3]);
  }
}
void nsref_phase_seconds(void* p, double* out5) {
  RefHandle* h = (RefHandle*)p;
  out5[0] = h->t_rhs; out5[1] = h->t_vel; out5[2] = h->t_div; out5[3] = h->t_phi; out5[4] = h->t_corr;
}

/* assembled operators built by the reference's own ConstructLHS: which 0 = LHS_V, 1 = LHS_phi */
int nsref_mat_nnz(void* p, int which) {
  RefHandle* h = (RefHandle*)p;
  int n, nnz; const int *rp, *c; const double* v;
  if (shim_mat_csr(which == 0 ? h->fs->LHS_V : h->fs->LHS_phi, &n, &nnz, &rp, &c, &v)) return -1;
  return nnz;
}
int nsref_mat_csr(void* p, int which, int* rowptr, int* col, double* val) {
  RefHandle* h = (RefHandle*)p;
  int n, nnz; const int *rp, *c; const double* v;
  if (shim_mat_csr(which == 0 ? h->fs->LHS_V : h->fs->LHS_phi, &n, &nnz, &rp, &c, &v)) return 1;
  std::memcpy(rowptr, rp, sizeof(int) * (n + 1));
  std::memcpy(col, c, sizeof(int) * nnz);
  std::memcpy(val, v, sizeof(double) * nnz);
  return 0;
}
/* which 0 = uSolver, 1 = phiSolver */
void nsref_ksp_stats(void* p, int which, int* its, int* matvecs, double* rnorm0, double* rnorm, int* converged,
                     double* seconds_total, long* solves, long* iterations_total) {
  RefHandle* h = (RefHandle*)p;
  shim_ksp_stats(which == 0 ? h->fs->uSolver : h->fs->phiSolver, its, matvecs, rnorm0, rnorm, converged, seconds_total,
                 solves, iterations_total);
}
void nsref_set_rtol(void* p, double rtol) {
  RefHandle* h = (RefHandle*)p;
  shim_ksp_set_rtol(h->fs->uSolver, rtol);
  shim_ksp_set_rtol(h->fs->phiSolver, rtol);
}
void nsref_export(void* p, int iter) { ((RefHandle*)p)->fs->ExportData(iter); }                        /* FS.cpp:569-615 */
double nsref_minmode(double a, double b) { return minmode(a, b); }                                     /* FS.cpp:671-674 */

}  /* extern "C" */
