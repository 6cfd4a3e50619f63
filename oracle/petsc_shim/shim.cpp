/* TEST INFRASTRUCTURE ONLY -- CPU stand-in for the 35 PETSc entry points the reference calls
 * (SURVEY.md section 8c).  Sequential Vec = plain array, Mat = CSR built from per-row lists,
 * KSPGMRES+PCILU -> ILU(0)-GMRES(30), KSPBCGSL+PCBJACOBI -> ILU(0)-BiCGStab(2) (one block on
 * one process), both left-preconditioned with PETSc's default stopping rule (App. C).
 *
 * Environment knobs (oracle only):
 *   NSREF_RTOL=<x>   override the relative tolerance of every KSP ("exact" mode, e.g. 1e-12)
 *   NSREF_LOG=1      one stderr line per KSPSolve: type, iterations, matvecs, norms, seconds
 */
#include "petscksp.h"
#include "../krylov.h"
#include <chrono>
#include <vector>
#include <utility>

using nsoracle::Csr;
using nsoracle::Ilu0;
using nsoracle::SolveOpts;
using nsoracle::SolveStats;

struct _shim_Vec { int n; double* a; };
struct _shim_NullSpace { bool has_const; };
struct _shim_Mat {
  int n = 0;
  std::vector<std::vector<std::pair<int, double>>> rows; /* before assembly */
  Csr csr;
  bool assembled = false;
  MatNullSpace nsp = nullptr;
};
struct _shim_PC { std::string type = "ilu"; };
struct _shim_KSP {
  Mat A = nullptr;
  std::string type = "gmres";
  _shim_PC pc;
  SolveOpts opts;
  Ilu0 ilu;
  bool is_setup = false;
  SolveStats last;
  double seconds_total = 0.0;
  long solves = 0, iterations_total = 0;
};

extern "C" {

PetscErrorCode PetscInitialize(int*, char***, const char*, const char*) { return 0; }

/* ---- Vec ---- */
PetscErrorCode VecCreateSeq(MPI_Comm, PetscInt n, Vec* v) {
  *v = new _shim_Vec{n, (double*)std::calloc((size_t)(n > 0 ? n : 1), sizeof(double))};
  return 0;
}
PetscErrorCode VecSet(Vec v, PetscScalar a) { for (int i = 0; i < v->n; i++) v->a[i] = a; return 0; }
PetscErrorCode VecAXPY(Vec y, PetscScalar a, Vec x) { for (int i = 0; i < y->n; i++) y->a[i] += a * x->a[i]; return 0; }
PetscErrorCode VecCopy(Vec x, Vec y) { std::memcpy(y->a, x->a, sizeof(double) * x->n); return 0; }
PetscErrorCode VecDestroy(Vec* v) { if (*v) { std::free((*v)->a); delete *v; *v = nullptr; } return 0; }
PetscErrorCode VecGetArray(Vec v, PetscScalar** a) { *a = v->a; return 0; }
PetscErrorCode VecRestoreArray(Vec, PetscScalar** a) { if (a) *a = nullptr; return 0; }
/* The reference never calls VecAssemblyBegin/End, so values land immediately (App. C). */
PetscErrorCode VecSetValues(Vec v, PetscInt n, const PetscInt ix[], const PetscScalar y[], InsertMode mode) {
  for (int k = 0; k < n; k++) {
    if (ix[k] < 0) continue;
    if (mode == ADD_VALUES) v->a[ix[k]] += y[k]; else v->a[ix[k]] = y[k];
  }
  return 0;
}
PetscErrorCode VecMin(Vec v, PetscInt* p, PetscReal* val) {
  int k = 0; for (int i = 1; i < v->n; i++) if (v->a[i] < v->a[k]) k = i;
  if (p) *p = k; if (val) *val = v->a[k]; return 0;
}
PetscErrorCode VecMax(Vec v, PetscInt* p, PetscReal* val) {
  int k = 0; for (int i = 1; i < v->n; i++) if (v->a[i] > v->a[k]) k = i;
  if (p) *p = k; if (val) *val = v->a[k]; return 0;
}

/* ---- Mat ---- */
PetscErrorCode MatCreate(MPI_Comm, Mat* A) { *A = new _shim_Mat; return 0; }
PetscErrorCode MatSetType(Mat, MatType) { return 0; }
PetscErrorCode MatSetSizes(Mat A, PetscInt, PetscInt, PetscInt M, PetscInt) { A->n = M; return 0; }
PetscErrorCode MatSetFromOptions(Mat) { return 0; }
PetscErrorCode MatSetUp(Mat A) { A->rows.assign(A->n, {}); return 0; }
PetscErrorCode MatSetValues(Mat A, PetscInt m, const PetscInt im[], PetscInt n, const PetscInt in[],
                            const PetscScalar v[], InsertMode mode) {
  for (int a = 0; a < m; a++) {
    if (im[a] < 0) continue;
    auto& row = A->rows[im[a]];
    for (int b = 0; b < n; b++) {
      if (in[b] < 0) continue;
      double x = v[a * n + b];
      bool found = false;
      for (auto& e : row) if (e.first == in[b]) { if (mode == ADD_VALUES) e.second += x; else e.second = x; found = true; break; }
      if (!found) row.emplace_back(in[b], x);
    }
  }
  return 0;
}
PetscErrorCode MatAssemblyBegin(Mat, MatAssemblyType) { return 0; }
PetscErrorCode MatAssemblyEnd(Mat A, MatAssemblyType) {
  Csr& c = A->csr;
  c.n = A->n;
  c.rowptr.assign(A->n + 1, 0);
  for (int i = 0; i < A->n; i++) c.rowptr[i + 1] = c.rowptr[i] + (int)A->rows[i].size();
  c.col.resize(c.rowptr[A->n]);
  c.val.resize(c.rowptr[A->n]);
  for (int i = 0; i < A->n; i++) {
    auto& row = A->rows[i];
    std::sort(row.begin(), row.end());
    for (size_t k = 0; k < row.size(); k++) { c.col[c.rowptr[i] + k] = row[k].first; c.val[c.rowptr[i] + k] = row[k].second; }
  }
  c.find_diag();
  A->assembled = true;
  std::vector<std::vector<std::pair<int, double>>>().swap(A->rows);
  return 0;
}
PetscErrorCode MatScale(Mat A, PetscScalar a) { for (auto& x : A->csr.val) x *= a; return 0; }
PetscErrorCode MatShift(Mat A, PetscScalar a) { for (int i = 0; i < A->n; i++) A->csr.val[A->csr.diag[i]] += a; return 0; }
PetscErrorCode MatMult(Mat A, Vec x, Vec y) { A->csr.mult(x->a, y->a); return 0; }
PetscErrorCode MatNullSpaceCreate(MPI_Comm, PetscBool has_cnst, PetscInt, const Vec[], MatNullSpace* sp) {
  *sp = new _shim_NullSpace{has_cnst == PETSC_TRUE}; return 0;
}
PetscErrorCode MatSetNullSpace(Mat A, MatNullSpace sp) { A->nsp = sp; return 0; }
PetscErrorCode MatSetTransposeNullSpace(Mat, MatNullSpace) { return 0; }
/* plain arithmetic mean over ids (FS.cpp:550) */
PetscErrorCode MatNullSpaceRemove(MatNullSpace sp, Vec v) {
  if (!sp || !sp->has_const || v->n == 0) return 0;
  double s = 0.0; for (int i = 0; i < v->n; i++) s += v->a[i];
  s /= v->n; for (int i = 0; i < v->n; i++) v->a[i] -= s;
  return 0;
}

/* ---- KSP ---- */
PetscErrorCode KSPCreate(MPI_Comm, KSP* ksp) { *ksp = new _shim_KSP; return 0; }
PetscErrorCode KSPSetOperators(KSP ksp, Mat A, Mat) { ksp->A = A; ksp->is_setup = false; return 0; }
PetscErrorCode KSPSetType(KSP ksp, KSPType t) { ksp->type = t; return 0; }
PetscErrorCode KSPGetPC(KSP ksp, PC* pc) { *pc = &ksp->pc; return 0; }
PetscErrorCode PCSetType(PC pc, PCType t) { pc->type = t; return 0; }
PetscErrorCode KSPSetTolerances(KSP ksp, PetscReal rtol, PetscReal abstol, PetscReal dtol, PetscInt maxits) {
  if (rtol != (PetscReal)PETSC_DEFAULT) ksp->opts.rtol = rtol;
  if (abstol != (PetscReal)PETSC_DEFAULT) ksp->opts.abstol = abstol;
  if (dtol != (PetscReal)PETSC_DEFAULT) ksp->opts.dtol = dtol;
  if (maxits != PETSC_DEFAULT) ksp->opts.maxits = maxits;
  return 0;
}
PetscErrorCode KSPSetUp(KSP ksp) {
  if (!ksp->A || !ksp->A->assembled) return 1;
  ksp->ilu.factor(ksp->A->csr);
  ksp->is_setup = true;
  return 0;
}
PetscErrorCode KSPSetInitialGuessNonzero(KSP ksp, PetscBool flg) { ksp->opts.nonzero_guess = (flg == PETSC_TRUE); return 0; }
PetscErrorCode KSPSolve(KSP ksp, Vec b, Vec x) {
  if (!ksp->is_setup) KSPSetUp(ksp);
  SolveOpts o = ksp->opts;
  if (const char* e = std::getenv("NSREF_RTOL")) o.rtol = std::atof(e);
  o.const_nullspace = ksp->A->nsp && ksp->A->nsp->has_const;
  auto t0 = std::chrono::steady_clock::now();
  if (ksp->type == "bcgsl") ksp->last = nsoracle::bcgsl_ilu(ksp->A->csr, ksp->ilu, b->a, x->a, o);
  else ksp->last = nsoracle::gmres_ilu(ksp->A->csr, ksp->ilu, b->a, x->a, o);
  double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  ksp->seconds_total += dt; ksp->solves++; ksp->iterations_total += ksp->last.iterations;
  if (std::getenv("NSREF_LOG"))
    std::fprintf(stderr, "[shim] KSPSolve %s n=%d its=%d matvecs=%d |Mb|=%.3e |Mr|=%.3e conv=%d %.4fs\n",
                 ksp->type.c_str(), ksp->A->n, ksp->last.iterations, ksp->last.matvecs, ksp->last.rnorm0,
                 ksp->last.rnorm, ksp->last.converged, dt);
  return 0;
}

/* ---- oracle-side introspection (not PETSc API) ---- */
int shim_vec_size(Vec v) { return v->n; }
double* shim_vec_data(Vec v) { return v->a; }
int shim_mat_csr(Mat A, int* n, int* nnz, const int** rowptr, const int** col, const double** val) {
  if (!A->assembled) return 1;
  *n = A->n; *nnz = (int)A->csr.col.size(); *rowptr = A->csr.rowptr.data(); *col = A->csr.col.data(); *val = A->csr.val.data();
  return 0;
}
void shim_ksp_stats(KSP ksp, int* its, int* matvecs, double* rnorm0, double* rnorm, int* converged,
                    double* seconds_total, long* solves, long* iterations_total) {
  if (its) *its = ksp->last.iterations; if (matvecs) *matvecs = ksp->last.matvecs;
  if (rnorm0) *rnorm0 = ksp->last.rnorm0; if (rnorm) *rnorm = ksp->last.rnorm;
  if (converged) *converged = ksp->last.converged; if (seconds_total) *seconds_total = ksp->seconds_total;
  if (solves) *solves = ksp->solves; if (iterations_total) *iterations_total = ksp->iterations_total;
}
void shim_ksp_set_rtol(KSP ksp, double rtol) { ksp->opts.rtol = rtol; }

}  /* extern "C" */
