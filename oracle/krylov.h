/* TEST INFRASTRUCTURE ONLY -- CPU oracle linear algebra (never linked into the product).
 *
 * CSR storage, ILU(0), left-preconditioned restarted GMRES and BiCGStab(l) with the
 * stopping rule PETSc uses by default (preconditioned residual norm relative to the
 * preconditioned right-hand side).  Shared by the PETSc stand-in
 * (petsc_shim/shim.cpp, which the unmodified reference sources link against).
 *
 * Pinning: the reference ships no tests, golden vectors or sample inputs, so parity is UNPINNED BY THE REFERENCE'S OWN TESTS.
 * What pins this oracle instead: everything outside the three KSPSolve calls is the reference's own code, compiled here from
 * /root/reference/SRC; the three solves are defined mathematically (LHS x = b; for the singular / inconsistent LHS_phi the
 * bordered system A x + c ILU0(A) 1 = b, mean x = 0) and this file's solvers are checked against an independent sparse direct
 * solver on the reference's assembled matrices (tests/test_host_and_oracle.py::test_oracle_linear_solves_against_sparse_direct,
 * ::test_outflow_fixed_point_of_the_reference_solve).  Real PETSc is not installed here; its ILU(0) in natural ordering is the
 * same unique factorisation, its Krylov paths differ from these only within the tolerance.
 */
#ifndef NSORACLE_KRYLOV_H
#define NSORACLE_KRYLOV_H
#include <vector>

namespace nsoracle {

struct Csr {
  int n = 0;
  std::vector<int> rowptr, col;
  std::vector<double> val;
  std::vector<int> diag;  /* index of the diagonal entry of each row */
  void mult(const double* x, double* y) const;
  void find_diag();
};

struct Ilu0 {
  const Csr* A = nullptr;
  std::vector<double> lu; /* same pattern as A: strict lower = L (unit diag), upper incl. diag = U */
  void factor(const Csr& A);
  void solve(const double* b, double* x) const; /* x = (LU)^-1 b; b and x may alias */
};

struct SolveStats {
  int iterations = 0;     /* Krylov iterations (GMRES: inner steps; BCGSL: matvec pairs) */
  int matvecs = 0;
  double rnorm0 = 0.0;    /* preconditioned norm of the rhs used as reference */
  double rnorm = 0.0;     /* final preconditioned residual norm */
  int converged = 0;
};

struct SolveOpts {
  double rtol = 1e-8, abstol = 1e-50, dtol = 1e4;
  int maxits = 10000;
  bool nonzero_guess = false;
  bool const_nullspace = false; /* project the constant out of every preconditioned vector */
  int restart = 30;             /* GMRES */
  int ell = 2;                  /* BiCGStab(l) */
};

SolveStats gmres_ilu(const Csr& A, const Ilu0& M, const double* b, double* x, const SolveOpts& o);
SolveStats bcgsl_ilu(const Csr& A, const Ilu0& M, const double* b, double* x, const SolveOpts& o);

}  // namespace nsoracle
#endif
