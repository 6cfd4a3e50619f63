// compat/petsc.h -- lets the reference's own MAIN_Solver.cpp (which includes "FluidSolver.h" and calls
// PetscInitialize, reference SRC/MAIN_Solver.cpp:9-10) compile UNMODIFIED against this host library.
// The GPU path does not use PETSc; the call is a no-op that keeps argc/argv untouched.
#ifndef NSB200_COMPAT_PETSC_H
#define NSB200_COMPAT_PETSC_H
typedef int PetscErrorCode;
static inline PetscErrorCode PetscInitialize(int*, char***, const char*, const char*) { return 0; }
#endif
