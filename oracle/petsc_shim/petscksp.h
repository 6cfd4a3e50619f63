/* TEST INFRASTRUCTURE ONLY -- KSP/PC part of the PETSc stand-in (see petsc.h).
 * Call sites: FS.cpp:61-82 (configuration), FS.cpp:54, FS.cpp:547-551 (KSPSolve x3). */
#ifndef NSORACLE_PETSCKSP_SHIM_H
#define NSORACLE_PETSCKSP_SHIM_H
#include "petsc.h"
extern "C" {
PetscErrorCode KSPCreate(MPI_Comm comm, KSP* ksp);
PetscErrorCode KSPSetOperators(KSP ksp, Mat A, Mat P);
PetscErrorCode KSPSetType(KSP ksp, KSPType t);
PetscErrorCode KSPGetPC(KSP ksp, PC* pc);
PetscErrorCode PCSetType(PC pc, PCType t);
PetscErrorCode KSPSetTolerances(KSP ksp, PetscReal rtol, PetscReal abstol, PetscReal dtol, PetscInt maxits);
PetscErrorCode KSPSetUp(KSP ksp);
PetscErrorCode KSPSetInitialGuessNonzero(KSP ksp, PetscBool flg);
PetscErrorCode KSPSolve(KSP ksp, Vec b, Vec x);
}
#endif
