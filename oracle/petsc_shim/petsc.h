/* TEST INFRASTRUCTURE ONLY -- part of the CPU oracle, never linked into the product.
 *
 * Declarations-only stand-in for the subset of PETSc that the reference
 * (shivams15/navierstokessolver, SRC/FluidSolver.{h,cpp}, SRC/MAIN_Solver.cpp) calls.
 * PETSc itself is not installed in this image, so the unmodified
 * reference sources are compiled against this header and linked with shim.cpp.
 * Call sites covered: FS.h:11-16, FS.cpp:37-43,63-69,75-81,121-144,156,162,335-340,
 * 347-359,375,485-506,526-528,547-557,577-586,618-622, MAIN:10.
 *
 * The reference relies on petsc.h dragging in a number of standard headers
 * (it uses stringstream, pow, abs, min, printf, sprintf without including them).
 */
#ifndef NSORACLE_PETSC_SHIM_H
#define NSORACLE_PETSC_SHIM_H

#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <sstream>
#include <algorithm>
#include <string>

typedef int PetscErrorCode;
typedef int PetscInt;
typedef double PetscScalar;
typedef double PetscReal;
typedef int MPI_Comm;
typedef enum { PETSC_FALSE = 0, PETSC_TRUE = 1 } PetscBool;
typedef enum { NOT_SET_VALUES = 0, INSERT_VALUES = 1, ADD_VALUES = 2 } InsertMode;
typedef enum { MAT_FLUSH_ASSEMBLY = 1, MAT_FINAL_ASSEMBLY = 0 } MatAssemblyType;

#define PETSC_COMM_SELF 1
#define PETSC_DECIDE (-1)
#define PETSC_DEFAULT (-2)

typedef const char* MatType;
typedef const char* KSPType;
typedef const char* PCType;
#define MATAIJ "aij"
#define KSPGMRES "gmres"
#define KSPBCGSL "bcgsl"
#define PCILU "ilu"
#define PCBJACOBI "bjacobi"

typedef struct _shim_Vec* Vec;
typedef struct _shim_Mat* Mat;
typedef struct _shim_KSP* KSP;
typedef struct _shim_PC* PC;
typedef struct _shim_NullSpace* MatNullSpace;

extern "C" {
PetscErrorCode PetscInitialize(int* argc, char*** argv, const char* file, const char* help);

PetscErrorCode VecCreateSeq(MPI_Comm comm, PetscInt n, Vec* v);
PetscErrorCode VecSet(Vec v, PetscScalar a);
PetscErrorCode VecAXPY(Vec y, PetscScalar a, Vec x);
PetscErrorCode VecCopy(Vec x, Vec y);
PetscErrorCode VecDestroy(Vec* v);
PetscErrorCode VecGetArray(Vec v, PetscScalar** a);
PetscErrorCode VecRestoreArray(Vec v, PetscScalar** a);
PetscErrorCode VecSetValues(Vec v, PetscInt n, const PetscInt ix[], const PetscScalar y[], InsertMode mode);
PetscErrorCode VecMin(Vec v, PetscInt* p, PetscReal* val);
PetscErrorCode VecMax(Vec v, PetscInt* p, PetscReal* val);

PetscErrorCode MatCreate(MPI_Comm comm, Mat* A);
PetscErrorCode MatSetType(Mat A, MatType t);
PetscErrorCode MatSetSizes(Mat A, PetscInt m, PetscInt n, PetscInt M, PetscInt N);
PetscErrorCode MatSetFromOptions(Mat A);
PetscErrorCode MatSetUp(Mat A);
PetscErrorCode MatSetValues(Mat A, PetscInt m, const PetscInt im[], PetscInt n, const PetscInt in[],
                            const PetscScalar v[], InsertMode mode);
PetscErrorCode MatAssemblyBegin(Mat A, MatAssemblyType t);
PetscErrorCode MatAssemblyEnd(Mat A, MatAssemblyType t);
PetscErrorCode MatScale(Mat A, PetscScalar a);
PetscErrorCode MatShift(Mat A, PetscScalar a);
PetscErrorCode MatMult(Mat A, Vec x, Vec y);
PetscErrorCode MatNullSpaceCreate(MPI_Comm comm, PetscBool has_cnst, PetscInt n, const Vec vecs[], MatNullSpace* sp);
PetscErrorCode MatSetNullSpace(Mat A, MatNullSpace sp);
PetscErrorCode MatSetTransposeNullSpace(Mat A, MatNullSpace sp);
PetscErrorCode MatNullSpaceRemove(MatNullSpace sp, Vec v);
}
#endif
