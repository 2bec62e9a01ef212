/* Typedef-only stand-in for <petscsnes.h> (oracle build, TEST INFRASTRUCTURE ONLY).  PETSc is absent from the
 * reference snapshot (third_party/petsc submodule missing); the reference's in-house solver stack
 * (d4est_solver_multigrid.c, d4est_solver_fcg_improved.c, d4est_krylov_pc_multigrid.c ...) only needs these names
 * to COMPILE: it includes Solver/d4est_solver_petsc.h for krylov_ctx_t (a KSP* member) and
 * Solver/d4est_krylov_pc.h, whose two static PCShell thunks mention the calls declared below.  None of them is
 * ever called by the oracle; link_stubs.c aborts if one is. */
#ifndef ORACLE_STUB_PETSCSNES_H
#define ORACLE_STUB_PETSCSNES_H
typedef int PetscErrorCode;
typedef int PetscInt;
typedef double PetscReal;
typedef double PetscScalar;
typedef struct oracle_petsc_pc_ *PC;
typedef struct oracle_petsc_vec_ *Vec;
typedef struct oracle_petsc_mat_ *Mat;
typedef struct oracle_petsc_ksp_ *KSP;
typedef struct oracle_petsc_snes_ *SNES;
#define CHKERRQ(ierr) do { if (ierr) return ierr; } while (0)
PetscErrorCode PCShellGetContext(PC pc, void **ctx);
PetscErrorCode VecGetArray(Vec v, double **a);
PetscErrorCode VecGetArrayRead(Vec v, const double **a);
PetscErrorCode VecRestoreArray(Vec v, double **a);
PetscErrorCode VecRestoreArrayRead(Vec v, const double **a);
#endif
