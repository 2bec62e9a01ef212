/* Link-time stubs for the serial oracle build of the d4est reference.
 * TEST INFRASTRUCTURE ONLY.  Symbols the compiled subset references but that
 * live in files needing MPI/HDF5. */
#include <stdio.h>
#include <stdlib.h>
#include <mpi.h>
int MPI_Comm_size(MPI_Comm c, int *s) { return sc_MPI_Comm_size(c, s); }
int MPI_Comm_rank(MPI_Comm c, int *r) { return sc_MPI_Comm_rank(c, r); }
void PPF_Print(MPI_Comm c, const char *fmt, ...) { (void)c; (void)fmt; }
void PTC_Print(MPI_Comm c, const char *fmt, ...) { (void)c; (void)fmt; }
void mpsort_mpi(void) { fprintf(stderr, "oracle stub: mpsort_mpi\n"); abort(); }
void *d4est_amr_smooth_pred_init(void) { fprintf(stderr, "oracle stub: smooth_pred needs HDF5\n"); abort(); return 0; }

/* Symbols that Solver/d4est_solver_multigrid.c and Solver/d4est_solver_fcg_improved.c reference for smoothers /
 * bottom solvers / checkpoints that need PETSc or HDF5 (absent): selecting one of them aborts. */
#define ORACLE_ABORT_STUB(name) void *name(void) { fprintf(stderr, "oracle stub: " #name " needs PETSc/HDF5\n"); abort(); return 0; }
ORACLE_ABORT_STUB(d4est_checkpoint_save)
ORACLE_ABORT_STUB(d4est_solver_multigrid_bottom_solver_krylov_petsc_destroy)
ORACLE_ABORT_STUB(d4est_solver_multigrid_bottom_solver_krylov_petsc_init)
ORACLE_ABORT_STUB(d4est_solver_multigrid_smoother_krylov_petsc_destroy)
ORACLE_ABORT_STUB(d4est_solver_multigrid_smoother_krylov_petsc_init)
ORACLE_ABORT_STUB(d4est_solver_multigrid_smoother_schwarz_destroy)
ORACLE_ABORT_STUB(d4est_solver_multigrid_smoother_schwarz_init)
ORACLE_ABORT_STUB(d4est_vtk_save_aux)
ORACLE_ABORT_STUB(PCShellGetContext)
ORACLE_ABORT_STUB(VecGetArray)
ORACLE_ABORT_STUB(VecGetArrayRead)
ORACLE_ABORT_STUB(VecRestoreArray)
ORACLE_ABORT_STUB(VecRestoreArrayRead)
