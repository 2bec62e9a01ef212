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
