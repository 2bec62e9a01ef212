/* Stub <mpi.h> for the serial oracle build: the reference pulls <mpi.h> through
 * third_party/PPF/ptools_ppf.h:18 and uses MPI_Comm_size/rank in its test mains.
 * Everything maps onto libsc's serial sc_MPI_* stubs (sc_mpi.c). */
#ifndef ORACLE_STUB_MPI_H
#define ORACLE_STUB_MPI_H
#include <sc.h>
typedef sc_MPI_Comm MPI_Comm;
#define MPI_COMM_WORLD sc_MPI_COMM_WORLD
int MPI_Comm_size(MPI_Comm c, int *s);
int MPI_Comm_rank(MPI_Comm c, int *r);
#endif
