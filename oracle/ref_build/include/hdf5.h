/* Typedef-only stand-in for <hdf5.h> (oracle build, TEST INFRASTRUCTURE ONLY): IO/d4est_checkpoint.h is included
 * by Solver/d4est_solver_fcg_improved.c for its optional checkpoint call, which the oracle never enables. */
#ifndef ORACLE_STUB_HDF5_H
#define ORACLE_STUB_HDF5_H
typedef long hid_t;
typedef unsigned long long hsize_t;
typedef int herr_t;
#define H5T_NATIVE_DOUBLE ((hid_t)1)
#define H5T_NATIVE_INT ((hid_t)2)
#endif
