/* Hand-written p4est configuration for the serial oracle build (see sc_config.h). */
#ifndef P4EST_CONFIG_H
#define P4EST_CONFIG_H
#define P4EST_PACKAGE "p4est"
#define P4EST_PACKAGE_NAME "p4est"
#define P4EST_PACKAGE_STRING "p4est oracle-serial"
#define P4EST_PACKAGE_VERSION "oracle-serial"
#define P4EST_PACKAGE_TARNAME "p4est"
#define P4EST_PACKAGE_BUGREPORT ""
#define P4EST_PACKAGE_URL ""
#define P4EST_VERSION "oracle-serial"
#define P4EST_VERSION_MAJOR 2
#define P4EST_VERSION_MINOR 0
#define P4EST_VERSION_POINT 0
#define P4EST_CC "gcc"
#define P4EST_CFLAGS "-O2"
#define P4EST_CPP "gcc -E"
#define P4EST_CPPFLAGS ""
#define P4EST_F77 "none"
#define P4EST_FFLAGS ""
#define P4EST_FC "none"
#define P4EST_FCFLAGS ""
#define P4EST_LDFLAGS ""
#define P4EST_LIBS "-lm -lz"
#define P4EST_BLAS_LIBS ""
#define P4EST_LAPACK_LIBS ""
#define P4EST_FLIBS ""
#define P4EST_ENABLE_BUILD_2D 1
#define P4EST_ENABLE_BUILD_3D 1
#define P4EST_ENABLE_BUILD_P6EST 1
#define P4EST_HAVE_ZLIB 1
#define P4EST_HAVE_UNISTD_H 1
#define P4EST_HAVE_FSYNC 1
#define P4EST_HAVE_ARPA_NET_H 1
#define P4EST_HAVE_NETINET_IN_H 1
#define P4EST_STDC_HEADERS 1
#endif
