/* Hand-written libsc configuration for the serial (no-MPI) oracle build of the
 * d4est reference.  TEST INFRASTRUCTURE ONLY.  The reference snapshot has no
 * generated configure script and autotools are absent in this image, so the
 * handful of macros sc/src reads (sc.c:1233-1246, sc.c:323) are stated here. */
#ifndef SC_CONFIG_H
#define SC_CONFIG_H
#define SC_PACKAGE "libsc"
#define SC_PACKAGE_NAME "libsc"
#define SC_PACKAGE_STRING "libsc oracle-serial"
#define SC_PACKAGE_VERSION "oracle-serial"
#define SC_PACKAGE_TARNAME "libsc"
#define SC_PACKAGE_BUGREPORT ""
#define SC_PACKAGE_URL ""
#define SC_VERSION "oracle-serial"
#define SC_VERSION_MAJOR 2
#define SC_VERSION_MINOR 0
#define SC_VERSION_POINT 0
#define SC_CC "gcc"
#define SC_CFLAGS "-O2"
#define SC_CPP "gcc -E"
#define SC_CPPFLAGS ""
#define SC_F77 "none"
#define SC_FFLAGS ""
#define SC_FC "none"
#define SC_FCFLAGS ""
#define SC_LDFLAGS ""
#define SC_LIBS "-lm -lz"
#define SC_BLAS_LIBS ""
#define SC_LAPACK_LIBS ""
#define SC_FLIBS ""
#define SC_HAVE_UNISTD_H 1
#define SC_HAVE_TIME_H 1
#define SC_HAVE_SYS_TYPES_H 1
#define SC_HAVE_SYS_TIME_H 1
#define SC_HAVE_SIGNAL_H 1
#define SC_HAVE_STRTOLL 1
#define SC_HAVE_POSIX_MEMALIGN 1
#define SC_HAVE_ANY_MEMALIGN 1
#define SC_ENABLE_MEMALIGN 1
#define SC_MEMALIGN 1
#define SC_MEMALIGN_BYTES 16
#define SC_HAVE_ZLIB 1
#define SC_SIZEOF_INT 4
#define SC_SIZEOF_LONG 8
#define SC_SIZEOF_LONG_LONG 8
#define SC_SIZEOF_UNSIGNED_LONG 8
#define SC_SIZEOF_UNSIGNED_LONG_LONG 8
#define SC_SIZEOF_VOID_P 8
#define SC_STDC_HEADERS 1
#endif
