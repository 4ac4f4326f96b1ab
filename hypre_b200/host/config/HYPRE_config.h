/* hypre_b200/host/config/HYPRE_config.h
 *
 * Hand-written configuration header used when the reference's host sources are compiled
 * where they lie under /root/reference/src (hypre_b200/host/Makefile for the product's
 * libHYPRE_host.so; oracle/Makefile for the checker's separately built libHYPRE_ref.so).
 * It stands in for the header that the reference's own configure step would generate
 * (template: src/config/HYPRE_config.h.in) for a `--without-MPI --with-openmp` host-only build:
 *   - HYPRE_SEQUENTIAL      : hypre_MPI_* resolve to the single-rank stubs in
 *                             src/utilities/mpistubs.c
 *   - HYPRE_USING_OPENMP    : OpenMP row loops (src/seq_mv/csr_matvec.c:673-848)
 *   - HYPRE_USING_HOST_MEMORY: every object lives in host memory
 *   - internal BLAS/LAPACK  : src/blas, src/lapack
 * HYPRE_Int = int, HYPRE_BigInt = int, HYPRE_Real = double
 * (src/utilities/HYPRE_utilities.h:35-61,105,131).
 */
#ifndef ORACLE_REF_HYPRE_CONFIG_H
#define ORACLE_REF_HYPRE_CONFIG_H

#define HAVE_INTTYPES_H 1
#define HAVE_STDINT_H 1
#define HAVE_STDIO_H 1
#define HAVE_STDLIB_H 1
#define HAVE_STRINGS_H 1
#define HAVE_STRING_H 1
#define HAVE_SYS_STAT_H 1
#define HAVE_SYS_TYPES_H 1
#define HAVE_UNISTD_H 1
#define STDC_HEADERS 1

#define HYPRE_RELEASE_NAME "hypre"
#define HYPRE_RELEASE_VERSION "2.32.0"
#define HYPRE_RELEASE_NUMBER 23200
#define HYPRE_RELEASE_DATE "2024/09/27"
#define HYPRE_RELEASE_TIME "00:00:00"
#define HYPRE_RELEASE_BUGS "https://github.com/hypre-space/hypre/issues"
#define HYPRE_DEVELOP_STRING "oracle-build"
#define HYPRE_DEVELOP_NUMBER 0
#define HYPRE_BRANCH_NAME "oracle"

#define HYPRE_LINUX 1
#define HYPRE_MAXDIM 3
#define HYPRE_FMANGLE 0
#define HYPRE_FMANGLE_BLAS 0
#define HYPRE_FMANGLE_LAPACK 0

#define HYPRE_SEQUENTIAL 1
#define HYPRE_TIMING 1
#define HYPRE_USING_OPENMP 1
#define HYPRE_USING_HOPSCOTCH 1
#define HYPRE_USING_HOST_MEMORY 1
#define HYPRE_USING_HYPRE_BLAS 1
#define HYPRE_USING_HYPRE_LAPACK 1

#endif
