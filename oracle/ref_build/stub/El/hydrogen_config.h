/* Hand-written stand-in for the file CMake would generate from
 * /root/reference/cmake/configure_files/hydrogen_config.h.in — CPU-only, no Aluminum, no GPU.
 * ORACLE BUILD ONLY. */
#ifndef HYDROGEN_CONFIG_H
#define HYDROGEN_CONFIG_H
#define HYDROGEN_GIT_VERSION "oracle-build"
#define HYDROGEN_LATEST_SHA "unknown"
#define HYDROGEN_LATEST_HYDROGEN_SHA "unknown"
#define HYDROGEN_VERSION_MAJOR "1"
#define HYDROGEN_VERSION_MINOR "5"
#define HYDROGEN_VERSION_PATCH "4"
#define HYDROGEN_VERSION "1.5.4"
#define HYDROGEN_RELEASE_BUILD
#define HYDROGEN_HAVE_LAPACK
#define HYDROGEN_BLAS_SUFFIX _
#define HYDROGEN_LAPACK_SUFFIX _
#define H_RESTRICT __restrict__
#define H_PRETTY_FUNCTION __PRETTY_FUNCTION__
#endif
