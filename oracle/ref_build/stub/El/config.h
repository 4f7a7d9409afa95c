/* Hand-written stand-in for cmake/configure_files/config.h.in. ORACLE BUILD ONLY. */
#ifndef EL_CONFIG_H
#define EL_CONFIG_H
#include "hydrogen_config.h"
#define EL_GIT_SHA1 "unknown"
#define EL_VERSION_MAJOR "1"
#define EL_VERSION_MINOR "5"
#define EL_VERSION_PATCH "4"
#define EL_CMAKE_BUILD_TYPE "Release"
#ifdef HYDROGEN_RELEASE_BUILD
#define EL_RELEASE
#endif
#define EL_CMAKE_C_COMPILER "gcc"
#define EL_CMAKE_CXX_COMPILER "g++"
#define EL_CXX_FLAGS "-O3"
#define EL_FORT_LOGICAL int
#define EL_FORT_TRUE 1
#define EL_FORT_FALSE 0
#define EL_MPI_C_COMPILER "stub"
#define EL_MPI_C_INCLUDE_PATH ""
#define EL_MPI_C_COMPILE_FLAGS ""
#define EL_MPI_C_LIBRARIES ""
#define EL_MPI_C_LINK_FLAGS ""
#define EL_MPI_CXX_COMPILER "stub"
#define EL_MPI_CXX_INCLUDE_PATH ""
#define EL_MPI_CXX_COMPILE_FLAGS ""
#define EL_MPI_CXX_LIBRARIES ""
#define EL_MPI_CXX_LINK_FLAGS ""
#define EL_MATH_LIBS "openblas"
#define EL_HAVE_OPENBLAS
#define EL_RESTRICT __restrict__
#define EL_HAVE_PRETTY_FUNCTION
#define EL_EXPORT
#define EL_LOCAL
#endif
