// Local triangular solve with many right-hand sides for sm_100a:
//   LEFT : op(A) X = alpha B        RIGHT: X op(A) = alpha B        (X overwrites B, column-major)
// Replaces gpu_blas::Trsm -> cublas{S,D}trsm as called by the reference's Trsm(Matrix<F,GPU>) and
// LocalTrsm (src/blas_like/level3/Trsm.cpp:26-50, 501-526; include/hydrogen/blas/GPU_BLAS_impl.hpp).
//
// Recursive blocking on the host side of the launch: the triangle is halved until a 32 x 32 diagonal
// block is left; every off-diagonal update is ONE call of this library's GEMM (tcgen05 3xTF32 /
// DMMA), so all but a 32/n fraction of the flops run on the tensor pipes.  The 32 x 32 diagonal
// solves run in trsm_base_kernel: the (padded) triangle sits in shared memory, every thread owns one
// right-hand side — a column of B staged through shared memory (LEFT) or a row of B read coalesced
// (RIGHT) — and substitutes with its 32 unknowns in registers (loops fully unrolled).
#include <cuda_runtime.h>
#include <cstdint>

#include "elb200_internal.h"
#include "trsm_rec.hpp"

namespace elb {
int check_cuda(cudaError_t e, const char* what);

namespace {

constexpr int NB = kTrsmBase;   // base-case triangle (32)
constexpr int kRhsPerCta = 128; // right-hand sides per CTA (= threads)

// E = op(A) restricted to the nb x nb block, padded to NB x NB with the identity: the padded rows
// solve to exactly 0 for zero right-hand sides, so every loop below has compile-time bounds.
template <typename T, bool TRANS, bool UNIT>
__device__ __forceinline__ void load_triangle(T (*sE)[NB + 1], const T* __restrict__ A, int64_t lda, int nb) {
    for (int idx = threadIdx.x; idx < NB * NB; idx += blockDim.x) {
        const int i = idx % NB, j = idx / NB;  // E(i,j)
        T v = (i == j) ? T(1) : T(0);
        if (i < nb && j < nb && !(UNIT && i == j)) v = TRANS ? A[j + int64_t(i) * lda] : A[i + int64_t(j) * lda];
        sE[i][j] = v;
    }
}

// LEFT: E X = B, B is nb x n. LOWER_EFF: E is lower triangular (forward substitution).
template <typename T, bool LOWER_EFF, bool TRANS, bool UNIT>
__global__ void __launch_bounds__(kRhsPerCta)
trsm_left_base_kernel(int nb, int64_t n, const T* __restrict__ A, int64_t lda, T* B, int64_t ldb) {
    __shared__ T sE[NB][NB + 1];
    __shared__ T sB[kRhsPerCta][NB + 1];
    load_triangle<T, TRANS, UNIT>(sE, A, lda, nb);
    const int64_t c0 = int64_t(blockIdx.x) * kRhsPerCta;
    // stage the nb x 128 tile of B: consecutive threads read consecutive rows of one column
    for (int idx = threadIdx.x; idx < NB * kRhsPerCta; idx += blockDim.x) {
        const int i = idx % NB, c = idx / NB;
        sB[c][i] = (i < nb && c0 + c < n) ? B[i + (c0 + c) * ldb] : T(0);
    }
    __syncthreads();
    T x[NB];
    const int t = threadIdx.x;
    if (LOWER_EFF) {
#pragma unroll
        for (int i = 0; i < NB; ++i) {
            T s = sB[t][i];
#pragma unroll
            for (int j = 0; j < i; ++j) s -= sE[i][j] * x[j];
            x[i] = UNIT ? s : s / sE[i][i];
        }
    } else {
#pragma unroll
        for (int i = NB - 1; i >= 0; --i) {
            T s = sB[t][i];
#pragma unroll
            for (int j = i + 1; j < NB; ++j) s -= sE[i][j] * x[j];
            x[i] = UNIT ? s : s / sE[i][i];
        }
    }
#pragma unroll
    for (int i = 0; i < NB; ++i) sB[t][i] = x[i];
    __syncthreads();
    for (int idx = threadIdx.x; idx < NB * kRhsPerCta; idx += blockDim.x) {
        const int i = idx % NB, c = idx / NB;
        if (i < nb && c0 + c < n) B[i + (c0 + c) * ldb] = sB[c][i];
    }
}

// RIGHT: X E = B, B is m x nb. UPPER_EFF: E is upper triangular (columns solved left to right).
template <typename T, bool UPPER_EFF, bool TRANS, bool UNIT>
__global__ void __launch_bounds__(kRhsPerCta)
trsm_right_base_kernel(int64_t m, int nb, const T* __restrict__ A, int64_t lda, T* B, int64_t ldb) {
    __shared__ T sE[NB][NB + 1];
    load_triangle<T, TRANS, UNIT>(sE, A, lda, nb);
    __syncthreads();
    const int64_t i = int64_t(blockIdx.x) * kRhsPerCta + threadIdx.x;
    if (i >= m) return;
    T x[NB];
#pragma unroll
    for (int j = 0; j < NB; ++j) x[j] = j < nb ? B[i + int64_t(j) * ldb] : T(0);  // row i: coalesced over threads
    if (UPPER_EFF) {
#pragma unroll
        for (int j = 0; j < NB; ++j) {
            T s = x[j];
#pragma unroll
            for (int l = 0; l < j; ++l) s -= x[l] * sE[l][j];
            x[j] = UNIT ? s : s / sE[j][j];
        }
    } else {
#pragma unroll
        for (int j = NB - 1; j >= 0; --j) {
            T s = x[j];
#pragma unroll
            for (int l = j + 1; l < NB; ++l) s -= x[l] * sE[l][j];
            x[j] = UNIT ? s : s / sE[j][j];
        }
    }
#pragma unroll
    for (int j = 0; j < NB; ++j)
        if (j < nb) B[i + int64_t(j) * ldb] = x[j];
}

template <typename T> struct GemmOf;
template <> struct GemmOf<float> {
    static int run(int ta, int tb, int64_t m, int64_t n, int64_t k, float al, const float* A, int64_t lda, const float* B,
                   int64_t ldb, float be, float* C, int64_t ldc, cudaStream_t s) {
        return elb200_gemm_f32(ta, tb, m, n, k, al, A, lda, B, ldb, be, C, ldc, s);
    }
};
template <> struct GemmOf<double> {
    static int run(int ta, int tb, int64_t m, int64_t n, int64_t k, double al, const double* A, int64_t lda, const double* B,
                   int64_t ldb, double be, double* C, int64_t ldc, cudaStream_t s) {
        return elb200_gemm_f64(ta, tb, m, n, k, al, A, lda, B, ldb, be, C, ldc, s);
    }
};

template <typename T, bool EFF, bool TRANS>
int launch_base(bool left, bool unit, int64_t m, int64_t n, const T* A, int64_t lda, T* B, int64_t ldb, cudaStream_t s) {
    const int64_t rhs = left ? n : m;
    const unsigned grid = unsigned((rhs + kRhsPerCta - 1) / kRhsPerCta);
    count_launch();
    if (left) {
        if (unit) trsm_left_base_kernel<T, EFF, TRANS, true><<<grid, kRhsPerCta, 0, s>>>(int(m), n, A, lda, B, ldb);
        else trsm_left_base_kernel<T, EFF, TRANS, false><<<grid, kRhsPerCta, 0, s>>>(int(m), n, A, lda, B, ldb);
    } else {
        if (unit) trsm_right_base_kernel<T, EFF, TRANS, true><<<grid, kRhsPerCta, 0, s>>>(m, int(n), A, lda, B, ldb);
        else trsm_right_base_kernel<T, EFF, TRANS, false><<<grid, kRhsPerCta, 0, s>>>(m, int(n), A, lda, B, ldb);
    }
    return check_cuda(cudaGetLastError(), "trsm_base_kernel");
}

// the backend trsm_recursive (trsm_rec.hpp) runs on: this library's GEMM + the base kernels, one stream
template <typename T>
struct DeviceBackend {
    cudaStream_t s;
    int gemm(int ta, int tb, int64_t m, int64_t n, int64_t k, T al, const T* A, int64_t lda, const T* B, int64_t ldb, T be, T* C,
             int64_t ldc) {
        return GemmOf<T>::run(ta, tb, m, n, k, al, A, lda, B, ldb, be, C, ldc, s);
    }
    int base(bool left, bool lower, bool trans, bool unit, int64_t m, int64_t n, const T* A, int64_t lda, T* B, int64_t ldb) {
        const bool eff_lower = lower != trans;
        const bool eff = left ? eff_lower : !eff_lower;  // LEFT kernels: "E lower"; RIGHT kernels: "E upper"
        if (eff) return trans ? launch_base<T, true, true>(left, unit, m, n, A, lda, B, ldb, s)
                              : launch_base<T, true, false>(left, unit, m, n, A, lda, B, ldb, s);
        return trans ? launch_base<T, false, true>(left, unit, m, n, A, lda, B, ldb, s)
                     : launch_base<T, false, false>(left, unit, m, n, A, lda, B, ldb, s);
    }
};

template <typename T>
int trsm_entry(int side, int uplo, int trans, int diag, int64_t m, int64_t n, T alpha, const T* A, int64_t lda, T* B,
               int64_t ldb, void* stream, int (*scale)(int64_t, int64_t, T, T*, int64_t, void*)) {
    if (side < 0 || side > 1 || uplo < 0 || uplo > 1 || trans < 0 || trans > 2 || diag < 0 || diag > 1 || m < 0 || n < 0)
        return ELB200_ERR_ARG;
    const int64_t na = side == 0 ? m : n;
    if (lda < (na > 1 ? na : 1) || ldb < (m > 1 ? m : 1)) return ELB200_ERR_ARG;
    if (m == 0 || n == 0) return ELB200_OK;
    if (!A || !B) return ELB200_ERR_ARG;
    int st = scale(m, n, alpha, B, ldb, stream);  // B *= alpha (alpha == 0: exact zeros; alpha == 1: nothing)
    if (st != ELB200_OK || alpha == T(0)) return st;
    DeviceBackend<T> be{static_cast<cudaStream_t>(stream)};
    return trsm_recursive<T>(be, side == 0, uplo == 0, trans != 0, diag == 1, m, n, A, lda, B, ldb);
}

}  // namespace
}  // namespace elb

using namespace elb;

extern "C" {

int elb200_trsm_f32(int side, int uplo, int trans, int diag, int64_t m, int64_t n, float alpha, const float* A, int64_t lda,
                    float* B, int64_t ldb, void* stream) {
    return trsm_entry<float>(side, uplo, trans, diag, m, n, alpha, A, lda, B, ldb, stream, elb200_scale2d_f32);
}
int elb200_trsm_f64(int side, int uplo, int trans, int diag, int64_t m, int64_t n, double alpha, const double* A, int64_t lda,
                    double* B, int64_t ldb, void* stream) {
    return trsm_entry<double>(side, uplo, trans, diag, m, n, alpha, A, lda, B, ldb, stream, elb200_scale2d_f64);
}

}  // extern "C"
