// Bring-up / measurement tool for the two local GEMM kernels (not part of the product path).
//   gemm_check correctness      : all (transA,transB) x edge shapes vs a naive fp64 device GEMM
//   gemm_check perf [n ...]     : CUDA-event timings of elb200 GEMMs and of cuBLAS (the reference's
//                                 LocalGemm backend) at the same sizes, same box, same run
// Build: see tools/Makefile. Links libelb200.so and cuBLAS (cuBLAS only for the comparison lines).
#include <cublas_v2.h>
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../include/elb200.h"

#define CK(x)                                                                          \
    do {                                                                               \
        cudaError_t e_ = (x);                                                          \
        if (e_ != cudaSuccess) {                                                       \
            std::printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
            std::exit(2);                                                              \
        }                                                                              \
    } while (0)

__device__ __forceinline__ uint32_t hash32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
    return x;
}
template <typename T>
__global__ void fill_uniform(T* p, int64_t rows, int64_t cols, int64_t ld, uint32_t seed, double lo,
                             double hi) {
    int64_t idx = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= ld * cols) return;
    int64_t i = idx % ld;
    if (i >= rows) { p[idx] = T(NAN); return; }  // padding rows must never be read
    uint32_t h = hash32(uint32_t(idx) * 2654435761u + seed);
    double u = (h >> 8) * (1.0 / 16777216.0);
    p[idx] = T(lo + (hi - lo) * u);
}
// C_ref(fp64) = alpha*op(A)*op(B) + beta*C0
template <typename T>
__global__ void ref_gemm(int ta, int tb, int m, int n, int k, double alpha, const T* A, int64_t lda,
                         const T* B, int64_t ldb, double beta, const T* C0, int64_t ldc, double* R) {
    int i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y * blockDim.y + threadIdx.y;
    if (i >= m || j >= n) return;
    double s = 0;
    for (int l = 0; l < k; ++l) {
        double a = ta ? double(A[int64_t(i) * lda + l]) : double(A[int64_t(l) * lda + i]);
        double b = tb ? double(B[int64_t(l) * ldb + j]) : double(B[int64_t(j) * ldb + l]);
        s += a * b;
    }
    double c = beta != 0 ? beta * double(C0[int64_t(j) * ldc + i]) : 0.0;
    R[int64_t(j) * m + i] = alpha * s + c;
}
template <typename T>
__global__ void err_kernel(int m, int n, const T* C, int64_t ldc, const double* R, double* acc) {
    int i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y * blockDim.y + threadIdx.y;
    if (i >= m || j >= n) return;
    double r = R[int64_t(j) * m + i], d = double(C[int64_t(j) * ldc + i]) - r;
    if (!(d == d)) d = 1e30;  // NaN in result
    atomicAdd(&acc[0], d * d);
    atomicAdd(&acc[1], r * r);
}

template <typename T>
static int call_gemm(int ta, int tb, int m, int n, int k, T alpha, const T* A, int64_t lda,
                     const T* B, int64_t ldb, T beta, T* C, int64_t ldc, int cg);
template <>
int call_gemm<float>(int ta, int tb, int m, int n, int k, float alpha, const float* A, int64_t lda,
                     const float* B, int64_t ldb, float beta, float* C, int64_t ldc, int cg) {
    return elb200_gemm_f32_ex(ta, tb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, cg, nullptr);
}
template <>
int call_gemm<double>(int ta, int tb, int m, int n, int k, double alpha, const double* A,
                      int64_t lda, const double* B, int64_t ldb, double beta, double* C,
                      int64_t ldc, int) {
    return elb200_gemm_f64(ta, tb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, nullptr);
}

template <typename T>
static bool check_one(const char* name, int cg, int ta, int tb, int m, int n, int k, double alpha,
                      double beta, int pad, double tol) {
    const int64_t a_rows = ta ? k : m, a_cols = ta ? m : k, b_rows = tb ? n : k, b_cols = tb ? k : n;
    const int64_t lda = a_rows + pad, ldb = b_rows + pad, ldc = m + pad;
    T *A, *B, *C, *C0; double *R, *acc;
    CK(cudaMalloc(&A, lda * a_cols * sizeof(T)));
    CK(cudaMalloc(&B, ldb * b_cols * sizeof(T)));
    CK(cudaMalloc(&C, ldc * n * sizeof(T)));
    CK(cudaMalloc(&C0, ldc * n * sizeof(T)));
    CK(cudaMalloc(&R, int64_t(m) * n * sizeof(double)));
    CK(cudaMalloc(&acc, 2 * sizeof(double)));
    CK(cudaMemset(acc, 0, 2 * sizeof(double)));
    auto fill = [&](T* p, int64_t rows, int64_t cols, int64_t ld, uint32_t seed) {
        int64_t tot = ld * cols;
        fill_uniform<T><<<unsigned((tot + 255) / 256), 256>>>(p, rows, cols, ld, seed, -1.0, 1.0);
    };
    fill(A, a_rows, a_cols, lda, 1); fill(B, b_rows, b_cols, ldb, 2); fill(C0, m, n, ldc, 3);
    CK(cudaMemcpy(C, C0, ldc * n * sizeof(T), cudaMemcpyDeviceToDevice));
    dim3 blk(16, 16), grd((m + 15) / 16, (n + 15) / 16);
    ref_gemm<T><<<grd, blk>>>(ta, tb, m, n, k, alpha, A, lda, B, ldb, beta, C0, ldc, R);
    CK(cudaDeviceSynchronize());
    int st = call_gemm<T>(ta, tb, m, n, k, T(alpha), A, lda, B, ldb, T(beta), C, ldc, cg);
    cudaError_t e = cudaDeviceSynchronize();
    if (st != 0 || e != cudaSuccess) {
        std::printf("FAIL %s cg=%d %c%c m=%d n=%d k=%d: status=%d (%s) cuda=%s\n", name, cg,
                    ta ? 'T' : 'N', tb ? 'T' : 'N', m, n, k, st, elb200_last_error(),
                    cudaGetErrorString(e));
        std::exit(3);  // context is likely poisoned
    }
    err_kernel<T><<<grd, blk>>>(m, n, C, ldc, R, acc);
    double h[2];
    CK(cudaMemcpy(h, acc, sizeof(h), cudaMemcpyDeviceToHost));
    double rel = std::sqrt(h[0] / (h[1] > 0 ? h[1] : 1));
    bool ok = rel <= tol * std::sqrt(double(k > 0 ? k : 1));
    std::printf("%s %s cg=%d %c%c m=%d n=%d k=%d pad=%d beta=%g relF=%.3e (bound %.1e)\n",
                ok ? "ok  " : "BAD ", name, cg, ta ? 'T' : 'N', tb ? 'T' : 'N', m, n, k, pad, beta,
                rel, tol * std::sqrt(double(k > 0 ? k : 1)));
    cudaFree(A); cudaFree(B); cudaFree(C); cudaFree(C0); cudaFree(R); cudaFree(acc);
    return ok;
}

static int correctness(const char* which) {
    struct Shape { int m, n, k; };
    const Shape shapes[] = {{128, 128, 32},  {256, 256, 64},  {128, 128, 8},   {256, 512, 96},
                            {300, 200, 100}, {129, 257, 33},  {1000, 1000, 1000}, {64, 40, 500},
                            {2000, 2000, 2000}, {4, 4, 4}, {513, 1, 77}};
    int bad = 0;
    const bool all = !strcmp(which, "all");
    for (int ta = 0; ta < 2; ++ta)
        for (int tb = 0; tb < 2; ++tb)
            for (const Shape& s : shapes)
                for (int v = 0; v < 2; ++v) {
                    const double beta = v ? -0.5 : 0.0;
                    const int pad = v ? 4 : 0;
                    if (all || !strcmp(which, "f32cg1"))
                        bad += !check_one<float>("f32", 1, ta, tb, s.m, s.n, s.k, 0.5, beta, pad, 1e-5);
                    if (all || !strcmp(which, "f32cg2"))
                        bad += !check_one<float>("f32", 2, ta, tb, s.m, s.n, s.k, 0.5, beta, pad, 1e-5);
                    if (all || !strcmp(which, "f64"))
                        bad += !check_one<double>("f64", 0, ta, tb, s.m, s.n, s.k, 0.5, beta, pad ? 3 : 0, 1e-13);
                }
    // operands that TMA cannot address directly (odd leading dimension): staged path
    if (all || !strncmp(which, "f32", 3))
        bad += !check_one<float>("f32", 2, 0, 1, 333, 222, 111, 0.5, -0.5, 1, 1e-5);
    std::printf("correctness(%s): %d bad\n", which, bad);
    return bad;
}

template <typename F>
static double time_ms(F&& f, int warm, int iters) {
    for (int i = 0; i < warm; ++i) f();
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(a));
    for (int i = 0; i < iters; ++i) f();
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    float ms = 0; CK(cudaEventElapsedTime(&ms, a, b));
    cudaEventDestroy(a); cudaEventDestroy(b);
    return ms / iters;
}

template <typename T>
static void perf_one(cublasHandle_t h, int ta, int tb, int m, int n, int k, bool with_cublas) {
    const int64_t a_rows = ta ? k : m, a_cols = ta ? m : k, b_rows = tb ? n : k, b_cols = tb ? k : n;
    T *A, *B, *C;
    CK(cudaMalloc(&A, a_rows * a_cols * sizeof(T)));
    CK(cudaMalloc(&B, b_rows * b_cols * sizeof(T)));
    CK(cudaMalloc(&C, int64_t(m) * n * sizeof(T)));
    fill_uniform<T><<<unsigned((a_rows * a_cols + 255) / 256), 256>>>(A, a_rows, a_cols, a_rows, 1, -0.1, 0.1);
    fill_uniform<T><<<unsigned((b_rows * b_cols + 255) / 256), 256>>>(B, b_rows, b_cols, b_rows, 2, -0.1, 0.1);
    fill_uniform<T><<<unsigned((int64_t(m) * n + 255) / 256), 256>>>(C, m, n, m, 3, -0.1, 0.1);
    const double flop = 2.0 * m * n * k;
    const int iters = flop > 5e12 ? 3 : 10;
    const bool f32 = sizeof(T) == 4;
    const char* tn = f32 ? "f32" : "f64";
    for (int cg = (f32 ? 1 : 0); cg <= (f32 ? 2 : 0); ++cg) {
        double ms = time_ms([&] { call_gemm<T>(ta, tb, m, n, k, T(0.5), A, a_rows, B, b_rows, T(-0.5), C, m, cg); }, 2, iters);
        std::printf("perf elb200 %s cg=%d %c%c m=%d n=%d k=%d : %.3f ms  %.1f TFLOP/s\n", tn, cg,
                    ta ? 'T' : 'N', tb ? 'T' : 'N', m, n, k, ms, flop / ms * 1e-9);
    }
    if (with_cublas) {
        cublasOperation_t oa = ta ? CUBLAS_OP_T : CUBLAS_OP_N, ob = tb ? CUBLAS_OP_T : CUBLAS_OP_N;
        if (f32) {
            float al = 0.5f, be = -0.5f;
            cublasSetMathMode(h, CUBLAS_DEFAULT_MATH);
            double ms = time_ms([&] { cublasSgemm(h, oa, ob, m, n, k, &al, (float*)A, a_rows, (float*)B, b_rows, &be, (float*)C, m); }, 2, iters);
            std::printf("perf cublasSgemm(default math, the reference's LocalGemm) m=%d n=%d k=%d : %.3f ms  %.1f TFLOP/s\n", m, n, k, ms, flop / ms * 1e-9);
            ms = time_ms([&] { cublasGemmEx(h, oa, ob, m, n, k, &al, A, CUDA_R_32F, a_rows, B, CUDA_R_32F, b_rows, &be, C, CUDA_R_32F, m, CUBLAS_COMPUTE_32F_FAST_TF32, CUBLAS_GEMM_DEFAULT); }, 2, iters);
            std::printf("perf cublasGemmEx(1xTF32, lossy) m=%d n=%d k=%d : %.3f ms  %.1f TFLOP/s\n", m, n, k, ms, flop / ms * 1e-9);
        } else {
            double al = 0.5, be = -0.5;
            double ms = time_ms([&] { cublasDgemm(h, oa, ob, m, n, k, &al, (double*)A, a_rows, (double*)B, b_rows, &be, (double*)C, m); }, 2, iters);
            std::printf("perf cublasDgemm(the reference's LocalGemm) m=%d n=%d k=%d : %.3f ms  %.1f TFLOP/s\n", m, n, k, ms, flop / ms * 1e-9);
        }
    }
    cudaFree(A); cudaFree(B); cudaFree(C);
}

int main(int argc, char** argv) {
    const char* mode = argc > 1 ? argv[1] : "correctness";
    if (!strcmp(mode, "correctness")) return correctness(argc > 2 ? argv[2] : "all") ? 1 : 0;
    if (!strcmp(mode, "perf")) {
        const char* dt = argc > 2 ? argv[2] : "both";
        cublasHandle_t h; cublasCreate(&h);
        std::vector<int> sizes;
        for (int i = 3; i < argc; ++i) sizes.push_back(atoi(argv[i]));
        if (sizes.empty()) sizes = {4096, 8192};
        for (int n : sizes) {
            if (strcmp(dt, "f64")) { perf_one<float>(h, 0, 1, n, n, n, true); perf_one<float>(h, 0, 0, n, n, n, false); }
            if (strcmp(dt, "f32")) { perf_one<double>(h, 0, 1, n, n, n, true); perf_one<double>(h, 0, 0, n, n, n, false); }
        }
        // SUMMA_C panel shape: rank-k update, C re-read every call
        if (strcmp(dt, "f64")) for (int k : {128, 512, 2048}) perf_one<float>(h, 0, 1, 16384, 16384, k, k == 128);
        if (strcmp(dt, "f32")) for (int k : {128, 512, 2048}) perf_one<double>(h, 0, 1, 16384, 8192, k, k == 128);
        cublasDestroy(h);
        return 0;
    }
    std::printf("usage: gemm_check correctness [all|f32cg1|f32cg2|f64] | perf [both|f32|f64] [n ...]\n");
    return 64;
}
