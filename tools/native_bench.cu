// native_bench — Python-free timings through the C ABI (starts in about a second on a fresh box, so it
// fits the tail end of a GPU budget).  CUDA-event or wall-clock times after warm-up; never run under a profiler.
//   native_bench e2e [n]        El::Gemm fp32 NN with HOST operands (elb200_dist_gemm_host_f32, pinned memory,
//                               H2D + D2H inside the timed region) on a 1 x 1 grid: the bench's `e2e` leg
//   native_bench f64 [n]        elb200_gemm_f64 NT: time + max relative difference to cublasDgemm
//   native_bench trapcheck      128-bit trapezoid kernel vs the scalar one vs the host definition, bit for bit
//   native_bench l3 [n]         trapezoid kernel GB/s, elb200_trsm TFLOP/s, Syrk / Gemm through the handle ABI
#include <cublas_v2.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "../include/elb200.h"

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { std::printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); std::exit(2); } } while (0)
#define EK(x) do { int s_ = (x); if (s_ != 0) { std::printf("elb200 error %d (%s) at %s:%d\n", s_, elb200_last_error(), __FILE__, __LINE__); std::exit(3); } } while (0)

template <typename T>
__global__ void fill_kernel(T* p, int64_t n, unsigned seed, double scale, int64_t diag_ld, double diag) {
    for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x) {
        unsigned h = unsigned(i) * 2654435761u ^ (seed * 40503u + unsigned(i >> 32));
        h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
        double v = (double(h & 0xFFFFFF) / double(0x1000000) - 0.5) * 2.0 * scale;
        if (diag_ld > 0 && i % (diag_ld + 1) == 0) v += diag;
        p[i] = T(v);
    }
}
template <typename T>
static void fill(T* p, int64_t n, unsigned seed, double scale, int64_t diag_ld = 0, double diag = 0) {
    fill_kernel<T><<<1184, 256>>>(p, n, seed, scale, diag_ld, diag);
    CK(cudaGetLastError());
}

template <typename F>
static double time_ms(F&& fn, int warm, int iters) {
    for (int i = 0; i < warm; ++i) fn();
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(a, nullptr));
    for (int i = 0; i < iters; ++i) fn();
    CK(cudaEventRecord(b, nullptr));
    CK(cudaEventSynchronize(b));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, a, b));
    cudaEventDestroy(a); cudaEventDestroy(b);
    return ms / iters;
}

static int run_e2e(int64_t n) {
    elb200_grid_t g;
    EK(elb200_grid_create(1, 1, 0, nullptr, 0, &g));
    float *A, *B, *C;
    const size_t bytes = size_t(n) * n * sizeof(float);
    CK(cudaMallocHost(&A, bytes)); CK(cudaMallocHost(&B, bytes)); CK(cudaMallocHost(&C, bytes));
    auto init = [&](float* p, unsigned seed) {
        std::vector<std::thread> th;
        const int nt = 16;
        for (int t = 0; t < nt; ++t)
            th.emplace_back([=] {
                const size_t lo = size_t(n) * n * t / nt, hi = size_t(n) * n * (t + 1) / nt;
                unsigned s = seed * 7919u + unsigned(t) * 104729u + 1u;
                for (size_t i = lo; i < hi; ++i) { s = s * 1664525u + 1013904223u; p[i] = (float(s >> 8) / 16777216.0f - 0.5f) * 0.2f; }
            });
        for (auto& x : th) x.join();
    };
    init(A, 1); init(B, 2); init(C, 3);
    auto call = [&] { EK(elb200_dist_gemm_host_f32(g, 0, 0, n, n, n, 0.5f, A, n, B, n, -0.5f, C, n)); };
    call();  // warm-up (pool growth, first-touch)
    call();
    double best = 1e30, sum = 0;
    const int reps = 3;
    for (int r = 0; r < reps; ++r) {
        const auto t0 = std::chrono::steady_clock::now();
        call();  // synchronous: C is complete in host memory on return
        const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        best = std::min(best, ms); sum += ms;
    }
    const double flop = 2.0 * double(n) * n * n;
    std::printf("e2e host-operand El::Gemm f32 n=%lld: mean %.2f ms (%.1f TFLOP/s), best %.2f ms (%.1f TFLOP/s); H2D %.2f GiB + D2H %.2f GiB per call\n",
                (long long)n, sum / reps, flop / (sum / reps) * 1e-9, best, flop / best * 1e-9, 3.0 * bytes / double(1 << 30), bytes / double(1 << 30));
    double chk = 0;
    for (size_t i = 0; i < size_t(n) * n; i += size_t(n) * n / 4096 + 1) chk += C[i];
    std::printf("e2e checksum (finite: %d) %.6e\n", int(std::isfinite(chk)), chk);
    elb200_grid_destroy(g);
    return std::isfinite(chk) ? 0 : 1;
}

static int run_f64(int64_t n) {
    double *A, *B, *C, *R;
    const size_t bytes = size_t(n) * n * sizeof(double);
    CK(cudaMalloc(&A, bytes)); CK(cudaMalloc(&B, bytes)); CK(cudaMalloc(&C, bytes)); CK(cudaMalloc(&R, bytes));
    fill(A, n * n, 1, 0.1); fill(B, n * n, 2, 0.1); fill(C, n * n, 3, 0.1);
    CK(cudaMemcpy(R, C, bytes, cudaMemcpyDeviceToDevice));
    cublasHandle_t h; cublasCreate(&h);
    const double al = 0.5, be = -0.5;
    EK(elb200_gemm_f64(0, 1, n, n, n, al, A, n, B, n, be, C, n, nullptr));
    cublasDgemm(h, CUBLAS_OP_N, CUBLAS_OP_T, int(n), int(n), int(n), &al, A, int(n), B, int(n), &be, R, int(n));
    CK(cudaDeviceSynchronize());
    std::vector<double> hc(size_t(n) * 64), hr(size_t(n) * 64);   // 64 columns spread over the matrix
    double num = 0, den = 0;
    for (int j = 0; j < 64; ++j) {
        const size_t col = size_t(j) * (n / 64);
        CK(cudaMemcpy(hc.data() + size_t(j) * n, C + col * n, n * sizeof(double), cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(hr.data() + size_t(j) * n, R + col * n, n * sizeof(double), cudaMemcpyDeviceToHost));
    }
    for (size_t i = 0; i < hc.size(); ++i) { num += (hc[i] - hr[i]) * (hc[i] - hr[i]); den += hr[i] * hr[i]; }
    const double rel = std::sqrt(num / den), bound = 1e-13 * std::sqrt(double(n));
    std::printf("f64 kernel n=%lld NT: rel. Frobenius difference to cublasDgemm %.3e (bound %.1e) %s\n",
                (long long)n, rel, bound, rel <= bound ? "ok" : "FAIL");
    const double flop = 2.0 * double(n) * n * n;
    double ms = time_ms([&] { elb200_gemm_f64(0, 1, n, n, n, al, A, n, B, n, be, C, n, nullptr); }, 1, 3);
    std::printf("f64 elb200_gemm_f64 NT n=%lld: %.3f ms  %.2f TFLOP/s\n", (long long)n, ms, flop / ms * 1e-9);
    ms = time_ms([&] { elb200_gemm_f64(0, 0, n, n, n, al, A, n, B, n, be, C, n, nullptr); }, 1, 3);
    std::printf("f64 elb200_gemm_f64 NN n=%lld: %.3f ms  %.2f TFLOP/s\n", (long long)n, ms, flop / ms * 1e-9);
    {
        ms = time_ms([&] { cublasDgemm(h, CUBLAS_OP_N, CUBLAS_OP_T, int(n), int(n), int(n), &al, A, int(n), B, int(n), &be, R, int(n)); }, 1, 3);
        std::printf("f64 cublasDgemm NT n=%lld: %.3f ms  %.2f TFLOP/s\n", (long long)n, ms, flop / ms * 1e-9);
    }
    cublasDestroy(h);
    return rel <= bound ? 0 : 1;
}

// native_bench f16 [n]: elb200_gemm_f16 (tcgen05 kind::f16) NT / NN / TN / TT against cublasGemmEx half (fp32 compute)
__global__ void fill_half_kernel(__half* p, int64_t n, unsigned seed) {
    for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x) {
        unsigned h = unsigned(i) * 2654435761u ^ (seed * 40503u + unsigned(i >> 32));
        h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
        p[i] = __float2half_rn((float(h & 0xFFFF) / 65536.0f - 0.5f) * 0.2f);
    }
}
static int run_f16(int64_t n) {
    __half *A, *B, *C, *R;
    const size_t bytes = size_t(n) * n * sizeof(__half);
    CK(cudaMalloc(&A, bytes)); CK(cudaMalloc(&B, bytes)); CK(cudaMalloc(&C, bytes)); CK(cudaMalloc(&R, bytes));
    fill_half_kernel<<<1184, 256>>>(A, n * n, 1); fill_half_kernel<<<1184, 256>>>(B, n * n, 2); fill_half_kernel<<<1184, 256>>>(C, n * n, 3);
    CK(cudaMemcpy(R, C, bytes, cudaMemcpyDeviceToDevice));
    cublasHandle_t h; cublasCreate(&h);
    const float al = 0.5f, be = -0.5f;
    int rc = 0;
    const double flop = 2.0 * double(n) * n * n;
    const cublasOperation_t ops[2] = {CUBLAS_OP_N, CUBLAS_OP_T};
    for (int ta = 0; ta < 2; ++ta)
        for (int tb = 0; tb < 2; ++tb) {
            CK(cudaMemcpy(C, R, bytes, cudaMemcpyDeviceToDevice));
            __half* Rc; CK(cudaMalloc(&Rc, bytes)); CK(cudaMemcpy(Rc, R, bytes, cudaMemcpyDeviceToDevice));
            EK(elb200_gemm_f16(ta, tb, n, n, n, al, A, n, B, n, be, C, n, nullptr));
            cublasGemmEx(h, ops[ta], ops[tb], int(n), int(n), int(n), &al, A, CUDA_R_16F, int(n), B, CUDA_R_16F, int(n), &be, Rc, CUDA_R_16F, int(n),
                         CUBLAS_COMPUTE_32F, CUBLAS_GEMM_DEFAULT);
            CK(cudaDeviceSynchronize());
            std::vector<__half> hc(size_t(n) * 16), hr(size_t(n) * 16);
            double num = 0, den = 0;
            for (int j = 0; j < 16; ++j) {
                const size_t col = size_t(j) * (n / 16);
                CK(cudaMemcpy(hc.data() + size_t(j) * n, C + col * n, n * sizeof(__half), cudaMemcpyDeviceToHost));
                CK(cudaMemcpy(hr.data() + size_t(j) * n, Rc + col * n, n * sizeof(__half), cudaMemcpyDeviceToHost));
            }
            for (size_t i = 0; i < hc.size(); ++i) { const double d = double(__half2float(hc[i])) - double(__half2float(hr[i])); num += d * d; den += double(__half2float(hr[i])) * __half2float(hr[i]); }
            const double rel = std::sqrt(num / den);
            if (!(rel <= 2e-3)) rc = 1;
            double ms = time_ms([&] { elb200_gemm_f16(ta, tb, n, n, n, al, A, n, B, n, be, C, n, nullptr); }, 2, 5);
            double msc = time_ms([&] { cublasGemmEx(h, ops[ta], ops[tb], int(n), int(n), int(n), &al, A, CUDA_R_16F, int(n), B, CUDA_R_16F, int(n), &be, Rc, CUDA_R_16F,
                                                    int(n), CUBLAS_COMPUTE_32F, CUBLAS_GEMM_DEFAULT); }, 2, 5);
            std::printf("f16 %c%c n=%lld: elb200_gemm_f16 %.3f ms %.1f TFLOP/s | cublasGemmEx(half, fp32 compute) %.3f ms %.1f TFLOP/s | rel. diff %.2e %s\n",
                        "NT"[ta], "NT"[tb], (long long)n, ms, flop / ms * 1e-9, msc, flop / msc * 1e-9, rel, rel <= 2e-3 ? "ok" : "FAIL");
            cudaFree(Rc);
        }
    cublasDestroy(h);
    return rc;
}

template <typename T>
static void l3_type(int64_t n, const char* tn) {
    const bool f32 = sizeof(T) == 4;
    T *X, *Y;
    const size_t bytes = size_t(n) * n * sizeof(T);
    CK(cudaMalloc(&X, bytes)); CK(cudaMalloc(&Y, bytes));
    fill(X, n * n, 5, 0.1, n, double(n));  // also the triangle of the Trsm: n on the diagonal
    fill(Y, n * n, 6, 0.1);
    auto trap = [&](int op, int uplo, T alpha, const T* x) {
        if (f32) return elb200_trapezoid_f32(op, uplo, 0, n, n, 0, 1, 0, 1, float(alpha), (const float*)x, n, (float*)Y, n, nullptr);
        return elb200_trapezoid_f64(op, uplo, 0, n, n, 0, 1, 0, 1, double(alpha), (const double*)x, n, (double*)Y, n, nullptr);
    };
    const double tri = double(n) * (n + 1) / 2 * sizeof(T);
    double ms = time_ms([&] { trap(ELB200_TRAP_AXPY, ELB200_LOWER, T(0.5), X); }, 2, 10);
    std::printf("l3 %s trapezoid AXPY  lower n=%lld: %.3f ms  %.0f GB/s (3 x |T| bytes)\n", tn, (long long)n, ms, 3 * tri / ms * 1e-6);
    ms = time_ms([&] { trap(ELB200_TRAP_SCALE, ELB200_UPPER, T(1.0001), nullptr); }, 2, 10);
    std::printf("l3 %s trapezoid SCALE upper n=%lld: %.3f ms  %.0f GB/s (2 x |T| bytes)\n", tn, (long long)n, ms, 2 * tri / ms * 1e-6);
    ms = time_ms([&] { trap(ELB200_TRAP_MAKE, ELB200_LOWER, T(0), nullptr); }, 2, 10);
    std::printf("l3 %s trapezoid MAKE  lower n=%lld: %.3f ms  %.0f GB/s (1 x outside bytes)\n", tn, (long long)n, ms, (double(n) * n * sizeof(T) - tri) / ms * 1e-6);
    // Trsm: X holds the triangle (diagonally dominant), Y the right-hand sides; n^2 * nrhs useful flops
    const int64_t nt = n / 2, nrhs = n / 2;
    fill(Y, n * n, 7, 0.1);
    auto trsm = [&](int side, int uplo, int trans) {
        if (f32) return elb200_trsm_f32(side, uplo, trans, 0, nt, nrhs, 1.0f, (const float*)X, n, (float*)Y, n, nullptr);
        return elb200_trsm_f64(side, uplo, trans, 0, nt, nrhs, 1.0, (const double*)X, n, (double*)Y, n, nullptr);
    };
    ms = time_ms([&] { trsm(ELB200_LEFT, ELB200_LOWER, 0); }, 1, 3);
    std::printf("l3 %s elb200_trsm LEFT LOWER N  %lld x %lld: %.3f ms  %.2f TFLOP/s\n", tn, (long long)nt, (long long)nrhs, ms, double(nt) * nt * nrhs / ms * 1e-9);
    ms = time_ms([&] { trsm(ELB200_RIGHT, ELB200_UPPER, 1); }, 1, 3);
    std::printf("l3 %s elb200_trsm RIGHT UPPER T %lld x %lld: %.3f ms  %.2f TFLOP/s\n", tn, (long long)nt, (long long)nrhs, ms, double(nt) * nt * nrhs / ms * 1e-9);
    cudaFree(X); cudaFree(Y);
}

static int run_l3(int64_t n) {
    l3_type<float>(n, "f32");
    l3_type<double>(n, "f64");
    // Syrk next to Gemm through the handle ABI (1 x 1 grid, fp32)
    elb200_grid_t g;
    EK(elb200_grid_create(1, 1, 0, nullptr, 0, &g));
    elb200_dm_t A, C;
    EK(elb200_dm_create(g, ELB200_F32, 0, 1, &A)); EK(elb200_dm_create(g, ELB200_F32, 0, 1, &C));
    EK(elb200_dm_resize(A, n, n)); EK(elb200_dm_resize(C, n, n));
    elb200_dm_info_t ia, ic;
    EK(elb200_dm_info(A, &ia)); EK(elb200_dm_info(C, &ic));
    EK(elb200_grid_sync(g));
    fill((float*)ia.buffer, ia.ldim * n, 8, 0.1); fill((float*)ic.buffer, ic.ldim * n, 9, 0.1);
    CK(cudaDeviceSynchronize());
    auto wall = [&](auto&& fn) {
        fn(); EK(elb200_grid_sync(g));
        const auto t0 = std::chrono::steady_clock::now();
        for (int i = 0; i < 3; ++i) fn();
        EK(elb200_grid_sync(g));
        return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count() / 3;
    };
    double ms = wall([&] { EK(elb200_dist_gemm(0, 1, 0.5, A, A, -0.5, C, 6)); });
    std::printf("l3 f32 Gemm NT (A A^T) n=%lld: %.2f ms  %.1f TFLOP/s\n", (long long)n, ms, 2.0 * n * n * n / ms * 1e-9);
    ms = wall([&] { EK(elb200_dist_syrk(ELB200_LOWER, 0, 0.5, A, -0.5, C)); });
    std::printf("l3 f32 Syrk LOWER N       n=%lld: %.2f ms  %.1f useful TFLOP/s (n^2 k)\n", (long long)n, ms, 1.0 * n * n * n / ms * 1e-9);
    elb200_dm_destroy(A); elb200_dm_destroy(C);
    elb200_grid_destroy(g);
    return 0;
}

// The 128-bit trapezoid kernel against the scalar one on identical data (the scalar path is taken when the
// base pointer is not 16-byte aligned): results must be bit-identical for every op / triangle / offset / index map.
template <typename T>
static int trapcheck_type(const char* tn) {
    const bool f32 = sizeof(T) == 4;
    const int64_t n = 1000, ld = 1000, total = ld * n;
    T *X, *Y0, *Ya, *Yb;   // Yb is used at +1 element: misaligned -> scalar kernel; its ld stays 1000
    CK(cudaMalloc(&X, (total + 4) * sizeof(T))); CK(cudaMalloc(&Y0, total * sizeof(T)));
    CK(cudaMalloc(&Ya, total * sizeof(T))); CK(cudaMalloc(&Yb, (total + 4) * sizeof(T)));
    fill(X, total + 4, 11, 1.0); fill(Y0, total, 12, 1.0);
    std::vector<T> ha(total), hb(total);
    int bad = 0, cases = 0;
    const int64_t maps[2][4] = {{0, 1, 0, 1}, {1, 2, 3, 4}};
    for (int op = 0; op < 3; ++op)
        for (int uplo = 0; uplo < 2; ++uplo)
            for (int64_t off : {int64_t(0), int64_t(1), int64_t(-5), int64_t(300)})
                for (const auto& mp : maps) {
                    CK(cudaMemcpy(Ya, Y0, total * sizeof(T), cudaMemcpyDeviceToDevice));
                    CK(cudaMemcpy(Yb + 1, Y0, total * sizeof(T), cudaMemcpyDeviceToDevice));
                    const double al = op == 0 ? -0.5 : 2.0;
                    int s1, s2;
                    if (f32) {
                        s1 = elb200_trapezoid_f32(op, uplo, off, n, n, mp[0], mp[1], mp[2], mp[3], float(al), (const float*)X, ld, (float*)Ya, ld, nullptr);
                        s2 = elb200_trapezoid_f32(op, uplo, off, n, n, mp[0], mp[1], mp[2], mp[3], float(al), (const float*)X, ld, (float*)(Yb + 1), ld, nullptr);
                    } else {
                        s1 = elb200_trapezoid_f64(op, uplo, off, n, n, mp[0], mp[1], mp[2], mp[3], al, (const double*)X, ld, (double*)Ya, ld, nullptr);
                        s2 = elb200_trapezoid_f64(op, uplo, off, n, n, mp[0], mp[1], mp[2], mp[3], al, (const double*)X, ld, (double*)(Yb + 1), ld, nullptr);
                    }
                    if (s1 || s2) { ++bad; continue; }
                    CK(cudaMemcpy(ha.data(), Ya, total * sizeof(T), cudaMemcpyDeviceToHost));
                    CK(cudaMemcpy(hb.data(), Yb + 1, total * sizeof(T), cudaMemcpyDeviceToHost));
                    if (std::memcmp(ha.data(), hb.data(), total * sizeof(T)) != 0) ++bad;
                    // and against the definition on the host for this case
                    std::vector<T> hy(total), hx(total);
                    CK(cudaMemcpy(hy.data(), Y0, total * sizeof(T), cudaMemcpyDeviceToHost));
                    CK(cudaMemcpy(hx.data(), X, total * sizeof(T), cudaMemcpyDeviceToHost));
                    int wrong = 0;
                    for (int64_t j = 0; j < n && !wrong; ++j)
                        for (int64_t i = 0; i < n; ++i) {
                            const int64_t gi = mp[0] + i * mp[1], gj = mp[2] + j * mp[3];
                            const bool in = uplo ? gi <= gj - off : gi >= gj - off;
                            T want = hy[i + j * ld];
                            if (op == 0 && in) want = T(al) * want;
                            if (op == 1 && in) want = T(double(want) + al * double(hx[i + j * ld]));   // alpha = 2: exact product
                            if (op == 2 && !in) want = T(0);
                            if (std::memcmp(&want, &ha[i + j * ld], sizeof(T)) != 0) { wrong = 1; break; }
                        }
                    bad += wrong;
                    ++cases;
                }
    std::printf("trapcheck %s: %d cases, vector vs scalar kernel vs host definition: %s\n", tn, cases, bad ? "FAIL" : "bit-identical");
    cudaFree(X); cudaFree(Y0); cudaFree(Ya); cudaFree(Yb);
    return bad;
}

int main(int argc, char** argv) {
    const char* mode = argc > 1 ? argv[1] : "";
    const int64_t n = argc > 2 ? std::atoll(argv[2]) : 0;
    if (!std::strcmp(mode, "e2e")) return run_e2e(n > 0 ? n : 16384);
    if (!std::strcmp(mode, "f64")) return run_f64(n > 0 ? n : 8192);
    if (!std::strcmp(mode, "f16")) return run_f16(n > 0 ? n : 8192);
    if (!std::strcmp(mode, "l3")) return run_l3(n > 0 ? n : 16384);
    if (!std::strcmp(mode, "trapcheck")) return trapcheck_type<float>("f32") + trapcheck_type<double>("f64");
    std::printf("usage: native_bench e2e|f64|f16|l3|trapcheck [n]\n");
    return 64;
}
