// C-ABI glue for the device layer: status strings, error capture, GEMM argument checks.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstring>
#include <atomic>
#include <mutex>

#include "elb200_internal.h"

namespace elb {

thread_local char g_last_error[512] = "";
std::atomic<unsigned long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

void set_last_error(const char* what, const char* detail) {
    std::snprintf(g_last_error, sizeof(g_last_error), "%s: %s", what, detail ? detail : "");
}

int check_cuda(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return ELB200_OK;
    set_last_error(what, cudaGetErrorString(e));
    return ELB200_ERR_CUDA;
}

int device_sm_count() {
    static std::atomic<int> cached[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    int n = cached[dev].load(std::memory_order_relaxed);
    if (n == 0) {
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
            n = 148;
        cached[dev].store(n, std::memory_order_relaxed);
    }
    return n;
}

namespace {

// Dimension / pointer checks shared by both GEMMs (the reference throws LogicError for
// non-conformal arguments, src/blas_like/level3/Gemm.cpp:199-238).
int check_gemm_args(int transA, int transB, int64_t m, int64_t n, int64_t k, const void* A,
                    int64_t lda, const void* B, int64_t ldb, const void* C, int64_t ldc) {
    if (transA < 0 || transA > 2 || transB < 0 || transB > 2) return ELB200_ERR_ARG;
    if (m < 0 || n < 0 || k < 0) return ELB200_ERR_ARG;
    const int64_t a_rows = transA ? k : m, b_rows = transB ? n : k;
    if (lda < (a_rows > 1 ? a_rows : 1) || ldb < (b_rows > 1 ? b_rows : 1) ||
        ldc < (m > 1 ? m : 1))
        return ELB200_ERR_ARG;
    if (m > 0 && n > 0 && !C) return ELB200_ERR_ARG;
    if (m > 0 && n > 0 && k > 0 && (!A || !B)) return ELB200_ERR_ARG;
    if (m > INT32_MAX || n > INT32_MAX || k > INT32_MAX) return ELB200_ERR_ARG;
    return ELB200_OK;
}

// ---- staging for elb200_gemm_f32_noalloc ----------------------------------------------------------
// Two regions of ELB200_NOALLOC_MAX_ELEMS floats (A, then B): 2 x 16 MiB of device memory that every GPU
// which loads this library holds until the process exits. Nothing in it has to start at zero: a staged
// operand is rewritten in full by the copy that precedes the GEMM, and the pad rows between an operand's
// height and its 4-element-rounded pitch lie outside the tensor map's extent, so TMA never reads them.
static __device__ __align__(256) float g_noalloc_scratch[2 * size_t(ELB200_NOALLOC_MAX_ELEMS)];

// Device address of the array on the CURRENT device, looked up once per device (the symbol lives in every
// context that loads the module, at a different address each). nullptr + last_error on failure.
float* noalloc_scratch_base() {
    static std::mutex mu;
    static float* base[64] = {nullptr};
    int dev = -1;
    if (check_cuda(cudaGetDevice(&dev), "cudaGetDevice")) return nullptr;
    if (dev < 0 || dev >= 64) { set_last_error("elb200_gemm_f32_noalloc", "device index out of range"); return nullptr; }
    std::lock_guard<std::mutex> lock(mu);
    if (!base[dev]) {
        void* p = nullptr;
        if (check_cuda(cudaGetSymbolAddress(&p, g_noalloc_scratch), "cudaGetSymbolAddress(g_noalloc_scratch)"))
            return nullptr;
        base[dev] = static_cast<float*>(p);
    }
    return base[dev];
}

}  // namespace
}  // namespace elb

using namespace elb;

extern "C" {

const char* elb200_status_string(int status) {
    switch (status) {
        case ELB200_OK: return "ok";
        case ELB200_ERR_CUDA: return "CUDA error";
        case ELB200_ERR_ARG: return "invalid argument";
        case ELB200_ERR_TENSORMAP: return "tensor map encode failed";
        case ELB200_ERR_NCCL: return "NCCL error";
        case ELB200_ERR_STATE: return "library not initialised";
        default: return "unknown status";
    }
}

const char* elb200_last_error(void) { return g_last_error; }
unsigned long long elb200_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

const char* elb200_build_info(void) {
    return "libelb200 sm_100a: gemm_f32 tcgen05 3xTF32 (TMA+TMEM, cta_group 1/2), "
           "gemm_f64 DMMA m8n8k4 (64x64x16 tiles, 4 CTAs/SM, cp.async ring), trsm (recursive GEMM + 32x32 base), blas1 vectorised, trapezoid, lattice pack/unpack, "
           "NVLink peer-memory AllGather / ReduceScatter (copy engines + stream memops), NCCL fallback";
}

int elb200_gemm_f32_ex(int transA, int transB, int64_t m, int64_t n, int64_t k, float alpha,
                       const float* A, int64_t lda, const float* B, int64_t ldb, float beta,
                       float* C, int64_t ldc, int cta_group, void* stream) {
    int st = check_gemm_args(transA, transB, m, n, k, A, lda, B, ldb, C, ldc);
    if (st != ELB200_OK) return st;
    if (m == 0 || n == 0) return ELB200_OK;
    if (k == 0 || alpha == 0.f) return elb200_scale2d_f32(m, n, beta, C, ldc, stream);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    // TMA needs 16-byte aligned bases and row pitches; stage non-conforming operands.
    const bool a_ok = (reinterpret_cast<uintptr_t>(A) & 15u) == 0 && (lda & 3) == 0;
    const bool b_ok = (reinterpret_cast<uintptr_t>(B) & 15u) == 0 && (ldb & 3) == 0;
    const float* Ause = A; const float* Buse = B;
    int64_t lda_use = lda, ldb_use = ldb;
    float* tmpA = nullptr; float* tmpB = nullptr;
    auto release = [&]() {  // stream-ordered: safe right after the launches that read them
        if (tmpA) cudaFreeAsync(tmpA, s);
        if (tmpB) cudaFreeAsync(tmpB, s);
        tmpA = tmpB = nullptr;
    };
    if (!a_ok) {
        const int64_t rows = transA ? k : m, cols = transA ? m : k;
        lda_use = (rows + 3) & ~int64_t(3);
        if (check_cuda(cudaMallocAsync(&tmpA, size_t(lda_use) * cols * 4, s), "cudaMallocAsync"))
            return ELB200_ERR_CUDA;
        st = elb200_copy2d_f32(rows, cols, A, 1, lda, tmpA, 1, lda_use, stream);
        if (st != ELB200_OK) { release(); return st; }
        Ause = tmpA;
    }
    if (!b_ok) {
        const int64_t rows = transB ? n : k, cols = transB ? k : n;
        ldb_use = (rows + 3) & ~int64_t(3);
        if (check_cuda(cudaMallocAsync(&tmpB, size_t(ldb_use) * cols * 4, s), "cudaMallocAsync")) {
            tmpB = nullptr;
            release();
            return ELB200_ERR_CUDA;
        }
        st = elb200_copy2d_f32(rows, cols, B, 1, ldb, tmpB, 1, ldb_use, stream);
        if (st != ELB200_OK) { release(); return st; }
        Buse = tmpB;
    }
    st = gemm_f32_tensor(transA != 0, transB != 0, m, n, k, alpha, Ause, lda_use, Buse, ldb_use,
                         beta, C, ldc, cta_group, s);
    if (st == ELB200_ERR_CUDA) set_last_error("gemm_f32 launch", cudaGetErrorString(cudaGetLastError()));
    release();
    return st;
}

// Allocation-free twin of elb200_gemm_f32_ex (no counterpart in the reference; see include/elb200.h).
// Misaligned operands are staged into the file-scope device array above instead of cudaMallocAsync
// temporaries. In host code `g_noalloc_scratch` is only a stand-in for the device symbol, so its device
// address is resolved once per device with cudaGetSymbolAddress and that pointer is what the kernels get.
int elb200_gemm_f32_noalloc(int transA, int transB, int64_t m, int64_t n, int64_t k, float alpha,
                            const float* A, int64_t lda, const float* B, int64_t ldb, float beta,
                            float* C, int64_t ldc, void* stream) {
    int st = check_gemm_args(transA, transB, m, n, k, A, lda, B, ldb, C, ldc);
    if (st != ELB200_OK) return st;
    if (m == 0 || n == 0) return ELB200_OK;
    if (k == 0 || alpha == 0.f) return elb200_scale2d_f32(m, n, beta, C, ldc, stream);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const bool a_ok = (reinterpret_cast<uintptr_t>(A) & 15u) == 0 && (lda & 3) == 0;
    const bool b_ok = (reinterpret_cast<uintptr_t>(B) & 15u) == 0 && (ldb & 3) == 0;
    const float* Ause = A; const float* Buse = B;
    int64_t lda_use = lda, ldb_use = ldb;
    if (!a_ok || !b_ok) {
        const int64_t a_rows = transA ? k : m, a_cols = transA ? m : k;
        const int64_t b_rows = transB ? n : k, b_cols = transB ? k : n;
        const int64_t a_ld = (a_rows + 3) & ~int64_t(3), b_ld = (b_rows + 3) & ~int64_t(3);
        // sizes are checked before anything is enqueued: a rejected call leaves the stream untouched
        if ((!a_ok && a_ld * a_cols > int64_t(ELB200_NOALLOC_MAX_ELEMS)) ||
            (!b_ok && b_ld * b_cols > int64_t(ELB200_NOALLOC_MAX_ELEMS))) {
            set_last_error("elb200_gemm_f32_noalloc",
                           "a misaligned operand needs more staging than ELB200_NOALLOC_MAX_ELEMS floats; "
                           "align it to 16 bytes (base and leading dimension) or use elb200_gemm_f32");
            return ELB200_ERR_ARG;
        }
        float* scratch = noalloc_scratch_base();
        if (!scratch) return ELB200_ERR_CUDA;
        if (!a_ok) {
            st = elb200_copy2d_f32(a_rows, a_cols, A, 1, lda, scratch, 1, a_ld, stream);
            if (st != ELB200_OK) return st;
            Ause = scratch; lda_use = a_ld;
        }
        if (!b_ok) {
            float* sb = scratch + ELB200_NOALLOC_MAX_ELEMS;
            st = elb200_copy2d_f32(b_rows, b_cols, B, 1, ldb, sb, 1, b_ld, stream);
            if (st != ELB200_OK) return st;
            Buse = sb; ldb_use = b_ld;
        }
    }
    st = gemm_f32_tensor(transA != 0, transB != 0, m, n, k, alpha, Ause, lda_use, Buse, ldb_use,
                         beta, C, ldc, 0, s);
    if (st == ELB200_ERR_CUDA) set_last_error("gemm_f32 launch", cudaGetErrorString(cudaGetLastError()));
    return st;
}

int elb200_gemm_f32(int transA, int transB, int64_t m, int64_t n, int64_t k, float alpha,
                    const float* A, int64_t lda, const float* B, int64_t ldb, float beta, float* C,
                    int64_t ldc, void* stream) {
    return elb200_gemm_f32_ex(transA, transB, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, 0,
                              stream);
}

// half operands and result, fp32 alpha / beta / accumulation
int elb200_gemm_f16(int transA, int transB, int64_t m, int64_t n, int64_t k, float alpha,
                    const void* A, int64_t lda, const void* B, int64_t ldb, float beta, void* C,
                    int64_t ldc, void* stream) {
    int st = check_gemm_args(transA, transB, m, n, k, A, lda, B, ldb, C, ldc);
    if (st != ELB200_OK) return st;
    if (m == 0 || n == 0) return ELB200_OK;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (k == 0 || alpha == 0.f) return elb200_scale2d_f16(m, n, beta, C, ldc, stream);
    st = gemm_f16_tensor(transA != 0, transB != 0, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, 0, s);
    if (st == ELB200_ERR_ARG)
        set_last_error("elb200_gemm_f16", "operands must be 16-byte aligned with leading dimensions that are multiples of 8 halfs");
    if (st == ELB200_ERR_CUDA) set_last_error("gemm_f16 launch", cudaGetErrorString(cudaGetLastError()));
    return st;
}

int elb200_gemm_f64(int transA, int transB, int64_t m, int64_t n, int64_t k, double alpha,
                    const double* A, int64_t lda, const double* B, int64_t ldb, double beta,
                    double* C, int64_t ldc, void* stream) {
    int st = check_gemm_args(transA, transB, m, n, k, A, lda, B, ldb, C, ldc);
    if (st != ELB200_OK) return st;
    if (m == 0 || n == 0) return ELB200_OK;
    if (k == 0 || alpha == 0.0) return elb200_scale2d_f64(m, n, beta, C, ldc, stream);
    st = gemm_f64_dmma(transA != 0, transB != 0, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc,
                       static_cast<cudaStream_t>(stream));
    if (st == ELB200_ERR_CUDA) set_last_error("gemm_f64 launch", cudaGetErrorString(cudaGetLastError()));
    return st;
}

}  // extern "C"
