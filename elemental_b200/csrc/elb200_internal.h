// Internal C++ declarations shared by the .cu translation units of libelb200.
#pragma once
#include <cuda_runtime.h>
#include <atomic>
#include <cstdint>

#include "../../include/elb200.h"

namespace elb {

int device_sm_count();

// cudaFuncSetAttribute (dynamic shared memory opt-in) is per DEVICE: one flag per device and kernel instantiation, so a
// process that drives several GPUs configures each of them (a process-wide `static bool` would skip the second device and
// its first launch there would fail).
struct PerDeviceOnce {
    std::atomic<bool> done[64];
    bool needs(int* dev_out) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { *dev_out = -1; return true; }
        *dev_out = dev;
        return !done[dev].load(std::memory_order_acquire);
    }
    void set(int dev) { if (dev >= 0) done[dev].store(true, std::memory_order_release); }
};
// every kernel launch of this library bumps a process-wide counter (bench.py's gpu_launches)
void count_launch();

int gemm_f32_tensor(int transA, int transB, int64_t m, int64_t n, int64_t k, float alpha,
                    const float* A, int64_t lda, const float* B, int64_t ldb, float beta, float* C,
                    int64_t ldc, int cta_group, cudaStream_t stream);

// half operands (passed as void* so that host-only translation units need no cuda_fp16.h), fp32 accumulate
int gemm_f16_tensor(int transA, int transB, int64_t m, int64_t n, int64_t k, float alpha,
                    const void* A, int64_t lda, const void* B, int64_t ldb, float beta, void* C,
                    int64_t ldc, int cta_group, cudaStream_t stream);

int gemm_f64_dmma(int transA, int transB, int64_t m, int64_t n, int64_t k, double alpha,
                  const double* A, int64_t lda, const double* B, int64_t ldb, double beta,
                  double* C, int64_t ldc, cudaStream_t stream);

}  // namespace elb
