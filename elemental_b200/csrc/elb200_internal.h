// Internal C++ declarations shared by the .cu translation units of libelb200.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

#include "../../include/elb200.h"

namespace elb {

int device_sm_count();
// every kernel launch of this library bumps a process-wide counter (bench.py's gpu_launches)
void count_launch();
// > 0: lattice (pack/unpack) kernels launched by this thread use at most this many CTAs
int lattice_cta_limit();
void set_lattice_cta_limit(int ctas);

int gemm_f32_tensor(int transA, int transB, int64_t m, int64_t n, int64_t k, float alpha,
                    const float* A, int64_t lda, const float* B, int64_t ldb, float beta, float* C,
                    int64_t ldc, int cta_group, cudaStream_t stream);

int gemm_f64_dmma(int transA, int transB, int64_t m, int64_t n, int64_t k, double alpha,
                  const double* A, int64_t lda, const double* B, int64_t ldb, double beta,
                  double* C, int64_t ldc, cudaStream_t stream);

}  // namespace elb
