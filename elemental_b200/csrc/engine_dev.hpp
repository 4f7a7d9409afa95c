// Dev<T>: element-type dispatch from the templated engine to the device C ABI (include/elb200.h).
// Every call throws EngineError on a non-zero status: there is no CPU fallback below this line.
#pragma once
#include <nccl.h>

#include "engine.hpp"

namespace elb {



template <> struct Dev<float> {
    static constexpr ncclDataType_t nccl = ncclFloat;
    static void gemm(int ta, int tb, int64_t m, int64_t n, int64_t k, float al, const float* A, int64_t lda,
                     const float* B, int64_t ldb, float be, float* C, int64_t ldc, cudaStream_t s) {
        throw_status(elb200_gemm_f32(ta, tb, m, n, k, al, A, lda, B, ldb, be, C, ldc, s), "elb200_gemm_f32");
    }
    static void copy2d(int64_t r, int64_t c, const float* src, int64_t srs, int64_t scs, float* dst,
                       int64_t drs, int64_t dcs, cudaStream_t s) {
        throw_status(elb200_copy2d_f32(r, c, src, srs, scs, dst, drs, dcs, s), "elb200_copy2d_f32");
    }
    static void transpose(int64_t r, int64_t c, const float* A, int64_t lda, float* B, int64_t ldb, cudaStream_t s) {
        throw_status(elb200_transpose_f32(r, c, A, lda, B, ldb, s), "elb200_transpose_f32");
    }
    static void axpy2d(int64_t r, int64_t c, float al, const float* X, int64_t xrs, int64_t xcs, float* Y,
                       int64_t yrs, int64_t ycs, cudaStream_t s) {
        throw_status(elb200_axpy2d_f32(r, c, al, X, xrs, xcs, Y, yrs, ycs, s), "elb200_axpy2d_f32");
    }
    static void axpy2d_transpose(int64_t r, int64_t c, float al, const float* A, int64_t lda, float* B,
                                 int64_t ldb, cudaStream_t s) {
        throw_status(elb200_axpy2d_transpose_f32(r, c, al, A, lda, B, ldb, s), "elb200_axpy2d_transpose_f32");
    }
    static void scale2d(int64_t r, int64_t c, float al, float* A, int64_t lda, cudaStream_t s) {
        throw_status(elb200_scale2d_f32(r, c, al, A, lda, s), "elb200_scale2d_f32");
    }
    static void fill2d(int64_t r, int64_t c, float v, float* A, int64_t lda, cudaStream_t s) {
        throw_status(elb200_fill2d_f32(r, c, v, A, lda, s), "elb200_fill2d_f32");
    }
    static void lattice(int mode, int nseg, const elb200_segment* segs, float* M, int64_t ldm, float* buf,
                        float al, cudaStream_t s) {
        throw_status(elb200_lattice_f32(mode, nseg, segs, M, ldm, buf, al, s), "elb200_lattice_f32");
    }
    static void peer_reduce_axpy(int nsrc, const float* const* srcs, int64_t off, const elb200_segment* seg, float* M,
                                 int64_t ldm, float al, cudaStream_t s) {
        throw_status(elb200_peer_reduce_axpy_f32(nsrc, srcs, off, seg, M, ldm, al, s), "elb200_peer_reduce_axpy_f32");
    }
    static void trapezoid(int op, int uplo, int64_t offset, int64_t r, int64_t c, int64_t row0, int64_t rstep, int64_t col0,
                          int64_t cstep, float al, const float* X, int64_t ldx, float* Y, int64_t ldy, cudaStream_t s) {
        throw_status(elb200_trapezoid_f32(op, uplo, offset, r, c, row0, rstep, col0, cstep, al, X, ldx, Y, ldy, s), "elb200_trapezoid_f32");
    }
    static void trsm(int side, int uplo, int trans, int diag, int64_t m, int64_t n, float al, const float* A, int64_t lda, float* B,
                     int64_t ldb, cudaStream_t s) {
        throw_status(elb200_trsm_f32(side, uplo, trans, diag, m, n, al, A, lda, B, ldb, s), "elb200_trsm_f32");
    }
};

template <> struct Dev<double> {
    static constexpr ncclDataType_t nccl = ncclDouble;
    static void gemm(int ta, int tb, int64_t m, int64_t n, int64_t k, double al, const double* A, int64_t lda,
                     const double* B, int64_t ldb, double be, double* C, int64_t ldc, cudaStream_t s) {
        throw_status(elb200_gemm_f64(ta, tb, m, n, k, al, A, lda, B, ldb, be, C, ldc, s), "elb200_gemm_f64");
    }
    static void copy2d(int64_t r, int64_t c, const double* src, int64_t srs, int64_t scs, double* dst,
                       int64_t drs, int64_t dcs, cudaStream_t s) {
        throw_status(elb200_copy2d_f64(r, c, src, srs, scs, dst, drs, dcs, s), "elb200_copy2d_f64");
    }
    static void transpose(int64_t r, int64_t c, const double* A, int64_t lda, double* B, int64_t ldb, cudaStream_t s) {
        throw_status(elb200_transpose_f64(r, c, A, lda, B, ldb, s), "elb200_transpose_f64");
    }
    static void axpy2d(int64_t r, int64_t c, double al, const double* X, int64_t xrs, int64_t xcs, double* Y,
                       int64_t yrs, int64_t ycs, cudaStream_t s) {
        throw_status(elb200_axpy2d_f64(r, c, al, X, xrs, xcs, Y, yrs, ycs, s), "elb200_axpy2d_f64");
    }
    static void axpy2d_transpose(int64_t r, int64_t c, double al, const double* A, int64_t lda, double* B,
                                 int64_t ldb, cudaStream_t s) {
        throw_status(elb200_axpy2d_transpose_f64(r, c, al, A, lda, B, ldb, s), "elb200_axpy2d_transpose_f64");
    }
    static void scale2d(int64_t r, int64_t c, double al, double* A, int64_t lda, cudaStream_t s) {
        throw_status(elb200_scale2d_f64(r, c, al, A, lda, s), "elb200_scale2d_f64");
    }
    static void fill2d(int64_t r, int64_t c, double v, double* A, int64_t lda, cudaStream_t s) {
        throw_status(elb200_fill2d_f64(r, c, v, A, lda, s), "elb200_fill2d_f64");
    }
    static void lattice(int mode, int nseg, const elb200_segment* segs, double* M, int64_t ldm, double* buf,
                        double al, cudaStream_t s) {
        throw_status(elb200_lattice_f64(mode, nseg, segs, M, ldm, buf, al, s), "elb200_lattice_f64");
    }
    static void peer_reduce_axpy(int nsrc, const double* const* srcs, int64_t off, const elb200_segment* seg, double* M,
                                 int64_t ldm, double al, cudaStream_t s) {
        throw_status(elb200_peer_reduce_axpy_f64(nsrc, srcs, off, seg, M, ldm, al, s), "elb200_peer_reduce_axpy_f64");
    }
    static void trapezoid(int op, int uplo, int64_t offset, int64_t r, int64_t c, int64_t row0, int64_t rstep, int64_t col0,
                          int64_t cstep, double al, const double* X, int64_t ldx, double* Y, int64_t ldy, cudaStream_t s) {
        throw_status(elb200_trapezoid_f64(op, uplo, offset, r, c, row0, rstep, col0, cstep, al, X, ldx, Y, ldy, s), "elb200_trapezoid_f64");
    }
    static void trsm(int side, int uplo, int trans, int diag, int64_t m, int64_t n, double al, const double* A, int64_t lda, double* B,
                     int64_t ldb, cudaStream_t s) {
        throw_status(elb200_trsm_f64(side, uplo, trans, diag, m, n, al, A, lda, B, ldb, s), "elb200_trsm_f64");
    }
};

// half: the C ABI takes the device pointers as void* and the scalars as float
template <> struct Dev<Half> {
    static constexpr ncclDataType_t nccl = ncclHalf;
    static void gemm(int ta, int tb, int64_t m, int64_t n, int64_t k, Half al, const Half* A, int64_t lda, const Half* B, int64_t ldb,
                     Half be, Half* C, int64_t ldc, cudaStream_t s) {
        throw_status(elb200_gemm_f16(ta, tb, m, n, k, float(al), A, lda, B, ldb, float(be), C, ldc, s), "elb200_gemm_f16");
    }
    static void copy2d(int64_t r, int64_t c, const Half* src, int64_t srs, int64_t scs, Half* dst, int64_t drs, int64_t dcs, cudaStream_t s) {
        throw_status(elb200_copy2d_f16(r, c, src, srs, scs, dst, drs, dcs, s), "elb200_copy2d_f16");
    }
    static void transpose(int64_t r, int64_t c, const Half* A, int64_t lda, Half* B, int64_t ldb, cudaStream_t s) {
        throw_status(elb200_transpose_f16(r, c, A, lda, B, ldb, s), "elb200_transpose_f16");
    }
    static void axpy2d(int64_t r, int64_t c, Half al, const Half* X, int64_t xrs, int64_t xcs, Half* Y, int64_t yrs, int64_t ycs, cudaStream_t s) {
        throw_status(elb200_axpy2d_f16(r, c, float(al), X, xrs, xcs, Y, yrs, ycs, s), "elb200_axpy2d_f16");
    }
    static void axpy2d_transpose(int64_t r, int64_t c, Half al, const Half* A, int64_t lda, Half* B, int64_t ldb, cudaStream_t s) {
        throw_status(elb200_axpy2d_transpose_f16(r, c, float(al), A, lda, B, ldb, s), "elb200_axpy2d_transpose_f16");
    }
    static void scale2d(int64_t r, int64_t c, Half al, Half* A, int64_t lda, cudaStream_t s) {
        throw_status(elb200_scale2d_f16(r, c, float(al), A, lda, s), "elb200_scale2d_f16");
    }
    static void fill2d(int64_t r, int64_t c, Half v, Half* A, int64_t lda, cudaStream_t s) {
        throw_status(elb200_fill2d_f16(r, c, float(v), A, lda, s), "elb200_fill2d_f16");
    }
    static void lattice(int mode, int nseg, const elb200_segment* segs, Half* M, int64_t ldm, Half* buf, Half al, cudaStream_t s) {
        throw_status(elb200_lattice_f16(mode, nseg, segs, M, ldm, buf, float(al), s), "elb200_lattice_f16");
    }
    static void peer_reduce_axpy(int nsrc, const Half* const* srcs, int64_t off, const elb200_segment* seg, Half* M, int64_t ldm, Half al,
                                 cudaStream_t s) {
        throw_status(elb200_peer_reduce_axpy_f16(nsrc, reinterpret_cast<const void* const*>(srcs), off, seg, M, ldm, float(al), s),
                     "elb200_peer_reduce_axpy_f16");
    }
};

// Stream-ordered scratch (the reference's simple_buffer<T,Device::GPU> on the CUB pool,
// include/hydrogen/utils/SimpleBuffer.hpp:23-123): allocated and freed on the caller's stream.
template <typename T>
struct Scratch {
    T* p = nullptr;
    cudaStream_t s;
    Scratch(int64_t elems, cudaStream_t stream) : s(stream) {
        if (elems > 0) cuda_check(cudaMallocAsync(reinterpret_cast<void**>(&p), size_t(elems) * sizeof(T), s), "cudaMallocAsync(scratch)");
    }
    ~Scratch() { if (p) cudaFreeAsync(p, s); }
    Scratch(const Scratch&) = delete;
    Scratch& operator=(const Scratch&) = delete;
};

}  // namespace elb
