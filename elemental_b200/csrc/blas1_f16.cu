// BLAS-1 entrywise kernels for IEEE half: copy / axpy / scale / fill / transpose on column-major
// matrices with general strides. HBM-bound like their fp32 / fp64 twins in blas1.cu; arithmetic is done in fp32 and
// rounded once (the reference's gpu_half_type routines: include/hydrogen/blas/gpu/{Copy,Axpy,Scale,Fill}.hpp
// instantiated for __half, src/hydrogen/blas/gpu/*.cu). 128-bit accesses (8 halfs) when both sides allow.
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>

#include "elb200_internal.h"

namespace elb {
int check_cuda(cudaError_t e, const char* what);
namespace {

enum : int { H_COPY = 0, H_AXPY = 1, H_SCALE = 2, H_FILL = 3 };

template <int OP>
__device__ __forceinline__ __half apply(float alpha, __half x, __half y) {
    if constexpr (OP == H_COPY) return x;
    else if constexpr (OP == H_AXPY) return __float2half_rn(fmaf(alpha, __half2float(x), __half2float(y)));
    else if constexpr (OP == H_SCALE) return alpha == 0.f ? __float2half_rn(0.f) : __float2half_rn(alpha * __half2float(y));
    else return __float2half_rn(alpha);
}

// unit row strides, 8 halfs per thread along the rows, grid-stride over columns
template <int OP>
__global__ void __launch_bounds__(256) panel_f16_vec(int64_t rows8, int64_t cols, const uint4* __restrict__ src, int64_t lds8,
                                                     uint4* dst, int64_t ldd8, float alpha) {
    const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= rows8) return;
    for (int64_t j = int64_t(blockIdx.y) * blockDim.y + threadIdx.y; j < cols; j += int64_t(gridDim.y) * blockDim.y) {
        uint4 xv = make_uint4(0, 0, 0, 0), yv = make_uint4(0, 0, 0, 0);
        if constexpr (OP == H_COPY || OP == H_AXPY) xv = src[i + j * lds8];
        if constexpr (OP == H_AXPY || OP == H_SCALE) yv = dst[i + j * ldd8];
        uint4 ov;
        const __half* x = reinterpret_cast<const __half*>(&xv);
        const __half* y = reinterpret_cast<const __half*>(&yv);
        __half* o = reinterpret_cast<__half*>(&ov);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = apply<OP>(alpha, x[e], y[e]);
        dst[i + j * ldd8] = ov;
    }
}

// general strides, one element per thread along the rows
template <int OP>
__global__ void __launch_bounds__(256) strided_f16(int64_t rows, int64_t cols, const __half* __restrict__ src, int64_t srs, int64_t scs,
                                                   __half* dst, int64_t drs, int64_t dcs, float alpha) {
    const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= rows) return;
    for (int64_t j = int64_t(blockIdx.y) * blockDim.y + threadIdx.y; j < cols; j += int64_t(gridDim.y) * blockDim.y) {
        __half x = __float2half_rn(0.f), y = __float2half_rn(0.f);
        if constexpr (OP == H_COPY || OP == H_AXPY) x = src[i * srs + j * scs];
        __half* d = dst + i * drs + j * dcs;
        if constexpr (OP == H_AXPY || OP == H_SCALE) y = *d;
        *d = apply<OP>(alpha, x, y);
    }
}

template <int OP>
int launch_f16(int64_t rows, int64_t cols, const __half* src, int64_t srs, int64_t scs, __half* dst, int64_t drs, int64_t dcs,
               float alpha, cudaStream_t s) {
    if (rows <= 0 || cols <= 0) return ELB200_OK;
    const bool needs_src = OP == H_COPY || OP == H_AXPY;
    bool vec = drs == 1 && rows % 8 == 0 && dcs % 8 == 0 && (reinterpret_cast<uintptr_t>(dst) & 15u) == 0;
    if (needs_src) vec = vec && srs == 1 && scs % 8 == 0 && (reinterpret_cast<uintptr_t>(src) & 15u) == 0;
    const int64_t r = vec ? rows / 8 : rows;
    int bx = 32;
    while (bx < 256 && bx < r) bx <<= 1;
    const int by = 256 / bx;
    int64_t gy = (cols + by - 1) / by;
    const int64_t cap = int64_t(device_sm_count()) * 16;
    if (gy > cap) gy = cap;
    dim3 grid(static_cast<unsigned>((r + bx - 1) / bx), static_cast<unsigned>(gy));
    count_launch();
    if (vec)
        panel_f16_vec<OP><<<grid, dim3(bx, by), 0, s>>>(r, cols, reinterpret_cast<const uint4*>(src), scs / 8, reinterpret_cast<uint4*>(dst),
                                                        dcs / 8, alpha);
    else
        strided_f16<OP><<<grid, dim3(bx, by), 0, s>>>(rows, cols, src, srs, scs, dst, drs, dcs, alpha);
    return check_cuda(cudaGetLastError(), "blas1_f16 kernel");
}

// B(cols x rows) = A(rows x cols)^T through a padded 32 x 32 tile: both sides coalesced
__global__ void __launch_bounds__(256) transpose_f16(int64_t rows, int64_t cols, const __half* __restrict__ A, int64_t lda, __half* B,
                                                     int64_t ldb) {
    __shared__ __half tile[32][34];
    const int64_t tiles_r = (rows + 31) / 32, tiles_c = (cols + 31) / 32;
    for (int64_t t = blockIdx.x; t < tiles_r * tiles_c; t += gridDim.x) {
        const int64_t r0 = (t % tiles_r) * 32, c0 = (t / tiles_r) * 32;
        for (int jj = threadIdx.y; jj < 32; jj += blockDim.y) {
            const int64_t i = r0 + threadIdx.x, j = c0 + jj;
            if (i < rows && j < cols) tile[jj][threadIdx.x] = A[i + j * lda];
        }
        __syncthreads();
        for (int ii = threadIdx.y; ii < 32; ii += blockDim.y) {
            const int64_t j = c0 + threadIdx.x, i = r0 + ii;   // B(j, i) = A(i, j)
            if (i < rows && j < cols) B[j + i * ldb] = tile[threadIdx.x][ii];
        }
        __syncthreads();
    }
}

// Table-driven pack / unpack / axpy-unpack (the half instantiation of elb200_lattice_*): one launch moves every segment.
// blockIdx.z = segment; threads walk the packed block (rows down a column are consecutive in the staging buffer and,
// when row_step == 1, in the matrix).
struct SegTableH {
    elb200_segment seg[ELB200_MAX_SEGMENTS];
};
template <int MODE>
__global__ void __launch_bounds__(256) lattice_f16(const __grid_constant__ SegTableH tab, __half* M, int64_t ldm, __half* buf, float alpha) {
    const elb200_segment sg = tab.seg[blockIdx.z];
    for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < sg.rows; i += int64_t(gridDim.x) * blockDim.x)
        for (int64_t j = int64_t(blockIdx.y) * blockDim.y + threadIdx.y; j < sg.cols; j += int64_t(gridDim.y) * blockDim.y) {
            __half* mp = M + sg.row0 + i * sg.row_step + (sg.col0 + j * sg.col_step) * ldm;
            __half* bp = buf + sg.buf_offset + i + j * sg.rows;
            if constexpr (MODE == ELB200_PACK) *bp = *mp;
            else if constexpr (MODE == ELB200_UNPACK) *mp = *bp;
            else *mp = __float2half_rn(fmaf(alpha, __half2float(*bp), __half2float(*mp)));
        }
}

struct PeerSrcsH {
    const __half* p[ELB200_MAX_SEGMENTS];
};
// M lattice += alpha * (sum over the sources, in array order, accumulated in fp32): receiver side of a ReduceScatter
__global__ void __launch_bounds__(256) peer_reduce_axpy_f16(const __grid_constant__ PeerSrcsH srcs, int nsrc, int64_t off,
                                                            const elb200_segment sg, __half* M, int64_t ldm, float alpha) {
    for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < sg.rows; i += int64_t(gridDim.x) * blockDim.x)
        for (int64_t j = int64_t(blockIdx.y) * blockDim.y + threadIdx.y; j < sg.cols; j += int64_t(gridDim.y) * blockDim.y) {
            const int64_t e = off + i + j * sg.rows;  // like the fp32 / fp64 kernels: the unpack segment's buf_offset is not used
            float acc = 0.f;
            for (int q = 0; q < nsrc; ++q) acc += __half2float(srcs.p[q][e]);
            __half* mp = M + sg.row0 + i * sg.row_step + (sg.col0 + j * sg.col_step) * ldm;
            *mp = __float2half_rn(fmaf(alpha, acc, __half2float(*mp)));
        }
}

// B(rows x cols) += alpha * A(cols x rows)^T through a padded tile
__global__ void __launch_bounds__(256) axpy_transpose_f16(int64_t rows, int64_t cols, float alpha, const __half* __restrict__ A, int64_t lda,
                                                          __half* B, int64_t ldb) {
    __shared__ __half tile[32][34];
    const int64_t tiles_r = (rows + 31) / 32, tiles_c = (cols + 31) / 32;
    for (int64_t t = blockIdx.x; t < tiles_r * tiles_c; t += gridDim.x) {
        const int64_t r0 = (t % tiles_r) * 32, c0 = (t / tiles_r) * 32;   // tile of B: rows r0.., cols c0..
        for (int ii = threadIdx.y; ii < 32; ii += blockDim.y) {            // A(j, i): j contiguous
            const int64_t j = c0 + threadIdx.x, i = r0 + ii;
            if (i < rows && j < cols) tile[ii][threadIdx.x] = A[j + i * lda];
        }
        __syncthreads();
        for (int jj = threadIdx.y; jj < 32; jj += blockDim.y) {
            const int64_t i = r0 + threadIdx.x, j = c0 + jj;
            if (i < rows && j < cols) {
                __half* bp = B + i + j * ldb;
                *bp = __float2half_rn(fmaf(alpha, __half2float(tile[threadIdx.x][jj]), __half2float(*bp)));
            }
        }
        __syncthreads();
    }
}

dim3 grid_for(int64_t rows, int64_t cols, int nz, dim3* block) {
    int bx = 32;
    while (bx < 256 && bx < rows) bx <<= 1;
    const int by = 256 / bx;
    *block = dim3(bx, by);
    int64_t gx = (rows + bx - 1) / bx, gy = (cols + by - 1) / by;
    const int64_t cap = int64_t(device_sm_count()) * 16;
    if (gx > cap) gx = cap;
    if (gy > cap / (gx < cap ? gx : cap) + 1) gy = cap / gx + 1;
    if (gy > 65535) gy = 65535;
    return dim3(unsigned(gx), unsigned(gy), unsigned(nz));
}

}  // namespace
}  // namespace elb

using namespace elb;

extern "C" {

int elb200_copy2d_f16(int64_t rows, int64_t cols, const void* src, int64_t srs, int64_t scs, void* dst, int64_t drs, int64_t dcs,
                      void* stream) {
    if (rows < 0 || cols < 0 || (rows > 0 && cols > 0 && (!src || !dst))) return ELB200_ERR_ARG;
    return launch_f16<H_COPY>(rows, cols, static_cast<const __half*>(src), srs, scs, static_cast<__half*>(dst), drs, dcs, 0.f,
                              static_cast<cudaStream_t>(stream));
}
int elb200_axpy2d_f16(int64_t rows, int64_t cols, float alpha, const void* X, int64_t xrs, int64_t xcs, void* Y, int64_t yrs,
                      int64_t ycs, void* stream) {
    if (rows < 0 || cols < 0 || (rows > 0 && cols > 0 && (!X || !Y))) return ELB200_ERR_ARG;
    return launch_f16<H_AXPY>(rows, cols, static_cast<const __half*>(X), xrs, xcs, static_cast<__half*>(Y), yrs, ycs, alpha,
                              static_cast<cudaStream_t>(stream));
}
int elb200_scale2d_f16(int64_t rows, int64_t cols, float alpha, void* A, int64_t lda, void* stream) {
    if (rows < 0 || cols < 0 || lda < (rows > 1 ? rows : 1) || (rows > 0 && cols > 0 && !A)) return ELB200_ERR_ARG;
    if (alpha == 1.f) return ELB200_OK;
    return launch_f16<H_SCALE>(rows, cols, nullptr, 1, 0, static_cast<__half*>(A), 1, lda, alpha, static_cast<cudaStream_t>(stream));
}
int elb200_fill2d_f16(int64_t rows, int64_t cols, float value, void* A, int64_t lda, void* stream) {
    if (rows < 0 || cols < 0 || lda < (rows > 1 ? rows : 1) || (rows > 0 && cols > 0 && !A)) return ELB200_ERR_ARG;
    return launch_f16<H_FILL>(rows, cols, nullptr, 1, 0, static_cast<__half*>(A), 1, lda, value, static_cast<cudaStream_t>(stream));
}
int elb200_transpose_f16(int64_t rows, int64_t cols, const void* A, int64_t lda, void* B, int64_t ldb, void* stream) {
    if (rows < 0 || cols < 0 || lda < (rows > 1 ? rows : 1) || ldb < (cols > 1 ? cols : 1)) return ELB200_ERR_ARG;
    if (rows == 0 || cols == 0) return ELB200_OK;
    if (!A || !B) return ELB200_ERR_ARG;
    const int64_t tiles = ((rows + 31) / 32) * ((cols + 31) / 32);
    const int64_t cap = int64_t(device_sm_count()) * 8;
    count_launch();
    transpose_f16<<<static_cast<unsigned>(tiles < cap ? tiles : cap), dim3(32, 8), 0, static_cast<cudaStream_t>(stream)>>>(
        rows, cols, static_cast<const __half*>(A), lda, static_cast<__half*>(B), ldb);
    return check_cuda(cudaGetLastError(), "transpose_f16");
}

int elb200_axpy2d_transpose_f16(int64_t rows, int64_t cols, float alpha, const void* A, int64_t lda, void* B, int64_t ldb, void* stream) {
    if (rows < 0 || cols < 0 || lda < (cols > 1 ? cols : 1) || ldb < (rows > 1 ? rows : 1)) return ELB200_ERR_ARG;
    if (rows == 0 || cols == 0) return ELB200_OK;
    if (!A || !B) return ELB200_ERR_ARG;
    const int64_t tiles = ((rows + 31) / 32) * ((cols + 31) / 32);
    const int64_t cap = int64_t(device_sm_count()) * 8;
    count_launch();
    axpy_transpose_f16<<<static_cast<unsigned>(tiles < cap ? tiles : cap), dim3(32, 8), 0, static_cast<cudaStream_t>(stream)>>>(
        rows, cols, alpha, static_cast<const __half*>(A), lda, static_cast<__half*>(B), ldb);
    return check_cuda(cudaGetLastError(), "axpy_transpose_f16");
}

int elb200_lattice_f16(int mode, int nseg, const elb200_segment* segs, void* M, int64_t ldm, void* buf, float alpha, void* stream) {
    if (mode < ELB200_PACK || mode > ELB200_UNPACK_AXPY || nseg < 0 || nseg > ELB200_MAX_SEGMENTS || (nseg > 0 && !segs)) return ELB200_ERR_ARG;
    SegTableH tab;
    int n = 0;
    int64_t max_rows = 0, max_cols = 0;
    for (int q = 0; q < nseg; ++q) {
        const elb200_segment& sg = segs[q];
        if (sg.rows < 0 || sg.cols < 0 || sg.row_step < 1 || sg.col_step < 1 || sg.row0 < 0 || sg.col0 < 0) return ELB200_ERR_ARG;
        if (sg.rows == 0 || sg.cols == 0) continue;
        tab.seg[n++] = sg;
        if (sg.rows > max_rows) max_rows = sg.rows;
        if (sg.cols > max_cols) max_cols = sg.cols;
    }
    if (n == 0) return ELB200_OK;
    if (!M || !buf) return ELB200_ERR_ARG;
    dim3 block;
    const dim3 grid = grid_for(max_rows, max_cols, n, &block);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    count_launch();
    if (mode == ELB200_PACK) lattice_f16<ELB200_PACK><<<grid, block, 0, s>>>(tab, static_cast<__half*>(M), ldm, static_cast<__half*>(buf), alpha);
    else if (mode == ELB200_UNPACK) lattice_f16<ELB200_UNPACK><<<grid, block, 0, s>>>(tab, static_cast<__half*>(M), ldm, static_cast<__half*>(buf), alpha);
    else lattice_f16<ELB200_UNPACK_AXPY><<<grid, block, 0, s>>>(tab, static_cast<__half*>(M), ldm, static_cast<__half*>(buf), alpha);
    return check_cuda(cudaGetLastError(), "lattice_f16");
}

int elb200_peer_reduce_axpy_f16(int nsrc, const void* const* srcs, int64_t src_offset, const elb200_segment* seg, void* M, int64_t ldm,
                                float alpha, void* stream) {
    if (nsrc < 1 || nsrc > ELB200_MAX_SEGMENTS || !srcs || !seg || seg->rows < 0 || seg->cols < 0 || seg->row_step < 1 || seg->col_step < 1)
        return ELB200_ERR_ARG;
    if (seg->rows == 0 || seg->cols == 0) return ELB200_OK;
    if (!M) return ELB200_ERR_ARG;
    PeerSrcsH ps;
    for (int q = 0; q < nsrc; ++q) ps.p[q] = static_cast<const __half*>(srcs[q]);
    dim3 block;
    const dim3 grid = grid_for(seg->rows, seg->cols, 1, &block);
    count_launch();
    peer_reduce_axpy_f16<<<grid, block, 0, static_cast<cudaStream_t>(stream)>>>(ps, nsrc, src_offset, *seg, static_cast<__half*>(M), ldm, alpha);
    return check_cuda(cudaGetLastError(), "peer_reduce_axpy_f16");
}

}  // extern "C"
