// BLAS-1 entrywise kernels of the El::Gemm path, written for HBM bandwidth on sm_100a:
// copy / axpy / scale / fill over column-major panels (128-bit accesses when alignment allows),
// 32x32 shared-memory transposes, and the table-driven lattice pack/unpack that moves every
// per-peer portion of an element-cyclic redistribution in ONE launch.
//
// Reference kernels replaced (all naive 32x8 tiles, no vector accesses):
//   src/hydrogen/blas/gpu/{Copy,Axpy,Scale,Fill,Transpose}.cu and the cublas?geam calls in
//   include/hydrogen/blas/GPU_BLAS_impl.hpp:110-198,254-273.
#include <cuda_runtime.h>
#include <algorithm>
#include <cstdint>
#include <type_traits>

#include "elb200_internal.h"

namespace elb {
int check_cuda(cudaError_t e, const char* what);

namespace {

template <typename T> struct VecOf;
template <> struct VecOf<float> { using type = float4; static constexpr int N = 4; };
template <> struct VecOf<double> { using type = double2; static constexpr int N = 2; };

__device__ __forceinline__ float4 vaxpy(float a, float4 x, float4 y) {
    return make_float4(fmaf(a, x.x, y.x), fmaf(a, x.y, y.y), fmaf(a, x.z, y.z), fmaf(a, x.w, y.w));
}
__device__ __forceinline__ double2 vaxpy(double a, double2 x, double2 y) {
    return make_double2(fma(a, x.x, y.x), fma(a, x.y, y.y));
}
__device__ __forceinline__ float4 vscale(float a, float4 x) {
    return make_float4(a * x.x, a * x.y, a * x.z, a * x.w);
}
__device__ __forceinline__ double2 vscale(double a, double2 x) { return make_double2(a * x.x, a * x.y); }
__device__ __forceinline__ float4 vsplat(float a) { return make_float4(a, a, a, a); }
__device__ __forceinline__ double2 vsplat(double a) { return make_double2(a, a); }

enum : int { OP_COPY = 0, OP_AXPY = 1, OP_SCALE = 2, OP_FILL = 3 };

// Column-contiguous panels: element (i,j) at base + i + j*ld. One thread handles one V-wide
// row chunk of kCols consecutive columns (kCols independent loads in flight).
constexpr int kColsPerThread = 4;

template <typename T, typename V, int OP>
__global__ void __launch_bounds__(256)
panel_kernel(int64_t rows_v, int64_t cols, const V* __restrict__ src, int64_t lds_v, V* dst,
             int64_t ldd_v, T alpha) {
    const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    const int64_t j0 = (int64_t(blockIdx.y) * blockDim.y + threadIdx.y) * kColsPerThread;
    if (i >= rows_v) return;
    V x[kColsPerThread], y[kColsPerThread];
#pragma unroll
    for (int c = 0; c < kColsPerThread; ++c) {
        const int64_t j = j0 + c;
        if (j < cols) {
            if constexpr (OP == OP_COPY || OP == OP_AXPY) x[c] = src[i + j * lds_v];
            if constexpr (OP == OP_AXPY || OP == OP_SCALE) y[c] = dst[i + j * ldd_v];
        }
    }
#pragma unroll
    for (int c = 0; c < kColsPerThread; ++c) {
        const int64_t j = j0 + c;
        if (j < cols) {
            V o;
            if constexpr (OP == OP_COPY) o = x[c];
            else if constexpr (OP == OP_AXPY) {
                if constexpr (std::is_same<V, T>::value) o = fma(alpha, x[c], y[c]);
                else o = vaxpy(alpha, x[c], y[c]);
            } else if constexpr (OP == OP_SCALE) {
                if constexpr (std::is_same<V, T>::value) o = alpha * y[c];
                else o = vscale(alpha, y[c]);
            } else {
                if constexpr (std::is_same<V, T>::value) o = alpha;
                else o = vsplat(alpha);
            }
            dst[i + j * ldd_v] = o;
        }
    }
}

template <typename T, int OP>
int launch_panel(int64_t rows, int64_t cols, const T* src, int64_t lds, T* dst, int64_t ldd,
                 T alpha, cudaStream_t s) {
    if (rows <= 0 || cols <= 0) return ELB200_OK;
    using V = typename VecOf<T>::type;
    constexpr int N = VecOf<T>::N;
    const bool needs_src = (OP == OP_COPY || OP == OP_AXPY);
    bool vec = (rows % N == 0) && (ldd % N == 0) && ((reinterpret_cast<uintptr_t>(dst) & 15u) == 0);
    if (needs_src) vec = vec && (lds % N == 0) && ((reinterpret_cast<uintptr_t>(src) & 15u) == 0);
    const int64_t rows_v = vec ? rows / N : rows;
    // threads along rows: up to 256; remaining threads of the block along columns
    int bx = 32;
    while (bx < 256 && bx < rows_v) bx <<= 1;
    const int by = 256 / bx;
    dim3 block(bx, by);
    const int64_t gx = (rows_v + bx - 1) / bx;
    const int64_t gy = (cols + int64_t(by) * kColsPerThread - 1) / (int64_t(by) * kColsPerThread);
    if (gy > 65535) {  // very wide & short: fall back to per-slab launches
        const int64_t slab = int64_t(65535) * by * kColsPerThread;
        for (int64_t c0 = 0; c0 < cols; c0 += slab) {
            const int64_t cc = (cols - c0 < slab) ? cols - c0 : slab;
            int st = launch_panel<T, OP>(rows, cc, needs_src ? src + c0 * lds : src, lds,
                                         dst + c0 * ldd, ldd, alpha, s);
            if (st != ELB200_OK) return st;
        }
        return ELB200_OK;
    }
    dim3 grid(static_cast<unsigned>(gx), static_cast<unsigned>(gy));
    count_launch();
    if (vec)
        panel_kernel<T, V, OP><<<grid, block, 0, s>>>(rows_v, cols, reinterpret_cast<const V*>(src),
                                                      lds / N, reinterpret_cast<V*>(dst), ldd / N,
                                                      alpha);
    else
        panel_kernel<T, T, OP><<<grid, block, 0, s>>>(rows_v, cols, src, lds, dst, ldd, alpha);
    return check_cuda(cudaGetLastError(), "panel_kernel");
}

// General 2-D strides (non-unit row stride on either side): scalar, threads along rows.
template <typename T, bool AXPY>
__global__ void __launch_bounds__(256)
strided_kernel(int64_t rows, int64_t cols, const T* __restrict__ src, int64_t srs, int64_t scs,
               T* dst, int64_t drs, int64_t dcs, T alpha) {
    const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= rows) return;
    for (int64_t j = int64_t(blockIdx.y) * blockDim.y + threadIdx.y; j < cols;
         j += int64_t(gridDim.y) * blockDim.y) {
        const T x = src[i * srs + j * scs];
        T* d = dst + i * drs + j * dcs;
        if constexpr (AXPY) *d = fma(alpha, x, *d);
        else *d = x;
    }
}

template <typename T, bool AXPY>
int launch_strided(int64_t rows, int64_t cols, const T* src, int64_t srs, int64_t scs, T* dst,
                   int64_t drs, int64_t dcs, T alpha, cudaStream_t s) {
    if (rows <= 0 || cols <= 0) return ELB200_OK;
    int bx = 32;
    while (bx < 256 && bx < rows) bx <<= 1;
    const int by = 256 / bx;
    int64_t gy = (cols + by - 1) / by;
    if (gy > 4096) gy = 4096;
    dim3 grid(static_cast<unsigned>((rows + bx - 1) / bx), static_cast<unsigned>(gy));
    count_launch();
    strided_kernel<T, AXPY><<<grid, dim3(bx, by), 0, s>>>(rows, cols, src, srs, scs, dst, drs, dcs,
                                                          alpha);
    return check_cuda(cudaGetLastError(), "strided_kernel");
}

// B(cols x rows) (+)= alpha * A(rows x cols)^T through a padded 32x32 smem tile; both global
// sides are coalesced 128-byte rows.
template <typename T, bool AXPY>
__global__ void __launch_bounds__(256)
transpose_kernel(int64_t rows, int64_t cols, const T* __restrict__ A, int64_t lda, T* B,
                 int64_t ldb, T alpha) {
    __shared__ T tile[32][33];
    const int64_t r0 = int64_t(blockIdx.x) * 32, c0 = int64_t(blockIdx.y) * 32;
#pragma unroll
    for (int jj = 0; jj < 32; jj += 8) {
        const int64_t r = r0 + threadIdx.x, c = c0 + threadIdx.y + jj;
        if (r < rows && c < cols) tile[threadIdx.y + jj][threadIdx.x] = A[r + c * lda];
    }
    __syncthreads();
#pragma unroll
    for (int jj = 0; jj < 32; jj += 8) {
        // B(c, r) = A(r, c): B row index = c (contiguous), B column index = r
        const int64_t c = c0 + threadIdx.x, r = r0 + threadIdx.y + jj;
        if (r < rows && c < cols) {
            const T v = tile[threadIdx.x][threadIdx.y + jj];
            T* d = B + c + r * ldb;
            if constexpr (AXPY) *d = fma(alpha, v, *d);
            else *d = v;
        }
    }
}

// 64 x 64 tile, 16-byte global accesses on both sides (4 fp32 / 2 fp64 per thread and access):
// enough bytes in flight per SM to reach the HBM roof also for 4-byte elements.
template <typename T, bool AXPY>
__global__ void __launch_bounds__(256)
transpose_vec_kernel(int64_t rows, int64_t cols, const T* __restrict__ A, int64_t lda, T* B, int64_t ldb, T alpha) {
    using V = typename VecOf<T>::type;
    constexpr int N = VecOf<T>::N, TS = 64, TX = TS / N, TY = 256 / TX;
    __shared__ T tile[TS][TS + 1];
    const int64_t r0 = int64_t(blockIdx.x) * TS, c0 = int64_t(blockIdx.y) * TS;
    const int tx = threadIdx.x % TX, ty = threadIdx.x / TX;
#pragma unroll
    for (int cc = ty; cc < TS; cc += TY) {
        const int64_t r = r0 + tx * N, c = c0 + cc;
        if (r < rows && c < cols) {
            T e[N];
            *reinterpret_cast<V*>(e) = *reinterpret_cast<const V*>(A + r + c * lda);
#pragma unroll
            for (int i = 0; i < N; ++i) tile[cc][tx * N + i] = e[i];
        }
    }
    __syncthreads();
#pragma unroll
    for (int rr = ty; rr < TS; rr += TY) {
        const int64_t c = c0 + tx * N, r = r0 + rr;  // B(c, r) = A(r, c)
        if (r < rows && c < cols) {
            T e[N];
#pragma unroll
            for (int i = 0; i < N; ++i) e[i] = tile[tx * N + i][rr];
            V* d = reinterpret_cast<V*>(B + c + r * ldb);
            if constexpr (AXPY) *d = vaxpy(alpha, *reinterpret_cast<V*>(e), *d);
            else *d = *reinterpret_cast<V*>(e);
        }
    }
}

template <typename T, bool AXPY>
int launch_transpose(int64_t rows, int64_t cols, const T* A, int64_t lda, T* B, int64_t ldb,
                     T alpha, cudaStream_t s) {
    if (rows <= 0 || cols <= 0) return ELB200_OK;
    constexpr int NV = VecOf<T>::N;
    // (fp64 already reaches the HBM roof with the 32 x 32 tile below: measured 6633 vs 5401 GB/s)
    if (sizeof(T) == 4 && rows % NV == 0 && cols % NV == 0 && lda % NV == 0 && ldb % NV == 0 &&
        (reinterpret_cast<uintptr_t>(A) & 15u) == 0 && (reinterpret_cast<uintptr_t>(B) & 15u) == 0 &&
        (cols + 63) / 64 <= 65535) {
        count_launch();
        transpose_vec_kernel<T, AXPY><<<dim3(unsigned((rows + 63) / 64), unsigned((cols + 63) / 64)), 256, 0, s>>>(
            rows, cols, A, lda, B, ldb, alpha);
        return check_cuda(cudaGetLastError(), "transpose_vec_kernel");
    }
    const int64_t gx = (rows + 31) / 32, gy = (cols + 31) / 32;
    if (gy > 65535) {
        const int64_t slab = int64_t(65535) * 32;
        for (int64_t c0 = 0; c0 < cols; c0 += slab) {
            const int64_t cc = (cols - c0 < slab) ? cols - c0 : slab;
            int st = launch_transpose<T, AXPY>(rows, cc, A + c0 * lda, lda, B + c0, ldb, alpha, s);
            if (st != ELB200_OK) return st;
        }
        return ELB200_OK;
    }
    count_launch();
    transpose_kernel<T, AXPY><<<dim3(unsigned(gx), unsigned(gy)), dim3(32, 8), 0, s>>>(
        rows, cols, A, lda, B, ldb, alpha);
    return check_cuda(cudaGetLastError(), "transpose_kernel");
}

// ------------------------------------------------------------------ lattice pack / unpack
struct SegTable {
    elb200_segment seg[ELB200_MAX_SEGMENTS];
};

// grid: x = row tiles, y = column tiles (strided loop), z = segment.
// V = T (scalar) or the 16-byte vector of T (requires row_step == 1 and aligned rows).
template <typename T, typename V, int MODE>
__global__ void __launch_bounds__(256)
lattice_kernel(const __grid_constant__ SegTable tab, T* M, int64_t ldm, T* buf, T alpha) {
    constexpr int N = sizeof(V) / sizeof(T);
    const elb200_segment sg = tab.seg[blockIdx.z];
    const int64_t rows_v = sg.rows / N;
    for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < rows_v; i += int64_t(gridDim.x) * blockDim.x) {
        const int64_t row_off = (N == 1) ? sg.row0 + i * sg.row_step : sg.row0 + i * N;
        for (int64_t j = int64_t(blockIdx.y) * blockDim.y + threadIdx.y; j < sg.cols;
             j += int64_t(gridDim.y) * blockDim.y) {
            V* mp = reinterpret_cast<V*>(M + row_off + (sg.col0 + j * sg.col_step) * ldm);
            V* bp = reinterpret_cast<V*>(buf + sg.buf_offset + i * N + j * sg.rows);
            if constexpr (MODE == ELB200_PACK) *bp = *mp;
            else if constexpr (MODE == ELB200_UNPACK) *mp = *bp;
            else {
                if constexpr (N == 1) *mp = fma(alpha, *bp, *mp);
                else *mp = vaxpy(alpha, *bp, *mp);
            }
        }
    }
}

// Row-interleaved portions (ColStridedPack / ColStridedUnpack, Copy/util.hpp:167-230): segment q
// owns the rows I of the matrix with I mod S == residue_q.  Walking each segment on its own touches
// one element per S of every 32-byte sector; walking the MATRIX rows instead (thread <-> row I,
// segment looked up from I mod S) makes the matrix side fully coalesced and the staging side S
// interleaved streams of consecutive elements, so every sector moved is fully used.
struct InterleaveTable {
    int64_t buf_offset[ELB200_MAX_SEGMENTS];  // by residue; -1: nobody owns that residue
    int64_t seg_rows[ELB200_MAX_SEGMENTS];
    int64_t first_row[ELB200_MAX_SEGMENTS];
    int64_t S, col0, col_step, cols, rows_total;
};
template <typename T, int MODE>
__global__ void __launch_bounds__(256)
interleave_kernel(const __grid_constant__ InterleaveTable tab, T* M, int64_t ldm, T* buf, T alpha) {
    for (int64_t I = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; I < tab.rows_total; I += int64_t(gridDim.x) * blockDim.x) {
        const int res = int(I % tab.S);
        const int64_t off = tab.buf_offset[res];
        if (off < 0 || I < tab.first_row[res]) continue;
        const int64_t i = (I - tab.first_row[res]) / tab.S, nr = tab.seg_rows[res];
        if (i >= nr) continue;
        // four columns per trip: four independent loads in flight before the first store
        constexpr int U = 4;
        for (int64_t j0 = int64_t(blockIdx.y) * U; j0 < tab.cols; j0 += int64_t(gridDim.y) * U) {
            T v[U], w[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t j = j0 + u;
                if (j < tab.cols) {
                    const T* mp = M + I + (tab.col0 + j * tab.col_step) * ldm;
                    const T* bp = buf + off + i + j * nr;
                    if constexpr (MODE == ELB200_PACK) v[u] = *mp;
                    else v[u] = *bp;
                    if constexpr (MODE == ELB200_UNPACK_AXPY) w[u] = *mp;
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t j = j0 + u;
                if (j < tab.cols) {
                    T* mp = M + I + (tab.col0 + j * tab.col_step) * ldm;
                    T* bp = buf + off + i + j * nr;
                    if constexpr (MODE == ELB200_PACK) *bp = v[u];
                    else if constexpr (MODE == ELB200_UNPACK) *mp = v[u];
                    else *mp = fma(alpha, v[u], w[u]);
                }
            }
        }
    }
}
// true (and the launch done) when the segment set is a row interleave
template <typename T>
bool try_interleave(int mode, int nseg, const elb200_segment* segs, T* M, int64_t ldm, T* buf, T alpha, cudaStream_t s,
                    int* status) {
    if (nseg < 2) return false;
    const int64_t S = segs[0].row_step;
    if (S < 2 || S > ELB200_MAX_SEGMENTS) return false;
    InterleaveTable tab;
    for (int q = 0; q < ELB200_MAX_SEGMENTS; ++q) { tab.buf_offset[q] = -1; tab.seg_rows[q] = 0; tab.first_row[q] = 0; }
    tab.S = S; tab.col0 = segs[0].col0; tab.col_step = segs[0].col_step; tab.cols = segs[0].cols; tab.rows_total = 0;
    int live = 0;
    for (int k = 0; k < nseg; ++k) {
        const elb200_segment& g = segs[k];
        if (g.rows == 0 || g.cols == 0) continue;
        if (g.row_step != S || g.col0 != tab.col0 || g.col_step != tab.col_step || g.cols != tab.cols) return false;
        const int res = int(g.row0 % S);
        if (tab.buf_offset[res] >= 0) return false;  // two segments on one residue: not an interleave
        tab.buf_offset[res] = g.buf_offset; tab.seg_rows[res] = g.rows; tab.first_row[res] = g.row0;
        tab.rows_total = std::max(tab.rows_total, g.row0 + (g.rows - 1) * S + 1);
        ++live;
    }
    if (live < 2 || tab.cols <= 0) return false;
    const int64_t gx = std::min<int64_t>((tab.rows_total + 255) / 256, 1024);
    const int64_t gy = std::min<int64_t>((tab.cols + 3) / 4, std::max<int64_t>(1, 16 * device_sm_count() / gx));
    dim3 grid(static_cast<unsigned>(gx), static_cast<unsigned>(gy));
    count_launch();
    if (mode == ELB200_PACK) interleave_kernel<T, ELB200_PACK><<<grid, 256, 0, s>>>(tab, M, ldm, buf, alpha);
    else if (mode == ELB200_UNPACK) interleave_kernel<T, ELB200_UNPACK><<<grid, 256, 0, s>>>(tab, M, ldm, buf, alpha);
    else interleave_kernel<T, ELB200_UNPACK_AXPY><<<grid, 256, 0, s>>>(tab, M, ldm, buf, alpha);
    *status = check_cuda(cudaGetLastError(), "interleave_kernel");
    return true;
}

template <typename T>
int launch_lattice(int mode, int nseg, const elb200_segment* segs, T* M, int64_t ldm, T* buf,
                   T alpha, cudaStream_t s) {
    if (nseg < 0 || nseg > ELB200_MAX_SEGMENTS || (nseg > 0 && !segs)) return ELB200_ERR_ARG;
    if (mode < 0 || mode > 2) return ELB200_ERR_ARG;
    for (int k = 0; k < nseg; ++k)
        if (segs[k].rows < 0 || segs[k].cols < 0 || segs[k].row_step < 1 || segs[k].col_step < 1) return ELB200_ERR_ARG;
    {
        int st = ELB200_OK;
        if (try_interleave<T>(mode, nseg, segs, M, ldm, buf, alpha, s, &st)) return st;
    }
    using V = typename VecOf<T>::type;
    constexpr int N = VecOf<T>::N;
    SegTable tab;
    int n = 0;
    int64_t max_rows = 0, max_cols = 0;
    bool vec = (ldm % N == 0) && ((reinterpret_cast<uintptr_t>(M) & 15u) == 0) &&
               ((reinterpret_cast<uintptr_t>(buf) & 15u) == 0);
    for (int k = 0; k < nseg; ++k) {
        const elb200_segment& g = segs[k];
        if (g.rows < 0 || g.cols < 0 || g.row_step < 1 || g.col_step < 1) return ELB200_ERR_ARG;
        if (g.rows == 0 || g.cols == 0) continue;
        tab.seg[n++] = g;
        if (g.rows > max_rows) max_rows = g.rows;
        if (g.cols > max_cols) max_cols = g.cols;
        vec = vec && g.row_step == 1 && (g.rows % N == 0) && (g.row0 % N == 0) &&
              (g.buf_offset % N == 0);
    }
    if (n == 0) return ELB200_OK;
    const int64_t rows_v = vec ? max_rows / N : max_rows;
    int bx = 32;
    while (bx < 256 && bx < rows_v) bx <<= 1;
    const int by = 256 / bx;
    int64_t gx = (rows_v + bx - 1) / bx;
    int64_t gy = (max_cols + by - 1) / by;
    if (gy > 1024) gy = 1024;
    dim3 grid(static_cast<unsigned>(gx), static_cast<unsigned>(gy), static_cast<unsigned>(n));
    dim3 block(bx, by);
    count_launch();
#define ELB_LAT(MODE_)                                                                  \
    if (vec) lattice_kernel<T, V, MODE_><<<grid, block, 0, s>>>(tab, M, ldm, buf, alpha); \
    else lattice_kernel<T, T, MODE_><<<grid, block, 0, s>>>(tab, M, ldm, buf, alpha);
    if (mode == ELB200_PACK) { ELB_LAT(ELB200_PACK) }
    else if (mode == ELB200_UNPACK) { ELB_LAT(ELB200_UNPACK) }
    else { ELB_LAT(ELB200_UNPACK_AXPY) }
#undef ELB_LAT
    return check_cuda(cudaGetLastError(), "lattice_kernel");
}

// B-lattice += alpha * sum over sources of their packed block: the ReduceScatter of AxpyContract
// done by the RECEIVER with loads over NVLink peer mappings, fused with the axpy-unpack.
// Sources are summed in communicator order on every rank (deterministic).
template <typename T>
struct PeerSources {
    const T* p[ELB200_MAX_SEGMENTS];
    int n;
};
template <typename T>
__global__ void __launch_bounds__(256)
peer_reduce_axpy_kernel(const __grid_constant__ PeerSources<T> src, int64_t off, const elb200_segment sg, T* M,
                        int64_t ldm, T alpha) {
    for (int64_t j = blockIdx.y; j < sg.cols; j += gridDim.y)
        for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < sg.rows; i += int64_t(gridDim.x) * blockDim.x) {
            const int64_t e = off + i + j * sg.rows;
            T sum = T(0);
            for (int s = 0; s < src.n; ++s) sum += src.p[s][e];
            T* mp = M + sg.row0 + i * sg.row_step + (sg.col0 + j * sg.col_step) * ldm;
            *mp = fma(alpha, sum, *mp);
        }
}
template <typename T>
__global__ void __launch_bounds__(256)
peer_reduce_axpy_vec_kernel(const __grid_constant__ PeerSources<T> src, int64_t off, const elb200_segment sg, T* M,
                            int64_t ldm, T alpha) {
    using V = typename VecOf<T>::type;
    constexpr int N = VecOf<T>::N;
    const int64_t rows_v = sg.rows / N;
    for (int64_t j = blockIdx.y; j < sg.cols; j += gridDim.y)
        for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < rows_v; i += int64_t(gridDim.x) * blockDim.x) {
            const int64_t e = off + i * N + j * sg.rows;
            T acc[N];
#pragma unroll
            for (int t = 0; t < N; ++t) acc[t] = T(0);
            for (int s = 0; s < src.n; ++s) {
                T v[N];
                *reinterpret_cast<V*>(v) = *reinterpret_cast<const V*>(src.p[s] + e);
#pragma unroll
                for (int t = 0; t < N; ++t) acc[t] += v[t];
            }
            V* mp = reinterpret_cast<V*>(M + sg.row0 + i * N + (sg.col0 + j * sg.col_step) * ldm);
            *mp = vaxpy(alpha, *reinterpret_cast<V*>(acc), *mp);
        }
}

template <typename T>
int peer_reduce_axpy(int nsrc, const T* const* srcs, int64_t off, const elb200_segment* seg, T* M, int64_t ldm, T alpha,
                     void* stream) {
    if (nsrc < 1 || nsrc > ELB200_MAX_SEGMENTS || !srcs || !seg || seg->rows < 0 || seg->cols < 0 || seg->row_step < 1 ||
        seg->col_step < 1)
        return ELB200_ERR_ARG;
    if (seg->rows == 0 || seg->cols == 0) return ELB200_OK;
    PeerSources<T> ps;
    ps.n = nsrc;
    for (int i = 0; i < nsrc; ++i) ps.p[i] = srcs[i];
    constexpr int NV = VecOf<T>::N;
    bool vec = seg->row_step == 1 && seg->rows % NV == 0 && seg->row0 % NV == 0 && off % NV == 0 && ldm % NV == 0 &&
               (reinterpret_cast<uintptr_t>(M) & 15u) == 0;
    for (int i = 0; i < nsrc; ++i) vec = vec && (reinterpret_cast<uintptr_t>(srcs[i]) & 15u) == 0;
    if (vec) {
        const int64_t rows_v = seg->rows / NV;
        const int64_t vx = std::min<int64_t>((rows_v + 255) / 256, 64);
        const int64_t vy = std::min<int64_t>(seg->cols, std::max<int64_t>(1, 16 * device_sm_count() / vx));
        count_launch();
        peer_reduce_axpy_vec_kernel<T><<<dim3(static_cast<unsigned>(vx), static_cast<unsigned>(vy)), 256, 0,
                                         static_cast<cudaStream_t>(stream)>>>(ps, off, *seg, M, ldm, alpha);
        return check_cuda(cudaGetLastError(), "peer_reduce_axpy_vec_kernel");
    }
    const int64_t gx = std::min<int64_t>((seg->rows + 255) / 256, 64);
    const int64_t gy = std::min<int64_t>(seg->cols, std::max<int64_t>(1, 16 * device_sm_count() / gx));
    count_launch();
    peer_reduce_axpy_kernel<T><<<dim3(static_cast<unsigned>(gx), static_cast<unsigned>(gy)), 256, 0,
                                 static_cast<cudaStream_t>(stream)>>>(ps, off, *seg, M, ldm, alpha);
    return check_cuda(cudaGetLastError(), "peer_reduce_axpy_kernel");
}

// Trapezoid-masked entrywise ops on the LOCAL part of an element-cyclic matrix: local (i,j) is global
// (row0 + i*row_step, col0 + j*col_step); the op touches the entries inside (SCALE, AXPY) or outside
// (MAKE: zero them) the trapezoid  LOWER: gi >= gj - offset,  UPPER: gi <= gj - offset.
// Replaces the host loops of ScaleTrapezoid / AxpyTrapezoid / LocalAxpyTrapezoid / MakeTrapezoidal
// (include/El/blas_like/level1/{ScaleTrapezoid,AxpyTrapezoid,MakeTrapezoidal}.hpp), which the
// reference runs per column on the CPU.  Threads run down the columns (coalesced); entries on the
// other side of the boundary are neither read nor written.
enum : int { TRAP_SCALE = 0, TRAP_AXPY = 1, TRAP_MAKE = 2 };
template <typename T, int OP>
__global__ void __launch_bounds__(256)
trapezoid_kernel(int64_t rows, int64_t cols, int upper, int64_t offset, int64_t row0, int64_t row_step, int64_t col0,
                 int64_t col_step, T alpha, const T* __restrict__ X, int64_t ldx, T* Y, int64_t ldy) {
    const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= rows) return;
    const int64_t gi = row0 + i * row_step;
    for (int64_t j = int64_t(blockIdx.y) * blockDim.y + threadIdx.y; j < cols; j += int64_t(gridDim.y) * blockDim.y) {
        const int64_t gj = col0 + j * col_step;
        const bool inside = upper ? (gi <= gj - offset) : (gi >= gj - offset);
        T* y = Y + i + j * ldy;
        if constexpr (OP == TRAP_SCALE) {
            if (inside) *y = alpha == T(0) ? T(0) : alpha * *y;
        } else if constexpr (OP == TRAP_AXPY) {
            if (inside) *y = fma(alpha, X[i + j * ldx], *y);
        } else {
            if (!inside) *y = T(0);
        }
    }
}

// 128-bit version: a thread owns N consecutive rows (float4 / double2) of a column. The mask is monotone down a
// column, so the group is uniform (all inside or all outside) unless the boundary cuts through it; only those
// groups — at most one per column — fall back to per-element work.
template <typename T, int OP>
__global__ void __launch_bounds__(256)
trapezoid_vec_kernel(int64_t rows_v, int64_t cols, int upper, int64_t offset, int64_t row0, int64_t row_step, int64_t col0,
                     int64_t col_step, T alpha, const T* __restrict__ X, int64_t ldx, T* Y, int64_t ldy) {
    using V = typename VecOf<T>::type;
    constexpr int N = VecOf<T>::N;
    const int64_t iv = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (iv >= rows_v) return;
    const int64_t i = iv * N;
    const int64_t g_first = row0 + i * row_step, g_last = row0 + (i + N - 1) * row_step;
    for (int64_t j = int64_t(blockIdx.y) * blockDim.y + threadIdx.y; j < cols; j += int64_t(gridDim.y) * blockDim.y) {
        const int64_t lim = col0 + j * col_step - offset;  // inside: upper gi <= lim, lower gi >= lim
        const bool in_first = upper ? (g_first <= lim) : (g_first >= lim);
        const bool in_last = upper ? (g_last <= lim) : (g_last >= lim);
        T* y = Y + i + j * ldy;
        if (in_first == in_last) {
            if (OP == TRAP_MAKE ? in_first : !in_first) continue;  // nothing to do for this group
            V* yv = reinterpret_cast<V*>(y);
            if constexpr (OP == TRAP_SCALE) {
                *yv = alpha == T(0) ? vsplat(T(0)) : vscale(alpha, *yv);
            } else if constexpr (OP == TRAP_AXPY) {
                *yv = vaxpy(alpha, *reinterpret_cast<const V*>(X + i + j * ldx), *yv);
            } else {
                *yv = vsplat(T(0));
            }
        } else {
#pragma unroll
            for (int e = 0; e < N; ++e) {
                const int64_t gi = g_first + e * row_step;
                const bool inside = upper ? (gi <= lim) : (gi >= lim);
                if constexpr (OP == TRAP_SCALE) {
                    if (inside) y[e] = alpha == T(0) ? T(0) : alpha * y[e];
                } else if constexpr (OP == TRAP_AXPY) {
                    if (inside) y[e] = fma(alpha, X[i + e + j * ldx], y[e]);
                } else {
                    if (!inside) y[e] = T(0);
                }
            }
        }
    }
}

template <typename T>
int launch_trapezoid(int op, int uplo, int64_t offset, int64_t rows, int64_t cols, int64_t row0, int64_t row_step,
                     int64_t col0, int64_t col_step, T alpha, const T* X, int64_t ldx, T* Y, int64_t ldy, void* stream) {
    if (op < TRAP_SCALE || op > TRAP_MAKE || uplo < 0 || uplo > 1 || rows < 0 || cols < 0 || row_step < 1 || col_step < 1 ||
        row0 < 0 || col0 < 0)
        return ELB200_ERR_ARG;
    if (rows == 0 || cols == 0) return ELB200_OK;
    if (!Y || ldy < rows || (op == TRAP_AXPY && (!X || ldx < rows))) return ELB200_ERR_ARG;
    if (op == TRAP_SCALE && alpha == T(1)) return ELB200_OK;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    constexpr int NV = VecOf<T>::N;
    const bool vec = rows % NV == 0 && ldy % NV == 0 && (reinterpret_cast<uintptr_t>(Y) & 15u) == 0 &&
                     (op != TRAP_AXPY || (ldx % NV == 0 && (reinterpret_cast<uintptr_t>(X) & 15u) == 0));
    const int64_t rows_t = vec ? rows / NV : rows;  // rows in units of one thread's work
    int bx = 32;
    while (bx < 256 && bx < rows_t) bx <<= 1;
    const int by = 256 / bx;
    const int64_t gx = (rows_t + bx - 1) / bx;
    int64_t gy = (cols + by - 1) / by;
    const int64_t want = std::max<int64_t>(1, 32 * int64_t(device_sm_count()) / gx);  // enough CTAs to fill the SMs, grid-stride beyond
    if (gy > want) gy = want;
    if (gy > 65535) gy = 65535;
    const dim3 grid(static_cast<unsigned>(gx), static_cast<unsigned>(gy)), block(bx, by);
    const int upper = uplo == 1;
    count_launch();
    if (vec) {
        if (op == TRAP_SCALE)
            trapezoid_vec_kernel<T, TRAP_SCALE><<<grid, block, 0, s>>>(rows_t, cols, upper, offset, row0, row_step, col0, col_step, alpha, X, ldx, Y, ldy);
        else if (op == TRAP_AXPY)
            trapezoid_vec_kernel<T, TRAP_AXPY><<<grid, block, 0, s>>>(rows_t, cols, upper, offset, row0, row_step, col0, col_step, alpha, X, ldx, Y, ldy);
        else
            trapezoid_vec_kernel<T, TRAP_MAKE><<<grid, block, 0, s>>>(rows_t, cols, upper, offset, row0, row_step, col0, col_step, alpha, X, ldx, Y, ldy);
        return check_cuda(cudaGetLastError(), "trapezoid_vec_kernel");
    }
    if (op == TRAP_SCALE)
        trapezoid_kernel<T, TRAP_SCALE><<<grid, block, 0, s>>>(rows, cols, upper, offset, row0, row_step, col0, col_step, alpha, X, ldx, Y, ldy);
    else if (op == TRAP_AXPY)
        trapezoid_kernel<T, TRAP_AXPY><<<grid, block, 0, s>>>(rows, cols, upper, offset, row0, row_step, col0, col_step, alpha, X, ldx, Y, ldy);
    else
        trapezoid_kernel<T, TRAP_MAKE><<<grid, block, 0, s>>>(rows, cols, upper, offset, row0, row_step, col0, col_step, alpha, X, ldx, Y, ldy);
    return check_cuda(cudaGetLastError(), "trapezoid_kernel");
}

template <typename T>
int copy2d(int64_t rows, int64_t cols, const T* src, int64_t srs, int64_t scs, T* dst, int64_t drs,
           int64_t dcs, void* stream) {
    if (rows < 0 || cols < 0) return ELB200_ERR_ARG;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (srs == 1 && drs == 1) return launch_panel<T, OP_COPY>(rows, cols, src, scs, dst, dcs, T(0), s);
    return launch_strided<T, false>(rows, cols, src, srs, scs, dst, drs, dcs, T(0), s);
}

template <typename T>
int axpy2d(int64_t rows, int64_t cols, T alpha, const T* X, int64_t xrs, int64_t xcs, T* Y,
           int64_t yrs, int64_t ycs, void* stream) {
    if (rows < 0 || cols < 0) return ELB200_ERR_ARG;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (xrs == 1 && yrs == 1) return launch_panel<T, OP_AXPY>(rows, cols, X, xcs, Y, ycs, alpha, s);
    return launch_strided<T, true>(rows, cols, X, xrs, xcs, Y, yrs, ycs, alpha, s);
}

template <typename T>
int scale2d(int64_t rows, int64_t cols, T alpha, T* A, int64_t lda, void* stream) {
    if (rows < 0 || cols < 0) return ELB200_ERR_ARG;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (alpha == T(1)) return ELB200_OK;
    if (alpha == T(0)) return launch_panel<T, OP_FILL>(rows, cols, static_cast<const T*>(nullptr), 0, A, lda, T(0), s);
    return launch_panel<T, OP_SCALE>(rows, cols, static_cast<const T*>(nullptr), 0, A, lda, alpha, s);
}

}  // namespace
}  // namespace elb

using namespace elb;

extern "C" {

int elb200_peer_reduce_axpy_f32(int nsrc, const float* const* srcs, int64_t src_offset, const elb200_segment* seg,
                                float* M, int64_t ldm, float alpha, void* stream) {
    return peer_reduce_axpy<float>(nsrc, srcs, src_offset, seg, M, ldm, alpha, stream);
}
int elb200_peer_reduce_axpy_f64(int nsrc, const double* const* srcs, int64_t src_offset, const elb200_segment* seg,
                                double* M, int64_t ldm, double alpha, void* stream) {
    return peer_reduce_axpy<double>(nsrc, srcs, src_offset, seg, M, ldm, alpha, stream);
}
int elb200_copy2d_f32(int64_t rows, int64_t cols, const float* src, int64_t srs, int64_t scs,
                      float* dst, int64_t drs, int64_t dcs, void* stream) {
    return copy2d<float>(rows, cols, src, srs, scs, dst, drs, dcs, stream);
}
int elb200_copy2d_f64(int64_t rows, int64_t cols, const double* src, int64_t srs, int64_t scs,
                      double* dst, int64_t drs, int64_t dcs, void* stream) {
    return copy2d<double>(rows, cols, src, srs, scs, dst, drs, dcs, stream);
}
int elb200_transpose_f32(int64_t rows, int64_t cols, const float* A, int64_t lda, float* B,
                         int64_t ldb, void* stream) {
    return launch_transpose<float, false>(rows, cols, A, lda, B, ldb, 0.f,
                                          static_cast<cudaStream_t>(stream));
}
int elb200_transpose_f64(int64_t rows, int64_t cols, const double* A, int64_t lda, double* B,
                         int64_t ldb, void* stream) {
    return launch_transpose<double, false>(rows, cols, A, lda, B, ldb, 0.0,
                                           static_cast<cudaStream_t>(stream));
}
int elb200_axpy2d_f32(int64_t rows, int64_t cols, float alpha, const float* X, int64_t xrs,
                      int64_t xcs, float* Y, int64_t yrs, int64_t ycs, void* stream) {
    return axpy2d<float>(rows, cols, alpha, X, xrs, xcs, Y, yrs, ycs, stream);
}
int elb200_axpy2d_f64(int64_t rows, int64_t cols, double alpha, const double* X, int64_t xrs,
                      int64_t xcs, double* Y, int64_t yrs, int64_t ycs, void* stream) {
    return axpy2d<double>(rows, cols, alpha, X, xrs, xcs, Y, yrs, ycs, stream);
}
// B(rows x cols) += alpha * A(cols x rows)^T
int elb200_axpy2d_transpose_f32(int64_t rows, int64_t cols, float alpha, const float* A,
                                int64_t lda, float* B, int64_t ldb, void* stream) {
    return launch_transpose<float, true>(cols, rows, A, lda, B, ldb, alpha,
                                         static_cast<cudaStream_t>(stream));
}
int elb200_axpy2d_transpose_f64(int64_t rows, int64_t cols, double alpha, const double* A,
                                int64_t lda, double* B, int64_t ldb, void* stream) {
    return launch_transpose<double, true>(cols, rows, A, lda, B, ldb, alpha,
                                          static_cast<cudaStream_t>(stream));
}
int elb200_scale2d_f32(int64_t rows, int64_t cols, float alpha, float* A, int64_t lda,
                       void* stream) {
    return scale2d<float>(rows, cols, alpha, A, lda, stream);
}
int elb200_scale2d_f64(int64_t rows, int64_t cols, double alpha, double* A, int64_t lda,
                       void* stream) {
    return scale2d<double>(rows, cols, alpha, A, lda, stream);
}
int elb200_fill2d_f32(int64_t rows, int64_t cols, float value, float* A, int64_t lda,
                      void* stream) {
    if (rows < 0 || cols < 0) return ELB200_ERR_ARG;
    return launch_panel<float, OP_FILL>(rows, cols, static_cast<const float*>(nullptr), 0, A, lda,
                                        value, static_cast<cudaStream_t>(stream));
}
int elb200_fill2d_f64(int64_t rows, int64_t cols, double value, double* A, int64_t lda,
                      void* stream) {
    if (rows < 0 || cols < 0) return ELB200_ERR_ARG;
    return launch_panel<double, OP_FILL>(rows, cols, static_cast<const double*>(nullptr), 0, A, lda,
                                         value, static_cast<cudaStream_t>(stream));
}
int elb200_lattice_f32(int mode, int nseg, const elb200_segment* segs, float* M, int64_t ldm,
                       float* buf, float alpha, void* stream) {
    return launch_lattice<float>(mode, nseg, segs, M, ldm, buf, alpha,
                                 static_cast<cudaStream_t>(stream));
}
int elb200_lattice_f64(int mode, int nseg, const elb200_segment* segs, double* M, int64_t ldm,
                       double* buf, double alpha, void* stream) {
    return launch_lattice<double>(mode, nseg, segs, M, ldm, buf, alpha,
                                  static_cast<cudaStream_t>(stream));
}

int elb200_trapezoid_f32(int op, int uplo, int64_t offset, int64_t rows, int64_t cols, int64_t row0, int64_t row_step,
                         int64_t col0, int64_t col_step, float alpha, const float* X, int64_t ldx, float* Y, int64_t ldy,
                         void* stream) {
    return launch_trapezoid<float>(op, uplo, offset, rows, cols, row0, row_step, col0, col_step, alpha, X, ldx, Y, ldy, stream);
}
int elb200_trapezoid_f64(int op, int uplo, int64_t offset, int64_t rows, int64_t cols, int64_t row0, int64_t row_step,
                         int64_t col0, int64_t col_step, double alpha, const double* X, int64_t ldx, double* Y, int64_t ldy,
                         void* stream) {
    return launch_trapezoid<double>(op, uplo, offset, rows, cols, row0, row_step, col0, col_step, alpha, X, ldx, Y, ldy, stream);
}

}  // extern "C"
