// FP64 GEMM on the DMMA pipe, TMA-fed (ELB200_F64_TMA=1; EXPERIMENTAL, off by default, not yet run on a GPU).
//
// Same tile (128 x 128 x 32), same mma.sync.m8n8k4.f64 warp layout and the same C^T = op(B)^T op(A)^T
// trick as gemm_f64.cu, but the operand tiles are landed by ONE elected thread with cp.async.bulk.tensor
// (TMA, SWIZZLE_128B) instead of 128 - 256 threads of cp.async: a producer warp and eight consumer warps
// hand stages over through full / empty mbarriers (the full barrier completes on the TMA's byte count).
// The dense swizzled tiles need a different fragment order than the padded ones to stay free of bank
// conflicts; the addressing lives in gemm_f64_layout.hpp and is checked on the CPU by
// tests/test_f64_tma_layout.py and modelled in tools/f64_tma_layout_check.py.
// Requirements (else the caller falls back to the cp.async kernels): 16-byte aligned A, B; even lda, ldb.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <mutex>

#include "elb200_internal.h"
#include "gemm_f64_layout.hpp"
#include "ptx.cuh"

namespace elb {
namespace {

using namespace f64tma;

constexpr int kStages = 3;
constexpr int kConsumerThreads = 256;
constexpr int kThreads = kConsumerThreads + 32;  // + one producer warp
constexpr int kStageBytes = 2 * kOperandBytes;   // A tile + B tile
constexpr int kSmemBytes = kStages * kStageBytes + 1024 /* 1 KiB alignment slack for the swizzle atoms */;

struct Params {
    double* C;
    int64_t ldc;
    int m, n, k;
    double alpha, beta;
    int tiles_m, tiles_n;
};

__device__ __forceinline__ void tile_coords(int t, int tiles_m, int tiles_n, int& tm, int& tn) {
    constexpr int kGroup = 8;
    const int per_group = kGroup * tiles_m;
    const int g = t / per_group;
    const int first_n = g * kGroup;
    const int gsize = min(kGroup, tiles_n - first_n);
    const int r = t - g * per_group;
    tn = first_n + r % gsize;
    tm = r / gsize;
}

__device__ __forceinline__ double lds64(uint32_t addr) {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void lds128(uint32_t addr, double& x, double& y) {
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(x), "=d"(y) : "r"(addr));
}

// TA: op(A) = A^T, i.e. A stored k-contiguous per m row; !TA: m-contiguous. TB: op(B) = B^T -> n-contiguous.
template <bool TA, bool TB>
__global__ void __maxnreg__(224)  // 288 threads x <= 224 registers fit the SM; __launch_bounds__(288) would cap ptxas at 168 and spill
gemm_f64_dmma_tma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const Params p) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t bars[2 * kStages];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = smem_u32(bars);
    auto full = [&](int s) { return bar_base + uint32_t(s) * 8u; };
    auto empty = [&](int s) { return bar_base + uint32_t(kStages + s) * 8u; };

    int tm, tn;
    tile_coords(blockIdx.x, p.tiles_m, p.tiles_n, tm, tn);
    const int m0 = tm * kBM, n0 = tn * kBN;
    const int num_kt = (p.k + kBK - 1) / kBK;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(full(s), 1);                      // the producer's arrive.expect_tx; TMA completes the bytes
            mbar_init(empty(s), kConsumerThreads / 32); // one arrive per consumer warp
        }
        fence_mbar_init();
    }
    __syncthreads();

    if (threadIdx.x >= kConsumerThreads) {
        // ---- producer warp: one elected lane issues every TMA
        if ((threadIdx.x & 31) == 0) {
            tma_prefetch_desc(&tmA);
            tma_prefetch_desc(&tmB);
            for (int kt = 0; kt < num_kt; ++kt) {
                const int s = kt % kStages;
                if (kt >= kStages) mbar_wait(empty(s), uint32_t((kt / kStages - 1) & 1));
                const uint32_t sa = smem_base + uint32_t(s) * kStageBytes, sb = sa + kOperandBytes;
                const int k0 = kt * kBK;
                mbar_expect_tx(full(s), kStageBytes);  // out-of-bounds parts of a box are zero-filled and still counted
                if constexpr (!TA) {
#pragma unroll
                    for (int b = 0; b < 8; ++b) tma_load_2d(&tmA, full(s), sa + b * 4096, m0 + 16 * b, k0);  // {mn, k}
                } else {
#pragma unroll
                    for (int b = 0; b < 2; ++b) tma_load_2d(&tmA, full(s), sa + b * 16384, k0 + 16 * b, m0);  // {k, mn}
                }
                if constexpr (TB) {
#pragma unroll
                    for (int b = 0; b < 8; ++b) tma_load_2d(&tmB, full(s), sb + b * 4096, n0 + 16 * b, k0);
                } else {
#pragma unroll
                    for (int b = 0; b < 2; ++b) tma_load_2d(&tmB, full(s), sb + b * 16384, k0 + 16 * b, n0);
                }
            }
        }
        return;
    }

    // ---- consumers: 8 warps as 2 (m) x 4 (n), each 64 x 32
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int wm = (warp & 1) * 64;
    const int wn = (warp >> 1) * 32;
    const uint32_t lr = uint32_t(lane >> 2), lk = uint32_t(lane & 3);
    // fragment rows this lane supplies: natural for MN-contiguous tiles, permuted for K-contiguous ones
    const uint32_t rowA = TA ? row_perm(lr) : lr;   // m index inside an 8-row group (op(A), the mma's B role)
    const uint32_t rowB = TB ? lr : row_perm(lr);   // n index inside an 8-row group (op(B), the mma's A role)

    // per-lane parts of the swizzled offsets (gemm_f64_layout.hpp); everything else below is an immediate
    const LaneMN a_mn = lane_mn(uint32_t(wm), lr, lk), b_mn = lane_mn(uint32_t(wn), lr, lk);
    const LaneK a_k = lane_k(uint32_t(wm), lr, lk), b_k = lane_k(uint32_t(wn), lr, lk);

    double acc[4][8][2];
#pragma unroll
    for (int ni = 0; ni < 4; ++ni)
#pragma unroll
        for (int mi = 0; mi < 8; ++mi) acc[ni][mi][0] = acc[ni][mi][1] = 0.0;

    for (int kt = 0; kt < num_kt; ++kt) {
        const int s = kt % kStages;
        mbar_wait(full(s), uint32_t((kt / kStages) & 1));
        const uint32_t sa = smem_base + uint32_t(s) * kStageBytes, sb = sa + kOperandBytes;
#pragma unroll
        for (int st = 0; st < kBK / 4; ++st) {  // mma k-steps; two of them cover one 8-k swizzle atom
            double bf[8], af[4];
#pragma unroll
            for (int mi = 0; mi < 8; ++mi) {
                if constexpr (TA) {
                    double x, y;  // the pair (k even, k odd) in one 16-byte load; this step uses one of them
                    lds128(sa + fast_k(a_k, uint32_t(mi), uint32_t(st)), x, y);
                    bf[mi] = (st & 1) ? y : x;
                } else {
                    bf[mi] = lds64(sa + fast_mn(a_mn, uint32_t(mi), uint32_t(st)));
                }
            }
#pragma unroll
            for (int ni = 0; ni < 4; ++ni) {
                if constexpr (TB) {
                    af[ni] = lds64(sb + fast_mn(b_mn, uint32_t(ni), uint32_t(st)));
                } else {
                    double x, y;
                    lds128(sb + fast_k(b_k, uint32_t(ni), uint32_t(st)), x, y);
                    af[ni] = (st & 1) ? y : x;
                }
            }
#pragma unroll
            for (int ni = 0; ni < 4; ++ni)
#pragma unroll
                for (int mi = 0; mi < 8; ++mi) dmma_m8n8k4(acc[ni][mi], af[ni], bf[mi]);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty(s));
    }

    // epilogue: the mma's D row is the lane's n fragment row, its two D columns are the m fragment rows 2*lk, 2*lk+1
    const bool vec_ok = !TA && ((p.ldc & 1) == 0) && ((reinterpret_cast<uintptr_t>(p.C) & 15u) == 0);
#pragma unroll
    for (int ni = 0; ni < 4; ++ni) {
        const int n = n0 + wn + ni * 8 + int(rowB);
        if (n >= p.n) continue;
        double* ccol = p.C + int64_t(n) * p.ldc;
#pragma unroll
        for (int mi = 0; mi < 8; ++mi) {
            const int mb = m0 + wm + mi * 8;
            const int ma = mb + int(TA ? row_perm(2 * lk) : 2 * lk), mc = mb + int(TA ? row_perm(2 * lk + 1) : 2 * lk + 1);
            double o0 = p.alpha * acc[ni][mi][0], o1 = p.alpha * acc[ni][mi][1];
            if (vec_ok && mc < p.m) {
                double2* ptr = reinterpret_cast<double2*>(ccol + ma);
                if (p.beta != 0.0) {
                    const double2 c = *ptr;
                    o0 = fma(p.beta, c.x, o0);
                    o1 = fma(p.beta, c.y, o1);
                }
                *ptr = make_double2(o0, o1);
            } else {
                if (ma < p.m) {
                    if (p.beta != 0.0) o0 = fma(p.beta, ccol[ma], o0);
                    ccol[ma] = o0;
                }
                if (mc < p.m) {
                    if (p.beta != 0.0) o1 = fma(p.beta, ccol[mc], o1);
                    ccol[mc] = o1;
                }
            }
        }
    }
}

// ------------------------------------------------------------------ host side
using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(ptr);
    });
    return fn;
}

// X is column-major with leading dimension ld; `inner` is the extent of its contiguous index, `outer` the other one.
// MN-contiguous use: box {16, 32}; K-contiguous use: box {16, 128}.
bool encode(CUtensorMap* tm, const double* X, int64_t inner, int64_t outer, int64_t ld, bool mn_contig) {
    EncodeTiledFn enc = encode_fn();
    if (!enc) return false;
    cuuint64_t dims[2] = {cuuint64_t(inner), cuuint64_t(outer)};
    cuuint64_t strides[1] = {cuuint64_t(ld) * 8};
    cuuint32_t box[2] = {16u, mn_contig ? 32u : 128u};
    cuuint32_t estr[2] = {1, 1};
    return enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, const_cast<double*>(X), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <bool TA, bool TB>
cudaError_t launch(const CUtensorMap& tmA, const CUtensorMap& tmB, const Params& p, cudaStream_t stream) {
    auto kern = gemm_f64_dmma_tma_kernel<TA, TB>;
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    count_launch();
    kern<<<dim3(unsigned(p.tiles_m) * unsigned(p.tiles_n)), kThreads, kSmemBytes, stream>>>(tmA, tmB, p);
    return cudaGetLastError();
}

}  // namespace

// Returns -1 when the operands do not meet the TMA requirements (the caller then uses the cp.async kernels).
int gemm_f64_dmma_tma(int transA, int transB, int64_t m, int64_t n, int64_t k, double alpha, const double* A, int64_t lda,
                      const double* B, int64_t ldb, double beta, double* C, int64_t ldc, cudaStream_t stream) {
    if ((lda & 1) || (ldb & 1) || (reinterpret_cast<uintptr_t>(A) & 15u) || (reinterpret_cast<uintptr_t>(B) & 15u)) return -1;
    const bool ta = transA != 0, tb = transB != 0;
    CUtensorMap tmA, tmB;
    // op(A)(m,k): 'N' -> A[k*lda + m] (m contiguous); 'T' -> A[m*lda + k] (k contiguous)
    if (!encode(&tmA, A, ta ? k : m, ta ? m : k, lda, !ta)) return -1;
    // op(B)(k,n): 'N' -> B[n*ldb + k] (k contiguous); 'T' -> B[k*ldb + n] (n contiguous)
    if (!encode(&tmB, B, tb ? n : k, tb ? k : n, ldb, tb)) return -1;
    Params p;
    p.C = C; p.ldc = ldc; p.m = int(m); p.n = int(n); p.k = int(k); p.alpha = alpha; p.beta = beta;
    p.tiles_m = int((m + kBM - 1) / kBM);
    p.tiles_n = int((n + kBN - 1) / kBN);
    cudaError_t e;
    if (!ta && !tb) e = launch<false, false>(tmA, tmB, p, stream);
    else if (!ta && tb) e = launch<false, true>(tmA, tmB, p, stream);
    else if (ta && !tb) e = launch<true, false>(tmA, tmB, p, stream);
    else e = launch<true, true>(tmA, tmB, p, stream);
    if (e != cudaSuccess) std::fprintf(stderr, "gemm_f64_dmma_tma: launch failed: %s\n", cudaGetErrorString(e));
    return e == cudaSuccess ? ELB200_OK : ELB200_ERR_CUDA;
}

}  // namespace elb
