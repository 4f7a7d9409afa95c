// FP64 GEMM for sm_100a on the DMMA pipe: C = alpha*op(A)*op(B) + beta*C, column-major.
//
// Replaces the cublasDgemm call of the reference's LocalGemm
// (reference: src/hydrogen/device/cuBLAS_API.cpp:282-300 via src/blas_like/level3/Gemm.cpp:163-186).
// tcgen05 has no f64 kind, so FP64 tensor math is the warp-level mma.sync.m8n8k4.f64 (SASS DMMA)
// fed by a 3-stage cp.async (LDGSTS) pipeline of 128 x 32 operand tiles (ELB_F64_BK / ELB_F64_STAGES;
// measured on B200: BK 32 x 3 stages 32.7 TF/s, BK 16 x 4 stages 31.5 TF/s at n = 16384).
//
// The kernel computes C^T = op(B)^T * op(A)^T so that each thread's accumulator pair is two
// CONSECUTIVE rows of one column of C: 16-byte loads/stores into the column-major C.
//   mma A-role (8x4, row): op(B)(k,n)  with n = lane/4, k = lane%4
//   mma B-role (4x8, col): op(A)(m,k)  with m = lane/4, k = lane%4
//   D (8x8): thread holds C(m = 2*(lane%4)+{0,1}, n = lane/4)
// CTA tile 128(m) x 128(n) x BK(k); 8 warps as 2(m) x 4(n), each 64 x 32: 32 DMMAs per k4 step
// against 12 conflict-free LDS.64.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdlib>

#include "ptx.cuh"
#include "elb200_internal.h"

namespace elb {
namespace {

#ifndef ELB_F64_BK
#define ELB_F64_BK 32
#endif
#ifndef ELB_F64_STAGES
#define ELB_F64_STAGES 3
#endif
constexpr int BM = 128, BN = 128, BK = ELB_F64_BK;
constexpr int kThreads = 256;
constexpr int kStages = ELB_F64_STAGES;
constexpr int kPitchMN = BM + 4;  // smem tile [k][mn]: pitch % 16 == 4 -> conflict-free fragment reads
constexpr int kPitchK = BK + 4;   // smem tile [mn][k]
constexpr int kOperandDoubles = (BK * kPitchMN > BM * kPitchK) ? BK * kPitchMN : BM * kPitchK;
constexpr int kStageDoubles = 2 * kOperandDoubles;
constexpr int kSmemBytes = kStages * kStageDoubles * 8;

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

// Copy one operand tile (128 mn x BK k) global -> smem with zero fill outside the matrix.
//  MN_CONTIG: X(mn,k) at X[k*ld + mn]  -> smem [k][mn] pitch kPitchMN
//  else     : X(mn,k) at X[mn*ld + k]  -> smem [mn][k] pitch kPitchK
template <bool MN_CONTIG, int VEC>
__device__ __forceinline__ void load_tile(const double* __restrict__ X, int64_t ld, int mn0, int k0,
                                          int mn_extent, int k_extent, uint32_t smem_tile) {
    constexpr int kChunks = (BM * BK) / VEC;
    constexpr int kPerThread = kChunks / kThreads;
#pragma unroll
    for (int i = 0; i < kPerThread; ++i) {
        const int c = threadIdx.x + i * kThreads;
        int mn, kk;
        uint32_t dst;
        if constexpr (MN_CONTIG) {
            constexpr int kPerRow = BM / VEC;
            kk = c / kPerRow;
            mn = (c % kPerRow) * VEC;
            dst = smem_tile + uint32_t(kk * kPitchMN + mn) * 8u;
        } else {
            constexpr int kPerRow = BK / VEC;
            mn = c / kPerRow;
            kk = (c % kPerRow) * VEC;
            dst = smem_tile + uint32_t(mn * kPitchK + kk) * 8u;
        }
        const int gmn = mn0 + mn, gk = k0 + kk;
        int valid;  // elements of this chunk inside the matrix
        if constexpr (MN_CONTIG) valid = (gk < k_extent) ? min(max(mn_extent - gmn, 0), VEC) : 0;
        else valid = (gmn < mn_extent) ? min(max(k_extent - gk, 0), VEC) : 0;
        const double* src = X;  // any valid address when nothing is read
        if (valid > 0) src = MN_CONTIG ? X + int64_t(gk) * ld + gmn : X + int64_t(gmn) * ld + gk;
        cp_async_zfill<VEC * 8>(dst, src, valid * 8);
    }
}

struct Params {
    const double* A;
    const double* B;
    double* C;
    int64_t lda, ldb, ldc;
    int m, n, k;
    double alpha, beta;
    int tiles_m, tiles_n;
};

template <bool TA, bool TB, int VEC>
__global__ void __launch_bounds__(kThreads, 1) gemm_f64_dmma_kernel(const Params p) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const uint32_t smem_base = smem_u32(smem_raw);
    const double* smem_d = reinterpret_cast<const double*>(smem_raw);

    int tm, tn;
    tile_coords(blockIdx.x, p.tiles_m, p.tiles_n, tm, tn);
    const int m0 = tm * BM, n0 = tn * BN;
    const int num_kt = (p.k + BK - 1) / BK;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int wm = (warp & 1) * 64;   // warp's m offset in the CTA tile
    const int wn = (warp >> 1) * 32;  // warp's n offset
    const int lr = lane >> 2, lk = lane & 3;

    // op(A)(m,k): 'N' -> A[k*lda+m] (m contiguous); 'T' -> A[m*lda+k]
    // op(B)(k,n): 'N' -> B[n*ldb+k] (k contiguous); 'T' -> B[k*ldb+n]
    auto issue_stage = [&](int kt, int s) {
        const uint32_t sa = smem_base + uint32_t(s * kStageDoubles) * 8u;
        const uint32_t sb = sa + uint32_t(kOperandDoubles) * 8u;
        load_tile<!TA, VEC>(p.A, p.lda, m0, kt * BK, p.m, p.k, sa);
        load_tile<TB, VEC>(p.B, p.ldb, n0, kt * BK, p.n, p.k, sb);
    };

    double acc[4][8][2];
#pragma unroll
    for (int ni = 0; ni < 4; ++ni)
#pragma unroll
        for (int mi = 0; mi < 8; ++mi) acc[ni][mi][0] = acc[ni][mi][1] = 0.0;

#pragma unroll
    for (int s = 0; s < kStages - 1; ++s) {
        if (s < num_kt) issue_stage(s, s);
        cp_async_commit();
    }

    for (int kt = 0; kt < num_kt; ++kt) {
        cp_async_wait<kStages - 2>();
        __syncthreads();
        {
            const int nk = kt + kStages - 1;
            if (nk < num_kt) issue_stage(nk, nk % kStages);
            cp_async_commit();
        }
        const double* sa = smem_d + (kt % kStages) * kStageDoubles;
        const double* sb = sa + kOperandDoubles;
#pragma unroll
        for (int k4 = 0; k4 < BK; k4 += 4) {
            double bf[8], af[4];
#pragma unroll
            for (int mi = 0; mi < 8; ++mi) {
                const int mm = wm + mi * 8 + lr, kk = k4 + lk;
                bf[mi] = TA ? sa[mm * kPitchK + kk] : sa[kk * kPitchMN + mm];
            }
#pragma unroll
            for (int ni = 0; ni < 4; ++ni) {
                const int nn = wn + ni * 8 + lr, kk = k4 + lk;
                af[ni] = TB ? sb[kk * kPitchMN + nn] : sb[nn * kPitchK + kk];
            }
#pragma unroll
            for (int ni = 0; ni < 4; ++ni)
#pragma unroll
                for (int mi = 0; mi < 8; ++mi) dmma_m8n8k4(acc[ni][mi], af[ni], bf[mi]);
        }
    }
    cp_async_wait<0>();

    // epilogue: thread owns C(m..m+1, n), m = m0+wm+mi*8+2*lk, n = n0+wn+ni*8+lr
    const bool vec_ok = ((p.ldc & 1) == 0) && ((reinterpret_cast<uintptr_t>(p.C) & 15u) == 0);
#pragma unroll
    for (int ni = 0; ni < 4; ++ni) {
        const int n = n0 + wn + ni * 8 + lr;
        if (n >= p.n) continue;
        double* ccol = p.C + int64_t(n) * p.ldc;
#pragma unroll
        for (int mi = 0; mi < 8; ++mi) {
            const int m = m0 + wm + mi * 8 + 2 * lk;
            if (m >= p.m) continue;
            double o0 = p.alpha * acc[ni][mi][0], o1 = p.alpha * acc[ni][mi][1];
            if (vec_ok && m + 1 < p.m) {
                double2* ptr = reinterpret_cast<double2*>(ccol + m);
                if (p.beta != 0.0) {
                    const double2 c = *ptr;
                    o0 = fma(p.beta, c.x, o0);
                    o1 = fma(p.beta, c.y, o1);
                }
                *ptr = make_double2(o0, o1);
            } else {
                if (p.beta != 0.0) o0 = fma(p.beta, ccol[m], o0);
                ccol[m] = o0;
                if (m + 1 < p.m) {
                    if (p.beta != 0.0) o1 = fma(p.beta, ccol[m + 1], o1);
                    ccol[m + 1] = o1;
                }
            }
        }
    }
}

template <bool TA, bool TB, int VEC>
cudaError_t launch(const Params& p, cudaStream_t stream) {
    auto kern = gemm_f64_dmma_kernel<TA, TB, VEC>;
    static bool configured = false;
    if (!configured) {
        cudaError_t e =
            cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    count_launch();
    kern<<<dim3(unsigned(p.tiles_m) * unsigned(p.tiles_n)), kThreads, kSmemBytes, stream>>>(p);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// Warp-specialised variant (ELB200_F64_WS=1). Measured on a B200 at n = 8192: 33.7 TFLOP/s against 32.2 for the
// kernel above (30.3 with two producer warps); off by default until the whole GPU suite has run with it.
// The ncu capture of the kernel above charges 6.4 % of the warp time to the CTA-wide barrier per
// k-tile and 6.2 % to the fragment loads right behind it (profiles/).  Here four
// producer warps issue every cp.async and eight consumer warps only do LDS + DMMA: a stage is
// handed over through a pair of mbarriers (full: cp.async.mbarrier.arrive of the 128 producer
// threads; empty: one arrive per consumer warp), so no warp ever waits for the slowest of the CTA,
// and the consumers prefetch the fragments of k4+4 while the DMMAs of k4 issue.
// Same tile, same shared-memory layout (conflict-free pitches), same epilogue as the kernel above.
// ---------------------------------------------------------------------------------------------
constexpr int kProducerThreads = 128;
constexpr int kThreadsWS = kThreads + kProducerThreads;

template <bool MN_CONTIG, int VEC>
__device__ __forceinline__ void load_tile_ws(const double* __restrict__ X, int64_t ld, int mn0, int k0, int mn_extent,
                                             int k_extent, uint32_t smem_tile, int ptid) {
    constexpr int kChunks = (BM * BK) / VEC;
    constexpr int kPerThread = kChunks / kProducerThreads;
#pragma unroll 4
    for (int i = 0; i < kPerThread; ++i) {
        const int c = ptid + i * kProducerThreads;
        int mn, kk;
        uint32_t dst;
        if constexpr (MN_CONTIG) {
            constexpr int kPerRow = BM / VEC;
            kk = c / kPerRow;
            mn = (c % kPerRow) * VEC;
            dst = smem_tile + uint32_t(kk * kPitchMN + mn) * 8u;
        } else {
            constexpr int kPerRow = BK / VEC;
            mn = c / kPerRow;
            kk = (c % kPerRow) * VEC;
            dst = smem_tile + uint32_t(mn * kPitchK + kk) * 8u;
        }
        const int gmn = mn0 + mn, gk = k0 + kk;
        int valid;
        if constexpr (MN_CONTIG) valid = (gk < k_extent) ? min(max(mn_extent - gmn, 0), VEC) : 0;
        else valid = (gmn < mn_extent) ? min(max(k_extent - gk, 0), VEC) : 0;
        const double* src = X;
        if (valid > 0) src = MN_CONTIG ? X + int64_t(gk) * ld + gmn : X + int64_t(gmn) * ld + gk;
        cp_async_zfill<VEC * 8>(dst, src, valid * 8);
    }
}

// all cp.async issued so far by this thread arrive on `bar` when they have landed (count not incremented)
__device__ __forceinline__ void cp_async_arrive_noinc(uint32_t bar) {
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}

template <bool TA, bool TB, int VEC>
__global__ void __launch_bounds__(kThreadsWS, 1) gemm_f64_dmma_ws_kernel(const Params p) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t bars[2 * kStages];  // full[0..kStages), empty[kStages..2*kStages)
    const uint32_t smem_base = smem_u32(smem_raw);
    const double* smem_d = reinterpret_cast<const double*>(smem_raw);
    const uint32_t bar_base = smem_u32(bars);
    auto full = [&](int s) { return bar_base + uint32_t(s) * 8u; };
    auto empty = [&](int s) { return bar_base + uint32_t(kStages + s) * 8u; };

    int tm, tn;
    tile_coords(blockIdx.x, p.tiles_m, p.tiles_n, tm, tn);
    const int m0 = tm * BM, n0 = tn * BN;
    const int num_kt = (p.k + BK - 1) / BK;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(full(s), kProducerThreads);
            mbar_init(empty(s), kThreads / 32);
        }
        fence_mbar_init();
    }
    __syncthreads();

    if (threadIdx.x >= kThreads) {
        // ---- producers
        const int ptid = threadIdx.x - kThreads;
        for (int kt = 0; kt < num_kt; ++kt) {
            const int s = kt % kStages;
            if (kt >= kStages) mbar_wait(empty(s), uint32_t((kt / kStages - 1) & 1));
            const uint32_t sa = smem_base + uint32_t(s * kStageDoubles) * 8u;
            const uint32_t sb = sa + uint32_t(kOperandDoubles) * 8u;
            load_tile_ws<!TA, VEC>(p.A, p.lda, m0, kt * BK, p.m, p.k, sa, ptid);
            load_tile_ws<TB, VEC>(p.B, p.ldb, n0, kt * BK, p.n, p.k, sb, ptid);
            cp_async_arrive_noinc(full(s));
        }
        cp_async_wait<0>();
        return;
    }

    // ---- consumers
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int wm = (warp & 1) * 64;
    const int wn = (warp >> 1) * 32;
    const int lr = lane >> 2, lk = lane & 3;

    double acc[4][8][2];
#pragma unroll
    for (int ni = 0; ni < 4; ++ni)
#pragma unroll
        for (int mi = 0; mi < 8; ++mi) acc[ni][mi][0] = acc[ni][mi][1] = 0.0;

    auto load_frags = [&](const double* sa, const double* sb, int k4, double (&bf)[8], double (&af)[4]) {
#pragma unroll
        for (int mi = 0; mi < 8; ++mi) {
            const int mm = wm + mi * 8 + lr, kk = k4 + lk;
            bf[mi] = TA ? sa[mm * kPitchK + kk] : sa[kk * kPitchMN + mm];
        }
#pragma unroll
        for (int ni = 0; ni < 4; ++ni) {
            const int nn = wn + ni * 8 + lr, kk = k4 + lk;
            af[ni] = TB ? sb[kk * kPitchMN + nn] : sb[nn * kPitchK + kk];
        }
    };

    for (int kt = 0; kt < num_kt; ++kt) {
        const int s = kt % kStages;
        mbar_wait(full(s), uint32_t((kt / kStages) & 1));
        const double* sa = smem_d + s * kStageDoubles;
        const double* sb = sa + kOperandDoubles;
        double bf[2][8], af[2][4];
        load_frags(sa, sb, 0, bf[0], af[0]);
#pragma unroll
        for (int k4 = 0; k4 < BK; k4 += 4) {
            const int cur = (k4 >> 2) & 1;
            if (k4 + 4 < BK) load_frags(sa, sb, k4 + 4, bf[cur ^ 1], af[cur ^ 1]);
#pragma unroll
            for (int ni = 0; ni < 4; ++ni)
#pragma unroll
                for (int mi = 0; mi < 8; ++mi) dmma_m8n8k4(acc[ni][mi], af[cur][ni], bf[cur][mi]);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty(s));  // this warp is done with the stage
    }

    const bool vec_ok = ((p.ldc & 1) == 0) && ((reinterpret_cast<uintptr_t>(p.C) & 15u) == 0);
#pragma unroll
    for (int ni = 0; ni < 4; ++ni) {
        const int n = n0 + wn + ni * 8 + lr;
        if (n >= p.n) continue;
        double* ccol = p.C + int64_t(n) * p.ldc;
#pragma unroll
        for (int mi = 0; mi < 8; ++mi) {
            const int m = m0 + wm + mi * 8 + 2 * lk;
            if (m >= p.m) continue;
            double o0 = p.alpha * acc[ni][mi][0], o1 = p.alpha * acc[ni][mi][1];
            if (vec_ok && m + 1 < p.m) {
                double2* ptr = reinterpret_cast<double2*>(ccol + m);
                if (p.beta != 0.0) {
                    const double2 c = *ptr;
                    o0 = fma(p.beta, c.x, o0);
                    o1 = fma(p.beta, c.y, o1);
                }
                *ptr = make_double2(o0, o1);
            } else {
                if (p.beta != 0.0) o0 = fma(p.beta, ccol[m], o0);
                ccol[m] = o0;
                if (m + 1 < p.m) {
                    if (p.beta != 0.0) o1 = fma(p.beta, ccol[m + 1], o1);
                    ccol[m + 1] = o1;
                }
            }
        }
    }
}

template <bool TA, bool TB, int VEC>
cudaError_t launch_ws(const Params& p, cudaStream_t stream) {
    auto kern = gemm_f64_dmma_ws_kernel<TA, TB, VEC>;
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    count_launch();
    kern<<<dim3(unsigned(p.tiles_m) * unsigned(p.tiles_n)), kThreadsWS, kSmemBytes, stream>>>(p);
    return cudaGetLastError();
}

}  // namespace

int gemm_f64_dmma(int transA, int transB, int64_t m, int64_t n, int64_t k, double alpha,
                  const double* A, int64_t lda, const double* B, int64_t ldb, double beta,
                  double* C, int64_t ldc, cudaStream_t stream) {
    Params p;
    p.A = A; p.B = B; p.C = C; p.lda = lda; p.ldb = ldb; p.ldc = ldc;
    p.m = int(m); p.n = int(n); p.k = int(k); p.alpha = alpha; p.beta = beta;
    p.tiles_m = int((m + BM - 1) / BM);
    p.tiles_n = int((n + BN - 1) / BN);
    // 16-byte cp.async needs even leading dimensions and 16B-aligned bases; else 8-byte copies.
    const bool aligned = ((lda & 1) == 0) && ((ldb & 1) == 0) &&
                         ((reinterpret_cast<uintptr_t>(A) & 15u) == 0) &&
                         ((reinterpret_cast<uintptr_t>(B) & 15u) == 0);
    const bool ta = transA != 0, tb = transB != 0;
    cudaError_t e;
#define ELB_DISPATCH(V)                                                \
    if (!ta && !tb) e = launch<false, false, V>(p, stream);           \
    else if (!ta && tb) e = launch<false, true, V>(p, stream);        \
    else if (ta && !tb) e = launch<true, false, V>(p, stream);        \
    else e = launch<true, true, V>(p, stream);
    static const bool tma = [] { const char* v = std::getenv("ELB200_F64_TMA"); return v && std::atoi(v) != 0; }();
    if (tma) {
        const int st = gemm_f64_dmma_tma(transA, transB, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, stream);
        if (st >= 0) return st;  // -1: operands not TMA-able (alignment), use the cp.async kernels
    }
    static const bool ws = [] { const char* v = std::getenv("ELB200_F64_WS"); return v && std::atoi(v) != 0; }();
    if (ws) {
#define ELB_DISPATCH_WS(V)                                                \
    if (!ta && !tb) e = launch_ws<false, false, V>(p, stream);           \
    else if (!ta && tb) e = launch_ws<false, true, V>(p, stream);        \
    else if (ta && !tb) e = launch_ws<true, false, V>(p, stream);        \
    else e = launch_ws<true, true, V>(p, stream);
        if (aligned) { ELB_DISPATCH_WS(2) } else { ELB_DISPATCH_WS(1) }
#undef ELB_DISPATCH_WS
        return e == cudaSuccess ? ELB200_OK : ELB200_ERR_CUDA;
    }
    if (aligned) { ELB_DISPATCH(2) } else { ELB_DISPATCH(1) }
#undef ELB_DISPATCH
    return e == cudaSuccess ? ELB200_OK : ELB200_ERR_CUDA;
}

}  // namespace elb
