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
#include <atomic>
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
// Multi-CTA-per-SM tiles (the default since round 2). The ncu capture of cublasDgemm on this GPU shows what its
// 36 TFLOP/s (DMMA pipe 96 % active, against 86 % for the 128 x 128 kernel above) rest on: SMALL tiles — 64 x 64 x 16,
// four warps of 32 x 32, 116 registers — so that THREE independent CTAs share an SM: while one sits in its per-k-tile
// barrier, waits for cp.async or runs its epilogue, the other two keep the pipe fed. One big CTA per SM (above) has
// nobody to cover those bubbles. Same math, same padded conflict-free shared-memory pitches, same C^T trick as above,
// with the tile shape, the stage count, the warp grid and the CTAs-per-SM target as template parameters:
//   Tile<64, 64, 16, S, 2, 2, 3>   the cuBLAS-like shape: 128 threads, 32 x 32 per warp, three CTAs per SM
//   Tile<128, 64, 16, S, 4, 2, 2>  256 threads, 32 x 32 per warp, two CTAs per SM (more reuse per byte from L2)
// Operand tiles are sized exactly per majorness ([k][mn + 4] or [mn][k + 4]) so that four stages of the NN / NT
// shapes still fit three times into the 227 KB of an SM.
// ---------------------------------------------------------------------------------------------
template <int BM_, int BN_, int BK_, int STAGES_, int WARPS_M_, int WARPS_N_, int MINB_>
struct Tile {
    static constexpr int BM = BM_, BN = BN_, BK = BK_, STAGES = STAGES_, WARPS_M = WARPS_M_, WARPS_N = WARPS_N_, MINB = MINB_;
    static constexpr int kThreads = 32 * WARPS_M * WARPS_N;
    static constexpr int WM = BM / WARPS_M, WN = BN / WARPS_N;  // warp tile
    static constexpr int MI = WM / 8, NI = WN / 8;              // 8 x 8 DMMA blocks per warp tile
    static_assert(WM % 8 == 0 && WN % 8 == 0 && BK % 4 == 0, "tile shape");
    // smem operand tile in doubles: MN-contiguous source -> [k][mn] with pitch mn + 4, else [mn][k] with pitch k + 4
    template <bool TA, bool TB> struct Smem {
        static constexpr int kA = !TA ? BK * (BM + 4) : BM * (BK + 4);
        static constexpr int kB = TB ? BK * (BN + 4) : BN * (BK + 4);
        static constexpr int kBytes = STAGES * (kA + kB) * 8;
    };
};

template <class C, bool MN_CONTIG, int MN, int VEC>
__device__ __forceinline__ void load_tile_t(const double* __restrict__ X, int64_t ld, int mn0, int k0, int mn_extent,
                                            int k_extent, uint32_t smem_tile) {
    constexpr int kChunks = (MN * C::BK) / VEC;
    constexpr int kIters = (kChunks + C::kThreads - 1) / C::kThreads;
#pragma unroll
    for (int i = 0; i < kIters; ++i) {
        const int c = threadIdx.x + i * C::kThreads;
        if (kChunks % C::kThreads != 0 && c >= kChunks) break;
        int mn, kk;
        uint32_t dst;
        if constexpr (MN_CONTIG) {
            constexpr int kPerRow = MN / VEC;
            kk = c / kPerRow;
            mn = (c % kPerRow) * VEC;
            dst = smem_tile + uint32_t(kk * (MN + 4) + mn) * 8u;
        } else {
            constexpr int kPerRow = C::BK / VEC;
            mn = c / kPerRow;
            kk = (c % kPerRow) * VEC;
            dst = smem_tile + uint32_t(mn * (C::BK + 4) + kk) * 8u;
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

template <class C, bool TA, bool TB, int VEC>
__global__ void __launch_bounds__(C::kThreads, C::MINB) gemm_f64_tile_kernel(const Params p) {
    constexpr int BM = C::BM, BN = C::BN, BK = C::BK, kStg = C::STAGES, MI = C::MI, NI = C::NI;
    constexpr int kADoubles = C::template Smem<TA, TB>::kA, kBDoubles = C::template Smem<TA, TB>::kB;
    constexpr int kStage = kADoubles + kBDoubles;
    constexpr int kPitchA = TA ? BK + 4 : BM + 4, kPitchB = TB ? BN + 4 : BK + 4;
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const uint32_t smem_base = smem_u32(smem_raw);
    const double* smem_d = reinterpret_cast<const double*>(smem_raw);

    int tm, tn;
    tile_coords(blockIdx.x, p.tiles_m, p.tiles_n, tm, tn);
    const int m0 = tm * BM, n0 = tn * BN;
    const int num_kt = (p.k + BK - 1) / BK;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int wm = (warp % C::WARPS_M) * C::WM;  // warp's m offset in the CTA tile
    const int wn = (warp / C::WARPS_M) * C::WN;  // warp's n offset
    const int lr = lane >> 2, lk = lane & 3;

    auto issue_stage = [&](int kt, int s) {
        const uint32_t sa = smem_base + uint32_t(s * kStage) * 8u;
        const uint32_t sb = sa + uint32_t(kADoubles) * 8u;
        load_tile_t<C, !TA, BM, VEC>(p.A, p.lda, m0, kt * BK, p.m, p.k, sa);
        load_tile_t<C, TB, BN, VEC>(p.B, p.ldb, n0, kt * BK, p.n, p.k, sb);
    };

    double acc[NI][MI][2];
#pragma unroll
    for (int ni = 0; ni < NI; ++ni)
#pragma unroll
        for (int mi = 0; mi < MI; ++mi) acc[ni][mi][0] = acc[ni][mi][1] = 0.0;

#pragma unroll
    for (int s = 0; s < kStg - 1; ++s) {
        if (s < num_kt) issue_stage(s, s);
        cp_async_commit();
    }

    int rd = 0, wr = kStg - 1;  // ring positions of the stage being read / refilled
    for (int kt = 0; kt < num_kt; ++kt) {
        cp_async_wait<kStg - 2>();
        __syncthreads();
        {
            const int nk = kt + kStg - 1;
            if (nk < num_kt) issue_stage(nk, wr);
            cp_async_commit();
        }
        const double* sa = smem_d + rd * kStage;
        const double* sb = sa + kADoubles;
#pragma unroll
        for (int k4 = 0; k4 < BK; k4 += 4) {
            double bf[MI], af[NI];
#pragma unroll
            for (int mi = 0; mi < MI; ++mi) {
                const int mm = wm + mi * 8 + lr, kk = k4 + lk;
                bf[mi] = TA ? sa[mm * kPitchA + kk] : sa[kk * kPitchA + mm];
            }
#pragma unroll
            for (int ni = 0; ni < NI; ++ni) {
                const int nn = wn + ni * 8 + lr, kk = k4 + lk;
                af[ni] = TB ? sb[kk * kPitchB + nn] : sb[nn * kPitchB + kk];
            }
#pragma unroll
            for (int ni = 0; ni < NI; ++ni)
#pragma unroll
                for (int mi = 0; mi < MI; ++mi) dmma_m8n8k4(acc[ni][mi], af[ni], bf[mi]);
        }
        rd = rd + 1 == kStg ? 0 : rd + 1;
        wr = wr + 1 == kStg ? 0 : wr + 1;
    }
    cp_async_wait<0>();

    // epilogue: thread owns C(m..m+1, n), m = m0+wm+mi*8+2*lk, n = n0+wn+ni*8+lr
    const bool vec_ok = ((p.ldc & 1) == 0) && ((reinterpret_cast<uintptr_t>(p.C) & 15u) == 0);
#pragma unroll
    for (int ni = 0; ni < NI; ++ni) {
        const int n = n0 + wn + ni * 8 + lr;
        if (n >= p.n) continue;
        double* ccol = p.C + int64_t(n) * p.ldc;
#pragma unroll
        for (int mi = 0; mi < MI; ++mi) {
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

// cudaFuncSetAttribute is per device: remember which devices have been configured for this instantiation
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

template <class C, bool TA, bool TB, int VEC>
cudaError_t launch_tile(Params p, cudaStream_t stream) {
    auto kern = gemm_f64_tile_kernel<C, TA, TB, VEC>;
    constexpr int smem = C::template Smem<TA, TB>::kBytes;
    static PerDeviceOnce once;
    int dev;
    if (once.needs(&dev)) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        if (e != cudaSuccess) return e;
        once.set(dev);
    }
    p.tiles_m = (p.m + C::BM - 1) / C::BM;
    p.tiles_n = (p.n + C::BN - 1) / C::BN;
    count_launch();
    kern<<<dim3(unsigned(p.tiles_m) * unsigned(p.tiles_n)), C::kThreads, smem, stream>>>(p);
    return cudaGetLastError();
}

// stage count per (tile, majorness): the most (<= 4) that still lets MINB CTAs share the SM's 227 KB (1 KB reserved per CTA)
template <int BM, int BN, int WMG, int WNG, int MINB, bool TA, bool TB, int VEC>
cudaError_t launch_tile_auto(const Params& p, cudaStream_t stream) {
    using C4 = Tile<BM, BN, 16, 4, WMG, WNG, MINB>;
    using C3 = Tile<BM, BN, 16, 3, WMG, WNG, MINB>;
    constexpr int kSmPerSm = 232448;
    if constexpr (MINB * (C4::template Smem<TA, TB>::kBytes + 1024) <= kSmPerSm) return launch_tile<C4, TA, TB, VEC>(p, stream);
    else return launch_tile<C3, TA, TB, VEC>(p, stream);
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
    // ELB200_F64_TILE: 64 = Tile<64,64,16> x 3 CTAs / SM (default), 12864 = Tile<128,64,16> x 2 CTAs / SM, 128 = the one-CTA-per-SM kernels
    static const int tile = [] { const char* v = std::getenv("ELB200_F64_TILE"); const int t = v ? std::atoi(v) : 0; return t == 128 || t == 12864 ? t : 64; }();
    static const bool ws = [] { const char* v = std::getenv("ELB200_F64_WS"); return v && std::atoi(v) != 0; }();
    if (tile != 128 && !ws) {
#define ELB_DISPATCH_T(BMv, BNv, WMG, WNG, MINB, V)                                                   \
    if (!ta && !tb) e = launch_tile_auto<BMv, BNv, WMG, WNG, MINB, false, false, V>(p, stream);       \
    else if (!ta && tb) e = launch_tile_auto<BMv, BNv, WMG, WNG, MINB, false, true, V>(p, stream);    \
    else if (ta && !tb) e = launch_tile_auto<BMv, BNv, WMG, WNG, MINB, true, false, V>(p, stream);    \
    else e = launch_tile_auto<BMv, BNv, WMG, WNG, MINB, true, true, V>(p, stream);
        if (tile == 64) { if (aligned) { ELB_DISPATCH_T(64, 64, 2, 2, 3, 2) } else { ELB_DISPATCH_T(64, 64, 2, 2, 3, 1) } }
        else { if (aligned) { ELB_DISPATCH_T(128, 64, 4, 2, 2, 2) } else { ELB_DISPATCH_T(128, 64, 4, 2, 2, 1) } }
#undef ELB_DISPATCH_T
        return e == cudaSuccess ? ELB200_OK : ELB200_ERR_CUDA;
    }
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
