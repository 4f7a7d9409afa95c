// FP64 GEMM for sm_100a on the DMMA pipe: C = alpha*op(A)*op(B) + beta*C, column-major.
//
// Replaces the cublasDgemm call of the reference's LocalGemm
// (reference: src/hydrogen/device/cuBLAS_API.cpp:282-300 via src/blas_like/level3/Gemm.cpp:163-186).
// tcgen05 has no f64 kind, so FP64 tensor math is the warp-level mma.sync.m8n8k4.f64 (SASS DMMA) fed by a
// cp.async (LDGSTS) ring of operand tiles. ncu puts the pipe's ceiling at 128 flop / clk / SM = 37.2 TFLOP/s.
//
// The kernel computes C^T = op(B)^T * op(A)^T so that each thread's accumulator pair is two
// CONSECUTIVE rows of one column of C: 16-byte loads/stores into the column-major C.
//   mma A-role (8x4, row): op(B)(k,n)  with n = lane/4, k = lane%4
//   mma B-role (4x8, col): op(A)(m,k)  with m = lane/4, k = lane%4
//   D (8x8): thread holds C(m = 2*(lane%4)+{0,1}, n = lane/4)
//
// History of the shape (all measured on B200, n = 8192 / 16384, cublasDgemm 35.9 / 36.1 TFLOP/s on the same box):
//   round 1  one 128 x 128 x 32 CTA of 8 warps per SM, 242 registers ............................ 32.2 / 32.7
//            the same, warp-specialised (4 cp.async producer warps + mbarriers) .................. 33.7 / 34.3
//            the same, TMA-fed: never launched ("too many resources": 288 threads x 208 registers + 198 KB)
//   round 2  64 x 64 x 16 tiles, 4 warps of 32 x 32, THREE CTAs per SM (the shape ncu showed cuBLAS using) 34.2 / 34.4
//            + cp.async addresses hoisted out of the k loop (4.6 -> 3.7 instructions per DMMA) .... 35.0 / 35.2
//            + FOUR CTAs per SM (<= 128 registers, 3 stages) — the default ..................... 35.2 / 35.4, DMMA pipe 94.9 %
// The one-CTA-per-SM kernels lost on every shape and were removed; profiles/r02_f64_* keeps their numbers.
#include <cuda_runtime.h>
#include <atomic>
#include <cstdint>
#include <cstdlib>

#include "ptx.cuh"
#include "elb200_internal.h"

namespace elb {
namespace {

// Raster order: groups of kGroup tile columns, tile rows fastest inside a group. With four 64 x 64 CTAs on each of 148 SMs the
// ~590 tiles in flight then form a window about kGroup wide and 590 / kGroup tall; 16 makes it 16 x 37 (8 was 8 x 74), i.e.
// fewer distinct A / B panels per resident tile and a higher L2 hit rate (ncu at n = 32768 with 8: 1.45 TB of DRAM reads per
// launch against 24 GB compulsory, 76 % L2 hits — far from HBM-bound at 740 GB/s, but it is energy).
__device__ __forceinline__ void tile_coords(int t, int tiles_m, int tiles_n, int& tm, int& tn) {
    constexpr int kGroup = 16;
    const int per_group = kGroup * tiles_m;
    const int g = t / per_group;
    const int first_n = g * kGroup;
    const int gsize = min(kGroup, tiles_n - first_n);
    const int r = t - g * per_group;
    tn = first_n + r % gsize;
    tm = r / gsize;
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

// ---------------------------------------------------------------------------------------------
// Multi-CTA-per-SM tiles. The ncu capture of cublasDgemm on this GPU (profiles/r02_ncu_f64_default_vs_cublas_details.txt)
// shows what its 36 TFLOP/s (DMMA pipe 96 % active, against 86 % for round 1's 128 x 128 kernel) rest on: SMALL tiles —
// 64 x 64 x 16, four warps of 32 x 32, 116 registers — so that several independent CTAs share an SM: while one sits in its
// per-k-tile barrier, waits for cp.async or runs its epilogue, the others keep the pipe fed. One big CTA per SM has nobody to
// cover those bubbles. Tile shape, stage count, warp grid and CTAs-per-SM target are template parameters:
//   Tile<64, 64, 16, 3, 2, 2, 4>   128 threads, 32 x 32 per warp, FOUR CTAs per SM (<= 128 registers): the default
//   Tile<64, 64, 16, S, 2, 2, 3>   the same with three CTAs per SM and, where it fits, a fourth stage
//   Tile<128, 64, 16, S, 4, 2, 2>  256 threads, two CTAs per SM (more reuse per byte from L2; slower: 33.2)
// Shared-memory tiles keep padded pitches ([k][mn + 4] or [mn][k + 4] doubles): every fragment load of every (N|T, N|T)
// case is bank-conflict free, and the operand tiles are sized exactly per majorness so that the stage count that still
// lets MINB CTAs share the SM's 227 KB is chosen at compile time.
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

// Per-thread cp.async plan for one operand tile, set up ONCE per CTA: the chunk a thread copies in iteration i differs
// from its chunk 0 only by a fixed stride (same mn, k row + i * kStepI — or same k, mn row + i * kStepI), so a k-tile
// costs one pointer increment and one cp.async per chunk instead of a division, a 64-bit multiply and two clamps each
// (the first version spent 3.6 integer / predicate instructions per DMMA on that; ncu, profiles/r02_*).
template <class C, bool MN_CONTIG, int MN, int VEC>
struct TileLoader {
    static constexpr int kChunks = (MN * C::BK) / VEC;
    static constexpr int kIters = kChunks / C::kThreads;
    static constexpr int kPerRow = MN_CONTIG ? MN / VEC : C::BK / VEC;
    static constexpr int kStepI = C::kThreads / kPerRow;  // rows (k rows or mn rows) between a thread's consecutive chunks
    static_assert(kChunks % C::kThreads == 0 && C::kThreads % kPerRow == 0, "tile / thread-count combination");
    static constexpr uint32_t kDstStep = uint32_t(kStepI) * uint32_t(MN_CONTIG ? MN + 4 : C::BK + 4) * 8u;

    const double* src;   // chunk 0 of the NEXT k-tile to issue
    int64_t step_i;      // elements between chunk i and chunk i + 1 in global memory
    int64_t step_k;      // elements from one k-tile to the next
    uint32_t dst0;       // byte offset of chunk 0 inside the operand tile
    int kk0;             // k offset of chunk 0 inside the tile
    int bytes_i[kIters]; // bytes of chunk i inside the matrix along mn (MN_CONTIG: the same for every i; else 0 or all)

    __device__ __forceinline__ void init(const double* X, int64_t ld, int mn0, int mn_extent) {
        const int r = threadIdx.x / kPerRow, c = (threadIdx.x % kPerRow) * VEC;
        if constexpr (MN_CONTIG) {
            kk0 = r;
            const int gmn = mn0 + c;
            const int b = min(max(mn_extent - gmn, 0), VEC) * 8;
#pragma unroll
            for (int i = 0; i < kIters; ++i) bytes_i[i] = b;
            src = X + int64_t(r) * ld + gmn;
            step_i = int64_t(kStepI) * ld;
            step_k = int64_t(C::BK) * ld;
            dst0 = uint32_t(r * (MN + 4) + c) * 8u;
        } else {
            kk0 = c;
#pragma unroll
            for (int i = 0; i < kIters; ++i) bytes_i[i] = (mn0 + r + i * kStepI < mn_extent) ? VEC * 8 : 0;
            src = X + int64_t(mn0 + r) * ld + c;
            step_i = int64_t(kStepI) * ld;
            step_k = C::BK;
            dst0 = uint32_t(r * (C::BK + 4) + c) * 8u;
        }
    }
    // Copy the tile whose first k index is k0 into smem_tile (zero fill outside the matrix), then advance to the next.
    // A chunk with src-size 0 reads nothing, so its (possibly out-of-range) address is never dereferenced — the
    // predicated-zfill idiom of cp.async.
    __device__ __forceinline__ void issue(uint32_t smem_tile, int k0, int k_extent) {
        const double* ptr = src;
        uint32_t dst = smem_tile + dst0;
        if (k0 + C::BK <= k_extent) {  // interior k-tile (all but the last): no k predicate at all
#pragma unroll
            for (int i = 0; i < kIters; ++i) {
                cp_async_zfill<VEC * 8>(dst, ptr, bytes_i[i]);
                ptr += step_i;
                dst += kDstStep;
            }
        } else if constexpr (MN_CONTIG) {
#pragma unroll
            for (int i = 0; i < kIters; ++i) {
                cp_async_zfill<VEC * 8>(dst, ptr, (k0 + kk0 + i * kStepI < k_extent) ? bytes_i[i] : 0);
                ptr += step_i;
                dst += kDstStep;
            }
        } else {
            const int kb = min(max(k_extent - (k0 + kk0), 0), VEC) * 8;
#pragma unroll
            for (int i = 0; i < kIters; ++i) {
                cp_async_zfill<VEC * 8>(dst, ptr, bytes_i[i] ? kb : 0);
                ptr += step_i;
                dst += kDstStep;
            }
        }
        src += step_k;
    }
};

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

    TileLoader<C, !TA, BM, VEC> load_a;
    TileLoader<C, TB, BN, VEC> load_b;
    load_a.init(p.A, p.lda, m0, p.m);
    load_b.init(p.B, p.ldb, n0, p.n);
    auto issue_stage = [&](int kt, int s) {  // k-tiles are issued in order: the loaders advance by one tile per call
        const uint32_t sa = smem_base + uint32_t(s * kStage) * 8u;
        const uint32_t sb = sa + uint32_t(kADoubles) * 8u;
        load_a.issue(sa, kt * BK, p.k);
        load_b.issue(sb, kt * BK, p.k);
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

}  // namespace

int gemm_f64_dmma(int transA, int transB, int64_t m, int64_t n, int64_t k, double alpha,
                  const double* A, int64_t lda, const double* B, int64_t ldb, double beta,
                  double* C, int64_t ldc, cudaStream_t stream) {
    Params p;
    p.A = A; p.B = B; p.C = C; p.lda = lda; p.ldb = ldb; p.ldc = ldc;
    p.m = int(m); p.n = int(n); p.k = int(k); p.alpha = alpha; p.beta = beta;
    p.tiles_m = p.tiles_n = 0;  // set per tile shape by launch_tile
    // 16-byte cp.async needs even leading dimensions and 16B-aligned bases; else 8-byte copies.
    const bool aligned = ((lda & 1) == 0) && ((ldb & 1) == 0) &&
                         ((reinterpret_cast<uintptr_t>(A) & 15u) == 0) &&
                         ((reinterpret_cast<uintptr_t>(B) & 15u) == 0);
    const bool ta = transA != 0, tb = transB != 0;
    cudaError_t e;
    // Tile<64,64,16> with FOUR CTAs per SM. Measured against the alternatives on the same box (profiles/r02_f64_tile_variants_*.log):
    // three CTAs with one more stage 35.2, Tile<128,64,16> with two CTAs 34.2, the round-1 128 x 128 x 32 one-CTA kernel 32.7,
    // this one 35.4-35.6 TFLOP/s at n = 16384; the losers were removed.
#define ELB_DISPATCH_T(BMv, BNv, WMG, WNG, MINB, V)                                                   \
    if (!ta && !tb) e = launch_tile_auto<BMv, BNv, WMG, WNG, MINB, false, false, V>(p, stream);       \
    else if (!ta && tb) e = launch_tile_auto<BMv, BNv, WMG, WNG, MINB, false, true, V>(p, stream);    \
    else if (ta && !tb) e = launch_tile_auto<BMv, BNv, WMG, WNG, MINB, true, false, V>(p, stream);    \
    else e = launch_tile_auto<BMv, BNv, WMG, WNG, MINB, true, true, V>(p, stream);
    if (aligned) { ELB_DISPATCH_T(64, 64, 2, 2, 4, 2) } else { ELB_DISPATCH_T(64, 64, 2, 2, 4, 1) }
#undef ELB_DISPATCH_T
    return e == cudaSuccess ? ELB200_OK : ELB200_ERR_CUDA;
}

}  // namespace elb
