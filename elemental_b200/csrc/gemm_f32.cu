// FP32 GEMM for sm_100a on the 5th-gen tensor cores: C = alpha*op(A)*op(B) + beta*C, column-major.
//
// Replaces the cublasSgemm call of the reference's LocalGemm
// (reference: src/hydrogen/device/cuBLAS_API.cpp:282-300 via src/blas_like/level3/Gemm.cpp:163-186).
//
// FP32 accuracy comes from a 3xTF32 split done INSIDE the kernel:
//   x = hi + lo,  hi = x with the low 13 mantissa bits cleared (what kind::tf32 reads from smem),
//   lo = x - hi (exact in fp32).   D += Alo*Bhi + Ahi*Blo + Ahi*Bhi   (fp32 accumulate in TMEM)
// The raw fp32 tile that TMA lands in shared memory IS the hi operand (the tensor core ignores the
// low 13 bits); four "splitter" warps compute the lo tile at the same (swizzled) byte offsets, so
// they never need to know the swizzle. No pre-pass over HBM, no extra workspace.
//
// Structure (one CTA, or one CTA pair with cta_group::2, per SM; persistent over output tiles):
//   warp 0      TMA producer        global -> smem (raw A, raw B), SWIZZLE_128B
//   warp 1      UMMA issuer         tcgen05.mma kind::tf32, accumulators in TMEM (2 buffers)
//   warps 2-5   splitters           raw -> lo planes in smem, fence.proxy.async, signal issuer
//   warps 6-9   epilogue            tcgen05.ld, alpha/beta, coalesced column-major stores
// Pipelines: full[s] (TMA->splitters), ready[s] (splitters->issuer), empty[s] (issuer->TMA),
//            tmem_full[a] (issuer->epilogue), tmem_empty[a] (epilogue->issuer).
#include <cuda_runtime.h>
#include <cuda.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <mutex>

#include "ptx.cuh"
#include "elb200_internal.h"

namespace elb {
namespace {

constexpr int kStages = 3;
constexpr int kBK = 32;                       // fp32 elements per k-block = one 128B swizzle row
constexpr int kTileMN = 128;                  // rows of A / rows of B held per CTA per stage
constexpr int kOperandBytes = kTileMN * kBK * 4;   // 16 KiB
constexpr int kStageBytes = 4 * kOperandBytes;     // A raw, B raw, A lo, B lo
constexpr int kOffAraw = 0, kOffBraw = kOperandBytes, kOffAlo = 2 * kOperandBytes,
              kOffBlo = 3 * kOperandBytes;
constexpr int kThreads = 320;
constexpr int kSplitThreads = 128;
constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;

enum : int { MAJOR_K = 0, MAJOR_MN = 1 };

// UMMA shared-memory matrix descriptor (cf. the PTX ISA "matrix descriptor" table).
// layout: 2 = SWIZZLE_128B (16B chunks), 1 = SWIZZLE_128B with 32B base chunks.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr, uint32_t lbo_bytes,
                                                   uint32_t sbo_bytes, uint32_t layout) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((saddr & 0x3FFFFu) >> 4);
    d |= static_cast<uint64_t>(lbo_bytes >> 4) << 16;
    d |= static_cast<uint64_t>(sbo_bytes >> 4) << 32;
    d |= 1ull << 46;  // descriptor version (Blackwell)
    d |= static_cast<uint64_t>(layout) << 61;
    return d;
}

// K-major tile  [mn-row r][32 k elems = 128B], SWIZZLE_128B:
//     SBO = 8 rows*128B between 8-row groups, LBO unused; one UMMA (K=8) = 32B inside the row.
// MN-major tile [mn-atom j (32 elems)][k-row r][128B]: 32-bit MN-major operands only exist in the
//     "128B swizzle, 32B base" layout (swizzle atom = 128B x 4 k-rows; TMA SWIZZLE_128B_ATOM_32B):
//     SBO = 4 rows*128B between k-atoms, LBO = 32 rows*128B between mn-atoms; one UMMA (K=8) = 8 rows.
template <int MAJOR>
__device__ __forceinline__ uint64_t operand_desc(uint32_t base, int kk) {
    if constexpr (MAJOR == MAJOR_MN)
        return make_smem_desc(base + kk * 1024, kBK * 128, 512, 1);
    else
        return make_smem_desc(base + kk * 32, 16, 1024, 2);
}

__device__ __forceinline__ void tile_coords(int t, int tiles_m, int tiles_n, int& tm, int& tn) {
    constexpr int kGroup = 8;  // n-tiles per raster group: A tiles reused 8x, B tiles ~(#clusters/8)x in L2
    const int per_group = kGroup * tiles_m;
    const int g = t / per_group;
    const int first_n = g * kGroup;
    const int gsize = min(kGroup, tiles_n - first_n);
    const int r = t - g * per_group;
    tn = first_n + r % gsize;
    tm = r / gsize;
}

struct Params {
    float* C;
    int64_t ldc;
    int m, n, k;
    float alpha, beta;
    int tiles_m, tiles_n;
};

template <int CG, int AMAJ, int BMAJ>
__global__ void __launch_bounds__(kThreads, 1)
gemm_f32_3xtf32_kernel(const __grid_constant__ CUtensorMap tmA,
                       const __grid_constant__ CUtensorMap tmB, const Params p) {
    constexpr int kUmmaM = 128 * CG;
    constexpr int kUmmaN = 128 * CG;
    constexpr uint32_t kTmemCols = 2 * kUmmaN;  // two accumulator buffers
    constexpr uint32_t kIdesc = (1u << 4) | (2u << 7) | (2u << 10) | (uint32_t(AMAJ) << 15) |
                                (uint32_t(BMAJ) << 16) | (uint32_t(kUmmaN >> 3) << 17) |
                                (uint32_t(kUmmaM >> 4) << 24);

    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = smem_base + kStages * kStageBytes;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto ready_bar = [&](int s) { return bar_base + 8u * (kStages + s); };
    auto empty_bar = [&](int s) { return bar_base + 8u * (2 * kStages + s); };
    auto tfull_bar = [&](int a) { return bar_base + 8u * (3 * kStages + a); };
    auto tempty_bar = [&](int a) { return bar_base + 8u * (3 * kStages + 2 + a); };
    const uint32_t tmem_slot = bar_base + 8u * (3 * kStages + 4);
    uint32_t* tmem_slot_ptr =
        reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t cta_rank = (CG == 2) ? cluster_ctarank() : 0u;
    const int cluster_id = blockIdx.x / CG;
    const int num_clusters = gridDim.x / CG;
    const int num_tiles = p.tiles_m * p.tiles_n;
    const int num_kb = (p.k + kBK - 1) / kBK;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(full_bar(s), 1);
            mbar_init(ready_bar(s), 4 * CG);
            mbar_init(empty_bar(s), 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(tfull_bar(a), 1);
            mbar_init(tempty_bar(a), 4 * CG);
        }
        fence_mbar_init();
    }
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
    }
    if (warp == 1) tmem_alloc<CG>(tmem_slot, kTmemCols);
    tc_fence_before_sync();
    if constexpr (CG == 2) cluster_sync_all(); else __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot_ptr;

    if (warp == 0) {
        // ------------------------------------------------------------ TMA producer
        if (lane == 0) {
            int s = 0; uint32_t ph = 0;
            for (int t = cluster_id; t < num_tiles; t += num_clusters) {
                int tm, tn;
                tile_coords(t, p.tiles_m, p.tiles_n, tm, tn);
                const int m0 = tm * kUmmaM + int(cta_rank) * kTileMN;
                const int n0 = tn * kUmmaN + int(cta_rank) * kTileMN;
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(empty_bar(s), ph ^ 1u);
                    const uint32_t st = smem_base + s * kStageBytes;
                    mbar_expect_tx(full_bar(s), 2 * kOperandBytes);
                    const int k0 = kb * kBK;
                    if constexpr (AMAJ == MAJOR_MN) {
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            tma_load_2d(&tmA, full_bar(s), st + kOffAraw + j * 4096, m0 + 32 * j, k0);
                    } else {
                        tma_load_2d(&tmA, full_bar(s), st + kOffAraw, k0, m0);
                    }
                    if constexpr (BMAJ == MAJOR_MN) {
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            tma_load_2d(&tmB, full_bar(s), st + kOffBraw + j * 4096, n0 + 32 * j, k0);
                    } else {
                        tma_load_2d(&tmB, full_bar(s), st + kOffBraw, k0, n0);
                    }
                    if (++s == kStages) { s = 0; ph ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ UMMA issuer (leader CTA)
        if (cta_rank == 0) {
            int s = 0; uint32_t ph = 0;
            int it = 0;
            for (int t = cluster_id; t < num_tiles; t += num_clusters, ++it) {
                const int acc = it & 1;
                const uint32_t acc_ph = (it >> 1) & 1;
                mbar_wait(tempty_bar(acc), acc_ph ^ 1u);
                tc_fence_after_sync();
                const uint32_t d_tmem = tmem_base + uint32_t(acc * kUmmaN);
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(ready_bar(s), ph);
                    tc_fence_after_sync();
                    if (lane == 0) {
                        const uint32_t st = smem_base + s * kStageBytes;
#pragma unroll
                        for (int kk = 0; kk < kBK / 8; ++kk) {
                            const uint64_t a_hi = operand_desc<AMAJ>(st + kOffAraw, kk);
                            const uint64_t a_lo = operand_desc<AMAJ>(st + kOffAlo, kk);
                            const uint64_t b_hi = operand_desc<BMAJ>(st + kOffBraw, kk);
                            const uint64_t b_lo = operand_desc<BMAJ>(st + kOffBlo, kk);
                            umma_tf32<CG>(d_tmem, a_lo, b_hi, kIdesc, (kb | kk) != 0 ? 1u : 0u);
                            umma_tf32<CG>(d_tmem, a_hi, b_lo, kIdesc, 1u);
                            umma_tf32<CG>(d_tmem, a_hi, b_hi, kIdesc, 1u);
                        }
                        umma_commit<CG>(empty_bar(s));
                        if (kb == num_kb - 1) umma_commit<CG>(tfull_bar(acc));
                    }
                    __syncwarp();
                    if (++s == kStages) { s = 0; ph ^= 1u; }
                }
            }
        }
    } else if (warp < 6) {
        // ------------------------------------------------------------ splitters: lo = x - trunc_tf32(x)
        const int tid = threadIdx.x - 64;
        int s = 0; uint32_t ph = 0;
        for (int t = cluster_id; t < num_tiles; t += num_clusters) {
            for (int kb = 0; kb < num_kb; ++kb) {
                mbar_wait(full_bar(s), ph);
                const uint32_t raw = smem_base + s * kStageBytes + kOffAraw + tid * 16;
                const uint32_t lo = raw + (kOffAlo - kOffAraw);
#pragma unroll
                for (int i0 = 0; i0 < (2 * kOperandBytes) / (kSplitThreads * 16); i0 += 4) {
                    uint32_t v[4][4];
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                                     : "=r"(v[i][0]), "=r"(v[i][1]), "=r"(v[i][2]), "=r"(v[i][3])
                                     : "r"(raw + (i0 + i) * kSplitThreads * 16));
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const uint32_t u = v[i][e];
                            const float x = __uint_as_float(u);
                            const float hi = __uint_as_float(u & 0xFFFFE000u);
                            // Inf/NaN: keep them in the hi product only (Inf - Inf would poison lo)
                            const float l = ((u & 0x7F800000u) == 0x7F800000u) ? 0.f : x - hi;
                            v[i][e] = __float_as_uint(l);
                        }
                        asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};"
                                     ::"r"(lo + (i0 + i) * kSplitThreads * 16), "r"(v[i][0]),
                                     "r"(v[i][1]), "r"(v[i][2]), "r"(v[i][3]) : "memory");
                    }
                }
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) {
                    if constexpr (CG == 2) mbar_arrive_cluster(ready_bar(s), 0);
                    else mbar_arrive(ready_bar(s));
                }
                if (++s == kStages) { s = 0; ph ^= 1u; }
            }
        }
    } else {
        // ------------------------------------------------------------ epilogue
        const int q = warp & 3;  // TMEM lane quadrant this warp may access
        int it = 0;
        for (int t = cluster_id; t < num_tiles; t += num_clusters, ++it) {
            int tm, tn;
            tile_coords(t, p.tiles_m, p.tiles_n, tm, tn);
            const int acc = it & 1;
            const uint32_t acc_ph = (it >> 1) & 1;
            const int row = tm * kUmmaM + int(cta_rank) * kTileMN + q * 32 + lane;
            const int n0 = tn * kUmmaN;
            mbar_wait(tfull_bar(acc), acc_ph);
            tc_fence_after_sync();
            const uint32_t t_addr = tmem_base + (uint32_t(q * 32) << 16) + uint32_t(acc * kUmmaN);
            const bool row_ok = row < p.m;
            float* crow = p.C + row;
#pragma unroll 1
            for (int c0 = 0; c0 < kUmmaN; c0 += 32) {
                if (n0 + c0 >= p.n) break;  // warp-uniform
                float cold[32];
                if (p.beta != 0.f) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const int col = n0 + c0 + j;
                        cold[j] = (row_ok && col < p.n) ? crow[int64_t(col) * p.ldc] : 0.f;
                    }
                }
                uint32_t v[32];
                tmem_ld_32x32b_x32(t_addr + uint32_t(c0), v);
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const int col = n0 + c0 + j;
                    float o = p.alpha * __uint_as_float(v[j]);
                    if (p.beta != 0.f) o = fmaf(p.beta, cold[j], o);
                    if (row_ok && col < p.n) crow[int64_t(col) * p.ldc] = o;
                }
            }
            tc_fence_before_sync();
            __syncwarp();
            if (lane == 0) {
                if constexpr (CG == 2) mbar_arrive_cluster(tempty_bar(acc), 0);
                else mbar_arrive(tempty_bar(acc));
            }
        }
    }

    tc_fence_before_sync();
    if constexpr (CG == 2) cluster_sync_all(); else __syncthreads();
    if (warp == 1) tmem_dealloc<CG>(tmem_base, kTmemCols);
}

// ------------------------------------------------------------------ host side
using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                   const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) ==
                cudaSuccess && qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(ptr);
    });
    return fn;
}

// Tensor map over a column-major matrix X (rows x cols, leading dimension ld, fp32).
//  as_mn_major: the GEMM's M/N index runs down X's columns (contiguous) -> box 32(mn) x 32(k)
//  else       : the GEMM's K index runs down X's columns (contiguous)   -> box 32(k)  x 128(mn)
bool encode_operand(CUtensorMap* tm, const float* X, int64_t rows, int64_t cols, int64_t ld,
                    bool as_mn_major) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) return false;
    cuuint64_t dims[2] = {cuuint64_t(rows), cuuint64_t(cols)};
    cuuint64_t strides[1] = {cuuint64_t(ld) * 4};
    cuuint32_t box[2] = {32u, as_mn_major ? 32u : 128u};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(X), dims, strides,
                     box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     as_mn_major ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS;
}

template <int CG, int AMAJ, int BMAJ>
cudaError_t launch(const CUtensorMap& tmA, const CUtensorMap& tmB, const Params& p, int num_sms,
                   cudaStream_t stream) {
    auto kern = gemm_f32_3xtf32_kernel<CG, AMAJ, BMAJ>;
    static PerDeviceOnce once;
    int dev;
    if (once.needs(&dev)) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             kSmemBytes);
        if (e != cudaSuccess) return e;
        once.set(dev);
    }
    const int num_tiles = p.tiles_m * p.tiles_n;
    int clusters = num_sms / CG;
    if (clusters > num_tiles) clusters = num_tiles;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(unsigned(clusters * CG));
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = kSmemBytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CG;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    count_launch();
    return cudaLaunchKernelEx(&cfg, kern, tmA, tmB, p);
}

}  // namespace

// transA/transB: 0 = 'N', 1 = 'T' (also 'C' for real types). cta_group: 1 or 2 (0 = default).
int gemm_f32_tensor(int transA, int transB, int64_t m, int64_t n, int64_t k, float alpha,
                    const float* A, int64_t lda, const float* B, int64_t ldb, float beta, float* C,
                    int64_t ldc, int cta_group, cudaStream_t stream) {
    if (cta_group == 0) cta_group = 2;
    // The TMEM accumulator truncates, so the error of one launch grows linearly with the length of
    // its accumulation chain (~5e-9 per k step, measured). Very long k is therefore cut into
    // chunks that accumulate through the epilogue's rounded fp32 FMA (beta = 1 after the first):
    // k = 262144 (BASELINE C5) then stays near 1e-4 instead of 1e-3 for < 1% more time. The bench
    // shape (k = 16384) is one launch.
    constexpr int64_t kchunk = 16384;
    if (k > kchunk + kchunk / 2) {
        for (int64_t k0 = 0; k0 < k; k0 += kchunk) {
            const int64_t kc = (k - k0 < kchunk + kchunk / 2) ? k - k0 : kchunk;
            const float* Ac = transA == 0 ? A + k0 * lda : A + k0;
            const float* Bc = transB == 0 ? B + k0 : B + k0 * ldb;
            const int st = gemm_f32_tensor(transA, transB, m, n, kc, alpha, Ac, lda, Bc, ldb, k0 == 0 ? beta : 1.0f, C, ldc, cta_group, stream);
            if (st != ELB200_OK) return st;
            if (kc != kchunk) break;
        }
        return ELB200_OK;
    }
    // A is (m x k) if N else (k x m), column-major. 'N' -> m contiguous -> MN-major.
    const bool a_mn = (transA == 0);
    // B is (k x n) if N else (n x k). 'N' -> k contiguous -> K-major; 'T' -> n contiguous -> MN-major.
    const bool b_mn = (transB != 0);
    CUtensorMap tmA, tmB;
    const int64_t a_rows = a_mn ? m : k, a_cols = a_mn ? k : m;
    const int64_t b_rows = b_mn ? n : k, b_cols = b_mn ? k : n;
    if (!encode_operand(&tmA, A, a_rows, a_cols, lda, a_mn)) return ELB200_ERR_TENSORMAP;
    if (!encode_operand(&tmB, B, b_rows, b_cols, ldb, b_mn)) return ELB200_ERR_TENSORMAP;
    Params p;
    p.C = C; p.ldc = ldc; p.m = int(m); p.n = int(n); p.k = int(k);
    p.alpha = alpha; p.beta = beta;
    const int tile = 128 * cta_group;
    p.tiles_m = int((m + tile - 1) / tile);
    p.tiles_n = int((n + tile - 1) / tile);
    const int num_sms = device_sm_count();
    cudaError_t e;
#define ELB_DISPATCH(CGV)                                                                   \
    if (a_mn && b_mn) e = launch<CGV, MAJOR_MN, MAJOR_MN>(tmA, tmB, p, num_sms, stream);    \
    else if (a_mn && !b_mn) e = launch<CGV, MAJOR_MN, MAJOR_K>(tmA, tmB, p, num_sms, stream); \
    else if (!a_mn && b_mn) e = launch<CGV, MAJOR_K, MAJOR_MN>(tmA, tmB, p, num_sms, stream); \
    else e = launch<CGV, MAJOR_K, MAJOR_K>(tmA, tmB, p, num_sms, stream);
    if (cta_group == 2) { ELB_DISPATCH(2) } else { ELB_DISPATCH(1) }
#undef ELB_DISPATCH
    return e == cudaSuccess ? ELB200_OK : ELB200_ERR_CUDA;
}

}  // namespace elb
