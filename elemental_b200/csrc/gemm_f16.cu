// FP16 GEMM for sm_100a on the 5th-gen tensor cores: C = alpha*op(A)*op(B) + beta*C, column-major, A / B / C in
// IEEE half, products accumulated in FP32 in TMEM, alpha / beta in FP32, one rounding to half in the epilogue.
//
// Replaces the cublasHgemm / cublasGemmEx(CUDA_R_16F) call behind the reference's gpu_half_type instantiation of Gemm
// (reference: src/hydrogen/device/cuBLAS_API.cpp:417 ADD_GEMM_IMPL(__half, H) via src/blas_like/level3/Gemm.cpp:455-462);
// SURVEY.md 8(f) N3. FP16 is the native tcgen05 type: ONE kind::f16 UMMA per k-slice (the fp32 kernel of gemm_f32.cu
// needs three kind::tf32 UMMAs plus splitter warps), so the same pipeline runs without the lo planes and with twice
// the stages.
//
// Structure (one CTA pair with cta_group::2 per 2 SMs; persistent over 256 x 256 output tiles):
//   warp 0      TMA producer        global -> smem (A, B), SWIZZLE_128B, 64 halfs (128 B) of k per row
//   warp 1      UMMA issuer         tcgen05.mma kind::f16, fp32 accumulators in TMEM (2 buffers)
//   warps 2-5   epilogue            tcgen05.ld, alpha/beta in fp32, half stores (coalesced down the columns of C)
// Pipelines: full[s] (TMA->issuer), empty[s] (issuer->TMA), tmem_full[a] (issuer->epilogue), tmem_empty[a] (epilogue->issuer).
// With cta_group::2 both CTAs' loads complete on the LEADER's full[s] (cp.async.bulk.tensor.cta_group::2), where the one
// issuer of the pair waits; the commit that frees a stage is multicast to both CTAs' empty[s].
#include <cuda_runtime.h>
#include <cuda.h>
#include <cuda_fp16.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <mutex>

#include "ptx.cuh"
#include "elb200_internal.h"

namespace elb {
namespace {

constexpr int kStages = 6;  // 7 fit (224 KiB) and measured the same: 1275 vs 1275 TFLOP/s at n = 8192
constexpr int kBK = 64;                       // halfs per k-block = one 128B swizzle row
constexpr int kTileMN = 128;                  // rows of A / rows of B held per CTA per stage
constexpr int kOperandBytes = kTileMN * kBK * 2;   // 16 KiB
constexpr int kStageBytes = 2 * kOperandBytes;     // A, B
constexpr int kOffAraw = 0, kOffBraw = kOperandBytes;
constexpr int kThreads = 192;  // TMA, UMMA, 4 epilogue warps
constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;
static_assert(kSmemBytes <= 227 * 1024, "stage ring exceeds the 227 KiB a CTA may use");

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

// K-major tile  [mn-row r][64 k halfs = 128B], SWIZZLE_128B:
//     SBO = 8 rows*128B between 8-row groups, LBO unused; one UMMA (K=16) = 32B inside the row.
// MN-major tile [mn-atom j (64 halfs)][k-row r][128B], SWIZZLE_128B (swizzle atom = 128B x 8 k-rows):
//     SBO = 8 rows*128B between k-groups, LBO = 64 rows*128B between mn-atoms; one UMMA (K=16) = 16 rows.
template <int MAJOR>
__device__ __forceinline__ uint64_t operand_desc(uint32_t base, int kk) {
    if constexpr (MAJOR == MAJOR_MN)
        return make_smem_desc(base + kk * 2048, kBK * 128, 1024, 2);
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
    __half* C;
    int64_t ldc;
    int m, n, k;
    float alpha, beta;
    int tiles_m, tiles_n;  // tiles_m counts cluster rows: kPairs output tiles of 128*CG rows each
};

// CG = CTAs per UMMA (cta_group), PAIRS = CTA pairs per cluster (1, or 2 with CG == 2: a 4-CTA cluster computing two
// 256 x 256 tiles that are neighbours in m and share their B tile, which is loaded ONCE from L2 and multicast).
template <int CG, int AMAJ, int BMAJ, int PAIRS>
__global__ void __launch_bounds__(kThreads, 1)
gemm_f16_kernel(const __grid_constant__ CUtensorMap tmA,
                       const __grid_constant__ CUtensorMap tmB, const Params p) {
    static_assert(PAIRS == 1 || (PAIRS == 2 && CG == 2), "multicast needs CTA pairs");
    constexpr int kCluster = CG * PAIRS;
    constexpr int kUmmaM = 128 * CG;
    constexpr int kUmmaN = 128 * CG;
    constexpr uint32_t kTmemCols = 2 * kUmmaN;  // two accumulator buffers
    constexpr uint32_t kIdesc = (1u << 4) | (0u << 7) | (0u << 10) | (uint32_t(AMAJ) << 15) |
                                (uint32_t(BMAJ) << 16) | (uint32_t(kUmmaN >> 3) << 17) |
                                (uint32_t(kUmmaM >> 4) << 24);

    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = smem_base + kStages * kStageBytes;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (kStages + s); };
    auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * kStages + a); };
    auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * kStages + 2 + a); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * kStages + 4);
    uint32_t* tmem_slot_ptr =
        reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t cluster_rank = (kCluster > 1) ? cluster_ctarank() : 0u;
    const uint32_t cta_rank = cluster_rank & uint32_t(CG - 1);  // rank inside the pair: 0 = leader
    const uint32_t pair = cluster_rank / CG;                     // which pair of the cluster
    const uint32_t leader = cluster_rank - cta_rank;             // cluster rank of this pair's leader
    const int cluster_id = blockIdx.x / kCluster;
    const int num_clusters = gridDim.x / kCluster;
    const int num_tiles = p.tiles_m * p.tiles_n;
    const int num_kb = (p.k + kBK - 1) / kBK;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(full_bar(s), 1);       // the leader's expect_tx arrival; the bytes landing in BOTH CTAs of the pair complete on it
            mbar_init(empty_bar(s), PAIRS);  // every pair of the cluster has consumed the stage (others' multicasts land in it too)
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
    if constexpr (kCluster > 1) cluster_sync_all(); else __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot_ptr;

    if (warp == 0) {
        // ------------------------------------------------------------ TMA producer (one per CTA; completions go to the pair leader)
        if (lane == 0) {
            auto load = [&](const CUtensorMap* tm, uint32_t bar, uint32_t dst, int c0, int c1) {
                if constexpr (CG == 2) tma_load_2d_pair(tm, bar, dst, c0, c1, leader);
                else tma_load_2d(tm, bar, dst, c0, c1);
            };
            int s = 0; uint32_t ph = 0;
            for (int t = cluster_id; t < num_tiles; t += num_clusters) {
                int tm, tn;
                tile_coords(t, p.tiles_m, p.tiles_n, tm, tn);
                const int m0 = (tm * PAIRS + int(pair)) * kUmmaM + int(cta_rank) * kTileMN;
                const int n0 = tn * kUmmaN + int(cta_rank) * kTileMN;
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(empty_bar(s), ph ^ 1u);
                    const uint32_t st = smem_base + s * kStageBytes;
                    if (cta_rank == 0) mbar_expect_tx(full_bar(s), CG * kStageBytes);
                    const int k0 = kb * kBK;
                    if constexpr (AMAJ == MAJOR_MN) {
#pragma unroll
                        for (int j = 0; j < 2; ++j)
                            load(&tmA, full_bar(s), st + kOffAraw + j * 8192, m0 + 64 * j, k0);
                    } else {
                        load(&tmA, full_bar(s), st + kOffAraw, k0, m0);
                    }
                    if constexpr (PAIRS == 2) {
                        // my 128 rows of B are also the 128 rows of the CTA with the same pair rank in the other pair: each of
                        // the two fetches one 64-row half (8 KiB, 1024-aligned: the same SWIZZLE_128B image as one box) for both
                        const uint16_t mask = uint16_t((1u << cta_rank) | (1u << (cta_rank + CG)));
                        const uint32_t dst = st + kOffBraw + pair * 8192;
                        const int nb = n0 + 64 * int(pair);
                        if constexpr (BMAJ == MAJOR_MN) tma_load_2d_pair_mc(&tmB, full_bar(s), dst, nb, k0, leader, mask);
                        else tma_load_2d_pair_mc(&tmB, full_bar(s), dst, k0, nb, leader, mask);
                    } else if constexpr (BMAJ == MAJOR_MN) {
#pragma unroll
                        for (int j = 0; j < 2; ++j)
                            load(&tmB, full_bar(s), st + kOffBraw + j * 8192, n0 + 64 * j, k0);
                    } else {
                        load(&tmB, full_bar(s), st + kOffBraw, k0, n0);
                    }
                    if (++s == kStages) { s = 0; ph ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ UMMA issuer (the pair's leader CTA)
        if (cta_rank == 0) {
            const uint16_t pair_mask = uint16_t(((1u << CG) - 1u) << leader);
            const uint16_t cluster_mask = uint16_t((1u << kCluster) - 1u);
            int s = 0; uint32_t ph = 0;
            int it = 0;
            for (int t = cluster_id; t < num_tiles; t += num_clusters, ++it) {
                const int acc = it & 1;
                const uint32_t acc_ph = (it >> 1) & 1;
                mbar_wait(tempty_bar(acc), acc_ph ^ 1u);
                tc_fence_after_sync();
                const uint32_t d_tmem = tmem_base + uint32_t(acc * kUmmaN);
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(full_bar(s), ph);
                    tc_fence_after_sync();
                    if (lane == 0) {
                        const uint32_t st = smem_base + s * kStageBytes;
#pragma unroll
                        for (int kk = 0; kk < kBK / 16; ++kk) {
                            const uint64_t a_d = operand_desc<AMAJ>(st + kOffAraw, kk);
                            const uint64_t b_d = operand_desc<BMAJ>(st + kOffBraw, kk);
                            umma_f16<CG>(d_tmem, a_d, b_d, kIdesc, (kb | kk) != 0 ? 1u : 0u);
                        }
                        umma_commit<CG>(empty_bar(s), cluster_mask);
                        if (kb == num_kb - 1) umma_commit<CG>(tfull_bar(acc), pair_mask);
                    }
                    __syncwarp();
                    if (++s == kStages) { s = 0; ph ^= 1u; }
                }
            }
        }
    } else {
        // ------------------------------------------------------------ epilogue (warps 2-5)
        const int q = warp & 3;  // TMEM lane quadrant this warp may access
        int it = 0;
        for (int t = cluster_id; t < num_tiles; t += num_clusters, ++it) {
            int tm, tn;
            tile_coords(t, p.tiles_m, p.tiles_n, tm, tn);
            const int acc = it & 1;
            const uint32_t acc_ph = (it >> 1) & 1;
            const int row = (tm * PAIRS + int(pair)) * kUmmaM + int(cta_rank) * kTileMN + q * 32 + lane;
            const int n0 = tn * kUmmaN;
            mbar_wait(tfull_bar(acc), acc_ph);
            tc_fence_after_sync();
            const uint32_t t_addr = tmem_base + (uint32_t(q * 32) << 16) + uint32_t(acc * kUmmaN);
            const bool row_ok = row < p.m;
            __half* crow = p.C + row;
#pragma unroll 1
            for (int c0 = 0; c0 < kUmmaN; c0 += 32) {
                if (n0 + c0 >= p.n) break;  // warp-uniform
                float cold[32];
                if (p.beta != 0.f) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const int col = n0 + c0 + j;
                        cold[j] = (row_ok && col < p.n) ? __half2float(crow[int64_t(col) * p.ldc]) : 0.f;
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
                    if (row_ok && col < p.n) crow[int64_t(col) * p.ldc] = __float2half_rn(o);
                }
            }
            tc_fence_before_sync();
            __syncwarp();
            if (lane == 0) {
                if constexpr (CG == 2) mbar_arrive_cluster(tempty_bar(acc), leader);
                else mbar_arrive(tempty_bar(acc));
            }
        }
    }

    tc_fence_before_sync();
    if constexpr (kCluster > 1) cluster_sync_all(); else __syncthreads();
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

// Tensor map over a column-major matrix X (rows x cols, leading dimension ld, half).
//  as_mn_major: the GEMM's M/N index runs down X's columns (contiguous) -> box 64(mn) x 64(k)
//  else       : the GEMM's K index runs down X's columns (contiguous)   -> box 64(k)  x 128(mn), or x 64(mn) when the
//               128 rows are fetched as two multicast halves (half_box)
bool encode_operand(CUtensorMap* tm, const __half* X, int64_t rows, int64_t cols, int64_t ld,
                    bool as_mn_major, bool half_box = false) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) return false;
    cuuint64_t dims[2] = {cuuint64_t(rows), cuuint64_t(cols)};
    cuuint64_t strides[1] = {cuuint64_t(ld) * 2};
    cuuint32_t box[2] = {64u, (as_mn_major || half_box) ? 64u : 128u};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<__half*>(X), dims, strides,
                     box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS;
}

template <int CG, int AMAJ, int BMAJ, int PAIRS>
cudaError_t launch(const CUtensorMap& tmA, const CUtensorMap& tmB, const Params& p, int num_sms,
                   cudaStream_t stream) {
    auto kern = gemm_f16_kernel<CG, AMAJ, BMAJ, PAIRS>;
    constexpr int kCluster = CG * PAIRS;
    static PerDeviceOnce once;
    int dev;
    if (once.needs(&dev)) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             kSmemBytes);
        if (e != cudaSuccess) return e;
        once.set(dev);
    }
    const int num_tiles = p.tiles_m * p.tiles_n;
    int clusters = num_sms / kCluster;
    if (clusters > num_tiles) clusters = num_tiles;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(unsigned(clusters * kCluster));
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = kSmemBytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = kCluster;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    count_launch();
    return cudaLaunchKernelEx(&cfg, kern, tmA, tmB, p);
}

}  // namespace

// transA/transB: 0 = 'N', 1 = 'T' (also 'C' for real types). cta_group: 1 or 2 (0 = default).
// Operands must be TMA-able: 16-byte aligned bases and leading dimensions that are multiples of 8 halfs
// (the engine's DistMat and El::Matrix pad their pitch to 16 bytes); anything else is ELB200_ERR_ARG.
int gemm_f16_tensor(int transA, int transB, int64_t m, int64_t n, int64_t k, float alpha,
                    const void* Av, int64_t lda, const void* Bv, int64_t ldb, float beta, void* Cv,
                    int64_t ldc, int cta_group, cudaStream_t stream) {
    const __half* A = static_cast<const __half*>(Av);
    const __half* B = static_cast<const __half*>(Bv);
    __half* C = static_cast<__half*>(Cv);
    if (cta_group == 0) cta_group = 2;
    if ((reinterpret_cast<uintptr_t>(A) & 15u) || (reinterpret_cast<uintptr_t>(B) & 15u) || (lda & 7) || (ldb & 7)) return ELB200_ERR_ARG;
    // A is (m x k) if N else (k x m), column-major. 'N' -> m contiguous -> MN-major.
    const bool a_mn = (transA == 0);
    // B is (k x n) if N else (n x k). 'N' -> k contiguous -> K-major; 'T' -> n contiguous -> MN-major.
    const bool b_mn = (transB != 0);
    CUtensorMap tmA, tmB;
    const int64_t a_rows = a_mn ? m : k, a_cols = a_mn ? k : m;
    const int64_t b_rows = b_mn ? n : k, b_cols = b_mn ? k : n;
    const int tile = 128 * cta_group;
    const int tiles_m = int((m + tile - 1) / tile);
    // Two tiles that are neighbours in m share their B tile: with ELB200_F16_MULTICAST=1 a 4-CTA cluster fetches it once
    // and multicasts it (L2 -> SM bytes per flop drop by a quarter). OPT-IN because it measured SLOWER on B200 — 790 vs
    // 1294 TFLOP/s at n = 8192, 819 vs 1349 at 16384 (profiles/r02_f16_multicast.log): the two pairs of a cluster then
    // advance in lockstep through every stage (each stage's empty barrier needs both pairs' commits, each full barrier
    // loads issued by all four CTAs), and at cluster size 4 L2 already merges most of the duplicate reads.
    static const bool mc_on = [] { const char* v = std::getenv("ELB200_F16_MULTICAST"); return v && std::atoi(v) != 0; }();
    const int pairs = (cta_group == 2 && tiles_m >= 2 && mc_on) ? 2 : 1;
    if (!encode_operand(&tmA, A, a_rows, a_cols, lda, a_mn)) return ELB200_ERR_TENSORMAP;
    if (!encode_operand(&tmB, B, b_rows, b_cols, ldb, b_mn, pairs == 2)) return ELB200_ERR_TENSORMAP;
    Params p;
    p.C = C; p.ldc = ldc; p.m = int(m); p.n = int(n); p.k = int(k);
    p.alpha = alpha; p.beta = beta;
    p.tiles_m = (tiles_m + pairs - 1) / pairs;
    p.tiles_n = int((n + tile - 1) / tile);
    const int num_sms = device_sm_count();
    cudaError_t e;
#define ELB_DISPATCH(CGV, PR)                                                                   \
    if (a_mn && b_mn) e = launch<CGV, MAJOR_MN, MAJOR_MN, PR>(tmA, tmB, p, num_sms, stream);    \
    else if (a_mn && !b_mn) e = launch<CGV, MAJOR_MN, MAJOR_K, PR>(tmA, tmB, p, num_sms, stream); \
    else if (!a_mn && b_mn) e = launch<CGV, MAJOR_K, MAJOR_MN, PR>(tmA, tmB, p, num_sms, stream); \
    else e = launch<CGV, MAJOR_K, MAJOR_K, PR>(tmA, tmB, p, num_sms, stream);
    if (pairs == 2) { ELB_DISPATCH(2, 2) } else if (cta_group == 2) { ELB_DISPATCH(2, 1) } else { ELB_DISPATCH(1, 1) }
#undef ELB_DISPATCH
    return e == cudaSuccess ? ELB200_OK : ELB200_ERR_CUDA;
}

}  // namespace elb
