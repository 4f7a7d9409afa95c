// Shared-memory addressing of the TMA-fed fp64 GEMM (gemm_f64_tma.cu), as plain host/device functions so
// that tests/test_f64_tma_layout.py can run the very same index arithmetic on the CPU (a model of the TMA
// box fill + 128-byte swizzle + the m8n8k4 fragment semantics) before the kernel meets a GPU.
//
// Stage layout of one operand tile (128 mn x 32 k doubles = 32 KiB, dense, TMA SWIZZLE_128B):
//   MN-contiguous operand (X(mn,k) at X[k*ld + mn]):  8 boxes of {16 mn doubles = 128 B} x {32 k rows};
//       box b = mn / 16 at byte b*4096, row = k
//   K-contiguous operand  (X(mn,k) at X[mn*ld + k]):  2 boxes of {16 k doubles = 128 B} x {128 mn rows};
//       box b = k / 16 at byte b*16384, row = mn
// Inside a box rows are 128 B and the 16-byte chunk index is XORed with (row & 7).
//
// Fragment order (tools/f64_tma_layout_check.py): k of step s (0..7 per 32-k tile) for lane lk is
//   k(s, lk) = 8*(s/2) + 2*lk + (s & 1)
// MN-contiguous tiles read row k directly (8-byte loads: rows {0,2,4,6} / {1,3,5,7} of a swizzle atom are
// conflict-free per half-warp); K-contiguous tiles read the PAIR (k = 8*(s/2) + 2*lk, +1) with one 16-byte
// load on even steps, with their 8 fragment rows permuted by kRowPerm (conflict-free per quarter-warp).
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define ELB_HD __host__ __device__ __forceinline__
#else
#define ELB_HD inline
#endif

namespace elb {
namespace f64tma {

constexpr int kBM = 128, kBN = 128, kBK = 32;
constexpr int kOperandBytes = kBM * kBK * 8;  // 32 KiB

// byte offset inside one box (rows of 128 B) of byte `b` of row `row` under the 128-byte swizzle
ELB_HD uint32_t swz(uint32_t row, uint32_t b) { return row * 128u + ((((b >> 4) ^ (row & 7u)) << 4) | (b & 15u)); }

// byte offset of element (mn, k) inside an operand's stage tile
ELB_HD uint32_t off_mn_contig(uint32_t mn, uint32_t k) { return (mn >> 4) * 4096u + swz(k, (mn & 15u) * 8u); }
ELB_HD uint32_t off_k_contig(uint32_t mn, uint32_t k) { return (k >> 4) * 16384u + swz(mn, (k & 15u) * 8u); }

// k index a lane supplies in step s of a k-tile
ELB_HD uint32_t k_of_step(uint32_t s, uint32_t lk) { return 8u * (s >> 1) + 2u * lk + (s & 1u); }

// fragment-row permutation of K-contiguous tiles: lane lr of an 8-row fragment handles row kRowPerm(lr)
ELB_HD uint32_t row_perm(uint32_t lr) { return ((lr & 1u) << 2) | (lr >> 1); }  // 0,4,1,5,2,6,3,7

// The same offsets split into a per-lane part (a handful of registers, computed once) and a part that depends
// only on the unrolled loop indices (immediates). `g0` = first row of the warp's 64- or 32-row slab (a multiple
// of 8), `i` = 8-row fragment index inside the slab, `st` = mma k-step (0..7), lr = lane / 4, lk = lane % 4.
//   MN-contiguous: element (mn = g0 + 8 i + lr, k = k_of_step(st, lk))
struct LaneMN { uint32_t x[2][2]; };  // [i & 1][st & 1]
ELB_HD LaneMN lane_mn(uint32_t g0, uint32_t lr, uint32_t lk) {
    LaneMN l;
    for (uint32_t ip = 0; ip < 2; ++ip)
        for (uint32_t par = 0; par < 2; ++par) {
            const uint32_t chunk = ((ip << 2) | (lr >> 1)) ^ (2u * lk + par);
            l.x[ip][par] = (g0 >> 4) * 4096u + (2u * lk + par) * 128u + (chunk << 4) + (lr & 1u) * 8u;
        }
    return l;
}
ELB_HD uint32_t fast_mn(const LaneMN& l, uint32_t i, uint32_t st) { return l.x[i & 1u][st & 1u] + (i >> 1) * 4096u + (st >> 1) * 1024u; }
//   K-contiguous: the 16-byte PAIR (mn = g0 + 8 i + row_perm(lr), k = k_of_step(st, lk) & ~1)
struct LaneK { uint32_t y[2]; };      // [(st >> 1) & 1]
ELB_HD LaneK lane_k(uint32_t g0, uint32_t lr, uint32_t lk) {
    LaneK l;
    const uint32_t r = row_perm(lr);
    for (uint32_t h = 0; h < 2; ++h) l.y[h] = (g0 + r) * 128u + ((((h << 2) | lk) ^ r) << 4);
    return l;
}
ELB_HD uint32_t fast_k(const LaneK& l, uint32_t i, uint32_t st) { return l.y[(st >> 1) & 1u] + i * 1024u + (st >> 2) * 16384u; }

}  // namespace f64tma
}  // namespace elb
