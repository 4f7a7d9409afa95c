// CPU emulation of the TMA-fed fp64 GEMM's shared-memory addressing (elemental_b200/csrc/gemm_f64_layout.hpp):
// a stage tile is filled the way TMA + SWIZZLE_128B would fill it, every consumer lane reads its fragments
// through the kernel's own offset functions, the m8n8k4 mma is applied with its documented fragment layout
// (A: lane holds (row lane/4, k lane%4); B: (k lane%4, col lane/4); D: (row lane/4, cols 2(lane%4), +1)), and
// the epilogue's (m, n) mapping scatters the accumulators — the result must be the tile product, for all four
// orientation pairs. Also checks that the offsets are a bijection onto the 32 KiB tile and that the
// lane-part / immediate split (fast_mn, fast_k) equals the plain functions.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

#include "../../elemental_b200/csrc/gemm_f64_layout.hpp"

using namespace elb::f64tma;

static double opA(int m, int k) { return std::sin(0.37 * m + 1.3 * k) + 0.01 * m; }
static double opB(int k, int n) { return std::cos(0.11 * n - 0.7 * k) - 0.02 * n; }

// shared-memory wavefronts of one warp-wide load: 8-byte loads are served per half-warp, 16-byte loads per
// quarter-warp; within such a group every bank may be asked for one distinct 4-byte word only
static int wavefronts(const uint32_t (&addr)[32], int bytes) {
    const int group = bytes == 8 ? 16 : 8;
    int worst = 0;
    for (int g0 = 0; g0 < 32; g0 += group) {
        int count[32] = {};
        uint32_t word[32][8];
        for (int l = g0; l < g0 + group; ++l)
            for (int w = 0; w < bytes / 4; ++w) {
                const uint32_t wd = addr[l] / 4 + w, bank = wd % 32;
                bool seen = false;
                for (int i = 0; i < count[bank]; ++i) seen |= word[bank][i] == wd;
                if (!seen) word[bank][count[bank]++] = wd;
            }
        for (int b = 0; b < 32; ++b) worst = count[b] > worst ? count[b] : worst;
    }
    return worst;
}

static int run(bool TA, bool TB) {
    int bad = 0;
    std::vector<uint8_t> sa(kOperandBytes, 0xFF), sb(kOperandBytes, 0xFF);
    std::vector<int> hitA(kOperandBytes / 8, 0), hitB(kOperandBytes / 8, 0);
    // ---- what TMA lands: op(A)(m,k) tile: MN-contiguous when !TA; op(B)(k,n): MN-contiguous (n) when TB
    for (int mn = 0; mn < kBM; ++mn)
        for (int k = 0; k < kBK; ++k) {
            const uint32_t oa = TA ? off_k_contig(mn, k) : off_mn_contig(mn, k);
            const uint32_t ob = TB ? off_mn_contig(mn, k) : off_k_contig(mn, k);
            if (oa % 8 || oa + 8 > uint32_t(kOperandBytes) || ob % 8 || ob + 8 > uint32_t(kOperandBytes)) { ++bad; continue; }
            ++hitA[oa / 8]; ++hitB[ob / 8];
            const double va = opA(mn, k), vb = opB(k, mn);
            std::memcpy(&sa[oa], &va, 8);
            std::memcpy(&sb[ob], &vb, 8);
        }
    for (int h : hitA) bad += h != 1;
    for (int h : hitB) bad += h != 1;
    // TMA writes rows of 128 contiguous bytes: inside a box row the swizzle may only permute 16-byte chunks
    for (int row = 0; row < 32; ++row)
        for (int b = 0; b < 128; ++b) bad += swz(row, b) / 128 != uint32_t(row) || (swz(row, b) & 15u) != uint32_t(b & 15);

    std::vector<double> C(kBM * kBN, 0.0);
    for (int warp = 0; warp < 8; ++warp) {
        const int wm = (warp & 1) * 64, wn = (warp >> 1) * 32;
        double acc[32][4][8][2] = {};
        for (int st = 0; st < kBK / 4; ++st) {
            double bf[32][8], af[32][4];
            for (int lane = 0; lane < 32; ++lane) {
                const uint32_t lr = lane >> 2, lk = lane & 3;
                const LaneMN a_mn = lane_mn(wm, lr, lk), b_mn = lane_mn(wn, lr, lk);
                const LaneK a_k = lane_k(wm, lr, lk), b_k = lane_k(wn, lr, lk);
                const uint32_t kk = k_of_step(st, lk);
                for (int mi = 0; mi < 8; ++mi) {
                    if (TA) {
                        const uint32_t o = fast_k(a_k, mi, st);
                        bad += o != off_k_contig(wm + mi * 8 + row_perm(lr), kk & ~1u) || o % 16;
                        double pair[2];
                        std::memcpy(pair, &sa[o], 16);
                        bf[lane][mi] = pair[st & 1];
                    } else {
                        const uint32_t o = fast_mn(a_mn, mi, st);
                        bad += o != off_mn_contig(wm + mi * 8 + lr, kk);
                        std::memcpy(&bf[lane][mi], &sa[o], 8);
                    }
                }
                for (int ni = 0; ni < 4; ++ni) {
                    if (TB) {
                        const uint32_t o = fast_mn(b_mn, ni, st);
                        bad += o != off_mn_contig(wn + ni * 8 + lr, kk);
                        std::memcpy(&af[lane][ni], &sb[o], 8);
                    } else {
                        const uint32_t o = fast_k(b_k, ni, st);
                        bad += o != off_k_contig(wn + ni * 8 + row_perm(lr), kk & ~1u) || o % 16;
                        double pair[2];
                        std::memcpy(pair, &sb[o], 16);
                        af[lane][ni] = pair[st & 1];
                    }
                }
            }
            // every fragment load of this step is one wavefront per half- / quarter-warp (no bank conflicts)
            for (int mi = 0; mi < 8; ++mi) {
                uint32_t ad[32];
                for (int lane = 0; lane < 32; ++lane)
                    ad[lane] = TA ? fast_k(lane_k(wm, lane >> 2, lane & 3), mi, st) : fast_mn(lane_mn(wm, lane >> 2, lane & 3), mi, st);
                bad += wavefronts(ad, TA ? 16 : 8) != 1;
            }
            for (int ni = 0; ni < 4; ++ni) {
                uint32_t ad[32];
                for (int lane = 0; lane < 32; ++lane)
                    ad[lane] = TB ? fast_mn(lane_mn(wn, lane >> 2, lane & 3), ni, st) : fast_k(lane_k(wn, lane >> 2, lane & 3), ni, st);
                bad += wavefronts(ad, TB ? 8 : 16) != 1;
            }
            // mma.sync.m8n8k4: D(r, c) += sum_q A(r, q) B(q, c); A(r, q) is lane 4r+q's a operand, B(q, c) lane 4c+q's b operand
            for (int ni = 0; ni < 4; ++ni)
                for (int mi = 0; mi < 8; ++mi)
                    for (int lane = 0; lane < 32; ++lane) {
                        const int r = lane >> 2;
                        for (int e = 0; e < 2; ++e) {
                            const int c = 2 * (lane & 3) + e;
                            double s = 0;
                            for (int q = 0; q < 4; ++q) s += af[4 * r + q][ni] * bf[4 * c + q][mi];
                            acc[lane][ni][mi][e] += s;
                        }
                    }
        }
        // epilogue mapping of gemm_f64_dmma_tma_kernel
        for (int lane = 0; lane < 32; ++lane) {
            const uint32_t lr = lane >> 2, lk = lane & 3;
            const uint32_t rowB = TB ? lr : row_perm(lr);
            for (int ni = 0; ni < 4; ++ni)
                for (int mi = 0; mi < 8; ++mi)
                    for (int e = 0; e < 2; ++e) {
                        const int n = wn + ni * 8 + int(rowB);
                        const int m = wm + mi * 8 + int(TA ? row_perm(2 * lk + e) : 2 * lk + e);
                        C[m + n * kBM] += acc[lane][ni][mi][e];
                    }
        }
    }
    double worst = 0;
    for (int n = 0; n < kBN; ++n)
        for (int m = 0; m < kBM; ++m) {
            double ref = 0;
            for (int k = 0; k < kBK; ++k) ref += opA(m, k) * opB(k, n);
            worst = std::fmax(worst, std::fabs(C[m + n * kBM] - ref));
        }
    if (!(worst < 1e-11)) ++bad;
    std::printf("TA=%d TB=%d: max |C - ref| = %.2e, %s\n", int(TA), int(TB), worst, bad ? "FAIL" : "ok");
    return bad;
}

int main() {
    int bad = 0;
    for (int ta = 0; ta < 2; ++ta)
        for (int tb = 0; tb < 2; ++tb) bad += run(ta, tb);
    std::printf(bad ? "F64_TMA_LAYOUT_FAIL %d\n" : "F64_TMA_LAYOUT_PASS\n", bad);
    return bad ? 1 : 0;
}
