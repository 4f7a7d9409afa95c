// Distributed Gemm: C = alpha op(A) op(B) + beta C over [MC,MR] DistMats — the SUMMA A/B/C/Dot
// families for all four orientation pairs (reference src/blas_like/level3/Gemm.cpp:273-302 and
// Gemm/{NN,NT,TN,TT}.hpp).
//
// B200-first differences from the reference's schedule (results agree within the GEMM tolerance,
// redistribution results are bit-identical):
//  * panels are `PanelWidth()` wide (a whole number of algorithmic blocks, default 2048 fp32 /
//    1024 fp64) instead of one block of 128: one pass over the local C then amortises enough k
//    for the tensor pipe to be the bound, not HBM (SURVEY.md §7 hard part 1);
//  * the local GEMM consumes panels in whichever majorness the gather produces ([MC,*], [*,MC],
//    [MR,*], [*,MR]) through its TMA/UMMA descriptors, so the reference's explicit panel
//    transposes (Transpose(B1, B1Trans_MR_STAR) etc.) disappear;
//  * a panel whose redistribution is the identity on this grid (1 x 1, or c == 1 / r == 1 along the
//    gathered dimension) is used in place, without a copy.
#include <algorithm>
#include <cstdlib>

#include "engine.hpp"
#include "engine_dev.hpp"

namespace elb {

// Optional per-launch timing of the local GEMM (bench.py's roofline leg): CUDA events on the
// launching stream around every LocalGemm kernel; durations are read back after the run.
namespace {
struct GemmProfile {
    bool on = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ev;
    double flop = 0;
} g_prof;
}  // namespace

void GemmProfileEnable(bool on) { g_prof.on = on; }
void GemmProfileCollect(double* ms, double* flop, int* launches) {
    double total = 0;
    for (auto& e : g_prof.ev) {
        float t = 0;
        cudaEventSynchronize(e.second);
        cudaEventElapsedTime(&t, e.first, e.second);
        total += t;
        cudaEventDestroy(e.first);
        cudaEventDestroy(e.second);
    }
    if (ms) *ms = total;
    if (flop) *flop = g_prof.flop;
    if (launches) *launches = int(g_prof.ev.size());
    g_prof.ev.clear();
    g_prof.flop = 0;
}

// one local GEMM launch on `s`, bracketed by events when the measurement hook is on
template <typename T>
void gemm_profiled(int ta, int tb, int64_t m, int64_t n, int64_t k, T alpha, const T* A, int64_t lda, const T* B,
                   int64_t ldb, T beta, T* C, int64_t ldc, cudaStream_t s) {
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    const bool prof = g_prof.on && m > 0 && n > 0 && k > 0;
    if (prof) {
        cuda_check(cudaEventCreate(&e0), "cudaEventCreate");
        cuda_check(cudaEventCreate(&e1), "cudaEventCreate");
        cuda_check(cudaEventRecord(e0, s), "cudaEventRecord");
    }
    Dev<T>::gemm(ta, tb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, s);
    if (prof) {
        cuda_check(cudaEventRecord(e1, s), "cudaEventRecord");
        g_prof.ev.emplace_back(e0, e1);
        g_prof.flop += 2.0 * double(m) * double(n) * double(k);
    }
}

template <typename T>
void LocalGemm(int ta, int tb, T alpha, const DistMat<T>& A, const DistMat<T>& B, T beta, DistMat<T>& C) {
    const int64_t m = C.lh(), n = C.lw();
    const int64_t ka = ta ? A.lh() : A.lw(), kb = tb ? B.lw() : B.lh();
    const int64_t ma = ta ? A.lw() : A.lh(), nb = tb ? B.lh() : B.lw();
    if (ka != kb || ma != m || nb != n)
        throw EngineError(ELB200_ERR_ARG, "LocalGemm: nonconformal local matrices");
    gemm_profiled<T>(ta, tb, m, n, ka, alpha, A.buf, A.ldim, B.buf, B.ldim, beta, C.buf, C.ldim, C.stream);
}

int ResolveGemmAlg(int, int, int64_t m, int64_t n, int64_t k, int alg) {
    // reference Gemm/NN.hpp:572-600 (same constants in NT/TN/TT)
    if (alg == GEMM_DEFAULT) {
        const double weightTowardsC = 2., weightAwayFromDot = 10.;
        if (weightAwayFromDot * m <= k && weightAwayFromDot * n <= k) alg = GEMM_SUMMA_DOT;
        else if (m <= n && weightTowardsC * m <= k) alg = GEMM_SUMMA_B;
        else if (n <= m && weightTowardsC * n <= k) alg = GEMM_SUMMA_A;
        else alg = GEMM_SUMMA_C;
    }
    // the multistream twins differ only in schedule; here every variant already pipelines
    if (alg == GEMM_SUMMA_A_MS) alg = GEMM_SUMMA_A;
    if (alg == GEMM_SUMMA_B_MS) alg = GEMM_SUMMA_B;
    if (alg == GEMM_SUMMA_C_MS) alg = GEMM_SUMMA_C;
    return alg;
}

namespace {


// src redistributed as (col,row) aligned with `like`; returns src itself when that
// redistribution is the identity on every rank (no data motion, no copy).
template <typename T>
const DistMat<T>& panel_as(const DistMat<T>& src, DistMat<T>& tmp, const DistSpec& like) {
    tmp.col_constrained = tmp.row_constrained = false;
    tmp.spec.col_align = tmp.spec.row_align = 0;
    tmp.AlignWith(like, true);
    Grid& g = *src.grid;
    const DistSpec target = ResolveCopySpec(g.shape, src.spec, tmp.spec, tmp.col_constrained, tmp.row_constrained);
    const CopyPlan plan = plan_copy(g.shape, g.me, src.spec, target);
    if (plan.kind == EX_LOCAL) {
        bool identity_everywhere = true;  // identical local shapes on every rank <=> strides/shifts agree
        if (dist_stride(g.shape, src.spec.col) != dist_stride(g.shape, target.col) ||
            dist_stride(g.shape, src.spec.row) != dist_stride(g.shape, target.row))
            identity_everywhere = false;
        for (int v = 0; v < g.size() && identity_everywhere; ++v) {
            const GridCoord q = coord_of_vc(g.shape, v);
            if (Shift(dist_rank(g.shape, src.spec.col, q), src.spec.col_align, dist_stride(g.shape, src.spec.col)) !=
                    Shift(dist_rank(g.shape, target.col, q), target.col_align, dist_stride(g.shape, target.col)) ||
                Shift(dist_rank(g.shape, src.spec.row, q), src.spec.row_align, dist_stride(g.shape, src.spec.row)) !=
                    Shift(dist_rank(g.shape, target.row, q), target.row_align, dist_stride(g.shape, target.row)))
                identity_everywhere = false;
        }
        if (identity_everywhere) return src;
    }
    Copy(src, tmp);
    return tmp;
}

struct Dims { int64_t m, n, k; };

// ---- stationary C, NN, aligned operands: zero-copy panels over NVLink ----------------------------
// Both panels of a k-range are AllGathers (A1 -> [MC,*] over the row communicator, B1 -> [*,MR] over
// the column communicator, Gemm/NN.hpp:363-377).  A sum over k does not care about the order of k,
// so the panel is laid out with the k indices PERMUTED such that every member's contribution is a
// pitched block: rows of the B panel are grouped by owner (member q's local rows are contiguous),
// and the columns of the A panel follow the same order.  Then each contribution is a handful of
// cudaMemcpy2DAsync transfers (copy engines) straight from the owner's local matrix into the
// panel inside every member's peer-mapped arena slot: no staging buffer, no pack or unpack kernel,
// no SM touched while the previous panel's GEMM runs.
template <typename T>
void SummaCDirectNN(T alpha, const DistMat<T>& A, const DistMat<T>& B, T beta, DistMat<T>& C, Dims d, int64_t W) {
    Grid& g = *C.grid;
    PeerArena& ar = *g.arena;
    const int64_t sz = int64_t(sizeof(T)), vec = 16 / sz;
    cudaStream_t cs = g.comm_stream;
    const std::vector<int> rowm = g.comm_members(MR), colm = g.comm_members(MC);
    const int ridx = g.me.mr, cidx = g.me.mc;
    const int64_t mloc = C.lh(), nloc = C.lw();
    const int64_t ldpa = std::max(vec, (mloc + vec - 1) / vec * vec);
    Ev ready[2], freed[2], start;
    PeerArena::Ticket ticketA[2], ticketB[2];
    bool live[2] = {false, false};
    start.record(C.stream);
    start.block(cs);
    const int64_t nbk = Grid::Blocksize();
    auto width = [&](int64_t p) { return p >= 2 ? W : std::max(nbk, (W >> (2 - p)) / nbk * nbk); };
    int64_t p = 0;
    for (int64_t k0 = 0, kb = 0; k0 < d.k; k0 += kb, ++p) {
        const int s = int(p & 1);
        kb = std::min(width(p), d.k - k0);
        if (live[s]) {  // the GEMM that read this slot two panels ago is done: hand the slot back
            freed[s].block(cs);
            ar.Release(ticketA[s], rowm, ridx, cs);
            ar.Release(ticketB[s], colm, cidx, cs);
            live[s] = false;
        }
        const DistMat<T> A1 = A.View(0, d.m, k0, k0 + kb);
        const DistMat<T> B1 = B.View(k0, k0 + kb, 0, d.n);
        const int64_t ldpb = std::max(vec, (kb + vec - 1) / vec * vec);
        const DirectPanelPlan plan = plan_direct_panel(g.shape, g.me, kb, A1.spec.row_align, B1.spec.col_align);
        std::vector<PeerArena::Copy2D> ca;
        for (const PanelCols& pc : plan.a_cols) {
            PeerArena::Copy2D cp;
            cp.dst_offset = size_t(pc.dst_col0 * ldpa * sz);
            cp.dpitch = size_t(pc.dst_step * ldpa * sz);
            cp.src = A1.buf + pc.src_col0 * A1.ldim;
            cp.spitch = size_t(pc.src_step * A1.ldim * sz);
            cp.width = size_t(mloc * sz);
            cp.height = size_t(pc.count);
            ca.push_back(cp);
        }
        ticketA[s] = ar.GatherDirect(Grid::arena_channel(MR), rowm, ridx, ca, cs);
        // my rows of the B panel: one block at row offset off[my mc]
        std::vector<PeerArena::Copy2D> cb(1);
        cb[0].dst_offset = size_t(plan.b_row0 * sz);
        cb[0].dpitch = size_t(ldpb * sz);
        cb[0].src = B1.buf;
        cb[0].spitch = size_t(B1.ldim * sz);
        cb[0].width = size_t(plan.b_rows * sz);
        cb[0].height = size_t(nloc);
        ticketB[s] = ar.GatherDirect(Grid::arena_channel(MC), colm, cidx, cb, cs);
        live[s] = true;
        ready[s].record(cs);
        ready[s].block(C.stream);
        gemm_profiled<T>(0, 0, mloc, nloc, kb, alpha, static_cast<const T*>(ticketA[s].data), ldpa,
                         static_cast<const T*>(ticketB[s].data), ldpb, p == 0 ? beta : T(1), C.buf, C.ldim, C.stream);
        freed[s].record(C.stream);
    }
    for (int s = 0; s < 2; ++s)
        if (live[s]) {
            freed[s].block(cs);
            ar.Release(ticketA[s], rowm, ridx, cs);
            ar.Release(ticketB[s], colm, cidx, cs);
        }
}

// ---- stationary C: C += alpha op(A)(:,kr) op(B)(kr,:) -----------------------------------------
template <typename T>
void SummaC(int ta, int tb, T alpha, const DistMat<T>& A, const DistMat<T>& B, T beta, DistMat<T>& C, Dims d) {
    // `beta` is applied by the FIRST panel's GEMM (C = alpha*a*b + beta*C), later panels accumulate:
    // the separate Scale(beta, C) pass of the reference (Gemm.cpp:282) — one more read and write of
    // the whole local C — is folded into the kernel epilogue.  beta == 0 never reads C.
    Grid& g = *C.grid;
    // on a 1 x 1 grid no panel moves: the whole k range is one local GEMM (C is read once)
    const int64_t W = g.size() == 1 ? std::max<int64_t>(d.k, 1) : Grid::PanelWidth(sizeof(T));
    if (g.size() > 1 && g.comm_stream && g.arena && !ta && !tb) {
        static const bool direct_on = [] { const char* e = std::getenv("ELB200_DIRECT"); return !(e && std::atoi(e) == 0); }();
        // the two panels live in arena slots: shrink the panel width until they fit
        const int64_t vec = 16 / int64_t(sizeof(T)), nbk = Grid::Blocksize();
        // sized from the LARGEST local extents on the grid, not this rank's: ragged matrices give ranks local
        // heights / widths that differ by one, and every rank must derive the same Wd and take the same branch
        // (the panels are collectives: a different kb or path on one rank would hang its peers)
        const int64_t per_k = std::max(MaxLength(d.m, g.shape.r) + vec, MaxLength(d.n, g.shape.c)) * int64_t(sizeof(T));
        int64_t Wd = std::min<int64_t>(W, (int64_t(g.arena->slot_bytes()) / per_k - vec) / nbk * nbk);
        if (direct_on && Wd >= std::max<int64_t>(nbk, 512) && A.spec.col_align == C.spec.col_align &&
            B.spec.row_align == C.spec.row_align) {
            SummaCDirectNN(alpha, A, B, beta, C, d, Wd);
            return;
        }
    }
    // An SM-resident NCCL gather fights the persistent GEMM for SMs (measured: slower than no
    // overlap at all), so the pipeline is used only with the copy-engine peer arena.
    if (g.size() == 1 || !g.comm_stream || !g.arena) {
        DistMat<T> PA(g, ta ? STAR : MC, ta ? MC : STAR, C.stream);
        DistMat<T> PB(g, tb ? MR : STAR, tb ? STAR : MR, C.stream);
        for (int64_t k0 = 0; k0 < d.k; k0 += W) {
            const int64_t kb = std::min(W, d.k - k0);
            DistMat<T> A1 = ta ? A.View(k0, k0 + kb, 0, d.m) : A.View(0, d.m, k0, k0 + kb);
            DistMat<T> B1 = tb ? B.View(0, d.n, k0, k0 + kb) : B.View(k0, k0 + kb, 0, d.n);
            const DistMat<T>& a = panel_as(A1, PA, C.spec);  // [MC,*] (or [*,MC]) aligned with C's rows
            const DistMat<T>& b = panel_as(B1, PB, C.spec);  // [*,MR] (or [MR,*]) aligned with C's cols
            LocalGemm(ta, tb, alpha, a, b, k0 == 0 ? beta : T(1), C);
        }
        return;
    }
    // Software pipeline over two panel slots: the gathers of panel p+1 run on the comm stream
    // (copy engines + small unpack kernels) while the local GEMM of panel p owns the SMs.
    //   ready[s]: panel in slot s is complete        (comm stream -> compute stream)
    //   freed[s]: the GEMM that read slot s is done   (compute stream -> comm stream)
    cudaStream_t cs = g.comm_stream;
    DistMat<T> PA[2] = {DistMat<T>(g, ta ? STAR : MC, ta ? MC : STAR, cs), DistMat<T>(g, ta ? STAR : MC, ta ? MC : STAR, cs)};
    DistMat<T> PB[2] = {DistMat<T>(g, tb ? MR : STAR, tb ? STAR : MR, cs), DistMat<T>(g, tb ? MR : STAR, tb ? STAR : MR, cs)};
    Ev ready[2], freed[2], start;  // RAII: released on every exit, a throwing redistribution included
    start.record(C.stream);        // A, B, C are ready in C.stream order
    start.block(cs);
    // ramp: the first gather cannot hide behind a GEMM, so the first two panels are narrower
    const int64_t nbk = Grid::Blocksize();
    auto width = [&](int64_t p) { return p >= 2 ? W : std::max(nbk, (W >> (2 - p)) / nbk * nbk); };
    int64_t p = 0;
    for (int64_t k0 = 0, kb = 0; k0 < d.k; k0 += kb, ++p) {
        const int s = int(p & 1);
        kb = std::min(width(p), d.k - k0);
        if (p >= 2) freed[s].block(cs);
        DistMat<T> A1 = ta ? A.View(k0, k0 + kb, 0, d.m) : A.View(0, d.m, k0, k0 + kb);
        DistMat<T> B1 = tb ? B.View(0, d.n, k0, k0 + kb) : B.View(k0, k0 + kb, 0, d.n);
        // (a cap on the CTAs of unpack kernels running underneath a GEMM was tried and removed: measured, a short full-width
        // unpack disturbs the GEMM less than a long narrow one — 41.3 ms vs 42.6 / 43.9 ms of GEMM time per step at N = 2)
        const DistMat<T>& a = panel_as(A1, PA[s], C.spec);
        const DistMat<T>& b = panel_as(B1, PB[s], C.spec);
        ready[s].record(cs);
        ready[s].block(C.stream);
        LocalGemm(ta, tb, alpha, a, b, p == 0 ? beta : T(1), C);
        freed[s].record(C.stream);
    }
    // the panels are freed in comm-stream order: not before the last GEMMs have read them
    for (int i = 0; i < 2 && i < p; ++i) freed[i].block(cs);
}

// Two-slot panel pipeline shared by SUMMA-A / SUMMA-B: the redistribution of panel p+1 runs on the comm stream
// (copy engines + small pack / unpack kernels) under the local GEMM and the reduce-scatter of panel p.
//   ready[s]: panel in slot s is complete        (comm stream -> compute stream)
//   freed[s]: the GEMM that read slot s is done   (compute stream -> comm stream)
struct PanelPipe {
    bool on;
    cudaStream_t cs, ms;  // panel (comm) stream, main (compute) stream
    Ev ready[2], freed[2], start;
    int64_t p = 0;
    PanelPipe(Grid& g, cudaStream_t main_stream) : on(false), cs(main_stream), ms(main_stream) {
        // Here BOTH streams run arena operations with flag waits at the same time (the reduce-scatter of panel p on the
        // compute stream, the exchange of panel p+1 on the comm stream), on different arena channels. Parity on 2 x 2 and
        // 2 x 4 B200s, C6 on 1 x 2: SUMMA-B 70.7 -> 66.8 ms, SUMMA-A unchanged; ELB200_AB_PIPELINE=0 serialises them again.
        static const bool overlap_on = [] { const char* e = std::getenv("ELB200_AB_PIPELINE"); return !(e && std::atoi(e) == 0); }();
        on = g.size() > 1 && g.comm_stream && g.arena && overlap_on;
        if (on) {
            cs = g.comm_stream;
            start.record(ms);  // the operands are ready in main-stream order
            start.block(cs);
        }
    }
    int slot() const { return on ? int(p & 1) : 0; }
    void before_panel() { if (on && p >= 2) freed[slot()].block(cs); }      // slot reusable?
    void panel_ready() { if (on) { ready[slot()].record(cs); ready[slot()].block(ms); } }
    void panel_consumed() { if (on) freed[slot()].record(ms); ++p; }
    void finish() { if (on) for (int i = 0; i < 2 && i < p; ++i) freed[i].block(cs); }  // panels die in comm-stream order
};

// ---- stationary A: per column panel of C, D1 = alpha op(A) op(B)(:,panel) summed over the
//      grid dimension that shares op(A)'s k index, then scattered into C1 --------------------
template <typename T>
void SummaA(int ta, int tb, T alpha, const DistMat<T>& A, const DistMat<T>& B, DistMat<T>& C, Dims d) {
    Grid& g = *C.grid;
    const int64_t W = Grid::PanelWidth(sizeof(T));
    // k runs over MR when A is used as is (k = A's columns), over MC when A is transposed
    const Dist kd = ta ? MC : MR;
    PanelPipe pipe(g, C.stream);
    DistMat<T> PB[2] = {DistMat<T>(g, tb ? STAR : kd, tb ? kd : STAR, pipe.cs), DistMat<T>(g, tb ? STAR : kd, tb ? kd : STAR, pipe.cs)};  // op(B) panel: k x nb or nb x k
    DistMat<T> D1(g, ta ? MR : MC, STAR, C.stream);                   // m x nb partial sums
    DistMat<T> D1s(g, MR, MC, C.stream);
    for (int64_t j0 = 0; j0 < d.n; j0 += W) {
        const int64_t nb = std::min(W, d.n - j0);
        DistMat<T> B1 = tb ? B.View(j0, j0 + nb, 0, d.k) : B.View(0, d.k, j0, j0 + nb);
        DistMat<T> C1 = C.View(0, d.m, j0, j0 + nb);
        pipe.before_panel();
        const DistMat<T>& b = panel_as(B1, PB[pipe.slot()], A.spec);
        pipe.panel_ready();
        D1.col_constrained = D1.row_constrained = false;
        D1.AlignWith(A.spec, true);
        D1.Resize(d.m, nb);
        LocalGemm(ta, tb, alpha, A, b, T(0), D1);
        pipe.panel_consumed();
        if (!ta) {
            AxpyContract(T(1), D1, C1);          // [MC,*] -> [MC,MR]: ReduceScatter over MR
        } else {
            D1s.col_constrained = D1s.row_constrained = false;
            Contract(D1, D1s);                   // [MR,*] -> [MR,MC]: ReduceScatter over MC
            Axpy(T(1), D1s, C1);                 // [MR,MC] -> [MC,MR] exchange + local axpy
        }
    }
    pipe.finish();
}

// ---- stationary B: per row panel of C -----------------------------------------------------------
template <typename T>
void SummaB(int ta, int tb, T alpha, const DistMat<T>& A, const DistMat<T>& B, DistMat<T>& C, Dims d) {
    Grid& g = *C.grid;
    const int64_t W = Grid::PanelWidth(sizeof(T));
    // k runs over MC when B is used as is (k = B's rows), over MR when B is transposed
    const Dist kd = tb ? MR : MC;
    PanelPipe pipe(g, C.stream);
    DistMat<T> PA[2] = {DistMat<T>(g, ta ? kd : STAR, ta ? STAR : kd, pipe.cs), DistMat<T>(g, ta ? kd : STAR, ta ? STAR : kd, pipe.cs)};  // op(A) panel: mb x k or k x mb
    DistMat<T> D1(g, STAR, tb ? MC : MR, C.stream);                   // mb x n partial sums
    DistMat<T> D1s(g, MR, MC, C.stream);
    for (int64_t i0 = 0; i0 < d.m; i0 += W) {
        const int64_t mb = std::min(W, d.m - i0);
        DistMat<T> A1 = ta ? A.View(0, d.k, i0, i0 + mb) : A.View(i0, i0 + mb, 0, d.k);
        DistMat<T> C1 = C.View(i0, i0 + mb, 0, d.n);
        pipe.before_panel();
        const DistMat<T>& a = panel_as(A1, PA[pipe.slot()], B.spec);
        pipe.panel_ready();
        D1.col_constrained = D1.row_constrained = false;
        D1.AlignWith(B.spec, true);
        D1.Resize(mb, d.n);
        LocalGemm(ta, tb, alpha, a, B, T(0), D1);
        pipe.panel_consumed();
        if (!tb) {
            AxpyContract(T(1), D1, C1);          // [*,MR] -> [MC,MR]: ReduceScatter over MC
        } else {
            D1s.col_constrained = D1s.row_constrained = false;
            Contract(D1, D1s);                   // [*,MC] -> [MR,MC]: ReduceScatter over MR
            Axpy(T(1), D1s, C1);
        }
    }
    pipe.finish();
}

// ---- Dot: k spread over all p ranks, C tiles reduce-scattered over the whole grid ---------------
template <typename T>
void SummaDot(int ta, int tb, T alpha, const DistMat<T>& A, const DistMat<T>& B, DistMat<T>& C, Dims d) {
    Grid& g = *C.grid;
    constexpr int64_t kBlock = 8192;  // output tile of the Dot variant (the reference walks 2000 x 2000 tiles, NN.hpp:492-510)
    if (g.size() == 1) {
        // one rank: every distribution is the whole matrix, the k-split and the reduction are
        // identities — multiply straight into C (beta was applied by the caller)
        gemm_profiled<T>(ta, tb, d.m, d.n, d.k, alpha, A.buf, A.ldim, B.buf, B.ldim, T(1), C.buf, C.ldim, C.stream);
        return;
    }
    // Tall-skinny regime proper (one C tile): the up-front all-to-all of A and B is cut into k ranges and
    // pipelined against the local GEMM over two slots, like the SUMMA-C panels — the exchange of range c+1
    // (comm stream: pack, copy-engine pulls, unpack) runs under the GEMM of range c, and the tile accumulates
    // in C11 across ranges before its single reduce-scatter. ELB200_DOT_CHUNK = local k per range (default
    // 4096 fp32 / 2048 fp64). C5 at 2 x 4: 25.2 -> 21.7 ms (1395 -> 1624 TFLOP/s) against the one-shot exchange, which was removed.
    static const int64_t chunk_env = [] { const char* e = std::getenv("ELB200_DOT_CHUNK"); return e ? std::atoll(e) : 0ll; }();
    const int64_t chunk_local = chunk_env > 0 ? chunk_env : (sizeof(T) == 4 ? 4096 : 2048);
    const int64_t chunk = chunk_local * g.size();  // global k per range
    if (g.comm_stream && d.m <= kBlock && d.n <= kBlock && d.k >= 2 * chunk) {
        cudaStream_t cs = g.comm_stream;
        DistMat<T> Aks[2] = {DistMat<T>(g, ta ? VC : STAR, ta ? STAR : VC, cs), DistMat<T>(g, ta ? VC : STAR, ta ? STAR : VC, cs)};
        DistMat<T> Bks[2] = {DistMat<T>(g, tb ? STAR : VC, tb ? VC : STAR, cs), DistMat<T>(g, tb ? STAR : VC, tb ? VC : STAR, cs)};
        DistMat<T> C11p(g, STAR, STAR, C.stream);
        C11p.Resize(d.m, d.n);
        Ev ready[2], freed[2], start;
        start.record(C.stream);  // A, B (and the freshly allocated C11) are ready in C.stream order
        start.block(cs);
        int64_t p = 0;
        for (int64_t k0 = 0; k0 < d.k; k0 += chunk, ++p) {
            const int sl = int(p & 1);
            const int64_t k1 = std::min(d.k, k0 + chunk);
            if (p >= 2) freed[sl].block(cs);  // the GEMM that read this slot two ranges ago is done
            const DistMat<T> A1 = ta ? A.View(k0, k1, 0, d.m) : A.View(0, d.m, k0, k1);
            const DistMat<T> B1 = tb ? B.View(0, d.n, k0, k1) : B.View(k0, k1, 0, d.n);
            Aks[sl].col_constrained = Aks[sl].row_constrained = false;
            Bks[sl].col_constrained = Bks[sl].row_constrained = false;
            Aks[sl].Align(0, 0); Bks[sl].Align(0, 0);
            Copy(A1, Aks[sl]);
            Copy(B1, Bks[sl]);
            ready[sl].record(cs);
            ready[sl].block(C.stream);
            LocalGemm(ta, tb, alpha, Aks[sl], Bks[sl], p == 0 ? T(0) : T(1), C11p);
            freed[sl].record(C.stream);
        }
        // the slots are freed in comm-stream order: not before the last GEMMs have read them
        for (int i = 0; i < 2 && i < p; ++i) freed[i].block(cs);
        AxpyContract(T(1), C11p, C);
        return;
    }
    DistMat<T> Ak(g, ta ? VC : STAR, ta ? STAR : VC, C.stream);  // k over VC
    DistMat<T> Bk(g, tb ? STAR : VC, tb ? VC : STAR, C.stream);
    Ak.Align(0, 0); Bk.Align(0, 0);
    Copy(A, Ak);
    Copy(B, Bk);
    DistMat<T> C11(g, STAR, STAR, C.stream);
    for (int64_t i0 = 0; i0 < d.m; i0 += kBlock) {
        const int64_t mb = std::min(kBlock, d.m - i0);
        DistMat<T> A1 = ta ? Ak.View(0, d.k, i0, i0 + mb) : Ak.View(i0, i0 + mb, 0, d.k);
        for (int64_t j0 = 0; j0 < d.n; j0 += kBlock) {
            const int64_t nb = std::min(kBlock, d.n - j0);
            DistMat<T> B1 = tb ? Bk.View(j0, j0 + nb, 0, d.k) : Bk.View(0, d.k, j0, j0 + nb);
            DistMat<T> Cv = C.View(i0, i0 + mb, j0, j0 + nb);
            C11.Resize(mb, nb);
            LocalGemm(ta, tb, alpha, A1, B1, T(0), C11);
            AxpyContract(T(1), C11, Cv);
        }
    }
}

}  // namespace

template <typename T>
void Gemm(int ta, int tb, T alpha, const DistMat<T>& Apre, const DistMat<T>& Bpre, T beta, DistMat<T>& Cpre, int alg) {
    ta = ta ? 1 : 0; tb = tb ? 1 : 0;
    Grid& g = *Cpre.grid;
    if (Apre.grid != &g || Bpre.grid != &g) throw EngineError(ELB200_ERR_ARG, "Gemm: matrices live on different grids");
    Dims d;
    d.m = ta ? Apre.spec.width : Apre.spec.height;
    d.k = ta ? Apre.spec.height : Apre.spec.width;
    d.n = tb ? Bpre.spec.height : Bpre.spec.width;
    const int64_t kb = tb ? Bpre.spec.width : Bpre.spec.height;
    if (d.k != kb || Cpre.spec.height != d.m || Cpre.spec.width != d.n)
        throw EngineError(ELB200_ERR_ARG, "Gemm: nonconformal matrices");
    if (alg == GEMM_CANNON) throw EngineError(ELB200_ERR_ARG, "Cannon_NN: Bad device (CPU only in the reference, Gemm/NN.hpp:30-31)");
    if (d.m == 0 || d.n == 0 || d.k == 0 || alpha == T(0)) { Scale(beta, Cpre); return; }
    alg = ResolveGemmAlg(ta, tb, d.m, d.n, d.k, alg);
    // Stationary-C on a [MC,MR] C applies beta inside its first local GEMM; every other variant
    // accumulates into C through AxpyContract, so C is scaled up front as in Gemm.cpp:282.
    const bool fold_beta = alg == GEMM_SUMMA_C && Cpre.spec.col == MC && Cpre.spec.row == MR;
    if (!fold_beta) Scale(beta, Cpre);

    // Read / read-write proxies to [MC,MR] (reference Proxy.hpp:236-257: pass-through when the
    // matrix already has the requested distribution).
    const bool a_ok = Apre.spec.col == MC && Apre.spec.row == MR;
    const bool b_ok = Bpre.spec.col == MC && Bpre.spec.row == MR;
    const bool c_ok = Cpre.spec.col == MC && Cpre.spec.row == MR;
    DistMat<T> Ap(g, MC, MR, Cpre.stream), Bp(g, MC, MR, Cpre.stream), Cp(g, MC, MR, Cpre.stream);
    const bool dot = alg == GEMM_SUMMA_DOT;
    if (!a_ok && !dot) Copy(Apre, Ap);
    if (!b_ok && !dot) Copy(Bpre, Bp);
    if (!c_ok) Copy(Cpre, Cp);
    const DistMat<T>& A = (a_ok || dot) ? Apre : Ap;
    const DistMat<T>& B = (b_ok || dot) ? Bpre : Bp;
    DistMat<T>& C = c_ok ? Cpre : Cp;

    switch (alg) {
        case GEMM_SUMMA_A: SummaA(ta, tb, alpha, A, B, C, d); break;
        case GEMM_SUMMA_B: SummaB(ta, tb, alpha, A, B, C, d); break;
        case GEMM_SUMMA_C: SummaC(ta, tb, alpha, A, B, fold_beta ? beta : T(1), C, d); break;
        case GEMM_SUMMA_DOT: SummaDot(ta, tb, alpha, A, B, C, d); break;
        default: throw EngineError(ELB200_ERR_ARG, "Gemm: unsupported GemmAlgorithm");
    }
    if (!c_ok) {
        DistMat<T> back = Cpre.View(0, d.m, 0, d.n);  // keeps Cpre's alignment and buffer
        Copy(C, back);
    }
}

// ---------------------------------------------------------------------------------------------
// Host-operand Gemm: C-block-major stationary-C with the whole k range per block.
//   h2d stream : A (once), then per column block of C its slices of B and C
//   comm stream: gather op(A) -> [MC,*] once, each B block -> [*,MR]   (peer arena / NCCL)
//   compute    : one full-k local GEMM per block (beta applied in the epilogue)
//   d2h stream : the finished C block
// Four block slots keep PCIe (both directions), NVLink and the tensor pipe busy at the same time.
// ---------------------------------------------------------------------------------------------
namespace {
template <typename T>
void copy_rect(T* dst, int64_t ldd, const T* src, int64_t lds, int64_t rows, int64_t cols, cudaMemcpyKind kind, cudaStream_t s) {
    if (rows <= 0 || cols <= 0) return;
    cuda_check(cudaMemcpy2DAsync(dst, size_t(ldd) * sizeof(T), src, size_t(lds) * sizeof(T), size_t(rows) * sizeof(T),
                                 size_t(cols), kind, s), "cudaMemcpy2DAsync");
}
// smallest k range worth a GEMM of its own in the host-operand Gemm (A ranges of block 0, B ranges of the split form)
int64_t host_min_krange() {
    static const int64_t v = [] { const char* e = std::getenv("ELB200_HOST_KRANGE"); const int64_t x = e ? std::atoll(e) : 0; return x > 0 ? x : int64_t(512); }();
    return v;
}
int64_t lcm64(int64_t a, int64_t b) { int64_t x = a, y = b; while (y) { int64_t t = x % y; x = y; y = t; } return a / x * b; }
}  // namespace

template <typename T>
void GemmHost(Grid& g, int ta, int tb, int64_t m, int64_t n, int64_t k, T alpha, const T* Ah, int64_t lda,
              const T* Bh, int64_t ldb, T beta, T* Ch, int64_t ldc) {
    ta = ta ? 1 : 0; tb = tb ? 1 : 0;
    if (m < 0 || n < 0 || k < 0) throw EngineError(ELB200_ERR_ARG, "GemmHost: negative dimension");
    if (m == 0 || n == 0) return;
    if (!Ch || (k > 0 && alpha != T(0) && (!Ah || !Bh))) throw EngineError(ELB200_ERR_ARG, "GemmHost: null operand");
    const GridShape gs = g.shape;
    const int64_t r = gs.r, c = gs.c;
    auto loc = [](int64_t n_, int64_t rank, int64_t stride) { return Length(n_, Shift(rank, 0, stride), stride); };
    const int64_t hA = ta ? k : m, wA = ta ? m : k, hB = tb ? n : k, wB = tb ? k : n;
    const int64_t lhA = loc(hA, g.me.mc, r), lwA = loc(wA, g.me.mr, c);
    const int64_t lhB = loc(hB, g.me.mc, r), lwB = loc(wB, g.me.mr, c);
    const int64_t lhC = loc(m, g.me.mc, r);
    if (lda < std::max<int64_t>(lhA, 1) || ldb < std::max<int64_t>(lhB, 1) || ldc < std::max<int64_t>(lhC, 1))
        throw EngineError(ELB200_ERR_ARG, "GemmHost: leading dimension smaller than the local height");
    const bool product = k > 0 && alpha != T(0);
    cudaStream_t s = g.stream, cs = g.comm_stream, hs = g.h2d_stream, ds = g.d2h_stream;

    // global column blocks of C: multiples of lcm(r,c)*256 so every block view keeps alignment 0
    const int64_t unit = lcm64(r, c) * 256;
    int64_t Wg = (n / 8 + unit - 1) / unit * unit;
    if (Wg < unit) Wg = unit;
    // Blocks of Wg columns, and a TAPERED tail: the last stretch is cut in halves down to one unit (…, Wg/2, Wg/4,
    // unit, unit). What is left to do after the last byte has gone up — the product and the download of the last
    // block — is pure latency on top of the PCIe-bound pipeline; with uniform blocks that was 1/8 of the product
    // plus 1/8 of the download (≈ 7 ms of 66 at n = 16384 fp32), with the taper it is one unit's worth (< 1 ms).
    constexpr bool taper = true;  // measured (B200, n = 16384 / 32768 fp32): 64.4 / 379.5 ms with the taper, 66.7 / 385.9 ms without
    std::vector<int64_t> starts;  // block j = columns [starts[j], starts[j+1])
    {
        int64_t at = 0;
        while (n - at > Wg) { starts.push_back(at); at += Wg; }
        while (taper && n - at > unit) {
            const int64_t w = std::max(unit, ((n - at) / 2 + unit - 1) / unit * unit);
            starts.push_back(at);
            at += w;
        }
        if (at < n) starts.push_back(at);
        starts.push_back(n);
    }
    const int64_t nblocks = int64_t(starts.size()) - 1;

    // Block slots. With two, the upload of block j had to wait for the download of block j-2, i.e. for a
    // whole GEMM + D2H, and the H2D stream — the bottleneck: 3 of the 4 bytes moved go up — idled between
    // blocks (timeline in DESIGN.md 3.6); four slots keep it busy back to back.
    constexpr int kSlots = 4;
    DistMat<T> A(g, MC, MR, s);
    DistMat<T> PA(g, ta ? STAR : MC, ta ? MC : STAR, cs);
    std::vector<DistMat<T>> Bblk, Cblk, Cin, PB;
    for (int i = 0; i < kSlots; ++i) {
        Bblk.emplace_back(g, MC, MR, s);
        Cblk.emplace_back(g, MC, MR, s);
        Cin.emplace_back(g, MC, MR, s);
        PB.emplace_back(g, tb ? MR : STAR, tb ? STAR : MR, cs);
    }
    // ELB200_HOST_BRANGES=R (default 4): blocks 1.. upload their B slice
    // in R k ranges and multiply each range as it lands, accumulating into the block with beta = 0 on the first;
    // the C slice goes up AFTER them into a landing block and is folded in by one axpy (C_j = acc + beta * C_j).
    // The product then trails the upload by one range instead of one block.
    constexpr int kMaxBRanges = 8;
    // default 4 since round 2: 64.4 -> 62.4 ms at n = 16384, 379.5 -> 370.8 ms at n = 32768 (fp32, one B200); 1 restores whole-slice uploads
    static const int b_ranges_env = [] { const char* e = std::getenv("ELB200_HOST_BRANGES"); return e ? std::atoi(e) : 4; }();
    const int b_ranges = product ? int(std::min<int64_t>(std::min(kMaxBRanges, std::max(1, b_ranges_env)), std::max<int64_t>(1, k / host_min_krange()))) : 1;
    const int64_t Kb = (k + b_ranges - 1) / b_ranges;
    Ev b_up[kSlots][kMaxBRanges], b_ready[kSlots][kMaxBRanges];
    const int64_t w0 = std::min(Wg, n);
    if (product) A.Resize(hA, wA);
    for (int i = 0; i < kSlots && i < nblocks; ++i) {
        if (product) Bblk[i].Resize(tb ? w0 : k, tb ? k : w0);
        Cblk[i].Resize(m, w0);
        if (b_ranges > 1 && beta != T(0)) Cin[i].Resize(m, w0);
    }
    Ev alloc, up[kSlots], ready[kSlots], done[kSlots], out[kSlots];
    alloc.record(s);
    alloc.block(hs);
    alloc.block(cs);
    alloc.block(ds);

    // op(A) as a panel aligned with C: [MC,*] (or [*,MC]); on a 1 x 1 grid A itself. A is NOT uploaded up front:
    // block 0's B and C slices go first and A follows in k ranges, each multiplied into block 0 as soon as it has
    // landed (and been gathered) — the first GEMM starts after ~1/8 of A instead of after all of it.
    DistSpec like;  // C's distribution: [MC,MR], alignment 0
    like.col = MC; like.row = MR; like.height = m; like.width = n;
    const bool single = g.size() == 1;
    if (product && !single) {
        PA.col_constrained = PA.row_constrained = false;
        PA.spec.col_align = PA.spec.row_align = 0;
        PA.AlignWith(like, true);
        PA.Resize(hA, wA);
    }
    const DistMat<T>& afull = single ? A : PA;
    constexpr int kMaxRanges = 8;
    const int64_t min_range = host_min_krange();
    const int nranges = product ? int(std::min<int64_t>(kMaxRanges, std::max<int64_t>(1, k / min_range))) : 0;
    const int64_t Kc = nranges ? (k + nranges - 1) / nranges : 0;
    Ev a_up[kMaxRanges], a_ready[kMaxRanges];

    for (int64_t j = 0; j < nblocks; ++j) {
        const int sl = int(j % kSlots);
        const int64_t J0 = starts[size_t(j)], J1 = starts[size_t(j) + 1], wj = J1 - J0;
        if (j >= kSlots) out[sl].block(hs);  // the slot's previous block has been multiplied and downloaded
        // local index ranges of the block: C columns and (B not transposed) B columns run over MR,
        // (B transposed) B rows run over MC
        const int64_t cc0 = loc(J0, g.me.mr, c), ccn = loc(J1, g.me.mr, c) - cc0;
        Cblk[sl].Resize(m, wj);
        if (b_ranges > 1 && j > 0) {
            const int64_t br0 = loc(J0, g.me.mc, r), brn = loc(J1, g.me.mc, r) - br0;  // B transposed: block rows over MC
            Bblk[sl].Resize(tb ? wj : k, tb ? k : wj);
            if (!single) {
                PB[sl].col_constrained = PB[sl].row_constrained = false;
                PB[sl].spec.col_align = PB[sl].spec.row_align = 0;
                PB[sl].AlignWith(Cblk[sl].spec, true);
                PB[sl].Resize(tb ? wj : k, tb ? k : wj);
            }
            const DistMat<T>& bfull = single ? Bblk[sl] : PB[sl];
            for (int q = 0; q < b_ranges; ++q) {
                const int64_t k0 = std::min(k, q * Kb), k1 = std::min(k, k0 + Kb);
                if (k1 <= k0) break;
                if (!tb) {  // rows k0:k1 of the slice: k runs over MC
                    const int64_t r0 = loc(k0, g.me.mc, r), rn = loc(k1, g.me.mc, r) - r0;
                    copy_rect(Bblk[sl].buf + r0, Bblk[sl].ldim, Bh + cc0 * ldb + r0, ldb, rn, ccn, cudaMemcpyHostToDevice, hs);
                } else {    // columns k0:k1 of the slice: k runs over MR
                    const int64_t c0 = loc(k0, g.me.mr, c), cn = loc(k1, g.me.mr, c) - c0;
                    copy_rect(Bblk[sl].buf + c0 * Bblk[sl].ldim, Bblk[sl].ldim, Bh + br0 + c0 * ldb, ldb, brn, cn, cudaMemcpyHostToDevice, hs);
                }
                b_up[sl][q].record(hs);
                b_up[sl][q].block(cs);
                if (!single) {
                    const DistMat<T> Bsrc = tb ? Bblk[sl].View(0, wj, k0, k1) : Bblk[sl].View(k0, k1, 0, wj);
                    DistMat<T> Bdst = tb ? PB[sl].View(0, wj, k0, k1) : PB[sl].View(k0, k1, 0, wj);
                    Copy(Bsrc, Bdst);
                }
                b_ready[sl][q].record(cs);
                b_ready[sl][q].block(s);
                const DistMat<T> av = ta ? afull.View(k0, k1, 0, m) : afull.View(0, m, k0, k1);
                const DistMat<T> bv = tb ? bfull.View(0, wj, k0, k1) : bfull.View(k0, k1, 0, wj);
                LocalGemm(ta, tb, alpha, av, bv, q == 0 ? T(0) : T(1), Cblk[sl]);
            }
            if (beta != T(0)) {
                Cin[sl].Resize(m, wj);
                copy_rect(Cin[sl].buf, Cin[sl].ldim, Ch + cc0 * ldc, ldc, lhC, ccn, cudaMemcpyHostToDevice, hs);
                up[sl].record(hs);
                up[sl].block(s);
                Axpy(beta, Cin[sl], Cblk[sl]);
            }
            done[sl].record(s);
            done[sl].block(ds);
            copy_rect(Ch + cc0 * ldc, ldc, Cblk[sl].buf, Cblk[sl].ldim, lhC, ccn, cudaMemcpyDeviceToHost, ds);
            out[sl].record(ds);
            continue;
        }
        // beta == 0: C's old contents are never read (they may hold NaNs) — nothing to upload
        if (beta != T(0) || !product) copy_rect(Cblk[sl].buf, Cblk[sl].ldim, Ch + cc0 * ldc, ldc, lhC, ccn, cudaMemcpyHostToDevice, hs);
        if (product) {
            if (!tb) {
                Bblk[sl].Resize(k, wj);
                copy_rect(Bblk[sl].buf, Bblk[sl].ldim, Bh + cc0 * ldb, ldb, lhB, ccn, cudaMemcpyHostToDevice, hs);
            } else {
                const int64_t br0 = loc(J0, g.me.mc, r), brn = loc(J1, g.me.mc, r) - br0;
                Bblk[sl].Resize(wj, k);
                copy_rect(Bblk[sl].buf, Bblk[sl].ldim, Bh + br0, ldb, brn, lwB, cudaMemcpyHostToDevice, hs);
            }
        }
        up[sl].record(hs);
        if (product) {
            up[sl].block(cs);
            const DistMat<T>& b = panel_as(Bblk[sl], PB[sl], Cblk[sl].spec);
            ready[sl].record(cs);
            ready[sl].block(s);
            if (j == 0) {
                for (int q = 0; q < nranges; ++q) {
                    const int64_t k0 = std::min(k, q * Kc), k1 = std::min(k, k0 + Kc);
                    if (k1 <= k0) break;
                    // this rank's part of A(:, k0:k1) (or A(k0:k1, :) when transposed): k runs over MR / MC
                    if (!ta) {
                        const int64_t c0 = loc(k0, g.me.mr, c), cn = loc(k1, g.me.mr, c) - c0;
                        copy_rect(A.buf + c0 * A.ldim, A.ldim, Ah + c0 * lda, lda, lhA, cn, cudaMemcpyHostToDevice, hs);
                    } else {
                        const int64_t r0 = loc(k0, g.me.mc, r), rn = loc(k1, g.me.mc, r) - r0;
                        copy_rect(A.buf + r0, A.ldim, Ah + r0, lda, rn, lwA, cudaMemcpyHostToDevice, hs);
                    }
                    a_up[q].record(hs);
                    a_up[q].block(cs);
                    if (!single) {
                        const DistMat<T> Asrc = ta ? A.View(k0, k1, 0, m) : A.View(0, m, k0, k1);
                        DistMat<T> Adst = ta ? PA.View(k0, k1, 0, m) : PA.View(0, m, k0, k1);
                        Copy(Asrc, Adst);
                    }
                    a_ready[q].record(cs);
                    a_ready[q].block(s);
                    const DistMat<T> av = ta ? afull.View(k0, k1, 0, m) : afull.View(0, m, k0, k1);
                    const DistMat<T> bv = tb ? b.View(0, wj, k0, k1) : b.View(k0, k1, 0, wj);
                    LocalGemm(ta, tb, alpha, av, bv, q == 0 ? beta : T(1), Cblk[sl]);
                }
            } else {
                LocalGemm(ta, tb, alpha, afull, b, beta, Cblk[sl]);
            }
        } else {
            up[sl].block(s);
            Scale(beta, Cblk[sl]);
        }
        done[sl].record(s);
        done[sl].block(ds);
        copy_rect(Ch + cc0 * ldc, ldc, Cblk[sl].buf, Cblk[sl].ldim, lhC, ccn, cudaMemcpyDeviceToHost, ds);
        out[sl].record(ds);
    }
    // join: the temporaries are freed in the order of their own streams, after their last readers
    for (int i = 0; i < kSlots && i < nblocks; ++i) { out[i].block(s); done[i].block(cs); }
    cuda_check(cudaStreamSynchronize(ds), "GemmHost: d2h");
    cuda_check(cudaStreamSynchronize(s), "GemmHost: compute");
    cuda_check(cudaStreamSynchronize(cs), "GemmHost: comm");
}

template <typename T>
const DistMat<T>& PanelAs(const DistMat<T>& src, DistMat<T>& tmp, const DistSpec& like) { return panel_as(src, tmp, like); }
template const DistMat<float>& PanelAs<float>(const DistMat<float>&, DistMat<float>&, const DistSpec&);
template const DistMat<double>& PanelAs<double>(const DistMat<double>&, DistMat<double>&, const DistSpec&);
template const DistMat<Half>& PanelAs<Half>(const DistMat<Half>&, DistMat<Half>&, const DistSpec&);

template void LocalGemm<float>(int, int, float, const DistMat<float>&, const DistMat<float>&, float, DistMat<float>&);
template void LocalGemm<double>(int, int, double, const DistMat<double>&, const DistMat<double>&, double, DistMat<double>&);
template void Gemm<float>(int, int, float, const DistMat<float>&, const DistMat<float>&, float, DistMat<float>&, int);
template void Gemm<double>(int, int, double, const DistMat<double>&, const DistMat<double>&, double, DistMat<double>&, int);
template void LocalGemm<Half>(int, int, Half, const DistMat<Half>&, const DistMat<Half>&, Half, DistMat<Half>&);
template void Gemm<Half>(int, int, Half, const DistMat<Half>&, const DistMat<Half>&, Half, DistMat<Half>&, int);
template void GemmHost<float>(Grid&, int, int, int64_t, int64_t, int64_t, float, const float*, int64_t, const float*, int64_t, float, float*, int64_t);
template void GemmHost<double>(Grid&, int, int, int64_t, int64_t, int64_t, double, const double*, int64_t, const double*, int64_t, double, double*, int64_t);

}  // namespace elb
