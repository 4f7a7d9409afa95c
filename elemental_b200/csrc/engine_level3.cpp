// Triangular rank-k family on top of the Gemm path's building blocks:
// ScaleTrapezoid / AxpyTrapezoid / MakeTrapezoidal / MakeSymmetric, LocalTrrk, Trrk, Syrk (= Herk for
// real types) and Symm (= Hemm for real types).  Reference: src/blas_like/level3/{Trrk,Syrk,Herk,
// Symm,Hemm}.cpp, Trrk/Local.hpp, Trrk/{NN,NT,TN,TT}.hpp, Syrk/{LN,LT,UN,UT}.hpp and the level-1
// trapezoid headers.  Everything here is panel AllGathers (engine_redist.cpp), the local GEMM
// kernels and ONE new elementwise kernel family (trapezoid_kernel, blas1.cu).
//
// B200-first differences from the reference's schedule (same results up to summation order):
//  * k is walked in PanelWidth() steps instead of Blocksize() = 128, as in SUMMA-C (engine_gemm.cpp);
//  * panels are consumed in the majorness the gather produces, so the reference's explicit
//    A1 -> [VR,*] -> Transpose -> [*,MR] chain of syrk::LN (Syrk/LN.hpp:40-47) is one AllGather of
//    A1 into [MR,*] and a GEMM with op(B) = T;
//  * LocalTrrk stops recursing at gridWidth * LocalTrrkBlocksize global rows like the reference, but
//    the default blocksize is 2048 (SetLocalTrrkBlocksize, ELB200_TRRK_BASE) instead of 64: fewer,
//    larger GEMMs for the tensor pipe; each diagonal block is computed in full into a temporary and
//    added under the trapezoid mask like the reference's LocalTrrkKernel (Trrk/Local.hpp:350-396);
//  * Symm symmetrises A once (MakeSymmetric: O(n^2) bytes) and runs the SUMMA Gemm (O(n^3) flops)
//    instead of the reference's LocalSymmetricAccumulate schedule (Symm/LL.hpp) — one code path,
//    one set of kernels, identical mathematics.
#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <memory>

#include "engine.hpp"
#include "engine_dev.hpp"

namespace elb {

// LocalTrrkBlocksize (reference src/blas_like/blocksizes.cpp:24-26,82-87: default 64, per scalar type;
// here one value, default 2048 or ELB200_TRRK_BASE): LocalTrrk stops recursing when
// C.Height() <= gridWidth * LocalTrrkBlocksize (Trrk/Local.hpp:800).
namespace {
std::atomic<int64_t>& trrk_bs() {
    static std::atomic<int64_t> v{[] { const char* e = std::getenv("ELB200_TRRK_BASE"); const int64_t x = e ? std::atoll(e) : 0; return x > 0 ? x : int64_t(2048); }()};
    return v;
}
}  // namespace
int64_t Grid::LocalTrrkBlocksize() { return trrk_bs().load(); }
void Grid::SetLocalTrrkBlocksize(int64_t nb) { trrk_bs().store(nb > 0 ? nb : 1); }

namespace {

enum { LOWER = 0, UPPER = 1 };

// trapezoid op on the local part of Y (and X, which must have Y's distribution and alignment)
template <typename T>
void trap_local(int op, int uplo, int64_t offset, T alpha, const DistMat<T>* X, DistMat<T>& Y) {
    const int64_t h = Y.lh(), w = Y.lw();
    if (h == 0 || w == 0) return;
    Dev<T>::trapezoid(op, uplo, offset, h, w, Y.col_shift(), Y.col_stride(), Y.row_shift(), Y.row_stride(), alpha,
                      X ? X->buf : nullptr, X ? X->ldim : 1, Y.buf, Y.ldim, Y.stream);
}

void check_uplo(int uplo, const char* who) {
    if (uplo != LOWER && uplo != UPPER) throw EngineError(ELB200_ERR_ARG, std::string(who) + ": uplo must be LOWER (0) or UPPER (1)");
}

template <typename T>
bool same_layout(const DistMat<T>& X, const DistMat<T>& Y) {
    return X.spec.col == Y.spec.col && X.spec.row == Y.spec.row && X.spec.col_align == Y.spec.col_align &&
           X.spec.row_align == Y.spec.row_align && X.spec.root == Y.spec.root;
}

// triangle of C += alpha op(A) op(B); A, B, C conformal and aligned (checked by LocalTrrk)
template <typename T>
void local_trrk_rec(int uplo, int ta, int tb, T alpha, const DistMat<T>& A, const DistMat<T>& B, DistMat<T>& C, int64_t base) {
    const int64_t n = C.spec.height;
    if (n == 0) return;
    const int64_t k = ta ? A.spec.height : A.spec.width;
    if (n <= base) {
        // diagonal block: full product into a temporary laid out like C, added under the mask
        // (reference LocalTrrkKernel: DTL / DBR + LocalAxpyTrapezoid, Trrk/Local.hpp:383-394)
        DistMat<T> D(*C.grid, MC, MR, C.stream);
        D.Align(C.spec.col_align, C.spec.row_align);
        D.Resize(n, n);
        LocalGemm(ta, tb, alpha, A, B, T(0), D);
        trap_local<T>(ELB200_TRAP_AXPY, uplo, 0, T(1), &D, C);
        return;
    }
    const int64_t half = n / 2;
    auto rows_of = [&](int64_t i0, int64_t i1) { return ta ? A.View(0, k, i0, i1) : A.View(i0, i1, 0, k); };
    auto cols_of = [&](int64_t j0, int64_t j1) { return tb ? B.View(j0, j1, 0, k) : B.View(0, k, j0, j1); };
    const DistMat<T> AT = rows_of(0, half), AB = rows_of(half, n);
    const DistMat<T> BL = cols_of(0, half), BR = cols_of(half, n);
    if (uplo == LOWER) {
        DistMat<T> CBL = C.View(half, n, 0, half);
        LocalGemm(ta, tb, alpha, AB, BL, T(1), CBL);
    } else {
        DistMat<T> CTR = C.View(0, half, half, n);
        LocalGemm(ta, tb, alpha, AT, BR, T(1), CTR);
    }
    DistMat<T> CTL = C.View(0, half, 0, half), CBR = C.View(half, n, half, n);
    local_trrk_rec(uplo, ta, tb, alpha, AT, BL, CTL, base);
    local_trrk_rec(uplo, ta, tb, alpha, AB, BR, CBR, base);
}

}  // namespace

template <typename T>
void ScaleTrapezoid(T alpha, int uplo, DistMat<T>& A, int64_t offset) {
    check_uplo(uplo, "ScaleTrapezoid");
    if (alpha == T(1)) return;  // ScaleTrapezoid.hpp:18-19
    trap_local<T>(ELB200_TRAP_SCALE, uplo, offset, alpha, nullptr, A);
}

template <typename T>
void MakeTrapezoidal(int uplo, DistMat<T>& A, int64_t offset) {
    check_uplo(uplo, "MakeTrapezoidal");
    trap_local<T>(ELB200_TRAP_MAKE, uplo, offset, T(0), nullptr, A);
}

template <typename T>
void AxpyTrapezoid(int uplo, T alpha, const DistMat<T>& X, DistMat<T>& Y, int64_t offset) {
    check_uplo(uplo, "AxpyTrapezoid");
    if (X.grid != Y.grid) throw EngineError(ELB200_ERR_ARG, "AxpyTrapezoid: matrices live on different grids");
    if (X.spec.height != Y.spec.height || X.spec.width != Y.spec.width)
        throw EngineError(ELB200_ERR_ARG, "AxpyTrapezoid: nonconformal dimensions");
    if (same_layout(X, Y)) {
        trap_local<T>(ELB200_TRAP_AXPY, uplo, offset, alpha, &X, Y);
        return;
    }
    // different distribution or alignment: bring X into Y's layout first (AxpyTrapezoid.hpp, the
    // ElementalMatrix overload copies X into a matrix constructed like Y)
    DistMat<T> X2(*Y.grid, Y.spec.col, Y.spec.row, Y.stream);
    X2.spec.root = Y.spec.root;
    X2.Align(Y.spec.col_align, Y.spec.row_align);
    Copy(X, X2);
    trap_local<T>(ELB200_TRAP_AXPY, uplo, offset, alpha, &X2, Y);
}

// A = tri(A) + strict-tri(A)^T  (MakeSymmetric.hpp:56-73)
template <typename T>
void MakeSymmetric(int uplo, DistMat<T>& A) {
    check_uplo(uplo, "MakeSymmetric");
    if (A.spec.height != A.spec.width) throw EngineError(ELB200_ERR_ARG, "MakeSymmetric: cannot make non-square matrix symmetric");
    MakeTrapezoidal(uplo, A, 0);
    DistMat<T> At(*A.grid, A.spec.col, A.spec.row, A.stream);
    At.spec.root = A.spec.root;
    At.Align(A.spec.col_align, A.spec.row_align);
    TransposeCopy(A, At);
    if (uplo == LOWER) trap_local<T>(ELB200_TRAP_AXPY, UPPER, 1, T(1), &At, A);
    else trap_local<T>(ELB200_TRAP_AXPY, LOWER, -1, T(1), &At, A);
}

template <typename T>
void LocalTrrk(int uplo, int ta, int tb, T alpha, const DistMat<T>& A, const DistMat<T>& B, T beta, DistMat<T>& C) {
    check_uplo(uplo, "LocalTrrk");
    ta = ta ? 1 : 0; tb = tb ? 1 : 0;
    if (A.grid != C.grid || B.grid != C.grid) throw EngineError(ELB200_ERR_ARG, "LocalTrrk: matrices live on different grids");
    // distributions: C [MC,MR]; A [MC,*] or (transposed) [*,MC]; B [*,MR] or (transposed) [MR,*]
    const bool dist_ok = C.spec.col == MC && C.spec.row == MR &&
                         (ta ? (A.spec.col == STAR && A.spec.row == MC) : (A.spec.col == MC && A.spec.row == STAR)) &&
                         (tb ? (B.spec.col == MR && B.spec.row == STAR) : (B.spec.col == STAR && B.spec.row == MR));
    if (!dist_ok) throw EngineError(ELB200_ERR_ARG, "LocalTrrk: needs C[MC,MR], A[MC,*]/[*,MC], B[*,MR]/[MR,*]");
    const int64_t n = C.spec.height;
    const int64_t ma = ta ? A.spec.width : A.spec.height, ka = ta ? A.spec.height : A.spec.width;
    const int64_t nb = tb ? B.spec.height : B.spec.width, kb = tb ? B.spec.width : B.spec.height;
    if (C.spec.width != n || ma != n || nb != n || ka != kb) throw EngineError(ELB200_ERR_ARG, "LocalTrrk: nonconformal matrices");
    // EnsureConformal (Trrk/Local.hpp:20-48): A and B must be aligned with C
    if ((ta ? A.spec.row_align : A.spec.col_align) != C.spec.col_align || (tb ? B.spec.col_align : B.spec.row_align) != C.spec.row_align)
        throw EngineError(ELB200_ERR_ARG, "LocalTrrk: A / B not aligned with C");
    ScaleTrapezoid(beta, uplo, C, 0);
    if (n == 0 || ka == 0 || alpha == T(0)) return;
    local_trrk_rec(uplo, ta, tb, alpha, A, B, C, int64_t(C.grid->shape.c) * Grid::LocalTrrkBlocksize());
}

template <typename T>
void Trrk(int uplo, int ta, int tb, T alpha, const DistMat<T>& Apre, const DistMat<T>& Bpre, T beta, DistMat<T>& Cpre) {
    check_uplo(uplo, "Trrk");
    ta = ta ? 1 : 0; tb = tb ? 1 : 0;
    Grid& g = *Cpre.grid;
    if (Apre.grid != &g || Bpre.grid != &g) throw EngineError(ELB200_ERR_ARG, "Trrk: matrices live on different grids");
    const int64_t n = Cpre.spec.height;
    const int64_t ma = ta ? Apre.spec.width : Apre.spec.height, k = ta ? Apre.spec.height : Apre.spec.width;
    const int64_t nb = tb ? Bpre.spec.height : Bpre.spec.width, kb = tb ? Bpre.spec.width : Bpre.spec.height;
    if (Cpre.spec.width != n || ma != n || nb != n || k != kb) throw EngineError(ELB200_ERR_ARG, "Trrk: nonconformal matrices");
    // read / read-write proxies to [MC,MR] (Trrk/NN.hpp:27-31)
    const bool a_ok = Apre.spec.col == MC && Apre.spec.row == MR;
    const bool b_ok = Bpre.spec.col == MC && Bpre.spec.row == MR;
    const bool c_ok = Cpre.spec.col == MC && Cpre.spec.row == MR;
    DistMat<T> Ap(g, MC, MR, Cpre.stream), Bp(g, MC, MR, Cpre.stream), Cp(g, MC, MR, Cpre.stream);
    const bool product = n > 0 && k > 0 && alpha != T(0);
    if (!a_ok && product) Copy(Apre, Ap);
    if (!b_ok && product && &Bpre != &Apre) Copy(Bpre, Bp);
    if (!c_ok) Copy(Cpre, Cp);
    const DistMat<T>& A = a_ok ? Apre : Ap;
    const DistMat<T>& B = b_ok ? Bpre : (&Bpre == &Apre ? Ap : Bp);
    DistMat<T>& C = c_ok ? Cpre : Cp;

    ScaleTrapezoid(beta, uplo, C, 0);  // Trrk.cpp:107
    if (product) {
        // one pass over the local triangle of C per panel: panels as wide as SUMMA-C's
        const int64_t W = g.size() == 1 ? k : Grid::PanelWidth(sizeof(T));
        // With the copy-engine peer arena the gathers of panel p+1 run on the comm stream under the local
        // products of panel p (two panel slots, ready / freed events), as in SUMMA-C (engine_gemm.cpp); without
        // it (one rank, NCCL only) everything is on C's stream.
        const bool pipe = g.size() > 1 && g.comm_stream && g.arena;
        cudaStream_t ps = pipe ? g.comm_stream : C.stream;
        DistMat<T> PA[2] = {DistMat<T>(g, ta ? STAR : MC, ta ? MC : STAR, ps), DistMat<T>(g, ta ? STAR : MC, ta ? MC : STAR, ps)};
        DistMat<T> PB[2] = {DistMat<T>(g, tb ? MR : STAR, tb ? STAR : MR, ps), DistMat<T>(g, tb ? MR : STAR, tb ? STAR : MR, ps)};
        std::unique_ptr<Ev[]> ev(pipe ? new Ev[5] : nullptr);  // ready[2], freed[2], start
        Ev* ready = ev.get();
        Ev* freed = pipe ? ev.get() + 2 : nullptr;
        if (pipe) {
            ev[4].record(C.stream);  // A, B, C are ready in C.stream order
            ev[4].block(ps);
        }
        int64_t p = 0;
        for (int64_t k0 = 0; k0 < k; k0 += W, ++p) {
            const int sl = pipe ? int(p & 1) : 0;
            const int64_t kw = std::min(W, k - k0);
            if (pipe && p >= 2) freed[sl].block(ps);
            const DistMat<T> A1 = ta ? A.View(k0, k0 + kw, 0, n) : A.View(0, n, k0, k0 + kw);
            const DistMat<T> B1 = tb ? B.View(0, n, k0, k0 + kw) : B.View(k0, k0 + kw, 0, n);
            const DistMat<T>& a = PanelAs(A1, PA[sl], C.spec);
            const DistMat<T>& b = PanelAs(B1, PB[sl], C.spec);
            if (pipe) {
                ready[sl].record(ps);
                ready[sl].block(C.stream);
            }
            // a / b may be A1 / B1 themselves when the gather is the identity on this grid (PanelAs), so
            // the recursion is entered directly instead of through LocalTrrk's distribution checks
            local_trrk_rec(uplo, ta, tb, alpha, a, b, C, int64_t(g.shape.c) * Grid::LocalTrrkBlocksize());
            if (pipe) freed[sl].record(C.stream);
        }
        // the panels are freed in comm-stream order: not before the last products have read them
        for (int i = 0; pipe && i < 2 && i < p; ++i) freed[i].block(ps);
    }
    if (!c_ok) {
        DistMat<T> back = Cpre.View(0, n, 0, n);
        Copy(C, back);
    }
}

// triangle of C = alpha A A^T + beta C (trans = 0) or alpha A^T A + beta C (trans != 0)
template <typename T>
void Syrk(int uplo, int trans, T alpha, const DistMat<T>& A, T beta, DistMat<T>& C) {
    check_uplo(uplo, "Syrk");
    const int64_t n = trans ? A.spec.width : A.spec.height;
    if (C.spec.height != n || C.spec.width != n) throw EngineError(ELB200_ERR_ARG, "Syrk: nonconformal matrices");  // Syrk.cpp:108-117
    Trrk(uplo, trans ? 1 : 0, trans ? 0 : 1, alpha, A, A, beta, C);
}

// C = alpha A B + beta C (side LEFT) or alpha B A + beta C (RIGHT), A symmetric, `uplo` triangle stored
template <typename T>
void Symm(int side, int uplo, T alpha, const DistMat<T>& A, const DistMat<T>& B, T beta, DistMat<T>& C) {
    check_uplo(uplo, "Symm");
    if (side != 0 && side != 1) throw EngineError(ELB200_ERR_ARG, "Symm: side must be LEFT (0) or RIGHT (1)");
    Grid& g = *C.grid;
    if (A.grid != &g || B.grid != &g) throw EngineError(ELB200_ERR_ARG, "Symm: matrices live on different grids");
    const int64_t n = side == 0 ? C.spec.height : C.spec.width;
    if (A.spec.height != n || A.spec.width != n || B.spec.height != C.spec.height || B.spec.width != C.spec.width)
        throw EngineError(ELB200_ERR_ARG, "Symm: nonconformal matrices");
    if (n == 0 || alpha == T(0)) { Scale(beta, C); return; }
    DistMat<T> Afull(g, MC, MR, C.stream);
    Copy(A, Afull);
    MakeSymmetric(uplo, Afull);
    if (side == 0) Gemm(0, 0, alpha, Afull, B, beta, C, GEMM_DEFAULT);
    else Gemm(0, 0, alpha, B, Afull, beta, C, GEMM_DEFAULT);
}

// ---- Trsm -------------------------------------------------------------------------------------------
// LocalTrsm (Trsm.cpp:501-526): A replicated ([*,*]), X distributed over the dimension that is NOT solved
// ([*,V] for LEFT, [U,*] for RIGHT): every rank solves its own right-hand sides, no communication.
template <typename T>
void LocalTrsm(int side, int uplo, int trans, int diag, T alpha, const DistMat<T>& A, DistMat<T>& X) {
    check_uplo(uplo, "LocalTrsm");
    if ((side != 0 && side != 1) || (diag != 0 && diag != 1)) throw EngineError(ELB200_ERR_ARG, "LocalTrsm: bad side / diag");
    if (A.grid != X.grid) throw EngineError(ELB200_ERR_ARG, "LocalTrsm: matrices live on different grids");
    if (A.spec.col != STAR || A.spec.row != STAR || A.spec.height != A.spec.width)
        throw EngineError(ELB200_ERR_ARG, "LocalTrsm: A must be a square [*,*] matrix");
    if ((side == 0 && X.spec.col != STAR) || (side == 1 && X.spec.row != STAR))
        throw EngineError(ELB200_ERR_ARG, "LocalTrsm: Dist of RHS must conform with that of triangle");  // Trsm.cpp:513-517
    if (A.spec.height != (side == 0 ? X.spec.height : X.spec.width)) throw EngineError(ELB200_ERR_ARG, "LocalTrsm: nonconformal matrices");
    Dev<T>::trsm(side, uplo, trans ? 1 : 0, diag, X.lh(), X.lw(), alpha, A.buf, A.ldim, X.buf, X.ldim, X.stream);
}

// Trsm (Trsm.cpp:129-440): the many-right-hand-sides schedule of trsm::{LLN,LLT,LUN,LUT}Large and
// trsm::{RLN,RLT,RUN,RUT} for every side / uplo / orientation, written once: walk the diagonal blocks in
// the direction of the substitution; per block  A11 -> [*,*],  X1 -> [*,VR] (LEFT) / [VC,*] (RIGHT),
// LocalTrsm,  X1 -> [*,MR] / [MC,*] and back into X,  then ONE LocalGemm updates all remaining rows
// (columns) of X with the panel of op(A) next to the block (LLN.hpp:37-68).  Blocks are 8 algorithmic
// blocksizes wide (1024 by default) instead of one: the update GEMM gets a k deep enough for the tensor
// pipe and X2 is re-read 8 x less often; the wider diagonal solve is itself recursive GEMMs (trsm.cu).
template <typename T>
void Trsm(int side, int uplo, int trans, int diag, T alpha, const DistMat<T>& Apre, DistMat<T>& Bpre) {
    check_uplo(uplo, "Trsm");
    if ((side != 0 && side != 1) || (diag != 0 && diag != 1)) throw EngineError(ELB200_ERR_ARG, "Trsm: bad side / diag");
    trans = trans ? 1 : 0;
    Grid& g = *Bpre.grid;
    if (Apre.grid != &g) throw EngineError(ELB200_ERR_ARG, "Trsm: matrices live on different grids");
    const int64_t na = Apre.spec.height;
    if (Apre.spec.width != na) throw EngineError(ELB200_ERR_ARG, "Trsm: A must be square");             // Trsm.cpp:146-147
    if (na != (side == 0 ? Bpre.spec.height : Bpre.spec.width)) throw EngineError(ELB200_ERR_ARG, "Trsm: nonconformal Trsm");
    Scale(alpha, Bpre);                                                                                   // Trsm.cpp:158
    if (Bpre.spec.height == 0 || Bpre.spec.width == 0 || alpha == T(0)) return;

    const bool a_ok = Apre.spec.col == MC && Apre.spec.row == MR, b_ok = Bpre.spec.col == MC && Bpre.spec.row == MR;
    DistMat<T> Ap(g, MC, MR, Bpre.stream), Bp(g, MC, MR, Bpre.stream);
    if (!a_ok) Copy(Apre, Ap);
    if (!b_ok) Copy(Bpre, Bp);
    const DistMat<T>& A = a_ok ? Apre : Ap;
    DistMat<T>& X = b_ok ? Bpre : Bp;
    const int64_t m = X.spec.height, n = X.spec.width;

    const int64_t W = 8 * int64_t(Grid::Blocksize());
    const bool eff_lower = (uplo == LOWER) != (trans != 0);       // op(A) is lower triangular
    const bool forward = side == 0 ? eff_lower : !eff_lower;      // LEFT: lower -> top down; RIGHT: upper -> left to right
    cudaStream_t s = X.stream;
    DistMat<T> A11s(g, STAR, STAR, s);
    DistMat<T> X1v(g, side == 0 ? STAR : VC, side == 0 ? VR : STAR, s);   // every rank solves its share of the right-hand sides
    DistMat<T> X1p(g, side == 0 ? STAR : MC, side == 0 ? MR : STAR, s);   // the solved block as a GEMM panel
    DistMat<T> PA(g, side == 0 ? (trans ? STAR : MC) : (trans ? MR : STAR), side == 0 ? (trans ? MC : STAR) : (trans ? STAR : MR), s);
    const int64_t nblocks = (na + W - 1) / W;
    for (int64_t b = 0; b < nblocks; ++b) {
        const int64_t blk = forward ? b : nblocks - 1 - b;
        const int64_t k0 = blk * W, k1 = std::min(na, k0 + W);
        const int64_t r0 = forward ? k1 : 0, r1 = forward ? na : k0;     // what is still to be updated
        const DistMat<T> A11 = A.View(k0, k1, k0, k1);
        A11s.col_constrained = A11s.row_constrained = false;
        Copy(A11, A11s);
        DistMat<T> X1 = side == 0 ? X.View(k0, k1, 0, n) : X.View(0, m, k0, k1);
        X1v.col_constrained = X1v.row_constrained = false;
        Copy(X1, X1v);
        LocalTrsm(side, uplo, trans, diag, T(1), A11s, X1v);
        X1p.col_constrained = X1p.row_constrained = false;
        X1p.spec.col_align = X1p.spec.row_align = 0;
        X1p.AlignWith(X.spec, true);
        Copy(X1v, X1p);
        Copy(X1p, X1);
        if (r1 <= r0) continue;
        if (side == 0) {
            DistMat<T> X2 = X.View(r0, r1, 0, n);
            // op(A)(rest, block): A(rest, block) as [MC,*], or (transposed) A(block, rest) as [*,MC]
            const DistMat<T> Ap1 = trans ? A.View(k0, k1, r0, r1) : A.View(r0, r1, k0, k1);
            const DistMat<T>& a = PanelAs(Ap1, PA, X2.spec);
            LocalGemm(trans, 0, T(-1), a, X1p, T(1), X2);
        } else {
            DistMat<T> X2 = X.View(0, m, r0, r1);
            // op(A)(block, rest): A(block, rest) as [*,MR], or (transposed) A(rest, block) as [MR,*]
            const DistMat<T> Ap1 = trans ? A.View(r0, r1, k0, k1) : A.View(k0, k1, r0, r1);
            const DistMat<T>& a = PanelAs(Ap1, PA, X2.spec);
            LocalGemm(0, trans, T(-1), X1p, a, T(1), X2);
        }
    }
    if (!b_ok) {
        DistMat<T> back = Bpre.View(0, m, 0, n);
        Copy(X, back);
    }
}

#define ELB_INST3(T)                                                                                          \
    template void ScaleTrapezoid<T>(T, int, DistMat<T>&, int64_t);                                            \
    template void AxpyTrapezoid<T>(int, T, const DistMat<T>&, DistMat<T>&, int64_t);                          \
    template void MakeTrapezoidal<T>(int, DistMat<T>&, int64_t);                                              \
    template void MakeSymmetric<T>(int, DistMat<T>&);                                                         \
    template void LocalTrrk<T>(int, int, int, T, const DistMat<T>&, const DistMat<T>&, T, DistMat<T>&);       \
    template void Trrk<T>(int, int, int, T, const DistMat<T>&, const DistMat<T>&, T, DistMat<T>&);            \
    template void Syrk<T>(int, int, T, const DistMat<T>&, T, DistMat<T>&);                                    \
    template void Symm<T>(int, int, T, const DistMat<T>&, const DistMat<T>&, T, DistMat<T>&);                 \
    template void LocalTrsm<T>(int, int, int, int, T, const DistMat<T>&, DistMat<T>&);                         \
    template void Trsm<T>(int, int, int, int, T, const DistMat<T>&, DistMat<T>&);
ELB_INST3(float)
ELB_INST3(double)
#undef ELB_INST3

}  // namespace elb
