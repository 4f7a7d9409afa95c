// el_level3_check — the triangular rank-k family of the El.hpp facade written the way a Hydrogen
// caller writes it (include/El/blas_like/level3.hpp:92-126,239-252,292-313,499-536): Syrk / Herk /
// Trrk / LocalTrrk / Symm / Hemm / Trsm / LocalTrsm and the trapezoid level-1 routines, float and double, each checked
// against El::Gemm on the same operands (Syrk = triangle of A A^T, Symm = Gemm with the symmetrised
// A, ...).  Compiled by `make`; run on a GPU box (tests/test_gpu_level3.py) and, linked against the
// hostsim engine, on CPU (tests/test_hostsim.py).
#include <El.hpp>

#include <cstdio>

using namespace El;

template <typename T>
static int exercise(const Grid& g) {
    int bad = 0;
    const double tol = sizeof(T) == 4 ? 2e-4 : 1e-11;
    const Int n = 150, k = 70, m2 = 60;
    SetLocalTrrkBlocksize<T>(32);  // several recursion levels at this size
    if (LocalTrrkBlocksize<T>() != 32) ++bad;
    DistMatrix<T> A(g), C(g), Full(g), Ref(g), Diff(g);
    Uniform(A, n, k, T(0), T(1), 11);
    Uniform(C, n, n, T(0), T(1), 12);
    for (UpperOrLower uplo : {LOWER, UPPER}) {
        // Syrk(uplo, NORMAL): triangle of 2 A A^T + 3 C; the other triangle keeps C
        DistMatrix<T> S(C);
        Syrk(uplo, NORMAL, T(2), A, T(3), S);
        Ref = C;
        Gemm(NORMAL, TRANSPOSE, T(2), A, A, T(3), Ref);
        Full = Ref;
        MakeTrapezoidal(uplo, Full);                       // wanted triangle of the Gemm result ...
        DistMatrix<T> Keep(C);
        MakeTrapezoidal(uplo == LOWER ? UPPER : LOWER, Keep, uplo == LOWER ? 1 : -1);  // ... + strict other triangle of C
        Axpy(T(1), Keep, Full);
        Diff = S;
        Axpy(T(-1), Full, Diff);
        if (FrobeniusNorm(Diff) > tol * FrobeniusNorm(Full)) ++bad;
        // Herk(TRANSPOSE) with resize == triangle of A^T A
        DistMatrix<T> H(g);
        Herk(uplo, TRANSPOSE, T(1), A, H);
        if (H.Height() != k || H.Width() != k) ++bad;
        Gemm(TRANSPOSE, NORMAL, T(1), A, A, Ref);
        MakeTrapezoidal(uplo, Ref);
        Diff = H;
        Axpy(T(-1), Ref, Diff);
        if (FrobeniusNorm(Diff) > tol * FrobeniusNorm(Ref)) ++bad;
        // Trrk(uplo, TRANSPOSE, NORMAL) on k x n operands and the matching LocalTrrk overload on panels
        DistMatrix<T> At(g), Bk(g);
        Transpose(A, At);                                  // k x n
        Uniform(Bk, k, n, T(0), T(1), 13);
        DistMatrix<T> R1(C), R2(C);
        Trrk(uplo, TRANSPOSE, NORMAL, T(0.5), At, Bk, T(-0.5), R1);
        DistMatrix<T, STAR, MC> At_STAR_MC(g);
        DistMatrix<T, STAR, MR> Bk_STAR_MR(g);
        At_STAR_MC.AlignWith(R2);
        Bk_STAR_MR.AlignWith(R2);
        At_STAR_MC = At;
        Bk_STAR_MR = Bk;
        LocalTrrk(uplo, TRANSPOSE, T(0.5), At_STAR_MC, Bk_STAR_MR, T(-0.5), R2);
        Diff = R1;
        Axpy(T(-1), R2, Diff);
        if (FrobeniusNorm(Diff) > tol * FrobeniusNorm(R1)) ++bad;
        // Symm / Hemm: A symmetric given by one triangle
        DistMatrix<T> Sy(g), B(g), Cs(g), Cg(g);
        Uniform(Sy, n, n, T(0), T(1), 14);
        Uniform(B, n, m2, T(0), T(1), 15);
        Uniform(Cs, n, m2, T(0), T(1), 16);
        Cg = Cs;
        DistMatrix<T> SyFull(Sy);
        MakeSymmetric(uplo, SyFull);
        Symm(LEFT, uplo, T(2), Sy, B, T(3), Cs);
        Gemm(NORMAL, NORMAL, T(2), SyFull, B, T(3), Cg);
        Axpy(T(-1), Cg, Cs);
        if (FrobeniusNorm(Cs) > tol * FrobeniusNorm(Cg)) ++bad;
        DistMatrix<T> Br(g), Cr(g), Cgr(g);
        Uniform(Br, m2, n, T(0), T(1), 17);
        Uniform(Cr, m2, n, T(0), T(1), 18);
        Cgr = Cr;
        Hemm(RIGHT, uplo, T(2), Sy, Br, T(3), Cr);
        Gemm(NORMAL, NORMAL, T(2), Br, SyFull, T(3), Cgr);
        Axpy(T(-1), Cgr, Cr);
        if (FrobeniusNorm(Cr) > tol * FrobeniusNorm(Cgr)) ++bad;
        // ScaleTrapezoid(2) then AxpyTrapezoid(-1 x original) leaves the original inside the trapezoid too
        DistMatrix<T> Z(C);
        ScaleTrapezoid(T(2), uplo, Z, 1);
        AxpyTrapezoid(uplo, T(-1), C, Z, 1);
        Axpy(T(-1), C, Z);
        if (FrobeniusNorm(Z) != 0) ++bad;
    }
    // Trsm: B := alpha op(T)^-1 B and back again with Gemm on the explicit triangle: residual check
    {
        DistMatrix<T> Tm(g), Bx(g), B0(g), R(g);
        Uniform(Tm, n, n, T(0), T(1), 21);
        DistMatrix<T> Shift(g);
        Shift.Resize(n, n);
        Zero(Shift);
        for (Int j = 0; j < Shift.LocalWidth(); ++j)
            for (Int i = 0; i < Shift.LocalHeight(); ++i)
                if (Shift.GlobalRow(i) == Shift.GlobalCol(j)) Shift.SetLocal(i, j, T(n));
        Axpy(T(1), Shift, Tm);                              // well conditioned: n on the diagonal
        for (LeftOrRight side : {LEFT, RIGHT})
            for (UpperOrLower uplo : {LOWER, UPPER})
                for (Orientation o : {NORMAL, TRANSPOSE}) {
                    const Int bh = side == LEFT ? n : m2, bw = side == LEFT ? m2 : n;
                    Uniform(B0, bh, bw, T(0), T(1), 22);
                    Bx = B0;
                    Trsm(side, uplo, o, NON_UNIT, T(2), Tm, Bx);
                    DistMatrix<T> Tri(Tm);
                    MakeTrapezoidal(uplo, Tri);
                    R = B0;                                 // R = op(Tri) X - 2 B0  (or X op(Tri) - 2 B0)
                    if (side == LEFT) Gemm(o, NORMAL, T(1), Tri, Bx, T(-2), R);
                    else Gemm(NORMAL, o, T(1), Bx, Tri, T(-2), R);
                    if (FrobeniusNorm(R) > 50 * tol * FrobeniusNorm(B0)) ++bad;
                }
        // LocalTrsm on a replicated triangle and [*,VR] right-hand sides equals Trsm
        DistMatrix<T, STAR, STAR> Ts(g);
        Ts = Tm;
        DistMatrix<T, STAR, VR> Xv(g);
        Uniform(B0, n, m2, T(0), T(1), 23);
        Xv = B0;
        LocalTrsm(LEFT, LOWER, NORMAL, UNIT, T(1), Ts, Xv);
        Bx = B0;
        Trsm(LEFT, LOWER, NORMAL, UNIT, T(1), Tm, Bx, false, TRSM_LARGE);
        DistMatrix<T> Xm(g);
        Xm = Xv;
        Axpy(T(-1), Bx, Xm);
        if (!(FrobeniusNorm(Xm) <= 50 * tol * FrobeniusNorm(Bx))) ++bad;
    }
    // collective single-entry Get / Set / Update, and the resizing LocalGemm overload
    {
        DistMatrix<T> G(C);
        const T before = G.Get(7, 5);
        G.Set(7, 5, T(3));
        G.Update(7, 5, T(0.5));
        if (G.Get(7, 5) != T(3.5) || C.Get(7, 5) != before) ++bad;
        DistMatrix<T, MC, STAR> Am(g);
        DistMatrix<T, STAR, MR> Bm(g);
        Am = A;                                             // n x k
        DistMatrix<T> At(g);
        Transpose(A, At);                                   // k x n
        Bm = At;
        DistMatrix<T> P(g), Q(g);
        LocalGemm(NORMAL, NORMAL, T(1), Am, Bm, P);         // resized to n x n: A A^T
        Gemm(NORMAL, TRANSPOSE, T(1), A, A, Q);
        if (P.Height() != n || P.Width() != n) ++bad;
        Axpy(T(-1), Q, P);
        if (FrobeniusNorm(P) > tol * FrobeniusNorm(Q)) ++bad;
    }
    // errors: non-square C, NORMAL where a transposed panel is required
    try {
        DistMatrix<T> Bad(g);
        Bad.Resize(n, n + 1);
        Syrk(LOWER, NORMAL, T(1), A, T(0), Bad);
        ++bad;
    } catch (const std::logic_error&) {}
    g.Synchronize();
    SetLocalTrrkBlocksize<T>(2048);
    return bad;
}

int main(int argc, char** argv) {
    Environment env(argc, argv);
    Grid g;  // 1 x 1 on the current device
    int bad = exercise<float>(g) + exercise<double>(g);
    std::printf(bad ? "EL_LEVEL3_CHECK_FAIL %d\n" : "EL_LEVEL3_CHECK_PASS\n", bad);
    return bad ? 1 : 0;
}
