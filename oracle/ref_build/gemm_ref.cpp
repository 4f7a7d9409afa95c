// ORACLE BUILD ONLY (test infrastructure): a driver over the REFERENCE'S OWN CPU El::Gemm,
// compiled from /root/reference where it lies, single rank (oracle/ref_build/stub/mpi.h),
// linked against the OpenBLAS found in the image. Nothing here is on the product path.
//
//   gemm_ref --bench f32|f64 n         time El::Gemm(NORMAL,NORMAL,0.5,A,B,-0.5,C,GEMM_SUMMA_C), nb=128
//   gemm_ref --dist-golden <dir> <r>   multi-rank (ELB_STUB_NP processes): per-rank local matrices after all
//                                      121 redistributions and 20 El::Gemm runs on an r x (NP/r) grid
//   gemm_ref --level3-golden <dir> <r> multi-rank: per-rank outputs of the trapezoid routines, LocalTrrk, Trrk, Syrk, Symm
//   gemm_ref --golden <dir>            write inputs + reference outputs of a fixed list of small
//                                      cases (all orientations x GemmAlgorithm, ragged sizes) and
//                                      tables of the index functions, as raw little-endian files
#include <El.hpp>

#include <chrono>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <string>
#include <vector>

using namespace El;

template <typename T>
static void fill_lcg(Matrix<T>& M, unsigned seed) {
    // deterministic, implementation-independent input generator (also restated in tests)
    unsigned long long s = 0x9E3779B97F4A7C15ull * (seed + 1);
    for (Int j = 0; j < M.Width(); ++j)
        for (Int i = 0; i < M.Height(); ++i) {
            s = s * 6364136223846793005ull + 1442695040888963407ull;
            const double u = double((s >> 11) & ((1ull << 53) - 1)) / double(1ull << 53);
            M.Set(i, j, T(2.0 * u - 1.0));
        }
}

template <typename T>
static void dump(const std::string& path, const Matrix<T>& M) {
    std::ofstream f(path, std::ios::binary);
    for (Int j = 0; j < M.Width(); ++j)
        f.write(reinterpret_cast<const char*>(M.LockedBuffer(0, j)), sizeof(T) * M.Height());
}

template <typename T>
static double bench(Int n) {
    const Grid& g = Grid::Default();
    DistMatrix<T> A(g), B(g), C(g);
    Uniform(A, n, n, T(0), T(0.1));
    Uniform(B, n, n, T(0), T(0.1));
    Uniform(C, n, n, T(0), T(0.1));
    SetBlocksize(128);
    {  // warm the BLAS threads
        DistMatrix<T> a(g), b(g), c(g);
        Uniform(a, 512, 512); Uniform(b, 512, 512); Uniform(c, 512, 512);
        Gemm(NORMAL, NORMAL, T(0.5), a, b, T(-0.5), c, GEMM_SUMMA_C);
    }
    auto t0 = std::chrono::steady_clock::now();
    Gemm(NORMAL, NORMAL, T(0.5), A, B, T(-0.5), C, GEMM_SUMMA_C);
    auto t1 = std::chrono::steady_clock::now();
    return std::chrono::duration<double>(t1 - t0).count();
}

template <typename T>
static void golden_cases(const std::string& dir, const char* tname, std::ofstream& index) {
    const Grid& g = Grid::Default();
    struct Case { Int m, n, k; Int nb; double alpha, beta; };
    const Case cases[] = {{37, 29, 53, 16, 2.0, 3.0}, {64, 64, 64, 128, 0.5, -0.5}, {5, 130, 41, 32, 1.0, 0.0}};
    const Orientation os[2] = {NORMAL, TRANSPOSE};
    const GemmAlgorithm algs[] = {GEMM_DEFAULT, GEMM_SUMMA_A, GEMM_SUMMA_B, GEMM_SUMMA_C, GEMM_SUMMA_DOT};
    int id = 0;
    for (const Case& c : cases)
        for (int ia = 0; ia < 2; ++ia)
            for (int ib = 0; ib < 2; ++ib)
                for (GemmAlgorithm alg : algs) {
                    DistMatrix<T> A(g), B(g), C(g);
                    A.Resize(ia ? c.k : c.m, ia ? c.m : c.k);
                    B.Resize(ib ? c.n : c.k, ib ? c.k : c.n);
                    C.Resize(c.m, c.n);
                    fill_lcg(A.Matrix(), 3 * id + 0);
                    fill_lcg(B.Matrix(), 3 * id + 1);
                    fill_lcg(C.Matrix(), 3 * id + 2);
                    const std::string base = dir + "/" + tname + "_" + std::to_string(id);
                    dump(base + "_A.bin", A.LockedMatrix());
                    dump(base + "_B.bin", B.LockedMatrix());
                    dump(base + "_C0.bin", C.LockedMatrix());
                    SetBlocksize(c.nb);
                    Gemm(os[ia], os[ib], T(c.alpha), A, B, T(c.beta), C, alg);
                    dump(base + "_C.bin", C.LockedMatrix());
                    index << tname << " " << id << " " << ia << " " << ib << " " << int(alg) << " " << c.m << " "
                          << c.n << " " << c.k << " " << c.nb << " " << c.alpha << " " << c.beta << "\n";
                    ++id;
                }
}

// Complex instantiations (src/blas_like/level3/Gemm.cpp:504-506 via El/macros/Instantiate.h): the reference's own
// El::Gemm<Complex<float|double>> on one rank, orientations NORMAL / TRANSPOSE / ADJOINT, complex alpha and beta.
template <typename R>
static void fill_lcg_complex(Matrix<Complex<R>>& M, unsigned seed) {
    Matrix<R> re(M.Height(), M.Width()), im(M.Height(), M.Width());
    fill_lcg(re, seed);
    fill_lcg(im, seed + 1000);
    for (Int j = 0; j < M.Width(); ++j)
        for (Int i = 0; i < M.Height(); ++i) M.Set(i, j, Complex<R>(re.Get(i, j), im.Get(i, j)));
}
template <typename R>
static void golden_cases_complex(const std::string& dir, const char* tname, std::ofstream& index) {
    using T = Complex<R>;
    const Grid& g = Grid::Default();
    struct Case { Int m, n, k; Int nb; double ar, ai, br, bi; bool all; };
    const Case cases[] = {{37, 29, 53, 16, 2.0, -0.5, 3.0, 0.25, true}, {5, 130, 41, 32, 1.0, 0.0, 0.0, 0.0, false},
                          {64, 48, 40, 128, 0.5, 0.0, -0.5, 0.0, false}};
    const Orientation os[3] = {NORMAL, TRANSPOSE, ADJOINT};
    const GemmAlgorithm algs[] = {GEMM_DEFAULT, GEMM_SUMMA_A, GEMM_SUMMA_B, GEMM_SUMMA_C, GEMM_SUMMA_DOT};
    int id = 0;
    for (const Case& c : cases)
        for (int ia = 0; ia < 3; ++ia)
            for (int ib = 0; ib < 3; ++ib)
                for (GemmAlgorithm alg : algs) {
                    if (!c.all && !((ia == 0 || ia == 2) && (ib == 0 || ib == 2) && (alg == GEMM_DEFAULT || alg == GEMM_SUMMA_C))) continue;
                    DistMatrix<T> A(g), B(g), C(g);
                    A.Resize(ia ? c.k : c.m, ia ? c.m : c.k);
                    B.Resize(ib ? c.n : c.k, ib ? c.k : c.n);
                    C.Resize(c.m, c.n);
                    fill_lcg_complex(A.Matrix(), 3 * id + 0);
                    fill_lcg_complex(B.Matrix(), 3 * id + 1);
                    fill_lcg_complex(C.Matrix(), 3 * id + 2);
                    const std::string base = dir + "/" + tname + "_" + std::to_string(id);
                    if (id == 0) {
                        dump(base + "_A.bin", A.LockedMatrix());
                        dump(base + "_B.bin", B.LockedMatrix());
                        dump(base + "_C0.bin", C.LockedMatrix());
                    }
                    SetBlocksize(c.nb);
                    Gemm(os[ia], os[ib], T(R(c.ar), R(c.ai)), A, B, T(R(c.br), R(c.bi)), C, alg);
                    dump(base + "_C.bin", C.LockedMatrix());
                    index << tname << " " << id << " " << ia << " " << ib << " " << int(alg) << " " << c.m << " " << c.n << " "
                          << c.k << " " << c.nb << " " << c.ar << " " << c.ai << " " << c.br << " " << c.bi << "\n";
                    ++id;
                }
}

// ---------------------------------------------------------------------------------------------
// Multi-rank golden run (ELB_STUB_NP processes over stub/mpi_stub.cpp): what the REFERENCE puts
// on every rank after each (U,V) -> (U2,V2) redistribution and after El::Gemm on an r x c grid.
// Every rank appends to <dir>/rank<q>.txt (records) and <dir>/rank<q>.bin (column-major doubles).
// ---------------------------------------------------------------------------------------------
struct DistOut {
    std::ofstream txt, bin;
    long offset = 0;
    void blob(const Matrix<double>& M) {
        for (Int j = 0; j < M.Width(); ++j)
            bin.write(reinterpret_cast<const char*>(M.LockedBuffer(0, j)), sizeof(double) * M.Height());
        offset += long(M.Height()) * M.Width();
    }
};

static double entry(Int i, Int j, Int w) { return 1.0 + double(i) * w + j; }

template <Dist U, Dist V, Dist U2, Dist V2>
static void redist_case(const Grid& g, DistOut& out) {
    const Int h = 13, w = 11;
    constexpr bool circ = U == CIRC || U2 == CIRC;  // Gather / Scatter cases: "Q" records with the two roots
    DistMatrix<double, U, V> A(g);
    if (U == CIRC) A.SetRoot(2 % g.Size());
    A.Align(1 % A.ColStride(), 2 % A.RowStride());
    A.Resize(h, w);
    for (Int j = 0; j < A.LocalWidth(); ++j)
        for (Int i = 0; i < A.LocalHeight(); ++i) A.SetLocal(i, j, entry(A.GlobalRow(i), A.GlobalCol(j), w));
    DistMatrix<double, U2, V2> B(g);
    if (U2 == CIRC) B.SetRoot(1 % g.Size());
    B.Align(3 % B.ColStride(), 1 % B.RowStride());
    B = A;
    out.txt << (circ ? "Q " : "R ") << int(U) << " " << int(V) << " " << int(U2) << " " << int(V2) << " " << A.ColAlign() << " "
            << A.RowAlign() << " " << B.ColAlign() << " " << B.RowAlign() << " " << B.ColShift() << " " << B.RowShift()
            << " " << B.LocalHeight() << " " << B.LocalWidth() << " " << out.offset;
    if (circ) out.txt << " " << A.Root() << " " << B.Root();
    out.txt << "\n";
    out.blob(B.LockedMatrix());
}
template <Dist U, Dist V>
static void redist_from(const Grid& g, DistOut& out) {
    redist_case<U, V, MC, MR>(g, out);   redist_case<U, V, MR, MC>(g, out);
    redist_case<U, V, MC, STAR>(g, out); redist_case<U, V, STAR, MC>(g, out);
    redist_case<U, V, MR, STAR>(g, out); redist_case<U, V, STAR, MR>(g, out);
    redist_case<U, V, VC, STAR>(g, out); redist_case<U, V, STAR, VC>(g, out);
    redist_case<U, V, VR, STAR>(g, out); redist_case<U, V, STAR, VR>(g, out);
    redist_case<U, V, STAR, STAR>(g, out);
}

static void fill_global_lcg(DistMatrix<double>& A, Int h, Int w, unsigned seed) {
    Matrix<double> G(h, w);
    fill_lcg(G, seed);  // every rank generates the same global matrix and keeps its own entries
    A.Resize(h, w);
    for (Int j = 0; j < A.LocalWidth(); ++j)
        for (Int i = 0; i < A.LocalHeight(); ++i) A.SetLocal(i, j, G.Get(A.GlobalRow(i), A.GlobalCol(j)));
}

static void dist_golden(const std::string& dir, int gridHeight) {
    const int rank = mpi::Rank(mpi::COMM_WORLD);
    const Grid g(mpi::NewWorldComm(), gridHeight);
    DistOut out;
    out.txt.open(dir + "/rank" + std::to_string(rank) + ".txt");
    out.bin.open(dir + "/rank" + std::to_string(rank) + ".bin", std::ios::binary);
    out.txt << "I " << rank << " " << g.Height() << " " << g.Width() << " " << g.Row() << " " << g.Col() << " "
            << g.VCRank() << " " << g.VRRank() << "\n";
    redist_from<MC, MR>(g, out);   redist_from<MR, MC>(g, out);
    redist_from<MC, STAR>(g, out); redist_from<STAR, MC>(g, out);
    redist_from<MR, STAR>(g, out); redist_from<STAR, MR>(g, out);
    redist_from<VC, STAR>(g, out); redist_from<STAR, VC>(g, out);
    redist_from<VR, STAR>(g, out); redist_from<STAR, VR>(g, out);
    redist_from<STAR, STAR>(g, out);
    // [CIRC,CIRC] (copy::Gather / copy::Scatter): every distribution to the root, the root to every
    // distribution, and root to root
    redist_case<MC, MR, CIRC, CIRC>(g, out);   redist_case<MR, MC, CIRC, CIRC>(g, out);
    redist_case<MC, STAR, CIRC, CIRC>(g, out); redist_case<STAR, MC, CIRC, CIRC>(g, out);
    redist_case<MR, STAR, CIRC, CIRC>(g, out); redist_case<STAR, MR, CIRC, CIRC>(g, out);
    redist_case<VC, STAR, CIRC, CIRC>(g, out); redist_case<STAR, VC, CIRC, CIRC>(g, out);
    redist_case<VR, STAR, CIRC, CIRC>(g, out); redist_case<STAR, VR, CIRC, CIRC>(g, out);
    redist_case<STAR, STAR, CIRC, CIRC>(g, out);
    redist_from<CIRC, CIRC>(g, out);
    redist_case<CIRC, CIRC, CIRC, CIRC>(g, out);

    const Int m = 37, n = 29, k = 53, nb = 16;
    const Orientation os[2] = {NORMAL, TRANSPOSE};
    const GemmAlgorithm algs[] = {GEMM_DEFAULT, GEMM_SUMMA_A, GEMM_SUMMA_B, GEMM_SUMMA_C, GEMM_SUMMA_DOT};
    int id = 0;
    for (int ia = 0; ia < 2; ++ia)
        for (int ib = 0; ib < 2; ++ib)
            for (GemmAlgorithm alg : algs) {
                DistMatrix<double> A(g), B(g), C(g);
                A.Align(1 % g.Height(), 2 % g.Width());
                B.Align(0, 1 % g.Width());
                C.Align(1 % g.Height(), 0);
                fill_global_lcg(A, ia ? k : m, ia ? m : k, 3 * id + 0);
                fill_global_lcg(B, ib ? n : k, ib ? k : n, 3 * id + 1);
                fill_global_lcg(C, m, n, 3 * id + 2);
                SetBlocksize(nb);
                Gemm(os[ia], os[ib], 2.0, A, B, 3.0, C, alg);
                out.txt << "G " << id << " " << ia << " " << ib << " " << int(alg) << " " << m << " " << n << " " << k << " "
                        << nb << " " << C.ColAlign() << " " << C.RowAlign() << " " << C.ColShift() << " " << C.RowShift()
                        << " " << C.LocalHeight() << " " << C.LocalWidth() << " " << out.offset << "\n";
                out.blob(C.LockedMatrix());
                ++id;
            }
}

// ---------------------------------------------------------------------------------------------
// Triangular rank-k family, multi-rank like dist_golden: per-rank local outputs of
// the reference's ScaleTrapezoid / MakeTrapezoidal / AxpyTrapezoid / MakeSymmetric, the four
// LocalTrrk overloads, Trrk, Syrk and Symm.  Records "L kind uplo p1 p2 colAlign rowAlign colShift
// rowShift lh lw offset" in <dir>/l3rank<q>.txt, blobs in <dir>/l3rank<q>.bin.
// ---------------------------------------------------------------------------------------------
template <typename DM>
static void fill_dm_lcg(DM& A, Int h, Int w, unsigned seed) {
    Matrix<double> G(h, w);
    fill_lcg(G, seed);
    A.Resize(h, w);
    for (Int j = 0; j < A.LocalWidth(); ++j)
        for (Int i = 0; i < A.LocalHeight(); ++i) A.SetLocal(i, j, G.Get(A.GlobalRow(i), A.GlobalCol(j)));
}

template <typename DM>
static void l3_record(DistOut& out, int kind, int uplo, int p1, int p2, const DM& C) {
    out.txt << "L " << kind << " " << uplo << " " << p1 << " " << p2 << " " << C.ColAlign() << " " << C.RowAlign() << " "
            << C.ColShift() << " " << C.RowShift() << " " << C.LocalHeight() << " " << C.LocalWidth() << " " << out.offset << "\n";
    out.blob(C.LockedMatrix());
}

static void level3_golden(const std::string& dir, int gridHeight) {
    const int rank = mpi::Rank(mpi::COMM_WORLD);
    const Grid g(mpi::NewWorldComm(), gridHeight);
    DistOut out;
    out.txt.open(dir + "/l3rank" + std::to_string(rank) + ".txt");
    out.bin.open(dir + "/l3rank" + std::to_string(rank) + ".bin", std::ios::binary);
    const UpperOrLower uplos[2] = {LOWER, UPPER};
    const Orientation os[2] = {NORMAL, TRANSPOSE};
    const Int ca = 1 % g.Height(), ra = 2 % g.Width();
    // kinds 0-2: trapezoid level-1 routines on a 13 x 11 [MC,MR] matrix, offsets 0, 1, -2
    const int offsets[3] = {0, 1, -2};
    for (int u = 0; u < 2; ++u)
        for (int oi = 0; oi < 3; ++oi) {
            DistMatrix<double> A(g), X(g);
            A.Align(ca, ra); X.Align(ca, ra);
            fill_dm_lcg(A, 13, 11, 7001);
            ScaleTrapezoid(-0.5, uplos[u], A, offsets[oi]);
            l3_record(out, 0, u, offsets[oi], 0, A);
            fill_dm_lcg(A, 13, 11, 7001);
            MakeTrapezoidal(uplos[u], A, offsets[oi]);
            l3_record(out, 1, u, offsets[oi], 0, A);
        }
    for (int u = 0; u < 2; ++u)
        for (int oi = 0; oi < 3; ++oi) {
            DistMatrix<double> A(g), X(g);
            A.Align(ca, ra); X.Align(ca, ra);
            fill_dm_lcg(A, 13, 13, 7001);
            fill_dm_lcg(X, 13, 13, 7004);
            // UPPER with a negative offset is skipped (+2 instead of -2): the reference's LocalAxpyTrapezoid does not
            // clamp j+1-offset to the height (AxpyTrapezoid.hpp:78, unlike ScaleTrapezoid.hpp:74-75) and runs past
            // the end of the local columns there — an upstream defect, not behaviour to reproduce
            const int off = (u == 1 && offsets[oi] < 0) ? -offsets[oi] : offsets[oi];
            AxpyTrapezoid(uplos[u], 2.0, X, A, off);
            l3_record(out, 2, u, off, 0, A);
        }
    // kind 3: MakeSymmetric
    for (int u = 0; u < 2; ++u) {
        DistMatrix<double> A(g);
        A.Align(ca, ra);
        fill_dm_lcg(A, 13, 13, 7001);
        MakeSymmetric(uplos[u], A);
        l3_record(out, 3, u, 0, 0, A);
    }
    // kind 4: the four LocalTrrk overloads (Trrk/Local.hpp:688-890), alpha 2, beta 3, n 37, k 23
    const Int n = 37, k = 23;
    SetBlocksize(16);
    SetLocalTrrkBlocksize<double>(4);
    for (int u = 0; u < 2; ++u) {
        {
            DistMatrix<double> C(g); C.Align(ca, ra); fill_dm_lcg(C, n, n, 7003);
            DistMatrix<double, MC, STAR> A(g); A.AlignWith(C); fill_dm_lcg(A, n, k, 7001);
            DistMatrix<double, STAR, MR> B(g); B.AlignWith(C); fill_dm_lcg(B, k, n, 7002);
            LocalTrrk(uplos[u], 2.0, A, B, 3.0, C);
            l3_record(out, 4, u, 0, 0, C);
        }
        {
            DistMatrix<double> C(g); C.Align(ca, ra); fill_dm_lcg(C, n, n, 7003);
            DistMatrix<double, MC, STAR> A(g); A.AlignWith(C); fill_dm_lcg(A, n, k, 7001);
            DistMatrix<double, MR, STAR> B(g); B.AlignWith(C); fill_dm_lcg(B, n, k, 7002);
            LocalTrrk(uplos[u], TRANSPOSE, 2.0, A, B, 3.0, C);
            l3_record(out, 4, u, 0, 1, C);
        }
        {
            DistMatrix<double> C(g); C.Align(ca, ra); fill_dm_lcg(C, n, n, 7003);
            DistMatrix<double, STAR, MC> A(g); A.AlignWith(C); fill_dm_lcg(A, k, n, 7001);
            DistMatrix<double, STAR, MR> B(g); B.AlignWith(C); fill_dm_lcg(B, k, n, 7002);
            LocalTrrk(uplos[u], TRANSPOSE, 2.0, A, B, 3.0, C);
            l3_record(out, 4, u, 1, 0, C);
        }
        {
            DistMatrix<double> C(g); C.Align(ca, ra); fill_dm_lcg(C, n, n, 7003);
            DistMatrix<double, STAR, MC> A(g); A.AlignWith(C); fill_dm_lcg(A, k, n, 7001);
            DistMatrix<double, MR, STAR> B(g); B.AlignWith(C); fill_dm_lcg(B, n, k, 7002);
            LocalTrrk(uplos[u], TRANSPOSE, TRANSPOSE, 2.0, A, B, 3.0, C);
            l3_record(out, 4, u, 1, 1, C);
        }
    }
    // kind 5: Trrk, all orientations; kind 6: Syrk; kind 7: Symm.  [MC,MR] operands with alignments.
    for (int u = 0; u < 2; ++u)
        for (int ia = 0; ia < 2; ++ia)
            for (int ib = 0; ib < 2; ++ib) {
                DistMatrix<double> A(g), B(g), C(g);
                A.Align(ca, 0); B.Align(0, ra); C.Align(ca, ra);
                fill_dm_lcg(A, ia ? k : n, ia ? n : k, 7001);
                fill_dm_lcg(B, ib ? n : k, ib ? k : n, 7002);
                fill_dm_lcg(C, n, n, 7003);
                Trrk(uplos[u], os[ia], os[ib], 2.0, A, B, 3.0, C);
                l3_record(out, 5, u, ia, ib, C);
            }
    for (int u = 0; u < 2; ++u)
        for (int it = 0; it < 2; ++it) {
            DistMatrix<double> A(g), C(g);
            A.Align(ca, 0); C.Align(0, ra);
            fill_dm_lcg(A, it ? k : n, it ? n : k, 7001);
            fill_dm_lcg(C, n, n, 7003);
            Syrk(uplos[u], os[it], 2.0, A, 3.0, C);
            l3_record(out, 6, u, it, 0, C);
        }
    const Int m2 = 29;
    const LeftOrRight sides[2] = {LEFT, RIGHT};
    for (int sd = 0; sd < 2; ++sd)
        for (int u = 0; u < 2; ++u) {
            DistMatrix<double> A(g), B(g), C(g);
            A.Align(ca, ra); C.Align(0, ra);
            fill_dm_lcg(A, n, n, 7001);
            fill_dm_lcg(B, sd ? m2 : n, sd ? n : m2, 7002);
            fill_dm_lcg(C, sd ? m2 : n, sd ? n : m2, 7003);
            Symm(sides[sd], uplos[u], 2.0, A, B, 3.0, C);
            l3_record(out, 7, u, sd, 0, C);
        }
    // kind 8: Trsm, every side / uplo / orientation / diag (p1 = side, p2 = 2*orientation + diag); A = LCG + n on the
    // diagonal (well conditioned; for UNIT the strict triangle is scaled by 1/n instead), blocksize 16
    for (int sd = 0; sd < 2; ++sd)
        for (int u = 0; u < 2; ++u)
            for (int it = 0; it < 2; ++it)
                for (int dg = 0; dg < 2; ++dg) {
                    DistMatrix<double> A(g), B(g);
                    A.Align(ca, 0); B.Align(0, ra);
                    fill_dm_lcg(A, n, n, 7001);
                    if (dg) Scale(1.0 / double(n), A);
                    else ShiftDiagonal(A, double(n));
                    fill_dm_lcg(B, sd ? m2 : n, sd ? n : m2, 7002);
                    Trsm(sides[sd], uplos[u], os[it], dg ? UNIT : NON_UNIT, 2.0, A, B);
                    l3_record(out, 8, u, sd, 2 * it + dg, B);
                }
}

int main(int argc, char* argv[]) {
    Environment env(argc, argv);
    if (argc >= 4 && !std::strcmp(argv[1], "--dist-golden")) {
        dist_golden(argv[2], std::atoi(argv[3]));
        return 0;
    }
    if (argc >= 4 && !std::strcmp(argv[1], "--level3-golden")) {
        level3_golden(argv[2], std::atoi(argv[3]));
        return 0;
    }
    if (argc >= 4 && !std::strcmp(argv[1], "--bench")) {
        const Int n = std::atoi(argv[3]);
        const double t = !std::strcmp(argv[2], "f32") ? bench<float>(n) : bench<double>(n);
        std::printf("REF_GEMM_SECONDS %.6f n=%d dtype=%s\n", t, int(n), argv[2]);
        return 0;
    }
    if (argc >= 3 && !std::strcmp(argv[1], "--golden-complex")) {
        const std::string dir = argv[2];
        std::ofstream index(dir + "/index_complex.txt");
        golden_cases_complex<float>(dir, "c32", index);
        golden_cases_complex<double>(dir, "c64", index);
        return 0;
    }
    if (argc >= 3 && !std::strcmp(argv[1], "--golden")) {
        const std::string dir = argv[2];
        std::ofstream index(dir + "/index.txt");
        golden_cases<float>(dir, "f32", index);
        golden_cases<double>(dir, "f64", index);
        // index-function tables (include/El/core/indexing/impl.hpp) and Grid::DefaultHeight
        std::ofstream tab(dir + "/indexing.txt");
        for (Int stride = 1; stride <= 8; ++stride)
            for (Int align = 0; align < stride; ++align)
                for (Int rank = 0; rank < stride; ++rank)
                    for (Int n : {0, 1, 7, 64, 1000}) {
                        const Int shift = Shift(rank, align, stride);
                        tab << "L " << n << " " << rank << " " << align << " " << stride << " " << shift << " "
                            << Length(n, shift, stride) << " " << MaxLength(n, stride) << "\n";
                    }
        for (int p = 1; p <= 64; ++p) tab << "H " << p << " " << Grid::DefaultHeight(p) << "\n";
        return 0;
    }
    std::printf("usage: gemm_ref --bench f32|f64 n | --golden dir | --golden-complex dir | --dist-golden dir gridHeight (under ELB_STUB_NP)\n");
    return 64;
}
