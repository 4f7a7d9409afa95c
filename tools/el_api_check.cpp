// el_api_check — compile-and-link check of the El.hpp facade against libelb200: a translation unit
// written the way a Hydrogen / LBANN caller writes it (SURVEY.md 8(b) "Source API"), touching every
// name of the facade for float and double.  `make tools/el_api_check` building it is the test on a
// machine without a GPU; run on a GPU box it executes a small end-to-end pass and prints PASS.
#include <El.hpp>

#include <cstdio>
#include <vector>

using namespace El;

template <typename T>
static int exercise(const Grid& g) {
    int bad = 0;
    const Int m = 96, n = 80, k = 112;
    DistMatrix<T, MC, MR, ELEMENT, Device::GPU> A(g), B(g), C(g);
    Uniform(A, m, k, T(0), T(1), 1);
    Uniform(B, k, n, T(0), T(1), 2);
    Uniform(C, m, n, T(0), T(1), 3);
    DistMatrix<T> C0(C);                                    // copy constructor
    PushBlocksizeStack(32);
    Gemm(NORMAL, NORMAL, T(2), A, B, T(3), C, GEMM_SUMMA_C);
    PopBlocksizeStack();
    // the same product out of redistributed panels + LocalGemm (what SUMMA_C does internally)
    DistMatrix<T, MC, STAR> A_MC_STAR(g);
    DistMatrix<T, STAR, MR> B_STAR_MR(g);
    A_MC_STAR.AlignWith(C0);
    B_STAR_MR.AlignWith(C0);
    A_MC_STAR = A;
    B_STAR_MR = B;
    DistMatrix<T> D(C0);
    LocalGemm(NORMAL, NORMAL, T(2), A_MC_STAR, B_STAR_MR, T(3), D);
    Axpy(T(-1), C, D);
    if (FrobeniusNorm(D) > 1e-3 * FrobeniusNorm(C)) ++bad;
    // local matrices, views, SyncInfo
    auto Cloc = C.Matrix();
    const auto Aloc = A.LockedMatrix();
    SyncInfo<Device::GPU> si = SyncInfoFromMatrix(Cloc);
    Matrix<T> L1(Cloc.Height(), Cloc.Width()), L2;
    SetSyncInfo(L1, si);
    SetSyncInfo(L2, si);
    Copy(Cloc, L1);
    Transpose(L1, L2);
    Scale(T(0.5), L1);
    Axpy(T(1), Cloc, L1);
    Zero(L2);
    Fill(L2, T(1));
    auto V = L1(IR(0, L1.Height() / 2), ALL);
    if (V.Height() != L1.Height() / 2 || !V.Viewing() || Aloc.GetDevice() != Device::GPU) ++bad;
    Matrix<T> P;
    SetSyncInfo(P, si);
    if (Aloc.Width() > 0 && Aloc.Height() > 0) {
        Matrix<T> At;
        SetSyncInfo(At, si);
        Transpose(Aloc, At);
        Gemm(NORMAL, NORMAL, T(1), At, Aloc, P);           // local Gemm(Matrix) with resize
        if (P.Height() != Aloc.Width()) ++bad;
    }
    Synchronize(si);
    // views, index maps, single entries, contractions, transposes, rooted matrices
    auto Cv = C(IR(8, m), IR(4, n));
    if (Cv.Height() != m - 8 || !Cv.Viewing() || Cv.ColAlign() != (C.ColAlign() + 8) % C.ColStride()) ++bad;
    if (C.LocalHeight() > 0 && C.LocalWidth() > 0) {
        const T v = C.GetLocal(0, 0);
        C.SetLocal(0, 0, v);
        if (!C.IsLocal(C.GlobalRow(0), C.GlobalCol(0)) || C.LocalRow(C.GlobalRow(0)) != 0) ++bad;
    }
    DistMatrix<T, MR, MC> Ct(g);
    Transpose(C, Ct);
    DistMatrix<T, STAR, STAR> Full(g);
    Copy(C, Full);
    DistMatrix<T> Acc(g);
    Acc.Resize(m, n);
    Zero(Acc);
    AxpyContract(T(1), Full, Acc);                          // sums p replicas
    Contract(Full, Acc);
    DistMatrix<T, CIRC, CIRC> Root(g);
    Root.SetRoot(g.Size() - 1);
    Root = C;
    if (Root.Root() != g.Size() - 1 || (g.Rank() == g.Size() - 1 ? Root.LocalHeight() != m : Root.LocalHeight() != 0)) ++bad;
    DistMatrix<T, VC, STAR> Rows(g);
    Rows = Root;
    Gemm(TRANSPOSE, NORMAL, T(1), A, C0, Acc);              // resize overload
    if (Acc.Height() != k || Acc.Width() != n) ++bad;
    g.Synchronize();
    return bad;
}

// gpu_half_type: the same calls on DistMatrix<gpu_half_type> and Matrix<gpu_half_type>; the product is
// compared with the float run of the same (half-representable) inputs.
static int exercise_half(const Grid& g) {
    int bad = 0;
    const Int m = 96, n = 80, k = 112;
    DistMatrix<float> Af(g), Bf(g), Cf(g);
    Uniform(Af, m, k, 0.f, 1.f, 1);
    Uniform(Bf, k, n, 0.f, 1.f, 2);
    DistMatrix<gpu_half_type, MC, MR, ELEMENT, Device::GPU> A(g), B(g), C(g);
    A.Resize(m, k); B.Resize(k, n);
    std::vector<float> hf(size_t(Af.LocalHeight()) * Af.LocalWidth());
    std::vector<gpu_half_type> hh(hf.size());
    Af.GetLocalToHost(hf.data());
    for (size_t i = 0; i < hf.size(); ++i) { hh[i] = gpu_half_type(hf[i]); hf[i] = float(hh[i]); }
    A.SetLocalFromHost(hh.data()); Af.SetLocalFromHost(hf.data());
    hf.resize(size_t(Bf.LocalHeight()) * Bf.LocalWidth()); hh.resize(hf.size());
    Bf.GetLocalToHost(hf.data());
    for (size_t i = 0; i < hf.size(); ++i) { hh[i] = gpu_half_type(hf[i]); hf[i] = float(hh[i]); }
    B.SetLocalFromHost(hh.data()); Bf.SetLocalFromHost(hf.data());
    Gemm(NORMAL, NORMAL, gpu_half_type(2.f), A, B, C, GEMM_SUMMA_C);      // resize overload
    Gemm(NORMAL, NORMAL, 2.f, Af, Bf, Cf, GEMM_SUMMA_C);
    DistMatrix<gpu_half_type, STAR, VR> Cv(g);
    Cv = C;                                                                 // redistribution of a half matrix
    DistMatrix<gpu_half_type> C2(Cv);
    hh.resize(size_t(C2.LocalHeight()) * C2.LocalWidth()); hf.resize(hh.size());
    C2.GetLocalToHost(hh.data());
    Cf.GetLocalToHost(hf.data());
    double num = 0, den = 0;
    for (size_t i = 0; i < hf.size(); ++i) { const double d = double(float(hh[i])) - hf[i]; num += d * d; den += double(hf[i]) * hf[i]; }
    if (!(num <= 4e-3 * 4e-3 * den)) ++bad;
    auto Cl = C.Matrix();
    Matrix<gpu_half_type> L1(Cl.Height(), Cl.Width());
    SetSyncInfo(L1, SyncInfoFromMatrix(Cl));
    Copy(Cl, L1);
    Scale(gpu_half_type(0.5f), L1);
    Axpy(gpu_half_type(1.f), Cl, L1);
    Synchronize(SyncInfoFromMatrix(Cl));
    g.Synchronize();
    return bad;
}

// Complex<float> / Complex<double>: DistMatrix<Complex<R>>, El::Gemm with ADJOINT and complex scalars, Adjoint,
// redistribution, Axpy / Scale, and Gemm(Matrix<Complex<R>>), checked against a host evaluation of the definition.
template <typename R>
static int exercise_complex(const Grid& g) {
    typedef Complex<R> Cx;
    int bad = 0;
    const Int m = 45, n = 38, k = 52;
    auto fill = [](std::vector<Cx>& v, unsigned seed) {
        unsigned long long s = 0x9E3779B97F4A7C15ull * (seed + 1);
        for (Cx& z : v) {
            s = s * 6364136223846793005ull + 1442695040888963407ull; const R re = R(double((s >> 11) & ((1ull << 53) - 1)) / double(1ull << 53) - 0.5);
            s = s * 6364136223846793005ull + 1442695040888963407ull; const R im = R(double((s >> 11) & ((1ull << 53) - 1)) / double(1ull << 53) - 0.5);
            z = Cx(re, im);
        }
    };
    // 1 x 1 grid: the local matrix is the global one, column-major with ld = height
    std::vector<Cx> hA(size_t(k) * m), hB(size_t(k) * n), hC(size_t(m) * n), out(size_t(m) * n);
    fill(hA, 1); fill(hB, 2); fill(hC, 3);
    DistMatrix<Cx, MC, MR, ELEMENT, Device::GPU> A(k, m, g), B(k, n, g), C(m, n, g);   // A is k x m: used as A^H
    A.SetLocalFromHost(hA.data()); B.SetLocalFromHost(hB.data()); C.SetLocalFromHost(hC.data());
    const Cx alpha(R(2), R(-0.5)), beta(R(0.25), R(3));
    Gemm(ADJOINT, NORMAL, alpha, A, B, beta, C, GEMM_SUMMA_C);
    C.GetLocalToHost(out.data());
    double num = 0, den = 0;
    for (Int j = 0; j < n; ++j)
        for (Int i = 0; i < m; ++i) {
            std::complex<double> acc(0, 0);
            for (Int l = 0; l < k; ++l) acc += std::conj(std::complex<double>(hA[size_t(l) + size_t(i) * k])) * std::complex<double>(hB[size_t(l) + size_t(j) * k]);
            const std::complex<double> want = std::complex<double>(alpha) * acc + std::complex<double>(beta) * std::complex<double>(hC[size_t(i) + size_t(j) * m]);
            num += std::norm(std::complex<double>(out[size_t(i) + size_t(j) * m]) - want); den += std::norm(want);
        }
    const double tol = (sizeof(R) == 4 ? 1e-5 : 1e-13) * std::sqrt(2.0 * k);
    if (!(num <= tol * tol * den)) ++bad;
    // Adjoint into another distribution and back; Axpy / Scale with complex scalars
    DistMatrix<Cx, STAR, VR> Ah(g);
    DistMatrix<Cx> At(g), A2(g);
    Adjoint(A, At);                       // m x k
    Ah = At;                              // redistribution
    A2 = Ah;
    std::vector<Cx> back(size_t(m) * k);
    A2.GetLocalToHost(back.data());
    for (Int j = 0; j < k; ++j)
        for (Int i = 0; i < m; ++i)
            if (back[size_t(i) + size_t(j) * m] != std::conj(hA[size_t(j) + size_t(i) * k])) { ++bad; i = m; j = k; }
    Axpy(Cx(R(0), R(1)), At, A2);         // A2 = (1 + i) A^H
    Scale(Cx(R(0.5), R(-0.5)), A2);       // (0.5 - 0.5 i)(1 + i) = 1  ->  A2 = A^H again (up to rounding)
    A2.GetLocalToHost(back.data());
    for (Int j = 0; j < k && !bad; ++j)
        for (Int i = 0; i < m; ++i)
            if (std::abs(std::complex<double>(back[size_t(i) + size_t(j) * m]) - std::conj(std::complex<double>(hA[size_t(j) + size_t(i) * k]))) > (sizeof(R) == 4 ? 1e-6 : 1e-14)) { ++bad; break; }
    Zero(A2);
    // local interleaved matrices
    Matrix<Cx> La(k, m), Lb(k, n), Lc(m, n);   // LDim() may exceed the height (16-byte pitch)
    auto up = [&](Matrix<Cx>& M, const std::vector<Cx>& h) {
        cudaStream_t st = M.GetSyncInfo().Stream();
        return cudaMemcpy2DAsync(M.Buffer(), size_t(M.LDim()) * sizeof(Cx), h.data(), size_t(M.Height()) * sizeof(Cx), size_t(M.Height()) * sizeof(Cx),
                                 size_t(M.Width()), cudaMemcpyHostToDevice, st) == cudaSuccess && cudaStreamSynchronize(st) == cudaSuccess;
    };
    if (!up(La, hA) || !up(Lb, hB) || !up(Lc, hC)) ++bad;
    Gemm(ADJOINT, NORMAL, alpha, La, Lb, beta, Lc);
    Synchronize(SyncInfoFromMatrix(Lc));
    std::vector<Cx> out2(out.size());
    if (cudaMemcpy2DAsync(out2.data(), size_t(m) * sizeof(Cx), Lc.LockedBuffer(), size_t(Lc.LDim()) * sizeof(Cx), size_t(m) * sizeof(Cx), size_t(n),
                          cudaMemcpyDeviceToHost, Lc.GetSyncInfo().Stream()) != cudaSuccess || cudaStreamSynchronize(Lc.GetSyncInfo().Stream()) != cudaSuccess) ++bad;
    double d2 = 0;
    for (size_t i = 0; i < out.size(); ++i) d2 += std::norm(std::complex<double>(out2[i]) - std::complex<double>(out[i]));
    if (!(d2 <= 4 * tol * tol * den)) ++bad;
    g.Synchronize();
    return bad;
}

int main(int argc, char** argv) {
    Environment env(argc, argv);
    Grid g;  // 1 x 1 on the current device
    int bad = exercise<float>(g) + exercise<double>(g) + exercise_half(g) + exercise_complex<float>(g) + exercise_complex<double>(g);
    if (Blocksize() != 128 || Grid::DefaultHeight(8) != 2) ++bad;
    std::printf(bad ? "EL_API_CHECK_FAIL %d\n" : "EL_API_CHECK_PASS\n", bad);
    return bad ? 1 : 0;
}
