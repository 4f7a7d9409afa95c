"""ORACLE (test infrastructure, not product code): numpy restatement of the reference's
distributed El::Gemm — the SUMMA A/B/C/Dot loops for NN/NT/TN/TT exactly as written in
src/blas_like/level3/Gemm/{NN,NT,TN,TT}.hpp, over p virtual ranks, plus the plain global GEMM
they must all equal.  Local products use numpy's matmul (OpenBLAS dgemm/sgemm — the same kind
of un-pinned CPU BLAS the reference links, src/core/imports/blas/Gemm.hpp:369-452).
"""
from __future__ import annotations

import numpy as np

from .dist import MC, MR, STAR, VC, VR, VDM, Grid, axpy_contract, redistribute, transpose  # noqa: F401

NORMAL, TRANSPOSE, ADJOINT = 0, 1, 2
# values of the reference enum El::GemmAlgorithm (include/El/blas_like/level3.hpp:23-33)
(GEMM_DEFAULT, GEMM_SUMMA_A_MS, GEMM_SUMMA_A, GEMM_SUMMA_B_MS, GEMM_SUMMA_B, GEMM_SUMMA_C_MS, GEMM_SUMMA_C,
 GEMM_SUMMA_DOT, GEMM_CANNON) = range(9)


def op(X: np.ndarray, orient: int) -> np.ndarray:
    if orient == NORMAL:
        return X
    return X.conj().T if (orient == ADJOINT and np.iscomplexobj(X)) else X.T


def gemm_global(orientA, orientB, alpha, A, B, beta, C, acc=np.float64) -> np.ndarray:
    """C = alpha op(A) op(B) + beta C computed in `acc` precision on the global matrices: the value
    every distributed variant approximates. Result is returned in `acc`."""
    if np.iscomplexobj(A) or np.iscomplexobj(B) or np.iscomplexobj(C) or isinstance(alpha, complex) or isinstance(beta, complex):
        acc = np.complex128
    P = op(np.asarray(A, dtype=acc), orientA) @ op(np.asarray(B, dtype=acc), orientB)
    return acc(alpha) * P + (acc(beta) * np.asarray(C, dtype=acc) if beta != 0 else 0)


def resolve_alg(m: int, n: int, k: int, alg: int = GEMM_DEFAULT) -> int:
    """GEMM_DEFAULT heuristic (Gemm/NN.hpp:572-600; same in NT.hpp:494-511, TN.hpp:491-508,
    TT.hpp:411-422)."""
    if alg == GEMM_DEFAULT:
        if 10. * m <= k and 10. * n <= k:
            return GEMM_SUMMA_DOT
        if m <= n and 2. * m <= k:
            return GEMM_SUMMA_B
        if n <= m and 2. * n <= k:
            return GEMM_SUMMA_A
        return GEMM_SUMMA_C
    return {GEMM_SUMMA_A_MS: GEMM_SUMMA_A, GEMM_SUMMA_B_MS: GEMM_SUMMA_B, GEMM_SUMMA_C_MS: GEMM_SUMMA_C}.get(alg, alg)


def _local_gemm(oa, ob, alpha, A: VDM, B: VDM, beta, C: VDM) -> None:
    """LocalGemm (src/blas_like/level3/Gemm.cpp:318-426): per-rank BLAS gemm on local matrices."""
    for q in range(C.g.p):
        P = op(A.loc[q], oa) @ op(B.loc[q], ob)
        C.loc[q][...] = alpha * P + (beta * C.loc[q] if beta != 0 else 0)


def _zeros_like_dist(g: Grid, U, V, ca, ra, h, w, dtype) -> VDM:
    return VDM.zeros(g, U, V, ca, ra, h, w, dtype)


def _axpy_redist(alpha, X: VDM, Y: VDM) -> None:
    """Axpy between different distributions = Copy to Y's distribution, then local axpy
    (level1/Axpy.hpp:120-175)."""
    s = Y.spec
    X2 = redistribute(X, s.U, s.V, s.col_align, s.row_align)
    for q in range(Y.g.p):
        Y.loc[q] += alpha * X2.loc[q]


def _contract(A: VDM, U, V, ca, ra) -> VDM:
    """Contract (level1/Contract.hpp:14-99): zero B, AxpyContract(1, A, B)."""
    B = _zeros_like_dist(A.g, U, V, ca, ra, A.spec.height, A.spec.width, A.loc[0].dtype)
    axpy_contract(1, A, B)
    return B


def summa(orientA, orientB, alpha, A: VDM, B: VDM, C: VDM, alg: int, nb: int = 128, dot_block: int = 2000) -> None:
    """C += alpha op(A) op(B), C/A/B all [MC,MR] (the proxies of every SUMMA_* routine)."""
    g = C.g
    dt = C.loc[0].dtype
    m, n = C.spec.height, C.spec.width
    k = A.spec.width if orientA == NORMAL else A.spec.height
    ta, tb = orientA != NORMAL, orientB != NORMAL
    cs = C.spec
    if alg == GEMM_SUMMA_C:
        for k0 in range(0, k, nb):
            kb = min(nb, k - k0)
            if not ta and not tb:      # NN.hpp:371-384
                A1 = redistribute(A.view(0, m, k0, k0 + kb), MC, STAR, cs.col_align, 0)
                B1T = redistribute(transpose(B.view(k0, k0 + kb, 0, n)), MR, STAR, cs.row_align, 0)
                _local_gemm(NORMAL, TRANSPOSE, alpha, A1, B1T, 1, C)
            elif not ta and tb:        # NT.hpp:280-294
                A1 = redistribute(A.view(0, m, k0, k0 + kb), MC, STAR, cs.col_align, 0)
                B1 = redistribute(B.view(0, n, k0, k0 + kb), VR, STAR, cs.row_align, 0)
                B1T = redistribute(transpose(B1), STAR, MR, 0, cs.row_align)
                _local_gemm(NORMAL, NORMAL, alpha, A1, B1T, 1, C)
            elif ta and not tb:        # TN.hpp:278-291
                A1 = redistribute(A.view(k0, k0 + kb, 0, m), STAR, MC, 0, cs.col_align)
                B1T = redistribute(transpose(B.view(k0, k0 + kb, 0, n)), MR, STAR, cs.row_align, 0)
                _local_gemm(TRANSPOSE, TRANSPOSE, alpha, A1, B1T, 1, C)
            else:                      # TT.hpp:225-240
                A1 = redistribute(A.view(k0, k0 + kb, 0, m), STAR, MC, 0, cs.col_align)
                B1 = redistribute(B.view(0, n, k0, k0 + kb), VR, STAR, cs.row_align, 0)
                B1T = redistribute(transpose(B1), STAR, MR, 0, cs.row_align)
                _local_gemm(TRANSPOSE, NORMAL, alpha, A1, B1T, 1, C)
    elif alg == GEMM_SUMMA_A:
        a = A.spec
        for j0 in range(0, n, nb):
            w = min(nb, n - j0)
            C1 = C.view(0, m, j0, j0 + w)
            if not ta and not tb:      # NN.hpp:138-153
                B1 = redistribute(B.view(0, k, j0, j0 + w), VR, STAR, a.row_align, 0)
                B1T = redistribute(transpose(B1), STAR, MR, 0, a.row_align)
                D1 = _zeros_like_dist(g, MC, STAR, a.col_align, 0, m, w, dt)
                _local_gemm(NORMAL, TRANSPOSE, alpha, A, B1T, 0, D1)
                axpy_contract(1, D1, C1)
            elif not ta and tb:        # NT.hpp:46-58
                B1T = redistribute(transpose(B.view(j0, j0 + w, 0, k)), MR, STAR, a.row_align, 0)
                D1 = _zeros_like_dist(g, MC, STAR, a.col_align, 0, m, w, dt)
                _local_gemm(NORMAL, NORMAL, alpha, A, B1T, 0, D1)
                axpy_contract(1, D1, C1)
            elif ta and not tb:        # TN.hpp:46-60
                B1 = redistribute(B.view(0, k, j0, j0 + w), MC, STAR, a.col_align, 0)
                D1 = _zeros_like_dist(g, MR, STAR, a.row_align, 0, m, w, dt)
                _local_gemm(TRANSPOSE, NORMAL, alpha, A, B1, 0, D1)
                _axpy_redist(1, _contract(D1, MR, MC, a.row_align, 0), C1)
            else:                      # TT.hpp:45-60
                B1 = redistribute(B.view(j0, j0 + w, 0, k), STAR, MC, 0, a.col_align)
                D1 = _zeros_like_dist(g, MR, STAR, a.row_align, 0, m, w, dt)
                _local_gemm(TRANSPOSE, TRANSPOSE, alpha, A, B1, 0, D1)
                _axpy_redist(1, _contract(D1, MR, MC, a.row_align, 0), C1)
    elif alg == GEMM_SUMMA_B:
        b = B.spec
        for i0 in range(0, m, nb):
            h = min(nb, m - i0)
            C1 = C.view(i0, i0 + h, 0, n)
            if not ta and not tb:      # NN.hpp:255-269
                A1 = redistribute(A.view(i0, i0 + h, 0, k), STAR, MC, 0, b.col_align)
                D1T = _zeros_like_dist(g, MR, STAR, b.row_align, 0, n, h, dt)
                _local_gemm(TRANSPOSE, TRANSPOSE, alpha, B, A1, 0, D1T)
                S = _contract(D1T, MR, MC, C1.spec.row_align, C1.spec.col_align)  # TransposeAxpyContract.hpp:26-41
                for q in range(g.p):
                    C1.loc[q] += S.loc[q].T
            elif not ta and tb:        # NT.hpp:161-175
                A1T = redistribute(transpose(A.view(i0, i0 + h, 0, k)), MR, STAR, b.row_align, 0)
                D1 = _zeros_like_dist(g, STAR, MC, 0, b.col_align, h, n, dt)
                _local_gemm(TRANSPOSE, TRANSPOSE, alpha, A1T, B, 0, D1)
                _axpy_redist(1, _contract(D1, MR, MC, 0, b.col_align), C1)
            elif ta and not tb:        # TN.hpp:164-175 (alpha applied in the contract step)
                A1 = redistribute(A.view(0, k, i0, i0 + h), MC, STAR, b.col_align, 0)
                D1T = _zeros_like_dist(g, MR, STAR, b.row_align, 0, n, h, dt)
                _local_gemm(TRANSPOSE, NORMAL, 1, B, A1, 0, D1T)
                S = _contract(D1T, MR, MC, C1.spec.row_align, C1.spec.col_align)
                for q in range(g.p):
                    C1.loc[q] += alpha * S.loc[q].T
            else:                      # TT.hpp:136-151
                A1 = redistribute(A.view(0, k, i0, i0 + h), VR, STAR, b.row_align, 0)
                A1T = redistribute(transpose(A1), STAR, MR, 0, b.row_align)
                D1 = _zeros_like_dist(g, STAR, MC, 0, b.col_align, h, n, dt)
                _local_gemm(NORMAL, TRANSPOSE, alpha, A1T, B, 0, D1)
                _axpy_redist(1, _contract(D1, MR, MC, 0, b.col_align), C1)
    elif alg == GEMM_SUMMA_DOT:
        # NN.hpp:476-510, NT.hpp:383-417, TN.hpp:384-415, TT.hpp:301-332: k spread over VC
        Ak = redistribute(A, VC, STAR, 0, 0) if ta else redistribute(A, STAR, VC, 0, 0)
        Bk = redistribute(B, STAR, VC, 0, 0) if tb else redistribute(B, VC, STAR, 0, 0)
        for i0 in range(0, m, dot_block):
            h = min(dot_block, m - i0)
            A1 = Ak.view(0, k, i0, i0 + h) if ta else Ak.view(i0, i0 + h, 0, k)
            for j0 in range(0, n, dot_block):
                w = min(dot_block, n - j0)
                B1 = Bk.view(j0, j0 + w, 0, k) if tb else Bk.view(0, k, j0, j0 + w)
                C11 = _zeros_like_dist(g, STAR, STAR, 0, 0, h, w, dt)
                _local_gemm(orientA, orientB, alpha, A1, B1, 0, C11)
                axpy_contract(1, C11, C.view(i0, i0 + h, j0, j0 + w))
    else:
        raise ValueError("unsupported GemmAlgorithm")


def dist_gemm(g: Grid, orientA, orientB, alpha, A, B, beta, C, alg=GEMM_DEFAULT, nb=128,
              aligns=((0, 0), (0, 0), (0, 0))) -> np.ndarray:
    """El::Gemm(orientA, orientB, alpha, A, B, beta, C, alg) on global numpy matrices, executed
    the reference's way over g.p virtual ranks (Gemm.cpp:273-302: Scale(beta,C), then SUMMA).
    Returns the global C in C's dtype."""
    m, n = C.shape
    k = A.shape[1] if orientA == NORMAL else A.shape[0]
    Ad = VDM.from_global(g, A, MC, MR, *aligns[0])
    Bd = VDM.from_global(g, B, MC, MR, *aligns[1])
    Cd = VDM.from_global(g, np.array(C, copy=True), MC, MR, *aligns[2])
    for q in range(g.p):  # Scale(beta, C): beta == 0 -> Zero (level1/Scale.hpp:17-30)
        Cd.loc[q][...] = 0 if beta == 0 else beta * Cd.loc[q]
    if m and n and k and alpha != 0:
        summa(orientA, orientB, C.dtype.type(alpha), Ad, Bd, Cd, resolve_alg(m, n, k, alg), nb)
    return Cd.to_global()


def dist_gemm_complex(g: Grid, orientA, orientB, alpha, A, B, beta, C, alg=GEMM_DEFAULT, nb=128) -> np.ndarray:
    """El::Gemm for Complex<float|double> (src/blas_like/level3/Gemm.cpp:504-506) the way THIS path computes it
    (elemental_b200/csrc/capi_complex.cpp, "4M"): with op(A) = Ar' + i sa Ai', op(B) = Br' + i sb Bi' (sa, sb = -1 for
    ADJOINT) the product is four REAL El::Gemm calls on the planes — each restated by dist_gemm above and pinned to the
    reference's real outputs —
        T = Ar'Br' - sa sb Ai'Bi',   U = sb Ar'Bi' + sa Ai'Br',   C = alpha (T + iU) + beta C.
    Pinned against the reference's own complex outputs by tests/test_oracle.py (tests/golden/ref_gemm_complex_1x1.npz)."""
    real = np.float32 if np.dtype(C.dtype) == np.complex64 else np.float64
    oa, ob = (NORMAL if orientA == NORMAL else TRANSPOSE), (NORMAL if orientB == NORMAL else TRANSPOSE)
    sa, sb = (-1.0 if orientA == ADJOINT else 1.0), (-1.0 if orientB == ADJOINT else 1.0)
    Ar, Ai = np.ascontiguousarray(A.real, dtype=real), np.ascontiguousarray(A.imag, dtype=real)
    Br, Bi = np.ascontiguousarray(B.real, dtype=real), np.ascontiguousarray(B.imag, dtype=real)
    Z = np.zeros(C.shape, dtype=real)
    T = dist_gemm(g, oa, ob, 1.0, Ar, Br, 0.0, Z, alg, nb)
    T = dist_gemm(g, oa, ob, -sa * sb, Ai, Bi, 1.0, T, alg, nb)
    U = dist_gemm(g, oa, ob, sb, Ar, Bi, 0.0, Z, alg, nb)
    U = dist_gemm(g, oa, ob, sa, Ai, Br, 1.0, U, alg, nb)
    al, be = complex(alpha), complex(beta)
    Cr, Ci = C.real.astype(real), C.imag.astype(real)
    out_r = real(al.real) * T - real(al.imag) * U + (real(be.real) * Cr - real(be.imag) * Ci if be != 0 else 0)
    out_i = real(al.real) * U + real(al.imag) * T + (real(be.real) * Ci + real(be.imag) * Cr if be != 0 else 0)
    return (out_r + 1j * out_i).astype(C.dtype)


def summa_nnc_1x1(alpha, A: np.ndarray, B: np.ndarray, beta, C: np.ndarray, nb: int = 128) -> np.ndarray:
    """The reference's CPU El::Gemm(NORMAL,NORMAL,...,GEMM_SUMMA_C) on a 1 x 1 grid, without the
    virtual-rank bookkeeping: Scale(beta,C) (Gemm.cpp:282), then per nb-panel a local copy of A1
    (RowAllGather.hpp:35-38), a local transpose of B1 (Transpose/ColAllGather.hpp:26-32) and one
    BLAS rank-nb update (NN.hpp:371-384). Used as the timed CPU port when oracle/_ref is absent."""
    C = np.array(C, order="F", copy=True)
    C *= C.dtype.type(beta)
    k = A.shape[1]
    al = C.dtype.type(alpha)
    for k0 in range(0, k, nb):
        A1 = np.array(A[:, k0:k0 + nb], order="F")          # A1_MC_STAR = A1
        B1T = np.array(B[k0:k0 + nb, :].T, order="F")       # Transpose(B1, B1Trans_MR_STAR)
        C += al * (A1 @ B1T.T)                              # LocalGemm(NORMAL, TRANSPOSE, alpha, ., ., 1, C)
    return C


def frobenius_rel_err(X, ref) -> float:
    acc = np.complex128 if (np.iscomplexobj(X) or np.iscomplexobj(ref)) else np.float64
    ref = np.asarray(ref, dtype=acc)
    d = np.asarray(X, dtype=acc) - ref
    den = np.linalg.norm(ref)
    return float(np.linalg.norm(d) / (den if den > 0 else 1.0))


def gemm_tolerance(dtype, k: int) -> float:
    """north_star bound: relative Frobenius error <= 1e-13*sqrt(k) (fp64) / 1e-5*sqrt(k) (fp32)."""
    if np.dtype(dtype) == np.float16:
        # half: no bound in north_star and no golden vectors in the reference; products accumulate in fp32
        # on the device and are rounded to half (eps 9.8e-4) once per SUMMA panel / partial-sum reduction
        return 4e-3
    if np.dtype(dtype).kind == "c":
        # complex: a complex dot product of length k sums 2k real products per component
        return gemm_tolerance(np.float32 if np.dtype(dtype) == np.complex64 else np.float64, 2 * k)
    return (1e-13 if np.dtype(dtype) == np.float64 else 1e-5) * float(np.sqrt(max(k, 1)))


def uniform_matrix(rows: int, cols: int, seed: int, dtype, center=0.0, radius=0.1) -> np.ndarray:
    """Synthetic inputs: Uniform(center-radius, center+radius) like the reference's drivers
    (tests/blas_like/Gemm_Suite.cpp:158,170-172; src/matrices/random/independent/Uniform.cpp:18-57),
    from a counter-based generator keyed by (seed, i, j) so the GLOBAL matrix does not depend on
    the grid shape."""
    rng = np.random.Generator(np.random.Philox(key=seed))
    return np.asfortranarray((center - radius) + 2 * radius * rng.random((rows, cols))).astype(dtype)
