"""ORACLE (test infrastructure, not product code): numpy restatement of the reference's triangular
rank-k family over p virtual ranks — the trapezoid level-1 routines, LocalTrrk's recursion, the
Trrk / Syrk panel loops and Symm — plus the global definitions they must equal.

  scale_trapezoid / axpy_trapezoid / make_trapezoidal   include/El/blas_like/level1/ScaleTrapezoid.hpp:49-96,
                                                        AxpyTrapezoid.hpp, MakeTrapezoidal.hpp:45-86
  make_symmetric                                        level1/MakeSymmetric.hpp:56-73
  local_trrk                                            src/blas_like/level3/Trrk/Local.hpp:350-396, 688-890
  trrk                                                  src/blas_like/level3/Trrk.cpp:100-118, Trrk/{NN,NT,TN,TT}.hpp
  syrk                                                  src/blas_like/level3/Syrk.cpp:172-190, Syrk/{LN,LT,UN,UT}.hpp
  symm                                                  src/blas_like/level3/Symm.cpp:56-80

Pinned against the reference itself: tests/golden/ref_level3.npz holds per-rank outputs of the
reference's own CPU build (oracle/ref_build, multi-process MPI stub) for every routine here on six
grid shapes; tests/test_oracle.py::test_level3_* compares.  GEMM values inside are un-pinned BLAS
(SURVEY 8(c)), so Trrk / Syrk / Symm agree to summation-order rounding; the trapezoid routines and
MakeSymmetric move or scale entries only and are bit-exact.
"""
from __future__ import annotations

import numpy as np

from .dist import MC, MR, STAR, VDM, Grid, Spec, redistribute, transpose  # noqa: F401
from .gemm import NORMAL, _local_gemm, op  # noqa: F401

LOWER, UPPER = 0, 1   # El::UpperOrLower (include/El/core/types.hpp:511-515)
LEFT, RIGHT = 0, 1    # El::LeftOrRight (types.hpp:418-422)
LOCAL_TRRK_BLOCKSIZE = 64  # src/blas_like/blocksizes.cpp:26


def trapezoid_mask(h: int, w: int, uplo: int, offset: int = 0, i0: int = 0, istep: int = 1, j0: int = 0, jstep: int = 1):
    """True where global (i,j) = (i0 + iLoc*istep, j0 + jLoc*jstep) lies in the trapezoid: LOWER keeps
    i >= j - offset (ScaleTrapezoid.hpp:86-92), UPPER keeps i <= j - offset (:70-78)."""
    gi = (i0 + istep * np.arange(h))[:, None]
    gj = (j0 + jstep * np.arange(w))[None, :]
    return gi <= gj - offset if uplo == UPPER else gi >= gj - offset


def _local_mask(A: VDM, q: int, uplo: int, offset: int):
    s, g = A.spec, A.g
    lh, lw = A.loc[q].shape
    return trapezoid_mask(lh, lw, uplo, offset, s.col_shift(g, q), g.stride(s.U), s.row_shift(g, q), g.stride(s.V))


def scale_trapezoid(alpha, uplo: int, A: VDM, offset: int = 0) -> None:
    if alpha == 1:
        return
    for q in range(A.g.p):
        if A.loc[q].size:
            m = _local_mask(A, q, uplo, offset)
            A.loc[q][m] = A.loc[q][m] * A.loc[q].dtype.type(alpha)


def axpy_trapezoid(uplo: int, alpha, X: VDM, Y: VDM, offset: int = 0) -> None:
    """Y += alpha X inside the trapezoid; X is brought into Y's distribution first when it differs."""
    s = Y.spec
    if (X.spec.U, X.spec.V, X.spec.col_align, X.spec.row_align, X.spec.root) != (s.U, s.V, s.col_align, s.row_align, s.root):
        X = redistribute(X, s.U, s.V, s.col_align, s.row_align, s.root)
    for q in range(Y.g.p):
        if Y.loc[q].size:
            m = _local_mask(Y, q, uplo, offset)
            Y.loc[q][m] += Y.loc[q].dtype.type(alpha) * X.loc[q][m]


def make_trapezoidal(uplo: int, A: VDM, offset: int = 0) -> None:
    for q in range(A.g.p):
        if A.loc[q].size:
            A.loc[q][~_local_mask(A, q, uplo, offset)] = 0


def make_symmetric(uplo: int, A: VDM) -> None:
    """MakeTrapezoidal, transpose into the same distribution, add the strict opposite triangle
    (MakeSymmetric.hpp:63-72)."""
    assert A.spec.height == A.spec.width
    make_trapezoidal(uplo, A, 0)
    s = A.spec
    At = redistribute(transpose(A), s.U, s.V, s.col_align, s.row_align, s.root)
    if uplo == LOWER:
        axpy_trapezoid(UPPER, 1, At, A, 1)
    else:
        axpy_trapezoid(LOWER, 1, At, A, -1)


def _rows_of(A: VDM, oa: int, i0: int, i1: int) -> VDM:
    return A.view(i0, i1, 0, A.spec.width) if oa == NORMAL else A.view(0, A.spec.height, i0, i1)


def _cols_of(B: VDM, ob: int, j0: int, j1: int) -> VDM:
    return B.view(0, B.spec.height, j0, j1) if ob == NORMAL else B.view(j0, j1, 0, B.spec.width)


def _trrk_kernel(uplo, oa, ob, alpha, A: VDM, B: VDM, C: VDM) -> None:
    """LocalTrrkKernel (Trrk/Local.hpp:350-396): one split; the off-diagonal quadrant by LocalGemm, the two
    diagonal quadrants through temporaries DTL / DBR and LocalAxpyTrapezoid."""
    n = C.spec.height
    half = n // 2
    dt = C.loc[0].dtype
    AT, AB = _rows_of(A, oa, 0, half), _rows_of(A, oa, half, n)
    BL, BR = _cols_of(B, ob, 0, half), _cols_of(B, ob, half, n)
    if uplo == LOWER:
        _local_gemm(oa, ob, alpha, AB, BL, 1, C.view(half, n, 0, half))
    else:
        _local_gemm(oa, ob, alpha, AT, BR, 1, C.view(0, half, half, n))
    for (a, b, c) in ((AT, BL, C.view(0, half, 0, half)), (AB, BR, C.view(half, n, half, n))):
        D = VDM.zeros(C.g, MC, MR, c.spec.col_align, c.spec.row_align, c.spec.height, c.spec.width, dt)
        _local_gemm(oa, ob, alpha, a, b, 0, D)
        axpy_trapezoid(uplo, 1, D, c, 0)


def local_trrk(uplo, oa, ob, alpha, A: VDM, B: VDM, beta, C: VDM, blocksize: int = LOCAL_TRRK_BLOCKSIZE) -> None:
    """LocalTrrk (Trrk/Local.hpp:688-890): ScaleTrapezoid(beta), then recursion on the two diagonal halves
    with one LocalGemm on the off-diagonal corner, down to C.Height() < gridWidth * LocalTrrkBlocksize."""
    g = C.g
    scale_trapezoid(beta, uplo, C, 0)
    n = C.spec.height
    if n < g.c * blocksize:
        _trrk_kernel(uplo, oa, ob, alpha, A, B, C)
        return
    half = n // 2
    AT, AB = _rows_of(A, oa, 0, half), _rows_of(A, oa, half, n)
    BL, BR = _cols_of(B, ob, 0, half), _cols_of(B, ob, half, n)
    if uplo == LOWER:
        _local_gemm(oa, ob, alpha, AB, BL, 1, C.view(half, n, 0, half))
    else:
        _local_gemm(oa, ob, alpha, AT, BR, 1, C.view(0, half, half, n))
    local_trrk(uplo, oa, ob, alpha, AT, BL, 1, C.view(0, half, 0, half), blocksize)
    local_trrk(uplo, oa, ob, alpha, AB, BR, 1, C.view(half, n, half, n), blocksize)


def trrk(uplo, oa, ob, alpha, A: VDM, B: VDM, beta, C: VDM, nb: int = 128) -> None:
    """Trrk on [MC,MR] operands (Trrk.cpp:100-118): ScaleTrapezoid(beta, uplo, C), then per nb-wide k
    panel A1 -> [MC,*] (or [*,MC]), B1 -> [MR,*] transposed (or [*,MR]) aligned with C, LocalTrrk
    (Trrk/NN.hpp:45-61 and the NT / TN / TT twins)."""
    g = C.g
    oa, ob = (1 if oa else 0), (1 if ob else 0)
    n = C.spec.height
    k = A.spec.height if oa else A.spec.width
    scale_trapezoid(beta, uplo, C, 0)
    ca, ra = C.spec.col_align, C.spec.row_align
    for k0 in range(0, k, nb):
        kb = min(nb, k - k0)
        A1 = A.view(k0, k0 + kb, 0, n) if oa else A.view(0, n, k0, k0 + kb)
        B1 = B.view(0, n, k0, k0 + kb) if ob else B.view(k0, k0 + kb, 0, n)
        a = redistribute(A1, STAR, MC, 0, ca) if oa else redistribute(A1, MC, STAR, ca, 0)
        b = redistribute(B1, MR, STAR, ra, 0) if ob else redistribute(B1, STAR, MR, 0, ra)
        local_trrk(uplo, oa, ob, alpha, a, b, 1, C)


def syrk(uplo, orientation, alpha, A: VDM, beta, C: VDM, nb: int = 128) -> None:
    """Syrk (Syrk.cpp:172-190): ScaleTrapezoid(beta, uplo, C), then syrk::{LN,LT,UN,UT}: per panel
    A1 -> [MC,*] and A1^T -> [*,MR] (LN.hpp:40-47), or A1^T -> [MR,*] and A1 -> [*,MC] (LT.hpp:40-51),
    LocalTrrk into the triangle — i.e. Trrk with B = A and the opposite orientation."""
    t = 1 if orientation else 0
    trrk(uplo, t, 1 - t, alpha, A, A, beta, C, nb)


def symm(side, uplo, alpha, A: VDM, B: VDM, beta, C: VDM) -> None:
    """Symm (Symm.cpp:56-80): C = alpha sym(A) B + beta C (LEFT) or alpha B sym(A) + beta C (RIGHT), only the
    `uplo` triangle of A referenced.  The reference's symm::{LL,LU,RL,RU} accumulate the two triangles'
    contributions separately (Symm/LL.hpp); the restatement evaluates the same definition on the
    global matrices in the local dtype and is pinned against the reference's outputs."""
    g = C.g
    Ag = A.to_global()
    tri = np.where(trapezoid_mask(*Ag.shape, uplo, 0), Ag, 0)
    strict = np.where(trapezoid_mask(*Ag.shape, uplo, -1 if uplo == LOWER else 1), Ag, 0)
    S = tri + strict.T
    Bg, Cg = B.to_global(), C.to_global()
    dt = Cg.dtype.type
    P = S @ Bg if side == LEFT else Bg @ S
    R = dt(alpha) * P + (dt(beta) * Cg if beta != 0 else 0)
    s = C.spec
    out = VDM.from_global(g, np.asfortranarray(R.astype(Cg.dtype)), s.U, s.V, s.col_align, s.row_align, s.root)
    for q in range(g.p):
        C.loc[q][...] = out.loc[q]


# ----------------------------------------------------------------------------- global definitions
def trrk_global(uplo, oa, ob, alpha, A, B, beta, C, acc=np.longdouble) -> np.ndarray:
    """Triangle of C = alpha op(A) op(B) + beta C, the other triangle untouched; computed in `acc`."""
    Ca = np.asarray(C, dtype=acc)
    full = acc(alpha) * (op(np.asarray(A, dtype=acc), oa) @ op(np.asarray(B, dtype=acc), ob)) + acc(beta) * Ca
    return np.where(trapezoid_mask(*Ca.shape, uplo, 0), full, Ca)


def syrk_global(uplo, orientation, alpha, A, beta, C, acc=np.longdouble) -> np.ndarray:
    t = 1 if orientation else 0
    return trrk_global(uplo, t, 1 - t, alpha, A, A, beta, C, acc)


def symm_global(side, uplo, alpha, A, B, beta, C, acc=np.longdouble) -> np.ndarray:
    Aa = np.asarray(A, dtype=acc)
    S = np.where(trapezoid_mask(*Aa.shape, uplo, 0), Aa, 0) + np.where(trapezoid_mask(*Aa.shape, uplo, -1 if uplo == LOWER else 1), Aa, 0).T
    Ba, Ca = np.asarray(B, dtype=acc), np.asarray(C, dtype=acc)
    return acc(alpha) * (S @ Ba if side == LEFT else Ba @ S) + acc(beta) * Ca


# ----------------------------------------------------------------------------- Trsm
NON_UNIT, UNIT = 0, 1  # El::UnitOrNonUnit (types.hpp:489-493)


def _tri_solve(side, uplo, orient, diag, A, B):
    """op(A) X = B (LEFT) or X op(A) = B (RIGHT) by substitution in the operands' dtype (longdouble allowed);
    only the `uplo` triangle of A is read. What BLAS trsm computes (src/core/imports/blas/Trsm.hpp)."""
    A = np.asarray(A)
    E = np.tril(A) if uplo == LOWER else np.triu(A)
    e_lower = (uplo == LOWER) != bool(orient)      # op(A) is lower triangular
    if orient != NORMAL:
        E = E.T
    if side == RIGHT:                               # X E = B  <=>  E^T X^T = B^T
        return _tri_solve(LEFT, UPPER if e_lower else LOWER, NORMAL, diag, E.T, np.asarray(B).T).T
    n = E.shape[0]
    X = np.array(B, dtype=np.result_type(E.dtype, np.asarray(B).dtype), copy=True)
    for i in (range(n) if e_lower else range(n - 1, -1, -1)):
        if e_lower:
            X[i, :] -= E[i, :i] @ X[:i, :]
        else:
            X[i, :] -= E[i, i + 1:] @ X[i + 1:, :]
        if diag != UNIT:
            X[i, :] /= E[i, i]
    return X


def trsm_global(side, uplo, orient, diag, alpha, A, B, acc=np.longdouble) -> np.ndarray:
    """B := alpha op(A)^-1 B (LEFT) or alpha B op(A)^-1 (RIGHT), computed in `acc`."""
    Aa = np.asarray(A, dtype=acc)
    if diag == UNIT:
        Aa = Aa.copy()
        np.fill_diagonal(Aa, 1)
    return _tri_solve(side, uplo, orient, NON_UNIT, Aa, acc(alpha) * np.asarray(B, dtype=acc))


def trsm(side, uplo, orient, diag, alpha, A: VDM, B: VDM, nb: int = 128) -> None:
    """Trsm on [MC,MR] operands (Trsm.cpp:129-440): B *= alpha, then the block loop of trsm::LLNLarge
    (Trsm/LLN.hpp:37-68) and its LLT / LUN / LUT / RLN / RLT / RUN / RUT twins: per nb block in the direction
    of the substitution  A11 -> [*,*],  X1 -> [*,VR] (LEFT) / [VC,*] (RIGHT),  LocalTrsm on every rank's share,
    X1 -> [*,MR] / [MC,*] and back into X,  LocalGemm update of the remaining rows / columns."""
    from .dist import VC, VR
    g = B.g
    t = 1 if orient else 0
    dt = B.loc[0].dtype
    for q in range(g.p):
        B.loc[q] *= dt.type(alpha)
    na = A.spec.height
    m, n = B.spec.height, B.spec.width
    eff_lower = (uplo == LOWER) != bool(t)
    forward = eff_lower if side == LEFT else not eff_lower
    starts = list(range(0, na, nb))
    for k0 in (starts if forward else reversed(starts)):
        k1 = min(na, k0 + nb)
        r0, r1 = (k1, na) if forward else (0, k0)
        A11 = redistribute(A.view(k0, k1, k0, k1), STAR, STAR)
        X1 = B.view(k0, k1, 0, n) if side == LEFT else B.view(0, m, k0, k1)
        X1v = redistribute(X1, STAR, VR, 0, X1.spec.row_align) if side == LEFT else redistribute(X1, VC, STAR, X1.spec.col_align, 0)
        for q in range(g.p):
            if X1v.loc[q].size:
                X1v.loc[q][...] = _tri_solve(side, uplo, t, diag, _unit(A11.loc[q], diag), X1v.loc[q]).astype(dt)
        X1p = (redistribute(X1v, STAR, MR, 0, B.spec.row_align) if side == LEFT else redistribute(X1v, MC, STAR, B.spec.col_align, 0))
        back = redistribute(X1p, MC, MR, X1.spec.col_align, X1.spec.row_align)
        for q in range(g.p):
            X1.loc[q][...] = back.loc[q]
        if r1 <= r0:
            continue
        if side == LEFT:
            X2 = B.view(r0, r1, 0, n)
            Ap = A.view(k0, k1, r0, r1) if t else A.view(r0, r1, k0, k1)
            a = redistribute(Ap, STAR, MC, 0, X2.spec.col_align) if t else redistribute(Ap, MC, STAR, X2.spec.col_align, 0)
            _local_gemm(t, 0, -1, a, X1p, 1, X2)
        else:
            X2 = B.view(0, m, r0, r1)
            Ap = A.view(r0, r1, k0, k1) if t else A.view(k0, k1, r0, r1)
            a = redistribute(Ap, MR, STAR, X2.spec.row_align, 0) if t else redistribute(Ap, STAR, MR, 0, X2.spec.row_align)
            _local_gemm(0, t, -1, X1p, a, 1, X2)


def _unit(A11: np.ndarray, diag: int) -> np.ndarray:
    if diag != UNIT:
        return A11
    out = A11.copy()
    np.fill_diagonal(out, 1)
    return out
