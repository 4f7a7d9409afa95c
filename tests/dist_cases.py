"""Shared distributed parity cases, run by every rank of a grid (1 x 1 inside pytest, r x c under
torchrun via tests/dist_worker.py). Each case builds the same seeded GLOBAL matrices on every rank,
loads this rank's entries, runs the CUDA path through the C ABI handle API, and compares this
rank's local result with the oracle's (oracle.dist / oracle.gemm over virtual ranks).

The procedure mirrors the reference's own tests:
  * tests/core/DistMatrix.cpp:12-76  — every (U,V) <- (U',V') assignment must equal the [*,*]
    gather exactly (bit-exact here too);
  * tests/blas_like/BasicGemm.cpp:134-235 — GPU Gemm vs CPU recomputation (alpha=2, beta=3);
  * tests/blas_like/Gemm.cpp:52-231 — orientations x GemmAlgorithm x alignments.
"""
from __future__ import annotations

import ctypes as C
import itertools
import os

import numpy as np
import torch

from elemental_b200 import el
from oracle import dist as OD
from oracle import gemm as OG
from oracle import level3 as O3

NPDT = {torch.float32: np.float32, torch.float64: np.float64, torch.float16: np.float16}


def use_backend(module) -> None:
    """Swap the `el` interface the cases drive: elemental_b200.el (the CUDA library, default) or
    tests/hostsim/sim_el.py (the same ENGINE sources over host test doubles, for machines without a GPU)."""
    global el
    el = module


def _pin(t: torch.Tensor) -> torch.Tensor:
    return t.pin_memory() if torch.cuda.is_available() else t


def _grid_oracle(grid) -> OD.Grid:
    return OD.Grid(grid.Height(), grid.Width())


def check_redistributions(grid, dtype=torch.float64, m=23, n=17, pairs=None, seed=0) -> int:
    """B = A for (U,V) -> (U',V') with random alignments; exact equality of local matrices."""
    og = _grid_oracle(grid)
    rng = np.random.default_rng(seed)  # same stream on every rank
    npdt = NPDT[dtype]
    G = np.asfortranarray(rng.uniform(-1, 1, (m, n)).astype(npdt))
    count = 0
    for (U, V), (U2, V2) in (pairs or itertools.product(OD.PAIRS, OD.PAIRS)):
        ca, ra = int(rng.integers(og.stride(U))), int(rng.integers(og.stride(V)))
        ca2, ra2 = int(rng.integers(og.stride(U2))), int(rng.integers(og.stride(V2)))
        A = el.DistMatrix(grid, U, V, dtype).Align(ca, ra).from_global(G)
        B = el.DistMatrix(grid, U2, V2, dtype).Align(ca2, ra2)
        el.Copy(A, B)
        got = B.get_local()
        spec = OD.Spec(U2, V2, ca2, ra2, m, n)
        exp = OD.local_of(G, spec, og, grid.Rank())
        assert got.shape == exp.shape, ((U, V), (U2, V2), got.shape, exp.shape)
        assert np.array_equal(got, exp), f"Copy [{OD.NAMES[U]},{OD.NAMES[V]}]->[{OD.NAMES[U2]},{OD.NAMES[V2]}] differs"
        # unconstrained target adopts compatible alignments and must still hold the right entries
        B2 = el.DistMatrix(grid, U2, V2, dtype)
        el.Copy(A, B2)
        spec2 = OD.Spec(U2, V2, B2.ColAlign(), B2.RowAlign(), m, n)
        assert np.array_equal(B2.get_local(), OD.local_of(G, spec2, og, grid.Rank()))
        count += 1
    return count


def check_circ(grid, dtype=torch.float64, m=23, n=17) -> int:
    """[CIRC,CIRC] (copy::Gather / copy::Scatter, Copy/Gather.hpp, Copy/Scatter.hpp): gather every
    distribution to a root, scatter it back with other alignments, move the root, transpose through
    it, and round-trip the local matrix through host memory — all bit-exact."""
    og = _grid_oracle(grid)
    p, me = og.p, grid.Rank()
    rng = np.random.default_rng(7)
    npdt = NPDT[dtype]
    G = np.asfortranarray(rng.uniform(-1, 1, (m, n)).astype(npdt))
    count = 0
    for (U, V) in OD.PAIRS:
        ca, ra = int(rng.integers(og.stride(U))), int(rng.integers(og.stride(V)))
        root = int(rng.integers(p))
        A = el.DistMatrix(grid, U, V, dtype).Align(ca, ra).from_global(G)
        R = el.DistMatrix(grid, el.CIRC, el.CIRC, dtype).SetRoot(root)
        el.Copy(A, R)
        assert R.Root() == root and (R.LocalHeight(), R.LocalWidth()) == ((m, n) if me == root else (0, 0))
        if me == root:
            assert np.array_equal(R.get_local(), G), f"gather [{OD.NAMES[U]},{OD.NAMES[V]}] -> root {root}"
        ca2, ra2 = int(rng.integers(og.stride(U))), int(rng.integers(og.stride(V)))
        B = el.DistMatrix(grid, U, V, dtype).Align(ca2, ra2)
        el.Copy(R, B)
        exp = OD.local_of(G, OD.Spec(U, V, ca2, ra2, m, n), og, me)
        assert np.array_equal(B.get_local(), exp), f"scatter root {root} -> [{OD.NAMES[U]},{OD.NAMES[V]}]"
        count += 2
    # root -> another root, and a view of the rooted matrix scattered out
    R = el.DistMatrix(grid, el.CIRC, el.CIRC, dtype).SetRoot(0).from_global(G)
    R2 = el.DistMatrix(grid, el.CIRC, el.CIRC, dtype).SetRoot(p - 1)
    el.Copy(R, R2)
    if me == p - 1:
        assert np.array_equal(R2.get_local(), G)
    else:
        assert R2.LocalHeight() == 0
    V_ = R2.View(2, m - 3, 1, n - 2)
    B = el.DistMatrix(grid, el.MC, el.MR, dtype)
    el.Copy(V_, B)
    sub = np.asfortranarray(G[2:m - 3, 1:n - 2])
    assert np.array_equal(B.get_local(), OD.local_of(sub, OD.Spec(OD.MC, OD.MR, B.ColAlign(), B.RowAlign(), *sub.shape), og, me))
    T = el.DistMatrix(grid, el.VR, el.STAR, dtype)
    el.Transpose(R2, T)
    assert np.array_equal(T.get_local(), OD.local_of(np.asfortranarray(G.T), OD.Spec(OD.VR, OD.STAR, T.ColAlign(), T.RowAlign(), n, m), og, me))
    # CPU <-> GPU copy of the local matrix (ProxyDevice.hpp): pinned host tensors, ld > local height
    A = el.DistMatrix(grid, el.MC, el.MR, dtype).Align(1 % og.r, 0).from_global(G)
    lh, lw = A.LocalHeight(), A.LocalWidth()
    host = _pin(torch.full((max(lw, 1), lh + 3), -7.0, dtype=dtype))
    A.read_local_to_host(host)
    grid.sync()
    loc = OD.local_of(G, OD.Spec(OD.MC, OD.MR, 1 % og.r, 0, m, n), og, me)
    if lh and lw:
        assert np.array_equal(host[:lw, :lh].numpy().T, loc) and bool((host[:, lh:] == -7.0).all())
    host[:lw, :lh] *= 2
    A.write_local_from_host(host)
    grid.sync()
    assert np.array_equal(A.get_local(), 2 * loc)
    # Gemm with rooted operands goes through the [MC,MR] proxies (Proxy.hpp:236-257)
    k = 11
    Ga = np.asfortranarray(rng.uniform(-1, 1, (m, k)).astype(npdt))
    Gb = np.asfortranarray(rng.uniform(-1, 1, (k, n)).astype(npdt))
    Ar = el.DistMatrix(grid, el.CIRC, el.CIRC, dtype).SetRoot(p - 1).from_global(Ga)
    Bm = el.DistMatrix(grid, el.MC, el.MR, dtype).from_global(Gb)
    Cr = el.DistMatrix(grid, el.CIRC, el.CIRC, dtype).SetRoot(0).from_global(G)
    el.Gemm(el.NORMAL, el.NORMAL, 2.0, Ar, Bm, 3.0, Cr, el.GEMM_SUMMA_C)
    if me == 0:
        ref = OG.gemm_global(0, 0, 2.0, Ga, Gb, 3.0, G, np.longdouble)
        assert OG.frobenius_rel_err(Cr.get_local(), ref) <= OG.gemm_tolerance(npdt, k)
    return count + 5


def check_views_and_transpose(grid, dtype=torch.float32) -> None:
    og = _grid_oracle(grid)
    npdt = NPDT[dtype]
    rng = np.random.default_rng(1)
    G = np.asfortranarray(rng.uniform(-1, 1, (29, 31)).astype(npdt))
    A = el.DistMatrix(grid, el.MC, el.MR, dtype).Align(1 % og.r, 2 % og.c).from_global(G)
    V = A.View(3, 20, 5, 29)
    sub = G[3:20, 5:29]
    spec = OD.Spec(OD.MC, OD.MR, V.ColAlign(), V.RowAlign(), *sub.shape)
    assert np.array_equal(V.get_local(), OD.local_of(sub, spec, og, grid.Rank()))
    for (U2, V2) in OD.PAIRS:
        B = el.DistMatrix(grid, U2, V2, dtype)
        el.Transpose(V, B)
        spec = OD.Spec(U2, V2, B.ColAlign(), B.RowAlign(), sub.shape[1], sub.shape[0])
        assert np.array_equal(B.get_local(), OD.local_of(np.asfortranarray(sub.T), spec, og, grid.Rank()))


def check_contractions(grid, dtype=torch.float64) -> None:
    """AxpyContract / Contract / TransposeAxpyContract: every rank holds a DIFFERENT partial sum;
    the result is the oracle's sum over ranks (order-dependent rounding -> tolerance 4 ulp * p)."""
    og = _grid_oracle(grid)
    npdt = NPDT[dtype]
    p, me = og.p, grid.Rank()
    m, n = 19, 14
    rng = np.random.default_rng(2)
    eps = np.finfo(npdt).eps
    cases = [((OD.MC, OD.STAR), (OD.MC, OD.MR)), ((OD.STAR, OD.MR), (OD.MC, OD.MR)),
             ((OD.STAR, OD.STAR), (OD.MC, OD.MR)), ((OD.MR, OD.STAR), (OD.MR, OD.MC)),
             ((OD.STAR, OD.MC), (OD.MR, OD.MC)), ((OD.MC, OD.MR), (OD.MC, OD.MR))]
    for (UA, VA), (UB, VB) in cases:
        for aligned in (True, False):
            caB, raB = int(rng.integers(og.stride(UB))), int(rng.integers(og.stride(VB)))
            caA = caB if (UA == UB and aligned) else int(rng.integers(og.stride(UA)))
            raA = raB if (VA == VB and aligned) else int(rng.integers(og.stride(VA)))
            specA = OD.Spec(UA, VA, caA, raA, m, n)
            parts = [np.asfortranarray(rng.uniform(-1, 1, specA.local_shape(og, q)).astype(npdt)) for q in range(p)]
            GB = np.asfortranarray(rng.uniform(-1, 1, (m, n)).astype(npdt))
            A = el.DistMatrix(grid, UA, VA, dtype).Align(caA, raA).Resize(m, n)
            A.set_local(parts[me])
            B = el.DistMatrix(grid, UB, VB, dtype).Align(caB, raB).from_global(GB)
            el.AxpyContract(0.5, A, B)
            oA = OD.VDM(og, specA, [x.copy() for x in parts])
            oB = OD.VDM.from_global(og, GB, UB, VB, caB, raB)
            OD.axpy_contract(npdt(0.5), oA, oB)
            assert np.allclose(B.get_local(), oB.loc[me], rtol=0, atol=4 * p * eps), ((UA, VA), (UB, VB), aligned)
    # Contract and TransposeAxpyContract on the SUMMA_B / SUMMA_A temporaries
    specA = OD.Spec(OD.MR, OD.STAR, 0, 0, n, m)
    parts = [np.asfortranarray(rng.uniform(-1, 1, specA.local_shape(og, q)).astype(npdt)) for q in range(p)]
    A = el.DistMatrix(grid, el.MR, el.STAR, dtype).Align(0, 0).Resize(n, m)
    A.set_local(parts[me])
    S = el.DistMatrix(grid, el.MR, el.MC, dtype)
    el.Contract(A, S)
    total = OD.VDM(og, specA, parts).contributions()
    exp = OD.local_of(total, OD.Spec(OD.MR, OD.MC, S.ColAlign(), S.RowAlign(), n, m), og, me)
    assert np.allclose(S.get_local(), exp, rtol=0, atol=4 * p * eps)
    GC = np.asfortranarray(rng.uniform(-1, 1, (m, n)).astype(npdt))
    Cm = el.DistMatrix(grid, el.MC, el.MR, dtype).Align(1 % og.r, 1 % og.c).from_global(GC)
    el.TransposeAxpyContract(2.0, A, Cm)
    expC = OD.local_of(GC + npdt(2.0) * total.T, OD.Spec(OD.MC, OD.MR, 1 % og.r, 1 % og.c, m, n), og, me)
    assert np.allclose(Cm.get_local(), expC, rtol=0, atol=8 * p * eps)


def gemm_case(grid, dtype, oa, ob, alg, m, n, k, alpha=2.0, beta=3.0, aligns=None, nb=None, dists=None):
    """One El::Gemm call vs the oracle; returns the relative Frobenius error of the global result
    restricted to this rank's entries, and the bound."""
    og = _grid_oracle(grid)
    npdt = NPDT[dtype]
    me = grid.Rank()
    A = OG.uniform_matrix(*((m, k) if oa == 0 else (k, m)), 101, npdt, 0.0, 1.0)
    B = OG.uniform_matrix(*((k, n) if ob == 0 else (n, k)), 102, npdt, 0.0, 1.0)
    C = OG.uniform_matrix(m, n, 103, npdt, 0.0, 1.0)
    aligns = aligns or ((0, 0), (0, 0), (0, 0))
    dists = dists or ((el.MC, el.MR),) * 3
    dA = el.DistMatrix(grid, *dists[0], dtype).Align(aligns[0][0] % og.stride(dists[0][0]), aligns[0][1] % og.stride(dists[0][1])).from_global(A)
    dB = el.DistMatrix(grid, *dists[1], dtype).Align(aligns[1][0] % og.stride(dists[1][0]), aligns[1][1] % og.stride(dists[1][1])).from_global(B)
    dC = el.DistMatrix(grid, *dists[2], dtype).Align(aligns[2][0] % og.stride(dists[2][0]), aligns[2][1] % og.stride(dists[2][1])).from_global(C)
    if nb:
        el.PushBlocksizeStack(nb)
    try:
        el.Gemm(oa, ob, alpha, dA, dB, beta, dC, alg)
    finally:
        if nb:
            el.PopBlocksizeStack()
    got = dC.get_local()
    ref = OG.gemm_global(oa, ob, alpha, A, B, beta, C, np.longdouble)
    spec = OD.Spec(dists[2][0], dists[2][1], dC.ColAlign(), dC.RowAlign(), m, n)
    ref_loc = OD.local_of(ref, spec, og, me)
    # error measured against the norm of the whole result (the reference's Frobenius criterion)
    num = float(np.linalg.norm(got.astype(np.longdouble) - ref_loc))
    den = float(np.linalg.norm(ref)) / np.sqrt(og.p)
    return num / max(den, 1e-300), OG.gemm_tolerance(npdt, k)


def check_gemm_variants(grid, dtype, shapes=((70, 50, 90),), algs=(el.GEMM_SUMMA_A, el.GEMM_SUMMA_B, el.GEMM_SUMMA_C, el.GEMM_SUMMA_DOT, el.GEMM_DEFAULT),
                        nb=16) -> int:
    og = _grid_oracle(grid)
    cnt = 0
    for (m, n, k) in shapes:
        for oa, ob in itertools.product((0, 1), (0, 1)):
            for alg in algs:
                for al in (((0, 0), (0, 0), (0, 0)), ((1, 2), (0, 1), (1, 0))):
                    err, tol = gemm_case(grid, dtype, oa, ob, alg, m, n, k, aligns=al, nb=nb)
                    assert err <= tol, f"Gemm oa={oa} ob={ob} alg={alg} {m}x{n}x{k} aligns={al}: {err:.3e} > {tol:.3e}"
                    cnt += 1
    return cnt


def check_gemm_proxies(grid, dtype) -> None:
    """Inputs that are not [MC,MR] go through the read / read-write proxies (Proxy.hpp:236-320)."""
    combos = [((el.STAR, el.VC), (el.STAR, el.STAR), (el.MC, el.MR)),
              ((el.VC, el.STAR), (el.MR, el.MC), (el.STAR, el.VR)),
              ((el.STAR, el.STAR), (el.MC, el.MR), (el.MR, el.MC))]
    for d in combos:
        for alg in (el.GEMM_DEFAULT, el.GEMM_SUMMA_C, el.GEMM_SUMMA_DOT):
            err, tol = gemm_case(grid, dtype, 0, 0, alg, 45, 37, 64, dists=d, nb=16)
            assert err <= tol, (d, alg, err, tol)


def check_gemm_host(grid, dtype, m=60, k=100) -> int:
    """el.GemmHost (elb200_dist_gemm_host_*): operands in HOST memory, streamed in column blocks.
    Sizes give eight blocks on this grid (six of two units, a tapered one of one unit, a ragged last one; four block slots, so
    slots are reused), all orientations, beta = 0 with NaNs in
    C (must not propagate), k = 0 (pure scaling), padded leading dimensions."""
    import math
    og = _grid_oracle(grid)
    npdt = NPDT[dtype]
    me = grid.Rank()
    unit = (og.r * og.c // math.gcd(og.r, og.c)) * 256
    n = 13 * unit + 77  # 6 blocks of 2 units + the tapered tail (1 unit, then the ragged 77): more blocks than block slots
    zero = lambda h, w: OD.Spec(OD.MC, OD.MR, 0, 0, h, w)  # noqa: E731
    cnt = 0
    for oa, ob in itertools.product((0, 1), (0, 1)):
        for (alpha, beta, kk) in ((2.0, 3.0, k), (0.5, 0.0, k), (1.0, -0.5, 0)):
            A = OG.uniform_matrix(*((m, kk) if oa == 0 else (kk, m)), 201, npdt, 0.0, 1.0)
            B = OG.uniform_matrix(*((kk, n) if ob == 0 else (n, kk)), 202, npdt, 0.0, 1.0)
            C0 = OG.uniform_matrix(m, n, 203, npdt, 0.0, 1.0)
            ref = OG.gemm_global(oa, ob, alpha, A, B, beta, C0, np.longdouble)

            def host(G, pad):
                loc = OD.local_of(G, zero(*G.shape), og, me)
                ld = max(loc.shape[0], 1) + pad
                buf = _pin(torch.zeros((max(loc.shape[1], 1), ld), dtype=dtype))
                view = buf.t()[:loc.shape[0], :loc.shape[1]]  # column-major window, leading dimension ld
                view.copy_(torch.from_numpy(np.ascontiguousarray(loc)))
                return buf, view, ld
            bA, _, lda = host(A, 3)
            bB, _, ldb = host(B, 0)
            bC, vC, ldc = host(C0, 5)
            if beta == 0.0:
                vC.fill_(float("nan"))
            el.GemmHost(grid, oa, ob, m, n, kk, alpha, bA, lda, bB, ldb, beta, bC, ldc)
            got = vC.numpy()
            ref_loc = OD.local_of(ref, zero(m, n), og, me)
            num = float(np.linalg.norm(got.astype(np.longdouble) - ref_loc))
            den = float(np.linalg.norm(ref)) / np.sqrt(og.p)
            assert np.isfinite(got).all(), (oa, ob, alpha, beta, kk)
            assert num / max(den, 1e-300) <= OG.gemm_tolerance(npdt, max(kk, 1)), (oa, ob, alpha, beta, kk, num / den)
            cnt += 1
    return cnt


def check_gemm_pipeline(grid, dtype=torch.float64) -> int:
    """SUMMA with MANY panels per call (k = 1300 is several panels once ELB200_PANEL_K is 512 or less;
    with the default width it is one): exercises the two-slot gather / GEMM pipeline, the ramp widths,
    slot hand-back and the zero-copy direct panels (NN, aligned) as well as the staged ones
    (misaligned / transposed), and the multi-panel loops of SUMMA A / B."""
    og = _grid_oracle(grid)
    cnt = 0
    mis = ((1, 2), (0, 1), (1, 0))
    for (oa, ob, alg, al) in ((0, 0, el.GEMM_SUMMA_C, None), (0, 0, el.GEMM_SUMMA_C, mis), (1, 0, el.GEMM_SUMMA_C, None),
                              (0, 1, el.GEMM_SUMMA_C, mis), (1, 1, el.GEMM_SUMMA_C, None)):
        err, tol = gemm_case(grid, dtype, oa, ob, alg, 150, 130, 1300, alpha=0.5, beta=-0.5, aligns=al)
        assert err <= tol, (oa, ob, alg, al, err, tol)
        cnt += 1
    for (oa, ob, alg) in ((0, 0, el.GEMM_SUMMA_A), (1, 0, el.GEMM_SUMMA_A), (0, 1, el.GEMM_SUMMA_B), (1, 1, el.GEMM_SUMMA_B)):
        m, n, k = (90, 1300, 110) if alg == el.GEMM_SUMMA_A else (1300, 90, 110)
        err, tol = gemm_case(grid, dtype, oa, ob, alg, m, n, k, alpha=0.5, beta=-0.5, aligns=mis)
        assert err <= tol, (oa, ob, alg, err, tol)
        cnt += 1
    return cnt


def check_gemm_ragged_slot_boundary(grid, dtype=torch.float64) -> int:
    """SUMMA-C (NN, aligned) with m, n NOT divisible by the grid, so local heights / widths differ by one across ranks.
    Every rank must derive the same panel width and take the same branch (zero-copy direct panels or the staged
    fallback) — the decision is made from the largest local extents on the grid. tests/test_hostsim.py runs this with an
    arena slot sized so that a rank-local decision would split the grid (ELB200_ARENA_KB=311: 75 local rows fit 512
    panel columns, 76 do not)."""
    cnt = 0
    for (m, n, k) in ((151, 120, 1300), (151, 121, 1300), (150, 121, 700)):
        err, tol = gemm_case(grid, dtype, 0, 0, el.GEMM_SUMMA_C, m, n, k, alpha=0.5, beta=-0.5)
        assert err <= tol, (m, n, k, err, tol)
        cnt += 1
    return cnt


def check_trapezoids(grid, dtype=torch.float64) -> int:
    """ScaleTrapezoid / MakeTrapezoidal / AxpyTrapezoid / MakeSymmetric (level-1 trapezoid headers) on
    several distributions, alignments and offsets: data movement and exact products only (alpha is a
    power of two), so every result must equal the oracle's bit for bit."""
    og = _grid_oracle(grid)
    npdt = NPDT[dtype]
    me = grid.Rank()
    rng = np.random.default_rng(11)
    cnt = 0
    for (U, V), (h, w) in (((OD.MC, OD.MR), (23, 23)), ((OD.MC, OD.MR), (17, 29)), ((OD.VC, OD.STAR), (21, 13)),
                           ((OD.STAR, OD.VR), (14, 26)), ((OD.MR, OD.MC), (19, 19)), ((OD.STAR, OD.STAR), (9, 12))):
        for uplo in (el.LOWER, el.UPPER):
            for offset in (0, 1, -2, 5):
                ca, ra = int(rng.integers(og.stride(U))), int(rng.integers(og.stride(V)))
                G = np.asfortranarray(rng.uniform(-1, 1, (h, w)).astype(npdt))
                X = np.asfortranarray(rng.uniform(-1, 1, (h, w)).astype(npdt))
                mask = O3.trapezoid_mask(h, w, uplo, offset)
                spec = OD.Spec(U, V, ca, ra, h, w)
                A = el.DistMatrix(grid, U, V, dtype).Align(ca, ra).from_global(G)
                el.ScaleTrapezoid(-0.5, uplo, A, offset)
                exp = np.where(mask, npdt(-0.5) * G, G)
                assert np.array_equal(A.get_local(), OD.local_of(exp, spec, og, me)), ("ScaleTrapezoid", U, V, uplo, offset)
                el.ScaleTrapezoid(0.0, uplo, A, offset)
                exp = np.where(mask, npdt(0), G)
                assert np.array_equal(A.get_local(), OD.local_of(exp, spec, og, me)), ("ScaleTrapezoid 0", U, V, uplo, offset)
                A.from_global(G)
                el.MakeTrapezoidal(uplo, A, offset)
                assert np.array_equal(A.get_local(), OD.local_of(np.where(mask, G, npdt(0)), spec, og, me)), ("MakeTrapezoidal", U, V, uplo, offset)
                # AxpyTrapezoid with X in the same layout, then in another distribution / alignment
                A.from_global(G)
                Xd = el.DistMatrix(grid, U, V, dtype).Align(ca, ra).from_global(X)
                el.AxpyTrapezoid(uplo, 2.0, Xd, A, offset)
                exp = np.where(mask, G + npdt(2.0) * X, G)
                assert np.array_equal(A.get_local(), OD.local_of(exp, spec, og, me)), ("AxpyTrapezoid", U, V, uplo, offset)
                A.from_global(G)
                Xo = el.DistMatrix(grid, OD.MC, OD.MR, dtype).Align(1 % og.r, 1 % og.c).from_global(X)
                el.AxpyTrapezoid(uplo, 2.0, Xo, A, offset)
                assert np.array_equal(A.get_local(), OD.local_of(exp, spec, og, me)), ("AxpyTrapezoid redistributed", U, V, uplo, offset)
                cnt += 5
    for (U, V) in ((OD.MC, OD.MR), (OD.MR, OD.MC), (OD.VC, OD.STAR), (OD.STAR, OD.STAR)):
        for uplo in (el.LOWER, el.UPPER):
            n = 27
            ca, ra = int(rng.integers(og.stride(U))), int(rng.integers(og.stride(V)))
            G = np.asfortranarray(rng.uniform(-1, 1, (n, n)).astype(npdt))
            A = el.DistMatrix(grid, U, V, dtype).Align(ca, ra).from_global(G)
            el.MakeSymmetric(uplo, A)
            tri = np.where(O3.trapezoid_mask(n, n, uplo, 0), G, npdt(0))
            strict = np.where(O3.trapezoid_mask(n, n, uplo, -1 if uplo == el.LOWER else 1), G, npdt(0))
            exp = np.asfortranarray(tri + strict.T)
            assert np.array_equal(A.get_local(), OD.local_of(exp, OD.Spec(U, V, ca, ra, n, n), og, me)), ("MakeSymmetric", U, V, uplo)
            cnt += 1
    return cnt


def _tri_err(got_loc, ref, spec, og, me, npdt, k):
    ref_loc = OD.local_of(ref, spec, og, me)
    num = float(np.linalg.norm(got_loc.astype(np.longdouble) - ref_loc))
    den = float(np.linalg.norm(ref)) / np.sqrt(og.p)
    return num / max(den, 1e-300), OG.gemm_tolerance(npdt, max(k, 1))


def check_level3(grid, dtype=torch.float64, trrk_blocksize=8) -> int:
    """LocalTrrk / Trrk / Syrk / Herk / Symm / Hemm against the global definition (oracle.level3) within the
    Gemm bound; the triangle of C that the routine must not touch has to come back bit-identical.
    LocalTrrkBlocksize is lowered so the recursion of LocalTrrk (Trrk/Local.hpp:782-830) runs several
    levels at these sizes; nb = 16 gives several k panels only where the engine's panel width allows
    (ELB200_PANEL_K)."""
    og = _grid_oracle(grid)
    npdt = NPDT[dtype]
    me = grid.Rank()
    cnt = 0
    old_bs = el.LocalTrrkBlocksize()
    el.SetLocalTrrkBlocksize(trrk_blocksize)
    try:
        n, k = 67, 45
        C0 = OG.uniform_matrix(n, n, 303, npdt, 0.0, 1.0)
        # LocalTrrk on explicitly built panels, aligned with a misaligned C
        for uplo in (el.LOWER, el.UPPER):
            for ta, tb in itertools.product((0, 1), (0, 1)):
                ca, ra = 1 % og.r, 2 % og.c
                A = OG.uniform_matrix(*((k, n) if ta else (n, k)), 301, npdt, 0.0, 1.0)
                B = OG.uniform_matrix(*((n, k) if tb else (k, n)), 302, npdt, 0.0, 1.0)
                dA = (el.DistMatrix(grid, el.STAR, el.MC, dtype).Align(0, ca) if ta else el.DistMatrix(grid, el.MC, el.STAR, dtype).Align(ca, 0)).from_global(A)
                dB = (el.DistMatrix(grid, el.MR, el.STAR, dtype).Align(ra, 0) if tb else el.DistMatrix(grid, el.STAR, el.MR, dtype).Align(0, ra)).from_global(B)
                dC = el.DistMatrix(grid, el.MC, el.MR, dtype).Align(ca, ra).from_global(C0)
                el.LocalTrrk(uplo, ta, tb, 2.0, dA, dB, 3.0, dC)
                ref = O3.trrk_global(uplo, ta, tb, 2.0, A, B, 3.0, C0)
                spec = OD.Spec(OD.MC, OD.MR, ca, ra, n, n)
                got = dC.get_local()
                err, tol = _tri_err(got, ref, spec, og, me, npdt, k)
                assert err <= tol, ("LocalTrrk", uplo, ta, tb, err, tol)
                keep = ~OD.local_of(np.asfortranarray(O3.trapezoid_mask(n, n, uplo, 0)), spec, og, me).astype(bool)
                assert np.array_equal(got[keep], OD.local_of(C0, spec, og, me)[keep]), ("LocalTrrk touched the other triangle", uplo, ta, tb)
                cnt += 1
        # Trrk: all orientations, aligned / misaligned [MC,MR] operands, and operands that need proxies
        layouts = [(((el.MC, el.MR),) * 3, ((0, 0), (0, 0), (0, 0))), (((el.MC, el.MR),) * 3, ((1, 2), (0, 1), (1, 0))),
                   (((el.VC, el.STAR), (el.STAR, el.VR), (el.MR, el.MC)), ((1, 0), (0, 1), (0, 1)))]
        for uplo in (el.LOWER, el.UPPER):
            for ta, tb in itertools.product((0, 1), (0, 1)):
                for dists, aligns in layouts:
                    A = OG.uniform_matrix(*((k, n) if ta else (n, k)), 311, npdt, 0.0, 1.0)
                    B = OG.uniform_matrix(*((n, k) if tb else (k, n)), 312, npdt, 0.0, 1.0)
                    mk = lambda d, al, G: el.DistMatrix(grid, d[0], d[1], dtype).Align(al[0] % og.stride(d[0]), al[1] % og.stride(d[1])).from_global(G)  # noqa: E731
                    dA, dB, dC = mk(dists[0], aligns[0], A), mk(dists[1], aligns[1], B), mk(dists[2], aligns[2], C0)
                    el.PushBlocksizeStack(16)
                    try:
                        el.Trrk(uplo, ta, tb, 0.5, dA, dB, -0.5, dC)
                    finally:
                        el.PopBlocksizeStack()
                    ref = O3.trrk_global(uplo, ta, tb, 0.5, A, B, -0.5, C0)
                    spec = OD.Spec(dists[2][0], dists[2][1], dC.ColAlign(), dC.RowAlign(), n, n)
                    got = dC.get_local()
                    err, tol = _tri_err(got, ref, spec, og, me, npdt, k)
                    assert err <= tol, ("Trrk", uplo, ta, tb, dists, err, tol)
                    keep = ~OD.local_of(np.asfortranarray(O3.trapezoid_mask(n, n, uplo, 0)), spec, og, me).astype(bool)
                    assert np.array_equal(got[keep], OD.local_of(C0, spec, og, me)[keep]), ("Trrk touched the other triangle", uplo, ta, tb)
                    cnt += 1
        # Syrk / Herk: C = alpha A A^T + beta C (NORMAL) or alpha A^T A + beta C (TRANSPOSE)
        for uplo in (el.LOWER, el.UPPER):
            for trans in (0, 1):
                for fn in (el.Syrk, el.Herk):
                    A = OG.uniform_matrix(*((k, n) if trans else (n, k)), 321, npdt, 0.0, 1.0)
                    dA = el.DistMatrix(grid, el.MC, el.MR, dtype).Align(1 % og.r, 0).from_global(A)
                    dC = el.DistMatrix(grid, el.MC, el.MR, dtype).Align(0, 1 % og.c).from_global(C0)
                    fn(uplo, trans, 2.0, dA, 3.0, dC)
                    ref = O3.syrk_global(uplo, trans, 2.0, A, 3.0, C0)
                    spec = OD.Spec(OD.MC, OD.MR, 0, 1 % og.c, n, n)
                    err, tol = _tri_err(dC.get_local(), ref, spec, og, me, npdt, k)
                    assert err <= tol, ("Syrk", uplo, trans, err, tol)
                    cnt += 1
        # Symm / Hemm: only the `uplo` triangle of A may be referenced -> poison the other one
        m2 = 41
        for side in (el.LEFT, el.RIGHT):
            for uplo in (el.LOWER, el.UPPER):
                A = OG.uniform_matrix(n, n, 331, npdt, 0.0, 1.0)
                Apoison = np.where(O3.trapezoid_mask(n, n, uplo, 0), A, npdt(np.nan))
                Bm = OG.uniform_matrix(*((n, m2) if side == el.LEFT else (m2, n)), 332, npdt, 0.0, 1.0)
                Cm = OG.uniform_matrix(*Bm.shape, 333, npdt, 0.0, 1.0)
                dA = el.DistMatrix(grid, el.MC, el.MR, dtype).Align(1 % og.r, 1 % og.c).from_global(np.asfortranarray(Apoison))
                dB = el.DistMatrix(grid, el.MC, el.MR, dtype).from_global(Bm)
                dC = el.DistMatrix(grid, el.MC, el.MR, dtype).Align(0, 1 % og.c).from_global(Cm)
                (el.Symm if uplo == el.LOWER else el.Hemm)(side, uplo, 2.0, dA, dB, 3.0, dC)
                ref = O3.symm_global(side, uplo, 2.0, A, Bm, 3.0, Cm)
                spec = OD.Spec(OD.MC, OD.MR, 0, 1 % og.c, *Cm.shape)
                got = dC.get_local()
                assert np.isfinite(got).all(), ("Symm read the unreferenced triangle", side, uplo)
                err, tol = _tri_err(got, ref, spec, og, me, npdt, n)
                assert err <= tol, ("Symm", side, uplo, err, tol)
                cnt += 1
    finally:
        el.SetLocalTrrkBlocksize(old_bs)
    return cnt


def check_level3_golden(grid, dtype=torch.float64) -> int:
    """The REFERENCE'S OWN per-rank outputs (tests/golden/ref_level3.npz: its CPU build run multi-rank on
    this grid shape) against this library: trapezoid routines and MakeSymmetric bit for bit, LocalTrrk /
    Trrk / Syrk / Symm / Trsm within the fp64 Gemm bound.  Grids the reference was not run on return 0."""
    import os
    from oracle.make_golden import lcg_matrix
    assert dtype == torch.float64
    og = _grid_oracle(grid)
    gold = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_level3.npz"))
    tag = f"r{og.r}c{og.c}"
    me = grid.Rank()
    if f"{tag}_q{me}_L" not in gold:
        return 0
    blob = gold[f"{tag}_q{me}_bin"]
    ca, ra = 1 % og.r, 2 % og.c
    n, k, m2 = 37, 23, 29
    f64 = np.float64
    DM = lambda U, V, a, b, G: el.DistMatrix(grid, U, V, dtype).Align(a, b).from_global(np.asfortranarray(G))  # noqa: E731
    old_bs = el.LocalTrrkBlocksize()
    el.SetLocalTrrkBlocksize(4)
    el.PushBlocksizeStack(16)
    cnt = 0
    try:
        for (kind, uplo, p1, p2, caC, raC, cs, rs, lh, lw, off) in (tuple(int(x) for x in r) for r in gold[f"{tag}_q{me}_L"]):
            if kind in (0, 1):
                R = DM(el.MC, el.MR, ca, ra, lcg_matrix(13, 11, 7001, f64))
                (el.ScaleTrapezoid(-0.5, uplo, R, p1) if kind == 0 else el.MakeTrapezoidal(uplo, R, p1))
            elif kind == 2:
                R = DM(el.MC, el.MR, ca, ra, lcg_matrix(13, 13, 7001, f64))
                el.AxpyTrapezoid(uplo, 2.0, DM(el.MC, el.MR, ca, ra, lcg_matrix(13, 13, 7004, f64)), R, p1)
            elif kind == 3:
                R = DM(el.MC, el.MR, ca, ra, lcg_matrix(13, 13, 7001, f64))
                el.MakeSymmetric(uplo, R)
            elif kind == 4:
                R = DM(el.MC, el.MR, ca, ra, lcg_matrix(n, n, 7003, f64))
                A = DM(el.STAR, el.MC, 0, ca, lcg_matrix(k, n, 7001, f64)) if p1 else DM(el.MC, el.STAR, ca, 0, lcg_matrix(n, k, 7001, f64))
                B = DM(el.MR, el.STAR, ra, 0, lcg_matrix(n, k, 7002, f64)) if p2 else DM(el.STAR, el.MR, 0, ra, lcg_matrix(k, n, 7002, f64))
                el.LocalTrrk(uplo, p1, p2, 2.0, A, B, 3.0, R)
            elif kind == 5:
                A = DM(el.MC, el.MR, ca, 0, lcg_matrix(*((k, n) if p1 else (n, k)), 7001, f64))
                B = DM(el.MC, el.MR, 0, ra, lcg_matrix(*((n, k) if p2 else (k, n)), 7002, f64))
                R = DM(el.MC, el.MR, ca, ra, lcg_matrix(n, n, 7003, f64))
                el.Trrk(uplo, p1, p2, 2.0, A, B, 3.0, R)
            elif kind == 6:
                A = DM(el.MC, el.MR, ca, 0, lcg_matrix(*((k, n) if p1 else (n, k)), 7001, f64))
                R = DM(el.MC, el.MR, 0, ra, lcg_matrix(n, n, 7003, f64))
                el.Syrk(uplo, p1, 2.0, A, 3.0, R)
            elif kind == 7:
                shape = (m2, n) if p1 else (n, m2)
                A = DM(el.MC, el.MR, ca, ra, lcg_matrix(n, n, 7001, f64))
                B = DM(el.MC, el.MR, 0, 0, lcg_matrix(*shape, 7002, f64))
                R = DM(el.MC, el.MR, 0, ra, lcg_matrix(*shape, 7003, f64))
                el.Symm(p1, uplo, 2.0, A, B, 3.0, R)
            else:
                side, orient, diag = p1, p2 // 2, p2 % 2
                Ag = lcg_matrix(n, n, 7001, f64)
                Ag = Ag * (1.0 / float(n)) if diag else Ag + float(n) * np.eye(n)
                A = DM(el.MC, el.MR, ca, 0, Ag)
                R = DM(el.MC, el.MR, 0, ra, lcg_matrix(*((m2, n) if side else (n, m2)), 7002, f64))
                el.Trsm(side, uplo, orient, diag, 2.0, A, R)
            assert (R.ColAlign(), R.RowAlign(), R.ColShift(), R.RowShift()) == (caC, raC, cs, rs), (tag, kind, uplo, p1, p2)
            got = R.get_local()
            ref = blob[off:off + lh * lw].reshape((lh, lw), order="F")
            assert got.shape == ref.shape, (tag, kind, uplo, p1, p2)
            if kind <= 3:
                assert np.array_equal(got, ref), (tag, me, kind, uplo, p1, p2)
            else:
                assert OG.frobenius_rel_err(got, ref) <= OG.gemm_tolerance(f64, k), (tag, me, kind, uplo, p1, p2)
            cnt += 1
    finally:
        el.PopBlocksizeStack()
        el.SetLocalTrrkBlocksize(old_bs)
    return cnt


def check_dist_golden(grid, dtype=torch.float64) -> int:
    """The REFERENCE'S OWN per-rank outputs of tests/golden/ref_dist.npz (its CPU build run multi-rank on
    this grid shape: 121 (U,V)->(U2,V2) redistributions, 23 [CIRC,CIRC] gather / scatter cases, 20 El::Gemm
    runs) against this library, rank by rank: redistributions bit for bit, Gemm within the fp64 bound."""
    import os
    from oracle.make_golden import lcg_matrix
    assert dtype == torch.float64
    og = _grid_oracle(grid)
    gold = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_dist.npz"))
    tag = f"r{og.r}c{og.c}"
    me = grid.Rank()
    if f"{tag}_q{me}_R" not in gold:
        return 0
    blob = gold[f"{tag}_q{me}_bin"]
    ELD = {0: el.MC, 2: el.MR, 3: el.VC, 4: el.VR, 5: el.STAR, 6: el.CIRC}  # El::Dist enum value -> ours
    G = np.asfortranarray(np.arange(13 * 11, dtype=np.float64).reshape(13, 11) + 1.0)
    cnt = 0
    recs = [tuple(int(x) for x in r) + (0, 0) for r in gold[f"{tag}_q{me}_R"]] + [tuple(int(x) for x in r) for r in gold[f"{tag}_q{me}_Q"]]
    for (U, V, U2, V2, caA, raA, caB, raB, csB, rsB, lh, lw, off, rootA, rootB) in recs:
        A = el.DistMatrix(grid, ELD[U], ELD[V], dtype)
        if U == 6:
            A.SetRoot(rootA)
        A.Align(caA, raA).from_global(G)
        B = el.DistMatrix(grid, ELD[U2], ELD[V2], dtype)
        if U2 == 6:
            B.SetRoot(rootB)
        B.Align(caB, raB)
        el.Copy(A, B)
        got = B.get_local()
        assert got.shape == (lh, lw), (tag, me, U, V, U2, V2, got.shape, (lh, lw))
        if lh and lw:
            assert (B.ColShift(), B.RowShift()) == (csB, rsB)
        assert np.array_equal(got, blob[off:off + lh * lw].reshape((lh, lw), order="F")), (tag, me, U, V, U2, V2)
        cnt += 1
    aligns = ((1 % og.r, 2 % og.c), (0, 1 % og.c), (1 % og.r, 0))
    for (cid, oa, ob, alg, m, n, k, nb, caC, raC, cs, rs, lh, lw, off) in (tuple(int(x) for x in r) for r in gold[f"{tag}_q{me}_G"]):
        mk = lambda al, M: el.DistMatrix(grid, el.MC, el.MR, dtype).Align(*al).from_global(np.asfortranarray(M))  # noqa: E731
        dA = mk(aligns[0], lcg_matrix(*((k, m) if oa else (m, k)), 3 * cid + 0, np.float64))
        dB = mk(aligns[1], lcg_matrix(*((n, k) if ob else (k, n)), 3 * cid + 1, np.float64))
        dC = mk(aligns[2], lcg_matrix(m, n, 3 * cid + 2, np.float64))
        el.PushBlocksizeStack(nb)
        try:
            el.Gemm(oa, ob, 2.0, dA, dB, 3.0, dC, alg)
        finally:
            el.PopBlocksizeStack()
        got = dC.get_local()
        ref = blob[off:off + lh * lw].reshape((lh, lw), order="F")
        assert got.shape == ref.shape and (dC.ColShift(), dC.RowShift()) == (cs, rs)
        assert OG.frobenius_rel_err(got, ref) <= OG.gemm_tolerance(np.float64, k), (tag, me, cid)
        cnt += 1
    return cnt


def check_trsm(grid, dtype=torch.float64) -> int:
    """LocalTrsm and Trsm for every side / uplo / orientation / diag against the substitution in long double
    (oracle.level3.trsm_global). A is well conditioned (unit-scale entries + n on the diagonal), the triangle
    that must not be referenced is poisoned with NaN. Blocksize 8 -> blocks of 64: several diagonal blocks and
    update GEMMs at n = 150; m2 = 45 right-hand sides; B misaligned; one case with operands that need proxies."""
    og = _grid_oracle(grid)
    npdt = NPDT[dtype]
    me = grid.Rank()
    n, m2 = 150, 45
    cnt = 0
    A0 = OG.uniform_matrix(n, n, 401, npdt, 0.0, 1.0) + npdt(n) * np.eye(n, dtype=npdt)
    el.PushBlocksizeStack(8)
    try:
        for side, uplo, t, diag in itertools.product((el.LEFT, el.RIGHT), (el.LOWER, el.UPPER), (0, 1), (el.NON_UNIT, el.UNIT)):
            keep = O3.trapezoid_mask(n, n, uplo, 0)
            Apoison = np.asfortranarray(np.where(keep, A0, npdt(np.nan)))
            if diag == el.UNIT:   # the diagonal must not be read either
                Apoison[np.arange(n), np.arange(n)] = np.nan
            # UNIT makes the off-diagonal part dominant unless it is scaled down
            Ause = np.asfortranarray(np.where(np.eye(n, dtype=bool), Apoison, Apoison / npdt(n))) if diag == el.UNIT else Apoison
            Aref = np.where(keep, A0 / (npdt(n) if diag == el.UNIT else npdt(1)), 0)
            B0 = OG.uniform_matrix(*((n, m2) if side == el.LEFT else (m2, n)), 402, npdt, 0.0, 1.0)
            ref = O3.trsm_global(side, uplo, t, diag, 2.0, Aref, B0)
            proxies = (side == el.LEFT and uplo == el.UPPER and t == 1)
            dA = el.DistMatrix(grid, *((el.VC, el.STAR) if proxies else (el.MC, el.MR)), dtype).Align(1 % og.r if not proxies else 1 % og.p, 0).from_global(Ause)
            dB = el.DistMatrix(grid, *((el.STAR, el.VR) if proxies else (el.MC, el.MR)), dtype).Align(0, 2 % (og.p if proxies else og.c)).from_global(B0)
            el.Trsm(side, uplo, t, diag, 2.0, dA, dB)
            spec = OD.Spec(OD.STAR if proxies else OD.MC, OD.VR if proxies else OD.MR, dB.ColAlign(), dB.RowAlign(), *B0.shape)
            got = dB.get_local()
            assert np.isfinite(got).all(), ("Trsm read the unreferenced triangle / diagonal", side, uplo, t, diag)
            err, tol = _tri_err(got, ref, spec, og, me, npdt, n)
            assert err <= 4 * tol, ("Trsm", side, uplo, t, diag, err, tol)
            cnt += 1
            # LocalTrsm: A replicated, every rank solves its own right-hand sides
            for nl in ((40, 150) if (t == 0 or diag == el.NON_UNIT) else ()):   # 150: three levels of the local recursion
                Al = np.asfortranarray(Ause[:nl, :nl])
                Bl0 = OG.uniform_matrix(*((nl, m2) if side == el.LEFT else (m2, nl)), 403, npdt, 0.0, 1.0)
                dAl = el.DistMatrix(grid, el.STAR, el.STAR, dtype).from_global(Al)
                dX = el.DistMatrix(grid, *((el.STAR, el.VR) if side == el.LEFT else (el.VC, el.STAR)), dtype).from_global(Bl0)
                el.LocalTrsm(side, uplo, t, diag, 0.5, dAl, dX)
                refl = O3.trsm_global(side, uplo, t, diag, 0.5, Aref[:nl, :nl], Bl0)
                specl = OD.Spec(OD.STAR if side == el.LEFT else OD.VC, OD.VR if side == el.LEFT else OD.STAR, 0, 0, *Bl0.shape)
                got = dX.get_local()
                assert np.isfinite(got).all(), ("LocalTrsm read the unreferenced triangle / diagonal", side, uplo, t, diag, nl)
                errl, toll = _tri_err(got, refl, specl, og, me, npdt, nl)
                assert errl <= 4 * toll, ("LocalTrsm", side, uplo, t, diag, nl, errl, toll)
                cnt += 1
    finally:
        el.PopBlocksizeStack()
    return cnt


def check_fuzz(grid, dtype=torch.float64, iters=None, seed=None) -> int:
    """Randomised sweep (same random stream on every rank): Gemm with random sizes — including 0, 1 and sizes
    smaller than the grid —, orientations, algorithms, alignments, blocksizes and operand distributions;
    redistributions of random shapes between random pairs; Trrk / Syrk / Trsm with random shapes."""
    import os
    iters = int(os.environ.get("ELB_FUZZ_ITERS", "40")) if iters is None else iters
    seed = int(os.environ.get("ELB_FUZZ_SEED", "1234")) if seed is None else seed
    og = _grid_oracle(grid)
    npdt = NPDT[dtype]
    me = grid.Rank()
    rng = np.random.default_rng(seed)
    dims = [0, 1, 2, 3, 5, 8, 13, 31, 64, 77]
    algs = (el.GEMM_DEFAULT, el.GEMM_SUMMA_A, el.GEMM_SUMMA_B, el.GEMM_SUMMA_C, el.GEMM_SUMMA_DOT, el.GEMM_SUMMA_C_MS)
    cnt = 0
    for it in range(iters):
        m, n, k = (int(rng.choice(dims)) for _ in range(3))
        oa, ob = int(rng.integers(2)), int(rng.integers(2))
        alg = algs[int(rng.integers(len(algs)))]
        nb = int(rng.choice([1, 4, 16, 128]))
        dists = tuple(OD.PAIRS[int(rng.integers(len(OD.PAIRS)))] if rng.random() < 0.3 else (el.MC, el.MR) for _ in range(3))
        aligns = tuple((int(rng.integers(8)), int(rng.integers(8))) for _ in range(3))
        alpha, beta = float(rng.choice([0.0, 1.0, -0.5, 2.0])), float(rng.choice([0.0, 1.0, 3.0]))
        err, tol = gemm_case(grid, dtype, oa, ob, alg, m, n, k, alpha=alpha, beta=beta, aligns=aligns, nb=nb, dists=dists)
        assert err <= tol, ("fuzz Gemm", it, (m, n, k), oa, ob, alg, nb, dists, aligns, alpha, beta, err, tol)
        # a random redistribution of a random shape
        h, w = int(rng.choice(dims)), int(rng.choice(dims))
        (U, V), (U2, V2) = OD.PAIRS[int(rng.integers(len(OD.PAIRS)))], OD.PAIRS[int(rng.integers(len(OD.PAIRS)))]
        ca, ra = int(rng.integers(og.stride(U))), int(rng.integers(og.stride(V)))
        ca2, ra2 = int(rng.integers(og.stride(U2))), int(rng.integers(og.stride(V2)))
        G = np.asfortranarray(rng.uniform(-1, 1, (h, w)).astype(npdt))
        A = el.DistMatrix(grid, U, V, dtype).Align(ca, ra).from_global(G)
        B = el.DistMatrix(grid, U2, V2, dtype).Align(ca2, ra2)
        el.Copy(A, B)
        assert np.array_equal(B.get_local(), OD.local_of(G, OD.Spec(U2, V2, ca2, ra2, h, w), og, me)), ("fuzz Copy", it, (h, w), (U, V), (U2, V2))
        # triangular family on a random size
        nn, kk = int(rng.choice(dims)), int(rng.choice(dims))
        uplo, t = int(rng.integers(2)), int(rng.integers(2))
        A2 = OG.uniform_matrix(*((kk, nn) if t else (nn, kk)), 900 + it, npdt, 0.0, 1.0)
        C2 = OG.uniform_matrix(nn, nn, 950 + it, npdt, 0.0, 1.0)
        dA = el.DistMatrix(grid, el.MC, el.MR, dtype).Align(aligns[0][0] % og.r, aligns[0][1] % og.c).from_global(A2)
        dC = el.DistMatrix(grid, el.MC, el.MR, dtype).Align(aligns[2][0] % og.r, aligns[2][1] % og.c).from_global(C2)
        el.PushBlocksizeStack(nb)
        old_bs = el.LocalTrrkBlocksize()
        el.SetLocalTrrkBlocksize(int(rng.choice([1, 4, 2048])))
        try:
            el.Syrk(uplo, t, alpha, dA, beta, dC)
            ref = O3.syrk_global(uplo, t, alpha, A2, beta, C2)
            spec = OD.Spec(OD.MC, OD.MR, dC.ColAlign(), dC.RowAlign(), nn, nn)
            if nn:
                e2, t2 = _tri_err(dC.get_local(), ref, spec, og, me, npdt, kk)
                assert e2 <= t2, ("fuzz Syrk", it, nn, kk, uplo, t, nb, e2, t2)
            side, diag = int(rng.integers(2)), int(rng.integers(2))
            T0 = OG.uniform_matrix(nn, nn, 970 + it, npdt, 0.0, 1.0)
            T0 = T0 / npdt(max(nn, 1)) if diag else T0 + npdt(nn) * np.eye(nn, dtype=npdt)
            Bs = OG.uniform_matrix(*((nn, kk) if side == el.LEFT else (kk, nn)), 980 + it, npdt, 0.0, 1.0)
            dT = el.DistMatrix(grid, el.MC, el.MR, dtype).Align(aligns[1][0] % og.r, aligns[1][1] % og.c).from_global(np.asfortranarray(T0))
            dB = el.DistMatrix(grid, el.MC, el.MR, dtype).Align(aligns[2][0] % og.r, aligns[2][1] % og.c).from_global(Bs)
            el.Trsm(side, uplo, t, diag, 2.0, dT, dB)
            if nn and kk:
                ref = O3.trsm_global(side, uplo, t, diag, 2.0, T0, Bs)
                spec = OD.Spec(OD.MC, OD.MR, dB.ColAlign(), dB.RowAlign(), *Bs.shape)
                e3, t3 = _tri_err(dB.get_local(), ref, spec, og, me, npdt, nn)
                assert e3 <= 4 * t3, ("fuzz Trsm", it, nn, kk, side, uplo, t, diag, nb, e3, t3)
        finally:
            el.SetLocalTrrkBlocksize(old_bs)
            el.PopBlocksizeStack()
        cnt += 1
    return cnt


def check_level3_errors(grid, dtype=torch.float64) -> int:
    """Argument errors of the level-3 family come back as ELB200_ERR_ARG (the reference throws LogicError:
    "Nonconformal Syrk" Syrk.cpp:108-117, "A must be square" / "Nonconformal Trsm" Trsm.cpp:146-156, "Dist of RHS must
    conform with that of triangle" Trsm.cpp:513-517, EnsureConformal Trrk/Local.hpp:20-48, "Cannot make non-square
    matrix symmetric" MakeSymmetric.hpp:60-61) and leave the operands untouched."""
    og = _grid_oracle(grid)
    npdt = NPDT[dtype]
    me = grid.Rank()
    G = OG.uniform_matrix(12, 12, 501, npdt, 0.0, 1.0)
    R = OG.uniform_matrix(12, 9, 502, npdt, 0.0, 1.0)
    mk = lambda U, V, M: el.DistMatrix(grid, U, V, dtype).from_global(M)  # noqa: E731
    sq, rect = mk(el.MC, el.MR, G), mk(el.MC, el.MR, R)
    star = mk(el.STAR, el.STAR, G)
    cnt = 0

    def must_fail(fn, what):
        nonlocal cnt
        try:
            fn()
        except Exception as e:  # Elb200Error of whichever binding is active
            assert getattr(e, "status", None) == 2, (what, e)
            cnt += 1
            return
        raise AssertionError(f"{what}: no error raised")

    must_fail(lambda: el.Syrk(el.LOWER, el.NORMAL, 1.0, rect, 0.0, rect), "Syrk into a non-square C")
    must_fail(lambda: el.Syrk(el.LOWER, el.TRANSPOSE, 1.0, rect, 0.0, sq), "Syrk with n != C.Height()")
    must_fail(lambda: el.Syrk(2, el.NORMAL, 1.0, rect, 0.0, sq), "Syrk with a bad uplo")
    must_fail(lambda: el.Trrk(el.UPPER, el.NORMAL, el.NORMAL, 1.0, rect, rect, 0.0, sq), "Trrk with nonconformal A, B")
    must_fail(lambda: el.Symm(el.LEFT, el.LOWER, 1.0, rect, rect, 0.0, rect), "Symm with a non-square A")
    must_fail(lambda: el.Symm(el.RIGHT, el.LOWER, 1.0, sq, rect, 0.0, rect), "Symm RIGHT with A of the wrong size")
    must_fail(lambda: el.Trsm(el.LEFT, el.LOWER, el.NORMAL, el.NON_UNIT, 1.0, rect, rect), "Trsm with a non-square A")
    must_fail(lambda: el.Trsm(el.RIGHT, el.LOWER, el.NORMAL, el.NON_UNIT, 1.0, sq, rect), "Trsm RIGHT with A of the wrong size")
    must_fail(lambda: el.LocalTrsm(el.LEFT, el.LOWER, el.NORMAL, el.NON_UNIT, 1.0, sq, rect), "LocalTrsm with a distributed triangle")
    must_fail(lambda: el.LocalTrsm(el.LEFT, el.LOWER, el.NORMAL, el.NON_UNIT, 1.0, star, rect), "LocalTrsm LEFT with X distributed over rows")
    must_fail(lambda: el.MakeSymmetric(el.LOWER, rect), "MakeSymmetric of a non-square matrix")
    must_fail(lambda: el.AxpyTrapezoid(el.LOWER, 1.0, rect, sq), "AxpyTrapezoid with nonconformal X, Y")
    a_mc = el.DistMatrix(grid, el.MC, el.STAR, dtype).from_global(R)
    b_mr = el.DistMatrix(grid, el.STAR, el.MR, dtype).from_global(np.asfortranarray(R.T))
    must_fail(lambda: el.LocalTrrk(el.LOWER, el.TRANSPOSE, el.NORMAL, 1.0, a_mc, b_mr, 0.0, sq), "LocalTrrk with [MC,*] passed as transposed A")
    if og.r > 1:
        c_mis = el.DistMatrix(grid, el.MC, el.MR, dtype).Align(1, 0).from_global(G)
        must_fail(lambda: el.LocalTrrk(el.LOWER, el.NORMAL, el.NORMAL, 1.0, a_mc, b_mr, 0.0, c_mis), "LocalTrrk with A not aligned with C")
    # nothing was modified
    assert np.array_equal(sq.get_local(), OD.local_of(G, OD.Spec(OD.MC, OD.MR, 0, 0, 12, 12), og, me))
    assert np.array_equal(rect.get_local(), OD.local_of(R, OD.Spec(OD.MC, OD.MR, 0, 0, 12, 9), og, me))
    return cnt


def run_level3(grid) -> dict:
    """The triangular rank-k family and the comparisons with the reference's own per-rank outputs."""
    out = {}
    out["trapezoids"] = check_trapezoids(grid, torch.float64) + check_trapezoids(grid, torch.float32)
    out["level3"] = check_level3(grid, torch.float64) + check_level3(grid, torch.float32)
    out["trsm"] = check_trsm(grid, torch.float64) + check_trsm(grid, torch.float32)
    out["level3_errors"] = check_level3_errors(grid, torch.float64)
    out["level3_ref_golden"] = check_level3_golden(grid)
    out["dist_ref_golden"] = check_dist_golden(grid)
    out["fuzz"] = check_fuzz(grid, torch.float64, iters=30) + check_fuzz(grid, torch.float32, iters=30)
    return out


def check_gemm_dot_pipeline(grid, dtype=torch.float64) -> int:
    """SUMMA_Dot in its tall-skinny regime (m, n << k): with ELB200_DOT_CHUNK small enough (hostsim tests set 4;
    the default is 4096 / 2048 local k per range) the up-front all-to-all of A and B runs in k ranges pipelined
    against the local GEMM; all orientations, aligned and misaligned operands, ragged last range."""
    cnt = 0
    mis = ((1, 2), (0, 1), (1, 0))
    for oa, ob in itertools.product((0, 1), (0, 1)):
        for al in (None, mis):
            for alg in (el.GEMM_SUMMA_DOT, el.GEMM_DEFAULT):   # DEFAULT resolves to DOT here (10 m <= k, 10 n <= k)
                err, tol = gemm_case(grid, dtype, oa, ob, alg, 33, 27, 1111, alpha=0.5, beta=-0.5, aligns=al, nb=16)
                assert err <= tol, (oa, ob, alg, al, err, tol)
                cnt += 1
    return cnt


def check_collectives(grid, dtype=torch.float64) -> int:
    """The per-collective C-ABI exports (include/elb200.h: elb200_allgather / reduce_scatter / allreduce / alltoall /
    sendrecv / bcast — what mpi::AllGather<GPU> ... bind, src/core/imports/mpi/*.hpp) over each of the grid's four
    communicators (Grid::MCComm / MRComm / VCComm / VRComm), at a size on the latency path and a ragged one. Values
    are small integers, so every result — sums included, whatever the order — is exact. Device buffers are the local
    matrices of [*,*] DistMatrices (one column), which both back ends can fill and read."""
    lib, check = el.lib, el.check
    npdt = NPDT[dtype]
    code = {torch.float32: 0, torch.float64: 1, torch.float16: 3}[dtype]
    stream = lib.elb200_grid_stream(grid._h)
    keep, cnt = [], 0
    half = dtype == torch.float16  # integers up to 2048 are exact in half: smaller tags, sums stay exact

    def dev(values):
        a = np.asarray(values, dtype=npdt).reshape(-1, 1)
        D = el.DistMatrix(grid, el.STAR, el.STAR, dtype).Resize(a.shape[0], 1)
        D.set_local(a)
        keep.append(D)
        return D, D._info().buffer

    def val(rank, idx):
        return (rank % 64) * 2 + (np.asarray(idx) % 2) if half else rank * 1000 + (np.asarray(idx) % 997)

    expect_rank = (grid.MCRank(), grid.MRRank(), grid.VCRank(), grid.VRRank())
    expect_size = (grid.Height(), grid.Width(), grid.Size(), grid.Size())
    for which in range(4):
        ch = C.c_void_p()
        check(lib.elb200_grid_comm(grid._h, which, C.byref(ch)), "elb200_grid_comm")
        p, me = lib.elb200_comm_size(ch), lib.elb200_comm_rank(ch)
        assert (me, p) == (expect_rank[which], expect_size[which]), (which, me, p)
        for n in (7, 3001):
            i = np.arange(n)
            # allgather
            S, sp = dev(val(me, i)); R, rp = dev(np.zeros(n * p))
            check(lib.elb200_allgather(sp, rp, n, code, ch, stream), "elb200_allgather")
            assert np.array_equal(R.get_local()[:, 0], np.concatenate([val(q, i) for q in range(p)]).astype(npdt)), ("allgather", which, n)
            # reduce_scatter, out of place and in place (result at the start of the buffer), allreduce
            big = np.arange(n * p)
            S, sp = dev(val(me, big)); R, rp = dev(np.zeros(n))
            want = sum(val(q, big[me * n:(me + 1) * n]) for q in range(p)).astype(npdt)
            check(lib.elb200_reduce_scatter(sp, rp, n, code, ch, stream), "elb200_reduce_scatter")
            assert np.array_equal(R.get_local()[:, 0], want), ("reduce_scatter", which, n)
            check(lib.elb200_reduce_scatter(sp, sp, n, code, ch, stream), "elb200_reduce_scatter (in place)")
            assert np.array_equal(S.get_local()[:n, 0], want), ("reduce_scatter in place", which, n)
            S, sp = dev(val(me, big)); R, rp = dev(np.zeros(n * p))
            check(lib.elb200_allreduce(sp, rp, n * p, code, ch, stream), "elb200_allreduce")
            assert np.array_equal(R.get_local()[:, 0], sum(val(q, big) for q in range(p)).astype(npdt)), ("allreduce", which, n)
            # alltoall: block j of rank q's buffer carries the tag 8 q + j
            S, sp = dev(np.concatenate([val(me * 8 + j, i) for j in range(p)])); R, rp = dev(np.zeros(n * p))
            check(lib.elb200_alltoall(sp, rp, n, code, ch, stream), "elb200_alltoall")
            assert np.array_equal(R.get_local()[:, 0], np.concatenate([val(q * 8 + me, i) for q in range(p)]).astype(npdt)), ("alltoall", which, n)
            # sendrecv around the ring, bcast from the last rank
            S, sp = dev(val(me, i)); R, rp = dev(np.zeros(n))
            check(lib.elb200_sendrecv(sp, n, (me + 1) % p, rp, n, (me - 1) % p, code, ch, stream), "elb200_sendrecv")
            assert np.array_equal(R.get_local()[:, 0], val((me - 1) % p, i).astype(npdt)), ("sendrecv", which, n)
            check(lib.elb200_bcast(sp, n, code, p - 1, ch, stream), "elb200_bcast")
            assert np.array_equal(S.get_local()[:, 0], val(p - 1, i).astype(npdt)), ("bcast", which, n)
            cnt += 7
            grid.sync()
            keep.clear()
        check(lib.elb200_comm_free(ch), "elb200_comm_free")
    return cnt


def check_half(grid) -> int:
    """the same path for IEEE half — DistMatrix<half> redistributions (bit-exact), the collectives, and
    El::Gemm in every orientation and SUMMA variant (tcgen05 kind::f16 local products, fp32 accumulation, one rounding to half
    per panel / reduction), against the oracle on the exact half inputs."""
    dt = torch.float16
    cnt = check_redistributions(grid, dt, m=23, n=17, seed=9,
                                pairs=[((OD.MC, OD.MR), t) for t in OD.PAIRS] + [(t, (OD.MC, OD.MR)) for t in OD.PAIRS])
    cnt += check_collectives(grid, dt)
    for oa, ob in itertools.product((0, 1), (0, 1)):
        for alg in (el.GEMM_SUMMA_A, el.GEMM_SUMMA_B, el.GEMM_SUMMA_C, el.GEMM_SUMMA_DOT, el.GEMM_DEFAULT):
            err, tol = gemm_case(grid, dt, oa, ob, alg, 70, 50, 90, alpha=2.0, beta=0.5, nb=16)
            assert err <= tol, f"half Gemm oa={oa} ob={ob} alg={alg}: {err:.3e} > {tol:.3e}"
            cnt += 1
    err, tol = gemm_case(grid, dt, 0, 0, el.GEMM_SUMMA_C, 520, 390, 1100, alpha=0.5, beta=-0.5)
    assert err <= tol, ("half Gemm 520x390x1100", err, tol)
    return cnt + 1


def _complex_matrix(rows, cols, seed, npdt):
    real = np.float32 if np.dtype(npdt) == np.complex64 else np.float64
    return (OG.uniform_matrix(rows, cols, seed, real, 0.0, 1.0) + 1j * OG.uniform_matrix(rows, cols, seed + 50, real, 0.0, 1.0)).astype(npdt)


def complex_gemm_case(grid, npdt, oa, ob, alg, m, n, k, alpha=2.0 - 0.5j, beta=3.0 + 0.25j, aligns=None, nb=None, dists=None):
    """One El::Gemm on complex DistMatrices (planar storage, 4M on the real kernels) vs the global definition with
    conjugation; error restricted to this rank's entries like gemm_case."""
    og = _grid_oracle(grid)
    me = grid.Rank()
    A = _complex_matrix(*((m, k) if oa == 0 else (k, m)), 201, npdt)
    B = _complex_matrix(*((k, n) if ob == 0 else (n, k)), 202, npdt)
    C = _complex_matrix(m, n, 203, npdt)
    aligns = aligns or ((0, 0), (0, 0), (0, 0))
    dists = dists or ((el.MC, el.MR),) * 3
    mats = []
    for G, d, al in zip((A, B, C), dists, aligns):
        mats.append(el.ComplexDistMatrix(grid, d[0], d[1], npdt).Align(al[0] % og.stride(d[0]), al[1] % og.stride(d[1])).from_global(G))
    dA, dB, dC = mats
    if nb:
        el.PushBlocksizeStack(nb)
    try:
        el.Gemm(oa, ob, alpha, dA, dB, beta, dC, alg)
    finally:
        if nb:
            el.PopBlocksizeStack()
    got = dC.get_local()
    ref = OG.gemm_global(oa, ob, alpha, A, B, beta, C)
    spec = OD.Spec(dists[2][0], dists[2][1], dC.ColAlign(), dC.RowAlign(), m, n)
    ref_loc = OD.local_of(ref, spec, og, me)
    num = float(np.linalg.norm(got.astype(np.complex128) - ref_loc))
    den = float(np.linalg.norm(ref)) / np.sqrt(og.p)
    for M in mats:
        M.destroy()
    return num / max(den, 1e-300), OG.gemm_tolerance(npdt, k)


def check_complex(grid) -> int:
    """SURVEY 8(f) N3, complex: El::Gemm<Complex<float|double>> in every orientation (NORMAL / TRANSPOSE / ADJOINT) and SUMMA
    variant, real and complex alpha / beta, aligned and misaligned operands, beta == 0 over NaN; redistribution, Transpose /
    Adjoint, Axpy, Scale of complex DistMatrices (bit-exact where only data moves)."""
    og = _grid_oracle(grid)
    me = grid.Rank()
    cnt = 0
    for npdt in (np.complex64, np.complex128):
        for oa, ob in itertools.product((0, 1, 2), (0, 1, 2)):
            for alg in (el.GEMM_SUMMA_C, el.GEMM_DEFAULT) if (oa, ob) not in ((0, 0), (2, 0), (0, 2)) else (
                    el.GEMM_SUMMA_A, el.GEMM_SUMMA_B, el.GEMM_SUMMA_C, el.GEMM_SUMMA_DOT, el.GEMM_DEFAULT):
                err, tol = complex_gemm_case(grid, npdt, oa, ob, alg, 70, 50, 90, nb=16)
                assert err <= tol, f"complex Gemm {npdt.__name__} oa={oa} ob={ob} alg={alg}: {err:.3e} > {tol:.3e}"
                cnt += 1
        # real scalars (the no-temporary path), misaligned operands, other distributions through the proxies
        err, tol = complex_gemm_case(grid, npdt, 2, 0, el.GEMM_SUMMA_C, 61, 47, 83, alpha=0.5, beta=-0.5, aligns=((1, 2), (0, 1), (1, 0)), nb=16)
        assert err <= tol, ("complex Gemm real scalars", npdt, err, tol)
        err, tol = complex_gemm_case(grid, npdt, 0, 2, el.GEMM_DEFAULT, 40, 33, 150, alpha=1j, beta=0.0, dists=((el.VC, el.STAR), (el.MR, el.MC), (el.MC, el.MR)), nb=16)
        assert err <= tol, ("complex Gemm other distributions", npdt, err, tol)
        cnt += 2
        # redistribution (bit-exact), Adjoint, Axpy and Scale by a complex scalar
        G = _complex_matrix(37, 29, 211, npdt)
        X = el.ComplexDistMatrix(grid, el.MC, el.MR, npdt).Align(1 % og.stride(el.MC), 0).from_global(G)
        for (U, V) in ((el.MC, el.STAR), (el.STAR, el.VR), (el.MR, el.MC), (el.STAR, el.STAR)):
            Y = el.ComplexDistMatrix(grid, U, V, npdt)
            el.Copy(X, Y)
            spec = OD.Spec(U, V, Y.ColAlign(), Y.RowAlign(), 37, 29)
            assert np.array_equal(Y.get_local(), OD.local_of(G, spec, og, me)), ("complex Copy", U, V)
            Y.destroy()
            cnt += 1
        Y = el.ComplexDistMatrix(grid, el.MC, el.MR, npdt)
        el.ZTranspose(X, Y, conjugate=True)
        spec = OD.Spec(el.MC, el.MR, Y.ColAlign(), Y.RowAlign(), 29, 37)
        assert np.array_equal(Y.get_local(), OD.local_of(G.conj().T, spec, og, me)), "complex Adjoint"
        Z = el.ComplexDistMatrix(grid, el.MC, el.MR, npdt).from_global(_complex_matrix(29, 37, 212, npdt))
        H = _complex_matrix(29, 37, 212, npdt)
        el.ZAxpy(0.5 - 2j, Y, Z)
        el.ZScale(-1.5 + 0.25j, Z)
        want = (-1.5 + 0.25j) * (H.astype(np.complex128) + (0.5 - 2j) * G.conj().T.astype(np.complex128))
        spec = OD.Spec(el.MC, el.MR, Z.ColAlign(), Z.RowAlign(), 29, 37)
        got = Z.get_local().astype(np.complex128)
        eps = 1e-5 if npdt == np.complex64 else 1e-13
        assert np.abs(got - OD.local_of(want, spec, og, me)).max() <= eps * 20, "complex Axpy / Scale"
        # beta == 0 must not read C
        N = el.ComplexDistMatrix(grid, el.MC, el.MR, npdt).from_global(np.full((29, 29), np.nan + 1j * np.nan, dtype=npdt))
        el.Gemm(0, 2, 1.0, Y, Y, 0.0, N, el.GEMM_SUMMA_C)
        assert np.isfinite(N.get_local()).all(), "complex Gemm with beta == 0 read C"
        for M in (X, Y, Z, N):
            M.destroy()
        cnt += 3
        # degenerate shapes: k == 0 is C = beta C, an empty C is a no-op, 1 x 1 x 1
        for (m, n, k) in ((5, 4, 0), (0, 4, 3), (5, 0, 3), (1, 1, 1)):
            err, tol = complex_gemm_case(grid, npdt, 0, 0, el.GEMM_DEFAULT, m, n, k, alpha=2.0 - 1.0j, beta=0.5 + 0.5j)
            assert err <= max(tol, 1e-6), ("complex Gemm degenerate shape", npdt, m, n, k, err)
            cnt += 1
    return cnt


def check_complex_ref_golden(grid) -> int:
    """El::Gemm on complex [MC,MR] DistMatrices of THIS grid against the reference's own (grid-independent) outputs:
    tests/golden/ref_gemm_complex_1x1.npz, 122 runs of the reference's CPU El::Gemm<Complex<float|double>>."""
    from oracle.make_golden import lcg_complex_matrix
    og = _grid_oracle(grid)
    me = grid.Rank()
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_gemm_complex_1x1.npz"))
    cnt = 0
    for row in z["meta"]:
        t, cid, oa, ob, alg, m, n, k, nb = map(int, row[:9])
        alpha, beta = complex(row[9], row[10]), complex(row[11], row[12])
        dt = np.complex64 if t == 0 else np.complex128
        A = lcg_complex_matrix(k if oa else m, m if oa else k, 3 * cid, dt)
        B = lcg_complex_matrix(n if ob else k, k if ob else n, 3 * cid + 1, dt)
        C0 = lcg_complex_matrix(m, n, 3 * cid + 2, dt)
        dA = el.ComplexDistMatrix(grid, el.MC, el.MR, dt).from_global(A)
        dB = el.ComplexDistMatrix(grid, el.MC, el.MR, dt).from_global(B)
        dC = el.ComplexDistMatrix(grid, el.MC, el.MR, dt).from_global(C0)
        el.PushBlocksizeStack(nb)
        try:
            el.Gemm(oa, ob, alpha, dA, dB, beta, dC, alg)
        finally:
            el.PopBlocksizeStack()
        ref = z[f"{'c32' if t == 0 else 'c64'}_{cid}_C"]
        spec = OD.Spec(el.MC, el.MR, dC.ColAlign(), dC.RowAlign(), m, n)
        num = float(np.linalg.norm(dC.get_local().astype(np.complex128) - OD.local_of(ref, spec, og, me)))
        den = float(np.linalg.norm(ref)) / np.sqrt(og.p)
        assert num / max(den, 1e-300) <= OG.gemm_tolerance(dt, k), ("complex Gemm vs the reference's output", row)
        for M in (dA, dB, dC):
            M.destroy()
        cnt += 1
    return cnt


def run_all(grid, heavy=False) -> dict:
    out = {}
    out["redist_f64"] = check_redistributions(grid, torch.float64)
    out["redist_f32"] = check_redistributions(grid, torch.float32, m=64, n=48, seed=4,
                                              pairs=[((OD.MC, OD.MR), t) for t in OD.PAIRS] + [(t, (OD.MC, OD.MR)) for t in OD.PAIRS])
    out["circ"] = check_circ(grid, torch.float64) + check_circ(grid, torch.float32)
    check_views_and_transpose(grid, torch.float32)
    check_views_and_transpose(grid, torch.float64)
    check_contractions(grid, torch.float64)
    check_contractions(grid, torch.float32)
    out["collectives"] = check_collectives(grid, torch.float64) + check_collectives(grid, torch.float32)
    out["half"] = check_half(grid)
    out["complex"] = check_complex(grid)
    out["complex_ref_golden"] = check_complex_ref_golden(grid)
    out["gemm_f64"] = check_gemm_variants(grid, torch.float64)
    out["gemm_f32"] = check_gemm_variants(grid, torch.float32)
    check_gemm_proxies(grid, torch.float64)
    check_gemm_proxies(grid, torch.float32)
    out["gemm_host"] = check_gemm_host(grid, torch.float32) + check_gemm_host(grid, torch.float64)
    out["gemm_pipeline"] = check_gemm_pipeline(grid, torch.float64) + check_gemm_pipeline(grid, torch.float32)
    out["gemm_dot_pipeline"] = check_gemm_dot_pipeline(grid, torch.float64) + check_gemm_dot_pipeline(grid, torch.float32)
    if heavy:
        # panels wider than one algorithmic block, ragged edges, default blocksize
        for dt in (torch.float32, torch.float64):
            for (oa, ob, alg) in ((0, 0, el.GEMM_SUMMA_C), (0, 1, el.GEMM_SUMMA_C), (1, 0, el.GEMM_SUMMA_A), (1, 1, el.GEMM_SUMMA_B),
                                   (0, 0, el.GEMM_SUMMA_DOT)):
                err, tol = gemm_case(grid, dt, oa, ob, alg, 1100, 900, 2500, alpha=0.5, beta=-0.5)
                assert err <= tol, (dt, oa, ob, alg, err, tol)
    return out
