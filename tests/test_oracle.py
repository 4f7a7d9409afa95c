"""CPU tests that PIN the oracle: tests/golden/ holds outputs of the reference itself (its own CPU
El::Gemm compiled by oracle/ref_build, see oracle/make_golden.py); the numpy restatement under
oracle/ has to reproduce them, and has to agree with itself across grid shapes."""
import itertools
import os

import numpy as np
import pytest

from oracle import dist as OD
from oracle import gemm as OG
from oracle import level3 as O3
from oracle.make_golden import lcg_matrix

GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module")
def gold_gemm():
    return np.load(os.path.join(GOLD, "ref_gemm_1x1.npz"))


@pytest.fixture(scope="module")
def gold_index():
    return np.load(os.path.join(GOLD, "ref_indexing.npz"))


def test_index_functions_match_reference(gold_index):
    """El::Shift / Length / MaxLength exactly as the reference evaluates them (indexing/impl.hpp)."""
    for n, rank, align, stride, shift, length, maxlen in gold_index["L"]:
        assert OD.Shift(int(rank), int(align), int(stride)) == shift
        assert OD.Length(int(n), int(shift), int(stride)) == length
        assert OD.MaxLength(int(n), int(stride)) == maxlen


def test_default_grid_height_matches_reference(gold_index):
    for p, h in gold_index["H"]:
        assert OD.DefaultHeight(int(p)) == h, (p, h)


def test_lcg_generator_matches_reference_inputs(gold_gemm):
    """Case 0's stored inputs are what the reference driver fed El::Gemm; lcg_matrix restates them."""
    for t, dt in (("f32", np.float32), ("f64", np.float64)):
        for i, name in enumerate(("A", "B", "C0")):
            ref = gold_gemm[f"{t}_0_{name}"]
            assert np.array_equal(lcg_matrix(ref.shape[0], ref.shape[1], i, dt), ref)


def _cases(gold):
    for row in gold["meta"]:
        ti, cid, oa, ob, alg, m, n, k, nb = (int(x) for x in row[:9])
        yield ("f32", np.float32) if ti == 0 else ("f64", np.float64), cid, oa, ob, alg, m, n, k, nb, row[9], row[10]


def test_oracle_gemm_reproduces_reference_outputs(gold_gemm):
    """All 120 reference runs (3 shapes x 4 orientations x {DEFAULT, SUMMA_A, B, C, DOT} x 2 dtypes,
    ragged sizes, several blocksizes): the oracle on a 1 x 1 grid gives the reference's C within the
    north_star bound; in fp64 the two differ only by summation order inside the BLAS."""
    worst = {"f32": 0.0, "f64": 0.0}
    count = 0
    for (t, dt), cid, oa, ob, alg, m, n, k, nb, alpha, beta in _cases(gold_gemm):
        A = lcg_matrix(*((k, m) if oa else (m, k)), 3 * cid + 0, dt)
        B = lcg_matrix(*((n, k) if ob else (k, n)), 3 * cid + 1, dt)
        C0 = lcg_matrix(m, n, 3 * cid + 2, dt)
        got = OG.dist_gemm(OD.Grid(1, 1), oa, ob, alpha, A.astype(np.float64), B.astype(np.float64), beta,
                           C0.astype(np.float64), alg, nb=nb)
        err = OG.frobenius_rel_err(gold_gemm[f"{t}_{cid}_C"], got)
        assert err <= OG.gemm_tolerance(dt, k), (t, cid, oa, ob, alg, err)
        worst[t] = max(worst[t], err)
        count += 1
    assert count == 120
    assert worst["f64"] < 1e-14  # far inside the bound: same algorithm, same arithmetic
    assert worst["f32"] < 1e-6


def test_reference_outputs_are_grid_independent(gold_gemm):
    """The same reference outputs are reproduced when the oracle runs the SUMMA loops over 2x2,
    2x3 and 3x1 virtual grids with non-zero alignments: ties the multi-rank restatement to the
    single-rank reference runs."""
    rows = [c for c in _cases(gold_gemm) if c[0][0] == "f64" and c[1] < 20]
    for g in (OD.Grid(2, 2), OD.Grid(2, 3), OD.Grid(3, 1)):
        for (t, dt), cid, oa, ob, alg, m, n, k, nb, alpha, beta in rows:
            A = lcg_matrix(*((k, m) if oa else (m, k)), 3 * cid + 0, dt)
            B = lcg_matrix(*((n, k) if ob else (k, n)), 3 * cid + 1, dt)
            C0 = lcg_matrix(m, n, 3 * cid + 2, dt)
            got = OG.dist_gemm(g, oa, ob, alpha, A, B, beta, C0, alg, nb=nb,
                               aligns=((1 % g.r, 1 % g.c), (0, 2 % g.c), (1 % g.r, 0)))
            assert OG.frobenius_rel_err(gold_gemm[f"{t}_{cid}_C"], got) <= OG.gemm_tolerance(dt, k)


def test_resolve_alg_heuristic():
    """GEMM_DEFAULT choice (Gemm/NN.hpp:572-600)."""
    assert OG.resolve_alg(10, 10, 100) == OG.GEMM_SUMMA_DOT
    assert OG.resolve_alg(10, 50, 20) == OG.GEMM_SUMMA_B
    assert OG.resolve_alg(50, 10, 20) == OG.GEMM_SUMMA_A
    assert OG.resolve_alg(100, 100, 100) == OG.GEMM_SUMMA_C
    assert OG.resolve_alg(100, 100, 100, OG.GEMM_SUMMA_A_MS) == OG.GEMM_SUMMA_A


@pytest.mark.parametrize("shape", [(1, 1), (2, 2), (2, 3), (4, 1), (1, 3)])
def test_redistribute_all_pairs_is_exact(shape):
    """Every (U,V) -> (U',V') redistribution among the 11 legal element-cyclic pairs moves the same
    global matrix (bit-exact), including ragged sizes and non-zero alignments."""
    g = OD.Grid(*shape)
    pairs = [(OD.MC, OD.MR), (OD.MR, OD.MC), (OD.MC, OD.STAR), (OD.STAR, OD.MC), (OD.MR, OD.STAR), (OD.STAR, OD.MR),
             (OD.VC, OD.STAR), (OD.STAR, OD.VC), (OD.VR, OD.STAR), (OD.STAR, OD.VR), (OD.STAR, OD.STAR)]
    G = np.arange(13 * 11, dtype=np.float64).reshape(13, 11)
    for (U, V), (U2, V2) in itertools.product(pairs, pairs):
        ca, ra = 1 % g.stride(U), 2 % g.stride(V)
        A = OD.VDM.from_global(g, G, U, V, ca, ra)
        assert np.array_equal(A.to_global(), G)
        B = OD.redistribute(A, U2, V2, 1 % g.stride(U2), 0)
        assert np.array_equal(B.to_global(), G)
        for q in range(g.p):
            assert B.loc[q].shape == B.spec.local_shape(g, q)


def test_empty_and_degenerate_shapes():
    g = OD.Grid(2, 2)
    for h, w in ((0, 5), (5, 0), (0, 0), (1, 1)):
        A = OD.VDM.from_global(g, np.zeros((h, w)), OD.MC, OD.MR)
        assert A.to_global().shape == (h, w)
    C = OG.dist_gemm(g, 0, 0, 2.0, np.zeros((4, 0)), np.zeros((0, 3)), 3.0, np.ones((4, 3)))
    assert np.array_equal(C, 3.0 * np.ones((4, 3)))  # k = 0: only Scale(beta, C) happens
    C = OG.dist_gemm(g, 0, 0, 2.0, np.ones((4, 2)), np.ones((2, 3)), 0.0, np.full((4, 3), np.nan))
    assert np.array_equal(C, 4.0 * np.ones((4, 3)))  # beta = 0 overwrites, NaNs do not propagate


def test_summa_variants_agree_across_grids():
    rng = np.random.default_rng(5)
    m, n, k = 23, 31, 47
    for oa, ob in itertools.product((0, 1), (0, 1)):
        A = rng.standard_normal((k, m) if oa else (m, k))
        B = rng.standard_normal((n, k) if ob else (k, n))
        C0 = rng.standard_normal((m, n))
        ref = OG.gemm_global(oa, ob, 0.5, A, B, -0.5, C0)
        for g in (OD.Grid(1, 1), OD.Grid(2, 2), OD.Grid(3, 2)):
            for alg in (OG.GEMM_SUMMA_A, OG.GEMM_SUMMA_B, OG.GEMM_SUMMA_C, OG.GEMM_SUMMA_DOT):
                got = OG.dist_gemm(g, oa, ob, 0.5, A, B, -0.5, C0, alg, nb=8)
                assert OG.frobenius_rel_err(got, ref) < 1e-14, (oa, ob, g, alg)


def test_axpy_contract_sums_contributions():
    """AxpyContract (level1/AxpyContract.hpp): [MC,STAR] partial sums reduce-scattered into [MC,MR]."""
    g = OD.Grid(2, 3)
    rng = np.random.default_rng(1)
    parts = OD.VDM.zeros(g, OD.MC, OD.STAR, 0, 0, 9, 7, np.float64)
    for q in range(g.p):
        parts.loc[q][...] = rng.standard_normal(parts.loc[q].shape)
    total = parts.contributions()
    B0 = rng.standard_normal((9, 7))
    Bd = OD.VDM.from_global(g, B0, OD.MC, OD.MR)
    OD.axpy_contract(2.0, parts, Bd)
    assert np.allclose(Bd.to_global(), B0 + 2.0 * total, rtol=0, atol=1e-13)


# ---------------------------------------------------------------------------------------------
# Multi-rank pinning: tests/golden/ref_dist.npz holds what the REFERENCE ITSELF placed on every
# rank (run multi-process over oracle/ref_build/stub/mpi_stub.cpp, see oracle/make_golden.py).
# ---------------------------------------------------------------------------------------------
EL_DIST = {0: OD.MC, 2: OD.MR, 3: OD.VC, 4: OD.VR, 5: OD.STAR}  # El::Dist enum value -> oracle constant
DIST_TAGS = ["r1c2", "r2c1", "r2c2", "r2c3", "r3c2", "r2c4"]


@pytest.fixture(scope="module")
def gold_dist():
    return np.load(os.path.join(GOLD, "ref_dist.npz"))


def _ranks(gold, tag):
    r, c = int(tag[1]), int(tag[3])
    return OD.Grid(r, c), [q for q in range(r * c)]


@pytest.mark.parametrize("tag", DIST_TAGS)
def test_grid_rank_maps_match_reference(gold_dist, tag):
    """Row / Col / VCRank / VRRank of every process as El::Grid computes them (Grid.cpp:143-148)."""
    g, ranks = _ranks(gold_dist, tag)
    for q in ranks:
        rank, h, w, row, col, vc, vr = (int(x) for x in gold_dist[f"{tag}_q{q}_I"])
        assert (h, w) == (g.r, g.c)
        assert g.coords(vc) == (row, col)
        assert g.vc(row, col) == vc and g.vr(row, col) == vr


@pytest.mark.parametrize("tag", DIST_TAGS)
def test_redistribution_placement_matches_reference(gold_dist, tag):
    """All 121 (U,V)->(U2,V2) pairs with non-zero alignments: the oracle's per-rank local matrices,
    shifts and local shapes equal the reference's, bit for bit, on every rank."""
    g, ranks = _ranks(gold_dist, tag)
    G = np.arange(13 * 11, dtype=np.float64).reshape(13, 11) + 1.0
    checked = 0
    for q in ranks:
        vc = int(gold_dist[f"{tag}_q{q}_I"][5])
        blob = gold_dist[f"{tag}_q{q}_bin"]
        for (U, V, U2, V2, caA, raA, caB, raB, csB, rsB, lh, lw, off) in gold_dist[f"{tag}_q{q}_R"]:
            A = OD.VDM.from_global(g, G, EL_DIST[int(U)], EL_DIST[int(V)], int(caA), int(raA))
            B = OD.redistribute(A, EL_DIST[int(U2)], EL_DIST[int(V2)], int(caB), int(raB))
            assert (B.spec.col_shift(g, vc), B.spec.row_shift(g, vc)) == (csB, rsB)
            assert B.loc[vc].shape == (lh, lw)
            ref = blob[off:off + lh * lw].reshape((lh, lw), order="F")
            assert np.array_equal(B.loc[vc], ref), (tag, q, U, V, U2, V2)
            checked += 1
    assert checked == 121 * g.p


@pytest.mark.parametrize("tag", DIST_TAGS)
def test_circ_gather_scatter_matches_reference(gold_dist, tag):
    """[CIRC,CIRC] (copy::Gather / copy::Scatter): every distribution gathered to a root, the root
    scattered to every distribution, and root-to-root — 23 cases per grid with roots 2 mod p
    (source) and 1 mod p (target); per-rank local matrices equal the reference's bit for bit,
    and non-root ranks hold 0 x 0."""
    g, ranks = _ranks(gold_dist, tag)
    G = np.arange(13 * 11, dtype=np.float64).reshape(13, 11) + 1.0
    el = dict(EL_DIST); el[6] = OD.CIRC
    checked = 0
    for q in ranks:
        vc = int(gold_dist[f"{tag}_q{q}_I"][5])
        blob = gold_dist[f"{tag}_q{q}_bin"]
        for (U, V, U2, V2, caA, raA, caB, raB, csB, rsB, lh, lw, off, rootA, rootB) in gold_dist[f"{tag}_q{q}_Q"]:
            A = OD.VDM.from_global(g, G, el[int(U)], el[int(V)], int(caA), int(raA), int(rootA))
            B = OD.redistribute(A, el[int(U2)], el[int(V2)], int(caB), int(raB), int(rootB))
            assert B.loc[vc].shape == (lh, lw), (tag, q, U, U2)
            if int(U2) == 6:
                assert (lh, lw) == ((13, 11) if vc == int(rootB) else (0, 0))
            ref = blob[off:off + lh * lw].reshape((lh, lw), order="F")
            assert np.array_equal(B.loc[vc], ref), (tag, q, U, V, U2, V2)
            checked += 1
    assert checked == 23 * g.p


@pytest.mark.parametrize("tag", DIST_TAGS)
def test_distributed_gemm_matches_reference_per_rank(gold_dist, tag):
    """20 El::Gemm runs of the reference per grid (NN/NT/TN/TT x {DEFAULT, SUMMA_A, B, C, DOT},
    aligned A, B, C, nb = 16): the oracle's SUMMA restatement gives the same local C on each rank
    within the fp64 bound (summation order differs inside BLAS and the reductions)."""
    g, ranks = _ranks(gold_dist, tag)
    aligns = ((1 % g.r, 2 % g.c), (0, 1 % g.c), (1 % g.r, 0))
    cache = {}
    for q in ranks:
        vc = int(gold_dist[f"{tag}_q{q}_I"][5])
        blob = gold_dist[f"{tag}_q{q}_bin"]
        for (cid, oa, ob, alg, m, n, k, nb, caC, raC, cs, rs, lh, lw, off) in (tuple(int(x) for x in r) for r in gold_dist[f"{tag}_q{q}_G"]):
            if cid not in cache:
                A = lcg_matrix(*((k, m) if oa else (m, k)), 3 * cid + 0, np.float64)
                B = lcg_matrix(*((n, k) if ob else (k, n)), 3 * cid + 1, np.float64)
                C0 = lcg_matrix(m, n, 3 * cid + 2, np.float64)
                cache[cid] = OG.dist_gemm(g, oa, ob, 2.0, A, B, 3.0, C0, alg, nb=nb, aligns=aligns)
            spec = OD.Spec(OD.MC, OD.MR, caC, raC, m, n)
            assert (caC, raC) == aligns[2]
            mine = OD.local_of(cache[cid], spec, g, vc)
            ref = blob[off:off + lh * lw].reshape((lh, lw), order="F")
            assert mine.shape == ref.shape
            assert OG.frobenius_rel_err(mine, ref) <= OG.gemm_tolerance(np.float64, k), (tag, q, cid)


# ---------------------------------------------------------------------------------------------
# Triangular rank-k family: tests/golden/ref_level3.npz holds per-rank outputs of
# the reference's ScaleTrapezoid / MakeTrapezoidal / AxpyTrapezoid / MakeSymmetric, the four
# LocalTrrk overloads, Trrk, Syrk, Symm and Trsm (oracle/ref_build/gemm_ref.cpp::level3_golden).
# ---------------------------------------------------------------------------------------------
L3_TAGS = ["r1c1"] + DIST_TAGS
L3_N, L3_K, L3_M2 = 37, 23, 29


def level3_oracle_case(g, kind, uplo, p1, p2):
    """Re-run one record of ref_level3.npz through the oracle restatement; returns the result VDM."""
    ca, ra = 1 % g.r, 2 % g.c
    n, k, m2 = L3_N, L3_K, L3_M2
    f64 = np.float64
    if kind in (0, 1):
        A = OD.VDM.from_global(g, lcg_matrix(13, 11, 7001, f64), OD.MC, OD.MR, ca, ra)
        (O3.scale_trapezoid(-0.5, uplo, A, p1) if kind == 0 else O3.make_trapezoidal(uplo, A, p1))
        return A
    if kind == 2:
        A = OD.VDM.from_global(g, lcg_matrix(13, 13, 7001, f64), OD.MC, OD.MR, ca, ra)
        X = OD.VDM.from_global(g, lcg_matrix(13, 13, 7004, f64), OD.MC, OD.MR, ca, ra)
        O3.axpy_trapezoid(uplo, 2.0, X, A, p1)
        return A
    if kind == 3:
        A = OD.VDM.from_global(g, lcg_matrix(13, 13, 7001, f64), OD.MC, OD.MR, ca, ra)
        O3.make_symmetric(uplo, A)
        return A
    if kind == 4:
        C = OD.VDM.from_global(g, lcg_matrix(n, n, 7003, f64), OD.MC, OD.MR, ca, ra)
        A = (OD.VDM.from_global(g, lcg_matrix(k, n, 7001, f64), OD.STAR, OD.MC, 0, ca) if p1 else
             OD.VDM.from_global(g, lcg_matrix(n, k, 7001, f64), OD.MC, OD.STAR, ca, 0))
        B = (OD.VDM.from_global(g, lcg_matrix(n, k, 7002, f64), OD.MR, OD.STAR, ra, 0) if p2 else
             OD.VDM.from_global(g, lcg_matrix(k, n, 7002, f64), OD.STAR, OD.MR, 0, ra))
        O3.local_trrk(uplo, p1, p2, 2.0, A, B, 3.0, C, blocksize=4)
        return C
    if kind == 5:
        A = OD.VDM.from_global(g, lcg_matrix(*((k, n) if p1 else (n, k)), 7001, f64), OD.MC, OD.MR, ca, 0)
        B = OD.VDM.from_global(g, lcg_matrix(*((n, k) if p2 else (k, n)), 7002, f64), OD.MC, OD.MR, 0, ra)
        C = OD.VDM.from_global(g, lcg_matrix(n, n, 7003, f64), OD.MC, OD.MR, ca, ra)
        O3.trrk(uplo, p1, p2, 2.0, A, B, 3.0, C, nb=16)
        return C
    if kind == 6:
        A = OD.VDM.from_global(g, lcg_matrix(*((k, n) if p1 else (n, k)), 7001, f64), OD.MC, OD.MR, ca, 0)
        C = OD.VDM.from_global(g, lcg_matrix(n, n, 7003, f64), OD.MC, OD.MR, 0, ra)
        O3.syrk(uplo, p1, 2.0, A, 3.0, C, nb=16)
        return C
    if kind == 7:
        shape = (m2, n) if p1 else (n, m2)
        A = OD.VDM.from_global(g, lcg_matrix(n, n, 7001, f64), OD.MC, OD.MR, ca, ra)
        B = OD.VDM.from_global(g, lcg_matrix(*shape, 7002, f64), OD.MC, OD.MR, 0, 0)
        C = OD.VDM.from_global(g, lcg_matrix(*shape, 7003, f64), OD.MC, OD.MR, 0, ra)
        O3.symm(p1, uplo, 2.0, A, B, 3.0, C)
        return C
    if kind == 8:
        side, orient, diag = p1, p2 // 2, p2 % 2
        Ag = lcg_matrix(n, n, 7001, f64)
        Ag = Ag * (1.0 / float(n)) if diag else Ag + float(n) * np.eye(n)
        A = OD.VDM.from_global(g, np.asfortranarray(Ag), OD.MC, OD.MR, ca, 0)
        B = OD.VDM.from_global(g, lcg_matrix(*((m2, n) if side else (n, m2)), 7002, f64), OD.MC, OD.MR, 0, ra)
        O3.trsm(side, uplo, orient, diag, 2.0, A, B, nb=16)
        return B
    raise ValueError(kind)


@pytest.mark.parametrize("tag", L3_TAGS)
def test_level3_matches_reference_per_rank(tag):
    """Every record of the reference's level-3 run: the oracle's per-rank local result equals the
    reference's — bit for bit for the trapezoid routines and MakeSymmetric (kinds 0-3), within the fp64
    Gemm bound for LocalTrrk / Trrk / Syrk / Symm / Trsm (BLAS summation order)."""
    gold = np.load(os.path.join(GOLD, "ref_level3.npz"))
    g = OD.Grid(int(tag[1]), int(tag[3]))
    cache = {}
    checked = 0
    for q in range(g.p):
        blob = gold[f"{tag}_q{q}_bin"]
        for (kind, uplo, p1, p2, caC, raC, cs, rs, lh, lw, off) in (tuple(int(x) for x in r) for r in gold[f"{tag}_q{q}_L"]):
            key = (kind, uplo, p1, p2)
            if key not in cache:
                cache[key] = level3_oracle_case(g, kind, uplo, p1, p2)
            R = cache[key]
            assert (R.spec.col_align, R.spec.row_align) == (caC, raC), (tag, key)
            assert (R.spec.col_shift(g, q), R.spec.row_shift(g, q)) == (cs, rs)
            ref = blob[off:off + lh * lw].reshape((lh, lw), order="F")
            assert R.loc[q].shape == ref.shape, (tag, q, key)
            if kind <= 3:
                assert np.array_equal(R.loc[q], ref), (tag, q, key)
            else:
                assert OG.frobenius_rel_err(R.loc[q], ref) <= OG.gemm_tolerance(np.float64, L3_K), (tag, q, key)
            checked += 1
    assert checked == 60 * g.p


def test_level3_restatement_equals_global_definition():
    """LocalTrrk's recursion / Trrk's panel loop / Syrk (oracle.level3) against the plain global
    definitions on grids the reference was not run on, several blocksizes."""
    rng = np.random.default_rng(3)
    n, k = 41, 19
    for g in (OD.Grid(1, 3), OD.Grid(3, 3), OD.Grid(4, 2)):
        for uplo in (O3.LOWER, O3.UPPER):
            for oa, ob in itertools.product((0, 1), (0, 1)):
                A = rng.standard_normal((k, n) if oa else (n, k))
                B = rng.standard_normal((n, k) if ob else (k, n))
                C0 = rng.standard_normal((n, n))
                C = OD.VDM.from_global(g, C0, OD.MC, OD.MR, 1 % g.r, 1 % g.c)
                O3.trrk(uplo, oa, ob, 0.5, OD.VDM.from_global(g, A, OD.MC, OD.MR), OD.VDM.from_global(g, B, OD.MC, OD.MR, 0, 1 % g.c),
                        -0.5, C, nb=8)
                ref = O3.trrk_global(uplo, oa, ob, 0.5, A, B, -0.5, C0, np.float64)
                assert OG.frobenius_rel_err(C.to_global(), ref) < 1e-14, (g.r, g.c, uplo, oa, ob)


def test_complex_gemm_oracle_reproduces_reference_outputs():
    """SURVEY 8(f) N3, complex: the reference's own El::Gemm<Complex<float|double>> outputs (122 runs: NORMAL / TRANSPOSE /
    ADJOINT x five GemmAlgorithm values, complex alpha and beta; tests/golden/ref_gemm_complex_1x1.npz) are reproduced
    (a) by the global definition with conjugation and (b) by the 4M composition over the pinned real restatement
    (oracle.gemm.dist_gemm_complex) on 1 x 1, 2 x 2 and 2 x 3 virtual grids — the arithmetic the product path runs."""
    from oracle.make_golden import lcg_complex_matrix
    z = np.load(os.path.join(GOLD, "ref_gemm_complex_1x1.npz"))
    meta = z["meta"]
    assert len(meta) == 122
    for row in meta:
        t, cid, oa, ob, alg, m, n, k, nb = map(int, row[:9])
        alpha, beta = complex(row[9], row[10]), complex(row[11], row[12])
        dt = np.complex64 if t == 0 else np.complex128
        A = lcg_complex_matrix(k if oa else m, m if oa else k, 3 * cid, dt)
        B = lcg_complex_matrix(n if ob else k, k if ob else n, 3 * cid + 1, dt)
        C0 = lcg_complex_matrix(m, n, 3 * cid + 2, dt)
        ref = z[f"{'c32' if t == 0 else 'c64'}_{cid}_C"]
        tol = OG.gemm_tolerance(dt, k)
        assert OG.frobenius_rel_err(OG.gemm_global(oa, ob, alpha, A, B, beta, C0), ref) <= tol, row
        for shape in ((1, 1), (2, 2), (2, 3)):
            if shape != (1, 1) and cid % 3:
                continue  # every third case on the larger virtual grids keeps the CPU suite short
            got = OG.dist_gemm_complex(OD.Grid(*shape), oa, ob, alpha, A, B, beta, C0, alg, nb)
            assert got.dtype == dt
            assert OG.frobenius_rel_err(got, ref) <= tol, (row, shape)
