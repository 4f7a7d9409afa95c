"""CPU tests of the host logic behind the C ABI: the redistribution planner, the contract planner,
the index functions and the GEMM_DEFAULT heuristic of libelb200, checked against the oracle (and,
through tests/golden, against the reference's own values).  No GPU and no compute entry point is
touched: plans are executed on numpy arrays by tests/plan_sim.py."""
import itertools
import os

import numpy as np
import pytest

from elemental_b200 import _lib as L
from oracle import dist as OD
from oracle import gemm as OG
from tests import plan_sim as PS

PAIRS = [(L.MC, L.MR), (L.MR, L.MC), (L.MC, L.STAR), (L.STAR, L.MC), (L.MR, L.STAR), (L.STAR, L.MR),
         (L.VC, L.STAR), (L.STAR, L.VC), (L.VR, L.STAR), (L.STAR, L.VR), (L.STAR, L.STAR)]
GRIDS = [(1, 1), (1, 2), (2, 1), (2, 2), (2, 3), (3, 2), (2, 4), (4, 2), (1, 5), (3, 3)]
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_index_functions_match_reference_tables():
    tab = np.load(os.path.join(GOLD, "ref_indexing.npz"))
    for n, rank, align, stride, shift, length, maxlen in tab["L"]:
        assert L.lib.elb200_shift(int(rank), int(align), int(stride)) == shift
        assert L.lib.elb200_length(int(n), int(shift), int(stride)) == length
        assert L.lib.elb200_max_length(int(n), int(stride)) == maxlen
    for p, h in tab["H"]:
        assert L.lib.elb200_default_grid_height(int(p)) == h


def test_resolve_alg_matches_oracle():
    for m, n, k in itertools.product((1, 7, 10, 64, 100, 1000), repeat=3):
        for alg in range(8):
            assert L.lib.elb200_gemm_resolve_alg(0, 0, m, n, k, alg) == OG.resolve_alg(m, n, k, alg), (m, n, k, alg)


@pytest.mark.parametrize("shape", GRIDS)
def test_copy_plans_all_pairs(shape):
    """121 (U,V)->(U',V') pairs per grid, ragged 13 x 11 matrix, non-zero alignments on both sides:
    executing the plans of all ranks reproduces the oracle's redistribution bit for bit, every
    send has its matching receive, and all ranks choose the same exchange kind."""
    g = OD.Grid(*shape)
    G = (np.arange(13 * 11, dtype=np.float64).reshape(13, 11) + 1.0)
    for (U, V), (U2, V2) in itertools.product(PAIRS, PAIRS):
        for (ca, ra, ca2, ra2) in ((0, 0, 0, 0), (1, 2, 3, 1)):
            A = OD.VDM.from_global(g, G, U, V, ca, ra)
            B = PS.run_copy(g, A, U2, V2, ca2 % g.stride(U2), ra2 % g.stride(V2))
            ref = OD.redistribute(A, U2, V2, ca2, ra2)
            for q in range(g.p):
                assert np.array_equal(B.loc[q], ref.loc[q]), (shape, U, V, U2, V2, ca, ra, ca2, ra2, q)


@pytest.mark.parametrize("shape", [(2, 2), (2, 3), (2, 4)])
def test_copy_plans_degenerate_shapes(shape):
    """Matrices smaller than the grid (some ranks own nothing), single rows/columns, empty."""
    g = OD.Grid(*shape)
    for (h, w) in ((1, 1), (1, 9), (9, 1), (2, 3), (0, 4), (4, 0)):
        G = np.arange(h * w, dtype=np.float32).reshape(h, w) + 1
        for (U, V), (U2, V2) in itertools.product(PAIRS, PAIRS):
            A = OD.VDM.from_global(g, G, U, V, 1, 1)
            if h == 0 or w == 0:
                continue  # the engine returns before planning (engine_redist.cpp Copy)
            B = PS.run_copy(g, A, U2, V2, 0, 1 % g.stride(V2))
            assert np.array_equal(B.to_global(), G)


def test_allgather_fast_path_is_recognised():
    """[MC,MR] -> [MC,STAR] is an AllGather over the row (MR) communicator, [MC,MR] -> [STAR,MR]
    over the column (MC) communicator, [VC,STAR] -> [STAR,STAR] over the world — the collectives
    the reference uses (Copy/RowAllGather.hpp, Copy/ColAllGather.hpp, Copy/AllGather.hpp)."""
    r, c = 2, 3
    a = PS.spec_c(L.MC, L.MR, 0, 0, 50, 40)
    p = PS.copy_plan(r, c, 1, a, PS.spec_c(L.MC, L.STAR, 0, 0, 50, 40))
    assert (p.kind, p.comm_dist, p.nrecv) == (PS.EX_ALLGATHER, L.MR, c)
    assert p.portion == OD.MaxLength(50, r) * OD.MaxLength(40, c)  # RowAllGather.hpp:47-51
    p = PS.copy_plan(r, c, 1, a, PS.spec_c(L.STAR, L.MR, 0, 0, 50, 40))
    assert (p.kind, p.comm_dist, p.nrecv) == (PS.EX_ALLGATHER, L.MC, r)
    p = PS.copy_plan(r, c, 4, PS.spec_c(L.VC, L.STAR, 0, 0, 50, 40), PS.spec_c(L.STAR, L.STAR, 0, 0, 50, 40))
    assert (p.kind, p.comm_dist, p.nrecv) == (PS.EX_ALLGATHER, L.VC, r * c)
    # identical distribution and alignment: nothing leaves the rank
    p = PS.copy_plan(r, c, 3, a, a)
    assert p.kind == PS.EX_LOCAL and p.has_self


def test_plan_argument_errors():
    a = PS.spec_c(L.MC, L.MR, 0, 0, 5, 5)
    out = L.CopyPlan()
    import ctypes as C
    assert L.lib.elb200_plan_copy(0, 1, 0, C.byref(a), C.byref(a), C.byref(out)) != 0  # empty grid
    assert L.lib.elb200_plan_copy(2, 2, 4, C.byref(a), C.byref(a), C.byref(out)) != 0  # rank out of range
    b = PS.spec_c(L.MC, L.MR, 0, 0, 6, 5)
    assert L.lib.elb200_plan_copy(2, 2, 0, C.byref(a), C.byref(b), C.byref(out)) != 0  # nonconformal
    bad = PS.spec_c(L.MC, L.MC, 0, 0, 5, 5)
    assert L.lib.elb200_plan_copy(2, 2, 0, C.byref(bad), C.byref(a), C.byref(out)) != 0  # illegal pair


CONTRACTS = [  # (A dist, B dist): the partial-sum shapes SUMMA produces (AxpyContract.hpp:483-509)
    ((L.MC, L.STAR), (L.MC, L.MR)), ((L.STAR, L.MR), (L.MC, L.MR)), ((L.STAR, L.STAR), (L.MC, L.MR)),
    ((L.MR, L.STAR), (L.MR, L.MC)), ((L.STAR, L.MC), (L.MR, L.MC)), ((L.STAR, L.STAR), (L.VC, L.STAR)),
    ((L.STAR, L.STAR), (L.STAR, L.VR)), ((L.VC, L.STAR), (L.VC, L.STAR)), ((L.MC, L.MR), (L.MC, L.MR))]


@pytest.mark.parametrize("shape", [(1, 1), (1, 2), (2, 1), (2, 2), (2, 3), (3, 2), (2, 4)])
def test_contract_plans(shape):
    """B += alpha * sum over replicas: pack table + ReduceScatter + axpy-unpack equals the oracle's
    AxpyContract (sums of small integers: exact in any summation order)."""
    g = OD.Grid(*shape)
    rng = np.random.default_rng(7)
    for (UA, VA), (UB, VB) in CONTRACTS:
        for (cb, rb) in ((0, 0), (1, 2)):
            cb, rb = cb % g.stride(UB), rb % g.stride(VB)
            ca = cb if UA == UB else 0
            ra = rb if VA == VB else 0
            A = OD.VDM.zeros(g, UA, VA, ca, ra, 14, 9, np.float64)
            for q in range(g.p):
                A.loc[q][...] = rng.integers(-8, 9, size=A.loc[q].shape)
            B0 = rng.integers(-8, 9, size=(14, 9)).astype(np.float64)
            B = OD.VDM.from_global(g, B0, UB, VB, cb, rb)
            Bref = OD.VDM.from_global(g, B0, UB, VB, cb, rb)
            assert PS.run_axpy_contract(g, 2.0, A, B), "aligned contract must plan directly"
            OD.axpy_contract(2.0, A, Bref)
            for q in range(g.p):
                assert np.array_equal(B.loc[q], Bref.loc[q]), (shape, UA, VA, UB, VB, cb, rb, q)


def test_contract_plan_rejects_misaligned_pair():
    """Different alignment on a shared distribution: the planner reports it (valid == 0) and the
    engine realigns with a Copy first, as the reference's unaligned branch does
    (AxpyContract.hpp:219-261)."""
    g = OD.Grid(2, 2)
    A = OD.VDM.zeros(g, L.MC, L.STAR, 1, 0, 6, 6, np.float64)
    B = OD.VDM.zeros(g, L.MC, L.MR, 0, 0, 6, 6, np.float64)
    assert PS.run_axpy_contract(g, 1.0, A, B) is False


def test_summa_c_from_plans_matches_reference_golden():
    """The SUMMA-C loop (Gemm/NN.hpp:330-387) driven entirely by libelb200's plans on a 2 x 2
    virtual grid — panel redistributions through plan_copy, local products in numpy — lands on
    the REFERENCE's output for golden case f64/NN/SUMMA_C."""
    gold = np.load(os.path.join(GOLD, "ref_gemm_1x1.npz"))
    from oracle.make_golden import lcg_matrix
    row = next(r for r in gold["meta"] if int(r[0]) == 1 and int(r[2]) == 0 and int(r[3]) == 0 and int(r[4]) == OG.GEMM_SUMMA_C)
    cid, m, n, k, nb, alpha, beta = int(row[1]), int(row[5]), int(row[6]), int(row[7]), int(row[8]), row[9], row[10]
    A = lcg_matrix(m, k, 3 * cid, np.float64)
    B = lcg_matrix(k, n, 3 * cid + 1, np.float64)
    C0 = lcg_matrix(m, n, 3 * cid + 2, np.float64)
    g = OD.Grid(2, 2)
    Ad, Bd = OD.VDM.from_global(g, A, L.MC, L.MR), OD.VDM.from_global(g, B, L.MC, L.MR)
    Cd = OD.VDM.from_global(g, beta * C0, L.MC, L.MR)
    for k0 in range(0, k, nb):
        k1 = min(k0 + nb, k)
        A1 = Ad.view(0, m, k0, k1)
        B1 = Bd.view(k0, k1, 0, n)
        A1_mc_star = PS.run_copy(g, A1, L.MC, L.STAR, Cd.spec.col_align, 0)
        B1_star_mr = PS.run_copy(g, B1, L.STAR, L.MR, 0, Cd.spec.row_align)
        for q in range(g.p):
            Cd.loc[q] += alpha * (A1_mc_star.loc[q] @ B1_star_mr.loc[q])
    assert OG.frobenius_rel_err(Cd.to_global(), gold[f"f64_{cid}_C"]) <= OG.gemm_tolerance(np.float64, k)


EL_DIST = {0: L.MC, 2: L.MR, 3: L.VC, 4: L.VR, 5: L.STAR}  # El::Dist enum value -> engine constant


@pytest.mark.parametrize("tag", ["r1c2", "r2c1", "r2c2", "r2c3", "r3c2", "r2c4"])
def test_planner_reproduces_reference_placement(tag):
    """libelb200's plans against the REFERENCE's own multi-rank run (tests/golden/ref_dist.npz):
    executing plan_copy for all ranks lands every entry where Hydrogen put it — 121 pairs per grid,
    non-zero alignments, bit-exact, including the 2 x 4 grid of the 8-GPU configuration."""
    gold = np.load(os.path.join(GOLD, "ref_dist.npz"))
    g = OD.Grid(int(tag[1]), int(tag[3]))
    G = np.arange(13 * 11, dtype=np.float64).reshape(13, 11) + 1.0
    done = {}
    for q in range(g.p):
        vc = int(gold[f"{tag}_q{q}_I"][5])
        blob = gold[f"{tag}_q{q}_bin"]
        for (U, V, U2, V2, caA, raA, caB, raB, _cs, _rs, lh, lw, off) in (tuple(int(x) for x in r) for r in gold[f"{tag}_q{q}_R"]):
            key = (U, V, U2, V2)
            if key not in done:
                A = OD.VDM.from_global(g, G, EL_DIST[U], EL_DIST[V], caA, raA)
                done[key] = PS.run_copy(g, A, EL_DIST[U2], EL_DIST[V2], caB, raB)
            ref = blob[off:off + lh * lw].reshape((lh, lw), order="F")
            assert np.array_equal(done[key].loc[vc], ref), (tag, q, key)
    assert len(done) == 121
    # [CIRC,CIRC]: gather to / scatter from a root, root to root (23 cases per grid)
    el = dict(EL_DIST); el[6] = L.CIRC
    done = {}
    for q in range(g.p):
        vc = int(gold[f"{tag}_q{q}_I"][5])
        blob = gold[f"{tag}_q{q}_bin"]
        for (U, V, U2, V2, caA, raA, caB, raB, _cs, _rs, lh, lw, off, rootA, rootB) in (tuple(int(x) for x in r) for r in gold[f"{tag}_q{q}_Q"]):
            key = (U, V, U2, V2)
            if key not in done:
                A = OD.VDM.from_global(g, G, el[U], el[V], caA, raA, rootA)
                done[key] = PS.run_copy(g, A, el[U2], el[V2], caB, raB, rootB)
            ref = blob[off:off + lh * lw].reshape((lh, lw), order="F")
            assert done[key].loc[vc].shape == (lh, lw), (tag, q, key)
            assert np.array_equal(done[key].loc[vc], ref), (tag, q, key)
    assert len(done) == 23


@pytest.mark.parametrize("shape", [(1, 2), (2, 1), (2, 2), (2, 4), (4, 2), (2, 3), (3, 2), (3, 3), (4, 6)])
def test_direct_panel_layout(shape):
    """Zero-copy SUMMA-C panels (elb200_plan_direct_panel): every rank's strided contributions,
    applied for all members of its row / column communicator, build an A panel and a B panel whose
    product equals A1 * B1 restricted to that rank — for ragged panel widths, panel offsets that
    shift the alignments, and grids where r does not divide c."""
    import ctypes as C
    g = OD.Grid(*shape)
    r, c = shape
    rng = np.random.default_rng(13)
    m, n = 9, 7
    for (k0, kb, caA, raA, caB, raB) in ((0, 16, 0, 0, 0, 0), (5, 13, 1, 2, 3, 1), (3, 1, 0, 1, 1, 0), (7, r * c + 3, 1, 1, 0, 2)):
        K = k0 + kb + 2
        A = rng.integers(-4, 5, (m, K)).astype(np.float64)
        B = rng.integers(-4, 5, (K, n)).astype(np.float64)
        Ad = OD.VDM.from_global(g, A, OD.MC, OD.MR, caA, raA)
        Bd = OD.VDM.from_global(g, B, OD.MC, OD.MR, caB, raB)
        A1, B1 = Ad.view(0, m, k0, k0 + kb), Bd.view(k0, k0 + kb, 0, n)
        ref = A[:, k0:k0 + kb] @ B[k0:k0 + kb, :]
        plans = {}
        for q in range(g.p):
            row_off = (C.c_int64 * (r + 1))()
            a_cols = (C.c_int64 * (5 * 64))()
            ncls, b0, bn = C.c_int(), C.c_int64(), C.c_int64()
            assert L.lib.elb200_plan_direct_panel(r, c, q, kb, A1.spec.row_align, B1.spec.col_align, row_off, a_cols,
                                                  C.byref(ncls), C.byref(b0), C.byref(bn)) == 0
            plans[q] = (list(row_off), [tuple(a_cols[5 * i:5 * i + 5]) for i in range(ncls.value)], b0.value, bn.value)
            assert row_off[r] == kb and bn.value == B1.loc[q].shape[0]
        for q in range(g.p):
            mc, mr = g.coords(q)
            PA = np.full((A1.loc[q].shape[0], kb), np.nan)
            PB = np.full((kb, B1.loc[q].shape[1]), np.nan)
            for j in range(c):  # row communicator: everybody with my mc contributes columns of A1
                src = g.vc(mc, j)
                for (d0, ds, s0, ss, cnt) in plans[src][1]:
                    for v in range(cnt):
                        assert np.isnan(PA[:, d0 + v * ds]).all(), "two columns landed on the same place"
                        PA[:, d0 + v * ds] = A1.loc[src][:, s0 + v * ss]
            for i in range(r):  # column communicator: everybody with my mr contributes rows of B1
                src = g.vc(i, mr)
                _, _, b0, bn = plans[src]
                assert np.isnan(PB[b0:b0 + bn, :]).all()
                PB[b0:b0 + bn, :] = B1.loc[src]
            assert not np.isnan(PA).any() and not np.isnan(PB).any(), (shape, k0, kb, q)
            cspec = OD.Spec(OD.MC, OD.MR, caA, raB, m, n)  # C aligned with A's rows and B's columns
            assert np.array_equal(PA @ PB, OD.local_of(ref, cspec, g, q)), (shape, k0, kb, q)
