"""GPU parity of the triangular rank-k family: the trapezoid kernel through the device
C ABI, then ScaleTrapezoid / AxpyTrapezoid / MakeTrapezoidal / MakeSymmetric, LocalTrrk, Trrk, Syrk /
Herk, Symm / Hemm, Trsm / LocalTrsm through the handle ABI on a 1 x 1 grid (and on r x c grids under torchrun when
several GPUs are visible) — against the oracle (oracle/level3.py) and against the REFERENCE'S OWN
outputs (tests/golden/ref_level3.npz, ref_dist.npz).  Sorted after the Gemm-path test files on purpose.
"""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from oracle import gemm as OG
from oracle import level3 as O3

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SUF = {np.float32: "f32", np.float64: "f64"}


@pytest.fixture(scope="module")
def L():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from elemental_b200 import _lib
    return _lib


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("rows,cols,pad", [(1, 1, 0), (37, 53, 3), (256, 64, 0), (1000, 333, 4), (70, 3000, 1), (4099, 517, 5)])
def test_trapezoid_kernel(L, dtype, rows, cols, pad):
    """elb200_trapezoid_*: SCALE / AXPY / MAKE for both triangles, offsets, and element-cyclic index maps
    (row0 + i*row_step, col0 + j*col_step). alpha is a power of two, so every result is exact: bit-for-bit
    equality with numpy, padding rows (NaN) untouched."""
    fn = getattr(L.lib, f"elb200_trapezoid_{SUF[dtype]}")
    rng = np.random.default_rng(rows * 131 + cols)
    ld = rows + pad
    for (r0, rs, c0, cs) in ((0, 1, 0, 1), (1, 2, 3, 4), (5, 3, 0, 2)):
        for uplo in (L.LOWER, L.UPPER):
            for offset in (0, 1, -3, 40):
                mask = O3.trapezoid_mask(rows, cols, uplo, offset, r0, rs, c0, cs)
                Y0 = np.full((cols, ld), np.nan, dtype=dtype)
                Y0[:, :rows] = rng.uniform(-1, 1, (cols, rows)).astype(dtype)
                X0 = np.full((cols, ld), np.nan, dtype=dtype)
                X0[:, :rows] = rng.uniform(-1, 1, (cols, rows)).astype(dtype)
                for op, alpha in ((L.TRAP_SCALE, -0.5), (L.TRAP_SCALE, 0.0), (L.TRAP_AXPY, 2.0), (L.TRAP_MAKE, 0.0)):
                    y, x = torch.from_numpy(Y0.copy()).cuda(), torch.from_numpy(X0).cuda()
                    L.check(fn(op, uplo, offset, rows, cols, r0, rs, c0, cs, alpha, x.data_ptr(), ld, y.data_ptr(), ld, 0), "trapezoid")
                    torch.cuda.synchronize()
                    got = y.cpu().numpy()
                    Y = Y0[:, :rows].T
                    if op == L.TRAP_SCALE:
                        exp = np.where(mask, dtype(alpha) * Y if alpha != 0 else dtype(0), Y)
                    elif op == L.TRAP_AXPY:
                        exp = np.where(mask, Y + dtype(alpha) * X0[:, :rows].T, Y)
                    else:
                        exp = np.where(mask, Y, dtype(0))
                    assert np.array_equal(got[:, :rows].T, exp), (op, uplo, offset, r0, rs, c0, cs)
                    assert np.isnan(got[:, rows:]).all()
    # argument errors
    y = torch.zeros(16, dtype=torch.float32 if dtype == np.float32 else torch.float64, device="cuda")
    assert fn(L.TRAP_SCALE, 2, 0, 4, 4, 0, 1, 0, 1, 2.0, None, 4, y.data_ptr(), 4, 0) == 2   # bad uplo
    assert fn(L.TRAP_AXPY, 0, 0, 4, 4, 0, 1, 0, 1, 2.0, None, 4, y.data_ptr(), 4, 0) == 2    # AXPY without X
    assert fn(L.TRAP_MAKE, 0, 0, 4, 4, 0, 0, 0, 1, 0.0, None, 4, y.data_ptr(), 4, 0) == 2    # zero stride


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_trapezoid_routines_1x1(grid1, dtype):
    from tests import dist_cases as dc
    assert dc.check_trapezoids(grid1, dtype) == 248


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_level3_1x1(grid1, dtype):
    from tests import dist_cases as dc
    assert dc.check_level3(grid1, dtype) == 44


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_syrk_symm_large_properties(grid1, dtype):
    """Bench-class size (n = 4096, k = 2048; several LocalTrrk recursion levels with the default blocksize)
    through size-independent properties: Syrk(LOWER) equals the lower triangle of Gemm(A, A^T) and leaves the
    strict upper triangle of C bit-identical; Symm(LEFT, LOWER) equals Gemm with the symmetrised matrix."""
    from elemental_b200 import el
    n, k = 4096, 2048
    npdt = np.float32 if dtype == torch.float32 else np.float64
    gen = torch.Generator(device="cuda").manual_seed(5)
    with torch.cuda.stream(grid1.stream):
        A = el.DistMatrix(grid1, dtype=dtype).Resize(n, k)
        C = el.DistMatrix(grid1, dtype=dtype).Resize(n, n)
        G = el.DistMatrix(grid1, dtype=dtype).Resize(n, n)
        A.Matrix().copy_((torch.rand((n, k), device="cuda", dtype=dtype, generator=gen) - 0.5) * 0.2)
        C.Matrix().copy_((torch.rand((n, n), device="cuda", dtype=dtype, generator=gen) - 0.5) * 0.2)
        G.Matrix().copy_(C.Matrix())
        c0 = C.Matrix().clone()
    grid1.sync()
    el.Syrk(el.LOWER, el.NORMAL, 0.5, A, -0.5, C)
    el.Gemm(el.NORMAL, el.TRANSPOSE, 0.5, A, A, -0.5, G, el.GEMM_SUMMA_C)
    grid1.sync()
    with torch.cuda.stream(grid1.stream):
        c, g = C.Matrix(), G.Matrix()
        lower = torch.tril(torch.ones((n, n), dtype=torch.bool, device="cuda"))
        err = float(torch.linalg.norm((c - g)[lower].double()) / torch.linalg.norm(g[lower].double()))
        untouched = bool(torch.equal(c[~lower], c0[~lower]))
    grid1.sync()
    assert untouched, "Syrk wrote outside the triangle"
    assert err <= 2 * OG.gemm_tolerance(npdt, k), err
    # Symm: only the lower triangle of S referenced (upper poisoned with NaN)
    with torch.cuda.stream(grid1.stream):
        S = el.DistMatrix(grid1, dtype=dtype).Resize(n, n)
        B = el.DistMatrix(grid1, dtype=dtype).Resize(n, 512)
        C1 = el.DistMatrix(grid1, dtype=dtype).Resize(n, 512)
        C2 = el.DistMatrix(grid1, dtype=dtype).Resize(n, 512)
        s = (torch.rand((n, n), device="cuda", dtype=dtype, generator=gen) - 0.5) * 0.2
        sym = torch.tril(s) + torch.tril(s, -1).T
        S.Matrix().copy_(torch.where(lower, s, torch.full_like(s, float("nan"))))
        B.Matrix().copy_((torch.rand((n, 512), device="cuda", dtype=dtype, generator=gen) - 0.5) * 0.2)
        C1.Matrix().zero_()
        C2.Matrix().zero_()
        F = el.DistMatrix(grid1, dtype=dtype).Resize(n, n)
        F.Matrix().copy_(sym)
    grid1.sync()
    el.Symm(el.LEFT, el.LOWER, 1.0, S, B, 0.0, C1)
    el.Gemm(el.NORMAL, el.NORMAL, 1.0, F, B, 0.0, C2, el.GEMM_SUMMA_C)
    grid1.sync()
    with torch.cuda.stream(grid1.stream):
        e2 = float(torch.linalg.norm((C1.Matrix() - C2.Matrix()).double()) / torch.linalg.norm(C2.Matrix().double()))
        finite = bool(torch.isfinite(C1.Matrix()).all())
    grid1.sync()
    assert finite and e2 <= 2 * OG.gemm_tolerance(npdt, n), (finite, e2)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_trsm_kernel(L, dtype):
    """elb200_trsm_*: every side / uplo / orientation / diag at sizes that stop in the base kernel (<= 32), split
    once, and recurse several levels (1000), ragged, padded leading dimensions; the triangle (and UNIT diagonal)
    that must not be referenced holds NaN. Against the long-double substitution within 4x the Gemm bound."""
    import itertools
    fn = getattr(L.lib, f"elb200_trsm_{SUF[dtype]}")
    rng = np.random.default_rng(17)
    for (na, nrhs) in ((1, 5), (32, 130), (47, 300), (200, 333), (1000, 257)):
        combos = list(itertools.product((L.LEFT, L.RIGHT), (L.LOWER, L.UPPER), (0, 1), (L.NON_UNIT, L.UNIT)))
        if na == 1000:
            combos = combos[::3]
        A0 = rng.uniform(-1, 1, (na, na))
        for side, uplo, t, diag in combos:
            keep = O3.trapezoid_mask(na, na, uplo, 0)
            Aref = np.where(keep, A0 / na if diag == L.UNIT else A0 + na * np.eye(na), 0.0)
            Adev = np.where(keep, Aref, np.nan)
            if diag == L.UNIT:
                Adev[np.arange(na), np.arange(na)] = np.nan
            m, n = (na, nrhs) if side == L.LEFT else (nrhs, na)
            B0 = rng.uniform(-1, 1, (m, n)).astype(dtype)
            lda, ldb = na + 3, m + 5
            abuf = np.full((na, lda), np.nan, dtype=dtype); abuf[:, :na] = Adev.T.astype(dtype)   # abuf[j, i] = A(i, j)
            bbuf = np.full((n, ldb), np.nan, dtype=dtype); bbuf[:, :m] = B0.T
            a, b = torch.from_numpy(abuf).cuda(), torch.from_numpy(bbuf).cuda()
            L.check(fn(side, uplo, t, diag, m, n, 2.0, a.data_ptr(), lda, b.data_ptr(), ldb, 0), "trsm")
            torch.cuda.synchronize()
            got = b.cpu().numpy()
            assert np.isnan(got[:, m:]).all(), "padding rows of B were written"
            X = got[:, :m].T
            assert np.isfinite(X).all(), ("read the unreferenced triangle / diagonal", na, side, uplo, t, diag)
            ref = O3.trsm_global(side, uplo, t, diag, 2.0, Aref.astype(dtype), B0)
            err = OG.frobenius_rel_err(X, ref)
            assert err <= 4 * OG.gemm_tolerance(dtype, na), (na, nrhs, side, uplo, t, diag, err)
    x = torch.zeros(16, dtype=torch.float32 if dtype == np.float32 else torch.float64, device="cuda")
    assert fn(2, 0, 0, 0, 4, 4, 1.0, x.data_ptr(), 4, x.data_ptr(), 4, 0) == 2      # bad side
    assert fn(0, 0, 0, 0, 4, 4, 1.0, x.data_ptr(), 3, x.data_ptr(), 4, 0) == 2      # lda < m


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_trsm_1x1(grid1, dtype):
    from tests import dist_cases as dc
    assert dc.check_trsm(grid1, dtype) == 40


def test_level3_errors_1x1(grid1):
    from tests import dist_cases as dc
    assert dc.check_level3_errors(grid1, torch.float64) == 13


def test_level3_matches_reference_outputs_1x1(grid1):
    """This library against the reference's own outputs for the same inputs (its CPU build, 1 x 1)."""
    from tests import dist_cases as dc
    assert dc.check_level3_golden(grid1) == 60


def test_el_level3_check_facade():
    """include/El.hpp: Syrk / Herk / Trrk / LocalTrrk / Symm / Hemm / trapezoid routines as a Hydrogen caller
    writes them (tools/el_level3_check.cpp, built by `make`)."""
    exe = os.path.join(ROOT, "tools", "el_level3_check")
    if not os.path.exists(exe):
        pytest.fail(f"{exe} is missing: run `make` (or __graft_entry__.build()) first")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0 and "EL_LEVEL3_CHECK_PASS" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_gemm_host_streamed_a_ranges_1x1(grid1, dtype):
    """Host-operand Gemm with k large enough (2100 > 4 x 512) for A to be streamed in several k ranges behind the
    first block (engine_gemm.cpp::GemmHost), all orientations, beta = 0 with NaNs in C, k = 0."""
    from tests import dist_cases as dc
    assert dc.check_gemm_host(grid1, dtype, m=260, k=2100) == 12


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_randomised_sweep_1x1(grid1, dtype):
    """Random sizes (0, 1, ragged), orientations, algorithms, alignments, blocksizes, operand distributions for Gemm,
    Copy, Syrk and Trsm — the edge cases of every kernel entry (empty and one-element operands, k = 0, beta = 0)."""
    from tests import dist_cases as dc
    assert dc.check_fuzz(grid1, dtype, iters=60, seed=4321) == 60


@pytest.mark.parametrize("nproc,height", [(2, 1), (4, 2), (8, 2)])
def test_multi_gpu_level3_and_reference_goldens(nproc, height):
    """r x c grids, one process per GPU: the level-3 cases plus this library against the reference's own
    per-rank outputs (ref_dist.npz: 121 redistributions, 23 [CIRC,CIRC] cases, 20 Gemm runs; ref_level3.npz)."""
    if torch.cuda.device_count() < nproc:
        pytest.skip(f"needs {nproc} GPUs")
    port = 29700 + nproc * 10 + height
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(ROOT, "tests", "dist_worker.py"), "--height", str(height), "--level3"]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert res.returncode == 0, res.stdout[-4000:] + res.stderr[-4000:]
    assert "DIST_WORKER_PASS" in res.stdout
