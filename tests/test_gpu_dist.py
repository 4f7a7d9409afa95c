"""GPU parity of the distributed layer through the handle C ABI (Grid / DistMatrix / Gemm).

Single process = 1 x 1 grid (every redistribution degenerates to local lattice copies, every
SUMMA variant to local GEMMs + local contractions). With >= 2 visible GPUs the same cases run on
1x2 / 2x1 / 2x2 ... grids under torchrun (tests/dist_worker.py), one process per GPU over NCCL.
"""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_redistributions_1x1(grid1):
    from tests import dist_cases as dc
    assert dc.check_redistributions(grid1, torch.float64) == 121
    dc.check_views_and_transpose(grid1, torch.float32)
    dc.check_views_and_transpose(grid1, torch.float64)


def test_circ_gather_scatter_1x1(grid1):
    from tests import dist_cases as dc
    assert dc.check_circ(grid1, torch.float64) == 27
    assert dc.check_circ(grid1, torch.float32) == 27


def test_contractions_1x1(grid1):
    from tests import dist_cases as dc
    dc.check_contractions(grid1, torch.float64)
    dc.check_contractions(grid1, torch.float32)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_gemm_all_variants_1x1(grid1, dtype):
    from tests import dist_cases as dc
    assert dc.check_gemm_variants(grid1, dtype) == 40
    dc.check_gemm_proxies(grid1, dtype)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_gemm_reference_driver_config(grid1, dtype):
    """BASELINE config C1 shape class (tests/blas_like/Gemm.cpp defaults: m=n=k=2000 is too slow
    for the oracle in float128, so 600) with the driver's blocksize 96 and BasicGemm's alpha=2,
    beta=3, over every orientation and SUMMA variant."""
    from elemental_b200 import el
    from tests import dist_cases as dc
    for oa in (0, 1):
        for ob in (0, 1):
            for alg in (el.GEMM_SUMMA_A, el.GEMM_SUMMA_B, el.GEMM_SUMMA_C, el.GEMM_SUMMA_DOT):
                err, tol = dc.gemm_case(grid1, dtype, oa, ob, alg, 600, 500, 700, nb=96)
                assert err <= tol, (oa, ob, alg, err, tol)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_gemm_host_operands_1x1(grid1, dtype):
    """Host-operand Gemm (H2D | GEMM | D2H pipeline over column blocks) against the oracle."""
    from tests import dist_cases as dc
    assert dc.check_gemm_host(grid1, dtype) == 12


def test_gemm_errors(grid1):
    from elemental_b200 import el
    from elemental_b200._lib import Elb200Error
    A = el.DistMatrix(grid1, dtype=torch.float32).Resize(8, 9)
    B = el.DistMatrix(grid1, dtype=torch.float32).Resize(10, 7)  # nonconformal with A
    C = el.DistMatrix(grid1, dtype=torch.float32).Resize(8, 7)
    with pytest.raises(Elb200Error):  # reference: LogicError("Nonconformal matrices"), Gemm/NN.hpp:560-568
        el.Gemm(0, 0, 1.0, A, B, 0.0, C)
    B.Resize(9, 7)
    with pytest.raises(Elb200Error):  # Cannon_NN rejects GPU matrices (Gemm/NN.hpp:30-31)
        el.Gemm(0, 0, 1.0, A, B, 0.0, C, el.GEMM_CANNON)
    D = el.DistMatrix(grid1, dtype=torch.float64).Resize(8, 7)
    with pytest.raises(Elb200Error):
        el.Gemm(0, 0, 1.0, A, B, 0.0, D)


def test_gemm_roundtrip_property_large(grid1):
    """Size-independent property at a bench-class size (n=4096 fp32): Gemm is linear in alpha and
    additive over a split of k — C(k) == C(k[:h]) + C(k[h:]) within the fp32 bound."""
    from elemental_b200 import el
    from oracle import gemm as OG
    n, h = 4096, 1536
    gen = torch.Generator(device="cuda").manual_seed(9)
    with torch.cuda.stream(grid1.stream):
        A = el.DistMatrix(grid1, dtype=torch.float32).Resize(n, n)
        B = el.DistMatrix(grid1, dtype=torch.float32).Resize(n, n)
        A.Matrix().copy_((torch.rand((n, n), device="cuda", generator=gen) - 0.5) * 0.2)
        B.Matrix().copy_((torch.rand((n, n), device="cuda", generator=gen) - 0.5) * 0.2)
    C1 = el.DistMatrix(grid1, dtype=torch.float32).Resize(n, n)
    C2 = el.DistMatrix(grid1, dtype=torch.float32).Resize(n, n)
    el.Gemm(0, 0, 0.5, A, B, 0.0, C1, el.GEMM_SUMMA_C)
    el.Gemm(0, 0, 0.5, A.View(0, n, 0, h), B.View(0, h, 0, n), 0.0, C2, el.GEMM_SUMMA_C)
    el.Gemm(0, 0, 0.5, A.View(0, n, h, n), B.View(h, n, 0, n), 1.0, C2, el.GEMM_SUMMA_A)
    grid1.sync()
    with torch.cuda.stream(grid1.stream):
        c1, c2 = C1.Matrix().double(), C2.Matrix().double()
        ref = 0.5 * (A.Matrix().double() @ B.Matrix().double())
        e12 = float(torch.linalg.norm(c1 - c2) / torch.linalg.norm(ref))
        e1 = float(torch.linalg.norm(c1 - ref) / torch.linalg.norm(ref))
    grid1.sync()
    tol = OG.gemm_tolerance(np.float32, n)
    assert e1 <= tol and e12 <= 2 * tol, (e1, e12, tol)


@pytest.mark.parametrize("dtype,alg", [(torch.float32, 6), (torch.float64, 6), (torch.float32, 7)])
def test_gemm_full_size_properties(grid1, dtype, alg):
    """BASELINE configs at FULL size (C2: fp32 n = 16384 SUMMA_C; the per-GPU shape of C3 in fp64;
    SUMMA_Dot on the same operands) checked through size-independent properties, since no CPU
    oracle finishes n = 16384 in seconds:
      * checksum of checksums: C 1 == alpha A (B 1) + beta C0 1 (fp64 matrix-vector products),
      * 96 sampled rows of C against the fp64 product of those rows of A with B,
    both within the north-star bound (1e-5 sqrt(k) fp32, 1e-13 sqrt(k) fp64, relative Frobenius)."""
    from elemental_b200 import el
    from oracle import gemm as OG
    n, alpha, beta = 16384, 0.5, -0.5
    gen = torch.Generator(device="cuda").manual_seed(21)
    mats = []
    with torch.cuda.stream(grid1.stream):
        for _ in range(3):
            M = el.DistMatrix(grid1, dtype=dtype).Resize(n, n)
            M.Matrix().copy_((torch.rand((n, n), device="cuda", dtype=dtype, generator=gen) - 0.5) * 0.2)
            mats.append(M)
        A, B, Cm = mats
        a, b, c = A.Matrix(), B.Matrix(), Cm.Matrix()   # local == global on 1 x 1
        ones = torch.ones(n, device="cuda", dtype=torch.float64)
        rows = torch.randperm(n, device="cuda", generator=gen)[:96]
        c0_sum = c.double() @ ones
        c0_rows = c[rows].double()
    grid1.sync()
    el.Gemm(el.NORMAL, el.NORMAL, alpha, A, B, beta, Cm, alg)
    grid1.sync()
    with torch.cuda.stream(grid1.stream):
        want_sum = alpha * (a.double() @ (b.double() @ ones)) + beta * c0_sum
        got_sum = c.double() @ ones
        want_rows = alpha * (a[rows].double() @ b.double()) + beta * c0_rows
        got_rows = c[rows].double()
        e_sum = float(torch.linalg.norm(got_sum - want_sum) / torch.linalg.norm(want_sum))
        e_rows = float(torch.linalg.norm(got_rows - want_rows) / torch.linalg.norm(want_rows))
    grid1.sync()
    tol = OG.gemm_tolerance(np.float32 if dtype == torch.float32 else np.float64, n)
    assert e_rows <= tol and e_sum <= tol, (e_rows, e_sum, tol)
    for M in mats:
        M.destroy()


def test_half_1x1(grid1):
    """DistMatrix<half> — redistributions bit-exact, collectives, El::Gemm in all orientations / SUMMA variants
    (tcgen05 kind::f16 local products) against the oracle."""
    from tests import dist_cases as dc
    assert dc.check_half(grid1) == 99


def test_collective_exports_1x1(grid1):
    """Every per-collective export (elb200_allgather / reduce_scatter / allreduce / alltoall / sendrecv / bcast) over the
    grid's four communicators; on one rank each is its defined local copy (src/core/imports/mpi/AllGather.hpp:17,
    ReduceScatter.hpp:47-48). The non-degenerate forms run in tests/dist_worker.py and in bench.py's parity pass."""
    from tests import dist_cases as dc
    assert dc.check_collectives(grid1, torch.float64) == 56
    assert dc.check_collectives(grid1, torch.float32) == 56


def test_comm_api_and_python_mirror(grid1):
    """Communicator handles: world bootstrap on one rank, split / dup / free, rank / size queries, argument errors, and
    the el.AllGather ... wrappers on torch tensors."""
    import ctypes as C
    from elemental_b200 import el
    from elemental_b200._lib import lib
    w = C.c_void_p()
    assert lib.elb200_comm_world(C.byref(w)) == 0 and lib.elb200_comm_size(w) == 1 and lib.elb200_comm_rank(w) == 0
    h, d = C.c_void_p(), C.c_void_p()
    assert lib.elb200_comm_split(w, 3, 0, C.byref(h)) == 0 and lib.elb200_comm_size(h) == 1
    assert lib.elb200_comm_dup(h, C.byref(d)) == 0 and lib.elb200_comm_rank(d) == 0
    none = C.c_void_p(1)
    assert lib.elb200_comm_split(w, -1, 0, C.byref(none)) == 0 and not none.value      # MPI_UNDEFINED
    assert lib.elb200_comm_barrier(d) == 0
    assert lib.elb200_comm_free(d) == 0 and lib.elb200_comm_free(h) == 0
    g2 = C.c_void_p()
    assert lib.elb200_grid_create_on(w, 2, 1, -1, C.byref(g2)) == 2                       # height 2 does not divide 1 rank
    assert lib.elb200_grid_create_on(w, 1, 0, -1, C.byref(g2)) == 0                       # ROW_MAJOR 1 x 1
    assert lib.elb200_grid_destroy(g2) == 0
    mc = grid1.MCComm()
    assert (mc.Rank(), mc.Size(), grid1.VRComm().Size()) == (0, 1, 1)
    x = torch.arange(10, dtype=torch.float32, device="cuda")
    y = torch.zeros_like(x)
    torch.cuda.synchronize()
    el.AllGather(x, y, mc)
    el.AllReduce(x, y, grid1.VCComm())
    el.Broadcast(y, 0, grid1.MRComm())
    torch.cuda.synchronize()
    assert torch.equal(x, y)
    assert lib.elb200_sendrecv(x.data_ptr(), 10, 1, y.data_ptr(), 10, 0, 0, mc._h, None) == 2     # peer out of range
    assert lib.elb200_allgather(x.data_ptr(), y.data_ptr(), 10, 7, mc._h, None) == 2              # unknown dtype


def test_matrix_streams_and_multisync(grid1):
    """elb200_dm_set_stream / elb200_dm_stream (SetSyncInfo / SyncInfoFromMatrix) and the MakeMultiSync(C, A, B)
    ordering inside elb200_dist_gemm: operands written on their own streams right before the call are seen."""
    from elemental_b200 import el
    from elemental_b200._lib import lib
    from oracle import gemm as OG
    n = 256
    A = el.DistMatrix(grid1, dtype=torch.float64).Resize(n, n)
    B = el.DistMatrix(grid1, dtype=torch.float64).Resize(n, n)
    Cm = el.DistMatrix(grid1, dtype=torch.float64).Resize(n, n)
    grid1.sync()
    sa, sb = torch.cuda.Stream(), torch.cuda.Stream()
    assert lib.elb200_dm_stream(A._h) == grid1.stream.cuda_stream
    assert lib.elb200_dm_set_stream(A._h, sa.cuda_stream) == 0 and lib.elb200_dm_set_stream(B._h, sb.cuda_stream) == 0
    assert lib.elb200_dm_stream(A._h) == sa.cuda_stream
    a = OG.uniform_matrix(n, n, 5, np.float64)
    b = OG.uniform_matrix(n, n, 6, np.float64)
    with torch.cuda.stream(sa):
        big = torch.randn(4096, 4096, device="cuda") @ torch.randn(4096, 4096, device="cuda")   # keeps sa busy first
        A.Matrix().copy_(torch.from_numpy(a).cuda())
    with torch.cuda.stream(sb):
        B.Matrix().copy_(torch.from_numpy(b).cuda())
    el.Gemm(el.NORMAL, el.NORMAL, 1.0, A, B, 0.0, Cm)
    grid1.sync()
    err = OG.frobenius_rel_err(Cm.get_local(), a @ b)
    assert err <= OG.gemm_tolerance(np.float64, n), err
    assert lib.elb200_dm_set_stream(A._h, None) == 0 and lib.elb200_dm_stream(A._h) == grid1.stream.cuda_stream
    del big


@pytest.mark.parametrize("nproc,height", [(2, 1), (2, 2), (4, 2), (8, 2)])
def test_multi_gpu_grids(nproc, height):
    if torch.cuda.device_count() < nproc:
        pytest.skip(f"needs {nproc} GPUs")
    port = 29600 + nproc * 10 + height
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(ROOT, "tests", "dist_worker.py"), "--height", str(height)]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert res.returncode == 0, res.stdout[-4000:] + res.stderr[-4000:]
    assert "DIST_WORKER_PASS" in res.stdout
