"""GPU parity tests of the device C ABI (include/elb200.h) against the numpy oracle.

Bit-exact for every kernel without arithmetic (copy, transpose, fill, lattice pack/unpack);
exact-formula equality for axpy/scale (one fma / one multiply per entry); tolerance from
north_star for the GEMMs: relative Frobenius error <= 1e-5*sqrt(k) (fp32), 1e-13*sqrt(k) (fp64).
"""
import ctypes as C
import itertools

import numpy as np
import pytest
import torch

from oracle import dist as OD
from oracle import gemm as OG

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def L():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from elemental_b200 import _lib
    return _lib


_KEEP = []  # device operands stay alive until the test ends (the C ABI calls are asynchronous)


def dev(a):
    t = torch.from_numpy(np.ascontiguousarray(a)).cuda()
    _KEEP.append(t)
    return t


@pytest.fixture(autouse=True)
def _release_operands():
    yield
    if torch.cuda.is_available():
        torch.cuda.synchronize()
    _KEEP.clear()


def colmajor(rows, cols, ld, dtype, seed, fill=np.nan):
    """column-major (ld x cols) buffer, rows >= `rows` padded with NaN (must never be read)."""
    rng = np.random.default_rng(seed)
    buf = np.full((cols, ld), fill, dtype=dtype)
    buf[:, :rows] = rng.uniform(-1, 1, (cols, rows)).astype(dtype)
    return buf  # buf[j, i] == M(i, j)


SUF = {np.float32: "f32", np.float64: "f64"}


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("rows,cols,pad", [(1, 1, 0), (37, 5, 3), (128, 64, 0), (1000, 33, 4), (4096, 257, 0), (5, 70000, 1)])
def test_copy_axpy_scale_fill(L, dtype, rows, cols, pad):
    sfx = SUF[dtype]
    lds, ldd = rows + pad, rows + 2 * pad
    src = colmajor(rows, cols, lds, dtype, 1)
    dst0 = colmajor(rows, cols, ldd, dtype, 2, fill=7.0)
    a = dtype(0.375)
    # copy2d (unit row stride)
    d = dev(dst0)
    st = getattr(L.lib, f"elb200_copy2d_{sfx}")(rows, cols, dev(src).data_ptr(), 1, lds, d.data_ptr(), 1, ldd, None)
    assert st == 0
    exp = dst0.copy(); exp[:, :rows] = src[:, :rows]
    assert np.array_equal(d.cpu().numpy(), exp)
    # axpy2d: Y += a X, one fma per entry
    d = dev(dst0)
    st = getattr(L.lib, f"elb200_axpy2d_{sfx}")(rows, cols, a, dev(src).data_ptr(), 1, lds, d.data_ptr(), 1, ldd, None)
    assert st == 0
    exp = dst0.copy()
    exp[:, :rows] = (np.longdouble(a) * src[:, :rows].astype(np.longdouble) + dst0[:, :rows].astype(np.longdouble)).astype(dtype)
    assert np.array_equal(d.cpu().numpy(), exp)
    # scale2d
    d = dev(dst0)
    assert getattr(L.lib, f"elb200_scale2d_{sfx}")(rows, cols, a, d.data_ptr(), ldd, None) == 0
    exp = dst0.copy(); exp[:, :rows] = a * dst0[:, :rows]
    assert np.array_equal(d.cpu().numpy(), exp)
    # scale by 0 writes exact zeros even over NaN (El::Scale -> Zero)
    nanbuf = np.full((cols, ldd), np.nan, dtype=dtype)
    d = dev(nanbuf)
    assert getattr(L.lib, f"elb200_scale2d_{sfx}")(rows, cols, dtype(0), d.data_ptr(), ldd, None) == 0
    out = d.cpu().numpy()
    assert np.all(out[:, :rows] == 0) and (pad == 0 or np.all(np.isnan(out[:, rows:])))
    # fill2d
    d = dev(dst0)
    assert getattr(L.lib, f"elb200_fill2d_{sfx}")(rows, cols, dtype(2.5), d.data_ptr(), ldd, None) == 0
    exp = dst0.copy(); exp[:, :rows] = 2.5
    assert np.array_equal(d.cpu().numpy(), exp)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_copy_axpy_general_strides(L, dtype):
    """InterleaveMatrix / InterleaveMatrixUpdate with non-unit row strides (Copy/util.hpp:258-264)."""
    sfx = SUF[dtype]
    rng = np.random.default_rng(3)
    S = rng.uniform(-1, 1, 4000).astype(dtype)
    D0 = rng.uniform(-1, 1, 6000).astype(dtype)
    rows, cols = 23, 11
    srs, scs, drs, dcs = 3, 100, 2, 70
    d = dev(D0)
    assert getattr(L.lib, f"elb200_copy2d_{sfx}")(rows, cols, dev(S).data_ptr(), srs, scs, d.data_ptr(), drs, dcs, None) == 0
    exp = D0.copy()
    for j in range(cols):
        exp[j * dcs:j * dcs + rows * drs:drs] = S[j * scs:j * scs + rows * srs:srs]
    assert np.array_equal(d.cpu().numpy(), exp)
    d = dev(D0)
    a = dtype(-1.25)
    assert getattr(L.lib, f"elb200_axpy2d_{sfx}")(rows, cols, a, dev(S).data_ptr(), srs, scs, d.data_ptr(), drs, dcs, None) == 0
    exp = D0.copy()
    for j in range(cols):
        x = S[j * scs:j * scs + rows * srs:srs].astype(np.longdouble)
        y = D0[j * dcs:j * dcs + rows * drs:drs].astype(np.longdouble)
        exp[j * dcs:j * dcs + rows * drs:drs] = (np.longdouble(a) * x + y).astype(dtype)
    assert np.array_equal(d.cpu().numpy(), exp)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("rows,cols", [(1, 1), (31, 33), (64, 8192), (1000, 77), (3, 70001)])
def test_transpose(L, dtype, rows, cols):
    sfx = SUF[dtype]
    lda, ldb = rows + 1, cols + 2
    A = colmajor(rows, cols, lda, dtype, 5)
    B0 = np.full((rows, ldb), 9.0, dtype=dtype)
    d = dev(B0)
    assert getattr(L.lib, f"elb200_transpose_{sfx}")(rows, cols, dev(A).data_ptr(), lda, d.data_ptr(), ldb, None) == 0
    exp = B0.copy(); exp[:, :cols] = A[:, :rows].T  # B(c, r) = A(r, c)
    assert np.array_equal(d.cpu().numpy(), exp)
    # B(rows x cols) += alpha * A2(cols x rows)^T
    A2 = colmajor(cols, rows, cols + 3, dtype, 6)
    Bt = colmajor(rows, cols, rows + 1, dtype, 7, fill=1.0)
    d = dev(Bt)
    a = dtype(0.5)
    assert getattr(L.lib, f"elb200_axpy2d_transpose_{sfx}")(rows, cols, a, dev(A2).data_ptr(), cols + 3, d.data_ptr(), rows + 1, None) == 0
    exp = Bt.copy()
    exp[:, :rows] = (np.longdouble(a) * A2[:, :cols].T.astype(np.longdouble) + Bt[:, :rows].astype(np.longdouble)).astype(dtype)
    assert np.array_equal(d.cpu().numpy(), exp)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("rows,cols", [(132, 200), (64, 64), (4, 516)])
def test_transpose_vector_path(L, dtype, rows, cols):
    """16-byte-aligned operands take the 64 x 64 vectorised kernel (tiles with ragged edges)."""
    sfx = SUF[dtype]
    lda, ldb = rows + 4, cols + 8
    A = colmajor(rows, cols, lda, dtype, 15)
    B0 = np.full((rows, ldb), 9.0, dtype=dtype)
    d = dev(B0)
    assert getattr(L.lib, f"elb200_transpose_{sfx}")(rows, cols, dev(A).data_ptr(), lda, d.data_ptr(), ldb, None) == 0
    exp = B0.copy(); exp[:, :cols] = A[:, :rows].T
    assert np.array_equal(d.cpu().numpy(), exp)
    A2 = colmajor(cols, rows, cols + 4, dtype, 16)
    Bt = colmajor(rows, cols, rows + 4, dtype, 17, fill=1.0)
    d = dev(Bt)
    a = dtype(0.5)
    assert getattr(L.lib, f"elb200_axpy2d_transpose_{sfx}")(rows, cols, a, dev(A2).data_ptr(), cols + 4, d.data_ptr(), rows + 4, None) == 0
    exp = Bt.copy()
    exp[:, :rows] = (np.longdouble(a) * A2[:, :cols].T.astype(np.longdouble) + Bt[:, :rows].astype(np.longdouble)).astype(dtype)
    assert np.array_equal(d.cpu().numpy(), exp)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("aligned", [True, False])
def test_lattice_pack_unpack(L, dtype, aligned):
    """One launch moves every portion: RowStridedUnpack / ColStridedPack / StridedPack shapes
    (Copy/util.hpp:149-230,288-306,667-718), checked bit-exactly against the oracle's
    restatement of the staging layout."""
    sfx = SUF[dtype]
    fn = getattr(L.lib, f"elb200_lattice_{sfx}")
    rng = np.random.default_rng(11)
    H, W = (64, 40) if aligned else (61, 37)
    ld = H if aligned else H + 3
    M = np.zeros((W, ld), dtype=dtype); M[:, :H] = rng.uniform(-1, 1, (W, H)).astype(dtype)
    Mcm = M[:, :H].T  # (H x W) logical view
    cases = []
    stride = 4  # column-interleaved portions (RowStridedPack/Unpack)
    portion = H * OD.MaxLength(W, stride)
    cases.append([(0, 1, H, OD.Shift(k, 1, stride), stride, OD.Length(W, OD.Shift(k, 1, stride), stride), k * portion) for k in range(stride)])
    stride = 3  # row-interleaved portions (ColStridedPack/Unpack)
    portion = OD.MaxLength(H, stride) * W
    cases.append([(OD.Shift(k, 2, stride), stride, OD.Length(H, OD.Shift(k, 2, stride), stride), 0, 1, W, k * portion) for k in range(stride)])
    r, c = 2, 3  # 2-D strided portions (StridedPack, Scatter)
    portion = OD.MaxLength(H, r) * OD.MaxLength(W, c)
    cases.append([(i, r, OD.Length(H, i, r), j, c, OD.Length(W, j, c), (i + r * j) * portion) for j in range(c) for i in range(r)])
    for segs in cases:
        total = max(s[6] + s[2] * s[5] for s in segs) + 5
        arr = (L.Segment * len(segs))(*[L.Segment(*s) for s in segs])
        # pack
        exp = OD.apply_segments_pack(Mcm, segs, total)
        b = dev(np.zeros(total, dtype=dtype))
        assert fn(L.PACK, len(segs), arr, dev(M).data_ptr(), ld, b.data_ptr(), dtype(0), None) == 0
        assert np.array_equal(b.cpu().numpy(), exp)
        # unpack into a fresh matrix reproduces the lattice entries exactly
        m2 = dev(np.zeros_like(M))
        assert fn(L.UNPACK, len(segs), arr, m2.data_ptr(), ld, dev(exp).data_ptr(), dtype(0), None) == 0
        ref = np.zeros((H, W), dtype=dtype); OD.apply_segments_unpack(ref, segs, exp)
        assert np.array_equal(m2.cpu().numpy()[:, :H].T, ref)
        # axpy-unpack (AxpyContract.hpp:334-337)
        m3 = dev(M)
        a = dtype(0.75)
        assert fn(L.UNPACK_AXPY, len(segs), arr, m3.data_ptr(), ld, dev(exp).data_ptr(), a, None) == 0
        ref = Mcm.astype(np.longdouble).copy()
        for (r0, rs, nr, c0, cs, nc, off) in segs:
            ref[r0:r0 + rs * nr:rs, c0:c0 + cs * nc:cs] += np.longdouble(a) * exp[off:off + nr * nc].reshape((nr, nc), order="F").astype(np.longdouble)
        assert np.array_equal(m3.cpu().numpy()[:, :H].T, ref.astype(dtype))


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_peer_reduce_axpy(L, dtype):
    """Receiver-side ReduceScatter fused with the axpy-unpack (AxpyContract.hpp:318-337): the
    lattice of M gets alpha * (sum over sources of their packed block at the same offset).  On one
    GPU the sources are plain device buffers; on a grid they are peer mappings of the arena slots."""
    import ctypes as C
    sfx = SUF[dtype]
    fn = getattr(L.lib, f"elb200_peer_reduce_axpy_{sfx}")
    rng = np.random.default_rng(3)
    H, W, ld = 37, 29, 40
    for (seg, nsrc, off) in (((1, 2, 18, 3, 4, 6, 0), 3, 7), ((0, 1, 37, 0, 1, 29, 0), 4, 7), ((5, 3, 1, 2, 1, 1, 0), 2, 7),
                             ((4, 1, 32, 1, 2, 9, 0), 3, 8), ((0, 1, 36, 0, 1, 29, 0), 5, 0)):  # last two: 16-byte vector path
        r0, rs, nr, c0, cs, nc, _ = seg
        total = off + nr * nc + 3
        srcs = [rng.integers(-8, 9, total).astype(dtype) for _ in range(nsrc)]  # small ints: exact sums
        M = np.zeros((W, ld), dtype=dtype); M[:, :H] = rng.integers(-8, 9, (W, H)).astype(dtype)
        dsrc = [dev(x) for x in srcs]
        ptrs = (C.c_void_p * nsrc)(*[d.data_ptr() for d in dsrc])
        dM = dev(M)
        sg = L.Segment(*seg)
        assert fn(nsrc, ptrs, off, C.byref(sg), dM.data_ptr(), ld, dtype(2.0), None) == 0
        ref = M[:, :H].T.copy()
        blk = sum(x[off:off + nr * nc] for x in srcs).reshape((nr, nc), order="F")
        ref[r0:r0 + rs * nr:rs, c0:c0 + cs * nc:cs] += dtype(2.0) * blk
        got = dM.cpu().numpy()
        assert np.array_equal(got[:, :H].T, ref)
        assert np.array_equal(got[:, H:], M[:, H:])  # padding rows untouched
    bad = L.Segment(0, 0, 1, 0, 1, 1, 0)
    assert fn(1, (C.c_void_p * 1)(0), 0, C.byref(bad), 0, 1, dtype(1), None) != 0  # row_step < 1


GEMM_SHAPES = [(128, 128, 32), (256, 512, 96), (300, 200, 100), (129, 257, 33), (1000, 1000, 1000),
               (64, 40, 500), (4, 4, 4), (513, 1, 77), (2000, 2000, 64)]


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("ta,tb", list(itertools.product([0, 1], [0, 1])))
def test_local_gemm_matches_oracle(L, dtype, ta, tb):
    """gpu_blas::Gemm replacement vs the oracle's global gemm in extended precision, with the
    reference's BasicGemm scalars alpha=2, beta=3 (tests/blas_like/BasicGemm.cpp:249-253)."""
    sfx = SUF[dtype]
    for (m, n, k) in GEMM_SHAPES:
        for pad, beta in ((0, 0.0), (4, 3.0)):
            ar, ac = (k, m) if ta else (m, k)
            br, bc = (n, k) if tb else (k, n)
            A = colmajor(ar, ac, ar + pad, dtype, 21)
            B = colmajor(br, bc, br + pad, dtype, 22)
            C0 = colmajor(m, n, m + pad, dtype, 23, fill=np.nan if beta == 0 else 5.0)
            if beta == 0:
                C0[:, :m] = np.nan  # beta == 0 must not read C
            c, dA, dB = dev(C0), dev(A), dev(B)  # keep the device operands alive across the async call
            st = getattr(L.lib, f"elb200_gemm_{sfx}")(ta, tb, m, n, k, dtype(2.0), dA.data_ptr(), ar + pad,
                                                     dB.data_ptr(), br + pad, dtype(beta), c.data_ptr(), m + pad, None)
            assert st == 0, L.lib.elb200_last_error()
            torch.cuda.synchronize()
            out = c.cpu().numpy()
            ref = OG.gemm_global(ta, tb, 2.0, A[:, :ar].T, B[:, :br].T, beta, np.nan_to_num(C0[:, :m].T), np.longdouble)
            err = OG.frobenius_rel_err(out[:, :m].T, ref)
            assert err <= OG.gemm_tolerance(dtype, k), (m, n, k, ta, tb, err)
            if pad:
                assert np.all(out[:, m:] == 5.0)  # padding rows of C untouched


@pytest.mark.parametrize("ta,tb", list(itertools.product([0, 1], [0, 1])))
def test_gemm_f32_long_k_chunks(L, ta, tb):
    """k beyond 1.5 x 16384 is accumulated in chunks (bounded TMEM accumulation chains, BASELINE C5
    shape class k >> m, n): every orientation must address its chunks correctly, keep the beta
    semantics of the first chunk only, and stay well inside the fp32 bound (checked at a quarter of it)."""
    m, n, k = 200, 136, 70000
    g = torch.Generator(device="cuda").manual_seed(3)
    ar, ac = (k, m) if ta else (m, k)
    br, bc = (n, k) if tb else (k, n)
    A = (torch.rand((ac, ar), device="cuda", generator=g) - 0.5) * 0.2   # column-major ar x ac
    B = (torch.rand((bc, br), device="cuda", generator=g) - 0.5) * 0.2
    C0 = torch.rand((n, m), device="cuda", generator=g) * 0.1
    c = C0.clone()
    assert L.lib.elb200_gemm_f32(ta, tb, m, n, k, 2.0, A.data_ptr(), ar, B.data_ptr(), br, 3.0, c.data_ptr(), m, None) == 0
    torch.cuda.synchronize()
    # a (cols, rows) row-major tensor X is the column-major matrix X.t()
    opA = A.double() if ta else A.double().t()      # m x k
    opB = B.double() if tb else B.double().t()      # k x n
    ref = 2.0 * (opA @ opB) + 3.0 * C0.double().t()
    err = float(torch.linalg.norm(c.double().t() - ref) / torch.linalg.norm(ref))
    assert err <= 0.25 * OG.gemm_tolerance(np.float32, k), err


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_gemm_degenerate_and_errors(L, dtype):
    sfx = SUF[dtype]
    fn = getattr(L.lib, f"elb200_gemm_{sfx}")
    C0 = colmajor(8, 8, 8, dtype, 1)
    c = dev(C0)
    A = dev(colmajor(8, 8, 8, dtype, 2))
    # k == 0 -> C *= beta (Gemm.cpp:240-247)
    assert fn(0, 0, 8, 8, 0, dtype(1), A.data_ptr(), 8, A.data_ptr(), 8, dtype(0.5), c.data_ptr(), 8, None) == 0
    assert np.array_equal(c.cpu().numpy(), dtype(0.5) * C0)
    # empty output is a no-op; nonconformal leading dimension is an argument error
    assert fn(0, 0, 0, 8, 8, dtype(1), A.data_ptr(), 8, A.data_ptr(), 8, dtype(0), c.data_ptr(), 8, None) == 0
    assert fn(0, 0, 8, 8, 8, dtype(1), A.data_ptr(), 4, A.data_ptr(), 8, dtype(0), c.data_ptr(), 8, None) == 2
    assert fn(0, 5, 8, 8, 8, dtype(1), A.data_ptr(), 8, A.data_ptr(), 8, dtype(0), c.data_ptr(), 8, None) == 2


def test_gemm_f32_linearity_large(L):
    """Size-independent property at the bench size class: C(A,B1+B2) == C(A,B1) + C(A,B2) within
    the fp32 bound, and both cta_group variants agree within it (n = 4096)."""
    n = 4096
    g = torch.Generator(device="cuda").manual_seed(5)
    A = (torch.rand((n, n), device="cuda", generator=g) - 0.5) * 0.2
    B1 = (torch.rand((n, n), device="cuda", generator=g) - 0.5) * 0.2
    B2 = (torch.rand((n, n), device="cuda", generator=g) - 0.5) * 0.2
    outs = []
    for Bm, cg in ((B1, 2), (B2, 2), (B1 + B2, 2), (B1 + B2, 1)):
        c = torch.zeros((n, n), device="cuda")
        assert L.lib.elb200_gemm_f32_ex(0, 1, n, n, n, 1.0, A.data_ptr(), n, Bm.data_ptr(), n, 0.0, c.data_ptr(), n, cg, None) == 0
        outs.append(c.double())
    torch.cuda.synchronize()
    tol = OG.gemm_tolerance(np.float32, n)
    s = outs[0] + outs[1]
    assert float(torch.linalg.norm(outs[2] - s) / torch.linalg.norm(s)) <= 2 * tol
    assert float(torch.linalg.norm(outs[2] - outs[3]) / torch.linalg.norm(s)) <= tol
    # torch tensors are row-major: buffer X holds the column-major matrix X^T, so the call above
    # computed C_cm = A^T (B1+B2)  ->  C_buf = C_cm^T = (B1+B2)^T A.  fp64 reference of that:
    ref = (B1 + B2).double().t() @ A.double()
    assert float(torch.linalg.norm(outs[2] - ref) / torch.linalg.norm(ref)) <= tol
