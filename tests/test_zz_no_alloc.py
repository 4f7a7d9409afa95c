"""elb200_gemm_f32_noalloc — the allocation-free twin of elb200_gemm_f32 / elb200_gemm_f32_ex.

The entry point has no counterpart in the reference (it was added
for callers that cannot call cudaMalloc); its contract is "same arguments, same results as the existing
entry point, staging misaligned operands in a fixed device array instead of stream-ordered temporaries".
So the checker here is the EXISTING entry point on the same operands, plus the fp64 product as the oracle.

Collected last on purpose (`zz`). Passes alone in a fresh process and
after the other GPU tests (it shares no state with them: the library's scratch array is only touched by
this entry point).
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

NOALLOC_MAX_ELEMS = 4 * 1024 * 1024  # ELB200_NOALLOC_MAX_ELEMS in include/elb200.h


@pytest.fixture(scope="module")
def L():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from elemental_b200 import _lib
    import re, os
    hdr = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "elb200.h")).read()
    m = re.search(r"#define ELB200_NOALLOC_MAX_ELEMS \((\d+) \* (\d+) \* (\d+)\)", hdr)
    assert m and int(m.group(1)) * int(m.group(2)) * int(m.group(3)) == NOALLOC_MAX_ELEMS
    return _lib


def _operand(rows, cols, ld, offset, seed):
    """Column-major rows x cols matrix inside a flat device buffer, starting `offset` floats in (offset odd
    => base not 16-byte aligned), leading dimension ld. Returns (device buffer, device pointer, host matrix)."""
    rng = np.random.default_rng(seed)
    host = rng.uniform(-1, 1, (rows, cols)).astype(np.float32)
    flat = np.full(offset + ld * cols, np.nan, dtype=np.float32)  # padding is NaN: it must never be read
    view = flat[offset:offset + ld * cols].reshape(cols, ld)
    view[:, :rows] = host.T
    d = torch.from_numpy(flat).cuda()
    return d, d.data_ptr() + 4 * offset, host


def _run_pair(L, ta, tb, m, n, k, a_off, a_pad, b_off, b_pad, seed, alpha=0.75, beta=-1.25):
    a_rows, a_cols = (k, m) if ta else (m, k)
    b_rows, b_cols = (n, k) if tb else (k, n)
    dA, pA, hA = _operand(a_rows, a_cols, a_rows + a_pad, a_off, seed)
    dB, pB, hB = _operand(b_rows, b_cols, b_rows + b_pad, b_off, seed + 1)
    rng = np.random.default_rng(seed + 2)
    c0 = rng.uniform(-1, 1, (n, m)).astype(np.float32)  # c0[j, i] == C(i, j), ldc == m
    c_old, c_new = torch.from_numpy(c0).cuda(), torch.from_numpy(c0).cuda()
    st = L.lib.elb200_gemm_f32_ex(ta, tb, m, n, k, alpha, pA, a_rows + a_pad, pB, b_rows + b_pad, beta,
                                  c_old.data_ptr(), m, 0, None)
    assert st == 0, L.lib.elb200_last_error()
    st = L.lib.elb200_gemm_f32_noalloc(ta, tb, m, n, k, alpha, pA, a_rows + a_pad, pB, b_rows + b_pad, beta,
                                       c_new.data_ptr(), m, None)
    assert st == 0, L.lib.elb200_last_error()
    torch.cuda.synchronize()
    got_old, got_new = c_old.cpu().numpy().T, c_new.cpu().numpy().T
    opA = hA.T if ta else hA
    opB = hB.T if tb else hB
    ref = alpha * (opA.astype(np.float64) @ opB.astype(np.float64)) + beta * c0.T.astype(np.float64)
    tol = 1e-5 * np.sqrt(k)  # the GEMM tolerance of tests/test_gpu_kernels.py (north_star, fp32)
    err_vs_old = np.linalg.norm(got_new.astype(np.float64) - got_old) / np.linalg.norm(got_old)
    err_vs_ref = np.linalg.norm(got_new - ref) / np.linalg.norm(ref)
    assert np.isfinite(got_new).all()
    assert err_vs_old <= tol, (err_vs_old, tol)
    assert err_vs_ref <= tol, (err_vs_ref, tol)
    # same staging values, same kernel, same tile order: the two entry points agree bit for bit
    assert np.array_equal(got_new, got_old)
    del dA, dB


def test_noalloc_matches_existing_entry_on_two_calls_in_a_row(L):
    # call 1: both operands misaligned (odd base offset AND odd leading dimension), N/N
    _run_pair(L, 0, 0, 517, 389, 301, a_off=1, a_pad=2, b_off=3, b_pad=0, seed=11)
    # call 2, straight after, smaller and transposed: the scratch still holds call 1's (larger) operands
    _run_pair(L, 1, 1, 130, 255, 77, a_off=1, a_pad=0, b_off=0, b_pad=1, seed=21)
    # only one side staged, and an aligned pair (pure pass-through)
    _run_pair(L, 0, 1, 256, 128, 64, a_off=0, a_pad=0, b_off=1, b_pad=1, seed=31)
    _run_pair(L, 1, 0, 256, 128, 64, a_off=0, a_pad=0, b_off=0, b_pad=0, seed=41)


def test_noalloc_reserves_no_device_memory(L):
    """After a warm-up call (kernel images and the scratch symbol are resident), further calls leave the amount
    of free device memory unchanged: nothing is allocated, pooled or cached on the call path."""
    m, n, k = 300, 200, 100
    dA, pA, _ = _operand(m, k, m + 1, 1, 5)
    dB, pB, _ = _operand(k, n, k + 1, 1, 6)
    c = torch.zeros((n, m), dtype=torch.float32, device="cuda")
    def call():
        assert L.lib.elb200_gemm_f32_noalloc(0, 0, m, n, k, 1.0, pA, m + 1, pB, k + 1, 0.0, c.data_ptr(), m, None) == 0
    call()
    torch.cuda.synchronize()
    free0, _total = torch.cuda.mem_get_info()
    call(); call()
    torch.cuda.synchronize()
    free1, _total = torch.cuda.mem_get_info()
    assert free1 >= free0, (free0, free1)


def test_noalloc_rejects_oversize_misaligned_operands(L):
    """A misaligned operand beyond ELB200_NOALLOC_MAX_ELEMS floats is an argument error, raised before anything
    is enqueued (C is untouched); the same operand ALIGNED passes straight through whatever its size."""
    m, k, n = 2049, 2048, 8           # A staged as 2052 x 2048 = 4,202,496 floats > 4,194,304
    assert ((m + 3) // 4 * 4) * k > NOALLOC_MAX_ELEMS
    dA, pA, hA = _operand(m, k, m, 1, 7)          # base off by one float
    dB, pB, hB = _operand(k, n, k, 0, 8)
    c = torch.full((n, m), 3.0, dtype=torch.float32, device="cuda")
    st = L.lib.elb200_gemm_f32_noalloc(0, 0, m, n, k, 1.0, pA, m, pB, k, 0.0, c.data_ptr(), m, None)
    assert st == 2  # ELB200_ERR_ARG
    assert b"ELB200_NOALLOC_MAX_ELEMS" in L.lib.elb200_last_error()
    torch.cuda.synchronize()
    assert bool((c == 3.0).all())
    # aligned operand of the same size (ld padded to a multiple of 4, 16-byte aligned base): accepted
    dA2, pA2, hA2 = _operand(m, k, m + 3, 0, 9)
    st = L.lib.elb200_gemm_f32_noalloc(0, 0, m, n, k, 1.0, pA2, m + 3, pB, k, 0.0, c.data_ptr(), m, None)
    assert st == 0, L.lib.elb200_last_error()
    torch.cuda.synchronize()
    ref = hA2.astype(np.float64) @ hB.astype(np.float64)
    got = c.cpu().numpy().T
    assert np.linalg.norm(got - ref) / np.linalg.norm(ref) <= 1e-5 * np.sqrt(k)
