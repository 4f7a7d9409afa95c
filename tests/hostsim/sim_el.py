"""The `el` interface (elemental_b200/el.py) bound to the hostsim build of the ENGINE
(tests/hostsim/_build/libelb200_hostsim.so) — test infrastructure, see tests/hostsim/sim.hpp.

Same names and argument order as elemental_b200.el so that tests/dist_cases.py runs against either;
local matrices move through the C ABI's host copies (elb200_dm_write_local / elb200_dm_read_local)
as numpy arrays instead of torch CUDA tensors.  One process, one THREAD per rank: `run_grid`
starts r*c threads, each with its own Grid handle (ctypes releases the GIL inside library calls,
so ranks blocked in a collective do not block the others).
"""
from __future__ import annotations

import ctypes as C
import os
import sys
import threading
import traceback
from typing import Callable, Optional

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from elemental_b200 import _lib as L  # noqa: E402  (signatures and struct layouts of include/elb200.h)
from tests.hostsim import build as _build  # noqa: E402

lib = C.CDLL(_build.build(), mode=C.RTLD_LOCAL)
for _name, (_res, _args) in L.SIGNATURES.items():
    _fn = getattr(lib, _name)
    _fn.restype = _res
    _fn.argtypes = _args
assert b"hostsim" in lib.elb200_build_info()

MC, MR, VC, VR, STAR, CIRC = L.MC, L.MR, L.VC, L.VR, L.STAR, L.CIRC
F32, F64, F16 = L.F32, L.F64, L.F16
NORMAL, TRANSPOSE, ADJOINT = 0, 1, 2
(GEMM_DEFAULT, GEMM_SUMMA_A_MS, GEMM_SUMMA_A, GEMM_SUMMA_B_MS, GEMM_SUMMA_B, GEMM_SUMMA_C_MS, GEMM_SUMMA_C,
 GEMM_SUMMA_DOT, GEMM_CANNON) = range(9)
_DTYPES = {torch.float32: F32, torch.float64: F64, torch.float16: F16, np.dtype("float32"): F32, np.dtype("float64"): F64,
           np.dtype("float16"): F16, np.float32: F32, np.float64: F64, np.float16: F16, F32: F32, F64: F64, F16: F16}
_NP = {F32: np.float32, F64: np.float64, F16: np.float16}


class Elb200Error(RuntimeError):
    def __init__(self, status: int, where: str):
        self.status = status
        super().__init__(f"{where}: {lib.elb200_status_string(status).decode()} ({lib.elb200_last_error().decode()})")


def check(status: int, where: str) -> None:
    if status != 0:
        raise Elb200Error(status, where)


def Blocksize() -> int: return lib.elb200_blocksize()
def SetBlocksize(nb: int) -> None: lib.elb200_set_blocksize(int(nb))
def PushBlocksizeStack(nb: int) -> None: lib.elb200_push_blocksize(int(nb))
def PopBlocksizeStack() -> None: lib.elb200_pop_blocksize()
def DefaultGridHeight(p: int) -> int: return lib.elb200_default_grid_height(int(p))


def unique_id():
    buf = (C.c_ubyte * L.UNIQUE_ID_BYTES)()
    check(lib.elb200_unique_id(buf), "elb200_unique_id")
    return buf


class Grid:
    def __init__(self, height: int, rank: int, size: int, idbuf=None, device: int = 0):
        if size % height:
            raise ValueError("Grid height must divide the number of processes")
        self.height, self.width, self.size, self.rank, self.device = height, size // height, size, rank, device
        h = C.c_void_p()
        check(lib.elb200_grid_create(height, self.width, rank, idbuf if size > 1 else None, device, C.byref(h)), "Grid")
        self._h = h

    def Height(self): return self.height
    def Width(self): return self.width
    def Size(self): return self.size
    def Rank(self): return self.rank
    def VCRank(self): return self.rank
    def MCRank(self): return self.rank % self.height
    def MRRank(self): return self.rank // self.height
    def Row(self): return self.MCRank()
    def Col(self): return self.MRRank()
    def VRRank(self): return self.MRRank() + self.width * self.MCRank()

    def sync(self) -> None:
        check(lib.elb200_grid_sync(self._h), "Grid.sync")

    def close(self) -> None:
        if getattr(self, "_h", None):
            lib.elb200_grid_destroy(self._h)
            self._h = None


class DistMatrix:
    def __init__(self, grid: Grid, U: int = MC, V: int = MR, dtype=torch.float32, _handle=None):
        self.grid = grid
        self.dtype = _DTYPES[dtype]
        if _handle is None:
            h = C.c_void_p()
            check(lib.elb200_dm_create(grid._h, self.dtype, U, V, C.byref(h)), "DistMatrix")
            _handle = h
        self._h = _handle
        self._keep = None

    def _info(self) -> L.DmInfo:
        inf = L.DmInfo()
        check(lib.elb200_dm_info(self._h, C.byref(inf)), "DistMatrix.info")
        return inf

    def Height(self): return self._info().height
    def Width(self): return self._info().width
    def LocalHeight(self): return self._info().local_height
    def LocalWidth(self): return self._info().local_width
    def LDim(self): return self._info().ldim
    def ColAlign(self): return self._info().col_align
    def RowAlign(self): return self._info().row_align
    def ColShift(self): return self._info().col_shift
    def RowShift(self): return self._info().row_shift
    def ColStride(self): return self._info().col_stride
    def RowStride(self): return self._info().row_stride
    def ColDist(self): return self._info().col_dist
    def RowDist(self): return self._info().row_dist
    def Root(self): return self._info().root

    def Align(self, colAlign: int, rowAlign: int):
        check(lib.elb200_dm_align(self._h, colAlign, rowAlign), "Align"); return self

    def AlignWith(self, other: "DistMatrix"):
        check(lib.elb200_dm_align_with(self._h, other._h), "AlignWith"); return self

    def SetRoot(self, root: int):
        check(lib.elb200_dm_set_root(self._h, root), "SetRoot"); return self

    def Resize(self, height: int, width: int):
        check(lib.elb200_dm_resize(self._h, height, width), "Resize"); return self

    def View(self, i0, i1, j0, j1) -> "DistMatrix":
        h = C.c_void_p()
        check(lib.elb200_dm_view(self._h, i0, i1, j0, j1, C.byref(h)), "View")
        v = DistMatrix(self.grid, dtype=self.dtype, _handle=h)
        v._keep = self
        return v

    def write_local_from_host(self, host: torch.Tensor) -> None:
        check(lib.elb200_dm_write_local(self._h, host.data_ptr(), host.stride(0) if host.shape[0] > 1 else max(host.shape[1], 1)), "write_local")

    def read_local_to_host(self, host: torch.Tensor) -> None:
        check(lib.elb200_dm_read_local(self._h, host.data_ptr(), host.stride(0) if host.shape[0] > 1 else max(host.shape[1], 1)), "read_local")

    def set_local(self, t) -> None:
        inf = self._info()
        a = np.asfortranarray(np.asarray(t, dtype=_NP[self.dtype]))
        assert a.shape == (inf.local_height, inf.local_width), (a.shape, inf.local_height, inf.local_width)
        if a.size:
            check(lib.elb200_dm_write_local(self._h, a.ctypes.data, max(a.shape[0], 1)), "write_local")
            self.grid.sync()  # `a` must stay alive until the (asynchronous) copy has run

    def get_local(self) -> np.ndarray:
        inf = self._info()
        out = np.full((inf.local_height, inf.local_width), np.nan, dtype=_NP[self.dtype], order="F")
        if out.size:
            check(lib.elb200_dm_read_local(self._h, out.ctypes.data, max(out.shape[0], 1)), "read_local")
        self.grid.sync()
        return out

    def from_global(self, G: np.ndarray) -> "DistMatrix":
        self.Resize(G.shape[0], G.shape[1])
        inf = self._info()
        if inf.local_height == 0 or inf.local_width == 0:
            return self
        self.set_local(np.ascontiguousarray(G[inf.col_shift::inf.col_stride, inf.row_shift::inf.row_stride]))
        return self

    def destroy(self):
        if getattr(self, "_h", None):
            lib.elb200_dm_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass


def Copy(A, B):
    if isinstance(A, ComplexDistMatrix):
        return ZCopy(A, B)
    check(lib.elb200_dm_copy(A._h, B._h), "Copy")
def Transpose(A, B): check(lib.elb200_dm_transpose(A._h, B._h), "Transpose")
def Axpy(alpha, X, Y): check(lib.elb200_dm_axpy(alpha, X._h, Y._h), "Axpy")
def Scale(alpha, A): check(lib.elb200_dm_scale(alpha, A._h), "Scale")
def Zero(A): check(lib.elb200_dm_zero(A._h), "Zero")
def AxpyContract(alpha, A, B): check(lib.elb200_dm_axpy_contract(alpha, A._h, B._h), "AxpyContract")
def Contract(A, B): check(lib.elb200_dm_contract(A._h, B._h), "Contract")
def TransposeAxpyContract(alpha, A, B): check(lib.elb200_dm_transpose_axpy_contract(alpha, A._h, B._h), "TransposeAxpyContract")


def LocalGemm(orientA, orientB, alpha, A, B, beta, C_):
    check(lib.elb200_dm_local_gemm(orientA, orientB, alpha, A._h, B._h, beta, C_._h), "LocalGemm")


def Gemm(orientA, orientB, alpha, A, B, beta, C_, alg: int = GEMM_DEFAULT):
    if isinstance(A, ComplexDistMatrix):
        return ZGemm(orientA, orientB, alpha, A, B, beta, C_, alg)
    check(lib.elb200_dist_gemm(orientA, orientB, alpha, A._h, B._h, beta, C_._h, alg), "Gemm")


# ---- triangular rank-k family (include/El/blas_like/level3.hpp:92-126,239-252,292-313,499-536) ----
LOWER, UPPER = 0, 1  # El::UpperOrLower (include/El/core/types.hpp:511-515)
LEFT, RIGHT = 0, 1   # El::LeftOrRight  (types.hpp:418-422)


def LocalTrrkBlocksize() -> int:
    """El::LocalTrrkBlocksize<T> (src/blas_like/blocksizes.cpp:86-87); one value for both types here."""
    return lib.elb200_local_trrk_blocksize()


def SetLocalTrrkBlocksize(nb: int) -> None:
    lib.elb200_set_local_trrk_blocksize(int(nb))


def ScaleTrapezoid(alpha: float, uplo: int, A, offset: int = 0) -> None:
    """include/El/blas_like/level1/ScaleTrapezoid.hpp:49-96."""
    check(lib.elb200_dm_scale_trapezoid(alpha, uplo, A._h, offset), "ScaleTrapezoid")


def AxpyTrapezoid(uplo: int, alpha: float, X, Y, offset: int = 0) -> None:
    check(lib.elb200_dm_axpy_trapezoid(uplo, alpha, X._h, Y._h, offset), "AxpyTrapezoid")


def MakeTrapezoidal(uplo: int, A, offset: int = 0) -> None:
    check(lib.elb200_dm_make_trapezoidal(uplo, A._h, offset), "MakeTrapezoidal")


def MakeSymmetric(uplo: int, A) -> None:
    """level1/MakeSymmetric.hpp:56-73 (real types: MakeHermitian is the same)."""
    check(lib.elb200_dm_make_symmetric(uplo, A._h), "MakeSymmetric")


MakeHermitian = MakeSymmetric


def LocalTrrk(uplo: int, orientA: int, orientB: int, alpha, A, B, beta, C_) -> None:
    """Trrk/Local.hpp:688-890: triangle of C[MC,MR] = alpha op(A) op(B) + beta C on aligned panels."""
    check(lib.elb200_dm_local_trrk(uplo, orientA, orientB, alpha, A._h, B._h, beta, C_._h), "LocalTrrk")


def Trrk(uplo: int, orientA: int, orientB: int, alpha, A, B, beta, C_) -> None:
    """src/blas_like/level3/Trrk.cpp:100-118."""
    check(lib.elb200_dist_trrk(uplo, orientA, orientB, alpha, A._h, B._h, beta, C_._h), "Trrk")


def Syrk(uplo: int, orientation: int, alpha, A, beta, C_) -> None:
    """src/blas_like/level3/Syrk.cpp:172-190."""
    check(lib.elb200_dist_syrk(uplo, orientation, alpha, A._h, beta, C_._h), "Syrk")


Herk = Syrk  # real types (src/blas_like/level3/Herk.cpp: Syrk with conjugate = true)


def Symm(side: int, uplo: int, alpha, A, B, beta, C_) -> None:
    """src/blas_like/level3/Symm.cpp:56-80."""
    check(lib.elb200_dist_symm(side, uplo, alpha, A._h, B._h, beta, C_._h), "Symm")


Hemm = Symm


NON_UNIT, UNIT = 0, 1  # El::UnitOrNonUnit (include/El/core/types.hpp:489-493)
TRSM_DEFAULT, TRSM_LARGE, TRSM_MEDIUM, TRSM_SMALL = range(4)  # El::TrsmAlgorithm (level3.hpp:394-401)


def LocalTrsm(side: int, uplo: int, orientation: int, diag: int, alpha, A, X, checkIfSingular: bool = False) -> None:
    """src/blas_like/level3/Trsm.cpp:501-526: A is [*,*], X is [*,V] (LEFT) or [U,*] (RIGHT)."""
    check(lib.elb200_dm_local_trsm(side, uplo, orientation, diag, alpha, A._h, X._h), "LocalTrsm")


def Trsm(side: int, uplo: int, orientation: int, diag: int, alpha, A, B, checkIfSingular: bool = False,
         alg: int = TRSM_DEFAULT) -> None:
    """src/blas_like/level3/Trsm.cpp:474-499: B := alpha op(A)^-1 B (LEFT) or alpha B op(A)^-1 (RIGHT)."""
    check(lib.elb200_dist_trsm(side, uplo, orientation, diag, alpha, A._h, B._h, alg), "Trsm")


def GemmHost(grid, orientA, orientB, m, n, k, alpha, A: torch.Tensor, lda, B: torch.Tensor, ldb, beta, C_: torch.Tensor, ldc):
    fn = lib.elb200_dist_gemm_host_f32 if C_.dtype == torch.float32 else lib.elb200_dist_gemm_host_f64
    check(fn(grid._h, orientA, orientB, m, n, k, alpha, A.data_ptr(), lda, B.data_ptr(), ldb, beta, C_.data_ptr(), ldc), "GemmHost")


# complex instantiations: the same mirror file as the product binding, bound to the hostsim library
from elemental_b200 import _zdm as _zdm_mod  # noqa: E402

_z = _zdm_mod.bind(lib, check, L.DmInfo)
ComplexDistMatrix, ZCopy, ZTranspose, ZZero, ZScale, ZAxpy, ZAxpyContract, ZGemm = (
    _z.ComplexDistMatrix, _z.ZCopy, _z.ZTranspose, _z.ZZero, _z.ZScale, _z.ZAxpy, _z.ZAxpyContract, _z.ZGemm)


def run_grid(r: int, c: int, fn: Callable[["Grid"], object], timeout: Optional[float] = None) -> list:
    """Run fn(grid) on every rank of an r x c grid, one thread per rank; returns the per-rank
    results. Any exception on any rank aborts the process (the other ranks would wait for it)."""
    p = r * c
    idbuf = unique_id() if p > 1 else None
    results = [None] * p

    def body(rank: int):
        try:
            g = Grid(r, rank, p, idbuf)
            results[rank] = fn(g)
            g.sync()
            g.close()
        except BaseException:
            sys.stderr.write(f"[hostsim] rank {rank} of {r}x{c} failed:\n{traceback.format_exc()}\n")
            sys.stderr.flush()
            os._exit(3)

    threads = [threading.Thread(target=body, args=(q,), name=f"rank{q}") for q in range(p)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout)
        if t.is_alive():
            sys.stderr.write(f"[hostsim] {t.name} of {r}x{c} did not finish\n")
            os._exit(4)
    return results
