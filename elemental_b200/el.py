"""Python mirror of the reference's `El::` interface for the Gemm path, bound to libelb200.so.

Names, argument order and meaning follow the reference (file:line cited per item) so that the
parity tests read like the reference's own drivers (tests/blas_like/Gemm.cpp, BasicGemm.cpp,
tests/core/DistMatrix.cpp).  All compute happens in the CUDA library; this module only holds
handles.  torch is used for device memory views and for the process-group side channel that
ships the NCCL unique id.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Tuple

import numpy as np
import torch

from . import _lib
from ._lib import (CIRC, F16, F32, F64, MC, MR, STAR, VC, VR, check, lib)  # noqa: F401

NORMAL, TRANSPOSE, ADJOINT = 0, 1, 2  # El::Orientation (include/El/core/types.hpp)
# El::GemmAlgorithm, values as in the reference enum (include/El/blas_like/level3.hpp:23-33)
(GEMM_DEFAULT, GEMM_SUMMA_A_MS, GEMM_SUMMA_A, GEMM_SUMMA_B_MS, GEMM_SUMMA_B, GEMM_SUMMA_C_MS, GEMM_SUMMA_C,
 GEMM_SUMMA_DOT, GEMM_CANNON) = range(9)

_DTYPES = {torch.float32: F32, torch.float64: F64, torch.float16: F16, np.dtype("float32"): F32,
           np.dtype("float64"): F64, np.dtype("float16"): F16, "float32": F32, "float64": F64, "float16": F16,
           F32: F32, F64: F64, F16: F16}
_TORCH = {F32: torch.float32, F64: torch.float64, F16: torch.float16}


def Blocksize() -> int:
    """El::Blocksize (src/blas_like/blocksizes.cpp:35-39)."""
    return lib.elb200_blocksize()


def SetBlocksize(nb: int) -> None:
    lib.elb200_set_blocksize(int(nb))


def PushBlocksizeStack(nb: int) -> None:
    lib.elb200_push_blocksize(int(nb))


def PopBlocksizeStack() -> None:
    lib.elb200_pop_blocksize()


def DefaultGridHeight(p: int) -> int:
    """Grid::DefaultHeight (src/core/Grid.cpp:57-63)."""
    return lib.elb200_default_grid_height(int(p))


class _DevBuf:
    """__cuda_array_interface__ wrapper so torch can view a DistMatrix's local buffer in place."""

    def __init__(self, ptr: int, shape: Tuple[int, ...], typestr: str):
        self.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (ptr, False),
                                         "version": 2}


class Grid:
    """El::Grid(comm, height, COLUMN_MAJOR) (include/El/core/Grid.hpp:15-140).

    One process per GPU. `rank`/`size` default to torch.distributed's (or 0/1); the NCCL unique id
    is created on rank 0 and broadcast through the torch process group.
    """

    def __init__(self, height: Optional[int] = None, rank: Optional[int] = None,
                 size: Optional[int] = None, device: Optional[int] = None):
        import torch.distributed as dist
        if size is None:
            size = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
        if rank is None:
            rank = dist.get_rank() if dist.is_available() and dist.is_initialized() else 0
        if height is None:
            height = DefaultGridHeight(size)
        if size % height:
            raise ValueError("Grid height must divide the number of processes")  # Grid.cpp:87-90
        if device is None:
            device = int(os.environ.get("LOCAL_RANK", "0")) if size > 1 else torch.cuda.current_device()
        self.height, self.width, self.size, self.rank, self.device = height, size // height, size, rank, device
        idbuf = None
        if size > 1:
            raw = torch.zeros(_lib.UNIQUE_ID_BYTES, dtype=torch.uint8)
            if rank == 0:
                tmp = (C.c_ubyte * _lib.UNIQUE_ID_BYTES)()
                check(lib.elb200_unique_id(tmp), "elb200_unique_id")
                raw = torch.tensor(list(tmp), dtype=torch.uint8)
            if dist.get_backend() == "nccl":
                raw = raw.cuda(device)
            dist.broadcast(raw, src=0)
            idbuf = (C.c_ubyte * _lib.UNIQUE_ID_BYTES)(*raw.cpu().tolist())
        h = C.c_void_p()
        check(lib.elb200_grid_create(height, self.width, rank, idbuf, device, C.byref(h)), "Grid")
        self._h = h
        self.stream = torch.cuda.ExternalStream(lib.elb200_grid_stream(h), device=device)

    # accessors named as in the reference (Grid.hpp:40-75)
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

    # Grid::MCComm / MRComm / VCComm / VRComm (Grid.hpp:62-75); VRComm is built on first use (collective)
    def MCComm(self) -> "Comm": return _grid_comm(self, 0)
    def MRComm(self) -> "Comm": return _grid_comm(self, 1)
    def VCComm(self) -> "Comm": return _grid_comm(self, 2)
    def VRComm(self) -> "Comm": return _grid_comm(self, 3)

    def sync(self) -> None:
        check(lib.elb200_grid_sync(self._h), "Grid.sync")

    def close(self) -> None:
        for c in self.__dict__.pop("_comms", {}).values():
            c.free()
        if getattr(self, "_h", None):
            lib.elb200_grid_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Comm:
    """mpi::Comm (include/El/core/imports/mpi/comm.hpp): a handle to a libelb200 communicator — NCCL plus, for a grid's
    MC / MR / VC communicators, the NVLink peer arena. Obtained from Grid.MCComm() ... or Comm.world()."""

    def __init__(self, handle, grid=None):
        self._h, self._grid = handle, grid

    @staticmethod
    def world() -> "Comm":
        h = C.c_void_p()
        check(lib.elb200_comm_world(C.byref(h)), "mpi::COMM_WORLD")
        return Comm(h)

    def Rank(self) -> int: return lib.elb200_comm_rank(self._h)
    def Size(self) -> int: return lib.elb200_comm_size(self._h)

    def free(self) -> None:
        if self._h:
            lib.elb200_comm_free(self._h)
            self._h = None


def _grid_comm(grid: "Grid", which: int) -> Comm:
    cache = grid.__dict__.setdefault("_comms", {})
    if which not in cache:
        h = C.c_void_p()
        check(lib.elb200_grid_comm(grid._h, which, C.byref(h)), "Grid communicator")
        cache[which] = Comm(h, grid)
    return cache[which]


def _coll(fn, name, *args):
    check(fn(*args), name)


# mpi::AllGather / ReduceScatter / AllReduce / AllToAll / SendRecv / Broadcast on torch CUDA tensors
# (src/core/imports/mpi/*.hpp): equal counts, asynchronous on `stream` (a cudaStream_t as int, default stream 0).
def AllGather(sbuf: torch.Tensor, rbuf: torch.Tensor, comm: Comm, stream: int = 0) -> None:
    _coll(lib.elb200_allgather, "mpi::AllGather", sbuf.data_ptr(), rbuf.data_ptr(), sbuf.numel(), _DTYPES[sbuf.dtype], comm._h, stream)


def ReduceScatter(sbuf: torch.Tensor, rbuf: torch.Tensor, comm: Comm, stream: int = 0) -> None:
    _coll(lib.elb200_reduce_scatter, "mpi::ReduceScatter", sbuf.data_ptr(), rbuf.data_ptr(), rbuf.numel() if sbuf.data_ptr() != rbuf.data_ptr() else sbuf.numel() // comm.Size(),
          _DTYPES[sbuf.dtype], comm._h, stream)


def AllReduce(sbuf: torch.Tensor, rbuf: torch.Tensor, comm: Comm, stream: int = 0) -> None:
    _coll(lib.elb200_allreduce, "mpi::AllReduce", sbuf.data_ptr(), rbuf.data_ptr(), sbuf.numel(), _DTYPES[sbuf.dtype], comm._h, stream)


def AllToAll(sbuf: torch.Tensor, rbuf: torch.Tensor, comm: Comm, stream: int = 0) -> None:
    _coll(lib.elb200_alltoall, "mpi::AllToAll", sbuf.data_ptr(), rbuf.data_ptr(), sbuf.numel() // comm.Size(), _DTYPES[sbuf.dtype], comm._h, stream)


def SendRecv(sbuf: torch.Tensor, to: int, rbuf: torch.Tensor, frm: int, comm: Comm, stream: int = 0) -> None:
    _coll(lib.elb200_sendrecv, "mpi::SendRecv", sbuf.data_ptr(), sbuf.numel(), to, rbuf.data_ptr(), rbuf.numel(), frm, _DTYPES[sbuf.dtype], comm._h, stream)


def Broadcast(buf: torch.Tensor, root: int, comm: Comm, stream: int = 0) -> None:
    _coll(lib.elb200_bcast, "mpi::Broadcast", buf.data_ptr(), buf.numel(), _DTYPES[buf.dtype], root, comm._h, stream)


class DistMatrix:
    """El::DistMatrix<T,U,V,ELEMENT,Device::GPU> (include/El/core/DistMatrix/ElementMatrix.hpp)."""

    def __init__(self, grid: Grid, U: int = MC, V: int = MR, dtype=torch.float32, _handle=None):
        self.grid = grid
        self.dtype = _DTYPES[dtype]
        if _handle is None:
            h = C.c_void_p()
            check(lib.elb200_dm_create(grid._h, self.dtype, U, V, C.byref(h)), "DistMatrix")
            _handle = h
        self._h = _handle
        self._keep = None  # attached tensor, kept alive

    def _info(self) -> _lib.DmInfo:
        inf = _lib.DmInfo()
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

    def Align(self, colAlign: int, rowAlign: int):
        check(lib.elb200_dm_align(self._h, colAlign, rowAlign), "Align"); return self

    def AlignWith(self, other: "DistMatrix"):
        check(lib.elb200_dm_align_with(self._h, other._h), "AlignWith"); return self

    def Root(self): return self._info().root

    def SetRoot(self, root: int):
        """AbstractDistMatrix::SetRoot — the VC rank holding a [CIRC,CIRC] matrix."""
        check(lib.elb200_dm_set_root(self._h, root), "SetRoot")
        return self

    def write_local_from_host(self, host: torch.Tensor) -> None:
        """Local matrix <- host tensor laid out column-major as (LocalWidth, ld) rows (ld >= LocalHeight):
        the CPU -> GPU DistMatrix copy of the reference (ProxyDevice.hpp). Async on the grid stream."""
        if host.dim() != 2 or host.stride(1) != 1 or host.dtype != _TORCH[self.dtype] or host.is_cuda:
            raise ValueError("host tensor must be a contiguous-row CPU tensor of the matrix dtype")
        check(lib.elb200_dm_write_local(self._h, host.data_ptr(), host.stride(0) if host.shape[0] > 1 else max(host.shape[1], 1)), "write_local")

    def read_local_to_host(self, host: torch.Tensor) -> None:
        if host.dim() != 2 or host.stride(1) != 1 or host.dtype != _TORCH[self.dtype] or host.is_cuda:
            raise ValueError("host tensor must be a contiguous-row CPU tensor of the matrix dtype")
        check(lib.elb200_dm_read_local(self._h, host.data_ptr(), host.stride(0) if host.shape[0] > 1 else max(host.shape[1], 1)), "read_local")

    def Resize(self, height: int, width: int):
        check(lib.elb200_dm_resize(self._h, height, width), "Resize"); return self

    def Attach(self, height, width, colAlign, rowAlign, tensor: torch.Tensor, ldim: int):
        """ElementalMatrix::Attach: view a caller-owned column-major device buffer."""
        self._keep = tensor
        check(lib.elb200_dm_attach(self._h, height, width, colAlign, rowAlign, tensor.data_ptr(), ldim), "Attach")
        return self

    def View(self, i0, i1, j0, j1) -> "DistMatrix":
        """A(IR(i0,i1), IR(j0,j1)) (zero-copy)."""
        h = C.c_void_p()
        check(lib.elb200_dm_view(self._h, i0, i1, j0, j1, C.byref(h)), "View")
        v = DistMatrix(self.grid, dtype=self.dtype, _handle=h)
        v._keep = self
        return v

    def Matrix(self) -> torch.Tensor:
        """The local matrix as a (LocalHeight x LocalWidth) torch view of the device buffer."""
        inf = self._info()
        lh, lw, ld = inf.local_height, inf.local_width, inf.ldim
        dt = _TORCH[self.dtype]
        if lh == 0 or lw == 0 or not inf.buffer:
            return torch.empty((lh, lw), dtype=dt, device=f"cuda:{self.grid.device}")
        typestr = {F32: "<f4", F64: "<f8", F16: "<f2"}[self.dtype]
        flat = torch.as_tensor(_DevBuf(inf.buffer, (lw * ld,), typestr), device=f"cuda:{self.grid.device}")
        return flat.view(lw, ld)[:, :lh].t()  # column-major (ld x lw) seen as rows x cols

    def set_local(self, t) -> None:
        with torch.cuda.stream(self.grid.stream):
            self.Matrix().copy_(torch.as_tensor(t, dtype=_TORCH[self.dtype]).to(f"cuda:{self.grid.device}", non_blocking=True))

    def get_local(self) -> np.ndarray:
        with torch.cuda.stream(self.grid.stream):
            out = self.Matrix().contiguous().cpu()
        return out.numpy()

    def from_global(self, G: np.ndarray) -> "DistMatrix":
        """Resize to G's shape and load this rank's entries of the global matrix G."""
        self.Resize(G.shape[0], G.shape[1])
        inf = self._info()
        if inf.local_height == 0 or inf.local_width == 0:
            return self  # e.g. a [CIRC,CIRC] matrix away from its root
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


def Copy(A: DistMatrix, B: DistMatrix) -> None:
    """B = A (DistMatrix::operator=; src/core/DistMatrix/ElementMatrix/*.cpp)."""
    if isinstance(A, ComplexDistMatrix):
        return ZCopy(A, B)
    check(lib.elb200_dm_copy(A._h, B._h), "Copy")


def Transpose(A: DistMatrix, B: DistMatrix) -> None:
    """include/El/blas_like/level1/Transpose.hpp:200-250."""
    check(lib.elb200_dm_transpose(A._h, B._h), "Transpose")


def Axpy(alpha: float, X: DistMatrix, Y: DistMatrix) -> None:
    check(lib.elb200_dm_axpy(alpha, X._h, Y._h), "Axpy")


def Scale(alpha: float, A: DistMatrix) -> None:
    check(lib.elb200_dm_scale(alpha, A._h), "Scale")


def Zero(A: DistMatrix) -> None:
    check(lib.elb200_dm_zero(A._h), "Zero")


def AxpyContract(alpha: float, A: DistMatrix, B: DistMatrix) -> None:
    """include/El/blas_like/level1/AxpyContract.hpp:483-544."""
    check(lib.elb200_dm_axpy_contract(alpha, A._h, B._h), "AxpyContract")


def Contract(A: DistMatrix, B: DistMatrix) -> None:
    check(lib.elb200_dm_contract(A._h, B._h), "Contract")


def TransposeAxpyContract(alpha: float, A: DistMatrix, B: DistMatrix) -> None:
    check(lib.elb200_dm_transpose_axpy_contract(alpha, A._h, B._h), "TransposeAxpyContract")


def LocalGemm(orientA, orientB, alpha, A: DistMatrix, B: DistMatrix, beta, C_: DistMatrix) -> None:
    """src/blas_like/level3/Gemm.cpp:318-426."""
    check(lib.elb200_dm_local_gemm(orientA, orientB, alpha, A._h, B._h, beta, C_._h), "LocalGemm")


def Gemm(orientA, orientB, alpha, A: DistMatrix, B: DistMatrix, beta, C_: DistMatrix,
         alg: int = GEMM_DEFAULT) -> None:
    """El::Gemm(orientA, orientB, alpha, A, B, beta, C, alg) (src/blas_like/level3/Gemm.cpp:273-302)."""
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


def GemmHost(grid: "Grid", orientA, orientB, m: int, n: int, k: int, alpha, A: torch.Tensor, lda: int,
             B: torch.Tensor, ldb: int, beta, C_: torch.Tensor, ldc: int) -> None:
    """El::Gemm on HOST operands: A, B, C_ are CPU tensors (pinned recommended) holding this rank's
    column-major local parts of zero-aligned [MC,MR] matrices; they stream through the GPU in a
    H2D | gather | GEMM | D2H pipeline and C_ is complete on return (elb200_dist_gemm_host_*)."""
    for t in (A, B, C_):
        if t.is_cuda:
            raise ValueError("GemmHost takes host tensors; use Gemm for device-resident DistMatrices")
    if not (A.dtype == B.dtype == C_.dtype) or C_.dtype not in (torch.float32, torch.float64):
        raise TypeError("GemmHost: float32 or float64 operands of one dtype")
    fn = lib.elb200_dist_gemm_host_f32 if C_.dtype == torch.float32 else lib.elb200_dist_gemm_host_f64
    check(fn(grid._h, orientA, orientB, m, n, k, alpha, A.data_ptr(), lda, B.data_ptr(), ldb, beta,
             C_.data_ptr(), ldc), "GemmHost")


def gemm_local(orientA, orientB, alpha, A: torch.Tensor, lda, B: torch.Tensor, ldb, beta,
               C_: torch.Tensor, ldc, m, n, k, stream: int = 0, cta_group: int = 0) -> None:
    """gpu_blas::Gemm replacement on raw column-major device buffers (GPU_BLAS_impl.hpp:417-444)."""
    if C_.dtype == torch.float32:
        st = lib.elb200_gemm_f32_ex(orientA, orientB, m, n, k, alpha, A.data_ptr(), lda, B.data_ptr(), ldb,
                                    beta, C_.data_ptr(), ldc, cta_group, stream)
    else:
        st = lib.elb200_gemm_f64(orientA, orientB, m, n, k, alpha, A.data_ptr(), lda, B.data_ptr(), ldb,
                                 beta, C_.data_ptr(), ldc, stream)
    check(st, "gemm_local")


# ---- complex instantiations: planar storage, 4M on the real kernels (csrc/capi_complex.cpp) ----
from . import _zdm as _zdm_mod  # noqa: E402

_z = _zdm_mod.bind(lib, check, _lib.DmInfo)
ComplexDistMatrix, ZCopy, ZTranspose, ZZero, ZScale, ZAxpy, ZAxpyContract, ZGemm = (
    _z.ComplexDistMatrix, _z.ZCopy, _z.ZTranspose, _z.ZZero, _z.ZScale, _z.ZAxpy, _z.ZAxpyContract, _z.ZGemm)


def gemm_local_complex(orientA, orientB, alpha, A: torch.Tensor, lda, B: torch.Tensor, ldb, beta, C_: torch.Tensor, ldc, m, n, k,
                       stream: int = 0) -> None:
    """Gemm(Matrix<Complex<T>,Device::GPU>) on raw interleaved column-major device buffers (the cublasCgemm / Zgemm call of
    src/hydrogen/device/cuBLAS_API.cpp:419-427): elb200_gemm_c32 / _c64."""
    if C_.dtype not in (torch.complex64, torch.complex128) or not (A.dtype == B.dtype == C_.dtype):
        raise TypeError("gemm_local_complex: complex64 or complex128 operands of one dtype")
    R = C.c_float if C_.dtype == torch.complex64 else C.c_double
    al, be = complex(alpha), complex(beta)
    fn = lib.elb200_gemm_c32 if C_.dtype == torch.complex64 else lib.elb200_gemm_c64
    check(fn(orientA, orientB, m, n, k, (R * 2)(al.real, al.imag), A.data_ptr(), lda, B.data_ptr(), ldb, (R * 2)(be.real, be.imag),
             C_.data_ptr(), ldc, stream), "gemm_local_complex")
