"""El::DistMatrix<Complex<float|double>,U,V,ELEMENT,Device::GPU> and the routines of the hot path on it (SURVEY 8(f) N3;
reference: src/blas_like/level3/Gemm.cpp:504-506 and the complex instantiations under src/core/DistMatrix/ElementMatrix) —
Python mirror of the elb200_zdm_* C ABI (include/elb200.h, csrc/capi_complex.cpp: planar storage, 4M products on the real
kernels).  `bind(lib, check, DmInfo)` returns the class and functions bound to one loaded library, so that
elemental_b200.el (the product library) and tests/hostsim/sim_el.py (the host engine's test double) share this file.
Local matrices cross the boundary as interleaved numpy arrays through elb200_zdm_write_local / read_local."""
from __future__ import annotations

import ctypes as C
from types import SimpleNamespace

import numpy as np

MC, MR = 0, 1
F32, F64 = 0, 1
GEMM_DEFAULT = 0


def bind(lib, check, DmInfo) -> SimpleNamespace:
    class ComplexDistMatrix:
        def __init__(self, grid, U: int = MC, V: int = MR, dtype=np.complex64, _handle=None):
            self.grid = grid
            self.np_dtype = np.dtype(dtype)
            if self.np_dtype not in (np.dtype(np.complex64), np.dtype(np.complex128)):
                raise ValueError("ComplexDistMatrix: dtype must be complex64 or complex128")
            self.real_dtype = F32 if self.np_dtype == np.dtype(np.complex64) else F64
            if _handle is None:
                h = C.c_void_p()
                check(lib.elb200_zdm_create(grid._h, self.real_dtype, U, V, C.byref(h)), "ComplexDistMatrix")
                _handle = h
            self._h = _handle
            self._keep = None

        def _info(self):
            p = C.c_void_p()
            check(lib.elb200_zdm_plane(self._h, 0, C.byref(p)), "ComplexDistMatrix.plane")
            inf = DmInfo()
            check(lib.elb200_dm_info(p, C.byref(inf)), "ComplexDistMatrix.info")
            return inf

        def Height(self): return self._info().height
        def Width(self): return self._info().width
        def LocalHeight(self): return self._info().local_height
        def LocalWidth(self): return self._info().local_width
        def ColAlign(self): return self._info().col_align
        def RowAlign(self): return self._info().row_align
        def ColShift(self): return self._info().col_shift
        def RowShift(self): return self._info().row_shift
        def ColStride(self): return self._info().col_stride
        def RowStride(self): return self._info().row_stride

        def Resize(self, height: int, width: int):
            check(lib.elb200_zdm_resize(self._h, height, width), "Resize"); return self

        def Align(self, colAlign: int, rowAlign: int):
            check(lib.elb200_zdm_align(self._h, colAlign, rowAlign), "Align"); return self

        def AlignWith(self, other: "ComplexDistMatrix"):
            check(lib.elb200_zdm_align_with(self._h, other._h), "AlignWith"); return self

        def SetRoot(self, root: int):
            check(lib.elb200_zdm_set_root(self._h, root), "SetRoot"); return self

        def View(self, i0, i1, j0, j1) -> "ComplexDistMatrix":
            h = C.c_void_p()
            check(lib.elb200_zdm_view(self._h, i0, i1, j0, j1, C.byref(h)), "View")
            v = ComplexDistMatrix(self.grid, dtype=self.np_dtype, _handle=h)
            v._keep = self
            return v

        def set_local(self, local: np.ndarray) -> None:
            """Local matrix <- (LocalHeight x LocalWidth) complex array (interleaved on the way in)."""
            inf = self._info()
            a = np.asfortranarray(np.asarray(local, dtype=self.np_dtype).reshape(inf.local_height, inf.local_width))
            if a.size:
                check(lib.elb200_zdm_write_local(self._h, a.ctypes.data, max(inf.local_height, 1)), "write_local")
                self.grid.sync()  # `a` is a temporary: the copy must have left it

        def get_local(self) -> np.ndarray:
            inf = self._info()
            out = np.zeros((inf.local_height, inf.local_width), dtype=self.np_dtype, order="F")
            if out.size:
                check(lib.elb200_zdm_read_local(self._h, out.ctypes.data, max(inf.local_height, 1)), "read_local")
                self.grid.sync()
            return out

        def from_global(self, G: np.ndarray) -> "ComplexDistMatrix":
            self.Resize(G.shape[0], G.shape[1])
            inf = self._info()
            if inf.local_height and inf.local_width:
                self.set_local(G[inf.col_shift::inf.col_stride, inf.row_shift::inf.row_stride])
            return self

        def destroy(self):
            if getattr(self, "_h", None):
                lib.elb200_zdm_destroy(self._h)
                self._h = None

        def __del__(self):
            try:
                self.destroy()
            except Exception:
                pass

    def _c(z):
        z = complex(z)
        return z.real, z.imag

    def ZCopy(A, B) -> None:
        check(lib.elb200_zdm_copy(A._h, B._h), "Copy")

    def ZTranspose(A, B, conjugate: bool = False) -> None:
        """Transpose / Adjoint (include/El/blas_like/level1/Transpose.hpp, Adjoint.hpp)."""
        check(lib.elb200_zdm_transpose(A._h, B._h, 1 if conjugate else 0), "Transpose")

    def ZZero(A) -> None:
        check(lib.elb200_zdm_zero(A._h), "Zero")

    def ZScale(alpha, A) -> None:
        check(lib.elb200_zdm_scale(*_c(alpha), A._h), "Scale")

    def ZAxpy(alpha, X, Y) -> None:
        check(lib.elb200_zdm_axpy(*_c(alpha), X._h, Y._h), "Axpy")

    def ZAxpyContract(alpha, A, B) -> None:
        check(lib.elb200_zdm_axpy_contract(*_c(alpha), A._h, B._h), "AxpyContract")

    def ZGemm(orientA, orientB, alpha, A, B, beta, C_, alg: int = GEMM_DEFAULT) -> None:
        """El::Gemm for complex DistMatrices; orientation 2 = ADJOINT."""
        check(lib.elb200_zdm_gemm(orientA, orientB, *_c(alpha), A._h, B._h, *_c(beta), C_._h, alg), "Gemm")

    return SimpleNamespace(ComplexDistMatrix=ComplexDistMatrix, ZCopy=ZCopy, ZTranspose=ZTranspose, ZZero=ZZero, ZScale=ZScale,
                           ZAxpy=ZAxpy, ZAxpyContract=ZAxpyContract, ZGemm=ZGemm)
