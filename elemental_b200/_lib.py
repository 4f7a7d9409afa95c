"""ctypes binding of libelb200.so — the C ABI declared in include/elb200.h.

The library is the product: if it is missing the import fails loudly (there is no Python or
CPU fallback for any compute entry point).  `torch` is imported first on purpose so that the
NCCL and CUDA runtime libraries torch ships are the ones the process resolves.
"""
from __future__ import annotations

import ctypes as C
import os

import torch  # noqa: F401  (must precede the CDLL load: pins libnccl.so.2 / libcudart.so.12)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libelb200.so")

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} not found: build it with `make` (or `python -c 'import __graft_entry__ as g; "
        "g.build()'`). elemental_b200 has no fallback path without its CUDA library."
    )

lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)

MAX_SEGMENTS = 64
UNIQUE_ID_BYTES = 128
F32, F64, BYTE, F16 = 0, 1, 2, 3
MC, MR, VC, VR, STAR, CIRC = 0, 1, 2, 3, 4, 5
DIST_NAMES = {MC: "MC", MR: "MR", VC: "VC", VR: "VR", STAR: "STAR", CIRC: "CIRC"}
# El::GemmAlgorithm, values as in the reference enum (include/El/blas_like/level3.hpp:23-33)
(GEMM_DEFAULT, GEMM_SUMMA_A_MS, GEMM_SUMMA_A, GEMM_SUMMA_B_MS, GEMM_SUMMA_B, GEMM_SUMMA_C_MS, GEMM_SUMMA_C,
 GEMM_SUMMA_DOT, GEMM_CANNON) = range(9)
PACK, UNPACK, UNPACK_AXPY = 0, 1, 2
LOWER, UPPER = 0, 1  # El::UpperOrLower
LEFT, RIGHT = 0, 1   # El::LeftOrRight
TRAP_SCALE, TRAP_AXPY, TRAP_MAKE = 0, 1, 2
NON_UNIT, UNIT = 0, 1  # El::UnitOrNonUnit

i64 = C.c_int64


class Segment(C.Structure):
    _fields_ = [("row0", i64), ("row_step", i64), ("rows", i64),
                ("col0", i64), ("col_step", i64), ("cols", i64), ("buf_offset", i64)]


class DistSpec(C.Structure):
    _fields_ = [("col_dist", C.c_int), ("row_dist", C.c_int), ("col_align", i64),
                ("row_align", i64), ("height", i64), ("width", i64), ("root", C.c_int)]


class Message(C.Structure):
    _fields_ = [("peer", C.c_int), ("valid", C.c_int), ("seg", Segment)]


class CopyPlan(C.Structure):
    _fields_ = [("kind", C.c_int), ("comm_dist", C.c_int), ("portion", i64),
                ("nsend", C.c_int), ("nrecv", C.c_int), ("has_self", C.c_int),
                ("send_elems", i64), ("recv_elems", i64),
                ("self_send", Message), ("self_recv", Message),
                ("sends", Message * MAX_SEGMENTS), ("recvs", Message * MAX_SEGMENTS)]


class ContractPlan(C.Structure):
    _fields_ = [("valid", C.c_int), ("comm_dist", C.c_int), ("group", C.c_int),
                ("my_index", C.c_int), ("portion", i64), ("npack", C.c_int),
                ("packs", Message * MAX_SEGMENTS), ("unpack", Segment)]


class DmInfo(C.Structure):
    _fields_ = [("dtype", C.c_int), ("col_dist", C.c_int), ("row_dist", C.c_int),
                ("is_view", C.c_int), ("height", i64), ("width", i64), ("col_align", i64),
                ("row_align", i64), ("col_shift", i64), ("row_shift", i64), ("col_stride", i64),
                ("row_stride", i64), ("local_height", i64), ("local_width", i64), ("ldim", i64),
                ("buffer", C.c_void_p), ("root", C.c_int)]


vp = C.c_void_p
f32, f64, ci = C.c_float, C.c_double, C.c_int

# name -> (restype, argtypes); every symbol include/elb200.h declares is listed here and
# tests/test_abi.py checks the two stay in sync.
SIGNATURES = {
    "elb200_status_string": (C.c_char_p, [ci]),
    "elb200_last_error": (C.c_char_p, []),
    "elb200_build_info": (C.c_char_p, []),
    "elb200_launch_count": (C.c_ulonglong, []),
    "elb200_gemm_f32": (ci, [ci, ci, i64, i64, i64, f32, vp, i64, vp, i64, f32, vp, i64, vp]),
    "elb200_gemm_f64": (ci, [ci, ci, i64, i64, i64, f64, vp, i64, vp, i64, f64, vp, i64, vp]),
    "elb200_gemm_f16": (ci, [ci, ci, i64, i64, i64, f32, vp, i64, vp, i64, f32, vp, i64, vp]),
    "elb200_copy2d_f16": (ci, [i64, i64, vp, i64, i64, vp, i64, i64, vp]),
    "elb200_axpy2d_f16": (ci, [i64, i64, f32, vp, i64, i64, vp, i64, i64, vp]),
    "elb200_scale2d_f16": (ci, [i64, i64, f32, vp, i64, vp]),
    "elb200_fill2d_f16": (ci, [i64, i64, f32, vp, i64, vp]),
    "elb200_transpose_f16": (ci, [i64, i64, vp, i64, vp, i64, vp]),
    "elb200_axpy2d_transpose_f16": (ci, [i64, i64, f32, vp, i64, vp, i64, vp]),
    "elb200_lattice_f16": (ci, [ci, ci, C.POINTER(Segment), vp, i64, vp, f32, vp]),
    "elb200_peer_reduce_axpy_f16": (ci, [ci, C.POINTER(vp), i64, C.POINTER(Segment), vp, i64, f32, vp]),
    "elb200_gemm_f32_ex": (ci, [ci, ci, i64, i64, i64, f32, vp, i64, vp, i64, f32, vp, i64, ci, vp]),
    "elb200_gemm_f32_noalloc": (ci, [ci, ci, i64, i64, i64, f32, vp, i64, vp, i64, f32, vp, i64, vp]),
    "elb200_copy2d_f32": (ci, [i64, i64, vp, i64, i64, vp, i64, i64, vp]),
    "elb200_copy2d_f64": (ci, [i64, i64, vp, i64, i64, vp, i64, i64, vp]),
    "elb200_transpose_f32": (ci, [i64, i64, vp, i64, vp, i64, vp]),
    "elb200_transpose_f64": (ci, [i64, i64, vp, i64, vp, i64, vp]),
    "elb200_axpy2d_f32": (ci, [i64, i64, f32, vp, i64, i64, vp, i64, i64, vp]),
    "elb200_axpy2d_f64": (ci, [i64, i64, f64, vp, i64, i64, vp, i64, i64, vp]),
    "elb200_axpy2d_transpose_f32": (ci, [i64, i64, f32, vp, i64, vp, i64, vp]),
    "elb200_axpy2d_transpose_f64": (ci, [i64, i64, f64, vp, i64, vp, i64, vp]),
    "elb200_scale2d_f32": (ci, [i64, i64, f32, vp, i64, vp]),
    "elb200_scale2d_f64": (ci, [i64, i64, f64, vp, i64, vp]),
    "elb200_fill2d_f32": (ci, [i64, i64, f32, vp, i64, vp]),
    "elb200_fill2d_f64": (ci, [i64, i64, f64, vp, i64, vp]),
    "elb200_lattice_f32": (ci, [ci, ci, C.POINTER(Segment), vp, i64, vp, f32, vp]),
    "elb200_lattice_f64": (ci, [ci, ci, C.POINTER(Segment), vp, i64, vp, f64, vp]),
    "elb200_peer_reduce_axpy_f32": (ci, [ci, C.POINTER(vp), i64, C.POINTER(Segment), vp, i64, f32, vp]),
    "elb200_peer_reduce_axpy_f64": (ci, [ci, C.POINTER(vp), i64, C.POINTER(Segment), vp, i64, f64, vp]),
    "elb200_peer_allgather": (ci, [vp, ci, vp, i64, ci, C.POINTER(Segment), vp, i64, vp]),
    "elb200_zdm_create": (ci, [vp, ci, ci, ci, C.POINTER(vp)]),
    "elb200_zdm_destroy": (ci, [vp]),
    "elb200_zdm_plane": (ci, [vp, ci, C.POINTER(vp)]),
    "elb200_zdm_resize": (ci, [vp, i64, i64]),
    "elb200_zdm_align": (ci, [vp, i64, i64]),
    "elb200_zdm_align_with": (ci, [vp, vp]),
    "elb200_zdm_set_root": (ci, [vp, ci]),
    "elb200_zdm_set_stream": (ci, [vp, vp]),
    "elb200_zdm_view": (ci, [vp, i64, i64, i64, i64, C.POINTER(vp)]),
    "elb200_zdm_write_local": (ci, [vp, vp, i64]),
    "elb200_zdm_read_local": (ci, [vp, vp, i64]),
    "elb200_zdm_copy": (ci, [vp, vp]),
    "elb200_zdm_transpose": (ci, [vp, vp, ci]),
    "elb200_zdm_zero": (ci, [vp]),
    "elb200_zdm_scale": (ci, [f64, f64, vp]),
    "elb200_zdm_axpy": (ci, [f64, f64, vp, vp]),
    "elb200_zdm_axpy_contract": (ci, [f64, f64, vp, vp]),
    "elb200_zdm_gemm": (ci, [ci, ci, f64, f64, vp, vp, f64, f64, vp, ci]),
    "elb200_gemm_c32": (ci, [ci, ci, i64, i64, i64, C.POINTER(f32), vp, i64, vp, i64, C.POINTER(f32), vp, i64, vp]),
    "elb200_gemm_c64": (ci, [ci, ci, i64, i64, i64, C.POINTER(f64), vp, i64, vp, i64, C.POINTER(f64), vp, i64, vp]),
    "elb200_peer_exchange": (ci, [vp, ci, ci, C.POINTER(Segment), vp, i64, ci, C.POINTER(Segment), C.POINTER(ci), vp, i64, vp]),
    "elb200_peer_reduce_scatter_f32": (ci, [vp, ci, C.POINTER(Segment), vp, i64, C.POINTER(Segment), vp, i64, f32, vp]),
    "elb200_peer_reduce_scatter_f64": (ci, [vp, ci, C.POINTER(Segment), vp, i64, C.POINTER(Segment), vp, i64, f64, vp]),
    "elb200_trapezoid_f32": (ci, [ci, ci, i64, i64, i64, i64, i64, i64, i64, f32, vp, i64, vp, i64, vp]),
    "elb200_trapezoid_f64": (ci, [ci, ci, i64, i64, i64, i64, i64, i64, i64, f64, vp, i64, vp, i64, vp]),
    "elb200_trsm_f32": (ci, [ci, ci, ci, ci, i64, i64, f32, vp, i64, vp, i64, vp]),
    "elb200_trsm_f64": (ci, [ci, ci, ci, ci, i64, i64, f64, vp, i64, vp, i64, vp]),
    "elb200_plan_copy": (ci, [ci, ci, ci, C.POINTER(DistSpec), C.POINTER(DistSpec), C.POINTER(CopyPlan)]),
    "elb200_plan_contract": (ci, [ci, ci, ci, C.POINTER(DistSpec), C.POINTER(DistSpec), C.POINTER(ContractPlan)]),
    "elb200_plan_send_offset": (i64, [ci, ci, ci, C.POINTER(DistSpec), C.POINTER(DistSpec), ci]),
    "elb200_plan_direct_panel": (ci, [ci, ci, ci, i64, i64, i64, C.POINTER(i64), C.POINTER(i64), C.POINTER(ci),
                                      C.POINTER(i64), C.POINTER(i64)]),
    "elb200_shift": (i64, [i64, i64, i64]),
    "elb200_length": (i64, [i64, i64, i64]),
    "elb200_max_length": (i64, [i64, i64]),
    "elb200_default_grid_height": (ci, [ci]),
    "elb200_gemm_resolve_alg": (ci, [ci, ci, i64, i64, i64, ci]),
    "elb200_unique_id": (ci, [vp]),
    "elb200_grid_create": (ci, [ci, ci, ci, vp, ci, C.POINTER(vp)]),
    "elb200_grid_destroy": (ci, [vp]),
    "elb200_grid_info": (ci, [vp] + [C.POINTER(ci)] * 5),
    "elb200_grid_stream": (vp, [vp]),
    "elb200_grid_sync": (ci, [vp]),
    "elb200_comm_world": (ci, [C.POINTER(vp)]),
    "elb200_comm_create": (ci, [vp, ci, ci, C.POINTER(vp)]),
    "elb200_comm_rank": (ci, [vp]),
    "elb200_comm_size": (ci, [vp]),
    "elb200_comm_split": (ci, [vp, ci, ci, C.POINTER(vp)]),
    "elb200_comm_dup": (ci, [vp, C.POINTER(vp)]),
    "elb200_comm_free": (ci, [vp]),
    "elb200_comm_barrier": (ci, [vp]),
    "elb200_grid_create_on": (ci, [vp, ci, ci, ci, C.POINTER(vp)]),
    "elb200_grid_comm": (ci, [vp, ci, C.POINTER(vp)]),
    "elb200_allgather": (ci, [vp, vp, i64, ci, vp, vp]),
    "elb200_reduce_scatter": (ci, [vp, vp, i64, ci, vp, vp]),
    "elb200_allreduce": (ci, [vp, vp, i64, ci, vp, vp]),
    "elb200_alltoall": (ci, [vp, vp, i64, ci, vp, vp]),
    "elb200_sendrecv": (ci, [vp, i64, ci, vp, i64, ci, ci, vp, vp]),
    "elb200_bcast": (ci, [vp, i64, ci, ci, vp, vp]),
    "elb200_gemm_profile": (None, [ci]),
    "elb200_gemm_profile_collect": (ci, [C.POINTER(f64), C.POINTER(f64), C.POINTER(ci)]),
    "elb200_blocksize": (ci, []),
    "elb200_set_blocksize": (None, [ci]),
    "elb200_push_blocksize": (None, [ci]),
    "elb200_pop_blocksize": (None, []),
    "elb200_dm_create": (ci, [vp, ci, ci, ci, C.POINTER(vp)]),
    "elb200_dm_destroy": (ci, [vp]),
    "elb200_dm_info": (ci, [vp, C.POINTER(DmInfo)]),
    "elb200_dm_align": (ci, [vp, i64, i64]),
    "elb200_dm_align_with": (ci, [vp, vp]),
    "elb200_dm_align_cols": (ci, [vp, i64]),
    "elb200_dm_align_rows": (ci, [vp, i64]),
    "elb200_dm_set_stream": (ci, [vp, vp]),
    "elb200_dm_stream": (vp, [vp]),
    "elb200_dm_resize": (ci, [vp, i64, i64]),
    "elb200_dm_set_root": (ci, [vp, ci]),
    "elb200_dm_write_local": (ci, [vp, vp, i64]),
    "elb200_dm_read_local": (ci, [vp, vp, i64]),
    "elb200_dm_attach": (ci, [vp, i64, i64, i64, i64, vp, i64]),
    "elb200_dm_view": (ci, [vp, i64, i64, i64, i64, C.POINTER(vp)]),
    "elb200_dm_copy": (ci, [vp, vp]),
    "elb200_dm_transpose": (ci, [vp, vp]),
    "elb200_dm_axpy": (ci, [f64, vp, vp]),
    "elb200_dm_scale": (ci, [f64, vp]),
    "elb200_dm_zero": (ci, [vp]),
    "elb200_dm_axpy_contract": (ci, [f64, vp, vp]),
    "elb200_dm_contract": (ci, [vp, vp]),
    "elb200_dm_transpose_axpy_contract": (ci, [f64, vp, vp]),
    "elb200_dm_local_gemm": (ci, [ci, ci, f64, vp, vp, f64, vp]),
    "elb200_dist_gemm": (ci, [ci, ci, f64, vp, vp, f64, vp, ci]),
    "elb200_local_trrk_blocksize": (i64, []),
    "elb200_set_local_trrk_blocksize": (None, [i64]),
    "elb200_dm_scale_trapezoid": (ci, [f64, ci, vp, i64]),
    "elb200_dm_axpy_trapezoid": (ci, [ci, f64, vp, vp, i64]),
    "elb200_dm_make_trapezoidal": (ci, [ci, vp, i64]),
    "elb200_dm_make_symmetric": (ci, [ci, vp]),
    "elb200_dm_local_trrk": (ci, [ci, ci, ci, f64, vp, vp, f64, vp]),
    "elb200_dist_trrk": (ci, [ci, ci, ci, f64, vp, vp, f64, vp]),
    "elb200_dist_syrk": (ci, [ci, ci, f64, vp, f64, vp]),
    "elb200_dist_symm": (ci, [ci, ci, f64, vp, vp, f64, vp]),
    "elb200_dm_local_trsm": (ci, [ci, ci, ci, ci, f64, vp, vp]),
    "elb200_dist_trsm": (ci, [ci, ci, ci, ci, f64, vp, vp, ci]),
    "elb200_dist_gemm_host_f32": (ci, [vp, ci, ci, i64, i64, i64, f32, vp, i64, vp, i64, f32, vp, i64]),
    "elb200_dist_gemm_host_f64": (ci, [vp, ci, ci, i64, i64, i64, f64, vp, i64, vp, i64, f64, vp, i64]),
}

for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)  # AttributeError here == the .so does not export a declared symbol
    _fn.restype = _res
    _fn.argtypes = _args


class Elb200Error(RuntimeError):
    """Non-zero elb200_status. ARG errors mirror El::LogicError, the rest El::RuntimeError."""

    def __init__(self, status: int, where: str):
        self.status = status
        msg = lib.elb200_status_string(status).decode()
        detail = lib.elb200_last_error().decode()
        super().__init__(f"{where}: {msg} ({detail})" if detail else f"{where}: {msg}")


def check(status: int, where: str) -> None:
    if status != 0:
        raise Elb200Error(status, where)
