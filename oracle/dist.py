"""ORACLE (test infrastructure, not product code): CPU restatement in numpy of the reference's
element-cyclic distribution algebra and of the redistribution pipelines on the El::Gemm path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this package.  Every function cites the reference file:line it restates (paths relative to the
reference tree).  p "virtual ranks" live in one process: a collective is a loop over ranks.

Parity status: PINNED.  The reference stores no golden vectors for this path, so
the fixtures under tests/golden/ were produced by the reference's own CPU code compiled here
(oracle/ref_build -> oracle/_ref, see oracle/README.md): single-rank El::Gemm outputs and index
tables, and MULTI-RANK runs (1x2, 2x1, 2x2, 2x3, 3x2, 2x4 grids over a small multi-process MPI
stub) giving the per-rank local matrices after all 121 redistributions, after the 23 [CIRC,CIRC]
gather / scatter cases, and after El::Gemm.
tests/test_oracle.py checks this restatement against them: redistributions bit for bit on every
rank, Gemm within the north-star tolerance.
"""
from __future__ import annotations

from dataclasses import dataclass, replace
from typing import Dict, List, Tuple

import numpy as np

MC, MR, VC, VR, STAR, CIRC = 0, 1, 2, 3, 4, 5
NAMES = {MC: "MC", MR: "MR", VC: "VC", VR: "VR", STAR: "STAR", CIRC: "CIRC"}
# the 11 ELEMENT-wrap pairs reachable from [MC,MR] Gemm
PAIRS = [(MC, MR), (MC, STAR), (STAR, MC), (MR, STAR), (STAR, MR), (MR, MC), (VC, STAR),
         (STAR, VC), (VR, STAR), (STAR, VR), (STAR, STAR)]
# + [CIRC,CIRC]: the whole matrix on one VC rank `root` (src/core/DistMatrix/ElementMatrix/CIRC_CIRC.cpp;
# filled and emptied by copy::Gather / copy::Scatter, Copy/Gather.hpp, Copy/Scatter.hpp)
PAIRS_WITH_CIRC = PAIRS + [(CIRC, CIRC)]


# --------------------------------------------------------------------------- indexing
def Mod(a: int, b: int) -> int:
    return a % b  # python's % is already the non-negative modulus the reference's Mod() returns


def Shift(rank: int, align: int, stride: int) -> int:
    """include/El/core/indexing/impl.hpp:244-245  Shift_ = Mod(rank-align, stride)"""
    return Mod(rank - align, stride)


def Length(n: int, shift: int, stride: int) -> int:
    """include/El/core/indexing/impl.hpp:33-36  Length_"""
    return (n - shift - 1) // stride + 1 if n > shift else 0


def MaxLength(n: int, stride: int) -> int:
    """include/El/core/indexing/impl.hpp:63-66  MaxLength_"""
    return (n - 1) // stride + 1 if n > 0 else 0


def DefaultHeight(p: int) -> int:
    """src/core/Grid.cpp:57-63: first divisor of p at or above floor(sqrt(p))"""
    h = int(np.sqrt(float(p)))
    while p % h != 0:
        h += 1
    return h


@dataclass(frozen=True)
class Grid:
    """r x c process grid, COLUMN_MAJOR (src/core/Grid.cpp:95-148)."""
    r: int
    c: int

    @property
    def p(self) -> int:
        return self.r * self.c

    def coords(self, vc: int) -> Tuple[int, int]:
        return vc % self.r, vc // self.r  # (mcRank, mrRank): VC rank = mc + r*mr (Grid.cpp:147)

    def vc(self, mc: int, mr: int) -> int:
        return mc + self.r * mr

    def vr(self, mc: int, mr: int) -> int:
        return mr + self.c * mc  # Grid.cpp:148

    def stride(self, d: int) -> int:
        return {MC: self.r, MR: self.c, VC: self.p, VR: self.p, STAR: 1, CIRC: 1}[d]

    def rank(self, d: int, vc: int) -> int:
        mc, mr = self.coords(vc)
        return {MC: mc, MR: mr, VC: vc, VR: self.vr(mc, mr), STAR: 0, CIRC: 0}[d]


@dataclass(frozen=True)
class Spec:
    """DistData + dims of one DistMatrix (include/El/core/DistMatrix/AbstractDistMatrix.hpp)."""
    U: int
    V: int
    col_align: int = 0
    row_align: int = 0
    height: int = 0
    width: int = 0
    root: int = 0  # [CIRC,CIRC] only: the owning VC rank (AbstractDistMatrix::Root, CrossComm = VCComm)

    def owns(self, vc: int) -> bool:
        return self.U != CIRC or vc == self.root

    def col_shift(self, g: Grid, vc: int) -> int:
        return Shift(g.rank(self.U, vc), self.col_align, g.stride(self.U))

    def row_shift(self, g: Grid, vc: int) -> int:
        return Shift(g.rank(self.V, vc), self.row_align, g.stride(self.V))

    def local_shape(self, g: Grid, vc: int) -> Tuple[int, int]:
        if not self.owns(vc):
            return (0, 0)
        return (Length(self.height, self.col_shift(g, vc), g.stride(self.U)),
                Length(self.width, self.row_shift(g, vc), g.stride(self.V)))


def local_of(G: np.ndarray, spec: Spec, g: Grid, vc: int) -> np.ndarray:
    """Entries of the global matrix G that rank vc stores: local(iLoc,jLoc) = G(colShift +
    iLoc*colStride, rowShift + jLoc*rowStride) (GlobalIndex, indexing/impl.hpp:68-69)."""
    if not spec.owns(vc):
        return np.zeros((0, 0), dtype=G.dtype, order="F")
    return np.array(G[spec.col_shift(g, vc)::g.stride(spec.U), spec.row_shift(g, vc)::g.stride(spec.V)],
                    order="F")


class VDM:
    """A DistMatrix over p virtual ranks: spec + one local column-major array per VC rank."""

    def __init__(self, g: Grid, spec: Spec, locals_: List[np.ndarray]):
        self.g, self.spec, self.loc = g, spec, locals_

    @staticmethod
    def from_global(g: Grid, G: np.ndarray, U: int, V: int, col_align: int = 0, row_align: int = 0,
                    root: int = 0) -> "VDM":
        spec = Spec(U, V, col_align % g.stride(U), row_align % g.stride(V), G.shape[0], G.shape[1],
                    root if U == CIRC else 0)
        return VDM(g, spec, [local_of(G, spec, g, q) for q in range(g.p)])

    @staticmethod
    def zeros(g: Grid, U, V, col_align, row_align, h, w, dtype, root: int = 0) -> "VDM":
        spec = Spec(U, V, col_align % g.stride(U), row_align % g.stride(V), h, w, root if U == CIRC else 0)
        return VDM(g, spec, [np.zeros(spec.local_shape(g, q), dtype=dtype, order="F") for q in range(g.p)])

    def to_global(self) -> np.ndarray:
        """The [*,*] gather the reference's DistMatrix test compares against
        (tests/core/DistMatrix.cpp:30-50). Replicas must agree; the first owner is taken."""
        s, g = self.spec, self.g
        G = np.zeros((s.height, s.width), dtype=self.loc[0].dtype, order="F")
        for q in range(g.p):
            if not s.owns(q):
                continue
            G[s.col_shift(g, q)::g.stride(s.U), s.row_shift(g, q)::g.stride(s.V)] = self.loc[q]
        return G

    def contributions(self) -> np.ndarray:
        """Sum over EVERY rank of its local entries placed at their global positions: what a
        ReduceScatter over the replicated dimensions adds up (AxpyContract.hpp:146-479)."""
        s, g = self.spec, self.g
        G = np.zeros((s.height, s.width), dtype=self.loc[0].dtype, order="F")
        for q in range(g.p):
            if not s.owns(q):
                continue
            G[s.col_shift(g, q)::g.stride(s.U), s.row_shift(g, q)::g.stride(s.V)] += self.loc[q]
        return G

    def view(self, i0, i1, j0, j1) -> "VDM":
        """A(IR(i0,i1), IR(j0,j1)): alignments advance by the offset, local ranges follow
        (include/El/core/View.hpp; AbstractDistMatrix::operator())."""
        s, g = self.spec, self.g
        nspec = Spec(s.U, s.V, (s.col_align + i0) % g.stride(s.U), (s.row_align + j0) % g.stride(s.V),
                     i1 - i0, j1 - j0, s.root)
        out = []
        for q in range(g.p):
            r0 = Length(i0, s.col_shift(g, q), g.stride(s.U))
            c0 = Length(j0, s.row_shift(g, q), g.stride(s.V))
            lh, lw = nspec.local_shape(g, q)
            out.append(self.loc[q][r0:r0 + lh, c0:c0 + lw])  # numpy view: writes go through
        return VDM(g, nspec, out)


def redistribute(A: VDM, U: int, V: int, col_align: int = 0, row_align: int = 0, root: int = 0) -> VDM:
    """B = A for any (U,V)->(U',V'): the RESULT every copy:: routine must produce (exactly the
    entries of the global matrix each rank owns under the target distribution)."""
    return VDM.from_global(A.g, A.to_global(), U, V, col_align, row_align, root)


# --------------------------------------------------------------------------- pipelines, step by step
def row_all_gather(A: VDM) -> VDM:
    """copy::RowAllGather, aligned branch (Copy/RowAllGather.hpp:30-76): (U,V) -> (U,*).
    pack = contiguous local matrix into a maxLocalHeight*maxLocalWidth portion; AllGather over the
    row communicator; RowStridedUnpack (Copy/util.hpp:288-306): column j of portion k lands at
    column Shift(k,rowAlign,rowStride) + j*rowStride."""
    g, s = A.g, A.spec
    stride = g.stride(s.V)
    maxLH, maxLW = MaxLength(s.height, g.stride(s.U)), MaxLength(s.width, stride)
    portion = max(1, maxLH * maxLW)
    tspec = Spec(s.U, STAR, s.col_align, 0, s.height, s.width)
    out = []
    for q in range(g.p):
        lh = tspec.local_shape(g, q)[0]
        # members of q's row communicator = ranks that agree with q on everything U depends on
        members = [t for t in range(g.p) if g.rank(s.U, t) == g.rank(s.U, q) and _same_free(g, s, t, q)]
        members.sort(key=lambda t: g.rank(s.V, t))
        recv = np.zeros(portion * stride, dtype=A.loc[0].dtype)
        for k, t in enumerate(members):
            flat = A.loc[t].flatten(order="F")
            recv[k * portion:k * portion + flat.size] = flat
        B = np.zeros((lh, s.width), dtype=A.loc[0].dtype, order="F")
        for k in range(stride):
            shift = Shift(k, s.row_align, stride)
            lw = Length(s.width, shift, stride)
            B[:, shift::stride] = recv[k * portion:k * portion + lh * lw].reshape((lh, lw), order="F")
        out.append(B)
    return VDM(g, tspec, out)


def _same_free(g: Grid, s: Spec, t: int, q: int) -> bool:
    """t and q agree on the grid coordinates that neither U nor V depends on."""
    dep = {MC: (True, False), MR: (False, True), VC: (True, True), VR: (True, True), STAR: (False, False),
           CIRC: (True, True)}
    du, dv = dep[s.U], dep[s.V]
    (tmc, tmr), (qmc, qmr) = g.coords(t), g.coords(q)
    ok = True
    if not (du[0] or dv[0]):
        ok = ok and tmc == qmc
    if not (du[1] or dv[1]):
        ok = ok and tmr == qmr
    return ok


def col_all_gather(A: VDM) -> VDM:
    """copy::ColAllGather, aligned branch (Copy/ColAllGather.hpp:45-100): (U,V) -> (*,V);
    ColStridedUnpack (Copy/util.hpp:200-230) interleaves ROWS with stride colStride."""
    g, s = A.g, A.spec
    stride = g.stride(s.U)
    maxLH, maxLW = MaxLength(s.height, stride), MaxLength(s.width, g.stride(s.V))
    portion = max(1, maxLH * maxLW)
    tspec = Spec(STAR, s.V, 0, s.row_align, s.height, s.width)
    out = []
    for q in range(g.p):
        lw = tspec.local_shape(g, q)[1]
        members = [t for t in range(g.p) if g.rank(s.V, t) == g.rank(s.V, q) and _same_free(g, s, t, q)]
        members.sort(key=lambda t: g.rank(s.U, t))
        recv = np.zeros(portion * stride, dtype=A.loc[0].dtype)
        for k, t in enumerate(members):
            flat = A.loc[t].flatten(order="F")
            recv[k * portion:k * portion + flat.size] = flat
        B = np.zeros((s.height, lw), dtype=A.loc[0].dtype, order="F")
        for k in range(stride):
            shift = Shift(k, s.col_align, stride)
            lh = Length(s.height, shift, stride)
            B[shift::stride, :] = recv[k * portion:k * portion + lh * lw].reshape((lh, lw), order="F")
        out.append(B)
    return VDM(g, tspec, out)


def axpy_contract(alpha, A: VDM, B: VDM) -> None:
    """AxpyContract (AxpyContract.hpp:483-509): B += alpha * sum over the ranks that hold partial
    sums of the same entry ((U,*)->(U,V) RowScatter :265-423, (*,V)->(U,V) ColScatter :146-263,
    (*,*)->(U,V) Scatter :425-479). The ReduceScatter's summation order is unspecified in the
    reference (NCCL/MPI); here it is ascending VC rank."""
    g, sb = A.g, B.spec
    total = A.contributions()
    for q in range(g.p):
        B.loc[q] += (alpha * total[sb.col_shift(g, q)::g.stride(sb.U), sb.row_shift(g, q)::g.stride(sb.V)]).astype(B.loc[q].dtype)


def transpose(A: VDM) -> VDM:
    """Transpose(A[U,V]) is [V,U] with swapped alignments and transposed local matrices
    (level1/Transpose.hpp:200-250, Transpose/ColAllGather.hpp:15-33)."""
    s = A.spec
    return VDM(A.g, Spec(s.V, s.U, s.row_align, s.col_align, s.width, s.height),
               [np.array(l.T, order="F") for l in A.loc])


# --------------------------------------------------------------------------- staging-buffer layouts
def row_strided_pack(A_loc: np.ndarray, width: int, row_align: int, row_stride: int, portion: int) -> np.ndarray:
    """copy::util::RowStridedPack (Copy/util.hpp:149-165): portion k <- columns
    Shift(k,rowAlign,rowStride)::rowStride of A, packed with leading dimension = height."""
    h = A_loc.shape[0]
    buf = np.zeros(portion * row_stride, dtype=A_loc.dtype)
    for k in range(row_stride):
        shift = Shift(k, row_align, row_stride)
        lw = Length(width, shift, row_stride)
        buf[k * portion:k * portion + h * lw] = A_loc[:, shift::row_stride].flatten(order="F")
    return buf


def col_strided_pack(A_loc: np.ndarray, height: int, col_align: int, col_stride: int, portion: int) -> np.ndarray:
    """copy::util::ColStridedPack (Copy/util.hpp:167-198): portion k <- rows
    Shift(k,colAlign,colStride)::colStride of A."""
    w = A_loc.shape[1]
    buf = np.zeros(portion * col_stride, dtype=A_loc.dtype)
    for k in range(col_stride):
        shift = Shift(k, col_align, col_stride)
        lh = Length(height, shift, col_stride)
        buf[k * portion:k * portion + lh * w] = A_loc[shift::col_stride, :].flatten(order="F")
    return buf


def apply_segments_pack(M: np.ndarray, segs, total: int) -> np.ndarray:
    """What elb200_lattice_*(ELB200_PACK) must write: each segment's lattice, packed column-major
    with leading dimension = rows, at buf_offset (include/elb200.h)."""
    buf = np.zeros(total, dtype=M.dtype)
    for (r0, rs, nr, c0, cs, nc, off) in segs:
        if nr and nc:
            buf[off:off + nr * nc] = M[r0:r0 + rs * nr:rs, c0:c0 + cs * nc:cs].flatten(order="F")
    return buf


def apply_segments_unpack(M: np.ndarray, segs, buf: np.ndarray, alpha=None) -> None:
    for (r0, rs, nr, c0, cs, nc, off) in segs:
        if nr and nc:
            blk = buf[off:off + nr * nc].reshape((nr, nc), order="F")
            if alpha is None:
                M[r0:r0 + rs * nr:rs, c0:c0 + cs * nc:cs] = blk
            else:
                M[r0:r0 + rs * nr:rs, c0:c0 + cs * nc:cs] += (alpha * blk).astype(M.dtype)
