"""Host-side execution of libelb200's redistribution plans WITHOUT a GPU: the planner
(elb200_plan_copy / elb200_plan_contract, pure host code behind the C ABI) is called for every
virtual rank of an r x c grid and the resulting pack -> exchange -> unpack is carried out on numpy
arrays.  What the lattice kernel must do with a segment table is oracle.dist.apply_segments_*;
the exchange itself is the textbook AllGather / grouped send-recv / ReduceScatter.  The GPU
engine executes the very same plans with elb200_lattice_* and NCCL (engine_redist.cpp)."""
from __future__ import annotations

import ctypes as C
from typing import List

import numpy as np

from elemental_b200 import _lib as L
from oracle import dist as OD

EX_LOCAL, EX_ALLGATHER, EX_SENDRECV = 0, 1, 2


def spec_c(U, V, ca, ra, h, w, root=0) -> L.DistSpec:
    return L.DistSpec(U, V, ca, ra, h, w, root)


def seg_tuple(s: L.Segment):
    return (s.row0, s.row_step, s.rows, s.col0, s.col_step, s.cols, s.buf_offset)


def copy_plan(r, c, vc, A: L.DistSpec, B: L.DistSpec) -> L.CopyPlan:
    p = L.CopyPlan()
    L.check(L.lib.elb200_plan_copy(r, c, vc, C.byref(A), C.byref(B), C.byref(p)), "plan_copy")
    return p


def contract_plan(r, c, vc, A: L.DistSpec, B: L.DistSpec) -> L.ContractPlan:
    p = L.ContractPlan()
    L.check(L.lib.elb200_plan_contract(r, c, vc, C.byref(A), C.byref(B), C.byref(p)), "plan_contract")
    return p


def comm_index(g: OD.Grid, comm_dist: int, vc: int) -> int:
    """Rank of grid member vc inside the MC (column), MR (row) or VC (world) communicator —
    the ncclCommSplit keys of engine_core.cpp."""
    mc, mr = g.coords(vc)
    return {L.MC: mc, L.MR: mr, L.VC: vc}[comm_dist]


def comm_members(g: OD.Grid, comm_dist: int, vc: int) -> List[int]:
    mc, mr = g.coords(vc)
    if comm_dist == L.MC:
        return [g.vc(i, mr) for i in range(g.r)]
    if comm_dist == L.MR:
        return [g.vc(mc, j) for j in range(g.c)]
    return list(range(g.p))


def run_copy(g: OD.Grid, A: OD.VDM, U2, V2, ca2, ra2, root2=0) -> OD.VDM:
    """B[U2,V2] = A executed from the plans of all g.p ranks."""
    h, w = A.spec.height, A.spec.width
    sa = spec_c(A.spec.U, A.spec.V, A.spec.col_align, A.spec.row_align, h, w, A.spec.root)
    sb = spec_c(U2, V2, ca2, ra2, h, w, root2)
    B = OD.VDM.zeros(g, U2, V2, ca2, ra2, h, w, A.loc[0].dtype, root2)
    for q in range(g.p):
        B.loc[q][...] = np.nan
    plans = [copy_plan(g.r, g.c, q, sa, sb) for q in range(g.p)]
    kinds = {p.kind for p in plans}
    assert len(kinds) == 1, f"ranks disagree on the exchange kind: {kinds}"  # a collective needs all ranks
    kind = kinds.pop()

    def self_copy(q, p):
        if p.has_self:
            s, d = seg_tuple(p.self_send.seg), seg_tuple(p.self_recv.seg)
            assert (s[2], s[5]) == (d[2], d[5])
            blk = OD.apply_segments_pack(A.loc[q], [s[:6] + (0,)], s[2] * s[5])
            OD.apply_segments_unpack(B.loc[q], [d[:6] + (0,)], blk)

    if kind == EX_LOCAL:
        for q, p in enumerate(plans):
            self_copy(q, p)
        return B
    if kind == EX_ALLGATHER:
        slots = {}
        for q, p in enumerate(plans):
            assert p.nsend == 1
            s = seg_tuple(p.sends[0].seg)
            slots[q] = OD.apply_segments_pack(A.loc[q], [s[:6] + (0,)], p.portion)
        for q, p in enumerate(plans):
            members = comm_members(g, p.comm_dist, q)
            assert p.nrecv == len(members)
            gathered = np.concatenate([slots[m] for m in members])  # AllGather: slot i from comm rank i
            OD.apply_segments_unpack(B.loc[q], [seg_tuple(p.recvs[i].seg) for i in range(p.nrecv)], gathered)
        return B
    # grouped point-to-point: match every send with the receive posted for it
    sent = {}
    for q, p in enumerate(plans):
        # a receiver can locate its message inside the sender's staging on its own (peer-memory pull)
        for i in range(p.nsend):
            assert L.lib.elb200_plan_send_offset(g.r, g.c, q, C.byref(sa), C.byref(sb), p.sends[i].peer) == p.sends[i].seg.buf_offset
        segs = [seg_tuple(p.sends[i].seg) for i in range(p.nsend)]
        buf = OD.apply_segments_pack(A.loc[q], segs, p.send_elems)
        for i in range(p.nsend):
            m = p.sends[i]
            n = m.seg.rows * m.seg.cols
            assert (q, m.peer) not in sent, "one message per ordered pair"
            sent[(q, m.peer)] = buf[m.seg.buf_offset:m.seg.buf_offset + n]
    for q, p in enumerate(plans):
        rbuf = np.zeros(p.recv_elems, dtype=A.loc[0].dtype)
        for i in range(p.nrecv):
            m = p.recvs[i]
            n = m.seg.rows * m.seg.cols
            msg = sent.pop((m.peer, q))
            assert msg.size == n, "send and receive sizes must agree"
            rbuf[m.seg.buf_offset:m.seg.buf_offset + n] = msg
        OD.apply_segments_unpack(B.loc[q], [seg_tuple(p.recvs[i].seg) for i in range(p.nrecv)], rbuf)
        self_copy(q, p)
    assert not sent, f"unmatched sends: {list(sent)}"
    return B


def run_axpy_contract(g: OD.Grid, alpha, A: OD.VDM, B: OD.VDM) -> bool:
    """B += alpha * sum-over-replicas(A) from the contract plans; False if the planner says the
    pair needs a realigning Copy first (plan.valid == 0)."""
    h, w = A.spec.height, A.spec.width
    sa = spec_c(A.spec.U, A.spec.V, A.spec.col_align, A.spec.row_align, h, w)
    sb = spec_c(B.spec.U, B.spec.V, B.spec.col_align, B.spec.row_align, h, w)
    plans = [contract_plan(g.r, g.c, q, sa, sb) for q in range(g.p)]
    if not all(p.valid for p in plans):
        assert not any(p.valid for p in plans)
        return False
    staged = {}
    for q, p in enumerate(plans):
        assert p.npack == p.group and p.my_index == (comm_index(g, p.comm_dist, q) if p.group > 1 else 0)
        segs = [seg_tuple(p.packs[i].seg) for i in range(p.npack)]
        staged[q] = OD.apply_segments_pack(A.loc[q], segs, p.group * p.portion)
    for q, p in enumerate(plans):
        members = comm_members(g, p.comm_dist, q) if p.group > 1 else [q]
        i = p.my_index
        slot = sum(staged[m][i * p.portion:(i + 1) * p.portion] for m in members)  # ReduceScatter(sum)
        u = seg_tuple(p.unpack)
        OD.apply_segments_unpack(B.loc[q], [u[:6] + (0,)], slot, alpha=alpha)
    return True
