"""One rank of a multi-process CPU run (torch.distributed, backend gloo): the N>1 host path.

Every rank asks libelb200's planner (C ABI, host code) for ITS OWN plan only — as the GPU engine
does — packs with numpy, and exchanges with real collectives over the grid's column (MC), row
(MR) and world (VC) process groups: all_gather for EX_ALLGATHER, batched isend/irecv for
EX_SENDRECV, all_reduce + own slot for the ReduceScatter of AxpyContract.  It then runs SUMMA-C,
SUMMA-A and SUMMA-Dot panel loops from those pieces and checks the result against the oracle.
Launched by tests/test_plan_gloo.py with RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT set."""
from __future__ import annotations

import itertools
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from elemental_b200 import _lib as L  # noqa: E402
from oracle import dist as OD  # noqa: E402
from oracle import gemm as OG  # noqa: E402
from tests import plan_sim as PS  # noqa: E402


class Comms:
    def __init__(self, g: OD.Grid, me: int):
        self.g, self.me = g, me
        self.groups = {}
        # every rank must create every group, in the same order
        for mr in range(g.c):
            ranks = [g.vc(i, mr) for i in range(g.r)]
            grp = dist.new_group(ranks)
            if me in ranks:
                self.groups[L.MC] = (grp, ranks)
        for mc in range(g.r):
            ranks = [g.vc(mc, j) for j in range(g.c)]
            grp = dist.new_group(ranks)
            if me in ranks:
                self.groups[L.MR] = (grp, ranks)
        self.groups[L.VC] = (dist.group.WORLD, list(range(g.p)))


def local_spec(spec: OD.Spec) -> L.DistSpec:
    return PS.spec_c(spec.U, spec.V, spec.col_align, spec.row_align, spec.height, spec.width, spec.root)


def copy(cm: Comms, a_loc: np.ndarray, sa: OD.Spec, sb: OD.Spec) -> np.ndarray:
    g, me = cm.g, cm.me
    p = PS.copy_plan(g.r, g.c, me, local_spec(sa), local_spec(sb))
    b_loc = np.full(sb.local_shape(g, me), np.nan, dtype=a_loc.dtype, order="F")

    def self_copy():
        if p.has_self:
            s, d = PS.seg_tuple(p.self_send.seg), PS.seg_tuple(p.self_recv.seg)
            blk = OD.apply_segments_pack(a_loc, [s[:6] + (0,)], s[2] * s[5])
            OD.apply_segments_unpack(b_loc, [d[:6] + (0,)], blk)

    if p.kind == PS.EX_LOCAL:
        self_copy()
    elif p.kind == PS.EX_ALLGATHER:
        grp, ranks = cm.groups[p.comm_dist]
        s = PS.seg_tuple(p.sends[0].seg)
        slot = torch.from_numpy(OD.apply_segments_pack(a_loc, [s[:6] + (0,)], p.portion))
        outs = [torch.empty_like(slot) for _ in ranks]
        dist.all_gather(outs, slot, group=grp)
        gathered = torch.cat(outs).numpy()
        OD.apply_segments_unpack(b_loc, [PS.seg_tuple(p.recvs[i].seg) for i in range(p.nrecv)], gathered)
    else:
        segs = [PS.seg_tuple(p.sends[i].seg) for i in range(p.nsend)]
        sbuf = torch.from_numpy(OD.apply_segments_pack(a_loc, segs, p.send_elems))
        rbuf = torch.zeros(p.recv_elems, dtype=sbuf.dtype)
        ops = []
        for i in range(p.nsend):
            m = p.sends[i]
            n = m.seg.rows * m.seg.cols
            if n:
                ops.append(dist.P2POp(dist.isend, sbuf[m.seg.buf_offset:m.seg.buf_offset + n], m.peer))
        for i in range(p.nrecv):
            m = p.recvs[i]
            n = m.seg.rows * m.seg.cols
            if n:
                ops.append(dist.P2POp(dist.irecv, rbuf[m.seg.buf_offset:m.seg.buf_offset + n], m.peer))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        OD.apply_segments_unpack(b_loc, [PS.seg_tuple(p.recvs[i].seg) for i in range(p.nrecv)], rbuf.numpy())
        self_copy()
    return b_loc


def axpy_contract(cm: Comms, alpha, a_loc: np.ndarray, sa: OD.Spec, b_loc: np.ndarray, sb: OD.Spec) -> None:
    g, me = cm.g, cm.me
    p = PS.contract_plan(g.r, g.c, me, local_spec(sa), local_spec(sb))
    assert p.valid
    segs = [PS.seg_tuple(p.packs[i].seg) for i in range(p.npack)]
    stage = torch.from_numpy(OD.apply_segments_pack(a_loc, segs, p.group * p.portion))
    if p.group > 1:
        grp, _ranks = cm.groups[p.comm_dist]
        dist.all_reduce(stage, group=grp)  # gloo has no reduce_scatter: reduce all, keep own slot
    slot = stage.numpy()[p.my_index * p.portion:(p.my_index + 1) * p.portion]
    u = PS.seg_tuple(p.unpack)
    OD.apply_segments_unpack(b_loc, [u[:6] + (0,)], slot, alpha=alpha)


def view_local(g, me, loc, s: OD.Spec, i0, i1, j0, j1):
    ns = OD.Spec(s.U, s.V, (s.col_align + i0) % g.stride(s.U), (s.row_align + j0) % g.stride(s.V), i1 - i0, j1 - j0)
    r0 = OD.Length(i0, s.col_shift(g, me), g.stride(s.U))
    c0 = OD.Length(j0, s.row_shift(g, me), g.stride(s.V))
    lh, lw = ns.local_shape(g, me)
    return loc[r0:r0 + lh, c0:c0 + lw], ns


def main() -> int:
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    r = int(os.environ.get("ELB_GRID_HEIGHT", "1"))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    g = OD.Grid(r, world // r)
    cm = Comms(g, rank)
    checks = 0

    # 1. every redistribution pair, ragged size, non-zero alignments
    G = np.arange(13 * 11, dtype=np.float64).reshape(13, 11) + 1
    pairs = [(L.MC, L.MR), (L.MR, L.MC), (L.MC, L.STAR), (L.STAR, L.MC), (L.MR, L.STAR), (L.STAR, L.MR),
             (L.VC, L.STAR), (L.STAR, L.VC), (L.VR, L.STAR), (L.STAR, L.VR), (L.STAR, L.STAR)]
    for (U, V), (U2, V2) in itertools.product(pairs, pairs):
        sa = OD.Spec(U, V, 1 % g.stride(U), 2 % g.stride(V), 13, 11)
        sb = OD.Spec(U2, V2, 3 % g.stride(U2), 1 % g.stride(V2), 13, 11)
        got = copy(cm, OD.local_of(G, sa, g, rank), sa, sb)
        assert np.array_equal(got, OD.local_of(G, sb, g, rank)), (U, V, U2, V2)
        checks += 1

    # 1b. [CIRC,CIRC]: gather to a root, scatter from it, move the root (Copy/Gather.hpp, Scatter.hpp)
    for (U, V) in pairs:
        sa = OD.Spec(U, V, 1 % g.stride(U), 2 % g.stride(V), 13, 11)
        sr = OD.Spec(L.CIRC, L.CIRC, 0, 0, 13, 11, (world - 1))
        sr2 = OD.Spec(L.CIRC, L.CIRC, 0, 0, 13, 11, 0)
        rooted = copy(cm, OD.local_of(G, sa, g, rank), sa, sr)
        assert rooted.shape == ((13, 11) if rank == world - 1 else (0, 0))
        assert np.array_equal(rooted, OD.local_of(G, sr, g, rank))
        moved = copy(cm, rooted, sr, sr2)
        assert np.array_equal(moved, OD.local_of(G, sr2, g, rank))
        sb = OD.Spec(U, V, 3 % g.stride(U), 1 % g.stride(V), 13, 11)
        assert np.array_equal(copy(cm, moved, sr2, sb), OD.local_of(G, sb, g, rank)), (U, V)
        checks += 3

    # 2. SUMMA-C (NN.hpp:330-387): AllGather panels, local product
    rng = np.random.default_rng(11)  # same stream on every rank
    m, n, k, nb = 29, 23, 37, 8
    A, B, C0 = rng.standard_normal((m, k)), rng.standard_normal((k, n)), rng.standard_normal((m, n))
    ref = OG.gemm_global(0, 0, 0.5, A, B, -0.5, C0)
    s_mcmr = lambda h, w: OD.Spec(L.MC, L.MR, 0, 0, h, w)  # noqa: E731
    sA, sB, sC = s_mcmr(m, k), s_mcmr(k, n), s_mcmr(m, n)
    a_loc, b_loc = OD.local_of(A, sA, g, rank), OD.local_of(B, sB, g, rank)
    c_loc = -0.5 * OD.local_of(C0, sC, g, rank)
    for k0 in range(0, k, nb):
        k1 = min(k0 + nb, k)
        a1, sa1 = view_local(g, rank, a_loc, sA, 0, m, k0, k1)
        b1, sb1 = view_local(g, rank, b_loc, sB, k0, k1, 0, n)
        a1_mc_star = copy(cm, a1, sa1, OD.Spec(L.MC, L.STAR, 0, 0, m, k1 - k0))
        b1_star_mr = copy(cm, b1, sb1, OD.Spec(L.STAR, L.MR, 0, 0, k1 - k0, n))
        c_loc += 0.5 * (a1_mc_star @ b1_star_mr)
    assert OG.frobenius_rel_err(c_loc, OD.local_of(ref, sC, g, rank)) < 1e-14
    checks += 1

    # 3. SUMMA-A (NN.hpp:34-108): B1 -> [MR,*], D1[MC,*] = A * B1, RowScatter-contract into C1
    c_loc = -0.5 * OD.local_of(C0, sC, g, rank)
    for j0 in range(0, n, nb):
        j1 = min(j0 + nb, n)
        b1, sb1 = view_local(g, rank, b_loc, sB, 0, k, j0, j1)
        b1_vr = copy(cm, b1, sb1, OD.Spec(L.VR, L.STAR, 0, 0, k, j1 - j0))
        b1_mr = copy(cm, b1_vr, OD.Spec(L.VR, L.STAR, 0, 0, k, j1 - j0), OD.Spec(L.MR, L.STAR, 0, 0, k, j1 - j0))
        d1 = a_loc @ b1_mr  # [MC,*] partial sums over the row communicator
        c1, sc1 = view_local(g, rank, c_loc, sC, 0, m, j0, j1)
        axpy_contract(cm, 0.5, np.asfortranarray(d1), OD.Spec(L.MC, L.STAR, 0, 0, m, j1 - j0), c1, sc1)
    assert OG.frobenius_rel_err(c_loc, OD.local_of(ref, sC, g, rank)) < 1e-14
    checks += 1

    # 4. SUMMA-Dot (NN.hpp:426-482): A -> [*,VC], B -> [VC,*], full [*,*] partial product, Scatter-contract
    a_star_vc = copy(cm, a_loc, sA, OD.Spec(L.STAR, L.VC, 0, 0, m, k))
    b_vc_star = copy(cm, b_loc, sB, OD.Spec(L.VC, L.STAR, 0, 0, k, n))
    c_loc = -0.5 * OD.local_of(C0, sC, g, rank)
    axpy_contract(cm, 0.5, np.asfortranarray(a_star_vc @ b_vc_star), OD.Spec(L.STAR, L.STAR, 0, 0, m, n), c_loc, sC)
    assert OG.frobenius_rel_err(c_loc, OD.local_of(ref, sC, g, rank)) < 1e-14
    checks += 1

    dist.barrier()
    dist.destroy_process_group()
    print(f"GLOO_WORKER_PASS rank={rank} grid={g.r}x{g.c} checks={checks}", flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
