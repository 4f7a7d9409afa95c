// Distributed BLAS-1 of the Gemm path: Copy (any (U,V)->(U',V')), TransposeCopy, AxpyContract,
// Contract, TransposeAxpyContract, Axpy, Scale, Zero. One generic lattice-exchange engine
// (plan.hpp) replaces the reference's per-pair routines; AllGather and ReduceScatter shapes run
// on the matching NCCL collective over the grid's MC / MR / VC communicator.
#include <algorithm>
#include <cstdlib>

#include "engine.hpp"
#include "engine_dev.hpp"

namespace elb {
namespace {

template <typename T>
void run_lattice(int mode, const std::vector<Message>& msgs, T* M, int64_t ldm, T* buf, T alpha,
                 cudaStream_t s) {
    elb200_segment segs[ELB200_MAX_SEGMENTS];
    int n = 0;
    for (const Message& m : msgs) {
        if (m.count() == 0) continue;
        segs[n++] = m.seg;
        if (n == ELB200_MAX_SEGMENTS) { Dev<T>::lattice(mode, n, segs, M, ldm, buf, alpha, s); n = 0; }
    }
    if (n) Dev<T>::lattice(mode, n, segs, M, ldm, buf, alpha, s);
}

// dst lattice <- src lattice, both on this rank (same number of rows/cols by construction)
template <typename T>
void lattice_to_lattice(const elb200_segment& a, const T* A, int64_t lda, const elb200_segment& b, T* B,
                        int64_t ldb, cudaStream_t s) {
    Dev<T>::copy2d(a.rows, a.cols, A + a.row0 + a.col0 * lda, a.row_step, a.col_step * lda,
                   B + b.row0 + b.col0 * ldb, b.row_step, b.col_step * ldb, s);
}

// Unconstrained alignments of B follow A where the distributions relate (reference: the
// AlignColsAndResize / AlignAndResize calls at the top of every copy:: routine).
template <typename T>
void adopt_alignment(const DistMat<T>& A, DistMat<T>& B) {
    const DistSpec t = ResolveCopySpec(A.grid->shape, A.spec, B.spec, B.col_constrained, B.row_constrained);
    B.spec.col_align = t.col_align;
    B.spec.row_align = t.row_align;
}

}  // namespace

// Messages that keep less than this in an arena slot are pure latency and take the fused one-kernel forms of
// peer_kernels.cu (ELB200_FUSED_BYTES overrides the 4 MiB default, 0 disables them). Every rank evaluates the bound on
// rank-invariant extents, so all members of a collective choose the same form.
static int64_t fused_bytes() {
    static const int64_t v = [] { const char* e = std::getenv("ELB200_FUSED_BYTES"); return e ? std::atoll(e) : (4ll << 20); }();
    return v;
}
template <typename T> struct FusedReduce { static constexpr bool ok = false; };
template <> struct FusedReduce<float> {
    static constexpr bool ok = true;
    static int run(const elb200_peer_gather* pg, int n, const elb200_segment* packs, const float* A, int64_t lda, const elb200_segment* un, float* B,
                   int64_t ldb, float alpha, cudaStream_t s) { return elb200_peer_reduce_scatter_f32(pg, n, packs, A, lda, un, B, ldb, alpha, s); }
};
template <> struct FusedReduce<double> {
    static constexpr bool ok = true;
    static int run(const elb200_peer_gather* pg, int n, const elb200_segment* packs, const double* A, int64_t lda, const elb200_segment* un, double* B,
                   int64_t ldb, double alpha, cudaStream_t s) { return elb200_peer_reduce_scatter_f64(pg, n, packs, A, lda, un, B, ldb, alpha, s); }
};

template <typename T>
void Copy(const DistMat<T>& A, DistMat<T>& B) {
    if (A.grid != B.grid) throw EngineError(ELB200_ERR_ARG, "Copy: matrices live on different grids");
    Grid& g = *A.grid;
    adopt_alignment(A, B);
    B.Resize(A.spec.height, A.spec.width);
    if (A.spec.height == 0 || A.spec.width == 0) return;
    cudaStream_t s = B.stream;
    const CopyPlan plan = plan_copy(g.shape, g.me, A.spec, B.spec);
    T* Abuf = const_cast<T*>(A.buf);

    if (plan.kind == EX_LOCAL) {
        if (plan.has_self)
            lattice_to_lattice<T>(plan.self_send.seg, A.buf, A.ldim, plan.self_recv.seg, B.buf, B.ldim, s);
        return;
    }
    if (plan.kind == EX_ALLGATHER) {
        ncclComm_t comm = g.comm_for(plan.comm_dist);
        if (!comm) throw EngineError(ELB200_ERR_STATE, "Copy: grid has no NCCL communicator");
        const int gsz = int(plan.recvs.size());
        const int64_t mine = plan.sends[0].count();
        // ship straight out of A's local matrix when it is one contiguous full-size block
        const auto& s0 = plan.sends[0].seg;
        const bool contiguous = s0.row_step == 1 && s0.col_step == 1 && (A.ldim == s0.rows || s0.cols <= 1);
        if (g.arena && size_t(gsz) * plan.portion * sizeof(T) <= g.arena->slot_bytes()) {
            // NVLink peer-memory gather (peer.hpp): copy engines + stream memops, no SM involved,
            // so it runs underneath the persistent GEMM of the previous panel.
            const std::vector<int> members = g.comm_members(plan.comm_dist);
            const int idx = g.comm_index(plan.comm_dist);
            Scratch<T> slot(contiguous ? 0 : mine, s);
            const T* send = A.buf + s0.row0 + s0.col0 * A.ldim;
            if (!contiguous && mine > 0) {
                Message one = plan.sends[0];
                one.seg.buf_offset = 0;
                run_lattice<T>(ELB200_PACK, std::vector<Message>{one}, Abuf, A.ldim, slot.p, T(0), s);
                send = slot.p;
            }
            // Large blocks (>= 4 MiB per member; ELB200_AG_SPLIT_BYTES overrides, 0 disables): do not wait for the whole
            // gather. My own block goes straight from A into B (no copy into my own slot) while the peers' blocks are in
            // flight, and each peer's block is unpacked as soon as ITS arrival flag is up. The sends leave in ROTATION
            // (peer.hpp AllGatherRotate), one after the other at the full link rate, so the blocks land staggered and block d
            // is unpacked while block d + 1 is on the wire. Measured on 2 x 2 B200s (profiles/r02_dist_sweep_4gpu*.json):
            // 1 GiB [MC,MR] -> [MC,*] 0.541 -> 0.475 ms, 256 MiB 0.167 -> 0.141 ms; all sends at once (no rotation) gained
            // nothing and was removed. Small messages are latency-bound and keep the one-wait, one-launch form.
            static const int64_t split_bytes = [] { const char* e = std::getenv("ELB200_AG_SPLIT_BYTES"); return e ? std::atoll(e) : (4ll << 20); }();
            if (split_bytes > 0 && plan.portion * int64_t(sizeof(T)) >= split_bytes && plan.recvs[size_t(idx)].peer == g.vc_rank) {
                PeerArena::Ticket t = g.arena->AllGatherRotate(Grid::arena_channel(plan.comm_dist), members, idx, send,
                                                               size_t(mine) * sizeof(T), size_t(plan.portion) * sizeof(T), s);
                // the rotation leaves a copy lane forked from `s`: join it on EVERY exit, a throwing unpack included, so the
                // arena's fork state is never stale for the next operation and the pack scratch is not freed under the lane
                struct Join {
                    PeerArena* a; cudaStream_t st; bool done = false;
                    ~Join() { if (!done) { try { a->AllGatherEnd(st); } catch (...) {} } }
                } join{g.arena, s};
                if (mine > 0) lattice_to_lattice<T>(s0, A.buf, A.ldim, plan.recvs[size_t(idx)].seg, B.buf, B.ldim, s);
                for (int d = 1; d < gsz; ++d) {
                    const int i = (idx - d + gsz) % gsz;  // arrival order under the rotation
                    g.arena->WaitMember(t, members, idx, i, s);
                    run_lattice<T>(ELB200_UNPACK, std::vector<Message>{plan.recvs[size_t(i)]}, B.buf, B.ldim, static_cast<T*>(t.data), T(0), s);
                }
                g.arena->AllGatherEnd(s);  // my sends are done: `send` (A or the pack scratch) may be reused
                join.done = true;
                g.arena->Release(t, members, idx, s);
                return;
            }
            // Small blocks: the whole gather and its unpack in ONE kernel over peer memory (peer_kernels.cu) instead of ~30
            // stream operations — these sizes are pure latency (1 MiB on 1 x 2: 40 -> 25 us).
            if (int64_t(gsz) * plan.portion * int64_t(sizeof(T)) < fused_bytes() && gsz <= ELB200_MAX_SEGMENTS) {
                elb200_segment segs[ELB200_MAX_SEGMENTS];
                int nseg = 0;
                for (const Message& m : plan.recvs)
                    if (m.count() > 0) segs[nseg++] = m.seg;
                g.arena->FusedAllGather(Grid::arena_channel(plan.comm_dist), members, idx, int(sizeof(T)), send, mine, plan.portion, nseg, segs,
                                        B.buf, B.ldim, s);
                return;
            }
            PeerArena::Ticket t = g.arena->AllGather(Grid::arena_channel(plan.comm_dist), members, idx, send,
                                                     size_t(mine) * sizeof(T), size_t(plan.portion) * sizeof(T), s);
            run_lattice<T>(ELB200_UNPACK, plan.recvs, B.buf, B.ldim, static_cast<T*>(t.data), T(0), s);
            g.arena->Release(t, members, idx, s);
            return;
        }
        const bool direct = mine == plan.portion && (A.ldim == plan.sends[0].seg.rows || plan.sends[0].seg.cols <= 1);
        Scratch<T> stage((gsz + (direct ? 0 : 1)) * plan.portion, s);
        T* recv = stage.p;
        const T* send = A.buf;
        if (!direct) {
            T* slot = stage.p + int64_t(gsz) * plan.portion;
            if (mine > 0) {
                std::vector<Message> one{plan.sends[0]};
                run_lattice<T>(ELB200_PACK, one, Abuf, A.ldim, slot, T(0), s);
            }
            send = slot;
        }
        g.NcclBegin(s);
        nccl_check(ncclAllGather(send, recv, size_t(plan.portion), Dev<T>::nccl, comm, s), "ncclAllGather");
        g.NcclEnd(s);
        run_lattice<T>(ELB200_UNPACK, plan.recvs, B.buf, B.ldim, recv, T(0), s);
        return;
    }
    // general exchange: pack -> point-to-point -> unpack
    if (!g.world) throw EngineError(ELB200_ERR_STATE, "Copy: grid has no NCCL communicator");
    // NVLink pull (peer.hpp): every rank packs ITS sends straight into its own arena slot, signals,
    // and the receivers fetch their messages from the senders' slots with the copy engines — no
    // NCCL kernels, no SMs during the transfer. Taken when the largest possible send staging fits a
    // slot, by a bound every rank evaluates identically: a whole local matrix of A ...
    constexpr bool pull = true;  // measured at 2 x 4, 1 GiB [MC,MR] -> [*,VR]: 328 GB/s per GPU against 177 with NCCL point-to-point
    // ... times the number of ranks that hold a copy of each entry under B's distribution: a sender packs one
    // message per destination, so towards a replicated target (e.g. [MC,MR] -> [MR,*]) its staging is that many
    // times its local matrix.
    const size_t copies = B.spec.col == CIRC ? 1 : size_t(g.size()) / (size_t(dist_stride(g.shape, B.spec.col)) * size_t(dist_stride(g.shape, B.spec.row)));
    const size_t max_local_a = size_t(MaxLength(A.spec.height, dist_stride(g.shape, A.spec.col))) *
                               size_t(MaxLength(A.spec.width, dist_stride(g.shape, A.spec.row))) * sizeof(T) * std::max<size_t>(copies, 1);
    if (pull && g.arena && max_local_a <= g.arena->slot_bytes()) {
        const std::vector<int> members = g.comm_members(VC);
        const int idx = g.vc_rank;
        if (int64_t(max_local_a) < fused_bytes() && g.size() <= ELB200_MAX_SEGMENTS && plan.sends.size() <= size_t(ELB200_MAX_SEGMENTS) &&
            plan.recvs.size() <= size_t(ELB200_MAX_SEGMENTS)) {
            // small: pack into my slot, flags, and the unpack straight out of the senders' slots in ONE kernel
            elb200_segment snd[ELB200_MAX_SEGMENTS], rcv[ELB200_MAX_SEGMENTS];
            int from[ELB200_MAX_SEGMENTS];
            int ns = 0, nr = 0;
            for (const Message& m : plan.sends) snd[ns++] = m.seg;
            for (const Message& m : plan.recvs) {
                rcv[nr] = m.seg;
                // where the sender packed the message for me inside its slot
                rcv[nr].buf_offset = plan_send_offset(g.shape, coord_of_vc(g.shape, m.peer), A.spec, B.spec, g.vc_rank);
                from[nr++] = m.peer;  // VC rank == index among the world channel's members
            }
            const elb200_peer_gather pg = g.arena->FusedBegin(Grid::arena_channel(VC), members, idx, 0, size_t(plan.send_elems) * sizeof(T));
            throw_status(elb200_peer_exchange(&pg, int(sizeof(T)), ns, snd, A.buf, A.ldim, nr, rcv, from, B.buf, B.ldim, s), "elb200_peer_exchange");
            if (plan.has_self)
                lattice_to_lattice<T>(plan.self_send.seg, A.buf, A.ldim, plan.self_recv.seg, B.buf, B.ldim, s);
            return;
        }
        PeerArena::Ticket t = g.arena->Publish(Grid::arena_channel(VC), members, idx, size_t(plan.send_elems) * sizeof(T), s);
        run_lattice<T>(ELB200_PACK, plan.sends, Abuf, A.ldim, static_cast<T*>(t.data), T(0), s);
        g.arena->Signal(t, members, idx, s);
        Scratch<T> rbuf(plan.recv_elems, s);
        std::vector<PeerArena::PullOp> pulls;
        for (const Message& m : plan.recvs) {
            // where the sender packed the message for me inside its slot
            const int64_t off = plan_send_offset(g.shape, coord_of_vc(g.shape, m.peer), A.spec, B.spec, g.vc_rank);
            pulls.push_back({m.peer, size_t(off) * sizeof(T), rbuf.p + m.seg.buf_offset, size_t(m.count()) * sizeof(T)});
        }
        g.arena->Pull(t, pulls, s);
        g.arena->Release(t, members, idx, s);  // stream order: the acks leave after my pulls completed
        run_lattice<T>(ELB200_UNPACK, plan.recvs, B.buf, B.ldim, rbuf.p, T(0), s);
        if (plan.has_self)
            lattice_to_lattice<T>(plan.self_send.seg, A.buf, A.ldim, plan.self_recv.seg, B.buf, B.ldim, s);
        return;
    }
    Scratch<T> sbuf(plan.send_elems, s), rbuf(plan.recv_elems, s);
    run_lattice<T>(ELB200_PACK, plan.sends, Abuf, A.ldim, sbuf.p, T(0), s);
    g.NcclBegin(s);
    nccl_check(ncclGroupStart(), "ncclGroupStart");
    for (const Message& m : plan.sends)
        nccl_check(ncclSend(sbuf.p + m.seg.buf_offset, size_t(m.count()), Dev<T>::nccl, m.peer, g.world, s), "ncclSend");
    for (const Message& m : plan.recvs)
        nccl_check(ncclRecv(rbuf.p + m.seg.buf_offset, size_t(m.count()), Dev<T>::nccl, m.peer, g.world, s), "ncclRecv");
    nccl_check(ncclGroupEnd(), "ncclGroupEnd");
    g.NcclEnd(s);
    run_lattice<T>(ELB200_UNPACK, plan.recvs, B.buf, B.ldim, rbuf.p, T(0), s);
    if (plan.has_self)
        lattice_to_lattice<T>(plan.self_send.seg, A.buf, A.ldim, plan.self_recv.seg, B.buf, B.ldim, s);
}

// B = A^T. The transpose of A[U,V] is a [V,U] matrix whose local matrix is the transpose of A's
// (reference Transpose(ElementalMatrix), level1/Transpose.hpp:200-250: local transpose into a
// [V,U] temporary, then Copy).
template <typename T>
void TransposeCopy(const DistMat<T>& A, DistMat<T>& B) {
    Grid& g = *A.grid;
    cudaStream_t s = B.stream;
    const bool same = B.spec.col == A.spec.row && B.spec.row == A.spec.col && B.spec.root == A.spec.root &&
                      (!B.col_constrained || B.spec.col_align == A.spec.row_align) &&
                      (!B.row_constrained || B.spec.row_align == A.spec.col_align);
    if (same) {
        B.spec.col_align = A.spec.row_align;
        B.spec.row_align = A.spec.col_align;
        B.Resize(A.spec.width, A.spec.height);
        Dev<T>::transpose(A.lh(), A.lw(), A.buf, A.ldim, B.buf, B.ldim, s);
        return;
    }
    DistMat<T> At(g, A.spec.row, A.spec.col, s);
    At.spec.root = A.spec.root;
    At.Align(A.spec.row_align, A.spec.col_align);
    At.Resize(A.spec.width, A.spec.height);
    Dev<T>::transpose(A.lh(), A.lw(), A.buf, A.ldim, At.buf, At.ldim, s);
    Copy(At, B);
}

template <typename T>
void Axpy(T alpha, const DistMat<T>& X, DistMat<T>& Y) {
    if (X.spec.height != Y.spec.height || X.spec.width != Y.spec.width)
        throw EngineError(ELB200_ERR_ARG, "Axpy: nonconformal dimensions");
    if (X.spec.col == Y.spec.col && X.spec.row == Y.spec.row && X.spec.col_align == Y.spec.col_align &&
        X.spec.row_align == Y.spec.row_align) {
        Dev<T>::axpy2d(Y.lh(), Y.lw(), alpha, X.buf, 1, X.ldim, Y.buf, 1, Y.ldim, Y.stream);
        return;
    }
    // different distribution or alignment: redistribute X like Y first (reference Axpy on
    // ElementalMatrix, level1/Axpy.hpp:120-175, does the same through a temporary copy)
    DistMat<T> X2(*Y.grid, Y.spec.col, Y.spec.row, Y.stream);
    X2.Align(Y.spec.col_align, Y.spec.row_align);
    Copy(X, X2);
    Dev<T>::axpy2d(Y.lh(), Y.lw(), alpha, X2.buf, 1, X2.ldim, Y.buf, 1, Y.ldim, Y.stream);
}

template <typename T>
void Scale(T alpha, DistMat<T>& A) { Dev<T>::scale2d(A.lh(), A.lw(), alpha, A.buf, A.ldim, A.stream); }
template <typename T>
void Zero(DistMat<T>& A) { Dev<T>::fill2d(A.lh(), A.lw(), T(0), A.buf, A.ldim, A.stream); }

// B += alpha * (sum over A's replicas)  (reference AxpyContract, level1/AxpyContract.hpp:483-509)
template <typename T>
void AxpyContract(T alpha, const DistMat<T>& A, DistMat<T>& B) {
    Grid& g = *A.grid;
    cudaStream_t s = B.stream;
    if (A.spec.height != B.spec.height || A.spec.width != B.spec.width)
        throw EngineError(ELB200_ERR_ARG, "AxpyContract: nonconformal dimensions");
    ContractPlan plan = plan_contract(g.shape, g.me, A.spec, B.spec);
    if (!plan.valid) {
        // Same distributions but different alignment: realign every partial-sum layer first
        // (the reference's unaligned branches do this with a SendRecv, AxpyContract.hpp:219-261).
        const bool col_ok = A.spec.col == B.spec.col || A.spec.col == STAR;
        const bool row_ok = A.spec.row == B.spec.row || A.spec.row == STAR;
        if (!col_ok || !row_ok) throw EngineError(ELB200_ERR_ARG, "AxpyContract: incompatible distributions");
        DistMat<T> A2(g, A.spec.col, A.spec.row, s);
        A2.Align(A.spec.col == B.spec.col ? B.spec.col_align : 0, A.spec.row == B.spec.row ? B.spec.row_align : 0);
        Copy(A, A2);
        plan = plan_contract(g.shape, g.me, A2.spec, B.spec);
        if (!plan.valid) throw EngineError(ELB200_ERR_ARG, "AxpyContract: realignment failed");
        AxpyContract(alpha, A2, B);
        return;
    }
    if (A.spec.height == 0 || A.spec.width == 0) return;
    T* Abuf = const_cast<T*>(A.buf);
    if (plan.group == 1) {
        const elb200_segment& a = plan.packs[0].seg;
        Dev<T>::axpy2d(a.rows, a.cols, alpha, A.buf + a.row0 + a.col0 * A.ldim, a.row_step, a.col_step * A.ldim,
                       B.buf, 1, B.ldim, s);
        return;
    }
    if (g.arena && size_t(plan.group) * plan.portion * sizeof(T) <= g.arena->slot_bytes() && plan.group <= ELB200_MAX_SEGMENTS) {
        // ReduceScatter by the receiver over NVLink peer memory (peer.hpp pull mode): every member
        // packs its partial sums into its own arena slot, then ONE kernel per rank loads its portion
        // from all members' slots, sums them in communicator order and applies the axpy-unpack.
        const std::vector<int> members = g.comm_members(plan.comm_dist);
        const int idx = g.comm_index(plan.comm_dist);
        if constexpr (FusedReduce<T>::ok) {
            if (int64_t(plan.group) * plan.portion * int64_t(sizeof(T)) < fused_bytes() && idx == plan.my_index &&
                plan.packs.size() <= size_t(ELB200_MAX_SEGMENTS)) {
                // small: pack, flags, and the sum of my portion over the members' slots + axpy-unpack in ONE kernel
                elb200_segment pk[ELB200_MAX_SEGMENTS];
                int np = 0;
                for (const Message& m : plan.packs) pk[np++] = m.seg;
                const elb200_peer_gather pg = g.arena->FusedBegin(Grid::arena_channel(plan.comm_dist), members, idx, plan.portion,
                                                                  size_t(plan.group) * plan.portion * sizeof(T));
                throw_status(FusedReduce<T>::run(&pg, np, pk, A.buf, A.ldim, &plan.unpack, B.buf, B.ldim, alpha, s), "elb200_peer_reduce_scatter");
                return;
            }
        }
        PeerArena::Ticket t = g.arena->Publish(Grid::arena_channel(plan.comm_dist), members, idx,
                                               size_t(plan.group) * plan.portion * sizeof(T), s);
        run_lattice<T>(ELB200_PACK, plan.packs, Abuf, A.ldim, static_cast<T*>(t.data), T(0), s);
        g.arena->Signal(t, members, idx, s);
        const T* srcs[ELB200_MAX_SEGMENTS];
        for (int i = 0; i < plan.group; ++i) srcs[i] = static_cast<const T*>(g.arena->PeerSlot(members[i], t));
        Dev<T>::peer_reduce_axpy(plan.group, srcs, int64_t(plan.my_index) * plan.portion, &plan.unpack, B.buf, B.ldim, alpha, s);
        g.arena->Release(t, members, idx, s);
        return;
    }
    ncclComm_t comm = g.comm_for(plan.comm_dist);
    if (!comm) throw EngineError(ELB200_ERR_STATE, "AxpyContract: grid has no NCCL communicator");
    Scratch<T> stage(int64_t(plan.group + 1) * plan.portion, s);
    run_lattice<T>(ELB200_PACK, plan.packs, Abuf, A.ldim, stage.p, T(0), s);
    T* recv = stage.p + int64_t(plan.group) * plan.portion;
    g.NcclBegin(s);
    nccl_check(ncclReduceScatter(stage.p, recv, size_t(plan.portion), Dev<T>::nccl, ncclSum, comm, s), "ncclReduceScatter");
    g.NcclEnd(s);
    elb200_segment u = plan.unpack;
    u.buf_offset = int64_t(plan.group) * plan.portion;
    if (u.rows > 0 && u.cols > 0) Dev<T>::lattice(ELB200_UNPACK_AXPY, 1, &u, B.buf, B.ldim, stage.p, alpha, s);
}

// B = sum over A's replicas (reference Contract, level1/Contract.hpp:14-99)
template <typename T>
void Contract(const DistMat<T>& A, DistMat<T>& B) {
    if (A.spec.col == B.spec.col && !B.col_constrained) B.spec.col_align = A.spec.col_align;
    if (A.spec.row == B.spec.row && !B.row_constrained) B.spec.row_align = A.spec.row_align;
    B.Resize(A.spec.height, A.spec.width);
    Zero(B);
    AxpyContract(T(1), A, B);
}

// B += alpha * (sum over A's replicas)^T (reference TransposeAxpyContract.hpp:14-47)
template <typename T>
void TransposeAxpyContract(T alpha, const DistMat<T>& A, DistMat<T>& B) {
    Grid& g = *A.grid;
    cudaStream_t s = B.stream;
    DistMat<T> sum(g, B.spec.row, B.spec.col, s);
    sum.Align(B.spec.row_align, B.spec.col_align);
    if (A.spec.col == sum.spec.col && A.spec.row == sum.spec.row) Copy(A, sum);
    else Contract(A, sum);
    // B(rows x cols) += alpha * sum(cols x rows)^T on the local matrices
    Dev<T>::axpy2d_transpose(B.lh(), B.lw(), alpha, sum.buf, sum.ldim, B.buf, B.ldim, s);
}

#define ELB_INST(T)                                                          \
    template void Copy<T>(const DistMat<T>&, DistMat<T>&);                   \
    template void TransposeCopy<T>(const DistMat<T>&, DistMat<T>&);          \
    template void AxpyContract<T>(T, const DistMat<T>&, DistMat<T>&);        \
    template void Contract<T>(const DistMat<T>&, DistMat<T>&);               \
    template void TransposeAxpyContract<T>(T, const DistMat<T>&, DistMat<T>&); \
    template void Axpy<T>(T, const DistMat<T>&, DistMat<T>&);                \
    template void Scale<T>(T, DistMat<T>&);                                  \
    template void Zero<T>(DistMat<T>&);
ELB_INST(float)
ELB_INST(double)
ELB_INST(Half)
#undef ELB_INST

}  // namespace elb
