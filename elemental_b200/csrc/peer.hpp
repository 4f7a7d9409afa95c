// PeerArena: AllGather over NVLink peer memory with ZERO SMs.
//
// The reference gathers SUMMA panels with an NCCL (Aluminum) AllGather kernel
// (src/core/imports/mpi/AllGather.hpp:8-21 <- Copy/RowAllGather.hpp:58-75, ColAllGather.hpp).
// An SM-resident collective cannot overlap a persistent GEMM that owns every SM, so here each
// rank owns a symmetric staging arena in HBM, maps every peer's arena through CUDA IPC, and a
// gather is
//     copy engine: my block -> slot[my index] in EVERY member's arena   (cudaMemcpyAsync, NVLink)
//     copy engine: arrival flag -> every member                         (4-byte write, same stream)
//     stream memop: wait until every member's flag shows this sequence  (cuStreamWaitValue32)
// and, once the consumer (one unpack kernel) has run, an ack flag the other way so a slot is never
// overwritten before it was read.  Nothing spins on an SM and no kernel is launched for the
// transfer; the waits are stream memory operations, so only the calling stream blocks.
#pragma once
#include <cuda_runtime.h>
#include <nccl.h>

#include <cstdint>
#include <vector>

#include "../../include/elb200.h"

namespace elb {

class PeerArena {
public:
    enum { kChannels = 3, kSlots = 2, kMaxRanks = 64 };  // channel: 0 = MC comm, 1 = MR comm, 2 = VC (world)
    struct Ticket {
        void* data = nullptr;  // local gathered buffer: member i's block at i * portion_bytes
        int channel = 0, slot = 0;
        uint32_t seq = 0;
    };

    // Collective over all `size` ranks of `world` (one process per GPU, same node). Returns nullptr
    // — on every rank alike — when peer mapping is not possible; callers then use NCCL.
    static PeerArena* Create(ncclComm_t world, int rank, int size, cudaStream_t s);
    ~PeerArena();
    // Unmap every peer's arena (idempotent). The owner may free its memory only after all
    // importers have done this, so Grid teardown runs Detach -> barrier -> delete.
    void Detach();

    size_t slot_bytes() const { return slot_bytes_; }
    // Collective over `members` (VC ranks in communicator order; members[my_index] == this rank).
    // All members pass the same channel and portion_bytes; send_bytes <= portion_bytes may differ.
    Ticket AllGather(int channel, const std::vector<int>& members, int my_index, const void* send,
                     size_t send_bytes, size_t portion_bytes, cudaStream_t s);
    // The same gather in two halves, so that the consumer can start on the blocks that have landed while the
    // others are still in flight: Begin sends my block to every member (skip_self: not into my own slot — the
    // caller reads its own block from the source) and returns without waiting; WaitMember blocks `s` until
    // member i's block has arrived in my slot. AllGather == Begin(skip_self = false) + WaitMember for all.
    Ticket AllGatherBegin(int channel, const std::vector<int>& members, int my_index, const void* send, size_t send_bytes,
                          size_t portion_bytes, bool skip_self, cudaStream_t s);
    void WaitMember(const Ticket& t, const std::vector<int>& members, int my_index, int i, cudaStream_t s);
    // Rotation form for large blocks: my block goes to members my_index + 1, my_index + 2, ... one after the other on
    // ONE copy lane, each at the full link rate (every member's ingress carries exactly one flow at a time), so the
    // blocks land STAGGERED — member my_index - d's block arrives d-th — and the consumer unpacks block d while
    // block d + 1 is on the wire. `s` is NOT joined: work the caller enqueues on `s` next (its own block) runs
    // underneath the sends. AllGatherEnd(s) makes `s` wait for my sends; call it before the send buffer is reused.
    Ticket AllGatherRotate(int channel, const std::vector<int>& members, int my_index, const void* send, size_t send_bytes,
                           size_t portion_bytes, cudaStream_t s);
    void AllGatherEnd(cudaStream_t s) { join(s); }
    // Small messages: a whole collective in ONE kernel (peer_kernels.cu: elb200_peer_allgather / _exchange / _reduce_scatter) —
    // peer stores or loads, flags and the pack / unpack lattices instead of 20-40 stream operations. FusedBegin takes the
    // channel's next slot and sequence number, exactly like AllGather / Publish, and returns the descriptor those kernels
    // want; the kernel itself waits for the previous acks and acknowledges at its end (no Release). `bytes` = what the
    // operation keeps in one slot (checked against the slot size).
    elb200_peer_gather FusedBegin(int channel, const std::vector<int>& members, int my_index, int64_t portion_elems, size_t bytes);
    // the gather through it (recvs: UNPACK segments relative to the slot base)
    void FusedAllGather(int channel, const std::vector<int>& members, int my_index, int elem_bytes, const void* send, int64_t send_elems,
                        int64_t portion_elems, int nseg, const elb200_segment* recvs, void* M, int64_t ldm, cudaStream_t s);
    // Call after the last read of ticket.data has been enqueued on `s`.
    void Release(const Ticket& t, const std::vector<int>& members, int my_index, cudaStream_t s);

    // Direct mode (SUMMA panels): every member's contribution goes with pitched 2-D copy-engine
    // transfers straight to its final place inside each member's slot, which IS the panel the GEMM
    // reads — no staging, no pack, no unpack kernel. `copies` is the same for every destination
    // (all members lay the panel out identically). Release only after the panel has been consumed.
    struct Copy2D {
        size_t dst_offset;  // bytes from the slot base
        size_t dpitch;      // bytes
        const void* src;
        size_t spitch;      // bytes
        size_t width;       // bytes per row of the transfer
        size_t height;      // rows of the transfer
    };
    Ticket GatherDirect(int channel, const std::vector<int>& members, int my_index, const std::vector<Copy2D>& copies,
                        cudaStream_t s);

    // Pull mode (reductions): every member fills ITS OWN slot (t.data, `bytes` long) after Publish,
    // then Signal; afterwards each member reads all members' slots through the peer mapping
    // (PeerSlot) from a kernel — NVLink loads — and finally calls Release.
    Ticket Publish(int channel, const std::vector<int>& members, int my_index, size_t bytes, cudaStream_t s);
    void Signal(const Ticket& t, const std::vector<int>& members, int my_index, cudaStream_t s);
    const void* PeerSlot(int rank, const Ticket& t) const { return slot_base(rank, t.channel, t.slot); }
    // Fetch pieces of the members' published slots into local memory with the copy engines, the
    // transfers from different peers running concurrently (general exchange of Copy). Call between
    // Signal and Release.
    struct PullOp {
        int rank;           // VC rank whose slot is read
        size_t src_offset;  // bytes from its slot base
        void* dst;          // local device memory
        size_t bytes;
    };
    void Pull(const Ticket& t, const std::vector<PullOp>& ops, cudaStream_t s);

private:
    PeerArena() = default;
    uint32_t* flag(int rank, int kind, int channel, int slot, int src) const;
    char* slot_base(int rank, int channel, int slot) const;
    void write_flag(uint32_t* dst, uint32_t seq, cudaStream_t s);
    void wait_flag(uint32_t* local, uint32_t seq, cudaStream_t s);
    // Transfers to / from different peers are independent: they are spread over `lanes_` auxiliary
    // streams forked from the caller's stream and joined back into it, so several copy engines
    // and NVLink paths work at once instead of one peer after the other.
    cudaStream_t fork(int item, cudaStream_t s);  // stream for item i of the current fan-out
    void join(cudaStream_t s);                    // caller's stream waits for every lane used

    int rank_ = 0, size_ = 1;
    size_t slot_bytes_ = 0;
    std::vector<char*> arena_;       // per rank: base of its arena as mapped in THIS process
    std::vector<uint32_t*> flags_;   // per rank: base of its flag block
    uint32_t* consts_ = nullptr;     // consts_[i] == i, source of the 4-byte flag writes
    unsigned int* counters_ = nullptr;  // 2 words per (channel, slot) for the fused kernels, zero between launches
    uint64_t uses_[kChannels] = {0, 0, 0};
    enum { kLanes = 4 };
    int lanes_ = 0;  // 0: no fan-out, everything on the caller's stream
    cudaStream_t lane_[kLanes] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t lane_done_[kLanes] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t fork_ev_ = nullptr;
    bool lane_used_[kLanes] = {false, false, false, false};
    bool forked_ = false;
};

}  // namespace elb
