// PeerArena implementation: CUDA IPC mapping of every rank's staging arena, copy-engine
// transfers, stream memory operations for the cross-GPU handshake. See peer.hpp.
#include "peer.hpp"

#include <cuda.h>
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "engine.hpp"

namespace elb {
namespace {

enum { ARRIVE = 0, ACK = 1 };
constexpr size_t kFlagWords = size_t(2) * PeerArena::kChannels * PeerArena::kSlots * PeerArena::kMaxRanks;
constexpr uint32_t kSeqMask = 0xFFFFu;

typedef CUresult (*WaitValue32Fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);

WaitValue32Fn wait_value_fn() {
    static WaitValue32Fn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<WaitValue32Fn>(ptr);
    });
    return fn;
}

struct Hello {
    cudaIpcMemHandle_t arena, flags;
    unsigned long long host;  // hash of the hostname: IPC only works inside one node
    int ok;
    int pad;
};

unsigned long long host_hash() {
    char name[256] = {0};
    gethostname(name, sizeof(name) - 1);
    unsigned long long h = 1469598103934665603ull;
    for (const char* p = name; *p; ++p) h = (h ^ (unsigned char)*p) * 1099511628211ull;
    return h;
}

}  // namespace

uint32_t* PeerArena::flag(int rank, int kind, int channel, int slot, int src) const {
    return flags_[rank] + ((size_t(kind) * kChannels + channel) * kSlots + slot) * kMaxRanks + src;
}
char* PeerArena::slot_base(int rank, int channel, int slot) const {
    return arena_[rank] + (size_t(channel) * kSlots + slot) * slot_bytes_;
}

PeerArena* PeerArena::Create(ncclComm_t world, int rank, int size, cudaStream_t s) {
    const char* off = std::getenv("ELB200_P2P");
    if ((off && std::atoi(off) == 0) || size < 2 || size > kMaxRanks || !world) return nullptr;
    const char* mb = std::getenv("ELB200_ARENA_MB");
    const char* kb = std::getenv("ELB200_ARENA_KB");  // finer unit, for tests of the does-not-fit fallbacks
    const size_t slot_bytes = kb && std::atoll(kb) > 0 ? size_t(std::atoll(kb)) << 10
                                                       : size_t(mb && std::atoll(mb) > 0 ? std::atoll(mb) : 1024) << 20;

    PeerArena* pa = new PeerArena();
    pa->rank_ = rank; pa->size_ = size; pa->slot_bytes_ = slot_bytes;
    pa->arena_.assign(size, nullptr);
    pa->flags_.assign(size, nullptr);

    Hello mine;
    std::memset(&mine, 0, sizeof(mine));
    mine.host = host_hash();
    char* arena = nullptr;
    uint32_t* flags = nullptr;
    bool ok = wait_value_fn() != nullptr;
    ok = ok && cudaMalloc(&arena, size_t(kChannels) * kSlots * slot_bytes) == cudaSuccess;
    ok = ok && cudaMalloc(&flags, kFlagWords * sizeof(uint32_t)) == cudaSuccess;
    ok = ok && cudaMalloc(&pa->consts_, (kSeqMask + 1) * sizeof(uint32_t)) == cudaSuccess;
    ok = ok && cudaMalloc(&pa->counters_, size_t(kChannels) * kSlots * 2 * sizeof(unsigned int)) == cudaSuccess;
    ok = ok && cudaMemset(pa->counters_, 0, size_t(kChannels) * kSlots * 2 * sizeof(unsigned int)) == cudaSuccess;
    if (ok) {
        std::vector<uint32_t> iota(kSeqMask + 1);
        for (uint32_t i = 0; i <= kSeqMask; ++i) iota[i] = i;
        ok = cudaMemcpy(pa->consts_, iota.data(), iota.size() * sizeof(uint32_t), cudaMemcpyHostToDevice) == cudaSuccess;
        ok = ok && cudaMemset(flags, 0, kFlagWords * sizeof(uint32_t)) == cudaSuccess;
        ok = ok && cudaIpcGetMemHandle(&mine.arena, arena) == cudaSuccess;
        ok = ok && cudaIpcGetMemHandle(&mine.flags, flags) == cudaSuccess;
        ok = ok && cudaDeviceSynchronize() == cudaSuccess;
    }
    cudaGetLastError();
    mine.ok = ok ? 1 : 0;
    pa->arena_[rank] = arena;
    pa->flags_[rank] = flags;

    // exchange the handles (device-side AllGather of a few hundred bytes over the world comm)
    std::vector<Hello> all(size);
    Hello* dev = nullptr;
    cuda_check(cudaMalloc(&dev, sizeof(Hello) * (size + 1)), "PeerArena hello");
    cuda_check(cudaMemcpyAsync(dev + size, &mine, sizeof(Hello), cudaMemcpyHostToDevice, s), "PeerArena hello");
    nccl_check(ncclAllGather(dev + size, dev, sizeof(Hello), ncclChar, world, s), "PeerArena hello");
    cuda_check(cudaMemcpyAsync(all.data(), dev, sizeof(Hello) * size, cudaMemcpyDeviceToHost, s), "PeerArena hello");
    cuda_check(cudaStreamSynchronize(s), "PeerArena hello");

    int good = 1;
    for (int r = 0; r < size; ++r) good = good && all[r].ok && all[r].host == mine.host;
    if (good) {
        for (int r = 0; r < size && good; ++r) {
            if (r == rank) continue;
            void *pa_ptr = nullptr, *pf_ptr = nullptr;
            if (cudaIpcOpenMemHandle(&pa_ptr, all[r].arena, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess ||
                cudaIpcOpenMemHandle(&pf_ptr, all[r].flags, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess)
                good = 0;
            pa->arena_[r] = static_cast<char*>(pa_ptr);
            pa->flags_[r] = static_cast<uint32_t*>(pf_ptr);
        }
        cudaGetLastError();
    }
    // every rank must take the same decision: min-reduce the local verdicts
    int* verdict = reinterpret_cast<int*>(dev);
    cuda_check(cudaMemcpyAsync(verdict, &good, sizeof(int), cudaMemcpyHostToDevice, s), "PeerArena verdict");
    nccl_check(ncclAllReduce(verdict, verdict, 1, ncclInt, ncclMin, world, s), "PeerArena verdict");
    cuda_check(cudaMemcpyAsync(&good, verdict, sizeof(int), cudaMemcpyDeviceToHost, s), "PeerArena verdict");
    cuda_check(cudaStreamSynchronize(s), "PeerArena verdict");
    cudaFree(dev);
    if (!good) {
        if (rank == 0 && std::getenv("ELB200_VERBOSE"))
            std::fprintf(stderr, "[elb200] peer arena unavailable (IPC mapping failed); redistributions use NCCL\n");
        delete pa;
        return nullptr;
    }
    // auxiliary lanes for concurrent peer transfers (two ranks have one peer: nothing to fan out)
    pa->lanes_ = size < 3 ? 0 : kLanes;
    for (int i = 0; i < pa->lanes_; ++i) {
        cuda_check(cudaStreamCreateWithFlags(&pa->lane_[i], cudaStreamNonBlocking), "PeerArena lane");
        cuda_check(cudaEventCreateWithFlags(&pa->lane_done_[i], cudaEventDisableTiming), "PeerArena lane event");
    }
    if (pa->lanes_) cuda_check(cudaEventCreateWithFlags(&pa->fork_ev_, cudaEventDisableTiming), "PeerArena fork event");
    return pa;
}

cudaStream_t PeerArena::fork(int item, cudaStream_t s) {
    if (!lanes_) return s;
    if (!forked_) {
        cuda_check(cudaEventRecord(fork_ev_, s), "PeerArena fork");
        forked_ = true;
    }
    const int l = item % lanes_;
    if (!lane_used_[l]) {
        cuda_check(cudaStreamWaitEvent(lane_[l], fork_ev_, 0), "PeerArena fork wait");
        lane_used_[l] = true;
    }
    return lane_[l];
}

void PeerArena::join(cudaStream_t s) {
    for (int l = 0; l < lanes_; ++l) {
        if (!lane_used_[l]) continue;
        cuda_check(cudaEventRecord(lane_done_[l], lane_[l]), "PeerArena join");
        cuda_check(cudaStreamWaitEvent(s, lane_done_[l], 0), "PeerArena join wait");
        lane_used_[l] = false;
    }
    forked_ = false;
}

void PeerArena::Pull(const Ticket& t, const std::vector<PullOp>& ops, cudaStream_t s) {
    for (const PullOp& op : ops)
        if (op.bytes && op.src_offset + op.bytes > slot_bytes_)
            throw EngineError(ELB200_ERR_ARG, "PeerArena::Pull: read leaves the slot");
    int item = 0;
    for (const PullOp& op : ops) {
        if (!op.bytes) continue;
        cudaStream_t ls = fork(item++, s);  // one lane per source: the peers are read concurrently
        cuda_check(cudaMemcpyAsync(op.dst, slot_base(op.rank, t.channel, t.slot) + op.src_offset, op.bytes,
                                   cudaMemcpyDeviceToDevice, ls), "PeerArena pull");
    }
    join(s);
}

void PeerArena::Detach() {
    for (int r = 0; r < size_ && r < int(arena_.size()); ++r) {
        if (r == rank_) continue;
        if (arena_[r]) cudaIpcCloseMemHandle(arena_[r]);
        if (flags_[r]) cudaIpcCloseMemHandle(flags_[r]);
        arena_[r] = nullptr;
        flags_[r] = nullptr;
    }
    cudaGetLastError();
}

PeerArena::~PeerArena() {
    Detach();
    for (int i = 0; i < kLanes; ++i) {
        if (lane_[i]) cudaStreamDestroy(lane_[i]);
        if (lane_done_[i]) cudaEventDestroy(lane_done_[i]);
    }
    if (fork_ev_) cudaEventDestroy(fork_ev_);
    if (!arena_.empty() && arena_[rank_]) cudaFree(arena_[rank_]);
    if (!flags_.empty() && flags_[rank_]) cudaFree(flags_[rank_]);
    if (consts_) cudaFree(consts_);
    if (counters_) cudaFree(counters_);
    cudaGetLastError();
}

void PeerArena::write_flag(uint32_t* dst, uint32_t seq, cudaStream_t s) {
    // a 4-byte copy-engine write on the same stream as the payload copies: it starts after they
    // completed, so a peer that sees the flag sees the data
    cuda_check(cudaMemcpyAsync(dst, consts_ + (seq & kSeqMask), sizeof(uint32_t), cudaMemcpyDeviceToDevice, s),
               "PeerArena flag write");
}

void PeerArena::wait_flag(uint32_t* local, uint32_t seq, cudaStream_t s) {
    // Equality is exact here: a flag advances to seq+1 only after this rank acknowledged seq.
    const CUresult r = wait_value_fn()(reinterpret_cast<CUstream>(s), reinterpret_cast<CUdeviceptr>(local),
                                       seq & kSeqMask, CU_STREAM_WAIT_VALUE_EQ);
    if (r != CUDA_SUCCESS) throw EngineError(ELB200_ERR_CUDA, "cuStreamWaitValue32 failed");
}

PeerArena::Ticket PeerArena::AllGatherBegin(int channel, const std::vector<int>& members, int my_index, const void* send,
                                            size_t send_bytes, size_t portion_bytes, bool skip_self, cudaStream_t s) {
    const int g = int(members.size());
    if (portion_bytes * size_t(g) > slot_bytes_ || send_bytes > portion_bytes)
        throw EngineError(ELB200_ERR_ARG, "PeerArena::AllGather: message larger than an arena slot");
    const uint64_t n = uses_[channel]++;
    Ticket t;
    t.channel = channel;
    t.slot = int(n & 1);
    t.seq = uint32_t(n / 2 + 1);
    // 1. my previous block in each member's slot must have been consumed
    for (int i = 0; i < g; ++i)
        if (i != my_index) wait_flag(flag(rank_, ACK, channel, t.slot, members[i]), t.seq - 1, s);
    // 2. my block into every member's slot (mine included unless skip_self), 3. tell them: payload and flag
    // for one member stay on one lane (ordered), different members go out concurrently
    for (int i = 0; i < g; ++i) {
        if (i == my_index && skip_self) continue;
        cudaStream_t ls = fork(i, s);
        if (send_bytes)
            cuda_check(cudaMemcpyAsync(slot_base(members[i], channel, t.slot) + size_t(my_index) * portion_bytes, send,
                                       send_bytes, cudaMemcpyDeviceToDevice, ls), "PeerArena copy");
        if (i != my_index) write_flag(flag(members[i], ARRIVE, channel, t.slot, rank_), t.seq, ls);
    }
    join(s);
    t.data = slot_base(rank_, channel, t.slot);
    return t;
}

PeerArena::Ticket PeerArena::AllGatherRotate(int channel, const std::vector<int>& members, int my_index, const void* send,
                                             size_t send_bytes, size_t portion_bytes, cudaStream_t s) {
    const int g = int(members.size());
    if (portion_bytes * size_t(g) > slot_bytes_ || send_bytes > portion_bytes)
        throw EngineError(ELB200_ERR_ARG, "PeerArena::AllGather: message larger than an arena slot");
    const uint64_t n = uses_[channel]++;
    Ticket t;
    t.channel = channel;
    t.slot = int(n & 1);
    t.seq = uint32_t(n / 2 + 1);
    for (int i = 0; i < g; ++i)
        if (i != my_index) wait_flag(flag(rank_, ACK, channel, t.slot, members[i]), t.seq - 1, s);
    cudaStream_t ls = fork(0, s);
    for (int d = 1; d < g; ++d) {
        const int i = (my_index + d) % g;
        if (send_bytes)
            cuda_check(cudaMemcpyAsync(slot_base(members[i], channel, t.slot) + size_t(my_index) * portion_bytes, send,
                                       send_bytes, cudaMemcpyDeviceToDevice, ls), "PeerArena copy");
        write_flag(flag(members[i], ARRIVE, channel, t.slot, rank_), t.seq, ls);
    }
    t.data = slot_base(rank_, channel, t.slot);
    return t;
}

void PeerArena::WaitMember(const Ticket& t, const std::vector<int>& members, int my_index, int i, cudaStream_t s) {
    if (i != my_index) wait_flag(flag(rank_, ARRIVE, t.channel, t.slot, members[i]), t.seq, s);
}

PeerArena::Ticket PeerArena::AllGather(int channel, const std::vector<int>& members, int my_index, const void* send,
                                       size_t send_bytes, size_t portion_bytes, cudaStream_t s) {
    Ticket t = AllGatherBegin(channel, members, my_index, send, send_bytes, portion_bytes, false, s);
    // 4. wait for theirs
    for (int i = 0; i < int(members.size()); ++i) WaitMember(t, members, my_index, i, s);
    return t;
}

PeerArena::Ticket PeerArena::GatherDirect(int channel, const std::vector<int>& members, int my_index,
                                          const std::vector<Copy2D>& copies, cudaStream_t s) {
    const int g = int(members.size());
    // argument checks come BEFORE the sequence counter moves or a wait is enqueued: a rank that throws here has
    // not touched the collective state, so its peers are not left waiting on flags that never come
    for (const Copy2D& c : copies)
        if (c.width && c.height && c.dst_offset + (c.height - 1) * c.dpitch + c.width > slot_bytes_)
            throw EngineError(ELB200_ERR_ARG, "PeerArena::GatherDirect: transfer leaves the slot");
    const uint64_t n = uses_[channel]++;
    Ticket t;
    t.channel = channel;
    t.slot = int(n & 1);
    t.seq = uint32_t(n / 2 + 1);
    for (int i = 0; i < g; ++i)
        if (i != my_index) wait_flag(flag(rank_, ACK, channel, t.slot, members[i]), t.seq - 1, s);
    for (int i = 0; i < g; ++i) {
        cudaStream_t ls = fork(i, s);  // one lane per destination; its arrival flag follows its payload
        for (const Copy2D& c : copies) {
            if (c.width == 0 || c.height == 0) continue;
            cuda_check(cudaMemcpy2DAsync(slot_base(members[i], channel, t.slot) + c.dst_offset, c.dpitch, c.src, c.spitch,
                                         c.width, c.height, cudaMemcpyDeviceToDevice, ls), "PeerArena direct copy");
        }
        if (i != my_index) write_flag(flag(members[i], ARRIVE, channel, t.slot, rank_), t.seq, ls);
    }
    join(s);
    for (int i = 0; i < g; ++i)
        if (i != my_index) wait_flag(flag(rank_, ARRIVE, channel, t.slot, members[i]), t.seq, s);
    t.data = slot_base(rank_, channel, t.slot);
    return t;
}

PeerArena::Ticket PeerArena::Publish(int channel, const std::vector<int>& members, int my_index, size_t bytes,
                                     cudaStream_t s) {
    if (bytes > slot_bytes_) throw EngineError(ELB200_ERR_ARG, "PeerArena::Publish: buffer larger than an arena slot");
    const int g = int(members.size());
    const uint64_t n = uses_[channel]++;
    Ticket t;
    t.channel = channel;
    t.slot = int(n & 1);
    t.seq = uint32_t(n / 2 + 1);
    // every member must be done with the previous contents of this slot generation (an ack means
    // "I have finished reading all members' slots", so all acks == nobody reads mine any more)
    for (int i = 0; i < g; ++i)
        if (i != my_index) wait_flag(flag(rank_, ACK, channel, t.slot, members[i]), t.seq - 1, s);
    t.data = slot_base(rank_, channel, t.slot);
    return t;
}

void PeerArena::Signal(const Ticket& t, const std::vector<int>& members, int my_index, cudaStream_t s) {
    const int g = int(members.size());
    for (int i = 0; i < g; ++i)
        if (i != my_index) write_flag(flag(members[i], ARRIVE, t.channel, t.slot, rank_), t.seq, s);
    for (int i = 0; i < g; ++i)
        if (i != my_index) wait_flag(flag(rank_, ARRIVE, t.channel, t.slot, members[i]), t.seq, s);
}

elb200_peer_gather PeerArena::FusedBegin(int channel, const std::vector<int>& members, int my_index, int64_t portion_elems, size_t bytes) {
    const int g = int(members.size());
    if (bytes > slot_bytes_ || g > ELB200_MAX_SEGMENTS || g < 2)
        throw EngineError(ELB200_ERR_ARG, "PeerArena::FusedBegin: message larger than an arena slot");
    const uint64_t n = uses_[channel]++;
    const int slot = int(n & 1);
    const uint32_t seq = uint32_t(n / 2 + 1);
    elb200_peer_gather pg;
    std::memset(&pg, 0, sizeof(pg));
    pg.nmembers = g; pg.my_index = my_index; pg.portion = portion_elems;
    pg.seq = seq & kSeqMask; pg.seq_prev = (seq - 1) & kSeqMask;
    pg.counters = counters_ + (size_t(channel) * kSlots + slot) * 2;
    for (int i = 0; i < g; ++i) {
        pg.slots[i] = slot_base(members[i], channel, slot);
        pg.arrive_out[i] = flag(members[i], ARRIVE, channel, slot, rank_);
        pg.arrive_in[i] = flag(rank_, ARRIVE, channel, slot, members[i]);  // i == my_index: my own flag word, nobody else's
        pg.ack_in[i] = flag(rank_, ACK, channel, slot, members[i]);
        pg.ack_out[i] = flag(members[i], ACK, channel, slot, rank_);
    }
    return pg;
}

void PeerArena::FusedAllGather(int channel, const std::vector<int>& members, int my_index, int elem_bytes, const void* send,
                               int64_t send_elems, int64_t portion_elems, int nseg, const elb200_segment* recvs, void* M, int64_t ldm,
                               cudaStream_t s) {
    if (send_elems > portion_elems) throw EngineError(ELB200_ERR_ARG, "PeerArena::FusedAllGather: block larger than its portion");
    const elb200_peer_gather pg = FusedBegin(channel, members, my_index, portion_elems, size_t(portion_elems) * size_t(elem_bytes) * members.size());
    throw_status(elb200_peer_allgather(&pg, elem_bytes, send, send_elems, nseg, recvs, M, ldm, s), "elb200_peer_allgather");
}

void PeerArena::Release(const Ticket& t, const std::vector<int>& members, int my_index, cudaStream_t s) {
    const int g = int(members.size());
    for (int i = 0; i < g; ++i)
        if (i != my_index) write_flag(flag(members[i], ACK, t.channel, t.slot, rank_), t.seq, s);
}

}  // namespace elb
