// hostsim: test double of the NCCL calls the elb engine makes, for ranks that are THREADS of one
// process (see sim.hpp). Collectives are stream-ordered like NCCL's: each call enqueues a closure
// on the caller's stream which meets the other ranks' closures at a barrier, reads their buffers
// and writes its own result. Point-to-point messages go through per-(src,dst) mailboxes.
#include <nccl.h>

#include <algorithm>

#include <chrono>
#include <condition_variable>
#include <cstring>
#include <deque>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "sim.hpp"

namespace {

struct Msg {
    const void* src;
    size_t bytes;
    bool done = false;
    double vt = 0;  // sender's virtual time when it posted the message
};

struct Shared {
    int n = 0;
    std::mutex m;
    std::condition_variable cv;
    int arrived = 0;
    uint64_t gen = 0;
    // what every rank published for the collective in flight
    std::vector<const void*> send;
    std::vector<int> color, key;
    std::map<int, std::shared_ptr<Shared>> split;
    int joined = 0;
    // point-to-point
    std::map<std::pair<int, int>, std::deque<std::shared_ptr<Msg>>> box;

    explicit Shared(int n_) : n(n_), send(n_, nullptr), color(n_, 0), key(n_, 0) {}

    // virtual time: all participants leave a collective at max(arrival times) + its duration
    std::vector<double> vt;
    double vt_sync(int rank, double seconds, const char* who) {
        if (vt.size() != size_t(n)) vt.assign(size_t(n), 0.0);
        vt[size_t(rank)] = sim::vt_now();
        barrier(who);
        double t = 0;
        for (double x : vt) t = x > t ? x : t;
        barrier(who);
        return t + seconds;
    }
    void barrier(const char* who) {
        std::unique_lock<std::mutex> l(m);
        const uint64_t g = gen;
        if (++arrived == n) {
            arrived = 0;
            ++gen;
            cv.notify_all();
            return;
        }
        if (!cv.wait_for(l, std::chrono::duration<double>(sim::wait_timeout_seconds()), [&] { return gen != g; }))
            sim::fail("%s: only %d of %d ranks reached the collective (deadlock / mismatched call order?)", who, arrived, n);
    }
};

size_t dtype_size(ncclDataType_t t) {
    switch (t) {
        case ncclInt8: case ncclUint8: return 1;
        case ncclFloat16: case ncclBfloat16: return 2;
        case ncclInt32: case ncclUint32: case ncclFloat32: return 4;
        case ncclInt64: case ncclUint64: case ncclFloat64: return 8;
        default: sim::fail("hostsim NCCL: unsupported datatype %d", int(t));
    }
}

template <typename T>
void reduce_into(T* acc, const T* x, size_t n, ncclRedOp_t op) {
    for (size_t i = 0; i < n; ++i) {
        if (op == ncclSum) acc[i] += x[i];
        else if (op == ncclMin) acc[i] = x[i] < acc[i] ? x[i] : acc[i];
        else if (op == ncclMax) acc[i] = x[i] > acc[i] ? x[i] : acc[i];
        else sim::fail("hostsim NCCL: unsupported reduction %d", int(op));
    }
}
void reduce_bytes(void* acc, const void* x, size_t n, ncclDataType_t t, ncclRedOp_t op) {
    switch (t) {
        case ncclInt32: reduce_into(static_cast<int32_t*>(acc), static_cast<const int32_t*>(x), n, op); break;
        case ncclFloat32: reduce_into(static_cast<float*>(acc), static_cast<const float*>(x), n, op); break;
        case ncclFloat64: reduce_into(static_cast<double*>(acc), static_cast<const double*>(x), n, op); break;
        case ncclFloat16: reduce_into(static_cast<_Float16*>(acc), static_cast<const _Float16*>(x), n, op); break;
        default: sim::fail("hostsim NCCL: reduction on datatype %d not supported", int(t));
    }
}

std::mutex g_registry_mutex;
std::map<std::string, std::shared_ptr<Shared>> g_registry;
uint64_t g_next_id = 1;

struct P2P {
    bool is_send;
    const void* sbuf;
    void* rbuf;
    size_t bytes;
    int peer;
    struct ncclComm* comm;
    cudaStream_t stream;
};
thread_local int t_group_depth = 0;
thread_local std::vector<P2P> t_group_ops;

}  // namespace

struct ncclComm {
    std::shared_ptr<Shared> sh;
    int rank = 0;
};

namespace {

void run_p2p(const std::vector<P2P>& ops) {
    std::vector<std::shared_ptr<Msg>> mine;
    for (const P2P& o : ops) {
        if (!o.is_send) continue;
        sim::check_device_range(o.sbuf, o.bytes, "ncclSend");
        auto msg = std::make_shared<Msg>();
        msg->src = o.sbuf;
        msg->bytes = o.bytes;
        msg->vt = sim::vt_now();
        Shared& sh = *o.comm->sh;
        {
            std::lock_guard<std::mutex> l(sh.m);
            sh.box[{o.comm->rank, o.peer}].push_back(msg);
        }
        sh.cv.notify_all();
        mine.push_back(msg);
    }
    for (const P2P& o : ops) {
        if (o.is_send) continue;
        sim::check_device_range(o.rbuf, o.bytes, "ncclRecv");
        Shared& sh = *o.comm->sh;
        std::shared_ptr<Msg> msg;
        {
            std::unique_lock<std::mutex> l(sh.m);
            auto& q = sh.box[{o.peer, o.comm->rank}];
            if (!sh.cv.wait_for(l, std::chrono::duration<double>(sim::wait_timeout_seconds()), [&] { return !q.empty(); }))
                sim::fail("ncclRecv: rank %d never got a message from rank %d (deadlock?)", o.comm->rank, o.peer);
            msg = q.front();
            q.pop_front();
        }
        if (msg->bytes != o.bytes)
            sim::fail("ncclRecv: rank %d expects %zu bytes from rank %d, sender posted %zu", o.comm->rank, o.bytes, o.peer, msg->bytes);
        if (sim::vtime_on()) {  // a message starts when both ends have posted it
            sim::vt_at_least(msg->vt);
            sim::vt_at_least(sim::vt_now() + 10e-6 + double(o.bytes) / 300e9);
        }
        if (!(sim::dry_run() && o.bytes > sim::kDryKeep)) std::memcpy(o.rbuf, msg->src, o.bytes);
        {
            std::lock_guard<std::mutex> l(sh.m);
            msg->done = true;
        }
        sh.cv.notify_all();
    }
    size_t i = 0;
    for (const P2P& o : ops) {
        if (!o.is_send) continue;
        Shared& sh = *o.comm->sh;
        std::unique_lock<std::mutex> l(sh.m);
        auto& msg = mine[i++];
        if (!sh.cv.wait_for(l, std::chrono::duration<double>(sim::wait_timeout_seconds()), [&] { return msg->done; }))
            sim::fail("ncclSend: message from rank %d to rank %d was never received (deadlock?)", o.comm->rank, o.peer);
    }
}

ncclResult_t submit(const P2P& op) {
    if (t_group_depth > 0) {
        t_group_ops.push_back(op);
        return ncclSuccess;
    }
    std::vector<P2P> one{op};
    sim::enqueue(op.stream, [one] { run_p2p(one); });
    return ncclSuccess;
}

}  // namespace

extern "C" {

const char* ncclGetErrorString(ncclResult_t r) { return r == ncclSuccess ? "no error" : "hostsim NCCL error"; }

ncclResult_t ncclGetUniqueId(ncclUniqueId* id) {
    std::memset(id, 0, sizeof(*id));
    std::lock_guard<std::mutex> l(g_registry_mutex);
    const uint64_t v = g_next_id++;
    std::memcpy(id->internal, "hostsim", 7);
    std::memcpy(id->internal + 8, &v, sizeof(v));
    return ncclSuccess;
}

ncclResult_t ncclCommInitRank(ncclComm_t* comm, int nranks, ncclUniqueId id, int rank) {
    if (nranks < 1 || rank < 0 || rank >= nranks) return ncclInvalidArgument;
    std::shared_ptr<Shared> sh;
    {
        std::lock_guard<std::mutex> l(g_registry_mutex);
        auto& slot = g_registry[std::string(id.internal, sizeof(id.internal))];
        if (!slot) slot = std::make_shared<Shared>(nranks);
        sh = slot;
    }
    if (sh->n != nranks) return ncclInvalidArgument;
    ncclComm* c = new ncclComm();
    c->sh = sh;
    c->rank = rank;
    sh->barrier("ncclCommInitRank");
    *comm = c;
    return ncclSuccess;
}

ncclResult_t ncclCommSplit(ncclComm_t comm, int color, int key, ncclComm_t* out, ncclConfig_t*) {
    Shared& sh = *comm->sh;
    {
        std::lock_guard<std::mutex> l(sh.m);
        sh.color[comm->rank] = color;
        sh.key[comm->rank] = key;
    }
    sh.barrier("ncclCommSplit");
    // members of my colour ordered by (key, parent rank)
    std::vector<std::pair<int, int>> members;
    for (int q = 0; q < sh.n; ++q)
        if (sh.color[q] == color) members.push_back({sh.key[q], q});
    std::sort(members.begin(), members.end());
    int my = -1;
    for (size_t i = 0; i < members.size(); ++i)
        if (members[i].second == comm->rank) my = int(i);
    {
        std::lock_guard<std::mutex> l(sh.m);
        auto& slot = sh.split[color];
        if (!slot) slot = std::make_shared<Shared>(int(members.size()));
    }
    sh.barrier("ncclCommSplit");
    ncclComm* c = nullptr;
    if (color != NCCL_SPLIT_NOCOLOR) {  // NOCOLOR: this rank takes part in the call but joins no new communicator
        c = new ncclComm();
        std::lock_guard<std::mutex> l(sh.m);
        c->sh = sh.split[color];
        c->rank = my;
    }
    sh.barrier("ncclCommSplit");
    if (comm->rank == 0) {
        std::lock_guard<std::mutex> l(sh.m);
        sh.split.clear();
    }
    sh.barrier("ncclCommSplit");
    *out = c;
    return ncclSuccess;
}

ncclResult_t ncclCommDestroy(ncclComm_t comm) {
    delete comm;
    return ncclSuccess;
}

ncclResult_t ncclCommUserRank(const ncclComm_t comm, int* rank) { *rank = comm->rank; return ncclSuccess; }
ncclResult_t ncclCommCount(const ncclComm_t comm, int* count) { *count = comm->sh->n; return ncclSuccess; }

ncclResult_t ncclBroadcast(const void* sendbuf, void* recvbuf, size_t count, ncclDataType_t dt, int root, ncclComm_t comm,
                           cudaStream_t s) {
    if (root < 0 || root >= comm->sh->n) return ncclInvalidArgument;
    const size_t bytes = count * dtype_size(dt);
    sim::enqueue(s, [=] {
        Shared& sh = *comm->sh;
        if (bytes) {
            if (comm->rank == root) sim::check_device_range(sendbuf, bytes, "ncclBroadcast send");
            sim::check_device_range(recvbuf, bytes, "ncclBroadcast recv");
        }
        sh.send[comm->rank] = sendbuf;
        if (sim::vtime_on()) sim::vt_at_least(sh.vt_sync(comm->rank, 20e-6 + double(bytes) / 300e9, "ncclBroadcast"));
        sh.barrier("ncclBroadcast");
        const bool skip = sim::dry_run() && bytes > sim::kDryKeep;
        std::vector<char> tmp(skip ? 0 : bytes);
        if (!skip && bytes) std::memcpy(tmp.data(), sh.send[root], bytes);
        sh.barrier("ncclBroadcast");
        if (!skip && bytes) std::memcpy(recvbuf, tmp.data(), bytes);
        sh.barrier("ncclBroadcast");
    });
    return ncclSuccess;
}

ncclResult_t ncclAllGather(const void* sendbuf, void* recvbuf, size_t count, ncclDataType_t dt, ncclComm_t comm,
                           cudaStream_t s) {
    const size_t bytes = count * dtype_size(dt);
    sim::enqueue(s, [=] {
        Shared& sh = *comm->sh;
        if (bytes) {
            sim::check_device_range(sendbuf, bytes, "ncclAllGather send");
            sim::check_device_range(recvbuf, bytes * sh.n, "ncclAllGather recv");
        }
        sh.send[comm->rank] = sendbuf;
        if (sim::vtime_on()) sim::vt_at_least(sh.vt_sync(comm->rank, 20e-6 + double(bytes) * double(sh.n - 1) / 300e9, "ncclAllGather"));
        sh.barrier("ncclAllGather");
        const bool skip = sim::dry_run() && bytes > sim::kDryKeep;
        std::vector<char> tmp(skip ? 0 : bytes * sh.n);
        if (!skip) for (int q = 0; q < sh.n; ++q) std::memcpy(tmp.data() + size_t(q) * bytes, sh.send[q], bytes);
        sh.barrier("ncclAllGather");
        if (!skip) std::memcpy(recvbuf, tmp.data(), tmp.size());
        sh.barrier("ncclAllGather");
    });
    return ncclSuccess;
}

ncclResult_t ncclReduceScatter(const void* sendbuf, void* recvbuf, size_t recvcount, ncclDataType_t dt, ncclRedOp_t op,
                               ncclComm_t comm, cudaStream_t s) {
    const size_t bytes = recvcount * dtype_size(dt);
    sim::enqueue(s, [=] {
        Shared& sh = *comm->sh;
        if (bytes) {
            sim::check_device_range(sendbuf, bytes * sh.n, "ncclReduceScatter send");
            sim::check_device_range(recvbuf, bytes, "ncclReduceScatter recv");
        }
        sh.send[comm->rank] = sendbuf;
        if (sim::vtime_on()) sim::vt_at_least(sh.vt_sync(comm->rank, 20e-6 + double(bytes) * double(sh.n - 1) / 300e9, "ncclReduceScatter"));
        sh.barrier("ncclReduceScatter");
        const bool skip = sim::dry_run() && bytes > sim::kDryKeep;
        std::vector<char> acc(skip ? 0 : bytes);
        for (int q = 0; q < sh.n && !skip; ++q) {
            const char* src = static_cast<const char*>(sh.send[q]) + size_t(comm->rank) * bytes;
            if (q == 0) std::memcpy(acc.data(), src, bytes);
            else reduce_bytes(acc.data(), src, recvcount, dt, op);
        }
        sh.barrier("ncclReduceScatter");
        if (!skip) std::memcpy(recvbuf, acc.data(), bytes);
        sh.barrier("ncclReduceScatter");
    });
    return ncclSuccess;
}

ncclResult_t ncclAllReduce(const void* sendbuf, void* recvbuf, size_t count, ncclDataType_t dt, ncclRedOp_t op,
                           ncclComm_t comm, cudaStream_t s) {
    const size_t bytes = count * dtype_size(dt);
    sim::enqueue(s, [=] {
        Shared& sh = *comm->sh;
        if (bytes) {
            sim::check_device_range(sendbuf, bytes, "ncclAllReduce send");
            sim::check_device_range(recvbuf, bytes, "ncclAllReduce recv");
        }
        sh.send[comm->rank] = sendbuf;
        if (sim::vtime_on()) sim::vt_at_least(sh.vt_sync(comm->rank, 20e-6 + 2.0 * double(bytes) / 300e9, "ncclAllReduce"));
        sh.barrier("ncclAllReduce");
        const bool skip = sim::dry_run() && bytes > sim::kDryKeep;
        std::vector<char> acc(skip ? 0 : bytes);
        for (int q = 0; q < sh.n && !skip; ++q) {
            if (q == 0) std::memcpy(acc.data(), sh.send[q], bytes);
            else reduce_bytes(acc.data(), sh.send[q], count, dt, op);
        }
        sh.barrier("ncclAllReduce");
        if (!skip) std::memcpy(recvbuf, acc.data(), bytes);
        sh.barrier("ncclAllReduce");
    });
    return ncclSuccess;
}

ncclResult_t ncclGroupStart(void) {
    ++t_group_depth;
    return ncclSuccess;
}

ncclResult_t ncclGroupEnd(void) {
    if (t_group_depth <= 0) return ncclInvalidUsage;
    if (--t_group_depth > 0) return ncclSuccess;
    std::vector<P2P> ops;
    ops.swap(t_group_ops);
    if (ops.empty()) return ncclSuccess;
    for (const P2P& o : ops)
        if (o.stream != ops[0].stream) sim::fail("hostsim NCCL: one group spanning several streams is not modelled");
    sim::enqueue(ops[0].stream, [ops] { run_p2p(ops); });
    return ncclSuccess;
}

ncclResult_t ncclSend(const void* buf, size_t count, ncclDataType_t dt, int peer, ncclComm_t comm, cudaStream_t s) {
    if (peer < 0 || peer >= comm->sh->n) return ncclInvalidArgument;
    return submit(P2P{true, buf, nullptr, count * dtype_size(dt), peer, comm, s});
}
ncclResult_t ncclRecv(void* buf, size_t count, ncclDataType_t dt, int peer, ncclComm_t comm, cudaStream_t s) {
    if (peer < 0 || peer >= comm->sh->n) return ncclInvalidArgument;
    return submit(P2P{false, nullptr, buf, count * dtype_size(dt), peer, comm, s});
}

}  // extern "C"
