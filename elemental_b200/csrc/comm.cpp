// Communicators and the per-collective C ABI: what the reference's mpi::Comm (AluminumComm over an MPI
// communicator, include/El/core/imports/mpi/aluminum_comm.hpp:93-236) and its GPU collectives
//   mpi::AllGather / ReduceScatter / AllReduce / AllToAll / SendRecv / Broadcast  <Device::GPU>
// (src/core/imports/mpi/AllGather.hpp:8-21, ReduceScatter.hpp:36-52,149-165, AllReduce.hpp, AllToAll.hpp,
//  SendRecv.hpp, Broadcast.hpp) are replaced by. There is no MPI here: a communicator is an NCCL communicator
// plus, when it is one of a Grid's MC / MR / VC communicators, the grid's NVLink peer arena — the zero-SM
// transport the SUMMA pipelines use (peer.hpp) — so that code calling the collectives one by one (the reference's
// own copy:: routines, INTEGRATION.md level 2) reaches the same data path as the fused engine calls.
//
// Bootstrap of the process-wide world communicator (elb200_comm_world): RANK / WORLD_SIZE from the environment
// (torchrun's names), the NCCL unique id travels from rank 0 either through a file (ELB200_ID_FILE) or over TCP
// (MASTER_ADDR : ELB200_PORT, default MASTER_PORT + 29). The reference gets its world from MPI_Init_thread
// (src/core/imports/mpi.cpp:80-91).
#include <arpa/inet.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <sys/socket.h>
#include <unistd.h>

#include <chrono>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <memory>
#include <mutex>
#include <thread>

#include "capi_handles.hpp"
#include "engine.hpp"
#include "engine_dev.hpp"

namespace elb {
void set_last_error(const char* what, const char* detail);
}
using namespace elb;

// One communicator handle. `nccl` is null for a single-rank communicator. A handle obtained from a Grid borrows
// the grid's NCCL communicator (owns == false) and keeps the grid alive.
struct elb200_comm_s {
    ncclComm_t nccl = nullptr;
    int rank = 0, size = 1;
    bool owns = false;
    std::shared_ptr<Grid> grid;  // non-null: this is the grid's `dist` communicator
    Dist dist = VC;
    bool arena_ok = false;       // MC / MR / VC communicator of a grid that has a peer arena
    // NCCL does not allow concurrent operations on one communicator: calls made on different streams are chained
    // in issue order through this event (grid communicators use the grid's own chain, Grid::NcclBegin / NcclEnd)
    cudaEvent_t order = nullptr;
    bool recorded = false;
    ~elb200_comm_s() {
        if (owns && nccl) ncclCommDestroy(nccl);
        if (order) cudaEventDestroy(order);
    }
};

namespace {

template <typename F>
int guarded(const char* where, F&& fn) {
    try {
        fn();
        return ELB200_OK;
    } catch (const EngineError& e) {
        set_last_error(where, e.what());
        return e.status;
    } catch (const std::exception& e) {
        set_last_error(where, e.what());
        return ELB200_ERR_STATE;
    }
}

size_t elem_bytes(int dtype) {
    switch (dtype) {
        case ELB200_F32: return 4;
        case ELB200_F64: return 8;
        case ELB200_BYTE: return 1;
        case ELB200_F16: return 2;
        default: throw EngineError(ELB200_ERR_ARG, "unknown dtype (ELB200_F32, ELB200_F64, ELB200_F16 or ELB200_BYTE)");
    }
}
ncclDataType_t nccl_type(int dtype) { return dtype == ELB200_F32 ? ncclFloat : dtype == ELB200_F64 ? ncclDouble : dtype == ELB200_F16 ? ncclHalf : ncclChar; }

struct NcclOrder {  // RAII bracket around one NCCL call (group)
    elb200_comm_s& c;
    cudaStream_t s;
    NcclOrder(elb200_comm_s& comm, cudaStream_t stream) : c(comm), s(stream) {
        if (c.grid) { c.grid->NcclBegin(s); return; }
        if (c.recorded) cuda_check(cudaStreamWaitEvent(s, c.order, 0), "cudaStreamWaitEvent(nccl order)");
    }
    ~NcclOrder() {
        if (c.grid) { c.grid->NcclEnd(s); return; }
        if (!c.order && cudaEventCreateWithFlags(&c.order, cudaEventDisableTiming) != cudaSuccess) return;
        if (cudaEventRecord(c.order, s) == cudaSuccess) c.recorded = true;
    }
};


// ---- id rendezvous without MPI ---------------------------------------------------------------------------
int env_int(const char* name, int dflt) {
    const char* v = std::getenv(name);
    return v && *v ? std::atoi(v) : dflt;
}

void send_all(int fd, const void* p, size_t n) {
    const char* c = static_cast<const char*>(p);
    while (n) {
        const ssize_t w = ::send(fd, c, n, MSG_NOSIGNAL);
        if (w <= 0) throw EngineError(ELB200_ERR_STATE, "world bootstrap: send failed");
        c += w; n -= size_t(w);
    }
}
void recv_all(int fd, void* p, size_t n) {
    char* c = static_cast<char*>(p);
    while (n) {
        const ssize_t r = ::recv(fd, c, n, 0);
        if (r <= 0) throw EngineError(ELB200_ERR_STATE, "world bootstrap: recv failed");
        c += r; n -= size_t(r);
    }
}

void exchange_id_tcp(ncclUniqueId* id, int rank, int size) {
    const char* addr = std::getenv("MASTER_ADDR");
    if (!addr || !*addr) addr = "127.0.0.1";
    const int port = env_int("ELB200_PORT", env_int("MASTER_PORT", 29488) + 29);
    if (rank == 0) {
        const int ls = ::socket(AF_INET, SOCK_STREAM, 0);
        if (ls < 0) throw EngineError(ELB200_ERR_STATE, "world bootstrap: socket failed");
        int one = 1;
        ::setsockopt(ls, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
        sockaddr_in sa{};
        sa.sin_family = AF_INET;
        sa.sin_addr.s_addr = htonl(INADDR_ANY);
        sa.sin_port = htons(uint16_t(port));
        if (::bind(ls, reinterpret_cast<sockaddr*>(&sa), sizeof(sa)) != 0 || ::listen(ls, size) != 0) {
            ::close(ls);
            throw EngineError(ELB200_ERR_STATE, "world bootstrap: cannot listen on port " + std::to_string(port));
        }
        for (int i = 1; i < size; ++i) {
            const int fd = ::accept(ls, nullptr, nullptr);
            if (fd < 0) { ::close(ls); throw EngineError(ELB200_ERR_STATE, "world bootstrap: accept failed"); }
            try { send_all(fd, id, sizeof(*id)); } catch (...) { ::close(fd); ::close(ls); throw; }
            ::close(fd);
        }
        ::close(ls);
        return;
    }
    addrinfo hints{}, *res = nullptr;
    hints.ai_family = AF_INET;
    hints.ai_socktype = SOCK_STREAM;
    if (::getaddrinfo(addr, std::to_string(port).c_str(), &hints, &res) != 0 || !res)
        throw EngineError(ELB200_ERR_STATE, std::string("world bootstrap: cannot resolve ") + addr);
    const auto t0 = std::chrono::steady_clock::now();
    for (;;) {
        const int fd = ::socket(AF_INET, SOCK_STREAM, 0);
        if (fd >= 0 && ::connect(fd, res->ai_addr, res->ai_addrlen) == 0) {
            try { recv_all(fd, id, sizeof(*id)); } catch (...) { ::close(fd); ::freeaddrinfo(res); throw; }
            ::close(fd);
            break;
        }
        if (fd >= 0) ::close(fd);
        if (std::chrono::steady_clock::now() - t0 > std::chrono::seconds(120)) {
            ::freeaddrinfo(res);
            throw EngineError(ELB200_ERR_STATE, "world bootstrap: rank 0 did not answer within 120 s");
        }
        std::this_thread::sleep_for(std::chrono::milliseconds(50));
    }
    ::freeaddrinfo(res);
}

void exchange_id_file(ncclUniqueId* id, int rank, const char* path) {
    if (rank == 0) {
        const std::string tmp = std::string(path) + ".tmp";
        std::ofstream(tmp, std::ios::binary).write(reinterpret_cast<const char*>(id), sizeof(*id));
        std::rename(tmp.c_str(), path);  // atomic publish
        return;
    }
    for (int tries = 0;; ++tries) {
        std::ifstream f(path, std::ios::binary);
        if (f && f.read(reinterpret_cast<char*>(id), sizeof(*id))) return;
        if (tries > 1200) throw EngineError(ELB200_ERR_STATE, "world bootstrap: timed out waiting for ELB200_ID_FILE");
        std::this_thread::sleep_for(std::chrono::milliseconds(100));
    }
}

std::mutex g_world_mutex;
elb200_comm_s* g_world = nullptr;

elb200_comm_s* make_comm(const ncclUniqueId* id, int rank, int size) {
    if (size < 1 || rank < 0 || rank >= size) throw EngineError(ELB200_ERR_ARG, "comm: bad rank / size");
    std::unique_ptr<elb200_comm_s> c(new elb200_comm_s);
    c->rank = rank; c->size = size;
    if (size > 1) {
        if (!id) throw EngineError(ELB200_ERR_ARG, "comm: size > 1 needs an NCCL unique id");
        nccl_check(ncclCommInitRank(&c->nccl, size, *id, rank), "ncclCommInitRank");
        c->owns = true;
    }
    return c.release();
}

// arena transport of a grid communicator: members (VC ranks in communicator order) and my index
struct ArenaView {
    PeerArena* ar;
    int channel;
    std::vector<int> members;
    int idx;
};
bool arena_view(const elb200_comm_s& c, size_t slot_need, ArenaView* v) {
    if (!c.arena_ok || !c.grid->arena) return false;
    if (slot_need > c.grid->arena->slot_bytes()) return false;  // same decision on every rank: equal counts
    v->ar = c.grid->arena;
    v->channel = Grid::arena_channel(c.dist);
    v->members = c.grid->comm_members(c.dist);
    v->idx = c.grid->comm_index(c.dist);
    return true;
}

}  // namespace

extern "C" {

int elb200_comm_create(const void* id128, int rank, int size, elb200_comm_t* out) {
    if (!out) return ELB200_ERR_ARG;
    return guarded("elb200_comm_create", [&] { *out = make_comm(static_cast<const ncclUniqueId*>(id128), rank, size); });
}

int elb200_comm_world(elb200_comm_t* out) {
    if (!out) return ELB200_ERR_ARG;
    return guarded("elb200_comm_world", [&] {
        std::lock_guard<std::mutex> lock(g_world_mutex);
        if (!g_world) {
            const int rank = env_int("RANK", 0), size = env_int("WORLD_SIZE", 1);
            ncclUniqueId id;
            std::memset(&id, 0, sizeof(id));
            if (size > 1) {
                if (rank == 0) nccl_check(ncclGetUniqueId(&id), "ncclGetUniqueId");
                const char* file = std::getenv("ELB200_ID_FILE");
                if (file && *file) exchange_id_file(&id, rank, file);
                else exchange_id_tcp(&id, rank, size);
            }
            g_world = make_comm(&id, rank, size);
        }
        *out = g_world;
    });
}

int elb200_comm_rank(elb200_comm_t c) { return c ? c->rank : -1; }
int elb200_comm_size(elb200_comm_t c) { return c ? c->size : 0; }

int elb200_comm_split(elb200_comm_t c, int color, int key, elb200_comm_t* out) {
    if (!c || !out) return ELB200_ERR_ARG;
    return guarded("elb200_comm_split", [&] {
        *out = nullptr;
        if (c->size == 1) {
            if (color >= 0) *out = make_comm(nullptr, 0, 1);
            return;
        }
        ncclComm_t nc = nullptr;
        nccl_check(ncclCommSplit(c->nccl, color < 0 ? NCCL_SPLIT_NOCOLOR : color, key, &nc, nullptr), "ncclCommSplit");
        if (!nc) return;  // color < 0: this rank is in no new communicator (MPI_UNDEFINED)
        std::unique_ptr<elb200_comm_s> n(new elb200_comm_s);
        n->nccl = nc; n->owns = true;
        nccl_check(ncclCommUserRank(nc, &n->rank), "ncclCommUserRank");
        nccl_check(ncclCommCount(nc, &n->size), "ncclCommCount");
        *out = n.release();
    });
}

int elb200_comm_dup(elb200_comm_t c, elb200_comm_t* out) {
    if (!c) return ELB200_ERR_ARG;
    return elb200_comm_split(c, 0, c->rank, out);
}

int elb200_comm_free(elb200_comm_t c) {
    if (!c) return ELB200_OK;
    {
        std::lock_guard<std::mutex> lock(g_world_mutex);
        if (c == g_world) g_world = nullptr;
    }
    delete c;
    return ELB200_OK;
}

int elb200_comm_barrier(elb200_comm_t c) {
    if (!c) return ELB200_ERR_ARG;
    if (c->size == 1) return ELB200_OK;
    return guarded("elb200_comm_barrier", [&] {
        cudaStream_t s = c->grid ? c->grid->stream : nullptr;
        int* token = nullptr;
        cuda_check(cudaMalloc(reinterpret_cast<void**>(&token), sizeof(int)), "cudaMalloc(barrier)");
        cudaMemsetAsync(token, 0, sizeof(int), s);
        ncclResult_t r;
        {
            NcclOrder o(*c, s);
            r = ncclAllReduce(token, token, 1, ncclInt, ncclSum, c->nccl, s);
        }
        const cudaError_t e = cudaStreamSynchronize(s);
        cudaFree(token);
        nccl_check(r, "ncclAllReduce(barrier)");
        cuda_check(e, "barrier sync");
    });
}

// El::Grid(mpi::Comm comm, int height, GridOrder order) (include/El/core/Grid.hpp:19-20, src/core/Grid.cpp:74-223):
// every rank of `comm` calls; height <= 0 picks Grid::DefaultHeight; order 1 = COLUMN_MAJOR (rank = row + r*col,
// the reference default), 0 = ROW_MAJOR (rank = col + c*row). device < 0: the current device.
int elb200_grid_create_on(elb200_comm_t c, int height, int order, int device, elb200_grid_t* out) {
    if (!c || !out || (order != 0 && order != 1)) return ELB200_ERR_ARG;
    return guarded("elb200_grid_create_on", [&] {
        const int p = c->size;
        const int h = height > 0 ? height : elb200_default_grid_height(p);
        if (p % h != 0) throw EngineError(ELB200_ERR_ARG, "Grid: height must divide the number of processes (Grid.cpp:87-90)");
        const int w = p / h;
        const int vc = order == 1 ? c->rank : (c->rank / w) + h * (c->rank % w);
        int dev = device;
        if (dev < 0) cuda_check(cudaGetDevice(&dev), "cudaGetDevice");
        ncclUniqueId id;
        std::memset(&id, 0, sizeof(id));
        if (p > 1) {
            if (c->rank == 0) nccl_check(ncclGetUniqueId(&id), "ncclGetUniqueId");
            cuda_check(cudaSetDevice(dev), "cudaSetDevice");
            char* buf = nullptr;
            cuda_check(cudaMalloc(reinterpret_cast<void**>(&buf), sizeof(id)), "cudaMalloc(id)");
            cudaStream_t s = nullptr;
            cudaError_t e = cudaMemcpyAsync(buf, &id, sizeof(id), cudaMemcpyHostToDevice, s);
            ncclResult_t r = ncclSuccess;
            if (e == cudaSuccess) {
                NcclOrder o(*c, s);
                r = ncclBroadcast(buf, buf, sizeof(id), ncclChar, 0, c->nccl, s);
            }
            if (e == cudaSuccess && r == ncclSuccess) e = cudaMemcpyAsync(&id, buf, sizeof(id), cudaMemcpyDeviceToHost, s);
            if (e == cudaSuccess) e = cudaStreamSynchronize(s);
            cudaFree(buf);
            nccl_check(r, "ncclBroadcast(grid id)");
            cuda_check(e, "grid id broadcast");
        }
        std::unique_ptr<elb200_grid_s> g(new elb200_grid_s);
        g->g = std::make_shared<Grid>(h, w, vc, p > 1 ? &id : nullptr, dev);
        *out = g.release();
    });
}

// Grid::MCComm / MRComm / VCComm / VRComm (include/El/core/Grid.hpp:62-75). The handle borrows the grid's
// communicator and keeps the grid alive; free it with elb200_comm_free. which: 0 MC (my grid column, ranked by row),
// 1 MR (my grid row, ranked by column), 2 VC (all, column-major), 3 VR (all, row-major; created on first use —
// collective over the grid).
int elb200_grid_comm(elb200_grid_t gh, int which, elb200_comm_t* out) {
    if (!gh || !gh->g || !out || which < 0 || which > 3) return ELB200_ERR_ARG;
    return guarded("elb200_grid_comm", [&] {
        Grid& g = *gh->g;
        std::unique_ptr<elb200_comm_s> c(new elb200_comm_s);
        c->grid = gh->g;
        const int r = g.shape.r, cc = g.shape.c;
        if (which == 0) { c->dist = MC; c->rank = g.me.mc; c->size = r; c->nccl = r > 1 ? g.mc_comm : nullptr; }
        else if (which == 1) { c->dist = MR; c->rank = g.me.mr; c->size = cc; c->nccl = cc > 1 ? g.mr_comm : nullptr; }
        else if (which == 2) { c->dist = VC; c->rank = g.vc_rank; c->size = g.size(); c->nccl = g.size() > 1 ? g.world : nullptr; }
        else {
            c->dist = VR; c->rank = g.me.mr + cc * g.me.mc; c->size = g.size();
            if (g.size() > 1) {
                if (!g.vr_comm) {
                    if (!g.world) throw EngineError(ELB200_ERR_STATE, "Grid has no NCCL communicator");
                    nccl_check(ncclCommSplit(g.world, 0, c->rank, &g.vr_comm, nullptr), "ncclCommSplit(VR)");
                }
                c->nccl = g.vr_comm;
            }
        }
        if (c->size > 1 && !c->nccl) throw EngineError(ELB200_ERR_STATE, "Grid has no NCCL communicator");
        c->arena_ok = which <= 2 && g.arena != nullptr && c->size > 1;
        *out = c.release();
    });
}

// mpi::AllGather(sbuf, sc, rbuf, rc, comm, syncInfo) (src/core/imports/mpi/AllGather.hpp:8-21): rbuf[i*count ..] = rank i's sbuf
int elb200_allgather(const void* sbuf, void* rbuf, int64_t count, int dtype, elb200_comm_t c, void* stream) {
    if (!c || count < 0 || (count > 0 && (!sbuf || !rbuf))) return ELB200_ERR_ARG;
    return guarded("elb200_allgather", [&] {
        const size_t bytes = size_t(count) * elem_bytes(dtype);
        cudaStream_t s = static_cast<cudaStream_t>(stream);
        if (bytes == 0) return;
        if (c->size == 1) {  // AllGather.hpp:17: a local copy
            if (sbuf != rbuf) cuda_check(cudaMemcpyAsync(rbuf, sbuf, bytes, cudaMemcpyDeviceToDevice, s), "allgather copy");
            return;
        }
        ArenaView v;
        if (arena_view(*c, bytes * size_t(c->size), &v)) {
            PeerArena::Ticket t = v.ar->AllGather(v.channel, v.members, v.idx, sbuf, bytes, bytes, s);
            cuda_check(cudaMemcpyAsync(rbuf, t.data, bytes * size_t(c->size), cudaMemcpyDeviceToDevice, s), "allgather unstage");
            v.ar->Release(t, v.members, v.idx, s);
            return;
        }
        NcclOrder o(*c, s);
        nccl_check(ncclAllGather(sbuf, rbuf, size_t(count), nccl_type(dtype), c->nccl, s), "ncclAllGather");
    });
}

// mpi::ReduceScatter(sbuf, rbuf, rc, SUM, comm, syncInfo) (src/core/imports/mpi/ReduceScatter.hpp:36-52): sbuf holds
// size * count elements, rbuf receives the sum over ranks of block [rank*count, (rank+1)*count). In place (sbuf == rbuf,
// ReduceScatter.hpp:149-165) is allowed: the result lands at the START of the buffer, as in the reference.
int elb200_reduce_scatter(const void* sbuf, void* rbuf, int64_t count, int dtype, elb200_comm_t c, void* stream) {
    if (!c || count < 0 || (count > 0 && (!sbuf || !rbuf))) return ELB200_ERR_ARG;
    if (dtype != ELB200_F32 && dtype != ELB200_F64 && dtype != ELB200_F16) return ELB200_ERR_ARG;
    return guarded("elb200_reduce_scatter", [&] {
        const size_t esz = elem_bytes(dtype), bytes = size_t(count) * esz;
        cudaStream_t s = static_cast<cudaStream_t>(stream);
        if (bytes == 0) return;
        if (c->size == 1) {
            if (sbuf != rbuf) cuda_check(cudaMemcpyAsync(rbuf, sbuf, bytes, cudaMemcpyDeviceToDevice, s), "reduce_scatter copy");
            return;
        }
        ArenaView v;
        if (c->size <= ELB200_MAX_SEGMENTS && arena_view(*c, bytes * size_t(c->size), &v)) {
            // every member publishes its partial sums in its own slot; ONE kernel per rank then loads its block from all
            // members' slots over NVLink and sums them in communicator order (deterministic)
            PeerArena::Ticket t = v.ar->Publish(v.channel, v.members, v.idx, bytes * size_t(c->size), s);
            cuda_check(cudaMemcpyAsync(t.data, sbuf, bytes * size_t(c->size), cudaMemcpyDeviceToDevice, s), "reduce_scatter stage");
            v.ar->Signal(t, v.members, v.idx, s);
            elb200_segment seg{0, 1, count, 0, 1, 1, 0};
            const void* srcs[ELB200_MAX_SEGMENTS];
            for (int i = 0; i < c->size; ++i) srcs[i] = v.ar->PeerSlot(v.members[i], t);
            if (dtype == ELB200_F32) {
                Dev<float>::fill2d(count, 1, 0.f, static_cast<float*>(rbuf), count, s);
                Dev<float>::peer_reduce_axpy(c->size, reinterpret_cast<const float* const*>(srcs), int64_t(v.idx) * count, &seg,
                                             static_cast<float*>(rbuf), count, 1.f, s);
            } else if (dtype == ELB200_F16) {
                Dev<Half>::fill2d(count, 1, Half(0.f), static_cast<Half*>(rbuf), count, s);
                Dev<Half>::peer_reduce_axpy(c->size, reinterpret_cast<const Half* const*>(srcs), int64_t(v.idx) * count, &seg,
                                            static_cast<Half*>(rbuf), count, Half(1.f), s);
            } else {
                Dev<double>::fill2d(count, 1, 0.0, static_cast<double*>(rbuf), count, s);
                Dev<double>::peer_reduce_axpy(c->size, reinterpret_cast<const double* const*>(srcs), int64_t(v.idx) * count, &seg,
                                              static_cast<double*>(rbuf), count, 1.0, s);
            }
            v.ar->Release(t, v.members, v.idx, s);
            return;
        }
        NcclOrder o(*c, s);
        if (sbuf == rbuf) {
            // NCCL's in-place form leaves the result at block `rank`; the reference's leaves it at the start of the buffer
            char* mine = static_cast<char*>(rbuf) + size_t(c->rank) * bytes;
            nccl_check(ncclReduceScatter(sbuf, mine, size_t(count), nccl_type(dtype), ncclSum, c->nccl, s), "ncclReduceScatter");
            if (c->rank != 0) cuda_check(cudaMemcpyAsync(rbuf, mine, bytes, cudaMemcpyDeviceToDevice, s), "reduce_scatter move");
            return;
        }
        nccl_check(ncclReduceScatter(sbuf, rbuf, size_t(count), nccl_type(dtype), ncclSum, c->nccl, s), "ncclReduceScatter");
    });
}

// mpi::AllReduce(sbuf, rbuf, count, SUM, comm, syncInfo) (src/core/imports/mpi/AllReduce.hpp)
int elb200_allreduce(const void* sbuf, void* rbuf, int64_t count, int dtype, elb200_comm_t c, void* stream) {
    if (!c || count < 0 || (count > 0 && (!sbuf || !rbuf))) return ELB200_ERR_ARG;
    if (dtype != ELB200_F32 && dtype != ELB200_F64 && dtype != ELB200_F16) return ELB200_ERR_ARG;
    return guarded("elb200_allreduce", [&] {
        const size_t bytes = size_t(count) * elem_bytes(dtype);
        cudaStream_t s = static_cast<cudaStream_t>(stream);
        if (bytes == 0) return;
        if (c->size == 1) {
            if (sbuf != rbuf) cuda_check(cudaMemcpyAsync(rbuf, sbuf, bytes, cudaMemcpyDeviceToDevice, s), "allreduce copy");
            return;
        }
        NcclOrder o(*c, s);
        nccl_check(ncclAllReduce(sbuf, rbuf, size_t(count), nccl_type(dtype), ncclSum, c->nccl, s), "ncclAllReduce");
    });
}

// mpi::AllToAll(sbuf, sc, rbuf, rc, comm, syncInfo) (src/core/imports/mpi/AllToAll.hpp): block j of my sbuf goes to rank j,
// block i of my rbuf comes from rank i; `count` elements per block.
int elb200_alltoall(const void* sbuf, void* rbuf, int64_t count, int dtype, elb200_comm_t c, void* stream) {
    if (!c || count < 0 || (count > 0 && (!sbuf || !rbuf))) return ELB200_ERR_ARG;
    return guarded("elb200_alltoall", [&] {
        const size_t bytes = size_t(count) * elem_bytes(dtype);
        cudaStream_t s = static_cast<cudaStream_t>(stream);
        if (bytes == 0) return;
        if (c->size == 1) {
            if (sbuf != rbuf) cuda_check(cudaMemcpyAsync(rbuf, sbuf, bytes, cudaMemcpyDeviceToDevice, s), "alltoall copy");
            return;
        }
        if (sbuf == rbuf) throw EngineError(ELB200_ERR_ARG, "alltoall: in-place exchange is not supported");
        ArenaView v;
        if (arena_view(*c, bytes * size_t(c->size), &v)) {
            // publish my send buffer in my slot; every receiver fetches its block from each member's slot with the copy engines
            PeerArena::Ticket t = v.ar->Publish(v.channel, v.members, v.idx, bytes * size_t(c->size), s);
            cuda_check(cudaMemcpyAsync(t.data, sbuf, bytes * size_t(c->size), cudaMemcpyDeviceToDevice, s), "alltoall stage");
            v.ar->Signal(t, v.members, v.idx, s);
            std::vector<PeerArena::PullOp> pulls;
            for (int i = 0; i < c->size; ++i)
                pulls.push_back({v.members[i], size_t(v.idx) * bytes, static_cast<char*>(rbuf) + size_t(i) * bytes, bytes});
            v.ar->Pull(t, pulls, s);
            v.ar->Release(t, v.members, v.idx, s);
            return;
        }
        NcclOrder o(*c, s);
        nccl_check(ncclGroupStart(), "ncclGroupStart");
        for (int i = 0; i < c->size; ++i) {
            nccl_check(ncclSend(static_cast<const char*>(sbuf) + size_t(i) * bytes, size_t(count), nccl_type(dtype), i, c->nccl, s), "ncclSend");
            nccl_check(ncclRecv(static_cast<char*>(rbuf) + size_t(i) * bytes, size_t(count), nccl_type(dtype), i, c->nccl, s), "ncclRecv");
        }
        nccl_check(ncclGroupEnd(), "ncclGroupEnd");
    });
}

// mpi::SendRecv(sbuf, sc, to, rbuf, rc, from, comm, syncInfo) (src/core/imports/mpi/SendRecv.hpp)
int elb200_sendrecv(const void* sbuf, int64_t scount, int to, void* rbuf, int64_t rcount, int from, int dtype,
                    elb200_comm_t c, void* stream) {
    if (!c || scount < 0 || rcount < 0 || to < 0 || from < 0 || to >= c->size || from >= c->size) return ELB200_ERR_ARG;
    return guarded("elb200_sendrecv", [&] {
        const size_t esz = elem_bytes(dtype);
        cudaStream_t s = static_cast<cudaStream_t>(stream);
        if (to == c->rank && from == c->rank) {
            if (scount != rcount) throw EngineError(ELB200_ERR_ARG, "sendrecv to self: counts differ");
            if (scount && sbuf != rbuf) cuda_check(cudaMemcpyAsync(rbuf, sbuf, size_t(scount) * esz, cudaMemcpyDeviceToDevice, s), "sendrecv copy");
            return;
        }
        NcclOrder o(*c, s);
        nccl_check(ncclGroupStart(), "ncclGroupStart");
        if (scount) nccl_check(ncclSend(sbuf, size_t(scount), nccl_type(dtype), to, c->nccl, s), "ncclSend");
        if (rcount) nccl_check(ncclRecv(rbuf, size_t(rcount), nccl_type(dtype), from, c->nccl, s), "ncclRecv");
        nccl_check(ncclGroupEnd(), "ncclGroupEnd");
    });
}

// mpi::Broadcast(buf, count, root, comm, syncInfo) (src/core/imports/mpi/Broadcast.hpp)
int elb200_bcast(void* buf, int64_t count, int dtype, int root, elb200_comm_t c, void* stream) {
    if (!c || count < 0 || root < 0 || root >= c->size || (count > 0 && !buf)) return ELB200_ERR_ARG;
    return guarded("elb200_bcast", [&] {
        const size_t n = size_t(count);
        elem_bytes(dtype);
        if (n == 0 || c->size == 1) return;
        cudaStream_t s = static_cast<cudaStream_t>(stream);
        NcclOrder o(*c, s);
        nccl_check(ncclBroadcast(buf, buf, n, nccl_type(dtype), root, c->nccl, s), "ncclBroadcast");
    });
}

}  // extern "C"
