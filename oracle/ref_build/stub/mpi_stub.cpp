// ORACLE BUILD ONLY (test infrastructure). Implementation of stub/mpi.h: a deliberately simple
// MPI subset over a shared-memory file, for running the REFERENCE multi-rank on one box.
// Design: NP processes (ELB_STUB_NP / ELB_STUB_RANK) mmap ELB_STUB_SHM; for every ordered pair
// (src,dst) there is a single-producer single-consumer byte ring.  A message is a 24-byte header
// {context, tag, bytes} followed by the payload.  Every collective is built from point-to-point
// messages through the communicator's first member, reductions are applied in rank order.
// Correct, not fast: golden cases are kilobytes.
#include "mpi.h"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <cstdint>
#include <cstdio>
#include <deque>
#include <vector>

namespace {

constexpr size_t kRingBytes = size_t(8) << 20;
struct Ring {
    std::atomic<uint64_t> head;  // bytes consumed
    std::atomic<uint64_t> tail;  // bytes produced
    char pad[48];
    char data[kRingBytes];
};
struct Header { int ctx, tag; long bytes; long pad; };
struct Shared { std::atomic<int> ready; int np; char pad[56]; Ring rings[1]; };

struct Msg { int ctx, tag; std::vector<char> data; };
struct CommInfo { int ctx = 0; std::vector<int> ranks; int me = -1; std::vector<int> dims; bool live = false; };
struct Pending { bool live = false; bool recv = false; void* buf = nullptr; long cap = 0; int src = 0, tag = 0, comm = 0; };

int g_np = 1, g_rank = 0, g_inited = 0, g_finalized = 0, g_next_ctx = 16;
Shared* g_shm = nullptr;
std::vector<CommInfo> g_comms;            // handle -> info
std::vector<std::vector<int>> g_groups;   // handle -> world ranks
std::vector<MPI_User_function*> g_ops;    // handle - 100 -> user function
std::vector<std::deque<Msg>> g_unexpected;  // per source world rank
std::vector<Pending> g_pending;

inline long extent(MPI_Datatype t) { return long(t & 0xffffffffL); }
inline int kind(MPI_Datatype t) { return int(t >> 32); }
inline Ring& ring(int src, int dst) { return g_shm->rings[size_t(src) * g_np + dst]; }

[[noreturn]] void die(const char* what) {
    std::fprintf(stderr, "[mpi stub rank %d] %s\n", g_rank, what);
    std::fflush(stderr);
    _exit(86);
}

void ring_write(Ring& r, const void* src, size_t n) {
    const char* p = static_cast<const char*>(src);
    while (n) {
        const uint64_t head = r.head.load(std::memory_order_acquire), tail = r.tail.load(std::memory_order_relaxed);
        const size_t space = kRingBytes - size_t(tail - head);
        if (!space) { usleep(50); continue; }
        const size_t off = size_t(tail % kRingBytes);
        const size_t chunk = std::min({n, space, kRingBytes - off});
        memcpy(r.data + off, p, chunk);
        r.tail.store(tail + chunk, std::memory_order_release);
        p += chunk; n -= chunk;
    }
}
void ring_read(Ring& r, void* dst, size_t n) {
    char* p = static_cast<char*>(dst);
    int idle = 0;
    while (n) {
        const uint64_t tail = r.tail.load(std::memory_order_acquire), head = r.head.load(std::memory_order_relaxed);
        const size_t avail = size_t(tail - head);
        if (!avail) { usleep(50); if (++idle > 1200000) die("receive timed out (60 s): deadlock in the stub or the caller"); continue; }
        idle = 0;
        const size_t off = size_t(head % kRingBytes);
        const size_t chunk = std::min({n, avail, kRingBytes - off});
        memcpy(p, r.data + off, chunk);
        r.head.store(head + chunk, std::memory_order_release);
        p += chunk; n -= chunk;
    }
}
bool ring_has_data(Ring& r) { return r.tail.load(std::memory_order_acquire) != r.head.load(std::memory_order_relaxed); }

void raw_send(int dst_world, int ctx, int tag, const void* buf, long bytes) {
    if (dst_world == g_rank) {
        Msg m; m.ctx = ctx; m.tag = tag; m.data.assign(static_cast<const char*>(buf), static_cast<const char*>(buf) + bytes);
        g_unexpected[g_rank].push_back(std::move(m));
        return;
    }
    Header h{ctx, tag, bytes, 0};
    Ring& r = ring(g_rank, dst_world);
    ring_write(r, &h, sizeof(h));
    ring_write(r, buf, size_t(bytes));
}
// pull the next message from src's ring into the unexpected queue
void pump(int src_world) {
    Header h;
    Ring& r = ring(src_world, g_rank);
    ring_read(r, &h, sizeof(h));
    Msg m; m.ctx = h.ctx; m.tag = h.tag; m.data.resize(size_t(h.bytes));
    ring_read(r, m.data.data(), size_t(h.bytes));
    g_unexpected[src_world].push_back(std::move(m));
}
bool match_queued(int src_world, int ctx, int tag, void* buf, long cap, MPI_Status* st, bool consume) {
    std::deque<Msg>& q = g_unexpected[src_world];
    for (size_t i = 0; i < q.size(); ++i)
        if (q[i].ctx == ctx && (tag == MPI_ANY_TAG || q[i].tag == tag)) {
            if (st) { st->MPI_TAG = q[i].tag; st->MPI_ERROR = 0; st->count_bytes = int(q[i].data.size()); }
            if (consume) {
                if (long(q[i].data.size()) > cap) die("message longer than the receive buffer");
                memcpy(buf, q[i].data.data(), q[i].data.size());
                q.erase(q.begin() + long(i));
            }
            return true;
        }
    return false;
}
void raw_recv(int src_world, int ctx, int tag, void* buf, long cap, MPI_Status* st) {
    for (;;) {
        if (match_queued(src_world, ctx, tag, buf, cap, st, true)) return;
        if (src_world == g_rank) die("receive from self without a matching send");
        pump(src_world);
    }
}

CommInfo& comm(MPI_Comm c) {
    if (c <= 0 || c >= int(g_comms.size()) || !g_comms[c].live) die("invalid communicator");
    return g_comms[c];
}
int new_comm(int ctx, const std::vector<int>& ranks) {
    CommInfo ci; ci.ctx = ctx; ci.ranks = ranks; ci.live = true;
    for (size_t i = 0; i < ranks.size(); ++i) if (ranks[i] == g_rank) ci.me = int(i);
    g_comms.push_back(ci);
    return int(g_comms.size()) - 1;
}
int new_group(const std::vector<int>& ranks) {
    if (ranks.empty()) return MPI_GROUP_EMPTY;
    g_groups.push_back(ranks);
    return int(g_groups.size()) - 1;
}
const std::vector<int>& group(MPI_Group g) {
    static const std::vector<int> empty;
    if (g == MPI_GROUP_EMPTY || g == MPI_GROUP_NULL) return empty;
    if (g < 0 || g >= int(g_groups.size())) die("invalid group");
    return g_groups[g];
}

// internal tags live below every user tag
enum { T_BARRIER = -100, T_BCAST, T_GATHER, T_SCATTER, T_A2A, T_REDUCE, T_CTX, T_SPLIT, T_SR };

// all members agree on a fresh context id: max of the local counters
int agree_ctx(CommInfo& c) {
    int mine = g_next_ctx, best = mine;
    const int n = int(c.ranks.size());
    if (c.me == 0) {
        for (int i = 1; i < n; ++i) { int v; raw_recv(c.ranks[i], c.ctx, T_CTX, &v, sizeof(v), nullptr); best = std::max(best, v); }
        for (int i = 1; i < n; ++i) raw_send(c.ranks[i], c.ctx, T_CTX, &best, sizeof(best));
    } else {
        raw_send(c.ranks[0], c.ctx, T_CTX, &mine, sizeof(mine));
        raw_recv(c.ranks[0], c.ctx, T_CTX, &best, sizeof(best), nullptr);
    }
    g_next_ctx = best + 1;
    return best;
}

template <typename T>
void reduce_typed(const T* in, T* inout, long n, MPI_Op op) {
    for (long i = 0; i < n; ++i) {
        const T a = in[i], b = inout[i];
        switch (op) {
            case MPI_SUM: inout[i] = T(a + b); break;
            case MPI_PROD: inout[i] = T(a * b); break;
            case MPI_MIN: inout[i] = a < b ? a : b; break;
            case MPI_MAX: inout[i] = a > b ? a : b; break;
            case MPI_LAND: inout[i] = T((a != T(0)) && (b != T(0))); break;
            case MPI_LOR: inout[i] = T((a != T(0)) || (b != T(0))); break;
            case MPI_LXOR: inout[i] = T((a != T(0)) != (b != T(0))); break;
            default: die("unsupported reduction operator for this datatype");
        }
    }
}
template <typename T>
void reduce_bits(const T* in, T* inout, long n, MPI_Op op) {
    if (op == MPI_BAND) for (long i = 0; i < n; ++i) inout[i] = T(inout[i] & in[i]);
    else if (op == MPI_BOR) for (long i = 0; i < n; ++i) inout[i] = T(inout[i] | in[i]);
    else if (op == MPI_BXOR) for (long i = 0; i < n; ++i) inout[i] = T(inout[i] ^ in[i]);
    else reduce_typed(in, inout, n, op);
}
template <typename V>
void reduce_loc(const void* in_, void* inout_, long n, MPI_Op op) {
    struct P { V v; int i; };
    const P* in = static_cast<const P*>(in_);
    P* io = static_cast<P*>(inout_);
    for (long k = 0; k < n; ++k) {
        const bool take = op == MPI_MINLOC ? (in[k].v < io[k].v || (in[k].v == io[k].v && in[k].i < io[k].i))
                                           : (in[k].v > io[k].v || (in[k].v == io[k].v && in[k].i < io[k].i));
        if (take) io[k] = in[k];
    }
}
// inout = in (op) inout, `in` comes from the lower rank
void apply_op(const void* in, void* inout, int count, MPI_Datatype t, MPI_Op op) {
    if (op >= 100) { MPI_Datatype tt = t; int c = count; g_ops[op - 100](const_cast<void*>(in), inout, &c, &tt); return; }
    const long n = count;
    switch (kind(t)) {
        case 1: reduce_bits(static_cast<const signed char*>(in), static_cast<signed char*>(inout), n, op); break;
        case 2: reduce_bits(static_cast<const unsigned char*>(in), static_cast<unsigned char*>(inout), n, op); break;
        case 3: reduce_bits(static_cast<const short*>(in), static_cast<short*>(inout), n, op); break;
        case 4: reduce_bits(static_cast<const int*>(in), static_cast<int*>(inout), n, op); break;
        case 5: reduce_bits(static_cast<const unsigned*>(in), static_cast<unsigned*>(inout), n, op); break;
        case 6: reduce_bits(static_cast<const long*>(in), static_cast<long*>(inout), n, op); break;
        case 7: reduce_bits(static_cast<const unsigned long*>(in), static_cast<unsigned long*>(inout), n, op); break;
        case 8: reduce_typed(static_cast<const float*>(in), static_cast<float*>(inout), n, op); break;
        case 9: reduce_typed(static_cast<const double*>(in), static_cast<double*>(inout), n, op); break;
        case 10: reduce_typed(static_cast<const long double*>(in), static_cast<long double*>(inout), n, op); break;
        case 11: if (op != MPI_SUM) die("complex reduction other than SUM"); reduce_typed(static_cast<const float*>(in), static_cast<float*>(inout), 2 * n, op); break;
        case 12: if (op != MPI_SUM) die("complex reduction other than SUM"); reduce_typed(static_cast<const double*>(in), static_cast<double*>(inout), 2 * n, op); break;
        case 13: reduce_loc<float>(in, inout, n, op); break;
        case 14: reduce_loc<double>(in, inout, n, op); break;
        case 15: reduce_loc<long>(in, inout, n, op); break;
        default: die("built-in reduction on a derived datatype");
    }
}

void init_world() {
    if (g_inited) return;
    g_inited = 1;
    const char* np = getenv("ELB_STUB_NP");
    g_np = np ? atoi(np) : 1;
    if (g_np < 1) g_np = 1;
    const char* rk = getenv("ELB_STUB_RANK");
    g_rank = rk ? atoi(rk) : 0;
    g_unexpected.assign(size_t(g_np), {});
    g_comms.assign(3, CommInfo());
    g_groups.assign(2, {});
    std::vector<int> all(size_t(g_np), 0);
    for (int i = 0; i < g_np; ++i) all[size_t(i)] = i;
    g_comms[MPI_COMM_WORLD].ctx = 1; g_comms[MPI_COMM_WORLD].ranks = all; g_comms[MPI_COMM_WORLD].me = g_rank; g_comms[MPI_COMM_WORLD].live = true;
    g_comms[MPI_COMM_SELF].ctx = 2; g_comms[MPI_COMM_SELF].ranks = {g_rank}; g_comms[MPI_COMM_SELF].me = 0; g_comms[MPI_COMM_SELF].live = true;
    g_pending.assign(1, Pending());
    if (g_np == 1) return;
    const char* path = getenv("ELB_STUB_SHM");
    if (!path) die("ELB_STUB_NP > 1 needs ELB_STUB_SHM");
    const size_t bytes = sizeof(Shared) + sizeof(Ring) * (size_t(g_np) * g_np);
    int fd = -1;
    if (g_rank == 0) {
        fd = open(path, O_RDWR | O_CREAT | O_TRUNC, 0600);
        if (fd < 0 || ftruncate(fd, off_t(bytes)) != 0) die("cannot create the shared file");
    } else {
        for (int tries = 0; tries < 60000; ++tries) {
            fd = open(path, O_RDWR);
            struct stat sb;
            if (fd >= 0 && fstat(fd, &sb) == 0 && size_t(sb.st_size) == bytes) break;
            if (fd >= 0) { close(fd); fd = -1; }
            usleep(1000);
        }
        if (fd < 0) die("shared file never appeared");
    }
    void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (p == MAP_FAILED) die("mmap failed");
    g_shm = static_cast<Shared*>(p);  // a fresh file is zero-filled: heads == tails == 0
    if (g_rank == 0) { g_shm->np = g_np; g_shm->ready.store(1, std::memory_order_release); }
    else while (g_shm->ready.load(std::memory_order_acquire) != 1) usleep(1000);
}

}  // namespace

extern "C" {

int MPI_Init(int*, char***) { init_world(); return MPI_SUCCESS; }
int MPI_Init_thread(int*, char***, int required, int* provided) { init_world(); if (provided) *provided = required; return MPI_SUCCESS; }
int MPI_Query_thread(int* provided) { *provided = MPI_THREAD_MULTIPLE; return MPI_SUCCESS; }
int MPI_Initialized(int* flag) { *flag = g_inited; return MPI_SUCCESS; }
int MPI_Finalized(int* flag) { *flag = g_finalized; return MPI_SUCCESS; }
int MPI_Finalize(void) { if (g_inited && !g_finalized && g_np > 1) MPI_Barrier(MPI_COMM_WORLD); g_finalized = 1; return MPI_SUCCESS; }
int MPI_Abort(MPI_Comm, int code) { _exit(code ? code : 1); }
double MPI_Wtime(void) { struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return double(ts.tv_sec) + 1e-9 * double(ts.tv_nsec); }
int MPI_Error_string(int, char* s, int* len) { strcpy(s, "mpi stub error"); *len = int(strlen(s)); return MPI_SUCCESS; }
int MPI_Comm_set_errhandler(MPI_Comm, MPI_Errhandler) { return MPI_SUCCESS; }

/* ---- communicators and groups ---- */
int MPI_Comm_rank(MPI_Comm c, int* r) { *r = comm(c).me; return MPI_SUCCESS; }
int MPI_Comm_size(MPI_Comm c, int* s) { *s = int(comm(c).ranks.size()); return MPI_SUCCESS; }
int MPI_Comm_dup(MPI_Comm c, MPI_Comm* n) {
    CommInfo ci = comm(c);
    const int ctx = agree_ctx(ci);
    *n = new_comm(ctx, ci.ranks);
    g_comms[*n].dims = ci.dims;
    return MPI_SUCCESS;
}
int MPI_Comm_split(MPI_Comm c, int color, int key, MPI_Comm* n) {
    CommInfo ci = comm(c);
    const int sz = int(ci.ranks.size());
    std::vector<int> ck(size_t(2 * sz));
    int mine[2] = {color, key};
    MPI_Allgather(mine, 2, MPI_INT, ck.data(), 2, MPI_INT, c);
    const int ctx = agree_ctx(ci);
    if (color == MPI_UNDEFINED) { *n = MPI_COMM_NULL; return MPI_SUCCESS; }
    std::vector<std::pair<std::pair<int, int>, int>> mem;  // ((key, old rank), world rank)
    for (int i = 0; i < sz; ++i)
        if (ck[size_t(2 * i)] == color) mem.push_back({{ck[size_t(2 * i + 1)], i}, ci.ranks[size_t(i)]});
    std::sort(mem.begin(), mem.end());
    std::vector<int> ranks;
    for (auto& m : mem) ranks.push_back(m.second);
    *n = new_comm(ctx, ranks);
    return MPI_SUCCESS;
}
int MPI_Comm_create(MPI_Comm c, MPI_Group grp, MPI_Comm* n) {
    CommInfo ci = comm(c);
    const int ctx = agree_ctx(ci);
    const std::vector<int>& g = group(grp);
    if (std::find(g.begin(), g.end(), g_rank) == g.end()) { *n = MPI_COMM_NULL; return MPI_SUCCESS; }
    *n = new_comm(ctx, g);
    return MPI_SUCCESS;
}
int MPI_Comm_free(MPI_Comm* c) { if (*c > MPI_COMM_SELF && *c < int(g_comms.size())) g_comms[*c].live = false; *c = MPI_COMM_NULL; return MPI_SUCCESS; }
int MPI_Comm_group(MPI_Comm c, MPI_Group* g) { *g = new_group(comm(c).ranks); return MPI_SUCCESS; }
int MPI_Comm_compare(MPI_Comm a, MPI_Comm b, int* r) {
    if (a == b) *r = MPI_IDENT;
    else if (a == MPI_COMM_NULL || b == MPI_COMM_NULL) *r = MPI_UNEQUAL;
    else *r = comm(a).ranks == comm(b).ranks ? MPI_CONGRUENT : MPI_UNEQUAL;
    return MPI_SUCCESS;
}
int MPI_Cart_create(MPI_Comm c, int ndims, const int* dims, const int*, int, MPI_Comm* n) {
    CommInfo ci = comm(c);
    const int ctx = agree_ctx(ci);
    long total = 1;
    for (int i = 0; i < ndims; ++i) total *= dims[i];
    if (total != long(ci.ranks.size())) die("Cart_create: dims do not cover the communicator");
    *n = new_comm(ctx, ci.ranks);  // no reordering: cartesian rank == old rank, row-major coordinates
    g_comms[*n].dims.assign(dims, dims + ndims);
    return MPI_SUCCESS;
}
int MPI_Cart_sub(MPI_Comm c, const int* remain, MPI_Comm* n) {
    CommInfo ci = comm(c);
    const int nd = int(ci.dims.size());
    if (nd == 0) die("Cart_sub on a non-cartesian communicator");
    // coordinates of a rank (row-major: last dimension varies fastest, as MPI defines)
    auto coords = [&](int r) { std::vector<int> x(size_t(nd), 0); for (int d = nd - 1; d >= 0; --d) { x[size_t(d)] = r % ci.dims[size_t(d)]; r /= ci.dims[size_t(d)]; } return x; };
    const std::vector<int> mine = coords(ci.me);
    int color = 0, key = 0;
    for (int d = 0; d < nd; ++d) {
        if (remain[d]) key = key * ci.dims[size_t(d)] + mine[size_t(d)];
        else color = color * ci.dims[size_t(d)] + mine[size_t(d)];
    }
    MPI_Comm_split(c, color, key, n);
    std::vector<int> kept;
    for (int d = 0; d < nd; ++d) if (remain[d]) kept.push_back(ci.dims[size_t(d)]);
    if (*n != MPI_COMM_NULL) g_comms[*n].dims = kept;
    return MPI_SUCCESS;
}
int MPI_Group_size(MPI_Group g, int* s) { *s = int(group(g).size()); return MPI_SUCCESS; }
int MPI_Group_rank(MPI_Group g, int* r) {
    const std::vector<int>& v = group(g);
    auto it = std::find(v.begin(), v.end(), g_rank);
    *r = it == v.end() ? MPI_UNDEFINED : int(it - v.begin());
    return MPI_SUCCESS;
}
int MPI_Group_free(MPI_Group* g) { *g = MPI_GROUP_NULL; return MPI_SUCCESS; }
int MPI_Group_dup(MPI_Group g, MPI_Group* n) { *n = new_group(group(g)); return MPI_SUCCESS; }
int MPI_Group_incl(MPI_Group g, int n, const int* ranks, MPI_Group* o) {
    const std::vector<int>& v = group(g);
    std::vector<int> out;
    for (int i = 0; i < n; ++i) out.push_back(v[size_t(ranks[i])]);
    *o = new_group(out);
    return MPI_SUCCESS;
}
int MPI_Group_excl(MPI_Group g, int n, const int* ranks, MPI_Group* o) {
    const std::vector<int>& v = group(g);
    std::vector<int> out;
    for (int i = 0; i < int(v.size()); ++i) if (std::find(ranks, ranks + n, i) == ranks + n) out.push_back(v[size_t(i)]);
    *o = new_group(out);
    return MPI_SUCCESS;
}
int MPI_Group_union(MPI_Group a, MPI_Group b, MPI_Group* o) {
    std::vector<int> out = group(a);
    for (int r : group(b)) if (std::find(out.begin(), out.end(), r) == out.end()) out.push_back(r);
    *o = new_group(out);
    return MPI_SUCCESS;
}
int MPI_Group_difference(MPI_Group a, MPI_Group b, MPI_Group* o) {
    const std::vector<int>& vb = group(b);
    std::vector<int> out;
    for (int r : group(a)) if (std::find(vb.begin(), vb.end(), r) == vb.end()) out.push_back(r);
    *o = new_group(out);
    return MPI_SUCCESS;
}
int MPI_Group_compare(MPI_Group a, MPI_Group b, int* r) {
    std::vector<int> va = group(a), vb = group(b);
    if (va == vb) { *r = MPI_IDENT; return MPI_SUCCESS; }
    std::sort(va.begin(), va.end()); std::sort(vb.begin(), vb.end());
    *r = va == vb ? MPI_SIMILAR : MPI_UNEQUAL;
    return MPI_SUCCESS;
}
int MPI_Group_translate_ranks(MPI_Group a, int n, const int* r1, MPI_Group b, int* r2) {
    const std::vector<int>&va = group(a), &vb = group(b);
    for (int i = 0; i < n; ++i) {
        auto it = std::find(vb.begin(), vb.end(), va[size_t(r1[i])]);
        r2[i] = it == vb.end() ? MPI_UNDEFINED : int(it - vb.begin());
    }
    return MPI_SUCCESS;
}

/* ---- datatypes: derived types are opaque byte blocks ---- */
int MPI_Type_contiguous(int count, MPI_Datatype old, MPI_Datatype* n) { *n = count * extent(old); return MPI_SUCCESS; }
int MPI_Type_create_struct(int count, const int* bl, const MPI_Aint* disp, const MPI_Datatype* types, MPI_Datatype* n) {
    long ext = 0;
    for (int i = 0; i < count; ++i) ext = std::max(ext, long(disp[i]) + bl[i] * extent(types[i]));
    *n = ext;
    return MPI_SUCCESS;
}
int MPI_Type_create_resized(MPI_Datatype, MPI_Aint, MPI_Aint ext, MPI_Datatype* n) { *n = long(ext); return MPI_SUCCESS; }
int MPI_Type_commit(MPI_Datatype*) { return MPI_SUCCESS; }
int MPI_Type_free(MPI_Datatype* t) { *t = MPI_DATATYPE_NULL; return MPI_SUCCESS; }
int MPI_Get_address(const void* p, MPI_Aint* a) { *a = MPI_Aint(p); return MPI_SUCCESS; }
int MPI_Op_create(MPI_User_function* f, int, MPI_Op* op) { g_ops.push_back(f); *op = 100 + int(g_ops.size()) - 1; return MPI_SUCCESS; }
int MPI_Op_free(MPI_Op* op) { *op = 0; return MPI_SUCCESS; }
int MPI_Get_count(const MPI_Status* st, MPI_Datatype t, int* c) { *c = (st && extent(t) > 0) ? int(st->count_bytes / extent(t)) : 0; return MPI_SUCCESS; }

/* ---- point to point ---- */
int MPI_Send(const void* buf, int count, MPI_Datatype t, int dest, int tag, MPI_Comm c) {
    CommInfo& ci = comm(c);
    raw_send(ci.ranks[size_t(dest)], ci.ctx, tag, buf, long(count) * extent(t));
    return MPI_SUCCESS;
}
int MPI_Recv(void* buf, int count, MPI_Datatype t, int src, int tag, MPI_Comm c, MPI_Status* st) {
    CommInfo& ci = comm(c);
    if (src == MPI_ANY_SOURCE) die("MPI_ANY_SOURCE is not supported by the stub");
    raw_recv(ci.ranks[size_t(src)], ci.ctx, tag, buf, long(count) * extent(t), st);
    if (st) st->MPI_SOURCE = src;
    return MPI_SUCCESS;
}
int MPI_Isend(const void* b, int n, MPI_Datatype t, int d, int tag, MPI_Comm c, MPI_Request* r) { *r = -1; return MPI_Send(b, n, t, d, tag, c); }
int MPI_Issend(const void* b, int n, MPI_Datatype t, int d, int tag, MPI_Comm c, MPI_Request* r) { *r = -1; return MPI_Send(b, n, t, d, tag, c); }
int MPI_Irsend(const void* b, int n, MPI_Datatype t, int d, int tag, MPI_Comm c, MPI_Request* r) { *r = -1; return MPI_Send(b, n, t, d, tag, c); }
int MPI_Irecv(void* b, int n, MPI_Datatype t, int src, int tag, MPI_Comm c, MPI_Request* r) {
    Pending p; p.live = true; p.recv = true; p.buf = b; p.cap = long(n) * extent(t); p.src = src; p.tag = tag; p.comm = c;
    g_pending.push_back(p);
    *r = int(g_pending.size()) - 1;
    return MPI_SUCCESS;
}
int MPI_Wait(MPI_Request* r, MPI_Status* st) {
    if (*r > 0 && *r < int(g_pending.size()) && g_pending[size_t(*r)].live) {
        Pending& p = g_pending[size_t(*r)];
        CommInfo& ci = comm(p.comm);
        raw_recv(ci.ranks[size_t(p.src)], ci.ctx, p.tag, p.buf, p.cap, st);
        if (st) st->MPI_SOURCE = p.src;
        p.live = false;
    }
    *r = MPI_REQUEST_NULL;
    return MPI_SUCCESS;
}
int MPI_Waitall(int n, MPI_Request* r, MPI_Status* st) { for (int i = 0; i < n; ++i) MPI_Wait(&r[i], st ? &st[i] : nullptr); return MPI_SUCCESS; }
int MPI_Test(MPI_Request* r, int* flag, MPI_Status* st) { MPI_Wait(r, st); *flag = 1; return MPI_SUCCESS; }
int MPI_Iprobe(int src, int tag, MPI_Comm c, int* flag, MPI_Status* st) {
    CommInfo& ci = comm(c);
    if (src == MPI_ANY_SOURCE) die("MPI_ANY_SOURCE is not supported by the stub");
    const int w = ci.ranks[size_t(src)];
    while (w != g_rank && ring_has_data(ring(w, g_rank))) pump(w);
    *flag = match_queued(w, ci.ctx, tag, nullptr, 0, st, false) ? 1 : 0;
    if (*flag && st) st->MPI_SOURCE = src;
    return MPI_SUCCESS;
}
int MPI_Sendrecv(const void* sb, int sc, MPI_Datatype st_, int dest, int stag, void* rb, int rc, MPI_Datatype rt, int src,
                 int rtag, MPI_Comm c, MPI_Status* st) {
    MPI_Send(sb, sc, st_, dest, stag, c);  // buffered by the ring (messages here are far below its capacity)
    return MPI_Recv(rb, rc, rt, src, rtag, c, st);
}
int MPI_Sendrecv_replace(void* b, int n, MPI_Datatype t, int dest, int stag, int src, int rtag, MPI_Comm c, MPI_Status* st) {
    std::vector<char> tmp(static_cast<char*>(b), static_cast<char*>(b) + long(n) * extent(t));
    MPI_Send(tmp.data(), n, t, dest, stag, c);
    return MPI_Recv(b, n, t, src, rtag, c, st);
}

/* ---- collectives (through member 0, reductions in rank order) ---- */
int MPI_Barrier(MPI_Comm c) {
    CommInfo& ci = comm(c);
    const int n = int(ci.ranks.size());
    char z = 0;
    if (ci.me == 0) {
        for (int i = 1; i < n; ++i) raw_recv(ci.ranks[size_t(i)], ci.ctx, T_BARRIER, &z, 1, nullptr);
        for (int i = 1; i < n; ++i) raw_send(ci.ranks[size_t(i)], ci.ctx, T_BARRIER, &z, 1);
    } else {
        raw_send(ci.ranks[0], ci.ctx, T_BARRIER, &z, 1);
        raw_recv(ci.ranks[0], ci.ctx, T_BARRIER, &z, 1, nullptr);
    }
    return MPI_SUCCESS;
}
int MPI_Bcast(void* buf, int count, MPI_Datatype t, int root, MPI_Comm c) {
    CommInfo& ci = comm(c);
    const long bytes = long(count) * extent(t);
    if (ci.me == root) { for (int i = 0; i < int(ci.ranks.size()); ++i) if (i != root) raw_send(ci.ranks[size_t(i)], ci.ctx, T_BCAST, buf, bytes); }
    else raw_recv(ci.ranks[size_t(root)], ci.ctx, T_BCAST, buf, bytes, nullptr);
    return MPI_SUCCESS;
}
int MPI_Ibcast(void* b, int n, MPI_Datatype t, int root, MPI_Comm c, MPI_Request* r) { *r = -1; return MPI_Bcast(b, n, t, root, c); }
int MPI_Gatherv(const void* s, int sc, MPI_Datatype st, void* r, const int* rc, const int* displs, MPI_Datatype rt, int root, MPI_Comm c) {
    CommInfo& ci = comm(c);
    if (ci.me != root) { raw_send(ci.ranks[size_t(root)], ci.ctx, T_GATHER, s, long(sc) * extent(st)); return MPI_SUCCESS; }
    for (int i = 0; i < int(ci.ranks.size()); ++i) {
        char* dst = static_cast<char*>(r) + long(displs[i]) * extent(rt);
        if (i == root) { if (s != MPI_IN_PLACE) memcpy(dst, s, size_t(long(sc) * extent(st))); }
        else raw_recv(ci.ranks[size_t(i)], ci.ctx, T_GATHER, dst, long(rc[i]) * extent(rt), nullptr);
    }
    return MPI_SUCCESS;
}
int MPI_Gather(const void* s, int sc, MPI_Datatype st, void* r, int rc, MPI_Datatype rt, int root, MPI_Comm c) {
    const int n = int(comm(c).ranks.size());
    std::vector<int> cnt(size_t(n), rc), dis(size_t(n), 0);
    for (int i = 0; i < n; ++i) dis[size_t(i)] = i * rc;
    return MPI_Gatherv(s, sc, st, r, cnt.data(), dis.data(), rt, root, c);
}
int MPI_Igather(const void* s, int sc, MPI_Datatype st, void* r, int rc, MPI_Datatype rt, int root, MPI_Comm c, MPI_Request* q) { *q = -1; return MPI_Gather(s, sc, st, r, rc, rt, root, c); }
int MPI_Allgatherv(const void* s, int sc, MPI_Datatype st, void* r, const int* rc, const int* displs, MPI_Datatype rt, MPI_Comm c) {
    CommInfo& ci = comm(c);
    const int n = int(ci.ranks.size());
    const void* src = s;
    if (s == MPI_IN_PLACE) { src = static_cast<char*>(r) + long(displs[ci.me]) * extent(rt); sc = rc[ci.me]; st = rt; }
    if (ci.me == 0) { if (src != static_cast<char*>(r) + long(displs[0]) * extent(rt)) memmove(static_cast<char*>(r) + long(displs[0]) * extent(rt), src, size_t(long(sc) * extent(st))); }
    MPI_Gatherv(ci.me == 0 ? MPI_IN_PLACE : src, sc, st, r, rc, displs, rt, 0, c);
    long hi = 0;
    for (int i = 0; i < n; ++i) hi = std::max(hi, long(displs[i]) + rc[i]);
    return MPI_Bcast(r, int(hi), rt, 0, c);
}
int MPI_Allgather(const void* s, int sc, MPI_Datatype st, void* r, int rc, MPI_Datatype rt, MPI_Comm c) {
    const int n = int(comm(c).ranks.size());
    std::vector<int> cnt(size_t(n), rc), dis(size_t(n), 0);
    for (int i = 0; i < n; ++i) dis[size_t(i)] = i * rc;
    return MPI_Allgatherv(s, sc, st, r, cnt.data(), dis.data(), rt, c);
}
int MPI_Scatterv(const void* s, const int* sc, const int* displs, MPI_Datatype st, void* r, int rc, MPI_Datatype rt, int root, MPI_Comm c) {
    CommInfo& ci = comm(c);
    if (ci.me != root) { raw_recv(ci.ranks[size_t(root)], ci.ctx, T_SCATTER, r, long(rc) * extent(rt), nullptr); return MPI_SUCCESS; }
    for (int i = 0; i < int(ci.ranks.size()); ++i) {
        const char* src = static_cast<const char*>(s) + long(displs[i]) * extent(st);
        if (i == root) { if (r != MPI_IN_PLACE) memcpy(r, src, size_t(long(sc[i]) * extent(st))); }
        else raw_send(ci.ranks[size_t(i)], ci.ctx, T_SCATTER, src, long(sc[i]) * extent(st));
    }
    return MPI_SUCCESS;
}
int MPI_Scatter(const void* s, int sc, MPI_Datatype st, void* r, int rc, MPI_Datatype rt, int root, MPI_Comm c) {
    const int n = int(comm(c).ranks.size());
    std::vector<int> cnt(size_t(n), sc), dis(size_t(n), 0);
    for (int i = 0; i < n; ++i) dis[size_t(i)] = i * sc;
    return MPI_Scatterv(s, cnt.data(), dis.data(), st, r, rc, rt, root, c);
}
int MPI_Alltoallv(const void* s, const int* sc, const int* sd, MPI_Datatype st, void* r, const int* rc, const int* rd, MPI_Datatype rt, MPI_Comm c) {
    CommInfo& ci = comm(c);
    const int n = int(ci.ranks.size());
    for (int i = 0; i < n; ++i)
        raw_send(ci.ranks[size_t(i)], ci.ctx, T_A2A, static_cast<const char*>(s) + long(sd[i]) * extent(st), long(sc[i]) * extent(st));
    for (int i = 0; i < n; ++i)
        raw_recv(ci.ranks[size_t(i)], ci.ctx, T_A2A, static_cast<char*>(r) + long(rd[i]) * extent(rt), long(rc[i]) * extent(rt), nullptr);
    return MPI_SUCCESS;
}
int MPI_Alltoall(const void* s, int sc, MPI_Datatype st, void* r, int rc, MPI_Datatype rt, MPI_Comm c) {
    const int n = int(comm(c).ranks.size());
    std::vector<int> scnt(size_t(n), sc), sdis(size_t(n), 0), rcnt(size_t(n), rc), rdis(size_t(n), 0);
    for (int i = 0; i < n; ++i) { sdis[size_t(i)] = i * sc; rdis[size_t(i)] = i * rc; }
    return MPI_Alltoallv(s, scnt.data(), sdis.data(), st, r, rcnt.data(), rdis.data(), rt, c);
}
int MPI_Reduce(const void* s, void* r, int count, MPI_Datatype t, MPI_Op op, int root, MPI_Comm c) {
    CommInfo& ci = comm(c);
    const int n = int(ci.ranks.size());
    const long bytes = long(count) * extent(t);
    if (ci.me != root) { raw_send(ci.ranks[size_t(root)], ci.ctx, T_REDUCE, s, bytes); return MPI_SUCCESS; }
    // result = x_0 op x_1 op ... op x_{n-1}, folded from the highest rank down so that each step
    // is inout = in(lower rank) op inout, which is what MPI_User_function computes
    std::vector<std::vector<char>> parts(static_cast<size_t>(n));
    for (int i = 0; i < n; ++i) {
        parts[size_t(i)].resize(size_t(bytes));
        if (i == root) memcpy(parts[size_t(i)].data(), s == MPI_IN_PLACE ? r : s, size_t(bytes));
        else raw_recv(ci.ranks[size_t(i)], ci.ctx, T_REDUCE, parts[size_t(i)].data(), bytes, nullptr);
    }
    for (int i = n - 2; i >= 0; --i) apply_op(parts[size_t(i)].data(), parts[size_t(n - 1)].data(), count, t, op);
    memcpy(r, parts[size_t(n - 1)].data(), size_t(bytes));
    return MPI_SUCCESS;
}
int MPI_Allreduce(const void* s, void* r, int count, MPI_Datatype t, MPI_Op op, MPI_Comm c) {
    MPI_Reduce(s, r, count, t, op, 0, c);
    return MPI_Bcast(r, count, t, 0, c);
}
int MPI_Scan(const void* s, void* r, int count, MPI_Datatype t, MPI_Op op, MPI_Comm c) {
    CommInfo& ci = comm(c);
    const int n = int(ci.ranks.size());
    const long bytes = long(count) * extent(t);
    std::vector<char> all(size_t(bytes) * size_t(n));
    MPI_Allgather(s == MPI_IN_PLACE ? r : s, int(bytes), MPI_BYTE, all.data(), int(bytes), MPI_BYTE, c);
    std::vector<char> acc(all.begin() + long(ci.me) * bytes, all.begin() + long(ci.me + 1) * bytes);
    for (int i = ci.me - 1; i >= 0; --i) apply_op(all.data() + long(i) * bytes, acc.data(), count, t, op);
    memcpy(r, acc.data(), size_t(bytes));
    return MPI_SUCCESS;
}
int MPI_Reduce_scatter(const void* s, void* r, const int* rc, MPI_Datatype t, MPI_Op op, MPI_Comm c) {
    CommInfo& ci = comm(c);
    const int n = int(ci.ranks.size());
    long total = 0;
    std::vector<int> dis(static_cast<size_t>(n), 0);
    for (int i = 0; i < n; ++i) { dis[size_t(i)] = int(total); total += rc[i]; }
    std::vector<char> full(size_t(total * extent(t)));
    MPI_Reduce(s == MPI_IN_PLACE ? r : s, full.data(), int(total), t, op, 0, c);
    return MPI_Scatterv(full.data(), rc, dis.data(), t, r, rc[ci.me], t, 0, c);
}
int MPI_Reduce_scatter_block(const void* s, void* r, int rc, MPI_Datatype t, MPI_Op op, MPI_Comm c) {
    const int n = int(comm(c).ranks.size());
    std::vector<int> cnt(size_t(n), rc);
    return MPI_Reduce_scatter(s, r, cnt.data(), t, op, c);
}

}  // extern "C"
