// el_comm_check — the communicator side of the facade, written the way the reference's callers write it:
//     El::Grid g(El::mpi::COMM_WORLD, r);   g.MCComm() / MRComm() / VCComm() / VRComm();
//     mpi::AllGather / ReduceScatter / AllReduce / AllToAll / SendRecv / Broadcast on device buffers;
//     DistMatrixReadProxy / ReadWriteProxy / WriteProxy;  MakeMultiSync;  SyncInfoPool
// (include/El/core/Grid.hpp:15-75, src/core/imports/mpi/*.hpp, include/El/core/Proxy.hpp:148-760,
//  include/hydrogen/MultiSync.hpp, src/blas_like/level3/SyncInfoPool.hpp).
// One process per GPU (RANK / WORLD_SIZE / LOCAL_RANK / MASTER_ADDR / MASTER_PORT, e.g. under tools/elrun);
// a single process runs it on a 1 x 1 grid. tests/hostsim/facade_ranks.cpp runs the same function with the
// ranks of a grid as threads. Every collective result is compared with its definition on the host.
#include <El.hpp>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

using namespace El;

namespace comm_check {

#define NOTE_FAIL() (std::fprintf(stderr, "el_comm_check: check failed at line %d\n", __LINE__), 1)
#define CC(x) do { if ((x) != cudaSuccess) throw std::runtime_error("CUDA call failed: " #x); } while (0)

template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t n;
    explicit DevBuf(size_t count) : n(count) { CC(cudaMalloc(reinterpret_cast<void**>(&p), (count ? count : 1) * sizeof(T))); }
    ~DevBuf() { cudaFree(p); }
    void put(const std::vector<T>& h, cudaStream_t s) { CC(cudaMemcpyAsync(p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, s)); CC(cudaStreamSynchronize(s)); }
    std::vector<T> get(cudaStream_t s) const {
        std::vector<T> h(n);
        CC(cudaMemcpyAsync(h.data(), p, n * sizeof(T), cudaMemcpyDeviceToHost, s));
        CC(cudaStreamSynchronize(s));
        return h;
    }
};

// value rank r contributes at position i of a message: exact in float for the sizes used here
template <typename T> T val(int rank, size_t i) { return T(rank * 1000 + int(i % 997)); }

template <typename T>
int collectives(const mpi::Comm& comm, const SyncInfo<Device::GPU>& si, size_t count) {
    int bad = 0;
    const int p = comm.Size(), me = comm.Rank();
    cudaStream_t s = si.Stream();
    // AllGather
    {
        DevBuf<T> sb(count), rb(count * p);
        std::vector<T> h(count);
        for (size_t i = 0; i < count; ++i) h[i] = val<T>(me, i);
        sb.put(h, s);
        mpi::AllGather(sb.p, int(count), rb.p, int(count), comm, si);
        const std::vector<T> got = rb.get(s);
        for (int q = 0; q < p; ++q)
            for (size_t i = 0; i < count; ++i)
                if (got[q * count + i] != val<T>(q, i)) { bad += NOTE_FAIL(); q = p; break; }
    }
    // ReduceScatter (out of place, then in place with the result at the start of the buffer) and AllReduce
    {
        DevBuf<T> sb(count * p), rb(count);
        std::vector<T> h(count * p);
        for (size_t i = 0; i < count * p; ++i) h[i] = val<T>(me, i);
        sb.put(h, s);
        mpi::ReduceScatter(sb.p, rb.p, int(count), comm, si);
        std::vector<T> want(count);
        for (size_t i = 0; i < count; ++i) {
            T acc = 0;
            for (int q = 0; q < p; ++q) acc += val<T>(q, size_t(me) * count + i);
            want[i] = acc;
        }
        std::vector<T> got = rb.get(s);
        for (size_t i = 0; i < count; ++i) if (got[i] != want[i]) { bad += NOTE_FAIL(); break; }
        mpi::ReduceScatter(sb.p, int(count), comm, si);
        got = sb.get(s);
        for (size_t i = 0; i < count; ++i) if (got[i] != want[i]) { bad += NOTE_FAIL(); break; }
        sb.put(h, s);
        DevBuf<T> all(count * p);
        mpi::AllReduce(sb.p, all.p, int(count * p), mpi::SUM, comm, si);
        got = all.get(s);
        for (size_t i = 0; i < count * p; ++i) {
            T acc = 0;
            for (int q = 0; q < p; ++q) acc += val<T>(q, i);
            if (got[i] != acc) { bad += NOTE_FAIL(); break; }
        }
    }
    // AllToAll: block j of rank q's buffer holds val(q*64 + j, i)
    {
        DevBuf<T> sb(count * p), rb(count * p);
        std::vector<T> h(count * p);
        for (int j = 0; j < p; ++j)
            for (size_t i = 0; i < count; ++i) h[j * count + i] = val<T>(me * 64 + j, i);
        sb.put(h, s);
        mpi::AllToAll(sb.p, int(count), rb.p, int(count), comm, si);
        const std::vector<T> got = rb.get(s);
        for (int q = 0; q < p; ++q)
            for (size_t i = 0; i < count; ++i)
                if (got[q * count + i] != val<T>(q * 64 + me, i)) { bad += NOTE_FAIL(); q = p; break; }
    }
    // SendRecv around the ring, Broadcast from the last rank
    {
        DevBuf<T> sb(count), rb(count);
        std::vector<T> h(count);
        for (size_t i = 0; i < count; ++i) h[i] = val<T>(me, i);
        sb.put(h, s);
        const int to = (me + 1) % p, from = (me + p - 1) % p;
        mpi::SendRecv(sb.p, int(count), to, rb.p, int(count), from, comm, si);
        std::vector<T> got = rb.get(s);
        for (size_t i = 0; i < count; ++i) if (got[i] != val<T>(from, i)) { bad += NOTE_FAIL(); break; }
        mpi::Broadcast(sb.p, int(count), p - 1, comm, si);
        got = sb.get(s);
        for (size_t i = 0; i < count; ++i) if (got[i] != val<T>(p - 1, i)) { bad += NOTE_FAIL(); break; }
    }
    return bad;
}

template <typename T>
double max_abs_diff(const AbstractDistMatrix<T>& X, const AbstractDistMatrix<T>& Y) {
    DistMatrix<T, STAR, STAR> a(X), b(Y);
    std::vector<T> ha(size_t(a.LocalHeight()) * a.LocalWidth()), hb(ha.size());
    a.GetLocalToHost(ha.data());
    b.GetLocalToHost(hb.data());
    double d = 0;
    for (size_t i = 0; i < ha.size(); ++i) d = std::max(d, std::fabs(double(ha[i]) - double(hb[i])));
    return d;
}

// a routine written like the reference's SUMMA bodies: operands arrive as AbstractDistMatrix, proxies hand them over as [MC,MR]
template <typename T>
void ScaledGemmThroughProxies(T alpha, const AbstractDistMatrix<T>& APre, const AbstractDistMatrix<T>& BPre, AbstractDistMatrix<T>& CPre) {
    DistMatrixReadProxy<T, T, MC, MR, ELEMENT, Device::GPU> AProx(APre), BProx(BPre);
    DistMatrixReadWriteProxy<T, T, MC, MR, ELEMENT, Device::GPU> CProx(CPre);
    auto& A = AProx.GetLocked();
    auto& B = BProx.GetLocked();
    auto& C = CProx.Get();
    Gemm(NORMAL, NORMAL, alpha, A, B, T(1), C, GEMM_SUMMA_C);
}

template <typename T>
int exercise(const mpi::Comm& comm, int gridHeight) {
    int bad = 0;
    const int p = comm.Size(), me = comm.Rank();
    Grid g(comm, gridHeight);  // the reference's own constructor line (tests/blas_like/Gemm.cpp:258)
    const int r = g.Height(), c = g.Width();
    if (r != gridHeight || r * c != p || g.Rank() != me || g.VCRank() != me || g.Row() != me % r || g.Col() != me / r) bad += NOTE_FAIL();
    if (g.MCComm().Size() != r || g.MCComm().Rank() != g.Row() || g.MRComm().Size() != c || g.MRComm().Rank() != g.Col()) bad += NOTE_FAIL();
    if (g.VCComm().Size() != p || g.VCComm().Rank() != g.VCRank() || g.VRComm().Size() != p || g.VRComm().Rank() != g.VRRank()) bad += NOTE_FAIL();
    if (g.VRRank() != g.VCToVR(g.VCRank()) || g.VRToVC(g.VRRank()) != g.VCRank() || g.GCD() * g.LCM() != r * c) bad += NOTE_FAIL();
    if (&g.ColComm() != &g.MCComm() || &g.RowComm() != &g.MRComm() || &g.Comm() != &g.VCComm()) bad += NOTE_FAIL();

    SyncInfoPool<Device::GPU> pool(3);
    if (pool.Size() != 3) bad += NOTE_FAIL();
    const SyncInfo<Device::GPU>& s1 = pool.Next();
    const SyncInfo<Device::GPU>& s2 = pool.Next();
    if (&pool.Next() == &s2 || &pool.Next() != &s1) bad += NOTE_FAIL();  // circular: 1, 2, 0, 1
    pool.Reset();

    // collectives over every communicator of the grid, small (latency path) and a size that is not a multiple of anything
    for (const mpi::Comm* cm : {&g.MCComm(), &g.MRComm(), &g.VCComm(), &g.VRComm()})
        for (size_t count : {size_t(5), size_t(4099)}) bad += collectives<T>(*cm, s1, count);
    // Split / Dup / Free: halves of the world by parity, ranked in reverse
    {
        mpi::Comm half, dup;
        mpi::Split(comm, me % 2, p - me, half);
        const int expect_size = (p + (me % 2 == 0 ? 1 : 0)) / 2;
        if (half.Size() != expect_size) bad += NOTE_FAIL();
        bad += collectives<T>(half, s2, 33);
        mpi::Dup(half, dup);
        if (dup.Size() != half.Size() || dup.Rank() != half.Rank()) bad += NOTE_FAIL();
        mpi::Free(dup);
        mpi::Free(half);
        if (!half.IsNull()) bad += NOTE_FAIL();
    }

    // proxies
    const Int m = 70, n = 50, k = 90;
    DistMatrix<T> A(g), B(g), C(g), C2(g);
    Uniform(A, m, k, T(0), T(1), 11);
    Uniform(B, k, n, T(0), T(1), 12);
    Uniform(C, m, n, T(0), T(1), 13);
    C2 = C;
    {
        DistMatrixReadProxy<T, T, MC, MR> same(A);
        if (&same.GetLocked() != &A) bad += NOTE_FAIL();                                  // pass-through, no copy (Proxy.hpp:236-257)
        ElementalProxyCtrl ctrl;
        ctrl.colConstrain = true; ctrl.colAlign = (A.ColAlign() + 1) % A.ColStride();
        DistMatrixReadProxy<T, T, MC, MR> moved(A, ctrl);
        if (r > 1 && (&moved.GetLocked() == &A || moved.GetLocked().ColAlign() != ctrl.colAlign)) bad += NOTE_FAIL();
        if (max_abs_diff(moved.GetLocked(), A) != 0) bad += NOTE_FAIL();
        DistMatrixReadProxy<T, T, STAR, VR> other(A);
        if (other.GetLocked().ColDist() != STAR || other.GetLocked().RowDist() != VR || max_abs_diff(other.GetLocked(), A) != 0) bad += NOTE_FAIL();
    }
    {
        DistMatrix<T, VC, STAR> Cv(g);
        Cv = C;
        DistMatrix<T, STAR, VR> Av(g);
        Av = A;
        ScaledGemmThroughProxies(T(2), Av, B, Cv);     // A and C are not [MC,MR]: copies in, C written back
        Gemm(NORMAL, NORMAL, T(2), A, B, T(1), C2, GEMM_SUMMA_C);
        if (max_abs_diff(Cv, C2) > 1e-3) bad += NOTE_FAIL();
        DistMatrix<T, MR, MC> W(g);
        W.Resize(m, n);
        {
            DistMatrixWriteProxy<T, T, MC, MR> wp(W);  // not read, written back
            wp.Get() = C2;
        }
        if (max_abs_diff(W, C2) != 0) bad += NOTE_FAIL();
    }
    // MultiSync: operands on pool streams, product on C's; same result as on the grid stream
    {
        DistMatrix<T> A1(A), B1(B), C1(C);
        g.Synchronize();                                // the copies above ran on the grid's stream; SetSyncInfo does not order streams
        SetSyncInfo(A1, s1);
        SetSyncInfo(B1, s2);
        if (SyncInfoFromMatrix(A1).Stream() != s1.Stream() || A1.Matrix().GetSyncInfo().Stream() != s1.Stream()) bad += NOTE_FAIL();
        Scale(T(3), A1);                                // runs on s1
        Scale(T(0.5), B1);                              // runs on s2
        Gemm(NORMAL, NORMAL, T(1), A1, B1, T(0), C1);   // C1's stream waits for s1 and s2 (MakeMultiSync(C, A, B))
        DistMatrix<T> Cref(g);
        Gemm(NORMAL, NORMAL, T(1.5), A, B, Cref);
        if (max_abs_diff(C1, Cref) > 1e-3) bad += NOTE_FAIL();
        auto sync = MakeMultiSync(SyncInfoFromMatrix(C1), s1, s2);
        const SyncInfo<Device::GPU>& master = sync;
        if (master.Stream() != C1.Stream()) bad += NOTE_FAIL();
    }
    Synchronize(s1);
    Synchronize(s2);
    g.Synchronize();

    // ROW_MAJOR ordering of the owning ranks (Grid.cpp:143-148)
    {
        Grid gr(comm, gridHeight, ROW_MAJOR);
        if (gr.Order() != ROW_MAJOR || gr.Rank() != me || gr.Row() != me / c || gr.Col() != me % c || gr.VCRank() != gr.Row() + r * gr.Col()) bad += NOTE_FAIL();
        DistMatrix<T> X(gr);
        Uniform(X, 23, 17, T(0), T(1), 5);
        DistMatrix<T, STAR, STAR> Xs(X);
        if (Xs.LocalHeight() != 23 || std::fabs(double(Xs.Get(3, 4)) - double(X.Get(3, 4))) != 0) bad += NOTE_FAIL();
        gr.Synchronize();
    }
    return bad;
}

}  // namespace comm_check

int main(int argc, char** argv) {
    Environment env(argc, argv);
    int height = -1;
    for (int i = 1; i + 1 < argc; ++i)
        if (!std::strcmp(argv[i], "--gridHeight")) height = std::atoi(argv[i + 1]);
    try {
        const mpi::Comm& world = mpi::COMM_WORLD;
        const int p = mpi::Size(world);
        if (height <= 0) height = Grid::DefaultHeight(p);
        int bad = comm_check::exercise<float>(world, height) + comm_check::exercise<double>(world, height);
        if (mpi::Rank(world) == 0) {
            if (bad) std::printf("EL_COMM_CHECK_FAIL %d\n", bad);
            else std::printf("EL_COMM_CHECK_PASS %dx%d\n", height, p / height);
        }
        mpi::Barrier(world);
        return bad ? 1 : 0;
    } catch (const std::exception& e) {
        std::fprintf(stderr, "el_comm_check: %s\n", e.what());
        return 2;
    }
}
