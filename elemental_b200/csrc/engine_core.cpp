// Grid and DistMat<T>: construction, alignment, views, memory. See engine.hpp.
#include "engine.hpp"
#include "engine_dev.hpp"

#include <cstdlib>
#include <cstring>
#include <mutex>

namespace elb {

void throw_status(int status, const char* where) {
    if (status == ELB200_OK) return;
    throw EngineError(status, std::string(where) + ": " + elb200_status_string(status) + " [" +
                                  elb200_last_error() + "]");
}
void cuda_check(cudaError_t e, const char* where) {
    if (e != cudaSuccess)
        throw EngineError(ELB200_ERR_CUDA, std::string(where) + ": " + cudaGetErrorString(e));
}
void nccl_check(ncclResult_t r, const char* where) {
    if (r != ncclSuccess)
        throw EngineError(ELB200_ERR_NCCL, std::string(where) + ": " + ncclGetErrorString(r));
}

// ------------------------------------------------------------------ Half <-> float (round to nearest even)
uint16_t Half::from_float(float f) {
    uint32_t x;
    std::memcpy(&x, &f, 4);
    const uint32_t sign = (x >> 16) & 0x8000u;
    x &= 0x7FFFFFFFu;
    if (x >= 0x7F800000u) return uint16_t(sign | 0x7C00u | (x > 0x7F800000u ? 0x200u : 0u));  // inf / nan
    if (x >= 0x477FF000u) return uint16_t(sign | 0x7C00u);                                       // rounds to inf
    if (x < 0x33000001u) return uint16_t(sign);                                                  // rounds to zero
    int e = int(x >> 23) - 127;
    uint32_t m = (x & 0x7FFFFFu) | 0x800000u;
    int shift = e >= -14 ? 13 : 13 + (-14 - e);                                                  // subnormal halves shift more
    uint32_t h = m >> shift;
    const uint32_t rem = m & ((1u << shift) - 1u), halfway = 1u << (shift - 1);
    if (rem > halfway || (rem == halfway && (h & 1u))) ++h;
    if (e >= -14) h += uint32_t(e + 15 - 1) << 10;   // h carries the implicit bit: adding (exp - 1) << 10 lets a mantissa carry bump the exponent
    return uint16_t(sign | h);
}
float Half::to_float(uint16_t hb) {
    const uint32_t sign = uint32_t(hb & 0x8000u) << 16;
    uint32_t e = (hb >> 10) & 0x1Fu, m = hb & 0x3FFu, x;
    if (e == 0) {
        if (m == 0) x = sign;
        else {
            int sh = 0;
            while (!(m & 0x400u)) { m <<= 1; ++sh; }
            x = sign | uint32_t(127 - 15 - sh + 1) << 23 | (m & 0x3FFu) << 13;
        }
    } else if (e == 31) x = sign | 0x7F800000u | m << 13;
    else x = sign | (e + 112u) << 23 | m << 13;
    float f;
    std::memcpy(&f, &x, 4);
    return f;
}

// ------------------------------------------------------------------ blocksize stack
namespace {
std::mutex g_bs_mutex;
std::vector<int>& bs_stack() {
#ifdef ELB200_HOSTSIM  // tests/hostsim runs the ranks of a grid as threads of one process: one stack per rank
    static thread_local std::vector<int> s{128};
#else
    static std::vector<int> s{128};
#endif
    return s;
}
}  // namespace
int Grid::Blocksize() { std::lock_guard<std::mutex> l(g_bs_mutex); return bs_stack().back(); }
void Grid::SetBlocksize(int nb) { std::lock_guard<std::mutex> l(g_bs_mutex); bs_stack().back() = nb > 0 ? nb : 1; }
void Grid::PushBlocksize(int nb) { std::lock_guard<std::mutex> l(g_bs_mutex); bs_stack().push_back(nb > 0 ? nb : 1); }
void Grid::PopBlocksize() { std::lock_guard<std::mutex> l(g_bs_mutex); if (bs_stack().size() > 1) bs_stack().pop_back(); }

int64_t Grid::PanelWidth(int elem_bytes) {
    // ELB200_PANEL_K overrides. Default: 16 KiB of k per row, i.e. 4096 (fp32) / 2048 (fp64). Per
    // panel the local C is re-read and re-written once and the GEMM pays one launch ramp and one
    // partial last wave; measured on B200 inside El::Gemm: fp32 213 -> 221 TF/s going from 2048 to
    // 4096, fp64 28.0 -> 29.5 TF/s going from 1024 to 2048.
    static int64_t env = [] { const char* e = std::getenv("ELB200_PANEL_K"); return e ? std::atoll(e) : 0ll; }();
    int64_t w = env > 0 ? env : 16384 / (elem_bytes > 0 ? elem_bytes : 4);  // 16 KiB of k per row: 8192 half, 4096 fp32, 2048 fp64
    const int64_t nb = Blocksize();
    if (w < nb) w = nb;
    return (w / nb) * nb;  // whole algorithmic blocks
}

// ------------------------------------------------------------------ Grid
Grid::Grid(int height, int width, int vc, const ncclUniqueId* id, int dev) {
    if (height < 1 || width < 1 || vc < 0 || vc >= height * width)
        throw EngineError(ELB200_ERR_ARG, "Grid: bad shape/rank");
    shape.r = height; shape.c = width;
    vc_rank = vc;
    me = coord_of_vc(shape, vc);
    device = dev;
    if (dev >= 0) {
        cuda_check(cudaSetDevice(dev), "cudaSetDevice");
        cuda_check(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking), "cudaStreamCreate");
        cuda_check(cudaStreamCreateWithFlags(&comm_stream, cudaStreamNonBlocking), "cudaStreamCreate");
        cuda_check(cudaStreamCreateWithFlags(&h2d_stream, cudaStreamNonBlocking), "cudaStreamCreate");
        cuda_check(cudaStreamCreateWithFlags(&d2h_stream, cudaStreamNonBlocking), "cudaStreamCreate");
        // Every temporary (panels, staging, proxies) comes from the device's stream-ordered pool.
        // Its default release threshold is 0, i.e. all cached memory goes back to the driver at the
        // next synchronisation and a Gemm with multi-GiB temporaries pays the mapping cost again
        // on every call; keep it cached like the reference's CUB pool does (src/core/imports/cub.cpp).
        // ELB200_POOL_KEEP_MB bounds what is kept (default: unbounded).
        cudaMemPool_t pool = nullptr;
        if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess && pool) {
            const char* e = std::getenv("ELB200_POOL_KEEP_MB");
            uint64_t keep = e && std::atoll(e) >= 0 ? uint64_t(std::atoll(e)) << 20 : UINT64_MAX;
            cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        }
    }
    if (id && shape.size() > 1) {
        nccl_check(ncclCommInitRank(&world, shape.size(), *id, vc), "ncclCommInitRank");
        // MC communicator: ranks of my grid COLUMN ordered by mc; MR: my grid ROW ordered by mr
        // (reference Grid.cpp:134-141: CartSub of the column-major cartesian communicator).
        if (shape.r > 1 && shape.c > 1) {
            nccl_check(ncclCommSplit(world, me.mr, me.mc, &mc_comm, nullptr), "ncclCommSplit(MC)");
            nccl_check(ncclCommSplit(world, me.mc, me.mr, &mr_comm, nullptr), "ncclCommSplit(MR)");
        } else if (shape.c == 1) {
            mc_comm = world;  // one grid column: MC == world
        } else {
            mr_comm = world;  // one grid row: MR == world
        }
        if (dev >= 0) arena = PeerArena::Create(world, vc, shape.size(), stream);
    } else if (shape.size() > 1 && dev >= 0) {
        throw EngineError(ELB200_ERR_STATE, "Grid: p > 1 requires an NCCL unique id");
    }
}

void Grid::NcclBegin(cudaStream_t s) {
    if (nccl_recorded) cuda_check(cudaStreamWaitEvent(s, nccl_done, 0), "cudaStreamWaitEvent(nccl order)");
}
void Grid::NcclEnd(cudaStream_t s) {
    if (!nccl_done) cuda_check(cudaEventCreateWithFlags(&nccl_done, cudaEventDisableTiming), "cudaEventCreate(nccl order)");
    cuda_check(cudaEventRecord(nccl_done, s), "cudaEventRecord(nccl order)");
    nccl_recorded = true;
}

std::vector<int> Grid::comm_members(Dist d) const {
    std::vector<int> v;
    if (d == MC) for (int i = 0; i < shape.r; ++i) v.push_back(vc_of(shape, GridCoord{i, me.mr}));
    else if (d == MR) for (int j = 0; j < shape.c; ++j) v.push_back(vc_of(shape, GridCoord{me.mc, j}));
    else for (int q = 0; q < shape.size(); ++q) v.push_back(q);
    return v;
}

void Grid::Shutdown() {
    if (stream) cudaStreamSynchronize(stream);
    if (comm_stream) cudaStreamSynchronize(comm_stream);
    if (arena) {
        // grid destruction is collective (like the reference's Grid::~Grid freeing its communicators):
        // unmap the peers, wait until everybody has, then release the local arena
        arena->Detach();
        int* token = nullptr;
        if (world && cudaMalloc(&token, sizeof(int)) == cudaSuccess) {
            cudaMemsetAsync(token, 0, sizeof(int), stream);
            if (ncclAllReduce(token, token, 1, ncclInt, ncclSum, world, stream) == ncclSuccess) cudaStreamSynchronize(stream);
            cudaFree(token);
        }
        cudaGetLastError();
        delete arena;
        arena = nullptr;
    }
    if (vr_comm) ncclCommDestroy(vr_comm);
    vr_comm = nullptr;
    if (mc_comm && mc_comm != world) ncclCommDestroy(mc_comm);
    if (mr_comm && mr_comm != world) ncclCommDestroy(mr_comm);
    if (world) ncclCommDestroy(world);
    mc_comm = mr_comm = world = nullptr;
}

Grid::~Grid() {
    Shutdown();
    if (stream) cudaStreamDestroy(stream);
    if (comm_stream) cudaStreamDestroy(comm_stream);
    if (h2d_stream) cudaStreamDestroy(h2d_stream);
    if (d2h_stream) cudaStreamDestroy(d2h_stream);
    if (nccl_done) cudaEventDestroy(nccl_done);
}

// ------------------------------------------------------------------ DistMat
template <typename T>
DistMat<T>::DistMat(Grid& g, Dist col, Dist row, cudaStream_t s) : grid(&g), stream(s ? s : g.stream) {
    if (!valid_pair(col, row)) throw EngineError(ELB200_ERR_ARG, "DistMat: unsupported distribution pair");
    spec.col = col; spec.row = row;
}
template <typename T>
DistMat<T>::~DistMat() {
    if (owner && buf) cudaFreeAsync(buf, stream);
}
template <typename T>
DistMat<T>::DistMat(DistMat&& o) noexcept
    : grid(o.grid), spec(o.spec), col_constrained(o.col_constrained), row_constrained(o.row_constrained),
      buf(o.buf), ldim(o.ldim), owner(o.owner), capacity(o.capacity), stream(o.stream) {
    o.buf = nullptr; o.owner = false; o.capacity = 0;
}

template <typename T>
void DistMat<T>::AlignCols(int64_t a, bool constrain) {
    spec.col_align = Mod(a, col_stride());
    if (constrain) col_constrained = true;
}
template <typename T>
void DistMat<T>::AlignRows(int64_t a, bool constrain) {
    spec.row_align = Mod(a, row_stride());
    if (constrain) row_constrained = true;
}
template <typename T>
void DistMat<T>::Align(int64_t ca, int64_t ra, bool constrain) { AlignCols(ca, constrain); AlignRows(ra, constrain); }

// alignment of distribution `mine` implied by `theirs` aligned at `a` (exact or partial match)
bool implied_align(Dist mine, Dist theirs, int64_t a, const GridShape& g, int64_t* out) {
    if (mine == STAR || theirs == STAR) return false;
    if (mine == theirs) { *out = a; return true; }
    if (mine == VC && theirs == MC) { *out = a; return true; }
    if (mine == VR && theirs == MR) { *out = a; return true; }
    if (mine == MC && theirs == VC) { *out = a % g.r; return true; }
    if (mine == MR && theirs == VR) { *out = a % g.c; return true; }
    return false;
}

DistSpec ResolveCopySpec(const GridShape& g, const DistSpec& A, DistSpec B, bool col_constrained,
                         bool row_constrained) {
    int64_t a;
    B.height = A.height; B.width = A.width;
    if (!col_constrained) {
        if (implied_align(B.col, A.col, A.col_align, g, &a) || implied_align(B.col, A.row, A.row_align, g, &a))
            B.col_align = Mod(a, dist_stride(g, B.col));
        else B.col_align = 0;
    }
    if (!row_constrained) {
        if (implied_align(B.row, A.row, A.row_align, g, &a) || implied_align(B.row, A.col, A.col_align, g, &a))
            B.row_align = Mod(a, dist_stride(g, B.row));
        else B.row_align = 0;
    }
    return B;
}

template <typename T>
void DistMat<T>::AlignWith(const DistSpec& o, bool constrain) {
    int64_t a;
    if (implied_align(spec.col, o.col, o.col_align, grid->shape, &a) ||
        implied_align(spec.col, o.row, o.row_align, grid->shape, &a))
        AlignCols(a, constrain);
    if (implied_align(spec.row, o.row, o.row_align, grid->shape, &a) ||
        implied_align(spec.row, o.col, o.col_align, grid->shape, &a))
        AlignRows(a, constrain);
}

template <typename T>
void DistMat<T>::Resize(int64_t h, int64_t w) {
    if (h < 0 || w < 0) throw EngineError(ELB200_ERR_ARG, "Resize: negative dimension");
    if (!owner) {
        if (h != spec.height || w != spec.width) throw EngineError(ELB200_ERR_ARG, "Resize: cannot resize a view");
        return;
    }
    spec.height = h; spec.width = w;
    const int64_t vec = 16 / int64_t(sizeof(T));
    const int64_t nlh = lh(), nlw = lw();
    const int64_t nld = nlh > 0 ? (nlh + vec - 1) / vec * vec : vec;  // 16-byte column pitch (TMA, 128-bit access)
    const int64_t need = nld * (nlw > 0 ? nlw : 1);
    if (need > capacity) {
        if (buf) cuda_check(cudaFreeAsync(buf, stream), "cudaFreeAsync");
        buf = nullptr;
        cuda_check(cudaMallocAsync(reinterpret_cast<void**>(&buf), size_t(need) * sizeof(T), stream), "cudaMallocAsync");
        capacity = need;
    }
    ldim = nld;
}

template <typename T>
void DistMat<T>::SetRoot(int root) {
    if (root < 0 || root >= grid->size()) throw EngineError(ELB200_ERR_ARG, "SetRoot: root outside the grid");
    if (spec.col != CIRC) {
        if (root != 0) throw EngineError(ELB200_ERR_ARG, "SetRoot: only [CIRC,CIRC] matrices have a root");
        return;
    }
    if (root == spec.root) return;
    if (!owner) throw EngineError(ELB200_ERR_ARG, "SetRoot: cannot move the root of a view");
    spec.root = root;
    Resize(spec.height, spec.width);  // local extent changes from everything to nothing or back
}

template <typename T>
void DistMat<T>::WriteLocal(const T* host, int64_t ld) {
    const int64_t h = lh(), w = lw();
    if (h == 0 || w == 0) return;
    if (!host || ld < h) throw EngineError(ELB200_ERR_ARG, "WriteLocal: null buffer or ld < local height");
    cuda_check(cudaMemcpy2DAsync(buf, size_t(ldim) * sizeof(T), host, size_t(ld) * sizeof(T), size_t(h) * sizeof(T), size_t(w),
                                 cudaMemcpyHostToDevice, stream), "cudaMemcpy2DAsync(H2D)");
}
template <typename T>
void DistMat<T>::ReadLocal(T* host, int64_t ld) const {
    const int64_t h = lh(), w = lw();
    if (h == 0 || w == 0) return;
    if (!host || ld < h) throw EngineError(ELB200_ERR_ARG, "ReadLocal: null buffer or ld < local height");
    cuda_check(cudaMemcpy2DAsync(host, size_t(ld) * sizeof(T), buf, size_t(ldim) * sizeof(T), size_t(h) * sizeof(T), size_t(w),
                                 cudaMemcpyDeviceToHost, stream), "cudaMemcpy2DAsync(D2H)");
}

template <typename T>
void DistMat<T>::Attach(int64_t h, int64_t w, int64_t ca, int64_t ra, T* buffer, int64_t ld) {
    if (owner && buf) cuda_check(cudaFreeAsync(buf, stream), "cudaFreeAsync");
    owner = false; capacity = 0;
    spec.height = h; spec.width = w;
    Align(ca, ra, true);
    buf = buffer;
    ldim = ld > 0 ? ld : 1;
    if (ld < (lh() > 1 ? lh() : 1)) throw EngineError(ELB200_ERR_ARG, "Attach: ldim < local height");
}

template <typename T>
DistMat<T> DistMat<T>::View(int64_t i0, int64_t i1, int64_t j0, int64_t j1) const {
    if (i0 < 0 || j0 < 0 || i1 > spec.height || j1 > spec.width || i0 > i1 || j0 > j1)
        throw EngineError(ELB200_ERR_ARG, "View: range out of bounds");
    DistMat<T> v(*grid, spec.col, spec.row, stream);
    v.owner = false;
    v.spec.height = i1 - i0; v.spec.width = j1 - j0;
    v.spec.col_align = Mod(spec.col_align + i0, col_stride());
    v.spec.row_align = Mod(spec.row_align + j0, row_stride());
    v.spec.root = spec.root;
    v.col_constrained = v.row_constrained = true;
    const int64_t lr0 = Length(i0, col_shift(), col_stride());
    const int64_t lc0 = Length(j0, row_shift(), row_stride());
    v.buf = buf && owns_data(grid->shape, spec, grid->me) ? buf + lr0 + lc0 * ldim : buf;
    v.ldim = ldim;
    return v;
}

template class DistMat<float>;
template class DistMat<double>;
template class DistMat<Half>;

}  // namespace elb
