// El.hpp — header-only C++ facade over libelb200's handle ABI (include/elb200.h) with the
// reference's names, for the El::Gemm path only.  Code written against Hydrogen's
//   El::Grid / El::DistMatrix<T,U,V> / El::Gemm / El::LocalGemm / El::Copy / El::Transpose /
//   El::Axpy / El::Scale / El::Zero / El::AxpyContract / El::Contract / El::SetBlocksize ...
// (reference include/El/core/Grid.hpp:15-140, include/El/core/DistMatrix/ElementMatrix.hpp,
//  include/El/blas_like/level3.hpp:20-90, include/El/blas_like/level1/decl.hpp) compiles against
// this header and runs on the B200 engine.  Everything is Device::GPU: there is no CPU path.
// Non-zero statuses become the reference's exception types: ELB200_ERR_ARG -> std::logic_error
// (El::LogicError), anything else -> std::runtime_error (El::RuntimeError).
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime_api.h>

#include <cmath>
#include <complex>
#include <cstdint>
#include <cstdlib>
#include <exception>
#include <initializer_list>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <utility>
#include <vector>

#include "elb200.h"

namespace El {

typedef int Int;  // reference default (no EL_USE_64BIT_INTS)
template <typename Real> using Complex = std::complex<Real>;  // El::Complex<Real> (include/El/core/Element/Complex/decl.hpp)
typedef __half gpu_half_type;  // hydrogen::gpu_half_type (include/hydrogen/utils/HalfPrecision.hpp): IEEE half on the device

// Values as in the reference (include/El/core/types.hpp:208-217, :464-469; level3.hpp:23-33).
enum Dist { MC, MD, MR, VC, VR, STAR, CIRC };
enum DistWrap { ELEMENT, BLOCK };
enum Orientation { NORMAL, TRANSPOSE, ADJOINT };
enum GridOrder { ROW_MAJOR, COLUMN_MAJOR };
enum UpperOrLower { LOWER, UPPER };   // include/El/core/types.hpp:511-515
enum LeftOrRight { LEFT, RIGHT };     // types.hpp:418-422
enum UnitOrNonUnit { NON_UNIT, UNIT };  // types.hpp:489-493
enum TrsmAlgorithm { TRSM_DEFAULT, TRSM_LARGE, TRSM_MEDIUM, TRSM_SMALL };  // level3.hpp:394-401
// hydrogen::Device (include/hydrogen/Device.hpp): only GPU exists here — there is no CPU path
enum class Device : unsigned char { CPU, GPU };
enum GemmAlgorithm {
    GEMM_DEFAULT, GEMM_SUMMA_A_MS, GEMM_SUMMA_A, GEMM_SUMMA_B_MS, GEMM_SUMMA_B, GEMM_SUMMA_C_MS, GEMM_SUMMA_C,
    GEMM_SUMMA_DOT, GEMM_CANNON
};

// El::Initialize / Finalize / Environment (src/core/environment.cpp:215-372). The reference brings up the GPU
// (device = local rank modulo device count, src/hydrogen/device/GPU.cpp:29-66), MPI and the blocksize stack; here the
// device is chosen the same way (LOCAL_RANK first — torchrun / tools/elrun — then the launcher variables the reference
// reads), and the world communicator bootstraps lazily on the first use of mpi::COMM_WORLD (see namespace mpi below).
inline void Initialize(int& /*argc*/, char**& /*argv*/) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n < 1)
        throw std::runtime_error("El::Initialize: no CUDA device (there is no CPU path)");
    int local = 0;
    for (const char* name : {"LOCAL_RANK", "FLUX_TASK_LOCAL_ID", "SLURM_LOCALID", "MV2_COMM_WORLD_LOCAL_RANK", "OMPI_COMM_WORLD_LOCAL_RANK"}) {
        const char* v = std::getenv(name);
        if (v && *v) { local = std::atoi(v); break; }
    }
    if (cudaSetDevice(local % n) != cudaSuccess) throw std::runtime_error("El::Initialize: cudaSetDevice failed");
}
inline void Initialize() { int c = 0; char** v = nullptr; Initialize(c, v); }
inline void Finalize() {
    // tears down the world communicator if one was bootstrapped (the reference: Al::Finalize + MPI_Finalize)
    if (std::getenv("WORLD_SIZE") && std::atoi(std::getenv("WORLD_SIZE")) > 1) {
        elb200_comm_t w = nullptr;
        if (elb200_comm_world(&w) == ELB200_OK && w) elb200_comm_free(w);
    }
}
struct Environment {
    Environment() { Initialize(); }
    Environment(int& argc, char**& argv) { Initialize(argc, argv); }
    ~Environment() { Finalize(); }
};

// SyncInfo<Device::GPU> (include/hydrogen/device/gpu/cuda/SyncInfo.hpp:14-84): a (stream, event)
// handle. Every device entry of libelb200 takes the stream explicitly; the event is for callers
// that order their own streams against it the way MultiSync does.
template <Device D> struct SyncInfo;
template <> struct SyncInfo<Device::GPU> {
    SyncInfo() = default;
    SyncInfo(cudaStream_t s, cudaEvent_t e) : stream_(s), event_(e) {}
    cudaStream_t Stream() const { return stream_; }
    cudaEvent_t Event() const { return event_; }
    cudaStream_t stream_ = nullptr;
    cudaEvent_t event_ = nullptr;
};
inline void Synchronize(const SyncInfo<Device::GPU>& si) {
    if (cudaStreamSynchronize(si.Stream()) != cudaSuccess) throw std::runtime_error("El::Synchronize failed");
}
inline void AddSynchronizationPoint(const SyncInfo<Device::GPU>& si) {
    if (si.Event() && cudaEventRecord(si.Event(), si.Stream()) != cudaSuccess) throw std::runtime_error("cudaEventRecord failed");
}
// `b` waits for everything enqueued on `a` so far (hydrogen/SyncInfoBase.hpp AddSynchronizationPoint(a, b))
inline void AddSynchronizationPoint(const SyncInfo<Device::GPU>& a, const SyncInfo<Device::GPU>& b) {
    if (!a.Event()) throw std::logic_error("AddSynchronizationPoint: the master SyncInfo has no event");
    if (cudaEventRecord(a.Event(), a.Stream()) != cudaSuccess || cudaStreamWaitEvent(b.Stream(), a.Event(), 0) != cudaSuccess)
        throw std::runtime_error("AddSynchronizationPoint failed");
}

namespace detail {
inline void Check(int st, const char* where) {
    if (st == ELB200_OK) return;
    std::string msg = std::string(where) + ": " + elb200_status_string(st);
    const char* d = elb200_last_error();
    if (d && *d) msg += std::string(" (") + d + ")";
    if (st == ELB200_ERR_ARG) throw std::logic_error(msg);
    throw std::runtime_error(msg);
}
inline int ToEngineDist(Dist d) {
    switch (d) {
        case MC: return 0;
        case MR: return 1;
        case VC: return 2;
        case VR: return 3;
        case STAR: return 4;
        case CIRC: return 5;
        default: throw std::logic_error("El::Dist MD is outside the Gemm path");
    }
}
inline Dist FromEngineDist(int d) {
    switch (d) {
        case 0: return MC;
        case 1: return MR;
        case 2: return VC;
        case 3: return VR;
        case 4: return STAR;
        default: return CIRC;
    }
}
template <typename T> struct TypeCode;
template <> struct TypeCode<float> { static constexpr int value = ELB200_F32; };
template <> struct TypeCode<double> { static constexpr int value = ELB200_F64; };
template <> struct TypeCode<__half> { static constexpr int value = ELB200_F16; };
template <typename T> struct IsElement { static constexpr bool value = std::is_same<T, float>::value || std::is_same<T, double>::value || std::is_same<T, __half>::value; };
// Complex<float> / Complex<double>: storable in a Matrix (interleaved, the reference's layout); the routines that take
// them are the complex section at the end of this header
template <typename T> struct IsComplex { static constexpr bool value = false; };
template <typename R> struct IsComplex<std::complex<R>> { static constexpr bool value = std::is_same<R, float>::value || std::is_same<R, double>::value; };
template <typename T> struct IsStorable { static constexpr bool value = IsElement<T>::value || IsComplex<T>::value; };
}  // namespace detail

// ---------------------------------------------------------------------------------------------
// hydrogen::MultiSync / MakeMultiSync (include/hydrogen/MultiSync.hpp:31-78): RAII fork/join of streams.
// Construction: the master (first) stream waits for everything enqueued on the others so far;
// destruction: the others wait for the master. Converts to the master SyncInfo.
// ---------------------------------------------------------------------------------------------
template <Device... Ds>
class MultiSync {
    static_assert(sizeof...(Ds) >= 1, "MultiSync needs at least the master SyncInfo");

public:
    MultiSync(const SyncInfo<Ds>&... syncInfos) : infos_{syncInfos...} {
        for (size_t i = 1; i < infos_.size(); ++i)
            if (infos_[i].Stream() != infos_[0].Stream()) Wait(infos_[i], infos_[0]);   // MasterWaitOnAll
    }
    ~MultiSync() {
        for (size_t i = 1; i < infos_.size(); ++i)
            if (infos_[i].Stream() != infos_[0].Stream()) {
                try { Wait(infos_[0], infos_[i]); } catch (...) {}                      // AllWaitOnMaster
            }
    }
    MultiSync(const MultiSync&) = delete;
    MultiSync& operator=(const MultiSync&) = delete;
    MultiSync(MultiSync&& o) noexcept : infos_(std::move(o.infos_)) { o.infos_.clear(); }
    operator const SyncInfo<Device::GPU>&() const noexcept { return infos_[0]; }

private:
    // `waiter` waits for everything enqueued on `on` so far; uses on's event when it has one
    static void Wait(const SyncInfo<Device::GPU>& on, const SyncInfo<Device::GPU>& waiter) {
        cudaEvent_t e = on.Event();
        const bool own = e == nullptr;
        if (own && cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) throw std::runtime_error("MultiSync: cudaEventCreate failed");
        const bool ok = cudaEventRecord(e, on.Stream()) == cudaSuccess && cudaStreamWaitEvent(waiter.Stream(), e, 0) == cudaSuccess;
        if (own) cudaEventDestroy(e);
        if (!ok) throw std::runtime_error("MultiSync: stream ordering failed");
    }
    std::vector<SyncInfo<Device::GPU>> infos_;
};
template <Device... Ds>
MultiSync<Ds...> MakeMultiSync(const SyncInfo<Ds>&... syncInfos) { return MultiSync<Ds...>(syncInfos...); }

// SyncInfoPool<Device::GPU> (src/blas_like/level3/SyncInfoPool.hpp:23-197): a circular pool of (stream, event) pairs
// the pool owns; Next() advances and wraps, EnsureSize only grows. The reference sizes its Gemm pool from
// H_STREAMPOOL_SIZE (src/blas_like/level3/Gemm.cpp:24); the engine here pipelines on its own two streams whatever
// that value is, so the pool is for callers that schedule their own work this way.
template <Device D> class SyncInfoPool;
template <>
class SyncInfoPool<Device::GPU> {
public:
    SyncInfoPool() noexcept = default;
    explicit SyncInfoPool(size_t pool_size) { EnsureSize(pool_size); }
    ~SyncInfoPool() {
        for (auto& si : pool_) {
            if (si.Event()) cudaEventDestroy(si.Event());
            if (si.Stream()) cudaStreamDestroy(si.Stream());
        }
    }
    SyncInfoPool(const SyncInfoPool&) = delete;
    SyncInfoPool& operator=(const SyncInfoPool&) = delete;
    SyncInfoPool(SyncInfoPool&& o) noexcept : pool_(std::move(o.pool_)), pos_(o.pos_) { o.pool_.clear(); o.pos_ = 0; }
    SyncInfoPool& operator=(SyncInfoPool&& o) noexcept { Swap(o); return *this; }
    size_t Size() const noexcept { return pool_.size(); }
    void EnsureSize(size_t pool_size) {
        while (pool_.size() < pool_size) {
            cudaStream_t s = nullptr;
            cudaEvent_t e = nullptr;
            if (cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) != cudaSuccess ||
                cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) {
                if (s) cudaStreamDestroy(s);
                throw std::runtime_error("SyncInfoPool: cannot create a stream / event");
            }
            pool_.emplace_back(s, e);
        }
    }
    void Swap(SyncInfoPool& other) noexcept { std::swap(pool_, other.pool_); std::swap(pos_, other.pos_); }
    void Reset() const noexcept { pos_ = 0; }
    const SyncInfo<Device::GPU>& Next() const {
        if (pool_.empty()) throw std::runtime_error("SyncInfoPool: Cannot call Next() on empty pool.");
        if (++pos_ == pool_.size()) pos_ = 0;
        return pool_[pos_];
    }

private:
    std::vector<SyncInfo<Device::GPU>> pool_;
    mutable size_t pos_ = 0;
};

// ---------------------------------------------------------------------------------------------
// El::mpi — the slice of the reference's MPI wrapper layer the Gemm path's callers touch
// (include/El/core/imports/mpi.hpp, mpi/comm.hpp, aluminum_comm.hpp). There is no MPI underneath:
// a Comm is a handle to libelb200's communicator (NCCL + the grid's NVLink arena, include/elb200.h),
// COMM_WORLD bootstraps from RANK / WORLD_SIZE / MASTER_ADDR (or ELB200_ID_FILE) on first use.
// Comm is a cheap value type like the reference's; Free() releases what Split / Dup created.
// ---------------------------------------------------------------------------------------------
namespace mpi {
class Comm {
public:
    Comm() = default;  // COMM_NULL
    explicit Comm(elb200_comm_t h) : h_(h) {}
    static Comm World() { Comm c; c.world_ = true; return c; }
    elb200_comm_t Handle() const {
        if (world_ && !h_) detail::Check(elb200_comm_world(&h_), "mpi::COMM_WORLD");
        return h_;
    }
    int Rank() const { const elb200_comm_t h = Handle(); return h ? elb200_comm_rank(h) : -1; }
    int Size() const { const elb200_comm_t h = Handle(); return h ? elb200_comm_size(h) : 0; }
    bool IsNull() const { return !world_ && !h_; }
    void Reset() { h_ = nullptr; world_ = false; }

private:
    mutable elb200_comm_t h_ = nullptr;
    bool world_ = false;
};
inline const Comm COMM_WORLD = Comm::World();
inline const Comm COMM_NULL = Comm();
inline int Rank(const Comm& comm = COMM_WORLD) { return comm.Rank(); }
inline int Size(const Comm& comm = COMM_WORLD) { return comm.Size(); }
inline void Barrier(const Comm& comm = COMM_WORLD) { detail::Check(elb200_comm_barrier(comm.Handle()), "mpi::Barrier"); }
inline void Split(const Comm& comm, int color, int key, Comm& newComm) {
    elb200_comm_t h = nullptr;
    detail::Check(elb200_comm_split(comm.Handle(), color, key, &h), "mpi::Split");
    newComm = Comm(h);
}
inline void Dup(const Comm& comm, Comm& newComm) {
    elb200_comm_t h = nullptr;
    detail::Check(elb200_comm_dup(comm.Handle(), &h), "mpi::Dup");
    newComm = Comm(h);
}
inline void Free(Comm& comm) {
    if (!comm.IsNull()) detail::Check(elb200_comm_free(comm.Handle()), "mpi::Free");
    comm.Reset();
}
enum Op { SUM };
// Device collectives with the reference's argument order (src/core/imports/mpi/{AllGather,ReduceScatter,AllReduce,
// AllToAll,SendRecv,Broadcast}.hpp): equal counts, device pointers, asynchronous on syncInfo's stream.
template <typename T>
void AllGather(const T* sbuf, int sc, T* rbuf, int rc, const Comm& comm, const SyncInfo<Device::GPU>& si) {
    if (sc != rc) throw std::logic_error("mpi::AllGather: send and receive counts differ");
    detail::Check(elb200_allgather(sbuf, rbuf, sc, detail::TypeCode<T>::value, comm.Handle(), si.Stream()), "mpi::AllGather");
}
template <typename T>
void ReduceScatter(const T* sbuf, T* rbuf, int rc, Op, const Comm& comm, const SyncInfo<Device::GPU>& si) {
    detail::Check(elb200_reduce_scatter(sbuf, rbuf, rc, detail::TypeCode<T>::value, comm.Handle(), si.Stream()), "mpi::ReduceScatter");
}
template <typename T>
void ReduceScatter(const T* sbuf, T* rbuf, int rc, const Comm& comm, const SyncInfo<Device::GPU>& si) { ReduceScatter(sbuf, rbuf, rc, SUM, comm, si); }
template <typename T>
void ReduceScatter(T* buf, int rc, Op, const Comm& comm, const SyncInfo<Device::GPU>& si) {  // in place, result at the start
    detail::Check(elb200_reduce_scatter(buf, buf, rc, detail::TypeCode<T>::value, comm.Handle(), si.Stream()), "mpi::ReduceScatter");
}
template <typename T>
void ReduceScatter(T* buf, int rc, const Comm& comm, const SyncInfo<Device::GPU>& si) { ReduceScatter(buf, rc, SUM, comm, si); }
template <typename T>
void AllReduce(const T* sbuf, T* rbuf, int count, Op, const Comm& comm, const SyncInfo<Device::GPU>& si) {
    detail::Check(elb200_allreduce(sbuf, rbuf, count, detail::TypeCode<T>::value, comm.Handle(), si.Stream()), "mpi::AllReduce");
}
template <typename T>
void AllReduce(T* buf, int count, Op op, const Comm& comm, const SyncInfo<Device::GPU>& si) { AllReduce(buf, buf, count, op, comm, si); }
template <typename T>
void AllToAll(const T* sbuf, int sc, T* rbuf, int rc, const Comm& comm, const SyncInfo<Device::GPU>& si) {
    if (sc != rc) throw std::logic_error("mpi::AllToAll: send and receive counts differ");
    detail::Check(elb200_alltoall(sbuf, rbuf, sc, detail::TypeCode<T>::value, comm.Handle(), si.Stream()), "mpi::AllToAll");
}
template <typename T>
void SendRecv(const T* sbuf, int sc, int to, T* rbuf, int rc, int from, const Comm& comm, const SyncInfo<Device::GPU>& si) {
    detail::Check(elb200_sendrecv(sbuf, sc, to, rbuf, rc, from, detail::TypeCode<T>::value, comm.Handle(), si.Stream()), "mpi::SendRecv");
}
template <typename T>
void Broadcast(T* buf, int count, int root, const Comm& comm, const SyncInfo<Device::GPU>& si) {
    detail::Check(elb200_bcast(buf, count, detail::TypeCode<T>::value, root, comm.Handle(), si.Stream()), "mpi::Broadcast");
}
}  // namespace mpi

// Half-open index range, El::IR(a,b) / El::ALL (include/El/core/Range.hpp in the reference).
struct Range {
    Int beg, end;
};
inline Range IR(Int a, Int b) { return Range{a, b}; }
static const Range ALL{0, -1};

// ---------------------------------------------------------------------------------------------
// El::Matrix<T, Device::GPU> (include/El/core/Matrix/decl.hpp:299-520): column-major device
// matrix, owner or view, carrying the stream its operations run on (SyncInfo).
// ---------------------------------------------------------------------------------------------
template <typename T, Device D = Device::GPU>
class Matrix {
    static_assert(D == Device::GPU, "elemental_b200 has no CPU path: Matrix<T, Device::GPU> only");
    static_assert(detail::IsStorable<T>::value, "float, double, gpu_half_type, Complex<float> and Complex<double> only");

public:
    Matrix() = default;
    Matrix(Int height, Int width, Int ldim = 0) { Resize(height, width, ldim); }
    ~Matrix() { Release(); }
    Matrix(const Matrix&) = delete;  // copies are explicit (El::Copy), like every GPU transfer here
    Matrix& operator=(const Matrix&) = delete;
    Matrix(Matrix&& o) noexcept { *this = std::move(o); }
    Matrix& operator=(Matrix&& o) noexcept {
        if (this != &o) {
            Release();
            buf_ = o.buf_; h_ = o.h_; w_ = o.w_; ld_ = o.ld_; own_ = o.own_; cap_ = o.cap_; sync_ = o.sync_;
            o.buf_ = nullptr; o.own_ = false; o.cap_ = 0;
        }
        return *this;
    }
    Int Height() const { return h_; }
    Int Width() const { return w_; }
    Int LDim() const { return ld_; }
    bool Viewing() const { return !own_ && buf_ != nullptr; }
    bool IsEmpty() const { return h_ == 0 || w_ == 0; }
    Device GetDevice() const { return Device::GPU; }
    T* Buffer() { return buf_; }
    T* Buffer(Int i, Int j) { return buf_ + size_t(i) + size_t(j) * size_t(ld_); }
    const T* LockedBuffer() const { return buf_; }
    const T* LockedBuffer(Int i, Int j) const { return buf_ + size_t(i) + size_t(j) * size_t(ld_); }
    void Resize(Int height, Int width, Int ldim = 0) {
        if (height < 0 || width < 0) throw std::logic_error("Matrix::Resize: negative dimension");
        constexpr Int q = Int(16 / sizeof(T));
        const Int ld = ldim > 0 ? ldim : (height > 0 ? (height + q - 1) / q * q : q);  // 16-byte column pitch
        if (ld < height) throw std::logic_error("Matrix::Resize: ldim < height");
        if (!own_ && buf_) {
            if (height != h_ || width != w_) throw std::logic_error("Matrix::Resize: cannot resize a view");
            return;
        }
        const size_t need = size_t(ld) * size_t(width > 0 ? width : 1);
        if (need > cap_) {
            Release();
            if (cudaMallocAsync(reinterpret_cast<void**>(&buf_), need * sizeof(T), sync_.Stream()) != cudaSuccess)
                throw std::runtime_error("Matrix::Resize: device allocation failed");
            cap_ = need; own_ = true;
        }
        h_ = height; w_ = width; ld_ = ld;
    }
    // Attach / LockedAttach: borrow device memory (Matrix/decl.hpp Attach)
    void Attach(Int height, Int width, T* buffer, Int ldim) {
        Release();
        buf_ = buffer; h_ = height; w_ = width; ld_ = ldim > 0 ? ldim : 1; own_ = false; cap_ = 0;
    }
    void LockedAttach(Int height, Int width, const T* buffer, Int ldim) { Attach(height, width, const_cast<T*>(buffer), ldim); }
    // A(IR(i0,i1), IR(j0,j1)): zero-copy view
    Matrix operator()(Range I, Range J) const;
    SyncInfo<Device::GPU> GetSyncInfo() const { return sync_; }
    void SetSyncInfo(const SyncInfo<Device::GPU>& si) { sync_ = si; }

private:
    void Release() {
        if (own_ && buf_) cudaFreeAsync(buf_, sync_.Stream());
        buf_ = nullptr; own_ = false; cap_ = 0;
    }
    T* buf_ = nullptr;
    Int h_ = 0, w_ = 0, ld_ = 1;
    bool own_ = false;
    size_t cap_ = 0;
    SyncInfo<Device::GPU> sync_;
};
template <typename T, Device D>
Matrix<T, D> Matrix<T, D>::operator()(Range I, Range J) const {
    const Int i1 = I.end < 0 ? h_ : I.end, j1 = J.end < 0 ? w_ : J.end;
    if (I.beg < 0 || J.beg < 0 || i1 > h_ || j1 > w_ || I.beg > i1 || J.beg > j1) throw std::logic_error("Matrix view out of bounds");
    Matrix v;
    v.Attach(i1 - I.beg, j1 - J.beg, const_cast<T*>(LockedBuffer(I.beg, J.beg)), ld_);
    v.SetSyncInfo(sync_);
    return v;
}
template <typename T>
SyncInfo<Device::GPU> SyncInfoFromMatrix(const Matrix<T, Device::GPU>& A) { return A.GetSyncInfo(); }
template <typename T>
void SetSyncInfo(Matrix<T, Device::GPU>& A, const SyncInfo<Device::GPU>& si) { A.SetSyncInfo(si); }

namespace detail {
template <typename T> struct LocalOps;
template <> struct LocalOps<float> {
    static int gemm(int ta, int tb, long m, long n, long k, float al, const float* A, long lda, const float* B, long ldb, float be, float* C, long ldc, void* s) { return elb200_gemm_f32(ta, tb, m, n, k, al, A, lda, B, ldb, be, C, ldc, s); }
    static int copy(long r, long c, const float* X, long xr, long xc, float* Y, long yr, long yc, void* s) { return elb200_copy2d_f32(r, c, X, xr, xc, Y, yr, yc, s); }
    static int transpose(long r, long c, const float* A, long lda, float* B, long ldb, void* s) { return elb200_transpose_f32(r, c, A, lda, B, ldb, s); }
    static int axpy(long r, long c, float al, const float* X, long xr, long xc, float* Y, long yr, long yc, void* s) { return elb200_axpy2d_f32(r, c, al, X, xr, xc, Y, yr, yc, s); }
    static int scale(long r, long c, float al, float* A, long lda, void* s) { return elb200_scale2d_f32(r, c, al, A, lda, s); }
    static int fill(long r, long c, float v, float* A, long lda, void* s) { return elb200_fill2d_f32(r, c, v, A, lda, s); }
};
template <> struct LocalOps<double> {
    static int gemm(int ta, int tb, long m, long n, long k, double al, const double* A, long lda, const double* B, long ldb, double be, double* C, long ldc, void* s) { return elb200_gemm_f64(ta, tb, m, n, k, al, A, lda, B, ldb, be, C, ldc, s); }
    static int copy(long r, long c, const double* X, long xr, long xc, double* Y, long yr, long yc, void* s) { return elb200_copy2d_f64(r, c, X, xr, xc, Y, yr, yc, s); }
    static int transpose(long r, long c, const double* A, long lda, double* B, long ldb, void* s) { return elb200_transpose_f64(r, c, A, lda, B, ldb, s); }
    static int axpy(long r, long c, double al, const double* X, long xr, long xc, double* Y, long yr, long yc, void* s) { return elb200_axpy2d_f64(r, c, al, X, xr, xc, Y, yr, yc, s); }
    static int scale(long r, long c, double al, double* A, long lda, void* s) { return elb200_scale2d_f64(r, c, al, A, lda, s); }
    static int fill(long r, long c, double v, double* A, long lda, void* s) { return elb200_fill2d_f64(r, c, v, A, lda, s); }
};
template <> struct LocalOps<__half> {
    static int gemm(int ta, int tb, long m, long n, long k, __half al, const __half* A, long lda, const __half* B, long ldb, __half be, __half* C, long ldc, void* s) { return elb200_gemm_f16(ta, tb, m, n, k, float(al), A, lda, B, ldb, float(be), C, ldc, s); }
    static int copy(long r, long c, const __half* X, long xr, long xc, __half* Y, long yr, long yc, void* s) { return elb200_copy2d_f16(r, c, X, xr, xc, Y, yr, yc, s); }
    static int transpose(long r, long c, const __half* A, long lda, __half* B, long ldb, void* s) { return elb200_transpose_f16(r, c, A, lda, B, ldb, s); }
    static int axpy(long r, long c, __half al, const __half* X, long xr, long xc, __half* Y, long yr, long yc, void* s) { return elb200_axpy2d_f16(r, c, float(al), X, xr, xc, Y, yr, yc, s); }
    static int scale(long r, long c, __half al, __half* A, long lda, void* s) { return elb200_scale2d_f16(r, c, float(al), A, lda, s); }
    static int fill(long r, long c, __half v, __half* A, long lda, void* s) { return elb200_fill2d_f16(r, c, float(v), A, lda, s); }
};
}  // namespace detail

// Local (sequential) BLAS on Matrix<T,GPU>: the reference's Gemm(Matrix) -> gpu_blas::Gemm
// (Gemm.cpp:141-271) and the level-1 routines around it, asynchronous on C's / B's stream.
template <typename T>
void Gemm(Orientation orientA, Orientation orientB, T alpha, const Matrix<T>& A, const Matrix<T>& B, T beta, Matrix<T>& C) {
    const Int m = orientA == NORMAL ? A.Height() : A.Width(), ka = orientA == NORMAL ? A.Width() : A.Height();
    const Int n = orientB == NORMAL ? B.Width() : B.Height(), kb = orientB == NORMAL ? B.Height() : B.Width();
    if (ka != kb || C.Height() != m || C.Width() != n) throw std::logic_error("Gemm: nonconformal matrices");  // Gemm.cpp:199-238
    detail::Check(detail::LocalOps<T>::gemm(orientA != NORMAL, orientB != NORMAL, m, n, ka, alpha, A.LockedBuffer(), A.LDim(),
                                            B.LockedBuffer(), B.LDim(), beta, C.Buffer(), C.LDim(), C.GetSyncInfo().Stream()), "Gemm");
}
template <typename T>
void Gemm(Orientation orientA, Orientation orientB, T alpha, const Matrix<T>& A, const Matrix<T>& B, Matrix<T>& C) {
    C.Resize(orientA == NORMAL ? A.Height() : A.Width(), orientB == NORMAL ? B.Width() : B.Height());
    Gemm(orientA, orientB, alpha, A, B, T(0), C);
}
template <typename T>
void Copy(const Matrix<T>& A, Matrix<T>& B) {
    B.Resize(A.Height(), A.Width());
    detail::Check(detail::LocalOps<T>::copy(A.Height(), A.Width(), A.LockedBuffer(), 1, A.LDim(), B.Buffer(), 1, B.LDim(),
                                            B.GetSyncInfo().Stream()), "Copy");
}
template <typename T>
void Transpose(const Matrix<T>& A, Matrix<T>& B) {
    B.Resize(A.Width(), A.Height());
    detail::Check(detail::LocalOps<T>::transpose(A.Height(), A.Width(), A.LockedBuffer(), A.LDim(), B.Buffer(), B.LDim(),
                                                 B.GetSyncInfo().Stream()), "Transpose");
}
template <typename T>
void Axpy(T alpha, const Matrix<T>& X, Matrix<T>& Y) {
    if (X.Height() != Y.Height() || X.Width() != Y.Width()) throw std::logic_error("Axpy: nonconformal matrices");
    detail::Check(detail::LocalOps<T>::axpy(X.Height(), X.Width(), alpha, X.LockedBuffer(), 1, X.LDim(), Y.Buffer(), 1, Y.LDim(),
                                            Y.GetSyncInfo().Stream()), "Axpy");
}
template <typename T>
void Scale(T alpha, Matrix<T>& A) {
    detail::Check(detail::LocalOps<T>::scale(A.Height(), A.Width(), alpha, A.Buffer(), A.LDim(), A.GetSyncInfo().Stream()), "Scale");
}
template <typename T>
void Fill(Matrix<T>& A, T value) {
    detail::Check(detail::LocalOps<T>::fill(A.Height(), A.Width(), value, A.Buffer(), A.LDim(), A.GetSyncInfo().Stream()), "Fill");
}
template <typename T>
void Zero(Matrix<T>& A) { Fill(A, T(0)); }

// ---------------------------------------------------------------------------------------------
// El::Grid (include/El/core/Grid.hpp:15-140, src/core/Grid.cpp:74-223) — r x c process grid, one process
// per GPU, built from a communicator exactly as in the reference:
//     El::Grid g(El::mpi::COMM_WORLD, r);        El::Grid g(comm, r, El::ROW_MAJOR);       El::Grid g;
// The grid's own NCCL communicators (VC world + MC / MR splits) come from an id broadcast over `comm`.
// ---------------------------------------------------------------------------------------------
class Grid {
public:
    typedef std::vector<unsigned char> Id;
    static Id UniqueId() {
        Id id(ELB200_UNIQUE_ID_BYTES);
        detail::Check(elb200_unique_id(id.data()), "Grid::UniqueId");
        return id;
    }
    static int DefaultHeight(int gridSize) { return elb200_default_grid_height(gridSize); }

    // Grid(): over mpi::COMM_WORLD with the default height (Grid.cpp:65-72); one process -> 1 x 1
    Grid() { CreateOn(mpi::COMM_WORLD, -1, COLUMN_MAJOR); }
    explicit Grid(mpi::Comm comm, GridOrder order = COLUMN_MAJOR) { CreateOn(comm, -1, order); }
    explicit Grid(mpi::Comm comm, int height, GridOrder order = COLUMN_MAJOR) {
        if (height <= 0) throw std::logic_error("Grid: height must be positive");
        CreateOn(comm, height, order);
    }
    // Without a communicator: the caller ships the NCCL id itself; this process is VC rank `rank` of `size`
    Grid(const Id& id, int rank, int size, int height = -1) {
        const int h = height > 0 ? height : DefaultHeight(size);
        if (size % h != 0) throw std::logic_error("Grid: height must divide the number of processes");
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess) throw std::runtime_error("Grid: no CUDA device (there is no CPU path)");
        detail::Check(elb200_grid_create(h, size / h, rank, id.data(), dev, &h_), "Grid");
        Finish(rank, COLUMN_MAJOR);
    }
    ~Grid() {
        for (mpi::Comm* c : {&mc_, &mr_, &vc_comm_, &vr_comm_})
            if (!c->IsNull()) elb200_comm_free(c->Handle());
        if (h_) elb200_grid_destroy(h_);
    }
    Grid(const Grid&) = delete;             // non-copyable like the reference (Grid.hpp:141-145)
    Grid& operator=(const Grid&) = delete;

    // simple interface
    int Row() const { return mc_rank_; }
    int Col() const { return mr_rank_; }
    int Height() const { return height_; }
    int Width() const { return width_; }
    int Size() const { return height_ * width_; }
    int Rank() const { return owning_rank_; }
    GridOrder Order() const { return order_; }
    const mpi::Comm& ColComm() const { return MCComm(); }
    const mpi::Comm& RowComm() const { return MRComm(); }
    const mpi::Comm& Comm() const { return order_ == COLUMN_MAJOR ? VCComm() : VRComm(); }
    // distribution-based interface
    int MCRank() const { return mc_rank_; }
    int MRRank() const { return mr_rank_; }
    int VCRank() const { return vc_; }
    int VRRank() const { return mr_rank_ + width_ * mc_rank_; }
    int MCSize() const { return height_; }
    int MRSize() const { return width_; }
    int VCSize() const { return Size(); }
    int VRSize() const { return Size(); }
    const mpi::Comm& MCComm() const { return Lazy(mc_, ELB200_COMM_MC); }
    const mpi::Comm& MRComm() const { return Lazy(mr_, ELB200_COMM_MR); }
    const mpi::Comm& VCComm() const { return Lazy(vc_comm_, ELB200_COMM_VC); }
    const mpi::Comm& VRComm() const { return Lazy(vr_comm_, ELB200_COMM_VR); }  // first call is collective
    const mpi::Comm& OwningComm() const { return Comm(); }
    const mpi::Comm& ViewingComm() const { return Comm(); }
    int GCD() const { int a = height_, b = width_; while (b) { const int t = a % b; a = b; b = t; } return a; }
    int LCM() const { return height_ * width_ / GCD(); }
    bool InGrid() const { return true; }
    bool HaveViewers() const { return false; }
    int OwningRank() const { return owning_rank_; }
    int ViewingRank() const { return owning_rank_; }
    int VCToVR(int vcRank) const { return (vcRank / height_) + width_ * (vcRank % height_); }  // Grid.cpp:289-294
    int VRToVC(int vrRank) const { return (vrRank / width_) + height_ * (vrRank % width_); }
    int VCToViewing(int vcRank) const { return order_ == COLUMN_MAJOR ? vcRank : VCToVR(vcRank); }

    cudaStream_t Stream() const { return static_cast<cudaStream_t>(elb200_grid_stream(h_)); }
    void Synchronize() const { detail::Check(elb200_grid_sync(h_), "Grid::Synchronize"); }
    elb200_grid_t Handle() const { return h_; }

private:
    void CreateOn(const mpi::Comm& comm, int height, GridOrder order) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess) throw std::runtime_error("Grid: no CUDA device (there is no CPU path)");
        detail::Check(elb200_grid_create_on(comm.Handle(), height, order == COLUMN_MAJOR ? ELB200_COLUMN_MAJOR : ELB200_ROW_MAJOR, dev, &h_),
                      "Grid");
        Finish(comm.Rank(), order);
    }
    void Finish(int owning_rank, GridOrder order) {
        detail::Check(elb200_grid_info(h_, &height_, &width_, &vc_, &mc_rank_, &mr_rank_), "Grid");
        owning_rank_ = owning_rank;
        order_ = order;
    }
    const mpi::Comm& Lazy(mpi::Comm& slot, int which) const {
        if (slot.IsNull()) {
            elb200_comm_t c = nullptr;
            detail::Check(elb200_grid_comm(h_, which, &c), "Grid communicator");
            slot = mpi::Comm(c);
        }
        return slot;
    }
    elb200_grid_t h_ = nullptr;
    int height_ = 1, width_ = 1, vc_ = 0, mc_rank_ = 0, mr_rank_ = 0, owning_rank_ = 0;
    GridOrder order_ = COLUMN_MAJOR;
    mutable mpi::Comm mc_, mr_, vc_comm_, vr_comm_;
};
inline bool operator==(const Grid& A, const Grid& B) { return &A == &B; }
inline bool operator!=(const Grid& A, const Grid& B) { return &A != &B; }

// El::Blocksize stack (src/blas_like/blocksizes.cpp:35-67)
inline Int Blocksize() { return elb200_blocksize(); }
inline void SetBlocksize(Int nb) { elb200_set_blocksize(nb); }
inline void PushBlocksizeStack(Int nb) { elb200_push_blocksize(nb); }
inline void PopBlocksizeStack() { elb200_pop_blocksize(); }

// ---------------------------------------------------------------------------------------------
// El::AbstractDistMatrix<T> / ElementalMatrix<T> / DistMatrix<T,U,V,ELEMENT,Device::GPU>
// (include/El/core/DistMatrix/{AbstractDistMatrix,ElementMatrix}.hpp, ElementMatrix/*.hpp).
// The engine's matrix handle carries its distribution at run time, so everything that does not need the
// (U,V) pair at compile time lives in the abstract base — which is also what the reference's routines take
// (El::Gemm(.., const AbstractDistMatrix<T>& A, ..), Copy, the proxies). Owns its local matrix unless it is a view.
// ---------------------------------------------------------------------------------------------
template <typename T>
class AbstractDistMatrix {
    static_assert(detail::IsElement<T>::value, "float, double and gpu_half_type only");

public:
    virtual ~AbstractDistMatrix() { if (h_) elb200_dm_destroy(h_); }
    AbstractDistMatrix(const AbstractDistMatrix&) = delete;
    AbstractDistMatrix& operator=(const AbstractDistMatrix&) = delete;

    void Align(Int colAlign, Int rowAlign, bool /*constrain*/ = true) { detail::Check(elb200_dm_align(h_, colAlign, rowAlign), "Align"); }
    void AlignCols(Int colAlign, bool /*constrain*/ = true) { detail::Check(elb200_dm_align_cols(h_, colAlign), "AlignCols"); }
    void AlignRows(Int rowAlign, bool /*constrain*/ = true) { detail::Check(elb200_dm_align_rows(h_, rowAlign), "AlignRows"); }
    void AlignWith(const AbstractDistMatrix<T>& A) { detail::Check(elb200_dm_align_with(h_, A.Handle()), "AlignWith"); }
    void Resize(Int height, Int width) { detail::Check(elb200_dm_resize(h_, height, width), "Resize"); }
    // [CIRC,CIRC] only: the VC rank that holds the matrix (AbstractDistMatrix.hpp:58,113)
    void SetRoot(int root) { detail::Check(elb200_dm_set_root(h_, root), "SetRoot"); }
    int Root() const { return Info().root; }
    // Attach(height, width, grid, colAlign, rowAlign, buffer, ldim): borrow device memory
    void Attach(Int height, Int width, Int colAlign, Int rowAlign, T* deviceBuffer, Int ldim) {
        detail::Check(elb200_dm_attach(h_, height, width, colAlign, rowAlign, deviceBuffer, ldim), "Attach");
    }

    const El::Grid& Grid() const { return *grid_; }
    Int Height() const { return Int(Info().height); }
    Int Width() const { return Int(Info().width); }
    Int LocalHeight() const { return Int(Info().local_height); }
    Int LocalWidth() const { return Int(Info().local_width); }
    Int LDim() const { return Int(Info().ldim); }
    Int ColAlign() const { return Int(Info().col_align); }
    Int RowAlign() const { return Int(Info().row_align); }
    Int ColShift() const { return Int(Info().col_shift); }
    Int RowShift() const { return Int(Info().row_shift); }
    Int ColStride() const { return Int(Info().col_stride); }
    Int RowStride() const { return Int(Info().row_stride); }
    Dist ColDist() const { return detail::FromEngineDist(Info().col_dist); }
    Dist RowDist() const { return detail::FromEngineDist(Info().row_dist); }
    bool Viewing() const { return Info().is_view != 0; }
    Int GlobalRow(Int iLoc) const { const auto d = Info(); return Int(d.col_shift + iLoc * d.col_stride); }
    Int GlobalCol(Int jLoc) const { const auto d = Info(); return Int(d.row_shift + jLoc * d.row_stride); }
    T* Buffer() { return static_cast<T*>(Info().buffer); }              // DEVICE pointer
    const T* LockedBuffer() const { return static_cast<const T*>(Info().buffer); }
    elb200_dm_t Handle() const { return h_; }
    Device GetLocalDevice() const { return Device::GPU; }
    DistWrap Wrap() const { return ELEMENT; }
    bool Participating() const { return true; }  // every process of the grid takes part (no viewing-comm subsets)
    // communicators of the distribution (ElementMatrix/*.cpp comm tables, e.g. MC_MR.cpp:170-200)
    const mpi::Comm& ColComm() const { return CommOf(ColDist()); }
    const mpi::Comm& RowComm() const { return CommOf(RowDist()); }
    const mpi::Comm& DistComm() const { return grid_->VCComm(); }
    // local <-> global index maps (include/El/core/DistMatrix/AbstractDistMatrix.hpp)
    bool IsLocalRow(Int i) const { const auto d = Info(); return d.local_height > 0 && (i - d.col_shift) % d.col_stride == 0 && i >= d.col_shift; }
    bool IsLocalCol(Int j) const { const auto d = Info(); return d.local_width > 0 && (j - d.row_shift) % d.row_stride == 0 && j >= d.row_shift; }
    bool IsLocal(Int i, Int j) const { return IsLocalRow(i) && IsLocalCol(j); }
    Int LocalRow(Int i) const { const auto d = Info(); return Int((i - d.col_shift) / d.col_stride); }
    Int LocalCol(Int j) const { const auto d = Info(); return Int((j - d.row_shift) / d.row_stride); }
    // The local matrix as a zero-copy El::Matrix view carrying this matrix's stream (Matrix() / LockedMatrix()).
    El::Matrix<T, Device::GPU> Matrix() {
        const auto d = Info();
        El::Matrix<T, Device::GPU> m;
        m.Attach(Int(d.local_height), Int(d.local_width), static_cast<T*>(d.buffer), Int(d.ldim));
        m.SetSyncInfo(GetSyncInfo());
        return m;
    }
    El::Matrix<T, Device::GPU> LockedMatrix() const { return const_cast<AbstractDistMatrix*>(this)->Matrix(); }
    // SyncInfoFromMatrix(A.LockedMatrix()) / SetSyncInfo(A.Matrix(), si): the stream operations writing this matrix run on
    SyncInfo<Device::GPU> GetSyncInfo() const { return SyncInfo<Device::GPU>(Stream(), nullptr); }
    void SetSyncInfo(const SyncInfo<Device::GPU>& si) { detail::Check(elb200_dm_set_stream(h_, si.Stream()), "SetSyncInfo"); }
    cudaStream_t Stream() const { return static_cast<cudaStream_t>(elb200_dm_stream(h_)); }

    // Single local entries through the host: synchronous, for tests and setup code only
    // (the reference's GetLocal / SetLocal on a GPU matrix are the same blocking copies).
    T GetLocal(Int iLoc, Int jLoc) const {
        const auto d = Info();
        if (iLoc < 0 || jLoc < 0 || iLoc >= d.local_height || jLoc >= d.local_width) throw std::logic_error("GetLocal: index out of range");
        T v;
        CudaCheck(cudaMemcpyAsync(&v, static_cast<const T*>(d.buffer) + size_t(iLoc) + size_t(jLoc) * size_t(d.ldim), sizeof(T),
                                  cudaMemcpyDeviceToHost, Stream()));
        CudaCheck(cudaStreamSynchronize(Stream()));
        return v;
    }
    void SetLocal(Int iLoc, Int jLoc, T v) {
        const auto d = Info();
        if (iLoc < 0 || jLoc < 0 || iLoc >= d.local_height || jLoc >= d.local_width) throw std::logic_error("SetLocal: index out of range");
        CudaCheck(cudaMemcpyAsync(static_cast<T*>(d.buffer) + size_t(iLoc) + size_t(jLoc) * size_t(d.ldim), &v, sizeof(T),
                                  cudaMemcpyHostToDevice, Stream()));
        CudaCheck(cudaStreamSynchronize(Stream()));
    }
    // Single GLOBAL entries (AbstractDistMatrix::Get / Set / Update, src/core/DistMatrix/ElementMatrix.cpp). Get is
    // collective: every process of the grid calls it and receives the entry (a one-entry view gathered to [*,*]);
    // Set / Update are carried out by the owning process(es) and are no-ops elsewhere, as in the reference.
    T Get(Int i, Int j) const;
    void Set(Int i, Int j, T v) {
        if (IsLocal(i, j)) SetLocal(LocalRow(i), LocalCol(j), v);
    }
    void Update(Int i, Int j, T v) {
        if (IsLocal(i, j)) SetLocal(LocalRow(i), LocalCol(j), GetLocal(LocalRow(i), LocalCol(j)) + v);
    }
    // Host <-> local device matrix, column-major with leading dimension LocalHeight() on the host
    // side (the reference's CPU<->GPU Matrix copies, include/El/core/Matrix/impl_gpu.hpp).
    void SetLocalFromHost(const T* host) {
        const auto d = Info();
        if (d.local_height == 0 || d.local_width == 0) return;
        CudaCheck(cudaMemcpy2DAsync(d.buffer, size_t(d.ldim) * sizeof(T), host, size_t(d.local_height) * sizeof(T),
                                    size_t(d.local_height) * sizeof(T), size_t(d.local_width), cudaMemcpyHostToDevice, Stream()));
        CudaCheck(cudaStreamSynchronize(Stream()));
    }
    void GetLocalToHost(T* host) const {
        const auto d = Info();
        if (d.local_height == 0 || d.local_width == 0) return;
        CudaCheck(cudaMemcpy2DAsync(host, size_t(d.local_height) * sizeof(T), d.buffer, size_t(d.ldim) * sizeof(T),
                                    size_t(d.local_height) * sizeof(T), size_t(d.local_width), cudaMemcpyDeviceToHost, Stream()));
        CudaCheck(cudaStreamSynchronize(Stream()));
    }

protected:
    AbstractDistMatrix(const El::Grid& g, Dist U, Dist V) : grid_(&g) {
        detail::Check(elb200_dm_create(g.Handle(), detail::TypeCode<T>::value, detail::ToEngineDist(U), detail::ToEngineDist(V), &h_),
                      "DistMatrix");
    }
    AbstractDistMatrix(const El::Grid& g, std::nullptr_t) : grid_(&g) {}
    void MoveFrom(AbstractDistMatrix& o) noexcept {
        if (this != &o) { if (h_) elb200_dm_destroy(h_); grid_ = o.grid_; h_ = o.h_; o.h_ = nullptr; }
    }
    elb200_dm_info_t Info() const {
        elb200_dm_info_t d;
        detail::Check(elb200_dm_info(h_, &d), "DistMatrix::Info");
        return d;
    }
    const mpi::Comm& CommOf(Dist d) const {
        switch (d) {
            case MC: return grid_->MCComm();
            case MR: return grid_->MRComm();
            case VC: return grid_->VCComm();
            case VR: return grid_->VRComm();
            default: return mpi::COMM_NULL;  // STAR / CIRC: one member (the reference hands out COMM_SELF)
        }
    }
    static void CudaCheck(cudaError_t e) { if (e != cudaSuccess) throw std::runtime_error(cudaGetErrorString(e)); }
    const El::Grid* grid_;
    elb200_dm_t h_ = nullptr;
};

// ElementalMatrix<T> (include/El/core/DistMatrix/ElementMatrix.hpp): the element-cyclic wrap — the only one on the Gemm
// path — so it adds nothing to the abstract base here; the name exists so caller code compiles unchanged.
template <typename T>
class ElementalMatrix : public AbstractDistMatrix<T> {
protected:
    using AbstractDistMatrix<T>::AbstractDistMatrix;
};

template <typename T, Dist U = MC, Dist V = MR, DistWrap W = ELEMENT, Device D = Device::GPU>
class DistMatrix : public ElementalMatrix<T> {
    static_assert(W == ELEMENT, "BLOCK-wrapped distributions are outside the Gemm path");
    static_assert(D == Device::GPU, "elemental_b200 has no CPU path: Device::GPU only");
    typedef AbstractDistMatrix<T> base;

public:
    explicit DistMatrix(const El::Grid& g) : ElementalMatrix<T>(g, U, V) {}
    DistMatrix(Int height, Int width, const El::Grid& g) : DistMatrix(g) { this->Resize(height, width); }
    DistMatrix(DistMatrix&& o) noexcept : ElementalMatrix<T>(*o.grid_, nullptr) { this->MoveFrom(o); }
    DistMatrix& operator=(DistMatrix&& o) noexcept { this->MoveFrom(o); return *this; }
    // B = A across distributions (the operator= tables of src/core/DistMatrix/ElementMatrix/*.cpp)
    DistMatrix& operator=(const AbstractDistMatrix<T>& A) {
        if (static_cast<const base*>(this) != &A) detail::Check(elb200_dm_copy(A.Handle(), this->h_), "DistMatrix::operator=");
        return *this;
    }
    DistMatrix& operator=(const DistMatrix& A) { return *this = static_cast<const base&>(A); }
    DistMatrix(const DistMatrix& A) : DistMatrix(A.Grid()) { *this = A; }
    explicit DistMatrix(const AbstractDistMatrix<T>& A) : DistMatrix(A.Grid()) { *this = A; }
    // A(IR(i0,i1), IR(j0,j1)): zero-copy view, alignments advance with the offset
    DistMatrix operator()(Range I, Range J) const {
        const elb200_dm_info_t d = this->Info();
        const Int i1 = I.end < 0 ? Int(d.height) : I.end, j1 = J.end < 0 ? Int(d.width) : J.end;
        DistMatrix v(*this->grid_, nullptr);
        detail::Check(elb200_dm_view(this->h_, I.beg, i1, J.beg, j1, &v.h_), "DistMatrix::operator()");
        return v;
    }
    Dist ColDist() const { return U; }
    Dist RowDist() const { return V; }

private:
    DistMatrix(const El::Grid& g, std::nullptr_t) : ElementalMatrix<T>(g, nullptr) {}
};

template <typename T>
T AbstractDistMatrix<T>::Get(Int i, Int j) const {
    if (i < 0 || j < 0 || i >= Height() || j >= Width()) throw std::logic_error("Get: index out of range");
    DistMatrix<T, STAR, STAR> one(*grid_);
    elb200_dm_t v = nullptr;
    detail::Check(elb200_dm_view(h_, i, i + 1, j, j + 1, &v), "Get");
    const int st = elb200_dm_copy(v, one.Handle());
    elb200_dm_destroy(v);
    detail::Check(st, "Get");
    return one.GetLocal(0, 0);
}
template <typename T>
SyncInfo<Device::GPU> SyncInfoFromMatrix(const AbstractDistMatrix<T>& A) { return A.GetSyncInfo(); }
template <typename T>
void SetSyncInfo(AbstractDistMatrix<T>& A, const SyncInfo<Device::GPU>& si) { A.SetSyncInfo(si); }

// ---------------------------------------------------------------------------------------------
// Proxies (include/El/core/Proxy.hpp:148-320 read, :481-560 write, :694-760 read-write): hand a routine its
// operand as DistMatrix<T,U,V>. When the operand already has that distribution (and the alignments the
// control structure constrains) the proxy IS the operand — no copy, Proxy.hpp:236-257 — otherwise a
// redistributed copy, written back on destruction by the write / read-write forms (not while unwinding).
// Only S == T: the engine has no mixed-precision copies.
// ---------------------------------------------------------------------------------------------
struct ElementalProxyCtrl {
    bool colConstrain = false, rowConstrain = false, rootConstrain = false;
    Int colAlign = 0, rowAlign = 0, root = 0;
};
namespace detail {
template <typename T, Dist U, Dist V>
class ProxyBase {
public:
    typedef DistMatrix<T, U, V, ELEMENT, Device::GPU> proxType;
    const proxType& GetLocked() const { return *prox_; }
    proxType& Get() { return *prox_; }

protected:
    ProxyBase(AbstractDistMatrix<T>& A, const ElementalProxyCtrl& ctrl, bool read, bool write) : orig_(&A), write_(write) {
        if (A.ColDist() == U && A.RowDist() == V) {
            const bool colMis = ctrl.colConstrain && A.ColAlign() != ctrl.colAlign;
            const bool rowMis = ctrl.rowConstrain && A.RowAlign() != ctrl.rowAlign;
            const bool rootMis = ctrl.rootConstrain && A.Root() != ctrl.root;
            if (!colMis && !rowMis && !rootMis) {
                prox_ = static_cast<proxType*>(&A);  // same layout: the handle carries everything (U,V) would add
                return;
            }
        }
        made_ = true;
        prox_ = new proxType(A.Grid());
        try {
            if (ctrl.rootConstrain) prox_->SetRoot(ctrl.root);
            if (ctrl.colConstrain) prox_->AlignCols(ctrl.colAlign);
            if (ctrl.rowConstrain) prox_->AlignRows(ctrl.rowAlign);
            if (read) *prox_ = A;
            else prox_->Resize(A.Height(), A.Width());
        } catch (...) { delete prox_; throw; }
    }
    ~ProxyBase() noexcept(false) {
        if (!made_) return;
        const bool unwinding = std::uncaught_exceptions() > 0;
        if (write_ && !unwinding) {
            const int st = elb200_dm_copy(prox_->Handle(), orig_->Handle());
            delete prox_;
            Check(st, "proxy write-back");
            return;
        }
        delete prox_;
    }
    ProxyBase(const ProxyBase&) = delete;
    ProxyBase& operator=(const ProxyBase&) = delete;

private:
    AbstractDistMatrix<T>* orig_;
    proxType* prox_ = nullptr;
    bool made_ = false, write_;
};
}  // namespace detail

template <typename S, typename T, Dist U = MC, Dist V = MR, DistWrap W = ELEMENT, Device D = Device::GPU>
class DistMatrixReadProxy : public detail::ProxyBase<T, U, V> {
    static_assert(std::is_same<S, T>::value && W == ELEMENT && D == Device::GPU, "proxies: same type, ELEMENT wrap, Device::GPU");
public:
    DistMatrixReadProxy(const AbstractDistMatrix<S>& A, const ElementalProxyCtrl& ctrl = ElementalProxyCtrl())
        : detail::ProxyBase<T, U, V>(const_cast<AbstractDistMatrix<S>&>(A), ctrl, true, false) {}
};
template <typename S, typename T, Dist U = MC, Dist V = MR, DistWrap W = ELEMENT, Device D = Device::GPU>
class DistMatrixWriteProxy : public detail::ProxyBase<T, U, V> {
    static_assert(std::is_same<S, T>::value && W == ELEMENT && D == Device::GPU, "proxies: same type, ELEMENT wrap, Device::GPU");
public:
    DistMatrixWriteProxy(AbstractDistMatrix<S>& A, const ElementalProxyCtrl& ctrl = ElementalProxyCtrl())
        : detail::ProxyBase<T, U, V>(A, ctrl, false, true) {}
};
template <typename S, typename T, Dist U = MC, Dist V = MR, DistWrap W = ELEMENT, Device D = Device::GPU>
class DistMatrixReadWriteProxy : public detail::ProxyBase<T, U, V> {
    static_assert(std::is_same<S, T>::value && W == ELEMENT && D == Device::GPU, "proxies: same type, ELEMENT wrap, Device::GPU");
public:
    DistMatrixReadWriteProxy(AbstractDistMatrix<S>& A, const ElementalProxyCtrl& ctrl = ElementalProxyCtrl())
        : detail::ProxyBase<T, U, V>(A, ctrl, true, true) {}
};

// ---------------------------------------------------------------------------------------------
// BLAS-like level 1 on DistMatrix (include/El/blas_like/level1/decl.hpp in the reference)
// ---------------------------------------------------------------------------------------------
template <typename T>
void Copy(const AbstractDistMatrix<T>& A, AbstractDistMatrix<T>& B) {
    detail::Check(elb200_dm_copy(A.Handle(), B.Handle()), "Copy");
}
template <typename T>
void Transpose(const AbstractDistMatrix<T>& A, AbstractDistMatrix<T>& B) {
    detail::Check(elb200_dm_transpose(A.Handle(), B.Handle()), "Transpose");
}
template <typename T>
void Axpy(T alpha, const AbstractDistMatrix<T>& X, AbstractDistMatrix<T>& Y) {
    detail::Check(elb200_dm_axpy(double(alpha), X.Handle(), Y.Handle()), "Axpy");
}
template <typename T>
void Scale(T alpha, AbstractDistMatrix<T>& A) { detail::Check(elb200_dm_scale(double(alpha), A.Handle()), "Scale"); }
template <typename T>
void Zero(AbstractDistMatrix<T>& A) { detail::Check(elb200_dm_zero(A.Handle()), "Zero"); }
// B += alpha * (sum over the replicas of A)   (level1/AxpyContract.hpp:483-509)
template <typename T>
void AxpyContract(T alpha, const AbstractDistMatrix<T>& A, AbstractDistMatrix<T>& B) {
    detail::Check(elb200_dm_axpy_contract(double(alpha), A.Handle(), B.Handle()), "AxpyContract");
}
template <typename T>
void Contract(const AbstractDistMatrix<T>& A, AbstractDistMatrix<T>& B) {
    detail::Check(elb200_dm_contract(A.Handle(), B.Handle()), "Contract");
}
template <typename T>
void TransposeAxpyContract(T alpha, const AbstractDistMatrix<T>& A, AbstractDistMatrix<T>& B) {
    detail::Check(elb200_dm_transpose_axpy_contract(double(alpha), A.Handle(), B.Handle()), "TransposeAxpyContract");
}

// ---------------------------------------------------------------------------------------------
// BLAS-like level 3 (include/El/blas_like/level3.hpp:67-90)
// ---------------------------------------------------------------------------------------------
// C := alpha op(A) op(B) + beta C
template <typename T>
void Gemm(Orientation orientA, Orientation orientB, T alpha, const AbstractDistMatrix<T>& A,
          const AbstractDistMatrix<T>& B, T beta, AbstractDistMatrix<T>& C, GemmAlgorithm alg = GEMM_DEFAULT) {
    detail::Check(elb200_dist_gemm(int(orientA), int(orientB), double(alpha), A.Handle(), B.Handle(), double(beta),
                                   C.Handle(), int(alg)), "Gemm");
}
// C := alpha op(A) op(B), C resized to fit (Gemm.cpp:304-316)
template <typename T>
void Gemm(Orientation orientA, Orientation orientB, T alpha, const AbstractDistMatrix<T>& A,
          const AbstractDistMatrix<T>& B, AbstractDistMatrix<T>& C, GemmAlgorithm alg = GEMM_DEFAULT) {
    const Int m = orientA == NORMAL ? A.Height() : A.Width();
    const Int n = orientB == NORMAL ? B.Width() : B.Height();
    C.Resize(m, n);
    Gemm(orientA, orientB, alpha, A, B, T(0), C, alg);
}
// Local product of conformally distributed operands (Gemm.cpp:318-440)
template <typename T>
void LocalGemm(Orientation orientA, Orientation orientB, T alpha, const AbstractDistMatrix<T>& A,
               const AbstractDistMatrix<T>& B, T beta, AbstractDistMatrix<T>& C) {
    detail::Check(elb200_dm_local_gemm(int(orientA), int(orientB), double(alpha), A.Handle(), B.Handle(), double(beta),
                                       C.Handle()), "LocalGemm");
}

// C resized to fit, then C := alpha op(A) op(B) locally (Gemm.cpp:428-440)
template <typename T>
void LocalGemm(Orientation orientA, Orientation orientB, T alpha, const AbstractDistMatrix<T>& A,
               const AbstractDistMatrix<T>& B, AbstractDistMatrix<T>& C) {
    C.Resize(orientA == NORMAL ? A.Height() : A.Width(), orientB == NORMAL ? B.Width() : B.Height());
    LocalGemm(orientA, orientB, alpha, A, B, T(0), C);
}

// ---------------------------------------------------------------------------------------------
// Triangular rank-k family (include/El/blas_like/level3.hpp:92-126,239-252,292-313,499-536 and the
// level-1 trapezoid routines).  conjugate / ADJOINT equal the plain forms for the real types here.
// ---------------------------------------------------------------------------------------------
template <typename T> Int LocalTrrkBlocksize() { return Int(elb200_local_trrk_blocksize()); }
template <typename T> void SetLocalTrrkBlocksize(Int blocksize) { elb200_set_local_trrk_blocksize(blocksize); }

template <typename T, typename S>
void ScaleTrapezoid(S alpha, UpperOrLower uplo, AbstractDistMatrix<T>& A, Int offset = 0) {
    detail::Check(elb200_dm_scale_trapezoid(double(alpha), int(uplo), A.Handle(), offset), "ScaleTrapezoid");
}
template <typename T, typename S>
void AxpyTrapezoid(UpperOrLower uplo, S alpha, const AbstractDistMatrix<T>& X, AbstractDistMatrix<T>& Y, Int offset = 0) {
    detail::Check(elb200_dm_axpy_trapezoid(int(uplo), double(alpha), X.Handle(), Y.Handle(), offset), "AxpyTrapezoid");
}
template <typename T>
void MakeTrapezoidal(UpperOrLower uplo, AbstractDistMatrix<T>& A, Int offset = 0) {
    detail::Check(elb200_dm_make_trapezoidal(int(uplo), A.Handle(), offset), "MakeTrapezoidal");
}
template <typename T>
void MakeSymmetric(UpperOrLower uplo, AbstractDistMatrix<T>& A, bool /*conjugate*/ = false) {
    detail::Check(elb200_dm_make_symmetric(int(uplo), A.Handle()), "MakeSymmetric");
}
template <typename T>
void MakeHermitian(UpperOrLower uplo, AbstractDistMatrix<T>& A) { MakeSymmetric(uplo, A, true); }

// triangle of C := alpha op(A) op(B) + beta C  (Trrk.cpp:100-118)
template <typename T>
void Trrk(UpperOrLower uplo, Orientation orientA, Orientation orientB, T alpha, const AbstractDistMatrix<T>& A,
          const AbstractDistMatrix<T>& B, T beta, AbstractDistMatrix<T>& C) {
    detail::Check(elb200_dist_trrk(int(uplo), int(orientA), int(orientB), double(alpha), A.Handle(), B.Handle(), double(beta),
                                   C.Handle()), "Trrk");
}
// The four LocalTrrk overloads (level3.hpp:513-536; Trrk/Local.hpp:688-890): panels aligned with C.
template <typename T>
void LocalTrrk(UpperOrLower uplo, T alpha, const DistMatrix<T, MC, STAR>& A, const DistMatrix<T, STAR, MR>& B, T beta,
               DistMatrix<T, MC, MR>& C) {
    detail::Check(elb200_dm_local_trrk(int(uplo), 0, 0, double(alpha), A.Handle(), B.Handle(), double(beta), C.Handle()), "LocalTrrk");
}
template <typename T>
void LocalTrrk(UpperOrLower uplo, Orientation orientB, T alpha, const DistMatrix<T, MC, STAR>& A,
               const DistMatrix<T, MR, STAR>& B, T beta, DistMatrix<T, MC, MR>& C) {
    if (orientB == NORMAL) throw std::logic_error("LocalTrrk: B must be (Conjugate)Transpose'd");  // Local.hpp:74-75
    detail::Check(elb200_dm_local_trrk(int(uplo), 0, 1, double(alpha), A.Handle(), B.Handle(), double(beta), C.Handle()), "LocalTrrk");
}
template <typename T>
void LocalTrrk(UpperOrLower uplo, Orientation orientA, T alpha, const DistMatrix<T, STAR, MC>& A,
               const DistMatrix<T, STAR, MR>& B, T beta, DistMatrix<T, MC, MR>& C) {
    if (orientA == NORMAL) throw std::logic_error("LocalTrrk: A must be (Conjugate)Transpose'd");  // Local.hpp:89-90
    detail::Check(elb200_dm_local_trrk(int(uplo), 1, 0, double(alpha), A.Handle(), B.Handle(), double(beta), C.Handle()), "LocalTrrk");
}
template <typename T>
void LocalTrrk(UpperOrLower uplo, Orientation orientA, Orientation orientB, T alpha, const DistMatrix<T, STAR, MC>& A,
               const DistMatrix<T, MR, STAR>& B, T beta, DistMatrix<T, MC, MR>& C) {
    if (orientA == NORMAL || orientB == NORMAL) throw std::logic_error("LocalTrrk: A and B must be (Conjugate)Transpose'd");
    detail::Check(elb200_dm_local_trrk(int(uplo), 1, 1, double(alpha), A.Handle(), B.Handle(), double(beta), C.Handle()), "LocalTrrk");
}
// triangle of C := alpha A A^T + beta C (NORMAL) or alpha A^T A + beta C  (Syrk.cpp:172-190)
template <typename T>
void Syrk(UpperOrLower uplo, Orientation orientation, T alpha, const AbstractDistMatrix<T>& A, T beta,
          AbstractDistMatrix<T>& C, bool /*conjugate*/ = false) {
    detail::Check(elb200_dist_syrk(int(uplo), int(orientation), double(alpha), A.Handle(), double(beta), C.Handle()), "Syrk");
}
// C resized to n x n and zeroed first (Syrk.cpp:192-203)
template <typename T>
void Syrk(UpperOrLower uplo, Orientation orientation, T alpha, const AbstractDistMatrix<T>& A, AbstractDistMatrix<T>& C,
          bool conjugate = false) {
    const Int n = orientation == NORMAL ? A.Height() : A.Width();
    C.Resize(n, n);
    Zero(C);
    Syrk(uplo, orientation, alpha, A, T(0), C, conjugate);
}
template <typename T>
void Herk(UpperOrLower uplo, Orientation orientation, T alpha, const AbstractDistMatrix<T>& A, T beta, AbstractDistMatrix<T>& C) {
    Syrk(uplo, orientation, alpha, A, beta, C, true);
}
template <typename T>
void Herk(UpperOrLower uplo, Orientation orientation, T alpha, const AbstractDistMatrix<T>& A, AbstractDistMatrix<T>& C) {
    Syrk(uplo, orientation, alpha, A, C, true);
}
// C := alpha A B + beta C (LEFT) or alpha B A + beta C (RIGHT), A symmetric, `uplo` triangle referenced (Symm.cpp:56-80)
template <typename T>
void Symm(LeftOrRight side, UpperOrLower uplo, T alpha, const AbstractDistMatrix<T>& A, const AbstractDistMatrix<T>& B, T beta,
          AbstractDistMatrix<T>& C, bool /*conjugate*/ = false) {
    detail::Check(elb200_dist_symm(int(side), int(uplo), double(alpha), A.Handle(), B.Handle(), double(beta), C.Handle()), "Symm");
}
template <typename T>
void Hemm(LeftOrRight side, UpperOrLower uplo, T alpha, const AbstractDistMatrix<T>& A, const AbstractDistMatrix<T>& B, T beta,
          AbstractDistMatrix<T>& C) {
    Symm(side, uplo, alpha, A, B, beta, C, true);
}

// B := alpha op(A)^-1 B (LEFT) or alpha B op(A)^-1 (RIGHT)  (Trsm.cpp:474-499). checkIfSingular is accepted and not
// evaluated (the reference's GPU path, cuBLAS trsm, does not evaluate it either); every TrsmAlgorithm value runs the
// many-right-hand-sides schedule.
template <typename F>
void Trsm(LeftOrRight side, UpperOrLower uplo, Orientation orientation, UnitOrNonUnit diag, F alpha, const AbstractDistMatrix<F>& A,
          AbstractDistMatrix<F>& B, bool /*checkIfSingular*/ = false, TrsmAlgorithm alg = TRSM_DEFAULT) {
    detail::Check(elb200_dist_trsm(int(side), int(uplo), int(orientation), int(diag), double(alpha), A.Handle(), B.Handle(), int(alg)), "Trsm");
}
// A replicated, X distributed over the dimension that is not solved (Trsm.cpp:501-526)
template <typename F>
void LocalTrsm(LeftOrRight side, UpperOrLower uplo, Orientation orientation, UnitOrNonUnit diag, F alpha,
               const AbstractDistMatrix<F>& A, AbstractDistMatrix<F>& X, bool /*checkIfSingular*/ = false) {
    detail::Check(elb200_dm_local_trsm(int(side), int(uplo), int(orientation), int(diag), double(alpha), A.Handle(), X.Handle()), "LocalTrsm");
}

// ---------------------------------------------------------------------------------------------
// Helpers the reference's drivers use around Gemm
// ---------------------------------------------------------------------------------------------
namespace detail {
// counter-based uniform in [0,1): value depends on (seed, global i, global j) only, so the global
// matrix is the same on every grid shape
inline double Hash01(std::uint64_t seed, std::uint64_t i, std::uint64_t j) {
    std::uint64_t z = seed * 0x9E3779B97F4A7C15ull + i * 0xBF58476D1CE4E5B9ull + j * 0x94D049BB133111EBull;
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull; z ^= z >> 27; z *= 0x94D049BB133111EBull; z ^= z >> 31;
    return double(z >> 11) / double(1ull << 53);
}
}  // namespace detail

// A := height x width, entries Uniform(center - radius, center + radius)
// (src/matrices/random/independent/Uniform.cpp:18-57; generator differs: counter-based)
template <typename T>
void Uniform(AbstractDistMatrix<T>& A, Int height, Int width, T center = T(0), T radius = T(1), std::uint64_t seed = 21) {
    A.Resize(height, width);
    const Int lh = A.LocalHeight(), lw = A.LocalWidth();
    std::vector<T> host(size_t(lh) * size_t(lw));
    for (Int j = 0; j < lw; ++j)
        for (Int i = 0; i < lh; ++i)
            host[size_t(i) + size_t(j) * lh] =
                T(double(center) - double(radius) + 2.0 * double(radius) * detail::Hash01(seed, A.GlobalRow(i), A.GlobalCol(j)));
    A.SetLocalFromHost(host.data());
}

// Frobenius norm of a [MC,MR]-like (non-replicated) matrix gathered through [*,*]
template <typename T>
double FrobeniusNorm(const AbstractDistMatrix<T>& A) {
    DistMatrix<T, STAR, STAR> G(A.Grid());
    G = A;
    std::vector<T> host(size_t(G.LocalHeight()) * size_t(G.LocalWidth()));
    G.GetLocalToHost(host.data());
    long double s = 0;
    for (T v : host) s += (long double)v * (long double)v;
    return double(std::sqrt(s));
}


// ---------------------------------------------------------------------------------------------
// Complex instantiations: DistMatrix<Complex<float|double>,U,V,ELEMENT,Device::GPU>, El::Gemm, Copy, Transpose /
// Adjoint, Axpy, Scale, Zero on them, and Gemm(Matrix<Complex<T>>) — the reference's complex instantiations of
// src/blas_like/level3/Gemm.cpp:504-506 and the ElementMatrix / level-1 sources (El/macros/Instantiate.h). The engine
// stores a complex DistMatrix PLANAR and multiplies with four real products (elb200_zdm_*, csrc/capi_complex.cpp), so the
// local matrix is not addressable as one interleaved device buffer: local data crosses through SetLocalFromHost /
// GetLocalToHost (interleaved, ld = LocalHeight()), Get / Set, or the real planes.
// ---------------------------------------------------------------------------------------------
namespace detail {
template <typename R>
class ComplexDistMatrixBase {
public:
    typedef Complex<R> value_type;
    virtual ~ComplexDistMatrixBase() { if (h_) elb200_zdm_destroy(h_); }
    ComplexDistMatrixBase(const ComplexDistMatrixBase&) = delete;
    ComplexDistMatrixBase& operator=(const ComplexDistMatrixBase&) = delete;

    void Resize(Int height, Int width) { Check(elb200_zdm_resize(h_, height, width), "Resize"); }
    void Align(Int colAlign, Int rowAlign, bool = true) { Check(elb200_zdm_align(h_, colAlign, rowAlign), "Align"); }
    void AlignWith(const ComplexDistMatrixBase<R>& A) { Check(elb200_zdm_align_with(h_, A.Handle()), "AlignWith"); }
    void SetRoot(int root) { Check(elb200_zdm_set_root(h_, root), "SetRoot"); }
    void SetSyncInfo(const SyncInfo<Device::GPU>& si) { Check(elb200_zdm_set_stream(h_, si.Stream()), "SetSyncInfo"); }
    const El::Grid& Grid() const { return *grid_; }
    Int Height() const { return Int(Info().height); }
    Int Width() const { return Int(Info().width); }
    Int LocalHeight() const { return Int(Info().local_height); }
    Int LocalWidth() const { return Int(Info().local_width); }
    Int ColAlign() const { return Int(Info().col_align); }
    Int RowAlign() const { return Int(Info().row_align); }
    Int ColShift() const { return Int(Info().col_shift); }
    Int RowShift() const { return Int(Info().row_shift); }
    Int ColStride() const { return Int(Info().col_stride); }
    Int RowStride() const { return Int(Info().row_stride); }
    Dist ColDist() const { return FromEngineDist(Info().col_dist); }
    Dist RowDist() const { return FromEngineDist(Info().row_dist); }
    Int GlobalRow(Int iLoc) const { const auto d = Info(); return Int(d.col_shift + iLoc * d.col_stride); }
    Int GlobalCol(Int jLoc) const { const auto d = Info(); return Int(d.row_shift + jLoc * d.row_stride); }
    cudaStream_t Stream() const { return static_cast<cudaStream_t>(elb200_dm_stream(Plane(0))); }
    elb200_zdm_t Handle() const { return h_; }
    elb200_dm_t Plane(int imag) const { elb200_dm_t p = nullptr; Check(elb200_zdm_plane(h_, imag, &p), "Plane"); return p; }
    // interleaved host <-> local matrix, leading dimension LocalHeight() on the host side; synchronous
    void SetLocalFromHost(const Complex<R>* host) {
        Check(elb200_zdm_write_local(h_, host, LocalHeight() > 0 ? LocalHeight() : 1), "SetLocalFromHost");
        if (cudaStreamSynchronize(Stream()) != cudaSuccess) throw std::runtime_error("SetLocalFromHost: stream synchronisation failed");
    }
    void GetLocalToHost(Complex<R>* host) const {
        Check(elb200_zdm_read_local(h_, host, LocalHeight() > 0 ? LocalHeight() : 1), "GetLocalToHost");
        if (cudaStreamSynchronize(Stream()) != cudaSuccess) throw std::runtime_error("GetLocalToHost: stream synchronisation failed");
    }

protected:
    ComplexDistMatrixBase(const El::Grid& g, Dist U, Dist V) : grid_(&g) {
        Check(elb200_zdm_create(g.Handle(), TypeCode<R>::value, ToEngineDist(U), ToEngineDist(V), &h_), "DistMatrix<Complex>");
    }
    ComplexDistMatrixBase(const El::Grid& g, std::nullptr_t) : grid_(&g) {}
    void MoveFrom(ComplexDistMatrixBase& o) noexcept {
        if (this != &o) { if (h_) elb200_zdm_destroy(h_); grid_ = o.grid_; h_ = o.h_; o.h_ = nullptr; }
    }
    elb200_dm_info_t Info() const {
        elb200_dm_info_t d;
        Check(elb200_dm_info(Plane(0), &d), "DistMatrix<Complex>::Info");
        return d;
    }
    const El::Grid* grid_;
    elb200_zdm_t h_ = nullptr;
};
}  // namespace detail

template <typename R, Dist U, Dist V>
class DistMatrix<Complex<R>, U, V, ELEMENT, Device::GPU> : public detail::ComplexDistMatrixBase<R> {
    static_assert(std::is_same<R, float>::value || std::is_same<R, double>::value, "Complex<float> and Complex<double> only");

public:
    explicit DistMatrix(const El::Grid& g) : detail::ComplexDistMatrixBase<R>(g, U, V) {}
    DistMatrix(Int height, Int width, const El::Grid& g) : DistMatrix(g) { this->Resize(height, width); }
    DistMatrix(DistMatrix&& o) noexcept : detail::ComplexDistMatrixBase<R>(*o.grid_, nullptr) { this->MoveFrom(o); }
    DistMatrix& operator=(DistMatrix&& o) noexcept { this->MoveFrom(o); return *this; }
    // redistribution from any (U2,V2): the real path applied to both planes
    DistMatrix& operator=(const detail::ComplexDistMatrixBase<R>& A) {
        detail::Check(elb200_zdm_copy(A.Handle(), this->Handle()), "DistMatrix<Complex>::operator=");
        return *this;
    }
    DistMatrix& operator=(const DistMatrix& A) { return *this = static_cast<const detail::ComplexDistMatrixBase<R>&>(A); }
};

template <typename R>
void Copy(const detail::ComplexDistMatrixBase<R>& A, detail::ComplexDistMatrixBase<R>& B) {
    detail::Check(elb200_zdm_copy(A.Handle(), B.Handle()), "Copy");
}
template <typename R>
void Transpose(const detail::ComplexDistMatrixBase<R>& A, detail::ComplexDistMatrixBase<R>& B, bool conjugate = false) {
    detail::Check(elb200_zdm_transpose(A.Handle(), B.Handle(), conjugate ? 1 : 0), "Transpose");
}
template <typename R>
void Adjoint(const detail::ComplexDistMatrixBase<R>& A, detail::ComplexDistMatrixBase<R>& B) { Transpose(A, B, true); }
template <typename R>
void Zero(detail::ComplexDistMatrixBase<R>& A) { detail::Check(elb200_zdm_zero(A.Handle()), "Zero"); }
template <typename R>
void Scale(Complex<R> alpha, detail::ComplexDistMatrixBase<R>& A) {
    detail::Check(elb200_zdm_scale(double(alpha.real()), double(alpha.imag()), A.Handle()), "Scale");
}
template <typename R>
void Axpy(Complex<R> alpha, const detail::ComplexDistMatrixBase<R>& X, detail::ComplexDistMatrixBase<R>& Y) {
    detail::Check(elb200_zdm_axpy(double(alpha.real()), double(alpha.imag()), X.Handle(), Y.Handle()), "Axpy");
}
// C := alpha op(A) op(B) + beta C; ADJOINT is the conjugate transpose
template <typename R>
void Gemm(Orientation orientA, Orientation orientB, Complex<R> alpha, const detail::ComplexDistMatrixBase<R>& A,
          const detail::ComplexDistMatrixBase<R>& B, Complex<R> beta, detail::ComplexDistMatrixBase<R>& C, GemmAlgorithm alg = GEMM_DEFAULT) {
    detail::Check(elb200_zdm_gemm(int(orientA), int(orientB), double(alpha.real()), double(alpha.imag()), A.Handle(), B.Handle(),
                                  double(beta.real()), double(beta.imag()), C.Handle(), int(alg)), "Gemm");
}
template <typename R>
void Gemm(Orientation orientA, Orientation orientB, Complex<R> alpha, const detail::ComplexDistMatrixBase<R>& A,
          const detail::ComplexDistMatrixBase<R>& B, detail::ComplexDistMatrixBase<R>& C, GemmAlgorithm alg = GEMM_DEFAULT) {
    C.Resize(orientA == NORMAL ? A.Height() : A.Width(), orientB == NORMAL ? B.Width() : B.Height());
    Gemm(orientA, orientB, alpha, A, B, Complex<R>(0), C, alg);
}
// Gemm(Matrix<Complex<T>,Device::GPU>): interleaved local matrices (the cublasCgemm / Zgemm call of the reference)
namespace detail {
inline int LocalComplexGemm(int ta, int tb, long m, long n, long k, Complex<float> al, const Complex<float>* A, long lda, const Complex<float>* B,
                            long ldb, Complex<float> be, Complex<float>* C, long ldc, void* s) {
    const float a2[2] = {al.real(), al.imag()}, b2[2] = {be.real(), be.imag()};
    return elb200_gemm_c32(ta, tb, m, n, k, a2, A, lda, B, ldb, b2, C, ldc, s);
}
inline int LocalComplexGemm(int ta, int tb, long m, long n, long k, Complex<double> al, const Complex<double>* A, long lda, const Complex<double>* B,
                            long ldb, Complex<double> be, Complex<double>* C, long ldc, void* s) {
    const double a2[2] = {al.real(), al.imag()}, b2[2] = {be.real(), be.imag()};
    return elb200_gemm_c64(ta, tb, m, n, k, a2, A, lda, B, ldb, b2, C, ldc, s);
}
}  // namespace detail
template <typename R>
void Gemm(Orientation orientA, Orientation orientB, Complex<R> alpha, const Matrix<Complex<R>>& A, const Matrix<Complex<R>>& B, Complex<R> beta,
          Matrix<Complex<R>>& C) {
    const Int m = orientA == NORMAL ? A.Height() : A.Width(), ka = orientA == NORMAL ? A.Width() : A.Height();
    const Int n = orientB == NORMAL ? B.Width() : B.Height(), kb = orientB == NORMAL ? B.Height() : B.Width();
    if (ka != kb || C.Height() != m || C.Width() != n) throw std::logic_error("Gemm: nonconformal matrices");
    detail::Check(detail::LocalComplexGemm(int(orientA), int(orientB), m, n, ka, alpha, A.LockedBuffer(), A.LDim(), B.LockedBuffer(), B.LDim(),
                                           beta, C.Buffer(), C.LDim(), C.GetSyncInfo().Stream()), "Gemm");
}

}  // namespace El
