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
#include <cuda_runtime_api.h>

#include <cmath>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <utility>
#include <vector>

#include "elb200.h"

namespace El {

typedef int Int;  // reference default (no EL_USE_64BIT_INTS)

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

// El::Initialize / Finalize / Environment (src/core/environment.cpp:215-372). The reference brings up
// MPI, the GPU and the blocksize stack; here there is no MPI (the NCCL bootstrap happens in Grid),
// so these only make sure a device is usable and keep the call sites compiling.
inline void Initialize(int& /*argc*/, char**& /*argv*/) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n < 1)
        throw std::runtime_error("El::Initialize: no CUDA device (there is no CPU path)");
}
inline void Initialize() { int c = 0; char** v = nullptr; Initialize(c, v); }
inline void Finalize() {}
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
template <typename T> struct TypeCode;
template <> struct TypeCode<float> { static constexpr int value = ELB200_F32; };
template <> struct TypeCode<double> { static constexpr int value = ELB200_F64; };
}  // namespace detail

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
    static_assert(std::is_same<T, float>::value || std::is_same<T, double>::value, "float and double only");

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
        const Int ld = ldim > 0 ? ldim : (height > 0 ? (height + 3) / 4 * 4 : 4);  // 16-byte column pitch
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
// El::Grid — r x c process grid, COLUMN_MAJOR (VC rank = row + r*col), one process per GPU.
// The reference builds it from an MPI communicator; here the bootstrap is an NCCL unique id that
// rank 0 creates (Grid::UniqueId) and the caller ships to the other ranks by any side channel.
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

    // single-process 1 x 1 grid on the current device
    Grid() { Create(1, 1, 0, nullptr); }
    // `size` processes, this one is VC rank `rank`; height <= 0 selects DefaultHeight(size)
    Grid(const Id& id, int rank, int size, int height = -1) {
        const int h = height > 0 ? height : DefaultHeight(size);
        if (size % h != 0) throw std::logic_error("Grid: height must divide the number of processes");
        Create(h, size / h, rank, id.data());
    }
    ~Grid() { if (h_) elb200_grid_destroy(h_); }
    Grid(const Grid&) = delete;             // non-copyable like the reference (Grid.hpp:141-145)
    Grid& operator=(const Grid&) = delete;

    int Height() const { return height_; }
    int Width() const { return width_; }
    int Size() const { return height_ * width_; }
    int Rank() const { return vc_; }
    int VCRank() const { return vc_; }
    int VRRank() const { return mr_ + width_ * mc_; }
    int Row() const { return mc_; }
    int Col() const { return mr_; }
    int MCRank() const { return mc_; }
    int MRRank() const { return mr_; }
    GridOrder Order() const { return COLUMN_MAJOR; }
    cudaStream_t Stream() const { return static_cast<cudaStream_t>(elb200_grid_stream(h_)); }
    void Synchronize() const { detail::Check(elb200_grid_sync(h_), "Grid::Synchronize"); }
    elb200_grid_t Handle() const { return h_; }

private:
    void Create(int h, int w, int rank, const void* id) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess) throw std::runtime_error("Grid: no CUDA device (there is no CPU path)");
        detail::Check(elb200_grid_create(h, w, rank, id, dev, &h_), "Grid");
        detail::Check(elb200_grid_info(h_, &height_, &width_, &vc_, &mc_, &mr_), "Grid");
    }
    elb200_grid_t h_ = nullptr;
    int height_ = 1, width_ = 1, vc_ = 0, mc_ = 0, mr_ = 0;
};

// El::Blocksize stack (src/blas_like/blocksizes.cpp:35-67)
inline Int Blocksize() { return elb200_blocksize(); }
inline void SetBlocksize(Int nb) { elb200_set_blocksize(nb); }
inline void PushBlocksizeStack(Int nb) { elb200_push_blocksize(nb); }
inline void PopBlocksizeStack() { elb200_pop_blocksize(); }

// ---------------------------------------------------------------------------------------------
// El::DistMatrix<T,U,V> (ELEMENT wrap, Device::GPU).  Owns its local matrix unless it is a view.
// ---------------------------------------------------------------------------------------------
template <typename T, Dist U = MC, Dist V = MR, DistWrap W = ELEMENT, Device D = Device::GPU>
class DistMatrix {
    static_assert(std::is_same<T, float>::value || std::is_same<T, double>::value, "float and double only");
    static_assert(W == ELEMENT, "BLOCK-wrapped distributions are outside the Gemm path");
    static_assert(D == Device::GPU, "elemental_b200 has no CPU path: Device::GPU only");

public:
    explicit DistMatrix(const El::Grid& g) : grid_(&g) {
        detail::Check(elb200_dm_create(g.Handle(), detail::TypeCode<T>::value, detail::ToEngineDist(U),
                                       detail::ToEngineDist(V), &h_), "DistMatrix");
    }
    DistMatrix(Int height, Int width, const El::Grid& g) : DistMatrix(g) { Resize(height, width); }
    ~DistMatrix() { if (h_) elb200_dm_destroy(h_); }
    DistMatrix(DistMatrix&& o) noexcept : grid_(o.grid_), h_(o.h_) { o.h_ = nullptr; }
    DistMatrix& operator=(DistMatrix&& o) noexcept {
        if (this != &o) { if (h_) elb200_dm_destroy(h_); grid_ = o.grid_; h_ = o.h_; o.h_ = nullptr; }
        return *this;
    }
    // B = A across distributions (the operator= tables of src/core/DistMatrix/ElementMatrix/*.cpp)
    template <Dist U2, Dist V2>
    DistMatrix& operator=(const DistMatrix<T, U2, V2, W, D>& A) {
        detail::Check(elb200_dm_copy(A.Handle(), h_), "DistMatrix::operator=");
        return *this;
    }
    DistMatrix& operator=(const DistMatrix& A) {
        if (this != &A) detail::Check(elb200_dm_copy(A.Handle(), h_), "DistMatrix::operator=");
        return *this;
    }
    DistMatrix(const DistMatrix& A) : DistMatrix(*A.grid_) { *this = A; }

    void Align(Int colAlign, Int rowAlign) { detail::Check(elb200_dm_align(h_, colAlign, rowAlign), "Align"); }
    template <Dist U2, Dist V2>
    void AlignWith(const DistMatrix<T, U2, V2, W, D>& A) { detail::Check(elb200_dm_align_with(h_, A.Handle()), "AlignWith"); }
    void Resize(Int height, Int width) { detail::Check(elb200_dm_resize(h_, height, width), "Resize"); }
    // [CIRC,CIRC] only: the VC rank that holds the matrix (AbstractDistMatrix.hpp:58,113)
    void SetRoot(int root) { detail::Check(elb200_dm_set_root(h_, root), "SetRoot"); }
    int Root() const { return Info().root; }
    // Attach(height, width, grid, colAlign, rowAlign, buffer, ldim): borrow device memory
    void Attach(Int height, Int width, Int colAlign, Int rowAlign, T* deviceBuffer, Int ldim) {
        detail::Check(elb200_dm_attach(h_, height, width, colAlign, rowAlign, deviceBuffer, ldim), "Attach");
    }
    // A(IR(i0,i1), IR(j0,j1)): zero-copy view, alignments advance with the offset
    DistMatrix operator()(Range I, Range J) const {
        const elb200_dm_info_t d = Info();
        const Int i1 = I.end < 0 ? Int(d.height) : I.end, j1 = J.end < 0 ? Int(d.width) : J.end;
        DistMatrix v(*grid_, nullptr);
        detail::Check(elb200_dm_view(h_, I.beg, i1, J.beg, j1, &v.h_), "DistMatrix::operator()");
        return v;
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
    Dist ColDist() const { return U; }
    Dist RowDist() const { return V; }
    bool Viewing() const { return Info().is_view != 0; }
    Int GlobalRow(Int iLoc) const { const auto d = Info(); return Int(d.col_shift + iLoc * d.col_stride); }
    Int GlobalCol(Int jLoc) const { const auto d = Info(); return Int(d.row_shift + jLoc * d.row_stride); }
    T* Buffer() { return static_cast<T*>(Info().buffer); }              // DEVICE pointer
    const T* LockedBuffer() const { return static_cast<const T*>(Info().buffer); }
    elb200_dm_t Handle() const { return h_; }
    Device GetLocalDevice() const { return Device::GPU; }
    DistWrap Wrap() const { return W; }
    bool Participating() const { return true; }  // every process of the grid takes part (no viewing-comm subsets)
    // local <-> global index maps (include/El/core/DistMatrix/AbstractDistMatrix.hpp)
    bool IsLocalRow(Int i) const { const auto d = Info(); return d.local_height > 0 && (i - d.col_shift) % d.col_stride == 0 && i >= d.col_shift; }
    bool IsLocalCol(Int j) const { const auto d = Info(); return d.local_width > 0 && (j - d.row_shift) % d.row_stride == 0 && j >= d.row_shift; }
    bool IsLocal(Int i, Int j) const { return IsLocalRow(i) && IsLocalCol(j); }
    Int LocalRow(Int i) const { const auto d = Info(); return Int((i - d.col_shift) / d.col_stride); }
    Int LocalCol(Int j) const { const auto d = Info(); return Int((j - d.row_shift) / d.row_stride); }
    // The local matrix as a zero-copy El::Matrix view on the grid's stream (Matrix() / LockedMatrix()).
    El::Matrix<T, Device::GPU> Matrix() {
        const auto d = Info();
        El::Matrix<T, Device::GPU> m;
        m.Attach(Int(d.local_height), Int(d.local_width), static_cast<T*>(d.buffer), Int(d.ldim));
        m.SetSyncInfo(SyncInfo<Device::GPU>(grid_->Stream(), nullptr));
        return m;
    }
    El::Matrix<T, Device::GPU> LockedMatrix() const { return const_cast<DistMatrix*>(this)->Matrix(); }
    // Single local entries through the host: synchronous, for tests and setup code only
    // (the reference's GetLocal / SetLocal on a GPU matrix are the same blocking copies).
    T GetLocal(Int iLoc, Int jLoc) const {
        const auto d = Info();
        if (iLoc < 0 || jLoc < 0 || iLoc >= d.local_height || jLoc >= d.local_width) throw std::logic_error("GetLocal: index out of range");
        T v;
        CudaCheck(cudaMemcpyAsync(&v, static_cast<const T*>(d.buffer) + size_t(iLoc) + size_t(jLoc) * size_t(d.ldim), sizeof(T),
                                  cudaMemcpyDeviceToHost, grid_->Stream()));
        grid_->Synchronize();
        return v;
    }
    void SetLocal(Int iLoc, Int jLoc, T v) {
        const auto d = Info();
        if (iLoc < 0 || jLoc < 0 || iLoc >= d.local_height || jLoc >= d.local_width) throw std::logic_error("SetLocal: index out of range");
        CudaCheck(cudaMemcpyAsync(static_cast<T*>(d.buffer) + size_t(iLoc) + size_t(jLoc) * size_t(d.ldim), &v, sizeof(T),
                                  cudaMemcpyHostToDevice, grid_->Stream()));
        grid_->Synchronize();
    }

    // Single GLOBAL entries (AbstractDistMatrix::Get / Set / Update, src/core/DistMatrix/ElementMatrix.cpp). Get is
    // collective: every process of the grid calls it and receives the entry (a one-entry view gathered to [*,*]);
    // Set / Update are carried out by the owning process(es) and are no-ops elsewhere, as in the reference.
    T Get(Int i, Int j) const {
        if (i < 0 || j < 0 || i >= Height() || j >= Width()) throw std::logic_error("Get: index out of range");
        DistMatrix<T, STAR, STAR, W, D> one(*grid_);
        elb200_dm_t v = nullptr;
        detail::Check(elb200_dm_view(h_, i, i + 1, j, j + 1, &v), "Get");
        const int st = elb200_dm_copy(v, one.Handle());
        elb200_dm_destroy(v);
        detail::Check(st, "Get");
        return one.GetLocal(0, 0);
    }
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
                                    size_t(d.local_height) * sizeof(T), size_t(d.local_width), cudaMemcpyHostToDevice,
                                    grid_->Stream()));
        grid_->Synchronize();
    }
    void GetLocalToHost(T* host) const {
        const auto d = Info();
        if (d.local_height == 0 || d.local_width == 0) return;
        CudaCheck(cudaMemcpy2DAsync(host, size_t(d.local_height) * sizeof(T), d.buffer, size_t(d.ldim) * sizeof(T),
                                    size_t(d.local_height) * sizeof(T), size_t(d.local_width), cudaMemcpyDeviceToHost,
                                    grid_->Stream()));
        grid_->Synchronize();
    }

private:
    DistMatrix(const El::Grid& g, std::nullptr_t) : grid_(&g) {}
    elb200_dm_info_t Info() const {
        elb200_dm_info_t d;
        detail::Check(elb200_dm_info(h_, &d), "DistMatrix::Info");
        return d;
    }
    static void CudaCheck(cudaError_t e) { if (e != cudaSuccess) throw std::runtime_error(cudaGetErrorString(e)); }
    const El::Grid* grid_;
    elb200_dm_t h_ = nullptr;
};

// ---------------------------------------------------------------------------------------------
// BLAS-like level 1 on DistMatrix (include/El/blas_like/level1/decl.hpp in the reference)
// ---------------------------------------------------------------------------------------------
template <typename T, Dist U, Dist V, Dist U2, Dist V2>
void Copy(const DistMatrix<T, U, V>& A, DistMatrix<T, U2, V2>& B) {
    detail::Check(elb200_dm_copy(A.Handle(), B.Handle()), "Copy");
}
template <typename T, Dist U, Dist V, Dist U2, Dist V2>
void Transpose(const DistMatrix<T, U, V>& A, DistMatrix<T, U2, V2>& B) {
    detail::Check(elb200_dm_transpose(A.Handle(), B.Handle()), "Transpose");
}
template <typename T, Dist U, Dist V, Dist U2, Dist V2>
void Axpy(T alpha, const DistMatrix<T, U, V>& X, DistMatrix<T, U2, V2>& Y) {
    detail::Check(elb200_dm_axpy(double(alpha), X.Handle(), Y.Handle()), "Axpy");
}
template <typename T, Dist U, Dist V>
void Scale(T alpha, DistMatrix<T, U, V>& A) { detail::Check(elb200_dm_scale(double(alpha), A.Handle()), "Scale"); }
template <typename T, Dist U, Dist V>
void Zero(DistMatrix<T, U, V>& A) { detail::Check(elb200_dm_zero(A.Handle()), "Zero"); }
// B += alpha * (sum over the replicas of A)   (level1/AxpyContract.hpp:483-509)
template <typename T, Dist U, Dist V, Dist U2, Dist V2>
void AxpyContract(T alpha, const DistMatrix<T, U, V>& A, DistMatrix<T, U2, V2>& B) {
    detail::Check(elb200_dm_axpy_contract(double(alpha), A.Handle(), B.Handle()), "AxpyContract");
}
template <typename T, Dist U, Dist V, Dist U2, Dist V2>
void Contract(const DistMatrix<T, U, V>& A, DistMatrix<T, U2, V2>& B) {
    detail::Check(elb200_dm_contract(A.Handle(), B.Handle()), "Contract");
}
template <typename T, Dist U, Dist V, Dist U2, Dist V2>
void TransposeAxpyContract(T alpha, const DistMatrix<T, U, V>& A, DistMatrix<T, U2, V2>& B) {
    detail::Check(elb200_dm_transpose_axpy_contract(double(alpha), A.Handle(), B.Handle()), "TransposeAxpyContract");
}

// ---------------------------------------------------------------------------------------------
// BLAS-like level 3 (include/El/blas_like/level3.hpp:67-90)
// ---------------------------------------------------------------------------------------------
// C := alpha op(A) op(B) + beta C
template <typename T, Dist UA, Dist VA, Dist UB, Dist VB, Dist UC, Dist VC_>
void Gemm(Orientation orientA, Orientation orientB, T alpha, const DistMatrix<T, UA, VA>& A,
          const DistMatrix<T, UB, VB>& B, T beta, DistMatrix<T, UC, VC_>& C, GemmAlgorithm alg = GEMM_DEFAULT) {
    detail::Check(elb200_dist_gemm(int(orientA), int(orientB), double(alpha), A.Handle(), B.Handle(), double(beta),
                                   C.Handle(), int(alg)), "Gemm");
}
// C := alpha op(A) op(B), C resized to fit (Gemm.cpp:304-316)
template <typename T, Dist UA, Dist VA, Dist UB, Dist VB, Dist UC, Dist VC_>
void Gemm(Orientation orientA, Orientation orientB, T alpha, const DistMatrix<T, UA, VA>& A,
          const DistMatrix<T, UB, VB>& B, DistMatrix<T, UC, VC_>& C, GemmAlgorithm alg = GEMM_DEFAULT) {
    const Int m = orientA == NORMAL ? A.Height() : A.Width();
    const Int n = orientB == NORMAL ? B.Width() : B.Height();
    C.Resize(m, n);
    Gemm(orientA, orientB, alpha, A, B, T(0), C, alg);
}
// Local product of conformally distributed operands (Gemm.cpp:318-440)
template <typename T, Dist UA, Dist VA, Dist UB, Dist VB, Dist UC, Dist VC_>
void LocalGemm(Orientation orientA, Orientation orientB, T alpha, const DistMatrix<T, UA, VA>& A,
               const DistMatrix<T, UB, VB>& B, T beta, DistMatrix<T, UC, VC_>& C) {
    detail::Check(elb200_dm_local_gemm(int(orientA), int(orientB), double(alpha), A.Handle(), B.Handle(), double(beta),
                                       C.Handle()), "LocalGemm");
}

// C resized to fit, then C := alpha op(A) op(B) locally (Gemm.cpp:428-440)
template <typename T, Dist UA, Dist VA, Dist UB, Dist VB, Dist UC, Dist VC_>
void LocalGemm(Orientation orientA, Orientation orientB, T alpha, const DistMatrix<T, UA, VA>& A,
               const DistMatrix<T, UB, VB>& B, DistMatrix<T, UC, VC_>& C) {
    C.Resize(orientA == NORMAL ? A.Height() : A.Width(), orientB == NORMAL ? B.Width() : B.Height());
    LocalGemm(orientA, orientB, alpha, A, B, T(0), C);
}

// ---------------------------------------------------------------------------------------------
// Triangular rank-k family (include/El/blas_like/level3.hpp:92-126,239-252,292-313,499-536 and the
// level-1 trapezoid routines).  conjugate / ADJOINT equal the plain forms for the real types here.
// ---------------------------------------------------------------------------------------------
template <typename T> Int LocalTrrkBlocksize() { return Int(elb200_local_trrk_blocksize()); }
template <typename T> void SetLocalTrrkBlocksize(Int blocksize) { elb200_set_local_trrk_blocksize(blocksize); }

template <typename T, typename S, Dist U, Dist V>
void ScaleTrapezoid(S alpha, UpperOrLower uplo, DistMatrix<T, U, V>& A, Int offset = 0) {
    detail::Check(elb200_dm_scale_trapezoid(double(alpha), int(uplo), A.Handle(), offset), "ScaleTrapezoid");
}
template <typename T, typename S, Dist U, Dist V, Dist U2, Dist V2>
void AxpyTrapezoid(UpperOrLower uplo, S alpha, const DistMatrix<T, U, V>& X, DistMatrix<T, U2, V2>& Y, Int offset = 0) {
    detail::Check(elb200_dm_axpy_trapezoid(int(uplo), double(alpha), X.Handle(), Y.Handle(), offset), "AxpyTrapezoid");
}
template <typename T, Dist U, Dist V>
void MakeTrapezoidal(UpperOrLower uplo, DistMatrix<T, U, V>& A, Int offset = 0) {
    detail::Check(elb200_dm_make_trapezoidal(int(uplo), A.Handle(), offset), "MakeTrapezoidal");
}
template <typename T, Dist U, Dist V>
void MakeSymmetric(UpperOrLower uplo, DistMatrix<T, U, V>& A, bool /*conjugate*/ = false) {
    detail::Check(elb200_dm_make_symmetric(int(uplo), A.Handle()), "MakeSymmetric");
}
template <typename T, Dist U, Dist V>
void MakeHermitian(UpperOrLower uplo, DistMatrix<T, U, V>& A) { MakeSymmetric(uplo, A, true); }

// triangle of C := alpha op(A) op(B) + beta C  (Trrk.cpp:100-118)
template <typename T, Dist UA, Dist VA, Dist UB, Dist VB, Dist UC, Dist VC_>
void Trrk(UpperOrLower uplo, Orientation orientA, Orientation orientB, T alpha, const DistMatrix<T, UA, VA>& A,
          const DistMatrix<T, UB, VB>& B, T beta, DistMatrix<T, UC, VC_>& C) {
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
template <typename T, Dist UA, Dist VA, Dist UC, Dist VC_>
void Syrk(UpperOrLower uplo, Orientation orientation, T alpha, const DistMatrix<T, UA, VA>& A, T beta,
          DistMatrix<T, UC, VC_>& C, bool /*conjugate*/ = false) {
    detail::Check(elb200_dist_syrk(int(uplo), int(orientation), double(alpha), A.Handle(), double(beta), C.Handle()), "Syrk");
}
// C resized to n x n and zeroed first (Syrk.cpp:192-203)
template <typename T, Dist UA, Dist VA, Dist UC, Dist VC_>
void Syrk(UpperOrLower uplo, Orientation orientation, T alpha, const DistMatrix<T, UA, VA>& A, DistMatrix<T, UC, VC_>& C,
          bool conjugate = false) {
    const Int n = orientation == NORMAL ? A.Height() : A.Width();
    C.Resize(n, n);
    Zero(C);
    Syrk(uplo, orientation, alpha, A, T(0), C, conjugate);
}
template <typename T, Dist UA, Dist VA, Dist UC, Dist VC_>
void Herk(UpperOrLower uplo, Orientation orientation, T alpha, const DistMatrix<T, UA, VA>& A, T beta, DistMatrix<T, UC, VC_>& C) {
    Syrk(uplo, orientation, alpha, A, beta, C, true);
}
template <typename T, Dist UA, Dist VA, Dist UC, Dist VC_>
void Herk(UpperOrLower uplo, Orientation orientation, T alpha, const DistMatrix<T, UA, VA>& A, DistMatrix<T, UC, VC_>& C) {
    Syrk(uplo, orientation, alpha, A, C, true);
}
// C := alpha A B + beta C (LEFT) or alpha B A + beta C (RIGHT), A symmetric, `uplo` triangle referenced (Symm.cpp:56-80)
template <typename T, Dist UA, Dist VA, Dist UB, Dist VB, Dist UC, Dist VC_>
void Symm(LeftOrRight side, UpperOrLower uplo, T alpha, const DistMatrix<T, UA, VA>& A, const DistMatrix<T, UB, VB>& B, T beta,
          DistMatrix<T, UC, VC_>& C, bool /*conjugate*/ = false) {
    detail::Check(elb200_dist_symm(int(side), int(uplo), double(alpha), A.Handle(), B.Handle(), double(beta), C.Handle()), "Symm");
}
template <typename T, Dist UA, Dist VA, Dist UB, Dist VB, Dist UC, Dist VC_>
void Hemm(LeftOrRight side, UpperOrLower uplo, T alpha, const DistMatrix<T, UA, VA>& A, const DistMatrix<T, UB, VB>& B, T beta,
          DistMatrix<T, UC, VC_>& C) {
    Symm(side, uplo, alpha, A, B, beta, C, true);
}

// B := alpha op(A)^-1 B (LEFT) or alpha B op(A)^-1 (RIGHT)  (Trsm.cpp:474-499). checkIfSingular is accepted and not
// evaluated (the reference's GPU path, cuBLAS trsm, does not evaluate it either); every TrsmAlgorithm value runs the
// many-right-hand-sides schedule.
template <typename F, Dist UA, Dist VA, Dist UB, Dist VB>
void Trsm(LeftOrRight side, UpperOrLower uplo, Orientation orientation, UnitOrNonUnit diag, F alpha, const DistMatrix<F, UA, VA>& A,
          DistMatrix<F, UB, VB>& B, bool /*checkIfSingular*/ = false, TrsmAlgorithm alg = TRSM_DEFAULT) {
    detail::Check(elb200_dist_trsm(int(side), int(uplo), int(orientation), int(diag), double(alpha), A.Handle(), B.Handle(), int(alg)), "Trsm");
}
// A replicated, X distributed over the dimension that is not solved (Trsm.cpp:501-526)
template <typename F, Dist UX, Dist VX>
void LocalTrsm(LeftOrRight side, UpperOrLower uplo, Orientation orientation, UnitOrNonUnit diag, F alpha,
               const DistMatrix<F, STAR, STAR>& A, DistMatrix<F, UX, VX>& X, bool /*checkIfSingular*/ = false) {
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
template <typename T, Dist U, Dist V>
void Uniform(DistMatrix<T, U, V>& A, Int height, Int width, T center = T(0), T radius = T(1), std::uint64_t seed = 21) {
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
template <typename T, Dist U, Dist V>
double FrobeniusNorm(const DistMatrix<T, U, V>& A) {
    DistMatrix<T, STAR, STAR> G(A.Grid());
    G = A;
    std::vector<T> host(size_t(G.LocalHeight()) * size_t(G.LocalWidth()));
    G.GetLocalToHost(host.data());
    long double s = 0;
    for (T v : host) s += (long double)v * (long double)v;
    return double(std::sqrt(s));
}

}  // namespace El
