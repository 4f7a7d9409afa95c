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
enum Orientation { NORMAL, TRANSPOSE, ADJOINT };
enum GridOrder { ROW_MAJOR, COLUMN_MAJOR };
enum GemmAlgorithm {
    GEMM_DEFAULT, GEMM_SUMMA_A_MS, GEMM_SUMMA_A, GEMM_SUMMA_B_MS, GEMM_SUMMA_B, GEMM_SUMMA_C_MS, GEMM_SUMMA_C,
    GEMM_SUMMA_DOT, GEMM_CANNON
};

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
template <typename T, Dist U = MC, Dist V = MR>
class DistMatrix {
    static_assert(std::is_same<T, float>::value || std::is_same<T, double>::value, "float and double only");

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
    DistMatrix& operator=(const DistMatrix<T, U2, V2>& A) {
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
    void AlignWith(const DistMatrix<T, U2, V2>& A) { detail::Check(elb200_dm_align_with(h_, A.Handle()), "AlignWith"); }
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
