// Host engine: Grid, type-erased-by-template DistMat<T>, redistribution, contraction and the
// SUMMA drivers. This is the C++ that sits between the El:: facade (include/El.hpp) / the handle
// C ABI (elb200_grid_*, elb200_dm_*, elb200_dist_gemm) and the device C ABI (elb200_gemm_*,
// elb200_lattice_*, ...). One process per GPU; NCCL communicators for the grid's MC / MR / VC
// sub-communicators (reference: src/core/Grid.cpp:95-223 builds them with MPI_Cart_sub/Comm_split).
#pragma once
#include <cuda_runtime.h>
#include <nccl.h>

#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "peer.hpp"
#include "plan.hpp"

namespace elb {

struct EngineError : std::runtime_error {
    int status;
    EngineError(int st, const std::string& what) : std::runtime_error(what), status(st) {}
};
void throw_status(int status, const char* where);
void cuda_check(cudaError_t e, const char* where);

// RAII event: record on one stream, make another wait. Released on every exit of the schedule that owns it,
// a throwing redistribution in the middle of a pipeline included.
struct Ev {
    cudaEvent_t e = nullptr;
    Ev() { cuda_check(cudaEventCreateWithFlags(&e, cudaEventDisableTiming), "cudaEventCreate"); }
    ~Ev() { if (e) cudaEventDestroy(e); }
    Ev(const Ev&) = delete;
    Ev& operator=(const Ev&) = delete;
    void record(cudaStream_t s) { cuda_check(cudaEventRecord(e, s), "cudaEventRecord"); }
    void block(cudaStream_t s) { cuda_check(cudaStreamWaitEvent(s, e, 0), "cudaStreamWaitEvent"); }
};
void nccl_check(ncclResult_t r, const char* where);

// IEEE binary16 as the engine's HOST-side element / scalar type for half matrices (the reference's
// gpu_half_type = __half). Storage is the 16 bits; the only host arithmetic the engine needs on a scalar is construction
// from 0 / 1 / a double and comparison with zero, both through float. Device code sees the same bits as __half.
struct Half {
    uint16_t bits = 0;
    Half() = default;
    Half(float f) : bits(from_float(f)) {}
    Half(double d) : bits(from_float(float(d))) {}
    Half(int i) : bits(from_float(float(i))) {}
    operator float() const { return to_float(bits); }
    static uint16_t from_float(float f);
    static float to_float(uint16_t h);
};

// Mirrors El::GemmAlgorithm (reference include/El/blas_like/level3.hpp:22-35), same order and values.
enum GemmAlg : int {
    GEMM_DEFAULT = 0, GEMM_SUMMA_A_MS = 1, GEMM_SUMMA_A = 2, GEMM_SUMMA_B_MS = 3, GEMM_SUMMA_B = 4,
    GEMM_SUMMA_C_MS = 5, GEMM_SUMMA_C = 6, GEMM_SUMMA_DOT = 7, GEMM_CANNON = 8
};

class Grid {
public:
    // nccl_id == nullptr: a grid without communicators (p must be 1, or plan-only use).
    Grid(int height, int width, int vc_rank, const ncclUniqueId* nccl_id, int device);
    ~Grid();
    // Collective teardown (peer arena, NCCL communicators); idempotent. The streams stay usable,
    // so matrices of this grid can still be destroyed afterwards.
    void Shutdown();
    Grid(const Grid&) = delete;
    Grid& operator=(const Grid&) = delete;

    GridShape shape;
    GridCoord me;
    int vc_rank = 0;
    int device = 0;
    cudaStream_t stream = nullptr;       // compute stream (the reference's SyncInfo of C)
    cudaStream_t comm_stream = nullptr;  // panel prefetch stream (overlap with the local GEMM)
    cudaStream_t h2d_stream = nullptr, d2h_stream = nullptr;  // host-operand streaming (GemmHost)
    ncclComm_t world = nullptr, mc_comm = nullptr, mr_comm = nullptr;
    ncclComm_t vr_comm = nullptr;  // all ranks in VR (row-major) order; created on first use by elb200_grid_comm
    // NVLink peer-memory staging (zero-SM AllGather); nullptr -> NCCL is used instead.
    PeerArena* arena = nullptr;
    // VC ranks of the MC (grid column) / MR (grid row) / VC (world) communicator this rank belongs
    // to, in communicator order, and this rank's index in it.
    std::vector<int> comm_members(Dist d) const;
    int comm_index(Dist d) const { return d == MC ? me.mc : d == MR ? me.mr : vc_rank; }
    static int arena_channel(Dist d) { return d == MC ? 0 : d == MR ? 1 : 2; }

    // NCCL fallbacks issued from different streams (the pipelines put redistributions on the comm stream) are
    // serialised in issue order: NCCL does not allow concurrent operations on one communicator, and the
    // MC / MR / VC communicators may be the same object. Bracket every NCCL call (group) with these.
    void NcclBegin(cudaStream_t s);
    void NcclEnd(cudaStream_t s);
    cudaEvent_t nccl_done = nullptr;
    bool nccl_recorded = false;

    int size() const { return shape.size(); }
    ncclComm_t comm_for(Dist d) const { return d == MC ? mc_comm : d == MR ? mr_comm : world; }
    // Algorithmic blocksize stack (reference src/blas_like/blocksizes.cpp:35-67; default 128,
    // src/core/environment.cpp:313-315). Process-global like the reference's.
    static int Blocksize();
    static void SetBlocksize(int nb);
    static void PushBlocksize(int nb);
    static void PopBlocksize();
    // LocalTrrk recursion cut-off (reference LocalTrrkBlocksize<T> / SetLocalTrrkBlocksize<T>,
    // src/blas_like/blocksizes.cpp:82-87); default 2048 here, see engine_level3.cpp.
    static int64_t LocalTrrkBlocksize();
    static void SetLocalTrrkBlocksize(int64_t nb);
    // Internal SUMMA_C/A/B panel width: several algorithmic blocks are moved and multiplied per
    // step so that one pass over the local C amortises >= this many k.
    static int64_t PanelWidth(int elem_bytes);
};

template <typename T>
class DistMat {
public:
    DistMat(Grid& g, Dist col, Dist row, cudaStream_t s = nullptr);
    ~DistMat();
    DistMat(const DistMat&) = delete;
    DistMat& operator=(const DistMat&) = delete;
    DistMat(DistMat&& o) noexcept;

    Grid* grid;
    DistSpec spec;
    bool col_constrained = false, row_constrained = false;
    T* buf = nullptr;
    int64_t ldim = 1;
    bool owner = true;
    int64_t capacity = 0;  // elements owned
    cudaStream_t stream;

    int64_t lh() const { return local_height(grid->shape, spec, grid->me); }
    int64_t lw() const { return local_width(grid->shape, spec, grid->me); }
    int col_stride() const { return dist_stride(grid->shape, spec.col); }
    int row_stride() const { return dist_stride(grid->shape, spec.row); }
    int64_t col_shift() const { return Shift(dist_rank(grid->shape, spec.col, grid->me), spec.col_align, col_stride()); }
    int64_t row_shift() const { return Shift(dist_rank(grid->shape, spec.row, grid->me), spec.row_align, row_stride()); }

    void Align(int64_t col_align, int64_t row_align, bool constrain = true);
    void AlignCols(int64_t col_align, bool constrain = true);
    void AlignRows(int64_t row_align, bool constrain = true);
    // Adopt whichever alignments of `other` are meaningful for this distribution
    // (reference ElementalMatrix::AlignWith, src/core/DistMatrix/ElementMatrix.cpp).
    void AlignWith(const DistSpec& other, bool constrain = true);
    void Resize(int64_t height, int64_t width);
    // [CIRC,CIRC]: which VC rank holds the matrix (reference SetRoot); keeps the contents only if
    // the root does not change.
    void SetRoot(int root);
    // local matrix <-> host memory (column-major, ld >= local height), async on `stream`
    void WriteLocal(const T* host, int64_t ld);
    void ReadLocal(T* host, int64_t ld) const;
    void Attach(int64_t height, int64_t width, int64_t col_align, int64_t row_align, T* buffer, int64_t ld);
    // Zero-copy view of rows [i0,i1) x cols [j0,j1) (reference View/LockedView + operator()(IR,IR)).
    DistMat View(int64_t i0, int64_t i1, int64_t j0, int64_t j1) const;
};

// ---- distributed BLAS-1 (reference include/El/blas_like/level1/*.hpp) -------------------------
template <typename T> void Copy(const DistMat<T>& A, DistMat<T>& B);                // B = A
template <typename T> void TransposeCopy(const DistMat<T>& A, DistMat<T>& B);       // B = A^T
template <typename T> void AxpyContract(T alpha, const DistMat<T>& A, DistMat<T>& B);
template <typename T> void Contract(const DistMat<T>& A, DistMat<T>& B);
template <typename T> void TransposeAxpyContract(T alpha, const DistMat<T>& A, DistMat<T>& B);
template <typename T> void Axpy(T alpha, const DistMat<T>& X, DistMat<T>& Y);       // same dist
template <typename T> void Scale(T alpha, DistMat<T>& A);
template <typename T> void Zero(DistMat<T>& A);

// ---- BLAS-3 ------------------------------------------------------------------------------------
template <typename T>
void LocalGemm(int ta, int tb, T alpha, const DistMat<T>& A, const DistMat<T>& B, T beta, DistMat<T>& C);
template <typename T>
void Gemm(int ta, int tb, T alpha, const DistMat<T>& A, const DistMat<T>& B, T beta, DistMat<T>& C, int alg);
// C = alpha op(A) op(B) + beta C with every operand in HOST memory: the local parts of zero-aligned
// [MC,MR] matrices (column-major, leading dimensions lda/ldb/ldc) stream through the GPU — A stays
// resident, column blocks of B and C are uploaded, gathered, multiplied and downloaded in a
// pipeline over four streams. Returns when C is complete in host memory.
template <typename T>
void GemmHost(Grid& g, int ta, int tb, int64_t m, int64_t n, int64_t k, T alpha, const T* A, int64_t lda,
              const T* B, int64_t ldb, T beta, T* C, int64_t ldc);
// ---- triangular rank-k family (reference src/blas_like/level3/{Trrk,Syrk,Herk,Symm,Hemm}.cpp) -----
// uplo: 0 = LOWER, 1 = UPPER (El::UpperOrLower); side: 0 = LEFT, 1 = RIGHT (El::LeftOrRight).
template <typename T> void ScaleTrapezoid(T alpha, int uplo, DistMat<T>& A, int64_t offset);
template <typename T> void AxpyTrapezoid(int uplo, T alpha, const DistMat<T>& X, DistMat<T>& Y, int64_t offset);
template <typename T> void MakeTrapezoidal(int uplo, DistMat<T>& A, int64_t offset);
template <typename T> void MakeSymmetric(int uplo, DistMat<T>& A);
template <typename T>
void LocalTrrk(int uplo, int ta, int tb, T alpha, const DistMat<T>& A, const DistMat<T>& B, T beta, DistMat<T>& C);
template <typename T>
void Trrk(int uplo, int ta, int tb, T alpha, const DistMat<T>& A, const DistMat<T>& B, T beta, DistMat<T>& C);
template <typename T> void Syrk(int uplo, int trans, T alpha, const DistMat<T>& A, T beta, DistMat<T>& C);
template <typename T>
void Symm(int side, int uplo, T alpha, const DistMat<T>& A, const DistMat<T>& B, T beta, DistMat<T>& C);
// B := alpha op(A)^-1 B (side LEFT) or alpha B op(A)^-1 (RIGHT); diag: 0 = NON_UNIT, 1 = UNIT
// (reference src/blas_like/level3/Trsm.cpp:129-440, 501-526).
template <typename T> void LocalTrsm(int side, int uplo, int trans, int diag, T alpha, const DistMat<T>& A, DistMat<T>& X);
template <typename T> void Trsm(int side, int uplo, int trans, int diag, T alpha, const DistMat<T>& A, DistMat<T>& B);
// src redistributed like tmp's distribution, aligned with `like`; returns src itself when that
// redistribution is the identity on every rank (engine_gemm.cpp).
template <typename T> const DistMat<T>& PanelAs(const DistMat<T>& src, DistMat<T>& tmp, const DistSpec& like);
// Which SUMMA variant GEMM_DEFAULT resolves to (reference Gemm/NN.hpp:585-600 and twins).
int ResolveGemmAlg(int ta, int tb, int64_t m, int64_t n, int64_t k, int alg);
void GemmProfileEnable(bool on);
void GemmProfileCollect(double* ms, double* flop, int* launches);

bool implied_align(Dist mine, Dist theirs, int64_t a, const GridShape& g, int64_t* out);
// Distribution spec B ends up with after `B = A` (dims from A; unconstrained alignments adopt A's).
DistSpec ResolveCopySpec(const GridShape& g, const DistSpec& A, DistSpec B, bool col_constrained,
                         bool row_constrained);

// device-op dispatch by element type (thin wrappers over the C ABI; throw on error)
template <typename T> struct Dev;

}  // namespace elb
