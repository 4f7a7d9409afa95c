/* elb200.h — C ABI of libelb200.so: the B200 (sm_100a) device layer underneath the
 * Hydrogen/Elemental `El::Gemm` hot path.
 *
 * Every entry point below is what the reference binds at the same place in its call stack; the
 * reference interface each one replaces is cited as file:line relative to the reference tree.
 * Conventions (same as the reference's gpu_blas layer):
 *   - column-major matrices, leading dimensions and strides in ELEMENTS (not bytes);
 *   - alpha/beta by value on the host (reference: CUBLAS_POINTER_MODE_HOST,
 *     src/hydrogen/device/cuBLAS.cpp:66-68);
 *   - every call is asynchronous on the caller's stream (`stream` is a cudaStream_t passed as
 *     void*; the reference passes SyncInfo<Device::GPU>::Stream()); no hidden synchronisation;
 *   - the return value is an elb200_status; nothing throws across this boundary. The C++ facade
 *     (include/El.hpp) converts non-zero statuses into the reference's exception types.
 *   - there is NO CPU fallback: without a CUDA device every compute entry point returns
 *     ELB200_ERR_CUDA.
 */
#ifndef ELB200_H_
#define ELB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum elb200_status {
    ELB200_OK = 0,
    ELB200_ERR_CUDA = 1,      /* a CUDA runtime call or kernel launch failed            */
    ELB200_ERR_ARG = 2,       /* non-conformal dimensions / null pointer / bad enum     */
    ELB200_ERR_TENSORMAP = 3, /* cuTensorMapEncodeTiled rejected the operand            */
    ELB200_ERR_NCCL = 4,      /* an NCCL call failed                                    */
    ELB200_ERR_STATE = 5      /* library used before elb200 comm/grid initialisation    */
} elb200_status;

/* 'N' = 0, 'T' = 1, 'C' = 2 (== 'T' for real types); mirrors El::Orientation
 * (include/El/core/types.hpp) and hydrogen::TransposeMode. */
typedef enum elb200_op { ELB200_OP_N = 0, ELB200_OP_T = 1, ELB200_OP_C = 2 } elb200_op;

const char* elb200_status_string(int status);
/* Text of the last CUDA/NCCL error seen by the calling thread ("" if none). */
const char* elb200_last_error(void);
/* Compile-time facts: "sm_100a", kernel list. */
const char* elb200_build_info(void);
/* Number of CUDA kernels this library has launched in this process (all entry points). */
unsigned long long elb200_launch_count(void);

/* ------------------------------------------------------------------------------------------
 * BLAS-3: local GEMM.  C(m x n) = alpha * op(A) * op(B) + beta * C
 * Replaces gpu_blas::Gemm -> cublas{S,D}gemm
 *   (include/hydrogen/blas/GPU_BLAS_impl.hpp:417-444,1036-1051;
 *    src/hydrogen/device/cuBLAS_API.cpp:282-300), called from Gemm_impl<Device::GPU>
 *   (src/blas_like/level3/Gemm.cpp:163-186).
 * f32: tcgen05.mma kind::tf32 with an in-kernel 3xTF32 split (FP32-level accuracy), TMEM
 *      accumulators, TMA operand loads.  f64: mma.sync m8n8k4 f64 (DMMA) + cp.async pipeline.
 * k == 0 or alpha == 0 degenerate to C *= beta, like the reference (Gemm.cpp:240-247).
 * ---------------------------------------------------------------------------------------- */
int elb200_gemm_f32(int transA, int transB, int64_t m, int64_t n, int64_t k, float alpha,
                    const float* A, int64_t lda, const float* B, int64_t ldb, float beta, float* C,
                    int64_t ldc, void* stream);
int elb200_gemm_f64(int transA, int transB, int64_t m, int64_t n, int64_t k, double alpha,
                    const double* A, int64_t lda, const double* B, int64_t ldb, double beta,
                    double* C, int64_t ldc, void* stream);
/* IEEE half operands and result (device pointers to __half, passed as void* so that this header stays plain C), alpha /
 * beta and the accumulation in FP32, one rounding to half in the epilogue: tcgen05.mma kind::f16 — the native tensor-core
 * type, one UMMA per k-slice where the fp32 path needs three. Replaces the gpu_half_type instantiation of gpu_blas::Gemm
 * -> cublasHgemm / cublasGemmEx (src/hydrogen/device/cuBLAS_API.cpp:417 ADD_GEMM_IMPL(__half, H);
 * src/blas_like/level3/Gemm.cpp:455-462,505). SURVEY.md 8(f) N3. Operands must be 16-byte aligned with leading
 * dimensions that are multiples of 8 halfs (else ELB200_ERR_ARG). */
int elb200_gemm_f16(int transA, int transB, int64_t m, int64_t n, int64_t k, float alpha,
                    const void* A, int64_t lda, const void* B, int64_t ldb, float beta, void* C,
                    int64_t ldc, void* stream);
/* Tuning/diagnostic variant: cta_group = 1 (one SM per tile) or 2 (CTA pair, cta_group::2). */
int elb200_gemm_f32_ex(int transA, int transB, int64_t m, int64_t n, int64_t k, float alpha,
                       const float* A, int64_t lda, const float* B, int64_t ldb, float beta,
                       float* C, int64_t ldc, int cta_group, void* stream);
/* Allocation-free variant of elb200_gemm_f32. It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it). Same arguments, same results as
 * elb200_gemm_f32. Operands whose base or leading dimension is not 16-byte aligned (what the TMA-fed
 * kernel needs) are staged into a fixed device array owned by the library instead of stream-ordered
 * temporaries: the call makes no cudaMalloc* / cudaFree* and no allocating library call.
 * Limit: a misaligned operand may occupy at most ELB200_NOALLOC_MAX_ELEMS floats after its leading
 * dimension is rounded up to 4 (16 MiB per operand; 32 MiB of device memory per GPU for the life of the
 * process); larger ones are rejected with ELB200_ERR_ARG before anything is enqueued. Aligned operands
 * of any size pass straight through. All calls share the one array: calls that may overlap on the GPU
 * (different streams with no ordering between them) race on it — keep them on one stream or order them
 * with events. */
#define ELB200_NOALLOC_MAX_ELEMS (4 * 1024 * 1024)
int elb200_gemm_f32_noalloc(int transA, int transB, int64_t m, int64_t n, int64_t k, float alpha,
                            const float* A, int64_t lda, const float* B, int64_t ldb, float beta,
                            float* C, int64_t ldc, void* stream);

/* ------------------------------------------------------------------------------------------
 * BLAS-1 entrywise kernels (HBM-bound, 128-bit vectorised where strides allow).
 * ---------------------------------------------------------------------------------------- */
/* dst(i,j) = src(i,j) with general 2-D strides.
 * Replaces hydrogen::Copy_GPU_impl 2-D (include/hydrogen/blas/gpu/Copy.hpp:100-104,
 * src/hydrogen/blas/gpu/Copy.cu:26-73,156-169) and gpu::Copy2DIntraDevice
 * (include/hydrogen/device/gpu/cuda/CUDACopy.hpp:70-82) as used by copy::util::InterleaveMatrix
 * (include/El/blas_like/level1/Copy/util.hpp:245-265). */
int elb200_copy2d_f32(int64_t rows, int64_t cols, const float* src, int64_t src_row_stride,
                      int64_t src_col_stride, float* dst, int64_t dst_row_stride,
                      int64_t dst_col_stride, void* stream);
int elb200_copy2d_f64(int64_t rows, int64_t cols, const double* src, int64_t src_row_stride,
                      int64_t src_col_stride, double* dst, int64_t dst_row_stride,
                      int64_t dst_col_stride, void* stream);

/* The half (__half) instantiations of the entrywise kernels (the reference instantiates Copy / Axpy / Scale / Fill /
 * Transpose for gpu_half_type as well, src/hydrogen/blas/gpu/{Copy,Axpy,Scale,Fill,Transpose}.cu); scalars in FP32. */
int elb200_copy2d_f16(int64_t rows, int64_t cols, const void* src, int64_t src_row_stride, int64_t src_col_stride,
                      void* dst, int64_t dst_row_stride, int64_t dst_col_stride, void* stream);
int elb200_axpy2d_f16(int64_t rows, int64_t cols, float alpha, const void* X, int64_t x_row_stride, int64_t x_col_stride,
                      void* Y, int64_t y_row_stride, int64_t y_col_stride, void* stream);
int elb200_scale2d_f16(int64_t rows, int64_t cols, float alpha, void* A, int64_t lda, void* stream);
int elb200_fill2d_f16(int64_t rows, int64_t cols, float value, void* A, int64_t lda, void* stream);
int elb200_transpose_f16(int64_t rows, int64_t cols, const void* A, int64_t lda, void* B, int64_t ldb, void* stream);
int elb200_axpy2d_transpose_f16(int64_t rows, int64_t cols, float alpha, const void* A, int64_t lda, void* B, int64_t ldb,
                                void* stream);

/* B(cols x rows, ldb) = A(rows x cols, lda)^T.
 * Replaces Transpose_GPU_impl / Copy2DImpl(transA) -> cublas?geam
 * (include/hydrogen/blas/gpu/Transpose.hpp:40-44; GPU_BLAS_impl.hpp:178-198), reached from
 * El::Transpose(Matrix<T,GPU>) (include/El/blas_like/level1/Transpose.hpp:161-181). */
int elb200_transpose_f32(int64_t rows, int64_t cols, const float* A, int64_t lda, float* B,
                         int64_t ldb, void* stream);
int elb200_transpose_f64(int64_t rows, int64_t cols, const double* A, int64_t lda, double* B,
                         int64_t ldb, void* stream);

/* Y(i,j) += alpha * X(i,j), general 2-D strides.
 * Replaces Axpy_GPU_impl 2-D (include/hydrogen/blas/gpu/Axpy.hpp:51-55,
 * src/hydrogen/blas/gpu/Axpy.cu:84-150) = axpy::util::InterleaveMatrixUpdate
 * (include/El/blas_like/level1/Axpy/util.hpp:40-52) and Axpy2DImpl -> geam
 * (GPU_BLAS_impl.hpp:110-132). */
int elb200_axpy2d_f32(int64_t rows, int64_t cols, float alpha, const float* X,
                      int64_t x_row_stride, int64_t x_col_stride, float* Y, int64_t y_row_stride,
                      int64_t y_col_stride, void* stream);
int elb200_axpy2d_f64(int64_t rows, int64_t cols, double alpha, const double* X,
                      int64_t x_row_stride, int64_t x_col_stride, double* Y, int64_t y_row_stride,
                      int64_t y_col_stride, void* stream);

/* B(rows x cols, ldb) += alpha * A(cols x rows, lda)^T.
 * Replaces Axpy_GPU_impl(transpA,...) / Axpy2DImplTranspose -> geam
 * (include/hydrogen/blas/gpu/Axpy.hpp:92-96; GPU_BLAS_impl.hpp:134-158), used by
 * El::TransposeAxpy (include/El/blas_like/level1/TransposeAxpy.hpp). */
int elb200_axpy2d_transpose_f32(int64_t rows, int64_t cols, float alpha, const float* A,
                                int64_t lda, float* B, int64_t ldb, void* stream);
int elb200_axpy2d_transpose_f64(int64_t rows, int64_t cols, double alpha, const double* A,
                                int64_t lda, double* B, int64_t ldb, void* stream);

/* A(i,j) *= alpha.  alpha == 0 writes exact zeros (no NaN propagation), like El::Scale -> Zero
 * (include/El/blas_like/level1/Scale.hpp:17-30). Replaces Scale_GPU_impl 2-D / Scale2DImpl -> geam
 * (include/hydrogen/blas/gpu/Scale.hpp:78-82; GPU_BLAS_impl.hpp:254-273). */
int elb200_scale2d_f32(int64_t rows, int64_t cols, float alpha, float* A, int64_t lda,
                       void* stream);
int elb200_scale2d_f64(int64_t rows, int64_t cols, double alpha, double* A, int64_t lda,
                       void* stream);

/* A(i,j) = value.  Replaces Fill_GPU_impl (include/hydrogen/blas/gpu/Fill.hpp:23-25,
 * src/hydrogen/blas/gpu/Fill.cu:19-78) as used by simple_buffer(size, value) and El::Zero. */
int elb200_fill2d_f32(int64_t rows, int64_t cols, float value, float* A, int64_t lda,
                      void* stream);
int elb200_fill2d_f64(int64_t rows, int64_t cols, double value, double* A, int64_t lda,
                      void* stream);

/* ------------------------------------------------------------------------------------------
 * Redistribution pack/unpack: ONE launch moves every per-peer portion of an element-cyclic
 * redistribution between a local column-major matrix and a staging buffer.
 *
 * Each segment s describes a 2-D lattice of the local matrix
 *     M(row0 + i*row_step, col0 + j*col_step),  0 <= i < rows, 0 <= j < cols
 * stored packed (column-major, leading dimension = rows) at buf + buf_offset.
 * This single table-driven kernel replaces the reference's loops of cudaMemcpy2DAsync /
 * copy_2d_kernel launches in copy::util::{RowStridedPack,RowStridedUnpack,ColStridedPack,
 * ColStridedUnpack,PartialRow/ColStridedPack/Unpack,StridedPack,StridedUnpack}
 * (include/El/blas_like/level1/Copy/util.hpp:232-355,359-417,460-552,667-718).
 * ---------------------------------------------------------------------------------------- */
typedef struct elb200_segment {
    int64_t row0, row_step, rows; /* lattice rows of the local matrix        */
    int64_t col0, col_step, cols; /* lattice columns of the local matrix     */
    int64_t buf_offset;           /* element offset of the packed block      */
} elb200_segment;

#define ELB200_MAX_SEGMENTS 64

/* buf[seg] <- M lattice   (mode 0)
 * M lattice <- buf[seg]   (mode 1)
 * M lattice += alpha*buf  (mode 2; the axpy-unpack after ReduceScatter,
 *                          include/El/blas_like/level1/AxpyContract.hpp:334-337) */
typedef enum elb200_pack_mode {
    ELB200_PACK = 0,
    ELB200_UNPACK = 1,
    ELB200_UNPACK_AXPY = 2
} elb200_pack_mode;

int elb200_lattice_f32(int mode, int nseg, const elb200_segment* segs, float* M, int64_t ldm,
                       float* buf, float alpha, void* stream);
int elb200_lattice_f64(int mode, int nseg, const elb200_segment* segs, double* M, int64_t ldm,
                       double* buf, double alpha, void* stream);

/* half instantiations of the two entries above / below (sums accumulate in FP32) */
int elb200_lattice_f16(int mode, int nseg, const elb200_segment* segs, void* M, int64_t ldm, void* buf, float alpha,
                       void* stream);
int elb200_peer_reduce_axpy_f16(int nsrc, const void* const* srcs, int64_t src_offset, const elb200_segment* seg, void* M,
                                int64_t ldm, float alpha, void* stream);

/* M lattice += alpha * (srcs[0][off + e] + srcs[1][off + e] + ... ), e running over the packed
 * rows x cols block of `seg`: the receiver side of a ReduceScatter, fused with the axpy-unpack.
 * srcs are device pointers this GPU can load from — its own staging buffer and its peers' buffers
 * mapped over NVLink (CUDA IPC / peer access); the sum runs in array order on every caller.
 * Replaces mpi::ReduceScatter<GPU> + the update loop in axpy_contract::{RowScatter,ColScatter,
 * Scatter} (include/El/blas_like/level1/AxpyContract.hpp:318-337, 196-216, 455-476;
 * src/core/imports/mpi/ReduceScatter.hpp:36-52). nsrc <= ELB200_MAX_SEGMENTS. */
int elb200_peer_reduce_axpy_f32(int nsrc, const float* const* srcs, int64_t src_offset,
                                const elb200_segment* seg, float* M, int64_t ldm, float alpha,
                                void* stream);
int elb200_peer_reduce_axpy_f64(int nsrc, const double* const* srcs, int64_t src_offset,
                                const elb200_segment* seg, double* M, int64_t ldm, double alpha,
                                void* stream);


/* Fused small-message collectives over NVLink peer memory: ONE kernel per rank does the whole collective — waits for the
 * members' acks of the slot's previous use, moves the data with peer stores / loads, raises an arrival flag, waits for the
 * other members' flags, puts the result into the destination matrix and acknowledges — where the copy-engine forms take
 * 20-40 stream operations (57-122 us at 1 MiB on 2 x 4 B200s). The descriptor is filled by the engine's PeerArena (slots and
 * flags are device addresses mapped in this process, some of them peer memory); `counters` are two zeroed words of local
 * device memory that the kernels leave zeroed. Collective: every member launches the same operation with the same seq.
 *   elb200_peer_allgather       PUSH: my packed block into every member's slot, then UNPACK my slot into M through `segs`.
 *                               Replaces pack + mpi::AllGather + unpack of copy::RowAllGather / ColAllGather
 *                               (include/El/blas_like/level1/Copy/RowAllGather.hpp:53-75; src/core/imports/mpi/AllGather.hpp:8-21).
 *   elb200_peer_exchange        PULL: PACK `sends` (offsets from MY slot base) out of A, then UNPACK `recvs` into B, recvs[q]
 *                               read from the slot of member recv_member[q] at recvs[q].buf_offset (where that member packed
 *                               its message for me). Replaces pack + mpi::AllToAll / SendRecv + unpack of
 *                               copy::ColAllToAllDemote / Promote / Exchange (Copy/ColAllToAllDemote.hpp:17-80, Copy/Exchange.hpp;
 *                               src/core/imports/mpi/AllToAll.hpp, SendRecv.hpp).
 *   elb200_peer_reduce_scatter  PULL: PACK `packs` (member i's portion at i * portion) out of A, then B-lattice(unpack) +=
 *                               alpha * (sum over the members' slots, in member order, of the portion at my_index * portion).
 *                               Same arithmetic and order as elb200_peer_reduce_axpy. Replaces zero-fill + pack +
 *                               mpi::ReduceScatter + InterleaveMatrixUpdate of AxpyContract (level1/AxpyContract.hpp:318-337;
 *                               src/core/imports/mpi/ReduceScatter.hpp:36-52).
 * The first two only move data: elem_bytes is 2, 4 or 8. */
typedef struct elb200_peer_gather {
    int nmembers, my_index;
    int64_t portion;                         /* elements per member block inside a slot (allgather, reduce_scatter) */
    void* slots[ELB200_MAX_SEGMENTS];        /* slots[i]: base of the slot in member i's arena (slots[my_index]: mine) */
    uint32_t* arrive_out[ELB200_MAX_SEGMENTS]; /* in member i's flag block: "the data of <me> is in place"                   */
    uint32_t* arrive_in[ELB200_MAX_SEGMENTS];  /* local: "member i's data is in place"; [my_index]: my own CTAs' stores are  */
    uint32_t* ack_in[ELB200_MAX_SEGMENTS];     /* local: "member i has finished reading the previous contents"               */
    uint32_t* ack_out[ELB200_MAX_SEGMENTS];    /* in member i's flag block: "<me> has finished reading"                      */
    uint32_t seq, seq_prev;
    unsigned int* counters;
} elb200_peer_gather;
int elb200_peer_allgather(const elb200_peer_gather* pg, int elem_bytes, const void* send, int64_t send_elems, int nseg,
                          const elb200_segment* segs, void* M, int64_t ldm, void* stream);
int elb200_peer_exchange(const elb200_peer_gather* pg, int elem_bytes, int nsend, const elb200_segment* sends, const void* A,
                         int64_t lda, int nrecv, const elb200_segment* recvs, const int* recv_member, void* B, int64_t ldb,
                         void* stream);
int elb200_peer_reduce_scatter_f32(const elb200_peer_gather* pg, int npack, const elb200_segment* packs, const float* A, int64_t lda,
                                   const elb200_segment* unpack, float* B, int64_t ldb, float alpha, void* stream);
int elb200_peer_reduce_scatter_f64(const elb200_peer_gather* pg, int npack, const elb200_segment* packs, const double* A, int64_t lda,
                                   const elb200_segment* unpack, double* B, int64_t ldb, double alpha, void* stream);

/* Trapezoid-masked entrywise ops on the LOCAL part of an element-cyclic matrix. Local entry (i,j) of
 * the rows x cols column-major block Y is global entry (row0 + i*row_step, col0 + j*col_step); with
 * uplo = ELB200_LOWER the trapezoid is { gi >= gj - offset }, with ELB200_UPPER { gi <= gj - offset }
 * (the reference's convention: offset 0 includes the diagonal, UPPER with offset 1 / LOWER with
 * offset -1 are the strict triangles).
 *   op ELB200_TRAP_SCALE: Y *= alpha inside the trapezoid (alpha == 0 writes exact zeros)
 *   op ELB200_TRAP_AXPY : Y += alpha * X inside the trapezoid (X: same local shape, ldx)
 *   op ELB200_TRAP_MAKE : Y = 0 OUTSIDE the trapezoid
 * Replaces the per-column host loops of ScaleTrapezoid, AxpyTrapezoid / LocalAxpyTrapezoid and
 * MakeTrapezoidal (include/El/blas_like/level1/ScaleTrapezoid.hpp:49-96, AxpyTrapezoid.hpp,
 * MakeTrapezoidal.hpp:45-86) as used by LocalTrrk (src/blas_like/level3/Trrk/Local.hpp:350-396),
 * Syrk (src/blas_like/level3/Syrk.cpp:172-190) and MakeSymmetric (level1/MakeSymmetric.hpp:56-73). */
typedef enum elb200_uplo { ELB200_LOWER = 0, ELB200_UPPER = 1 } elb200_uplo;   /* El::UpperOrLower */
typedef enum elb200_side { ELB200_LEFT = 0, ELB200_RIGHT = 1 } elb200_side;    /* El::LeftOrRight  */
typedef enum elb200_trap_op { ELB200_TRAP_SCALE = 0, ELB200_TRAP_AXPY = 1, ELB200_TRAP_MAKE = 2 } elb200_trap_op;
int elb200_trapezoid_f32(int op, int uplo, int64_t offset, int64_t rows, int64_t cols, int64_t row0,
                         int64_t row_step, int64_t col0, int64_t col_step, float alpha, const float* X,
                         int64_t ldx, float* Y, int64_t ldy, void* stream);
int elb200_trapezoid_f64(int op, int uplo, int64_t offset, int64_t rows, int64_t cols, int64_t row0,
                         int64_t row_step, int64_t col0, int64_t col_step, double alpha, const double* X,
                         int64_t ldx, double* Y, int64_t ldy, void* stream);

/* Local triangular solve with many right-hand sides, in place:
 *   side ELB200_LEFT : op(A) X = alpha B, A is m x m      side ELB200_RIGHT: X op(A) = alpha B, A is n x n
 * B (m x n, column-major) is overwritten with X; only the `uplo` triangle of A is referenced; diag = 1
 * (El::UNIT) takes the diagonal as ones without reading it.  Replaces gpu_blas::Trsm -> cublas{S,D}trsm
 * (include/hydrogen/blas/GPU_BLAS_impl.hpp; src/blas_like/level3/Trsm.cpp:26-50), reached from
 * Trsm(Matrix<F,GPU>) and LocalTrsm (Trsm.cpp:501-526).  Recursive blocking: off-diagonal updates are
 * elb200_gemm_* calls, 32 x 32 diagonal blocks are solved by trsm_base_kernel. */
typedef enum elb200_diag { ELB200_NON_UNIT = 0, ELB200_UNIT = 1 } elb200_diag;  /* El::UnitOrNonUnit */
int elb200_trsm_f32(int side, int uplo, int trans, int diag, int64_t m, int64_t n, float alpha,
                    const float* A, int64_t lda, float* B, int64_t ldb, void* stream);
int elb200_trsm_f64(int side, int uplo, int trans, int diag, int64_t m, int64_t n, double alpha,
                    const double* A, int64_t lda, double* B, int64_t ldb, void* stream);

/* ------------------------------------------------------------------------------------------
 * Redistribution planner (pure host code; callable without a GPU).
 * For an r x c grid (COLUMN_MAJOR: VC rank = mc + r*mr, reference src/core/Grid.cpp:147-148) and
 * two element-cyclic distribution specs, returns what THIS rank packs, sends, receives and
 * unpacks for B = A (elb200_plan_copy) or B += sum-over-replicas(A) (elb200_plan_contract).
 * Distributions: 0=MC 1=MR 2=VC 3=VR 4=STAR 5=CIRC (reference El::Dist, include/El/core/types.hpp).
 * CIRC is valid only as [CIRC,CIRC]: the whole matrix on VC rank `root`
 * (src/core/DistMatrix/ElementMatrix/CIRC_CIRC.cpp; copy::Gather / copy::Scatter).
 * ---------------------------------------------------------------------------------------- */
typedef struct elb200_dist_spec {
    int col_dist, row_dist;
    int64_t col_align, row_align;
    int64_t height, width;
    int root; /* [CIRC,CIRC] only; 0 otherwise */
} elb200_dist_spec;

typedef struct elb200_message {
    int peer;   /* VC rank of the other side */
    int valid;  /* 0: slot unused */
    elb200_segment seg;
} elb200_message;

typedef struct elb200_copy_plan {
    int kind;       /* 0 local, 1 allgather, 2 grouped send/recv */
    int comm_dist;  /* allgather: communicator (0=MC column comm, 1=MR row comm, 2=VC world) */
    int64_t portion;
    int nsend, nrecv, has_self;
    int64_t send_elems, recv_elems;
    elb200_message self_send, self_recv;
    elb200_message sends[ELB200_MAX_SEGMENTS], recvs[ELB200_MAX_SEGMENTS];
} elb200_copy_plan;

typedef struct elb200_contract_plan {
    int valid, comm_dist, group, my_index;
    int64_t portion;
    int npack;
    elb200_message packs[ELB200_MAX_SEGMENTS];
    elb200_segment unpack;
} elb200_contract_plan;

int elb200_plan_copy(int grid_height, int grid_width, int vc_rank, const elb200_dist_spec* A,
                     const elb200_dist_spec* B, elb200_copy_plan* out);
int elb200_plan_contract(int grid_height, int grid_width, int vc_rank, const elb200_dist_spec* A,
                         const elb200_dist_spec* B, elb200_contract_plan* out);
/* Offset (elements) of the message rank `sender_vc` addresses to rank `dest_vc` inside the sender's
 * packed send staging of the general exchange of B = A (== sends[i].seg.buf_offset of the sender's
 * plan); lets a receiver pull its message out of the sender's peer-mapped staging. -1 on bad
 * arguments. Host-only. */
int64_t elb200_plan_send_offset(int grid_height, int grid_width, int sender_vc, const elb200_dist_spec* A,
                                const elb200_dist_spec* B, int dest_vc);
/* Layout of a zero-copy SUMMA-C panel (NN) for an r x c grid, seen from VC rank `vc_rank`: the k
 * indices of a kb-wide panel are permuted so that every owner's block is strided (see DESIGN.md
 * 3.5). row_off[r+1]: member q's B rows occupy panel rows [row_off[q], row_off[q+1]);
 * a_cols[5*i..]: {dst_col0, dst_step, src_col0, src_step, count} for this rank's i-th class of A1
 * columns (at most r classes); b_row0/b_rows: this rank's rows of the B panel. Host-only. */
int elb200_plan_direct_panel(int grid_height, int grid_width, int vc_rank, int64_t kb,
                             int64_t a_row_align, int64_t b_col_align, int64_t* row_off,
                             int64_t* a_cols, int* nclasses, int64_t* b_row0, int64_t* b_rows);
/* Shift_/Length_/MaxLength_ (include/El/core/indexing/impl.hpp:33-66,244-245) and
 * Grid::DefaultHeight (src/core/Grid.cpp:57-63), exported for the parity tests. */
int64_t elb200_shift(int64_t rank, int64_t align, int64_t stride);
int64_t elb200_length(int64_t n, int64_t shift, int64_t stride);
int64_t elb200_max_length(int64_t n, int64_t stride);
int elb200_default_grid_height(int grid_size);
/* GEMM_DEFAULT heuristic (src/blas_like/level3/Gemm/NN.hpp:572-600); returns a GemmAlgorithm. */
int elb200_gemm_resolve_alg(int transA, int transB, int64_t m, int64_t n, int64_t k, int alg);

/* ------------------------------------------------------------------------------------------
 * Handle API over the host engine: what the C++ facade (include/El.hpp) and the Python mirror
 * (elemental_b200/el.py) bind. One process per GPU; every rank of a grid calls collectively and
 * in the same order (same contract as the reference, SURVEY.md 8(b) "Threading").
 *   elb200_grid_t  <-> El::Grid        (include/El/core/Grid.hpp:15-140, src/core/Grid.cpp:74-223)
 *   elb200_dm_t    <-> El::DistMatrix<T,U,V,ELEMENT,Device::GPU>
 *                      (include/El/core/DistMatrix/ElementMatrix.hpp, AbstractDistMatrix.hpp)
 * ---------------------------------------------------------------------------------------- */
typedef struct elb200_grid_s* elb200_grid_t;
typedef struct elb200_dm_s* elb200_dm_t;
enum { ELB200_F32 = 0, ELB200_F64 = 1, ELB200_BYTE = 2 /* collectives, data movement only: no reductions */, ELB200_F16 = 3 /* IEEE half */ };
#define ELB200_UNIQUE_ID_BYTES 128

/* NCCL bootstrap: rank 0 calls elb200_unique_id and ships the 128 bytes to the other ranks by any
 * side channel (the reference gets its communicator from MPI, src/core/imports/mpi.cpp:80-91). */
int elb200_unique_id(void* out128);
/* height x width grid, this process is VC rank `vc_rank` and drives CUDA device `device`.
 * id128 may be NULL when height*width == 1. */
int elb200_grid_create(int height, int width, int vc_rank, const void* id128, int device,
                       elb200_grid_t* out);
int elb200_grid_destroy(elb200_grid_t g);
int elb200_grid_info(elb200_grid_t g, int* height, int* width, int* vc_rank, int* mc_rank,
                     int* mr_rank);
void* elb200_grid_stream(elb200_grid_t g); /* the grid's compute stream (cudaStream_t) */
int elb200_grid_sync(elb200_grid_t g);     /* cudaStreamSynchronize on it */

/* ------------------------------------------------------------------------------------------
 * Communicators and per-collective entry points: what the reference binds below its copy:: and
 * axpy_contract:: routines. A communicator stands for mpi::Comm (an AluminumComm over an MPI communicator,
 * include/El/core/imports/mpi/aluminum_comm.hpp:93-236); there is no MPI here — it is an NCCL communicator,
 * and when it is a Grid's MC / MR / VC communicator the collectives travel through the grid's NVLink peer
 * arena (copy engines + stream memory operations, no SM) whenever the message fits a slot,
 * with NCCL as the fallback (ELB200_P2P=0 forces NCCL). Equal counts on every rank, counts in
 * ELEMENTS of `dtype` (int64: the reference's `int` counts, src/core/imports/mpi/AllGather.hpp:11-12, cap a
 * call at 2^31 elements), device pointers, asynchronous on the caller's stream, every rank of the
 * communicator calls in the same order (SURVEY.md 8(b) "Threading").
 * ---------------------------------------------------------------------------------------- */
typedef struct elb200_comm_s* elb200_comm_t;
enum { ELB200_COMM_MC = 0, ELB200_COMM_MR = 1, ELB200_COMM_VC = 2, ELB200_COMM_VR = 3 };
enum { ELB200_ROW_MAJOR = 0, ELB200_COLUMN_MAJOR = 1 }; /* El::GridOrder (include/El/core/types.hpp) */

/* mpi::COMM_WORLD: bootstrapped once per process from RANK / WORLD_SIZE; the NCCL id travels from rank 0
 * through ELB200_ID_FILE (a path every rank can reach) or over TCP to MASTER_ADDR : ELB200_PORT (default
 * MASTER_PORT + 29). Replaces MPI_Init_thread + MPI_COMM_WORLD (src/core/imports/mpi.cpp:80-91). The handle is
 * owned by the library (elb200_comm_free on it tears the world down: El::Finalize). */
int elb200_comm_world(elb200_comm_t* out);
/* A communicator over an explicit NCCL id (elb200_unique_id) for callers with their own side channel. */
int elb200_comm_create(const void* id128, int rank, int size, elb200_comm_t* out);
int elb200_comm_rank(elb200_comm_t c);  /* mpi::Rank(comm) */
int elb200_comm_size(elb200_comm_t c);  /* mpi::Size(comm) */
/* mpi::Split(comm, color, key, newComm) (src/core/imports/mpi.cpp); color < 0 = MPI_UNDEFINED: *out = NULL */
int elb200_comm_split(elb200_comm_t c, int color, int key, elb200_comm_t* out);
int elb200_comm_dup(elb200_comm_t c, elb200_comm_t* out);   /* mpi::Dup */
int elb200_comm_free(elb200_comm_t c);                       /* mpi::Free */
int elb200_comm_barrier(elb200_comm_t c);                    /* mpi::Barrier: blocks the host */
/* El::Grid(mpi::Comm comm, int height, GridOrder order) (include/El/core/Grid.hpp:19-20,
 * src/core/Grid.cpp:74-223). Collective over `c`; height <= 0 = Grid::DefaultHeight(size); device < 0 = the
 * current CUDA device. The grid's own NCCL communicators are created from an id broadcast over `c`. */
int elb200_grid_create_on(elb200_comm_t c, int height, int order, int device, elb200_grid_t* out);
/* Grid::MCComm() / MRComm() / VCComm() / VRComm() (include/El/core/Grid.hpp:62-75): `which` is an
 * ELB200_COMM_* value. The handle borrows the grid's communicator and keeps the grid alive; release it with
 * elb200_comm_free. VRComm is built on first use (collective over the grid). */
int elb200_grid_comm(elb200_grid_t g, int which, elb200_comm_t* out);

/* mpi::AllGather(sbuf, sc, rbuf, rc, comm, syncInfo) (src/core/imports/mpi/AllGather.hpp:8-21):
 * rbuf[i*count .. (i+1)*count) = rank i's sbuf. */
int elb200_allgather(const void* sbuf, void* rbuf, int64_t count, int dtype, elb200_comm_t c, void* stream);
/* mpi::ReduceScatter(sbuf, rbuf, rc, mpi::SUM, comm, syncInfo) (ReduceScatter.hpp:36-52): sbuf holds
 * size*count elements, rbuf gets the sum over ranks of block `rank`. sbuf == rbuf is the in-place form
 * (ReduceScatter.hpp:149-165): the result lands at the start of the buffer. Over the arena the sum runs in
 * communicator order on every rank (deterministic); over NCCL the order is NCCL's. */
int elb200_reduce_scatter(const void* sbuf, void* rbuf, int64_t count, int dtype, elb200_comm_t c, void* stream);
/* mpi::AllReduce(sbuf, rbuf, count, mpi::SUM, comm, syncInfo) (src/core/imports/mpi/AllReduce.hpp) */
int elb200_allreduce(const void* sbuf, void* rbuf, int64_t count, int dtype, elb200_comm_t c, void* stream);
/* mpi::AllToAll(sbuf, sc, rbuf, rc, comm, syncInfo) (src/core/imports/mpi/AllToAll.hpp): block j of sbuf goes to
 * rank j, block i of rbuf comes from rank i, `count` elements each. */
int elb200_alltoall(const void* sbuf, void* rbuf, int64_t count, int dtype, elb200_comm_t c, void* stream);
/* mpi::SendRecv(sbuf, sc, to, rbuf, rc, from, comm, syncInfo) (src/core/imports/mpi/SendRecv.hpp) */
int elb200_sendrecv(const void* sbuf, int64_t scount, int to, void* rbuf, int64_t rcount, int from, int dtype,
                    elb200_comm_t c, void* stream);
/* mpi::Broadcast(buf, count, root, comm, syncInfo) (src/core/imports/mpi/Broadcast.hpp) */
int elb200_bcast(void* buf, int64_t count, int dtype, int root, elb200_comm_t c, void* stream);

/* Measurement hook (bench.py roofline leg): when enabled, every LocalGemm kernel launch is
 * bracketed by CUDA events on its own stream; collect() synchronises and returns the summed device
 * time (ms), the algorithmic flops (2*m*n*k per launch) and the launch count, then resets. */
void elb200_gemm_profile(int enable);
int elb200_gemm_profile_collect(double* total_ms, double* total_flop, int* launches);

/* El::Blocksize / SetBlocksize / PushBlocksizeStack / PopBlocksizeStack
 * (src/blas_like/blocksizes.cpp:35-67). */
int elb200_blocksize(void);
void elb200_set_blocksize(int nb);
void elb200_push_blocksize(int nb);
void elb200_pop_blocksize(void);

typedef struct elb200_dm_info_t {
    int dtype, col_dist, row_dist, is_view;
    int64_t height, width, col_align, row_align, col_shift, row_shift;
    int64_t col_stride, row_stride, local_height, local_width, ldim;
    void* buffer; /* device pointer to the local column-major matrix */
    int root;     /* owning VC rank of a [CIRC,CIRC] matrix (AbstractDistMatrix::Root) */
} elb200_dm_info_t;

int elb200_dm_create(elb200_grid_t g, int dtype, int col_dist, int row_dist, elb200_dm_t* out);
int elb200_dm_destroy(elb200_dm_t a);
int elb200_dm_info(elb200_dm_t a, elb200_dm_info_t* out);
int elb200_dm_align(elb200_dm_t a, int64_t col_align, int64_t row_align);
int elb200_dm_align_with(elb200_dm_t a, elb200_dm_t other);
int elb200_dm_align_cols(elb200_dm_t a, int64_t col_align); /* ElementalMatrix::AlignCols */
int elb200_dm_align_rows(elb200_dm_t a, int64_t row_align); /* ElementalMatrix::AlignRows */
/* SetSyncInfo / SyncInfoFromMatrix on the local matrix (include/hydrogen/device/gpu/cuda/SyncInfo.hpp:14-84): the
 * cudaStream_t that operations writing this matrix are enqueued on (NULL = back to the grid's compute stream).
 * elb200_dist_gemm orders the three operands' streams like MakeMultiSync(C, A, B) in Gemm_impl<GPU>
 * (src/blas_like/level3/Gemm.cpp:172-175; include/hydrogen/MultiSync.hpp:31-78). */
int elb200_dm_set_stream(elb200_dm_t a, void* stream);
void* elb200_dm_stream(elb200_dm_t a);
/* AbstractDistMatrix::SetRoot (include/El/core/DistMatrix/AbstractDistMatrix.hpp:58): the VC rank
 * that holds a [CIRC,CIRC] matrix. Collective metadata; every rank passes the same root. */
int elb200_dm_set_root(elb200_dm_t a, int root);
int elb200_dm_resize(elb200_dm_t a, int64_t height, int64_t width);
/* Local matrix <-> HOST memory (column-major, leading dimension ld >= local height), the data
 * movement of the reference's CPU <-> GPU DistMatrix copies (include/El/core/ProxyDevice.hpp,
 * Matrix<T,Device::GPU>(Matrix<T,Device::CPU> const&)). Asynchronous on the matrix's stream; pinned
 * host memory makes it a true overlap-able DMA. */
int elb200_dm_write_local(elb200_dm_t a, const void* host_src, int64_t ld);
int elb200_dm_read_local(elb200_dm_t a, void* host_dst, int64_t ld);
int elb200_dm_attach(elb200_dm_t a, int64_t height, int64_t width, int64_t col_align,
                     int64_t row_align, void* device_buffer, int64_t ldim);
int elb200_dm_view(elb200_dm_t a, int64_t i0, int64_t i1, int64_t j0, int64_t j1,
                   elb200_dm_t* out);
/* B = A for any pair of the 11 distributions and [CIRC,CIRC] (DistMatrix::operator=). */
int elb200_dm_copy(elb200_dm_t A, elb200_dm_t B);
int elb200_dm_transpose(elb200_dm_t A, elb200_dm_t B); /* B = A^T (El::Transpose) */
int elb200_dm_axpy(double alpha, elb200_dm_t X, elb200_dm_t Y);
int elb200_dm_scale(double alpha, elb200_dm_t A);
int elb200_dm_zero(elb200_dm_t A);
int elb200_dm_axpy_contract(double alpha, elb200_dm_t A, elb200_dm_t B);
int elb200_dm_contract(elb200_dm_t A, elb200_dm_t B);
int elb200_dm_transpose_axpy_contract(double alpha, elb200_dm_t A, elb200_dm_t B);
/* El::LocalGemm (src/blas_like/level3/Gemm.cpp:318-426) */
int elb200_dm_local_gemm(int transA, int transB, double alpha, elb200_dm_t A, elb200_dm_t B,
                         double beta, elb200_dm_t C);
/* El::Gemm(orientA, orientB, alpha, A, B, beta, C, alg) (src/blas_like/level3/Gemm.cpp:273-302).
 * alg: the reference's El::GemmAlgorithm values (include/El/blas_like/level3.hpp:23-33):
 *      0=DEFAULT 1=SUMMA_A_MS 2=SUMMA_A 3=SUMMA_B_MS 4=SUMMA_B 5=SUMMA_C_MS 6=SUMMA_C
 *      7=SUMMA_DOT 8=CANNON (the _MS variants run their plain twins; CANNON errors on GPU
 *      as in Gemm/NN.hpp:30-31). */
int elb200_dist_gemm(int transA, int transB, double alpha, elb200_dm_t A, elb200_dm_t B,
                     double beta, elb200_dm_t C, int alg);

/* ------------------------------------------------------------------------------------------
 * Triangular rank-k family built from the same blocks: panel AllGathers, LocalGemm
 * and the trapezoid kernels above. uplo / side are El::UpperOrLower / El::LeftOrRight values.
 *   ScaleTrapezoid / AxpyTrapezoid / MakeTrapezoidal / MakeSymmetric
 *       (include/El/blas_like/level1/{ScaleTrapezoid,AxpyTrapezoid,MakeTrapezoidal,MakeSymmetric}.hpp)
 *   LocalTrrk: triangle of C[MC,MR] += alpha op(A) op(B), A = [MC,*] (N) or [*,MC] (T), B = [*,MR] (N)
 *       or [MR,*] (T), aligned with C (src/blas_like/level3/Trrk/Local.hpp:688-890; the GPU
 *       instantiation Trrk.cpp:117-176 is the (T,N) case)
 *   Trrk : triangle of C = alpha op(A) op(B) + beta C        (src/blas_like/level3/Trrk.cpp:100-118)
 *   Syrk / Herk (real): triangle of C = alpha A A^T + beta C (N) or alpha A^T A + beta C (T)
 *       (src/blas_like/level3/Syrk.cpp:172-190, Herk.cpp)
 *   Symm / Hemm (real): C = alpha A B + beta C (LEFT) or alpha B A + beta C (RIGHT), A symmetric with
 *       only its `uplo` triangle referenced (src/blas_like/level3/Symm.cpp:56-80, Hemm.cpp)
 * Entries of C outside the triangle are never read or written by Trrk / Syrk.
 * ---------------------------------------------------------------------------------------- */
/* El::LocalTrrkBlocksize<T> / SetLocalTrrkBlocksize<T> (src/blas_like/blocksizes.cpp:82-87): LocalTrrk
 * recurses while C.Height() > gridWidth * blocksize. One value for both types; default 2048. */
int64_t elb200_local_trrk_blocksize(void);
void elb200_set_local_trrk_blocksize(int64_t nb);
int elb200_dm_scale_trapezoid(double alpha, int uplo, elb200_dm_t A, int64_t offset);
int elb200_dm_axpy_trapezoid(int uplo, double alpha, elb200_dm_t X, elb200_dm_t Y, int64_t offset);
int elb200_dm_make_trapezoidal(int uplo, elb200_dm_t A, int64_t offset);
int elb200_dm_make_symmetric(int uplo, elb200_dm_t A);
int elb200_dm_local_trrk(int uplo, int transA, int transB, double alpha, elb200_dm_t A, elb200_dm_t B,
                         double beta, elb200_dm_t C);
int elb200_dist_trrk(int uplo, int transA, int transB, double alpha, elb200_dm_t A, elb200_dm_t B,
                     double beta, elb200_dm_t C);
int elb200_dist_syrk(int uplo, int trans, double alpha, elb200_dm_t A, double beta, elb200_dm_t C);
int elb200_dist_symm(int side, int uplo, double alpha, elb200_dm_t A, elb200_dm_t B, double beta,
                     elb200_dm_t C);
/* El::Trsm(side, uplo, orientation, diag, alpha, A, B, checkIfSingular, alg) on DistMatrices
 * (src/blas_like/level3/Trsm.cpp:129-440, Trsm/{LLN,LLT,LUN,LUT,RLN,RLT,RUN,RUT}.hpp): B := alpha op(A)^-1 B
 * (LEFT) or alpha B op(A)^-1 (RIGHT).  alg: El::TrsmAlgorithm values (0 DEFAULT, 1 LARGE, 2 MEDIUM, 3 SMALL);
 * every value runs the many-right-hand-sides schedule (the *Large variants).  checkIfSingular is not
 * evaluated (the reference's GPU path does not evaluate it either: cuBLAS trsm has no such check).
 * El::LocalTrsm (Trsm.cpp:501-526): A is [*,*], X is [*,V] (LEFT) or [U,*] (RIGHT); purely local. */
int elb200_dm_local_trsm(int side, int uplo, int trans, int diag, double alpha, elb200_dm_t A, elb200_dm_t X);
int elb200_dist_trsm(int side, int uplo, int trans, int diag, double alpha, elb200_dm_t A, elb200_dm_t B, int alg);

/* El::Gemm with every operand in HOST memory (the reference reaches this case through its
 * CPU<->GPU proxies, include/El/core/ProxyDevice.hpp + Matrix<T,GPU> = Matrix<T,CPU> copies,
 * include/El/core/Matrix/impl_gpu.hpp, which upload everything, compute, download).
 * A, B, C are this rank's LOCAL parts of zero-aligned [MC,MR] matrices of global size
 * op(A) m x k, op(B) k x n, C m x n (column-major, ld in elements; pinned memory recommended).
 * A stays resident on the device, column blocks of B and C stream through a four-stream
 * pipeline (H2D | gather | GEMM | D2H). Synchronous: C is complete in host memory on return. */
int elb200_dist_gemm_host_f32(elb200_grid_t g, int transA, int transB, int64_t m, int64_t n, int64_t k,
                              float alpha, const float* A, int64_t lda, const float* B, int64_t ldb,
                              float beta, float* C, int64_t ldc);
int elb200_dist_gemm_host_f64(elb200_grid_t g, int transA, int transB, int64_t m, int64_t n, int64_t k,
                              double alpha, const double* A, int64_t lda, const double* B, int64_t ldb,
                              double beta, double* C, int64_t ldc);

/* ---- Complex instantiations: El::Gemm, redistributions and entrywise routines for Complex<float> /
 * Complex<double> (reference: src/blas_like/level3/Gemm.cpp:504-506 and the complex instantiations of
 * the sources under src/core/DistMatrix/ElementMatrix and src/blas_like/level1 via include/El/macros/Instantiate.h),
 * built on the real ABI ("4M").
 * A complex DistMatrix is stored PLANAR — a real and an imaginary DistMatrix (elb200_zdm_plane) with one distribution and
 * alignment — so every redistribution is the real path applied to both planes and El::Gemm is four real El::Gemm calls
 * plus entrywise combinations (csrc/capi_complex.cpp). The reference's interleaved (re, im) layout exists at the
 * boundary: elb200_zdm_write_local / read_local take interleaved HOST or DEVICE buffers (ld in complex elements), and
 * elb200_gemm_c32 / _c64 multiply interleaved local matrices (the Gemm(Matrix<Complex<T>,GPU>) -> cublasCgemm / Zgemm of
 * src/hydrogen/device/cuBLAS_API.cpp:419-427). real_dtype is ELB200_F32 or ELB200_F64; scalars are (re, im) pairs. */
typedef struct elb200_zdm_s* elb200_zdm_t;
int elb200_zdm_create(elb200_grid_t g, int real_dtype, int col_dist, int row_dist, elb200_zdm_t* out);
int elb200_zdm_destroy(elb200_zdm_t a);
int elb200_zdm_plane(elb200_zdm_t a, int imag, elb200_dm_t* out); /* borrowed: the real (0) / imaginary (1) DistMatrix */
int elb200_zdm_resize(elb200_zdm_t a, int64_t height, int64_t width);
int elb200_zdm_align(elb200_zdm_t a, int64_t col_align, int64_t row_align);
int elb200_zdm_align_with(elb200_zdm_t a, elb200_zdm_t other);
int elb200_zdm_set_root(elb200_zdm_t a, int root);
int elb200_zdm_set_stream(elb200_zdm_t a, void* stream);
int elb200_zdm_view(elb200_zdm_t a, int64_t i0, int64_t i1, int64_t j0, int64_t j1, elb200_zdm_t* out);
int elb200_zdm_write_local(elb200_zdm_t a, const void* interleaved, int64_t ld);
int elb200_zdm_read_local(elb200_zdm_t a, void* interleaved, int64_t ld);
int elb200_zdm_copy(elb200_zdm_t A, elb200_zdm_t B);                     /* B = A, any pair of distributions */
int elb200_zdm_transpose(elb200_zdm_t A, elb200_zdm_t B, int conjugate); /* Transpose / Adjoint (level1/Transpose.hpp) */
int elb200_zdm_zero(elb200_zdm_t A);
int elb200_zdm_scale(double alpha_re, double alpha_im, elb200_zdm_t A);
int elb200_zdm_axpy(double alpha_re, double alpha_im, elb200_zdm_t X, elb200_zdm_t Y);
int elb200_zdm_axpy_contract(double alpha_re, double alpha_im, elb200_zdm_t A, elb200_zdm_t B);
/* El::Gemm(orientA, orientB, alpha, A, B, beta, C, alg) for complex DistMatrices; ELB200_OP_C is the conjugate transpose. */
int elb200_zdm_gemm(int transA, int transB, double alpha_re, double alpha_im, elb200_zdm_t A, elb200_zdm_t B, double beta_re,
                    double beta_im, elb200_zdm_t C, int alg);
/* Local complex GEMM on interleaved column-major device matrices; alpha, beta point to (re, im). beta == 0 never reads C. */
int elb200_gemm_c32(int transA, int transB, int64_t m, int64_t n, int64_t k, const float* alpha, const void* A, int64_t lda,
                    const void* B, int64_t ldb, const float* beta, void* C, int64_t ldc, void* stream);
int elb200_gemm_c64(int transA, int transB, int64_t m, int64_t n, int64_t k, const double* alpha, const void* A, int64_t lda,
                    const void* B, int64_t ldb, const double* beta, void* C, int64_t ldc, void* stream);

#ifdef __cplusplus
} /* extern "C" */
#endif

#endif /* ELB200_H_ */
