// Element-cyclic distribution algebra and redistribution plans (pure host code, no CUDA).
//
// The reference implements every (U,V) -> (U',V') redistribution as its own routine
// (include/El/blas_like/level1/Copy/*.hpp, ~8,500 lines). Here ONE planner covers them all:
// for any pair of element-cyclic distributions over an r x c grid, the set of entries that rank s
// owns in the source AND rank q owns in the target is a 2-D lattice (Chinese remainder theorem on
// the row and on the column index separately), so every redistribution is "pack lattices ->
// exchange -> unpack lattices", with AllGather / ReduceScatter recognised as special exchanges.
//
// Index conventions follow the reference exactly:
//   Shift_(rank, align, stride) = (rank - align) mod stride     include/El/core/indexing/impl.hpp:244-245
//   Length_(n, shift, stride)   = n > shift ? (n-shift-1)/stride + 1 : 0     impl.hpp:33-36
//   MaxLength_(n, stride)       = n > 0 ? (n-1)/stride + 1 : 0               impl.hpp:63-66
//   VC rank = mc + r*mr, VR rank = mr + c*mc                                  src/core/Grid.cpp:147-148
#pragma once
#include <cstdint>
#include <vector>

#include "../../include/elb200.h"

namespace elb {

// CIRC: the whole matrix on one rank (`root`, a VC rank); only valid as [CIRC,CIRC]
// (reference DistMatrix<T,CIRC,CIRC>, src/core/DistMatrix/ElementMatrix/CIRC_CIRC.cpp).
enum Dist : int { MC = 0, MR = 1, VC = 2, VR = 3, STAR = 4, CIRC = 5 };

inline int64_t Mod(int64_t a, int64_t b) { int64_t r = a % b; return r < 0 ? r + b : r; }
inline int64_t Shift(int64_t rank, int64_t align, int64_t stride) { return Mod(rank - align, stride); }
inline int64_t Length(int64_t n, int64_t shift, int64_t stride) {
    return n > shift ? (n - shift - 1) / stride + 1 : 0;
}
inline int64_t MaxLength(int64_t n, int64_t stride) { return n > 0 ? (n - 1) / stride + 1 : 0; }

struct GridShape {
    int r = 1, c = 1;  // grid height (MC size) and width (MR size)
    int size() const { return r * c; }
};
struct GridCoord {
    int mc = 0, mr = 0;
};
inline GridCoord coord_of_vc(const GridShape& g, int vc) { return {vc % g.r, vc / g.r}; }
inline int vc_of(const GridShape& g, GridCoord q) { return q.mc + g.r * q.mr; }
inline int vr_of(const GridShape& g, GridCoord q) { return q.mr + g.c * q.mc; }

int dist_stride(const GridShape& g, Dist d);
int dist_rank(const GridShape& g, Dist d, GridCoord q);
bool valid_pair(Dist u, Dist v);  // the 11 ELEMENT-wrap pairs reachable from [MC,MR] Gemm + [CIRC,CIRC]

// Distribution metadata of one DistMatrix (the reference's DistData + dims,
// include/El/core/DistMatrix/AbstractDistMatrix.hpp).
struct DistSpec {
    Dist col = MC, row = MR;
    int64_t col_align = 0, row_align = 0;
    int64_t height = 0, width = 0;
    int root = 0;  // [CIRC,CIRC] only: VC rank that owns the matrix (reference Root(), CrossComm = VCComm)
};
inline bool owns_data(const GridShape& g, const DistSpec& d, GridCoord q) {
    return d.col != CIRC || vc_of(g, q) == d.root;
}
int64_t local_height(const GridShape& g, const DistSpec& d, GridCoord q);
int64_t local_width(const GridShape& g, const DistSpec& d, GridCoord q);

// One message of an exchange: the lattice it occupies in the local matrix of this rank and its
// place in the staging buffer. `peer` is a VC rank (== rank in the grid's world communicator).
struct Message {
    int peer = 0;
    elb200_segment seg{};  // rows/cols may be 0 (then the message does not exist)
    int64_t count() const { return seg.rows * seg.cols; }
};

enum ExchangeKind : int {
    EX_LOCAL = 0,      // only the self message is non-empty
    EX_ALLGATHER = 1,  // every rank of `comm_dist`'s communicator sends its whole local matrix
    EX_SENDRECV = 2    // general: grouped point-to-point over the world communicator
};

struct CopyPlan {
    ExchangeKind kind = EX_LOCAL;
    Dist comm_dist = STAR;  // EX_ALLGATHER: MC, MR or VC (communicator the gather runs over)
    int64_t portion = 0;    // EX_ALLGATHER: slot size (elements) = maxLocalHeight*maxLocalWidth
    std::vector<Message> sends;  // lattices of the SOURCE local matrix, buf_offset in send staging
    std::vector<Message> recvs;  // lattices of the TARGET local matrix, buf_offset in recv staging
    int64_t send_elems = 0, recv_elems = 0;
    bool has_self = false;
    Message self_send{}, self_recv{};  // same entries seen from both local matrices
};

// Plan of B = A for this rank (reference: Copy(ElementalMatrix, ElementalMatrix) dispatch,
// src/core/DistMatrix/ElementMatrix/*.cpp operator= tables).
CopyPlan plan_copy(const GridShape& g, GridCoord me, const DistSpec& A, const DistSpec& B);

// Offset (elements) inside rank `sender`'s packed send staging of its message to VC rank
// `dest_vc`, for the EX_SENDRECV plan of B = A: what a receiver needs to fetch its message from the
// sender's staging without asking (peer-memory pull, engine_redist.cpp).
int64_t plan_send_offset(const GridShape& g, GridCoord sender, const DistSpec& A, const DistSpec& B, int dest_vc);

// Plan of B += alpha * (sum over the replicas of A) for the partial-sum distributions of the
// reference's axpy_contract::{RowScatter,ColScatter,Scatter}
// (include/El/blas_like/level1/AxpyContract.hpp:146-479). A's col/row dist must each equal B's or
// be STAR, with matching alignment on the equal ones.
struct ContractPlan {
    bool valid = false;
    Dist comm_dist = STAR;  // MC, MR, VC (or STAR: nothing to reduce, plain axpy)
    int group = 1;          // communicator size
    int my_index = 0;       // this rank's index in it
    int64_t portion = 0;    // reduce-scatter slot size (elements)
    std::vector<Message> packs;  // one per group member k: lattice of A's local matrix
    elb200_segment unpack{};     // where this rank's reduced slot lands in B's local matrix
};
ContractPlan plan_contract(const GridShape& g, GridCoord me, const DistSpec& A, const DistSpec& B);

// Zero-copy SUMMA-C panels (NN): layout of a kb-wide k-panel whose k indices are permuted so that
// every owner's contribution is a strided block (engine_gemm.cpp SummaCDirectNN).
//   B panel (kb x nloc): rows grouped by their owner in the column communicator; member q's local
//     rows of B1 are rows [row_off[q], row_off[q+1]) of the panel.
//   A panel (mloc x kb): column for panel-relative index t sits where row t of the B panel sits.
//     This rank's local columns of A1 split into classes; class u maps local columns
//     src_col0 + v*src_step to panel columns dst_col0 + v*dst_step, v = 0..count-1.
struct PanelCols { int64_t dst_col0, dst_step, src_col0, src_step, count; };
struct DirectPanelPlan {
    std::vector<int64_t> row_off;   // r + 1 entries
    std::vector<PanelCols> a_cols;  // this rank's contribution to the A panel
    int64_t b_row0 = 0, b_rows = 0; // this rank's contribution to the B panel
};
// a_row_align: row alignment of the A1 view (k over MR); b_col_align: column alignment of the B1
// view (k over MC).
DirectPanelPlan plan_direct_panel(const GridShape& g, GridCoord me, int64_t kb, int64_t a_row_align,
                                  int64_t b_col_align);

}  // namespace elb
