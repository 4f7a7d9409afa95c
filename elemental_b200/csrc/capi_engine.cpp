// extern "C" handle API over the host engine + the planner entry points. See include/elb200.h.
#include <cmath>
#include <cstring>
#include <memory>

#include "capi_handles.hpp"
#include "engine.hpp"
#include "engine_dev.hpp"

namespace elb {
void set_last_error(const char* what, const char* detail);
}
using namespace elb;

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

bool same_dtype(elb200_dm_t a, elb200_dm_t b) { return a && b && a->dtype == b->dtype; }
cudaStream_t dm_stream(elb200_dm_t a) { return a->dtype == ELB200_F32 ? a->f->stream : a->dtype == ELB200_F64 ? a->d->stream : a->h->stream; }
bool known_dtype(int dt) { return dt == ELB200_F32 || dt == ELB200_F64 || dt == ELB200_F16; }
bool level3_dtype(elb200_dm_t a) { return a && (a->dtype == ELB200_F32 || a->dtype == ELB200_F64); }  // the triangular family: float / double only

// element type of a DistMat and the conversion of the ABI's double scalars to it
template <class M> struct ElemOf;
template <class T> struct ElemOf<DistMat<T>> { using type = T; };
template <class M> typename ElemOf<std::decay_t<M>>::type sc(const M&, double v) { return typename ElemOf<std::decay_t<M>>::type(v); }

// run `f` on the typed matrix (matrices) behind one, two or three handles of the same dtype
template <class F> void on(elb200_dm_t a, F&& f) {
    if (a->dtype == ELB200_F32) f(*a->f);
    else if (a->dtype == ELB200_F64) f(*a->d);
    else f(*a->h);
}
template <class F> void on2(elb200_dm_t a, elb200_dm_t b, F&& f) {
    if (a->dtype == ELB200_F32) f(*a->f, *b->f);
    else if (a->dtype == ELB200_F64) f(*a->d, *b->d);
    else f(*a->h, *b->h);
}
template <class F> void on3(elb200_dm_t a, elb200_dm_t b, elb200_dm_t c, F&& f) {
    if (a->dtype == ELB200_F32) f(*a->f, *b->f, *c->f);
    else if (a->dtype == ELB200_F64) f(*a->d, *b->d, *c->d);
    else f(*a->h, *b->h, *c->h);
}

// hydrogen::MultiSync (include/hydrogen/MultiSync.hpp:31-78): on construction the master stream waits for the others,
// on destruction the others wait for the master — all asynchronously with respect to the host.
struct StreamJoin {
    cudaStream_t master, other[2];
    StreamJoin(cudaStream_t m, cudaStream_t a, cudaStream_t b) : master(m), other{a, b} {
        if (other[1] == other[0]) other[1] = master;
        for (cudaStream_t o : other)
            if (o != master) { Ev e; e.record(o); e.block(master); }
    }
    ~StreamJoin() {
        for (cudaStream_t o : other)
            if (o != master) {
                cudaEvent_t e = nullptr;
                if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) continue;
                if (cudaEventRecord(e, master) == cudaSuccess) cudaStreamWaitEvent(o, e, 0);
                cudaEventDestroy(e);
            }
    }
};

DistSpec to_spec(const elb200_dist_spec& s) {
    DistSpec d;
    d.col = Dist(s.col_dist); d.row = Dist(s.row_dist);
    d.col_align = s.col_align; d.row_align = s.row_align;
    d.height = s.height; d.width = s.width;
    d.root = s.root;
    return d;
}
elb200_message to_msg(const Message& m) {
    elb200_message o;
    o.peer = m.peer; o.valid = 1; o.seg = m.seg;
    return o;
}
bool spec_ok(const GridShape& g, const elb200_dist_spec& s) {
    if (!valid_pair(Dist(s.col_dist), Dist(s.row_dist))) return false;
    if (s.height < 0 || s.width < 0) return false;
    if (Dist(s.col_dist) == CIRC ? (s.root < 0 || s.root >= g.size()) : s.root != 0) return false;
    if (s.col_align < 0 || s.col_align >= dist_stride(g, Dist(s.col_dist))) return false;
    if (s.row_align < 0 || s.row_align >= dist_stride(g, Dist(s.row_dist))) return false;
    return true;
}

}  // namespace

extern "C" {

int64_t elb200_shift(int64_t rank, int64_t align, int64_t stride) { return Shift(rank, align, stride); }
int64_t elb200_length(int64_t n, int64_t shift, int64_t stride) { return Length(n, shift, stride); }
int64_t elb200_max_length(int64_t n, int64_t stride) { return MaxLength(n, stride); }

int elb200_default_grid_height(int p) {
    // reference Grid::DefaultHeight, src/core/Grid.cpp:57-63
    if (p < 1) return 1;
    int h = int(std::sqrt(double(p)));
    if (h < 1) h = 1;
    while (p % h != 0) ++h;
    return h;
}

int elb200_gemm_resolve_alg(int ta, int tb, int64_t m, int64_t n, int64_t k, int alg) {
    return ResolveGemmAlg(ta, tb, m, n, k, alg);
}

int elb200_plan_copy(int r, int c, int vc, const elb200_dist_spec* A, const elb200_dist_spec* B,
                     elb200_copy_plan* out) {
    if (!A || !B || !out || r < 1 || c < 1 || vc < 0 || vc >= r * c) return ELB200_ERR_ARG;
    GridShape g; g.r = r; g.c = c;
    if (r * c > ELB200_MAX_SEGMENTS || !spec_ok(g, *A) || !spec_ok(g, *B)) return ELB200_ERR_ARG;
    if (A->height != B->height || A->width != B->width) return ELB200_ERR_ARG;
    const CopyPlan p = plan_copy(g, coord_of_vc(g, vc), to_spec(*A), to_spec(*B));
    std::memset(out, 0, sizeof(*out));
    out->kind = p.kind; out->comm_dist = p.comm_dist; out->portion = p.portion;
    out->nsend = int(p.sends.size()); out->nrecv = int(p.recvs.size());
    out->has_self = p.has_self ? 1 : 0;
    out->send_elems = p.send_elems; out->recv_elems = p.recv_elems;
    if (p.has_self) { out->self_send = to_msg(p.self_send); out->self_recv = to_msg(p.self_recv); }
    for (int i = 0; i < out->nsend; ++i) out->sends[i] = to_msg(p.sends[i]);
    for (int i = 0; i < out->nrecv; ++i) out->recvs[i] = to_msg(p.recvs[i]);
    return ELB200_OK;
}

int elb200_plan_contract(int r, int c, int vc, const elb200_dist_spec* A, const elb200_dist_spec* B,
                         elb200_contract_plan* out) {
    if (!A || !B || !out || r < 1 || c < 1 || vc < 0 || vc >= r * c) return ELB200_ERR_ARG;
    GridShape g; g.r = r; g.c = c;
    if (r * c > ELB200_MAX_SEGMENTS || !spec_ok(g, *A) || !spec_ok(g, *B)) return ELB200_ERR_ARG;
    const ContractPlan p = plan_contract(g, coord_of_vc(g, vc), to_spec(*A), to_spec(*B));
    std::memset(out, 0, sizeof(*out));
    out->valid = p.valid ? 1 : 0; out->comm_dist = p.comm_dist; out->group = p.group;
    out->my_index = p.my_index; out->portion = p.portion;
    out->npack = int(p.packs.size());
    for (int i = 0; i < out->npack; ++i) out->packs[i] = to_msg(p.packs[i]);
    out->unpack = p.unpack;
    return ELB200_OK;
}

int64_t elb200_plan_send_offset(int r, int c, int sender, const elb200_dist_spec* A, const elb200_dist_spec* B, int dest) {
    if (!A || !B || r < 1 || c < 1 || sender < 0 || sender >= r * c || dest < 0 || dest >= r * c) return -1;
    GridShape g; g.r = r; g.c = c;
    if (!spec_ok(g, *A) || !spec_ok(g, *B)) return -1;
    return plan_send_offset(g, coord_of_vc(g, sender), to_spec(*A), to_spec(*B), dest);
}

int elb200_plan_direct_panel(int r, int c, int vc, int64_t kb, int64_t a_row_align, int64_t b_col_align,
                             int64_t* row_off, int64_t* a_cols, int* nclasses, int64_t* b_row0, int64_t* b_rows) {
    if (r < 1 || c < 1 || vc < 0 || vc >= r * c || kb < 0 || !row_off || !a_cols || !nclasses || !b_row0 || !b_rows)
        return ELB200_ERR_ARG;
    if (r > ELB200_MAX_SEGMENTS) return ELB200_ERR_ARG;
    GridShape g; g.r = r; g.c = c;
    const DirectPanelPlan p = plan_direct_panel(g, coord_of_vc(g, vc), kb, Mod(a_row_align, c), Mod(b_col_align, r));
    for (size_t i = 0; i < p.row_off.size(); ++i) row_off[i] = p.row_off[i];
    *nclasses = int(p.a_cols.size());
    for (size_t i = 0; i < p.a_cols.size(); ++i) {
        const PanelCols& pc = p.a_cols[i];
        int64_t* o = a_cols + 5 * i;
        o[0] = pc.dst_col0; o[1] = pc.dst_step; o[2] = pc.src_col0; o[3] = pc.src_step; o[4] = pc.count;
    }
    *b_row0 = p.b_row0; *b_rows = p.b_rows;
    return ELB200_OK;
}

int elb200_unique_id(void* out128) {
    if (!out128) return ELB200_ERR_ARG;
    static_assert(sizeof(ncclUniqueId) <= ELB200_UNIQUE_ID_BYTES, "unique id size");
    ncclUniqueId id;
    ncclResult_t r = ncclGetUniqueId(&id);
    if (r != ncclSuccess) { set_last_error("ncclGetUniqueId", ncclGetErrorString(r)); return ELB200_ERR_NCCL; }
    std::memset(out128, 0, ELB200_UNIQUE_ID_BYTES);
    std::memcpy(out128, &id, sizeof(id));
    return ELB200_OK;
}

int elb200_grid_create(int height, int width, int vc, const void* id128, int device, elb200_grid_t* out) {
    if (!out) return ELB200_ERR_ARG;
    *out = nullptr;
    return guarded("elb200_grid_create", [&] {
        ncclUniqueId id;
        if (id128) std::memcpy(&id, id128, sizeof(id));
        auto h = std::make_unique<elb200_grid_s>();
        h->g = std::make_shared<Grid>(height, width, vc, id128 ? &id : nullptr, device);
        *out = h.release();
    });
}
int elb200_grid_destroy(elb200_grid_t g) {
    if (!g) return ELB200_OK;
    // Collective part now (peer arena unmap + barrier, NCCL communicators); the streams live on
    // until the last DistMatrix handle of this grid is destroyed.
    return guarded("elb200_grid_destroy", [&] { g->g->Shutdown(); delete g; });
}
int elb200_grid_info(elb200_grid_t g, int* h, int* w, int* vc, int* mc, int* mr) {
    if (!g) return ELB200_ERR_ARG;
    if (h) *h = g->g->shape.r;
    if (w) *w = g->g->shape.c;
    if (vc) *vc = g->g->vc_rank;
    if (mc) *mc = g->g->me.mc;
    if (mr) *mr = g->g->me.mr;
    return ELB200_OK;
}
void* elb200_grid_stream(elb200_grid_t g) { return g ? g->g->stream : nullptr; }
int elb200_grid_sync(elb200_grid_t g) {
    if (!g) return ELB200_ERR_ARG;
    return guarded("elb200_grid_sync", [&] { cuda_check(cudaStreamSynchronize(g->g->stream), "cudaStreamSynchronize"); });
}

void elb200_gemm_profile(int enable) { GemmProfileEnable(enable != 0); }
int elb200_gemm_profile_collect(double* total_ms, double* total_flop, int* launches) {
    return guarded("elb200_gemm_profile_collect", [&] { GemmProfileCollect(total_ms, total_flop, launches); });
}

int elb200_blocksize(void) { return Grid::Blocksize(); }
void elb200_set_blocksize(int nb) { Grid::SetBlocksize(nb); }
void elb200_push_blocksize(int nb) { Grid::PushBlocksize(nb); }
void elb200_pop_blocksize(void) { Grid::PopBlocksize(); }
int64_t elb200_local_trrk_blocksize(void) { return Grid::LocalTrrkBlocksize(); }
void elb200_set_local_trrk_blocksize(int64_t nb) { Grid::SetLocalTrrkBlocksize(nb); }

int elb200_dm_create(elb200_grid_t g, int dtype, int col, int row, elb200_dm_t* out) {
    if (!g || !out || !known_dtype(dtype)) return ELB200_ERR_ARG;
    *out = nullptr;
    return guarded("elb200_dm_create", [&] {
        auto h = std::make_unique<elb200_dm_s>();
        h->dtype = dtype;
        h->keep = g->g;
        if (dtype == ELB200_F32) h->f = std::make_unique<DistMat<float>>(*g->g, Dist(col), Dist(row));
        else if (dtype == ELB200_F64) h->d = std::make_unique<DistMat<double>>(*g->g, Dist(col), Dist(row));
        else h->h = std::make_unique<DistMat<Half>>(*g->g, Dist(col), Dist(row));
        *out = h.release();
    });
}
int elb200_dm_destroy(elb200_dm_t a) {
    if (!a) return ELB200_OK;
    return guarded("elb200_dm_destroy", [&] { delete a; });
}

// the triangular family below still dispatches over float / double only
#define ELB_ON(a, expr_f, expr_d) do { if ((a)->dtype == ELB200_F32) { auto& M = *(a)->f; expr_f; } else { auto& M = *(a)->d; expr_d; } } while (0)

int elb200_dm_info(elb200_dm_t a, elb200_dm_info_t* o) {
    if (!a || !o) return ELB200_ERR_ARG;
    on(a, [&](auto& M) {
        o->dtype = a->dtype; o->col_dist = M.spec.col; o->row_dist = M.spec.row; o->is_view = M.owner ? 0 : 1;
        o->height = M.spec.height; o->width = M.spec.width;
        o->col_align = M.spec.col_align; o->row_align = M.spec.row_align;
        o->col_shift = M.col_shift(); o->row_shift = M.row_shift();
        o->col_stride = M.col_stride(); o->row_stride = M.row_stride();
        o->local_height = M.lh(); o->local_width = M.lw(); o->ldim = M.ldim;
        o->buffer = M.buf;
        o->root = M.spec.root;
    });
    return ELB200_OK;
}
int elb200_dm_align(elb200_dm_t a, int64_t ca, int64_t ra) {
    if (!a) return ELB200_ERR_ARG;
    return guarded("elb200_dm_align", [&] { on(a, [&](auto& M) { M.Align(ca, ra); }); });
}
int elb200_dm_align_with(elb200_dm_t a, elb200_dm_t o) {
    if (!a || !o) return ELB200_ERR_ARG;
    return guarded("elb200_dm_align_with", [&] {
        DistSpec s;
        on(o, [&](auto& M) { s = M.spec; });
        on(a, [&](auto& M) { M.AlignWith(s); });
    });
}
int elb200_dm_set_root(elb200_dm_t a, int root) {
    if (!a) return ELB200_ERR_ARG;
    return guarded("elb200_dm_set_root", [&] { on(a, [&](auto& M) { M.SetRoot(root); }); });
}
int elb200_dm_write_local(elb200_dm_t a, const void* src, int64_t ld) {
    if (!a) return ELB200_ERR_ARG;
    return guarded("elb200_dm_write_local", [&] {
        on(a, [&](auto& M) { M.WriteLocal(static_cast<const typename ElemOf<std::decay_t<decltype(M)>>::type*>(src), ld); });
    });
}
int elb200_dm_read_local(elb200_dm_t a, void* dst, int64_t ld) {
    if (!a) return ELB200_ERR_ARG;
    return guarded("elb200_dm_read_local", [&] {
        on(a, [&](auto& M) { M.ReadLocal(static_cast<typename ElemOf<std::decay_t<decltype(M)>>::type*>(dst), ld); });
    });
}
int elb200_dm_resize(elb200_dm_t a, int64_t h, int64_t w) {
    if (!a) return ELB200_ERR_ARG;
    return guarded("elb200_dm_resize", [&] { on(a, [&](auto& M) { M.Resize(h, w); }); });
}
int elb200_dm_attach(elb200_dm_t a, int64_t h, int64_t w, int64_t ca, int64_t ra, void* buf, int64_t ld) {
    if (!a) return ELB200_ERR_ARG;
    return guarded("elb200_dm_attach", [&] {
        on(a, [&](auto& M) { M.Attach(h, w, ca, ra, static_cast<typename ElemOf<std::decay_t<decltype(M)>>::type*>(buf), ld); });
    });
}
int elb200_dm_view(elb200_dm_t a, int64_t i0, int64_t i1, int64_t j0, int64_t j1, elb200_dm_t* out) {
    if (!a || !out) return ELB200_ERR_ARG;
    *out = nullptr;
    return guarded("elb200_dm_view", [&] {
        auto h = std::make_unique<elb200_dm_s>();
        h->dtype = a->dtype;
        h->keep = a->keep;
        if (a->dtype == ELB200_F32) h->f = std::make_unique<DistMat<float>>(a->f->View(i0, i1, j0, j1));
        else if (a->dtype == ELB200_F64) h->d = std::make_unique<DistMat<double>>(a->d->View(i0, i1, j0, j1));
        else h->h = std::make_unique<DistMat<Half>>(a->h->View(i0, i1, j0, j1));
        *out = h.release();
    });
}

#define ELB_BINARY(NAME, CALL_F, CALL_D)                                                  \
    if (!same_dtype(A, B) || !level3_dtype(A)) return ELB200_ERR_ARG;                     \
    return guarded(NAME, [&] { if (A->dtype == ELB200_F32) { CALL_F; } else { CALL_D; } });

int elb200_dm_copy(elb200_dm_t A, elb200_dm_t B) {
    if (!same_dtype(A, B)) return ELB200_ERR_ARG;
    return guarded("elb200_dm_copy", [&] { on2(A, B, [&](auto& a, auto& b) { Copy(a, b); }); });
}
int elb200_dm_transpose(elb200_dm_t A, elb200_dm_t B) {
    if (!same_dtype(A, B)) return ELB200_ERR_ARG;
    return guarded("elb200_dm_transpose", [&] { on2(A, B, [&](auto& a, auto& b) { TransposeCopy(a, b); }); });
}
int elb200_dm_axpy(double alpha, elb200_dm_t A, elb200_dm_t B) {
    if (!same_dtype(A, B)) return ELB200_ERR_ARG;
    return guarded("elb200_dm_axpy", [&] { on2(A, B, [&](auto& a, auto& b) { Axpy(sc(a, alpha), a, b); }); });
}
int elb200_dm_axpy_contract(double alpha, elb200_dm_t A, elb200_dm_t B) {
    if (!same_dtype(A, B)) return ELB200_ERR_ARG;
    return guarded("elb200_dm_axpy_contract", [&] { on2(A, B, [&](auto& a, auto& b) { AxpyContract(sc(a, alpha), a, b); }); });
}
int elb200_dm_contract(elb200_dm_t A, elb200_dm_t B) {
    if (!same_dtype(A, B)) return ELB200_ERR_ARG;
    return guarded("elb200_dm_contract", [&] { on2(A, B, [&](auto& a, auto& b) { Contract(a, b); }); });
}
int elb200_dm_transpose_axpy_contract(double alpha, elb200_dm_t A, elb200_dm_t B) {
    if (!same_dtype(A, B)) return ELB200_ERR_ARG;
    return guarded("elb200_dm_transpose_axpy_contract", [&] { on2(A, B, [&](auto& a, auto& b) { TransposeAxpyContract(sc(a, alpha), a, b); }); });
}

int elb200_dm_scale(double alpha, elb200_dm_t A) {
    if (!A) return ELB200_ERR_ARG;
    return guarded("elb200_dm_scale", [&] { on(A, [&](auto& M) { Scale(sc(M, alpha), M); }); });
}
int elb200_dm_zero(elb200_dm_t A) {
    if (!A) return ELB200_ERR_ARG;
    return guarded("elb200_dm_zero", [&] { on(A, [&](auto& M) { Zero(M); }); });
}
int elb200_dm_local_gemm(int ta, int tb, double alpha, elb200_dm_t A, elb200_dm_t B, double beta, elb200_dm_t C) {
    if (!same_dtype(A, B) || !same_dtype(A, C)) return ELB200_ERR_ARG;
    return guarded("elb200_dm_local_gemm", [&] {
        on3(A, B, C, [&](auto& a, auto& b, auto& c) { LocalGemm(ta, tb, sc(a, alpha), a, b, sc(a, beta), c); });
    });
}
int elb200_dist_gemm(int ta, int tb, double alpha, elb200_dm_t A, elb200_dm_t B, double beta, elb200_dm_t C, int alg) {
    if (!same_dtype(A, B) || !same_dtype(A, C)) return ELB200_ERR_ARG;
    if (ta < 0 || ta > 2 || tb < 0 || tb > 2 || alg < 0 || alg > 8) return ELB200_ERR_ARG;
    return guarded("elb200_dist_gemm", [&] {
        // MakeMultiSync(SyncInfoFromMatrix(C), SyncInfoFromMatrix(A), SyncInfoFromMatrix(B)) of Gemm_impl<GPU>
        // (src/blas_like/level3/Gemm.cpp:172-175): the product runs on C's stream, which first waits for A's and B's;
        // afterwards their streams wait for it.
        StreamJoin join(dm_stream(C), dm_stream(A), dm_stream(B));
        on3(A, B, C, [&](auto& a, auto& b, auto& c) { Gemm(ta, tb, sc(a, alpha), a, b, sc(a, beta), c, alg); });
    });
}

// SetSyncInfo / SyncInfoFromMatrix on a DistMatrix's local matrix (include/El/core/Matrix/decl.hpp, hydrogen SyncInfo):
// the stream every operation WRITING this matrix is enqueued on. Defaults to the grid's compute stream.
int elb200_dm_set_stream(elb200_dm_t a, void* stream) {
    if (!a) return ELB200_ERR_ARG;
    cudaStream_t s = stream ? static_cast<cudaStream_t>(stream) : a->grid()->stream;
    on(a, [&](auto& M) { M.stream = s; });
    return ELB200_OK;
}
void* elb200_dm_stream(elb200_dm_t a) { return a ? dm_stream(a) : nullptr; }
int elb200_dm_align_cols(elb200_dm_t a, int64_t ca) {
    if (!a) return ELB200_ERR_ARG;
    return guarded("elb200_dm_align_cols", [&] { on(a, [&](auto& M) { M.AlignCols(ca); }); });
}
int elb200_dm_align_rows(elb200_dm_t a, int64_t ra) {
    if (!a) return ELB200_ERR_ARG;
    return guarded("elb200_dm_align_rows", [&] { on(a, [&](auto& M) { M.AlignRows(ra); }); });
}

int elb200_dist_gemm_host_f32(elb200_grid_t g, int ta, int tb, int64_t m, int64_t n, int64_t k, float alpha,
                              const float* A, int64_t lda, const float* B, int64_t ldb, float beta, float* C,
                              int64_t ldc) {
    if (!g || ta < 0 || ta > 2 || tb < 0 || tb > 2) return ELB200_ERR_ARG;
    return guarded("elb200_dist_gemm_host_f32", [&] { GemmHost<float>(*g->g, ta, tb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc); });
}
int elb200_dist_gemm_host_f64(elb200_grid_t g, int ta, int tb, int64_t m, int64_t n, int64_t k, double alpha,
                              const double* A, int64_t lda, const double* B, int64_t ldb, double beta, double* C,
                              int64_t ldc) {
    if (!g || ta < 0 || ta > 2 || tb < 0 || tb > 2) return ELB200_ERR_ARG;
    return guarded("elb200_dist_gemm_host_f64", [&] { GemmHost<double>(*g->g, ta, tb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc); });
}

// ---- triangular rank-k family (engine_level3.cpp) ---------------------------------------------------
int elb200_dm_scale_trapezoid(double alpha, int uplo, elb200_dm_t A, int64_t offset) {
    if (!level3_dtype(A)) return ELB200_ERR_ARG;
    return guarded("elb200_dm_scale_trapezoid", [&] { ELB_ON(A, ScaleTrapezoid(float(alpha), uplo, M, offset), ScaleTrapezoid(alpha, uplo, M, offset)); });
}
int elb200_dm_make_trapezoidal(int uplo, elb200_dm_t A, int64_t offset) {
    if (!level3_dtype(A)) return ELB200_ERR_ARG;
    return guarded("elb200_dm_make_trapezoidal", [&] { ELB_ON(A, MakeTrapezoidal(uplo, M, offset), MakeTrapezoidal(uplo, M, offset)); });
}
int elb200_dm_make_symmetric(int uplo, elb200_dm_t A) {
    if (!level3_dtype(A)) return ELB200_ERR_ARG;
    return guarded("elb200_dm_make_symmetric", [&] { ELB_ON(A, MakeSymmetric(uplo, M), MakeSymmetric(uplo, M)); });
}
int elb200_dm_axpy_trapezoid(int uplo, double alpha, elb200_dm_t A, elb200_dm_t B, int64_t offset) {
    ELB_BINARY("elb200_dm_axpy_trapezoid", AxpyTrapezoid(uplo, float(alpha), *A->f, *B->f, offset), AxpyTrapezoid(uplo, alpha, *A->d, *B->d, offset))
}
int elb200_dm_local_trrk(int uplo, int ta, int tb, double alpha, elb200_dm_t A, elb200_dm_t B, double beta, elb200_dm_t C) {
    if (!same_dtype(A, B) || !same_dtype(A, C) || !level3_dtype(A) || ta < 0 || ta > 2 || tb < 0 || tb > 2) return ELB200_ERR_ARG;
    return guarded("elb200_dm_local_trrk", [&] {
        if (A->dtype == ELB200_F32) LocalTrrk(uplo, ta, tb, float(alpha), *A->f, *B->f, float(beta), *C->f);
        else LocalTrrk(uplo, ta, tb, alpha, *A->d, *B->d, beta, *C->d);
    });
}
int elb200_dist_trrk(int uplo, int ta, int tb, double alpha, elb200_dm_t A, elb200_dm_t B, double beta, elb200_dm_t C) {
    if (!same_dtype(A, B) || !same_dtype(A, C) || !level3_dtype(A) || ta < 0 || ta > 2 || tb < 0 || tb > 2) return ELB200_ERR_ARG;
    return guarded("elb200_dist_trrk", [&] {
        if (A->dtype == ELB200_F32) Trrk(uplo, ta, tb, float(alpha), *A->f, *B->f, float(beta), *C->f);
        else Trrk(uplo, ta, tb, alpha, *A->d, *B->d, beta, *C->d);
    });
}
int elb200_dist_syrk(int uplo, int trans, double alpha, elb200_dm_t A, double beta, elb200_dm_t C) {
    if (!same_dtype(A, C) || !level3_dtype(A) || trans < 0 || trans > 2) return ELB200_ERR_ARG;
    return guarded("elb200_dist_syrk", [&] {
        if (A->dtype == ELB200_F32) Syrk(uplo, trans, float(alpha), *A->f, float(beta), *C->f);
        else Syrk(uplo, trans, alpha, *A->d, beta, *C->d);
    });
}
int elb200_dist_symm(int side, int uplo, double alpha, elb200_dm_t A, elb200_dm_t B, double beta, elb200_dm_t C) {
    if (!same_dtype(A, B) || !same_dtype(A, C) || !level3_dtype(A)) return ELB200_ERR_ARG;
    return guarded("elb200_dist_symm", [&] {
        if (A->dtype == ELB200_F32) Symm(side, uplo, float(alpha), *A->f, *B->f, float(beta), *C->f);
        else Symm(side, uplo, alpha, *A->d, *B->d, beta, *C->d);
    });
}

int elb200_dm_local_trsm(int side, int uplo, int trans, int diag, double alpha, elb200_dm_t A, elb200_dm_t B) {
    if (trans < 0 || trans > 2) return ELB200_ERR_ARG;
    ELB_BINARY("elb200_dm_local_trsm", LocalTrsm(side, uplo, trans, diag, float(alpha), *A->f, *B->f), LocalTrsm(side, uplo, trans, diag, alpha, *A->d, *B->d))
}
int elb200_dist_trsm(int side, int uplo, int trans, int diag, double alpha, elb200_dm_t A, elb200_dm_t B, int alg) {
    if (trans < 0 || trans > 2 || alg < 0 || alg > 3) return ELB200_ERR_ARG;
    ELB_BINARY("elb200_dist_trsm", Trsm(side, uplo, trans, diag, float(alpha), *A->f, *B->f), Trsm(side, uplo, trans, diag, alpha, *A->d, *B->d))
}

}  // extern "C"
