// Complex instantiations of the hot path — El::Gemm, the redistributions and the entrywise routines for
// Complex<float> / Complex<double> (reference: src/blas_like/level3/Gemm.cpp:504-506 and the EL_ENABLE complex
// instantiations behind include/El/macros/Instantiate.h) — built ON TOP of the real C ABI, "4M":
//
//   * a complex DistMatrix is stored PLANAR: a real and an imaginary DistMatrix with one distribution and alignment.
//     Every redistribution, contraction, view and alignment rule is then the real path — the one that is pinned to the
//     reference's per-rank outputs and measured on NVLink — applied to both planes; nothing new touches the wire.
//   * with op(A) = Ar' + i sa Ai', op(B) = Br' + i sb Bi' (sa, sb = -1 for ADJOINT):
//         T = Ar'Br' - sa sb Ai'Bi',   U = sb Ar'Bi' + sa Ai'Br',   C = alpha (T + iU) + beta C
//     is four real El::Gemm calls (the tcgen05 3xTF32 / DMMA kernels, SUMMA A/B/C/Dot with their pipelines) — exactly the
//     8 m n k real flops of a complex product — plus entrywise combinations; real alpha and beta need no temporary.
//   * the reference's interleaved (std::complex) layout exists at the boundary only: elb200_zdm_write_local / read_local
//     and the local elb200_gemm_c32 / _c64 de-interleave with the strided copy kernel.
// No new kernel: everything here calls exported elb200_* entry points, so the hostsim doubles cover it unchanged.
#include <cuda_runtime.h>

#include <memory>

#include "capi_handles.hpp"

namespace elb {
void set_last_error(const char* what, const char* detail);
}

struct elb200_zdm_s {
    int real_dtype = ELB200_F32;
    elb200_dm_t re = nullptr, im = nullptr;  // owned
    ~elb200_zdm_s() {
        elb200_dm_destroy(re);
        elb200_dm_destroy(im);
    }
};

namespace {

#define ZTRY(expr)                         \
    do {                                   \
        const int st_ = (expr);            \
        if (st_ != ELB200_OK) return st_;  \
    } while (0)

bool same_kind(elb200_zdm_t a, elb200_zdm_t b) { return a && b && a->real_dtype == b->real_dtype; }
int real_orient(int t) { return t == ELB200_OP_N ? ELB200_OP_N : ELB200_OP_T; }
double conj_sign(int t) { return t == ELB200_OP_C ? -1.0 : 1.0; }

// a real temporary with C's distribution, alignment and size
int temp_like(elb200_dm_t like, elb200_dm_t* out) {
    elb200_dm_info_t inf;
    ZTRY(elb200_dm_info(like, &inf));
    elb200_grid_s g{like->keep};
    ZTRY(elb200_dm_create(&g, inf.dtype, inf.col_dist, inf.row_dist, out));
    int st = elb200_dm_set_stream(*out, elb200_dm_stream(like));
    if (st == ELB200_OK && inf.col_dist == 5 /* CIRC */) st = elb200_dm_set_root(*out, inf.root);
    if (st == ELB200_OK) st = elb200_dm_align_with(*out, like);
    if (st == ELB200_OK) st = elb200_dm_resize(*out, inf.height, inf.width);
    if (st != ELB200_OK) { elb200_dm_destroy(*out); *out = nullptr; }
    return st;
}
struct Temp {  // RAII for the above
    elb200_dm_t h = nullptr;
    ~Temp() { elb200_dm_destroy(h); }
};

// Y (both planes) = (br + i bi) * Y + (ar + i ai) * (T + i U); T, U real matrices distributed like Y's planes (may be null
// when alpha == 0). beta == 0 never reads Y.
int combine(double ar, double ai, elb200_dm_t T, elb200_dm_t U, double br, double bi, elb200_zdm_t Y) {
    Temp old_re;
    if (bi != 0.0) {
        ZTRY(temp_like(Y->re, &old_re.h));
        ZTRY(elb200_dm_copy(Y->re, old_re.h));
    }
    if (br == 0.0 && bi == 0.0) {
        ZTRY(elb200_dm_zero(Y->re));
        ZTRY(elb200_dm_zero(Y->im));
    } else {
        if (br != 1.0) ZTRY(elb200_dm_scale(br, Y->re));
        if (bi != 0.0) ZTRY(elb200_dm_axpy(-bi, Y->im, Y->re));  // the old imaginary part
        if (br != 1.0) ZTRY(elb200_dm_scale(br, Y->im));
        if (bi != 0.0) ZTRY(elb200_dm_axpy(bi, old_re.h, Y->im));
    }
    if (T && U) {
        if (ar != 0.0) { ZTRY(elb200_dm_axpy(ar, T, Y->re)); ZTRY(elb200_dm_axpy(ar, U, Y->im)); }
        if (ai != 0.0) { ZTRY(elb200_dm_axpy(-ai, U, Y->re)); ZTRY(elb200_dm_axpy(ai, T, Y->im)); }
    }
    return ELB200_OK;
}

// the strided copy of the real ABI, by dtype
int copy2d(int dt, int64_t r, int64_t c, const void* src, int64_t srs, int64_t scs, void* dst, int64_t drs, int64_t dcs, void* s) {
    return dt == ELB200_F32 ? elb200_copy2d_f32(r, c, static_cast<const float*>(src), srs, scs, static_cast<float*>(dst), drs, dcs, s)
                            : elb200_copy2d_f64(r, c, static_cast<const double*>(src), srs, scs, static_cast<double*>(dst), drs, dcs, s);
}
size_t real_bytes(int dt) { return dt == ELB200_F32 ? sizeof(float) : sizeof(double); }

struct DevScratch {  // stream-ordered device scratch (the engine's pool)
    void* p = nullptr;
    cudaStream_t s;
    DevScratch(size_t bytes, cudaStream_t st) : s(st) {
        if (bytes && cudaMallocAsync(&p, bytes, s) != cudaSuccess) p = nullptr;
    }
    ~DevScratch() { if (p) cudaFreeAsync(p, s); }
};

// local matrix of a complex DistMatrix <-> interleaved buffer (host or device), through a device staging block
int move_local(elb200_zdm_t a, void* user, int64_t ld, bool write) {
    elb200_dm_info_t ir, ii;
    ZTRY(elb200_dm_info(a->re, &ir));
    ZTRY(elb200_dm_info(a->im, &ii));
    const int64_t lh = ir.local_height, lw = ir.local_width;
    if (lh == 0 || lw == 0) return ELB200_OK;
    if (!user || ld < lh) { elb::set_last_error("elb200_zdm_*_local", "null buffer or ld < local height"); return ELB200_ERR_ARG; }
    const size_t rb = real_bytes(a->real_dtype);
    cudaStream_t s = static_cast<cudaStream_t>(elb200_dm_stream(a->re));
    DevScratch stage(size_t(2 * lh * lw) * rb, s);
    if (!stage.p) { elb::set_last_error("elb200_zdm_*_local", "cudaMallocAsync failed"); return ELB200_ERR_CUDA; }
    char* st = static_cast<char*>(stage.p);
    if (write) {
        if (cudaMemcpy2DAsync(st, size_t(2 * lh) * rb, user, size_t(2 * ld) * rb, size_t(2 * lh) * rb, size_t(lw), cudaMemcpyDefault, s) != cudaSuccess)
            return ELB200_ERR_CUDA;
        ZTRY(copy2d(a->real_dtype, lh, lw, st, 2, 2 * lh, ir.buffer, 1, ir.ldim, s));
        ZTRY(copy2d(a->real_dtype, lh, lw, st + rb, 2, 2 * lh, ii.buffer, 1, ii.ldim, s));
    } else {
        ZTRY(copy2d(a->real_dtype, lh, lw, ir.buffer, 1, ir.ldim, st, 2, 2 * lh, s));
        ZTRY(copy2d(a->real_dtype, lh, lw, ii.buffer, 1, ii.ldim, st + rb, 2, 2 * lh, s));
        if (cudaMemcpy2DAsync(user, size_t(2 * ld) * rb, st, size_t(2 * lh) * rb, size_t(2 * lh) * rb, size_t(lw), cudaMemcpyDefault, s) != cudaSuccess)
            return ELB200_ERR_CUDA;
    }
    return ELB200_OK;
}

// ---- local complex GEMM on interleaved operands -----------------------------------------------------------------
template <typename R>
struct RealAbi;
template <>
struct RealAbi<float> {
    static constexpr int dt = ELB200_F32;
    static int gemm(int ta, int tb, int64_t m, int64_t n, int64_t k, float al, const float* A, int64_t lda, const float* B, int64_t ldb, float be,
                    float* C, int64_t ldc, void* s) { return elb200_gemm_f32(ta, tb, m, n, k, al, A, lda, B, ldb, be, C, ldc, s); }
    static int axpy(int64_t r, int64_t c, float al, const float* X, int64_t ldx, float* Y, int64_t ldy, void* s) { return elb200_axpy2d_f32(r, c, al, X, 1, ldx, Y, 1, ldy, s); }
    static int scale(int64_t r, int64_t c, float al, float* A, int64_t lda, void* s) { return elb200_scale2d_f32(r, c, al, A, lda, s); }
    static int fill(int64_t r, int64_t c, float v, float* A, int64_t lda, void* s) { return elb200_fill2d_f32(r, c, v, A, lda, s); }
};
template <>
struct RealAbi<double> {
    static constexpr int dt = ELB200_F64;
    static int gemm(int ta, int tb, int64_t m, int64_t n, int64_t k, double al, const double* A, int64_t lda, const double* B, int64_t ldb, double be,
                    double* C, int64_t ldc, void* s) { return elb200_gemm_f64(ta, tb, m, n, k, al, A, lda, B, ldb, be, C, ldc, s); }
    static int axpy(int64_t r, int64_t c, double al, const double* X, int64_t ldx, double* Y, int64_t ldy, void* s) { return elb200_axpy2d_f64(r, c, al, X, 1, ldx, Y, 1, ldy, s); }
    static int scale(int64_t r, int64_t c, double al, double* A, int64_t lda, void* s) { return elb200_scale2d_f64(r, c, al, A, lda, s); }
    static int fill(int64_t r, int64_t c, double v, double* A, int64_t lda, void* s) { return elb200_fill2d_f64(r, c, v, A, lda, s); }
};

template <typename R>
int gemm_complex(int ta, int tb, int64_t m, int64_t n, int64_t k, const R* alpha, const void* A, int64_t lda, const void* B, int64_t ldb,
                 const R* beta, void* C, int64_t ldc, void* stream) {
    using Abi = RealAbi<R>;
    if (ta < 0 || ta > 2 || tb < 0 || tb > 2 || m < 0 || n < 0 || k < 0 || !alpha || !beta) return ELB200_ERR_ARG;
    const int64_t ar_ = ta == ELB200_OP_N ? m : k, ac_ = ta == ELB200_OP_N ? k : m;  // A as stored
    const int64_t br_ = tb == ELB200_OP_N ? k : n, bc_ = tb == ELB200_OP_N ? n : k;
    if (lda < (ar_ > 1 ? ar_ : 1) || ldb < (br_ > 1 ? br_ : 1) || ldc < (m > 1 ? m : 1)) return ELB200_ERR_ARG;
    if (m == 0 || n == 0) return ELB200_OK;
    if ((!A || !B) && k > 0) return ELB200_ERR_ARG;
    if (!C) return ELB200_ERR_ARG;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const R al_r = alpha[0], al_i = alpha[1], be_r = beta[0], be_i = beta[1];
    const bool have_prod = k > 0 && (al_r != R(0) || al_i != R(0));
    const bool have_c = be_r != R(0) || be_i != R(0);
    // 16-byte aligned pitches for the planes (the TMA path of the real kernels wants them)
    auto pitch = [](int64_t rows) { const int64_t v = 16 / int64_t(sizeof(R)); return (rows + v - 1) / v * v > 0 ? (rows + v - 1) / v * v : v; };
    const int64_t pa = pitch(ar_), pb = pitch(br_), pc = pitch(m);
    const size_t na = have_prod ? size_t(pa * ac_) : 0, nb = have_prod ? size_t(pb * bc_) : 0, nc = size_t(pc * n);
    // planes: Ar Ai Br Bi | T U | Cr Ci
    DevScratch buf((2 * na + 2 * nb + 4 * nc) * sizeof(R), s);
    if (!buf.p) { elb::set_last_error("elb200_gemm_c*", "cudaMallocAsync failed"); return ELB200_ERR_CUDA; }
    R* Ar = static_cast<R*>(buf.p);
    R* Ai = Ar + na;
    R* Br = Ai + na;
    R* Bi = Br + nb;
    R* T = Bi + nb;
    R* U = T + nc;
    R* Cr = U + nc;
    R* Ci = Cr + nc;
    const R* Az = static_cast<const R*>(A);
    const R* Bz = static_cast<const R*>(B);
    R* Cz = static_cast<R*>(C);
    if (have_prod) {
        ZTRY(copy2d(Abi::dt, ar_, ac_, Az, 2, 2 * lda, Ar, 1, pa, s));
        ZTRY(copy2d(Abi::dt, ar_, ac_, Az + 1, 2, 2 * lda, Ai, 1, pa, s));
        ZTRY(copy2d(Abi::dt, br_, bc_, Bz, 2, 2 * ldb, Br, 1, pb, s));
        ZTRY(copy2d(Abi::dt, br_, bc_, Bz + 1, 2, 2 * ldb, Bi, 1, pb, s));
        const int oa = real_orient(ta), ob = real_orient(tb);
        const R sa = R(conj_sign(ta)), sb = R(conj_sign(tb));
        ZTRY(Abi::gemm(oa, ob, m, n, k, R(1), Ar, pa, Br, pb, R(0), T, pc, s));
        ZTRY(Abi::gemm(oa, ob, m, n, k, -sa * sb, Ai, pa, Bi, pb, R(1), T, pc, s));
        ZTRY(Abi::gemm(oa, ob, m, n, k, sb, Ar, pa, Bi, pb, R(0), U, pc, s));
        ZTRY(Abi::gemm(oa, ob, m, n, k, sa, Ai, pa, Br, pb, R(1), U, pc, s));
    }
    // new planes: Nr = br Cr - bi Ci + ar T - ai U,  Ni = br Ci + bi Cr + ar U + ai T  (beta == 0 never reads C)
    R* Nr = Cr;
    R* Ni = Ci;
    if (have_c) {
        // the old planes are loaded into Cr / Ci and turned into beta * C in place (one spare plane when beta is complex)
        ZTRY(copy2d(Abi::dt, m, n, Cz, 2, 2 * ldc, Cr, 1, pc, s));
        ZTRY(copy2d(Abi::dt, m, n, Cz + 1, 2, 2 * ldc, Ci, 1, pc, s));
        if (be_i != R(0)) {
            // the new imaginary plane needs the OLD real plane and vice versa: build br Ci + bi Cr in a spare first
            DevScratch spare(nc * sizeof(R), s);
            if (!spare.p) return ELB200_ERR_CUDA;
            R* Sp = static_cast<R*>(spare.p);
            ZTRY(copy2d(Abi::dt, m, n, Ci, 1, pc, Sp, 1, pc, s));            // Sp = Ci
            ZTRY(Abi::scale(m, n, be_r, Sp, pc, s));                           // Sp = br Ci
            ZTRY(Abi::axpy(m, n, be_i, Cr, pc, Sp, pc, s));                    // Sp += bi Cr
            ZTRY(Abi::scale(m, n, be_r, Cr, pc, s));                           // Cr = br Cr
            ZTRY(Abi::axpy(m, n, -be_i, Ci, pc, Cr, pc, s));                   // Cr -= bi Ci (old Ci)
            ZTRY(copy2d(Abi::dt, m, n, Sp, 1, pc, Ci, 1, pc, s));            // Ci = Sp
        } else if (be_r != R(1)) {
            ZTRY(Abi::scale(m, n, be_r, Cr, pc, s));
            ZTRY(Abi::scale(m, n, be_r, Ci, pc, s));
        }
    } else {
        ZTRY(Abi::fill(m, n, R(0), Cr, pc, s));
        ZTRY(Abi::fill(m, n, R(0), Ci, pc, s));
    }
    if (have_prod) {
        if (al_r != R(0)) { ZTRY(Abi::axpy(m, n, al_r, T, pc, Nr, pc, s)); ZTRY(Abi::axpy(m, n, al_r, U, pc, Ni, pc, s)); }
        if (al_i != R(0)) { ZTRY(Abi::axpy(m, n, -al_i, U, pc, Nr, pc, s)); ZTRY(Abi::axpy(m, n, al_i, T, pc, Ni, pc, s)); }
    }
    ZTRY(copy2d(Abi::dt, m, n, Nr, 1, pc, Cz, 2, 2 * ldc, s));
    ZTRY(copy2d(Abi::dt, m, n, Ni, 1, pc, Cz + 1, 2, 2 * ldc, s));
    return ELB200_OK;
}

}  // namespace

extern "C" {

int elb200_zdm_create(elb200_grid_t g, int real_dtype, int col_dist, int row_dist, elb200_zdm_t* out) {
    if (!g || !out || (real_dtype != ELB200_F32 && real_dtype != ELB200_F64)) return ELB200_ERR_ARG;
    *out = nullptr;
    auto h = std::make_unique<elb200_zdm_s>();
    h->real_dtype = real_dtype;
    ZTRY(elb200_dm_create(g, real_dtype, col_dist, row_dist, &h->re));
    ZTRY(elb200_dm_create(g, real_dtype, col_dist, row_dist, &h->im));
    *out = h.release();
    return ELB200_OK;
}
int elb200_zdm_destroy(elb200_zdm_t a) {
    delete a;
    return ELB200_OK;
}
int elb200_zdm_plane(elb200_zdm_t a, int imag, elb200_dm_t* out) {
    if (!a || !out) return ELB200_ERR_ARG;
    *out = imag ? a->im : a->re;
    return ELB200_OK;
}
int elb200_zdm_resize(elb200_zdm_t a, int64_t h, int64_t w) {
    if (!a) return ELB200_ERR_ARG;
    ZTRY(elb200_dm_resize(a->re, h, w));
    return elb200_dm_resize(a->im, h, w);
}
int elb200_zdm_align(elb200_zdm_t a, int64_t ca, int64_t ra) {
    if (!a) return ELB200_ERR_ARG;
    ZTRY(elb200_dm_align(a->re, ca, ra));
    return elb200_dm_align(a->im, ca, ra);
}
int elb200_zdm_align_with(elb200_zdm_t a, elb200_zdm_t o) {
    if (!a || !o) return ELB200_ERR_ARG;
    ZTRY(elb200_dm_align_with(a->re, o->re));
    return elb200_dm_align_with(a->im, o->re);
}
int elb200_zdm_set_root(elb200_zdm_t a, int root) {
    if (!a) return ELB200_ERR_ARG;
    ZTRY(elb200_dm_set_root(a->re, root));
    return elb200_dm_set_root(a->im, root);
}
int elb200_zdm_set_stream(elb200_zdm_t a, void* stream) {
    if (!a) return ELB200_ERR_ARG;
    ZTRY(elb200_dm_set_stream(a->re, stream));
    return elb200_dm_set_stream(a->im, stream);
}
int elb200_zdm_view(elb200_zdm_t a, int64_t i0, int64_t i1, int64_t j0, int64_t j1, elb200_zdm_t* out) {
    if (!a || !out) return ELB200_ERR_ARG;
    *out = nullptr;
    auto h = std::make_unique<elb200_zdm_s>();
    h->real_dtype = a->real_dtype;
    ZTRY(elb200_dm_view(a->re, i0, i1, j0, j1, &h->re));
    ZTRY(elb200_dm_view(a->im, i0, i1, j0, j1, &h->im));
    *out = h.release();
    return ELB200_OK;
}
int elb200_zdm_write_local(elb200_zdm_t a, const void* interleaved, int64_t ld) {
    if (!a) return ELB200_ERR_ARG;
    return move_local(a, const_cast<void*>(interleaved), ld, true);
}
int elb200_zdm_read_local(elb200_zdm_t a, void* interleaved, int64_t ld) {
    if (!a) return ELB200_ERR_ARG;
    return move_local(a, interleaved, ld, false);
}
int elb200_zdm_copy(elb200_zdm_t A, elb200_zdm_t B) {
    if (!same_kind(A, B)) return ELB200_ERR_ARG;
    ZTRY(elb200_dm_copy(A->re, B->re));
    return elb200_dm_copy(A->im, B->im);
}
int elb200_zdm_transpose(elb200_zdm_t A, elb200_zdm_t B, int conjugate) {
    if (!same_kind(A, B)) return ELB200_ERR_ARG;
    ZTRY(elb200_dm_transpose(A->re, B->re));
    ZTRY(elb200_dm_transpose(A->im, B->im));
    return conjugate ? elb200_dm_scale(-1.0, B->im) : ELB200_OK;
}
int elb200_zdm_zero(elb200_zdm_t A) {
    if (!A) return ELB200_ERR_ARG;
    ZTRY(elb200_dm_zero(A->re));
    return elb200_dm_zero(A->im);
}
int elb200_zdm_scale(double ar, double ai, elb200_zdm_t A) {
    if (!A) return ELB200_ERR_ARG;
    return combine(0.0, 0.0, nullptr, nullptr, ar, ai, A);
}
int elb200_zdm_axpy(double ar, double ai, elb200_zdm_t X, elb200_zdm_t Y) {
    if (!same_kind(X, Y)) return ELB200_ERR_ARG;
    // Y += (ar + i ai)(Xr + i Xi); elb200_dm_axpy redistributes X like Y when the distributions differ
    if (ar != 0.0) { ZTRY(elb200_dm_axpy(ar, X->re, Y->re)); ZTRY(elb200_dm_axpy(ar, X->im, Y->im)); }
    if (ai != 0.0) { ZTRY(elb200_dm_axpy(-ai, X->im, Y->re)); ZTRY(elb200_dm_axpy(ai, X->re, Y->im)); }
    return ELB200_OK;
}
int elb200_zdm_axpy_contract(double ar, double ai, elb200_zdm_t A, elb200_zdm_t B) {
    if (!same_kind(A, B)) return ELB200_ERR_ARG;
    if (ar != 0.0) { ZTRY(elb200_dm_axpy_contract(ar, A->re, B->re)); ZTRY(elb200_dm_axpy_contract(ar, A->im, B->im)); }
    if (ai != 0.0) { ZTRY(elb200_dm_axpy_contract(-ai, A->im, B->re)); ZTRY(elb200_dm_axpy_contract(ai, A->re, B->im)); }
    return ELB200_OK;
}

int elb200_zdm_gemm(int ta, int tb, double ar, double ai, elb200_zdm_t A, elb200_zdm_t B, double br, double bi, elb200_zdm_t C, int alg) {
    if (!same_kind(A, B) || !same_kind(A, C)) return ELB200_ERR_ARG;
    if (ta < 0 || ta > 2 || tb < 0 || tb > 2 || alg < 0 || alg > 8) return ELB200_ERR_ARG;
    const int oa = real_orient(ta), ob = real_orient(tb);
    const double sa = conj_sign(ta), sb = conj_sign(tb);
    if (ai == 0.0 && bi == 0.0) {
        // real alpha and beta: the four products go straight into C's planes
        ZTRY(elb200_dist_gemm(oa, ob, ar, A->re, B->re, br, C->re, alg));
        ZTRY(elb200_dist_gemm(oa, ob, -ar * sa * sb, A->im, B->im, 1.0, C->re, alg));
        ZTRY(elb200_dist_gemm(oa, ob, ar * sb, A->re, B->im, br, C->im, alg));
        return elb200_dist_gemm(oa, ob, ar * sa, A->im, B->re, 1.0, C->im, alg);
    }
    Temp T, U;
    ZTRY(temp_like(C->re, &T.h));
    ZTRY(temp_like(C->re, &U.h));
    ZTRY(elb200_dist_gemm(oa, ob, 1.0, A->re, B->re, 0.0, T.h, alg));
    ZTRY(elb200_dist_gemm(oa, ob, -sa * sb, A->im, B->im, 1.0, T.h, alg));
    ZTRY(elb200_dist_gemm(oa, ob, sb, A->re, B->im, 0.0, U.h, alg));
    ZTRY(elb200_dist_gemm(oa, ob, sa, A->im, B->re, 1.0, U.h, alg));
    return combine(ar, ai, T.h, U.h, br, bi, C);
}

int elb200_gemm_c32(int ta, int tb, int64_t m, int64_t n, int64_t k, const float* alpha, const void* A, int64_t lda, const void* B, int64_t ldb,
                    const float* beta, void* C, int64_t ldc, void* stream) {
    return gemm_complex<float>(ta, tb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, stream);
}
int elb200_gemm_c64(int ta, int tb, int64_t m, int64_t n, int64_t k, const double* alpha, const void* A, int64_t lda, const void* B, int64_t ldb,
                    const double* beta, void* C, int64_t ldc, void* stream) {
    return gemm_complex<double>(ta, tb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, stream);
}

}  // extern "C"
