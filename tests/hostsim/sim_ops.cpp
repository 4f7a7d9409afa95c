// hostsim: test doubles of libelb200's DEVICE entry points (the kernels of blas1.cu, gemm_f32.cu,
// gemm_f64.cu and the glue of capi_core.cu), same C signatures and argument checks, executed as
// stream-ordered host closures (see sim.hpp). They exist so that the host ENGINE can be run without
// a GPU in tests; they say nothing about the CUDA kernels, which are checked on a B200 by the
// `-m gpu` tests. Every access is bounds-checked against the simulated device allocations.
#include <atomic>
#include <cstdio>
#include <cstring>
#include <chrono>
#include <thread>
#include <vector>

#include "../../include/elb200.h"
#include "../../elemental_b200/csrc/trsm_rec.hpp"
#include "sim.hpp"

namespace elb {

thread_local char g_last_error[512] = "";
std::atomic<unsigned long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
void set_last_error(const char* what, const char* detail) {
    std::snprintf(g_last_error, sizeof(g_last_error), "%s: %s", what, detail ? detail : "");
}
int device_sm_count() { return 148; }

}  // namespace elb

namespace {

using elb::count_launch;

cudaStream_t st(void* s) { return static_cast<cudaStream_t>(s); }

// byte span touched by a rows x cols lattice with element strides (rs, cs) starting at p
template <typename T>
void check_lattice(const T* p, int64_t rows, int64_t cols, int64_t rs, int64_t cs, const char* what) {
    if (rows <= 0 || cols <= 0) return;
    if (rs < 0 || cs < 0) sim::fail("%s: negative stride", what);
    const int64_t last = (rows - 1) * rs + (cols - 1) * cs;
    sim::check_device_range(p, size_t(last + 1) * sizeof(T), what);
}

template <typename T>
int copy2d(int64_t rows, int64_t cols, const T* src, int64_t srs, int64_t scs, T* dst, int64_t drs, int64_t dcs, void* s) {
    if (rows < 0 || cols < 0) return ELB200_ERR_ARG;
    if (rows == 0 || cols == 0) return ELB200_OK;
    count_launch();
    sim::enqueue(st(s), [=] {
        check_lattice(src, rows, cols, srs, scs, "copy2d src");
        check_lattice(dst, rows, cols, drs, dcs, "copy2d dst");
        sim::vt_op(sim::RES_HBM, sim::kLaunch + 2.0 * double(rows) * double(cols) * sizeof(T) / sim::kRateHbm);
        if (sim::dry_run()) return;
        for (int64_t j = 0; j < cols; ++j)
            for (int64_t i = 0; i < rows; ++i) dst[i * drs + j * dcs] = src[i * srs + j * scs];
    });
    return ELB200_OK;
}

template <typename T>
int axpy2d(int64_t rows, int64_t cols, T alpha, const T* X, int64_t xrs, int64_t xcs, T* Y, int64_t yrs, int64_t ycs, void* s) {
    if (rows < 0 || cols < 0) return ELB200_ERR_ARG;
    if (rows == 0 || cols == 0) return ELB200_OK;
    count_launch();
    sim::enqueue(st(s), [=] {
        check_lattice(X, rows, cols, xrs, xcs, "axpy2d X");
        check_lattice(Y, rows, cols, yrs, ycs, "axpy2d Y");
        sim::vt_op(sim::RES_HBM, sim::kLaunch + 3.0 * double(rows) * double(cols) * sizeof(T) / sim::kRateHbm);
        if (sim::dry_run()) return;
        for (int64_t j = 0; j < cols; ++j)
            for (int64_t i = 0; i < rows; ++i) Y[i * yrs + j * ycs] += alpha * X[i * xrs + j * xcs];
    });
    return ELB200_OK;
}

// B(cols x rows, ldb) = A(rows x cols, lda)^T, or B += alpha * A^T
template <typename T, bool AXPY>
int transpose(int64_t rows, int64_t cols, const T* A, int64_t lda, T* B, int64_t ldb, T alpha, void* s) {
    if (rows < 0 || cols < 0 || lda < (rows > 1 ? rows : 1) || ldb < (cols > 1 ? cols : 1)) return ELB200_ERR_ARG;
    if (rows == 0 || cols == 0) return ELB200_OK;
    count_launch();
    sim::enqueue(st(s), [=] {
        check_lattice(A, rows, cols, 1, lda, "transpose A");
        check_lattice(B, cols, rows, 1, ldb, "transpose B");
        sim::vt_op(sim::RES_HBM, sim::kLaunch + (AXPY ? 3.0 : 2.0) * double(rows) * double(cols) * sizeof(T) / sim::kRateHbm);
        if (sim::dry_run()) return;
        for (int64_t j = 0; j < cols; ++j)
            for (int64_t i = 0; i < rows; ++i) {
                if (AXPY) B[j + i * ldb] += alpha * A[i + j * lda];
                else B[j + i * ldb] = A[i + j * lda];
            }
    });
    return ELB200_OK;
}

template <typename T>
int fill2d(int64_t rows, int64_t cols, T v, T* A, int64_t lda, void* s) {
    if (rows < 0 || cols < 0) return ELB200_ERR_ARG;
    if (rows == 0 || cols == 0) return ELB200_OK;
    count_launch();
    sim::enqueue(st(s), [=] {
        check_lattice(A, rows, cols, 1, lda, "fill2d");
        sim::vt_op(sim::RES_HBM, sim::kLaunch + 1.0 * double(rows) * double(cols) * sizeof(T) / sim::kRateHbm);
        if (sim::dry_run()) return;
        for (int64_t j = 0; j < cols; ++j)
            for (int64_t i = 0; i < rows; ++i) A[i + j * lda] = v;
    });
    return ELB200_OK;
}

template <typename T>
int scale2d(int64_t rows, int64_t cols, T alpha, T* A, int64_t lda, void* s) {
    if (rows < 0 || cols < 0) return ELB200_ERR_ARG;
    if (alpha == T(1)) return ELB200_OK;
    if (alpha == T(0)) return fill2d<T>(rows, cols, T(0), A, lda, s);
    if (rows == 0 || cols == 0) return ELB200_OK;
    count_launch();
    sim::enqueue(st(s), [=] {
        check_lattice(A, rows, cols, 1, lda, "scale2d");
        sim::vt_op(sim::RES_HBM, sim::kLaunch + 2.0 * double(rows) * double(cols) * sizeof(T) / sim::kRateHbm);
        if (sim::dry_run()) return;
        for (int64_t j = 0; j < cols; ++j)
            for (int64_t i = 0; i < rows; ++i) A[i + j * lda] *= alpha;
    });
    return ELB200_OK;
}

template <typename T>
int lattice(int mode, int nseg, const elb200_segment* segs, T* M, int64_t ldm, T* buf, T alpha, void* s) {
    if (mode < ELB200_PACK || mode > ELB200_UNPACK_AXPY || nseg < 0 || nseg > ELB200_MAX_SEGMENTS || (nseg > 0 && !segs))
        return ELB200_ERR_ARG;
    std::vector<elb200_segment> tab;
    for (int i = 0; i < nseg; ++i) {
        const elb200_segment& g = segs[i];
        if (g.rows < 0 || g.cols < 0 || g.row_step < 1 || g.col_step < 1 || g.row0 < 0 || g.col0 < 0 || g.buf_offset < 0)
            return ELB200_ERR_ARG;
        if (g.rows > 0 && g.cols > 0) tab.push_back(g);
    }
    if (tab.empty()) return ELB200_OK;
    if (!M || !buf || ldm < 1) return ELB200_ERR_ARG;
    count_launch();
    sim::enqueue(st(s), [=] {
        double elems = 0;
        for (const elb200_segment& g : tab) elems += double(g.rows) * double(g.cols);
        sim::vt_op(sim::RES_HBM, sim::kLaunch + (mode == ELB200_UNPACK_AXPY ? 3.0 : 2.0) * elems * sizeof(T) / sim::kRateHbm);
        for (const elb200_segment& g : tab) {
            check_lattice(M + g.row0 + g.col0 * ldm, g.rows, g.cols, g.row_step, g.col_step * ldm, "lattice M");
            sim::check_device_range(buf + g.buf_offset, size_t(g.rows) * size_t(g.cols) * sizeof(T), "lattice staging");
            if (sim::dry_run()) continue;
            for (int64_t j = 0; j < g.cols; ++j)
                for (int64_t i = 0; i < g.rows; ++i) {
                    T* mp = M + g.row0 + i * g.row_step + (g.col0 + j * g.col_step) * ldm;
                    T* bp = buf + g.buf_offset + i + j * g.rows;
                    if (mode == ELB200_PACK) *bp = *mp;
                    else if (mode == ELB200_UNPACK) *mp = *bp;
                    else *mp += alpha * *bp;
                }
        }
    });
    return ELB200_OK;
}

template <typename T>
int peer_reduce_axpy(int nsrc, const T* const* srcs, int64_t off, const elb200_segment* seg, T* M, int64_t ldm, T alpha, void* s) {
    if (nsrc < 1 || nsrc > ELB200_MAX_SEGMENTS || !srcs || !seg || seg->rows < 0 || seg->cols < 0 || seg->row_step < 1 ||
        seg->col_step < 1)
        return ELB200_ERR_ARG;
    if (seg->rows == 0 || seg->cols == 0) return ELB200_OK;
    std::vector<const T*> src(srcs, srcs + nsrc);
    const elb200_segment g = *seg;
    count_launch();
    sim::enqueue(st(s), [=] {
        check_lattice(M + g.row0 + g.col0 * ldm, g.rows, g.cols, g.row_step, g.col_step * ldm, "peer_reduce_axpy M");
        for (const T* p : src) sim::check_device_range(p + off, size_t(g.rows) * size_t(g.cols) * sizeof(T), "peer_reduce_axpy source");
        {
            const double bytes = double(g.rows) * double(g.cols) * sizeof(T);
            const double t_link = double(src.size() > 1 ? src.size() - 1 : 0) * bytes / sim::kRateNvlink, t_hbm = 3.0 * bytes / sim::kRateHbm;
            sim::vt_op(sim::RES_HBM, sim::kLaunch + (t_link > t_hbm ? t_link : t_hbm));
        }
        if (sim::dry_run()) return;
        for (int64_t j = 0; j < g.cols; ++j)
            for (int64_t i = 0; i < g.rows; ++i) {
                const int64_t e = off + i + j * g.rows;
                T sum = T(0);
                for (const T* p : src) sum += p[e];
                T* mp = M + g.row0 + i * g.row_step + (g.col0 + j * g.col_step) * ldm;
                *mp += alpha * sum;
            }
    });
    return ELB200_OK;
}

template <typename T>
int trapezoid(int op, int uplo, int64_t offset, int64_t rows, int64_t cols, int64_t row0, int64_t row_step, int64_t col0,
              int64_t col_step, T alpha, const T* X, int64_t ldx, T* Y, int64_t ldy, void* s) {
    if (op < ELB200_TRAP_SCALE || op > ELB200_TRAP_MAKE || uplo < 0 || uplo > 1 || rows < 0 || cols < 0 || row_step < 1 ||
        col_step < 1 || row0 < 0 || col0 < 0)
        return ELB200_ERR_ARG;
    if (rows == 0 || cols == 0) return ELB200_OK;
    if (!Y || ldy < rows || (op == ELB200_TRAP_AXPY && (!X || ldx < rows))) return ELB200_ERR_ARG;
    if (op == ELB200_TRAP_SCALE && alpha == T(1)) return ELB200_OK;
    count_launch();
    sim::enqueue(st(s), [=] {
        check_lattice(Y, rows, cols, 1, ldy, "trapezoid Y");
        if (op == ELB200_TRAP_AXPY) check_lattice(X, rows, cols, 1, ldx, "trapezoid X");
        sim::vt_op(sim::RES_HBM, sim::kLaunch + (op == ELB200_TRAP_AXPY ? 1.5 : op == ELB200_TRAP_SCALE ? 1.0 : 0.5) * double(rows) * double(cols) * sizeof(T) / sim::kRateHbm);
        if (sim::dry_run()) return;
        for (int64_t j = 0; j < cols; ++j)
            for (int64_t i = 0; i < rows; ++i) {
                const int64_t gi = row0 + i * row_step, gj = col0 + j * col_step;
                const bool inside = uplo == ELB200_UPPER ? gi <= gj - offset : gi >= gj - offset;
                T* y = Y + i + j * ldy;
                if (op == ELB200_TRAP_SCALE) { if (inside) *y = alpha == T(0) ? T(0) : alpha * *y; }
                else if (op == ELB200_TRAP_AXPY) { if (inside) *y += alpha * X[i + j * ldx]; }
                else if (!inside) *y = T(0);
            }
    });
    return ELB200_OK;
}

// Base-case solve (triangle <= kTrsmBase): straightforward substitution in double precision, one right-hand
// side at a time, enqueued like a kernel. alpha has been applied by the caller.
template <typename T>
int trsm_base(bool left, bool lower, bool tr, bool unit, int64_t m, int64_t n, const T* A, int64_t lda, T* B, int64_t ldb, void* s) {
    const int64_t na = left ? m : n;
    if (na > elb::kTrsmBase) sim::fail("trsm base case called with a %lld x %lld triangle", (long long)na, (long long)na);
    count_launch();
    sim::enqueue(st(s), [=] {
        check_lattice(A, na, na, 1, lda, "trsm A");
        check_lattice(B, m, n, 1, ldb, "trsm B");
        sim::vt_op(sim::RES_TENSOR, sim::kLaunch + double(na) * double(na) * double(left ? n : m) / 4e12);  // CUDA-core substitution
        if (sim::dry_run()) return;
        const bool eff_lower = lower != tr;
        // E(i,j) = op(A)(i,j), read only inside the stored triangle (and never on a UNIT diagonal)
        auto E = [&](int64_t i, int64_t j) { return double(tr ? A[j + i * lda] : A[i + j * lda]); };
        std::vector<double> x(static_cast<size_t>(na));
        const int64_t nrhs = left ? n : m;
        // LEFT: E x = b. RIGHT: x^T E = b^T  <=>  E^T x = b, i.e. the same substitution with E transposed.
        const bool fwd = left ? eff_lower : !eff_lower;
        for (int64_t r = 0; r < nrhs; ++r) {
            auto b = [&](int64_t i) -> T& { return left ? B[i + r * ldb] : B[r + i * ldb]; };
            for (int64_t t = 0; t < na; ++t) {
                const int64_t i = fwd ? t : na - 1 - t;
                double sum = double(b(i));
                if (fwd) for (int64_t j = 0; j < i; ++j) sum -= (left ? E(i, j) : E(j, i)) * x[j];
                else for (int64_t j = i + 1; j < na; ++j) sum -= (left ? E(i, j) : E(j, i)) * x[j];
                x[i] = unit ? sum : sum / E(i, i);
            }
            for (int64_t i = 0; i < na; ++i) b(i) = T(x[i]);
        }
    });
    return ELB200_OK;
}

template <typename T>
int gemm(int ta, int tb, int64_t m, int64_t n, int64_t k, T alpha, const T* A, int64_t lda, const T* B, int64_t ldb, T beta,
         T* C, int64_t ldc, void* s);

// The REAL recursion of the product (elemental_b200/csrc/trsm_rec.hpp) over the doubles above.
template <typename T>
struct SimBackend {
    void* s;
    int gemm(int ta, int tb, int64_t m, int64_t n, int64_t k, T al, const T* A, int64_t lda, const T* B, int64_t ldb, T be, T* C,
             int64_t ldc) { return ::gemm<T>(ta, tb, m, n, k, al, A, lda, B, ldb, be, C, ldc, s); }
    int base(bool left, bool lower, bool trans, bool unit, int64_t m, int64_t n, const T* A, int64_t lda, T* B, int64_t ldb) {
        return trsm_base<T>(left, lower, trans, unit, m, n, A, lda, B, ldb, s);
    }
};

template <typename T>
int trsm(int side, int uplo, int trans, int diag, int64_t m, int64_t n, T alpha, const T* A, int64_t lda, T* B, int64_t ldb, void* s) {
    if (side < 0 || side > 1 || uplo < 0 || uplo > 1 || trans < 0 || trans > 2 || diag < 0 || diag > 1 || m < 0 || n < 0)
        return ELB200_ERR_ARG;
    const int64_t na = side == 0 ? m : n;
    if (lda < (na > 1 ? na : 1) || ldb < (m > 1 ? m : 1)) return ELB200_ERR_ARG;
    if (m == 0 || n == 0) return ELB200_OK;
    if (!A || !B) return ELB200_ERR_ARG;
    int stt = scale2d<T>(m, n, alpha, B, ldb, s);
    if (stt != ELB200_OK || alpha == T(0)) return stt;
    SimBackend<T> be{s};
    return elb::trsm_recursive<T>(be, side == 0, uplo == ELB200_LOWER, trans != 0, diag == ELB200_UNIT, m, n, A, lda, B, ldb);
}

int check_gemm_args(int transA, int transB, int64_t m, int64_t n, int64_t k, const void* A, int64_t lda, const void* B,
                    int64_t ldb, const void* C, int64_t ldc) {
    if (transA < 0 || transA > 2 || transB < 0 || transB > 2) return ELB200_ERR_ARG;
    if (m < 0 || n < 0 || k < 0) return ELB200_ERR_ARG;
    const int64_t a_rows = transA ? k : m, b_rows = transB ? n : k;
    if (lda < (a_rows > 1 ? a_rows : 1) || ldb < (b_rows > 1 ? b_rows : 1) || ldc < (m > 1 ? m : 1)) return ELB200_ERR_ARG;
    if (m > 0 && n > 0 && !C) return ELB200_ERR_ARG;
    if (m > 0 && n > 0 && k > 0 && (!A || !B)) return ELB200_ERR_ARG;
    return ELB200_OK;
}

// plain triple loop in double precision (axpy form over packed column-major op(A))
template <typename T>
int gemm(int ta, int tb, int64_t m, int64_t n, int64_t k, T alpha, const T* A, int64_t lda, const T* B, int64_t ldb, T beta,
         T* C, int64_t ldc, void* s) {
    int stt = check_gemm_args(ta, tb, m, n, k, A, lda, B, ldb, C, ldc);
    if (stt != ELB200_OK) return stt;
    if (m == 0 || n == 0) return ELB200_OK;
    if (k == 0 || alpha == T(0)) return scale2d<T>(m, n, beta, C, ldc, s);
    count_launch();
    sim::enqueue(st(s), [=] {
        check_lattice(A, ta ? k : m, ta ? m : k, 1, lda, "gemm A");
        check_lattice(B, tb ? n : k, tb ? k : n, 1, ldb, "gemm B");
        check_lattice(C, m, n, 1, ldc, "gemm C");
        sim::vt_op(sim::RES_TENSOR, 2.0 * sim::kLaunch + 2.0 * double(m) * double(n) * double(k) / (sizeof(T) == 4 ? sim::kRateTensorF32 : sim::kRateTensorF64));
        if (sim::dry_run()) return;
        std::vector<double> a(static_cast<size_t>(m * k));
        std::vector<double> acc(static_cast<size_t>(m), 0.0);
        for (int64_t l = 0; l < k; ++l)
            for (int64_t i = 0; i < m; ++i) a[size_t(i) + size_t(l) * m] = double(ta ? A[l + i * lda] : A[i + l * lda]);
        for (int64_t j = 0; j < n; ++j) {
            std::fill(acc.begin(), acc.end(), 0.0);
            for (int64_t l = 0; l < k; ++l) {
                const double b = double(tb ? B[j + l * ldb] : B[l + j * ldb]);
                const double* col = a.data() + size_t(l) * m;
                for (int64_t i = 0; i < m; ++i) acc[i] += col[i] * b;
            }
            for (int64_t i = 0; i < m; ++i) {
                T* c = C + i + j * ldc;
                *c = beta == T(0) ? T(double(alpha) * acc[i]) : T(double(alpha) * acc[i] + double(beta) * double(*c));
            }
        }
    });
    return ELB200_OK;
}

}  // namespace

extern "C" {

const char* elb200_status_string(int status) {
    switch (status) {
        case ELB200_OK: return "ok";
        case ELB200_ERR_CUDA: return "CUDA error";
        case ELB200_ERR_ARG: return "invalid argument";
        case ELB200_ERR_TENSORMAP: return "tensor map encode failed";
        case ELB200_ERR_NCCL: return "NCCL error";
        case ELB200_ERR_STATE: return "library not initialised";
        default: return "unknown status";
    }
}
const char* elb200_last_error(void) { return elb::g_last_error; }
unsigned long long elb200_launch_count(void) { return elb::g_launches.load(std::memory_order_relaxed); }
const char* elb200_build_info(void) { return "hostsim: host test double of libelb200's device layer (tests only, no GPU code)"; }

int elb200_gemm_f32(int ta, int tb, int64_t m, int64_t n, int64_t k, float al, const float* A, int64_t lda, const float* B,
                    int64_t ldb, float be, float* C, int64_t ldc, void* s) {
    return gemm<float>(ta, tb, m, n, k, al, A, lda, B, ldb, be, C, ldc, s);
}
int elb200_gemm_f32_ex(int ta, int tb, int64_t m, int64_t n, int64_t k, float al, const float* A, int64_t lda, const float* B,
                       int64_t ldb, float be, float* C, int64_t ldc, int, void* s) {
    return gemm<float>(ta, tb, m, n, k, al, A, lda, B, ldb, be, C, ldc, s);
}
int elb200_gemm_f32_noalloc(int ta, int tb, int64_t m, int64_t n, int64_t k, float al, const float* A, int64_t lda, const float* B,
                            int64_t ldb, float be, float* C, int64_t ldc, void* s) {
    return gemm<float>(ta, tb, m, n, k, al, A, lda, B, ldb, be, C, ldc, s);
}
// half entries: g++'s _Float16 is IEEE binary16; arithmetic in float like the device kernels
typedef _Float16 simhalf;
int elb200_gemm_f16(int ta, int tb, int64_t m, int64_t n, int64_t k, float al, const void* Av, int64_t lda, const void* Bv, int64_t ldb,
                    float be, void* Cv, int64_t ldc, void* s) {
    const simhalf* A = static_cast<const simhalf*>(Av); const simhalf* B = static_cast<const simhalf*>(Bv); simhalf* C = static_cast<simhalf*>(Cv);
    sim::enqueue(static_cast<cudaStream_t>(s), [=] {
        for (int64_t j = 0; j < n; ++j)
            for (int64_t i = 0; i < m; ++i) {
                float acc = 0.f;
                for (int64_t l = 0; l < k; ++l) acc += float(ta ? A[l + i * lda] : A[i + l * lda]) * float(tb ? B[j + l * ldb] : B[l + j * ldb]);
                const float old = be == 0.f ? 0.f : be * float(C[i + j * ldc]);
                C[i + j * ldc] = simhalf(al * acc + old);
            }
    });
    return ELB200_OK;
}
int elb200_copy2d_f16(int64_t r, int64_t c, const void* src, int64_t srs, int64_t scs, void* dst, int64_t drs, int64_t dcs, void* s) {
    const simhalf* X = static_cast<const simhalf*>(src); simhalf* Y = static_cast<simhalf*>(dst);
    sim::enqueue(static_cast<cudaStream_t>(s), [=] { for (int64_t j = 0; j < c; ++j) for (int64_t i = 0; i < r; ++i) Y[i * drs + j * dcs] = X[i * srs + j * scs]; });
    return ELB200_OK;
}
int elb200_axpy2d_f16(int64_t r, int64_t c, float al, const void* Xv, int64_t xrs, int64_t xcs, void* Yv, int64_t yrs, int64_t ycs, void* s) {
    const simhalf* X = static_cast<const simhalf*>(Xv); simhalf* Y = static_cast<simhalf*>(Yv);
    sim::enqueue(static_cast<cudaStream_t>(s), [=] { for (int64_t j = 0; j < c; ++j) for (int64_t i = 0; i < r; ++i) Y[i * yrs + j * ycs] = simhalf(al * float(X[i * xrs + j * xcs]) + float(Y[i * yrs + j * ycs])); });
    return ELB200_OK;
}
int elb200_scale2d_f16(int64_t r, int64_t c, float al, void* Av, int64_t lda, void* s) {
    simhalf* A = static_cast<simhalf*>(Av);
    sim::enqueue(static_cast<cudaStream_t>(s), [=] { for (int64_t j = 0; j < c; ++j) for (int64_t i = 0; i < r; ++i) A[i + j * lda] = al == 0.f ? simhalf(0.f) : simhalf(al * float(A[i + j * lda])); });
    return ELB200_OK;
}
int elb200_fill2d_f16(int64_t r, int64_t c, float v, void* Av, int64_t lda, void* s) {
    simhalf* A = static_cast<simhalf*>(Av);
    sim::enqueue(static_cast<cudaStream_t>(s), [=] { for (int64_t j = 0; j < c; ++j) for (int64_t i = 0; i < r; ++i) A[i + j * lda] = simhalf(v); });
    return ELB200_OK;
}
int elb200_transpose_f16(int64_t r, int64_t c, const void* Av, int64_t lda, void* Bv, int64_t ldb, void* s) {
    const simhalf* A = static_cast<const simhalf*>(Av); simhalf* B = static_cast<simhalf*>(Bv);
    sim::enqueue(static_cast<cudaStream_t>(s), [=] { for (int64_t j = 0; j < c; ++j) for (int64_t i = 0; i < r; ++i) B[j + i * ldb] = A[i + j * lda]; });
    return ELB200_OK;
}
int elb200_axpy2d_transpose_f16(int64_t r, int64_t c, float al, const void* Av, int64_t lda, void* Bv, int64_t ldb, void* s) {
    const simhalf* A = static_cast<const simhalf*>(Av); simhalf* B = static_cast<simhalf*>(Bv);
    sim::enqueue(static_cast<cudaStream_t>(s), [=] { for (int64_t j = 0; j < c; ++j) for (int64_t i = 0; i < r; ++i) B[i + j * ldb] = simhalf(al * float(A[j + i * lda]) + float(B[i + j * ldb])); });
    return ELB200_OK;
}
int elb200_lattice_f16(int mode, int nseg, const elb200_segment* segs, void* Mv, int64_t ldm, void* bufv, float al, void* s) {
    simhalf* M = static_cast<simhalf*>(Mv); simhalf* buf = static_cast<simhalf*>(bufv);
    std::vector<elb200_segment> sg(segs, segs + nseg);
    sim::enqueue(static_cast<cudaStream_t>(s), [=] {
        for (const elb200_segment& g : sg)
            for (int64_t j = 0; j < g.cols; ++j)
                for (int64_t i = 0; i < g.rows; ++i) {
                    simhalf* mp = M + g.row0 + i * g.row_step + (g.col0 + j * g.col_step) * ldm;
                    simhalf* bp = buf + g.buf_offset + i + j * g.rows;
                    if (mode == ELB200_PACK) *bp = *mp;
                    else if (mode == ELB200_UNPACK) *mp = *bp;
                    else *mp = simhalf(al * float(*bp) + float(*mp));
                }
    });
    return ELB200_OK;
}
int elb200_peer_reduce_axpy_f16(int nsrc, const void* const* srcs, int64_t off, const elb200_segment* seg, void* Mv, int64_t ldm, float al, void* s) {
    simhalf* M = static_cast<simhalf*>(Mv);
    std::vector<const simhalf*> src(nsrc);
    for (int q = 0; q < nsrc; ++q) src[q] = static_cast<const simhalf*>(srcs[q]);
    const elb200_segment g = *seg;
    sim::enqueue(static_cast<cudaStream_t>(s), [=] {
        for (int64_t j = 0; j < g.cols; ++j)
            for (int64_t i = 0; i < g.rows; ++i) {
                float acc = 0.f;
                for (const simhalf* p : src) acc += float(p[off + i + j * g.rows]);
                simhalf* mp = M + g.row0 + i * g.row_step + (g.col0 + j * g.col_step) * ldm;
                *mp = simhalf(al * acc + float(*mp));
            }
    });
    return ELB200_OK;
}
// the fused one-kernel collectives of peer_kernels.cu as host routines on the caller's stream: same protocol, same flags
// (device memory is host memory here)
namespace {
void fused_spin(const uint32_t* f, uint32_t want, const char* what) {
    const auto t0 = std::chrono::steady_clock::now();
    while (__atomic_load_n(f, __ATOMIC_ACQUIRE) != want) {
        if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > sim::wait_timeout_seconds())
            sim::fail("fused peer collective: %s flag never reached %u (deadlock?)", what, want);
        std::this_thread::yield();
    }
}
void fused_wait_acks(const elb200_peer_gather& pg) {
    for (int i = 0; i < pg.nmembers; ++i)
        if (i != pg.my_index) fused_spin(pg.ack_in[i], pg.seq_prev, "ack");
}
void fused_publish_and_wait(const elb200_peer_gather& pg) {
    for (int i = 0; i < pg.nmembers; ++i)
        __atomic_store_n(i == pg.my_index ? pg.arrive_in[i] : pg.arrive_out[i], pg.seq, __ATOMIC_RELEASE);
    for (int i = 0; i < pg.nmembers; ++i) fused_spin(pg.arrive_in[i], pg.seq, "arrive");
}
void fused_acknowledge(const elb200_peer_gather& pg) {
    for (int i = 0; i < pg.nmembers; ++i)
        if (i != pg.my_index) __atomic_store_n(pg.ack_out[i], pg.seq, __ATOMIC_RELEASE);
}
// modelled time: one launch, the payload over NVLink, two flag round trips
void fused_vt(double bytes) { sim::vt_op(sim::RES_HBM, sim::kLaunch + bytes / sim::kRateNvlink + 2.0 * sim::kFlagLatency); }
bool fused_args_ok(const elb200_peer_gather* pg, int elem_bytes) {
    return pg && pg->nmembers >= 2 && pg->nmembers <= ELB200_MAX_SEGMENTS && pg->my_index >= 0 && pg->my_index < pg->nmembers && pg->counters &&
           (elem_bytes == 2 || elem_bytes == 4 || elem_bytes == 8);
}
}  // namespace
int elb200_peer_allgather(const elb200_peer_gather* pgp, int elem_bytes, const void* send, int64_t send_elems, int nseg,
                          const elb200_segment* segs, void* M, int64_t ldm, void* s) {
    if (!fused_args_ok(pgp, elem_bytes) || send_elems > pgp->portion) return ELB200_ERR_ARG;
    const elb200_peer_gather pg = *pgp;
    std::vector<elb200_segment> sg(segs, segs + nseg);
    sim::enqueue(static_cast<cudaStream_t>(s), [=] {
        const size_t eb = size_t(elem_bytes);
        fused_vt(double(send_elems) * double(eb) * pg.nmembers);
        fused_wait_acks(pg);
        for (int i = 0; i < pg.nmembers; ++i) {
            char* dst = static_cast<char*>(pg.slots[i]) + size_t(pg.my_index) * size_t(pg.portion) * eb;
            if (send_elems) { sim::check_device_range(dst, size_t(send_elems) * eb, "fused gather slot"); std::memcpy(dst, send, size_t(send_elems) * eb); }
        }
        fused_publish_and_wait(pg);
        const char* buf = static_cast<const char*>(pg.slots[pg.my_index]);
        for (const elb200_segment& g : sg)
            for (int64_t j = 0; j < g.cols; ++j)
                for (int64_t i = 0; i < g.rows; ++i)
                    std::memcpy(static_cast<char*>(M) + size_t(g.row0 + i * g.row_step + (g.col0 + j * g.col_step) * ldm) * eb,
                                buf + size_t(g.buf_offset + i + j * g.rows) * eb, eb);
        fused_acknowledge(pg);
    });
    return ELB200_OK;
}
int elb200_peer_exchange(const elb200_peer_gather* pgp, int elem_bytes, int nsend, const elb200_segment* sends, const void* A, int64_t lda,
                         int nrecv, const elb200_segment* recvs, const int* recv_member, void* B, int64_t ldb, void* s) {
    if (!fused_args_ok(pgp, elem_bytes) || nsend < 0 || nrecv < 0 || nsend > ELB200_MAX_SEGMENTS || nrecv > ELB200_MAX_SEGMENTS) return ELB200_ERR_ARG;
    const elb200_peer_gather pg = *pgp;
    std::vector<elb200_segment> snd(sends, sends + nsend), rcv(recvs, recvs + nrecv);
    std::vector<int> from(recv_member, recv_member + nrecv);
    for (int m : from)
        if (m < 0 || m >= pg.nmembers) return ELB200_ERR_ARG;
    sim::enqueue(static_cast<cudaStream_t>(s), [=] {
        const size_t eb = size_t(elem_bytes);
        double bytes = 0;
        for (const elb200_segment& g : rcv) bytes += double(g.rows) * double(g.cols) * double(eb);
        fused_vt(bytes);
        fused_wait_acks(pg);
        char* mine = static_cast<char*>(pg.slots[pg.my_index]);
        for (const elb200_segment& g : snd) {
            if (g.rows > 0 && g.cols > 0) sim::check_device_range(mine + size_t(g.buf_offset) * eb, size_t(g.rows * g.cols) * eb, "fused exchange slot");
            for (int64_t j = 0; j < g.cols; ++j)
                for (int64_t i = 0; i < g.rows; ++i)
                    std::memcpy(mine + size_t(g.buf_offset + i + j * g.rows) * eb,
                                static_cast<const char*>(A) + size_t(g.row0 + i * g.row_step + (g.col0 + j * g.col_step) * lda) * eb, eb);
        }
        fused_publish_and_wait(pg);
        for (size_t q = 0; q < rcv.size(); ++q) {
            const elb200_segment& g = rcv[q];
            const char* src = static_cast<const char*>(pg.slots[from[q]]);
            if (g.rows > 0 && g.cols > 0) sim::check_device_range(src + size_t(g.buf_offset) * eb, size_t(g.rows * g.cols) * eb, "fused exchange peer slot");
            for (int64_t j = 0; j < g.cols; ++j)
                for (int64_t i = 0; i < g.rows; ++i)
                    std::memcpy(static_cast<char*>(B) + size_t(g.row0 + i * g.row_step + (g.col0 + j * g.col_step) * ldb) * eb,
                                src + size_t(g.buf_offset + i + j * g.rows) * eb, eb);
        }
        fused_acknowledge(pg);
    });
    return ELB200_OK;
}
extern "C++" {
namespace {
template <typename T>
int fused_reduce_scatter(const elb200_peer_gather* pgp, int npack, const elb200_segment* packs, const T* A, int64_t lda, const elb200_segment* unpack,
                         T* B, int64_t ldb, T alpha, void* s) {
    if (!fused_args_ok(pgp, int(sizeof(T))) || npack < 0 || npack > ELB200_MAX_SEGMENTS || !unpack || unpack->rows * unpack->cols > pgp->portion)
        return ELB200_ERR_ARG;
    const elb200_peer_gather pg = *pgp;
    std::vector<elb200_segment> pk(packs, packs + npack);
    const elb200_segment un = *unpack;
    sim::enqueue(static_cast<cudaStream_t>(s), [=] {
        fused_vt(double(un.rows) * double(un.cols) * sizeof(T) * pg.nmembers);
        fused_wait_acks(pg);
        T* mine = static_cast<T*>(pg.slots[pg.my_index]);
        for (const elb200_segment& g : pk) {
            if (g.rows > 0 && g.cols > 0) sim::check_device_range(mine + g.buf_offset, size_t(g.rows * g.cols) * sizeof(T), "fused reduce slot");
            for (int64_t j = 0; j < g.cols; ++j)
                for (int64_t i = 0; i < g.rows; ++i) mine[g.buf_offset + i + j * g.rows] = A[g.row0 + i * g.row_step + (g.col0 + j * g.col_step) * lda];
        }
        fused_publish_and_wait(pg);
        const int64_t off = int64_t(pg.my_index) * pg.portion;
        for (int64_t j = 0; j < un.cols; ++j)
            for (int64_t i = 0; i < un.rows; ++i) {
                T sum = T(0);
                for (int q = 0; q < pg.nmembers; ++q) sum += static_cast<const T*>(pg.slots[q])[off + i + j * un.rows];
                T* mp = B + un.row0 + i * un.row_step + (un.col0 + j * un.col_step) * ldb;
                *mp += alpha * sum;
            }
        fused_acknowledge(pg);
    });
    return ELB200_OK;
}
}  // namespace
}  // extern "C++"
int elb200_peer_reduce_scatter_f32(const elb200_peer_gather* pg, int npack, const elb200_segment* packs, const float* A, int64_t lda,
                                   const elb200_segment* unpack, float* B, int64_t ldb, float alpha, void* s) {
    return fused_reduce_scatter<float>(pg, npack, packs, A, lda, unpack, B, ldb, alpha, s);
}
int elb200_peer_reduce_scatter_f64(const elb200_peer_gather* pg, int npack, const elb200_segment* packs, const double* A, int64_t lda,
                                   const elb200_segment* unpack, double* B, int64_t ldb, double alpha, void* s) {
    return fused_reduce_scatter<double>(pg, npack, packs, A, lda, unpack, B, ldb, alpha, s);
}
int elb200_gemm_f64(int ta, int tb, int64_t m, int64_t n, int64_t k, double al, const double* A, int64_t lda, const double* B,
                    int64_t ldb, double be, double* C, int64_t ldc, void* s) {
    return gemm<double>(ta, tb, m, n, k, al, A, lda, B, ldb, be, C, ldc, s);
}

#define SIM_BLAS1(SUF, T)                                                                                                       \
    int elb200_copy2d_##SUF(int64_t r, int64_t c, const T* src, int64_t srs, int64_t scs, T* dst, int64_t drs, int64_t dcs,     \
                            void* s) { return copy2d<T>(r, c, src, srs, scs, dst, drs, dcs, s); }                              \
    int elb200_transpose_##SUF(int64_t r, int64_t c, const T* A, int64_t lda, T* B, int64_t ldb, void* s) {                     \
        return transpose<T, false>(r, c, A, lda, B, ldb, T(0), s); }                                                            \
    int elb200_axpy2d_##SUF(int64_t r, int64_t c, T al, const T* X, int64_t xrs, int64_t xcs, T* Y, int64_t yrs, int64_t ycs,   \
                            void* s) { return axpy2d<T>(r, c, al, X, xrs, xcs, Y, yrs, ycs, s); }                              \
    int elb200_axpy2d_transpose_##SUF(int64_t r, int64_t c, T al, const T* A, int64_t lda, T* B, int64_t ldb, void* s) {        \
        return transpose<T, true>(c, r, A, lda, B, ldb, al, s); }                                                               \
    int elb200_scale2d_##SUF(int64_t r, int64_t c, T al, T* A, int64_t lda, void* s) { return scale2d<T>(r, c, al, A, lda, s); } \
    int elb200_fill2d_##SUF(int64_t r, int64_t c, T v, T* A, int64_t lda, void* s) { return fill2d<T>(r, c, v, A, lda, s); }    \
    int elb200_lattice_##SUF(int mode, int nseg, const elb200_segment* segs, T* M, int64_t ldm, T* buf, T al, void* s) {        \
        return lattice<T>(mode, nseg, segs, M, ldm, buf, al, s); }                                                              \
    int elb200_peer_reduce_axpy_##SUF(int nsrc, const T* const* srcs, int64_t off, const elb200_segment* seg, T* M,             \
                                      int64_t ldm, T al, void* s) { return peer_reduce_axpy<T>(nsrc, srcs, off, seg, M, ldm, al, s); }
int elb200_trapezoid_f32(int op, int uplo, int64_t off, int64_t r, int64_t c, int64_t r0, int64_t rs, int64_t c0, int64_t cs, float al,
                         const float* X, int64_t ldx, float* Y, int64_t ldy, void* s) {
    return trapezoid<float>(op, uplo, off, r, c, r0, rs, c0, cs, al, X, ldx, Y, ldy, s);
}
int elb200_trapezoid_f64(int op, int uplo, int64_t off, int64_t r, int64_t c, int64_t r0, int64_t rs, int64_t c0, int64_t cs, double al,
                         const double* X, int64_t ldx, double* Y, int64_t ldy, void* s) {
    return trapezoid<double>(op, uplo, off, r, c, r0, rs, c0, cs, al, X, ldx, Y, ldy, s);
}
int elb200_trsm_f32(int side, int uplo, int trans, int diag, int64_t m, int64_t n, float al, const float* A, int64_t lda, float* B,
                    int64_t ldb, void* s) { return trsm<float>(side, uplo, trans, diag, m, n, al, A, lda, B, ldb, s); }
int elb200_trsm_f64(int side, int uplo, int trans, int diag, int64_t m, int64_t n, double al, const double* A, int64_t lda, double* B,
                    int64_t ldb, void* s) { return trsm<double>(side, uplo, trans, diag, m, n, al, A, lda, B, ldb, s); }
SIM_BLAS1(f32, float)
SIM_BLAS1(f64, double)
#undef SIM_BLAS1

}  // extern "C"
