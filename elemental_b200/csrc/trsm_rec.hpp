// Host-side recursion of the local triangular solve (see trsm.cu): halve the triangle down to
// kTrsmBase x kTrsmBase diagonal blocks; every off-diagonal update is one GEMM.  Written against a
// backend (gemm + base-case solve) so that tests/hostsim runs this very code over its host doubles.
//   LEFT : op(A) X = B      RIGHT: X op(A) = B      (alpha already applied to B; X overwrites B)
// `lower` is the STORED triangle of A, `trans` the orientation; op(A) is lower iff lower != trans.
#pragma once
#include <cstdint>

namespace elb {

constexpr int kTrsmBase = 32;

// Backend: int gemm(int ta, int tb, int64_t m, int64_t n, int64_t k, T alpha, const T* A, int64_t lda,
//                   const T* B, int64_t ldb, T beta, T* C, int64_t ldc);
//          int base(bool left, bool lower, bool trans, bool unit, int64_t m, int64_t n,
//                   const T* A, int64_t lda, T* B, int64_t ldb);      // triangle <= kTrsmBase
template <typename T, typename Backend>
int trsm_recursive(Backend& be, bool left, bool lower, bool trans, bool unit, int64_t m, int64_t n, const T* A, int64_t lda,
                   T* B, int64_t ldb) {
    const int64_t na = left ? m : n;
    if (na <= kTrsmBase) return be.base(left, lower, trans, unit, m, n, A, lda, B, ldb);
    const bool eff_lower = lower != trans;
    const int64_t n1 = ((na / 2 + kTrsmBase - 1) / kTrsmBase) * kTrsmBase, n2 = na - n1;
    const T* A11 = A;
    const T* A21 = A + n1;
    const T* A12 = A + n1 * lda;
    const T* A22 = A + n1 + n1 * lda;
    int st;
    if (left) {
        T* B1 = B;
        T* B2 = B + n1;
        if (eff_lower) {  // X1 = solve(A11, B1); B2 -= op(A)(2,1) X1; X2 = solve(A22, B2)
            if ((st = trsm_recursive<T>(be, left, lower, trans, unit, n1, n, A11, lda, B1, ldb))) return st;
            st = trans ? be.gemm(1, 0, n2, n, n1, T(-1), A12, lda, B1, ldb, T(1), B2, ldb)
                       : be.gemm(0, 0, n2, n, n1, T(-1), A21, lda, B1, ldb, T(1), B2, ldb);
            if (st) return st;
            return trsm_recursive<T>(be, left, lower, trans, unit, n2, n, A22, lda, B2, ldb);
        }
        if ((st = trsm_recursive<T>(be, left, lower, trans, unit, n2, n, A22, lda, B2, ldb))) return st;
        st = trans ? be.gemm(1, 0, n1, n, n2, T(-1), A21, lda, B2, ldb, T(1), B1, ldb)
                   : be.gemm(0, 0, n1, n, n2, T(-1), A12, lda, B2, ldb, T(1), B1, ldb);
        if (st) return st;
        return trsm_recursive<T>(be, left, lower, trans, unit, n1, n, A11, lda, B1, ldb);
    }
    T* B1 = B;
    T* B2 = B + n1 * ldb;
    if (!eff_lower) {  // op(A) upper: X1 = solve(B1, A11); B2 -= X1 op(A)(1,2); X2 = solve(B2, A22)
        if ((st = trsm_recursive<T>(be, left, lower, trans, unit, m, n1, A11, lda, B1, ldb))) return st;
        st = trans ? be.gemm(0, 1, m, n2, n1, T(-1), B1, ldb, A21, lda, T(1), B2, ldb)
                   : be.gemm(0, 0, m, n2, n1, T(-1), B1, ldb, A12, lda, T(1), B2, ldb);
        if (st) return st;
        return trsm_recursive<T>(be, left, lower, trans, unit, m, n2, A22, lda, B2, ldb);
    }
    if ((st = trsm_recursive<T>(be, left, lower, trans, unit, m, n2, A22, lda, B2, ldb))) return st;
    st = trans ? be.gemm(0, 1, m, n1, n2, T(-1), B2, ldb, A12, lda, T(1), B1, ldb)
               : be.gemm(0, 0, m, n1, n2, T(-1), B2, ldb, A21, lda, T(1), B1, ldb);
    if (st) return st;
    return trsm_recursive<T>(be, left, lower, trans, unit, m, n1, A11, lda, B1, ldb);
}

}  // namespace elb
