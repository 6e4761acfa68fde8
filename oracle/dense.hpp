// TEST INFRASTRUCTURE ONLY -- part of the CPU oracle (see oracle/README.md).
// Never linked into, imported by, or executed from the product path.
//
// Minimal column-major dense double matrix with exactly the Eigen3 operations the
// reference hot path calls:
//   MatrixXd * MatrixXd, transpose, +, -, scalar *,
//   .inverse()            == PartialPivLU(S).inverse()      (ekf.cpp:301, ukf.cpp:378)
//   .llt().matrixL()      == unblocked left-looking LLT     (ukf.cpp:280)
//   conservativeResizeLike                                   (ekf.cpp:271-278, ukf.cpp:238-242)
// Eigen itself is NOT on disk; this restates its published algorithms
// (Eigen 3.3.7: LLT.h llt_inplace<Lower>::unblocked, PartialPivLU.h unblocked_lu,
// DenseBase::conservativeResizeLike). Summation is the natural k-ascending order.
// Compile with -ffp-contract=off so no FMA contraction happens anywhere.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <vector>

namespace orc {

struct Mat {
    int r = 0, c = 0;
    std::vector<double> d;  // column-major, like Eigen::MatrixXd

    Mat() {}
    Mat(int rows, int cols) : r(rows), c(cols), d((size_t)rows * cols, 0.0) {}
    static Mat zero(int rows, int cols) { return Mat(rows, cols); }
    static Mat identity(int rows, int cols) {
        Mat m(rows, cols);
        for (int i = 0; i < std::min(rows, cols); ++i) m(i, i) = 1.0;
        return m;
    }
    static Mat constant(int rows, int cols, double v) {
        Mat m(rows, cols);
        std::fill(m.d.begin(), m.d.end(), v);
        return m;
    }
    double& operator()(int i, int j) { return d[(size_t)j * r + i]; }
    double operator()(int i, int j) const { return d[(size_t)j * r + i]; }
    // vector-style access (column vectors)
    double& operator()(int i) { return d[i]; }
    double operator()(int i) const { return d[i]; }
    int size() const { return r * c; }
    void fill(double v) { std::fill(d.begin(), d.end(), v); }
};

inline Mat transpose(const Mat& a) {
    Mat t(a.c, a.r);
    for (int j = 0; j < a.c; ++j)
        for (int i = 0; i < a.r; ++i) t(j, i) = a(i, j);
    return t;
}

// C = A * B, each output element accumulated over k ascending, mul then add (no FMA).
inline Mat matmul(const Mat& a, const Mat& b) {
    Mat o(a.r, b.c);
    for (int j = 0; j < b.c; ++j)
        for (int i = 0; i < a.r; ++i) {
            double s = 0.0;
            for (int k = 0; k < a.c; ++k) s += a(i, k) * b(k, j);
            o(i, j) = s;
        }
    return o;
}

inline Mat add(const Mat& a, const Mat& b) {
    Mat o(a.r, a.c);
    for (size_t i = 0; i < o.d.size(); ++i) o.d[i] = a.d[i] + b.d[i];
    return o;
}
inline Mat sub(const Mat& a, const Mat& b) {
    Mat o(a.r, a.c);
    for (size_t i = 0; i < o.d.size(); ++i) o.d[i] = a.d[i] - b.d[i];
    return o;
}
inline Mat scale(const Mat& a, double s) {
    Mat o(a.r, a.c);
    for (size_t i = 0; i < o.d.size(); ++i) o.d[i] = a.d[i] * s;
    return o;
}

// DenseBase::conservativeResizeLike(other): result has other's shape; the old top-left
// block is kept, every new entry is taken from `other`.
inline void conservative_resize_like(Mat& m, const Mat& other) {
    Mat o = other;
    int rr = std::min(m.r, other.r), cc = std::min(m.c, other.c);
    for (int j = 0; j < cc; ++j)
        for (int i = 0; i < rr; ++i) o(i, j) = m(i, j);
    m = o;
}

// PartialPivLU unblocked_lu (Eigen 3.3.7 PartialPivLU.h): for each column k pick the
// first row with the largest |a(i,k)|, i >= k; swap; scale the sub-column by 1/pivot is
// NOT what Eigen does -- it divides (`lu.col(k).tail(rrows) /= lu.coeff(k,k)`); then
// rank-1 update of the trailing block. A zero pivot is skipped silently (no error).
// Returns the row transpositions; `lu` is overwritten with L (unit, below) and U.
inline void partial_piv_lu(Mat& lu, std::vector<int>& transp) {
    const int n = lu.r;
    transp.assign(n, 0);
    for (int k = 0; k < n; ++k) {
        int piv = k;
        double best = std::fabs(lu(k, k));
        for (int i = k + 1; i < n; ++i) {
            double v = std::fabs(lu(i, k));
            if (v > best) { best = v; piv = i; }
        }
        transp[k] = piv;
        if (best != 0.0) {
            if (piv != k)
                for (int j = 0; j < n; ++j) std::swap(lu(k, j), lu(piv, j));
            const double p = lu(k, k);
            for (int i = k + 1; i < n; ++i) lu(i, k) /= p;
        }
        for (int j = k + 1; j < n; ++j) {
            const double ukj = lu(k, j);
            for (int i = k + 1; i < n; ++i) lu(i, j) -= lu(i, k) * ukj;
        }
    }
}

// MatrixBase::inverse() for dynamic sizes == PartialPivLU(m).solve(Identity):
// permute I, forward-substitute with unit-lower L, back-substitute with U.
inline Mat inverse(const Mat& m) {
    const int n = m.r;
    Mat lu = m;
    std::vector<int> tr;
    partial_piv_lu(lu, tr);
    Mat x = Mat::identity(n, n);
    for (int k = 0; k < n; ++k)
        if (tr[k] != k)
            for (int j = 0; j < n; ++j) std::swap(x(k, j), x(tr[k], j));
    for (int j = 0; j < n; ++j) {
        for (int i = 0; i < n; ++i) {  // L y = b (unit diagonal)
            double s = x(i, j);
            for (int k = 0; k < i; ++k) s -= lu(i, k) * x(k, j);
            x(i, j) = s;
        }
        for (int i = n - 1; i >= 0; --i) {  // U x = y
            double s = x(i, j);
            for (int k = i + 1; k < n; ++k) s -= lu(i, k) * x(k, j);
            x(i, j) = s / lu(i, i);
        }
    }
    return x;
}

// LLT<MatrixXd, Lower>: llt_inplace<Lower>::unblocked (Eigen 3.3.7 LLT.h), left-looking.
// On the first pivot x <= 0 at column k it returns k and leaves column k.. untouched
// (they still hold the input's lower triangle); Eigen's `.llt()` ignores the failure and
// `.matrixL()` hands back that half-factored lower triangle. Eigen
// only uses this unblocked routine for size < 32; larger sizes run a blocked variant whose
// SUCCESS result agrees to rounding but whose failure state differs -- `fail_k` >= 0 on a
// size >= 32 matrix therefore means "outside the pinned contract".
// Returns L as a dense matrix with a zero strict upper triangle.
inline Mat llt_lower(const Mat& a, int* fail_k) {
    const int n = a.r;
    Mat m = a;
    int fk = -1;
    for (int k = 0; k < n; ++k) {
        double x = m(k, k);
        if (k > 0) {
            double s = 0.0;
            for (int j = 0; j < k; ++j) s += m(k, j) * m(k, j);
            x -= s;
        }
        if (x <= 0.0) { fk = k; break; }
        x = std::sqrt(x);
        m(k, k) = x;
        for (int i = k + 1; i < n; ++i) {
            double s = 0.0;
            for (int j = 0; j < k; ++j) s += m(i, j) * m(k, j);
            m(i, k) = (m(i, k) - s) / x;
        }
    }
    for (int j = 1; j < n; ++j)
        for (int i = 0; i < j; ++i) m(i, j) = 0.0;
    if (fail_k) *fail_k = fk;
    return m;
}

}  // namespace orc
