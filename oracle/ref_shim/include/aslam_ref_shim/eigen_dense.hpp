// TEST INFRASTRUCTURE ONLY -- Eigen3 stand-in used to compile the UNMODIFIED reference sources
// (/root/reference/awesome_slam/src/{ekf,ukf,sensor_landmark}/*.cpp and include/awesome_slam/*.h) into
// oracle/_ref/ (see oracle/ref_shim/README.md). Never part of the product.
//
// Eigen3 is an un-vendored, un-pinned third-party dependency of the reference (implied 3.3.7: Ubuntu 20.04 / ROS
// Noetic, README.md:3-5) and is absent from this image. This header provides exactly the entry points the reference
// text uses (ekf.h:49-51, ekf.cpp:49-311, ukf.h:56-82, ukf.cpp:49-392, common.h:46-108, tools.h:53-59,
// structures.h:44-134, sensor_landmark.cpp:192-298) as an EAGER value library on top of oracle/dense.hpp:
//   * every operator returns an evaluated matrix; products nested in a sum or another product become temporaries,
//     which is what Eigen's evaluators do for these expressions (product operands are `nested_eval`-ed, and
//     `dst = a + b*c` / `dst += b*c` go through a temporary or `dst = a; dst += b*c` -- same rounding);
//   * `scalar * v * v.transpose()` evaluates (scalar*v) first and forms the outer product column by column
//     (outer_product_selector_run: dst.col(j) = rhs(j) * actual_lhs) -> entry (i,j) = (s*v_i) * v_j;
//   * a float scalar times a double matrix is promoted to double first (ScalarBinaryOpTraits);
//   * summation inside a product is k-ascending from 0, multiply then add, no FMA. Eigen's own kernels (GEBP,
//     GEMV, blocked PartialPivLU above size 16) reorder partial sums depending on size AND operand alignment, so
//     Eigen's inner arithmetic is not bit-reproducible even against itself; this is the one part of `_ref` that is
//     restated, and DESIGN.md section 5 says so.
// Compile with -ffp-contract=off.
#pragma once
#include <algorithm>
#include <cassert>
#include <cmath>
#include <complex>
#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <iostream>
#include <limits>
#include <string>
#include <type_traits>
#include <utility>
#include <vector>

#include "../../../dense.hpp"

namespace Eigen {

typedef std::ptrdiff_t Index;
const int Dynamic = -1;
enum DecompositionOptions { ComputeFullU = 0x04, ComputeThinU = 0x08, ComputeFullV = 0x10, ComputeThinV = 0x20 };

class Dyn;
class LLTShim;
class ColPivQRShim;

// Writable view of a rectangular part of a Dyn (head / col / row / corners).
class BlockRef {
  public:
    BlockRef(Dyn& p, Index r0, Index c0, Index nr, Index nc) : p_(p), r0_(r0), c0_(c0), nr_(nr), nc_(nc) {}
    BlockRef& operator=(const Dyn& v);
    BlockRef& operator=(const BlockRef& v);
    operator Dyn() const;
    double& operator()(Index i);
    double& operator()(Index i, Index j);
    Index rows() const { return nr_; }
    Index cols() const { return nc_; }
    Index size() const { return nr_ * nc_; }
    Dyn transpose() const;

  private:
    Dyn& p_;
    Index r0_, c0_, nr_, nc_;
};

// `m << a, b, c;` fills in row-major order, as Eigen's CommaInitializer does.
class CommaInit {
  public:
    CommaInit(Dyn& m, double first);
    template <class T>
    CommaInit& operator,(const T& v) {
        put((double)v);
        return *this;
    }

  private:
    void put(double v);
    Dyn& m_;
    Index k_ = 0;
};

class Dyn {
  public:
    orc::Mat m;

    Dyn() {}
    explicit Dyn(const orc::Mat& a) : m(a) {}
    Dyn(Index r, Index c, int) : m((int)r, (int)c) {}

    Index rows() const { return m.r; }
    Index cols() const { return m.c; }
    Index size() const { return (Index)m.r * m.c; }

    double& operator()(Index i, Index j) { return m((int)i, (int)j); }
    double operator()(Index i, Index j) const { return m((int)i, (int)j); }
    double& operator()(Index i) { return m.d[(size_t)i]; }
    double operator()(Index i) const { return m.d[(size_t)i]; }
    double& operator[](Index i) { return m.d[(size_t)i]; }
    double operator[](Index i) const { return m.d[(size_t)i]; }

    void fill(double v) { m.fill(v); }
    Dyn transpose() const { return Dyn(orc::transpose(m)); }
    Dyn inverse() const { return Dyn(orc::inverse(m)); }  // dynamic size: PartialPivLU(m).inverse()
    LLTShim llt() const;
    ColPivQRShim colPivHouseholderQr() const;

    // ArrayWrapper / MatrixWrapper round trip of sensor_landmark.cpp:266
    const Dyn& array() const { return *this; }
    const Dyn& matrix() const { return *this; }
    Dyn asDiagonal() const {
        const int n = (int)size();
        Dyn d(n, n, 0);
        for (int i = 0; i < n; ++i) d(i, i) = (*this)(i);
        return d;
    }

    // DenseBase::conservativeResizeLike(other)
    void conservativeResizeLike(const Dyn& other) { orc::conservative_resize_like(m, other.m); }

    BlockRef head(Index n) { return BlockRef(*this, 0, 0, n, 1); }
    Dyn head(Index n) const { return part(0, 0, n, 1); }
    BlockRef col(Index j) { return BlockRef(*this, 0, j, rows(), 1); }
    Dyn col(Index j) const { return part(0, j, rows(), 1); }
    BlockRef row(Index i) { return BlockRef(*this, i, 0, 1, cols()); }
    Dyn row(Index i) const { return part(i, 0, 1, cols()); }
    BlockRef topLeftCorner(Index r, Index c) { return BlockRef(*this, 0, 0, r, c); }
    BlockRef bottomRightCorner(Index r, Index c) { return BlockRef(*this, rows() - r, cols() - c, r, c); }

    Dyn part(Index r0, Index c0, Index nr, Index nc) const {
        Dyn o(nr, nc, 0);
        for (Index j = 0; j < nc; ++j)
            for (Index i = 0; i < nr; ++i) o(i, j) = (*this)(r0 + i, c0 + j);
        return o;
    }

    template <class T>
    CommaInit operator<<(const T& v) {
        return CommaInit(*this, (double)v);
    }

    Dyn& operator+=(const Dyn& o) {
        for (size_t i = 0; i < m.d.size(); ++i) m.d[i] = m.d[i] + o.m.d[i];
        return *this;
    }
    Dyn& operator-=(const Dyn& o) {
        for (size_t i = 0; i < m.d.size(); ++i) m.d[i] = m.d[i] - o.m.d[i];
        return *this;
    }
};

inline CommaInit::CommaInit(Dyn& m, double first) : m_(m) { put(first); }
inline void CommaInit::put(double v) {
    const Index c = m_.cols();
    m_(k_ / c, k_ % c) = v;
    ++k_;
}

inline BlockRef& BlockRef::operator=(const Dyn& v) {
    assert(v.rows() == nr_ && v.cols() == nc_);
    for (Index j = 0; j < nc_; ++j)
        for (Index i = 0; i < nr_; ++i) p_(r0_ + i, c0_ + j) = v(i, j);
    return *this;
}
inline BlockRef& BlockRef::operator=(const BlockRef& v) { return *this = (Dyn)v; }
inline BlockRef::operator Dyn() const { return p_.part(r0_, c0_, nr_, nc_); }
inline double& BlockRef::operator()(Index i) { return nc_ == 1 ? p_(r0_ + i, c0_) : p_(r0_, c0_ + i); }
inline double& BlockRef::operator()(Index i, Index j) { return p_(r0_ + i, c0_ + j); }
inline Dyn BlockRef::transpose() const { return ((Dyn)*this).transpose(); }

// ---- operators (non-template on purpose: Matrix<> binds directly, BlockRef through its conversion) ----
inline Dyn operator+(const Dyn& a, const Dyn& b) { return Dyn(orc::add(a.m, b.m)); }
inline Dyn operator-(const Dyn& a, const Dyn& b) { return Dyn(orc::sub(a.m, b.m)); }
inline Dyn operator*(const Dyn& a, double s) { return Dyn(orc::scale(a.m, s)); }
inline Dyn operator*(double s, const Dyn& a) {
    Dyn o(a.rows(), a.cols(), 0);
    for (size_t i = 0; i < o.m.d.size(); ++i) o.m.d[i] = s * a.m.d[i];
    return o;
}
inline Dyn operator*(const Dyn& a, const Dyn& b) {
    assert(a.cols() == b.rows());
    if (a.cols() == 1) {  // outer product: entry (i,j) = b(0,j) * a(i,0), no accumulation
        Dyn o(a.rows(), b.cols(), 0);
        for (Index j = 0; j < b.cols(); ++j)
            for (Index i = 0; i < a.rows(); ++i) o(i, j) = b(0, j) * a(i, 0);
        return o;
    }
    return Dyn(orc::matmul(a.m, b.m));
}

// LLT<MatrixXd, Lower>: compute() never throws; matrixL() returns the lower triangle of whatever the in-place
// factorisation left behind (half-factored after a non-positive pivot) -- orc::llt_lower.
class LLTShim {
  public:
    explicit LLTShim(const Dyn& a) : l_(orc::llt_lower(a.m, &fail_)) {}
    Dyn matrixL() const { return Dyn(l_); }
    int failedAt() const { return fail_; }

  private:
    int fail_ = -1;
    orc::Mat l_;
};
inline LLTShim Dyn::llt() const { return LLTShim(*this); }

template <class Scalar, int R, int C>
class Matrix : public Dyn {
    static_assert(std::is_same<Scalar, double>::value, "the reference path only uses double matrices");

  public:
    Matrix() {
        if (R != Dynamic && C != Dynamic) m = orc::Mat(R, C);
    }
    // Vector(size) for dynamic vectors
    template <class T, class = typename std::enable_if<std::is_integral<T>::value>::type>
    explicit Matrix(T n) {
        static_assert(C == 1 || R == 1, "size constructor is for vectors");
        m = C == 1 ? orc::Mat((int)n, 1) : orc::Mat(1, (int)n);
    }
    // Matrix(rows, cols) for dynamic matrices, Vector2d(x, y) for fixed 2-vectors
    template <class A, class B, class = typename std::enable_if<std::is_arithmetic<A>::value && std::is_arithmetic<B>::value>::type>
    Matrix(const A& a, const B& b) {
        if (R == 2 && C == 1) {
            m = orc::Mat(2, 1);
            m.d[0] = (double)a;
            m.d[1] = (double)b;
        } else {
            m = orc::Mat((int)a, (int)b);
        }
    }
    Matrix(const Dyn& o) : Dyn(o) {}
    Matrix(const BlockRef& b) : Dyn((Dyn)b) {}
    Matrix& operator=(const Dyn& o) {
        m = o.m;
        return *this;
    }
    Matrix& operator=(const BlockRef& b) {
        m = ((Dyn)b).m;
        return *this;
    }

    static Matrix Zero(Index n) { return Matrix(Dyn(C == 1 ? n : 1, C == 1 ? 1 : n, 0)); }
    static Matrix Zero(Index r, Index c) { return Matrix(Dyn(r, c, 0)); }
    static Matrix Ones(Index n) {
        Matrix o = Zero(n);
        o.fill(1.0);
        return o;
    }
    static Matrix Identity(Index r, Index c) { return Matrix(Dyn(orc::Mat::identity((int)r, (int)c))); }
};

typedef Matrix<double, Dynamic, Dynamic> MatrixXd;
typedef Matrix<double, Dynamic, 1> VectorXd;
typedef Matrix<double, 2, 1> Vector2d;
typedef Matrix<double, 4, 4> Matrix4d;

}  // namespace Eigen
