// TEST INFRASTRUCTURE ONLY -- the three Eigen decompositions the reference's laser front end calls
// (sensor_landmark.cpp:251 JacobiSVD, :272 SelfAdjointEigenSolver, :286 colPivHouseholderQr().solve), as stand-ins
// for compiling the unmodified reference into oracle/_ref (see eigen_dense.hpp). Eigen is absent from this image;
// these follow the published methods (one-sided Jacobi SVD, cyclic Jacobi eigen-solver, Householder QR with column
// pivoting), not Eigen's code. Results agree with Eigen's to rounding; the front end publishes floats.
#pragma once
#include "eigen_dense.hpp"

namespace Eigen {

// Thin SVD of an n x c matrix (n >= c) by one-sided (Hestenes) Jacobi rotations on the columns;
// singular values sorted in decreasing order, as JacobiSVD returns them.
template <class MatrixType>
class JacobiSVD {
  public:
    JacobiSVD(const Dyn& a, unsigned int = 0) {
        const Index n = a.rows(), c = a.cols();
        Dyn u = a;
        Dyn v(c, c, 0);
        for (Index i = 0; i < c; ++i) v(i, i) = 1.0;
        for (int sweep = 0; sweep < 60; ++sweep) {
            bool rotated = false;
            for (Index p = 0; p < c - 1; ++p)
                for (Index q = p + 1; q < c; ++q) {
                    double alpha = 0, beta = 0, gamma = 0;
                    for (Index i = 0; i < n; ++i) {
                        alpha += u(i, p) * u(i, p);
                        beta += u(i, q) * u(i, q);
                        gamma += u(i, p) * u(i, q);
                    }
                    if (gamma == 0.0 || std::fabs(gamma) <= 2.2e-16 * std::sqrt(alpha * beta)) continue;
                    rotated = true;
                    const double zeta = (beta - alpha) / (2.0 * gamma);
                    const double t = (zeta >= 0 ? 1.0 : -1.0) / (std::fabs(zeta) + std::sqrt(1.0 + zeta * zeta));
                    const double cs = 1.0 / std::sqrt(1.0 + t * t), sn = cs * t;
                    for (Index i = 0; i < n; ++i) {
                        const double up = u(i, p), uq = u(i, q);
                        u(i, p) = cs * up - sn * uq;
                        u(i, q) = sn * up + cs * uq;
                    }
                    for (Index i = 0; i < c; ++i) {
                        const double vp = v(i, p), vq = v(i, q);
                        v(i, p) = cs * vp - sn * vq;
                        v(i, q) = sn * vp + cs * vq;
                    }
                }
            if (!rotated) break;
        }
        std::vector<double> s((size_t)c);
        std::vector<Index> order((size_t)c);
        for (Index j = 0; j < c; ++j) {
            double nrm = 0;
            for (Index i = 0; i < n; ++i) nrm += u(i, j) * u(i, j);
            s[(size_t)j] = std::sqrt(nrm);
            order[(size_t)j] = j;
        }
        std::stable_sort(order.begin(), order.end(), [&](Index x, Index y) { return s[(size_t)x] > s[(size_t)y]; });
        sv_ = Dyn(c, 1, 0);
        u_ = Dyn(n, c, 0);
        v_ = Dyn(c, c, 0);
        for (Index j = 0; j < c; ++j) {
            const Index src = order[(size_t)j];
            const double sj = s[(size_t)src];
            sv_(j) = sj;
            for (Index i = 0; i < n; ++i) u_(i, j) = sj > 0 ? u(i, src) / sj : 0.0;
            for (Index i = 0; i < c; ++i) v_(i, j) = v(i, src);
        }
    }
    const MatrixType& singularValues() const { return sv_; }
    const MatrixType& matrixU() const { return u_; }
    const MatrixType& matrixV() const { return v_; }

  private:
    MatrixType sv_, u_, v_;
};

// Symmetric eigen-decomposition (reads the lower triangle, like Eigen): cyclic Jacobi, eigenvalues ascending,
// eigenvectors as unit columns.
template <class MatrixType>
class SelfAdjointEigenSolver {
  public:
    explicit SelfAdjointEigenSolver(const Dyn& q) {
        const Index n = q.rows();
        Dyn a(n, n, 0), v(n, n, 0);
        for (Index i = 0; i < n; ++i) {
            v(i, i) = 1.0;
            for (Index j = 0; j <= i; ++j) a(i, j) = a(j, i) = q(i, j);
        }
        for (int sweep = 0; sweep < 60; ++sweep) {
            double off = 0;
            for (Index i = 0; i < n; ++i)
                for (Index j = i + 1; j < n; ++j) off += std::fabs(a(i, j));
            if (off == 0.0) break;
            for (Index p = 0; p < n - 1; ++p)
                for (Index r = p + 1; r < n; ++r) {
                    if (a(p, r) == 0.0) continue;
                    const double g = 100.0 * std::fabs(a(p, r));
                    if (sweep > 3 && std::fabs(a(p, p)) + g == std::fabs(a(p, p)) &&
                        std::fabs(a(r, r)) + g == std::fabs(a(r, r))) {
                        a(p, r) = a(r, p) = 0.0;
                        continue;
                    }
                    const double theta = (a(r, r) - a(p, p)) / (2.0 * a(p, r));
                    const double t = (theta >= 0 ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0));
                    const double c = 1.0 / std::sqrt(t * t + 1.0), s = t * c;
                    for (Index k = 0; k < n; ++k) {
                        const double akp = a(k, p), akr = a(k, r);
                        a(k, p) = c * akp - s * akr;
                        a(k, r) = s * akp + c * akr;
                    }
                    for (Index k = 0; k < n; ++k) {
                        const double apk = a(p, k), ark = a(r, k);
                        a(p, k) = c * apk - s * ark;
                        a(r, k) = s * apk + c * ark;
                    }
                    a(p, r) = a(r, p) = 0.0;
                    for (Index k = 0; k < n; ++k) {
                        const double vkp = v(k, p), vkr = v(k, r);
                        v(k, p) = c * vkp - s * vkr;
                        v(k, r) = s * vkp + c * vkr;
                    }
                }
        }
        std::vector<Index> order((size_t)n);
        for (Index i = 0; i < n; ++i) order[(size_t)i] = i;
        std::stable_sort(order.begin(), order.end(), [&](Index x, Index y) { return a(x, x) < a(y, y); });
        w_ = Dyn(n, 1, 0);
        v_ = Dyn(n, n, 0);
        for (Index j = 0; j < n; ++j) {
            w_(j) = a(order[(size_t)j], order[(size_t)j]);
            for (Index i = 0; i < n; ++i) v_(i, j) = v(i, order[(size_t)j]);
        }
    }
    const MatrixType& eigenvalues() const { return w_; }
    const MatrixType& eigenvectors() const { return v_; }

  private:
    MatrixType w_, v_;
};

// Householder QR with column pivoting; solve() for a square, full-rank system.
class ColPivQRShim {
  public:
    explicit ColPivQRShim(const Dyn& a) : qr_(a) {
        const Index n = a.rows(), c = a.cols();
        perm_.resize((size_t)c);
        for (Index j = 0; j < c; ++j) perm_[(size_t)j] = j;
        beta_.assign((size_t)std::min(n, c), 0.0);
        for (Index k = 0; k < std::min(n, c); ++k) {
            Index best = k;
            double bn = -1;
            for (Index j = k; j < c; ++j) {
                double s = 0;
                for (Index i = k; i < n; ++i) s += qr_(i, j) * qr_(i, j);
                if (s > bn) {
                    bn = s;
                    best = j;
                }
            }
            if (best != k) {
                for (Index i = 0; i < n; ++i) std::swap(qr_(i, k), qr_(i, best));
                std::swap(perm_[(size_t)k], perm_[(size_t)best]);
            }
            double nrm = std::sqrt(bn);
            if (nrm == 0.0) continue;
            const double x0 = qr_(k, k);
            const double alpha = x0 >= 0 ? -nrm : nrm;
            const double v0 = x0 - alpha;  // v = x - alpha e1, stored scaled so that v(0) = 1
            for (Index i = k + 1; i < n; ++i) qr_(i, k) /= v0;
            beta_[(size_t)k] = -v0 / alpha;  // 2 / (v'v) with v(0) = 1
            qr_(k, k) = alpha;
            for (Index j = k + 1; j < c; ++j) {
                double s = qr_(k, j);
                for (Index i = k + 1; i < n; ++i) s += qr_(i, k) * qr_(i, j);
                s *= beta_[(size_t)k];
                qr_(k, j) -= s;
                for (Index i = k + 1; i < n; ++i) qr_(i, j) -= s * qr_(i, k);
            }
        }
    }
    Dyn solve(const Dyn& b) const {
        const Index n = qr_.rows(), c = qr_.cols();
        Dyn y = b;
        for (Index col = 0; col < b.cols(); ++col) {
            for (Index k = 0; k < std::min(n, c); ++k) {  // y <- Q' y
                if (beta_[(size_t)k] == 0.0) continue;
                double s = y(k, col);
                for (Index i = k + 1; i < n; ++i) s += qr_(i, k) * y(i, col);
                s *= beta_[(size_t)k];
                y(k, col) -= s;
                for (Index i = k + 1; i < n; ++i) y(i, col) -= s * qr_(i, k);
            }
        }
        Dyn x(c, b.cols(), 0);
        for (Index col = 0; col < b.cols(); ++col) {
            std::vector<double> z((size_t)c);
            for (Index i = c - 1; i >= 0; --i) {
                double s = y(i, col);
                for (Index k = i + 1; k < c; ++k) s -= qr_(i, k) * z[(size_t)k];
                z[(size_t)i] = s / qr_(i, i);
            }
            for (Index i = 0; i < c; ++i) x(perm_[(size_t)i], col) = z[(size_t)i];
        }
        return x;
    }

  private:
    Dyn qr_;
    std::vector<Index> perm_;
    std::vector<double> beta_;
};
inline ColPivQRShim Dyn::colPivHouseholderQr() const { return ColPivQRShim(*this); }

}  // namespace Eigen
