// dense.hpp — the Eigen surface the drop-in headers need (MatrixXd / VectorXd as plain
// containers).  If a real Eigen is on the include path it is used; otherwise this minimal
// stand-in supplies the few members the reference-compatible signatures touch
// (reference: include/element.hpp:6-8, include/solver.hpp:3).  Containers only — every
// number is produced by the CUDA library behind include/csim_b200.h.
#pragma once
#if defined(CSIM_USE_REAL_EIGEN) && __has_include(<Eigen/Dense>)
#include <Eigen/Dense>
#else
#include <algorithm>
#include <cmath>
#include <complex>
#include <cstddef>
#include <vector>

namespace Eigen {
using Index = std::ptrdiff_t;

template <class S>
class DenseVector {
public:
    DenseVector() = default;
    explicit DenseVector(Index n) : v_(static_cast<std::size_t>(n)) {}
    static DenseVector Zero(Index n) {
        DenseVector r(n);
        r.setZero();
        return r;
    }
    Index size() const { return static_cast<Index>(v_.size()); }
    Index rows() const { return size(); }
    S& operator()(Index i) { return v_[static_cast<std::size_t>(i)]; }
    const S& operator()(Index i) const { return v_[static_cast<std::size_t>(i)]; }
    S* data() { return v_.data(); }
    const S* data() const { return v_.data(); }
    void setZero() { std::fill(v_.begin(), v_.end(), S(0)); }
    void setZero(Index n) { v_.assign(static_cast<std::size_t>(n), S(0)); }
    void resize(Index n) { v_.resize(static_cast<std::size_t>(n)); }
    bool allFinite() const {
        for (const S& e : v_)
            if (!std::isfinite(std::abs(e))) return false;
        return true;
    }

private:
    std::vector<S> v_;
};

template <class S>
class DenseMatrix {  // row-major
public:
    DenseMatrix() = default;
    DenseMatrix(Index r, Index c) : r_(r), c_(c), v_(static_cast<std::size_t>(r * c)) {}
    static DenseMatrix Zero(Index r, Index c) {
        DenseMatrix m(r, c);
        m.setZero();
        return m;
    }
    Index rows() const { return r_; }
    Index cols() const { return c_; }
    S& operator()(Index i, Index j) { return v_[static_cast<std::size_t>(i * c_ + j)]; }
    const S& operator()(Index i, Index j) const { return v_[static_cast<std::size_t>(i * c_ + j)]; }
    S* data() { return v_.data(); }
    const S* data() const { return v_.data(); }
    void setZero() { std::fill(v_.begin(), v_.end(), S(0)); }

private:
    Index r_ = 0, c_ = 0;
    std::vector<S> v_;
};

using VectorXd = DenseVector<double>;
using MatrixXd = DenseMatrix<double>;
using VectorXcd = DenseVector<std::complex<double>>;
using MatrixXcd = DenseMatrix<std::complex<double>>;
}  // namespace Eigen
#endif
