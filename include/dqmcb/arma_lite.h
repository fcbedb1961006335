// dqmcb/arma_lite.h -- the small subset of Armadillo's types that the reference's hot-path interface
// exposes (arma::mat / vec / imat / uword: dense, column-major, value semantics), for hosts without
// Armadillo.  Layout-identical to Armadillo (memptr(), n_rows, n_cols, n_elem, column-major), so a
// build that has the real library can define DQMCB_USE_ARMADILLO and get <armadillo> instead.
//
// This is container plumbing only: no linear algebra lives here (that is the CUDA library's job);
// the two numeric helpers at the bottom (symmetric eigen-decomposition for exp(-dtau K), used once
// at model construction like arma::expmat in source/model.cpp:32-35) are start-up code.
#pragma once
#ifdef DQMCB_USE_ARMADILLO
#include <armadillo>
#else
#include <algorithm>
#include <cmath>
#include <complex>
#include <cstddef>
#include <cstring>
#include <initializer_list>
#include <stdexcept>
#include <vector>

namespace arma {

typedef unsigned long long uword;

template <typename T>
class Mat {
   public:
    uword n_rows = 0, n_cols = 0, n_elem = 0;
    Mat() = default;
    Mat(uword r, uword c) { set_size(r, c); }
    Mat(std::initializer_list<std::initializer_list<T>> rows) {
        set_size(rows.size(), rows.size() ? rows.begin()->size() : 0);
        uword i = 0;
        for (const auto& row : rows) {
            uword j = 0;
            for (const T& v : row) (*this)(i, j++) = v;
            ++i;
        }
    }
    void set_size(uword r, uword c) {
        n_rows = r;
        n_cols = c;
        n_elem = r * c;
        data_.assign(n_elem, T());
    }
    void zeros() { std::fill(data_.begin(), data_.end(), T()); }
    void zeros(uword r, uword c) { set_size(r, c); }
    void eye(uword r, uword c) {
        set_size(r, c);
        for (uword i = 0; i < std::min(r, c); ++i) (*this)(i, i) = T(1);
    }
    void fill(T v) { std::fill(data_.begin(), data_.end(), v); }
    T* memptr() { return data_.data(); }
    const T* memptr() const { return data_.data(); }
    T& operator()(uword i, uword j) { return data_[i + j * n_rows]; }
    const T& operator()(uword i, uword j) const { return data_[i + j * n_rows]; }
    T& operator()(uword i) { return data_[i]; }
    const T& operator()(uword i) const { return data_[i]; }
    T& at(uword i, uword j) {
        if (i >= n_rows || j >= n_cols) throw std::out_of_range("Mat::at(): index out of bounds");
        return data_[i + j * n_rows];
    }
    bool is_empty() const { return n_elem == 0; }

   private:
    std::vector<T> data_;
};

template <typename T>
class Col : public Mat<T> {
   public:
    Col() = default;
    explicit Col(uword n) : Mat<T>(n, 1) {}
    void set_size(uword n) { Mat<T>::set_size(n, 1); }
    void zeros(uword n) { Mat<T>::set_size(n, 1); }
    using Mat<T>::zeros;
};

// (n_rows, n_cols, n_slices), column-major with slices consecutive -- arma::Cube's memory order
template <typename T>
class Cube {
   public:
    uword n_rows = 0, n_cols = 0, n_slices = 0, n_elem = 0;
    Cube() = default;
    Cube(uword r, uword c, uword s) { set_size(r, c, s); }
    void set_size(uword r, uword c, uword s) {
        n_rows = r;
        n_cols = c;
        n_slices = s;
        n_elem = r * c * s;
        data_.assign(n_elem, T());
    }
    void zeros() { std::fill(data_.begin(), data_.end(), T()); }
    T* memptr() { return data_.data(); }
    const T* memptr() const { return data_.data(); }
    T* slice_memptr(uword s) { return data_.data() + s * n_rows * n_cols; }
    const T* slice_memptr(uword s) const { return data_.data() + s * n_rows * n_cols; }
    T& operator()(uword i, uword j, uword k) { return data_[i + n_rows * (j + n_cols * k)]; }
    const T& operator()(uword i, uword j, uword k) const { return data_[i + n_rows * (j + n_cols * k)]; }

   private:
    std::vector<T> data_;
};

typedef Mat<double> mat;
typedef Col<double> vec;
typedef Mat<long long> imat;  // arma::imat is 64-bit
typedef Col<uword> uvec;
typedef std::complex<double> cx_double;
typedef Cube<double> cube;
typedef Cube<cx_double> cx_cube;

}  // namespace arma
#endif  // DQMCB_USE_ARMADILLO

namespace dqmcb {

// Symmetric eigen-decomposition by cyclic Jacobi rotations: A = V diag(w) V^T (A is n x n, column-major).
// Used once per model to form exp(-+dtau K) and exp(-+dtau K / 2); K is real symmetric
// (source/model.cpp:39-60), so exp(s K) = V diag(exp(s w)) V^T equals arma::expmat(s K) to rounding.
inline void sym_eig_jacobi(const arma::mat& A, std::vector<double>& w, arma::mat& V) {
    const std::size_t n = A.n_rows;
    arma::mat a = A;
    V.set_size(n, n);
    for (std::size_t i = 0; i < n; ++i) V(i, i) = 1.0;
    for (int sweep = 0; sweep < 100; ++sweep) {
        double off = 0.0;
        for (std::size_t p = 0; p < n; ++p)
            for (std::size_t q = p + 1; q < n; ++q) off += a(p, q) * a(p, q);
        if (off < 1e-30) break;
        for (std::size_t p = 0; p < n; ++p)
            for (std::size_t q = p + 1; q < n; ++q) {
                const double apq = a(p, q);
                if (std::fabs(apq) < 1e-300) continue;
                const double theta = (a(q, q) - a(p, p)) / (2.0 * apq);
                const double t = (theta >= 0 ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0));
                const double c = 1.0 / std::sqrt(t * t + 1.0), s = t * c;
                for (std::size_t k = 0; k < n; ++k) {
                    const double akp = a(k, p), akq = a(k, q);
                    a(k, p) = c * akp - s * akq;
                    a(k, q) = s * akp + c * akq;
                }
                for (std::size_t k = 0; k < n; ++k) {
                    const double apk = a(p, k), aqk = a(q, k);
                    a(p, k) = c * apk - s * aqk;
                    a(q, k) = s * apk + c * aqk;
                }
                for (std::size_t k = 0; k < n; ++k) {
                    const double vkp = V(k, p), vkq = V(k, q);
                    V(k, p) = c * vkp - s * vkq;
                    V(k, q) = s * vkp + c * vkq;
                }
            }
    }
    w.resize(n);
    for (std::size_t i = 0; i < n; ++i) w[i] = a(i, i);
}

// exp(s * K) for symmetric K from its eigen-decomposition
inline arma::mat expm_sym(const std::vector<double>& w, const arma::mat& V, double s) {
    const std::size_t n = V.n_rows;
    arma::mat E(n, n);
    std::vector<double> ew(n);
    for (std::size_t k = 0; k < n; ++k) ew[k] = std::exp(s * w[k]);
    for (std::size_t j = 0; j < n; ++j)
        for (std::size_t i = 0; i < n; ++i) {
            double acc = 0.0;
            for (std::size_t k = 0; k < n; ++k) acc += V(i, k) * ew[k] * V(j, k);
            E(i, j) = acc;
        }
    return E;
}

}  // namespace dqmcb
