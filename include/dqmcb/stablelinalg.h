// dqmcb/stablelinalg.h -- host mirror of `namespace stablelinalg` (reference include/stablelinalg.h,
// source/stablelinalg.cpp): the LDR type and the free functions, same names and signatures.  Each call
// forwards to the batched CUDA implementation (batch = 1) through the C ABI; the device engine's
// sweeps never go through here (they keep the factors on the GPU), this is the unit-level surface.
#pragma once
#include <vector>

#include "arma_lite.h"
#include "detail.h"

namespace stablelinalg {

class LDR {
   private:
    arma::mat L_;
    arma::vec d_;
    arma::mat R_;

   public:
    LDR() = default;
    LDR(const arma::mat& L, const arma::vec& d, const arma::mat& R) : L_(L), d_(d), R_(R) {}
    const arma::mat& get_L() const { return L_; }
    const arma::vec& get_d() const { return d_; }
    const arma::mat& get_R() const { return R_; }
    // L * diag(d) * R: O(N^3) convenience for checks only (reference include/stablelinalg.h:29)
    arma::mat to_matrix() const {
        const arma::uword n = L_.n_rows;
        arma::mat M(n, R_.n_cols);
        for (arma::uword j = 0; j < R_.n_cols; ++j)
            for (arma::uword k = 0; k < n; ++k) {
                const double s = d_(k) * R_(k, j);
                for (arma::uword i = 0; i < n; ++i) M(i, j) += L_(i, k) * s;
            }
        return M;
    }
    int n_rows() const { return (int)L_.n_rows; }
    int n_cols() const { return (int)R_.n_cols; }
};

// CUDA device used by the stand-alone functions below (default 0)
inline int& device() {
    static int dev = 0;
    return dev;
}

// elementwise helpers (source/stablelinalg.cpp:9-27); O(N^2) glue, fused away inside the kernels
inline arma::mat diag_mul_mat(const arma::vec& diag, const arma::mat& m) {
    arma::mat o = m;
    for (arma::uword j = 0; j < m.n_cols; ++j)
        for (arma::uword i = 0; i < m.n_rows; ++i) o(i, j) = diag(i) * m(i, j);
    return o;
}
inline arma::mat mat_mul_diag(const arma::mat& m, const arma::vec& diag) {
    arma::mat o = m;
    for (arma::uword j = 0; j < m.n_cols; ++j)
        for (arma::uword i = 0; i < m.n_rows; ++i) o(i, j) = m(i, j) * diag(j);
    return o;
}
inline arma::mat I_minus_mat(const arma::mat& m) {
    arma::mat o = m;
    for (arma::uword j = 0; j < m.n_cols; ++j)
        for (arma::uword i = 0; i < m.n_rows; ++i) o(i, j) = ((i == j) ? 1.0 : 0.0) - m(i, j);
    return o;
}

namespace detail {
inline LDR make(int n) { return LDR(arma::mat(n, n), arma::vec(n), arma::mat(n, n)); }
inline double* w(const arma::mat& m) { return const_cast<double*>(m.memptr()); }
}  // namespace detail

// source/stablelinalg.cpp:35-55
inline LDR to_LDR(const arma::mat& M) {
    const int n = (int)M.n_rows;
    LDR o = detail::make(n);
    dqmcb::check(dqmcb_to_LDR(device(), n, 1, M.memptr(), detail::w(o.get_L()), detail::w(o.get_d()), detail::w(o.get_R())), nullptr,
                 "QR decomposition failed in to_LDR");
    return o;
}
// source/stablelinalg.cpp:57-67
inline LDR ldr_mul_mat(const LDR& F, const arma::mat& M) {
    const int n = (int)M.n_rows;
    LDR o = detail::make(n);
    dqmcb::check(dqmcb_ldr_mul_mat(device(), n, 1, F.get_L().memptr(), F.get_d().memptr(), F.get_R().memptr(), M.memptr(),
                                   detail::w(o.get_L()), detail::w(o.get_d()), detail::w(o.get_R())),
                 nullptr, "ldr_mul_mat");
    return o;
}
// source/stablelinalg.cpp:69-79
inline LDR mat_mul_ldr(const arma::mat& M, const LDR& F) {
    const int n = (int)M.n_rows;
    LDR o = detail::make(n);
    dqmcb::check(dqmcb_mat_mul_ldr(device(), n, 1, M.memptr(), F.get_L().memptr(), F.get_d().memptr(), F.get_R().memptr(),
                                   detail::w(o.get_L()), detail::w(o.get_d()), detail::w(o.get_R())),
                 nullptr, "mat_mul_ldr");
    return o;
}
// source/stablelinalg.cpp:81-92
inline LDR ldr_mul_ldr(const LDR& F1, const LDR& F2) {
    const int n = F1.n_rows();
    LDR o = detail::make(n);
    dqmcb::check(dqmcb_ldr_mul_ldr(device(), n, 1, F1.get_L().memptr(), F1.get_d().memptr(), F1.get_R().memptr(), F2.get_L().memptr(),
                                   F2.get_d().memptr(), F2.get_R().memptr(), detail::w(o.get_L()), detail::w(o.get_d()),
                                   detail::w(o.get_R())),
                 nullptr, "ldr_mul_ldr");
    return o;
}
// source/stablelinalg.cpp:94-126
inline arma::mat inv_I_plus_ldr(const LDR& F, double& log_det_M_out) {
    const int n = F.n_rows();
    arma::mat G(n, n);
    dqmcb::check(dqmcb_inv_I_plus_ldr(device(), n, 1, F.get_L().memptr(), F.get_d().memptr(), F.get_R().memptr(), G.memptr(),
                                      &log_det_M_out),
                 nullptr, "inv_I_plus_ldr");
    return G;
}
// source/stablelinalg.cpp:128-158
inline arma::mat inv_I_plus_ldr_mul_ldr(const LDR& F1, const LDR& F2) {
    const int n = F1.n_rows();
    arma::mat G(n, n);
    dqmcb::check(dqmcb_inv_I_plus_ldr_mul_ldr(device(), n, 1, F1.get_L().memptr(), F1.get_d().memptr(), F1.get_R().memptr(),
                                              F2.get_L().memptr(), F2.get_d().memptr(), F2.get_R().memptr(), G.memptr()),
                 nullptr, "inv_I_plus_ldr_mul_ldr");
    return G;
}
// source/stablelinalg.cpp:160-190
inline arma::mat inv_invldr_plus_ldr(const LDR& F1, const LDR& F2) {
    const int n = F1.n_rows();
    arma::mat G(n, n);
    dqmcb::check(dqmcb_inv_invldr_plus_ldr(device(), n, 1, F1.get_L().memptr(), F1.get_d().memptr(), F1.get_R().memptr(),
                                           F2.get_L().memptr(), F2.get_d().memptr(), F2.get_R().memptr(), G.memptr()),
                 nullptr, "inv_invldr_plus_ldr");
    return G;
}

}  // namespace stablelinalg
