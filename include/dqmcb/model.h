// dqmcb/model.h -- host mirror of `class AttractiveHubbard` and `namespace Observables`
// (reference include/model.h, source/model.cpp:1-288).  Same constructor and member signatures.
//
// The object owns a dqmcb_ctx (the device engine): fields, the current equal-time Green's function,
// the LDR stack and the B-matrix factors live on the GPU.  Host members (`fields()`, the matrices
// returned by expK()...) are views for code written against the reference's interface.
// Extension over the reference: [simulation] n_walkers (default 1) batches that many independent
// chains in the context; the reference-style single-chain interface then addresses walker 0 and the
// batched state is reachable through ctx().
#pragma once
#include <cmath>
#include <utility>
#include <vector>

#include "arma_lite.h"
#include "detail.h"
#include "field.h"
#include "lattice.h"
#include "stablelinalg.h"
#include "stackngf.h"
#include "utility.h"

class AttractiveHubbard {
   private:
    double t_;
    double mu_;
    double g_;      // sqrt(dtau * |U| / 2), source/model.cpp:27
    double alpha_;  // -1, source/model.cpp:28
    int ns_;
    int nt_;
    arma::mat expK_, invexpK_, expKhalf_, invexpKhalf_;
    GHQField fields_;
    utility::random& rng_;
    dqmcb_ctx* ctx_ = nullptr;
    int n_walkers_ = 1;
    unsigned long long synced_version_ = ~0ull;

    // source/model.cpp:39-60 (assignment, not accumulation: for L = 2 the +-x bonds coincide)
    arma::mat build_K_matrix(const Lattice& lat) {
        const int n = lat.n_sites();
        arma::mat K(n, n);
        for (int i = 0; i < n; ++i) {
            K(i, i) = -mu_;
            const int nx = lat.site_neighbors(i, {1, 0}, 0);
            K(i, nx) = -t_;
            K(nx, i) = -t_;
            const int ny = lat.site_neighbors(i, {0, 1}, 0);
            K(i, ny) = -t_;
            K(ny, i) = -t_;
        }
        return K;
    }

   public:
    AttractiveHubbard(const utility::parameters& params, const Lattice& lat, utility::random& rng, double replica_beta) : rng_(rng) {
        t_ = params.getDouble("hubbard", "t");
        mu_ = params.getDouble("hubbard", "mu");
        ns_ = lat.n_cells();
        nt_ = (int)params.getDouble("simulation", "nt");
        const double U = params.getDouble("hubbard", "U");
        const double dtau = replica_beta / nt_;
        fields_ = GHQField(nt_, ns_, rng);  // by value: the caller's generator is not advanced (Q1)
        g_ = std::sqrt(0.5 * std::fabs(U) * dtau);
        alpha_ = -1.0;
        // exp(-+dtau K), exp(-+dtau K/2): K is real symmetric, so the eigen form equals arma::expmat
        const arma::mat K = build_K_matrix(lat);
        std::vector<double> w;
        arma::mat V;
        dqmcb::sym_eig_jacobi(K, w, V);
        expK_ = dqmcb::expm_sym(w, V, -dtau);
        invexpK_ = dqmcb::expm_sym(w, V, dtau);
        expKhalf_ = dqmcb::expm_sym(w, V, -0.5 * dtau);
        invexpKhalf_ = dqmcb::expm_sym(w, V, 0.5 * dtau);
        // device engine
        n_walkers_ = params.getInt("simulation", "n_walkers", 1);
        dqmcb_config cfg;
        cfg.n_sites = ns_;
        cfg.nt = nt_;
        cfg.n_stab = params.getInt("simulation", "n_stab");
        cfg.n_walkers = n_walkers_;
        cfg.device = params.getInt("simulation", "device", 0);
        cfg.delay_k = params.getInt("simulation", "delay_k", 0);
        cfg.g = g_;
        cfg.alpha = alpha_;
        for (int s = 0; s < 4; ++s) {
            cfg.eta[s] = fields_.eta(s);
            cfg.gamma[s] = fields_.gamma(s);
        }
        dqmcb::check(dqmcb_create(&cfg, expK_.memptr(), invexpK_.memptr(), &ctx_), nullptr, "dqmcb_create");
        dqmcb::check(dqmcb_set_half_expK(ctx_, expKhalf_.memptr(), invexpKhalf_.memptr()), ctx_, "dqmcb_set_half_expK");
    }
    ~AttractiveHubbard() { dqmcb_destroy(ctx_); }
    AttractiveHubbard(const AttractiveHubbard&) = delete;
    AttractiveHubbard& operator=(const AttractiveHubbard&) = delete;

    const arma::mat& expK(int) const { return expK_; }
    const arma::mat& invexpK(int) const { return invexpK_; }
    const arma::mat& expK_half(int) const { return expKhalf_; }
    const arma::mat& invexpK_half(int) const { return invexpKhalf_; }

    // source/model.cpp:62-84 (host view; the kernels use an 8-entry LUT of the same expressions)
    arma::vec expV(int l, int) {
        arma::vec v(ns_);
        for (int i = 0; i < ns_; ++i) v(i) = std::exp(g_ * fields_.eta(fields_.single_val(l, i)));
        return v;
    }
    arma::vec invexpV(int l, int) {
        arma::vec v(ns_);
        for (int i = 0; i < ns_; ++i) v(i) = std::exp(-g_ * fields_.eta(fields_.single_val(l, i)));
        return v;
    }

    GHQField& fields() { return fields_; }
    int nt() const { return nt_; }
    int ns() const { return ns_; }
    int n_flavor() const { return 1; }
    utility::random& rng() { return rng_; }

    // source/model.cpp:90-97
    double det_ratio(arma::mat& gtt, double delta, int i) {
        const double det_flv = 1.0 + (1.0 - gtt(i, i)) * delta;
        return det_flv * det_flv;
    }
    // source/model.cpp:99-107
    std::pair<double, double> bosonic_ratio(int new_field, int old_field) {
        const double d_eta = fields_.eta(new_field) - fields_.eta(old_field);
        const double boson = std::exp(alpha_ * g_ * d_eta);
        return {boson, 1.0 / boson - 1.0};
    }
    // source/model.cpp:109-122: O(1) scalar on the caller's GF, as in the reference.  (The Metropolis
    // kernel evaluates the same expression from its device-resident G; dqmcb_local_update_ratio exposes it.)
    std::pair<double, double> local_update_ratio(std::vector<GF>& greens, int l, int field_idx, int new_field) {
        const int old_field = fields_.single_val(l, field_idx);
        const double gammaR = fields_.gamma(new_field) / fields_.gamma(old_field);
        const auto br = bosonic_ratio(new_field, old_field);
        const double detR = det_ratio(greens[0].Gtt[l + 1], br.second, field_idx);
        return {gammaR * br.first * detR, br.second};
    }
    // source/model.cpp:124-138: rank-1 Sherman-Morrison, executed on the device on the caller's Gtt[l+1]
    void update_greens_local(std::vector<GF>& greens, double delta, int l, int field_idx) {
        arma::mat& G = greens[0].Gtt[l + 1];
        upload_G(G);
        dqmcb::check(dqmcb_update_greens_local(ctx_, 0, delta, field_idx), ctx_, "update_greens_local");
        download_G(G);
    }
    // source/model.cpp:140-159: the field sum runs on the device; log_det_M is taken from the caller's GF
    double global_action(const std::vector<GF>& greens) {
        sync_fields_to_device();
        std::vector<double> S(n_walkers_), ld(n_walkers_);
        dqmcb::check(dqmcb_global_action(ctx_, S.data()), ctx_, "global_action");
        dqmcb::check(dqmcb_get_logdet(ctx_, ld.data()), ctx_, "get_logdet");
        return S[0] + 2.0 * ld[0] - 2.0 * greens[0].log_det_M;
    }

    // ---- device plumbing used by DQMC / update (not in the reference) ----
    dqmcb_ctx* ctx() { return ctx_; }
    int n_walkers() const { return n_walkers_; }
    double g() const { return g_; }
    // push host fields (walker 0; every walker when `all` and the contexts batches copies) if edited
    void sync_fields_to_device(bool force = false) {
        if (!force && synced_version_ == fields_.version()) return;
        dqmcb::check(dqmcb_set_fields_i64(ctx_, 0, fields_.fields().memptr()), ctx_, "set_fields");
        synced_version_ = fields_.version();
    }
    void refresh_fields_from_device() {
        dqmcb::check(dqmcb_get_fields_i64(ctx_, 0, fields_.fields_mut().memptr()), ctx_, "get_fields");
        synced_version_ = fields_.version();
    }
    // single-walker G transfer helpers (walker 0 of the batch)
    void upload_G(const arma::mat& G) {
        if (n_walkers_ == 1) {
            dqmcb::check(dqmcb_set_G(ctx_, G.memptr()), ctx_, "set_G");
            return;
        }
        std::vector<double> all((size_t)n_walkers_ * ns_ * ns_);
        dqmcb::check(dqmcb_get_G(ctx_, all.data()), ctx_, "get_G");
        std::copy(G.memptr(), G.memptr() + (size_t)ns_ * ns_, all.begin());
        dqmcb::check(dqmcb_set_G(ctx_, all.data()), ctx_, "set_G");
    }
    void download_G(arma::mat& G) {
        std::vector<double> all((size_t)n_walkers_ * ns_ * ns_);
        dqmcb::check(dqmcb_get_G(ctx_, all.data()), ctx_, "get_G");
        G.set_size(ns_, ns_);
        std::copy(all.begin(), all.begin() + (size_t)ns_ * ns_, G.memptr());
    }
};

// Equal-time observables (source/model.cpp:165-288) from greens[0].Gtt[0], computed by the fused
// measurement kernel.  The Lattice argument is kept for signature compatibility.
namespace Observables {
namespace detail {
inline void measure(AttractiveHubbard& model, const std::vector<GF>& greens, double* scalars3, arma::mat* corr) {
    model.upload_G(greens[0].Gtt[0]);
    const int W = model.n_walkers(), n = model.ns();
    std::vector<double> sc((size_t)3 * W), cc;
    if (corr) cc.resize((size_t)W * n * n);
    dqmcb::check(dqmcb_measure_equal_time(model.ctx(), sc.data(), corr ? cc.data() : nullptr), model.ctx(), "measure_equal_time");
    for (int k = 0; k < 3; ++k) scalars3[k] = sc[k];
    if (corr) {
        corr->set_size(n, n);
        std::copy(cc.begin(), cc.begin() + (size_t)n * n, corr->memptr());
    }
}
}  // namespace detail
// The reference's free functions take (greens, lat); the device engine also needs the model that owns
// the context, passed as a leading argument.
inline double calculate_density(AttractiveHubbard& model, const std::vector<GF>& greens, const Lattice&) {
    double s[3];
    detail::measure(model, greens, s, nullptr);
    return s[0];
}
inline double calculate_doubleOccupancy(AttractiveHubbard& model, const std::vector<GF>& greens, const Lattice&) {
    double s[3];
    detail::measure(model, greens, s, nullptr);
    return s[1];
}
inline double calculate_swavePairing(AttractiveHubbard& model, const std::vector<GF>& greens, const Lattice&) {
    double s[3];
    detail::measure(model, greens, s, nullptr);
    return s[2];
}
inline arma::mat calculate_densityCorr(AttractiveHubbard& model, const std::vector<GF>& greens, const Lattice&) {
    double s[3];
    arma::mat c;
    detail::measure(model, greens, s, &c);
    return c;
}
// source/model.cpp:290-392.  Evaluated on the device from the arrays of the last DQMC::sweep_unequalTime (the
// host copies in `greens` are mirrors of the same data): (N, N, nt+1) cubes of walker 0.
namespace detail {
inline arma::cube tau_cube(AttractiveHubbard& model, const Lattice& lat, int which) {
    const int W = model.n_walkers(), n = model.ns(), nt1 = model.nt() + 1;
    std::vector<double> buf((size_t)W * nt1 * n * n);
    double* outs[3] = {nullptr, nullptr, nullptr};
    outs[which] = buf.data();
    dqmcb::check(dqmcb_measure_unequal_time(model.ctx(), lat.L1(), lat.L2(), outs[0], outs[1], outs[2]), model.ctx(), "measure_unequal_time");
    arma::cube c(n, n, nt1);
    std::copy(buf.begin(), buf.begin() + (size_t)nt1 * n * n, c.memptr());
    return c;
}
}  // namespace detail
inline arma::cube calculate_greenTau(AttractiveHubbard& model, const std::vector<GF>&, const Lattice& lat) { return detail::tau_cube(model, lat, 0); }
inline arma::cube calculate_doublonTau(AttractiveHubbard& model, const std::vector<GF>&, const Lattice& lat) { return detail::tau_cube(model, lat, 1); }
inline arma::cube calculate_currxxTau(AttractiveHubbard& model, const std::vector<GF>&, const Lattice& lat) { return detail::tau_cube(model, lat, 2); }
}  // namespace Observables
