// dqmcb/dqmc.h -- host mirror of `class DQMC` (reference include/dqmc.h, source/dqmc.cpp): the sweep
// engine.  Same constructor and public members; the work runs in the CUDA engine owned by the model.
//
// Host-view contract (the device is authoritative while sweeping):
//   * init_stacks / init_greenfunctions return host copies of what the device computed;
//   * after sweep_0_to_beta   greens[0].Gtt[nt] and log_det_M are current (reference: source/dqmc.cpp:383),
//     after sweep_beta_to_0   greens[0].Gtt[0]  and log_det_M are current (source/dqmc.cpp:443);
//     the intermediate Gtt[l] of the reference (touched two-adjacent-at-a-time) are not materialised;
//   * with [simulation] isMeasureUnequalTime, sweep_unequalTime fills greens[0].Gtt / Gt0 / G0t [0..nt]
//     from the device (source/dqmc.cpp:458-514) and half_warp transforms all three (source/dqmc.cpp:303-312);
//   * model.fields() and every stack entry are refreshed after each sweep (mirror_stacks(false)
//     turns the stack copy off for throughput runs);
//   * the random stream is drawn from model.rng() with the reference's own libstdc++ calls
//     (std::shuffle, by-value uniform_int, bernoulli's generate_canonical), so a given seed yields the
//     reference's trajectory.
#pragma once
#include <algorithm>
#include <cmath>
#include <iostream>
#include <numeric>
#include <random>
#include <vector>

#include "model.h"
#include "stablelinalg.h"
#include "stackngf.h"
#include "utility.h"

namespace dqmcb {

// Draws of `n_calls` consecutive update::local_update calls (source/update.cpp:10-25) taken from
// `gen` exactly as the reference takes them: shuffle of the site order, proposal index from a COPY
// of the generator (include/field.h:76-83 takes utility::random by value), then the canonical double
// that rng.bernoulli(p) compares with p.  Appends to perm / k / u in visiting order.
inline void draw_tape(std::mt19937& gen, int nv, int n_calls, std::vector<uint8_t>& perm, std::vector<uint8_t>& k, std::vector<double>& u) {
    std::vector<int> order(nv);
    for (int c = 0; c < n_calls; ++c) {
        std::iota(order.begin(), order.end(), 0);
        std::shuffle(order.begin(), order.end(), gen);
        for (int idx = 0; idx < nv; ++idx) {
            perm.push_back((uint8_t)order[idx]);
            {
                std::mt19937 copy = gen;
                std::uniform_int_distribution<int> dist(0, 2);
                k.push_back((uint8_t)dist(copy));
            }
            u.push_back(std::generate_canonical<double, std::numeric_limits<double>::digits>(gen));
        }
    }
}

}  // namespace dqmcb

class DQMC {
   private:
    AttractiveHubbard& model_;
    bool isUnequalTime_;
    int nt_, n_stab_, n_stack_;
    bool mirror_stacks_ = true;
    double warned_err_ = 0.0;
    double unequal_err_ = 0.0;
    std::vector<std::mt19937> extra_gens_;  // walkers 1.. of a batch (walker 0 follows model.rng())

    dqmcb_ctx* ctx() { return model_.ctx(); }

    void load_tape(int n_calls) {
        const int W = model_.n_walkers(), nv = model_.ns();
        std::vector<uint8_t> perm, k;
        std::vector<double> u;
        perm.reserve((size_t)W * n_calls * nv);
        k.reserve(perm.capacity());
        u.reserve(perm.capacity());
        // walker 0 follows model.rng() (the reference's stream); further walkers of a batch use
        // generators derived from it so that chains are independent
        dqmcb::draw_tape(model_.rng().get_generator(), nv, n_calls, perm, k, u);
        if ((int)extra_gens_.size() < W - 1)
            for (int w = (int)extra_gens_.size() + 1; w < W; ++w) extra_gens_.emplace_back(104729u + 7919u * (unsigned)w);
        for (int w = 1; w < W; ++w) dqmcb::draw_tape(extra_gens_[w - 1], nv, n_calls, perm, k, u);
        dqmcb::check(dqmcb_set_tape(ctx(), n_calls, perm.data(), k.data(), u.data()), ctx(), "set_tape");
    }
    void refresh_stack(LDRStack& stack) {
        const int n = model_.ns(), W = model_.n_walkers();
        if (stack.size() != (size_t)n_stack_) stack = LDRStack(n_stack_);
        std::vector<double> L((size_t)W * n * n), R((size_t)W * n * n), d((size_t)W * n);
        for (int i = 0; i < n_stack_; ++i) {
            dqmcb::check(dqmcb_get_ldr(ctx(), i, L.data(), d.data(), R.data()), ctx(), "get_ldr");
            arma::mat Lm(n, n), Rm(n, n);
            arma::vec dv(n);
            std::copy(L.begin(), L.begin() + (size_t)n * n, Lm.memptr());
            std::copy(R.begin(), R.begin() + (size_t)n * n, Rm.memptr());
            std::copy(d.begin(), d.begin() + n, dv.memptr());
            stack.set(i, stablelinalg::LDR(Lm, dv, Rm));
        }
    }
    void after_sweep(std::vector<GF>& greens, std::vector<LDRStack>& stacks, int l_current) {
        if (greens[0].Gtt.size() != (size_t)nt_ + 1) greens[0].Gtt.resize(nt_ + 1);
        model_.download_G(greens[0].Gtt[l_current]);
        std::vector<double> ld(model_.n_walkers());
        dqmcb::check(dqmcb_get_logdet(ctx(), ld.data()), ctx(), "get_logdet");
        greens[0].log_det_M = ld[0];
        model_.refresh_fields_from_device();
        if (mirror_stacks_) refresh_stack(stacks[0]);
        // source/dqmc.cpp:390-393: the reference warns per stabilisation; here once per sweep when it grows
        const double e = max_err();
        if (e > 1e-6 && e > warned_err_) {
            std::cerr << "Warning: large precision error in stabilised Green's function: " << e << std::endl;
            warned_err_ = e;
        }
    }

   public:
    DQMC(const utility::parameters& params, AttractiveHubbard& model) : model_(model) {
        nt_ = params.getInt("simulation", "nt");
        n_stab_ = params.getInt("simulation", "n_stab");
        n_stack_ = (int)std::ceil(static_cast<double>(nt_) / n_stab_);  // source/dqmc.cpp:10
        isUnequalTime_ = params.getBool("simulation", "isMeasureUnequalTime", false);
        if (n_stack_ != dqmcb_n_stack(ctx())) throw std::invalid_argument("DQMC: nt / n_stab differ from the model's context");
    }

    // include/dqmc.h:80-82 (walker 0; all walkers through dqmcb_get_stats on model.ctx())
    double acc_rate() { return stat(0); }
    double max_err() { return stat(1); }
    double mean_err() { return stat(2); }
    double stat(int which) {
        const int W = model_.n_walkers();
        std::vector<double> a(W), m(W), e(W);
        dqmcb::check(dqmcb_get_stats(ctx(), a.data(), m.data(), e.data()), ctx(), "get_stats");
        return which == 0 ? a[0] : (which == 1 ? m[0] : e[0]);
    }
    void mirror_stacks(bool on) { mirror_stacks_ = on; }
    bool is_unequal_time() const { return isUnequalTime_; }
    int n_stack() const { return n_stack_; }

    // source/dqmc.cpp:43-59
    LDRStack init_stacks(int /*flv*/) {
        model_.sync_fields_to_device(true);
        dqmcb::check(dqmcb_init_stacks(ctx()), ctx(), "init_stacks");
        LDRStack s(n_stack_);
        refresh_stack(s);
        return s;
    }
    // source/dqmc.cpp:61-72: G(0) = [I + stack[0]]^-1 from the stack that is passed in
    GF init_greenfunctions(LDRStack& propagation_stack) {
        const int n = model_.ns(), W = model_.n_walkers();
        if (W == 1) {  // honour a caller-modified stack[0]
            const stablelinalg::LDR& F = propagation_stack[0];
            dqmcb::check(dqmcb_set_ldr(ctx(), 0, F.get_L().memptr(), F.get_d().memptr(), F.get_R().memptr()), ctx(), "set_ldr");
        }
        dqmcb::check(dqmcb_init_greenfunctions(ctx()), ctx(), "init_greenfunctions");
        GF greens;
        greens.Gtt.resize(nt_ + 1);
        greens.Gt0.resize(nt_ + 1);
        greens.G0t.resize(nt_ + 1);
        model_.download_G(greens.Gtt[0]);
        std::vector<double> ld(W);
        dqmcb::check(dqmcb_get_logdet(ctx(), ld.data()), ctx(), "get_logdet");
        greens.log_det_M = ld[0];
        (void)n;
        return greens;
    }
    // source/dqmc.cpp:288-315.  Equal-time: Gtt[0] of GF_asymm is half-warped.  With isMeasureUnequalTime the
    // three arrays of the device's last sweep_unequalTime are half-warped for every t < nt (the reference
    // transforms GF_asymm's host copies, which sweep_unequalTime below mirrors from the same device data).
    void half_warp(std::vector<GF>& GF_tosymm, std::vector<GF>& GF_asymm) {
        model_.upload_G(GF_asymm[0].Gtt[0]);
        const int n = model_.ns(), W = model_.n_walkers();
        std::vector<double> out((size_t)W * n * n);
        dqmcb::check(dqmcb_half_wrap(ctx(), out.data()), ctx(), "half_wrap");
        if (GF_tosymm[0].Gtt.empty()) GF_tosymm[0].Gtt.resize(nt_ + 1);
        GF_tosymm[0].Gtt[0].set_size(n, n);
        std::copy(out.begin(), out.begin() + (size_t)n * n, GF_tosymm[0].Gtt[0].memptr());
        if (!isUnequalTime_) return;
        GF_tosymm[0].Gtt.resize(nt_ + 1);
        GF_tosymm[0].Gt0.resize(nt_ + 1);
        GF_tosymm[0].G0t.resize(nt_ + 1);
        for (int which = 0; which < 3; ++which) {
            std::vector<arma::mat>& dst = which == 0 ? GF_tosymm[0].Gtt : (which == 1 ? GF_tosymm[0].Gt0 : GF_tosymm[0].G0t);
            for (int t = (which == 0 ? 1 : 0); t < nt_; ++t) {
                dqmcb::check(dqmcb_half_wrap_unequal(ctx(), which, t, out.data()), ctx(), "half_wrap_unequal");
                dst[t].set_size(n, n);
                std::copy(out.begin(), out.begin() + (size_t)n * n, dst[t].memptr());
            }
        }
    }
    // source/dqmc.cpp:337-396
    void sweep_0_to_beta(std::vector<GF>& greens, std::vector<LDRStack>& propagation_stacks) {
        load_tape(nt_);
        dqmcb::check(dqmcb_sweep_0_to_beta(ctx()), ctx(), "sweep_0_to_beta");
        after_sweep(greens, propagation_stacks, nt_);
    }
    // source/dqmc.cpp:398-456
    void sweep_beta_to_0(std::vector<GF>& greens, std::vector<LDRStack>& propagation_stacks) {
        load_tape(nt_);
        dqmcb::check(dqmcb_sweep_beta_to_0(ctx()), ctx(), "sweep_beta_to_0");
        after_sweep(greens, propagation_stacks, 0);
    }
    // source/dqmc.cpp:458-514: no-op unless [simulation] isMeasureUnequalTime; otherwise the device computes
    // G(tau,tau), G(tau,0), G(0,tau) for tau = 0..nt (stabilised at every block end) and walker 0's are
    // mirrored into greens[0].Gtt / Gt0 / G0t like the reference fills them.  The stack is read-only.
    void sweep_unequalTime(std::vector<GF>& greens, std::vector<LDRStack>&) {
        if (!isUnequalTime_) return;
        dqmcb::check(dqmcb_sweep_unequal_time(ctx()), ctx(), "sweep_unequal_time");
        const int n = model_.ns(), W = model_.n_walkers();
        std::vector<double> buf((size_t)W * n * n);
        greens[0].Gtt.resize(nt_ + 1);
        greens[0].Gt0.resize(nt_ + 1);
        greens[0].G0t.resize(nt_ + 1);
        for (int which = 0; which < 3; ++which) {
            std::vector<arma::mat>& dst = which == 0 ? greens[0].Gtt : (which == 1 ? greens[0].Gt0 : greens[0].G0t);
            for (int l = 0; l <= nt_; ++l) {
                dqmcb::check(dqmcb_get_G_unequal(ctx(), which, l, buf.data()), ctx(), "get_G_unequal");
                dst[l].set_size(n, n);
                std::copy(buf.begin(), buf.begin() + (size_t)n * n, dst[l].memptr());
            }
        }
        std::vector<double> ld(W), err(W);
        dqmcb::check(dqmcb_get_logdet(ctx(), ld.data()), ctx(), "get_logdet");
        greens[0].log_det_M = ld[0];
        dqmcb::check(dqmcb_get_unequal_err(ctx(), err.data()), ctx(), "get_unequal_err");
        unequal_err_ = err[0];
        if (err[0] > 1e-6)  // source/dqmc.cpp:505-508
            std::cerr << "WARNING: GF unequaltime precision > 1e-6. Reduce n_stab or increasing nt. Error: " << err[0] << std::endl;
    }
    double unequal_err() const { return unequal_err_; }
};
