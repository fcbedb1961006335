// dqmcb/stackngf.h -- host mirror of `struct GF` and `class LDRStack` (reference include/stackngf.h).
// Same layout and behaviour (move-only stack, bounds-checked access throwing std::out_of_range).
// With the device engine these are host views: the engine refreshes Gtt[l] / stack entries it has
// recomputed (see dqmcb/dqmc.h for which entries are kept current).
#pragma once
#include <stdexcept>
#include <vector>

#include "arma_lite.h"
#include "stablelinalg.h"

struct GF {
    std::vector<arma::mat> Gtt;
    std::vector<arma::mat> Gt0;  // G(tau, 0)   (unequal-time: not populated by this build)
    std::vector<arma::mat> G0t;  // G(0, tau)
    double log_det_M;
    GF() : log_det_M(0.0) {}
};

class LDRStack {
   private:
    size_t n_stack_ = 0;
    std::vector<stablelinalg::LDR> stack_;

   public:
    LDRStack() = default;
    explicit LDRStack(size_t n_stack) : n_stack_(n_stack), stack_(n_stack) {}
    LDRStack(LDRStack&& other) noexcept : n_stack_(other.n_stack_), stack_(std::move(other.stack_)) {}
    LDRStack& operator=(LDRStack&& other) noexcept {
        if (this != &other) {
            n_stack_ = other.n_stack_;
            stack_ = std::move(other.stack_);
        }
        return *this;
    }
    LDRStack(const LDRStack&) = delete;
    LDRStack& operator=(const LDRStack&) = delete;
    stablelinalg::LDR& operator[](size_t idx) {
        if (idx >= n_stack_) throw std::out_of_range("LDR Stack index out of bounds");
        return stack_[idx];
    }
    const stablelinalg::LDR& operator[](size_t idx) const {
        if (idx >= n_stack_) throw std::out_of_range("LDR Stack index out of bounds");
        return stack_[idx];
    }
    constexpr size_t size() const noexcept { return n_stack_; }
    void set(size_t idx, const stablelinalg::LDR& ldr) {
        if (idx >= n_stack_) throw std::out_of_range("LDR Stack index out of bounds");
        stack_[idx] = ldr;
    }
};
