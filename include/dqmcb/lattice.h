// dqmcb/lattice.h -- the part of the reference's `Lattice` (include/lattice.h:24-107) that the sweep
// path consumes: dimensions and the periodic neighbour rule used to build K (source/model.cpp:39-60).
// k-points / site positions / save_info belong to the measurement layer (out of scope).
#pragma once
#include <array>
#include <stdexcept>
#include <vector>

#include "utility.h"

class Lattice {
   private:
    std::array<double, 2> a1_, a2_;
    std::vector<std::array<double, 2>> orbs_;
    int L1_, L2_, n_orb_;

   public:
    Lattice(const utility::parameters& params, const std::array<double, 2>& a1, const std::array<double, 2>& a2,
            const std::vector<std::array<double, 2>>& orbs)
        : a1_(a1), a2_(a2), orbs_(orbs), L1_(params.getInt("Lattice", "L1")), L2_(params.getInt("Lattice", "L2")),
          n_orb_(static_cast<int>(orbs.size())) {
        if (L1_ <= 0 || L2_ <= 0 || n_orb_ == 0) throw std::invalid_argument("Bad lattice dims");  // include/lattice.h:33
    }
    int n_cells() const noexcept { return L1_ * L2_; }
    int n_sites() const noexcept { return L1_ * L2_ * n_orb_; }
    const int& L1() const noexcept { return L1_; }
    const int& L2() const noexcept { return L2_; }
    const int& n_orb() const noexcept { return n_orb_; }
    // include/lattice.h:100-107
    int site_neighbors(int idx, std::array<int, 2> delta, int orb) const {
        const int cell = idx / n_orb_;
        const int ux = cell % L1_, uy = cell / L1_;
        const int tx = ((ux + delta[0]) % L1_ + L1_) % L1_;
        const int ty = ((uy + delta[1]) % L2_ + L2_) % L2_;
        return (ty * L1_ + tx) * n_orb_ + orb;
    }
};
