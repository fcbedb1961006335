// dqmcb/lattice.h -- the part of the reference's `Lattice` (include/lattice.h:24-107) that the sweep
// path and the equal-time measurement pipeline consume: dimensions, the periodic neighbour rule used
// to build K (source/model.cpp:39-60), lattice vectors and the k-point grid (include/lattice.h:36-49).
// Site positions / save_info belong to the on-disk layer.
#pragma once
#include <array>
#include <cmath>
#include <stdexcept>
#include <vector>

#include "utility.h"

class Lattice {
   private:
    std::array<double, 2> a1_, a2_, b1_, b2_;
    std::vector<std::array<double, 2>> orbs_;
    int L1_, L2_, n_orb_;
    std::vector<std::array<double, 2>> k_points_;

   public:
    Lattice(const utility::parameters& params, const std::array<double, 2>& a1, const std::array<double, 2>& a2,
            const std::vector<std::array<double, 2>>& orbs)
        : a1_(a1), a2_(a2), orbs_(orbs), L1_(params.getInt("Lattice", "L1")), L2_(params.getInt("Lattice", "L2")),
          n_orb_(static_cast<int>(orbs.size())) {
        if (L1_ <= 0 || L2_ <= 0 || n_orb_ == 0) throw std::invalid_argument("Bad lattice dims");  // include/lattice.h:33
        const double det = a1_[0] * a2_[1] - a1_[1] * a2_[0];
        if (std::abs(det) < 1e-12) throw std::invalid_argument("Singular lattice");  // include/lattice.h:37
        const double two_pi = 2 * 3.14159265358979323846;
        b1_ = {two_pi * a2_[1] / det / L1_, -two_pi * a2_[0] / det / L1_};
        b2_ = {-two_pi * a1_[1] / det / L2_, two_pi * a1_[0] / det / L2_};
        // q = -L/2+1 ... L/2 in each direction, k-point (n, m) at index n*L2 + m (include/lattice.h:41-49)
        k_points_.reserve((size_t)L1_ * L2_);
        for (int n = 0; n < L1_; ++n)
            for (int m = 0; m < L2_; ++m) {
                const int qx = n - L1_ / 2 + 1, qy = m - L2_ / 2 + 1;
                k_points_.push_back({qx * b1_[0] + qy * b2_[0], qx * b1_[1] + qy * b2_[1]});
            }
    }
    const std::array<double, 2>& a1() const noexcept { return a1_; }
    const std::array<double, 2>& a2() const noexcept { return a2_; }
    const std::array<double, 2>& b1() const noexcept { return b1_; }
    const std::array<double, 2>& b2() const noexcept { return b2_; }
    const std::vector<std::array<double, 2>>& k_points() const noexcept { return k_points_; }
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
