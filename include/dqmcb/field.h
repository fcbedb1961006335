// dqmcb/field.h -- host mirror of `class GHQField` (reference include/field.h:13-84): the 4-state
// Gauss-Hermite HS field.  Same members and semantics, including the by-value RNG arguments (the
// caller's generator is not advanced by the ctor or by propose_new_field).  The
// authoritative copy during sweeps lives on the device as int8 [nt][N]; this object is its host view
// (`version()` lets the engine notice host-side edits such as replica exchange's set_fields).
#pragma once
#include <cmath>
#include <random>

#include "arma_lite.h"
#include "utility.h"

class GHQField {
   private:
    arma::vec gamma_;
    arma::vec eta_;
    arma::imat proposal_;
    arma::imat fields_;
    unsigned long long version_ = 0;

   public:
    GHQField() = default;
    GHQField(int nt, int nv, utility::random rng) {
        gamma_.set_size(4);
        eta_.set_size(4);
        const double s6 = std::sqrt(6.0);
        const double g[4] = {1.0 - s6 / 3.0, 1.0 + s6 / 3.0, 1.0 + s6 / 3.0, 1.0 - s6 / 3.0};
        const double e[4] = {-std::sqrt(2.0 * (3.0 + s6)), -std::sqrt(2.0 * (3.0 - s6)), std::sqrt(2.0 * (3.0 - s6)),
                             std::sqrt(2.0 * (3.0 + s6))};
        for (int i = 0; i < 4; ++i) {
            gamma_(i) = g[i];
            eta_(i) = e[i];
        }
        proposal_.set_size(4, 3);
        for (int o = 0; o < 4; ++o)
            for (int k = 0, v = 0; v < 4; ++v)
                if (v != o) proposal_(o, k++) = v;  // the k-th element of {0,1,2,3} \ {old}, ascending
        fields_.set_size(nt, nv);
        std::uniform_int_distribution<int> dist(0, 3);
        for (arma::uword i = 0; i < fields_.n_elem; ++i) fields_(i) = dist(rng.get_generator());
    }
    double gamma(int f) const { return gamma_(f); }
    double eta(int f) const { return eta_(f); }
    int single_val(int time_slice, int field_idx) const { return (int)fields_(time_slice, field_idx); }
    const arma::imat& fields() const { return fields_; }
    int nv() const { return (int)fields_.n_cols; }
    void set_single_field(int l, int i, int new_value) {
        fields_(l, i) = new_value;
        ++version_;
    }
    void set_fields(const arma::imat& new_fields) {
        fields_ = new_fields;
        ++version_;
    }
    int propose_new_field(int old_field, utility::random rng) const {
        std::uniform_int_distribution<int> dist(0, 2);
        const int k = dist(rng.get_generator());
        return (int)proposal_(old_field, k);
    }
    // ---- additions for the device-backed engine ----
    unsigned long long version() const { return version_; }
    arma::imat& fields_mut() { return fields_; }  // engine-side refresh after device sweeps (no version bump)
};
