// dqmcb/measurementh5.h -- host mirror of the reference's measurement layer (include/measurementh5.h):
// namespace transform (pbc_shortest :13-17, chi_site_to_chi_r :20-76, chi_r_to_chi_k :78-116) and
// class MeasurementManager (:120-361), same names and call signatures, so that source/main.cpp:114-119
// and :165-170 compile against it.
//
// What is different underneath:
//  * the transforms run on the GPU (dqmcb_chi_site_to_chi_r / dqmcb_chi_r_to_chi_k);
//  * `measure_device(model)` is the batched path: one asynchronous call measures the four built-in
//    observables of source/main.cpp:116-119 for every walker of the model's context and accumulates
//    them on the device (dqmcb_bin_measure); `accumulate(lat)` then fetches 3 + L1*L2 doubles.
//    `measure(greens, lat)` keeps the reference's generic behaviour for user-registered calculators
//    (host accumulation of whatever the std::function returns) and can be mixed with it;
//  * results go to a dqmcb::BinStore (path, dims, row-major payload -- the tree of results/data_<rank>.h5,
//    include/measurementh5.h:294-361).  With -DDQMCB_USE_HDF5 the same tree is also written through the
//    HDF5 C API exactly as the reference does (file results/data_<rank>.h5); without HDF5 (this image)
//    `store().save_text(path)` dumps it as text.
//
// chi_r_to_chi_k: the reference files k-point index kidx = n*L2 + m under (kidx / L1, kidx % L2), which
// is (n, m) only on square lattices and out of bounds otherwise; (n, m) is used here.
#pragma once
#include <complex>
#include <cstdio>
#include <functional>
#include <string>
#include <vector>

#ifdef DQMCB_USE_HDF5
#include <hdf5.h>
#include <sys/stat.h>
#endif

#include "arma_lite.h"
#include "detail.h"
#include "lattice.h"
#include "model.h"
#include "stackngf.h"
#include "utility.h"

namespace dqmcb {

// One recorded dataset: dims as HDF5 shows them, payload in C (row-major) order.
struct BinDataset {
    std::string path;
    std::vector<unsigned long long> dims;
    std::vector<double> data;
};

class BinStore {
    std::vector<BinDataset> sets_;

   public:
    const std::vector<BinDataset>& datasets() const { return sets_; }
    const BinDataset* find(const std::string& path) const {
        for (const auto& d : sets_)
            if (d.path == path) return &d;
        return nullptr;
    }
    void put_scalar(const std::string& path, double v) { sets_.push_back({path, {1ull}, {v}}); }
    // arma::cube (r, c, s) -> dataset dims (r, c, s), element [i][j][k] = cube(i,j,k) (include/h5utils.h:50-79)
    void put_cube(const std::string& path, const arma::cube& c) {
        BinDataset d{path, {c.n_rows, c.n_cols, c.n_slices}, {}};
        d.data.resize(c.n_elem);
        for (arma::uword i = 0; i < c.n_rows; ++i)
            for (arma::uword j = 0; j < c.n_cols; ++j)
                for (arma::uword k = 0; k < c.n_slices; ++k) d.data[(i * c.n_cols + j) * c.n_slices + k] = c(i, j, k);
        sets_.push_back(std::move(d));
    }
    // complex cube -> dims (r, c, s, 2), trailing (re, im) (include/h5utils.h:81-119)
    void put_complex_cube(const std::string& path, const arma::cx_cube& c) {
        BinDataset d{path, {c.n_rows, c.n_cols, c.n_slices, 2ull}, {}};
        d.data.resize(2 * c.n_elem);
        for (arma::uword i = 0; i < c.n_rows; ++i)
            for (arma::uword j = 0; j < c.n_cols; ++j)
                for (arma::uword k = 0; k < c.n_slices; ++k) {
                    const size_t o = 2 * ((i * c.n_cols + j) * c.n_slices + k);
                    d.data[o] = c(i, j, k).real();
                    d.data[o + 1] = c(i, j, k).imag();
                }
        sets_.push_back(std::move(d));
    }
    // "<path> <rank> <dims...>\n<values...>\n" per dataset
    bool save_text(const std::string& file) const {
        std::FILE* f = std::fopen(file.c_str(), "w");
        if (!f) return false;
        for (const auto& d : sets_) {
            std::fprintf(f, "%s %zu", d.path.c_str(), d.dims.size());
            for (auto v : d.dims) std::fprintf(f, " %llu", v);
            std::fprintf(f, "\n");
            for (double v : d.data) std::fprintf(f, "%.17g ", v);
            std::fprintf(f, "\n");
        }
        std::fclose(f);
        return true;
    }
};

}  // namespace dqmcb

namespace transform {

inline int pbc_shortest(int d, int L) { return dqmcb_pbc_shortest(d, L); }

namespace detail {
inline void need_one_orbital(const Lattice& lat) {
    if (lat.n_orb() != 1) throw std::invalid_argument("transform: the device transforms support one orbital per cell (source/main.cpp:87)");
}
}  // namespace detail

// chi_site (N, N, n_tau) -> chi_r (L1, L2, n_tau), include/measurementh5.h:20-66
inline arma::cube chi_site_to_chi_r(const arma::cube& chi_site, const Lattice& lat, int device = 0) {
    detail::need_one_orbital(lat);
    arma::cube out(lat.L1(), lat.L2(), chi_site.n_slices);
    dqmcb::check(dqmcb_chi_site_to_chi_r(device, lat.L1(), lat.L2(), (int)chi_site.n_slices, chi_site.memptr(), out.memptr()), nullptr,
                 "chi_site_to_chi_r");
    return out;
}
// equal-time overload, include/measurementh5.h:68-76
inline arma::cube chi_site_to_chi_r(const arma::mat& chi_site, const Lattice& lat, int device = 0) {
    detail::need_one_orbital(lat);
    arma::cube out(lat.L1(), lat.L2(), 1);
    dqmcb::check(dqmcb_chi_site_to_chi_r(device, lat.L1(), lat.L2(), 1, chi_site.memptr(), out.memptr()), nullptr, "chi_site_to_chi_r");
    return out;
}
// chi_r (L1, L2, n) -> chi_k (L1, L2, n) complex, include/measurementh5.h:78-116
inline arma::cx_cube chi_r_to_chi_k(const arma::cube& chi_r, const Lattice& lat, int device = 0) {
    detail::need_one_orbital(lat);
    arma::cx_cube out(lat.L1(), lat.L2(), chi_r.n_slices);
    static_assert(sizeof(arma::cx_double) == 2 * sizeof(double), "interleaved complex layout");
    dqmcb::check(dqmcb_chi_r_to_chi_k(device, lat.L1(), lat.L2(), (int)chi_r.n_slices, chi_r.memptr(),
                                      reinterpret_cast<double*>(out.memptr())),
                 nullptr, "chi_r_to_chi_k");
    return out;
}

}  // namespace transform

class MeasurementManager {
   public:
    using ScalarFn = std::function<double(const std::vector<GF>&, const Lattice&)>;
    using EqTimeFn = std::function<arma::mat(const std::vector<GF>&, const Lattice&)>;
    using UneqTimeFn = std::function<arma::cube(const std::vector<GF>&, const Lattice&)>;

   private:
    struct Scalar {
        std::string name;
        ScalarFn fn;
        double sum = 0.0;
    };
    struct EqTime {
        std::string name;
        EqTimeFn fn;
        arma::mat sum;
    };
    struct UneqTime {
        std::string name;
        UneqTimeFn fn;
        arma::cube sum;
    };
    std::vector<Scalar> scalars_;
    std::vector<EqTime> eqtime_;
    std::vector<UneqTime> uneqtime_;
    long long dev_tau_count_ = 0;
    int rank_;
    bool unequal_;
    int host_count_ = 0;      // measure() calls in the open bin
    long long dev_count_ = 0; // samples the device bin holds
    AttractiveHubbard* dev_model_ = nullptr;
    int current_bin_ = 0;
    dqmcb::BinStore store_;
#ifdef DQMCB_USE_HDF5
    hid_t file_id_ = -1;
#endif

   public:
    // the communicator argument of the reference is kept for source compatibility and unused: every rank
    // writes its own file there too (include/measurementh5.h:288-292)
    template <typename Comm>
    MeasurementManager(const utility::parameters& params, Comm, int rank)
        : rank_(rank), unequal_(params.getBool("simulation", "isMeasureUnequalTime")) {}
    ~MeasurementManager() {
#ifdef DQMCB_USE_HDF5
        if (file_id_ >= 0) H5Fclose(file_id_);
#endif
    }

    void addScalar(const std::string& name, ScalarFn fn) { scalars_.push_back({name, std::move(fn), 0.0}); }
    void addEqualTime(const std::string& name, EqTimeFn fn) { eqtime_.push_back({name, std::move(fn), arma::mat()}); }
    void addUnequalTime(const std::string& name, UneqTimeFn fn) {  // ignored unless isMeasureUnequalTime (:181-187)
        if (!unequal_) return;
        uneqtime_.push_back({name, std::move(fn), arma::cube()});
    }

    // reference semantics: one sample per call from the host Green's functions (include/measurementh5.h:189-229)
    void measure(const std::vector<GF>& greens, const Lattice& lat) {
        for (auto& s : scalars_) s.sum += s.fn(greens, lat);
        for (auto& e : eqtime_) {
            arma::mat r = e.fn(greens, lat);
            if (e.sum.n_elem == 0) {
                e.sum = r;
            } else {
                for (arma::uword i = 0; i < r.n_elem; ++i) e.sum(i) += r(i);
            }
        }
        if (unequal_) {
            for (auto& u : uneqtime_) {
                arma::cube r = u.fn(greens, lat);
                if (u.sum.n_elem == 0) {
                    u.sum = r;
                } else {
                    for (arma::uword i = 0; i < r.n_elem; ++i) u.sum.memptr()[i] += r.memptr()[i];
                }
            }
        }
        ++host_count_;
    }

    // batched device path: density, doubleOcc, swave and densityCorr of every walker of `model`, from the
    // Green's functions resident on the GPU; W samples per call, nothing leaves the device
    void measure_device(AttractiveHubbard& model, const Lattice& lat) {
        if (dev_model_ != &model) {
            dqmcb::check(dqmcb_bin_configure(model.ctx(), lat.L1(), lat.L2()), model.ctx(), "bin_configure");
            dev_model_ = &model;
            dev_count_ = 0;
        }
        dqmcb::check(dqmcb_bin_measure(model.ctx()), model.ctx(), "bin_measure");
        dev_count_ += model.n_walkers();
        if (unequal_) {  // greenTau, doublonTau, currxxTau of the sweep_unequalTime that preceded this call
            dqmcb::check(dqmcb_bin_measure_unequal(model.ctx()), model.ctx(), "bin_measure_unequal");
            dev_tau_count_ += model.n_walkers();
        }
    }

    // include/measurementh5.h:231-274 (+ saveToHDF5 :276-361)
    void accumulate(const Lattice& lat) {
        const std::string bin_r = "/bin_" + std::to_string(current_bin_), bin_k = "/binK_" + std::to_string(current_bin_);
        if (host_count_ > 0) {
            for (auto& s : scalars_) {
                store_.put_scalar(bin_r + "/scalar/" + s.name, s.sum / host_count_);
                s.sum = 0.0;
            }
            for (auto& e : eqtime_) {
                if (e.sum.n_elem == 0) continue;
                for (arma::uword i = 0; i < e.sum.n_elem; ++i) e.sum(i) /= host_count_;
                emit_equal_time(bin_r, bin_k, e.name, transform::chi_site_to_chi_r(e.sum, lat));
                e.sum.zeros();
            }
            for (auto& u : uneqtime_) {
                if (u.sum.n_elem == 0) continue;
                for (arma::uword i = 0; i < u.sum.n_elem; ++i) u.sum.memptr()[i] /= host_count_;
                emit_cube(bin_r, bin_k, "/unequaltime/", u.name, transform::chi_site_to_chi_r(u.sum, lat));
                u.sum.zeros();
            }
        }
        if (dev_model_ && dev_count_ > 0) {
            const int n = lat.L1() * lat.L2();
            double sc[3];
            arma::cube chi_r(lat.L1(), lat.L2(), 1);
            long long cnt = 0;
            dqmcb::check(dqmcb_bin_fetch(dev_model_->ctx(), sc, chi_r.memptr(), &cnt, 1), dev_model_->ctx(), "bin_fetch");
            static const char* names[3] = {"density", "doubleOcc", "swave"};  // source/main.cpp:116-118
            for (int k = 0; k < 3; ++k) store_.put_scalar(bin_r + "/scalar/" + names[k], sc[k] / (double)cnt);
            for (int i = 0; i < n; ++i) chi_r.memptr()[i] /= (double)cnt;
            emit_equal_time(bin_r, bin_k, "densityCorr", chi_r);
            dev_count_ = 0;
            if (dev_tau_count_ > 0) {
                const int nt1 = dev_model_->nt() + 1;
                std::vector<double> sums((size_t)3 * nt1 * n);
                long long c2 = 0;
                dqmcb::check(dqmcb_bin_fetch_unequal(dev_model_->ctx(), sums.data(), &c2, 1), dev_model_->ctx(), "bin_fetch_unequal");
                static const char* tnames[3] = {"greenTau", "doublonTau", "currxxTau"};  // source/main.cpp:120-122
                for (int k = 0; k < 3; ++k) {
                    arma::cube cube(lat.L1(), lat.L2(), nt1);
                    for (size_t i = 0; i < (size_t)nt1 * n; ++i) cube.memptr()[i] = sums[(size_t)k * nt1 * n + i] / (double)c2;
                    emit_cube(bin_r, bin_k, "/unequaltime/", tnames[k], cube);
                }
                dev_tau_count_ = 0;
            }
        }
#ifdef DQMCB_USE_HDF5
        write_bin_hdf5(bin_r, bin_k);
#endif
        host_count_ = 0;
        ++current_bin_;
    }

    int current_bin() const { return current_bin_; }
    const dqmcb::BinStore& store() const { return store_; }
    std::string file_name() const { return "results/data_" + std::to_string(rank_) + ".h5"; }  // :290

   private:
    size_t hdf5_written_ = 0;
    void emit_equal_time(const std::string& bin_r, const std::string& bin_k, const std::string& name, const arma::cube& chi_r) {
        emit_cube(bin_r, bin_k, "/equaltime/", name, chi_r);
    }
    void emit_cube(const std::string& bin_r, const std::string& bin_k, const std::string& group, const std::string& name,
                   const arma::cube& chi_r) {
        store_.put_cube(bin_r + group + name, chi_r);
        // k-space needs the lattice only for L1, L2: rebuild the k-grid on the device from the cube's shape
        arma::cx_cube chi_k(chi_r.n_rows, chi_r.n_cols, chi_r.n_slices);
        dqmcb::check(dqmcb_chi_r_to_chi_k(0, (int)chi_r.n_rows, (int)chi_r.n_cols, (int)chi_r.n_slices, chi_r.memptr(),
                                          reinterpret_cast<double*>(chi_k.memptr())),
                     nullptr, "chi_r_to_chi_k");
        store_.put_complex_cube(bin_k + group + name, chi_k);
    }
#ifdef DQMCB_USE_HDF5
    // the reference's group tree (include/measurementh5.h:294-312): every bin has scalar / equaltime /
    // unequaltime under /bin_<n> and equaltime / unequaltime under /binK_<n>, datasets below them
    void write_bin_hdf5(const std::string& bin_r, const std::string& bin_k) {
        if (file_id_ < 0) {
            struct stat info;
            if (stat("results", &info) != 0) mkdir("results", 0755);
            file_id_ = H5Fcreate(file_name().c_str(), H5F_ACC_TRUNC, H5P_DEFAULT, H5P_DEFAULT);
            if (file_id_ < 0) throw std::runtime_error("Failed to create HDF5 file: " + file_name());
        }
        const char* groups[7] = {"", "/scalar", "/equaltime", "/unequaltime", "", "/equaltime", "/unequaltime"};
        for (int g = 0; g < 7; ++g) {
            const std::string p = (g < 4 ? bin_r : bin_k) + groups[g];
            H5Gclose(H5Gcreate2(file_id_, p.c_str(), H5P_DEFAULT, H5P_DEFAULT, H5P_DEFAULT));
        }
        const auto& sets = store_.datasets();
        for (; hdf5_written_ < sets.size(); ++hdf5_written_) {
            const auto& d = sets[hdf5_written_];
            std::vector<hsize_t> dims(d.dims.begin(), d.dims.end());
            const hid_t sp = H5Screate_simple((int)dims.size(), dims.data(), NULL);
            const hid_t ds = H5Dcreate2(file_id_, d.path.c_str(), H5T_NATIVE_DOUBLE, sp, H5P_DEFAULT, H5P_DEFAULT, H5P_DEFAULT);
            const herr_t st = H5Dwrite(ds, H5T_NATIVE_DOUBLE, H5S_ALL, H5S_ALL, H5P_DEFAULT, d.data.data());
            H5Dclose(ds);
            H5Sclose(sp);
            if (st < 0) throw std::runtime_error("Failed to write " + d.path + " to HDF5 file");
        }
    }
#endif
};
