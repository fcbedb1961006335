// oracle/ref_measure.cpp  --  TEST INFRASTRUCTURE, not product code.
//
// C-ABI harness around the reference's UNMODIFIED measurement pipeline
// (/root/reference/include/measurementh5.h: transform::chi_site_to_chi_r :20-66,
// transform::chi_r_to_chi_k :78-116, MeasurementManager::measure/accumulate/saveToHDF5 :189-361),
// compiled where it lies against oracle/shim/{armadillo,mpi.h,hdf5.h}.  The HDF5 shim records every
// dataset the reference would have written (path, dims, payload) instead of creating a file, so the
// parity tests see exactly the tree of results/data_<rank>.h5.
//
// Wiring follows source/main.cpp:114-119,165-170: three scalars + densityCorr, measure() after each
// sweep, accumulate() per bin.  Only tests/ and tests/golden/make_golden.py load this.
#include <armadillo>
#include <mpi.h>

#include <cstdio>
#include <fstream>
#include <memory>
#include <unistd.h>

#include <utility.h>
#include <lattice.h>
#include <stackngf.h>
#include <model.h>
#include <measurementh5.h>

namespace {

struct MeasSim {
    std::unique_ptr<utility::parameters> params;
    std::unique_ptr<Lattice> lat;
    std::unique_ptr<MeasurementManager> mgr;
    std::string cwd_before, workdir;
};

arma::mat mat_in2(int r, int c, const double* p) {
    arma::mat m(r, c);
    std::memcpy(m.memptr(), p, sizeof(double) * r * c);
    return m;
}

}  // namespace

extern "C" {

// MeasurementManager wired as in source/main.cpp:114-119 for an L1 x L2 square lattice
void* refm_create_ex(int L1, int L2, int rank, int unequal);
void* refm_create(int L1, int L2, int rank) { return refm_create_ex(L1, L2, rank, 0); }
// unequal != 0: isMeasureUnequalTime = true and greenTau / doublonTau / currxxTau registered (source/main.cpp:120-122)
void* refm_create_ex(int L1, int L2, int rank, int unequal) {
    char path[] = "/tmp/dqmc_refm_params_XXXXXX";
    int fd = mkstemp(path);
    if (fd < 0) return nullptr;
    {
        std::ofstream f(path);
        f << "[Lattice]\nL1 = " << L1 << "\nL2 = " << L2 << "\n[simulation]\nisMeasureUnequalTime = " << (unequal ? "true" : "false") << "\n";
    }
    close(fd);
    MeasSim* s = new MeasSim();
    try {
        s->params.reset(new utility::parameters(path));
        std::remove(path);
        std::array<double, 2> a1{{1.0, 0.0}}, a2{{0.0, 1.0}};
        std::vector<std::array<double, 2>> orbs{{{0.0, 0.0}}};
        s->lat.reset(new Lattice(*s->params, a1, a2, orbs));
        s->mgr.reset(new MeasurementManager(*s->params, MPI_COMM_WORLD, rank));
        s->mgr->addScalar("density", Observables::calculate_density);
        s->mgr->addScalar("doubleOcc", Observables::calculate_doubleOccupancy);
        s->mgr->addScalar("swave", Observables::calculate_swavePairing);
        s->mgr->addEqualTime("densityCorr", Observables::calculate_densityCorr);
        s->mgr->addUnequalTime("greenTau", Observables::calculate_greenTau);
        s->mgr->addUnequalTime("doublonTau", Observables::calculate_doublonTau);
        s->mgr->addUnequalTime("currxxTau", Observables::calculate_currxxTau);
        h5shim::store() = h5shim::State();
    } catch (const std::exception& e) {
        std::fprintf(stderr, "refm_create: %s\n", e.what());
        std::remove(path);
        delete s;
        return nullptr;
    }
    return s;
}
void refm_destroy(void* h) { delete (MeasSim*)h; }

// MeasurementManager::measure on greens[0].Gtt[0] = G (N x N column-major)
void refm_measure(void* h, const double* G) {
    MeasSim* s = (MeasSim*)h;
    const int n = s->lat->n_sites();
    std::vector<GF> greens(1);
    greens[0].Gtt.resize(1);
    greens[0].Gtt[0] = mat_in2(n, n, G);
    s->mgr->measure(greens, *s->lat);
}

// measure with the full unequal-time GF: Gtt, Gt0, G0t = [n_tau][N*N] column-major each
void refm_measure_unequal(void* h, int n_tau, const double* Gtt, const double* Gt0, const double* G0t) {
    MeasSim* s = (MeasSim*)h;
    const int n = s->lat->n_sites();
    std::vector<GF> greens(1);
    greens[0].Gtt.resize(n_tau);
    greens[0].Gt0.resize(n_tau);
    greens[0].G0t.resize(n_tau);
    for (int t = 0; t < n_tau; ++t) {
        greens[0].Gtt[t] = mat_in2(n, n, Gtt + (size_t)t * n * n);
        greens[0].Gt0[t] = mat_in2(n, n, Gt0 + (size_t)t * n * n);
        greens[0].G0t[t] = mat_in2(n, n, G0t + (size_t)t * n * n);
    }
    s->mgr->measure(greens, *s->lat);
}

// MeasurementManager::accumulate; saveToHDF5 creates ./results (measurementh5.h:278-287), so the call
// runs inside a scratch directory
void refm_accumulate(void* h) {
    MeasSim* s = (MeasSim*)h;
    char cwd[4096];
    if (!getcwd(cwd, sizeof(cwd))) cwd[0] = 0;
    char tmpl[] = "/tmp/dqmc_refm_XXXXXX";
    const char* d = mkdtemp(tmpl);
    if (d && chdir(d) != 0) d = nullptr;
    s->mgr->accumulate(*s->lat);
    if (d) {
        if (cwd[0] && chdir(cwd) != 0) std::fprintf(stderr, "refm_accumulate: cannot return to %s\n", cwd);
        std::string r = std::string(d) + "/results";
        rmdir(r.c_str());
        rmdir(d);
    }
}

// recorded datasets ------------------------------------------------------------------------------
int refm_n_datasets() { return (int)h5shim::store().sets.size(); }
int refm_dataset_info(int i, char* path, int path_cap, long long* dims /*[4]*/, int* rank) {
    const auto& d = h5shim::store().sets.at(i);
    std::snprintf(path, path_cap, "%s", d.path.c_str());
    *rank = (int)d.dims.size();
    for (int k = 0; k < 4; ++k) dims[k] = k < (int)d.dims.size() ? (long long)d.dims[k] : 1;
    return (int)d.data.size();
}
void refm_dataset_data(int i, double* out) {
    const auto& d = h5shim::store().sets.at(i);
    std::memcpy(out, d.data.data(), d.data.size() * sizeof(double));
}
int refm_file_name(char* out, int cap) {
    const auto& f = h5shim::store().files;
    std::snprintf(out, cap, "%s", f.empty() ? "" : f.back().c_str());
    return (int)f.size();
}

// stand-alone transforms (cube forms, n_tau slices) ----------------------------------------------------
// chi_site: [n_tau][N*N] column-major slices -> out in arma::cube memory order (L1, L2, n_tau) column-major
void refm_chi_site_to_chi_r(void* h, int n_tau, const double* chi_site, double* out) {
    MeasSim* s = (MeasSim*)h;
    const int n = s->lat->n_sites();
    arma::cube c(n, n, n_tau);
    for (int t = 0; t < n_tau; ++t) c.slice(t) = mat_in2(n, n, chi_site + (size_t)t * n * n);
    arma::cube r = transform::chi_site_to_chi_r(c, *s->lat);
    std::memcpy(out, r.memptr(), r.n_elem * sizeof(double));
}
// chi_r: (L1, L2, n_tau) column-major -> out interleaved (re, im) in arma::cx_cube memory order
void refm_chi_r_to_chi_k(void* h, int n_tau, const double* chi_r, double* out) {
    MeasSim* s = (MeasSim*)h;
    const int L1 = s->lat->L1(), L2 = s->lat->L2();
    arma::cube c(L1, L2, n_tau);
    for (int t = 0; t < n_tau; ++t) c.slice(t) = mat_in2(L1, L2, chi_r + (size_t)t * L1 * L2);
    arma::cx_cube k = transform::chi_r_to_chi_k(c, *s->lat);
    std::memcpy(out, k.memptr(), k.n_elem * 2 * sizeof(double));
}
int refm_pbc_shortest(int d, int L) { return transform::pbc_shortest(d, L); }

}  // extern "C"
