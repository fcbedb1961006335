// oracle/ref_harness.cpp  --  TEST INFRASTRUCTURE, not product code.
//
// C-ABI harness around the *unmodified* reference translation units
// (/root/reference/source/{model,dqmc,update,stablelinalg}.cpp), which
// oracle/Makefile compiles where they lie against oracle/shim/{armadillo,mpi.h}
// into oracle/_ref/libdqmc_ref.so.  Replaces source/main.cpp (which needs HDF5):
// it injects the seed (main.cpp:37 uses wall-clock time), exposes every step of
// the sweep hot path so parity tests can be teacher-forced, and times the
// reference's own sweep loop for bench.py's cpu_baseline / --impl reference.
//
// Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may load it.
#include <armadillo>
#include <mpi.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <memory>
#include <random>
#include <thread>
#include <unistd.h>

// dqmc.h / model.h keep the propagation helpers and g_ private; the harness needs
// them for step-wise parity.  Access specifiers do not change object layout.
#define private public
#include <utility.h>
#include <lattice.h>
#include <stablelinalg.h>
#include <stackngf.h>
#include <field.h>
#include <model.h>
#include <dqmc.h>
#include <update.h>
#undef private

namespace {

struct Sim {
    std::unique_ptr<utility::parameters> params;
    std::unique_ptr<utility::random> rng;
    std::unique_ptr<Lattice> lat;
    std::unique_ptr<AttractiveHubbard> model;
    std::unique_ptr<DQMC> sim;
    std::vector<LDRStack> stacks;
    std::vector<GF> greens;
    int exchange_attempt = 0;
    int exchange_accepted = 0;
};

void copy_out(const arma::mat& m, double* out) { std::memcpy(out, m.memptr(), m.n_elem * sizeof(double)); }
arma::mat mat_in(int n, const double* p) {
    arma::mat m(n, n);
    std::memcpy(m.memptr(), p, sizeof(double) * n * n);
    return m;
}
arma::vec vec_in(int n, const double* p) {
    arma::vec v(n);
    std::memcpy(v.memptr(), p, sizeof(double) * n);
    return v;
}
stablelinalg::LDR ldr_in(int n, const double* L, const double* d, const double* R) {
    return stablelinalg::LDR(mat_in(n, L), vec_in(n, d), mat_in(n, R));
}
void ldr_out(const stablelinalg::LDR& F, double* L, double* d, double* R) {
    copy_out(F.get_L(), L);
    std::memcpy(d, F.get_d().memptr(), F.get_d().n_elem * sizeof(double));
    copy_out(F.get_R(), R);
}

}  // namespace

extern "C" {

void ref_set_blas_threads(int n) { scipy_openblas_set_num_threads(n); }

void* ref_create_ex(int L1, int L2, double U, double t, double mu, double beta, int nt, int n_stab, unsigned seed, int unequal);
void* ref_create(int L1, int L2, double U, double t, double mu, double beta, int nt, int n_stab, unsigned seed) {
    return ref_create_ex(L1, L2, U, t, mu, beta, nt, n_stab, seed, 0);
}
// unequal != 0: [simulation] isMeasureUnequalTime = true (sweep_unequalTime active, source/dqmc.cpp:458-514)
void* ref_create_ex(int L1, int L2, double U, double t, double mu, double beta, int nt, int n_stab, unsigned seed, int unequal) {
    scipy_openblas_set_num_threads(1);  // reference links mkl_sequential (CMakeLists.txt:27)
    char path[] = "/tmp/dqmc_ref_params_XXXXXX";
    int fd = mkstemp(path);
    if (fd < 0) return nullptr;
    {
        std::ofstream f(path);
        f.precision(17);
        f << "[Lattice]\nL1 = " << L1 << "\nL2 = " << L2 << "\n[hubbard]\nU = " << U << "\nt = " << t
          << "\nmu = " << mu << "\n[simulation]\nbeta = " << beta << "\nnt = " << nt
          << "\nn_therms = 0\nn_sweeps = 1\nn_bins = 1\nn_stab = " << n_stab
          << "\nsymmetric = false\nisMeasureUnequalTime = " << (unequal ? "true" : "false")
          << "\n[ParallelTempering]\nenabled = false\n";
    }
    close(fd);
    Sim* s = new Sim();
    try {
        s->params.reset(new utility::parameters(path));
        std::remove(path);
        s->rng.reset(new utility::random(seed));
        std::array<double, 2> a1{{1.0, 0.0}}, a2{{0.0, 1.0}};
        std::vector<std::array<double, 2>> orbs{{{0.0, 0.0}}};
        s->lat.reset(new Lattice(*s->params, a1, a2, orbs));
        s->model.reset(new AttractiveHubbard(*s->params, *s->lat, *s->rng, beta));
        s->sim.reset(new DQMC(*s->params, *s->model));
        const int nfl = s->model->n_flavor();
        s->stacks.resize(nfl);
        s->greens.resize(nfl);
        for (int flv = 0; flv < nfl; ++flv) {
            s->stacks[flv] = s->sim->init_stacks(flv);
            s->greens[flv] = s->sim->init_greenfunctions(s->stacks[flv]);
        }
    } catch (const std::exception& e) {
        std::fprintf(stderr, "ref_create: %s\n", e.what());
        std::remove(path);
        delete s;
        return nullptr;
    }
    return s;
}

void ref_destroy(void* h) { delete (Sim*)h; }

int ref_ns(void* h) { return ((Sim*)h)->model->ns(); }
int ref_nt(void* h) { return ((Sim*)h)->model->nt(); }
int ref_n_stack(void* h) { return ((Sim*)h)->sim->n_stack_; }
double ref_g(void* h) { return ((Sim*)h)->model->g_; }

// re-seed the generator (state as if freshly constructed with this seed)
void ref_reseed(void* h, unsigned seed) { ((Sim*)h)->rng->get_generator().seed(seed); }

void ref_get_fields(void* h, long long* out) {  // nt x N, column-major (include/field.h:52)
    const arma::imat& f = ((Sim*)h)->model->fields().fields();
    std::memcpy(out, f.memptr(), f.n_elem * sizeof(long long));
}
void ref_set_fields(void* h, const long long* in) {
    Sim* s = (Sim*)h;
    arma::imat f(s->model->nt(), s->model->ns());
    std::memcpy(f.memptr(), in, f.n_elem * sizeof(long long));
    s->model->fields().set_fields(f);
}
// rebuild stacks + G(0) from the current fields (source/dqmc.cpp:43-72)
void ref_init(void* h) {
    Sim* s = (Sim*)h;
    for (int flv = 0; flv < s->model->n_flavor(); ++flv) {
        s->stacks[flv] = s->sim->init_stacks(flv);
        s->greens[flv] = s->sim->init_greenfunctions(s->stacks[flv]);
    }
}
void ref_get_expK(void* h, double* out) { copy_out(((Sim*)h)->model->expK(0), out); }
void ref_get_invexpK(void* h, double* out) { copy_out(((Sim*)h)->model->invexpK(0), out); }
void ref_get_expK_half(void* h, double* out) { copy_out(((Sim*)h)->model->expK_half(0), out); }
void ref_get_invexpK_half(void* h, double* out) { copy_out(((Sim*)h)->model->invexpK_half(0), out); }

int ref_get_G(void* h, int l, double* out) {
    Sim* s = (Sim*)h;
    const arma::mat& g = s->greens[0].Gtt[l];
    if (g.n_elem == 0) return 1;
    copy_out(g, out);
    return 0;
}
void ref_set_G(void* h, int l, const double* in) {
    Sim* s = (Sim*)h;
    s->greens[0].Gtt[l] = mat_in(s->model->ns(), in);
}
double ref_logdet(void* h) { return ((Sim*)h)->greens[0].log_det_M; }
void ref_get_stack(void* h, int i, double* L, double* d, double* R) { ldr_out(((Sim*)h)->stacks[0][i], L, d, R); }
void ref_set_stack(void* h, int i, const double* L, const double* d, const double* R) {
    Sim* s = (Sim*)h;
    s->stacks[0][i] = ldr_in(s->model->ns(), L, d, R);
}

void ref_sweep_fwd(void* h) {
    Sim* s = (Sim*)h;
    s->sim->sweep_0_to_beta(s->greens, s->stacks);
}
void ref_sweep_bwd(void* h) {
    Sim* s = (Sim*)h;
    s->sim->sweep_beta_to_0(s->greens, s->stacks);
}
// DQMC::sweep_unequalTime (source/dqmc.cpp:458-514); which: 0 Gtt, 1 Gt0, 2 G0t
void ref_sweep_unequal(void* h) {
    Sim* s = (Sim*)h;
    s->sim->sweep_unequalTime(s->greens, s->stacks);
}
int ref_get_G_unequal(void* h, int which, int l, double* out) {
    Sim* s = (Sim*)h;
    const std::vector<arma::mat>& v = which == 0 ? s->greens[0].Gtt : (which == 1 ? s->greens[0].Gt0 : s->greens[0].G0t);
    if (l < 0 || l >= (int)v.size() || v[l].n_elem == 0) return 1;
    copy_out(v[l], out);
    return 0;
}
// DQMC::half_warp with unequal-time on (source/dqmc.cpp:288-315): out[3][nt][N*N] = sym Gtt, Gt0, G0t for t < nt
void ref_half_wrap_unequal(void* h, double* out) {
    Sim* s = (Sim*)h;
    std::vector<GF> sym = s->greens;
    s->sim->half_warp(sym, s->greens);
    const int nt = s->model->nt(), nn = s->model->ns() * s->model->ns();
    for (int t = 0; t < nt; ++t) {
        copy_out(sym[0].Gtt[t], out + ((size_t)0 * nt + t) * nn);
        copy_out(sym[0].Gt0[t], out + ((size_t)1 * nt + t) * nn);
        copy_out(sym[0].G0t[t], out + ((size_t)2 * nt + t) * nn);
    }
}
// Observables::calculate_{greenTau,doublonTau,currxxTau} (source/model.cpp:290-392): out[n_tau][N*N] each
void ref_observables_tau(void* h, double* green, double* doublon, double* currxx) {
    Sim* s = (Sim*)h;
    const size_t nn = (size_t)s->model->ns() * s->model->ns();
    const arma::cube g = Observables::calculate_greenTau(s->greens, *s->lat);
    const arma::cube d = Observables::calculate_doublonTau(s->greens, *s->lat);
    const arma::cube c = Observables::calculate_currxxTau(s->greens, *s->lat);
    for (arma::uword t = 0; t < g.n_slices; ++t) {
        copy_out(g.slice(t), green + t * nn);
        copy_out(d.slice(t), doublon + t * nn);
        copy_out(c.slice(t), currxx + t * nn);
    }
}

// utility::parameters (include/utility.h:50-276) on a file: kind 0 getString, 1 getInt, 2 getDouble, 3 getBool,
// 4 getDoubleVector (num[0] = count, values follow), 5 hasSection, 6 hasKey.  Returns 0, or 1 on std::runtime_error.
int ref_params_query(const char* path, const char* section, const char* key, int kind, double* num, char* str, int cap) {
    try {
        utility::parameters p(path);
        switch (kind) {
            case 0: std::snprintf(str, cap, "%s", p.getString(section, key).c_str()); break;
            case 1: num[0] = p.getInt(section, key); break;
            case 2: num[0] = p.getDouble(section, key); break;
            case 3: num[0] = p.getBool(section, key) ? 1.0 : 0.0; break;
            case 4: {
                const std::vector<double> v = p.getDoubleVector(section, key);
                num[0] = (double)v.size();
                for (size_t i = 0; i < v.size() && i < 30; ++i) num[1 + i] = v[i];
                break;
            }
            case 5: num[0] = p.hasSection(section) ? 1.0 : 0.0; break;
            default: num[0] = p.hasKey(section, key) ? 1.0 : 0.0; break;
        }
    } catch (const std::runtime_error&) {
        return 1;
    }
    return 0;
}

double ref_acc_rate(void* h) { return ((Sim*)h)->sim->acc_rate(); }
double ref_max_err(void* h) { return ((Sim*)h)->sim->max_err(); }
double ref_mean_err(void* h) { return ((Sim*)h)->sim->mean_err(); }

// --- single steps ------------------------------------------------------------------------
double ref_local_update(void* h, int l) {  // source/update.cpp:5-32 on Gtt[l+1]
    Sim* s = (Sim*)h;
    double acc = 0.0;
    update::local_update(s->model->rng(), *s->model, s->greens, l, acc);
    return acc;
}
void ref_local_update_ratio(void* h, int l, int i, int new_field, double* R, double* delta) {
    Sim* s = (Sim*)h;
    auto rd = s->model->local_update_ratio(s->greens, l, i, new_field);
    *R = rd.first;
    *delta = rd.second;
}
void ref_update_greens_local(void* h, double delta, int l, int i) {
    Sim* s = (Sim*)h;
    s->model->update_greens_local(s->greens, delta, l, i);
}
void ref_wrap_fwd(void* h, int l) { ((Sim*)h)->sim->propagate_GF_forward(((Sim*)h)->greens, l); }
void ref_wrap_bwd(void* h, int l) { ((Sim*)h)->sim->propagate_GF_backward(((Sim*)h)->greens, l); }
void ref_Bbar(void* h, int i_stack, double* out) { copy_out(((Sim*)h)->sim->calculate_Bbar(i_stack, 0), out); }
void ref_expV(void* h, int l, double* out) {
    arma::vec v = ((Sim*)h)->model->expV(l, 0);
    std::memcpy(out, v.memptr(), v.n_elem * sizeof(double));
}
void ref_update_stack_fwd(void* h, int i_stack, const double* Bbar) {
    Sim* s = (Sim*)h;
    arma::mat B = mat_in(s->model->ns(), Bbar);
    s->sim->update_stack_forward(s->stacks[0], B, i_stack);
}
void ref_update_stack_bwd(void* h, int i_stack, const double* Bbar) {
    Sim* s = (Sim*)h;
    arma::mat B = mat_in(s->model->ns(), Bbar);
    s->sim->update_stack_backward(s->stacks[0], B, i_stack);
}
void ref_stabilize_fwd(void* h, int l) { ((Sim*)h)->sim->stabilize_GF_forward(((Sim*)h)->greens[0], ((Sim*)h)->stacks[0], l); }
void ref_stabilize_bwd(void* h, int l) { ((Sim*)h)->sim->stabilize_GF_backward(((Sim*)h)->greens[0], ((Sim*)h)->stacks[0], l); }
void ref_half_wrap(void* h, double* out) {  // source/dqmc.cpp:288-315, equal-time part
    Sim* s = (Sim*)h;
    std::vector<GF> sym = s->greens;
    s->sim->half_warp(sym, s->greens);
    copy_out(sym[0].Gtt[0], out);
}

double ref_global_action(void* h) { return ((Sim*)h)->model->global_action(((Sim*)h)->greens); }

// out3 = {density, doubleOcc, swave}; corr = N x N column-major densityCorr (source/model.cpp:167-288)
void ref_observables(void* h, double* out3, double* corr) {
    Sim* s = (Sim*)h;
    out3[0] = Observables::calculate_density(s->greens, *s->lat);
    out3[1] = Observables::calculate_doubleOccupancy(s->greens, *s->lat);
    out3[2] = Observables::calculate_swavePairing(s->greens, *s->lat);
    if (corr) copy_out(Observables::calculate_densityCorr(s->greens, *s->lat), corr);
}

// --- RNG tape ------------------------------------------------------------------------------
// Replays, on a COPY of the generator, exactly the draws that `n_calls` consecutive
// update::local_update calls will make (source/update.cpp:10-25, include/field.h:76-83,
// include/utility.h:34-37): perm = shuffled site order, k = proposal index 0..2 (drawn from a
// by-value copy, so it does not advance the stream), u = the canonical double that
// bernoulli_distribution compares with p.  If advance != 0 the real generator is advanced too.
void ref_rng_tape(void* h, int n_calls, int advance, int* perm, int* k, double* u) {
    Sim* s = (Sim*)h;
    const int nv = s->model->ns();
    std::mt19937 g = s->rng->get_generator();
    std::vector<int> order(nv);
    for (int c = 0; c < n_calls; ++c) {
        for (int i = 0; i < nv; ++i) order[i] = i;
        std::shuffle(order.begin(), order.end(), g);
        for (int idx = 0; idx < nv; ++idx) {
            perm[c * nv + idx] = order[idx];
            std::mt19937 gcopy = g;
            std::uniform_int_distribution<int> dist(0, 2);
            k[c * nv + idx] = dist(gcopy);
            u[c * nv + idx] = std::generate_canonical<double, std::numeric_limits<double>::digits>(g);
        }
    }
    if (advance) s->rng->get_generator() = g;
}
// the canonical double the NEXT rng.bernoulli() call would compare with p (include/utility.h:34-37),
// without advancing the generator (replica-exchange decider, source/update.cpp:96)
double ref_rng_peek_canonical(void* h) {
    std::mt19937 g = ((Sim*)h)->rng->get_generator();
    return std::generate_canonical<double, std::numeric_limits<double>::digits>(g);
}
// raw 32-bit outputs of a fresh mt19937(seed) (for testing device-side generators)
void ref_mt19937_raw(unsigned seed, int n, unsigned* out) {
    std::mt19937 g(seed);
    for (int i = 0; i < n; ++i) out[i] = (unsigned)g();
}

// --- stablelinalg free functions on raw buffers (column-major N x N) -------------------------
int ref_to_LDR(int n, const double* M, double* L, double* d, double* R) {
    try {
        ldr_out(stablelinalg::to_LDR(mat_in(n, M)), L, d, R);
    } catch (...) {
        return 1;
    }
    return 0;
}
void ref_mat_mul_ldr(int n, const double* M, const double* L, const double* d, const double* R, double* Lo, double* dO, double* Ro) {
    ldr_out(stablelinalg::mat_mul_ldr(mat_in(n, M), ldr_in(n, L, d, R)), Lo, dO, Ro);
}
void ref_ldr_mul_mat(int n, const double* L, const double* d, const double* R, const double* M, double* Lo, double* dO, double* Ro) {
    ldr_out(stablelinalg::ldr_mul_mat(ldr_in(n, L, d, R), mat_in(n, M)), Lo, dO, Ro);
}
void ref_ldr_mul_ldr(int n, const double* L1, const double* d1, const double* R1, const double* L2, const double* d2, const double* R2,
                     double* Lo, double* dO, double* Ro) {
    ldr_out(stablelinalg::ldr_mul_ldr(ldr_in(n, L1, d1, R1), ldr_in(n, L2, d2, R2)), Lo, dO, Ro);
}
double ref_inv_I_plus_ldr(int n, const double* L, const double* d, const double* R, double* G) {
    double logdet = 0.0;
    copy_out(stablelinalg::inv_I_plus_ldr(ldr_in(n, L, d, R), logdet), G);
    return logdet;
}
void ref_inv_I_plus_ldr_mul_ldr(int n, const double* L1, const double* d1, const double* R1, const double* L2, const double* d2,
                                const double* R2, double* G) {
    copy_out(stablelinalg::inv_I_plus_ldr_mul_ldr(ldr_in(n, L1, d1, R1), ldr_in(n, L2, d2, R2)), G);
}
void ref_inv_invldr_plus_ldr(int n, const double* L1, const double* d1, const double* R1, const double* L2, const double* d2,
                             const double* R2, double* G) {
    copy_out(stablelinalg::inv_invldr_plus_ldr(ldr_in(n, L1, d1, R1), ldr_in(n, L2, d2, R2)), G);
}

int ref_partner_rank(int rank, int world_size, int attempt) { return update::partner_rank(rank, world_size, attempt); }

// --- replica exchange between in-process "ranks" (threads) -----------------------------------
// Runs the reference's update::replica_exchange (source/update.cpp:47-117) once on every
// handle, rank r = index r.  accepted_out[r] = the `accept` outcome as visible through the
// rank-0 counter only (reference semantics), so callers should compare fields instead.
void ref_replica_exchange(void** handles, int world_size) {
    std::vector<std::thread> th;
    for (int r = 0; r < world_size; ++r) {
        th.emplace_back([=] {
            shim_mpi_set_rank(r, world_size);
            Sim* s = (Sim*)handles[r];
            update::replica_exchange(r, world_size, *s->rng, s->exchange_attempt, s->exchange_accepted, *s->model, *s->sim,
                                     s->greens, s->stacks);
        });
    }
    for (auto& t : th) t.join();
}
int ref_exchange_attempt(void* h) { return ((Sim*)h)->exchange_attempt; }
int ref_exchange_accepted(void* h) { return ((Sim*)h)->exchange_accepted; }

// --- timing: the reference's own sweep loop (source/main.cpp:130-133) on P threads ------------
// Every handle is an independent chain (stand-in for `mpirun -np P`: non-PT ranks never
// communicate during sweeps).  Returns wall seconds (steady_clock, as main.cpp:129-135).
double ref_run_sweeps(void** handles, int n_handles, int n_sweeps) {
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> th;
    for (int r = 0; r < n_handles; ++r) {
        th.emplace_back([=] {
            Sim* s = (Sim*)handles[r];
            for (int i = 0; i < n_sweeps; ++i) {
                s->sim->sweep_0_to_beta(s->greens, s->stacks);
                s->sim->sweep_beta_to_0(s->greens, s->stacks);
            }
        });
    }
    for (auto& t : th) t.join();
    const auto t1 = std::chrono::steady_clock::now();
    return std::chrono::duration<double>(t1 - t0).count();
}

}  // extern "C"
