// tests/cpp/host_mirror_test.cpp -- a reference-style driver written against the C++ host mirror
// (include/dqmcb/*.h): the same object wiring as the reference's source/main.cpp:85-106,130-133
// (Lattice -> AttractiveHubbard -> DQMC -> init_stacks / init_greenfunctions -> sweeps), with an
// injected seed.  Results are dumped as "name count\nvalues...\n" records for tests/test_cpp_mirror.py,
// which compares them with the reference-generated golden fixtures.
//
//   host_mirror_test traj  L1 L2 U t mu beta nt n_stab seed n_sweeps out.txt
//   host_mirror_test exchange out.txt         (4-replica beta ladder of tests/golden/exchange.npz)
//   host_mirror_test api out.txt              (stablelinalg / containers / error behaviour)
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include <dqmcb/dqmc.h>
#include <dqmcb/field.h>
#include <dqmcb/lattice.h>
#include <dqmcb/model.h>
#include <dqmcb/stablelinalg.h>
#include <dqmcb/stackngf.h>
#include <dqmcb/update.h>
#include <dqmcb/utility.h>

static void dump(FILE* f, const std::string& name, const double* v, size_t n) {
    fprintf(f, "%s %zu\n", name.c_str(), n);
    for (size_t i = 0; i < n; ++i) fprintf(f, "%.17g ", v[i]);
    fprintf(f, "\n");
}
static void dump_mat(FILE* f, const std::string& name, const arma::mat& m) { dump(f, name, m.memptr(), m.n_elem); }
static void dump_fields(FILE* f, const std::string& name, const arma::imat& m) {
    std::vector<double> v(m.n_elem);
    for (size_t i = 0; i < m.n_elem; ++i) v[i] = (double)m(i);
    dump(f, name, v.data(), v.size());
}

static utility::parameters make_params(int L1, int L2, double U, double t, double mu, double beta, int nt, int n_stab) {
    utility::parameters p;
    p.set("Lattice", "L1", L1).set("Lattice", "L2", L2);
    p.set("hubbard", "U", U).set("hubbard", "t", t).set("hubbard", "mu", mu);
    p.set("simulation", "beta", beta).set("simulation", "nt", nt).set("simulation", "n_stab", n_stab);
    p.set("simulation", "isMeasureUnequalTime", "false");
    return p;
}

static int run_traj(int argc, char** argv) {
    if (argc < 13) return 2;
    const int L1 = atoi(argv[2]), L2 = atoi(argv[3]);
    const double U = atof(argv[4]), t = atof(argv[5]), mu = atof(argv[6]), beta = atof(argv[7]);
    const int nt = atoi(argv[8]), n_stab = atoi(argv[9]);
    const unsigned seed = (unsigned)atol(argv[10]);
    const int n_sweeps = atoi(argv[11]);
    FILE* f = fopen(argv[12], "w");
    if (!f) return 3;
    utility::parameters params = make_params(L1, L2, U, t, mu, beta, nt, n_stab);
    utility::random rng(seed);
    Lattice lat(params, {1.0, 0.0}, {0.0, 1.0}, {{0.0, 0.0}});
    AttractiveHubbard model(params, lat, rng, beta);
    DQMC sim(params, model);
    std::vector<LDRStack> stacks(model.n_flavor());
    std::vector<GF> greens(model.n_flavor());
    stacks[0] = sim.init_stacks(0);
    greens[0] = sim.init_greenfunctions(stacks[0]);
    dump_fields(f, "fields0", model.fields().fields());
    dump_mat(f, "expK", model.expK(0));
    dump_mat(f, "G0", greens[0].Gtt[0]);
    dump(f, "logdet0", &greens[0].log_det_M, 1);
    const double S0 = model.global_action(greens);
    dump(f, "action0", &S0, 1);
    {
        double obs[3] = {Observables::calculate_density(model, greens, lat), Observables::calculate_doubleOccupancy(model, greens, lat),
                         Observables::calculate_swavePairing(model, greens, lat)};
        dump(f, "obs0", obs, 3);
        dump_mat(f, "corr0", Observables::calculate_densityCorr(model, greens, lat));
        std::vector<GF> sym(1);
        sim.half_warp(sym, greens);
        dump_mat(f, "half_wrap0", sym[0].Gtt[0]);
        dump_mat(f, "stack0_prod0", stacks[0][0].to_matrix());
    }
    for (int s = 0; s < n_sweeps; ++s) {
        sim.sweep_0_to_beta(greens, stacks);
        dump_fields(f, "traj_fields_" + std::to_string(2 * s), model.fields().fields());
        dump_mat(f, "traj_G_" + std::to_string(2 * s), greens[0].Gtt[nt]);
        sim.sweep_beta_to_0(greens, stacks);
        sim.sweep_unequalTime(greens, stacks);
        dump_fields(f, "traj_fields_" + std::to_string(2 * s + 1), model.fields().fields());
        dump_mat(f, "traj_G_" + std::to_string(2 * s + 1), greens[0].Gtt[0]);
    }
    const double st[3] = {sim.acc_rate(), sim.max_err(), sim.mean_err()};
    dump(f, "stats", st, 3);
    const double S1 = model.global_action(greens);
    dump(f, "traj_action", &S1, 1);
    dump(f, "traj_logdet", &greens[0].log_det_M, 1);
    dump_mat(f, "traj_stack_prod0", stacks[0][0].to_matrix());
    // teacher-forced single-slice call through update::local_update on a wrapped copy is covered by the
    // Python C-ABI tests; here: one extra slice from the current state to exercise the entry point
    {
        greens[0].Gtt[1] = greens[0].Gtt[0];
        double acc = -1.0;
        update::local_update(model.rng(), model, greens, 0, acc);
        dump(f, "slice_acc", &acc, 1);
    }
    fclose(f);
    return 0;
}

static int run_exchange(int argc, char** argv) {
    if (argc < 3) return 2;
    const double betas[4] = {2.0, 1.6, 1.2, 0.8};
    const int L = 4, nt = 16, n_stab = 4, ws = 4;
    update::LocalTransport tp;
    update::transport() = &tp;
    std::vector<std::vector<double>> out_fields(2 * ws), out_logdet(2 * ws);
    std::vector<int> accepted(ws, 0);
    std::vector<std::thread> th;
    for (int r = 0; r < ws; ++r) {
        th.emplace_back([&, r] {
            utility::parameters params = make_params(L, L, 4.0, 1.0, 0.0, betas[r], nt, n_stab);
            utility::random rng(300 + r);
            Lattice lat(params, {1.0, 0.0}, {0.0, 1.0}, {{0.0, 0.0}});
            AttractiveHubbard model(params, lat, rng, betas[r]);
            DQMC sim(params, model);
            std::vector<LDRStack> stacks(1);
            std::vector<GF> greens(1);
            stacks[0] = sim.init_stacks(0);
            greens[0] = sim.init_greenfunctions(stacks[0]);
            sim.sweep_0_to_beta(greens, stacks);
            sim.sweep_beta_to_0(greens, stacks);
            int attempt = 0, acc = 0;
            for (int a = 0; a < 2; ++a) {
                update::replica_exchange(r, ws, rng, attempt, acc, model, sim, greens, stacks);
                const arma::imat& fm = model.fields().fields();
                std::vector<double> v(fm.n_elem);
                for (size_t i = 0; i < fm.n_elem; ++i) v[i] = (double)fm(i);
                out_fields[a * ws + r] = v;
                out_logdet[a * ws + r] = {greens[0].log_det_M};
            }
            accepted[r] = acc;
        });
    }
    for (auto& t : th) t.join();
    FILE* f = fopen(argv[2], "w");
    if (!f) return 3;
    for (int a = 0; a < 2; ++a)
        for (int r = 0; r < ws; ++r) {
            dump(f, "fields_after_" + std::to_string(a) + "_" + std::to_string(r), out_fields[a * ws + r].data(), out_fields[a * ws + r].size());
            dump(f, "logdet_after_" + std::to_string(a) + "_" + std::to_string(r), out_logdet[a * ws + r].data(), 1);
        }
    fclose(f);
    return 0;
}

static int run_api(int argc, char** argv) {
    if (argc < 3) return 2;
    FILE* f = fopen(argv[2], "w");
    if (!f) return 3;
    const int n = 20;
    arma::mat M(n, n), M2(n, n);
    std::mt19937 g(5);
    std::normal_distribution<double> nd;
    for (int j = 0; j < n; ++j)
        for (int i = 0; i < n; ++i) {
            M(i, j) = nd(g) * std::exp(0.4 * (i - j));
            M2(i, j) = nd(g);
        }
    dump_mat(f, "M", M);
    dump_mat(f, "M2", M2);
    stablelinalg::LDR F = stablelinalg::to_LDR(M);
    dump_mat(f, "to_LDR_prod", F.to_matrix());
    dump(f, "to_LDR_d", F.get_d().memptr(), n);
    stablelinalg::LDR F2 = stablelinalg::to_LDR(M2);
    dump_mat(f, "mat_mul_ldr", stablelinalg::mat_mul_ldr(M2, F).to_matrix());
    dump_mat(f, "ldr_mul_mat", stablelinalg::ldr_mul_mat(F, M2).to_matrix());
    dump_mat(f, "ldr_mul_ldr", stablelinalg::ldr_mul_ldr(F, F2).to_matrix());
    double ld = 0.0;
    dump_mat(f, "inv_I_plus_ldr", stablelinalg::inv_I_plus_ldr(F, ld));
    dump(f, "inv_I_plus_ldr_logdet", &ld, 1);
    dump_mat(f, "inv_I_plus_ldr_mul_ldr", stablelinalg::inv_I_plus_ldr_mul_ldr(F, F2));
    dump_mat(f, "inv_invldr_plus_ldr", stablelinalg::inv_invldr_plus_ldr(F, F2));
    // container / error behaviour of the reference (include/stackngf.h:61-74)
    double flags[4] = {0, 0, 0, 0};
    LDRStack st(3);
    try {
        st[3];
    } catch (const std::out_of_range&) {
        flags[0] = 1;
    }
    try {
        st.set(7, F);
    } catch (const std::out_of_range&) {
        flags[1] = 1;
    }
    try {
        utility::parameters p;
        p.getInt("simulation", "nt");
    } catch (const std::runtime_error&) {
        flags[2] = 1;
    }
    try {
        utility::parameters p;
        p.set("Lattice", "L1", 0).set("Lattice", "L2", 4);
        Lattice lat(p, {1.0, 0.0}, {0.0, 1.0}, {{0.0, 0.0}});
    } catch (const std::invalid_argument&) {
        flags[3] = 1;
    }
    dump(f, "exceptions", flags, 4);
    double pr[8];
    for (int r = 0; r < 8; ++r) pr[r] = update::partner_rank(r, 8, 1);
    dump(f, "partner_rank_ws8_a1", pr, 8);
    fclose(f);
    return 0;
}

int main(int argc, char** argv) {
    if (argc < 2) return 2;
    try {
        if (!strcmp(argv[1], "traj")) return run_traj(argc, argv);
        if (!strcmp(argv[1], "exchange")) return run_exchange(argc, argv);
        if (!strcmp(argv[1], "api")) return run_api(argc, argv);
    } catch (const std::exception& e) {
        fprintf(stderr, "host_mirror_test: %s\n", e.what());
        return 1;
    }
    return 2;
}
