/*
 * dqmc_b200.h -- C ABI of the B200-native DQMC sweep hot path.
 *
 * This is the drop-in boundary for the reference's Armadillo/MKL path (kfkq/DQMC).  The
 * reference has no FFI of its own; the boundary is its C++ surface, so each
 * entry point below cites the reference function it replaces (paths relative to the reference
 * root).  The C++ host mirror in include/dqmcb/ (same class / function names as the reference's
 * headers) and the Python binding dqmc_b200/capi.py both sit on top of exactly these symbols.
 *
 * Conventions
 *   - every call returns 0 on success, a DQMCB_E* code otherwise; dqmcb_last_error() gives text;
 *   - all matrices are FP64, column-major, dense N x N (arma::mat layout);
 *   - one context = W independent walkers (Markov chains) of one model on one GPU; all batched
 *     calls act on every walker; host buffers are walker-major ([W][...]);
 *   - work is enqueued on the context's CUDA stream (dqmcb_set_stream); only the get_* / host
 *     buffer calls synchronise;
 *   - there is NO CPU fallback: without a CUDA device every compute call fails with DQMCB_ECUDA.
 */
#ifndef DQMC_B200_H
#define DQMC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DQMCB_OK 0
#define DQMCB_EINVAL 1   /* bad argument (std::invalid_argument / std::out_of_range on the C++ side) */
#define DQMCB_ECUDA 2    /* CUDA runtime failure or no device */
#define DQMCB_ENUMERIC 3 /* factorisation failed (std::runtime_error, source/stablelinalg.cpp:43-45) */
#define DQMCB_ESTATE 4   /* call order violated (e.g. sweep before init) */

typedef struct dqmcb_ctx dqmcb_ctx;

/* RNG modes for the Metropolis loop */
#define DQMCB_RNG_TAPE 0   /* host-supplied tape (parity runs: replay of the reference's mt19937 stream) */
#define DQMCB_RNG_PHILOX 1 /* Philox4x32-10 keyed on (seed, walker, visit counter, site) generated in-kernel */
#define DQMCB_RNG_MT19937 2 /* device-resident std::mt19937 per walker reproducing libstdc++'s
                               shuffle / uniform_int / bernoulli draws (source/update.cpp:10-25) */

typedef struct {
    int n_sites;   /* N = L1*L2 (model.ns(), include/model.h:49) */
    int nt;        /* number of time slices */
    int n_stab;    /* stabilisation interval (source/dqmc.cpp:9) */
    int n_walkers; /* W independent chains batched on this GPU */
    int device;    /* CUDA device ordinal */
    int delay_k;   /* 0/1: rank-1 Sherman-Morrison per accepted flip (reference behaviour);
                      k>1: delayed rank-k updates (algebraically identical) */
    double g;      /* sqrt(|U| dtau / 2), source/model.cpp:27 */
    double alpha;  /* -1 for the attractive model, source/model.cpp:28 */
    double eta[4];   /* include/field.h:40-43 */
    double gamma[4]; /* include/field.h:35-38 */
} dqmcb_config;

/* ---- lifetime ------------------------------------------------------------------------------- */
/* AttractiveHubbard ctor + DQMC ctor (source/model.cpp:3-36, source/dqmc.cpp:5-36).  expK / invexpK
 * are host N x N column-major exp(-/+ dtau K); K itself is built on the host (source/model.cpp:39-60).
 * For N = 64 the matrices are tested for the product form Ey (x) Ex of the 8 x 8 square lattice (site x + 8 y,
 * include/lattice.h:100-107); if both reproduce to 1e-14 the fast path applies them as 8 x 8 factors (a quarter of the
 * flops of the dense product, results equal to rounding order), otherwise -- or with env DQMCB_NO_KRON=1 -- densely. */
int dqmcb_create(const dqmcb_config* cfg, const double* expK, const double* invexpK, dqmcb_ctx** out);
void dqmcb_destroy(dqmcb_ctx* ctx);
const char* dqmcb_last_error(const dqmcb_ctx* ctx); /* ctx may be NULL: last create() error */
int dqmcb_set_stream(dqmcb_ctx* ctx, void* cuda_stream);
int dqmcb_synchronize(dqmcb_ctx* ctx);
int dqmcb_n_stack(const dqmcb_ctx* ctx); /* ceil(nt / n_stab), source/dqmc.cpp:10 */
/* kernels launched by this context so far (sweep, init and measurement kernels) */
long long dqmcb_launch_count(const dqmcb_ctx* ctx);
/* walker groups the sweeps are split into (independent chains on separate streams so that partial
 * waves of one group's kernels are filled by another's; env DQMCB_STREAMS, default 2) */
int dqmcb_n_groups(const dqmcb_ctx* ctx);
/* half-wrap matrices exp(-/+ dtau K / 2) for dqmcb_half_wrap (source/model.cpp:34-35) */
int dqmcb_set_half_expK(dqmcb_ctx* ctx, const double* expK_half, const double* invexpK_half);

/* ---- HS field (GHQField, include/field.h) ----------------------------------------------------- */
/* Native device layout: int8 [W][nt][N].  The *_i64 forms take the reference's arma::imat layout
 * (int64, nt x N column-major, include/field.h:52) for ONE walker. */
int dqmcb_set_fields(dqmcb_ctx* ctx, const int8_t* fields);
int dqmcb_get_fields(dqmcb_ctx* ctx, int8_t* fields);
int dqmcb_set_fields_i64(dqmcb_ctx* ctx, int walker, const long long* fields);
int dqmcb_get_fields_i64(dqmcb_ctx* ctx, int walker, long long* fields);

/* ---- RNG ---------------------------------------------------------------------------------------- */
int dqmcb_rng_mode(dqmcb_ctx* ctx, int mode, uint64_t seed);
/* mode MT19937: one std::mt19937 seed per walker (utility::random ctor, include/utility.h:27); selects the mode.
 * dqmcb_rng_mode(ctx, DQMCB_RNG_MT19937, seed) seeds walker w with seed + w.  The stream is consumed in call
 * order exactly as the reference consumes it (N/2 + 2N words per local_update for even N). */
int dqmcb_rng_seed_mt19937(dqmcb_ctx* ctx, const uint32_t* seeds);
/* mode TAPE: draws for the next `n_visits` local_update calls of every walker, in visiting order:
 * perm[W][n_visits][N] site order, knew[W][n_visits][N] proposal index 0..2, u[W][n_visits][N]. */
int dqmcb_set_tape(dqmcb_ctx* ctx, int n_visits, const uint8_t* perm, const uint8_t* knew, const double* u);
/* Pipelined variant: upload the NEXT tape (pinned host buffers, which must stay valid until the commit) on a copy
 * stream while the current step's kernels run; dqmcb_tape_commit switches to it in stream order. */
int dqmcb_set_tape_async(dqmcb_ctx* ctx, int n_visits, const uint8_t* perm, const uint8_t* knew, const double* u);
int dqmcb_tape_commit(dqmcb_ctx* ctx);

/* ---- stacks and Green's functions (source/dqmc.cpp:43-72) -------------------------------------- */
int dqmcb_init_stacks(dqmcb_ctx* ctx);         /* DQMC::init_stacks            */
int dqmcb_init_greenfunctions(dqmcb_ctx* ctx); /* DQMC::init_greenfunctions -> G(0), log_det_M */

/* ---- single steps of the sweep (teacher-forced parity; DQMC's private helpers) ----------------- */
int dqmcb_wrap_forward(dqmcb_ctx* ctx, int l);  /* propagate_GF_forward,  source/dqmc.cpp:113-132 */
int dqmcb_wrap_backward(dqmcb_ctx* ctx, int l); /* propagate_GF_backward, source/dqmc.cpp:169-187 */
int dqmcb_local_update(dqmcb_ctx* ctx, int l);  /* update::local_update,  source/update.cpp:5-32  */
int dqmcb_block_product(dqmcb_ctx* ctx, int i_stack); /* calculate_Bbar, source/dqmc.cpp:88-105 (into ctx Bbar) */
int dqmcb_get_block_product(dqmcb_ctx* ctx, double* Bbar /* [W][N*N] */);
/* update_stack_* + stabilize_GF_* + check_error at the end of block i_stack (source/dqmc.cpp:369-394,
 * 429-454), using the Bbar computed from the current fields. */
int dqmcb_stabilize_forward(dqmcb_ctx* ctx, int l);
int dqmcb_stabilize_backward(dqmcb_ctx* ctx, int l);

/* ---- whole sweeps (source/dqmc.cpp:337-396, 398-456) -------------------------------------------- */
int dqmcb_sweep_0_to_beta(dqmcb_ctx* ctx);
int dqmcb_sweep_beta_to_0(dqmcb_ctx* ctx);

/* ---- state access -------------------------------------------------------------------------------- */
int dqmcb_get_G(dqmcb_ctx* ctx, double* G /* [W][N*N] */); /* current equal-time G of every walker */
int dqmcb_set_G(dqmcb_ctx* ctx, const double* G);
int dqmcb_get_ldr(dqmcb_ctx* ctx, int i_stack, double* L, double* d, double* R); /* [W][N*N],[W][N],[W][N*N] */
int dqmcb_set_ldr(dqmcb_ctx* ctx, int i_stack, const double* L, const double* d, const double* R);
int dqmcb_get_logdet(dqmcb_ctx* ctx, double* logdet /* [W] */); /* GF::log_det_M */
/* acc_rate(), max_err(), mean_err() of include/dqmc.h:80-82, per walker ([W] each, any may be NULL) */
int dqmcb_get_stats(dqmcb_ctx* ctx, double* acc_rate, double* max_err, double* mean_err);
int dqmcb_reset_stats(dqmcb_ctx* ctx);
/* raw device pointers for zero-copy interop (NCCL field swaps, torch views) */
#define DQMCB_PTR_FIELDS 0
#define DQMCB_PTR_G 1
#define DQMCB_PTR_LOGDET 2
#define DQMCB_PTR_OBS 3
#define DQMCB_PTR_STACK_L 4 /* f64 [W][n_stack][N*N]: LDRStack factors L (include/stackngf.h:34-77), for device-side snapshots */
#define DQMCB_PTR_STACK_D 5 /* f64 [W][n_stack][N] */
#define DQMCB_PTR_STACK_R 6 /* f64 [W][n_stack][N*N] */
void* dqmcb_device_ptr(dqmcb_ctx* ctx, int which);

/* ---- measurements / action ------------------------------------------------------------------------ */
/* Observables::calculate_{density,doubleOccupancy,swavePairing,densityCorr} (source/model.cpp:167-288)
 * from the current G: scalars[W][3], corr[W][N*N] (may be NULL). */
int dqmcb_measure_equal_time(dqmcb_ctx* ctx, double* scalars, double* corr);
/* Kinetic energy per site and spin correlation from the current G.  NO REFERENCE COUNTERPART (source/model.cpp:167-288 has
 * neither; BASELINE's north_star names both): by Wick with G_up = G_down = G, ekin[W] = (2t/N) sum_<ij> (G_ij + G_ji) over the
 * +x / +y bonds of the periodic L1 x L2 lattice, and szz_r[W][L1*L2] (index dx + L1*dy) = (1/N) sum_i <Sz_i Sz_{i+r}> with
 * <Sz_i Sz_j> = (delta_ij - G_ji) G_ij / 2.  The structure factor is S(q) = sum_r cos(q.r) szz_r.  Parity: pinned on the
 * analytic U = 0 result and on a numpy restatement (tests/test_gpu_measure.py), not on reference outputs. */
int dqmcb_measure_kinetic_spin(dqmcb_ctx* ctx, int L1, int L2, double t, double* ekin, double* szz_r);
/* ---- measurement bins on the device (include/measurementh5.h) ------------------------------------------
 * MeasurementManager::measure (:189-229) / accumulate (:231-274) for the observables source/main.cpp:116-119
 * registers (density, doubleOcc, swave, densityCorr), for all W walkers at once and without leaving the
 * device: each measure adds W samples to the bin; the correlation is accumulated already folded onto
 * displacements (transform::chi_site_to_chi_r, :20-66 -- the fold is linear), so a bin is 3 + L1*L2 doubles.
 * Lattice as source/main.cpp:84-88 (a1=(1,0), a2=(0,1), one orbital, site = uy*L1 + ux). */
int dqmcb_bin_configure(dqmcb_ctx* ctx, int L1, int L2);  /* L1*L2 == n_sites; clears the bin */
int dqmcb_bin_measure(dqmcb_ctx* ctx);                    /* asynchronous on the context's stream */
/* SUMS over the samples since the last reset (divide by *n_samples; sums add across ranks):
 * scalar_sums[3] = density, doubleOcc, swave; chi_r_sum[L1*L2], element (bx,by) at bx + L1*by, the
 * memory order of the arma::cube the reference writes.  Any pointer may be NULL. */
int dqmcb_bin_fetch(dqmcb_ctx* ctx, double* scalar_sums, double* chi_r_sum, long long* n_samples, int reset);
/* transform::chi_site_to_chi_r (include/measurementh5.h:20-66) for n_mats column-major N x N matrices
 * (the slices of the reference's cube; host buffers): chi_r[n_mats][L1*L2]. */
int dqmcb_chi_site_to_chi_r(int device, int L1, int L2, int n_mats, const double* chi_site, double* chi_r);
/* transform::chi_r_to_chi_k (include/measurementh5.h:78-116), k-points of include/lattice.h:36-49:
 * chi_k[n_mats][L1*L2][2] (re, im), element (kx,ky) at kx + L1*ky.  For L1 != L2 the reference indexes
 * its output out of bounds; k-point (n, m) is stored at (n, m) here, identical for L1 == L2. */
int dqmcb_chi_r_to_chi_k(int device, int L1, int L2, int n_mats, const double* chi_r, double* chi_k);
/* transform::pbc_shortest (include/measurementh5.h:13-17); host arithmetic, needs no device */
int dqmcb_pbc_shortest(int d, int L);

/* ---- unequal-time sweep -----------------------------------------------------------
 * DQMC::sweep_unequalTime (source/dqmc.cpp:458-514: propagate_unequalTime_GF_forward :223-246,
 * propagate_Bt0_Bbt :248-262, stabilize_unequalTime :264-280) for every walker.  Call after
 * dqmcb_sweep_beta_to_0 (needs G(0,0) and the right-product stack); fills G(tau,tau), G(tau,0), G(0,tau)
 * for tau = 0..nt on the device, leaves the stack untouched and updates logdet at tau = beta. */
int dqmcb_sweep_unequal_time(dqmcb_ctx* ctx);
/* which: 0 Gtt[l], 1 Gt0[l], 2 G0t[l] (include/stackngf.h:15-22), l in 0..nt: out[W][N*N] column-major */
int dqmcb_get_G_unequal(dqmcb_ctx* ctx, int which, int l, double* out);
/* largest check_error of the sweep's stabilisations per walker (the reference warns above 1e-6, :505-508) */
int dqmcb_get_unequal_err(dqmcb_ctx* ctx, double* err_max);
/* DQMC::half_warp for the unequal-time arrays (source/dqmc.cpp:303-312): invexpKhalf X[which][l] expKhalf */
int dqmcb_half_wrap_unequal(dqmcb_ctx* ctx, int which, int l, double* Gsym);
/* Observables::calculate_{greenTau,doublonTau,currxxTau} (source/model.cpp:290-392), site-indexed:
 * each out[W][nt+1][N*N] column-major slices of the reference's cube; any may be NULL */
int dqmcb_measure_unequal_time(dqmcb_ctx* ctx, int L1, int L2, double* green, double* doublon, double* currxx);
/* the same three observables for all walkers, folded onto displacements and accumulated on the device
 * (MeasurementManager::measure for source/main.cpp:120-122); needs dqmcb_bin_configure.
 * fetch: chi_r_sums[3][nt+1][L1*L2] (greenTau, doublonTau, currxxTau; element (bx,by) at bx + L1*by), SUMS */
int dqmcb_bin_measure_unequal(dqmcb_ctx* ctx);
int dqmcb_bin_fetch_unequal(dqmcb_ctx* ctx, double* chi_r_sums, long long* n_samples, int reset);

/* AttractiveHubbard::global_action, source/model.cpp:140-159: S[W] */
int dqmcb_global_action(dqmcb_ctx* ctx, double* S);
/* DQMC::half_warp, equal-time entry (source/dqmc.cpp:288-315): Gsym[W][N*N] = invexpKhalf G expKhalf */
int dqmcb_half_wrap(dqmcb_ctx* ctx, double* Gsym);

/* ---- end-to-end convenience: host fields in, one full sweep, host results out ------------------------ */
/* set_fields + init_stacks + init_greenfunctions happen only if `reinit` != 0; then n_sweeps x
 * (sweep_0_to_beta; sweep_beta_to_0); then fields, scalars (and G if non-NULL) are copied back. */
int dqmcb_run_sweeps_host(dqmcb_ctx* ctx, int reinit, const int8_t* fields_in, int n_sweeps, int8_t* fields_out,
                          double* scalars_out, double* G_out);

/* ---- stand-alone batched stablelinalg (source/stablelinalg.cpp), host buffers, `batch` problems ---- */
int dqmcb_to_LDR(int device, int n, int batch, const double* M, double* L, double* d, double* R);
int dqmcb_mat_mul_ldr(int device, int n, int batch, const double* M, const double* L, const double* d, const double* R,
                      double* Lo, double* dO, double* Ro);
int dqmcb_ldr_mul_mat(int device, int n, int batch, const double* L, const double* d, const double* R, const double* M,
                      double* Lo, double* dO, double* Ro);
int dqmcb_ldr_mul_ldr(int device, int n, int batch, const double* L1, const double* d1, const double* R1, const double* L2,
                      const double* d2, const double* R2, double* Lo, double* dO, double* Ro);
int dqmcb_inv_I_plus_ldr(int device, int n, int batch, const double* L, const double* d, const double* R, double* G,
                         double* logdet);
int dqmcb_inv_I_plus_ldr_mul_ldr(int device, int n, int batch, const double* L1, const double* d1, const double* R1,
                                 const double* L2, const double* d2, const double* R2, double* G);

/* Allocation-free variant of dqmcb_to_LDR.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * dM, dL, dd, dR are DEVICE pointers owned by the caller (batch column-major n x n matrices / n-vectors, as dqmcb_to_LDR);
 * the kernel's scratch lives in a static device array of DQMCB_NOALLOC_SCRATCH_BYTES that every GPU loading the library
 * holds until exit.  The call allocates and frees nothing and does not synchronise: it is asynchronous on `stream`
 * (a cudaStream_t, NULL = default stream).  Limit: 256 + batch * 3 * n * (n + 1 - n % 2) * 8 bytes must fit in the array
 * for n > ~96 (smaller n keep their work tiles in shared memory: any batch); beyond it DQMCB_EINVAL with a message in
 * dqmcb_last_error(NULL).  All calls share the one array: calls that overlap on the GPU (different streams) race. */
#define DQMCB_NOALLOC_SCRATCH_BYTES (8u << 20)
int dqmcb_to_LDR_noalloc(int device, int n, int batch, const double* dM, double* dL, double* dd, double* dR, void* stream);

/* stablelinalg::inv_invldr_plus_ldr (source/stablelinalg.cpp:160-190; used by the unequal-time sweep) */
int dqmcb_inv_invldr_plus_ldr(int device, int n, int batch, const double* L1, const double* d1, const double* R1,
                              const double* L2, const double* d2, const double* R2, double* G);

/* ---- single-proposal model calls (teacher-forced parity of source/model.cpp:109-138) ------------------ */
/* AttractiveHubbard::local_update_ratio on the current G of every walker: R[W], delta[W] */
int dqmcb_local_update_ratio(dqmcb_ctx* ctx, int l, int site, int new_field, double* R, double* delta);
/* AttractiveHubbard::update_greens_local on one walker's current G */
int dqmcb_update_greens_local(dqmcb_ctx* ctx, int walker, double delta, int site);
/* GHQField::set_single_field (include/field.h:68-70) on the device copy */
int dqmcb_set_single_field(dqmcb_ctx* ctx, int walker, int l, int site, int value);

/* ---- per-kernel-class device timing (CUDA events on the context's stream), for bench.py's roofline ---- */
#define DQMCB_K_PROPAGATE 0 /* wrap + Metropolis kernels */
#define DQMCB_K_STABILIZE 1 /* Bbar + stack update + stabilised G kernels */
#define DQMCB_K_INIT 2      /* init_stacks / init_greenfunctions */
#define DQMCB_K_OTHER 3     /* measurement, action, half wrap */
#define DQMCB_K_CLASSES 4
int dqmcb_timing_enable(dqmcb_ctx* ctx, int on); /* turning it on resets the counters */
int dqmcb_timing_get(dqmcb_ctx* ctx, double* ms /* [4] */, long long* launches /* [4] */);

/* update::partner_rank, source/update.cpp:34-45 (pure host arithmetic) */
int dqmcb_partner_rank(int rank, int world_size, int exchange_attempt);

/* library build info, e.g. "dqmc_b200 sm_100a" */
const char* dqmcb_version(void);

#ifdef __cplusplus
}
#endif
#endif /* DQMC_B200_H */
