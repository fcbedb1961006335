// dqmcb_internal.cuh -- shared declarations of the sweep engine (device parameter block, context).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/dqmc_b200.h"

namespace dq {

constexpr int kThreads = 256;        // threads per CTA of the generic kernels
constexpr int kMaxSmem = 227 * 1024; // opt-in dynamic shared memory per CTA on sm_100a

// Everything a kernel needs; passed by value (lives in constant bank / param space).
// leading dimension of the delayed-update factors U (n x K) and V (K x n) in shared memory: n + 4 where n is a multiple of 8,
// so that the tensor-core fragments of the rank-K flush (lanes spread over 4 consecutive K-indices) hit distinct banks
__host__ __device__ inline int delay_ld(int n) { return (n % 8 == 0) ? n + 4 : n; }

struct Params {
    int N, LD, nt, n_stab, n_stack, W;
    int smem_mats_prop;  // how many N x LD matrices of the propagate kernel live in SMEM (0..2)
    int smem_mats_stab;  // ... of the stabilisation kernels (0 or 3)
    int smem_stage;      // 1: no resident tiles but one shared-memory staging tile for the factorisations
    int smem_rhs;        // 1: beside the staging tile, a 24-column right-hand-side panel for cta_lu_solve_panels (N = 144)
    int smem_arena;      // 1: neither fits (N = 256): shared memory holds the operand panels of the TMA-staged product (gemm_tma.cuh)
    int delay_k;         // > 1: delayed rank-k Metropolis updates in the generic propagate kernel
    int w0, w1;          // walker range [w0, w1) of this launch (walker groups on separate streams)
    const double* expK;
    const double* invexpK;
    const double* expK_half;
    const double* invexpK_half;
    // exp(-dtau K) = Ey (x) Ex on a square lattice with site index x + L1 y (K = Kx (+) Ky): 8 x 8 factors, column-major,
    // [0..63] = Ex, [64..127] = Ey, [128..255] the same for exp(+dtau K); null unless N = 64 and the host verified the
    // structure of both to 1e-14 (fast-path Bbar chain and wraps)
    const double* kron;
    // the same structure for ANY lattice L1 x L2 = N with L1, L2 <= 16 (generic engine: wraps and block products through
    // cta_kron_apply): four 16 x 16 column-major factor tables [Fx, Fy of exp(-dtau K); Fx, Fy of exp(+dtau K)], verified
    // by the host to reproduce both matrices to 1e-14 and to be symmetric; null otherwise (dense products)
    const double* kg;
    int kL1, kL2;
    int8_t* fields;   // [W][nt][N]
    double* G;        // [W][N*N]   current equal-time Green's function, column-major, ld = N
    double* Bbar;     // [W][N*N]
    double* stackL;   // [W][n_stack][N*N]
    double* stackD;   // [W][n_stack][N]
    double* stackR;   // [W][n_stack][N*N]
    double* ws;       // [W][3][N*LD]  global workspace when the tiles do not fit in SMEM
    double* logdet;   // [W]
    // RNG
    int rng_mode;
    uint64_t seed;
    unsigned long long visit_base;  // visits (local_update calls) consumed so far
    const uint8_t* tape_perm;       // [W][tape_visits][N]
    const uint8_t* tape_k;
    const double* tape_u;
    int tape_visits;
    // statistics
    unsigned long long* accepted;  // [W]
    double* err_max;               // [W]
    double* err_sum;               // [W]
    unsigned long long* err_cnt;   // [W]
    int* fail;                     // [1] first failing info code
    // lookup tables (flip ratio / B-matrix diagonals), computed on the host with the reference's formulas
    double lut_delta[16];  // [old*4+new]  exp(g d_eta) - 1       source/model.cpp:103-105
    double lut_gb[16];     // [old*4+new]  gammaR * bosonR       source/model.cpp:117,121
    double lut_ev[4];      // exp(+g eta)                         source/model.cpp:69
    double lut_iev[4];     // exp(-g eta)                         source/model.cpp:81
    double lut_action[4];  // alpha g eta(s) + log gamma(s)       source/model.cpp:153-154
    long long* prof;       // development aid (DQMCB_PHASE_PROF=1): clock64 stamps of CTA 0, else null
};

// phase stamp k of CTA 0 / thread 0 (no-op unless the context was created with DQMCB_PHASE_PROF=1)
#define DQ_PH(p, k)                                                                         \
    do {                                                                                    \
        asm volatile("" ::: "memory");                                                      \
        if ((p).prof && blockIdx.x == 0 && threadIdx.x == 0) (p).prof[k] = clock64();       \
        asm volatile("" ::: "memory");                                                      \
    } while (0)

}  // namespace dq

struct dqmcb_ctx {
    dqmcb_config cfg;
    dq::Params p;
    cudaStream_t stream = nullptr;
    std::string err;
    bool stacks_ready = false, green_ready = false;
    // owned device allocations
    double *d_expK = nullptr, *d_invexpK = nullptr, *d_expKh = nullptr, *d_invexpKh = nullptr, *d_kron = nullptr, *d_kg = nullptr;
    uint8_t *d_tape_perm = nullptr, *d_tape_k = nullptr;
    double* d_tape_u = nullptr;
    size_t tape_cap = 0;  // visits allocated
    // second tape set + copy stream: dqmcb_set_tape_async uploads the next tape while the sweeps of the current
    // one run; dqmcb_tape_commit makes it current
    uint8_t *d_tape2_perm = nullptr, *d_tape2_k = nullptr;
    double* d_tape2_u = nullptr;
    size_t tape2_cap = 0;
    int tape2_visits = 0;          // visits uploaded into the pending set (0 = nothing pending)
    cudaStream_t cstream = nullptr;
    cudaEvent_t ev_tape = nullptr, ev_free = nullptr;
    int tape_pos = 0;     // next visit to consume (TAPE mode)
    double* d_obs = nullptr;  // [W][4]
    double* d_scratchNN = nullptr;  // [W][N*N] for measurement outputs
    uint32_t* d_mt = nullptr;       // [W][625] mt19937 state + index (MT19937 mode)
    bool mt_attr_set = false;       // k_mt_fill's dynamic shared-memory limit raised on this context's device
    // measurement bin on the device (dqmcb_bin_*): per-walker vectors and the running sums
    int bin_L1 = 0, bin_L2 = 0;
    double* d_binpart = nullptr;    // [W][3 + N]
    double* d_binacc = nullptr;     // [3 + N]
    long long bin_samples = 0;
    // unequal-time sweep (dqmcb_sweep_unequal_time): Gtt/Gt0/G0t for every slice, running left product, error
    double *d_uneq = nullptr, *d_bt0 = nullptr, *d_uneq_err = nullptr;
    bool uneq_ready = false;
    double *d_tbinpart = nullptr, *d_tbinacc = nullptr;  // [W][3][nt+1][N], [3][nt+1][N]
    long long tbin_samples = 0;
    size_t smem_prop = 0, smem_stab = 0;
    // optional per-launch timing
    // walker groups: independent chains split over `n_groups` streams so that the partial last wave of
    // one group's kernel overlaps with the other groups' kernels (DQMCB_STREAMS, default 2)
    int n_groups = 1;
    cudaStream_t gstream[8] = {};
    cudaEvent_t gdone[8] = {};
    cudaEvent_t ev_fork = nullptr;
    long long n_launches = 0;  // kernels launched by this context so far (dqmcb_launch_count)
    int gthreads = dq::kThreads;  // threads per CTA of the generic sweep kernels (64 on small lattices)
    bool use_stab2 = false;      // N = 64 with Kronecker factors: second-generation stabilisation kernel (stab2.cuh)
    bool force_generic = false;  // DQMCB_FORCE_GENERIC=1: bypass the compile-time-N fast path (parity cross-checks)
    bool timing = false;
    struct Span { cudaEvent_t a, b; int cls; };
    std::vector<Span> spans;
    double t_ms[DQMCB_K_CLASSES] = {0, 0, 0, 0};
    long long t_n[DQMCB_K_CLASSES] = {0, 0, 0, 0};
};
