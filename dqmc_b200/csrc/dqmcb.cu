// dqmcb.cu -- kernels and C ABI (include/dqmc_b200.h) of the B200-native DQMC sweep hot path.
//
// One CTA per walker (Markov chain); W walkers batched per launch.  See DESIGN.md for the kernel
// inventory and rooflines.  Compiled for sm_100a only; there is no CPU fallback in this library.
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <string.h>
#include <stdlib.h>

#include <algorithm>
#include <string>
#include <vector>

#include "dqmcb_internal.cuh"
#include "engine.cuh"
#include "fast64.cuh"
#include "measure.cuh"
#include "mt19937.cuh"

using namespace dq;

// =====================================================================================================
// kernels (generic N)
// =====================================================================================================

// Propagate + Metropolis over `n_slices` consecutive slices starting at l_first, in sweep direction
// dir (+1: wrap then flip, source/dqmc.cpp:357-366; -1: flip then wrap, source/dqmc.cpp:416-426).
// flags: bit0 = do wraps, bit1 = do flips.
// BIG: instantiation for N >= 128 (large-tile tensor-core products, one CTA per SM); the other one keeps its register count
// for the small lattices, where two CTAs share an SM.
template <bool BIG>
__global__ void __launch_bounds__(kThreads, BIG ? 1 : 2) k_propagate(Params p, int l_first, int n_slices, int dir, unsigned long long visit0, int flags) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int w = blockIdx.x, n = p.N;
    // small arrays
    double* s_ev = reinterpret_cast<double*>(smem);
    double* s_iev = s_ev + n;
    double* s_u = s_iev + n;
    double* s_ucol = s_u + n;
    double* s_vrow = s_ucol + n;
    int* s_perm = reinterpret_cast<int*>(s_vrow + n);
    int* s_k = s_perm + n;
    uint32_t* s_key = reinterpret_cast<uint32_t*>(s_k + n);
    int8_t* s_f = reinterpret_cast<int8_t*>(s_key + n);
    size_t off = (size_t)(5 * n) * 8 + (size_t)(3 * n) * 4 + (size_t)n;
    off = (off + 15) & ~(size_t)15;
    double* tiles = reinterpret_cast<double*>(smem + off);
    double* gws = p.ws + (size_t)w * 3 * n * p.LD;
    double* g = (p.smem_mats_prop >= 1) ? tiles : gws;
    double* t = (p.smem_mats_prop >= 2) ? tiles + (size_t)n * p.LD : gws + (size_t)n * p.LD;
    // delayed-update factors U, V (K x n each) + running diagonal (2 copies), behind the resident tiles
    double* dU = tiles + (size_t)p.smem_mats_prop * n * p.LD;
    double* dV = dU + (size_t)p.delay_k * delay_ld(n);
    double* s_diag = dV + (size_t)p.delay_k * delay_ld(n);
    const int ld = p.LD;
    double* Gg = p.G + (size_t)w * n * n;
    cta_copy(g, ld, Gg, n, n);
    __syncthreads();
    int accepted = 0;
    for (int s = 0; s < n_slices; ++s) {
        const int l = l_first + dir * s;
        int8_t* fl = p.fields + ((size_t)w * p.nt + l) * n;
        for (int i = threadIdx.x; i < n; i += blockDim.x) s_f[i] = fl[i];
        __syncthreads();
        if (dir > 0 && (flags & 1)) {
            for (int i = threadIdx.x; i < n; i += blockDim.x) {
                s_ev[i] = p.lut_ev[s_f[i]];
                s_iev[i] = p.lut_iev[s_f[i]];
            }
            __syncthreads();
            dev_wrap_forward<BIG>(p, n, g, ld, t, ld, s_ev, s_iev);
        }
        if (flags & 2) {
            load_visit_draws(p, w, visit0 + s, s_perm, s_k, s_u, s_key);
            if (p.delay_k > 1)
                accepted += dev_local_update_delayed<BIG>(p, n, g, ld, s_f, s_perm, s_k, s_u, s_diag, dU, dV, p.delay_k);
            else
                accepted += dev_local_update(p, n, g, ld, s_f, s_perm, s_k, s_u, s_ucol, s_vrow);
            __syncthreads();
            for (int i = threadIdx.x; i < n; i += blockDim.x) fl[i] = s_f[i];
        }
        if (dir < 0 && (flags & 1)) {
            for (int i = threadIdx.x; i < n; i += blockDim.x) {
                s_ev[i] = p.lut_ev[s_f[i]];
                s_iev[i] = p.lut_iev[s_f[i]];
            }
            __syncthreads();
            dev_wrap_backward<BIG>(p, n, g, ld, t, ld, s_ev, s_iev);
        }
    }
    __syncthreads();
    cta_copy(Gg, n, g, ld, n);
    if (threadIdx.x == 0) p.accepted[w] += (unsigned long long)accepted;
}

// calculate_Bbar into the global Bbar buffer (stand-alone step for parity tests)
__global__ void __launch_bounds__(kThreads) k_bbar(Params p, int i_stack) {
    extern __shared__ __align__(16) unsigned char smem[];
    Work wk;
    carve(wk, p, blockIdx.x, p.smem_mats_stab, smem);
    dev_bbar(p, wk, blockIdx.x, i_stack);
    __syncthreads();
    cta_copy(p.Bbar + (size_t)blockIdx.x * p.N * p.N, p.N, wk.b[0], wk.ld[0], p.N);
}

__device__ inline double* stkL(const Params& p, int w, int i) { return p.stackL + ((size_t)w * p.n_stack + i) * p.N * p.N; }
__device__ inline double* stkR(const Params& p, int w, int i) { return p.stackR + ((size_t)w * p.n_stack + i) * p.N * p.N; }
__device__ inline double* stkD(const Params& p, int w, int i) { return p.stackD + ((size_t)w * p.n_stack + i) * p.N; }

__device__ inline void store_d(double* dst, const double* src, int n) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) dst[i] = src[i];
}
#include "unequal.cuh"  // needs the stack accessors above

// End-of-block work of the sweeps (source/dqmc.cpp:369-394 forward, 429-454 backward):
// Bbar from the current fields, stack update, stabilised G, check_error against the wrapped G.
// `part`: 0 = everything in one launch.  N = 256 runs it as three launches (1: block product and the product with the
// neighbouring stack factor, 2: the pivoted QR, 3: the rest) because the two halves want opposite shared-memory carve-outs:
// the TMA-staged products and solves need ~200 KB of operand panels, while the QR -- whose trailing matrix fits in L1
// for its last ~150 steps -- loses 30% when that shared memory leaves it no L1.  All state that crosses a part boundary
// lives in the walker's global work tiles (b0..b2) and in the stack entry itself.
__global__ void __launch_bounds__(kThreads) k_stabilize(Params p, int l, int dir, int part) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int w = blockIdx.x, n = p.N;
    Work wk;
    carve(wk, p, w, p.smem_mats_stab, smem);
    const int i = l / p.n_stab;
    const bool do1 = part == 0 || part == 1, do2 = part == 0 || part == 2, do3 = part == 0 || part == 3;
    double *b0 = wk.b[0], *b1 = wk.b[1], *b2 = wk.b[2];
    const int l0 = wk.ld[0], l1 = wk.ld[1], l2 = wk.ld[2];
    const bool first = (dir > 0) ? (i == 0) : (i == p.n_stack - 1);  // the stack entry is to_LDR(Bbar) itself
    if (do1) {
        DQ_PH(p, 0);
        dev_bbar(p, wk, w, i);  // -> b[0]
        DQ_PH(p, 1);
        __syncthreads();
        if (!first) {
            if (dir > 0) {
                // update_stack_forward (source/dqmc.cpp:134-146): mat_mul_ldr (source/stablelinalg.cpp:69-79): (Bbar L) D -> QR
                store_d(wk.sv[0], stkD(p, w, i - 1), n);
                __syncthreads();
                dev_gemm<false>(wk, b1, l1, b0, 1, l0, stkL(p, w, i - 1), 1, n, n, nullptr, wk.sv[0]);
            } else {
                // update_stack_backward (source/dqmc.cpp:189-201): ldr_mul_mat (source/stablelinalg.cpp:57-67): D (R Bbar) -> QR
                store_d(wk.sv[0], stkD(p, w, i + 1), n);
                __syncthreads();
                dev_gemm<false>(wk, b1, l1, stkR(p, w, i + 1), 1, n, b0, 1, l0, n, wk.sv[0], nullptr);
            }
            __syncthreads();
        }
        DQ_PH(p, 2);
    }
    if (do2) {
        if (first)
            dev_to_ldr(wk, n, b0, l0, b1, l1, b2, l2, wk.dnew);
        else
            dev_to_ldr(wk, n, b1, l1, b0, l0, b2, l2, wk.dnew);
        store_d(stkD(p, w, i), wk.dnew, n);
        __syncthreads();
        DQ_PH(p, 3);
    }
    if (!do3) return;
    if (first) {
        cta_copy(stkL(p, w, i), n, b1, l1, n);
        cta_copy(stkR(p, w, i), n, b2, l2, n);
    } else if (dir > 0) {
        // R <- R_new R_old
        dev_gemm<false>(wk, stkR(p, w, i), n, b2, 1, l2, stkR(p, w, i - 1), 1, n, n, nullptr, nullptr);
        cta_copy(stkL(p, w, i), n, b0, l0, n);
    } else {
        // L <- L_old Q
        dev_gemm<false>(wk, stkL(p, w, i), n, stkL(p, w, i + 1), 1, n, b0, 1, l0, n, nullptr, nullptr);
        cta_copy(stkR(p, w, i), n, b2, l2, n);
    }
    __syncthreads();  // stack entry i visible to the whole CTA (global writes + barrier)
    DQ_PH(p, 4);
    // stabilize_GF_* (source/dqmc.cpp:148-161, 203-215)
    const double* Gnew;
    int ldn;
    const bool at_end = (dir > 0) ? (l == p.nt - 1) : (l == 0);
    if (at_end) {
        const double ld_ = dev_inv_I_plus_ldr(p, wk, n, stkL(p, w, i), n, stkD(p, w, i), stkR(p, w, i), n);
        if (threadIdx.x == 0) p.logdet[w] = ld_;
        Gnew = b1;
        ldn = l1;
    } else if (dir > 0) {
        dev_inv_I_plus_ldr_mul_ldr(p, wk, n, stkL(p, w, i), stkD(p, w, i), stkR(p, w, i), stkL(p, w, i + 1), stkD(p, w, i + 1),
                                   stkR(p, w, i + 1), n);
        Gnew = b2;
        ldn = l2;
    } else {
        dev_inv_I_plus_ldr_mul_ldr(p, wk, n, stkL(p, w, i - 1), stkD(p, w, i - 1), stkR(p, w, i - 1), stkL(p, w, i), stkD(p, w, i),
                                   stkR(p, w, i), n);
        Gnew = b2;
        ldn = l2;
    }
    // check_error (source/dqmc.cpp:317-329) + install the stabilised G
    double* Gg = p.G + (size_t)w * n * n;
    double emax = 0.0;
    for (int e0 = threadIdx.x; e0 < n * n; e0 += 4 * blockDim.x) {  // 8 loads in flight per thread
        double v[4], g[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int e = e0 + u * blockDim.x;
            const bool ok = e < n * n;
            v[u] = ok ? Gnew[e % n + (size_t)(e / n) * ldn] : 0.0;
            g[u] = ok ? Gg[e] : 0.0;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int e = e0 + u * blockDim.x;
            if (e < n * n) {
                emax = fmax(emax, fabs(g[u] - v[u]));
                Gg[e] = v[u];
            }
        }
    }
    emax = cta_reduce_max(emax, wk.red);
    DQ_PH(p, 13);
    if (threadIdx.x == 0) {
        p.err_max[w] = fmax(p.err_max[w], emax);
        p.err_sum[w] += emax;
        p.err_cnt[w] += 1ull;
    }
}

// DQMC::init_stacks (source/dqmc.cpp:43-59): right-hand partial products, top block first
__global__ void __launch_bounds__(kThreads) k_init_stacks(Params p) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int w = blockIdx.x, n = p.N;
    Work wk;
    carve(wk, p, w, p.smem_mats_stab, smem);
    double *b0 = wk.b[0], *b1 = wk.b[1], *b2 = wk.b[2];
    const int l0 = wk.ld[0], l1 = wk.ld[1], l2 = wk.ld[2];
    for (int i = p.n_stack - 1; i >= 0; --i) {
        dev_bbar(p, wk, w, i);
        __syncthreads();
        // Bbar_ldr = to_LDR(Bbar): Q -> b1, Rn -> b2, d -> sv[2]
        dev_to_ldr(wk, n, b0, l0, b1, l1, b2, l2, wk.sv[2]);
        if (i == p.n_stack - 1) {
            cta_copy(stkL(p, w, i), n, b1, l1, n);
            cta_copy(stkR(p, w, i), n, b2, l2, n);
            store_d(stkD(p, w, i), wk.sv[2], n);
        } else {
            // ldr_mul_ldr(stack[i+1], Bbar_ldr) (source/stablelinalg.cpp:81-92)
            store_d(wk.sv[0], stkD(p, w, i + 1), n);
            __syncthreads();
            // M' = D1 (R1 L2) D2 -> b0
            dev_gemm<false>(wk, b0, l0, stkR(p, w, i + 1), 1, n, b1, 1, l1, n, wk.sv[0], wk.sv[2]);
            // keep R2 (b2) : factor b0 -> Q in b1 (L2 no longer needed), Rn in global R slot as temp
            double* Rtmp = stkR(p, w, i);
            dev_to_ldr(wk, n, b0, l0, b1, l1, Rtmp, n, wk.dnew);
            // L = L1 Q
            dev_gemm<false>(wk, stkL(p, w, i), n, stkL(p, w, i + 1), 1, n, b1, 1, l1, n, nullptr, nullptr);
            // R = R_new R2  (Rtmp -> b0 first, then product into the global slot)
            cta_copy(b0, l0, Rtmp, n, n);
            __syncthreads();
            dev_gemm<false>(wk, Rtmp, n, b0, 1, l0, b2, 1, l2, n, nullptr, nullptr);
            store_d(stkD(p, w, i), wk.dnew, n);
        }
        __syncthreads();
    }
}

// DQMC::init_greenfunctions (source/dqmc.cpp:61-72): G(0) = [I + stack[0]]^-1, log_det_M
__global__ void __launch_bounds__(kThreads) k_init_green(Params p) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int w = blockIdx.x, n = p.N;
    Work wk;
    carve(wk, p, w, p.smem_mats_stab, smem);
    const double ld_ = dev_inv_I_plus_ldr(p, wk, n, stkL(p, w, 0), n, stkD(p, w, 0), stkR(p, w, 0), n);
    if (threadIdx.x == 0) p.logdet[w] = ld_;
    cta_copy(p.G + (size_t)w * n * n, n, wk.b[1], wk.ld[1], n);
}

// Observables::calculate_{density,doubleOccupancy,swavePairing,densityCorr} (source/model.cpp:167-288)
__global__ void __launch_bounds__(kThreads) k_measure(Params p, double* obs /*[W][4]*/, double* corr /*[W][N*N] or null*/) {
    __shared__ double red[64];
    const int w = blockIdx.x, n = p.N;
    const double* G = p.G + (size_t)w * n * n;
    double dsum = 0.0, d2sum = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const double gc = 1.0 - G[i + (size_t)i * n];
        dsum += gc + gc;
        d2sum += gc * gc;
    }
    dsum = cta_reduce_sum(dsum, red);
    d2sum = cta_reduce_sum(d2sum, red);
    double ssum = 0.0;
    for (int e = threadIdx.x; e < n * n; e += blockDim.x) {
        const int j = e % n, i = e / n;
        const double gc = ((i == j) ? 1.0 : 0.0) - G[j + (size_t)i * n];
        ssum += gc * gc;
    }
    ssum = cta_reduce_sum(ssum, red);
    const double navg = dsum / n;  // n_avg of calculate_densityCorr equals the density
    if (threadIdx.x == 0) {
        obs[w * 4 + 0] = dsum / n;
        obs[w * 4 + 1] = d2sum / n;
        obs[w * 4 + 2] = ssum / n;
        obs[w * 4 + 3] = 0.0;
    }
    if (corr) {
        double* C = corr + (size_t)w * n * n;
        for (int e = threadIdx.x; e < n * n; e += blockDim.x) {
            const int i = e % n, j = e / n;
            const double ni = 2.0 * (1.0 - G[i + (size_t)i * n]);
            const double nj = 2.0 * (1.0 - G[j + (size_t)j * n]);
            const double ex = 2.0 * (1.0 - G[j + (size_t)i * n]) * G[i + (size_t)j * n];  // Q5: 1.0, not delta_ij
            C[i + (size_t)j * n] = ni * nj + ex - navg * navg;
        }
    }
}

// AttractiveHubbard::global_action (source/model.cpp:140-159)
__global__ void __launch_bounds__(kThreads) k_action(Params p, double* S) {
    __shared__ double red[64];
    const int w = blockIdx.x;
    const int8_t* f = p.fields + (size_t)w * p.nt * p.N;
    double part = 0.0;
    for (int e = threadIdx.x; e < p.nt * p.N; e += blockDim.x) part += p.lut_action[f[e]];
    part = cta_reduce_sum(part, red);
    if (threadIdx.x == 0) S[w] = -2.0 * p.logdet[w] - part;
}

// DQMC::half_warp, equal-time entry (source/dqmc.cpp:288-315)
__global__ void __launch_bounds__(kThreads) k_half_wrap(Params p, double* out) {
    const int w = blockIdx.x, n = p.N;
    double* gws = p.ws + (size_t)w * 3 * n * p.LD;
    cta_gemm<false>(gws, n, p.invexpK_half, 1, n, p.G + (size_t)w * n * n, 1, n, n, nullptr, nullptr);
    __syncthreads();
    cta_gemm<false>(out + (size_t)w * n * n, n, gws, 1, n, p.expK_half, 1, n, n, nullptr, nullptr);
}

// stand-alone batched stablelinalg ops on global buffers (ld = n).  op: 0 to_LDR, 1 mat_mul_ldr,
// 2 ldr_mul_mat, 3 ldr_mul_ldr, 4 inv_I_plus_ldr, 5 inv_I_plus_ldr_mul_ldr
struct LdrArgs {
    const double *M, *L1, *d1, *R1, *L2, *d2, *R2;
    double *Lo, *dO, *Ro, *G, *logdet;
};
__global__ void __launch_bounds__(kThreads) k_ldr_op(Params p, int op, LdrArgs a) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int w = blockIdx.x, n = p.N;
    const size_t nn = (size_t)n * n;
    Work wk;
    carve(wk, p, w, p.smem_mats_stab, smem);
    double *b0 = wk.b[0], *b1 = wk.b[1], *b2 = wk.b[2];
    const int l0 = wk.ld[0], l1 = wk.ld[1], l2 = wk.ld[2];
    const double* M = a.M ? a.M + w * nn : nullptr;
    const double* L1 = a.L1 ? a.L1 + w * nn : nullptr;
    const double* R1 = a.R1 ? a.R1 + w * nn : nullptr;
    const double* d1 = a.d1 ? a.d1 + (size_t)w * n : nullptr;
    const double* L2 = a.L2 ? a.L2 + w * nn : nullptr;
    const double* R2 = a.R2 ? a.R2 + w * nn : nullptr;
    const double* d2 = a.d2 ? a.d2 + (size_t)w * n : nullptr;
    double* Lo = a.Lo ? a.Lo + w * nn : nullptr;
    double* Ro = a.Ro ? a.Ro + w * nn : nullptr;
    double* dO = a.dO ? a.dO + (size_t)w * n : nullptr;
    double* G = a.G ? a.G + w * nn : nullptr;
    if (op == 0) {
        cta_copy(b0, l0, M, n, n);
        dev_to_ldr(wk, n, b0, l0, b1, l1, b2, l2, wk.dnew);
        cta_copy(Lo, n, b1, l1, n);
        cta_copy(Ro, n, b2, l2, n);
        store_d(dO, wk.dnew, n);
    } else if (op == 1) {  // mat_mul_ldr(M, F1)
        store_d(wk.sv[0], d1, n);
        __syncthreads();
        dev_gemm<false>(wk, b1, l1, M, 1, n, L1, 1, n, n, nullptr, wk.sv[0]);
        dev_to_ldr(wk, n, b1, l1, b0, l0, b2, l2, wk.dnew);
        dev_gemm<false>(wk, Ro, n, b2, 1, l2, R1, 1, n, n, nullptr, nullptr);
        cta_copy(Lo, n, b0, l0, n);
        store_d(dO, wk.dnew, n);
    } else if (op == 2) {  // ldr_mul_mat(F1, M)
        store_d(wk.sv[0], d1, n);
        __syncthreads();
        dev_gemm<false>(wk, b1, l1, R1, 1, n, M, 1, n, n, wk.sv[0], nullptr);
        dev_to_ldr(wk, n, b1, l1, b0, l0, b2, l2, wk.dnew);
        dev_gemm<false>(wk, Lo, n, L1, 1, n, b0, 1, l0, n, nullptr, nullptr);
        cta_copy(Ro, n, b2, l2, n);
        store_d(dO, wk.dnew, n);
    } else if (op == 3) {  // ldr_mul_ldr(F1, F2)
        store_d(wk.sv[0], d1, n);
        store_d(wk.sv[2], d2, n);
        __syncthreads();
        dev_gemm<false>(wk, b0, l0, R1, 1, n, L2, 1, n, n, wk.sv[0], wk.sv[2]);
        dev_to_ldr(wk, n, b0, l0, b1, l1, b2, l2, wk.dnew);
        dev_gemm<false>(wk, Lo, n, L1, 1, n, b1, 1, l1, n, nullptr, nullptr);
        dev_gemm<false>(wk, Ro, n, b2, 1, l2, R2, 1, n, n, nullptr, nullptr);
        store_d(dO, wk.dnew, n);
    } else if (op == 4) {
        const double ld_ = dev_inv_I_plus_ldr(p, wk, n, L1, n, d1, R1, n);
        if (threadIdx.x == 0 && a.logdet) a.logdet[w] = ld_;
        cta_copy(G, n, b1, l1, n);
    } else if (op == 5 || op == 6) {
        dev_inv_I_plus_ldr_mul_ldr(p, wk, n, L1, d1, R1, L2, d2, R2, n, op == 6);
        cta_copy(G, n, b2, l2, n);
    }
}

// AttractiveHubbard::local_update_ratio (source/model.cpp:109-122) for one proposal of every walker
__global__ void k_ratio(Params p, int l, int site, int newf, double* out /*[W][2]*/) {
    const int w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= p.W) return;
    const int oldf = p.fields[((size_t)w * p.nt + l) * p.N + site];
    const double gii = p.G[(size_t)w * p.N * p.N + site + (size_t)site * p.N];
    const double delta = p.lut_delta[oldf * 4 + newf];
    const double r = 1.0 + (1.0 - gii) * delta;
    out[2 * w] = p.lut_gb[oldf * 4 + newf] * (r * r);
    out[2 * w + 1] = delta;
}

// AttractiveHubbard::update_greens_local (source/model.cpp:124-138) on walker `w`: rank-1 Sherman-Morrison
__global__ void __launch_bounds__(kThreads) k_rank1(Params p, int w, double delta, int site) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int n = p.N;
    double* su = reinterpret_cast<double*>(smem);
    double* sv = su + n;
    double* G = p.G + (size_t)w * n * n;
    const double pref = delta / (1.0 + (1.0 - G[site + (size_t)site * n]) * delta);
    for (int t = threadIdx.x; t < n; t += blockDim.x) {
        su[t] = __dmul_rn(pref, G[t + (size_t)site * n]);
        sv[t] = G[site + (size_t)t * n] - ((t == site) ? 1.0 : 0.0);
    }
    __syncthreads();
    for (int e = threadIdx.x; e < n * n; e += blockDim.x) G[e] = __dadd_rn(G[e], __dmul_rn(su[e % n], sv[e / n]));
}

// =====================================================================================================
// host side
// =====================================================================================================
static thread_local std::string g_create_err;

#define CK(call)                                                                                  \
    do {                                                                                          \
        cudaError_t e_ = (call);                                                                  \
        if (e_ != cudaSuccess) {                                                                  \
            ctx->err = std::string(#call) + ": " + cudaGetErrorString(e_);                        \
            return DQMCB_ECUDA;                                                                   \
        }                                                                                         \
    } while (0)

// Shared-memory plan of the stabilisation kernels for an n-site walker: leading dimension of the work tiles and which of
// {three resident tiles, one staging tile, the TMA operand arena of gemm_tma.cuh} shared memory holds.  Returns the bytes.
static size_t plan_stab_smem(Params& p, int n) {
    p.LD = (n % 2 == 0) ? n + 1 : n;  // odd leading dimension: conflict-free row access in SMEM
    p.smem_arena = 0;
    const size_t sb = ((small_bytes(n) + 15) & ~(size_t)15);
    const char* nt_ = getenv("DQMCB_NO_TMA_GEMM");
    // N = 256: no tile of the walker fits in shared memory, so the work tiles stay in global memory with an EVEN leading
    // dimension (16-byte-aligned columns for cp.async.bulk; n + 2 rather than n: with columns exactly 2 KB apart the
    // column-wise factorisations pile onto a few L2 slices, measured +30% on the QR) and shared memory holds the operand
    // panels of gemm_tma.cuh
    if (n == kGT_N && sb + (size_t)n * p.LD * 8 > (size_t)kMaxSmem && sb + kGemmTmaArenaBytes <= (size_t)kMaxSmem && !(nt_ && nt_[0] == (char)49)) {
        p.LD = n + 2;
        p.smem_arena = 1;
    }
    const size_t tile = (size_t)n * p.LD * 8;
    p.smem_mats_stab = (sb + 3 * tile <= (size_t)kMaxSmem) ? 3 : 0;
    p.smem_stage = (!p.smem_arena && p.smem_mats_stab == 0 && sb + tile <= (size_t)kMaxSmem) ? 1 : 0;
    // beside the staging tile: a right-hand-side panel for the triangular solves (n = 144: 28 KB)
    const size_t rhs = (size_t)(n + 4) * kRhsPanelCols * 8;
    const char* np_ = getenv("DQMCB_NO_RHS_PANEL");
    p.smem_rhs = (p.smem_stage && (n & 15) == 0 && n % kRhsPanelCols == 0 && (n & 7) == 0 && sb + tile + rhs <= (size_t)kMaxSmem &&
                  !(np_ && np_[0] == (char)49)) ? 1 : 0;
    return sb + (p.smem_mats_stab + p.smem_stage) * tile + (p.smem_arena ? kGemmTmaArenaBytes : 0) + (p.smem_rhs ? rhs : 0);
}

static size_t prop_small_bytes(int n) {
    size_t off = (size_t)(5 * n) * 8 + (size_t)(3 * n) * 4 + (size_t)n;
    return (off + 15) & ~(size_t)15;
}

static void fill_luts(const dqmcb_config& c, Params& p) {
    for (int o = 0; o < 4; ++o) {
        for (int nw = 0; nw < 4; ++nw) {
            // AttractiveHubbard::bosonic_ratio / local_update_ratio (source/model.cpp:99-122)
            const double d_eta = c.eta[nw] - c.eta[o];
            const double boson = std::exp(c.alpha * c.g * d_eta);
            const double delta = (1.0 / boson) - 1.0;
            const double gammaR = c.gamma[nw] / c.gamma[o];
            p.lut_delta[o * 4 + nw] = delta;
            p.lut_gb[o * 4 + nw] = gammaR * boson;
        }
        p.lut_ev[o] = std::exp(c.g * c.eta[o]);   // source/model.cpp:69
        p.lut_iev[o] = std::exp(-c.g * c.eta[o]);  // source/model.cpp:81
        p.lut_action[o] = c.alpha * c.g * c.eta[o] + std::log(c.gamma[o]);
    }
}

struct TimeScope {
    dqmcb_ctx* c;
    dqmcb_ctx::Span sp;
    bool on;
    TimeScope(dqmcb_ctx* ctx, int cls) : c(ctx), on(ctx->timing) {
        if (!on) return;
        sp.cls = cls;
        cudaEventCreate(&sp.a);
        cudaEventCreate(&sp.b);
        cudaEventRecord(sp.a, c->stream);
    }
    ~TimeScope() {
        if (!on) return;
        cudaEventRecord(sp.b, c->stream);
        c->spans.push_back(sp);
    }
};

static void drain_spans(dqmcb_ctx* ctx) {
    for (auto& sp : ctx->spans) {
        cudaEventSynchronize(sp.b);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, sp.a, sp.b);
        ctx->t_ms[sp.cls] += ms;
        ctx->t_n[sp.cls] += 1;
        cudaEventDestroy(sp.a);
        cudaEventDestroy(sp.b);
    }
    ctx->spans.clear();
}

extern "C" {

int dqmcb_timing_enable(dqmcb_ctx* ctx, int on) {
    if (!ctx) return DQMCB_EINVAL;
    cudaSetDevice(ctx->cfg.device);
    drain_spans(ctx);
    ctx->timing = on != 0;
    if (on)
        for (int i = 0; i < DQMCB_K_CLASSES; ++i) {
            ctx->t_ms[i] = 0.0;
            ctx->t_n[i] = 0;
        }
    return DQMCB_OK;
}
int dqmcb_timing_get(dqmcb_ctx* ctx, double* ms, long long* launches) {
    if (!ctx) return DQMCB_EINVAL;
    cudaSetDevice(ctx->cfg.device);
    drain_spans(ctx);
    for (int i = 0; i < DQMCB_K_CLASSES; ++i) {
        if (ms) ms[i] = ctx->t_ms[i];
        if (launches) launches[i] = ctx->t_n[i];
    }
    return DQMCB_OK;
}

const char* dqmcb_version(void) { return "dqmc_b200 0.1 (sm_100a)"; }

const char* dqmcb_last_error(const dqmcb_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

int dqmcb_partner_rank(int rank, int world_size, int exchange_attempt) {
    // update::partner_rank (source/update.cpp:34-45): periodic even/odd pairing
    const bool even_attempt = (exchange_attempt % 2 == 0);
    const bool even_rank = (rank % 2 == 0);
    const int off = (even_attempt == even_rank) ? 1 : -1;
    return (rank + off + world_size) % world_size;
}

int dqmcb_create(const dqmcb_config* cfg, const double* expK, const double* invexpK, dqmcb_ctx** out) {
    if (!cfg || !expK || !invexpK || !out) {
        g_create_err = "null argument";
        return DQMCB_EINVAL;
    }
    if (cfg->n_sites <= 0 || cfg->n_sites > 256 || cfg->nt <= 0 || cfg->n_stab <= 0 || cfg->n_walkers <= 0) {
        g_create_err = "bad dimensions (need 0 < n_sites <= 256, nt > 0, n_stab > 0, n_walkers > 0)";
        return DQMCB_EINVAL;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        g_create_err = "no CUDA device: dqmc_b200 has no CPU fallback";
        return DQMCB_ECUDA;
    }
    if (cfg->device < 0 || cfg->device >= ndev) {
        g_create_err = "bad device ordinal";
        return DQMCB_EINVAL;
    }
    dqmcb_ctx* ctx = new dqmcb_ctx();
    ctx->cfg = *cfg;
    Params& p = ctx->p;
    memset(&p, 0, sizeof(p));
    const int N = cfg->n_sites, W = cfg->n_walkers, nt = cfg->nt;
    p.N = N;
    const size_t stab_smem = plan_stab_smem(p, N);
    p.nt = nt;
    p.n_stab = cfg->n_stab;
    p.n_stack = (nt + cfg->n_stab - 1) / cfg->n_stab;
    p.W = W;
    p.w0 = 0;
    p.w1 = W;
    p.rng_mode = DQMCB_RNG_PHILOX;
    p.seed = 0x5DEECE66Dull;
    fill_luts(*cfg, p);
#define CKC(call)                                                          \
    do {                                                                   \
        cudaError_t e_ = (call);                                           \
        if (e_ != cudaSuccess) {                                           \
            g_create_err = std::string(#call) + ": " + cudaGetErrorString(e_); \
            dqmcb_destroy(ctx);                                            \
            return DQMCB_ECUDA;                                            \
        }                                                                  \
    } while (0)
    CKC(cudaSetDevice(cfg->device));
    const size_t nn = (size_t)N * N;
    CKC(cudaMalloc(&ctx->d_expK, nn * 8));
    CKC(cudaMalloc(&ctx->d_invexpK, nn * 8));
    CKC(cudaMalloc(&ctx->d_expKh, nn * 8));
    CKC(cudaMalloc(&ctx->d_invexpKh, nn * 8));
    CKC(cudaMemcpy(ctx->d_expK, expK, nn * 8, cudaMemcpyHostToDevice));
    CKC(cudaMemcpy(ctx->d_invexpK, invexpK, nn * 8, cudaMemcpyHostToDevice));
    CKC(cudaMemset(ctx->d_expKh, 0, nn * 8));
    CKC(cudaMemset(ctx->d_invexpKh, 0, nn * 8));
    p.expK = ctx->d_expK;
    p.invexpK = ctx->d_invexpK;
    p.expK_half = ctx->d_expKh;
    p.invexpK_half = ctx->d_invexpKh;
    p.kron = nullptr;
    {
        // Kronecker structure of exp(-dtau K) for the 8 x 8 lattice (site i = x + 8 y): E = Ey (x) Ex with
        // Ex = E[(0,x),(0,x')] and Ey = E[(y,0),(y',0)] / E[0,0]; accepted only if it reproduces E to 1e-14 of its
        // largest entry (any other geometry, anisotropy that breaks the product form, or a user matrix: dense path)
        const char* nk = getenv("DQMCB_NO_KRON");
        if (N == 64 && !(nk && nk[0] == (char)49) && expK[0] != 0.0 && invexpK[0] != 0.0) {
            double f[256];  // Ex, Ey (of expK), Exi, Eyi (of invexpK), 8 x 8 column-major each
            bool ok = true;
            for (int which = 0; which < 2 && ok; ++which) {
                const double* E = which ? invexpK : expK;
                double* fx = f + 128 * which;
                double* fy = fx + 64;
                for (int b = 0; b < 8; ++b)
                    for (int a = 0; a < 8; ++a) {
                        fx[a + 8 * b] = E[(size_t)a + (size_t)b * N];                         // Fx(a, b)
                        fy[a + 8 * b] = E[(size_t)(8 * a) + (size_t)(8 * b) * N] / E[0];      // Fy(a, b)
                    }
                double emax = 0.0, dmax = 0.0;
                for (int c = 0; c < N; ++c)
                    for (int r = 0; r < N; ++r) {
                        const double e = E[(size_t)r + (size_t)c * N];
                        const double k = fy[(r >> 3) + 8 * (c >> 3)] * fx[(r & 7) + 8 * (c & 7)];
                        emax = std::max(emax, std::fabs(e));
                        dmax = std::max(dmax, std::fabs(e - k));
                    }
                ok = dmax <= 1e-14 * emax;
                // the row-wise passes of the wraps (kron_right16) and the transposed chain of the backward stabilisation feed
                // F(gr, tg) where F(tg, gr) is meant: they need SYMMETRIC factors.  A matrix of product form that is not
                // symmetric (a caller-supplied exp(-dtau K)) takes the dense path.
                double smax = 0.0;
                for (int b = 0; b < 8; ++b)
                    for (int a = 0; a < 8; ++a) {
                        smax = std::max(smax, std::fabs(fx[a + 8 * b] - fx[b + 8 * a]));
                        smax = std::max(smax, std::fabs(fy[a + 8 * b] - fy[b + 8 * a]));
                    }
                ok = ok && smax <= 1e-14 * std::max(1.0, emax / std::fabs(E[0])) * std::fabs(fx[0]);
            }
            if (ok) {
                CKC(cudaMalloc(&ctx->d_kron, sizeof(f)));
                CKC(cudaMemcpy(ctx->d_kron, f, sizeof(f), cudaMemcpyHostToDevice));
                p.kron = ctx->d_kron;
            }
        }
    }
    p.kg = nullptr;
    p.kL1 = p.kL2 = 0;
    {
        // Generic Kronecker structure: try every factorisation N = L1 L2 (site i = x + L1 y; most balanced first) and
        // accept the first whose factors Fx = E[(x,0),(x',0)], Fy = E[(0,y),(0,y')] / E[0,0] reproduce exp(-dtau K) AND
        // exp(+dtau K) to 1e-14 of their largest entry and are symmetric (the right products use the transposed view).
        const char* nk = getenv("DQMCB_NO_KRON");
        std::vector<int> cand;
        // Lattices below 7 x 7 keep the dense products: the run-time-masked routine (no compile-time extents there) costs more
        // than 4 N^3 flops of plain FMA tiles at that size (measured at 6 x 6: propagate 1.00 -> 0.76 ms, stabilize 0.59 -> 0.50 ms
        // per launch; at 10 x 10 and 14 x 14 the structured products win by 10-15%).
        for (int a = 1; a <= 16 && N > 48; ++a)
            if (N % a == 0 && N / a <= 16) cand.push_back(a);
        std::sort(cand.begin(), cand.end(), [&](int a, int b) { return std::abs(a - N / a) < std::abs(b - N / b) || (std::abs(a - N / a) == std::abs(b - N / b) && a > b); });
        for (size_t ci = 0; ci < cand.size() && !p.kg && !(nk && nk[0] == (char)49) && expK[0] != 0.0 && invexpK[0] != 0.0; ++ci) {
            const int L1 = cand[ci], L2 = N / L1;
            std::vector<double> f(4 * 256, 0.0);
            bool ok = true;
            for (int which = 0; which < 2 && ok; ++which) {
                const double* E = which ? invexpK : expK;
                double* fx = f.data() + 512 * which;
                double* fy = fx + 256;
                for (int b = 0; b < L1; ++b)
                    for (int a = 0; a < L1; ++a) fx[a + 16 * b] = E[(size_t)a + (size_t)b * N];
                for (int b = 0; b < L2; ++b)
                    for (int a = 0; a < L2; ++a) fy[a + 16 * b] = E[(size_t)(L1 * a) + (size_t)(L1 * b) * N] / E[0];
                double emax = 0.0, dmax = 0.0, smax = 0.0;
                for (int c = 0; c < N && ok; ++c)
                    for (int r = 0; r < N; ++r) {
                        const double e = E[(size_t)r + (size_t)c * N];
                        const double k = fy[(r / L1) + 16 * (c / L1)] * fx[(r % L1) + 16 * (c % L1)];
                        emax = std::max(emax, std::fabs(e));
                        dmax = std::max(dmax, std::fabs(e - k));
                    }
                for (int b = 0; b < 16; ++b)
                    for (int a = 0; a < 16; ++a) {
                        smax = std::max(smax, std::fabs(fx[a + 16 * b] - fx[b + 16 * a]));
                        smax = std::max(smax, std::fabs(fy[a + 16 * b] - fy[b + 16 * a]));
                    }
                ok = dmax <= 1e-14 * emax && smax <= 1e-14 * emax;
            }
            if (ok) {
                CKC(cudaMalloc(&ctx->d_kg, f.size() * sizeof(double)));
                CKC(cudaMemcpy(ctx->d_kg, f.data(), f.size() * sizeof(double), cudaMemcpyHostToDevice));
                p.kg = ctx->d_kg;
                p.kL1 = L1;
                p.kL2 = L2;
            }
        }
    }
    CKC(cudaMalloc(&p.fields, (size_t)W * nt * N));
    CKC(cudaMemset(p.fields, 0, (size_t)W * nt * N));
    CKC(cudaMalloc(&p.G, W * nn * 8));
    CKC(cudaMalloc(&p.Bbar, W * nn * 8));
    CKC(cudaMalloc(&p.stackL, (size_t)W * p.n_stack * nn * 8));
    CKC(cudaMalloc(&p.stackR, (size_t)W * p.n_stack * nn * 8));
    CKC(cudaMalloc(&p.stackD, (size_t)W * p.n_stack * N * 8));
    CKC(cudaMalloc(&p.ws, (size_t)W * 3 * N * p.LD * 8));
    CKC(cudaMalloc(&p.logdet, (size_t)W * 8));
    CKC(cudaMalloc(&p.accepted, (size_t)W * 8));
    CKC(cudaMalloc(&p.err_max, (size_t)W * 8));
    CKC(cudaMalloc(&p.err_sum, (size_t)W * 8));
    CKC(cudaMalloc(&p.err_cnt, (size_t)W * 8));
    CKC(cudaMalloc(&p.fail, sizeof(int)));
    CKC(cudaMalloc(&ctx->d_obs, (size_t)W * 4 * 8));
    CKC(cudaMalloc(&ctx->d_scratchNN, W * nn * 8));
    CKC(cudaMemset(p.G, 0, W * nn * 8));
    CKC(cudaMemset(p.logdet, 0, (size_t)W * 8));
    CKC(cudaMemset(p.accepted, 0, (size_t)W * 8));
    CKC(cudaMemset(p.err_max, 0, (size_t)W * 8));
    CKC(cudaMemset(p.err_sum, 0, (size_t)W * 8));
    CKC(cudaMemset(p.err_cnt, 0, (size_t)W * 8));
    CKC(cudaMemset(p.fail, 0, sizeof(int)));
    // shared-memory plan
    const size_t tile = (size_t)N * p.LD * 8;
    const size_t sb = ((small_bytes(N) + 15) & ~(size_t)15);
    ctx->smem_stab = stab_smem;
    p.delay_k = (cfg->delay_k > 1) ? std::min(cfg->delay_k, 64) : 0;
    const size_t pb = prop_small_bytes(N);
    const size_t dly = p.delay_k ? ((size_t)2 * p.delay_k * delay_ld(N) + 2 * N) * 8 : 0;
    p.smem_mats_prop = (pb + dly + 2 * tile <= (size_t)kMaxSmem) ? 2 : ((pb + dly + tile <= (size_t)kMaxSmem) ? 1 : 0);
    ctx->smem_prop = pb + dly + p.smem_mats_prop * tile;
    if (ctx->smem_prop > (size_t)kMaxSmem) {
        g_create_err = "delay_k too large for shared memory at this N";
        dqmcb_destroy(ctx);
        return DQMCB_EINVAL;
    }
    CKC(cudaFuncSetAttribute(k_propagate<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ctx->smem_prop));
    CKC(cudaFuncSetAttribute(k_propagate<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ctx->smem_prop));
    CKC(cudaFuncSetAttribute(k_bbar, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ctx->smem_stab));
    CKC(cudaFuncSetAttribute(k_stabilize, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ctx->smem_stab));
    CKC(cudaFuncSetAttribute(k_init_stacks, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ctx->smem_stab));
    CKC(cudaFuncSetAttribute(k_init_green, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ctx->smem_stab));
    CKC(cudaFuncSetAttribute(k_ldr_op, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ctx->smem_stab));
    CKC(cudaFuncSetAttribute(k_uneq_block, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ctx->smem_stab));
    { const char* fg = getenv("DQMCB_FORCE_GENERIC"); ctx->force_generic = fg && fg[0] == (char)49; }
    {
        // threads per walker of the generic sweep kernels: on small lattices a step of a factorisation has only a few dozen
        // elements per column, so 256 threads mostly wait at barriers; 64 threads per walker let four to eight walkers share
        // an SM and interleave their latency chains (C0, N = 36: measured in profiles/r02_generic_threads_c0.json)
        const char* gt = getenv("DQMCB_GENERIC_THREADS");
        int t = gt ? atoi(gt) : (N <= 48 ? 64 : kThreads);
        ctx->gthreads = (t == 32 || t == 64 || t == 128 || t == 256) ? t : kThreads;
    }
    {
        const char* pp = getenv("DQMCB_PHASE_PROF");
        if (pp && pp[0] == (char)49) {
            CKC(cudaMalloc(&p.prof, 64 * sizeof(long long)));
            CKC(cudaMemset(p.prof, 0, 64 * sizeof(long long)));
            CKC(cudaMemcpyToSymbol(dq::g_dq_prof, &p.prof, sizeof(p.prof)));
        }
    }
    CKC(fast64_setup(ctx));
    {
        const char* s2e = getenv("DQMCB_STAB2");  // "0": keep the first-generation stabilisation kernel (A/B comparisons)
        ctx->use_stab2 = N == 64 && p.kron != nullptr && cfg->n_stab <= dq::s2::MAX_SLICES && !ctx->force_generic &&
                         !(s2e && s2e[0] == (char)48);
    }
    {
        const char* gs = getenv("DQMCB_STREAMS");
        int ng = gs ? atoi(gs) : 2;
        ng = std::max(1, std::min(ng, 8));
        if (W < 64 * ng || !fast64_available(ctx)) ng = 1;  // grouping only pays with many walkers per group
        ctx->n_groups = ng;
        if (ng > 1) {
            for (int g = 0; g < ng; ++g) {
                CKC(cudaStreamCreateWithFlags(&ctx->gstream[g], cudaStreamNonBlocking));
                CKC(cudaEventCreateWithFlags(&ctx->gdone[g], cudaEventDisableTiming));
            }
            CKC(cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
        }
    }
#undef CKC
    *out = ctx;
    return DQMCB_OK;
}

void dqmcb_destroy(dqmcb_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->cfg.device);
    Params& p = ctx->p;
    cudaFree(ctx->d_expK);
    cudaFree(ctx->d_invexpK);
    cudaFree(ctx->d_expKh);
    cudaFree(ctx->d_invexpKh);
    cudaFree(ctx->d_kron);
    cudaFree(ctx->d_kg);
    cudaFree(p.fields);
    cudaFree(p.G);
    cudaFree(p.Bbar);
    cudaFree(p.stackL);
    cudaFree(p.stackR);
    cudaFree(p.stackD);
    cudaFree(p.ws);
    cudaFree(p.logdet);
    cudaFree(p.accepted);
    cudaFree(p.err_max);
    cudaFree(p.err_sum);
    cudaFree(p.err_cnt);
    cudaFree(p.fail);
    cudaFree(ctx->d_obs);
    cudaFree(ctx->d_tape2_perm);
    cudaFree(ctx->d_tape2_k);
    cudaFree(ctx->d_tape2_u);
    if (ctx->cstream) cudaStreamDestroy(ctx->cstream);
    if (ctx->ev_tape) cudaEventDestroy(ctx->ev_tape);
    if (ctx->ev_free) cudaEventDestroy(ctx->ev_free);
    cudaFree(ctx->d_binpart);
    cudaFree(ctx->d_binacc);
    cudaFree(ctx->d_uneq);
    cudaFree(ctx->d_bt0);
    cudaFree(ctx->d_uneq_err);
    cudaFree(ctx->d_tbinpart);
    cudaFree(ctx->d_tbinacc);
    cudaFree(ctx->d_scratchNN);
    cudaFree(ctx->d_tape_perm);
    cudaFree(ctx->d_tape_k);
    cudaFree(ctx->d_tape_u);
    cudaFree(ctx->d_mt);
    for (int g = 0; g < 8; ++g) {
        if (ctx->gstream[g]) cudaStreamDestroy(ctx->gstream[g]);
        if (ctx->gdone[g]) cudaEventDestroy(ctx->gdone[g]);
    }
    if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
    if (p.prof) {
        long long h[64];
        cudaDeviceSynchronize();
        if (cudaMemcpy(h, p.prof, sizeof(h), cudaMemcpyDeviceToHost) == cudaSuccess) {
            fprintf(stderr, "[dqmcb phase prof] clock64 offsets from stamp 0, last instrumented launch (CTA 0):");
            for (int i = 1; i < 40; ++i)
                if (h[i]) fprintf(stderr, " %d:%lld", i, h[i] - h[0]);
            fprintf(stderr, "\n");
            if (h[44])  // accumulated over every kf_propagate launch, walker 0 of CTA 0
                fprintf(stderr, "[dqmcb phase prof] propagate (walker 0, per slice): wrap %lld clk, Metropolis %lld clk, draws+rest %lld clk, accepts %.1f of N\n",
                        h[40] / h[44], h[41] / h[44], (h[43] - h[40] - h[41]) / h[44], (double)h[42] / (double)h[44]);
        }
        cudaFree(p.prof);
    }
    drain_spans(ctx);
    delete ctx;
}

#define NEED(ctx)                     \
    if (!(ctx)) return DQMCB_EINVAL;  \
    CK(cudaSetDevice((ctx)->cfg.device))

int dqmcb_set_stream(dqmcb_ctx* ctx, void* s) {
    if (!ctx) return DQMCB_EINVAL;
    ctx->stream = (cudaStream_t)s;
    return DQMCB_OK;
}
int dqmcb_synchronize(dqmcb_ctx* ctx) {
    NEED(ctx);
    CK(cudaStreamSynchronize(ctx->stream));
    int fail = 0;
    CK(cudaMemcpy(&fail, ctx->p.fail, sizeof(int), cudaMemcpyDeviceToHost));
    if (fail) {
        ctx->err = "LU factorisation hit an exactly singular pivot (code " + std::to_string(fail) + ")";
        cudaMemset(ctx->p.fail, 0, sizeof(int));
        return DQMCB_ENUMERIC;
    }
    return DQMCB_OK;
}
int dqmcb_n_stack(const dqmcb_ctx* ctx) { return ctx ? ctx->p.n_stack : -1; }
long long dqmcb_launch_count(const dqmcb_ctx* ctx) { return ctx ? ctx->n_launches : -1; }
int dqmcb_n_groups(const dqmcb_ctx* ctx) { return ctx ? ctx->n_groups : -1; }

int dqmcb_set_half_expK(dqmcb_ctx* ctx, const double* a, const double* b) {
    NEED(ctx);
    const size_t nn = (size_t)ctx->p.N * ctx->p.N;
    CK(cudaMemcpy(ctx->d_expKh, a, nn * 8, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(ctx->d_invexpKh, b, nn * 8, cudaMemcpyHostToDevice));
    return DQMCB_OK;
}

// ---- fields -----------------------------------------------------------------------------------------
int dqmcb_set_fields(dqmcb_ctx* ctx, const int8_t* f) {
    NEED(ctx);
    const size_t sz = (size_t)ctx->p.W * ctx->p.nt * ctx->p.N;
    for (size_t i = 0; i < sz; ++i)
        if (f[i] < 0 || f[i] > 3) {
            ctx->err = "field value outside {0,1,2,3}";
            return DQMCB_EINVAL;
        }
    CK(cudaMemcpyAsync(ctx->p.fields, f, sz, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->stacks_ready = ctx->green_ready = false;
    return DQMCB_OK;
}
int dqmcb_get_fields(dqmcb_ctx* ctx, int8_t* f) {
    NEED(ctx);
    const size_t sz = (size_t)ctx->p.W * ctx->p.nt * ctx->p.N;
    CK(cudaMemcpyAsync(f, ctx->p.fields, sz, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return DQMCB_OK;
}
int dqmcb_set_fields_i64(dqmcb_ctx* ctx, int walker, const long long* f) {
    NEED(ctx);
    const int nt = ctx->p.nt, N = ctx->p.N;
    if (walker < 0 || walker >= ctx->p.W) return DQMCB_EINVAL;
    std::vector<int8_t> tmp((size_t)nt * N);
    for (int s = 0; s < N; ++s)
        for (int l = 0; l < nt; ++l) {
            const long long v = f[l + (size_t)s * nt];  // arma::imat(nt, N) column-major
            if (v < 0 || v > 3) {
                ctx->err = "field value outside {0,1,2,3}";
                return DQMCB_EINVAL;
            }
            tmp[(size_t)l * N + s] = (int8_t)v;
        }
    CK(cudaMemcpy(ctx->p.fields + (size_t)walker * nt * N, tmp.data(), tmp.size(), cudaMemcpyHostToDevice));
    ctx->stacks_ready = ctx->green_ready = false;
    return DQMCB_OK;
}
int dqmcb_get_fields_i64(dqmcb_ctx* ctx, int walker, long long* f) {
    NEED(ctx);
    const int nt = ctx->p.nt, N = ctx->p.N;
    if (walker < 0 || walker >= ctx->p.W) return DQMCB_EINVAL;
    std::vector<int8_t> tmp((size_t)nt * N);
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemcpy(tmp.data(), ctx->p.fields + (size_t)walker * nt * N, tmp.size(), cudaMemcpyDeviceToHost));
    for (int s = 0; s < N; ++s)
        for (int l = 0; l < nt; ++l) f[l + (size_t)s * nt] = tmp[(size_t)l * N + s];
    return DQMCB_OK;
}

// ---- RNG ----------------------------------------------------------------------------------------------
static int mt_seed(dqmcb_ctx* ctx, const uint32_t* seeds) {
    const Params& p = ctx->p;
    if (!ctx->d_mt) CK(cudaMalloc(&ctx->d_mt, (size_t)p.W * 625 * sizeof(uint32_t)));
    uint32_t* d_seeds = nullptr;
    CK(cudaMalloc(&d_seeds, (size_t)p.W * sizeof(uint32_t)));
    CK(cudaMemcpy(d_seeds, seeds, (size_t)p.W * sizeof(uint32_t), cudaMemcpyHostToDevice));
    k_mt_seed<<<(p.W + 127) / 128, 128, 0, ctx->stream>>>(p.W, d_seeds, ctx->d_mt);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(ctx->stream));
    cudaFree(d_seeds);
    ctx->p.rng_mode = DQMCB_RNG_MT19937;
    ctx->p.visit_base = 0;
    ctx->tape_pos = 0;
    ctx->p.tape_visits = 0;
    return DQMCB_OK;
}
int dqmcb_rng_mode(dqmcb_ctx* ctx, int mode, uint64_t seed) {
    NEED(ctx);
    if (mode != DQMCB_RNG_TAPE && mode != DQMCB_RNG_PHILOX && mode != DQMCB_RNG_MT19937) return DQMCB_EINVAL;
    if (mode == DQMCB_RNG_MT19937) {  // walker w gets std::mt19937(seed + w)
        std::vector<uint32_t> seeds(ctx->p.W);
        for (int w = 0; w < ctx->p.W; ++w) seeds[w] = (uint32_t)(seed + (uint64_t)w);
        return mt_seed(ctx, seeds.data());
    }
    ctx->p.rng_mode = mode;
    ctx->p.seed = seed;
    ctx->p.visit_base = 0;
    ctx->tape_pos = 0;
    ctx->p.tape_visits = 0;
    return DQMCB_OK;
}
int dqmcb_rng_seed_mt19937(dqmcb_ctx* ctx, const uint32_t* seeds) {
    NEED(ctx);
    if (!seeds) return DQMCB_EINVAL;
    return mt_seed(ctx, seeds);
}
// MT19937 mode: regenerate the next `n` visits of every walker's libstdc++ stream into the tape buffers
static int mt_refill(dqmcb_ctx* ctx, int n) {
    const Params& p = ctx->p;
    const size_t cnt = (size_t)p.W * n * p.N;
    if ((size_t)n > ctx->tape_cap) {
        CK(cudaStreamSynchronize(ctx->stream));
        cudaFree(ctx->d_tape_perm);
        cudaFree(ctx->d_tape_k);
        cudaFree(ctx->d_tape_u);
        CK(cudaMalloc(&ctx->d_tape_perm, cnt));
        CK(cudaMalloc(&ctx->d_tape_k, cnt));
        CK(cudaMalloc(&ctx->d_tape_u, cnt * 8));
        ctx->tape_cap = n;
    }
    const size_t smem = (size_t)624 * kMtWalkers * sizeof(uint32_t) + (size_t)p.N * kMtWalkers;
    if (!ctx->mt_attr_set) {  // per context (= per device): the attribute belongs to the device the context lives on
        CK(cudaFuncSetAttribute(k_mt_fill, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(624 * kMtWalkers * 4 + 256 * kMtWalkers)));
        ctx->mt_attr_set = true;
    }
    k_mt_fill<<<(p.W + kMtWalkers - 1) / kMtWalkers, kMtWalkers, smem, ctx->stream>>>(p.W, p.N, n, ctx->d_mt, ctx->d_tape_perm, ctx->d_tape_k,
                                                                                     ctx->d_tape_u);
    CK(cudaGetLastError());
    ctx->n_launches += 1;
    ctx->p.tape_perm = ctx->d_tape_perm;
    ctx->p.tape_k = ctx->d_tape_k;
    ctx->p.tape_u = ctx->d_tape_u;
    ctx->p.tape_visits = n;
    ctx->tape_pos = 0;
    ctx->p.visit_base = 0;
    return DQMCB_OK;
}
int dqmcb_set_tape(dqmcb_ctx* ctx, int n_visits, const uint8_t* perm, const uint8_t* knew, const double* u) {
    NEED(ctx);
    if (n_visits <= 0 || !perm || !knew || !u) return DQMCB_EINVAL;
    const size_t cnt = (size_t)ctx->p.W * n_visits * ctx->p.N;
    CK(cudaStreamSynchronize(ctx->stream));
    if ((size_t)n_visits > ctx->tape_cap) {
        cudaFree(ctx->d_tape_perm);
        cudaFree(ctx->d_tape_k);
        cudaFree(ctx->d_tape_u);
        CK(cudaMalloc(&ctx->d_tape_perm, cnt));
        CK(cudaMalloc(&ctx->d_tape_k, cnt));
        CK(cudaMalloc(&ctx->d_tape_u, cnt * 8));
        ctx->tape_cap = n_visits;
    }
    CK(cudaMemcpy(ctx->d_tape_perm, perm, cnt, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(ctx->d_tape_k, knew, cnt, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(ctx->d_tape_u, u, cnt * 8, cudaMemcpyHostToDevice));
    ctx->p.tape_perm = ctx->d_tape_perm;
    ctx->p.tape_k = ctx->d_tape_k;
    ctx->p.tape_u = ctx->d_tape_u;
    ctx->p.tape_visits = n_visits;
    ctx->p.rng_mode = DQMCB_RNG_TAPE;
    ctx->tape_pos = 0;
    ctx->p.visit_base = 0;
    return DQMCB_OK;
}

// Asynchronous variant for pipelined end-to-end runs: the tape of the NEXT step is uploaded (from pinned host
// memory, on a copy stream) into the inactive buffer set while the kernels of the current step still read the
// active one; dqmcb_tape_commit then switches sets in stream order.
int dqmcb_set_tape_async(dqmcb_ctx* ctx, int n_visits, const uint8_t* perm, const uint8_t* knew, const double* u) {
    NEED(ctx);
    if (n_visits <= 0 || !perm || !knew || !u) return DQMCB_EINVAL;
    const size_t cnt = (size_t)ctx->p.W * n_visits * ctx->p.N;
    if (!ctx->cstream) {
        CK(cudaStreamCreateWithFlags(&ctx->cstream, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&ctx->ev_tape, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&ctx->ev_free, cudaEventDisableTiming));
        CK(cudaEventRecord(ctx->ev_free, ctx->stream));
    }
    if ((size_t)n_visits > ctx->tape2_cap) {
        CK(cudaStreamSynchronize(ctx->cstream));
        cudaFree(ctx->d_tape2_perm);
        cudaFree(ctx->d_tape2_k);
        cudaFree(ctx->d_tape2_u);
        CK(cudaMalloc(&ctx->d_tape2_perm, cnt));
        CK(cudaMalloc(&ctx->d_tape2_k, cnt));
        CK(cudaMalloc(&ctx->d_tape2_u, cnt * 8));
        ctx->tape2_cap = n_visits;
    }
    // the inactive set was the active one until the last commit: its readers precede ev_free in ctx->stream
    CK(cudaStreamWaitEvent(ctx->cstream, ctx->ev_free, 0));
    CK(cudaMemcpyAsync(ctx->d_tape2_perm, perm, cnt, cudaMemcpyHostToDevice, ctx->cstream));
    CK(cudaMemcpyAsync(ctx->d_tape2_k, knew, cnt, cudaMemcpyHostToDevice, ctx->cstream));
    CK(cudaMemcpyAsync(ctx->d_tape2_u, u, cnt * 8, cudaMemcpyHostToDevice, ctx->cstream));
    CK(cudaEventRecord(ctx->ev_tape, ctx->cstream));
    ctx->tape2_visits = n_visits;
    return DQMCB_OK;
}
int dqmcb_tape_commit(dqmcb_ctx* ctx) {
    NEED(ctx);
    if (ctx->tape2_visits <= 0) {
        ctx->err = "dqmcb_tape_commit: no tape pending (call dqmcb_set_tape_async first)";
        return DQMCB_ESTATE;
    }
    CK(cudaEventRecord(ctx->ev_free, ctx->stream));           // everything that reads the old set is before this
    CK(cudaStreamWaitEvent(ctx->stream, ctx->ev_tape, 0));    // the new set is complete before anything after this
    std::swap(ctx->d_tape_perm, ctx->d_tape2_perm);
    std::swap(ctx->d_tape_k, ctx->d_tape2_k);
    std::swap(ctx->d_tape_u, ctx->d_tape2_u);
    std::swap(ctx->tape_cap, ctx->tape2_cap);
    ctx->p.tape_perm = ctx->d_tape_perm;
    ctx->p.tape_k = ctx->d_tape_k;
    ctx->p.tape_u = ctx->d_tape_u;
    ctx->p.tape_visits = ctx->tape2_visits;
    ctx->tape2_visits = 0;
    ctx->p.rng_mode = DQMCB_RNG_TAPE;
    ctx->tape_pos = 0;
    ctx->p.visit_base = 0;
    return DQMCB_OK;
}

// consume `n` visits: returns the first visit index to hand to the kernel
static int take_visits(dqmcb_ctx* ctx, int n, unsigned long long* first) {
    if (ctx->p.rng_mode == DQMCB_RNG_MT19937) {  // the stream is consumed strictly in call order: generate exactly n visits
        if (int rc = mt_refill(ctx, n)) return rc;
        *first = 0;
        ctx->tape_pos = n;
        return DQMCB_OK;
    }
    if (ctx->p.rng_mode == DQMCB_RNG_TAPE) {
        if (ctx->tape_pos + n > ctx->p.tape_visits) {
            ctx->err = "RNG tape exhausted: call dqmcb_set_tape with enough visits";
            return DQMCB_ESTATE;
        }
        *first = (unsigned long long)ctx->tape_pos;
        ctx->tape_pos += n;
    } else {
        *first = ctx->p.visit_base;
        ctx->p.visit_base += n;
    }
    return DQMCB_OK;
}

// ---- launches -------------------------------------------------------------------------------------------
static int launch_propagate(dqmcb_ctx* ctx, int l_first, int n_slices, int dir, int flags) {
    unsigned long long v0 = 0;
    if (flags & 2) {
        int rc = take_visits(ctx, n_slices, &v0);
        if (rc) return rc;
    }
    Params p = ctx->p;
    if (p.rng_mode != DQMCB_RNG_TAPE) p.visit_base = 0;
    TimeScope ts(ctx, DQMCB_K_PROPAGATE);
    ctx->n_launches += 1;
    if (fast64_available(ctx)) return fast64_propagate(ctx, p, l_first, n_slices, dir, v0, flags);
    if (p.N >= 128)
        k_propagate<true><<<p.W, ctx->gthreads, ctx->smem_prop, ctx->stream>>>(p, l_first, n_slices, dir, v0, flags);
    else
        k_propagate<false><<<p.W, ctx->gthreads, ctx->smem_prop, ctx->stream>>>(p, l_first, n_slices, dir, v0, flags);
    CK(cudaGetLastError());
    return DQMCB_OK;
}

int dqmcb_init_stacks(dqmcb_ctx* ctx) {
    NEED(ctx);
    TimeScope ts(ctx, DQMCB_K_INIT);
    ctx->n_launches += 1;
    if (fast64_available(ctx)) fast_init_stacks(ctx); else k_init_stacks<<<ctx->p.W, ctx->gthreads, ctx->smem_stab, ctx->stream>>>(ctx->p);
    CK(cudaGetLastError());
    ctx->stacks_ready = true;
    return DQMCB_OK;
}
int dqmcb_init_greenfunctions(dqmcb_ctx* ctx) {
    NEED(ctx);
    if (!ctx->stacks_ready) {
        ctx->err = "init_greenfunctions before init_stacks";
        return DQMCB_ESTATE;
    }
    TimeScope ts(ctx, DQMCB_K_INIT);
    ctx->n_launches += 1;
    if (fast64_available(ctx)) fast_init_green(ctx); else k_init_green<<<ctx->p.W, ctx->gthreads, ctx->smem_stab, ctx->stream>>>(ctx->p);
    CK(cudaGetLastError());
    ctx->green_ready = true;
    return DQMCB_OK;
}
static int check_l(dqmcb_ctx* ctx, int l) {
    if (l < 0 || l >= ctx->p.nt) {
        ctx->err = "time slice out of range";
        return DQMCB_EINVAL;
    }
    return DQMCB_OK;
}
int dqmcb_wrap_forward(dqmcb_ctx* ctx, int l) {
    NEED(ctx);
    if (int rc = check_l(ctx, l)) return rc;
    return launch_propagate(ctx, l, 1, +1, 1);
}
int dqmcb_wrap_backward(dqmcb_ctx* ctx, int l) {
    NEED(ctx);
    if (int rc = check_l(ctx, l)) return rc;
    return launch_propagate(ctx, l, 1, -1, 1);
}
int dqmcb_local_update(dqmcb_ctx* ctx, int l) {
    NEED(ctx);
    if (int rc = check_l(ctx, l)) return rc;
    return launch_propagate(ctx, l, 1, +1, 2);
}
int dqmcb_block_product(dqmcb_ctx* ctx, int i_stack) {
    NEED(ctx);
    if (i_stack < 0 || i_stack >= ctx->p.n_stack) {
        ctx->err = "LDR Stack index out of bounds";
        return DQMCB_EINVAL;
    }
    if (fast64_available(ctx)) fast_bbar(ctx, i_stack); else k_bbar<<<ctx->p.W, ctx->gthreads, ctx->smem_stab, ctx->stream>>>(ctx->p, i_stack);
    CK(cudaGetLastError());
    return DQMCB_OK;
}
int dqmcb_get_block_product(dqmcb_ctx* ctx, double* out) {
    NEED(ctx);
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemcpy(out, ctx->p.Bbar, (size_t)ctx->p.W * ctx->p.N * ctx->p.N * 8, cudaMemcpyDeviceToHost));
    return DQMCB_OK;
}
static int launch_stabilize(dqmcb_ctx* ctx, int l, int dir) {
    TimeScope ts(ctx, DQMCB_K_STABILIZE);
    ctx->n_launches += 1;
    if (fast64_available(ctx)) {
        fast_stabilize(ctx, l, dir);
    } else if (ctx->p.smem_arena) {
        // N = 256: three launches (see k_stabilize); the middle one without the operand arena, i.e. with a large L1
        Params p2 = ctx->p;
        p2.smem_arena = 0;
        const size_t sb = ctx->smem_stab - kGemmTmaArenaBytes;
        k_stabilize<<<ctx->p.W, ctx->gthreads, ctx->smem_stab, ctx->stream>>>(ctx->p, l, dir, 1);
        k_stabilize<<<ctx->p.W, ctx->gthreads, sb, ctx->stream>>>(p2, l, dir, 2);
        k_stabilize<<<ctx->p.W, ctx->gthreads, ctx->smem_stab, ctx->stream>>>(ctx->p, l, dir, 3);
        ctx->n_launches += 2;
    } else {
        k_stabilize<<<ctx->p.W, ctx->gthreads, ctx->smem_stab, ctx->stream>>>(ctx->p, l, dir, 0);
    }
    CK(cudaGetLastError());
    return DQMCB_OK;
}
int dqmcb_stabilize_forward(dqmcb_ctx* ctx, int l) {
    NEED(ctx);
    if (int rc = check_l(ctx, l)) return rc;
    return launch_stabilize(ctx, l, +1);
}
int dqmcb_stabilize_backward(dqmcb_ctx* ctx, int l) {
    NEED(ctx);
    if (int rc = check_l(ctx, l)) return rc;
    return launch_stabilize(ctx, l, -1);
}

// Whole half sweep with the walkers split into groups on separate streams (fast path, timing off):
// the chains are independent, so a group's propagate -> stabilise chain only depends on itself and the
// partial last wave of one kernel is filled by another group's kernels.  Forks from / joins to the
// context's stream, so callers see ordinary stream semantics.
static int grouped_sweep(dqmcb_ctx* ctx, int dir) {
    const Params& p0 = ctx->p;
    const int ng = ctx->n_groups;
    unsigned long long v0 = 0;
    if (int rc = take_visits(ctx, p0.nt, &v0)) return rc;
    CK(cudaEventRecord(ctx->ev_fork, ctx->stream));
    for (int g = 0; g < ng; ++g) CK(cudaStreamWaitEvent(ctx->gstream[g], ctx->ev_fork, 0));
    for (int b = 0; b < p0.n_stack; ++b) {
        const int i = (dir > 0) ? b : p0.n_stack - 1 - b;
        const int l0 = i * p0.n_stab, l1 = std::min(l0 + p0.n_stab, p0.nt);
        for (int g = 0; g < ng; ++g) {
            Params p = p0;
            // group boundaries at multiples of the propagate kernel's walkers per CTA
            const int per = ((p0.W + ng - 1) / ng + dq::kPropWPC - 1) / dq::kPropWPC * dq::kPropWPC;
            p.w0 = std::min(g * per, p0.W);
            p.w1 = std::min(p.w0 + per, p0.W);
            if (p.w0 >= p.w1) continue;
            if (p.rng_mode != DQMCB_RNG_TAPE) p.visit_base = 0;
            // visits are consumed in sweep order: slice l is visit (l) forward, (nt-1-l) backward
            const unsigned long long vfirst = v0 + (unsigned long long)((dir > 0) ? l0 : (p0.nt - l1));
            cudaStream_t gs = ctx->gstream[g];
#ifdef DQMCB_DEBUG_GROUPS  // development aid (compile-time): synchronise after every launch to localise a faulting kernel
#define DQ_GROUP_CHECK(what)                                                                                                          \
    do {                                                                                                                              \
        const cudaError_t e_ = cudaStreamSynchronize(gs);                                                                             \
        if (e_ != cudaSuccess) {                                                                                                      \
            fprintf(stderr, "[dqmcb debug] %s failed: %s dir=%d block=%d group=%d w=[%d,%d)\n", what, cudaGetErrorString(e_), dir, i, g, p.w0, p.w1); \
            return DQMCB_ECUDA;                                                                                                       \
        }                                                                                                                             \
    } while (0)
#else
#define DQ_GROUP_CHECK(what) do { } while (0)
#endif
            if (fast64_propagate(ctx, p, (dir > 0) ? l0 : l1 - 1, l1 - l0, dir, vfirst, 3, gs)) return DQMCB_ECUDA;
            DQ_GROUP_CHECK("propagate");
            fast_stabilize(ctx, (dir > 0) ? l1 - 1 : l0, dir, p.w0, p.w1, gs);
            DQ_GROUP_CHECK("stabilize");
#undef DQ_GROUP_CHECK
            ctx->n_launches += 2;
        }
    }
    CK(cudaGetLastError());
    for (int g = 0; g < ng; ++g) {
        CK(cudaEventRecord(ctx->gdone[g], ctx->gstream[g]));
        CK(cudaStreamWaitEvent(ctx->stream, ctx->gdone[g], 0));
    }
    return DQMCB_OK;
}

int dqmcb_sweep_0_to_beta(dqmcb_ctx* ctx) {
    NEED(ctx);
    if (!ctx->green_ready) {
        ctx->err = "sweep before init_stacks/init_greenfunctions";
        return DQMCB_ESTATE;
    }
    if (ctx->n_groups > 1 && !ctx->timing && fast64_available(ctx)) return grouped_sweep(ctx, +1);
    const Params& p = ctx->p;
    for (int i = 0; i < p.n_stack; ++i) {
        const int l0 = i * p.n_stab, l1 = std::min(l0 + p.n_stab, p.nt);
        if (int rc = launch_propagate(ctx, l0, l1 - l0, +1, 3)) return rc;
        if (int rc = launch_stabilize(ctx, l1 - 1, +1)) return rc;
    }
    return DQMCB_OK;
}
int dqmcb_sweep_beta_to_0(dqmcb_ctx* ctx) {
    NEED(ctx);
    if (!ctx->green_ready) {
        ctx->err = "sweep before init_stacks/init_greenfunctions";
        return DQMCB_ESTATE;
    }
    if (ctx->n_groups > 1 && !ctx->timing && fast64_available(ctx)) return grouped_sweep(ctx, -1);
    const Params& p = ctx->p;
    for (int i = p.n_stack - 1; i >= 0; --i) {
        const int l0 = i * p.n_stab, l1 = std::min(l0 + p.n_stab, p.nt);
        if (int rc = launch_propagate(ctx, l1 - 1, l1 - l0, -1, 3)) return rc;
        if (int rc = launch_stabilize(ctx, l0, -1)) return rc;
    }
    return DQMCB_OK;
}

// ---- state access -------------------------------------------------------------------------------------------
int dqmcb_get_G(dqmcb_ctx* ctx, double* G) {
    NEED(ctx);
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemcpy(G, ctx->p.G, (size_t)ctx->p.W * ctx->p.N * ctx->p.N * 8, cudaMemcpyDeviceToHost));
    return dqmcb_synchronize(ctx);
}
int dqmcb_set_G(dqmcb_ctx* ctx, const double* G) {
    NEED(ctx);
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemcpy(ctx->p.G, G, (size_t)ctx->p.W * ctx->p.N * ctx->p.N * 8, cudaMemcpyHostToDevice));
    ctx->green_ready = true;
    return DQMCB_OK;
}
int dqmcb_get_ldr(dqmcb_ctx* ctx, int i, double* L, double* d, double* R) {
    NEED(ctx);
    const Params& p = ctx->p;
    if (i < 0 || i >= p.n_stack) {
        ctx->err = "LDR Stack index out of bounds";
        return DQMCB_EINVAL;
    }
    CK(cudaStreamSynchronize(ctx->stream));
    const size_t nn = (size_t)p.N * p.N;
    for (int w = 0; w < p.W; ++w) {
        const size_t so = (size_t)w * p.n_stack + i;
        if (L) CK(cudaMemcpy(L + w * nn, p.stackL + so * nn, nn * 8, cudaMemcpyDeviceToHost));
        if (R) CK(cudaMemcpy(R + w * nn, p.stackR + so * nn, nn * 8, cudaMemcpyDeviceToHost));
        if (d) CK(cudaMemcpy(d + (size_t)w * p.N, p.stackD + so * p.N, (size_t)p.N * 8, cudaMemcpyDeviceToHost));
    }
    return dqmcb_synchronize(ctx);
}
int dqmcb_set_ldr(dqmcb_ctx* ctx, int i, const double* L, const double* d, const double* R) {
    NEED(ctx);
    const Params& p = ctx->p;
    if (i < 0 || i >= p.n_stack) {
        ctx->err = "LDR Stack index out of bounds";
        return DQMCB_EINVAL;
    }
    CK(cudaStreamSynchronize(ctx->stream));
    const size_t nn = (size_t)p.N * p.N;
    for (int w = 0; w < p.W; ++w) {
        const size_t so = (size_t)w * p.n_stack + i;
        CK(cudaMemcpy(p.stackL + so * nn, L + w * nn, nn * 8, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(p.stackR + so * nn, R + w * nn, nn * 8, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(p.stackD + so * p.N, d + (size_t)w * p.N, (size_t)p.N * 8, cudaMemcpyHostToDevice));
    }
    ctx->stacks_ready = true;
    return DQMCB_OK;
}
int dqmcb_get_logdet(dqmcb_ctx* ctx, double* out) {
    NEED(ctx);
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemcpy(out, ctx->p.logdet, (size_t)ctx->p.W * 8, cudaMemcpyDeviceToHost));
    return DQMCB_OK;
}
int dqmcb_get_stats(dqmcb_ctx* ctx, double* acc_rate, double* max_err, double* mean_err) {
    NEED(ctx);
    const Params& p = ctx->p;
    CK(cudaStreamSynchronize(ctx->stream));
    std::vector<unsigned long long> acc(p.W), cnt(p.W);
    std::vector<double> emax(p.W), esum(p.W);
    CK(cudaMemcpy(acc.data(), p.accepted, (size_t)p.W * 8, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(cnt.data(), p.err_cnt, (size_t)p.W * 8, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(emax.data(), p.err_max, (size_t)p.W * 8, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(esum.data(), p.err_sum, (size_t)p.W * 8, cudaMemcpyDeviceToHost));
    for (int w = 0; w < p.W; ++w) {
        // acc_rate_ accumulates accepted/N/nt per slice (source/dqmc.cpp:366, update.cpp:31)
        if (acc_rate) acc_rate[w] = (double)acc[w] / ((double)p.N * (double)p.nt);
        if (max_err) max_err[w] = emax[w];
        if (mean_err) mean_err[w] = cnt[w] ? esum[w] / (double)cnt[w] : 0.0;
    }
    return DQMCB_OK;
}
int dqmcb_reset_stats(dqmcb_ctx* ctx) {
    NEED(ctx);
    const Params& p = ctx->p;
    CK(cudaMemsetAsync(p.accepted, 0, (size_t)p.W * 8, ctx->stream));
    CK(cudaMemsetAsync(p.err_max, 0, (size_t)p.W * 8, ctx->stream));
    CK(cudaMemsetAsync(p.err_sum, 0, (size_t)p.W * 8, ctx->stream));
    CK(cudaMemsetAsync(p.err_cnt, 0, (size_t)p.W * 8, ctx->stream));
    return DQMCB_OK;
}
void* dqmcb_device_ptr(dqmcb_ctx* ctx, int which) {
    if (!ctx) return nullptr;
    switch (which) {
        case DQMCB_PTR_FIELDS: return ctx->p.fields;
        case DQMCB_PTR_G: return ctx->p.G;
        case DQMCB_PTR_LOGDET: return ctx->p.logdet;
        case DQMCB_PTR_OBS: return ctx->d_obs;
        case DQMCB_PTR_STACK_L: return ctx->p.stackL;
        case DQMCB_PTR_STACK_D: return ctx->p.stackD;
        case DQMCB_PTR_STACK_R: return ctx->p.stackR;
        default: return nullptr;
    }
}

// ---- measurements ------------------------------------------------------------------------------------------------
int dqmcb_measure_equal_time(dqmcb_ctx* ctx, double* scalars, double* corr) {
    NEED(ctx);
    const Params& p = ctx->p;
    {
        TimeScope ts(ctx, DQMCB_K_OTHER);
        k_measure<<<p.W, kThreads, 0, ctx->stream>>>(p, ctx->d_obs, corr ? ctx->d_scratchNN : nullptr);
    }
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(ctx->stream));
    std::vector<double> o((size_t)p.W * 4);
    CK(cudaMemcpy(o.data(), ctx->d_obs, o.size() * 8, cudaMemcpyDeviceToHost));
    if (scalars)
        for (int w = 0; w < p.W; ++w)
            for (int k = 0; k < 3; ++k) scalars[w * 3 + k] = o[w * 4 + k];
    if (corr) CK(cudaMemcpy(corr, ctx->d_scratchNN, (size_t)p.W * p.N * p.N * 8, cudaMemcpyDeviceToHost));
    return DQMCB_OK;
}
// kinetic energy per site and displacement-folded Sz-Sz correlation of every walker (no reference counterpart: measure.cuh)
int dqmcb_measure_kinetic_spin(dqmcb_ctx* ctx, int L1, int L2, double t, double* ekin, double* szz_r) {
    NEED(ctx);
    const Params& p = ctx->p;
    if (L1 <= 0 || L2 <= 0 || L1 * L2 != p.N || !ekin || !szz_r) {
        ctx->err = "dqmcb_measure_kinetic_spin: L1*L2 must equal n_sites; outputs must not be null";
        return DQMCB_EINVAL;
    }
    double* d_e = ctx->d_obs;           // [W] of the [W][4] scalar scratch
    double* d_s = ctx->d_scratchNN;     // [W][N] of the [W][N*N] matrix scratch
    {
        TimeScope ts(ctx, DQMCB_K_OTHER);
        ctx->n_launches += 1;
        k_kinetic_spin<<<p.W, kThreads, 0, ctx->stream>>>(p, L1, L2, t, d_e, d_s);
    }
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(ekin, d_e, (size_t)p.W * 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(szz_r, d_s, (size_t)p.W * p.N * 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return DQMCB_OK;
}
// ---- measurement bins (include/measurementh5.h:189-274) -----------------------------------------------
int dqmcb_bin_configure(dqmcb_ctx* ctx, int L1, int L2) {
    NEED(ctx);
    const Params& p = ctx->p;
    if (L1 <= 0 || L2 <= 0 || L1 * L2 != p.N) {
        ctx->err = "dqmcb_bin_configure: L1*L2 must equal n_sites";
        return DQMCB_EINVAL;
    }
    const size_t len = 3 + (size_t)p.N;
    if (!ctx->d_binpart) CK(cudaMalloc(&ctx->d_binpart, (size_t)p.W * len * 8));
    if (!ctx->d_binacc) CK(cudaMalloc(&ctx->d_binacc, len * 8));
    CK(cudaMemsetAsync(ctx->d_binacc, 0, len * 8, ctx->stream));
    ctx->bin_L1 = L1;
    ctx->bin_L2 = L2;
    ctx->bin_samples = 0;
    return DQMCB_OK;
}
int dqmcb_bin_measure(dqmcb_ctx* ctx) {
    NEED(ctx);
    const Params& p = ctx->p;
    if (!ctx->d_binacc) {
        ctx->err = "dqmcb_bin_measure: call dqmcb_bin_configure first";
        return DQMCB_ESTATE;
    }
    const int len = 3 + p.N;
    {
        TimeScope ts(ctx, DQMCB_K_OTHER);
        k_bin_measure<<<p.W, kThreads, 0, ctx->stream>>>(p, ctx->bin_L1, ctx->bin_L2, ctx->d_binpart);
        k_bin_reduce<<<(len + 127) / 128, 128, 0, ctx->stream>>>(p.W, len, ctx->d_binpart, ctx->d_binacc);
    }
    ctx->n_launches += 2;
    CK(cudaGetLastError());
    ctx->bin_samples += p.W;
    return DQMCB_OK;
}
int dqmcb_bin_fetch(dqmcb_ctx* ctx, double* scalar_sums, double* chi_r_sum, long long* n_samples, int reset) {
    NEED(ctx);
    const Params& p = ctx->p;
    if (!ctx->d_binacc) {
        ctx->err = "dqmcb_bin_fetch: call dqmcb_bin_configure first";
        return DQMCB_ESTATE;
    }
    const size_t len = 3 + (size_t)p.N;
    std::vector<double> h(len);
    CK(cudaMemcpyAsync(h.data(), ctx->d_binacc, len * 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (scalar_sums)
        for (int k = 0; k < 3; ++k) scalar_sums[k] = h[k];
    if (chi_r_sum)
        for (int k = 0; k < p.N; ++k) chi_r_sum[k] = h[3 + k];
    if (n_samples) *n_samples = ctx->bin_samples;
    if (reset) {
        CK(cudaMemsetAsync(ctx->d_binacc, 0, len * 8, ctx->stream));
        ctx->bin_samples = 0;
    }
    return DQMCB_OK;
}

// stand-alone transforms on host buffers (transform::chi_site_to_chi_r / chi_r_to_chi_k)
static int run_transform(int device, int L1, int L2, int n_mats, const double* in, size_t in_per, double* out, size_t out_per,
                         int which) {
    if (L1 <= 0 || L2 <= 0 || n_mats < 0 || (n_mats > 0 && (!in || !out))) return DQMCB_EINVAL;
    if (n_mats == 0) return DQMCB_OK;
    if (cudaSetDevice(device) != cudaSuccess) return DQMCB_ECUDA;
    double *d_in = nullptr, *d_out = nullptr;
    int rc = DQMCB_OK;
    if (cudaMalloc(&d_in, in_per * n_mats * 8) != cudaSuccess || cudaMalloc(&d_out, out_per * n_mats * 8) != cudaSuccess) rc = DQMCB_ECUDA;
    if (rc == DQMCB_OK && cudaMemcpy(d_in, in, in_per * n_mats * 8, cudaMemcpyHostToDevice) != cudaSuccess) rc = DQMCB_ECUDA;
    if (rc == DQMCB_OK) {
        if (which == 0)
            k_chi_site_to_chi_r<<<n_mats, 128>>>(L1, L2, d_in, d_out);
        else
            k_chi_r_to_chi_k<<<n_mats, 128>>>(L1, L2, d_in, d_out);
        if (cudaGetLastError() != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) rc = DQMCB_ECUDA;
    }
    if (rc == DQMCB_OK && cudaMemcpy(out, d_out, out_per * n_mats * 8, cudaMemcpyDeviceToHost) != cudaSuccess) rc = DQMCB_ECUDA;
    cudaFree(d_in);
    cudaFree(d_out);
    if (rc != DQMCB_OK) g_create_err = "dqmcb transform: CUDA failure (no device?)";
    return rc;
}
int dqmcb_chi_site_to_chi_r(int device, int L1, int L2, int n_mats, const double* chi_site, double* chi_r) {
    const size_t n = (size_t)L1 * L2;
    return run_transform(device, L1, L2, n_mats, chi_site, n * n, chi_r, n, 0);
}
int dqmcb_chi_r_to_chi_k(int device, int L1, int L2, int n_mats, const double* chi_r, double* chi_k) {
    const size_t n = (size_t)L1 * L2;
    return run_transform(device, L1, L2, n_mats, chi_r, n, chi_k, 2 * n, 1);
}
int dqmcb_pbc_shortest(int d, int L) { return dq::pbc_shortest(d, L); }

// ---- unequal-time sweep (source/dqmc.cpp:223-280,458-514) ---------------------------------------------
static UneqBufs uneq_bufs(dqmcb_ctx* ctx) { return UneqBufs{ctx->d_uneq, ctx->d_bt0, ctx->d_uneq_err}; }

int dqmcb_sweep_unequal_time(dqmcb_ctx* ctx) {
    NEED(ctx);
    const Params& p = ctx->p;
    if (!ctx->stacks_ready || !ctx->green_ready) {
        ctx->err = "dqmcb_sweep_unequal_time: needs the state after sweep_beta_to_0 (stack + G(0,0))";
        return DQMCB_ESTATE;
    }
    const size_t nn = (size_t)p.N * p.N;
    if (!ctx->d_uneq) {
        CK(cudaMalloc(&ctx->d_uneq, (size_t)p.W * 3 * (p.nt + 1) * nn * 8));
        CK(cudaMalloc(&ctx->d_bt0, (size_t)p.W * (2 * nn + p.N) * 8));
        CK(cudaMalloc(&ctx->d_uneq_err, (size_t)p.W * 8));
    }
    for (int i = 0; i < p.n_stack; ++i) {
        TimeScope ts(ctx, DQMCB_K_OTHER);
        k_uneq_block<<<p.W, kThreads, ctx->smem_stab, ctx->stream>>>(p, uneq_bufs(ctx), i);
        ctx->n_launches += 1;
    }
    CK(cudaGetLastError());
    ctx->uneq_ready = true;
    return DQMCB_OK;
}
#define NEED_UNEQ(ctx)                                                              \
    if (!(ctx)->uneq_ready) {                                                       \
        (ctx)->err = "unequal-time data missing: call dqmcb_sweep_unequal_time first"; \
        return DQMCB_ESTATE;                                                        \
    }
int dqmcb_get_G_unequal(dqmcb_ctx* ctx, int which, int l, double* out) {
    NEED(ctx);
    NEED_UNEQ(ctx);
    const Params& p = ctx->p;
    if (which < 0 || which > 2 || l < 0 || l > p.nt || !out) {
        ctx->err = "dqmcb_get_G_unequal: which in 0..2, l in 0..nt";
        return DQMCB_EINVAL;
    }
    const size_t nn = (size_t)p.N * p.N;
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemcpy2D(out, nn * 8, ctx->d_uneq + ((size_t)which * (p.nt + 1) + l) * nn, (size_t)3 * (p.nt + 1) * nn * 8, nn * 8, p.W,
                    cudaMemcpyDeviceToHost));
    return DQMCB_OK;
}
int dqmcb_get_unequal_err(dqmcb_ctx* ctx, double* err_max) {
    NEED(ctx);
    NEED_UNEQ(ctx);
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemcpy(err_max, ctx->d_uneq_err, (size_t)ctx->p.W * 8, cudaMemcpyDeviceToHost));
    return DQMCB_OK;
}
int dqmcb_half_wrap_unequal(dqmcb_ctx* ctx, int which, int l, double* Gsym) {
    NEED(ctx);
    NEED_UNEQ(ctx);
    const Params& p = ctx->p;
    if (which < 0 || which > 2 || l < 0 || l > p.nt || !Gsym) {
        ctx->err = "dqmcb_half_wrap_unequal: which in 0..2, l in 0..nt";
        return DQMCB_EINVAL;
    }
    k_uneq_half_wrap<<<p.W, kThreads, 0, ctx->stream>>>(p, uneq_bufs(ctx), which, l, ctx->d_scratchNN);
    ctx->n_launches += 1;
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemcpy(Gsym, ctx->d_scratchNN, (size_t)p.W * p.N * p.N * 8, cudaMemcpyDeviceToHost));
    return DQMCB_OK;
}
int dqmcb_measure_unequal_time(dqmcb_ctx* ctx, int L1, int L2, double* green, double* doublon, double* currxx) {
    NEED(ctx);
    NEED_UNEQ(ctx);
    const Params& p = ctx->p;
    if (L1 <= 0 || L2 <= 0 || L1 * L2 != p.N) {
        ctx->err = "dqmcb_measure_unequal_time: L1*L2 must equal n_sites";
        return DQMCB_EINVAL;
    }
    const size_t per = (size_t)p.W * (p.nt + 1) * p.N * p.N;
    double* outs[3] = {green, doublon, currxx};
    double* d[3] = {nullptr, nullptr, nullptr};
    int rc = DQMCB_OK;
    for (int k = 0; k < 3 && rc == DQMCB_OK; ++k)
        if (outs[k] && cudaMalloc(&d[k], per * 8) != cudaSuccess) rc = DQMCB_ECUDA;
    if (rc == DQMCB_OK) {
        k_uneq_obs_site<<<dim3(p.nt + 1, p.W), kThreads, 0, ctx->stream>>>(p, uneq_bufs(ctx), L1, d[0], d[1], d[2]);
        ctx->n_launches += 1;
        if (cudaGetLastError() != cudaSuccess || cudaStreamSynchronize(ctx->stream) != cudaSuccess) rc = DQMCB_ECUDA;
    }
    for (int k = 0; k < 3; ++k) {
        if (rc == DQMCB_OK && outs[k] && cudaMemcpy(outs[k], d[k], per * 8, cudaMemcpyDeviceToHost) != cudaSuccess) rc = DQMCB_ECUDA;
        cudaFree(d[k]);
    }
    if (rc != DQMCB_OK) ctx->err = "dqmcb_measure_unequal_time: CUDA failure";
    return rc;
}
// MeasurementManager::measure for the three unequal-time observables of source/main.cpp:120-122, all walkers,
// folded onto displacements and accumulated on the device (needs dqmcb_bin_configure for L1, L2)
int dqmcb_bin_measure_unequal(dqmcb_ctx* ctx) {
    NEED(ctx);
    NEED_UNEQ(ctx);
    const Params& p = ctx->p;
    if (!ctx->d_binacc) {
        ctx->err = "dqmcb_bin_measure_unequal: call dqmcb_bin_configure first";
        return DQMCB_ESTATE;
    }
    const int len = 3 * (p.nt + 1) * p.N;
    if (!ctx->d_tbinpart) {
        CK(cudaMalloc(&ctx->d_tbinpart, (size_t)p.W * len * 8));
        CK(cudaMalloc(&ctx->d_tbinacc, (size_t)len * 8));
        CK(cudaMemsetAsync(ctx->d_tbinacc, 0, (size_t)len * 8, ctx->stream));
        ctx->tbin_samples = 0;
    }
    k_uneq_bin<<<dim3(p.nt + 1, p.W), 128, 0, ctx->stream>>>(p, uneq_bufs(ctx), ctx->bin_L1, ctx->bin_L2, ctx->d_tbinpart);
    k_bin_reduce<<<(len + 127) / 128, 128, 0, ctx->stream>>>(p.W, len, ctx->d_tbinpart, ctx->d_tbinacc);
    ctx->n_launches += 2;
    CK(cudaGetLastError());
    ctx->tbin_samples += p.W;
    return DQMCB_OK;
}
int dqmcb_bin_fetch_unequal(dqmcb_ctx* ctx, double* chi_r_sums, long long* n_samples, int reset) {
    NEED(ctx);
    const Params& p = ctx->p;
    if (!ctx->d_tbinacc) {
        ctx->err = "dqmcb_bin_fetch_unequal: nothing measured";
        return DQMCB_ESTATE;
    }
    const size_t len = (size_t)3 * (p.nt + 1) * p.N;
    CK(cudaStreamSynchronize(ctx->stream));
    if (chi_r_sums) CK(cudaMemcpy(chi_r_sums, ctx->d_tbinacc, len * 8, cudaMemcpyDeviceToHost));
    if (n_samples) *n_samples = ctx->tbin_samples;
    if (reset) {
        CK(cudaMemsetAsync(ctx->d_tbinacc, 0, len * 8, ctx->stream));
        ctx->tbin_samples = 0;
    }
    return DQMCB_OK;
}

int dqmcb_global_action(dqmcb_ctx* ctx, double* S) {
    NEED(ctx);
    const Params& p = ctx->p;
    k_action<<<p.W, kThreads, 0, ctx->stream>>>(p, ctx->d_obs);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemcpy(S, ctx->d_obs, (size_t)p.W * 8, cudaMemcpyDeviceToHost));
    return DQMCB_OK;
}
int dqmcb_half_wrap(dqmcb_ctx* ctx, double* Gsym) {
    NEED(ctx);
    const Params& p = ctx->p;
    k_half_wrap<<<p.W, kThreads, 0, ctx->stream>>>(p, ctx->d_scratchNN);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemcpy(Gsym, ctx->d_scratchNN, (size_t)p.W * p.N * p.N * 8, cudaMemcpyDeviceToHost));
    return DQMCB_OK;
}

int dqmcb_run_sweeps_host(dqmcb_ctx* ctx, int reinit, const int8_t* fields_in, int n_sweeps, int8_t* fields_out, double* scalars_out,
                          double* G_out) {
    NEED(ctx);
    const Params& p = ctx->p;
    const size_t fsz = (size_t)p.W * p.nt * p.N;
    if (fields_in) {
        CK(cudaMemcpyAsync(p.fields, fields_in, fsz, cudaMemcpyHostToDevice, ctx->stream));
        ctx->stacks_ready = ctx->green_ready = false;
    }
    if (reinit || !ctx->green_ready) {
        if (int rc = dqmcb_init_stacks(ctx)) return rc;
        if (int rc = dqmcb_init_greenfunctions(ctx)) return rc;
    }
    for (int s = 0; s < n_sweeps; ++s) {
        if (int rc = dqmcb_sweep_0_to_beta(ctx)) return rc;
        if (int rc = dqmcb_sweep_beta_to_0(ctx)) return rc;
    }
    if (scalars_out) {
        TimeScope ts(ctx, DQMCB_K_OTHER);
        ctx->n_launches += 1;
        k_measure<<<p.W, kThreads, 0, ctx->stream>>>(p, ctx->d_obs, nullptr);
        CK(cudaGetLastError());
    }
    if (fields_out) CK(cudaMemcpyAsync(fields_out, p.fields, fsz, cudaMemcpyDeviceToHost, ctx->stream));
    if (G_out) CK(cudaMemcpyAsync(G_out, p.G, (size_t)p.W * p.N * p.N * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if (scalars_out) {
        // d_obs rows are [density, doubleOcc, swave, pad]; copy as W x 4 then compact on the host
        std::vector<double> o((size_t)p.W * 4);
        CK(cudaMemcpyAsync(o.data(), ctx->d_obs, o.size() * 8, cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        for (int w = 0; w < p.W; ++w)
            for (int k = 0; k < 3; ++k) scalars_out[w * 3 + k] = o[w * 4 + k];
    }
    return dqmcb_synchronize(ctx);
}

// ---- stand-alone batched stablelinalg ---------------------------------------------------------------------------
static int run_ldr_op(int device, int n, int batch, int op, const double* M, const double* L1, const double* d1, const double* R1,
                      const double* L2, const double* d2, const double* R2, double* Lo, double* dO, double* Ro, double* G,
                      double* logdet) {
    if (n <= 0 || n > 256 || batch <= 0) return DQMCB_EINVAL;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        g_create_err = "no CUDA device: dqmc_b200 has no CPU fallback";
        return DQMCB_ECUDA;
    }
    if (cudaSetDevice(device) != cudaSuccess) return DQMCB_ECUDA;
    Params p;
    memset(&p, 0, sizeof(p));
    p.N = n;
    const size_t smem = plan_stab_smem(p, n);
    p.W = batch;
    const size_t nn = (size_t)n * n;
    std::vector<void*> allocs;
    int rc = DQMCB_OK;
    auto up = [&](const double* h, size_t cnt) -> double* {
        if (!h) return nullptr;
        double* dptr = nullptr;
        if (cudaMalloc(&dptr, cnt * 8) != cudaSuccess) {
            rc = DQMCB_ECUDA;
            return nullptr;
        }
        allocs.push_back(dptr);
        if (cudaMemcpy(dptr, h, cnt * 8, cudaMemcpyHostToDevice) != cudaSuccess) rc = DQMCB_ECUDA;
        return dptr;
    };
    auto mk = [&](bool want, size_t cnt) -> double* {
        if (!want) return nullptr;
        double* dptr = nullptr;
        if (cudaMalloc(&dptr, cnt * 8) != cudaSuccess) {
            rc = DQMCB_ECUDA;
            return nullptr;
        }
        allocs.push_back(dptr);
        return dptr;
    };
    LdrArgs a;
    a.M = up(M, batch * nn);
    a.L1 = up(L1, batch * nn);
    a.d1 = up(d1, (size_t)batch * n);
    a.R1 = up(R1, batch * nn);
    a.L2 = up(L2, batch * nn);
    a.d2 = up(d2, (size_t)batch * n);
    a.R2 = up(R2, batch * nn);
    a.Lo = mk(Lo, batch * nn);
    a.dO = mk(dO, (size_t)batch * n);
    a.Ro = mk(Ro, batch * nn);
    a.G = mk(G, batch * nn);
    a.logdet = mk(logdet, batch);
    p.ws = mk(true, (size_t)batch * 3 * n * p.LD);
    int* fail = nullptr;
    if (cudaMalloc(&fail, sizeof(int)) != cudaSuccess) rc = DQMCB_ECUDA;
    if (rc == DQMCB_OK) {
        cudaMemset(fail, 0, sizeof(int));
        p.fail = fail;
        cudaFuncSetAttribute(k_ldr_op, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k_ldr_op<<<batch, kThreads, smem>>>(p, op, a);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) {
            g_create_err = cudaGetErrorString(e);
            rc = DQMCB_ECUDA;
        }
    }
    if (rc == DQMCB_OK) {
        int f = 0;
        cudaMemcpy(&f, fail, sizeof(int), cudaMemcpyDeviceToHost);
        if (f) rc = DQMCB_ENUMERIC;
        if (Lo) cudaMemcpy(Lo, a.Lo, batch * nn * 8, cudaMemcpyDeviceToHost);
        if (dO) cudaMemcpy(dO, a.dO, (size_t)batch * n * 8, cudaMemcpyDeviceToHost);
        if (Ro) cudaMemcpy(Ro, a.Ro, batch * nn * 8, cudaMemcpyDeviceToHost);
        if (G) cudaMemcpy(G, a.G, batch * nn * 8, cudaMemcpyDeviceToHost);
        if (logdet) cudaMemcpy(logdet, a.logdet, (size_t)batch * 8, cudaMemcpyDeviceToHost);
    }
    for (void* q : allocs) cudaFree(q);
    cudaFree(fail);
    return rc;
}

// ---- allocation-free variant of dqmcb_to_LDR (no counterpart in the reference) ----
// Scratch of k_ldr_op (the three per-matrix work tiles it keeps in global memory when they do not fit in shared memory,
// n > ~96, and the failure flag) lives in a file-scope device array instead of a cudaMalloc'ed buffer: usable where the
// caller may not allocate (CUDA graph capture, serving loops).  Every GPU that loads the library holds these
// DQMCB_NOALLOC_SCRATCH_BYTES until exit.  All calls share the array: calls that overlap on the GPU race.
static __device__ __align__(256) unsigned char g_noalloc_scratch[DQMCB_NOALLOC_SCRATCH_BYTES];
// the array's NAME is a host-side stand-in in host code; its device address is taken once per device, at set-up
static unsigned char* g_noalloc_ptr[64] = {};
static size_t g_noalloc_attr[64] = {};  // dynamic shared-memory limit set for k_ldr_op so far

int dqmcb_to_LDR_noalloc(int device, int n, int batch, const double* dM, double* dL, double* dd, double* dR, void* stream) {
    if (n <= 0 || n > 256 || batch <= 0 || !dM || !dL || !dd || !dR || device < 0 || device >= 64) {
        g_create_err = "dqmcb_to_LDR_noalloc: bad argument (need 0 < n <= 256, batch > 0, device pointers, 0 <= device < 64)";
        return DQMCB_EINVAL;
    }
    Params p;
    memset(&p, 0, sizeof(p));
    p.N = n;
    p.LD = (n % 2 == 0) ? n + 1 : n;
    p.W = batch;
    const size_t tile = (size_t)n * p.LD * 8;
    const size_t sb = ((small_bytes(n) + 15) & ~(size_t)15);
    p.smem_mats_stab = (sb + 3 * tile <= (size_t)kMaxSmem) ? 3 : 0;
    p.smem_stage = (p.smem_mats_stab == 0 && sb + tile <= (size_t)kMaxSmem) ? 1 : 0;
    const size_t smem = sb + (p.smem_mats_stab + p.smem_stage) * tile;
    // layout of the static scratch: [0, 256) failure flag, then batch x 3 work tiles (touched only when they are not in SMEM)
    const size_t need = 256 + (p.smem_mats_stab == 3 ? 0 : (size_t)batch * 3 * tile);
    if (need > (size_t)DQMCB_NOALLOC_SCRATCH_BYTES) {
        g_create_err = "dqmcb_to_LDR_noalloc: batch x n needs " + std::to_string(need) + " bytes of scratch, the static array holds " +
                       std::to_string((size_t)DQMCB_NOALLOC_SCRATCH_BYTES) + " (DQMCB_NOALLOC_SCRATCH_BYTES): split the batch or use dqmcb_to_LDR";
        return DQMCB_EINVAL;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        g_create_err = "no CUDA device: dqmc_b200 has no CPU fallback";
        return DQMCB_ECUDA;
    }
    if (device >= ndev || cudaSetDevice(device) != cudaSuccess) {
        g_create_err = "dqmcb_to_LDR_noalloc: bad device ordinal";
        return DQMCB_ECUDA;
    }
    if (!g_noalloc_ptr[device]) {  // set-up, once per device, with that device current
        unsigned char* ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch);
        if (e != cudaSuccess || !ptr) {
            g_create_err = std::string("cudaGetSymbolAddress(g_noalloc_scratch): ") + cudaGetErrorString(e);
            return DQMCB_ECUDA;
        }
        g_noalloc_ptr[device] = ptr;
    }
    unsigned char* scratch = g_noalloc_ptr[device];
    cudaStream_t st = (cudaStream_t)stream;
    // the array keeps the previous call's contents: the failure flag must start at zero (the tiles are written before read)
    if (cudaMemsetAsync(scratch, 0, 256, st) != cudaSuccess) {
        g_create_err = "dqmcb_to_LDR_noalloc: cudaMemsetAsync failed";
        return DQMCB_ECUDA;
    }
    p.fail = reinterpret_cast<int*>(scratch);
    p.ws = reinterpret_cast<double*>(scratch + 256);
    LdrArgs a;
    memset(&a, 0, sizeof(a));
    a.M = dM;
    a.Lo = dL;
    a.dO = dd;
    a.Ro = dR;
    if (smem > g_noalloc_attr[device]) {  // raise the opt-in dynamic shared-memory limit of the kernel when this size needs more
        if (cudaFuncSetAttribute(k_ldr_op, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
            g_create_err = "dqmcb_to_LDR_noalloc: cudaFuncSetAttribute failed";
            return DQMCB_ECUDA;
        }
        g_noalloc_attr[device] = smem;
    }
    k_ldr_op<<<batch, kThreads, smem, st>>>(p, 0, a);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        g_create_err = std::string("dqmcb_to_LDR_noalloc launch: ") + cudaGetErrorString(e);
        return DQMCB_ECUDA;
    }
    return DQMCB_OK;  // asynchronous on `stream`; the pivoted QR has no numerical failure mode
}

int dqmcb_to_LDR(int device, int n, int batch, const double* M, double* L, double* d, double* R) {
    return run_ldr_op(device, n, batch, 0, M, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, L, d, R, nullptr, nullptr);
}
int dqmcb_mat_mul_ldr(int device, int n, int batch, const double* M, const double* L, const double* d, const double* R, double* Lo,
                      double* dO, double* Ro) {
    return run_ldr_op(device, n, batch, 1, M, L, d, R, nullptr, nullptr, nullptr, Lo, dO, Ro, nullptr, nullptr);
}
int dqmcb_ldr_mul_mat(int device, int n, int batch, const double* L, const double* d, const double* R, const double* M, double* Lo,
                      double* dO, double* Ro) {
    return run_ldr_op(device, n, batch, 2, M, L, d, R, nullptr, nullptr, nullptr, Lo, dO, Ro, nullptr, nullptr);
}
int dqmcb_ldr_mul_ldr(int device, int n, int batch, const double* L1, const double* d1, const double* R1, const double* L2,
                      const double* d2, const double* R2, double* Lo, double* dO, double* Ro) {
    return run_ldr_op(device, n, batch, 3, nullptr, L1, d1, R1, L2, d2, R2, Lo, dO, Ro, nullptr, nullptr);
}
int dqmcb_inv_I_plus_ldr(int device, int n, int batch, const double* L, const double* d, const double* R, double* G, double* logdet) {
    return run_ldr_op(device, n, batch, 4, nullptr, L, d, R, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, G, logdet);
}
int dqmcb_inv_I_plus_ldr_mul_ldr(int device, int n, int batch, const double* L1, const double* d1, const double* R1, const double* L2,
                                 const double* d2, const double* R2, double* G) {
    return run_ldr_op(device, n, batch, 5, nullptr, L1, d1, R1, L2, d2, R2, nullptr, nullptr, nullptr, G, nullptr);
}
int dqmcb_inv_invldr_plus_ldr(int device, int n, int batch, const double* L1, const double* d1, const double* R1, const double* L2,
                              const double* d2, const double* R2, double* G) {
    return run_ldr_op(device, n, batch, 6, nullptr, L1, d1, R1, L2, d2, R2, nullptr, nullptr, nullptr, G, nullptr);
}

int dqmcb_local_update_ratio(dqmcb_ctx* ctx, int l, int site, int new_field, double* R, double* delta) {
    NEED(ctx);
    const Params& p = ctx->p;
    if (int rc = check_l(ctx, l)) return rc;
    if (site < 0 || site >= p.N || new_field < 0 || new_field > 3) {
        ctx->err = "site or field value out of range";
        return DQMCB_EINVAL;
    }
    k_ratio<<<(p.W + 127) / 128, 128, 0, ctx->stream>>>(p, l, site, new_field, ctx->d_scratchNN);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(ctx->stream));
    std::vector<double> o((size_t)2 * p.W);
    CK(cudaMemcpy(o.data(), ctx->d_scratchNN, o.size() * 8, cudaMemcpyDeviceToHost));
    for (int w = 0; w < p.W; ++w) {
        if (R) R[w] = o[2 * w];
        if (delta) delta[w] = o[2 * w + 1];
    }
    return DQMCB_OK;
}
int dqmcb_update_greens_local(dqmcb_ctx* ctx, int walker, double delta, int site) {
    NEED(ctx);
    const Params& p = ctx->p;
    if (walker < 0 || walker >= p.W || site < 0 || site >= p.N) {
        ctx->err = "walker or site out of range";
        return DQMCB_EINVAL;
    }
    k_rank1<<<1, kThreads, 2 * p.N * sizeof(double), ctx->stream>>>(p, walker, delta, site);
    CK(cudaGetLastError());
    return DQMCB_OK;
}
int dqmcb_set_single_field(dqmcb_ctx* ctx, int walker, int l, int site, int value) {
    NEED(ctx);
    const Params& p = ctx->p;
    if (walker < 0 || walker >= p.W || l < 0 || l >= p.nt || site < 0 || site >= p.N || value < 0 || value > 3) {
        ctx->err = "set_single_field: argument out of range";
        return DQMCB_EINVAL;
    }
    const int8_t v = (int8_t)value;
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemcpy(p.fields + ((size_t)walker * p.nt + l) * p.N + site, &v, 1, cudaMemcpyHostToDevice));
    return DQMCB_OK;
}

}  // extern "C"
