// Unequal-time sweep and its observables.
//
// Reference: DQMC::sweep_unequalTime (source/dqmc.cpp:458-514) with
// propagate_unequalTime_GF_forward (:223-246), propagate_Bt0_Bbt (:248-262), stabilize_unequalTime
// (:264-280); Observables::calculate_{greenTau,doublonTau,currxxTau} (source/model.cpp:290-392).
// Runs after sweep_beta_to_0: p.G = Gtt[0], the stack holds the right products B(beta, tau).
//
// Layout: U = [W][3][nt+1][N*N] column-major matrices (which = 0 Gtt, 1 Gt0, 2 G0t); the running
// left product Bt0 = B(tau, 0) is one LDR per walker, bt0 = [W][2*N*N + N] (L, R, d).  One CTA per
// walker, one launch per stabilisation block, built from the generic engine (any N <= 256): the
// unequal-time path is off by default in the reference (examples/parameters.in:18) and is not a
// throughput target; it reuses the wrap products, dev_bbar, dev_to_ldr and the stabilised inverses.
#pragma once
#include "engine.cuh"

namespace dq {

struct UneqBufs {
    double* U;      // [W][3][nt+1][N*N]
    double* bt0;    // [W][2*N*N + N]
    double* err;    // [W] running max of the stabilisation error (check_error, source/dqmc.cpp:317-329)
};

__device__ __forceinline__ double* uneq_mat(const Params& p, const UneqBufs& u, int w, int which, int l) {
    return u.U + ((((size_t)w * 3 + which) * (p.nt + 1)) + l) * (size_t)p.N * p.N;
}

// dst (global, ld n) <- sign * src (tile) [or I - src when one_minus]; returns max |dst_old - new| of this thread
__device__ inline double uneq_install(double* dst, int n, const double* src, int lds, double sign, bool one_minus) {
    double emax = 0.0;
    for (int e = threadIdx.x; e < n * n; e += blockDim.x) {
        const int r = e % n, c = e / n;
        double v = src[r + (size_t)c * lds];
        v = one_minus ? (((r == c) ? 1.0 : 0.0) - v) : sign * v;
        emax = fmax(emax, fabs(dst[e] - v));
        dst[e] = v;
    }
    return emax;
}

__global__ void __launch_bounds__(kThreads) k_uneq_block(Params p, UneqBufs u, int i_stack) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int w = blockIdx.x, n = p.N;
    Work wk;
    carve(wk, p, w, p.smem_mats_stab, smem);
    double *b0 = wk.b[0], *b1 = wk.b[1], *b2 = wk.b[2];
    const int l0 = wk.ld[0], l1 = wk.ld[1], l2 = wk.ld[2];
    const int lb = i_stack * p.n_stab, le = min(lb + p.n_stab, p.nt);
    const int8_t* f = p.fields + (size_t)w * p.nt * n;
    if (i_stack == 0) {  // tau = 0 (source/dqmc.cpp:235-240): G(tau,0) = G(0,0), G(0,tau) = -(I - G(0,0))
        const double* G = p.G + (size_t)w * n * n;
        double *gtt = uneq_mat(p, u, w, 0, 0), *gt0 = uneq_mat(p, u, w, 1, 0), *g0t = uneq_mat(p, u, w, 2, 0);
        for (int e = threadIdx.x; e < n * n; e += blockDim.x) {
            const double v = G[e];
            gtt[e] = v;
            gt0[e] = v;
            g0t[e] = v - ((e % n == e / n) ? 1.0 : 0.0);
        }
        if (threadIdx.x == 0) u.err[w] = 0.0;
        __syncthreads();
    }
    // naive propagation through the block (source/dqmc.cpp:241-243), B_l = diag(ev) expK, B_l^-1 = invexpK diag(iev)
    for (int l = lb; l < le; ++l) {
        __syncthreads();
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            wk.sv[0][i] = p.lut_ev[f[(size_t)l * n + i]];
            wk.sv[1][i] = p.lut_iev[f[(size_t)l * n + i]];
        }
        __syncthreads();
        cta_gemm<false>(b0, l0, p.expK, 1, n, uneq_mat(p, u, w, 0, l), 1, n, n, nullptr, nullptr);
        cta_gemm<false>(uneq_mat(p, u, w, 1, l + 1), n, p.expK, 1, n, uneq_mat(p, u, w, 1, l), 1, n, n, wk.sv[0], nullptr);
        cta_gemm<false>(uneq_mat(p, u, w, 2, l + 1), n, uneq_mat(p, u, w, 2, l), 1, n, p.invexpK, 1, n, n, nullptr, wk.sv[1]);
        __syncthreads();
        cta_gemm<false>(uneq_mat(p, u, w, 0, l + 1), n, b0, 1, l0, p.invexpK, 1, n, n, wk.sv[0], wk.sv[1]);
    }
    __syncthreads();
    const int l = le - 1;  // block end: stabilise
    double* BL = u.bt0 + (size_t)w * (2 * (size_t)n * n + n);
    double* BR = BL + (size_t)n * n;
    double* Bd = BR + (size_t)n * n;
    // propagate_Bt0_Bbt (source/dqmc.cpp:248-262)
    dev_bbar(p, wk, w, i_stack);  // -> b0
    __syncthreads();
    if (i_stack == 0) {
        dev_to_ldr(wk, n, b0, l0, b1, l1, b2, l2, wk.dnew);
        __syncthreads();
        cta_copy(BL, n, b1, l1, n);
        cta_copy(BR, n, b2, l2, n);
    } else {  // mat_mul_ldr (source/stablelinalg.cpp:69-79): (Bbar L) D -> QR ; R <- R_new R_old
        store_d(wk.sv[0], Bd, n);
        __syncthreads();
        cta_gemm<false>(b1, l1, b0, 1, l0, BL, 1, n, n, nullptr, wk.sv[0]);
        dev_to_ldr(wk, n, b1, l1, b0, l0, b2, l2, wk.dnew);
        __syncthreads();
        cta_gemm<false>(b1, l1, b2, 1, l2, BR, 1, n, n, nullptr, nullptr);
        __syncthreads();
        cta_copy(BL, n, b0, l0, n);
        cta_copy(BR, n, b1, l1, n);
    }
    store_d(Bd, wk.dnew, n);
    __syncthreads();
    // stabilize_unequalTime (source/dqmc.cpp:264-280)
    double emax = 0.0;
    double *gtt = uneq_mat(p, u, w, 0, l + 1), *gt0 = uneq_mat(p, u, w, 1, l + 1), *g0t = uneq_mat(p, u, w, 2, l + 1);
    if (l == p.nt - 1) {
        const double ld_ = dev_inv_I_plus_ldr(p, wk, n, BL, n, Bd, BR, n);  // -> b1
        if (threadIdx.x == 0) p.logdet[w] = ld_;
        emax = fmax(emax, uneq_install(gtt, n, b1, l1, 1.0, false));   // G(beta,beta) = G(0,0)
        emax = fmax(emax, uneq_install(gt0, n, b1, l1, 1.0, true));    // G(beta,0) = I - G(0,0)
        emax = fmax(emax, uneq_install(g0t, n, b1, l1, -1.0, false));  // G(0,beta) = -G(0,0)
    } else {
        const int j = i_stack + 1;  // Bbt = B(beta, tau) = stack[i+1]
        dev_inv_I_plus_ldr_mul_ldr(p, wk, n, BL, Bd, BR, stkL(p, w, j), stkD(p, w, j), stkR(p, w, j), n, 0);
        emax = fmax(emax, uneq_install(gtt, n, b2, l2, 1.0, false));
        __syncthreads();
        dev_inv_I_plus_ldr_mul_ldr(p, wk, n, BL, Bd, BR, stkL(p, w, j), stkD(p, w, j), stkR(p, w, j), n, 1);
        emax = fmax(emax, uneq_install(gt0, n, b2, l2, 1.0, false));
        __syncthreads();
        dev_inv_I_plus_ldr_mul_ldr(p, wk, n, stkL(p, w, j), stkD(p, w, j), stkR(p, w, j), BL, Bd, BR, n, 1);
        emax = fmax(emax, uneq_install(g0t, n, b2, l2, -1.0, false));
    }
    emax = cta_reduce_max(emax, wk.red);
    if (threadIdx.x == 0) u.err[w] = fmax(u.err[w], emax);
}

// DQMC::half_warp for the unequal-time arrays (source/dqmc.cpp:303-312): out[w] = invexpKhalf X[which][l] expKhalf
__global__ void __launch_bounds__(kThreads) k_uneq_half_wrap(Params p, UneqBufs u, int which, int l, double* out) {
    const int w = blockIdx.x, n = p.N;
    double* gws = p.ws + (size_t)w * 3 * n * p.LD;
    cta_gemm<false>(gws, n, p.invexpK_half, 1, n, uneq_mat(p, u, w, which, l), 1, n, n, nullptr, nullptr);
    __syncthreads();
    cta_gemm<false>(out + (size_t)w * n * n, n, gws, 1, n, p.expK_half, 1, n, n, nullptr, nullptr);
}

// element (i, j) of greenTau / doublonTau / currxxTau at one tau (source/model.cpp:290-392; both spin
// species share G in the attractive model: "up + dn" = 2x, "up * dn" = square).  X(r, c) = X[r + c*n].
struct TauObs {
    double green, doublon, currxx;
};
__device__ __forceinline__ TauObs tau_obs(int i, int j, int n, int L1, const double* Gtt, const double* G00, const double* Gt0,
                                          const double* G0t) {
    TauObs o;
    const double g = Gt0[i + (size_t)j * n];
    o.green = g + g;
    o.doublon = g * g;
    const int ix = (i / L1) * L1 + (i % L1 + 1) % L1;  // lat.site_neighbors(i, {1,0}, 0), include/lattice.h:100-107
    const int jx = (j / L1) * L1 + (j % L1 + 1) % L1;
#define DQ_X(M, r, c) M[(r) + (size_t)(c) * n]
    const double d1i = DQ_X(Gtt, ix, i) + DQ_X(Gtt, ix, i), d2i = DQ_X(Gtt, i, ix) + DQ_X(Gtt, i, ix);
    const double d1j = DQ_X(G00, jx, j) + DQ_X(G00, jx, j), d2j = DQ_X(G00, j, jx) + DQ_X(G00, j, jx);
    const double c1 = DQ_X(G0t, jx, i) * DQ_X(Gt0, ix, j) + DQ_X(G0t, jx, i) * DQ_X(Gt0, ix, j);
    const double c2 = DQ_X(G0t, j, i) * DQ_X(Gt0, ix, jx) + DQ_X(G0t, j, i) * DQ_X(Gt0, ix, jx);
    const double c3 = DQ_X(G0t, jx, ix) * DQ_X(Gt0, i, j) + DQ_X(G0t, jx, ix) * DQ_X(Gt0, i, j);
    const double c4 = DQ_X(G0t, j, ix) * DQ_X(Gt0, i, jx) + DQ_X(G0t, j, ix) * DQ_X(Gt0, i, jx);
#undef DQ_X
    const double t1 = d1i * d1j - c1, t2 = d1i * d2j - c2, t3 = d2i * d1j - c3, t4 = d2i * d2j - c4;
    o.currxx = -(t1 - t2 - t3 + t4);
    return o;
}

// site-indexed cubes, one CTA per (tau, walker): out_x[w][tau][i + j*n]; any output may be null
__global__ void __launch_bounds__(kThreads) k_uneq_obs_site(Params p, UneqBufs u, int L1, double* out_g, double* out_d, double* out_c) {
    const int tau = blockIdx.x, w = blockIdx.y, n = p.N;
    const double *Gtt = uneq_mat(p, u, w, 0, tau), *G00 = uneq_mat(p, u, w, 0, 0), *Gt0 = uneq_mat(p, u, w, 1, tau),
                 *G0t = uneq_mat(p, u, w, 2, tau);
    const size_t off = ((size_t)w * (p.nt + 1) + tau) * n * n;
    for (int e = threadIdx.x; e < n * n; e += blockDim.x) {
        const TauObs o = tau_obs(e % n, e / n, n, L1, Gtt, G00, Gt0, G0t);
        if (out_g) out_g[off + e] = o.green;
        if (out_d) out_d[off + e] = o.doublon;
        if (out_c) out_c[off + e] = o.currxx;
    }
}

// folded straight onto displacements (transform::chi_site_to_chi_r, include/measurementh5.h:20-66), one CTA per
// (tau, walker): part[w][obs][tau][bx + L1*by], obs = 0 greenTau, 1 doublonTau, 2 currxxTau
__global__ void __launch_bounds__(kThreads) k_uneq_bin(Params p, UneqBufs u, int L1, int L2, double* __restrict__ part) {
    const int tau = blockIdx.x, w = blockIdx.y, n = p.N, nt1 = p.nt + 1;
    const double *Gtt = uneq_mat(p, u, w, 0, tau), *G00 = uneq_mat(p, u, w, 0, 0), *Gt0 = uneq_mat(p, u, w, 1, tau),
                 *G0t = uneq_mat(p, u, w, 2, tau);
    double* out = part + (size_t)w * 3 * nt1 * n;
    for (int b = threadIdx.x; b < n; b += blockDim.x) {
        const int bx = b % L1, by = b / L1;
        double ag = 0.0, ad = 0.0, ac = 0.0;
        for (int i = 0; i < n; ++i) {
            const TauObs o = tau_obs(i, site_at_bin(i, bx, by, L1, L2), n, L1, Gtt, G00, Gt0, G0t);
            ag += o.green / n;
            ad += o.doublon / n;
            ac += o.currxx / n;
        }
        out[((size_t)0 * nt1 + tau) * n + b] = ag;
        out[((size_t)1 * nt1 + tau) * n + b] = ad;
        out[((size_t)2 * nt1 + tau) * n + b] = ac;
    }
}

}  // namespace dq
