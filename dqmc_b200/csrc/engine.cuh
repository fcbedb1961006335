// engine.cuh -- generic-N device code of the sweep engine (one CTA per walker).
//
// Device restatement of the reference's equal-time sweep (source/dqmc.cpp, source/update.cpp,
// source/model.cpp, source/stablelinalg.cpp) on top of the CTA-level primitives in linalg.cuh.
// Works for any N; tiles live in shared memory when they fit, otherwise in an L2-resident
// global workspace.  The N = 64 fast path (DMMA wrap, register-resident flips) lives in
// fast64.cuh and is parity-tested against this path as well as against the oracle.
#pragma once
#include "dqmcb_internal.cuh"
#include "linalg.cuh"
#include "gemm_tma.cuh"

namespace dq {

// --------------------------------------------------------------------------------------------------
// per-CTA workspace carved out of dynamic shared memory (+ global fallback for the big tiles)
// --------------------------------------------------------------------------------------------------
constexpr int kRhsPanelCols = 24;  // right-hand-side panel of the N = 144 solves: 144 = 6 x 24, 148 x 24 doubles = 28 KB beside the staging tile

struct Work {
    double* b[3];  // N x LD tiles
    int ld[3];
    double* stage;  // one shared-memory tile for the matrix being factorised when b[] live in global memory (N = 144)
    int lds;
    double *tau, *vn1, *vn2, *sv[4], *dnew, *prow, *jrow, *red;
    int *jpvt, *ipiv, *info;
    unsigned char* arena;  // operand panels of cta_gemm_tma256 (N = 256), else null
    double* rhs_panel;     // 24-column right-hand-side panel of cta_lu_solve_panels (N = 144), else null
    int rhs_ld;
};

__host__ __device__ inline size_t small_bytes(int N) {
    // tau vn1 vn2 sv0-3 dnew prow jrow = 10 double vectors, red[64], jpvt ipiv int vectors, info
    return (size_t)(10 * N + 64) * sizeof(double) + (size_t)(2 * N + 4) * sizeof(int);
}

__device__ inline void carve(Work& wk, const Params& p, int walker, int nmats_smem, unsigned char* smem) {
    const int N = p.N;
    double* s = reinterpret_cast<double*>(smem);
    wk.tau = s;
    wk.vn1 = s + N;
    wk.vn2 = s + 2 * N;
    for (int i = 0; i < 4; ++i) wk.sv[i] = s + (3 + i) * N;
    wk.dnew = s + 7 * N;
    wk.prow = s + 8 * N;
    wk.jrow = s + 9 * N;
    wk.red = s + 10 * N;
    int* ip = reinterpret_cast<int*>(s + 10 * N + 64);
    wk.jpvt = ip;
    ip += N;
    wk.ipiv = ip;
    ip += N;
    wk.info = ip;
    double* tiles = reinterpret_cast<double*>(smem + ((small_bytes(N) + 15) & ~(size_t)15));
    wk.stage = (nmats_smem == 0 && p.smem_stage) ? tiles : nullptr;
    wk.lds = p.LD;
    wk.arena = (nmats_smem == 0 && !p.smem_stage && p.smem_arena) ? reinterpret_cast<unsigned char*>(tiles) : nullptr;
    wk.rhs_ld = N + 4;
    wk.rhs_panel = (wk.stage && p.smem_rhs) ? tiles + (size_t)N * p.LD : nullptr;
    double* gws = p.ws + (size_t)walker * 3 * N * p.LD;
    for (int i = 0; i < 3; ++i) {
        if (i < nmats_smem) {
            wk.b[i] = tiles + (size_t)i * N * p.LD;
            wk.ld[i] = p.LD;
        } else {
            wk.b[i] = gws + (size_t)i * N * p.LD;
            wk.ld[i] = p.LD;
        }
    }
}

// Product of the stabilisation kernels: the TMA-staged kernel when the walker has its shared-memory arena (N = 256, both
// operands column-major with 16-byte-aligned columns), else the register-ring / FMA product of linalg.cuh.
template <bool ACC>
__device__ __forceinline__ void dev_gemm(const Work& wk, double* C, int ldc, const double* A, int rsA, int csA, const double* B, int rsB,
                                         int csB, int n, const double* rs, const double* cs) {
    if (wk.arena && n == kGT_N && blockDim.x == 256 && rsA == 1 && rsB == 1 && !(csA & 1) && !(csB & 1) &&
        !((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(B)) & 15))
        cta_gemm_tma256(C, ldc, A, csA, B, csB, rs, cs, ACC, wk.arena);
    else
        cta_gemm<ACC>(C, ldc, A, rsA, csA, B, rsB, csB, n, rs, cs);
}

// shared-memory staging of cta_lu_blocked's panel and block row: the operand arena, when the walker has one (N = 256)
__device__ __forceinline__ double* lu_staging(const Work& wk) { return wk.arena ? reinterpret_cast<double*>(wk.arena + 64) : nullptr; }

// Solve with a cta_lu factorisation on a right-hand side written with its rows already interchanged.
__device__ __forceinline__ void dev_lu_solve_prepermuted(const Work& wk, double* A, int lda, int n, double* Bm, int ldb) {
    if (wk.arena && n == kGT_N && blockDim.x == 256 && !(lda & 1) && !(ldb & 1) &&
        !((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(Bm)) & 15))
        cta_lu_solve_tma256(A, lda, Bm, ldb, wk.arena);
    else if (wk.rhs_panel && (n & 15) == 0 && n % kRhsPanelCols == 0)
        cta_lu_solve_panels<kRhsPanelCols>(A, lda, n, Bm, ldb, wk.rhs_panel, wk.rhs_ld, wk.prow);
    else
        cta_lu_solve(A, lda, n, wk.ipiv, Bm, 1, ldb, true);
}

__device__ inline double cta_reduce_sum(double v, double* red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    double t = 0.0;
    for (int i = 0; i < nw; ++i) t += red[i];
    return t;
}
__device__ inline double cta_reduce_max(double v, double* red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    double t = red[0];
    for (int i = 1; i < nw; ++i) t = fmax(t, red[i]);
    return t;
}

// --------------------------------------------------------------------------------------------------
// stablelinalg::to_LDR (source/stablelinalg.cpp:35-55).  A is factorised in place (destroyed);
// Q -> Qo, D^-1 R P^T -> Ro (both distinct from A), |diag R| -> dout (shared vector).
// --------------------------------------------------------------------------------------------------
// Factorisations sweep the whole matrix once per column: when the tiles live in global memory (N = 144)
// the matrix is first staged into the single shared-memory tile that fits.
__device__ inline void stage_in(Work& wk, int n, double*& F, int& ldf) {
    if (wk.stage && F != wk.stage) {
        __syncthreads();
        cta_copy(wk.stage, wk.lds, F, ldf, n);
        __syncthreads();
        F = wk.stage;
        ldf = wk.lds;
    }
}

// (benchmarks/qrdev compiles this with DQ_TOLDR_PROF to stamp the phases; no code otherwise)
#ifdef DQ_TOLDR_PROF
__device__ long long g_toldr_prof[8];
#define DQ_TL(k) do { __syncthreads(); if (threadIdx.x == 0 && blockIdx.x == 0) g_toldr_prof[k] = clock64(); } while (0)
#else
#define DQ_TL(k) do { } while (0)
#endif
__device__ void dev_to_ldr(Work& wk, int n, double* A, int lda, double* Qo, int ldq, double* Ro, int ldr, double* dout) {
    __syncthreads();
    DQ_TL(0);
    stage_in(wk, n, A, lda);
    DQ_TL(1);
    int* list = reinterpret_cast<int*>(wk.jrow);  // n doubles = 2 n ints: active-column list and its inverse map
    cta_qrp(A, lda, n, wk.tau, wk.jpvt, wk.vn1, wk.vn2, list, list + n, wk.prow);  // prow: second buffer of the partial norms
    DQ_TL(2);
    // column jpvt[j] of A holds R(0..j, j) (and the reflector below): d_j = |R(j,j)|, and R_final = (D^-1 R) P^T puts
    // that column back at index jpvt[j] -- where it already is
    for (int j = threadIdx.x; j < n; j += blockDim.x) {
        const int c = wk.jpvt[j];
        const double dj = fabs(A[j + (size_t)c * lda]);
        dout[j] = dj;
        wk.vn1[j] = 1.0 / dj;  // D_inv_vec (vn1 is free after the factorisation)
        wk.ipiv[c] = j;        // step at which column c was chosen (scratch)
    }
    __syncthreads();
    for (int e = threadIdx.x; e < n * n; e += blockDim.x) {
        const int i = e % n, c = e / n;
        Ro[i + (size_t)c * ldr] = (i <= wk.ipiv[c]) ? wk.vn1[i] * A[i + (size_t)c * lda] : 0.0;
    }
    DQ_TL(3);
    cta_form_q(Qo, ldq, A, lda, wk.tau, n, wk.jpvt);
    __syncthreads();
    DQ_TL(4);
}

// split d into D_large^-1 (or D_large) and D_small (source/stablelinalg.cpp:100-108)
__device__ inline void split_d(const double* d, int n, double* dl_inv, double* ds) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const double v = d[i];
        if (v >= 1.0) {
            dl_inv[i] = 1.0 / v;
            ds[i] = 1.0;
        } else {
            dl_inv[i] = 1.0;
            ds[i] = v;
        }
    }
}

__device__ inline void set_diag(double* B, int ld, int n, const double* dvec) {
    for (int e = threadIdx.x; e < n * n; e += blockDim.x) {
        const int r = e % n, c = e / n;
        B[r + (size_t)c * ld] = (r == c) ? dvec[r] : 0.0;
    }
}

__device__ inline void note_fail(const Params& p, const Work& wk, int code) {
    if (threadIdx.x == 0 && *wk.info != 0) atomicCAS(p.fail, 0, code);
}

// --------------------------------------------------------------------------------------------------
// stablelinalg::inv_I_plus_ldr (source/stablelinalg.cpp:94-126).  Result G in wk.b[1]; returns logdet.
// --------------------------------------------------------------------------------------------------
__device__ double dev_inv_I_plus_ldr(const Params& p, Work& wk, int n, const double* L, int ldl, const double* d, const double* R,
                                     int ldr) {
    double* b0 = wk.b[0];
    double* b1 = wk.b[1];
    double* b2 = wk.b[2];
    const int l0 = wk.ld[0], l1 = wk.ld[1], l2 = wk.ld[2];
    __syncthreads();
    split_d(d, n, wk.sv[0], wk.sv[1]);
    cta_copy(b0, l0, R, ldr, n);
    __syncthreads();
    double part = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) part += (d[i] >= 1.0) ? log(d[i]) : 0.0;  // sum log D_large
    const double logdl = cta_reduce_sum(part, wk.red);
    double* f0 = b0;
    int lf0 = l0;
    stage_in(wk, n, f0, lf0);
    cta_lu(f0, lf0, n, wk.ipiv, wk.prow, wk.jrow, wk.info, lu_staging(wk));
    note_fail(p, wk, 10);
    set_diag(b1, l1, n, wk.sv[0]);
    __syncthreads();
    cta_lu_solve(f0, lf0, n, wk.ipiv, b1, 1, l1);  // X = R^-1 D_large^-1
    __syncthreads();
    for (int e = threadIdx.x; e < n * n; e += blockDim.x) {  // M = X + L D_small ; b0 <- M^T
        const int r = e % n, c = e / n;
        const double m = b1[r + (size_t)c * l1] + L[r + (size_t)c * ldl] * wk.sv[1][c];
        b2[r + (size_t)c * l2] = m;
        b0[c + (size_t)r * l0] = m;
    }
    __syncthreads();
    f0 = b0;
    lf0 = l0;
    stage_in(wk, n, f0, lf0);
    cta_lu(f0, lf0, n, wk.ipiv, wk.prow, wk.jrow, wk.info, lu_staging(wk));
    note_fail(p, wk, 11);
    part = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) part += log(fabs(f0[i + (size_t)i * lf0]));
    const double logm = cta_reduce_sum(part, wk.red);
    // solve M^T Gt = X^T in place on X viewed transposed: the result is G column-major in b1
    cta_lu_solve(f0, lf0, n, wk.ipiv, b1, l1, 1);
    __syncthreads();
    return logdl + logm;
}

// --------------------------------------------------------------------------------------------------
// stablelinalg::inv_I_plus_ldr_mul_ldr (source/stablelinalg.cpp:128-158).  Result G in wk.b[2].
// --------------------------------------------------------------------------------------------------
__device__ void dev_inv_I_plus_ldr_mul_ldr(const Params& p, Work& wk, int n, const double* L1, const double* d1, const double* R1,
                                           const double* L2, const double* d2, const double* R2, int ldg, int rhs_mode = 0) {
    double* b0 = wk.b[0];
    double* b1 = wk.b[1];
    double* b2 = wk.b[2];
    const int l0 = wk.ld[0], l1 = wk.ld[1], l2 = wk.ld[2];
    __syncthreads();
    split_d(d1, n, wk.sv[0], wk.sv[1]);
    split_d(d2, n, wk.sv[2], wk.sv[3]);
    double* f0 = wk.stage ? wk.stage : b0;  // factorise R2 directly in the staging tile when there is one
    int lf0 = wk.stage ? wk.lds : l0;
    cta_copy(f0, lf0, R2, ldg, n);
    __syncthreads();
    DQ_PH(p, 5);
    cta_lu(f0, lf0, n, wk.ipiv, wk.prow, wk.jrow, wk.info, lu_staging(wk));
    DQ_PH(p, 6);
    note_fail(p, wk, 12);
    // right-hand side D2_large^-1 written with its rows already interchanged: (P D)(i, c) = D(perm[i], c)
    cta_perm_from_ipiv(wk.ipiv, wk.jpvt, n);
    for (int e = threadIdx.x; e < n * n; e += blockDim.x) {
        const int r = e % n, c = e / n;
        b1[r + (size_t)c * l1] = (wk.jpvt[r] == c) ? wk.sv[2][c] : 0.0;
    }
    __syncthreads();
    dev_lu_solve_prepermuted(wk, f0, lf0, n, b1, l1);  // X = R2^-1 D2_large^-1
    __syncthreads();
    DQ_PH(p, 7);
    // TermA = D1l^-1 (L1^T X)
    if (wk.arena && n == kGT_N && blockDim.x == 256) {  // L1^T materialised in b0 (free between the two factorisations): both operands column-major for the staged product
        cta_transpose256(b0, l0, L1, ldg, wk.arena);
        dev_gemm<false>(wk, b2, l2, b0, 1, l0, b1, 1, l1, n, wk.sv[0], nullptr);
        __syncthreads();
    } else {
        cta_gemm<false>(b2, l2, L1, ldg, 1, b1, 1, l1, n, wk.sv[0], nullptr);
    }
    // b0 = L2 D2s
    cta_copy(b0, l0, L2, ldg, n, nullptr, wk.sv[3]);
    __syncthreads();
    // M = TermA + D1s (R1 (L2 D2s))
    dev_gemm<true>(wk, b2, l2, R1, 1, ldg, b0, 1, l0, n, wk.sv[1], nullptr);
    __syncthreads();
    double* f2 = b2;
    int lf2 = l2;
    stage_in(wk, n, f2, lf2);
    DQ_PH(p, 9);
    cta_lu(f2, lf2, n, wk.ipiv, wk.prow, wk.jrow, wk.info, lu_staging(wk));
    DQ_PH(p, 10);
    note_fail(p, wk, 13);
    // RHS = D1l^-1 L1^T  -> b0      (rhs_mode 1: RHS = D1s R1, stablelinalg::inv_invldr_plus_ldr,
    //                                  source/stablelinalg.cpp:160-190 -- same M, different right-hand side),
    // built after the factorisation so that it can be written with its rows already interchanged
    cta_perm_from_ipiv(wk.ipiv, wk.jpvt, n);
    for (int e0 = threadIdx.x; e0 < n * n; e0 += 4 * blockDim.x) {  // four loads in flight per thread
        double v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int e = e0 + u * blockDim.x;
            const int r = e % n, c = e / n;
            const int pr = wk.jpvt[r];
            v[u] = (e < n * n) ? (rhs_mode ? wk.sv[1][pr] * R1[pr + (size_t)c * ldg] : wk.sv[0][pr] * L1[c + (size_t)pr * ldg]) : 0.0;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int e = e0 + u * blockDim.x;
            if (e < n * n) b0[e % n + (size_t)(e / n) * l0] = v[u];
        }
    }
    __syncthreads();
    dev_lu_solve_prepermuted(wk, f2, lf2, n, b0, l0);  // Y = M^-1 RHS
    __syncthreads();
    DQ_PH(p, 11);
    dev_gemm<false>(wk, b2, l2, b1, 1, l1, b0, 1, l0, n, nullptr, nullptr);  // G = X Y
    __syncthreads();
}

// --------------------------------------------------------------------------------------------------
// B-matrix helpers
// --------------------------------------------------------------------------------------------------
// DQMC::calculate_Bbar (source/dqmc.cpp:88-105): product of the block's B_l = diag(e^{V_l}) expK,
// slices ascending, from the CURRENT fields.  Result in wk.b[0] (uses wk.b[1] as ping-pong).
__device__ void dev_bbar(const Params& p, Work& wk, int walker, int i_stack) {
    const int n = p.N;
    const int l0 = i_stack * p.n_stab;
    const int l1 = min(l0 + p.n_stab, p.nt);
    const int8_t* f = p.fields + ((size_t)walker * p.nt) * n;
    if (p.kg) {
        // Kronecker chain in place in b[0]: B_l X = diag(ev_l) (Fy (x) Fx) X; a warp keeps its column groups for the whole
        // chain, so the slices need no barrier for the matrix itself
        if (wk.arena && n == kGT_N && p.kL1 == 16 && p.kL2 == 16 && blockDim.x == 256 && l1 - l0 <= 2 * kGT_KP * kGT_LDS / kGT_N) {
            // N = 256: the chain acts on every column separately, so each warp takes 8 columns of a 64-column panel in shared
            // memory through ALL slices of the block (no block barrier, no L2 round trip per slice) and writes them out once
            double* panel = reinterpret_cast<double*>(wk.arena + 64);
            double* sc = panel + (size_t)kGT_LDS * kGT_CP;  // exp(g eta) of every slice of the block
            const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
            __syncthreads();
            for (int e = threadIdx.x; e < (l1 - l0) * n; e += blockDim.x) sc[e] = p.lut_ev[f[(size_t)l0 * n + e]];
            __syncthreads();
            double* mine = panel + (size_t)(8 * warp) * kGT_LDS;
            for (int c0 = 8 * warp; c0 < n; c0 += kGT_CP) {
                for (int q = 0; q < 8; ++q)  // B_{l0} = diag(ev) expK, this warp's columns
                    for (int r = lane; r < n; r += 32) mine[r + (size_t)q * kGT_LDS] = sc[r] * p.expK[r + (size_t)(c0 + q) * n];
                __syncwarp();
                for (int l = 1; l < l1 - l0; ++l)
                    cta_kron_apply_ct<16, 16, false, 4>(panel, kGT_LDS, p.kg, p.kg + 256, nullptr, sc + (size_t)l * n, warp);
                for (int q = 0; q < 8; ++q)
                    for (int r = lane; r < n; r += 32) wk.b[0][r + (size_t)(c0 + q) * wk.ld[0]] = mine[r + (size_t)q * kGT_LDS];
                __syncwarp();
            }
            __syncthreads();
            return;
        }
        // (N = 144: the chain runs in the shared-memory staging tile, idle at this point, and is copied out at the end)
        double* X = wk.stage ? wk.stage : wk.b[0];
        const int ldx = wk.stage ? wk.lds : wk.ld[0];
        __syncthreads();
        for (int e = threadIdx.x; e < n * n; e += blockDim.x) {  // B_{l0} = diag(ev) expK
            const int r = e % n, c = e / n;
            X[r + (size_t)c * ldx] = p.lut_ev[f[(size_t)l0 * n + r]] * p.expK[e];
        }
        __syncthreads();
        // the slice's scaling exp(g eta) is staged in shared memory, double buffered: ONE block barrier per slice
        for (int l = l0 + 1; l < l1; ++l) {
            double* ev = wk.sv[l & 1];
            for (int i = threadIdx.x; i < n; i += blockDim.x) ev[i] = p.lut_ev[f[(size_t)l * n + i]];
            __syncthreads();
            cta_kron<false, 4>(X, ldx, n, p.kL1, p.kL2, p.kg, p.kg + 256, nullptr, ev);
        }
        if (wk.stage) {
            __syncthreads();
            cta_copy(wk.b[0], wk.ld[0], X, ldx, n);
            __syncthreads();
        }
        return;
    }
    double* cur = ((l1 - l0) & 1) ? wk.b[0] : wk.b[1];  // so that the last product lands in b[0]
    double* oth = ((l1 - l0) & 1) ? wk.b[1] : wk.b[0];
    const int ldc = ((l1 - l0) & 1) ? wk.ld[0] : wk.ld[1];
    const int ldo = ((l1 - l0) & 1) ? wk.ld[1] : wk.ld[0];
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x) wk.sv[0][i] = p.lut_ev[f[(size_t)l0 * n + i]];
    __syncthreads();
    cta_copy(cur, ldc, p.expK, n, n, wk.sv[0], nullptr);  // B_{l0} (reference multiplies by the identity: exact)
    __syncthreads();
    int ldcur = ldc, ldoth = ldo;
    for (int l = l0 + 1; l < l1; ++l) {
        for (int i = threadIdx.x; i < n; i += blockDim.x) wk.sv[0][i] = p.lut_ev[f[(size_t)l * n + i]];
        __syncthreads();
        // B_l * Bbar = diag(ev) (expK * Bbar)
        cta_gemm<false>(oth, ldoth, p.expK, 1, n, cur, 1, ldcur, n, wk.sv[0], nullptr);
        __syncthreads();
        double* t = cur;
        cur = oth;
        oth = t;
        const int tl = ldcur;
        ldcur = ldoth;
        ldoth = tl;
    }
}

// propagate_GF_forward: G <- B_l G B_l^-1 (source/dqmc.cpp:113-132).  G in g (ld ldg), scratch t.
template <bool BIG = true>
__device__ void dev_wrap_forward(const Params& p, int n, double* g, int ldg, double* t, int ldt, const double* ev, const double* iev) {
    if (p.kg) {  // G <- diag(ev) (E G) : left product;  then (.) Ei diag(iev) : left product on the transposed view
        cta_kron<false, BIG ? 4 : 2>(g, ldg, n, p.kL1, p.kL2, p.kg, p.kg + 256, nullptr, ev);
        __syncthreads();
        cta_kron<true, BIG ? 4 : 2>(g, ldg, n, p.kL1, p.kL2, p.kg + 512, p.kg + 768, nullptr, iev);
        __syncthreads();
        return;
    }
    cta_gemm<false, BIG>(t, ldt, p.expK, 1, n, g, 1, ldg, n, nullptr, nullptr);
    __syncthreads();
    cta_gemm<false, BIG>(g, ldg, t, 1, ldt, p.invexpK, 1, n, n, ev, iev);
    __syncthreads();
}
// propagate_GF_backward: G <- B_l^-1 G B_l (source/dqmc.cpp:169-187)
template <bool BIG = true>
__device__ void dev_wrap_backward(const Params& p, int n, double* g, int ldg, double* t, int ldt, const double* ev, const double* iev) {
    if (p.kg) {  // G <- Ei (diag(iev) G) : left product with pre-scaling;  then (. diag(ev)) E on the transposed view
        cta_kron<false, BIG ? 4 : 2>(g, ldg, n, p.kL1, p.kL2, p.kg + 512, p.kg + 768, iev, nullptr);
        __syncthreads();
        cta_kron<true, BIG ? 4 : 2>(g, ldg, n, p.kL1, p.kL2, p.kg, p.kg + 256, ev, nullptr);
        __syncthreads();
        return;
    }
    // t = invexpK (diag(iev) G)   [reference: (invexpK diag(iev)) G — the same product up to rounding order]
    for (int e = threadIdx.x; e < n * n; e += blockDim.x) {
        const int r = e % n, c = e / n;
        g[r + (size_t)c * ldg] *= iev[r];
    }
    __syncthreads();
    cta_gemm<false, BIG>(t, ldt, p.invexpK, 1, n, g, 1, ldg, n, nullptr, nullptr);
    __syncthreads();
    // g = (t diag(ev)) expK
    for (int e = threadIdx.x; e < n * n; e += blockDim.x) {
        const int r = e % n, c = e / n;
        t[r + (size_t)c * ldt] *= ev[c];
    }
    __syncthreads();
    cta_gemm<false, BIG>(g, ldg, t, 1, ldt, p.expK, 1, n, n, nullptr, nullptr);
    __syncthreads();
}

// --------------------------------------------------------------------------------------------------
// RNG: Philox4x32-10 (throughput mode)
// --------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += 0x9E3779B9u;
        k.y += 0xBB67AE85u;
    }
    return c;
}

// Fill the per-slice draw tables for one local_update call ("visit") of one walker.
//   s_perm[N] visiting order, s_k[N] proposal index 0..2, s_u[N] uniform in [0,1)
// skey: N uint32 scratch.
__device__ void load_visit_draws(const Params& p, int walker, unsigned long long visit, int* s_perm, int* s_k, double* s_u,
                                 uint32_t* skey) {
    const int n = p.N;
    if (p.rng_mode == DQMCB_RNG_PHILOX) {
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            const uint4 r = philox4x32_10(make_uint4((uint32_t)i, (uint32_t)visit, (uint32_t)(visit >> 32), (uint32_t)walker),
                                          make_uint2((uint32_t)p.seed, (uint32_t)(p.seed >> 32)));
            skey[i] = r.x;
            s_k[i] = (int)__umulhi(r.y, 3u);
            const unsigned long long bits = ((unsigned long long)r.w << 32) | r.z;
            s_u[i] = (double)(bits >> 11) * (1.0 / 9007199254740992.0);
        }
        __syncthreads();
        for (int i = threadIdx.x; i < n; i += blockDim.x) {  // rank of key i = position in the random order
            const uint32_t ki = skey[i];
            int rank = 0;
            for (int j = 0; j < n; ++j) {
                const uint32_t kj = skey[j];
                rank += (kj < ki) || (kj == ki && j < i);
            }
            s_perm[rank] = i;
        }
    } else {  // TAPE / MT19937: draws were staged in global memory in visiting order
        const size_t off = ((size_t)walker * p.tape_visits + (size_t)(visit - p.visit_base)) * n;
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            s_perm[i] = p.tape_perm[off + i];
            s_k[i] = p.tape_k[off + i];
            s_u[i] = p.tape_u[off + i];
        }
    }
    __syncthreads();
}

// --------------------------------------------------------------------------------------------------
// update::local_update (source/update.cpp:5-32) with AttractiveHubbard::local_update_ratio /
// update_greens_local (source/model.cpp:109-138): sequential Metropolis over the shuffled sites of
// one slice, rank-1 Sherman-Morrison on accept.  Every thread evaluates the (identical) accept
// decision, so no barrier is needed for rejected proposals.  The update uses separate multiply
// and add (no FMA contraction), i.e. exactly the reference's `G += prefactor*U*V` rounding.
// --------------------------------------------------------------------------------------------------
__device__ int dev_local_update(const Params& p, int n, double* g, int ldg, int8_t* s_f, const int* s_perm, const int* s_k,
                                const double* s_u, double* s_ucol, double* s_vrow) {
    int accepted = 0;
    for (int idx = 0; idx < n; ++idx) {
        const int i = s_perm[idx];
        const int oldf = s_f[i];
        const int newf = s_k[idx] + (s_k[idx] >= oldf);  // k-th element of {0,1,2,3}\{old}, include/field.h:45-48
        const double gii = g[i + (size_t)i * ldg];
        const double delta = p.lut_delta[oldf * 4 + newf];
        const double r = 1.0 + (1.0 - gii) * delta;
        const double R = p.lut_gb[oldf * 4 + newf] * (r * r);
        const double prob = fmin(1.0, fabs(R));
        if (s_u[idx] < prob) {
            ++accepted;
            const double pref = delta / r;
            for (int t = threadIdx.x; t < n; t += blockDim.x) {
                s_ucol[t] = __dmul_rn(pref, g[t + (size_t)i * ldg]);
                s_vrow[t] = g[i + (size_t)t * ldg] - ((t == i) ? 1.0 : 0.0);
            }
            __syncthreads();
            for (int e = threadIdx.x; e < n * n; e += blockDim.x) {
                const int rr = e % n, cc = e / n;
                double* ge = g + rr + (size_t)cc * ldg;
                *ge = __dadd_rn(*ge, __dmul_rn(s_ucol[rr], s_vrow[cc]));
            }
            if (threadIdx.x == 0) s_f[i] = (int8_t)newf;
            __syncthreads();
        }
    }
    return accepted;
}

// --------------------------------------------------------------------------------------------------
// Delayed rank-k variant of the same loop (not in the reference, which lists it as a TODO in its
// README; algebraically identical to the rank-1 sequence: SURVEY.md appendix A).  G = G0 + U V with
// U (n x m), V (m x n), m <= K accepted flips pending: a proposal needs only the running diagonal;
// an accept forms the current column / row from G0 and the pending factors (2 n m flops instead of a
// 2 n^2 sweep over G) and G0 is touched once per K accepts by a rank-K product.  Pays off where G does
// not sit in registers: N = 144 (G in shared memory) and N = 256 (G in L2 / HBM).
//   dU[m*ldu + r] = U(r, m),  dV[m*ldu + c] = V(m, c)  (ldu = delay_ld(n)),  s_diag[n] running diag(G).
// --------------------------------------------------------------------------------------------------
template <bool BIG = true>
__device__ int dev_local_update_delayed(const Params& p, int n, double* g, int ldg, int8_t* s_f, const int* s_perm, const int* s_k,
                                        const double* s_u, double* s_diag, double* dU, double* dV, int K) {
    int accepted = 0, cnt = 0;
    const int ldu = delay_ld(n);
    // running diagonal, double-buffered: an accepting thread writes the other copy, so a warp that is
    // still deciding this (or an earlier, rejected) proposal never sees a half-updated diagonal
    double* dcur = s_diag;
    double* dnxt = s_diag + n;
    for (int t = threadIdx.x; t < n; t += blockDim.x) dcur[t] = g[t + (size_t)t * ldg];
    __syncthreads();
    for (int idx = 0; idx < n; ++idx) {
        const int i = s_perm[idx];
        const int oldf = s_f[i];
        const int newf = s_k[idx] + (s_k[idx] >= oldf);
        const double gii = dcur[i];
        const double delta = p.lut_delta[oldf * 4 + newf];
        const double r = 1.0 + (1.0 - gii) * delta;
        const double R = p.lut_gb[oldf * 4 + newf] * (r * r);
        const double prob = fmin(1.0, fabs(R));
        if (s_u[idx] < prob) {
            ++accepted;
            const double pref = delta * rcp_newton(r);  // <= 1 ulp from delta / r, off the IEEE division's ~150-cycle chain
            for (int t = threadIdx.x; t < n; t += blockDim.x) {
                double uc = g[t + (size_t)i * ldg], vr = g[i + (size_t)t * ldg];
                // pending factors, two independent accumulation chains per sum (the loads of a step overlap the FMAs of the other)
                double uc1 = 0.0, vr1 = 0.0;
                int m = 0;
                for (; m + 1 < cnt; m += 2) {
                    uc = fma(dU[m * ldu + t], dV[m * ldu + i], uc);
                    vr = fma(dU[m * ldu + i], dV[m * ldu + t], vr);
                    uc1 = fma(dU[(m + 1) * ldu + t], dV[(m + 1) * ldu + i], uc1);
                    vr1 = fma(dU[(m + 1) * ldu + i], dV[(m + 1) * ldu + t], vr1);
                }
                if (m < cnt) {
                    uc = fma(dU[m * ldu + t], dV[m * ldu + i], uc);
                    vr = fma(dU[m * ldu + i], dV[m * ldu + t], vr);
                }
                uc += uc1;
                vr += vr1;
                if (t == i) vr -= 1.0;
                const double us = pref * uc;
                dU[cnt * ldu + t] = us;
                dV[cnt * ldu + t] = vr;
                dnxt[t] = fma(us, vr, dcur[t]);
            }
            ++cnt;
            double* sw = dcur;
            dcur = dnxt;
            dnxt = sw;
            __syncthreads();
            if (threadIdx.x == 0) s_f[i] = (int8_t)newf;
            if (cnt == K) {
                cta_gemm<true, BIG>(g, ldg, dU, 1, ldu, dV, ldu, 1, n, nullptr, nullptr, cnt);
                cnt = 0;
                __syncthreads();
            }
        }
    }
    if (cnt) {
        cta_gemm<true, BIG>(g, ldg, dU, 1, ldu, dV, ldu, 1, n, nullptr, nullptr, cnt);
        __syncthreads();
    }
    return accepted;
}

}  // namespace dq
