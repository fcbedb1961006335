// fast64.cuh -- compile-time-N fast path of the sweep engine (N = 32 or 64; the headline config
// C1 is N = 64).  Same mathematics and operation order as engine.cuh (which remains the any-N path
// and the second parity reference), but
//   * every dense product runs on the FP64 tensor cores (mma.sync m8n8k4 f64 -> DMMA.8x8x4) from
//     bank-conflict-free shared-memory tiles (tile.cuh),
//   * HBM operands written by earlier kernels are staged with TMA bulk copies (cp.async.bulk),
//   * QR / LU / triangular solves are fully unrolled for N,
//   * the propagate kernel keeps G in registers (DMMA accumulator layout) across wrap and flips,
//     batches several walkers per CTA so exp(-+dtau K) is shared in SMEM, and emits the block
//     product Bbar as a tail phase.
#pragma once
#include "dqmcb_internal.cuh"
#include "engine.cuh"
#include "tile.cuh"

namespace dq {

// ---------------------------------------------------------------------------------------------------
// shared-memory plan of the stabilisation kernels: small vectors, 2 mbarriers, 3 tiles
// ---------------------------------------------------------------------------------------------------
template <int N>
struct StabSmem {
    static constexpr size_t SMALL = ((size_t)(12 * N + 4 + 64) * 8 + (size_t)(2 * N + 4) * 4 + 15) & ~(size_t)15;
    static constexpr size_t BARS = 16;
    static constexpr size_t TOTAL = SMALL + BARS + 3 * Tile<N>::BYTES;
};

template <int N>
struct FWork {
    double* t[3];
    double *tau, *vn1, *vn2, *sv[4], *dnew, *prow, *jrow, *red, *lbuf;
    int *jpvt, *ipiv, *info;
    uint64_t* bar;
    uint32_t parity;
};

template <int N>
__device__ inline void fcarve(FWork<N>& wk, unsigned char* smem) {
    double* s = reinterpret_cast<double*>(smem);
    wk.tau = s;
    wk.vn1 = s + N;
    wk.vn2 = s + 2 * N;
    for (int i = 0; i < 4; ++i) wk.sv[i] = s + (3 + i) * N;
    wk.dnew = s + 7 * N;
    wk.prow = s + 8 * N;
    wk.jrow = s + 9 * N;
    wk.red = s + 10 * N;
    wk.lbuf = s + 10 * N + 64;  // 2 (N + 2) doubles: double-buffered multiplier column of lu_tile
    int* ip = reinterpret_cast<int*>(s + 12 * N + 4 + 64);
    wk.jpvt = ip;
    wk.ipiv = ip + N;
    wk.info = ip + 2 * N;
    wk.bar = reinterpret_cast<uint64_t*>(smem + StabSmem<N>::SMALL);
    double* tiles = reinterpret_cast<double*>(smem + StabSmem<N>::SMALL + StabSmem<N>::BARS);
    for (int i = 0; i < 3; ++i) wk.t[i] = tiles + (size_t)i * Tile<N>::ELEMS;
    wk.parity = 0;
    if (threadIdx.x == 0) {
        mbar_init(wk.bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
        fence_proxy_async();
    }
    __syncthreads();
}

// Blocking TMA fetch of a dense global N x N matrix (written by an EARLIER kernel) into a tile.
template <int N>
__device__ inline void tile_fetch(FWork<N>& wk, double* dst, const double* __restrict__ src) {
    fence_proxy_async();  // order this thread's earlier generic accesses to `dst` before the async writes
    tile_load_async<N>(dst, src, wk.bar);
    mbar_wait(wk.bar, wk.parity);
    wk.parity ^= 1;
}

// to_LDR on tile A (destroyed): Q -> tile Qo, D^-1 R P^T -> tile Ro, |diag R| -> dout
template <int N>
__device__ void f_to_ldr(const Params& p, FWork<N>& wk, double* A, double* Qo, double* Ro, double* dout) {
    constexpr int LDT = Tile<N>::LDT;
    __syncthreads();
    DQ_PH(p, 20);
    qrp_tile<N>(A, wk.tau, wk.jpvt, wk.vn1, wk.sv[1]);  // sv[1..2] (contiguous, dead here): the published pivot column
    DQ_PH(p, 21);
    // column jpvt[j] of A holds R(0..j, j) (and the reflector below): d_j = |R(j,j)|, and
    // R_final = (D^-1 R) P^T puts that column back at index jpvt[j] -- where it already is
    for (int j = threadIdx.x; j < N; j += blockDim.x) {
        const int c = wk.jpvt[j];
        const double dj = fabs(A[j + (size_t)c * LDT]);
        dout[j] = dj;
        wk.vn2[j] = 1.0 / dj;
        wk.ipiv[c] = j;  // step at which column c was chosen (scratch)
    }
    __syncthreads();
    for (int e = threadIdx.x; e < N * N; e += blockDim.x) {
        const int i = e % N, c = e / N;
        Ro[i + (size_t)c * LDT] = (i <= wk.ipiv[c]) ? wk.vn2[i] * A[i + (size_t)c * LDT] : 0.0;
    }
    DQ_PH(p, 22);
    form_q_tile<N>(Qo, A, wk.tau, wk.jpvt);
    __syncthreads();
    DQ_PH(p, 23);
}

template <int N>
__device__ inline void f_set_diag(double* B, const double* dvec) {
    constexpr int LDT = Tile<N>::LDT;
    for (int e = threadIdx.x; e < N * N; e += blockDim.x) {
        const int r = e % N, c = e / N;
        B[r + (size_t)c * LDT] = (r == c) ? dvec[r] : 0.0;
    }
}

template <int N>
__device__ inline void f_note_fail(const Params& p, const FWork<N>& wk, int code) {
    if (threadIdx.x == 0 && *wk.info != 0) atomicCAS(p.fail, 0, code);
}

// inv_I_plus_ldr (source/stablelinalg.cpp:94-126); L, d, R dense global (ld = N).  G -> wk.t[2].
template <int N>
__device__ double f_inv_I_plus_ldr(const Params& p, FWork<N>& wk, const double* L, const double* d, const double* R) {
    constexpr int LDT = Tile<N>::LDT;
    double *t0 = wk.t[0], *t1 = wk.t[1], *t2 = wk.t[2];
    __syncthreads();
    split_d(d, N, wk.sv[0], wk.sv[1]);
    tile_load_global<N>(t1, R);
    double part = 0.0;
    for (int i = threadIdx.x; i < N; i += blockDim.x) part += (d[i] >= 1.0) ? log(d[i]) : 0.0;
    const double logdl = cta_reduce_sum(part, wk.red);  // contains barriers
    lu_tile<N>(t1, reinterpret_cast<int*>(wk.jrow), wk.prow, wk.info, wk.lbuf);
    f_note_fail<N>(p, wk, 10);
    f_set_diag<N>(t2, wk.sv[0]);
    __syncthreads();
    lu_solve_tile<N>(t1, reinterpret_cast<int*>(wk.jrow), wk.prow, t2, 1, LDT);  // X
    __syncthreads();
    for (int e = threadIdx.x; e < N * N; e += blockDim.x) {  // t0 <- M^T,  M = X + L D_small
        const int r = e % N, c = e / N;
        t0[c + (size_t)r * LDT] = t2[r + (size_t)c * LDT] + L[r + (size_t)c * N] * wk.sv[1][c];
    }
    __syncthreads();
    lu_tile<N>(t0, reinterpret_cast<int*>(wk.jrow), wk.prow, wk.info, wk.lbuf);
    f_note_fail<N>(p, wk, 11);
    part = 0.0;
    for (int i = threadIdx.x; i < N; i += blockDim.x) part += log(fabs(t0[i + (size_t)i * LDT]));
    const double logm = cta_reduce_sum(part, wk.red);
    lu_solve_tile<N>(t0, reinterpret_cast<int*>(wk.jrow), wk.prow, t2, LDT, 1);  // M^T Gt = X^T, in place on X viewed transposed
    __syncthreads();
    return logdl + logm;
}

// inv_I_plus_ldr_mul_ldr (source/stablelinalg.cpp:128-158); all six inputs dense global (ld = N).
// G -> wk.t[2].
template <int N>
__device__ void f_inv_I_plus_ldr_mul_ldr(const Params& p, FWork<N>& wk, const double* L1, const double* d1, const double* R1,
                                         const double* L2, const double* d2, const double* R2) {
    constexpr int LDT = Tile<N>::LDT;
    double *t0 = wk.t[0], *t1 = wk.t[1], *t2 = wk.t[2];
    __syncthreads();
    split_d(d1, N, wk.sv[0], wk.sv[1]);
    split_d(d2, N, wk.sv[2], wk.sv[3]);
    tile_load_global<N>(t0, R2);
    __syncthreads();
    DQ_PH(p, 5);
    lu_tile<N>(t0, reinterpret_cast<int*>(wk.jrow), wk.prow, wk.info, wk.lbuf);
    DQ_PH(p, 6);
    f_note_fail<N>(p, wk, 12);
    f_set_diag<N>(t1, wk.sv[2]);
    __syncthreads();
    lu_solve_tile<N>(t0, reinterpret_cast<int*>(wk.jrow), wk.prow, t1, 1, LDT);  // X = R2^-1 D2l^-1  (t1)
    __syncthreads();
    DQ_PH(p, 7);
    // M = D1l^-1 (L1^T X) + D1s (R1 (L2 D2s))   (t2).  The column scaling by D2s is applied to the
    // product instead of to L2 first: identical in exact arithmetic, no cancellation involved.
    cta_mma<N, false>(t2, LDT, L1, N, 1, t1, 1, LDT, wk.sv[0], nullptr);
    __syncthreads();
    cta_mma<N, true>(t2, LDT, R1, 1, N, L2, 1, N, wk.sv[1], wk.sv[3]);
    DQ_PH(p, 8);
    // RHS = D1l^-1 L1^T  (t0; the LU of R2 is dead)
    for (int e = threadIdx.x; e < N * N; e += blockDim.x) {
        const int c = e % N, r = e / N;
        t0[r + (size_t)c * LDT] = wk.sv[0][r] * L1[c + (size_t)r * N];
    }
    __syncthreads();
    DQ_PH(p, 9);
    lu_tile<N>(t2, reinterpret_cast<int*>(wk.jrow), wk.prow, wk.info, wk.lbuf);
    DQ_PH(p, 10);
    f_note_fail<N>(p, wk, 13);
    lu_solve_tile<N>(t2, reinterpret_cast<int*>(wk.jrow), wk.prow, t0, 1, LDT);  // Y (t0)
    __syncthreads();
    DQ_PH(p, 11);
    cta_mma<N, false>(t2, LDT, t1, 1, LDT, t0, 1, LDT, nullptr, nullptr);  // G = X Y
    __syncthreads();
    DQ_PH(p, 12);
}

}  // namespace dq

namespace dq {

// calculate_Bbar (source/dqmc.cpp:88-105) on the tensor cores: expK staged once into t[2], running
// product ping-pongs between t[0] and t[1]; result in t[0].
template <int N>
__device__ void f_bbar(const Params& p, FWork<N>& wk, int walker, int i_stack) {
    constexpr int LDT = Tile<N>::LDT;
    const int l0 = i_stack * p.n_stab;
    const int l1 = min(l0 + p.n_stab, p.nt);
    const int8_t* f = p.fields + ((size_t)walker * p.nt) * N;
    const int m = l1 - l0;
    // Kronecker chain (8 x 8 lattice): in place in t[0], exp(g eta) of the whole block staged in the idle tile t[1]
    const bool kron = (N == 64) && p.kron != nullptr && (m * N <= Tile<N>::ELEMS);
    double* cur = (kron || (m & 1)) ? wk.t[0] : wk.t[1];
    double* oth = (kron || (m & 1)) ? wk.t[1] : wk.t[0];
    __syncthreads();
    // in-place chain: exp(g eta) of every slice of the block is staged up front in the idle tile t[1] (one exposure of
    // the HBM latency of the field bytes instead of one per slice)
    if (kron)
        for (int e = threadIdx.x; e < m * N; e += blockDim.x) wk.t[1][e] = p.lut_ev[f[(size_t)l0 * N + e]];
    tile_fetch<N>(wk, wk.t[2], p.expK);
    for (int i = threadIdx.x; i < N; i += blockDim.x) wk.sv[0][i] = p.lut_ev[f[(size_t)l0 * N + i]];
    __syncthreads();
    for (int e = threadIdx.x; e < N * N; e += blockDim.x) {
        const int r = e % N, c = e / N;
        cur[r + (size_t)c * LDT] = wk.sv[0][r] * wk.t[2][r + (size_t)c * LDT];
    }
    __syncthreads();
    if (kron) {
        // exp(-dtau K) = Ey (x) Ex: B_l X = diag(ev_l) (Ey (x) I)(I (x) Ex) X as two in-place passes of 8 x 8 tensor-core
        // products with the factor held in registers (256 DMMA.8x8x4 per slice instead of 1024):
        //   pass 1: every 8 x 8 block (rows 8y.., columns 8cb..) <- Ex * block;
        //   pass 2: every column c, viewed as the 8 x 8 matrix W(x, y') = X[x + 8 y', c], <- W Ey^T, scaled by ev.
        // A task's inputs are fully loaded (warp-collective DMMA) before its outputs are stored: in place is safe.
        // Both passes act on each COLUMN separately (V <- Ex V Ey^T for the 8 x 8 matrix V(x, y) = X[x + 8 y, c]), so warp q
        // owns the columns [8q, 8q+8) for the whole chain: no block barrier between passes or slices, only __syncwarp.
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, gr = lane >> 2, tg = lane & 3;
        const double ex0 = p.kron[gr + 8 * tg], ex1 = p.kron[gr + 8 * (tg + 4)];            // A fragment: Ex(gr, k)
        const double ey0 = p.kron[64 + gr + 8 * tg], ey1 = p.kron[64 + gr + 8 * (tg + 4)];  // B fragment: Ey^T(k, gr) = Ey(gr, k)
        double* X = cur;
        for (int l = l0 + 1; l < l1; ++l) {
            const double* evl = wk.t[1] + (size_t)(l - l0) * N;
            // all loads of a pass are issued before its first DMMA (8 independent tasks per warp)
            constexpr int NW = kThreads / 32, NT = 64 / NW;
            double b0[NT], b1[NT], c0[NT], c1[NT];
#pragma unroll
            for (int u = 0; u < NT; ++u) {
                const int y = u, cb = warp;  // this warp's 8 columns, every row block
                const double* bp = X + (8 * y + tg) + (size_t)(8 * cb + gr) * LDT;
                b0[u] = bp[0];
                b1[u] = bp[4];
            }
#pragma unroll
            for (int u = 0; u < NT; ++u) {
                c0[u] = c1[u] = 0.0;
                dmma884(c0[u], c1[u], ex0, b0[u]);
            }
#pragma unroll
            for (int u = 0; u < NT; ++u) dmma884(c0[u], c1[u], ex1, b1[u]);
#pragma unroll
            for (int u = 0; u < NT; ++u) {
                const int y = u, cb = warp;
                double* cp = X + (8 * y + gr) + (size_t)(8 * cb + 2 * tg) * LDT;
                cp[0] = c0[u];
                cp[LDT] = c1[u];
            }
            __syncwarp();
#pragma unroll
            for (int u = 0; u < NT; ++u) {
                const double* colp = X + (size_t)(NT * warp + u) * LDT;
                b0[u] = colp[gr + 8 * tg];
                b1[u] = colp[gr + 8 * (tg + 4)];
            }
#pragma unroll
            for (int u = 0; u < NT; ++u) {
                c0[u] = c1[u] = 0.0;
                dmma884(c0[u], c1[u], b0[u], ey0);
            }
#pragma unroll
            for (int u = 0; u < NT; ++u) dmma884(c0[u], c1[u], b1[u], ey1);
            {
                const int r0 = gr + 16 * tg;  // row x + 8 y of the output element (x = gr, y = 2 tg + j)
                const double e0 = evl[r0], e1 = evl[r0 + 8];
#pragma unroll
                for (int u = 0; u < NT; ++u) {
                    double* colp = X + (size_t)(NT * warp + u) * LDT;
                    colp[r0] = e0 * c0[u];
                    colp[r0 + 8] = e1 * c1[u];
                }
            }
            __syncwarp();
        }
        return;
    }
    for (int l = l0 + 1; l < l1; ++l) {
        for (int i = threadIdx.x; i < N; i += blockDim.x) wk.sv[0][i] = p.lut_ev[f[(size_t)l * N + i]];
        __syncthreads();
        cta_mma<N, false>(oth, LDT, wk.t[2], 1, LDT, cur, 1, LDT, wk.sv[0], nullptr);
        __syncthreads();
        double* tmp = cur;
        cur = oth;
        oth = tmp;
    }
}

__device__ inline double* fstkL(const Params& p, int w, int i) { return p.stackL + ((size_t)w * p.n_stack + i) * p.N * p.N; }
__device__ inline double* fstkR(const Params& p, int w, int i) { return p.stackR + ((size_t)w * p.n_stack + i) * p.N * p.N; }
__device__ inline double* fstkD(const Params& p, int w, int i) { return p.stackD + ((size_t)w * p.n_stack + i) * p.N; }
__device__ inline void fstore_d(double* dst, const double* src, int n) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) dst[i] = src[i];
}

template <int N>
__global__ void __launch_bounds__(kThreads, 2) kf_bbar(Params p, int i_stack) {
    extern __shared__ __align__(16) unsigned char smem[];
    FWork<N> wk;
    fcarve<N>(wk, smem);
    f_bbar<N>(p, wk, blockIdx.x, i_stack);
    __syncthreads();
    tile_store_global<N>(p.Bbar + (size_t)blockIdx.x * N * N, wk.t[0]);
}

// End-of-block work (source/dqmc.cpp:369-394 / 429-454), fast path.  Same structure as k_stabilize.
template <int N>
__global__ void __launch_bounds__(kThreads, 2) kf_stabilize(Params p, int l, int dir) {
    extern __shared__ __align__(16) unsigned char smem[];
    constexpr int LDT = Tile<N>::LDT;
    const int w = p.w0 + blockIdx.x;
    FWork<N> wk;
    fcarve<N>(wk, smem);
    const int i = l / p.n_stab;
    DQ_PH(p, 0);
    // The two neighbouring stack entries (the one the new entry is built from and the one it is combined
    // with for G) stream from HBM: pull them into L2 now, behind the Bbar chain and the QR.
    if (threadIdx.x == 0) {
        constexpr uint32_t MB_ = (uint32_t)(N * N * sizeof(double));
        if (i > 0) {
            bulk_prefetch_l2(fstkL(p, w, i - 1), MB_);
            bulk_prefetch_l2(fstkR(p, w, i - 1), MB_);
        }
        if (i + 1 < p.n_stack) {
            bulk_prefetch_l2(fstkL(p, w, i + 1), MB_);
            bulk_prefetch_l2(fstkR(p, w, i + 1), MB_);
        }
    }
    f_bbar<N>(p, wk, w, i);  // -> t[0]
    double *t0 = wk.t[0], *t1 = wk.t[1], *t2 = wk.t[2];
    __syncthreads();
    DQ_PH(p, 1);
    const bool first = (dir > 0) ? (i == 0) : (i == p.n_stack - 1);
    if (first) {
        f_to_ldr<N>(p, wk, t0, t1, t2, wk.dnew);
        tile_store_global<N>(fstkL(p, w, i), t1);
        tile_store_global<N>(fstkR(p, w, i), t2);
    } else if (dir > 0) {
        // mat_mul_ldr: (Bbar L_prev) D_prev -> QR ; R <- R_new R_prev
        fstore_d(wk.sv[0], fstkD(p, w, i - 1), N);
        tile_fetch<N>(wk, t1, fstkL(p, w, i - 1));
        __syncthreads();
        cta_mma<N, false>(t2, LDT, t0, 1, LDT, t1, 1, LDT, nullptr, wk.sv[0]);
        f_to_ldr<N>(p, wk, t2, t0, t1, wk.dnew);
        fence_proxy_async();
        tile_load_async<N>(t2, fstkR(p, w, i - 1), wk.bar);  // in flight while L is written out
        DQ_PH(p, 24);
        tile_store_global<N>(fstkL(p, w, i), t0);
        mbar_wait(wk.bar, wk.parity);
        wk.parity ^= 1;
        __syncthreads();
        DQ_PH(p, 25);
        cta_mma<N, false>(fstkR(p, w, i), N, t1, 1, LDT, t2, 1, LDT, nullptr, nullptr);
        DQ_PH(p, 26);
    } else {
        // ldr_mul_mat: D_next (R_next Bbar) -> QR ; L <- L_next Q
        fstore_d(wk.sv[0], fstkD(p, w, i + 1), N);
        tile_fetch<N>(wk, t1, fstkR(p, w, i + 1));
        __syncthreads();
        cta_mma<N, false>(t2, LDT, t1, 1, LDT, t0, 1, LDT, wk.sv[0], nullptr);
        f_to_ldr<N>(p, wk, t2, t0, t1, wk.dnew);
        fence_proxy_async();
        tile_load_async<N>(t2, fstkL(p, w, i + 1), wk.bar);  // in flight while R is written out
        tile_store_global<N>(fstkR(p, w, i), t1);
        mbar_wait(wk.bar, wk.parity);
        wk.parity ^= 1;
        __syncthreads();
        cta_mma<N, false>(fstkL(p, w, i), N, t2, 1, LDT, t0, 1, LDT, nullptr, nullptr);
    }
    fstore_d(fstkD(p, w, i), wk.dnew, N);
    __syncthreads();
    DQ_PH(p, 4);
    const bool at_end = (dir > 0) ? (l == p.nt - 1) : (l == 0);
    if (at_end) {
        const double ld_ = f_inv_I_plus_ldr<N>(p, wk, fstkL(p, w, i), fstkD(p, w, i), fstkR(p, w, i));
        if (threadIdx.x == 0) p.logdet[w] = ld_;
    } else if (dir > 0) {
        f_inv_I_plus_ldr_mul_ldr<N>(p, wk, fstkL(p, w, i), fstkD(p, w, i), fstkR(p, w, i), fstkL(p, w, i + 1), fstkD(p, w, i + 1),
                                    fstkR(p, w, i + 1));
    } else {
        f_inv_I_plus_ldr_mul_ldr<N>(p, wk, fstkL(p, w, i - 1), fstkD(p, w, i - 1), fstkR(p, w, i - 1), fstkL(p, w, i), fstkD(p, w, i),
                                    fstkR(p, w, i));
    }
    // check_error + install
    double* Gg = p.G + (size_t)w * N * N;
    double emax = 0.0;
    {
        constexpr int PER = N * N / 2 / kThreads;  // double2 elements per thread, all loads issued up front
        static_assert((N * N / 2) % kThreads == 0, "check_error tiling");
        double2 old[PER];
#pragma unroll
        for (int u = 0; u < PER; ++u) old[u] = reinterpret_cast<const double2*>(Gg)[threadIdx.x + u * kThreads];
#pragma unroll
        for (int u = 0; u < PER; ++u) {
            const int e2 = threadIdx.x + u * kThreads;
            const int r = (e2 % (N / 2)) * 2, c = e2 / (N / 2);
            const double2 v = *reinterpret_cast<const double2*>(t2 + r + (size_t)c * LDT);
            emax = fmax(emax, fmax(fabs(old[u].x - v.x), fabs(old[u].y - v.y)));
            reinterpret_cast<double2*>(Gg)[e2] = v;
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

// init_stacks (source/dqmc.cpp:43-59), fast path
template <int N>
__global__ void __launch_bounds__(kThreads, 2) kf_init_stacks(Params p) {
    extern __shared__ __align__(16) unsigned char smem[];
    constexpr int LDT = Tile<N>::LDT;
    const int w = blockIdx.x;
    FWork<N> wk;
    fcarve<N>(wk, smem);
    double *t0 = wk.t[0], *t1 = wk.t[1], *t2 = wk.t[2];
    for (int i = p.n_stack - 1; i >= 0; --i) {
        f_bbar<N>(p, wk, w, i);
        __syncthreads();
        f_to_ldr<N>(p, wk, t0, t1, t2, wk.sv[2]);  // Bbar_ldr: L2 = t1, R2 = t2, d2 = sv[2]
        if (i == p.n_stack - 1) {
            tile_store_global<N>(fstkL(p, w, i), t1);
            tile_store_global<N>(fstkR(p, w, i), t2);
            fstore_d(fstkD(p, w, i), wk.sv[2], N);
        } else {
            // ldr_mul_ldr(stack[i+1], Bbar_ldr): M' = D1 (R1 L2) D2 -> QR ; L = L1 Q ; R = R_new R2
            fstore_d(wk.sv[0], fstkD(p, w, i + 1), N);
            __syncthreads();
            cta_mma<N, false>(t0, LDT, fstkR(p, w, i + 1), 1, N, t1, 1, LDT, wk.sv[0], wk.sv[2]);
            // park R2 in the (not yet written) global R slot of entry i while the tiles are busy
            tile_store_global<N>(fstkR(p, w, i), t2);
            f_to_ldr<N>(p, wk, t0, t1, t2, wk.dnew);  // Q = t1, R_new = t2
            cta_mma<N, false>(fstkL(p, w, i), N, fstkL(p, w, i + 1), 1, N, t1, 1, LDT, nullptr, nullptr);
            tile_load_global<N>(t0, fstkR(p, w, i));  // R2 back (same CTA wrote it: generic proxy)
            __syncthreads();
            cta_mma<N, false>(fstkR(p, w, i), N, t2, 1, LDT, t0, 1, LDT, nullptr, nullptr);
            fstore_d(fstkD(p, w, i), wk.dnew, N);
        }
        __syncthreads();
    }
}

template <int N>
__global__ void __launch_bounds__(kThreads, 2) kf_init_green(Params p) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int w = blockIdx.x;
    FWork<N> wk;
    fcarve<N>(wk, smem);
    const double ld_ = f_inv_I_plus_ldr<N>(p, wk, fstkL(p, w, 0), fstkD(p, w, 0), fstkR(p, w, 0));
    if (threadIdx.x == 0) p.logdet[w] = ld_;
    tile_store_global<N>(p.G + (size_t)w * N * N, wk.t[2]);
}

}  // namespace dq

#include "stab2.cuh"
#include "prop64.cuh"

// =====================================================================================================
// host-side dispatch of the fast path
// =====================================================================================================
static inline bool fast_n(int n) { return n == 64 || n == 32; }
static inline bool fast64_available(const dqmcb_ctx* ctx) { return fast_n(ctx->p.N) && !ctx->force_generic; }

#define FAST_DISPATCH(N_, EXPR)             \
    do {                                    \
        if ((N_) == 64) {                   \
            constexpr int FN = 64;          \
            EXPR;                           \
        } else {                            \
            constexpr int FN = 32;          \
            EXPR;                           \
        }                                   \
    } while (0)

template <int N>
static cudaError_t fast_setup_n() {
    cudaError_t e;
    const int sm = (int)dq::StabSmem<N>::TOTAL;
    if ((e = cudaFuncSetAttribute(dq::kf_bbar<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(dq::kf_stabilize<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(dq::kf_init_stacks<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(dq::kf_init_green<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm)) != cudaSuccess) return e;
    if (N == 64 && (e = dq::s2::setup()) != cudaSuccess) return e;
    return dq::fast_prop_setup<N>();
}

static inline cudaError_t fast64_setup(dqmcb_ctx* ctx) {
    if (!fast_n(ctx->p.N)) return cudaSuccess;
    cudaError_t e = cudaSuccess;
    FAST_DISPATCH(ctx->p.N, e = fast_setup_n<FN>());
    return e;
}

static inline void fast_bbar(dqmcb_ctx* ctx, int i_stack) {
    const dq::Params& p = ctx->p;
    FAST_DISPATCH(p.N, (dq::kf_bbar<FN><<<p.W, dq::kThreads, dq::StabSmem<FN>::TOTAL, ctx->stream>>>(p, i_stack)));
}
static inline void fast_stabilize(dqmcb_ctx* ctx, int l, int dir, int w0 = 0, int w1 = -1, cudaStream_t st = nullptr) {
    dq::Params p = ctx->p;
    p.w0 = w0;
    p.w1 = (w1 < 0) ? p.W : w1;
    if (!st) st = ctx->stream;
    if (ctx->use_stab2) {  // warp pair per walker, four walkers per SM (stab2.cuh)
        if (dir > 0)
            dq::s2::kf2_stabilize<1><<<p.w1 - p.w0, dq::s2::NTH, dq::s2::Smem::TOTAL, st>>>(p, l);
        else
            dq::s2::kf2_stabilize<-1><<<p.w1 - p.w0, dq::s2::NTH, dq::s2::Smem::TOTAL, st>>>(p, l);
        return;
    }
    FAST_DISPATCH(p.N, (dq::kf_stabilize<FN><<<p.w1 - p.w0, dq::kThreads, dq::StabSmem<FN>::TOTAL, st>>>(p, l, dir)));
}
static inline void fast_init_stacks(dqmcb_ctx* ctx) {
    const dq::Params& p = ctx->p;
    FAST_DISPATCH(p.N, (dq::kf_init_stacks<FN><<<p.W, dq::kThreads, dq::StabSmem<FN>::TOTAL, ctx->stream>>>(p)));
}
static inline void fast_init_green(dqmcb_ctx* ctx) {
    const dq::Params& p = ctx->p;
    FAST_DISPATCH(p.N, (dq::kf_init_green<FN><<<p.W, dq::kThreads, dq::StabSmem<FN>::TOTAL, ctx->stream>>>(p)));
}
static inline int fast64_propagate(dqmcb_ctx* ctx, const dq::Params& p, int l_first, int n_slices, int dir, unsigned long long v0, int flags,
                                   cudaStream_t st = nullptr) {
    if (!st) st = ctx->stream;
    FAST_DISPATCH(p.N, dq::fast_prop_launch<FN>(st, p, l_first, n_slices, dir, v0, flags));
    return cudaGetLastError() == cudaSuccess ? DQMCB_OK : DQMCB_ECUDA;
}
