// tile.cuh -- compile-time-N building blocks of the fast path: FP64 tensor-core (DMMA) products on
// shared-memory tiles, TMA bulk staging of HBM operands, and unrolled QR / LU / triangular solves.
//
// Tile layout: column-major N x N with leading dimension LDT = N + 4 doubles.  With LDT = 4 (mod 8)
// every mma.sync.m8n8k4.f64 fragment load (A: row = lane/4, k = lane%4; B: k = lane%4, col = lane/4;
// also for transposed operands) touches 16 distinct 8-byte banks per half warp: conflict-free.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "linalg.cuh"

namespace dq {

// development aid: set through DQMCB_PHASE_PROF=1 (see dqmcb.cu); stamps from inside the tile routines
__constant__ long long* g_dq_prof = nullptr;
#define DQ_PHG(k)                                                                    \
    do {                                                                             \
        if (g_dq_prof && blockIdx.x == 0 && (threadIdx.x & 31) == 0) g_dq_prof[k] = clock64(); \
    } while (0)

template <int N>
struct Tile {
    static constexpr int LDT = N + 4;
    static constexpr int ELEMS = N * LDT;
    static constexpr size_t BYTES = (size_t)ELEMS * sizeof(double);
};

// D(8x8) += A(8x4) * B(4x8), FP64 tensor core (SASS: DMMA.8x8x4)
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// Warp-level product: acc[(8 MB) x (8 NB)] += A[r0.., 0..K) * B[0..K), c0..)
//   A(r,k) at A[r*rsA + k*csA],  B(k,c) at B[k*rsB + c*csB]   (shared or global, any strides)
// C fragment element j of block (mb, nb): row r0 + 8 mb + lane/4, col c0 + 8 nb + 2 (lane%4) + j.
template <int MB, int NB, int K>
__device__ __forceinline__ void warp_mma(double (&acc)[MB][NB][2], const double* __restrict__ A, int rsA, int csA,
                                         const double* __restrict__ B, int rsB, int csB, int r0, int c0) {
    const int lane = threadIdx.x & 31, gr = lane >> 2, tg = lane & 3;
    const double* ap = A + (size_t)(r0 + gr) * rsA + (size_t)tg * csA;
    const double* bp = B + (size_t)tg * rsB + (size_t)(c0 + gr) * csB;
    double a[2][MB], b[2][NB];
#pragma unroll
    for (int m = 0; m < MB; ++m) a[0][m] = ap[(size_t)(8 * m) * rsA];
#pragma unroll
    for (int n = 0; n < NB; ++n) b[0][n] = bp[(size_t)(8 * n) * csB];
#pragma unroll
    for (int k0 = 0; k0 < K; k0 += 4) {
        const int cur = (k0 >> 2) & 1, nxt = cur ^ 1;
        if (k0 + 4 < K) {
#pragma unroll
            for (int m = 0; m < MB; ++m) a[nxt][m] = ap[(size_t)(8 * m) * rsA + (size_t)(k0 + 4) * csA];
#pragma unroll
            for (int n = 0; n < NB; ++n) b[nxt][n] = bp[(size_t)(k0 + 4) * rsB + (size_t)(8 * n) * csB];
        }
#pragma unroll
        for (int m = 0; m < MB; ++m)
#pragma unroll
            for (int n = 0; n < NB; ++n) dmma884(acc[m][n][0], acc[m][n][1], a[cur][m], b[cur][n]);
    }
}

// Same product restricted to the n-blocks [NB0, NB0 + NBC) of a wider accumulator array (keeps the
// B-fragment double buffer small when a warp owns a wide tile).
template <int MB, int NBTOT, int NB0, int NBC, int K>
__device__ __forceinline__ void warp_mma_cols(double (&acc)[MB][NBTOT][2], const double* __restrict__ A, int rsA, int csA,
                                              const double* __restrict__ B, int rsB, int csB, int r0) {
    const int lane = threadIdx.x & 31, gr = lane >> 2, tg = lane & 3;
    const double* ap = A + (size_t)(r0 + gr) * rsA + (size_t)tg * csA;
    const double* bp = B + (size_t)tg * rsB + (size_t)(8 * NB0 + gr) * csB;
    double a[2][MB], b[2][NBC];
#pragma unroll
    for (int m = 0; m < MB; ++m) a[0][m] = ap[(size_t)(8 * m) * rsA];
#pragma unroll
    for (int n = 0; n < NBC; ++n) b[0][n] = bp[(size_t)(8 * n) * csB];
#pragma unroll
    for (int k0 = 0; k0 < K; k0 += 4) {
        const int cur = (k0 >> 2) & 1, nxt = cur ^ 1;
        if (k0 + 4 < K) {
#pragma unroll
            for (int m = 0; m < MB; ++m) a[nxt][m] = ap[(size_t)(8 * m) * rsA + (size_t)(k0 + 4) * csA];
#pragma unroll
            for (int n = 0; n < NBC; ++n) b[nxt][n] = bp[(size_t)(k0 + 4) * rsB + (size_t)(8 * n) * csB];
        }
#pragma unroll
        for (int m = 0; m < MB; ++m)
#pragma unroll
            for (int n = 0; n < NBC; ++n) dmma884(acc[m][NB0 + n][0], acc[m][NB0 + n][1], a[cur][m], b[cur][n]);
    }
}

template <int MB, int NB>
__device__ __forceinline__ void acc_zero(double (&acc)[MB][NB][2]) {
#pragma unroll
    for (int m = 0; m < MB; ++m)
#pragma unroll
        for (int n = 0; n < NB; ++n) acc[m][n][0] = acc[m][n][1] = 0.0;
}

// Store a warp's accumulator tile: C(r,c) = rs[r] * acc * cs[c] (+ C if ACCUM), column-major ld.
template <int MB, int NB, bool ACCUM>
__device__ __forceinline__ void acc_store(const double (&acc)[MB][NB][2], double* __restrict__ C, int ldc, int r0, int c0,
                                          const double* __restrict__ rs, const double* __restrict__ cs) {
    const int lane = threadIdx.x & 31, gr = lane >> 2, tg = lane & 3;
#pragma unroll
    for (int m = 0; m < MB; ++m) {
        const int r = r0 + 8 * m + gr;
        const double rsv = rs ? rs[r] : 1.0;
#pragma unroll
        for (int n = 0; n < NB; ++n) {
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int c = c0 + 8 * n + 2 * tg + j;
                double v = acc[m][n][j];
                if (rs) v *= rsv;
                if (cs) v *= cs[c];
                double* dst = C + r + (size_t)c * ldc;
                if (ACCUM) v += *dst;
                *dst = v;
            }
        }
    }
}

// CTA-level N x N product with NWARPS warps (256 threads -> 8 warps: 16 x 32 tile per warp for N = 64).
// C = diag(rs) (A B) diag(cs) [+ C].  Caller synchronises before and after.
template <int N, bool ACCUM>
__device__ __forceinline__ void cta_mma(double* __restrict__ C, int ldc, const double* __restrict__ A, int rsA, int csA,
                                        const double* __restrict__ B, int rsB, int csB, const double* __restrict__ rs,
                                        const double* __restrict__ cs) {
    static_assert(N % 32 == 0, "cta_mma tiling assumes N multiple of 32");
    constexpr int MB = 2, NB = 4;  // 16 x 32 per warp
    constexpr int TR = N / 16, TC = N / 32;
    const int warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    for (int t = warp; t < TR * TC; t += nwarps) {
        const int r0 = (t % TR) * 16, c0 = (t / TR) * 32;
        double acc[MB][NB][2];
        acc_zero(acc);
        warp_mma<MB, NB, N>(acc, A, rsA, csA, B, rsB, csB, r0, c0);
        acc_store<MB, NB, ACCUM>(acc, C, ldc, r0, c0, rs, cs);
    }
}

// ------------------------------------------------------------------------------------------------
// TMA bulk staging (cp.async.bulk, SASS UBLKCP): one 1-D bulk copy per matrix column from a dense
// column-major global matrix (ld = N) into a padded shared tile (ld = LDT), completion on an mbarrier.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE;\n"
        "bra WAIT_LOOP;\n"
        "DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }

// Issue the copy of an N x N global matrix (ld = N, 16-byte aligned) into tile `dst` (ld = LDT).
// Called by all threads of the CTA; threads 0..N-1 each issue one column.  The caller must have
// made prior generic-proxy accesses to `dst` visible (barrier + fence_proxy_async) and waits with
// mbar_wait(bar, parity) before reading.
template <int N>
__device__ __forceinline__ void tile_load_async(double* dst, const double* __restrict__ src, uint64_t* bar) {
    constexpr int LDT = Tile<N>::LDT;
    if (threadIdx.x == 0) mbar_expect_tx(bar, (uint32_t)(N * N * sizeof(double)));
    __syncthreads();
    if (threadIdx.x < N) bulk_g2s(dst + (size_t)threadIdx.x * LDT, src + (size_t)threadIdx.x * N, (uint32_t)(N * sizeof(double)), bar);
}

// plain (synchronous) tile <-> global copies, vectorised 16 B
template <int N>
__device__ __forceinline__ void tile_store_global(double* __restrict__ dst, const double* __restrict__ tile) {
    constexpr int LDT = Tile<N>::LDT;
    for (int e = threadIdx.x; e < N * N / 2; e += blockDim.x) {
        const int r = (e % (N / 2)) * 2, c = e / (N / 2);
        const double2 v = *reinterpret_cast<const double2*>(tile + r + (size_t)c * LDT);
        *reinterpret_cast<double2*>(dst + r + (size_t)c * N) = v;
    }
}
template <int N>
__device__ __forceinline__ void tile_load_global(double* __restrict__ tile, const double* __restrict__ src, const double* rs = nullptr,
                                                 const double* cs = nullptr) {
    constexpr int LDT = Tile<N>::LDT;
    for (int e = threadIdx.x; e < N * N / 2; e += blockDim.x) {
        const int r = (e % (N / 2)) * 2, c = e / (N / 2);
        double2 v = *reinterpret_cast<const double2*>(src + r + (size_t)c * N);
        if (rs) {
            v.x *= rs[r];
            v.y *= rs[r + 1];
        }
        if (cs) {
            v.x *= cs[c];
            v.y *= cs[c];
        }
        *reinterpret_cast<double2*>(tile + r + (size_t)c * LDT) = v;
    }
}

// ------------------------------------------------------------------------------------------------
// Column-pivoted Householder QR, compile-time N in {32, 64}, REGISTER resident: warp q owns the
// columns [8q, 8q+8) for the whole factorisation (lane = rows lane, lane+32: 16 doubles per lane).
//
//   * pivoting is implicit: columns never move; jpvt[j] = the column chosen at step j (LAPACK's
//     jpvt) and a chosen column is frozen.  On exit column jpvt[j] of A holds R(0..j, j) in rows
//     <= j and the Householder vector of step j below (unit diagonal implicit), tau[j] as in dgeqp3.
//   * pivot rule: largest partial column norm |a(j:, c)|, as in dgeqp3/dlaqp2.  LAPACK tracks these
//     norms by down-dating and recomputes them exactly whenever cancellation threatens (which, on
//     the strongly graded matrices of the stabilisation, is most of the time); here they are simply
//     recomputed exactly every step (one extra transposing reduction), so no safeguard logic
//     exists.  The pivot order can differ from LAPACK's only on floating-point near-ties.
//   * two block barriers per step; the reflector application needs no shared-memory traffic for the
//     matrix and reduces the eight dot products of a warp with one transposing butterfly
//     (9 shuffles instead of 40).
// norm2[N]: squared partial norms (-1 = column already chosen); vfull[N]: dense reflector;
// wbuf[N]: tau * (v^T a_c).
// ------------------------------------------------------------------------------------------------
template <int N>
__device__ void qrp_tile(double* __restrict__ A, double* __restrict__ tau, int* __restrict__ jpvt, double* __restrict__ norm2,
                         double* __restrict__ vfull, double* __restrict__ wbuf) {
    static_assert(N == 32 || N == 64, "qrp_tile: one warp per 8-column block, two rows per lane");
    constexpr int LDT = Tile<N>::LDT;
    constexpr bool HB = N > 32;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int rA = lane, rB = lane + 32;
    const int cbase = 8 * warp;
    const bool owner = cbase < N;
    const int mycol = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);  // see warp_sum8_transposed
    double ca[8], cb[8];
    unsigned done = 0;  // bit q: column cbase + q already chosen
    if (owner) {
        double s[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const double* cp = A + (size_t)(cbase + q) * LDT;
            ca[q] = cp[rA];
            cb[q] = HB ? cp[rB] : 0.0;
            s[q] = fma(ca[q], ca[q], cb[q] * cb[q]);
        }
        const double tot = warp_sum8_transposed(s);
        if ((lane & 3) == 0) norm2[cbase + mycol] = tot;
    }
    __syncthreads();
    for (int j = 0; j < N; ++j) {
        // ---- pivot (every warp computes it: no barrier needed to publish it)
        const int pc1 = lane + 32;
        const double n0 = norm2[lane], n1 = HB ? norm2[pc1] : -1.0;
        const int bi = warp_argmax2(n0, lane, n0 >= 0.0, n1, pc1, HB && n1 >= 0.0);
        if (owner && (bi >> 3) == warp) {
            // ---- dlarfg on rows j.. of the pivot column
            const int qs = bi & 7;
            double x = 0.0, y = 0.0;
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                x = (q == qs) ? ca[q] : x;
                y = (q == qs) ? cb[q] : y;
            }
            const double alpha = __shfl_sync(DQ_FULL, (j < 32) ? x : y, j & 31);
            const double xs = (rA > j) ? x : 0.0, ys = (HB && rB > j) ? y : 0.0;
            const double xn2 = warp_sum(fma(xs, xs, ys * ys));
            double t = 0.0;
            if (xn2 != 0.0) {
                const double beta = -copysign(sqrt(fma(alpha, alpha, xn2)), alpha);
                t = (beta - alpha) / beta;
                const double scal = 1.0 / (alpha - beta);
                if (rA > j) x *= scal;
                if (rB > j) y *= scal;
                if (rA == j) x = beta;
                if (rB == j) y = beta;
            }
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                ca[q] = (q == qs) ? x : ca[q];
                cb[q] = (q == qs) ? y : cb[q];
            }
            done |= 1u << qs;
            vfull[rA] = (rA > j) ? x : ((rA == j) ? 1.0 : 0.0);
            if (HB) vfull[rB] = (rB > j) ? y : ((rB == j) ? 1.0 : 0.0);
            if (lane == 0) {
                tau[j] = t;
                jpvt[j] = bi;
            }
        }
        __syncthreads();
        // retire the pivot column only now: every warp derives the pivot of this step from norm2 on its own, so
        // the entry must not change before all of them have passed the barrier above (a warp held up for a few
        // hundred cycles -- e.g. by a co-resident CTA in a tensor-core phase -- would otherwise see the column
        // already retired, elect a different pivot and run a second, conflicting reflector generation)
        if (owner && (bi >> 3) == warp && lane == 0) norm2[bi] = -1.0;
        if (owner && done != 0xffu) {
            const double t = tau[j];
            const double vA = vfull[rA], vB = HB ? vfull[rB] : 0.0;
            double pr[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) pr[q] = fma(vA, ca[q], vB * cb[q]);
            const double tot = warp_sum8_transposed(pr);
            if ((lane & 3) == 0) wbuf[cbase + mycol] = t * tot;
            __syncwarp();
            double s[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const double w = ((done >> q) & 1u) ? 0.0 : wbuf[cbase + q];
                ca[q] = fma(-w, vA, ca[q]);
                cb[q] = fma(-w, vB, cb[q]);
                const double ea = (rA > j) ? ca[q] : 0.0, eb = (HB && rB > j) ? cb[q] : 0.0;
                s[q] = fma(ea, ea, eb * eb);
            }
            const double n2 = warp_sum8_transposed(s);  // exact |a(j+1:, c)|^2
            if ((lane & 3) == 0 && !((done >> mycol) & 1u)) norm2[cbase + mycol] = n2;
        }
        __syncthreads();
    }
    if (owner) {
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            double* cp = A + (size_t)(cbase + q) * LDT;
            cp[rA] = ca[q];
            if (HB) cp[rB] = cb[q];
        }
    }
    __syncthreads();
}

// Q = H_0 ... H_{N-1} explicitly (out of place), column-per-warp, barrier-free (see cta_form_q).
template <int N>
__device__ void form_q_tile(double* __restrict__ Q, const double* __restrict__ V, const double* __restrict__ tau,
                            const int* __restrict__ jpvt) {
    static_assert(N <= 64, "form_q_tile holds two rows per lane");
    constexpr int LDT = Tile<N>::LDT;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const int rA = lane, rB = lane + 32;
    const bool hasA = rA < N, hasB = rB < N;
    // four columns per warp pass: c, c+1, c+2, c+3 share every reflector load
    for (int cb = 4 * warp; cb < N; cb += 4 * nwarps) {
        double qa[4], qb[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            qa[u] = (rA == cb + u) ? 1.0 : 0.0;
            qb[u] = (rB == cb + u) ? 1.0 : 0.0;
        }
        const int top = min(cb + 3, N - 1);
        for (int i = top; i >= 0; --i) {
            const double t = tau[i];
            const double* v = V + (size_t)jpvt[i] * LDT;  // implicit pivoting: reflector i sits in the column chosen at step i
            const double vA = (rA > i && hasA) ? v[rA] : ((rA == i) ? 1.0 : 0.0);
            const double vB = (rB > i && hasB) ? v[rB] : ((rB == i) ? 1.0 : 0.0);
            double s[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) s[u] = fma(vA, qa[u], vB * qb[u]);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
                for (int u = 0; u < 4; ++u) s[u] += __shfl_xor_sync(DQ_FULL, s[u], o);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                // H_i only touches columns >= i (e_c is untouched by H_i for i > c)
                const double w = (cb + u >= i) ? t * s[u] : 0.0;
                qa[u] = fma(-w, vA, qa[u]);
                qb[u] = fma(-w, vB, qb[u]);
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (cb + u < N) {
                if (hasA) Q[rA + (size_t)(cb + u) * LDT] = qa[u];
                if (hasB) Q[rB + (size_t)(cb + u) * LDT] = qb[u];
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// LU with partial pivoting (dgetf2 semantics: arma::solve / arma::log_det -> dgetrf), compile-time
// N in {32, 64}, REGISTER resident: warp q owns the columns [8q, 8q+8) (lane = rows lane, lane+32).
//
// Row interchanges are implicit while factorising: a pivoted row is frozen where it is, `perm[j]`
// records the row chosen at step j, and the final store writes row perm[i] to position i, which
// yields exactly dgetrf's in-place layout (unit-lower multipliers below the diagonal, U on and above
// it) -- same pivots, same multipliers, same arithmetic as the interchanging form.
// ONE block barrier per step: the warp owning column j finds the pivot among the still-active rows
// (integer REDUX argmax), publishes the multiplier column through a double-buffered shared vector,
// and every warp applies the rank-1 update to its own register columns; the pivot-row values are
// fetched with one shuffle per column.
// Outputs: factors in A (dgetrf layout), perm[i] = original row at position i, rdiag[j] = 1/U(j,j),
// ipiv is not produced (the solve uses perm).  lbuf: 2 N doubles of scratch.
// ------------------------------------------------------------------------------------------------
template <int N>
__device__ void lu_tile(double* __restrict__ A, int* __restrict__ perm, double* __restrict__ rdiag, int* __restrict__ info,
                        double* __restrict__ lbuf) {
    static_assert(N == 32 || N == 64, "lu_tile: one warp per 8-column block, two rows per lane");
    constexpr int LDT = Tile<N>::LDT;
    constexpr bool HB = N > 32;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int rA = lane, rB = lane + 32;
    const int cbase = 8 * warp;
    const bool owner = cbase < N;
    double ca[8], cb[8];
    bool actA = true, actB = HB;  // my rows not yet chosen as pivot rows
    int posA = 0, posB = 0;       // step at which my rows were chosen = their final position
    if (threadIdx.x == 0) *info = 0;
    if (owner) {
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const double* cp = A + (size_t)(cbase + q) * LDT;
            ca[q] = cp[rA];
            cb[q] = HB ? cp[rB] : 0.0;
        }
    }
    __syncthreads();  // everybody has read A; *info initialised
    for (int j = 0; j < N; ++j) {
        double* lb = lbuf + (j & 1) * (N + 2);  // [0..N-1] multipliers, [N] pivot row (as double)
        if (owner && (j >> 3) == warp) {
            const int qs = j & 7;
            double x = 0.0, y = 0.0;
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                x = (q == qs) ? ca[q] : x;
                y = (q == qs) ? cb[q] : y;
            }
            const int p = warp_argmax2(fabs(x), rA, actA, fabs(y), rB, actB);
            const double piv = __shfl_sync(DQ_FULL, (p < 32) ? x : y, p & 31);
            const double rinv = (piv != 0.0) ? 1.0 / piv : 0.0;
            // multipliers of the rows that stay active; they are also the L entries of column j
            const double lA = (actA && rA != p) ? x * rinv : 0.0;
            const double lB = (actB && rB != p) ? y * rinv : 0.0;
            lb[rA] = lA;
            if (HB) lb[rB] = lB;
            if (actA && rA != p) x = lA;
            if (actB && rB != p) y = lB;
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                ca[q] = (q == qs) ? x : ca[q];
                cb[q] = (q == qs) ? y : cb[q];
            }
            if (lane == 0) {
                lb[N] = (double)p;
                perm[j] = p;
                rdiag[j] = rinv;
                if (piv == 0.0 && *info == 0) *info = j + 1;
            }
        }
        __syncthreads();
        const int p = (int)lb[N];
        if (rA == p) {
            actA = false;
            posA = j;
        }
        if (rB == p) {
            actB = false;
            posB = j;
        }
        if (owner && cbase + 7 > j) {
            const double lA = lb[rA], lB = HB ? lb[rB] : 0.0;  // zero for frozen rows and for row p
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const double u = __shfl_sync(DQ_FULL, (p < 32) ? ca[q] : cb[q], p & 31);
                if (cbase + q > j) {
                    ca[q] = fma(-lA, u, ca[q]);
                    cb[q] = fma(-lB, u, cb[q]);
                }
            }
        }
    }
    // store in dgetrf's row order: original row r goes to position pos(r)
    __syncthreads();
    if (owner) {
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            double* cp = A + (size_t)(cbase + q) * LDT;
            cp[posA] = ca[q];
            if (HB) cp[posB] = cb[q];
        }
    }
    __syncthreads();
}

// dgetrs 'N' in place on an N x N right-hand side: element (i, c) at Bm[i*rsB + c*csB].
// Two adjacent lanes own one right-hand-side column (even / odd rows, N/2 values each in registers;
// threads 0..2N-1, i.e. four warps for N = 64).  The triangular factors are read as warp-wide
// broadcasts; the only exchange per elimination step is one shuffle between the two lanes of a
// column.  Uses perm / rdiag from lu_tile.  Caller synchronises before and after.
template <int N>
__device__ __noinline__ void lu_solve_tile(const double* __restrict__ A, const int* __restrict__ perm, const double* __restrict__ rdiag,
                                           double* __restrict__ Bm, int rsB, int csB) {
    static_assert(N % 2 == 0 && N <= 64, "lu_solve_tile layout");
    constexpr int LDT = Tile<N>::LDT;
    constexpr int H = N / 2;
    const int t = threadIdx.x;
    if (t >= 2 * N) return;  // whole warps only (2N is a multiple of 32 for N = 32, 64)
    const int c = t >> 1, h = t & 1, lane = t & 31;
    const int pair = lane & ~1;
    double x[H];  // x[m] = row 2m + h
    double* col = Bm + (size_t)c * csB;
#pragma unroll
    for (int m = 0; m < H; ++m) x[m] = col[(size_t)perm[2 * m + h] * rsB];
    // forward: L y = P b (unit diagonal)
#pragma unroll
    for (int k = 0; k < N - 1; ++k) {
        const double xk = __shfl_sync(DQ_FULL, x[k >> 1], pair | (k & 1));
        const double* lk = A + (size_t)k * LDT + h;
        if ((k & 1) == 0) {
            if (h) x[k >> 1] = fma(-lk[k], xk, x[k >> 1]);  // row k+1 (odd lane)
        }
#pragma unroll
        for (int m = (k >> 1) + 1; m < H; ++m) x[m] = fma(-lk[2 * m], xk, x[m]);
    }
    // backward: U x = y
#pragma unroll
    for (int k = N - 1; k >= 0; --k) {
        const double xk = __shfl_sync(DQ_FULL, x[k >> 1] * rdiag[k], pair | (k & 1));
        if (h == (k & 1)) x[k >> 1] = xk;
        const double* uk = A + (size_t)k * LDT + h;
#pragma unroll
        for (int m = 0; m < (k >> 1); ++m) x[m] = fma(-uk[2 * m], xk, x[m]);
        if (k & 1) {
            if (!h) x[k >> 1] = fma(-uk[k - 1], xk, x[k >> 1]);  // row k-1 (even lane)
        }
    }
#pragma unroll
    for (int m = 0; m < H; ++m) col[(size_t)(2 * m + h) * rsB] = x[m];
}

}  // namespace dq
