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
//
// PANEL form: the warp owning columns [8 pb, 8 pb + 8) factorises its whole panel on its own -- pivot
// search (integer REDUX argmax), reciprocal, multipliers and the rank-1 updates of its remaining panel
// columns need warp-level exchange only -- and publishes the eight multiplier columns and pivot rows
// through a double-buffered shared panel; ONE block barrier per panel (N/8 per factorisation instead of
// N), after which every warp to the right applies the eight rank-1 updates to its own register columns
// in step order (so every element sees the same operations in the same order as the column-at-a-time
// form).  The panel buffer lives in the storage of tile A itself, which is dead between the initial
// load and the final store.
// Outputs: factors in A (dgetrf layout), perm[i] = original row at position i, rdiag[j] = 1/U(j,j),
// ipiv is not produced (the solve uses perm).  lbuf: >= 16 ints of scratch (pivot rows of two panels).
// ------------------------------------------------------------------------------------------------
template <int N>
__device__ void lu_tile(double* __restrict__ A, int* __restrict__ perm, double* __restrict__ rdiag, int* __restrict__ info,
                        double* __restrict__ lbuf) {
    static_assert(N == 32 || N == 64, "lu_tile: one warp per 8-column block, two rows per lane");
    constexpr int LDT = Tile<N>::LDT;
    constexpr bool HB = N > 32;
    static_assert(2 * 8 * N <= Tile<N>::ELEMS, "panel buffers must fit in the tile");
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int rA = lane, rB = lane + 32;
    const int cbase = 8 * warp;
    const bool owner = cbase < N;
    int* pvbuf = reinterpret_cast<int*>(lbuf);
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
    __syncthreads();  // everybody has read A (its storage now carries the panels); *info initialised
    for (int pb = 0; pb < N / 8; ++pb) {
        double* Lp = A + (size_t)(pb & 1) * 8 * N;  // [jj][row] multipliers of step 8 pb + jj
        int* pv = pvbuf + (pb & 1) * 8;             // pivot row of step 8 pb + jj
        if (warp == pb) {
#pragma unroll
            for (int jj = 0; jj < 8; ++jj) {
                const int j = 8 * pb + jj;
                const double x = ca[jj], y = cb[jj];
                const int p = warp_argmax2(fabs(x), rA, actA, fabs(y), rB, actB);
                const double piv = __shfl_sync(DQ_FULL, (p < 32) ? x : y, p & 31);
                const double rinv = (piv != 0.0) ? 1.0 / piv : 0.0;
                // multipliers of the rows that stay active; they are also the L entries of column j
                const bool kA = actA && rA != p, kB = actB && rB != p;
                const double lA = kA ? x * rinv : 0.0;
                const double lB = kB ? y * rinv : 0.0;
                Lp[jj * N + rA] = lA;
                if (HB) Lp[jj * N + rB] = lB;
                if (kA) ca[jj] = lA;
                if (kB) cb[jj] = lB;
                if (rA == p) {
                    actA = false;
                    posA = j;
                }
                if (rB == p) {
                    actB = false;
                    posB = j;
                }
                if (lane == 0) {
                    pv[jj] = p;
                    perm[j] = p;
                    rdiag[j] = rinv;
                    if (piv == 0.0 && *info == 0) *info = j + 1;
                }
#pragma unroll
                for (int q = jj + 1; q < 8; ++q) {
                    const double u = __shfl_sync(DQ_FULL, (p < 32) ? ca[q] : cb[q], p & 31);
                    ca[q] = fma(-lA, u, ca[q]);
                    cb[q] = fma(-lB, u, cb[q]);
                }
            }
        }
        __syncthreads();
        if (warp != pb) {
            int pj[8];
#pragma unroll
            for (int jj = 0; jj < 8; ++jj) {
                pj[jj] = pv[jj];
                if (rA == pj[jj]) {
                    actA = false;
                    posA = 8 * pb + jj;
                }
                if (rB == pj[jj]) {
                    actB = false;
                    posB = 8 * pb + jj;
                }
            }
            if (owner && warp > pb) {
                double lA[8], lB[8];  // zero for rows frozen before the step and for its pivot row
#pragma unroll
                for (int jj = 0; jj < 8; ++jj) {
                    lA[jj] = Lp[jj * N + rA];
                    lB[jj] = HB ? Lp[jj * N + rB] : 0.0;
                }
#pragma unroll
                for (int jj = 0; jj < 8; ++jj) {
                    const int p = pj[jj];
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        const double u = __shfl_sync(DQ_FULL, (p < 32) ? ca[q] : cb[q], p & 31);
                        ca[q] = fma(-lA[jj], u, ca[q]);
                        cb[q] = fma(-lB[jj], u, cb[q]);
                    }
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

// dgetrs 'N' in place on an N x N right-hand side: element (i, c) at Bm[i*rsB + c*csB].  BLOCKED by
// 8 rows: the 8 x 8 diagonal block is solved in registers, one thread per right-hand-side column
// (seven dependent FMAs), and the remaining block rows are updated with rank-8 FP64 tensor-core
// products (two DMMA.8x8x4 per 8 x 8 block, the warp's B fragment loaded once per block row);
// two block barriers per block row -- 4 N / 8 per solve instead of a 2 N step serial chain.
// Uses perm / rdiag from lu_tile.  All threads of the CTA must call; caller synchronises before and after.
template <int N>
__device__ __noinline__ void lu_solve_tile(const double* __restrict__ A, const int* __restrict__ perm, const double* __restrict__ rdiag,
                                           double* __restrict__ Bm, int rsB, int csB) {
    static_assert(N == 32 || N == 64, "lu_solve_tile layout");
    constexpr int LDT = Tile<N>::LDT;
    constexpr int NBK = N / 8;
    const int t = threadIdx.x, nthr = blockDim.x;
    const int lane = t & 31, warp = t >> 5, nwarps = nthr >> 5, gr = lane >> 2, tg = lane & 3;
    // ---- b <- P b: every element is read before any is written.  The unit-stride index of the
    // right-hand side runs across the lanes (fast), eight entries of the other index per thread.
    {
        constexpr int NPARTS = kThreads / N;
        constexpr int PASSES = (N + NPARTS * 8 - 1) / (NPARTS * 8);
        const int f = t % N, part = t / N;
        const bool rows_fast = (rsB == 1);
        double v[PASSES][8];
#pragma unroll
        for (int ps = 0; ps < PASSES; ++ps) {
            const int s0 = (ps * NPARTS + part) * 8;
            if (s0 < N) {
#pragma unroll
                for (int i = 0; i < 8; ++i)
                    v[ps][i] = rows_fast ? Bm[(size_t)perm[f] + (size_t)(s0 + i) * csB] : Bm[(size_t)perm[s0 + i] * rsB + (size_t)f];
            }
        }
        __syncthreads();
#pragma unroll
        for (int ps = 0; ps < PASSES; ++ps) {
            const int s0 = (ps * NPARTS + part) * 8;
            if (s0 < N) {
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    if (rows_fast)
                        Bm[(size_t)f + (size_t)(s0 + i) * csB] = v[ps][i];
                    else
                        Bm[(size_t)(s0 + i) * rsB + (size_t)f] = v[ps][i];
                }
            }
        }
        __syncthreads();
    }
    // ---- forward: L y = P b (unit diagonal)
    for (int kb = 0; kb < NBK; ++kb) {
        const int k0 = 8 * kb;
        if (t < N) {
            double* col = Bm + (size_t)k0 * rsB + (size_t)t * csB;
            double x[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) x[i] = col[(size_t)i * rsB];
#pragma unroll
            for (int j = 0; j < 7; ++j) {
                const double* lj = A + k0 + (size_t)(k0 + j) * LDT;
#pragma unroll
                for (int i = j + 1; i < 8; ++i) x[i] = fma(-lj[i], x[j], x[i]);
            }
#pragma unroll
            for (int i = 1; i < 8; ++i) col[(size_t)i * rsB] = x[i];
        }
        if (kb + 1 == NBK) break;
        __syncthreads();
        const int ntask = (NBK - 1 - kb) * NBK;
        for (int task = warp; task < ntask; task += nwarps) {
            const int cb = task % NBK, rb = kb + 1 + task / NBK;
            const double* ap = A + (8 * rb + gr) + (size_t)(k0 + tg) * LDT;
            const double* bp = Bm + (size_t)(k0 + tg) * rsB + (size_t)(8 * cb + gr) * csB;
            double* cp = Bm + (size_t)(8 * rb + gr) * rsB + (size_t)(8 * cb + 2 * tg) * csB;
            const double a0 = -ap[0], a1 = -ap[4 * LDT];
            const double b0 = bp[0], b1 = bp[(size_t)4 * rsB];
            double c0 = cp[0], c1 = cp[csB];
            dmma884(c0, c1, a0, b0);
            dmma884(c0, c1, a1, b1);
            cp[0] = c0;
            cp[csB] = c1;
        }
        __syncthreads();
    }
    __syncthreads();
    // ---- backward: U x = y
    for (int kb = NBK - 1; kb >= 0; --kb) {
        const int k0 = 8 * kb;
        if (t < N) {
            double* col = Bm + (size_t)k0 * rsB + (size_t)t * csB;
            double x[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) x[i] = col[(size_t)i * rsB];
#pragma unroll
            for (int j = 7; j >= 0; --j) {
                const double* uj = A + k0 + (size_t)(k0 + j) * LDT;
                x[j] *= rdiag[k0 + j];
#pragma unroll
                for (int i = 0; i < j; ++i) x[i] = fma(-uj[i], x[j], x[i]);
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) col[(size_t)i * rsB] = x[i];
        }
        if (kb == 0) break;
        __syncthreads();
        const int ntask = kb * NBK;
        for (int task = warp; task < ntask; task += nwarps) {
            const int cb = task % NBK, rb = task / NBK;
            const double* ap = A + (8 * rb + gr) + (size_t)(k0 + tg) * LDT;
            const double* bp = Bm + (size_t)(k0 + tg) * rsB + (size_t)(8 * cb + gr) * csB;
            double* cp = Bm + (size_t)(8 * rb + gr) * rsB + (size_t)(8 * cb + 2 * tg) * csB;
            const double a0 = -ap[0], a1 = -ap[4 * LDT];
            const double b0 = bp[0], b1 = bp[(size_t)4 * rsB];
            double c0 = cp[0], c1 = cp[csB];
            dmma884(c0, c1, a0, b0);
            dmma884(c0, c1, a1, b1);
            cp[0] = c0;
            cp[csB] = c1;
        }
        __syncthreads();
    }
}

}  // namespace dq
