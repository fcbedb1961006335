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
// TMA bulk prefetch of `bytes` (multiple of 16) of global memory into L2 (SASS UBLKPF): issued by one thread
// long before a matrix is needed, so that the later fetch / operand loads hit L2 instead of HBM.
__device__ __forceinline__ void bulk_prefetch_l2(const void* gsrc, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;\n" ::"l"(gsrc), "r"(bytes) : "memory");
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
// Column-pivoted Householder QR, compile-time N in {32, 64}, REGISTER resident.
//
// Layout: TPC = 256 / N adjacent lanes own one column (RPT = N / TPC consecutive rows each: 16 doubles
// per thread for N = 64), so a column dot product is RPT thread-local FMAs plus log2(TPC) shuffle
// stages, and the whole CTA works on all N columns at once.
//
//   * pivoting is implicit: columns never move; jpvt[j] = the column chosen at step j (LAPACK's
//     jpvt) and a chosen column is frozen.  On exit column jpvt[j] of A holds R(0..j, j) in rows
//     <= j and the Householder vector of step j below (unit diagonal implicit), tau[j] as in dgeqp3.
//   * pivot rule: largest partial column norm |a(j:, c)|, as in dgeqp3/dlaqp2.  LAPACK tracks these
//     norms by down-dating and recomputes them exactly whenever cancellation threatens (which, on
//     the strongly graded matrices of the stabilisation, is most of the time); here they are simply
//     recomputed exactly every step, so no safeguard logic exists.  The pivot order can differ from
//     LAPACK's only on floating-point near-ties.
//   * a finished row is RETIRED: at step j the entry R(j, c) of every trailing column goes straight to
//     its place in the shared tile (whose storage is dead while the matrix lives in registers) and the
//     register is zeroed.  Registers therefore hold exact zeros in rows <= j, and neither the dot
//     products nor the norms nor the published pivot column need row masks.  The step loop is unrolled
//     over the RPT rows of a thread, so "row j" is a compile-time register index.
//   * the owners of the elected column publish it RAW; beta, tau and the scaling of the reflector are
//     derived from it by every thread on its own (branch-free sqrt / reciprocal), overlapping the dot
//     products.  The weight tau (v^T a_c) is computed redundantly by the TPC lanes of a column: no
//     shared-memory exchange besides the pivot column.  Two block barriers per step.
// norm2[N]: squared partial norms (-1 = column already chosen); xraw[(RPT + 2) TPC]: raw pivot column,
// chunk of RPT rows per lane-of-column padded to RPT + 2 (conflict-free 16-byte loads).
// ------------------------------------------------------------------------------------------------
template <int N>
__device__ void qrp_tile(double* __restrict__ A, double* __restrict__ tau, int* __restrict__ jpvt, double* __restrict__ norm2,
                         double* __restrict__ xraw) {
    static_assert(N == 32 || N == 64, "qrp_tile: 256 threads, 256 / N lanes per column");
    constexpr int LDT = Tile<N>::LDT;
    constexpr bool HB = N > 32;
    constexpr int TPC = kThreads / N;  // lanes per column
    constexpr int RPT = N / TPC;       // rows per lane
    constexpr int XS = RPT + 2;        // padded chunk stride of the published column
    static_assert(RPT % 2 == 0, "16-byte loads of the published column");
    const int lane = threadIdx.x & 31;
    const int c = threadIdx.x / TPC, sub = threadIdx.x % TPC;
    const int grp = lane & ~(TPC - 1);
    const unsigned gmask = ((1u << TPC) - 1u) << grp;
    double a[RPT];
    {
        const double* cp = A + (size_t)c * LDT + sub * RPT;
        double s0 = 0.0, s1 = 0.0;
#pragma unroll
        for (int k = 0; k < RPT; k += 2) {
            const double2 v = *reinterpret_cast<const double2*>(cp + k);
            a[k] = v.x;
            a[k + 1] = v.y;
            s0 = fma(v.x, v.x, s0);
            s1 = fma(v.y, v.y, s1);
        }
        double sq = s0 + s1;
#pragma unroll
        for (int o = 1; o < TPC; o <<= 1) sq += __shfl_xor_sync(gmask, sq, o);
        if (sub == 0) norm2[c] = sq;
    }
    bool done = false;    // my column already chosen
    int mystep = N;       // ... at this step
    double myscal = 0.0;  // ... and the scaling of its Householder vector
    __syncthreads();
    for (int jb = 0; jb < TPC; ++jb) {
#pragma unroll
        for (int k = 0; k < RPT; ++k) {
            const int j = jb * RPT + k;  // row j = register k of the lanes with sub == jb
            // ---- pivot (every warp computes it: no barrier needed to publish it)
            const int pc1 = lane + 32;
            const double n0 = norm2[lane], n1 = HB ? norm2[pc1] : -1.0;
            const int bi = warp_argmax2_64(n0, lane, n0 >= 0.0, n1, pc1, HB && n1 >= 0.0);
            const double n2 = norm2[bi];  // exact |a(j:, bi)|^2 (recomputed after the previous reflector)
            const bool mine = (c == bi);
            if (mine) {
                // publish the RAW column (rows < j are exact zeros); everything else about the reflector is
                // derived from it by the readers
                double* xp = xraw + sub * XS;
#pragma unroll
                for (int k2 = 0; k2 < RPT; k2 += 2) *reinterpret_cast<double2*>(xp + k2) = make_double2(a[k2], a[k2 + 1]);
                if (sub == 0) jpvt[j] = bi;
                done = true;
                mystep = j;
            }
            __syncthreads();
            // retire the pivot column only now: every warp derives the pivot of this step from norm2 on its own, so
            // the entry must not change before all of them have passed the barrier above (a warp held up for a few
            // hundred cycles -- e.g. by a co-resident CTA in a tensor-core phase -- would otherwise see the column
            // already retired and elect a different pivot)
            if (mine && sub == 0) norm2[bi] = -1.0;
            if (!done || mine) {
                const double alpha = xraw[jb * XS + k];
                // ---- dlarfg scalars (every thread; a latency chain that overlaps the dot products):
                // beta = -sign(alpha) nrm, tau = (beta - alpha) / beta = 1 + |alpha| / nrm,
                // v = x / (alpha - beta) = x sign(alpha) / (|alpha| + nrm) below the diagonal.
                // A column that is already zero below the diagonal is still reflected (tau = 2, v = e_j: the sign
                // of row j flips, LAPACK would take H = I) -- the same factorisation up to the sign freedom of
                // Q / R rows; only an exactly zero column takes H = I.
                const double nn = fmax(n2, alpha * alpha);
                const bool refl = nn > 0.0;
                double hrs;
                const double nrm = sqrt_newton(refl ? nn : 1.0, &hrs);
                const double beta = refl ? -copysign(nrm, alpha) : alpha;
                const double tt = refl ? fma(fabs(alpha), hrs + hrs, 1.0) : 0.0;
                const double scal = refl ? copysign(rcp_newton(fabs(alpha) + nrm), alpha) : 0.0;
                if (!mine) {
                    const double* xp = xraw + sub * XS;
                    double x[RPT];
#pragma unroll
                    for (int k2 = 0; k2 < RPT; k2 += 2) {
                        const double2 v = *reinterpret_cast<const double2*>(xp + k2);
                        x[k2] = v.x;
                        x[k2 + 1] = v.y;
                    }
                    if (sub == jb) x[k] = 0.0;  // row j itself enters through a(j, c) below
                    double d0 = 0.0, d1 = 0.0, d2 = 0.0, d3 = 0.0;
#pragma unroll
                    for (int k2 = 0; k2 < RPT; k2 += 4) {
                        d0 = fma(x[k2], a[k2], d0);
                        d1 = fma(x[k2 + 1], a[k2 + 1], d1);
                        d2 = fma(x[k2 + 2], a[k2 + 2], d2);
                        d3 = fma(x[k2 + 3], a[k2 + 3], d3);
                    }
                    double dot = (d0 + d1) + (d2 + d3);  // x(j+1:)^T a(j+1:, c), this lane's rows
#pragma unroll
                    for (int o = 1; o < TPC; o <<= 1) dot += __shfl_xor_sync(gmask, dot, o);
                    const double aj = __shfl_sync(gmask, a[k], grp | jb);  // a(j, c)
                    const double w = tt * fma(scal, dot, aj);               // tau (v^T a_c)
                    const double ws = w * scal;
#pragma unroll
                    for (int k2 = 0; k2 < RPT; ++k2) a[k2] = fma(-ws, x[k2], a[k2]);
                    if (sub == jb) {
                        A[j + (size_t)c * LDT] = a[k] - w;  // R(j, c) is final: retire the row
                        a[k] = 0.0;
                    }
                    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
#pragma unroll
                    for (int k2 = 0; k2 < RPT; k2 += 4) {
                        s0 = fma(a[k2], a[k2], s0);
                        s1 = fma(a[k2 + 1], a[k2 + 1], s1);
                        s2 = fma(a[k2 + 2], a[k2 + 2], s2);
                        s3 = fma(a[k2 + 3], a[k2 + 3], s3);
                    }
                    double sq = (s0 + s1) + (s2 + s3);  // exact |a(j+1:, c)|^2
#pragma unroll
                    for (int o = 1; o < TPC; o <<= 1) sq += __shfl_xor_sync(gmask, sq, o);
                    if (sub == 0) norm2[c] = sq;
                } else {
                    // my column is the pivot column: it becomes (R(0..j-1, j) [already in the tile], beta, v(j+1:));
                    // the scaling of v = x / (alpha - beta) is applied when the column is stored at the end
                    myscal = scal;
                    if (sub == jb) {
                        A[j + (size_t)c * LDT] = beta;
                        a[k] = 0.0;
                    }
                    if (sub == 0) tau[j] = tt;
                }
            }
            __syncthreads();
        }
    }
    // rows below the diagonal entry of each column come from the registers (the rows above were retired into the tile)
    {
        double* cp = A + (size_t)c * LDT + sub * RPT;
#pragma unroll
        for (int k = 0; k < RPT; ++k)
            if (sub * RPT + k > mystep) cp[k] = a[k] * myscal;
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
    // A column's cost grows with its index (column c takes reflectors c .. 0): the passes pair the short
    // column groups of a warp with long ones (4 w, then N - 4 - 4 w, ...), so that every warp applies
    // about the same number of reflectors and the closing barrier waits for nobody in particular.
    const int ngroups = N / 4;
    for (int gi = warp; gi < ngroups; gi += nwarps) {
        const int pass = gi / nwarps;
        const int cb = 4 * ((pass & 1) ? (pass + 1) * nwarps - 1 - warp : gi);
        if (cb >= N || cb < 0) continue;
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
            int bad = 0;
#pragma unroll
            for (int jj = 0; jj < 8; ++jj) {
                const int j = 8 * pb + jj;
                const double x = ca[jj], y = cb[jj];
                const int p = warp_argmax2_64(fabs(x), rA, actA, fabs(y), rB, actB);
                const double piv = __shfl_sync(DQ_FULL, (p < 32) ? x : y, p & 31);
                const double rinv = (piv != 0.0) ? rcp_newton(piv) : 0.0;
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
                }
                bad = (piv == 0.0 && bad == 0) ? j + 1 : bad;  // first exactly singular pivot of this panel
#pragma unroll
                for (int q = jj + 1; q < 8; ++q) {
                    const double u = __shfl_sync(DQ_FULL, (p < 32) ? ca[q] : cb[q], p & 31);
                    ca[q] = fma(-lA, u, ca[q]);
                    cb[q] = fma(-lB, u, cb[q]);
                }
            }
            if (bad != 0) {  // warp-uniform, rare; panels run one after the other, so the first one to report wins
                if (lane == 0 && *info == 0) *info = bad;
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

// dgetrs 'N' in place on an N x N right-hand side: element (i, c) at Bm[i*rsB + c*csB].
//
// The N right-hand-side columns are independent, so warp q solves the columns [8q, 8q+8) completely on
// its own -- NO block barrier, no shared-memory traffic for the right-hand side: the warp's 64 x 8 block
// lives in registers in the FP64 tensor-core accumulator layout (row block rb: lane (gr, tg) holds rows
// 8 rb + gr, columns 2 tg, 2 tg + 1).  Per 8-row block: the 8 x 8 triangular system is solved by
// substitution across the lanes (one shuffle + one FMA per row step), the solved block is turned into a
// DMMA B fragment with four shuffles, and every remaining row block gets its rank-8 update as two
// DMMA.8x8x4 straight on the register accumulators (A fragments = blocks of L / U from the shared tile).
// Backward substitution keeps the unscaled values on the dependent chain (U entries pre-multiplied by the
// reciprocal pivots, which does not depend on the data).  Uses perm / rdiag from lu_tile.  All threads
// of the CTA may call; the caller synchronises before and after.
template <int N>
__device__ __forceinline__ void lu_solve_tile(const double* __restrict__ A, const int* __restrict__ perm, const double* __restrict__ rdiag,
                                           double* __restrict__ Bm, int rsB, int csB) {
    static_assert(N == 32 || N == 64, "lu_solve_tile layout");
    constexpr int LDT = Tile<N>::LDT;
    constexpr int NBK = N / 8;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, gr = lane >> 2, tg = lane & 3;
    if (8 * warp >= N) return;  // whole warps only
    double* bcol = Bm + (size_t)(8 * warp + 2 * tg) * csB;
    double c[NBK][2];
    // ---- b <- P b while loading (the warp reads and writes only its own columns)
#pragma unroll
    for (int rb = 0; rb < NBK; ++rb) {
        const size_t ro = (size_t)perm[8 * rb + gr] * rsB;
        c[rb][0] = bcol[ro];
        c[rb][1] = bcol[ro + csB];
    }
    const int src0 = (tg << 2) | (gr >> 1), src1 = ((tg + 4) << 2) | (gr >> 1);
    const bool odd = gr & 1;
    // ---- forward: L y = P b (unit diagonal)
#pragma unroll
    for (int kb = 0; kb < NBK; ++kb) {
        const int k0 = 8 * kb;
        double l[7];
#pragma unroll
        for (int jj = 0; jj < 7; ++jj) l[jj] = A[(k0 + gr) + (size_t)(k0 + jj) * LDT];  // row gr of the diagonal block
#pragma unroll
        for (int jj = 0; jj < 7; ++jj) {
            const double v0 = __shfl_sync(DQ_FULL, c[kb][0], (jj << 2) | tg);
            const double v1 = __shfl_sync(DQ_FULL, c[kb][1], (jj << 2) | tg);
            if (gr > jj) {
                c[kb][0] = fma(-l[jj], v0, c[kb][0]);
                c[kb][1] = fma(-l[jj], v1, c[kb][1]);
            }
        }
        if (kb + 1 < NBK) {
            const double e00 = __shfl_sync(DQ_FULL, c[kb][0], src0), e01 = __shfl_sync(DQ_FULL, c[kb][1], src0);
            const double e10 = __shfl_sync(DQ_FULL, c[kb][0], src1), e11 = __shfl_sync(DQ_FULL, c[kb][1], src1);
            const double b0 = odd ? e01 : e00, b1 = odd ? e11 : e10;
#pragma unroll
            for (int rb = kb + 1; rb < NBK; ++rb) {
                const double* ap = A + (8 * rb + gr) + (size_t)(k0 + tg) * LDT;
                dmma884(c[rb][0], c[rb][1], -ap[0], b0);
                dmma884(c[rb][0], c[rb][1], -ap[4 * LDT], b1);
            }
        }
    }
    // ---- backward: U x = y
#pragma unroll
    for (int kb = NBK - 1; kb >= 0; --kb) {
        const int k0 = 8 * kb;
        const double rd = rdiag[k0 + gr];
        double u[8];
#pragma unroll
        for (int jj = 1; jj < 8; ++jj) u[jj] = A[(k0 + gr) + (size_t)(k0 + jj) * LDT] * rdiag[k0 + jj];  // U(gr, jj) / U(jj, jj)
#pragma unroll
        for (int jj = 7; jj >= 1; --jj) {
            const double v0 = __shfl_sync(DQ_FULL, c[kb][0], (jj << 2) | tg);  // y_jj before its division by the pivot
            const double v1 = __shfl_sync(DQ_FULL, c[kb][1], (jj << 2) | tg);
            if (gr < jj) {
                c[kb][0] = fma(-u[jj], v0, c[kb][0]);
                c[kb][1] = fma(-u[jj], v1, c[kb][1]);
            }
        }
        c[kb][0] *= rd;
        c[kb][1] *= rd;
        if (kb > 0) {
            const double e00 = __shfl_sync(DQ_FULL, c[kb][0], src0), e01 = __shfl_sync(DQ_FULL, c[kb][1], src0);
            const double e10 = __shfl_sync(DQ_FULL, c[kb][0], src1), e11 = __shfl_sync(DQ_FULL, c[kb][1], src1);
            const double b0 = odd ? e01 : e00, b1 = odd ? e11 : e10;
#pragma unroll
            for (int rb = 0; rb < kb; ++rb) {
                const double* ap = A + (8 * rb + gr) + (size_t)(k0 + tg) * LDT;
                dmma884(c[rb][0], c[rb][1], -ap[0], b0);
                dmma884(c[rb][0], c[rb][1], -ap[4 * LDT], b1);
            }
        }
    }
#pragma unroll
    for (int rb = 0; rb < NBK; ++rb) {
        const size_t ro = (size_t)(8 * rb + gr) * rsB;
        bcol[ro] = c[rb][0];
        bcol[ro + csB] = c[rb][1];
    }
}

}  // namespace dq
