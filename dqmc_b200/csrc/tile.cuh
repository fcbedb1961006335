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
// Column-pivoted Householder QR (dlaqp2 semantics, see linalg.cuh::cta_qrp) for compile-time N <= 64,
// N multiple of 32 not required; each lane owns rows lane and lane+32 of a column.
// ------------------------------------------------------------------------------------------------
template <int N>
__device__ void qrp_tile(double* __restrict__ A, double* __restrict__ tau, int* __restrict__ jpvt, double* __restrict__ vn1,
                         double* __restrict__ vn2) {
    static_assert(N <= 64, "qrp_tile holds two rows per lane");
    constexpr int LDT = Tile<N>::LDT;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const double tol3z = 1.0536712127723509e-08;
    const int rA = lane, rB = lane + 32;
    const bool hasB = rB < N, hasA = rA < N;
    for (int c = warp; c < N; c += nwarps) {
        const double x = hasA ? A[rA + (size_t)c * LDT] : 0.0, y = hasB ? A[rB + (size_t)c * LDT] : 0.0;
        const double s = warp_sum(fma(x, x, y * y));
        if (lane == 0) {
            const double nr = sqrt(s);
            vn1[c] = nr;
            vn2[c] = nr;
            jpvt[c] = c;
        }
    }
    __syncthreads();
    for (int j = 0; j < N; ++j) {
        double* colj = A + (size_t)j * LDT;
        if (warp == 0) {
            double best = -1.0;
            int bi = j;
            {
                const int c0 = j + lane, c1 = j + lane + 32;
                if (c0 < N) {
                    best = vn1[c0];
                    bi = c0;
                }
                if (c1 < N) {
                    const double v = vn1[c1];
                    if (v > best) {
                        best = v;
                        bi = c1;
                    }
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const double ob = __shfl_xor_sync(DQ_FULL, best, o);
                const int oi = __shfl_xor_sync(DQ_FULL, bi, o);
                if (ob > best || (ob == best && oi < bi)) {
                    best = ob;
                    bi = oi;
                }
            }
            double x = hasA ? colj[rA] : 0.0, y = hasB ? colj[rB] : 0.0;
            if (bi != j) {
                double* colp = A + (size_t)bi * LDT;
                const double px = hasA ? colp[rA] : 0.0, py = hasB ? colp[rB] : 0.0;
                if (hasA) colp[rA] = x;
                if (hasB) colp[rB] = y;
                x = px;
                y = py;
                if (lane == 0) {
                    const int tj = jpvt[j];
                    jpvt[j] = jpvt[bi];
                    jpvt[bi] = tj;
                    vn1[bi] = vn1[j];
                    vn2[bi] = vn2[j];
                }
            }
            // dlarfg on rows j.. of the pivot column (held in x, y)
            const double alpha = __shfl_sync(DQ_FULL, (j < 32) ? x : y, j & 31);
            const double xs = (rA > j) ? x : 0.0, ys = (rB > j && hasB) ? y : 0.0;
            const double xnorm = sqrt(warp_sum(fma(xs, xs, ys * ys)));
            double t = 0.0;
            if (xnorm != 0.0) {
                const double beta = -copysign(hypot(alpha, xnorm), alpha);
                t = (beta - alpha) / beta;
                const double scal = 1.0 / (alpha - beta);
                if (rA > j) x *= scal;
                if (rB > j) y *= scal;
                if (rA == j) x = beta;
                if (rB == j) y = beta;
            }
            if (hasA) colj[rA] = x;
            if (hasB) colj[rB] = y;
            if (lane == 0) tau[j] = t;
        }
        __syncthreads();
        const double t = tau[j];
        const double vA = (rA > j && hasA) ? colj[rA] : 0.0, vB = (rB > j && hasB) ? colj[rB] : 0.0;
        // two columns per pass for ILP on the shuffle reductions
        for (int c = j + 1 + 2 * warp; c < N; c += 2 * nwarps) {
            const bool two = c + 1 < N;
            double* c0p = A + (size_t)c * LDT;
            double* c1p = A + (size_t)(two ? c + 1 : c) * LDT;
            double a0 = hasA ? c0p[rA] : 0.0, b0 = hasB ? c0p[rB] : 0.0;
            double a1 = hasA ? c1p[rA] : 0.0, b1 = hasB ? c1p[rB] : 0.0;
            double s0 = fma(vA, a0, vB * b0), s1 = fma(vA, a1, vB * b1);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                s0 += __shfl_xor_sync(DQ_FULL, s0, o);
                s1 += __shfl_xor_sync(DQ_FULL, s1, o);
            }
            const double aj0 = __shfl_sync(DQ_FULL, (j < 32) ? a0 : b0, j & 31);
            const double aj1 = __shfl_sync(DQ_FULL, (j < 32) ? a1 : b1, j & 31);
            const double w0 = t * (aj0 + s0), w1 = t * (aj1 + s1);
            a0 = fma(-w0, vA, a0);
            b0 = fma(-w0, vB, b0);
            a1 = fma(-w1, vA, a1);
            b1 = fma(-w1, vB, b1);
            const double an0 = aj0 - w0, an1 = aj1 - w1;
            if (rA == j) {
                a0 = an0;
                a1 = an1;
            }
            if (rB == j) {
                b0 = an0;
                b1 = an1;
            }
            if (rA >= j && hasA) {
                c0p[rA] = a0;
                if (two) c1p[rA] = a1;
            }
            if (rB >= j && hasB) {
                c0p[rB] = b0;
                if (two) c1p[rB] = b1;
            }
            // partial-norm down-date: lane 0 -> column c, lane 1 -> column c+1
            int recompute = 0;
            if (lane < 2 && (lane == 0 || two)) {
                const int cc = c + lane;
                const double anew = lane ? an1 : an0;
                const double v1 = vn1[cc];
                if (v1 != 0.0) {
                    double temp = fabs(anew) / v1;
                    temp = fmax(1.0 - temp * temp, 0.0);
                    const double q = v1 / vn2[cc];
                    if (temp * (q * q) <= tol3z)
                        recompute = 1;
                    else
                        vn1[cc] = v1 * sqrt(temp);
                }
            }
            const unsigned rmask = __ballot_sync(DQ_FULL, recompute);
            if (rmask) {
                const double xs0 = (rA > j) ? a0 : 0.0, ys0 = (rB > j && hasB) ? b0 : 0.0;
                const double xs1 = (rA > j) ? a1 : 0.0, ys1 = (rB > j && hasB) ? b1 : 0.0;
                const double n0 = sqrt(warp_sum(fma(xs0, xs0, ys0 * ys0)));
                const double n1 = sqrt(warp_sum(fma(xs1, xs1, ys1 * ys1)));
                if (lane == 0 && (rmask & 1)) {
                    vn1[c] = n0;
                    vn2[c] = n0;
                }
                if (lane == 1 && (rmask & 2)) {
                    vn1[c + 1] = n1;
                    vn2[c + 1] = n1;
                }
            }
        }
        __syncthreads();
    }
}

// Q = H_0 ... H_{N-1} explicitly (out of place), column-per-warp, barrier-free (see cta_form_q).
template <int N>
__device__ void form_q_tile(double* __restrict__ Q, const double* __restrict__ V, const double* __restrict__ tau) {
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
            const double* v = V + (size_t)i * LDT;
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
// LU with partial pivoting (dgetf2 semantics, see linalg.cuh::cta_lu), compile-time N <= 64.
// ------------------------------------------------------------------------------------------------
template <int N>
__device__ void lu_tile(double* __restrict__ A, int* __restrict__ ipiv, double* __restrict__ prow, double* __restrict__ jrow,
                        int* __restrict__ info) {
    static_assert(N <= 64, "lu_tile holds two rows per lane");
    constexpr int LDT = Tile<N>::LDT;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const int rA = lane, rB = lane + 32;
    const bool hasA = rA < N, hasB = rB < N;
    if (threadIdx.x == 0) *info = 0;
    int prev_p = 0;
    double prev_rinv = 0.0, prev_jrowj = 0.0;
    for (int j = 0; j <= N; ++j) {
        if (warp == 0) {
            if (j > 0) {  // deferred scaling of column j-1 (rows >= j)
                double* colq = A + (size_t)(j - 1) * LDT;
                if (rA >= j && hasA) colq[rA] = ((rA == prev_p) ? prev_jrowj : colq[rA]) * prev_rinv;
                if (rB >= j && hasB) colq[rB] = ((rB == prev_p) ? prev_jrowj : colq[rB]) * prev_rinv;
            }
            if (j < N) {
                const double* colj = A + (size_t)j * LDT;
                double best = -1.0;
                int bi = j;
                if (rA >= j && hasA) {
                    best = fabs(colj[rA]);
                    bi = rA;
                }
                if (rB >= j && hasB) {
                    const double v = fabs(colj[rB]);
                    if (v > best) {
                        best = v;
                        bi = rB;
                    }
                }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    const double ob = __shfl_xor_sync(DQ_FULL, best, o);
                    const int oi = __shfl_xor_sync(DQ_FULL, bi, o);
                    if (ob > best || (ob == best && oi < bi)) {
                        best = ob;
                        bi = oi;
                    }
                }
                if (lane == 0) {
                    ipiv[j] = bi;
                    if (best == 0.0 && *info == 0) *info = j + 1;
                }
            }
        }
        if (j == N) break;
        __syncthreads();
        const int p = ipiv[j];
        for (int c = threadIdx.x; c < N; c += blockDim.x) {
            prow[c] = A[p + (size_t)c * LDT];
            jrow[c] = A[j + (size_t)c * LDT];
        }
        __syncthreads();
        const double piv = prow[j];
        const double rinv = (piv != 0.0) ? 1.0 / piv : 0.0;
        const double jrowj = jrow[j];
        for (int c = threadIdx.x; c < N; c += blockDim.x) {
            A[j + (size_t)c * LDT] = prow[c];
            if (p != j && c < j) A[p + (size_t)c * LDT] = jrow[c];
        }
        const double* colj = A + (size_t)j * LDT;
        const double lA = (rA > j && hasA) ? ((rA == p) ? jrowj : colj[rA]) * rinv : 0.0;
        const double lB = (rB > j && hasB) ? ((rB == p) ? jrowj : colj[rB]) * rinv : 0.0;
        for (int c = j + 1 + warp; c < N; c += nwarps) {
            double* colc = A + (size_t)c * LDT;
            const double pc = prow[c], jc = jrow[c];
            if (rA > j && hasA) colc[rA] = fma(-lA, pc, (rA == p) ? jc : colc[rA]);
            if (rB > j && hasB) colc[rB] = fma(-lB, pc, (rB == p) ? jc : colc[rB]);
        }
        prev_p = p;
        prev_rinv = rinv;
        prev_jrowj = jrowj;
        __syncthreads();
    }
    __syncthreads();
}

// dgetrs 'N' in place on an N x N right-hand side: element (i, c) at Bm[i*rsB + c*csB].
// A team of 4 lanes owns one RHS column (rows i = 4 m + q in registers); requires blockDim = 4 N.
template <int N>
__device__ void lu_solve_tile(const double* __restrict__ A, const int* __restrict__ ipiv, double* __restrict__ Bm, int rsB, int csB) {
    static_assert(N % 4 == 0 && N <= 64, "lu_solve_tile layout");
    constexpr int LDT = Tile<N>::LDT;
    constexpr int M = N / 4;
    const int lane = threadIdx.x & 31;
    const int q = threadIdx.x & 3, c = threadIdx.x >> 2;
    if (c < N) {  // row interchanges, one thread per column does the (cheap, serial) dlaswp
        if (q == 0) {
            double* col = Bm + (size_t)c * csB;
            for (int j = 0; j < N; ++j) {
                const int p = ipiv[j];
                if (p != j) {
                    const double tmp = col[(size_t)j * rsB];
                    col[(size_t)j * rsB] = col[(size_t)p * rsB];
                    col[(size_t)p * rsB] = tmp;
                }
            }
        }
    }
    __syncwarp();
    const bool act = c < N;
    double x[M];
#pragma unroll
    for (int m = 0; m < M; ++m) x[m] = act ? Bm[(size_t)(4 * m + q) * rsB + (size_t)c * csB] : 0.0;
    const int team_base = lane & ~3;
    // forward: L y = b (unit diagonal)
#pragma unroll
    for (int k = 0; k < N - 1; ++k) {
        const double xk = __shfl_sync(DQ_FULL, x[k >> 2], team_base | (k & 3));
        const double* lk = A + (size_t)k * LDT;
#pragma unroll
        for (int m = (k + 1) >> 2; m < M; ++m) {
            const int i = 4 * m + q;
            if (i > k) x[m] = fma(-lk[i], xk, x[m]);
        }
    }
    // backward: U x = y
#pragma unroll
    for (int k = N - 1; k >= 0; --k) {
        const double* uk = A + (size_t)k * LDT;
        double xk = __shfl_sync(DQ_FULL, x[k >> 2], team_base | (k & 3));
        xk = xk / uk[k];
        if ((k & 3) == q) x[k >> 2] = xk;
#pragma unroll
        for (int m = 0; m <= (k >> 2); ++m) {
            const int i = 4 * m + q;
            if (i < k) x[m] = fma(-uk[i], xk, x[m]);
        }
    }
    if (act) {
#pragma unroll
        for (int m = 0; m < M; ++m) Bm[(size_t)(4 * m + q) * rsB + (size_t)c * csB] = x[m];
    }
}

}  // namespace dq
