// gemm_tma.cuh -- N = 256 product of the stabilisation kernels with both operands staged in shared memory by TMA.
//
// The matrices of one 16 x 16-lattice walker (512 KB each) do not fit in an SM; the register-ring product of linalg.cuh
// (cta_gemm_dmma_big) therefore gathers every DMMA fragment from L2 with 8-byte loads: each warp re-reads the whole B
// panel and a 32-row strip of A for each of its tiles, 8.4 MB of L2 traffic per product in 32-byte sectors.  Here one
// thread streams the operands with cp.async.bulk (SASS UBLKCP), whole 2 KB columns at a time:
//   * a 256 x 64 panel of B stays in shared memory for a 256 x 64 panel of C (one copy per panel instead of one per warp);
//   * A streams through two 256 x 16 k-panels (double-buffered, completion on mbarriers), shared by all 8 warps;
//   * each warp owns 32 rows x 64 columns of the C panel: 64 accumulators per thread, every A fragment feeds 8 DMMA,
//     every B fragment 4.
// L2 -> SM traffic per product: 4 x (512 KB of A + 128 KB of B) = 2.6 MB, in full lines.
// Shared tiles are padded to LDS = 260 doubles (4 mod 8): conflict-free fragment loads (tile.cuh).
// Requires: n == 256, blockDim.x == 256, A / B / C column-major with 16-byte-aligned columns, C aliasing neither operand.
#pragma once
#include "tile.cuh"

namespace dq {

#ifdef DQ_TMA_PROF
__device__ long long g_tma_prof[8];
#define DQ_TP(k, t0) do { if (threadIdx.x == 0 && blockIdx.x == 0) g_tma_prof[k] += clock64() - (t0); } while (0)
#define DQ_T0() clock64()
#else
#define DQ_TP(k, t0) do { } while (0)
#define DQ_T0() 0ll
#endif

constexpr int kGT_N = 256, kGT_LDS = 260, kGT_KP = 16, kGT_CP = 64;
constexpr size_t kGemmTmaPanelBytes = (size_t)kGT_LDS * (kGT_CP + 2 * kGT_KP) * sizeof(double);
constexpr size_t kGemmTmaArenaBytes = 64 + kGemmTmaPanelBytes + kGT_N * sizeof(double);  // barriers, panels, reciprocal diagonal of the solve

__device__ __forceinline__ void mbar_inval(uint64_t* bar) { asm volatile("mbarrier.inval.shared::cta.b64 [%0];\n" ::"r"(smem_u32(bar)) : "memory"); }

// C(r,c) [+]= rs[r] * (A B)(r,c) * cs[c]
__device__ __noinline__ void cta_gemm_tma256(double* __restrict__ C, int ldc, const double* __restrict__ A, int lda,
                                             const double* __restrict__ B, int ldb, const double* __restrict__ rs,
                                             const double* __restrict__ cs, bool accumulate, unsigned char* arena) {
    constexpr int N = kGT_N, LDS = kGT_LDS, KP = kGT_KP, CP = kGT_CP;
    uint64_t* bar = reinterpret_cast<uint64_t*>(arena);  // [0]: B panel, [1], [2]: A k-panels
    double* Bs = reinterpret_cast<double*>(arena + 64);
    double* As[2] = {Bs + (size_t)LDS * CP, Bs + (size_t)LDS * (CP + KP)};
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, gr = lane >> 2, tg = lane & 3;
    const int r0 = 32 * warp;
    __syncthreads();  // every earlier generic-proxy use of the arena is over
    if (tid == 0) {
        mbar_init(bar, 1);
        mbar_init(bar + 1, 1);
        mbar_init(bar + 2, 1);
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    }
    __syncthreads();
    uint32_t pB = 0, pA0 = 0, pA1 = 0;
    const bool preload = accumulate && !rs && !cs;
    for (int c0 = 0; c0 < N; c0 += CP) {
        if (tid == 0) {
            fence_proxy_async();
            mbar_expect_tx(bar, (uint32_t)(CP * N * sizeof(double)));
            for (int c = 0; c < CP; ++c) bulk_g2s(Bs + (size_t)c * LDS, B + (size_t)(c0 + c) * ldb, (uint32_t)(N * sizeof(double)), bar);
            mbar_expect_tx(bar + 1, (uint32_t)(KP * N * sizeof(double)));
            for (int k = 0; k < KP; ++k) bulk_g2s(As[0] + (size_t)k * LDS, A + (size_t)k * lda, (uint32_t)(N * sizeof(double)), bar + 1);
        }
        double acc[4][8][2];
#pragma unroll
        for (int m = 0; m < 4; ++m)
#pragma unroll
            for (int q = 0; q < 8; ++q)
#pragma unroll
                for (int j = 0; j < 2; ++j) acc[m][q][j] = preload ? C[(r0 + 8 * m + gr) + (size_t)(c0 + 8 * q + 2 * tg + j) * ldc] : 0.0;
        mbar_wait(bar, pB);
        pB ^= 1;
        for (int kp = 0; kp < N / KP; ++kp) {
            const int buf = kp & 1;
            if (tid == 0 && kp + 1 < N / KP) {  // the other buffer was released by the barrier that ended the previous iteration
                fence_proxy_async();
                mbar_expect_tx(bar + 1 + (buf ^ 1), (uint32_t)(KP * N * sizeof(double)));
                const double* src = A + (size_t)(kp + 1) * KP * lda;
                double* dst = As[buf ^ 1];
                for (int k = 0; k < KP; ++k) bulk_g2s(dst + (size_t)k * LDS, src + (size_t)k * lda, (uint32_t)(N * sizeof(double)), bar + 1 + (buf ^ 1));
            }
            if (buf == 0) {
                mbar_wait(bar + 1, pA0);
                pA0 ^= 1;
            } else {
                mbar_wait(bar + 2, pA1);
                pA1 ^= 1;
            }
            const double* ap = As[buf] + (r0 + gr) + (size_t)tg * LDS;
            const double* bp = Bs + (kp * KP + tg) + (size_t)gr * LDS;
#pragma unroll
            for (int ks = 0; ks < KP / 4; ++ks) {
                double a[4], b[8];
#pragma unroll
                for (int m = 0; m < 4; ++m) a[m] = ap[8 * m + (size_t)(4 * ks) * LDS];
#pragma unroll
                for (int q = 0; q < 8; ++q) b[q] = bp[4 * ks + (size_t)(8 * q) * LDS];
#pragma unroll
                for (int m = 0; m < 4; ++m)
#pragma unroll
                    for (int q = 0; q < 8; ++q) dq_dmma884(acc[m][q][0], acc[m][q][1], a[m], b[q]);
            }
            __syncthreads();  // all warps are done with As[buf] (and, after the last k-panel, with Bs)
        }
        // scalings, then (accumulating with scalings) all loads of C before the first store: independent L2 accesses instead of
        // a load-add-store chain per element
#pragma unroll
        for (int m = 0; m < 4; ++m) {
            const int r = r0 + 8 * m + gr;
            const double rsv = rs ? rs[r] : 1.0;
#pragma unroll
            for (int q = 0; q < 8; ++q)
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    const int c = c0 + 8 * q + 2 * tg + j;
                    double v = acc[m][q][j];
                    if (rs) v *= rsv;
                    if (cs) v *= cs[c];
                    if (accumulate && !preload) v += C[r + (size_t)c * ldc];
                    acc[m][q][j] = v;
                }
        }
#pragma unroll
        for (int m = 0; m < 4; ++m)
#pragma unroll
            for (int q = 0; q < 8; ++q)
#pragma unroll
                for (int j = 0; j < 2; ++j) C[(r0 + 8 * m + gr) + (size_t)(c0 + 8 * q + 2 * tg + j) * ldc] = acc[m][q][j];
    }
    __syncthreads();
    if (tid == 0) {
        mbar_inval(bar);
        mbar_inval(bar + 1);
        mbar_inval(bar + 2);
    }
    __syncthreads();
}

// ------------------------------------------------------------------------------------------------
// Solve (L U) X = B in place for N = 256 (dgetrs 'N' on a right-hand side whose rows are already interchanged), same
// arithmetic as cta_lu_solve_blocked (16-row block steps: 16 x 16 diagonal block by substitution, one thread per column;
// the other rows by a rank-16 tensor-core update with C preloaded into the accumulators) but organised around shared
// memory: the right-hand-side columns are independent, so they are taken 64 at a time and such a 256 x 64 panel stays
// in shared memory for its whole forward and backward substitution, while the 16-column panels of L and U stream in by
// cp.async.bulk (double-buffered, one step ahead).  L2 traffic per solve: 4 x 0.5 MB of factors + the right-hand side
// once in, once out (the blocked version re-reads and re-writes the trailing right-hand side at each of its 32 steps).
// A: cta_lu output, column-major, lda even, 16-byte aligned; Bm column-major (ldb), 256 right-hand sides.
// ------------------------------------------------------------------------------------------------
__device__ __noinline__ void cta_lu_solve_tma256(const double* __restrict__ A, int lda, double* __restrict__ Bm, int ldb,
                                                 unsigned char* arena) {
    constexpr int N = kGT_N, LDS = kGT_LDS, CP = kGT_CP, NBLK = N / 16;
    uint64_t* bar = reinterpret_cast<uint64_t*>(arena);  // [1], [2]: factor panels
    double* Xs = reinterpret_cast<double*>(arena + 64);
    double* rinv = reinterpret_cast<double*>(arena + 64 + kGemmTmaPanelBytes);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, gr = lane >> 2, tg = lane & 3;
    constexpr int ISSUER = 224;  // lane 0 of the last warp: it has no part in the substitution phase, which its copies overlap
    __syncthreads();
    if (tid == ISSUER) {
        mbar_init(bar + 1, 1);
        mbar_init(bar + 2, 1);
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    }
    rinv[tid] = 1.0 / A[(size_t)tid + (size_t)tid * lda];  // 16 dependent divisions per backward step otherwise
    __syncthreads();
    // step s of a right-hand-side panel: s < 16 forward with block kb = s (rows k0.. of L's columns k0..k0+15),
    // s >= 16 backward with kb = 31 - s (rows 0..k0+15 of U's columns)
#define DQ_SOLVE_ISSUE(g_)                                                                                                    \
    do {                                                                                                                      \
        const int s_ = (g_)&31, buf_ = (g_)&1;                                                                                \
        const bool fwd_ = s_ < NBLK;                                                                                          \
        const int k0_ = (fwd_ ? s_ : 2 * NBLK - 1 - s_) << 4;                                                                 \
        const int row0_ = fwd_ ? k0_ : 0, rows_ = fwd_ ? N - k0_ : k0_ + 16;                                                  \
        double* dst_ = Xs + (size_t)LDS * (CP + kGT_KP * buf_);                                                               \
        fence_proxy_async();                                                                                                  \
        mbar_expect_tx(bar + 1 + buf_, (uint32_t)(16 * rows_ * sizeof(double)));                                              \
        for (int k = 0; k < 16; ++k)                                                                                          \
            bulk_g2s(dst_ + (size_t)k * LDS, A + (size_t)row0_ + (size_t)(k0_ + k) * lda, (uint32_t)(rows_ * sizeof(double)), \
                     bar + 1 + buf_);                                                                                         \
    } while (0)
    uint32_t par0 = 0, par1 = 0;
    const int total = (N / CP) * 2 * NBLK;
    if (tid == ISSUER) DQ_SOLVE_ISSUE(0);
    for (int g = 0; g < total; ++g) {
        const int s = g & 31, buf = g & 1, c0 = (g >> 5) * CP;
        const bool fwd = s < NBLK;
        const int k0 = (fwd ? s : 2 * NBLK - 1 - s) << 4;
        if (s == 0) {  // bring the next 64 right-hand sides in (the previous panel was stored behind a barrier)
            for (int e = tid; e < CP * (N / 2); e += blockDim.x) {
                const int c = e / (N / 2), r2 = e % (N / 2);
                *reinterpret_cast<double2*>(Xs + 2 * r2 + (size_t)c * LDS) = *reinterpret_cast<const double2*>(Bm + 2 * r2 + (size_t)(c0 + c) * ldb);
            }
        }
        [[maybe_unused]] long long tp_ = DQ_T0();
        if (buf == 0) {
            mbar_wait(bar + 1, par0);
            par0 ^= 1;
        } else {
            mbar_wait(bar + 2, par1);
            par1 ^= 1;
        }
        DQ_TP(1, tp_);
        __syncthreads();  // right-hand sides of the previous step's update (or the panel load) are in place
        tp_ = DQ_T0();
        const double* P = Xs + (size_t)LDS * (CP + kGT_KP * buf);
        const int d0 = fwd ? 0 : k0;  // row of the diagonal block inside the staged panel
        if (tid == ISSUER && g + 1 < total) DQ_SOLVE_ISSUE(g + 1);  // its buffer was released by the barrier that ended step g - 1
        if (tid < CP) {
            double* col = Xs + k0 + (size_t)tid * LDS;
            double x[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) x[i] = col[i];
            if (fwd) {
#pragma unroll
                for (int j = 0; j < 15; ++j) {
                    const double* lj = P + d0 + (size_t)j * LDS;
#pragma unroll
                    for (int i = j + 1; i < 16; ++i) x[i] = fma(-lj[i], x[j], x[i]);
                }
#pragma unroll
                for (int i = 1; i < 16; ++i) col[i] = x[i];
            } else {
#pragma unroll
                for (int j = 15; j >= 0; --j) {
                    const double* uj = P + d0 + (size_t)j * LDS;
                    x[j] = x[j] * rinv[k0 + j];
#pragma unroll
                    for (int i = 0; i < j; ++i) x[i] = fma(-uj[i], x[j], x[i]);
                }
#pragma unroll
                for (int i = 0; i < 16; ++i) col[i] = x[i];
            }
        }
        __syncthreads();
        DQ_TP(2, tp_);
        tp_ = DQ_T0();
        // rank-16 update of the rows still to be solved: forward rows k0+16.., backward rows 0..k0-1
        const int ur0 = fwd ? k0 + 16 : 0, um = fwd ? N - k0 - 16 : k0;
        const int pr0 = fwd ? 16 : 0;  // row of ur0 inside the staged panel
        const int tr = um >> 4;
        if (warp < tr) {
            // one warp per 16-row strip, all 64 columns: the fragments of the solved block are the same for every strip
            double b[4][8];
#pragma unroll
            for (int kb = 0; kb < 4; ++kb)
#pragma unroll
                for (int q = 0; q < 8; ++q) b[kb][q] = Xs[(k0 + 4 * kb + tg) + (size_t)(8 * q + gr) * LDS];
            for (int strip = warp; strip < tr; strip += 8) {
                const int rt = strip << 4;
                double acc[2][8][2], a[4][2];
#pragma unroll
                for (int mm = 0; mm < 2; ++mm)
#pragma unroll
                    for (int q = 0; q < 8; ++q)
#pragma unroll
                        for (int j = 0; j < 2; ++j) acc[mm][q][j] = Xs[(ur0 + rt + 8 * mm + gr) + (size_t)(8 * q + 2 * tg + j) * LDS];
#pragma unroll
                for (int kb = 0; kb < 4; ++kb) {
                    a[kb][0] = -P[(pr0 + rt + gr) + (size_t)(4 * kb + tg) * LDS];
                    a[kb][1] = -P[(pr0 + rt + 8 + gr) + (size_t)(4 * kb + tg) * LDS];
                }
#pragma unroll
                for (int kb = 0; kb < 4; ++kb)
#pragma unroll
                    for (int mm = 0; mm < 2; ++mm)
#pragma unroll
                        for (int q = 0; q < 8; ++q) dq_dmma884(acc[mm][q][0], acc[mm][q][1], a[kb][mm], b[kb][q]);
#pragma unroll
                for (int mm = 0; mm < 2; ++mm)
#pragma unroll
                    for (int q = 0; q < 8; ++q)
#pragma unroll
                        for (int j = 0; j < 2; ++j) Xs[(ur0 + rt + 8 * mm + gr) + (size_t)(8 * q + 2 * tg + j) * LDS] = acc[mm][q][j];
            }
        }
        if (s == 2 * NBLK - 1) {  // panel solved: write it back
            __syncthreads();
            for (int e = tid; e < CP * (N / 2); e += blockDim.x) {
                const int c = e / (N / 2), r2 = e % (N / 2);
                *reinterpret_cast<double2*>(Bm + 2 * r2 + (size_t)(c0 + c) * ldb) = *reinterpret_cast<const double2*>(Xs + 2 * r2 + (size_t)c * LDS);
            }
        }
        __syncthreads();  // Ps[buf] and the updated rows are settled before the next step
        DQ_TP(3, tp_);
    }
    if (tid == ISSUER) {
        mbar_inval(bar + 1);
        mbar_inval(bar + 2);
    }
    __syncthreads();
#undef DQ_SOLVE_ISSUE
}

// D = S^T for 256 x 256 column-major matrices, 32 x 32 blocks through padded shared-memory tiles (one per warp, in the
// arena): both the global reads and the global writes are 256-byte rows of a block.
__device__ __forceinline__ void cta_transpose256(double* __restrict__ D, int ldd, const double* __restrict__ S, int lds, unsigned char* arena) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double* t = reinterpret_cast<double*>(arena + 64) + (size_t)warp * 32 * 33;
    __syncthreads();
    for (int blk = warp; blk < 64; blk += 8) {
        const int r0 = (blk & 7) * 32, c0 = (blk >> 3) * 32;
#pragma unroll 8
        for (int k = 0; k < 32; ++k) t[lane + 33 * k] = S[(r0 + lane) + (size_t)(c0 + k) * lds];  // t[r][c]
        __syncwarp();
#pragma unroll 8
        for (int k = 0; k < 32; ++k) D[(c0 + lane) + (size_t)(r0 + k) * ldd] = t[k + 33 * lane];  // D(c, r) = S(r, c)
        __syncwarp();
    }
    __syncthreads();
}

}  // namespace dq
