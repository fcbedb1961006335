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

constexpr int kGT_N = 256, kGT_LDS = 260, kGT_KP = 16, kGT_CP = 64;
constexpr size_t kGemmTmaArenaBytes = 64 + (size_t)kGT_LDS * (kGT_CP + 2 * kGT_KP) * sizeof(double);

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
                    double* dst = C + r + (size_t)c * ldc;
                    if (accumulate && !preload) v += *dst;
                    *dst = v;
                }
        }
    }
    __syncthreads();
    if (tid == 0) {
        mbar_inval(bar);
        mbar_inval(bar + 1);
        mbar_inval(bar + 2);
    }
    __syncthreads();
}

}  // namespace dq
