// stab2.cuh -- end-of-block stabilisation for the 8 x 8 lattice (N = 64), second generation.
//
// Replaces, like kf_stabilize (fast64.cuh): calculate_Bbar + update_stack_* (mat_mul_ldr / ldr_mul_mat -> to_LDR) +
// stabilize_GF_* (inv_I_plus_ldr[_mul_ldr]) + check_error  (source/dqmc.cpp:88-105,134-161,189-215,317-329;
// source/stablelinalg.cpp:35-158).
//
// Design.  kf_stabilize spreads ONE walker over 256 threads; its factorisations (pivoted QR, LU) are then chains of
// barriers, shuffles and reductions that leave the FP64 pipe 85 % idle, and registers (256 x 128) cap an SM at two
// walkers, so nothing else runs under those waits (ncu: barrier stall 30 %, 23 % warps active).  Here a walker is a
// WARP PAIR (64 threads, one CTA), four walkers per SM:
//   * factorisations are THREAD-SERIAL over a whole column / row held in registers (64 doubles per thread):
//       QR     thread c owns column c: the reflector dot product, the update and the new exact column norm are
//              thread-local FMA loops -- no shuffle at all; per step one elected column is published through shared
//              memory (two 64-thread barriers);
//       Q      thread c builds column c of Q = H_0 .. H_c e_c from the published reflectors (broadcast loads);
//       LU     thread r owns row r (dgetf2 order, implicit row interchange): each warp elects its pivot candidate with
//              two REDUX and publishes that row speculatively, ONE barrier per column;
//       solve  thread c owns right-hand side c: forward / backward substitution without any synchronisation;
//     the inner loops run over 8-row register blocks with compile-time bounds (the row inside a block is selected with
//     predicated moves), which keeps the code at ~1/6 of a full unroll;
//   * the block product is never formed: the Kronecker chain B_l = diag(e^V) (Ey (x) Ex) is applied straight to
//     L_prev D_prev (forward) or, transposed, to (D_next R_next)^T (backward) in place on the walker's single tile --
//     one GEMM and one operand fetch fewer than (Bbar L) D, same flops as forming Bbar;
//   * the four remaining dense products run on the FP64 tensor cores (DMMA.8x8x4) with the left operand in the tile and
//     the right operand streamed from L2 in 8-column chunks by cp.async (double buffered);
//   * one 34 KB tile + 18 KB of buffers per walker -> 4 CTAs / SM, 255 registers per thread.
// Needs the Kronecker factors (dqmcb_create verified product form AND symmetry); any other model takes kf_stabilize.
#pragma once
#include "dqmcb_internal.cuh"
#include "engine.cuh"
#include "tile.cuh"

namespace dq {
namespace s2 {

constexpr int N = 64;
constexpr int LDT = 68;   // tile leading dimension (= 4 mod 8: conflict-free DMMA fragments)
constexpr int NTH = 64;   // threads per walker
constexpr int CW = 8;     // columns per streamed chunk
constexpr int MAX_SLICES = 24;

struct Smem {
    static constexpr size_t TILE = (size_t)N * LDT * 8;              // 34816
    static constexpr size_t CHUNK = (size_t)CW * LDT * 8;            // 4352
    static constexpr size_t ARENA = 3 * CHUNK;                       // 2 B chunks + conversion buffer; aliased by phase
    static constexpr size_t SMALL = (size_t)9 * N * 8 + (size_t)3 * N * 4 + 16;
    static constexpr size_t TOTAL = TILE + ARENA + SMALL;
};
static_assert((size_t)MAX_SLICES * N * 8 <= Smem::ARENA, "slice scalings must fit in the arena");

struct Wk {
    double* T;       // tile
    double* arena;   // phase-dependent
    double *tau, *dnew, *dinv, *sv[4], *rdiag, *red;
    int *jpvt, *perm, *info;
};

__device__ __forceinline__ void carve(Wk& wk, unsigned char* smem) {
    wk.T = reinterpret_cast<double*>(smem);
    wk.arena = reinterpret_cast<double*>(smem + Smem::TILE);
    double* s = reinterpret_cast<double*>(smem + Smem::TILE + Smem::ARENA);
    wk.tau = s;
    wk.dnew = s + N;
    wk.dinv = s + 2 * N;
    for (int i = 0; i < 4; ++i) wk.sv[i] = s + (3 + i) * N;
    wk.rdiag = s + 7 * N;
    wk.red = s + 8 * N;
    int* ip = reinterpret_cast<int*>(s + 9 * N);
    wk.jpvt = ip;
    wk.perm = ip + N;
    wk.info = ip + 2 * N;
}

// element jj (run time, 0..7) of an 8-register block: a select tree of depth three (a chain of seven dependent
// 64-bit selects sat on the critical path of every substitution / elimination step)
__device__ __forceinline__ double sel8(const double (&a)[N], int base, int jj) {
    const bool b0 = jj & 1, b1 = jj & 2, b2 = jj & 4;
    const double t0 = b0 ? a[base + 1] : a[base], t1 = b0 ? a[base + 3] : a[base + 2];
    const double t2 = b0 ? a[base + 5] : a[base + 4], t3 = b0 ? a[base + 7] : a[base + 6];
    const double u0 = b1 ? t1 : t0, u1 = b1 ? t3 : t2;
    return b2 ? u1 : u0;
}
__device__ __forceinline__ void put8(double (&a)[N], int base, int jj, double val) {
#pragma unroll
    for (int q = 0; q < 8; ++q) a[base + q] = (jj == q) ? val : a[base + q];
}

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int NLEFT>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;\n" ::"n"(NLEFT) : "memory");
}

// ---------------------------------------------------------------------------------------------------------------
// dense global (ld = N) -> tile, T(r, c) = src(r, c) * rs[r] * cs[c]  (rs / cs may be null); coalesced 16-byte loads.
// Operands may have been written earlier in this launch by other threads of the CTA (new stack entry, scratch): every
// such read goes to L2 (ld.global.cg), never through the non-coherent path.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void load_tile(double* __restrict__ T, const double* __restrict__ src, const double* rs, const double* cs) {
    const int tid = threadIdx.x;
#pragma unroll 1
    for (int it0 = 0; it0 < 32; it0 += 8) {
        double2 v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int e = (it0 + u) * NTH + tid;  // pair index: 32 pairs per column
            v[u] = __ldcg(reinterpret_cast<const double2*>(src + 2 * (e & 31) + (size_t)(e >> 5) * N));
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int e = (it0 + u) * NTH + tid;
            const int r = 2 * (e & 31), c = e >> 5;
            double2 x = v[u];
            if (rs) {
                x.x *= rs[r];
                x.y *= rs[r + 1];
            }
            if (cs) {
                const double s = cs[c];
                x.x *= s;
                x.y *= s;
            }
            *reinterpret_cast<double2*>(T + r + (size_t)c * LDT) = x;
        }
    }
}
// dense global (ld = N) -> tile TRANSPOSED, T(c, r) = src(r, c) * rs[r] * cs[c]
__device__ __forceinline__ void load_tile_t(double* __restrict__ T, const double* __restrict__ src, const double* rs, const double* cs) {
    const int tid = threadIdx.x;
#pragma unroll 1
    for (int c0 = 0; c0 < N; c0 += 8) {
        double v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) v[u] = __ldcg(src + tid + (size_t)(c0 + u) * N);
        const double rsv = rs ? rs[tid] : 1.0;
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            double x = v[u] * rsv;
            if (cs) x *= cs[c0 + u];
            T[(c0 + u) + (size_t)tid * LDT] = x;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Kronecker chain in place on the tile: for s = 0 .. m-1:  X <- diag(sc[s]) (Ey (x) Ex) X.
// Both passes act on every column on its own (V <- Ex V Ey^T for V(x, y) = X[x + 8 y, c]); warp w owns columns
// [32 w, 32 w + 32) for the whole chain: no block barrier, __syncwarp only.  Two 8-column groups are in flight per
// warp for instruction-level parallelism.  Same fragment / store pattern as f_bbar (fast64.cuh).
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void chain(double* __restrict__ X, const double* __restrict__ sc, int m, const double* __restrict__ kron) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, gr = lane >> 2, tg = lane & 3;
    const double ex0 = kron[gr + 8 * tg], ex1 = kron[gr + 8 * (tg + 4)];            // A fragment: Ex(gr, k)
    const double ey0 = kron[64 + gr + 8 * tg], ey1 = kron[64 + gr + 8 * (tg + 4)];  // B fragment: Ey^T(k, gr) = Ey(gr, k)
    const int r0 = gr + 16 * tg;
#pragma unroll 1
    for (int s = 0; s < m; ++s) {
        const double e0 = sc[s * N + r0], e1 = sc[s * N + r0 + 8];
#pragma unroll 1
        for (int gp = 0; gp < 2; ++gp) {
            double b0[2][8], b1[2][8], c0[2][8], c1[2][8];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int cb = 4 * warp + 2 * gp + h;  // 8-column block
#pragma unroll
                for (int y = 0; y < 8; ++y) {
                    const double* bp = X + (8 * y + tg) + (size_t)(8 * cb + gr) * LDT;
                    b0[h][y] = bp[0];
                    b1[h][y] = bp[4];
                }
            }
#pragma unroll
            for (int h = 0; h < 2; ++h)
#pragma unroll
                for (int y = 0; y < 8; ++y) {
                    c0[h][y] = c1[h][y] = 0.0;
                    dmma884(c0[h][y], c1[h][y], ex0, b0[h][y]);
                }
#pragma unroll
            for (int h = 0; h < 2; ++h)
#pragma unroll
                for (int y = 0; y < 8; ++y) dmma884(c0[h][y], c1[h][y], ex1, b1[h][y]);
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int cb = 4 * warp + 2 * gp + h;
#pragma unroll
                for (int y = 0; y < 8; ++y) {
                    double* cp = X + (8 * y + gr) + (size_t)(8 * cb + 2 * tg) * LDT;
                    cp[0] = c0[h][y];
                    cp[LDT] = c1[h][y];
                }
            }
            __syncwarp();
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int cb = 4 * warp + 2 * gp + h;
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const double* colp = X + (size_t)(8 * cb + u) * LDT;
                    b0[h][u] = colp[gr + 8 * tg];
                    b1[h][u] = colp[gr + 8 * (tg + 4)];
                }
            }
#pragma unroll
            for (int h = 0; h < 2; ++h)
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    c0[h][u] = c1[h][u] = 0.0;
                    dmma884(c0[h][u], c1[h][u], b0[h][u], ey0);
                }
#pragma unroll
            for (int h = 0; h < 2; ++h)
#pragma unroll
                for (int u = 0; u < 8; ++u) dmma884(c0[h][u], c1[h][u], b1[h][u], ey1);
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int cb = 4 * warp + 2 * gp + h;
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    double* colp = X + (size_t)(8 * cb + u) * LDT;
                    colp[r0] = e0 * c0[h][u];
                    colp[r0 + 8] = e1 * c1[h][u];
                }
            }
            __syncwarp();
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Column-pivoted Householder QR of the matrix in the tile, thread c = column c, all 64 rows in registers.
//   TR = false: M(r, c) = T[r + LDT c];  TR = true: the tile holds M^T, M(r, c) = T[c + LDT r].
// Same conventions as qrp_tile (tile.cuh): implicit pivoting (jpvt[j] = column chosen at step j), exact partial norms
// recomputed every step, a finished row is retired into the tile and zeroed in registers, always-reflect rule,
// branch-free Newton sqrt / reciprocal.  On exit: M(j, c) = R(j, c) for j <= step(c), the scaled reflector of step
// step(c) below; tau[]; jpvt[]; returns step(c) of the calling thread's column.
// arena: sn[64] squared partial norms (-1 = chosen), xr[2][64] published pivot column (double buffered).
// ---------------------------------------------------------------------------------------------------------------
template <bool TR, int JB>
__device__ __forceinline__ void qr_block(const Wk& wk, double (&a)[N], bool& done, int& mystep, double& myscal);

template <bool TR>
__device__ __forceinline__ int qr_cols(Wk& wk, double (&a)[N]) {
    double* T = wk.T;
    double* sn = wk.arena;
    const int tid = threadIdx.x;
    {
        double s0 = 0.0, s1 = 0.0;
#pragma unroll
        for (int r = 0; r < N; r += 2) {
            if (TR) {
                a[r] = T[tid + (size_t)r * LDT];
                a[r + 1] = T[tid + (size_t)(r + 1) * LDT];
            } else {
                const double2 v = *reinterpret_cast<const double2*>(T + r + (size_t)tid * LDT);
                a[r] = v.x;
                a[r + 1] = v.y;
            }
            s0 = fma(a[r], a[r], s0);
            s1 = fma(a[r + 1], a[r + 1], s1);
        }
        sn[tid] = s0 + s1;
    }
    bool done = false;
    int mystep = N;
    double myscal = 0.0;
    __syncthreads();
    // the 8-row register blocks are instantiated one by one (template parameter): a `#pragma unroll` over the block loop
    // is not always honoured around the rolled inner loop, and a single run-time index sends the whole array to local memory
    qr_block<TR, 0>(wk, a, done, mystep, myscal);
    qr_block<TR, 1>(wk, a, done, mystep, myscal);
    qr_block<TR, 2>(wk, a, done, mystep, myscal);
    qr_block<TR, 3>(wk, a, done, mystep, myscal);
    qr_block<TR, 4>(wk, a, done, mystep, myscal);
    qr_block<TR, 5>(wk, a, done, mystep, myscal);
    qr_block<TR, 6>(wk, a, done, mystep, myscal);
    qr_block<TR, 7>(wk, a, done, mystep, myscal);
    // the reflector of every column (rows below its diagonal entry), scaled, from the registers its thread still holds:
    // all threads store in parallel here instead of one lane per step inside the loop
#pragma unroll
    for (int r = 0; r < N; ++r)
        if (r > mystep) T[TR ? (tid + (size_t)r * LDT) : (r + (size_t)tid * LDT)] = a[r] * myscal;
    __syncthreads();
    return mystep;
}

template <bool TR, int JB>
__device__ __forceinline__ void qr_block(const Wk& wk, double (&a)[N], bool& done, int& mystep, double& myscal) {
    double* T = wk.T;
    double* sn = wk.arena;
    double* xrb = wk.arena + N;
    const int tid = threadIdx.x, lane = tid & 31;
    {
        constexpr int rb = 8 * JB;  // rows below rb are retired (exact zeros in registers)
#pragma unroll 1
        for (int jj = 0; jj < 8; ++jj) {
            const int j = rb + jj;
            double* xr = xrb + (j & 1) * N;
            const double n0 = sn[lane], n1 = sn[lane + 32];
            const int bi = warp_argmax2_64(n0, lane, n0 >= 0.0, n1, lane + 32, n1 >= 0.0);
            const double n2 = sn[bi];
            const bool mine = (tid == bi);
            if (mine) {
#pragma unroll
                for (int r = rb; r < N; r += 2) *reinterpret_cast<double2*>(xr + r) = make_double2(a[r], a[r + 1]);
                wk.jpvt[j] = bi;
                done = true;
                mystep = j;
            }
            __syncthreads();
            if (mine) sn[bi] = -1.0;  // after the barrier: every warp has elected this step's pivot (see qrp_tile)
            if (!done || mine) {
                const double alpha = xr[j];
                const double nn = fmax(n2, alpha * alpha);
                const bool refl = nn > 0.0;
                double hrs;
                const double nrm = sqrt_newton(refl ? nn : 1.0, &hrs);
                const double beta = refl ? -copysign(nrm, alpha) : alpha;
                const double tt = refl ? fma(fabs(alpha), hrs + hrs, 1.0) : 0.0;
                const double scal = refl ? copysign(rcp_newton(fabs(alpha) + nrm), alpha) : 0.0;
                if (!mine) {
                    // x(j+1:)^T a(j+1:, c): rows rb .. j-1 are zero on both sides, row j is masked out
                    double d0 = 0.0, d1 = 0.0, d2 = 0.0, d3 = 0.0;
#pragma unroll
                    for (int r = rb; r < rb + 8; r += 4) {
                        const double2 xa = *reinterpret_cast<const double2*>(xr + r), xb = *reinterpret_cast<const double2*>(xr + r + 2);
                        d0 = fma((r == j) ? 0.0 : xa.x, a[r], d0);
                        d1 = fma((r + 1 == j) ? 0.0 : xa.y, a[r + 1], d1);
                        d2 = fma((r + 2 == j) ? 0.0 : xb.x, a[r + 2], d2);
                        d3 = fma((r + 3 == j) ? 0.0 : xb.y, a[r + 3], d3);
                    }
#pragma unroll
                    for (int r = rb + 8; r < N; r += 4) {
                        const double2 xa = *reinterpret_cast<const double2*>(xr + r), xb = *reinterpret_cast<const double2*>(xr + r + 2);
                        d0 = fma(xa.x, a[r], d0);
                        d1 = fma(xa.y, a[r + 1], d1);
                        d2 = fma(xb.x, a[r + 2], d2);
                        d3 = fma(xb.y, a[r + 3], d3);
                    }
                    const double dot = (d0 + d1) + (d2 + d3);
                    const double aj = sel8(a, rb, jj);
                    const double w = tt * fma(scal, dot, aj);  // tau (v^T a_c)
                    const double ws = w * scal;
                    T[TR ? (tid + (size_t)j * LDT) : (j + (size_t)tid * LDT)] = aj - w;  // R(j, c) is final: retire the row
                    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
#pragma unroll
                    for (int r = rb; r < rb + 8; r += 4) {
                        const double2 xa = *reinterpret_cast<const double2*>(xr + r), xb = *reinterpret_cast<const double2*>(xr + r + 2);
                        a[r] = (r == j) ? 0.0 : fma(-ws, xa.x, a[r]);
                        a[r + 1] = (r + 1 == j) ? 0.0 : fma(-ws, xa.y, a[r + 1]);
                        a[r + 2] = (r + 2 == j) ? 0.0 : fma(-ws, xb.x, a[r + 2]);
                        a[r + 3] = (r + 3 == j) ? 0.0 : fma(-ws, xb.y, a[r + 3]);
                        s0 = fma(a[r], a[r], s0);
                        s1 = fma(a[r + 1], a[r + 1], s1);
                        s2 = fma(a[r + 2], a[r + 2], s2);
                        s3 = fma(a[r + 3], a[r + 3], s3);
                    }
#pragma unroll
                    for (int r = rb + 8; r < N; r += 4) {
                        const double2 xa = *reinterpret_cast<const double2*>(xr + r), xb = *reinterpret_cast<const double2*>(xr + r + 2);
                        a[r] = fma(-ws, xa.x, a[r]);
                        a[r + 1] = fma(-ws, xa.y, a[r + 1]);
                        a[r + 2] = fma(-ws, xb.x, a[r + 2]);
                        a[r + 3] = fma(-ws, xb.y, a[r + 3]);
                        s0 = fma(a[r], a[r], s0);
                        s1 = fma(a[r + 1], a[r + 1], s1);
                        s2 = fma(a[r + 2], a[r + 2], s2);
                        s3 = fma(a[r + 3], a[r + 3], s3);
                    }
                    sn[tid] = (s0 + s1) + (s2 + s3);  // exact |a(j+1:, c)|^2
                } else {
                    // my column is the pivot column: R(0..j-1, c) is in the tile already; beta on the diagonal; the
                    // reflector v = x / (alpha - beta) below it is stored after the last step (a[] is frozen from here on)
                    T[TR ? (tid + (size_t)j * LDT) : (j + (size_t)tid * LDT)] = beta;
                    wk.tau[j] = tt;
                    myscal = scal;
                }
            }
            __syncthreads();
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Column tid of Q = H_0 ... H_63 (dorgqr's result): q = H_0 .. H_tid e_tid; reflector i lives in column jpvt[i] of M,
// rows > i (unit diagonal implicit), read with broadcast loads.  The rows above 8 * (i / 8) of q are still zero when
// H_i is applied, so the loops start there.
// ---------------------------------------------------------------------------------------------------------------
template <bool TR, int IB>
__device__ __forceinline__ void form_q_block(const Wk& wk, double (&q)[N]) {
    const double* T = wk.T;
    const int tid = threadIdx.x;
    {
        constexpr int rb = 8 * IB;
#pragma unroll 1
        for (int ii = 7; ii >= 0; --ii) {
            const int i = rb + ii;
            if (i <= tid) {
                const int col = wk.jpvt[i];
                const double t = wk.tau[i];
                const double* vp = TR ? (T + col) : (T + (size_t)col * LDT);
                constexpr int VS = TR ? LDT : 1;
                double d0 = 0.0, d1 = 0.0;
#pragma unroll
                for (int r = rb; r < rb + 8; ++r) {
                    const double raw = vp[(size_t)r * VS];
                    const double v = (r > i) ? raw : ((r == i) ? 1.0 : 0.0);
                    if (r & 1) d1 = fma(v, q[r], d1); else d0 = fma(v, q[r], d0);
                }
#pragma unroll
                for (int r = rb + 8; r < N; ++r) {
                    const double v = vp[(size_t)r * VS];
                    if (r & 1) d1 = fma(v, q[r], d1); else d0 = fma(v, q[r], d0);
                }
                const double w = t * (d0 + d1);
#pragma unroll
                for (int r = rb; r < rb + 8; ++r) {
                    const double raw = vp[(size_t)r * VS];
                    const double v = (r > i) ? raw : ((r == i) ? 1.0 : 0.0);
                    q[r] = fma(-w, v, q[r]);
                }
#pragma unroll
                for (int r = rb + 8; r < N; ++r) q[r] = fma(-w, vp[(size_t)r * VS], q[r]);
            }
        }
    }
}
template <bool TR>
__device__ __forceinline__ void form_q_col(const Wk& wk, double (&q)[N]) {
    const int tid = threadIdx.x;
#pragma unroll
    for (int r = 0; r < N; ++r) q[r] = (r == tid) ? 1.0 : 0.0;
    form_q_block<TR, 7>(wk, q);
    form_q_block<TR, 6>(wk, q);
    form_q_block<TR, 5>(wk, q);
    form_q_block<TR, 4>(wk, q);
    form_q_block<TR, 3>(wk, q);
    form_q_block<TR, 2>(wk, q);
    form_q_block<TR, 1>(wk, q);
    form_q_block<TR, 0>(wk, q);
}

// ---------------------------------------------------------------------------------------------------------------
// LU with partial pivoting, thread r = row r in registers (dgetf2 order; rows never move: `pos` = step at which the
// row became the pivot row = its position in dgetrf's layout).  Per column: every warp elects its best active row with
// two REDUX and publishes that row (entries right of the column) and its pivot value speculatively; after ONE barrier
// every thread takes the better of the two candidates.  Buffers double buffered by column parity.
// arena: prow[2 parity][2 warp][64] doubles, cand[2 parity][2 warp][2] doubles (|pivot| or -1, pivot), cidx ints.
// On exit: factors in the tile in dgetrf's layout (row pos), perm[pos] = original row, rdiag[j] = 1 / U(j, j).
// ---------------------------------------------------------------------------------------------------------------
template <int JB>
__device__ __forceinline__ void lu_block(const Wk& wk, double (&a)[N], bool& act, int& pos, double& v, int& bl, bool& any) {
    double* prow = wk.arena;               // [2][2][64]
    double* cand = wk.arena + 4 * N;       // [2][2][2]
    int* cidx = reinterpret_cast<int*>(wk.arena + 4 * N + 8);  // [2][2]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int cb = 8 * JB;
#pragma unroll 1
    for (int jj = 0; jj < 8; ++jj) {
        const int j = cb + jj;
        const int par = j & 1;
        // (v, bl, any): this thread's entry of column j and the row its warp elected, found during step j - 1
        if (any && tid == bl) {
            double* pr = prow + (size_t)(par * 2 + warp) * N;
#pragma unroll
            for (int k = cb; k < cb + 8; k += 2)
                *reinterpret_cast<double2*>(pr + k) = make_double2((k > j) ? a[k] : 0.0, (k + 1 > j) ? a[k + 1] : 0.0);
#pragma unroll
            for (int k = cb + 8; k < N; k += 2) *reinterpret_cast<double2*>(pr + k) = make_double2(a[k], a[k + 1]);
            cand[(par * 2 + warp) * 2] = fabs(v);
            cand[(par * 2 + warp) * 2 + 1] = v;
            cidx[par * 2 + warp] = tid;
        }
        if (!any && lane == 0) {
            cand[(par * 2 + warp) * 2] = -1.0;
            cand[(par * 2 + warp) * 2 + 1] = 0.0;
            cidx[par * 2 + warp] = -1;
        }
        __syncthreads();
        const double m0 = cand[(par * 2 + 0) * 2], m1 = cand[(par * 2 + 1) * 2];
        const int ws = (m1 > m0) ? 1 : 0;  // ties: the lower row index (warp 0), as idamax
        const double piv = cand[(par * 2 + ws) * 2 + 1];
        const int p = cidx[par * 2 + ws];
        const double rinv = (piv != 0.0) ? rcp_newton(piv) : 0.0;
        const double* pr = prow + (size_t)(par * 2 + ws) * N;
        const bool upd = act && tid != p;
        const double l = upd ? v * rinv : 0.0;
        if (tid == p) {
            act = false;
            pos = j;
            wk.rdiag[j] = rinv;
            wk.perm[j] = p;
            if (piv == 0.0 && *wk.info == 0) *wk.info = j + 1;
        }
        if (upd) put8(a, cb, jj, l);
        // the block's own columns and the first column right of it, then the election of the next pivot candidate
        // (two REDUX) overlaps the update of the remaining columns.  l = 0 for rows that are not updated.
#pragma unroll
        for (int k = cb; k < ((cb + 10 < N) ? cb + 10 : N); k += 2) {
            const double2 u = *reinterpret_cast<const double2*>(pr + k);
            a[k] = fma(-l, u.x, a[k]);
            a[k + 1] = fma(-l, u.y, a[k + 1]);
        }
        if (j + 1 < N) {
            v = (jj < 7) ? sel8(a, cb, (jj + 1) & 7) : a[(cb + 8 < N) ? cb + 8 : 0];
            any = __any_sync(DQ_FULL, act);
            bl = warp_argmax2_64(fabs(v), tid, act, 0.0, 0, false);  // row index 0..63 (valid iff any)
        }
#pragma unroll
        for (int k = cb + 10; k < N; k += 2) {
            const double2 u = *reinterpret_cast<const double2*>(pr + k);
            a[k] = fma(-l, u.x, a[k]);
            a[k + 1] = fma(-l, u.y, a[k + 1]);
        }
    }
}
__device__ __forceinline__ void lu_rows(Wk& wk, double (&a)[N]) {
    const int tid = threadIdx.x;
    bool act = true;
    int pos = 0;
    if (tid == 0) *wk.info = 0;
    double v = a[0];
    bool any = true;
    int bl = warp_argmax2_64(fabs(v), tid, true, 0.0, 0, false);
    lu_block<0>(wk, a, act, pos, v, bl, any);
    lu_block<1>(wk, a, act, pos, v, bl, any);
    lu_block<2>(wk, a, act, pos, v, bl, any);
    lu_block<3>(wk, a, act, pos, v, bl, any);
    lu_block<4>(wk, a, act, pos, v, bl, any);
    lu_block<5>(wk, a, act, pos, v, bl, any);
    lu_block<6>(wk, a, act, pos, v, bl, any);
    lu_block<7>(wk, a, act, pos, v, bl, any);
    __syncthreads();  // nobody reads the tile's previous contents any more
#pragma unroll
    for (int k = 0; k < N; ++k) wk.T[pos + (size_t)k * LDT] = a[k];
    __syncthreads();
}

// dgetrs 'N' for the right-hand side column held by this thread (x[i] = (P b)[i] on entry), factors from lu_rows.
template <int KB>
__device__ __forceinline__ void solve_fwd_block(const Wk& wk, double (&x)[N]) {
    const double* T = wk.T;
    constexpr int cb = 8 * KB;
    double xk = x[cb];
#pragma unroll 1
    for (int kk = 0; kk < 8; ++kk) {
        const int k = cb + kk;
        const double* lp = T + (size_t)k * LDT;
#pragma unroll
        for (int i = cb; i < cb + 8; i += 2) {
            const double2 l = *reinterpret_cast<const double2*>(lp + i);
            x[i] = fma(-((i > k) ? l.x : 0.0), xk, x[i]);
            x[i + 1] = fma(-((i + 1 > k) ? l.y : 0.0), xk, x[i + 1]);
        }
        // x[k + 1] is final now: select it while the rows below the block are updated
        const double xn = sel8(x, cb, (kk + 1) & 7);
#pragma unroll
        for (int i = cb + 8; i < N; i += 2) {
            const double2 l = *reinterpret_cast<const double2*>(lp + i);
            x[i] = fma(-l.x, xk, x[i]);
            x[i + 1] = fma(-l.y, xk, x[i + 1]);
        }
        xk = xn;
    }
}
template <int KB>
__device__ __forceinline__ void solve_bwd_block(const Wk& wk, double (&x)[N]) {
    const double* T = wk.T;
    constexpr int cb = 8 * KB;
    double xk = x[cb + 7] * wk.rdiag[cb + 7];
#pragma unroll 1
    for (int kk = 7; kk >= 0; --kk) {
        const int k = cb + kk;
        put8(x, cb, kk, xk);
        const double* up = T + (size_t)k * LDT;
#pragma unroll
        for (int i = cb; i < cb + 8; i += 2) {
            const double2 u = *reinterpret_cast<const double2*>(up + i);
            x[i] = fma(-((i < k) ? u.x : 0.0), xk, x[i]);
            x[i + 1] = fma(-((i + 1 < k) ? u.y : 0.0), xk, x[i + 1]);
        }
        const int kn = (kk + 7) & 7;  // kk - 1 (the value is unused after kk = 0)
        const double xn = sel8(x, cb, kn) * wk.rdiag[cb + kn];
#pragma unroll
        for (int i = 0; i < cb; i += 2) {
            const double2 u = *reinterpret_cast<const double2*>(up + i);
            x[i] = fma(-u.x, xk, x[i]);
            x[i + 1] = fma(-u.y, xk, x[i + 1]);
        }
        xk = xn;
    }
}
__device__ __forceinline__ void solve_col(const Wk& wk, double (&x)[N]) {
    // forward: L y = P b (unit diagonal)
    solve_fwd_block<0>(wk, x);
    solve_fwd_block<1>(wk, x);
    solve_fwd_block<2>(wk, x);
    solve_fwd_block<3>(wk, x);
    solve_fwd_block<4>(wk, x);
    solve_fwd_block<5>(wk, x);
    solve_fwd_block<6>(wk, x);
    solve_fwd_block<7>(wk, x);
    // backward: U x = y
    solve_bwd_block<7>(wk, x);
    solve_bwd_block<6>(wk, x);
    solve_bwd_block<5>(wk, x);
    solve_bwd_block<4>(wk, x);
    solve_bwd_block<3>(wk, x);
    solve_bwd_block<2>(wk, x);
    solve_bwd_block<1>(wk, x);
    solve_bwd_block<0>(wk, x);
}

// ---------------------------------------------------------------------------------------------------------------
// C = op(A) B on the FP64 tensor cores, A in the tile (AT: A(r, k) = T[k + LDT r], else T[r + LDT k]), B dense in
// global memory (ld = N, L2 resident) streamed in 8-column chunks by cp.async into a double buffer, optional scaling
// of B's columns by bcs[].  Warp w produces rows [32 w, 32 w + 32) of every chunk: acc[mb][j] = C(32 w + 8 mb + gr,
// 8 ch + 2 tg + j), handed to `sink(ch, acc)`.  arena: chunk buffers 0, 1 (conversion buffer 2 is the sink's).
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void issue_chunk(double* __restrict__ buf, const double* __restrict__ B, int ch) {
    const int tid = threadIdx.x;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int idx = tid + NTH * q;  // 256 pieces of 16 bytes
        const int cc = idx >> 5, k2 = (idx & 31) * 2;
        cp_async16(buf + k2 + (size_t)cc * LDT, B + k2 + (size_t)(CW * ch + cc) * N);
    }
    cp_async_commit();
}

// Sinks (no functors: a lambda capturing the 64-double register array by reference takes its address and the compiler
// then keeps the array in local memory for the whole kernel):
//   SINK_ROWS   m[] (thread r <- C(r, 8 ch .. 8 ch + 7)) += chunk, through the conversion buffer (arena chunk 2)
//   SINK_STORE  dst(r, c) = C(r, c)                       dense global, ld = N
//   SINK_G      as SINK_STORE, and emax = max(emax, |dst_old - C|)   (check_error, source/dqmc.cpp:317-329)
enum { SINK_ROWS = 0, SINK_STORE = 1, SINK_G = 2 };

template <bool AT, int SINK>
__device__ __forceinline__ void gemm_stream(const Wk& wk, const double* __restrict__ B, const double* __restrict__ bcs, double (&m)[N],
                                            double* __restrict__ dst, double& emax) {
    const double* T = wk.T;
    double* bufs = wk.arena;
    double* cv = wk.arena + 2 * CW * LDT;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, gr = lane >> 2, tg = lane & 3;
    issue_chunk(bufs, B, 0);
    issue_chunk(bufs + CW * LDT, B, 1);
    // SINK_ROWS needs the chunk index at compile time (register index of m[]); the storing sinks keep the loop rolled
#pragma unroll(SINK == SINK_ROWS ? N / CW : 1)
    for (int ch = 0; ch < N / CW; ++ch) {
        if (ch + 1 < N / CW) cp_async_wait<1>(); else cp_async_wait<0>();
        __syncthreads();  // chunk ch has landed for every thread; the conversion buffer of chunk ch - 1 is consumed
        const double* bc = bufs + (size_t)(ch & 1) * CW * LDT;
        double acc[4][2];
#pragma unroll
        for (int mb = 0; mb < 4; ++mb) acc[mb][0] = acc[mb][1] = 0.0;
        const double bs = bcs ? bcs[CW * ch + gr] : 1.0;
#pragma unroll
        for (int k0 = 0; k0 < N; k0 += 4) {
            double b = bc[(k0 + tg) + (size_t)gr * LDT];
            if (bcs) b *= bs;
#pragma unroll
            for (int mb = 0; mb < 4; ++mb) {
                const int r = 32 * warp + 8 * mb + gr;
                const double av = AT ? T[(k0 + tg) + (size_t)r * LDT] : T[r + (size_t)(k0 + tg) * LDT];
                dmma884(acc[mb][0], acc[mb][1], av, b);
            }
        }
        if (SINK == SINK_ROWS) {
#pragma unroll
            for (int mb = 0; mb < 4; ++mb) {
                const int r = 32 * warp + 8 * mb + gr;
                cv[r + (size_t)(2 * tg) * LDT] = acc[mb][0];
                cv[r + (size_t)(2 * tg + 1) * LDT] = acc[mb][1];
            }
            __syncthreads();
#pragma unroll
            for (int cc = 0; cc < CW; ++cc) m[CW * ch + cc] += cv[threadIdx.x + (size_t)cc * LDT];
        } else {
#pragma unroll
            for (int mb = 0; mb < 4; ++mb) {
                const int r = 32 * warp + 8 * mb + gr;
                double* g0 = dst + r + (size_t)(CW * ch + 2 * tg) * N;
                if (SINK == SINK_G) emax = fmax(emax, fmax(fabs(g0[0] - acc[mb][0]), fabs(g0[N] - acc[mb][1])));
                g0[0] = acc[mb][0];
                g0[N] = acc[mb][1];
            }
        }
        __syncthreads();  // everybody is done with chunk buffer ch & 1
        if (ch + 2 < N / CW) issue_chunk(bufs + (size_t)(ch & 1) * CW * LDT, B, ch + 2);
    }
}

__device__ __forceinline__ double block_sum(const Wk& wk, double v) {
    v = warp_sum(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) wk.red[threadIdx.x >> 5] = v;
    __syncthreads();
    return wk.red[0] + wk.red[1];
}
__device__ __forceinline__ double block_max(const Wk& wk, double v) {
    v = warp_max(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) wk.red[threadIdx.x >> 5] = v;
    __syncthreads();
    return fmax(wk.red[0], wk.red[1]);
}

__device__ __forceinline__ void note_fail(const Params& p, const Wk& wk, int code) {
    if (threadIdx.x == 0 && *wk.info != 0) atomicCAS(p.fail, 0, code);
}

// ---------------------------------------------------------------------------------------------------------------
// The kernel.  DIR = +1: sweep_0_to_beta block end (source/dqmc.cpp:369-394); DIR = -1: sweep_beta_to_0 block start
// (source/dqmc.cpp:429-454).
// ---------------------------------------------------------------------------------------------------------------
template <int DIR>
__global__ void __launch_bounds__(NTH, 4) kf2_stabilize(Params p, int l) {
    extern __shared__ __align__(16) unsigned char smem[];
    constexpr bool TR = DIR < 0;  // backward: the tile holds the TRANSPOSE of the matrix being factorised
    const int tid = threadIdx.x;
    const int w = p.w0 + blockIdx.x;
    Wk wk;
    carve(wk, smem);
    const int i = l / p.n_stab;
    const int l0 = i * p.n_stab, l1 = min(l0 + p.n_stab, p.nt), m = l1 - l0;
    const bool first = (DIR > 0) ? (i == 0) : (i == p.n_stack - 1);
    const bool at_end = (DIR > 0) ? (l == p.nt - 1) : (l == 0);
    const int iprev = (DIR > 0) ? i - 1 : i + 1;  // the entry the new one is built from
    const int inext = (DIR > 0) ? i + 1 : i - 1;  // the entry it is combined with for G
    double* wsX = p.ws + (size_t)w * 3 * N * p.LD;  // per-walker global scratch, 3 slots of >= N * N doubles
    double* wsY = wsX + (size_t)N * p.LD;
    DQ_PH(p, 0);
    if (tid == 0) {  // pull the stack entries this launch needs later into L2 (TMA bulk prefetch)
        constexpr uint32_t MB_ = (uint32_t)(N * N * sizeof(double));
        if (!first) bulk_prefetch_l2((DIR > 0) ? fstkR(p, w, iprev) : fstkL(p, w, iprev), MB_);
        if (!at_end) {
            bulk_prefetch_l2(fstkL(p, w, inext), MB_);
            bulk_prefetch_l2(fstkR(p, w, inext), MB_);
        }
        bulk_prefetch_l2(p.G + (size_t)w * N * N, MB_);
    }
    // ---- slice scalings of the chain: forward  X <- diag(ev_l) E X, l ascending;
    //      backward (transposed)  X^T <- E diag(ev_l) X^T, l descending = initial scaling by ev_{l1-1}, then
    //      [E, diag(ev_{l1-2})], ..., [E, diag(ev_{l0})], [E, 1]
    {
        const int8_t* f = p.fields + ((size_t)w * p.nt) * N;
        double* sc = wk.arena;
        for (int s = 0; s < m; ++s) {
            double v;
            if (DIR > 0) v = p.lut_ev[f[(size_t)(l0 + s) * N + tid]];
            else v = (s < m - 1) ? p.lut_ev[f[(size_t)(l1 - 2 - s) * N + tid]] : 1.0;
            sc[s * N + tid] = v;
        }
        if (DIR < 0) wk.sv[1][tid] = p.lut_ev[f[(size_t)(l1 - 1) * N + tid]];
        if (!first) wk.sv[0][tid] = fstkD(p, w, iprev)[tid];
    }
    __syncthreads();
    // ---- X0 into the tile
    if (first) {
        for (int e = tid; e < N * N; e += NTH) {
            const int r = e & 63, c = e >> 6;
            wk.T[r + (size_t)c * LDT] = (r == c) ? ((DIR > 0) ? 1.0 : wk.sv[1][r]) : 0.0;
        }
    } else if (DIR > 0) {
        load_tile(wk.T, fstkL(p, w, iprev), nullptr, wk.sv[0]);  // L_prev D_prev
    } else {
        load_tile_t(wk.T, fstkR(p, w, iprev), wk.sv[0], wk.sv[1]);  // (D_next R_next)^T, rows scaled by ev_{l1-1}
    }
    __syncthreads();
    DQ_PH(p, 1);
    chain(wk.T, wk.arena, m, p.kron);
    __syncthreads();
    DQ_PH(p, 2);
    // ---- to_LDR (source/stablelinalg.cpp:35-55)
    double v[N];
    const int mystep = qr_cols<TR>(wk, v);
    DQ_PH(p, 3);
    {
        const int c = wk.jpvt[tid];
        const double dj = fabs(wk.T[TR ? (c + (size_t)tid * LDT) : (tid + (size_t)c * LDT)]);
        wk.dnew[tid] = dj;
        wk.dinv[tid] = 1.0 / dj;
        fstkD(p, w, i)[tid] = dj;
    }
    __syncthreads();
    form_q_col<TR>(wk, v);
    // forward: L_i = Q.  backward: L_i = L_next Q (Q parked in global scratch as the streamed operand), or Q itself.
    {
        double* qdst = (DIR > 0 || first) ? fstkL(p, w, i) : wsX;
#pragma unroll
        for (int r = 0; r < N; r += 2) *reinterpret_cast<double2*>(qdst + r + (size_t)tid * N) = make_double2(v[r], v[r + 1]);
    }
    __syncthreads();  // every reflector has been read
    DQ_PH(p, 4);
    // R_final = (D^-1 R) P^T: column c keeps rows <= step(c)
    {
#pragma unroll
        for (int r = 0; r < N; ++r) {
            const double raw = wk.T[TR ? (tid + (size_t)r * LDT) : (r + (size_t)tid * LDT)];
            v[r] = (r <= mystep) ? raw * wk.dinv[r] : 0.0;
        }
        if (DIR < 0 || first) {  // R_i = R_final
            double* rdst = fstkR(p, w, i);
#pragma unroll
            for (int r = 0; r < N; r += 2) *reinterpret_cast<double2*>(rdst + r + (size_t)tid * N) = make_double2(v[r], v[r + 1]);
        } else {
#pragma unroll
            for (int r = 0; r < N; r += 2) *reinterpret_cast<double2*>(wk.T + r + (size_t)tid * LDT) = make_double2(v[r], v[r + 1]);
        }
    }
    __threadfence();
    __syncthreads();
    double emax = 0.0;
    if (!first) {
        if (DIR > 0) {
            // R_i = R_final R_prev
            gemm_stream<false, SINK_STORE>(wk, fstkR(p, w, iprev), nullptr, v, fstkR(p, w, i), emax);
        } else {
            // L_i = L_next Q
            load_tile(wk.T, fstkL(p, w, iprev), nullptr, nullptr);
            gemm_stream<false, SINK_STORE>(wk, wsX, nullptr, v, fstkL(p, w, i), emax);
        }
        __threadfence();
        __syncthreads();
    }
    DQ_PH(p, 5);

    // ================= stabilised Green's function =================
    // at_end:  G = [I + F]^-1, F = stack[i]                      (inv_I_plus_ldr, source/stablelinalg.cpp:94-126)
    // else:    G = [I + F1 F2]^-1, (F1, F2) = (stack[i], stack[i+1]) forward, (stack[i-1], stack[i]) backward
    //                                                            (inv_I_plus_ldr_mul_ldr, source/stablelinalg.cpp:128-158)
    const int i1 = at_end ? i : ((DIR > 0) ? i : i - 1);
    const int i2 = at_end ? i : ((DIR > 0) ? i + 1 : i);
    const double *L1 = fstkL(p, w, i1), *R1 = fstkR(p, w, i1), *L2 = fstkL(p, w, i2), *R2 = fstkR(p, w, i2);
    {
        const double d1 = __ldcg(fstkD(p, w, i1) + tid), d2 = __ldcg(fstkD(p, w, i2) + tid);
        wk.sv[0][tid] = (d1 >= 1.0) ? 1.0 / d1 : 1.0;  // D1_large^-1
        wk.sv[1][tid] = (d1 >= 1.0) ? 1.0 : d1;        // D1_small
        wk.sv[2][tid] = (d2 >= 1.0) ? 1.0 / d2 : 1.0;
        wk.sv[3][tid] = (d2 >= 1.0) ? 1.0 : d2;
    }
    double logdet = 0.0;
    if (at_end) logdet = block_sum(wk, (wk.dnew[tid] >= 1.0) ? log(wk.dnew[tid]) : 0.0);  // contains barriers
    else __syncthreads();
    double* Gg = p.G + (size_t)w * N * N;
#pragma unroll 1
    for (int rep = 0; rep < 2; ++rep) {
        // ---- the matrix to factorise, row tid in registers
        if (rep == 0) {
            const double* Rs = at_end ? R1 : R2;
#pragma unroll
            for (int k = 0; k < N; ++k) v[k] = __ldcg(Rs + tid + (size_t)k * N);
        } else if (at_end) {
            // M = X + L D_small; row tid of M^T = column tid of M: v holds X(:, tid) from the first solve
            const double ds = wk.sv[1][tid];
#pragma unroll
            for (int k = 0; k < N; k += 2) {
                const double2 lv = __ldcg(reinterpret_cast<const double2*>(L1 + k + (size_t)tid * N));
                v[k] = v[k] + lv.x * ds;
                v[k + 1] = v[k + 1] + lv.y * ds;
            }
        } else {
            // M = D1l^-1 (L1^T X) + D1s (R1 (L2 D2s)), assembled row by row from two streamed products
#pragma unroll
            for (int k = 0; k < N; ++k) v[k] = 0.0;
            load_tile(wk.T, L1, nullptr, wk.sv[0]);  // T = L1 D1l^-1, used transposed
            gemm_stream<true, SINK_ROWS>(wk, wsX, nullptr, v, nullptr, emax);
            load_tile(wk.T, R1, wk.sv[1], nullptr);  // T = D1s R1
            gemm_stream<false, SINK_ROWS>(wk, L2, wk.sv[3], v, nullptr, emax);
        }
        DQ_PH(p, 6 + 4 * rep);
        lu_rows(wk, v);
        note_fail(p, wk, 12 + rep);
        DQ_PH(p, 7 + 4 * rep);
        if (rep == 1 && at_end) {
            const double lm = block_sum(wk, log(fabs(wk.T[tid + (size_t)tid * LDT])));
            logdet += lm;
        }
        // ---- right-hand side column tid, rows interchanged while loading
        if (rep == 0) {
            const double dl = at_end ? wk.sv[0][tid] : wk.sv[2][tid];  // D_large^-1 of the factor whose R is being inverted
#pragma unroll
            for (int k = 0; k < N; ++k) v[k] = (wk.perm[k] == tid) ? dl : 0.0;
        } else if (at_end) {
            // M^T G^T = X^T: column tid of X^T = row tid of X
#pragma unroll
            for (int k = 0; k < N; ++k) v[k] = __ldcg(wsX + tid + (size_t)wk.perm[k] * N);
        } else {
            // RHS = D1l^-1 L1^T: (row r, column tid) = D1l^-1[r] L1(tid, r)
#pragma unroll
            for (int k = 0; k < N; ++k) {
                const int pr = wk.perm[k];
                v[k] = wk.sv[0][pr] * __ldcg(L1 + tid + (size_t)pr * N);
            }
        }
        solve_col(wk, v);
        DQ_PH(p, 8 + 4 * rep);
        if (rep == 0) {
            // X = R^-1 D_large^-1, column tid -> global scratch (streamed operand of the products / rows for at_end)
#pragma unroll
            for (int k = 0; k < N; k += 2) *reinterpret_cast<double2*>(wsX + k + (size_t)tid * N) = make_double2(v[k], v[k + 1]);
            __threadfence();
            __syncthreads();
        }
    }
    // ---- G, check_error (source/dqmc.cpp:317-329), install
    if (at_end) {
        // v = column tid of G^T = row tid of G
#pragma unroll
        for (int k = 0; k < N; ++k) {
            double* gp = Gg + tid + (size_t)k * N;
            emax = fmax(emax, fabs(*gp - v[k]));
            *gp = v[k];
        }
    } else {
        // G = X Y: Y(:, tid) = v -> global scratch, X -> tile
#pragma unroll
        for (int k = 0; k < N; k += 2) *reinterpret_cast<double2*>(wsY + k + (size_t)tid * N) = make_double2(v[k], v[k + 1]);
        __threadfence();
        __syncthreads();
        load_tile(wk.T, wsX, nullptr, nullptr);
        gemm_stream<false, SINK_G>(wk, wsY, nullptr, v, Gg, emax);
    }
    emax = block_max(wk, emax);
    DQ_PH(p, 14);
    if (tid == 0) {
        if (at_end) p.logdet[w] = logdet;
        p.err_max[w] = fmax(p.err_max[w], emax);
        p.err_sum[w] += emax;
        p.err_cnt[w] += 1ull;
    }
}

static inline cudaError_t setup() {
    cudaError_t e;
    if ((e = cudaFuncSetAttribute(kf2_stabilize<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Smem::TOTAL)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(kf2_stabilize<-1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Smem::TOTAL)) != cudaSuccess) return e;
    return cudaSuccess;
}

}  // namespace s2
}  // namespace dq
