// Equal-time measurement pipeline on the device.
//
// Reference: include/measurementh5.h -- MeasurementManager::measure (:189-229) sums, per sweep, the
// three scalars and the N x N connected density correlation of Observables:: (source/model.cpp:167-288);
// accumulate (:231-274) divides by the count and, on the way to disk, folds site pairs onto
// displacements (transform::chi_site_to_chi_r :20-66) and Fourier-transforms (chi_r_to_chi_k :78-116).
//
// Here: the fold is linear, so each walker's correlation is folded straight from G into L1*L2
// displacement bins (the N x N matrix is never materialised: 8N^2 bytes read, 8N written per walker
// instead of 8N^2 written + read again), a second kernel adds the W per-walker vectors into the bin
// accumulator in walker order (deterministic, no floating-point atomics), and only 3 + L1*L2 doubles
// per bin ever cross PCIe or NVLink.  Geometry as in source/main.cpp:84-88: a1 = (1,0), a2 = (0,1),
// one orbital, site i = uy*L1 + ux.
#pragma once
#include "dqmcb_internal.cuh"
#include "linalg.cuh"

namespace dq {

// transform::pbc_shortest, include/measurementh5.h:13-17
__host__ __device__ inline int pbc_shortest(int d, int L) {
    if (d > L / 2) d -= L;
    if (d <= -L / 2) d += L;
    return d;
}

// site j whose displacement from site i falls in bin (bx, by) = (dx + L1/2 - 1, dy + L2/2 - 1).
// pbc_shortest is a bijection between the L residues and its L output values, so j is unique.
__device__ __forceinline__ int site_at_bin(int i, int bx, int by, int L1, int L2) {
    const int cxi = i % L1, cyi = i / L1;
    const int ddx = bx - (L1 / 2 - 1), ddy = by - (L2 / 2 - 1);
    const int cxj = ((cxi + ddx) % L1 + L1) % L1;
    const int cyj = ((cyi + ddy) % L2 + L2) % L2;
    return cyj * L1 + cxj;
}

// One CTA per walker: part[w] = { density, doubleOcc, swave, chi_r[L1*L2] } of the walker's current G.
// chi_r[bx + L1*by] = (1/N) sum_i C(i, j(i; bx, by)),  C = calculate_densityCorr (source/model.cpp:258-288).
__global__ void __launch_bounds__(kThreads) k_bin_measure(Params p, int L1, int L2, double* __restrict__ part) {
    __shared__ double red[64];
    __shared__ double s_navg;
    const int w = blockIdx.x, n = p.N;
    const double* G = p.G + (size_t)w * n * n;
    double* out = part + (size_t)w * (3 + n);
    double dsum = 0.0, d2sum = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const double gc = 1.0 - G[i + (size_t)i * n];
        dsum += gc + gc;
        d2sum += gc * gc;
    }
    dsum = cta_reduce_sum(dsum, red);
    d2sum = cta_reduce_sum(d2sum, red);
    double ssum = 0.0;
    for (int e = threadIdx.x; e < n * n; e += blockDim.x) {
        const int j = e % n, i = e / n;
        const double gc = ((i == j) ? 1.0 : 0.0) - G[j + (size_t)i * n];
        ssum += gc * gc;
    }
    ssum = cta_reduce_sum(ssum, red);
    if (threadIdx.x == 0) {
        out[0] = dsum / n;
        out[1] = d2sum / n;
        out[2] = ssum / n;
        s_navg = dsum / n;  // n_avg of calculate_densityCorr equals the density
    }
    __syncthreads();
    const double navg = s_navg;
    // bins x sites: a group of `tpb` adjacent threads shares one bin and splits the sites
    int tpb = 1;
    while (tpb * 2 * n <= (int)blockDim.x && tpb < 32) tpb *= 2;
    const int sub = threadIdx.x % tpb;
    for (int b0 = 0; b0 < n; b0 += blockDim.x / tpb) {
        const int b = b0 + threadIdx.x / tpb;
        double acc = 0.0;
        if (b < n) {
            const int bx = b % L1, by = b / L1;
            for (int i = sub; i < n; i += tpb) {
                const int j = site_at_bin(i, bx, by, L1, L2);
                const double ni = 2.0 * (1.0 - G[i + (size_t)i * n]);
                const double nj = 2.0 * (1.0 - G[j + (size_t)j * n]);
                const double ex = 2.0 * (1.0 - G[j + (size_t)i * n]) * G[i + (size_t)j * n];  // Q5: 1.0, not delta_ij
                acc += (ni * nj + ex - navg * navg) / n;
            }
        }
        for (int o = tpb >> 1; o > 0; o >>= 1) acc += __shfl_xor_sync(DQ_FULL, acc, o);
        if (b < n && sub == 0) out[3 + b] = acc;
    }
}

// acc[k] += sum_w part[w][k] in walker order (one thread per component: coalesced, deterministic)
__global__ void k_bin_reduce(int W, int len, const double* __restrict__ part, double* __restrict__ acc) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= len) return;
    double s = 0.0;
    for (int w = 0; w < W; ++w) s += part[(size_t)w * len + k];
    acc[k] += s;
}

// transform::chi_site_to_chi_r for a batch of N x N column-major matrices (one CTA each):
// chi_r[m][bx + L1*by] = (1/N) sum_i chi_site[m][i + N*j(i; bx, by)]
__global__ void k_chi_site_to_chi_r(int L1, int L2, const double* __restrict__ chi_site, double* __restrict__ chi_r) {
    const int n = L1 * L2;
    const double* C = chi_site + (size_t)blockIdx.x * n * n;
    double* out = chi_r + (size_t)blockIdx.x * n;
    for (int b = threadIdx.x; b < n; b += blockDim.x) {
        const int bx = b % L1, by = b / L1;
        double acc = 0.0;
        for (int i = 0; i < n; ++i) acc += C[i + (size_t)site_at_bin(i, bx, by, L1, L2) * n] / n;
        out[b] = acc;
    }
}

// transform::chi_r_to_chi_k for a batch of L1 x L2 slices (one CTA each), k-points of
// include/lattice.h:36-49: chi_k[m][n + L1*mm] = sum_{x,y} chi_r[m][x + L1*y] exp(-i k_{n,mm} . r_{x,y}),
// interleaved (re, im).  The reference files k-point kidx = n*L2 + mm under (kidx / L1, kidx % L2):
// (n, mm) for L1 == L2, out of bounds otherwise; (n, mm) is used here.
__global__ void k_chi_r_to_chi_k(int L1, int L2, const double* __restrict__ chi_r, double* __restrict__ chi_k) {
    const int n = L1 * L2;
    const double* R = chi_r + (size_t)blockIdx.x * n;
    double* out = chi_k + (size_t)blockIdx.x * n * 2;
    const double b1x = 2.0 * M_PI / L1, b2y = 2.0 * M_PI / L2;
    for (int kidx = threadIdx.x; kidx < n; kidx += blockDim.x) {
        const int kn = kidx / L2, km = kidx % L2;
        const double kx = (kn - L1 / 2 + 1) * b1x, ky = (km - L2 / 2 + 1) * b2y;
        double re = 0.0, im = 0.0;
        for (int x = 0; x < L1; ++x)
            for (int y = 0; y < L2; ++y) {
                const double phase = kx * (double)(x - (L1 / 2 - 1)) + ky * (double)(y - (L2 / 2 - 1));
                double sn, cs;
                sincos(phase, &sn, &cs);
                const double v = R[x + L1 * y];
                re += v * cs;
                im -= v * sn;
            }
        out[2 * (kn + L1 * km)] = re;
        out[2 * (kn + L1 * km) + 1] = im;
    }
}

// ------------------------------------------------------------------------------------------------
// Kinetic energy and spin correlation (north_star's correctness clause names them; the reference has NO code for either:
// source/model.cpp:167-288 stops at density / double occupancy / s-wave / density correlation).  Per HS configuration the
// two spin species share G (attractive model, charge-channel decoupling), <c+_i c_j> = delta_ij - G(j, i), so by Wick:
//   E_kin / N = (2 t / N) sum_i sum_{d in {+x, +y}} [ G(i + d, i) + G(i, i + d) ]          (both spins, PBC, site x + L1 y)
//   <Sz_i Sz_j> = (1/2) (delta_ij - G(j, i)) G(i, j)
//   szz_r(dx, dy) = (1/N) sum_i <Sz_i Sz_{i + (dx, dy)}>          (displacement fold; S(q) = sum_r cos(q r) szz_r on the host)
// One CTA per walker; thread r owns displacement r and sums over the N origins (deterministic, no atomics).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) k_kinetic_spin(Params p, int L1, int L2, double t, double* __restrict__ ekin /*[W]*/,
                                                           double* __restrict__ szz /*[W][N]*/) {
    __shared__ double red[64];
    const int w = blockIdx.x, n = p.N;
    const double* G = p.G + (size_t)w * n * n;
    double part = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int x = i % L1, y = i / L1;
        const int jx = (x + 1) % L1 + L1 * y, jy = x + L1 * ((y + 1) % L2);
        part += G[jx + (size_t)i * n] + G[i + (size_t)jx * n] + G[jy + (size_t)i * n] + G[i + (size_t)jy * n];
    }
    const double tot = cta_reduce_sum(part, red);
    if (threadIdx.x == 0) ekin[w] = 2.0 * t * tot / n;
    for (int r = threadIdx.x; r < n; r += blockDim.x) {
        const int dx = r % L1, dy = r / L1;
        double s = 0.0;
        for (int i = 0; i < n; ++i) {
            const int x = i % L1, y = i / L1;
            const int j = (x + dx) % L1 + L1 * ((y + dy) % L2);
            s += 0.5 * (((i == j) ? 1.0 : 0.0) - G[j + (size_t)i * n]) * G[i + (size_t)j * n];
        }
        szz[(size_t)w * n + r] = s / n;
    }
}

}  // namespace dq
