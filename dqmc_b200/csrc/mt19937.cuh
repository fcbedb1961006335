// Device-resident std::mt19937 per walker (DQMCB_RNG_MT19937): the libstdc++ random stream of the reference
// regenerated on the GPU, so that a seed reproduces the reference's trajectory with no host tape at all.
//
// What is reproduced (restated and pinned in oracle/dqmc_oracle.py: MT19937, _lemire,
// std_shuffle, rng_tape): per update::local_update call (source/update.cpp:10-25)
//   * std::shuffle of the site order (libstdc++ bits/stl_algo.h: two swap positions per draw, Lemire's
//     multiply-shift uniform_int with rejection, bits/uniform_int_dist.h);
//   * per site: the proposal index from a BY-VALUE COPY of the generator (include/field.h:76: it re-uses the first
//     word of the next draw) and bernoulli's generate_canonical<double,53> (two 32-bit words).
// One thread per walker; the 624-word state and the permutation live in shared memory, transposed so that the
// 64 walkers of a CTA hit 64 different banks.  Output goes to the context's tape buffers in visiting order.
#pragma once
#include "dqmcb_internal.cuh"

namespace dq {

constexpr int kMtWalkers = 64;  // walkers (threads) per CTA

struct MtGen {
    uint32_t* st;  // shared: st[i * kMtWalkers] = word i of this walker's state
    int idx;
    __device__ __forceinline__ uint32_t& at(int i) { return st[i * kMtWalkers]; }
    __device__ void twist() {
        for (int i = 0; i < 624; ++i) {
            const uint32_t y = (at(i) & 0x80000000u) | (at((i + 1) % 624) & 0x7fffffffu);
            at(i) = at((i + 397) % 624) ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
        }
        idx = 0;
    }
    __device__ __forceinline__ uint32_t next() {
        if (idx >= 624) twist();
        uint32_t y = at(idx++);
        y ^= y >> 11;
        y ^= (y << 7) & 0x9d2c5680u;
        y ^= (y << 15) & 0xefc60000u;
        y ^= y >> 18;
        return y;
    }
    // uniform_int_distribution over [0, range) for a 32-bit URNG (libstdc++ _S_nd)
    __device__ __forceinline__ uint32_t lemire(uint32_t range) {
        unsigned long long product = (unsigned long long)next() * range;
        uint32_t low = (uint32_t)product;
        if (low < range) {
            const uint32_t threshold = (0u - range) % range;
            while (low < threshold) {
                product = (unsigned long long)next() * range;
                low = (uint32_t)product;
            }
        }
        return (uint32_t)(product >> 32);
    }
};

// std::mt19937(seed) for every walker: mt[w][0..623] state, mt[w][624] = position (624 = "twist before use")
__global__ void k_mt_seed(int W, const uint32_t* __restrict__ seeds, uint32_t* __restrict__ mt) {
    const int w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= W) return;
    uint32_t* s = mt + (size_t)w * 625;
    uint32_t x = seeds[w];
    s[0] = x;
    for (uint32_t i = 1; i < 624; ++i) {
        x = 1812433253u * (x ^ (x >> 30)) + i;
        s[i] = x;
    }
    s[624] = 624u;
}

// draws of `n_visits` consecutive local_update calls of every walker -> tape[w][visit][site]
__global__ void __launch_bounds__(kMtWalkers) k_mt_fill(int W, int N, int n_visits, uint32_t* __restrict__ mt, uint8_t* __restrict__ tperm,
                                                        uint8_t* __restrict__ tk, double* __restrict__ tu) {
    extern __shared__ __align__(16) unsigned char smem[];
    uint32_t* st_all = reinterpret_cast<uint32_t*>(smem);                      // [624][64]
    uint8_t* perm_all = smem + (size_t)624 * kMtWalkers * sizeof(uint32_t);    // [N][64]
    const int t = threadIdx.x, w = blockIdx.x * kMtWalkers + t;
    if (w >= W) return;
    MtGen g{st_all + t, 0};
    uint32_t* gs = mt + (size_t)w * 625;
    for (int i = 0; i < 624; ++i) g.at(i) = gs[i];
    g.idx = (int)gs[624];
    uint8_t* a = perm_all + t;  // a[i * kMtWalkers]
    for (int v = 0; v < n_visits; ++v) {
        const size_t off = ((size_t)w * n_visits + v) * N;
        // std::shuffle(0..N-1) (N < 65536: the pairwise branch)
        for (int i = 0; i < N; ++i) a[i * kMtWalkers] = (uint8_t)i;
        if (N >= 2) {
            int i = 1;
            if ((N & 1) == 0) {
                const int j = (int)g.lemire(2u);
                const uint8_t tmp = a[i * kMtWalkers];
                a[i * kMtWalkers] = a[j * kMtWalkers];
                a[j * kMtWalkers] = tmp;
                ++i;
            }
            while (i < N) {
                const uint32_t sr = (uint32_t)i + 1u;
                const uint32_t x = g.lemire(sr * (sr + 1u));
                const int p0 = (int)(x / (sr + 1u)), p1 = (int)(x % (sr + 1u));
                uint8_t tmp = a[i * kMtWalkers];
                a[i * kMtWalkers] = a[p0 * kMtWalkers];
                a[p0 * kMtWalkers] = tmp;
                ++i;
                tmp = a[i * kMtWalkers];
                a[i * kMtWalkers] = a[p1 * kMtWalkers];
                a[p1 * kMtWalkers] = tmp;
                ++i;
            }
        }
        for (int idx = 0; idx < N; ++idx) {
            const uint32_t x = g.next(), y = g.next();
            unsigned long long prod = (unsigned long long)x * 3u;  // the by-value copy sees x first
            if ((uint32_t)prod < 3u && (uint32_t)prod < (0u - 3u) % 3u) prod = (unsigned long long)y * 3u;
            double val = ((double)x + (double)y * 4294967296.0) * (1.0 / 18446744073709551616.0);
            if (val >= 1.0) val = 0.99999999999999988897769753748;  // nextafter(1, 0)
            tperm[off + idx] = a[idx * kMtWalkers];
            tk[off + idx] = (uint8_t)(prod >> 32);
            tu[off + idx] = val;
        }
    }
    for (int i = 0; i < 624; ++i) gs[i] = g.at(i);
    gs[624] = (uint32_t)g.idx;
}

}  // namespace dq
