// Development harness: to_LDR (pivoted QR + R scaling + Q formation) of the generic engine alone, in a translation unit that
// compiles in well under a minute.  extern "C" qr_mini(n, batch, dM, dL, dd, dR, ws): device pointers, ws = batch*3*n*LD doubles.
#define DQ_TOLDR_PROF 1
#include <cstdio>
#include <cstdlib>
#include "../../dqmc_b200/csrc/dqmcb_internal.cuh"
#include "../../dqmc_b200/csrc/engine.cuh"
using namespace dq;
__global__ void __launch_bounds__(kThreads) k_mini(Params p, const double* M, double* Lo, double* dO, double* Ro) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int w = blockIdx.x, n = p.N;
    const size_t nn = (size_t)n * n;
    Work wk;
    carve(wk, p, w, p.smem_mats_stab, smem);
    cta_copy(wk.b[0], wk.ld[0], M + w * nn, n, n);
    dev_to_ldr(wk, n, wk.b[0], wk.ld[0], wk.b[1], wk.ld[1], wk.b[2], wk.ld[2], wk.dnew);
    cta_copy(Lo + w * nn, n, wk.b[1], wk.ld[1], n);
    cta_copy(Ro + w * nn, n, wk.b[2], wk.ld[2], n);
    for (int i = threadIdx.x; i < n; i += blockDim.x) dO[(size_t)w * n + i] = wk.dnew[i];
}
__global__ void __launch_bounds__(512) k_mini512(Params p, const double* M, double* Lo, double* dO, double* Ro) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int w = blockIdx.x, n = p.N;
    const size_t nn = (size_t)n * n;
    Work wk;
    carve(wk, p, w, p.smem_mats_stab, smem);
    cta_copy(wk.b[0], wk.ld[0], M + w * nn, n, n);
    dev_to_ldr(wk, n, wk.b[0], wk.ld[0], wk.b[1], wk.ld[1], wk.b[2], wk.ld[2], wk.dnew);
    cta_copy(Lo + w * nn, n, wk.b[1], wk.ld[1], n);
    cta_copy(Ro + w * nn, n, wk.b[2], wk.ld[2], n);
    for (int i = threadIdx.x; i < n; i += blockDim.x) dO[(size_t)w * n + i] = wk.dnew[i];
}
extern "C" int qr_mini(int n, int batch, const double* dM, double* dL, double* dd, double* dR, double* ws) {
    Params p;
    memset(&p, 0, sizeof(p));
    p.N = n;
    p.LD = (n % 2 == 0) ? n + 1 : n;
    p.W = batch;
    p.ws = ws;
    const size_t tile = (size_t)n * p.LD * 8, sb = ((small_bytes(n) + 15) & ~(size_t)15);
    p.smem_mats_stab = (sb + 3 * tile <= (size_t)kMaxSmem) ? 3 : 0;
    p.smem_stage = (p.smem_mats_stab == 0 && sb + tile <= (size_t)kMaxSmem) ? 1 : 0;
    const size_t smem = sb + (p.smem_mats_stab + p.smem_stage) * tile;
    const char* t5 = getenv("QR_MINI_512");
    if (t5 && t5[0] == (char)49) {
        cudaFuncSetAttribute(k_mini512, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k_mini512<<<batch, 512, smem>>>(p, dM, dL, dd, dR);
    } else {
        cudaFuncSetAttribute(k_mini, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k_mini<<<batch, kThreads, smem>>>(p, dM, dL, dd, dR);
    }
    cudaDeviceSynchronize();
    long long h[8];
    cudaMemcpyFromSymbol(h, g_toldr_prof, sizeof(h));
    printf("to_LDR phases (CTA 0, cycles): stage_in %lld, QR %lld, d + R %lld, form Q %lld\n", h[1] - h[0], h[2] - h[1], h[3] - h[2], h[4] - h[3]);
    return (int)cudaGetLastError();
}
