// Development aid: stand-alone timing / correctness harness of gemm_tma.cuh (N = 256), one CTA per matrix, 148 matrices.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -o harness harness.cu && ./harness
#define DQ_TMA_PROF 1
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../dqmc_b200/csrc/dqmcb_internal.cuh"
#include "../../dqmc_b200/csrc/gemm_tma.cuh"
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)
constexpr int N = 256;
__global__ void __launch_bounds__(256, 1) k_gemm(double* C, const double* A, const double* B, int ld, int variant, long long* cyc) {
    extern __shared__ __align__(128) unsigned char smem[];
    const size_t o = (size_t)blockIdx.x * N * ld;
    const long long t0 = clock64();
    if (variant == 0) dq::cta_gemm_tma256(C + o, ld, A + o, ld, B + o, ld, nullptr, nullptr, false, smem);
    else dq::cta_gemm<false>(C + o, ld, A + o, 1, ld, B + o, 1, ld, N, nullptr, nullptr);
    __syncthreads();
    if (threadIdx.x == 0) cyc[blockIdx.x] = clock64() - t0;
}
__global__ void __launch_bounds__(256, 1) k_solve(const double* A, double* B, int ld, int variant, long long* cyc) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ int ipiv[N];
    const size_t o = (size_t)blockIdx.x * N * ld;
    for (int i = threadIdx.x; i < N; i += blockDim.x) ipiv[i] = i;
    __syncthreads();
    const long long t0 = clock64();
    if (variant == 0) dq::cta_lu_solve_tma256(A + o, ld, B + o, ld, smem);
    else dq::cta_lu_solve(A + o, ld, N, ipiv, B + o, 1, ld, true);
    __syncthreads();
    if (threadIdx.x == 0) cyc[blockIdx.x] = clock64() - t0;
}
__global__ void k_ref_gemm(double* C, const double* A, const double* B, int ld) {
    const size_t o = (size_t)blockIdx.x * N * ld;
    for (int e = threadIdx.x; e < N * N; e += blockDim.x) {
        const int r = e % N, c = e / N;
        double s = 0;
        for (int k = 0; k < N; ++k) s += A[o + r + (size_t)k * ld] * B[o + k + (size_t)c * ld];
        C[o + r + (size_t)c * ld] = s;
    }
}
// residual of (L U) X = B0 with unit-lower L and upper U packed in A
__global__ void k_ref_lux(double* R, const double* A, const double* X, int ld) {
    const size_t o = (size_t)blockIdx.x * N * ld;
    for (int e = threadIdx.x; e < N * N; e += blockDim.x) {
        const int r = e % N, c = e / N;
        double s = 0;  // (L (U X))(r, c) = sum_k L(r,k) (UX)(k,c): do it in two stages on the fly (O(n^2) per element is fine here)
        for (int k = 0; k <= r; ++k) {
            double ux = 0;
            for (int j = k; j < N; ++j) ux += A[o + k + (size_t)j * ld] * X[o + j + (size_t)c * ld];
            s += (k == r ? 1.0 : A[o + r + (size_t)k * ld]) * ux;
        }
        R[o + r + (size_t)c * ld] = s;
    }
}
int main(int argc, char** argv) {
    const int W = 148, nb = argc > 1 ? atoi(argv[1]) : 1;
    for (int ld : {256, 258}) {
        const size_t cnt = (size_t)W * N * ld;
        std::vector<double> hA(cnt), hB(cnt);
        srand(1);
        for (size_t i = 0; i < cnt; ++i) { hA[i] = rand() / (double)RAND_MAX - 0.5; hB[i] = rand() / (double)RAND_MAX - 0.5; }
        // make A a well-conditioned LU packing for the solve: dominant diagonal
        std::vector<double> hLU = hA;
        for (int w = 0; w < W; ++w) for (int i = 0; i < N; ++i) hLU[(size_t)w * N * ld + i + (size_t)i * ld] = 8.0 + hA[(size_t)w * N * ld + i + (size_t)i * ld];
        for (size_t i = 0; i < cnt; ++i) hLU[i] *= 0.25;
        double *A, *B, *C, *Cr, *LU, *X;
        long long* cyc;
        CK(cudaMalloc(&A, cnt * 8)); CK(cudaMalloc(&B, cnt * 8)); CK(cudaMalloc(&C, cnt * 8)); CK(cudaMalloc(&Cr, cnt * 8)); CK(cudaMalloc(&LU, cnt * 8)); CK(cudaMalloc(&X, cnt * 8));
        CK(cudaMalloc(&cyc, W * 8));
        CK(cudaMemcpy(A, hA.data(), cnt * 8, cudaMemcpyHostToDevice)); CK(cudaMemcpy(B, hB.data(), cnt * 8, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(LU, hLU.data(), cnt * 8, cudaMemcpyHostToDevice));
        const int smem = (int)dq::kGemmTmaArenaBytes;
        CK(cudaFuncSetAttribute(k_gemm, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        CK(cudaFuncSetAttribute(k_solve, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        k_ref_gemm<<<W, 256>>>(Cr, A, B, ld);
        CK(cudaDeviceSynchronize());
        std::vector<double> hC(cnt), hR(cnt);
        CK(cudaMemcpy(hR.data(), Cr, cnt * 8, cudaMemcpyDeviceToHost));
        for (int variant = 0; variant < 2; ++variant) {
            cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
            CK(cudaMemset(C, 0, cnt * 8));
            k_gemm<<<W, 256, smem>>>(C, A, B, ld, variant, cyc);
            CK(cudaDeviceSynchronize());
            cudaEventRecord(e0);
            for (int i = 0; i < nb; ++i) k_gemm<<<W, 256, smem>>>(C, A, B, ld, variant, cyc);
            cudaEventRecord(e1);
            CK(cudaDeviceSynchronize());
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            CK(cudaMemcpy(hC.data(), C, cnt * 8, cudaMemcpyDeviceToHost));
            double err = 0;
            for (int w = 0; w < W; ++w) for (int c = 0; c < N; ++c) for (int r = 0; r < N; ++r) { const size_t i = (size_t)w * N * ld + r + (size_t)c * ld; err = fmax(err, fabs(hC[i] - hR[i])); }
            long long hc[148]; CK(cudaMemcpy(hc, cyc, W * 8, cudaMemcpyDeviceToHost));
            long long mx = 0; for (int w = 0; w < W; ++w) mx = hc[w] > mx ? hc[w] : mx;
            printf("ld=%d gemm %s: %.3f ms/launch, max cycles %lld, %.1f TFLOP/s, max|err| %.2e\n", ld, variant ? "ring" : "tma ", ms / nb, mx, 2.0 * N * N * N * W / (ms / nb * 1e-3) / 1e12, err);
        }
        for (int variant = 0; variant < 2; ++variant) {
            cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
            CK(cudaMemcpy(X, B, cnt * 8, cudaMemcpyDeviceToDevice));
            k_solve<<<W, 256, smem>>>(LU, X, ld, variant, cyc);
            CK(cudaDeviceSynchronize());
            k_ref_lux<<<W, 256>>>(Cr, LU, X, ld);
            CK(cudaDeviceSynchronize());
            CK(cudaMemcpy(hC.data(), Cr, cnt * 8, cudaMemcpyDeviceToHost));
            double err = 0;
            for (int w = 0; w < W; ++w) for (int c = 0; c < N; ++c) for (int r = 0; r < N; ++r) { const size_t i = (size_t)w * N * ld + r + (size_t)c * ld; err = fmax(err, fabs(hC[i] - hB[i])); }
            cudaEventRecord(e0);
            for (int i = 0; i < nb; ++i) k_solve<<<W, 256, smem>>>(LU, X, ld, variant, cyc);  // (solving the solution again: same work)
            cudaEventRecord(e1);
            CK(cudaDeviceSynchronize());
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            long long hc[148]; CK(cudaMemcpy(hc, cyc, W * 8, cudaMemcpyDeviceToHost));
            long long mx = 0; for (int w = 0; w < W; ++w) mx = hc[w] > mx ? hc[w] : mx;
            printf("ld=%d solve %s: %.3f ms/launch, max cycles %lld, residual %.2e\n", ld, variant ? "blocked" : "tma    ", ms / nb, mx, err);
            if (variant == 0) {
                long long hp[8]; CK(cudaMemcpyFromSymbol(hp, dq::g_tma_prof, sizeof(hp)));
                printf("   per call (CTA 0, thread 0): issue %lld, wait %lld, substitution %lld, update %lld cycles\n", hp[0] / (nb + 1), hp[1] / (nb + 1), hp[2] / (nb + 1), hp[3] / (nb + 1));
                long long z[8] = {0}; CK(cudaMemcpyToSymbol(dq::g_tma_prof, z, sizeof(z)));
            }
        }
        cudaFree(A); cudaFree(B); cudaFree(C); cudaFree(Cr); cudaFree(LU); cudaFree(X); cudaFree(cyc);
    }
    return 0;
}
