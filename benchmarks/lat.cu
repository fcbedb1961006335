// benchmarks/lat.cu -- dependent-issue latencies (SM cycles) of the FP64 / shuffle / shared-memory
// operations that sit on the serial chains of the factorisation kernels.  One warp, clock64-timed.
#include <cstdio>
#include <cuda_runtime.h>
#define ITER 512
template <int OP>
__global__ void k(double* out, long long* cyc, double seed) {
    __shared__ double sm[64];
    sm[threadIdx.x] = seed + threadIdx.x;
    sm[threadIdx.x + 32] = seed;
    __syncwarp();
    double x = seed + threadIdx.x * 1e-3, y = 1.0000001;
    int idx = threadIdx.x;
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < ITER; ++i) {
        if (OP == 0) x = fma(x, y, 1e-9);
        if (OP == 1) x = x + y;
        if (OP == 2) x = x * y;
        if (OP == 3) x = (x > y) ? x * 0.5 : x + 1.0;          // DSETP + select-ish
        if (OP == 4) x += __shfl_xor_sync(0xffffffffu, x, 1);  // 64-bit shuffle + DADD
        if (OP == 5) x = 1.0 / (x + 1.0);
        if (OP == 6) x = sqrt(x + 2.0);
        if (OP == 7) idx = __reduce_max_sync(0xffffffffu, (unsigned)idx) + threadIdx.x;
        if (OP == 8) { idx = (int)sm[idx & 63] & 63; }          // dependent LDS (+ cvt)
        if (OP == 9) idx = __shfl_xor_sync(0xffffffffu, idx, 1) + 1;  // 32-bit shuffle + IADD
        if (OP == 10) x = fmax(x * y, 0.5);
        if (OP == 11) { asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(x), "+d"(y) : "d"(1e-9), "d"(1e-9)); }
    }
    long long t1 = clock64();
    out[threadIdx.x] = x + y + idx;
    if (threadIdx.x == 0) cyc[OP] = t1 - t0;
}
int main() {
    double* out; long long* cyc;
    cudaMalloc(&out, 256); cudaMalloc(&cyc, 16 * 8); cudaMemset(cyc, 0, 128);
    const char* names[] = {"dfma", "dadd", "dmul", "dsetp_sel_mul", "shfl64_dadd", "ddiv_dadd", "dsqrt_dadd", "redux_max_iadd", "lds_dependent", "shfl32_iadd", "dmul_dmax", "dmma884"};
#define RUN(OP) k<OP><<<1, 32>>>(out, cyc, 1.5); k<OP><<<1, 32>>>(out, cyc, 1.5);
    RUN(0) RUN(1) RUN(2) RUN(3) RUN(4) RUN(5) RUN(6) RUN(7) RUN(8) RUN(9) RUN(10) RUN(11)
    cudaDeviceSynchronize();
    long long h[16]; cudaMemcpy(h, cyc, 128, cudaMemcpyDeviceToHost);
    printf("{");
    for (int i = 0; i < 12; ++i) printf("%s\"%s_cycles\": %.1f", i ? ", " : "", names[i], (double)h[i] / ITER);
    printf("}\n");
    return 0;
}
