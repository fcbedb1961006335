// benchmarks/tput.cu -- issue rate of DFMA / DMMA.8x8x4 / broadcast LDS.128 on ONE SM as a function of warps per
// scheduler (1, 2, 4) and independent chains per thread (ILP): what a warp-pair-per-walker kernel can reach.
// Output: cycles per warp-instruction per scheduler (SMSP).
#include <cstdio>
#include <cuda_runtime.h>
#define ITER 256
template <int OP, int ILP>
__global__ void k(double* out, long long* cyc, double seed) {
    __shared__ __align__(16) double sm[512];
    for (int i = threadIdx.x; i < 512; i += blockDim.x) sm[i] = seed + i;
    __syncthreads();
    double x[ILP], y[ILP];
#pragma unroll
    for (int j = 0; j < ILP; ++j) { x[j] = seed + threadIdx.x * 1e-3 + j; y[j] = 1e-9 * j; }
    const double m = 1.0000001;
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < ITER; ++i) {
#pragma unroll
        for (int j = 0; j < ILP; ++j) {
            if (OP == 0) x[j] = fma(x[j], m, 1e-9);
            if (OP == 1) asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(x[j]), "+d"(y[j]) : "d"(m), "d"(m));
            if (OP == 2) { const double2 v = *reinterpret_cast<const double2*>(sm + ((i * 2 + j * 4) & 510)); x[j] = fma(v.x, m, x[j]); y[j] = fma(v.y, m, y[j]); }
        }
    }
    long long t1 = clock64();
    double s = 0; 
#pragma unroll
    for (int j = 0; j < ILP; ++j) s += x[j] + y[j];
    out[threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
template <int OP, int ILP>
void run(const char* name, double* out, long long* cyc) {
    for (int wps = 1; wps <= 4; wps *= 2) {
        k<OP, ILP><<<1, 128 * wps>>>(out, cyc, 1.5);
        k<OP, ILP><<<1, 128 * wps>>>(out, cyc, 1.5);
        cudaDeviceSynchronize();
        long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
        const double inst = (double)ITER * ILP * wps * (OP == 2 ? 3 : 1);
        printf("\"%s_ilp%d_wps%d\": %.2f, ", name, ILP, wps, (double)h / inst);
    }
    printf("\n");
}
int main() {
    double* out; long long* cyc;
    cudaMalloc(&out, 8192); cudaMalloc(&cyc, 64);
    printf("{\n");
    run<0, 1>("dfma", out, cyc); run<0, 2>("dfma", out, cyc); run<0, 4>("dfma", out, cyc); run<0, 8>("dfma", out, cyc); run<0, 16>("dfma", out, cyc);
    run<1, 1>("dmma", out, cyc); run<1, 2>("dmma", out, cyc); run<1, 4>("dmma", out, cyc); run<1, 8>("dmma", out, cyc);
    run<2, 4>("lds128_2fma", out, cyc); run<2, 8>("lds128_2fma", out, cyc);
    printf("\"unit\": \"cycles per warp-instruction per scheduler\"}\n");
    return 0;
}
