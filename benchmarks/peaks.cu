// benchmarks/peaks.cu -- measures the FP64 and shared-memory denominators the DQMC rooflines need
// and that MEASURED_PEAKS.json does not carry (it has HBM GB/s and bf16 only):
//   fp64 DFMA (vector pipe), fp64 DMMA (mma.sync m8n8k4 / m16n8k8 / m16n8k16 f64), SMEM read and
//   read+write streaming bandwidth.  Prints one JSON object.  Build: nvcc -arch=sm_100a.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

#define CHECK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

__global__ void k_dfma(double* out, int iters) {
    double a[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 1e-3 + i;
    const double b = 1.0000001, c = 1e-9;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) a[i] = fma(a[i], b, c);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void dmma1688(double* c, const double* a, const double* b) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                 : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3]) : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
}
__device__ __forceinline__ void dmma16816(double* c, const double* a, const double* b) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};\n"
                 : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
                 : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]), "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
}

template <int NACC>
__global__ void k_dmma884(double* out, int iters) {
    double c[NACC][2];
#pragma unroll
    for (int i = 0; i < NACC; ++i) c[i][0] = c[i][1] = 0.0;
    const double a = 1.0 + threadIdx.x * 1e-6, b = 1.0 - threadIdx.x * 1e-6;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < NACC; ++i) dmma884(c[i][0], c[i][1], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int NACC>
__global__ void k_dmma1688(double* out, int iters) {
    double c[NACC][4];
#pragma unroll
    for (int i = 0; i < NACC; ++i) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.0;
    double a[4], b[2];
    for (int i = 0; i < 4; ++i) a[i] = 1.0 + threadIdx.x * 1e-6 * (i + 1);
    for (int i = 0; i < 2; ++i) b[i] = 1.0 - threadIdx.x * 1e-6 * (i + 1);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < NACC; ++i) dmma1688(c[i], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int NACC>
__global__ void k_dmma16816(double* out, int iters) {
    double c[NACC][4];
#pragma unroll
    for (int i = 0; i < NACC; ++i) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.0;
    double a[8], b[4];
    for (int i = 0; i < 8; ++i) a[i] = 1.0 + threadIdx.x * 1e-6 * (i + 1);
    for (int i = 0; i < 4; ++i) b[i] = 1.0 - threadIdx.x * 1e-6 * (i + 1);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < NACC; ++i) dmma16816(c[i], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// SMEM streaming: each thread reads (and optionally writes back) 16-byte words, conflict-free
template <bool WRITE>
__global__ void k_smem(double* out, int iters, int words /* double2 words in the buffer */) {
    extern __shared__ double2 sm[];
    for (int i = threadIdx.x; i < words; i += blockDim.x) sm[i] = make_double2(i, -i);
    __syncthreads();
    double2 acc = make_double2(0, 0);
    for (int it = 0; it < iters; ++it) {
#pragma unroll 8
        for (int i = threadIdx.x; i < words; i += blockDim.x) {
            double2 v = sm[i];
            if (WRITE) {
                v.x += 1.0;
                sm[i] = v;
            } else {
                acc.x += v.x;
                acc.y += v.y;
            }
        }
        if (WRITE) __syncthreads();
    }
    if (WRITE) acc = sm[threadIdx.x];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc.x + acc.y;
}

template <typename F>
static double time_ms(F f, int reps = 5) {
    cudaEvent_t e0, e1;
    CHECK(cudaEventCreate(&e0));
    CHECK(cudaEventCreate(&e1));
    f();
    CHECK(cudaDeviceSynchronize());
    double best = 1e30;
    for (int r = 0; r < reps; ++r) {
        CHECK(cudaEventRecord(e0));
        f();
        CHECK(cudaEventRecord(e1));
        CHECK(cudaEventSynchronize(e1));
        float ms;
        CHECK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    CHECK(cudaGetLastError());
    return best;
}

int main() {
    cudaDeviceProp prop;
    CHECK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount;
    double* out;
    CHECK(cudaMalloc(&out, sizeof(double) * sms * 16 * 1024));
    const int iters = 20000;
    printf("{\"gpu\": \"%s\", \"sms\": %d", prop.name, sms);
    {
        const int blocks = sms * 4, threads = 512;
        double ms = time_ms([&] { k_dfma<<<blocks, threads>>>(out, iters); });
        double flops = 2.0 * 16 * iters * (double)blocks * threads;
        printf(", \"fp64_dfma_tflops\": %.2f", flops / ms * 1e-9);
    }
    {
        const int blocks = sms * 4, threads = 256;
        double ms = time_ms([&] { k_dmma884<8><<<blocks, threads>>>(out, iters); });
        double flops = 2.0 * 8 * 8 * 4 * 8 * iters * (double)blocks * (threads / 32);
        printf(", \"fp64_dmma_m8n8k4_tflops\": %.2f", flops / ms * 1e-9);
        ms = time_ms([&] { k_dmma884<8><<<sms, 128>>>(out, iters); });
        flops = 2.0 * 8 * 8 * 4 * 8 * iters * (double)sms * 4;
        printf(", \"fp64_dmma_m8n8k4_4warps_per_sm_tflops\": %.2f", flops / ms * 1e-9);
        ms = time_ms([&] { k_dmma884<16><<<sms, 128>>>(out, iters); });
        flops = 2.0 * 8 * 8 * 4 * 16 * iters * (double)sms * 4;
        printf(", \"fp64_dmma_m8n8k4_4warps_16acc_tflops\": %.2f", flops / ms * 1e-9);
        ms = time_ms([&] { k_dmma884<1><<<sms, 128>>>(out, iters); });
        // dependent chain: latency per DMMA in ns
        printf(", \"dmma_m8n8k4_dependent_latency_ns\": %.2f", ms * 1e6 / iters);
    }
    {
        const int blocks = sms * 4, threads = 256;
        double ms = time_ms([&] { k_dmma1688<8><<<blocks, threads>>>(out, iters); });
        double flops = 2.0 * 16 * 8 * 8 * 8 * iters * (double)blocks * (threads / 32);
        printf(", \"fp64_dmma_m16n8k8_tflops\": %.2f", flops / ms * 1e-9);
        ms = time_ms([&] { k_dmma16816<8><<<blocks, threads>>>(out, iters); });
        flops = 2.0 * 16 * 8 * 16 * 8 * iters * (double)blocks * (threads / 32);
        printf(", \"fp64_dmma_m16n8k16_tflops\": %.2f", flops / ms * 1e-9);
        ms = time_ms([&] { k_dmma16816<8><<<sms, 128>>>(out, iters); });
        flops = 2.0 * 16 * 8 * 16 * 8 * iters * (double)sms * 4;
        printf(", \"fp64_dmma_m16n8k16_4warps_per_sm_tflops\": %.2f", flops / ms * 1e-9);
    }
    {
        const int words = 8192;  // 128 KB
        const size_t bytes = words * sizeof(double2);
        CHECK(cudaFuncSetAttribute(k_smem<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
        CHECK(cudaFuncSetAttribute(k_smem<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
        const int it2 = 2000;
        double ms = time_ms([&] { k_smem<false><<<sms, 1024, bytes>>>(out, it2, words); });
        printf(", \"smem_read_gbs\": %.1f", (double)bytes * it2 * sms / ms * 1e-6);
        ms = time_ms([&] { k_smem<true><<<sms, 1024, bytes>>>(out, it2, words); });
        printf(", \"smem_read_write_gbs\": %.1f", 2.0 * bytes * it2 * sms / ms * 1e-6);
    }
    int clk = 0;
    cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf(", \"sm_clock_khz_max\": %d}\n", clk);
    return 0;
}
