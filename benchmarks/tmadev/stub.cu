// Development aid: compiles gemm_tma.cuh alone (seconds) to read ptxas resource usage / SASS.
#include "../../dqmc_b200/csrc/dqmcb_internal.cuh"
#include "../../dqmc_b200/csrc/gemm_tma.cuh"
extern "C" __global__ void __launch_bounds__(256, 1) k_gemm(double* C, const double* A, const double* B, const double* rs, const double* cs, int acc) {
    extern __shared__ __align__(128) unsigned char smem[];
    dq::cta_gemm_tma256(C, 256, A, 256, B, 256, rs, cs, acc != 0, smem);
}
extern "C" __global__ void __launch_bounds__(256, 1) k_solve(const double* A, double* B) {
    extern __shared__ __align__(128) unsigned char smem[];
    dq::cta_lu_solve_tma256(A, 256, B, 256, smem);
}
