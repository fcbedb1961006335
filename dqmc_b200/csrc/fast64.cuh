// fast64.cuh -- N = 64 fast path (placeholder until the DMMA / register-resident kernels land).
#pragma once
#include "dqmcb_internal.cuh"

static inline cudaError_t fast64_setup(dqmcb_ctx*) { return cudaSuccess; }
static inline bool fast64_available(const dqmcb_ctx*) { return false; }
static inline int fast64_propagate(dqmcb_ctx*, const dq::Params&, int, int, int, unsigned long long, int) { return DQMCB_ESTATE; }
