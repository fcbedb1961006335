#!/bin/bash
# Development aid: C3 with and without the TMA-staged N = 256 product / solve (gemm_tma.cuh); parity first.
mkdir -p gpurun_out/r2z
timeout 600 python -m pytest tests/test_gpu_parity.py -q -k "stablelinalg" 2>&1 | tail -15 > gpurun_out/r2z/stablelinalg.log
timeout 900 python -m pytest tests/test_gpu_full_configs.py -q -k "c3" 2>&1 | tail -5 > gpurun_out/r2z/c3_parity.log
for v in 0 1; do
  DQMCB_NO_TMA_GEMM=$v timeout 600 python benchmarks/phase_prof_big.py 16 148 32 > gpurun_out/r2z/phase_c3_notma$v.log 2>&1
  DQMCB_NO_TMA_GEMM=$v timeout 900 python bench.py --workload C3 --steps 5 --warmup 3 --therm 5 --no-cpu-baseline 2>/dev/null > gpurun_out/r2z/bench_c3_notma$v.json
done
