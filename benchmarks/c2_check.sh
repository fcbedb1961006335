#!/bin/bash
# Development aid: N = 144 parity, phase profile and a short C2 bench, with and without the right-hand-side panel solve.
mkdir -p gpurun_out/r2z
timeout 1200 python -m pytest tests/test_gpu_full_configs.py tests/test_gpu_parity.py -q -k "c2 or C2s or stablelinalg or global_memory" 2>&1 | tail -6 > gpurun_out/r2z/c2_parity.log
for v in 0 1; do
  DQMCB_NO_RHS_PANEL=$v timeout 600 python benchmarks/phase_prof_big.py 12 148 16 > gpurun_out/r2z/phase_c2_nopanel$v.log 2>&1
  DQMCB_NO_RHS_PANEL=$v timeout 900 python bench.py --workload C2 --steps 10 --warmup 3 --therm 5 --no-cpu-baseline 2>/dev/null > gpurun_out/r2z/bench_c2_nopanel$v.json
done
