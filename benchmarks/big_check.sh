#!/bin/bash
# Development aid: parity of the large-lattice engine (full-config goldens, stablelinalg, block products), then phase profiles and short benches.
mkdir -p gpurun_out/r2z
timeout 1200 python -m pytest tests/test_gpu_full_configs.py tests/test_gpu_parity.py -q -k "c2 or c3 or C2s or C3s or stablelinalg or global_memory" 2>&1 | tail -6 > gpurun_out/r2z/big_parity.log
timeout 600 python benchmarks/phase_prof_big.py 16 148 32 > gpurun_out/r2z/phase_c3.log 2>&1
timeout 600 python benchmarks/phase_prof_big.py 12 148 16 > gpurun_out/r2z/phase_c2.log 2>&1
timeout 900 python bench.py --workload C3 --steps 5 --warmup 3 --therm 5 --no-cpu-baseline 2>/dev/null > gpurun_out/r2z/bench_c3.json
timeout 900 python bench.py --workload C2 --steps 10 --warmup 3 --therm 5 --no-cpu-baseline 2>/dev/null > gpurun_out/r2z/bench_c2.json
