#!/bin/bash
# Round-2 validation on one B200: the whole GPU suite, then the bench lines of the four single-GPU workloads.
mkdir -p gpurun_out/r2f
timeout 2400 python -m pytest tests -q -m gpu 2>&1 | tail -8 > gpurun_out/r2f/pytest_gpu.log
python bench.py --workload C3 --steps 6 --warmup 3 --therm 20 > gpurun_out/r2f/bench_C3.json 2> gpurun_out/r2f/bench_C3.err
python bench.py --workload C3 --n-stab 5 --steps 4 --warmup 3 --therm 10 --no-cpu-baseline > gpurun_out/r2f/bench_C3_nstab5.json 2> gpurun_out/r2f/bench_C3_nstab5.err
python bench.py --workload C0 --steps 40 --warmup 5 > gpurun_out/r2f/bench_C0.json 2> gpurun_out/r2f/bench_C0.err
