#!/bin/bash
# Round-2 validation on one B200: the whole GPU suite, smoke(), then the bench lines of the single-GPU workloads
# (C1 exactly as the driver runs it: no flags; then the reference arm; then C2, C3 at both n_stab, C0).
mkdir -p gpurun_out/r2f
timeout 2400 python -m pytest tests -q -m gpu 2>&1 | tail -8 > gpurun_out/r2f/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2f/smoke.log 2>&1
python bench.py > gpurun_out/r2f/bench_C1.json 2> gpurun_out/r2f/bench_C1.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2f/bench_C1_reference_arm.json 2> gpurun_out/r2f/bench_C1_reference_arm.err
python bench.py --workload C2 --steps 15 --warmup 3 > gpurun_out/r2f/bench_C2.json 2> gpurun_out/r2f/bench_C2.err
python bench.py --workload C3 --steps 6 --warmup 3 --therm 20 > gpurun_out/r2f/bench_C3.json 2> gpurun_out/r2f/bench_C3.err
python bench.py --workload C3 --n-stab 5 --steps 4 --warmup 3 --therm 10 > gpurun_out/r2f/bench_C3_nstab5.json 2> gpurun_out/r2f/bench_C3_nstab5.err
python bench.py --workload C0 --steps 40 --warmup 5 > gpurun_out/r2f/bench_C0.json 2> gpurun_out/r2f/bench_C0.err
