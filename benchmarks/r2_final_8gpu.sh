#!/bin/bash
# Round-2 evidence on 8 GPUs of one box: the multi-GPU test, C4 on both ladders, the C1 weak-scaling line.
mkdir -p gpurun_out/r2g
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_gpu_multi.py -q 2>&1 | tail -5 > gpurun_out/r2g/pytest_multi.log
timeout 600 $TR --master-port 29611 benchmarks/replica_exchange_run.py --ladder beta --sweeps 60 --every 4 > gpurun_out/r2g/c4_beta.json 2> gpurun_out/r2g/c4_beta.err
timeout 600 $TR --master-port 29612 benchmarks/replica_exchange_run.py --ladder U --beta0 8 --nt 80 --sweeps 60 --every 4 > gpurun_out/r2g/c4_U.json 2> gpurun_out/r2g/c4_U.err
timeout 900 $TR --master-port 29613 bench.py --gpus $N --steps 60 --warmup 5 > gpurun_out/r2g/bench_C1_${N}gpu.json 2> gpurun_out/r2g/bench_C1_${N}gpu.err
