#!/bin/bash
# Run on the GPU box: ncu --set full of the three launches of one mid-sweep stabilisation at N = 256 (148 walkers), reduced to text.
out=gpurun_out/r2n; mkdir -p $out
timeout 300 python benchmarks/prof_run_big.py 16 148 32 30 > $out/plain.log 2>&1 || exit 1
for part in 1 2 3; do
  bash benchmarks/ncu_box.sh $out c3_stab_part$part k_stabilize $((2 + part)) 1 -- python benchmarks/prof_run_big.py 16 148 32 30
done
ls -la $out
