#!/bin/bash
# Development aid: C0 (or $1) throughput of the generic engine as a function of threads per walker.
wl=${1:-C0}
for t in 256 128 64 32; do
  DQMCB_GENERIC_THREADS=$t python bench.py --workload $wl --steps 20 --warmup 3 --therm 10 --no-cpu-baseline 2>/dev/null > /tmp/gt_$t.json
  python - $t <<'PY'
import json, sys
d = json.load(open(f"/tmp/gt_{sys.argv[1]}.json")); r = d["roofline"]
print(json.dumps({"generic_threads": int(sys.argv[1]), "workload": d["config"]["workload"][:2], "site_flips_per_sec": d["value"], "ms_per_step": d["ms_per_step"],
                  "stabilize_ms": r["avg_launch_ms"], "propagate_ms": r["other_kernel"]["avg_launch_ms"], "max_wrap_err": d["max_wrap_err"]}))
PY
done
