#!/bin/bash
# Development aid: C1 throughput as a function of the number of walker groups on streams (DQMCB_STREAMS).
for g in 1 2 3 4 6; do
  DQMCB_STREAMS=$g python bench.py --workload C1 --steps 40 --warmup 5 --therm 10 --no-cpu-baseline 2>/dev/null > /tmp/sg_$g.json
  python - $g <<'PY'
import json, sys
d = json.load(open(f"/tmp/sg_{sys.argv[1]}.json"))
print(json.dumps({"streams": int(sys.argv[1]), "site_flips_per_sec": d["value"], "ms_per_step": d["ms_per_step"], "e2e": d["e2e"]["value"]}))
PY
done
