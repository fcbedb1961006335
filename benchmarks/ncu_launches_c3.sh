#!/bin/bash
# Run on the GPU box: ncu launch list (gpu__time_duration.sum) of a short C3 bench, reduced to a per-kernel summary.
mkdir -p gpurun_out/r2n
python bench.py --workload C3 --steps 1 --warmup 1 --therm 0 --no-cpu-baseline > gpurun_out/r2n/c3_plain.json 2>/dev/null || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/r2n/launches_c3.csv \
    python bench.py --workload C3 --steps 1 --warmup 1 --therm 0 --no-cpu-baseline > gpurun_out/r2n/c3_ncu.log 2>&1
python - <<'PY'
import csv, collections, json
rows = [r for r in csv.reader(open("gpurun_out/r2n/launches_c3.csv")) if len(r) > 10]
h = rows[0]; ik = h.index("Kernel Name"); iv = h.index("Metric Value"); iu = h.index("Metric Unit"); ig = h.index("Grid Size"); ism = h.index("Section Name") if "Section Name" in h else None
agg = collections.OrderedDict()
for r in rows[1:]:
    try:
        v = float(r[iv].replace(",", ""))
    except ValueError:
        continue
    u = r[iu]
    us = v / 1e3 if u in ("ns", "nsecond") else (v if u in ("us", "usecond") else v * 1e3 if u in ("ms", "msecond") else v)
    k = r[ik].split("(")[0]
    a = agg.setdefault(k, [0, 0.0])
    a[0] += 1; a[1] += us
tot = sum(a[1] for a in agg.values())
out = {"command": "bench.py --workload C3 --steps 1 --warmup 1 --therm 0 --no-cpu-baseline under ncu --metrics gpu__time_duration.sum --clock-control none (per-launch times are cold-cache and serialised: compare SHARES, not absolutes)",
       "total_us": tot, "kernels": {k: {"launches": a[0], "total_us": a[1], "mean_us": a[1] / a[0], "share": a[1] / tot} for k, a in sorted(agg.items(), key=lambda x: -x[1][1])}}
json.dump(out, open("gpurun_out/r2n/launch_summary_c3.json", "w"), indent=1)
for k, v in list(out["kernels"].items())[:8]:
    print(f"{k[:60]:60s} {v['launches']:5d} launches  {v['mean_us']:10.1f} us  share {v['share']:.3f}")
PY
rm -f gpurun_out/r2n/launches_c3.csv.tmp
