"""Condense the text outputs of benchmarks/ncu_box.sh (raw metric CSVs) and an ncu launch list into profiles/:
    python benchmarks/ncu_condense.py <dir with *.raw.csv> <launches.csv> <round tag>
writes profiles/<tag>_ncu_summary.json (per capture: the kept counters, DRAM bytes per launch) and
profiles/<tag>_launch_summary_c1.json (per kernel: launches, total and mean device time, share of the step); also refreshes
profiles/ncu_summary.json["C1"] (the file bench.py reads roofline.traffic from)."""
import collections
import csv
import glob
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "launch__waves_per_multiprocessor", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor_subpipe_dmma.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "launch__grid_size", "launch__block_size", "l1tex__throughput.avg.pct_of_peak_sustained_active", "lts__t_sector_hit_rate.pct",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "smsp__pcsamp_warps_issue_stalled_barrier", "smsp__pcsamp_warps_issue_stalled_long_scoreboard",
        "smsp__pcsamp_warps_issue_stalled_short_scoreboard", "smsp__pcsamp_warps_issue_stalled_wait", "smsp__pcsamp_warps_issue_stalled_selected",
        "smsp__pcsamp_warps_issue_stalled_math_pipe_throttle", "smsp__pcsamp_warps_issue_stalled_mio_throttle",
        "smsp__pcsamp_warps_issue_stalled_no_instructions", "smsp__pcsamp_warps_issue_stalled_not_selected"]
SCALE = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0}


def one(path):
    rows = list(csv.reader(open(path)))
    h, u, v = rows[0], rows[1], rows[2]
    d = {}
    for i, n in enumerate(h):
        if n in KEEP:
            try:
                d[n] = {"value": float(v[i].replace(",", "")), "unit": u[i]}
            except ValueError:
                pass
    d["kernel"] = v[h.index("Kernel Name")]
    b = 0.0
    for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        b += d[k]["value"] * SCALE.get(d[k]["unit"], 1.0)
    d["dram_bytes_per_launch"] = b
    st = {k.split("stalled_")[1]: x["value"] for k, x in d.items() if isinstance(x, dict) and "pcsamp" in k}
    tot = sum(st.values()) or 1.0
    d["stall_share_of_listed_pct"] = {k: round(100 * x / tot, 1) for k, x in sorted(st.items(), key=lambda t: -t[1])}
    return d


def main():
    src, launches, tag = sys.argv[1], sys.argv[2], sys.argv[3]
    out = {}
    for f in sorted(glob.glob(os.path.join(src, "*.raw.csv"))):
        out[os.path.basename(f)[:-8]] = one(f)
    json.dump(out, open(os.path.join(ROOT, "profiles", f"{tag}_ncu_summary.json"), "w"), indent=1)
    legacy = os.path.join(ROOT, "profiles", "ncu_summary.json")
    try:
        old = json.load(open(legacy))
    except Exception:
        old = {}
    if "c1_stab2" in out and "c1_prop" in out:
        old["C1"] = {"stabilize": dict(out["c1_stab2"], capture=f"{tag}: ncu --set full, 1 launch, W=1024, 8x8, n_stab=10, DQMCB_STREAMS=1 (kf2_stabilize)"),
                     "propagate": dict(out["c1_prop"], capture=f"{tag}: ncu --set full, 1 launch, W=1024, 8x8, n_stab=10, DQMCB_STREAMS=1")}
    for key, a, b in (("C3", "c3_stab", "c3_prop"), ("C2", "c2_stab", "c2_prop")):
        if a in out and b in out:
            old[key] = {"stabilize": out[a], "propagate": out[b]}
    json.dump(old, open(legacy, "w"), indent=1)
    # launch list
    agg = collections.defaultdict(lambda: [0, 0.0])
    rows = list(csv.reader(l for l in open(launches) if not l.startswith("==")))
    h = rows[0]
    ki, mi, vi, ui = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("Metric Unit")
    for r in rows[1:]:
        if len(r) <= vi or r[mi] != "gpu__time_duration.sum":
            continue
        t = float(r[vi].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3}.get(r[ui], 1.0)
        name = r[ki].split("(")[0].replace("void ", "").replace("dq::", "")
        agg[name][0] += 1
        agg[name][1] += t
    tot = sum(v[1] for v in agg.values()) or 1.0
    summ = {k: {"launches": v[0], "total_us": round(v[1], 1), "mean_us": round(v[1] / v[0], 2), "share_pct": round(100 * v[1] / tot, 2)}
            for k, v in sorted(agg.items(), key=lambda t: -t[1][1])}
    json.dump({"command": "python bench.py --steps 2 --warmup 3 --therm 2 --no-cpu-baseline (under ncu --metrics gpu__time_duration.sum "
                          "--clock-control none -c 600: cold-cache, serialised launches: compare SHARES, not absolutes)", "kernels": summ},
              open(os.path.join(ROOT, "profiles", f"{tag}_launch_summary_c1.json"), "w"), indent=1)
    print(json.dumps(summ, indent=1)[:1500])


main()
