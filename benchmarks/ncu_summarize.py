"""Condense `ncu --set full` reports of the two fast-path kernels into profiles/ncu_summary.json (the file
bench.py reads `roofline.traffic` from) and per-kernel hot-line listings.
usage: python benchmarks/ncu_summarize.py <stabilize.ncu-rep> <propagate.ncu-rep> <tag>"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "launch__waves_per_multiprocessor", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor_subpipe_dmma.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "launch__grid_size",
        "launch__block_size", "l1tex__throughput.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"]
SCALE = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0}


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    h, u, v = rows[0], rows[1], rows[2]
    d = {}
    for i, n in enumerate(h):
        if n in KEEP:
            d[n] = {"value": float(v[i].replace(",", "")), "unit": u[i]}
    d["kernel"] = v[h.index("Kernel Name")]
    b = 0.0
    for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        b += d[k]["value"] * SCALE.get(d[k]["unit"], 1.0)
    d["dram_bytes_per_launch"] = b
    return d


def main():
    stab, prop, tag = sys.argv[1], sys.argv[2], sys.argv[3]
    s = {"C1": {"stabilize": raw(stab), "propagate": raw(prop)}}
    for k, rep in (("stabilize", stab), ("propagate", prop)):
        s["C1"][k]["capture"] = ("ncu --set full --clock-control none --import-source on, 1 launch, W=1024 walkers, C1 (8x8, nt=80, "
                                 "n_stab=10), DQMCB_STREAMS=1; report %s" % os.path.basename(rep))
    with open(os.path.join(ROOT, "profiles", "ncu_summary.json"), "w") as f:
        json.dump(s, f, indent=1)
    for k, rep, rx in (("kf_stabilize", stab, "kf_stabilize"), ("kf_propagate", prop, "kf_propagate")):
        out = subprocess.run([sys.executable, os.path.join(ROOT, "benchmarks", "ncu_lines.py"), rep, rx, "60"], capture_output=True, text=True).stdout
        with open(os.path.join(ROOT, "profiles", "r01_%s_%s_hotlines.txt" % (tag, k)), "w") as f:
            f.write(out)
    print(json.dumps({k: {"ms": v["gpu__time_duration.sum"]["value"], "dram_MB": v["dram_bytes_per_launch"] / 1e6} for k, v in s["C1"].items()}))


main()
