"""Aggregate an `ncu --page source --csv --print-source cuda,sass` export per CUDA source line and
per source FUNCTION-ish region (file + line range), with the dominant stall reasons.
usage: python benchmarks/ncu_lines.py report.ncu-rep kernel_regex [top]"""
import collections
import csv
import io
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", kern],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
fname, hdr, data = "?", None, []
stalls_tot = collections.Counter()
for r in rows:
    if len(r) >= 2 and r[0] == "File Name":
        fname = r[1].split("/")[-1]
        continue
    if len(r) > 8 and r[0] == "Line No":
        hdr = r
        si, ie = hdr.index("# Samples"), hdr.index("Instructions Executed")
        st = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
        continue
    if hdr is None or len(r) < len(hdr) or r[0] == "":
        continue
    try:
        s, n = int(r[si]), int(r[ie])
    except ValueError:
        continue
    sd = {}
    for i, h in st:
        try:
            v = int(r[i])
        except ValueError:
            v = 0
        if v:
            sd[h[6:]] = v
            stalls_tot[h[6:]] += v
    data.append((s, n, fname, int(r[0]) if r[0].isdigit() else -1, r[1].strip()[:90], sd))
tot = sum(d[0] for d in data) or 1
toti = sum(d[1] for d in data) or 1
print(f"kernel {kern}: samples {tot} warp-instructions {toti}")
print("stall reasons (all samples):", ", ".join(f"{k} {100 * v / tot:.1f}%" for k, v in stalls_tot.most_common(8)))
byfile = collections.Counter()
for d in data:
    byfile[d[2]] += d[0]
print("by file:", ", ".join(f"{k} {100 * v / tot:.1f}%" for k, v in byfile.most_common()))
for d in sorted(data, key=lambda x: -x[0])[:top]:
    s = ",".join(f"{k}:{100 * v / max(d[0], 1):.0f}" for k, v in sorted(d[5].items(), key=lambda x: -x[1])[:2])
    print(f"{100 * d[0] / tot:5.1f}%s {100 * d[1] / toti:5.1f}%i {d[2]}:{d[3]:>4} [{s}] {d[4]}")
