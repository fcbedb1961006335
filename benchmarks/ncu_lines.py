"""Aggregate an `ncu --page source --csv --print-source cuda,sass` export per CUDA source line.
usage: python benchmarks/ncu_lines.py report.ncu-rep kernel_regex [top]"""
import csv, subprocess, sys, io
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", kern],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
fname = "?"
data = []
hdr = None
for r in rows:
    if len(r) >= 2 and r[0] == "File Name":
        fname = r[1].split("/")[-1]
        continue
    if len(r) > 8 and r[0] == "Line No":
        hdr = r
        continue
    if hdr is None or len(r) < len(hdr) or r[0] == "":
        continue
    try:
        si = hdr.index("# Samples"); ie = hdr.index("Instructions Executed")
        data.append((int(r[si]), int(r[ie]), fname, r[0], r[1].strip()[:100]))
    except Exception:
        pass
tot = sum(d[0] for d in data) or 1
toti = sum(d[1] for d in data) or 1
print(f"kernel {kern}: samples {tot} warp-instructions {toti}")
for d in sorted(data, reverse=True)[:top]:
    print(f"{100*d[0]/tot:5.1f}%s {100*d[1]/toti:5.1f}%i {d[2]}:{d[3]:>4} {d[4]}")
