#!/bin/bash
# Run on the GPU box: capture one launch of a kernel with ncu --set full, reduce the report to text (raw metrics CSV of the
# kept counters + hot source lines) under gpurun_out/<dir>/ and delete the .ncu-rep (gpurun brings back at most 64 MiB).
# usage: ncu_box.sh <outdir> <tag> <kernel-regex> <skip> <count> -- <command...>
out=$1; tag=$2; kern=$3; skip=$4; cnt=$5; shift 6
mkdir -p "$out"
ncu --set full --clock-control none --import-source on -k "regex:$kern" -s "$skip" -c "$cnt" -o "$out/$tag" "$@" > "$out/$tag.ncu.log" 2>&1
ncu -i "$out/$tag.ncu-rep" --page raw --csv > "$out/$tag.raw.csv" 2>/dev/null
python benchmarks/ncu_lines.py "$out/$tag.ncu-rep" "$kern" 40 > "$out/$tag.hotlines.txt" 2>&1
rm -f "$out/$tag.ncu-rep"
