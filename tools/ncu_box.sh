#!/bin/bash
# usage (on the GPU box, under gpurun): tools/ncu_box.sh <tag> <kernel-regex> <count> <command...>
# Runs the command once plainly, then under ncu --set full; keeps only text summaries in gpurun_out/ (a .ncu-rep of a few
# kernels is tens of MiB and gpurun_out/ is capped at 64 MiB): the raw-page summary of every captured kernel and the
# gzipped source page (SASS, per-instruction counts and stall samples).
tag=$1; re=$2; cnt=$3; shift 3
"$@" > /dev/null 2>&1 || { echo "plain run failed"; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"$re" -c "$cnt" -o /tmp/$tag -f "$@" > gpurun_out/${tag}_ncu.log 2>&1
python tools/ncu_summary.py /tmp/$tag.ncu-rep > gpurun_out/${tag}_summary.txt 2>&1
ncu -i /tmp/$tag.ncu-rep --page source --csv --print-source sass 2>/dev/null | gzip -9 > gpurun_out/${tag}_source.csv.gz
ls -la /tmp/$tag.ncu-rep gpurun_out/${tag}_*
