#!/bin/bash
# usage (on the GPU box): tools/ncu_capture.sh <name> <kernel regex> <launch count> -- <command...>
# One `ncu --set full` pass over the matching kernels; the .ncu-rep is turned into the raw-page CSV (+ the per-kernel
# summary tools/ncu_summary.py prints) and deleted when it is larger than 20 MB (gpurun_out/ carries <= 64 MiB back).
name=$1; rx=$2; cnt=$3; shift 3; shift
mkdir -p gpurun_out
ncu --set full --clock-control none --import-source on -k "regex:$rx" -c "$cnt" -o "gpurun_out/$name" -f "$@" > "gpurun_out/$name.log" 2>&1
ncu -i "gpurun_out/$name.ncu-rep" --page raw --csv > "gpurun_out/$name.raw.csv" 2>> "gpurun_out/$name.log"
sz=$(stat -c %s "gpurun_out/$name.ncu-rep" 2>/dev/null || echo 0)
if [ "$sz" -gt 20000000 ]; then rm -f "gpurun_out/$name.ncu-rep"; fi
python tools/ncu_summary.py "gpurun_out/$name.raw.csv" > "gpurun_out/$name.summary.txt" 2>&1
