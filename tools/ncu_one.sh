#!/bin/bash
# Usage: tools/ncu_one.sh NAME KERNEL_REGEX SKIP -- command...
# One `ncu --set full` capture of the SKIP+1-th launch matching KERNEL_REGEX; keeps the text pages
# (details + per-line source counters) under gpurun_out/ and drops the large .ncu-rep.
name=$1; regex=$2; skip=$3; shift 4
rep=/tmp/$name.ncu-rep
ncu --set full --clock-control none --import-source on -k regex:$regex -s $skip -c 1 -o /tmp/$name -f "$@" > /dev/null 2>&1
ncu -i $rep --page details > gpurun_out/${name}_ncu_full_details.txt 2>&1
ncu -i $rep --page source --csv --print-source sass > gpurun_out/${name}_ncu_source_sass.csv 2>/dev/null
ls -la gpurun_out/${name}_*
