#!/bin/bash
# Raw metrics + source page of the fused correlation kernel on scripts/corr_fused_timing.py (batch of 16 pairs = launch 13 on):
# bash scripts/ncu_corr_fused.sh <kernel regex> <tag>.  The report stays on the box; the CSVs come back gzipped.
set -u
RX=$1; TAG=$2
mkdir -p gpurun_out /tmp/ncu_rep
ncu --set full --clock-control none --import-source on -k regex:"$RX" -s 13 -c 1 \
    -o /tmp/ncu_rep/$TAG -f python scripts/corr_fused_timing.py > gpurun_out/${TAG}_ncu.log 2>&1
ncu -i /tmp/ncu_rep/$TAG.ncu-rep --page source --csv > gpurun_out/${TAG}_source.csv 2>/dev/null
ncu -i /tmp/ncu_rep/$TAG.ncu-rep --page raw --csv > gpurun_out/${TAG}_raw.csv 2>/dev/null
gzip -f gpurun_out/${TAG}_source.csv gpurun_out/${TAG}_raw.csv
ls -la gpurun_out | grep $TAG
