#!/bin/bash
# Raw metrics + source page of one kernel family under any command:
#   bash scripts/ncu_kernel.sh <kernel regex> <tag> <launch skip> <command...>
# The report stays on the box; the CSVs come back gzipped in gpurun_out/.
set -u
RX=$1; TAG=$2; SKIP=$3; shift 3
mkdir -p gpurun_out /tmp/ncu_rep
ncu --set full --clock-control none --import-source on -k regex:"$RX" -s $SKIP -c 1 -o /tmp/ncu_rep/$TAG -f "$@" > gpurun_out/${TAG}_ncu.log 2>&1
ncu -i /tmp/ncu_rep/$TAG.ncu-rep --page source --csv > gpurun_out/${TAG}_source.csv 2>/dev/null
ncu -i /tmp/ncu_rep/$TAG.ncu-rep --page raw --csv > gpurun_out/${TAG}_raw.csv 2>/dev/null
gzip -f gpurun_out/${TAG}_source.csv gpurun_out/${TAG}_raw.csv
ls -la gpurun_out | grep $TAG
