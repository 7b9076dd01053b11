#!/bin/bash
# Source page (per-instruction stall samples) + raw metrics of ONE kernel family: bash scripts/ncu_one.sh <regex> <tag> [launch-skip]
# on scripts/post_chain_ms.py (both sides on one stream).  The report stays on the box; the CSVs come back gzipped.
set -u
RX=$1; TAG=$2; SKIP=${3:-0}
mkdir -p gpurun_out /tmp/ncu_rep
STEREO_POST_SERIAL=1 ncu --set full --clock-control none --import-source on -k regex:"$RX" -s $SKIP -c 2 \
    -o /tmp/ncu_rep/$TAG -f python scripts/post_chain_ms.py 228 > gpurun_out/${TAG}_ncu.log 2>&1
ncu -i /tmp/ncu_rep/$TAG.ncu-rep --page source --csv > gpurun_out/${TAG}_source.csv 2>/dev/null
ncu -i /tmp/ncu_rep/$TAG.ncu-rep --page raw --csv > gpurun_out/${TAG}_raw.csv 2>/dev/null
gzip -f gpurun_out/${TAG}_source.csv gpurun_out/${TAG}_raw.csv
ls -la gpurun_out | grep $TAG
