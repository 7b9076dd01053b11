#!/bin/bash
# ncu captures of one round (run on the GPU box through gpurun; every command has exited 0 without ncu before):
#   launch list of the bench command, --set full captures of every kernel family, summaries as CSV/markdown.
# Usage: bash scripts/ncu_capture.sh <tag>      (writes gpurun_out/<tag>_*)
set -u
TAG=${1:-r02}
OUT=gpurun_out
REP=/tmp/ncu_rep        # the reports stay on the box (tens of MB each): only their CSV / markdown summaries come back
mkdir -p $OUT $REP
BENCH="python bench.py --steps 2 --warmup 3 --pairs 4 --e2e-pairs 4 --no-cpu-baseline"
$BENCH > $OUT/${TAG}_bench_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $OUT/${TAG}_launches.csv $BENCH > $OUT/${TAG}_ncu_launch.log 2>&1
python scripts/ncu_targets.py > $OUT/${TAG}_targets_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on \
    -k regex:"cbca_fused|sgm_ring|sgm_prep|wta_vec|scale_kernel|volume_moments|mean2d|median2d" -c 40 \
    -o $REP/${TAG}_post -f python scripts/ncu_targets.py > $OUT/${TAG}_ncu_post.log 2>&1
ncu --set full --clock-control none --import-source on \
    -k regex:"corr_sweep|corr_both|conv3x3_tc3|conv1_tc|conv_tcr|lr_check" -c 40 \
    -o $REP/${TAG}_trunk -f python scripts/ncu_targets.py > $OUT/${TAG}_ncu_trunk.log 2>&1
python scripts/ncu_summary.py --json $OUT/${TAG}_ncu_metrics.json $REP/${TAG}_post.ncu-rep $REP/${TAG}_trunk.ncu-rep > $OUT/${TAG}_ncu_table.md 2> $OUT/${TAG}_ncu_summary.err
ncu -i $REP/${TAG}_post.ncu-rep --page source --csv -k regex:cbca_fused > $OUT/${TAG}_cbca_source.csv 2>/dev/null
ncu -i $REP/${TAG}_post.ncu-rep --page source --csv -k regex:sgm_ring_h > $OUT/${TAG}_sgm_h_source.csv 2>/dev/null
ncu -i $REP/${TAG}_trunk.ncu-rep --page source --csv -k regex:"corr_sweep|corr_both" > $OUT/${TAG}_corr_source.csv 2>/dev/null
ncu -i $REP/${TAG}_post.ncu-rep --page raw --csv > $OUT/${TAG}_post_raw.csv 2>/dev/null
ncu -i $REP/${TAG}_trunk.ncu-rep --page raw --csv > $OUT/${TAG}_trunk_raw.csv 2>/dev/null
gzip -f $OUT/${TAG}_*_source.csv $OUT/${TAG}_*_raw.csv
ls -la $OUT $REP | tail -30; du -sh $OUT
