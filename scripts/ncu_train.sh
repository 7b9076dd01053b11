#!/bin/bash
# ncu evidence of the training step (run on the GPU box through gpurun): launch list of `bench.py --config train_dp` and a
# --set full capture of its kernel families; the report stays on the box, the summaries come back in gpurun_out/.
# Usage: bash scripts/ncu_train.sh <tag>
set -u
TAG=${1:-r02}
OUT=gpurun_out
REP=/tmp/ncu_rep
mkdir -p $OUT $REP
BENCH="python bench.py --config train_dp --steps 2 --warmup 3 --no-cpu-baseline"
$BENCH > $OUT/${TAG}_train_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file $OUT/${TAG}_train_launches.csv $BENCH > $OUT/${TAG}_train_ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on \
    -k regex:"wgrad_tc|wgrad_reduce|conv3x3_tc3|conv_ffma|train_moments4|train_bn_bwd|train_bn_apply4|nhwc_to_planar|train_conv_bwd_w_small|train_sum_partials" \
    -s 270 -c 90 -o $REP/${TAG}_train -f $BENCH > $OUT/${TAG}_train_ncu_full.log 2>&1
python scripts/ncu_summary.py $REP/${TAG}_train.ncu-rep > $OUT/${TAG}_train_ncu_table.md 2> $OUT/${TAG}_train_ncu_summary.err
ncu -i $REP/${TAG}_train.ncu-rep --page source --csv -k regex:wgrad_tc > $OUT/${TAG}_wgrad_source.csv 2>/dev/null
ncu -i $REP/${TAG}_train.ncu-rep --page raw --csv > $OUT/${TAG}_train_raw.csv 2>/dev/null
gzip -f $OUT/${TAG}_wgrad_source.csv $OUT/${TAG}_train_raw.csv
ls -la $OUT | grep ${TAG}_train; ls -la $OUT | grep wgrad
