#!/bin/bash
# ncu launch list (gpu__time_duration only) of one short bench run.  usage: gpu_launchlist.sh <tag> <shape> <particles>
R=$1; SH=$2; P=$3
mkdir -p gpurun_out
SHORT="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --shape $SH --particles $P"
timeout 600 $SHORT > gpurun_out/plain_ll_$R.log 2>&1 || { echo plain failed; tail -3 gpurun_out/plain_ll_$R.log; exit 1; }
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_rejuvenate -c 500 --csv \
    --log-file gpurun_out/launches_$R.csv $SHORT > gpurun_out/ncu_ll_$R.log 2>&1
echo "launch list exit $?"; wc -l gpurun_out/launches_$R.csv
