#!/bin/bash
# `ncu --set full` captures of k_rejuvenate: usage gpu_ncu2.sh <tag> <shape> <particles> [launch-skip...]
mkdir -p gpurun_out
R=$1; SH=$2; P=$3; shift; shift; shift
SHORT="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --shape $SH --particles $P"
timeout 300 $SHORT > gpurun_out/plain_$R.log 2>&1 || { echo "plain run failed"; tail -3 gpurun_out/plain_$R.log; exit 1; }
for S in "$@"; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_rejuvenate -s $S -c 1 \
      -f -o gpurun_out/rejuv_${R}_s$S $SHORT > gpurun_out/ncu_full_${R}_s$S.log 2>&1
  echo "capture -s $S exit $?"; tail -1 gpurun_out/ncu_full_${R}_s$S.log
done
ls -la gpurun_out/*.ncu-rep
