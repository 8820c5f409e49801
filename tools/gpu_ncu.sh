#!/bin/bash
# one `ncu --set full` capture of k_rejuvenate (launch ~60 of a run); extra bench.py arguments after the tag
mkdir -p gpurun_out
R=${1:-ncu}; shift
SHORT="python bench.py --steps 1 --warmup 3 --no-cpu-baseline $*"
timeout 300 $SHORT > gpurun_out/plain_$R.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_rejuvenate -s 60 -c 1 \
    -f -o gpurun_out/rejuv_$R $SHORT > gpurun_out/ncu_full_$R.log 2>&1
echo "full capture exit $?"; tail -2 gpurun_out/ncu_full_$R.log
