#!/bin/bash
# `ncu --set full` of k_rejuvenate at PF step t = 60 of the first timed-shape run, sphere and cube (after a plain run exited 0)
R=${1:-r02z}; mkdir -p gpurun_out
SHORT="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 600 $SHORT > gpurun_out/plain_$R.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_rejuvenate -s 85 -c 1 -f -o gpurun_out/rejuv_sphere_$R $SHORT > gpurun_out/ncu_sphere_$R.log 2>&1; echo "sphere exit $?"
SHORTC="$SHORT --shape cube --particles 262144"
timeout 600 $SHORTC > gpurun_out/plainc_$R.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_rejuvenate -s 85 -c 1 -f -o gpurun_out/rejuv_cube_$R $SHORTC > gpurun_out/ncu_cube_$R.log 2>&1; echo "cube exit $?"
