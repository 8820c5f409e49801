#!/bin/bash
# bench lines of BASELINE configs 0-2 (+ smoke, smoke diag).  usage: gpu_cfg.sh <tag> [extra args for config 2]
R=${1:-c}; shift
mkdir -p gpurun_out
timeout 600 python __graft_entry__.py smoke > gpurun_out/smoke_$R.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke_$R.log
timeout 600 python tools/diag_smoke.py > gpurun_out/diag_smoke_$R.log 2>&1; echo "diag exit $?"; tail -4 gpurun_out/diag_smoke_$R.log
for c in 0 1; do
  timeout 900 python bench.py --config $c --steps 3 --warmup 2 > gpurun_out/config${c}_$R.json 2> gpurun_out/config${c}_$R.err; echo "config $c exit $?"
  cut -c1-700 gpurun_out/config${c}_$R.json; tail -3 gpurun_out/config${c}_$R.err
done
timeout 1500 python bench.py --config 2 --steps 1 --warmup 1 "$@" > gpurun_out/config2_$R.json 2> gpurun_out/config2_$R.err; echo "config 2 exit $?"
cut -c1-900 gpurun_out/config2_$R.json; tail -3 gpurun_out/config2_$R.err
