#!/bin/bash
# quick GPU iteration: parity tests + short bench (no ncu)
set -u
mkdir -p gpurun_out
R=${1:-q}
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$R.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/pytest_gpu_$R.log
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$R.json 2> gpurun_out/bench_$R.err; echo "bench exit $?"; cat gpurun_out/bench_$R.json; tail -5 gpurun_out/bench_$R.err
