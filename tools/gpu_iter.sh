#!/bin/bash
# quick GPU iteration: selected parity tests + short bench lines (no ncu).  usage: gpu_iter.sh <tag> [pytest -k expr] [shapes...]
set -u
mkdir -p gpurun_out
R=${1:-q}; K=${2:-}; shift; shift
if [ -n "$K" ]; then
  timeout 1500 python -m pytest tests -m gpu -x -q -s -k "$K" > gpurun_out/pytest_gpu_$R.log 2>&1; echo "pytest exit $?"
  grep -E "^T4|^T1|passed|failed|Error|assert" gpurun_out/pytest_gpu_$R.log | tail -15
fi
for shape in "$@"; do
  P=1048576; [ "$shape" = cube ] && P=262144
  timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --shape $shape --particles $P > gpurun_out/bench_${shape}_$R.json 2> gpurun_out/bench_${shape}_$R.err
  echo "bench $shape exit $?"; python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/bench_${shape}_$R.json").read().strip().split("\n")[-1])
    print("$shape: value %.2f M  ms/step %.1f  rejuv avg %.3f ms  frac %.4f  shares %s  post_e %.5f" % (d["value"] / 1e6, d["ms_per_step"], d["roofline"]["avg_launch_ms"], d["roofline"]["frac"], {k: round(v, 3) for k, v in d["roofline"]["share_of_step"].items()}, d["posterior_e_mean"]))
except Exception as e:
    print("no bench line", e); print(open("gpurun_out/bench_${shape}_$R.err").read()[-800:])
PY
done
