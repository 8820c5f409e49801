#!/bin/bash
# One gpurun call: GPU parity tests, the default bench line, the ncu launch list of the short bench command and one
# `--set full` capture of the dominant kernel.  Everything lands in gpurun_out/.
set -u
mkdir -p gpurun_out
R=${1:-r01}
SHORT="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
(which julia; python -c "import pybullet" ) > gpurun_out/probe_$R.log 2>&1
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi_$R.log 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$R.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu_$R.log
tail -5 gpurun_out/pytest_gpu_$R.log
timeout 600 python __graft_entry__.py smoke > gpurun_out/smoke_$R.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke_$R.log
timeout 900 python bench.py > gpurun_out/bench_$R.json 2> gpurun_out/bench_$R.err; echo "bench exit $?"; cat gpurun_out/bench_$R.json
timeout 600 python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/bench_ref_$R.json 2>> gpurun_out/bench_$R.err; echo "ref exit $?"; cat gpurun_out/bench_ref_$R.json
timeout 600 $SHORT > gpurun_out/plain_$R.log 2>&1 &&
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -s 2900 -c 1000 --csv \
    --log-file gpurun_out/launches_$R.csv $SHORT > gpurun_out/ncu_launches_$R.log 2>&1
echo "launch list exit $?"
[ -n "${NO_FULL:-}" ] && exit 0   # the full capture then goes in its own call: tools/gpu_ncu.sh <tag>
timeout 600 $SHORT > gpurun_out/plain2_$R.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:k_rejuvenate -s 60 -c 2 \
    -f -o gpurun_out/rejuv_$R $SHORT > gpurun_out/ncu_full_$R.log 2>&1
echo "full capture exit $?"
ls -la gpurun_out
