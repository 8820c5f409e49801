#!/bin/bash
# round 2, call A: parity tests (incl. T4 vs float64, T5), T5 table, smoke
set -u
mkdir -p gpurun_out
(which julia; python -c "import pybullet") > gpurun_out/probe_r02.log 2>&1
timeout 2400 python -m pytest tests -m gpu -q -s > gpurun_out/pytest_gpu_r02a.log 2>&1; echo "pytest exit $?"
grep -E "^T4|^T1|r = |passed|failed|Error|assert" gpurun_out/pytest_gpu_r02a.log | tail -40
timeout 600 python __graft_entry__.py smoke > gpurun_out/smoke_r02a.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke_r02a.log
