#!/bin/bash
# end-of-round verification on one GPU: the exact commands the driver runs + a cube line
R=${1:-v}
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_$R.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/pytest_gpu_$R.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_$R.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/smoke_$R.log
timeout 900 python bench.py > gpurun_out/bench_$R.json 2> gpurun_out/bench_$R.err; echo "bench exit $?"; cut -c1-260 gpurun_out/bench_$R.json
timeout 900 python bench.py --shape cube --particles 262144 --steps 2 --warmup 3 > gpurun_out/bench_cube_$R.json 2> gpurun_out/bench_cube_$R.err; echo "cube exit $?"; cut -c1-200 gpurun_out/bench_cube_$R.json
