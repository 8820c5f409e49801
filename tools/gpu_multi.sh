#!/bin/bash
# N-GPU bench of the sharded filter (torchrun, NCCL) + a small cube run on one GPU
N=${1:-2}; R=${2:-m}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus $N --steps 2 --warmup 3 > gpurun_out/bench_${N}gpu_$R.json 2> gpurun_out/bench_${N}gpu_$R.err
echo "bench $N gpus exit $?"; grep '^{' gpurun_out/bench_${N}gpu_$R.json | cut -c1-900; tail -3 gpurun_out/bench_${N}gpu_$R.err
[ -n "${SKIP_CUBE:-}" ] || timeout 600 python bench.py --shape cube --particles 131072 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/bench_cube_$R.json 2> gpurun_out/bench_cube_$R.err
echo "cube exit $?"; cut -c1-1600 gpurun_out/bench_cube_$R.json; tail -3 gpurun_out/bench_cube_$R.err
