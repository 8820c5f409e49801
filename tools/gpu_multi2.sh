#!/bin/bash
# N-GPU checks: NCCL-rank parity test + N-GPU bench lines.  usage: gpu_multi2.sh <N> <tag> [shape particles]...
N=${1:-2}; R=${2:-m}; shift; shift
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_$R.log 2>&1
timeout 900 python -m pytest tests -m gpu -x -q -s -k "nccl_ranks" > gpurun_out/pytest_nccl_$R.log 2>&1; echo "pytest nccl exit $?"; tail -5 gpurun_out/pytest_nccl_$R.log
while [ $# -ge 2 ]; do
  SH=$1; P=$2; shift; shift
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps ${STEPS:-2} --warmup 3 --shape $SH --particles $P > gpurun_out/bench_${N}gpu_${SH}_$R.json 2> gpurun_out/bench_${N}gpu_${SH}_$R.err
  echo "bench $N gpus $SH exit $?"; grep '^{' gpurun_out/bench_${N}gpu_${SH}_$R.json | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print({k: d[k] for k in ('value','ms_per_step','n_gpus','posterior_e_mean','check')}, d['roofline']['share_of_step'])
"; tail -4 gpurun_out/bench_${N}gpu_${SH}_$R.err
done
