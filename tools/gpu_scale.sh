#!/bin/bash
# sphere bench line on N GPUs (torchrun, NCCL + peer stores inside libsmcphys) + the NCCL-rank parity test.  usage: gpu_scale.sh <N> <tag>
N=$1; R=$2
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "nccl_ranks" > gpurun_out/pytest_nccl_${N}gpu_$R.log 2>&1; echo "pytest nccl exit $?"; tail -1 gpurun_out/pytest_nccl_${N}gpu_$R.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/scale_${N}gpu_$R.json 2> gpurun_out/scale_${N}gpu_$R.err
echo "bench $N exit $?"; grep '^{' gpurun_out/scale_${N}gpu_$R.json | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print({k: d[k] for k in ('value','ms_per_step','n_gpus','check')})"
