#!/bin/bash
# bench the default library and every smc_physics_b200/libsmcphys_v*.so tuning variant (short run, no ncu)
mkdir -p gpurun_out
R=${1:-var}
if [ -n "${RUN_TESTS:-}" ]; then
  timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$R.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu_$R.log
fi
for so in smc_physics_b200/libsmcphys.so smc_physics_b200/libsmcphys_v*.so; do
  [ -f "$so" ] || continue
  echo "== $so"
  SMCPHYS_SO=$PWD/$so timeout 150 python bench.py --steps 1 --warmup 3 --no-cpu-baseline ${BENCH_ARGS:-} 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print(json.dumps({k: d[k] for k in ('value','ms_per_step','posterior_e_mean')}), d['roofline']['avg_launch_ms'], d['roofline']['share_of_step'])
    else: print(l, end='')
" | tee -a gpurun_out/variants_$R.log
done
if [ -n "${NCU_SO:-}" ]; then
  export SMCPHYS_SO=$PWD/$NCU_SO
  SHORT="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
  timeout 600 $SHORT > gpurun_out/plain_$R.log 2>&1 &&
  timeout 1500 ncu --set full --clock-control none --import-source on -k regex:k_rejuvenate -s 60 -c 1 \
      -f -o gpurun_out/rejuv_$R $SHORT > gpurun_out/ncu_full_$R.log 2>&1
  echo "full capture exit $?"
fi
