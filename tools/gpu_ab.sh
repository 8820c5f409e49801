#!/bin/bash
# lineage-controlled A/B of k_rejuvenate over the default library and every libsmcphys_v*.so variant
mkdir -p gpurun_out
for so in smc_physics_b200/libsmcphys.so smc_physics_b200/libsmcphys_v*.so; do
  [ -f "$so" ] || continue
  echo "== $so"
  SMCPHYS_SO=$PWD/$so timeout 300 python tools/bench_kernel.py "$@" 2>&1 | tail -1
done | tee gpurun_out/ab_${TAG:-x}.log
