#!/bin/bash
# Round-2 measurement suite on one B200: parity tests, smoke, bench lines (default + reference arm + cube + configs 0-2),
# ncu launch list + full captures (only after the same command exited 0 without ncu), T5 tables.  Everything -> gpurun_out/.
set -u
R=${1:-r02}
mkdir -p gpurun_out
(which julia; python -c "import pybullet") > gpurun_out/probe_$R.log 2>&1
timeout 2400 python -m pytest tests -m gpu -q -s > gpurun_out/pytest_gpu_$R.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/pytest_gpu_$R.log
timeout 600 python __graft_entry__.py smoke > gpurun_out/smoke_$R.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/smoke_$R.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_$R.json 2> gpurun_out/bench_$R.err; echo "bench exit $?"; cut -c1-300 gpurun_out/bench_$R.json
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference_$R.json 2>> gpurun_out/bench_$R.err; echo "ref exit $?"
timeout 900 python bench.py --shape cube --particles 262144 --steps 2 --warmup 3 > gpurun_out/bench_cube_$R.json 2> gpurun_out/bench_cube_$R.err; echo "cube exit $?"; cut -c1-200 gpurun_out/bench_cube_$R.json
for c in 0 1; do
  timeout 900 python bench.py --config $c --steps 3 --warmup 2 > gpurun_out/config${c}_$R.json 2> gpurun_out/config${c}_$R.err; echo "config $c exit $?"
done
timeout 1800 python bench.py --config 2 --steps 1 --warmup 1 > gpurun_out/config2_$R.json 2> gpurun_out/config2_$R.err; echo "config 2 exit $?"; cut -c1-200 gpurun_out/config2_$R.json
timeout 900 python tools/t5_exp1.py --seeds 16 --out gpurun_out/t5_systematic_$R.json > gpurun_out/t5_systematic_$R.log 2>&1; echo "t5 exit $?"
timeout 900 python tools/t5_exp1.py --seeds 16 --resampler 1 --out gpurun_out/t5_multinomial_$R.json > gpurun_out/t5_multinomial_$R.log 2>&1; echo "t5 multinomial exit $?"; tail -6 gpurun_out/t5_multinomial_$R.log
SHORT="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 600 $SHORT > gpurun_out/plain_$R.log 2>&1 &&
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 7000 --csv \
    --log-file gpurun_out/launches_$R.csv $SHORT > gpurun_out/ncu_launches_$R.log 2>&1
echo "launch list exit $?"
timeout 600 $SHORT > gpurun_out/plain2_$R.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:k_rejuvenate -s 85 -c 1 \
    -f -o gpurun_out/rejuv_sphere_$R $SHORT > gpurun_out/ncu_full_sphere_$R.log 2>&1
echo "sphere capture exit $?"
SHORTC="$SHORT --shape cube --particles 262144"
timeout 600 $SHORTC > gpurun_out/plain3_$R.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:k_rejuvenate -s 85 -c 1 \
    -f -o gpurun_out/rejuv_cube_$R $SHORTC > gpurun_out/ncu_full_cube_$R.log 2>&1
echo "cube capture exit $?"
ls gpurun_out | wc -l
