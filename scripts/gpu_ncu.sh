#!/bin/bash
# ncu: launch list of a short bench run + full capture of the decode kernels (both variants).
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --layers 4 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain_tma.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_tma.csv $CMD > gpurun_out/ncu_launch_tma.log 2>&1
echo "launchlist rc=$?"
$CMD > gpurun_out/plain_tma2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:paged_decode -s 12 -c 2 -o gpurun_out/prof_tma -f $CMD > gpurun_out/ncu_full_tma.log 2>&1
echo "full tma rc=$?"
CMD2="$CMD --variant 2"
$CMD2 > gpurun_out/plain_cp.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:paged_decode -s 12 -c 2 -o gpurun_out/prof_cp -f $CMD2 > gpurun_out/ncu_full_cp.log 2>&1
echo "full cp rc=$?"
ls -la gpurun_out
