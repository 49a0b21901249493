#!/bin/bash
# Round-2 evidence on the final build: ncu launch list of a short default bench run, full captures of the cooperative
# stream-K decode kernel at cfg3 / cfg2 (bf16) and cfg4 (fp8), of the cache write and of the plan kernel.
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 3 --layers 4 --no-e2e --no-cpu-baseline --no-extras --no-verify --no-graph"
$B > gpurun_out/r2_plain_cfg3.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_cfg3.csv $B > gpurun_out/r2_ncu_launch.log 2>&1
echo "launchlist rc=$?"
for wl in cfg3 cfg2 cfg4; do
  $B --workload $wl > gpurun_out/r2_plain_$wl.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:"paged_decode_sk3" -s 24 -c 1 -o gpurun_out/r2_prof_$wl -f $B --workload $wl > gpurun_out/r2_ncu_full_$wl.log 2>&1
  echo "full $wl rc=$?"
done
ncu --set full --clock-control none -k regex:"cache_write|sk3_plan" -s 70 -c 2 -o gpurun_out/r2_prof_small -f $B > gpurun_out/r2_ncu_small.log 2>&1
echo "small rc=$?"
ls -la gpurun_out | grep r2_ | head -20
