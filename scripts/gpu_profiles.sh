#!/bin/bash
# Round-1 evidence: launch list of a short default bench run + full ncu captures of the stream-K
# decode kernel (bf16 cfg3, fp8 cfg4) and of the cache-write / reduce kernels.
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --layers 4 --no-e2e --no-cpu-baseline --no-graph"
$CMD > gpurun_out/r1_plain_cfg3.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r1_launches_cfg3.csv $CMD > gpurun_out/r1_ncu_launch.log 2>&1
echo "launchlist rc=$?"
$CMD > gpurun_out/r1_plain_cfg3b.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"paged_decode_streamk|streamk_reduce|cache_write" -s 15 -c 3 -o gpurun_out/r1_prof_cfg3 -f $CMD > gpurun_out/r1_ncu_full_cfg3.log 2>&1
echo "full cfg3 rc=$?"
CMD4="$CMD --workload cfg4"
$CMD4 > gpurun_out/r1_plain_cfg4.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"paged_decode_streamk" -s 5 -c 1 -o gpurun_out/r1_prof_cfg4 -f $CMD4 > gpurun_out/r1_ncu_full_cfg4.log 2>&1
echo "full cfg4 rc=$?"
