#!/bin/bash
# ncu full capture of one decode variant: VARIANT=3 bash scripts/gpu_ncu1.sh
mkdir -p gpurun_out
V=${VARIANT:-3}
CMD="python bench.py --steps 2 --warmup 3 --layers 4 --no-e2e --no-cpu-baseline --variant $V ${BENCH_ARGS}"
$CMD > gpurun_out/plain_v$V.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:paged_decode -s 12 -c 1 -o gpurun_out/prof_v$V -f $CMD > gpurun_out/ncu_full_v$V.log 2>&1
echo "full v$V rc=$?"
