#!/bin/bash
mkdir -p gpurun_out
for mode in "0 0" "1 0" "0 1"; do
  set -- $mode
  B200ATTN_SK2_DEBUG_INTERLEAVE=$1 CONTIG=$2 timeout 200 python scripts/sk2_trace.py cfg3 8 1 3 > gpurun_out/sk2_trace_i$1_c$2.json 2> gpurun_out/sk2_trace.err; echo "trace interleave=$1 contig=$2 rc=$?"
  cut -c1-900 gpurun_out/sk2_trace_i$1_c$2.json; tail -2 gpurun_out/sk2_trace.err | cut -c1-300
done
