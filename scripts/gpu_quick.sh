#!/bin/bash
# quick A/B pass: comparators + timeline only (no tests). usage: gpu_quick.sh <tag> [workloads...]
mkdir -p gpurun_out
tag=${1:-q}; shift
wls=${@:-cfg3}
for wl in $wls; do
  timeout 300 python scripts/compare_vllm_backends.py $wl 16 > gpurun_out/${tag}_cmp_$wl.jsonl 2> gpurun_out/${tag}_cmp_$wl.err; echo "compare $wl rc=$?"
  cut -c1-200 gpurun_out/${tag}_cmp_$wl.jsonl | grep -v "flash-attn"; tail -1 gpurun_out/${tag}_cmp_$wl.err | cut -c1-200
  CALIB_STEPS=${CALIB_STEPS:-30} timeout 200 python scripts/sk_timeline.py $wl 8 1 3 > gpurun_out/${tag}_timeline_$wl.json 2> gpurun_out/${tag}_timeline.err; echo "timeline $wl rc=$?"
  cut -c1-640 gpurun_out/${tag}_timeline_$wl.json | head -4; tail -2 gpurun_out/${tag}_timeline.err | cut -c1-300
done
