#!/bin/bash
# round-2 GPU passes. usage: gpu_r2.sh <mode>
#   decode  : decode parity tests (fail fast) + comparators + timeline
#   full    : whole GPU suite + bench + comparators
mkdir -p gpurun_out
mode=${1:-decode}
tag=${2:-r2}
if [ "$mode" = "decode" ]; then
  timeout 900 python -m pytest tests/test_gpu_decode.py -x -q -m gpu --timeout 120 > gpurun_out/${tag}_tests.log 2>&1; echo "tests rc=$?"
  tail -15 gpurun_out/${tag}_tests.log
else
  timeout 1200 python -m pytest tests -x -q -m gpu --timeout 300 > gpurun_out/${tag}_tests.log 2>&1; echo "tests rc=$?"
  tail -8 gpurun_out/${tag}_tests.log
  timeout 600 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
  tail -c 4500 gpurun_out/${tag}_bench.json; tail -12 gpurun_out/${tag}_bench.err
fi
for wl in cfg3 cfg2 cfg5; do
  timeout 300 python scripts/compare_vllm_backends.py $wl 16 > gpurun_out/${tag}_cmp_$wl.jsonl 2> gpurun_out/${tag}_cmp_$wl.err; echo "compare $wl rc=$?"
  cut -c1-230 gpurun_out/${tag}_cmp_$wl.jsonl | grep -v "flash-attn v[34]"; tail -2 gpurun_out/${tag}_cmp_$wl.err | cut -c1-300
done
for wl in cfg3 cfg2; do
  timeout 200 python scripts/sk_timeline.py $wl 8 1 2 > gpurun_out/${tag}_timeline_$wl.json 2> gpurun_out/${tag}_timeline.err; echo "timeline $wl rc=$?"
  cut -c1-700 gpurun_out/${tag}_timeline_$wl.json | head -3; tail -2 gpurun_out/${tag}_timeline.err | cut -c1-300
done
