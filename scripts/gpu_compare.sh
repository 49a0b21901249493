#!/bin/bash
mkdir -p gpurun_out
for wl in cfg3 cfg5 cfg2 cfg4; do
  timeout 300 python scripts/compare_vllm_backends.py $wl 8 > gpurun_out/compare_$wl.json 2> gpurun_out/compare_$wl.err; echo "compare $wl rc=$?"
  cat gpurun_out/compare_$wl.json | cut -c1-420
done
timeout 400 python scripts/prefill_bench.py > gpurun_out/prefill_bench.json 2> gpurun_out/prefill_bench.err; echo "prefill rc=$?"
cat gpurun_out/prefill_bench.json | cut -c1-300; tail -3 gpurun_out/prefill_bench.err | cut -c1-300
