#!/bin/bash
# iteration: GPU tests, stream-K trace A/B, short bench
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x --timeout 180 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/pytest_gpu.log
for wl in ${WORKLOADS:-cfg3 cfg5}; do
  timeout 300 python scripts/sk_trace.py $wl 8 > gpurun_out/sk_trace_$wl.log 2>&1; echo "trace $wl rc=$?"
  cat gpurun_out/sk_trace_$wl.log | tail -8
done
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_iter.json 2> gpurun_out/bench_iter.err; echo "bench rc=$?"
cat gpurun_out/bench_iter.json
