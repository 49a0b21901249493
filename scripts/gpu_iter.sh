#!/bin/bash
# Iteration pass: GPU parity tests then short bench runs of the variants given in $VARIANTS.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout 180 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/pytest_gpu.log
for v in ${VARIANTS:-3}; do
  for wl in ${WORKLOADS:-cfg3}; do
    timeout 300 python bench.py --steps 5 --warmup 3 --variant $v --workload $wl --no-e2e --no-cpu-baseline ${BENCH_ARGS} > gpurun_out/bench_${wl}_v$v.log 2>&1
    echo "bench $wl v$v rc=$?"
    python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_${wl}_v$v.log").read().strip().splitlines()[-1])
    r=d["roofline"]; print("  $wl v$v: %.0f tok/s, kernel %.1f us, %.0f GB/s (%.3f of measured, %.3f of 8TB/s), splits %d, step share %.2f" % (d["value"], r["kernel_avg_us"], r["achieved"], r["frac"], r["frac_of_8TBps"], r["num_splits"], r["kernel_share_of_step"]))
except Exception as e:
    print("  parse failed", e); print(open("gpurun_out/bench_${wl}_v$v.log").read()[-1500:])
PY
  done
done
