#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_decode.py -m gpu -q -x --timeout 300 > gpurun_out/pytest_decode.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_decode.log
for calib in 8 0; do
  B200ATTN_SK2_CALIB=$calib timeout 200 python scripts/sk2_trace.py cfg3 8 1 2 > gpurun_out/sk2_trace_calib$calib.json 2> gpurun_out/sk2_trace.err; echo "trace calib=$calib rc=$?"
  cut -c1-700 gpurun_out/sk2_trace_calib$calib.json; tail -2 gpurun_out/sk2_trace.err | cut -c1-300
done
for wl in cfg3 cfg2 cfg5; do
  timeout 300 python scripts/compare_vllm_backends.py $wl 8 2> gpurun_out/compare_$wl.err | head -7 > gpurun_out/compare_$wl.json; echo "compare $wl rc=$?"
  cut -c1-250 gpurun_out/compare_$wl.json; tail -2 gpurun_out/compare_$wl.err | cut -c1-300
done
