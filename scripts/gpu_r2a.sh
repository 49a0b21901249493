#!/bin/bash
# round 2, first pass: stream-K v2 parity + A/B against v1 and the stock kernels + per-warp timeline
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_decode.py -m gpu -q -x --timeout 300 > gpurun_out/pytest_decode.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_decode.log
for wl in cfg3 cfg2 cfg5; do
  timeout 300 python scripts/compare_vllm_backends.py $wl 8 > gpurun_out/compare_$wl.json 2> gpurun_out/compare_$wl.err; echo "compare $wl rc=$?"
  cut -c1-330 gpurun_out/compare_$wl.json; tail -2 gpurun_out/compare_$wl.err | cut -c1-300
done
for wl in cfg3 cfg2; do
  timeout 200 python scripts/sk2_trace.py $wl 8 > gpurun_out/sk2_trace_$wl.json 2> gpurun_out/sk2_trace_$wl.err; echo "trace $wl rc=$?"
  cat gpurun_out/sk2_trace_$wl.json; tail -2 gpurun_out/sk2_trace_$wl.err | cut -c1-300
done
B200ATTN_PDL=0 timeout 300 python scripts/compare_vllm_backends.py cfg3 8 2>/dev/null | head -6 | cut -c1-330 | tee gpurun_out/compare_cfg3_nopdl.json
