#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_decode.py tests/test_gpu_backend.py tests/test_gpu_e2e_vllm.py -x -q -m gpu --timeout 900 2>&1 | tail -4
for wl in cfg3 cfg2 cfg5; do timeout 300 python scripts/compare_vllm_backends.py $wl 16 2>/dev/null | grep "planned + early KV\|trtllm\|cache write + decode per" | cut -c1-200; done
timeout 300 python bench.py --no-e2e --no-cpu-baseline --no-extras 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
r = d['roofline']
print('bench: value', round(d['value']), 'ms/step', round(d['ms_per_step'], 3), 'decode us', round(r['kernel_avg_us'], 2), 'frac', round(r['frac'], 4), 'of 8TB/s', round(r['frac_of_8TBps'], 4), 'traffic', r['traffic'], 'verify ok', d['verify']['bit_exact'], d['verify']['fast_path_ok'])"
CALIB_STEPS=30 timeout 200 python scripts/sk_timeline.py cfg3 8 1 2 2>/dev/null | grep -v "sm_end_corr" | cut -c1-420
