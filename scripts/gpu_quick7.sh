#!/bin/bash
for i in 1 2; do
timeout 300 python bench.py --no-e2e --no-cpu-baseline --no-extras 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
r = d['roofline']
print('bench: value', round(d['value']), 'ms/step', round(d['ms_per_step'], 3), 'decode us', round(r['kernel_avg_us'], 2), 'frac', round(r['frac'], 4), 'of 8TB/s', round(r['frac_of_8TBps'], 4), 'verify ok', d['verify']['bit_exact'], d['verify']['fast_path_ok'])"
done
CALIB_STEPS=30 timeout 200 python scripts/sk_timeline.py cfg3 8 1 2 2>/dev/null | grep -v "sm_end_corr" | cut -c1-420
timeout 300 python scripts/compare_vllm_backends.py cfg3 16 2>/dev/null | grep "planned + early KV\|trtllm" | cut -c1-200
