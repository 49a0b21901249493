#!/bin/bash
# the committed evidence of the final build: bench (default + cfg2 / cfg5), comparators, split-K sweep, timelines
mkdir -p gpurun_out
timeout 600 python bench.py > gpurun_out/r2_final_bench.json 2> gpurun_out/r2_final_bench.err; echo "bench rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/r2_final_bench.json').read().strip().splitlines()[-1]); r=d['roofline']
print('cfg3', round(d['value']), 'tok/s', round(d['ms_per_step'],3), 'ms; decode us', round(r['kernel_avg_us'],2), 'frac', round(r['frac'],4), 'of8', round(r['frac_of_8TBps'],4), 'traffic', r['traffic'], 'e2e', round(d['e2e']['value']), 'cpu', round(d['cpu_baseline']['value'],2), 'ragged us', round(d['extras']['ragged']['decode_plus_reduce_us'],2), 'verify', d['verify']['bit_exact'])"
for wl in cfg3 cfg2 cfg5; do
  timeout 300 python scripts/compare_vllm_backends.py $wl 16 > gpurun_out/r2z_cmp_$wl.jsonl 2> gpurun_out/r2z_cmp_$wl.err
  grep "planned + early KV\|trtllm" gpurun_out/r2z_cmp_$wl.jsonl | cut -c1-200
done
for wl in cfg2 cfg5; do timeout 300 python bench.py --workload $wl --no-e2e --no-cpu-baseline --no-extras > gpurun_out/r2_final_bench_$wl.json 2>/dev/null; python -c "
import json
d=json.loads(open('gpurun_out/r2_final_bench_$wl.json').read().strip().splitlines()[-1]); r=d['roofline']
print('$wl', round(d['value']), 'tok/s decode us', round(r['kernel_avg_us'],2), 'frac', round(r['frac'],4), 'of8', round(r['frac_of_8TBps'],4))"; done
CALIB_STEPS=30 timeout 200 python scripts/sk_timeline.py cfg3 8 1 3 > gpurun_out/r2_final_timeline_cfg3.json 2>/dev/null
CALIB_STEPS=60 timeout 200 python scripts/sk_timeline.py cfg2 8 1 3 > gpurun_out/r2_final_timeline_cfg2.json 2>/dev/null
timeout 900 python scripts/splitk_sweep.py cfg5 8 > gpurun_out/sweep_cfg5.json 2> gpurun_out/sweep_cfg5.err; echo "sweep rc=$?"
grep "cooperative" gpurun_out/sweep_cfg5.json | cut -c1-200
