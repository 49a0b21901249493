#!/bin/bash
# final single-GPU evidence pass on the final build
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -x -q -m gpu --timeout 900 > gpurun_out/r2_final_tests.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/r2_final_tests.log
timeout 600 python bench.py > gpurun_out/r2_final_bench.json 2> gpurun_out/r2_final_bench.err; echo "bench rc=$?"; tail -c 1200 gpurun_out/r2_final_bench.json; tail -8 gpurun_out/r2_final_bench.err
for wl in cfg3 cfg2 cfg5 cfg4; do
  timeout 300 python scripts/compare_vllm_backends.py $wl 16 > gpurun_out/r2z_cmp_$wl.jsonl 2> gpurun_out/r2z_cmp_$wl.err; echo "compare $wl rc=$?"
  grep "planned + early KV\|trtllm\|v2\|v1 (round" gpurun_out/r2z_cmp_$wl.jsonl | cut -c1-200
done
timeout 900 python scripts/splitk_sweep.py cfg5 8 > gpurun_out/sweep_cfg5.json 2> gpurun_out/sweep_cfg5.err; echo "sweep rc=$?"
grep "cooperative" gpurun_out/sweep_cfg5.json | cut -c1-200
for wl in cfg2 cfg4 cfg5; do timeout 300 python bench.py --workload $wl --no-e2e --no-cpu-baseline --no-extras > gpurun_out/r2_final_bench_$wl.json 2>/dev/null; python -c "
import json,sys
d=json.loads(open('gpurun_out/r2_final_bench_$wl.json').read().strip().splitlines()[-1]); r=d['roofline']
print('$wl', round(d['value']), 'tok/s decode us', round(r['kernel_avg_us'],2), 'frac', round(r['frac'],4), 'of8', round(r['frac_of_8TBps'],4))"; done
CALIB_STEPS=30 timeout 200 python scripts/sk_timeline.py cfg3 8 1 3 > gpurun_out/r2_final_timeline_cfg3.json 2>/dev/null
CALIB_STEPS=30 timeout 200 python scripts/sk_timeline.py cfg2 8 1 3 > gpurun_out/r2_final_timeline_cfg2.json 2>/dev/null
bash scripts/gpu_profiles_r2.sh 2>&1 | grep "rc="
