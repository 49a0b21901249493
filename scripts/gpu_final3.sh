#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -x -q -m gpu --timeout 900 > gpurun_out/r2_final_tests.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/r2_final_tests.log
timeout 600 python bench.py > gpurun_out/r2_final_bench.json 2> gpurun_out/r2_final_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/r2_final_bench.err
timeout 300 python bench.py --workload cfg4 --no-e2e --no-cpu-baseline --no-extras > gpurun_out/r2_final_bench_cfg4.json 2>/dev/null
python -c "
import json
for wl in ('','_cfg4'):
    d=json.loads(open('gpurun_out/r2_final_bench%s.json' % wl).read().strip().splitlines()[-1]); r=d['roofline']
    print(wl or 'cfg3', round(d['value']), 'tok/s', round(d['ms_per_step'],3), 'ms; decode us', round(r['kernel_avg_us'],2), 'frac', round(r['frac'],4), 'of8', round(r['frac_of_8TBps'],4), r['kernel'][:40], 'traffic', r['traffic'])"
timeout 300 python scripts/compare_vllm_backends.py cfg4 16 2>/dev/null > gpurun_out/r2z_cmp_cfg4.jsonl; cut -c1-180 gpurun_out/r2z_cmp_cfg4.jsonl | head -4
timeout 900 python scripts/splitk_sweep.py cfg5 8 > gpurun_out/sweep_cfg5.json 2> gpurun_out/sweep_cfg5.err; echo "sweep rc=$?"
grep "cooperative" gpurun_out/sweep_cfg5.json | cut -c1-200
