#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/e2e_vllm.py --omni --backend CUSTOM --kv-cache-dtype fp8 --kv-scales per-head --out gpurun_out/omni_perhead.json > gpurun_out/omni_perhead.log 2>&1; echo "per-head rc=$?"
grep -v "warn\|Warn\|WARNING\|INFO" gpurun_out/omni_perhead.log | tail -12 | cut -c1-900
timeout 600 python scripts/e2e_vllm.py --omni --backend CUSTOM --out gpurun_out/omni_bf16.json > gpurun_out/omni_bf16.log 2>&1; echo "bf16 rc=$?"
python - <<'PY'
import json
try:
    a=json.load(open('gpurun_out/omni_perhead.json')); b=json.load(open('gpurun_out/omni_bf16.json'))
    n=same=0; worst=0
    for ra,rb in zip(a['results'],b['results']):
        for i,(ta,tb) in enumerate(zip(ra['tokens'],rb['tokens'])):
            n+=1
            if ta!=tb: break
            same+=1; worst=max(worst,abs(ra['logprobs'][i]-rb['logprobs'][i]))
    print("per-head fp8 vs bf16 cache:", same, n, worst, a.get('quant_routing'), a.get('impl_scale_numel'), a.get('kv_scales'))
except Exception as e: print("compare failed", e)
PY
