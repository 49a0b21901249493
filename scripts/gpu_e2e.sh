#!/bin/bash
# e2e smoke inside vLLM: CUSTOM vs FLASH_ATTN on the same random-init model (eager, CUDA graphs, fp8 KV)
mkdir -p gpurun_out
run() { # name, args...
  local name=$1; shift
  timeout 600 python scripts/e2e_vllm.py "$@" --out gpurun_out/e2e_$name.json > gpurun_out/e2e_$name.log 2>&1; echo "e2e $name rc=$?"
}
run custom_eager --backend CUSTOM
run fa_eager --backend FLASH_ATTN
python scripts/e2e_compare.py gpurun_out/e2e_custom_eager.json gpurun_out/e2e_fa_eager.json | tee gpurun_out/e2e_cmp_eager.json; echo "cmp rc=$?"
run custom_graphs --backend CUSTOM --graphs
python scripts/e2e_compare.py gpurun_out/e2e_custom_graphs.json gpurun_out/e2e_fa_eager.json | tee gpurun_out/e2e_cmp_graphs.json; echo "cmp rc=$?"
run custom_fp8 --backend CUSTOM --kv-cache-dtype fp8
run fa_fp8 --backend FLASH_ATTN --kv-cache-dtype fp8
python scripts/e2e_compare.py gpurun_out/e2e_custom_fp8.json gpurun_out/e2e_fa_fp8.json 0.15 | tee gpurun_out/e2e_cmp_fp8.json; echo "cmp rc=$?"
tail -3 gpurun_out/e2e_custom_graphs.log | cut -c1-400
tail -3 gpurun_out/e2e_fa_fp8.log | cut -c1-400
