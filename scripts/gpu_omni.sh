#!/bin/bash
mkdir -p gpurun_out
for be in CUSTOM FLASH_ATTN; do
  timeout 600 python scripts/e2e_vllm.py --omni --backend $be --out gpurun_out/omni_$be.json > gpurun_out/omni_$be.log 2>&1; echo "omni $be rc=$?"
  tail -25 gpurun_out/omni_$be.log | cut -c1-600
done
