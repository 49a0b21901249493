#!/bin/bash
mkdir -p gpurun_out
export B200ATTN_SK3_POOL=0
CALIB_STEPS=30 timeout 200 python scripts/sk_timeline.py cfg3 8 1 3 > gpurun_out/r2j_timeline_cfg3.json 2> gpurun_out/r2j_timeline.err; echo "timeline rc=$?"
grep -v sm_end_correlation gpurun_out/r2j_timeline_cfg3.json | cut -c1-600; tail -2 gpurun_out/r2j_timeline.err
timeout 300 python scripts/compare_vllm_backends.py cfg3 16 2>/dev/null | head -3 | cut -c1-200
