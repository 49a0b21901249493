#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_decode.py -x -q -m gpu --timeout 120 > gpurun_out/r2k_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r2k_tests.log
for pf in 4 0 12; do
  echo "== L2PF=$pf"
  B200ATTN_SK3_L2PF=$pf timeout 300 python scripts/compare_vllm_backends.py cfg3 16 2>/dev/null | head -4 | cut -c1-190
done
for wl in cfg2 cfg5; do B200ATTN_SK3_L2PF=4 timeout 300 python scripts/compare_vllm_backends.py $wl 16 2>/dev/null | head -3 | cut -c1-190; done
CALIB_STEPS=30 timeout 200 python scripts/sk_timeline.py cfg3 8 1 2 2>/dev/null | grep -v sm_end_corr | cut -c1-420
