#!/bin/bash
# final single-GPU pass: the test files after test_gpu_e2e_vllm (the earlier run stopped there), final sweep / compare,
# ncu re-capture of the final kernel
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_e2e_vllm.py tests/test_gpu_peer_allreduce.py tests/test_gpu_prefill.py -x -q -m gpu --timeout 900 -s 2>&1 | grep -v "Warning\|warn" | tail -12
bash scripts/gpu_profiles_r2.sh 2>&1 | grep "rc="
for wl in cfg3 cfg2 cfg5; do
  timeout 300 python scripts/compare_vllm_backends.py $wl 16 > gpurun_out/r2z_cmp_$wl.jsonl 2> gpurun_out/r2z_cmp_$wl.err; echo "compare $wl rc=$?"
  grep "planned + early KV\|trtllm" gpurun_out/r2z_cmp_$wl.jsonl | cut -c1-200
done
timeout 900 python scripts/splitk_sweep.py cfg5 8 > gpurun_out/sweep_cfg5.json 2> gpurun_out/sweep_cfg5.err; echo "sweep rc=$?"
grep "cooperative" gpurun_out/sweep_cfg5.json | cut -c1-200
