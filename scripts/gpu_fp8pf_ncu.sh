#!/bin/bash
# launch list of the fp8 prefill paths (8k case): mma.sync kernel, then staging kernels + tcgen05 kernel
mkdir -p gpurun_out
timeout 240 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"stage_fp8|paged_prefill" -c 28 --csv \
  --log-file gpurun_out/fp8pf_launches.csv python scripts/prefill_bench.py --fp8-only > gpurun_out/fp8pf_ncu.log 2>&1
echo "ncu rc=$?"; tail -5 gpurun_out/fp8pf_launches.csv | cut -c1-300
