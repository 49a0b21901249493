#!/bin/bash
# First GPU pass: smoke, GPU parity tests, short bench, ncu launch list.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/rc.txt
timeout 900 python -m pytest tests -m gpu -q -x --timeout 120 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/rc.txt
tail -5 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_cfg3.log 2>&1; echo "bench rc=$?" | tee -a gpurun_out/rc.txt
tail -3 gpurun_out/bench_cfg3.log
timeout 300 python bench.py --steps 5 --warmup 3 --variant 2 --no-e2e --no-cpu-baseline > gpurun_out/bench_cfg3_cpasync.log 2>&1; echo "bench cpasync rc=$?" | tee -a gpurun_out/rc.txt
tail -1 gpurun_out/bench_cfg3_cpasync.log
