#!/bin/bash
# 1-GPU pass: full GPU parity suite, golden vectors from vLLM's stock kernels, one default bench line.
mkdir -p gpurun_out/golden
timeout 1500 python -m pytest tests -m gpu -q --timeout 300 ${K:+-k "$K"} > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python tests/golden/gen_vllm_gpu_golden.py gpurun_out/golden > gpurun_out/golden.log 2>&1; echo "golden rc=$?"
tail -8 gpurun_out/golden.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_cfg3.json 2> gpurun_out/bench_cfg3.err; echo "bench rc=$?"
cat gpurun_out/bench_cfg3.json | cut -c1-1500
