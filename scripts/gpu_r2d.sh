#!/bin/bash
# round-2 GPU pass d: decode parity after the parallel plan kernel, bench.py (new contract) at N=1, comparators
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_decode.py -x -q -m gpu > gpurun_out/r2d_tests.log 2>&1; echo "tests rc=$?" 
tail -5 gpurun_out/r2d_tests.log
timeout 600 python bench.py > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err; echo "bench rc=$?"
tail -c 6000 gpurun_out/r2d_bench.json; tail -20 gpurun_out/r2d_bench.err
timeout 300 python scripts/compare_vllm_backends.py cfg3 48 > gpurun_out/r2d_cmp_cfg3.jsonl 2>&1; cat gpurun_out/r2d_cmp_cfg3.jsonl | cut -c1-400
