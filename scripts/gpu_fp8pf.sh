#!/bin/bash
# staged fp8 prefill on the tcgen05 kernel: parity tests, backend tests, timing against the mma.sync fp8 kernel
mkdir -p gpurun_out
timeout 420 python -m pytest tests/test_gpu_prefill.py tests/test_gpu_backend.py -x -q -m gpu > gpurun_out/fp8pf_tests.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/fp8pf_tests.log
tail -15 gpurun_out/fp8pf_tests.log
timeout 200 python scripts/prefill_bench.py --fp8-only > gpurun_out/fp8pf_bench.jsonl 2> gpurun_out/fp8pf_bench.err; echo "bench rc=$?"
cat gpurun_out/fp8pf_bench.jsonl; tail -3 gpurun_out/fp8pf_bench.err
