#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu --timeout 600 > gpurun_out/r2p_tests.log 2>&1; echo "tests rc=$?"
tail -8 gpurun_out/r2p_tests.log
grep -n "PASSED\|vllm-omni runner\|per-head fp8\|stock-vs-stock\|eager:\|graphs:" gpurun_out/r2p_tests.log | head
