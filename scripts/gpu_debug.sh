#!/bin/bash
mkdir -p gpurun_out
T='tests/test_gpu_decode.py::test_named_shapes_full_context'
timeout 300 python -m pytest "$T" -x -q -m gpu --timeout 120 2>&1 | tail -5; echo "plain rc=$?"
B200ATTN_SK3_OWNER_FAST_ONLY=1 timeout 300 python -m pytest "$T" -x -q -m gpu --timeout 120 2>&1 | tail -3; echo "owner-fast-only"
B200ATTN_SK3_POOL=0 timeout 300 python -m pytest "$T" -x -q -m gpu --timeout 120 2>&1 | tail -3; echo "pool 0"
timeout 600 compute-sanitizer --tool memcheck --print-limit 5 python -m pytest "$T" -x -q -m gpu --timeout 500 -k "cfg2" > gpurun_out/dbg_memcheck.log 2>&1; echo "memcheck rc=$?"
grep -E "ERROR SUMMARY|Invalid|at .*sk3|passed|failed" gpurun_out/dbg_memcheck.log | head -20
timeout 600 compute-sanitizer --tool racecheck --print-limit 5 python -m pytest "$T" -x -q -m gpu --timeout 500 -k "cfg2" > gpurun_out/dbg_racecheck.log 2>&1; echo "racecheck rc=$?"
grep -E "RACECHECK SUMMARY|hazard|at .*sk3|passed|failed" gpurun_out/dbg_racecheck.log | head -20
timeout 600 python -m pytest tests/test_gpu_decode.py -x -q -m gpu --timeout 120 2>&1 | tail -5
