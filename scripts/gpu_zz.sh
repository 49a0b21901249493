#!/bin/bash
# static-scratch entry point: alone in a fresh process, then after the whole GPU suite; then smoke + bench on HEAD
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/zz_alone.log 2>&1; echo "zz alone rc=$?" | tee -a gpurun_out/zz_alone.log
tail -3 gpurun_out/zz_alone.log
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/zz_suite.log 2>&1; echo "suite rc=$?" | tee -a gpurun_out/zz_suite.log
tail -5 gpurun_out/zz_suite.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/zz_smoke.log 2>&1; tail -2 gpurun_out/zz_smoke.log
timeout 300 python bench.py > gpurun_out/zz_bench.json 2> gpurun_out/zz_bench.err; tail -1 gpurun_out/zz_bench.json
