#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_decode.py -x -q -m gpu --timeout 300 2>&1 | tail -4
for wl in cfg3 cfg2 cfg5; do timeout 300 python scripts/compare_vllm_backends.py $wl 16 2>/dev/null | grep -v "flash-attn\|v1\|default call\|once per step\"" | cut -c1-200; done
timeout 900 python scripts/splitk_sweep.py cfg5 8 2>/dev/null | grep "cooperative\|fixed-split 16\|fixed-split 8\|fixed-split 4\"\|fixed-split 2\"" | cut -c1-200
