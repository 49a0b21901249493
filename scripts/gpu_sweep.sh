#!/bin/bash
mkdir -p gpurun_out
timeout 900 python scripts/splitk_sweep.py cfg5 8 > gpurun_out/sweep_cfg5.json 2> gpurun_out/sweep_cfg5.err; echo "sweep rc=$?"
grep "cooperative\|fixed-split 16\|fixed-split 8\"" gpurun_out/sweep_cfg5.json | cut -c1-200; tail -2 gpurun_out/sweep_cfg5.err
