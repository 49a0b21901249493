#!/bin/bash
mkdir -p gpurun_out
N=${1:-2}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 5 > gpurun_out/r2_scale_n$N.json 2> gpurun_out/r2_scale_n$N.err; echo "bench N=$N rc=$?"
tail -c 3500 gpurun_out/r2_scale_n$N.json; tail -8 gpurun_out/r2_scale_n$N.err | cut -c1-300
