#!/bin/bash
# N-GPU checks (run with gpurun --gpus N): bench at TP=N, and KV-head replication in both modes.
# Every run has its own short timeout (a hung rank must not eat the box time).
N=${N:-2}
T=${T:-240}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
if [ -z "$SKIP_PLAIN" ]; then
timeout -k 5 $T $TR bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err; echo "bench n$N rc=$?"
tail -2 gpurun_out/scale_n$N.err | cut -c1-300; cat gpurun_out/scale_n$N.json
fi
if [ -n "$REP" ]; then
for mode in replicate split; do
  timeout -k 5 $T $TR bench.py --gpus $N --steps 10 --warmup 3 --total-kv-heads $((N/2)) --replica-mode $mode > gpurun_out/scale_n${N}_rep_$mode.json 2> gpurun_out/scale_n${N}_rep_$mode.err; echo "bench n$N $mode rc=$?"
  tail -2 gpurun_out/scale_n${N}_rep_$mode.err | cut -c1-300; cat gpurun_out/scale_n${N}_rep_$mode.json
done
fi
