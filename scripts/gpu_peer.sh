#!/bin/bash
# N-GPU pass (gpurun --gpus N): peer all-reduce check/timing over real NVLink, then bench with the peer kernel and with NCCL.
N=${N:-2}
T=${T:-240}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_n$N.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port"
timeout -k 5 $T $TR 29541 scripts/peer_check.py gpurun_out/peer_check_n$N.json > gpurun_out/peer_check_n$N.log 2>&1; echo "peer_check rc=$?"
tail -5 gpurun_out/peer_check_n$N.log | cut -c1-400
cat gpurun_out/peer_check_n$N.json 2>/dev/null | tr -d '\n ' | cut -c1-3000; echo
for ar in peer nccl; do
  timeout -k 5 $T $TR 29542 bench.py --gpus $N --steps 10 --warmup 3 --allreduce $ar > gpurun_out/scale_n${N}_$ar.json 2> gpurun_out/scale_n${N}_$ar.err; echo "bench n$N $ar rc=$?"
  tail -2 gpurun_out/scale_n${N}_$ar.err | cut -c1-300; cut -c1-400 gpurun_out/scale_n${N}_$ar.json; grep -o '"collective": "[^"]*"' gpurun_out/scale_n${N}_$ar.json
done
