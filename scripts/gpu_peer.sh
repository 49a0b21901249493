#!/bin/bash
# N-GPU pass (gpurun --gpus N): peer-memory collectives over real NVLink — protocol tests, check/timing against NCCL,
# then bench with the peer kernels and with NCCL (plain TP and KV-head replica split).
N=${N:-2}
T=${T:-200}
mkdir -p gpurun_out/golden
nvidia-smi topo -m > gpurun_out/topo_n$N.txt 2>&1
if [ -n "$GOLDEN" ]; then timeout 300 python tests/golden/gen_vllm_gpu_golden.py gpurun_out/golden > gpurun_out/golden.log 2>&1; echo "golden rc=$?"; tail -4 gpurun_out/golden.log; fi
[ -n "$SKIP_PYTEST" ] || timeout 300 python -m pytest tests -m gpu -q --timeout 120 -k "peer or fixtures" > gpurun_out/pytest_peer.log 2>&1; echo "pytest peer rc=$?"; tail -3 gpurun_out/pytest_peer.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port"
[ -n "$SKIP_CHECK" ] || timeout -k 5 $T $TR 29541 scripts/peer_check.py gpurun_out/peer_check_n$N.json > gpurun_out/peer_check_n$N.log 2>&1; echo "peer_check rc=$?"
tail -5 gpurun_out/peer_check_n$N.log | cut -c1-400
cat gpurun_out/peer_check_n$N.json 2>/dev/null | tr -d '\n ' | cut -c1-3000; echo
for ar in ${BENCH_ARS:-peer nccl}; do
  timeout -k 5 $T $TR 29542 bench.py --gpus $N --steps 10 --warmup 3 --allreduce $ar > gpurun_out/scale_n${N}_$ar.json 2> gpurun_out/scale_n${N}_$ar.err; echo "bench n$N $ar rc=$?"
  tail -2 gpurun_out/scale_n${N}_$ar.err | cut -c1-300; cut -c1-330 gpurun_out/scale_n${N}_$ar.json; grep -o '"collective": "[^"]*"' gpurun_out/scale_n${N}_$ar.json; grep -o '"e2e": {"value": [0-9.]*' gpurun_out/scale_n${N}_$ar.json
done
if [ -z "$SKIP_REP" ]; then
for xm in ${REPMODES:-peer}; do
  timeout -k 5 $T $TR 29543 bench.py --gpus $N --steps 10 --warmup 3 --total-kv-heads $((N/2)) --replica-mode split --replica-exchange $xm --allreduce $xm > gpurun_out/scale_n${N}_rep_$xm.json 2> gpurun_out/scale_n${N}_rep_$xm.err; echo "bench n$N replica-split $xm rc=$?"
  tail -2 gpurun_out/scale_n${N}_rep_$xm.err | cut -c1-300; cut -c1-330 gpurun_out/scale_n${N}_rep_$xm.json; grep -o '"replica_exchange": "[^"]*"' gpurun_out/scale_n${N}_rep_$xm.json; grep -o '"e2e": {"value": [0-9.]*' gpurun_out/scale_n${N}_rep_$xm.json
done
fi
