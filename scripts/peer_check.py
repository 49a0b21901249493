#!/usr/bin/env python
"""Multi-process check + timing of the peer-memory all-reduce (one rank per GPU, torchrun):

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
      scripts/peer_check.py [out.json]

For every transport that can be set up (torch symmetric memory, CUDA IPC) and every schedule (one-shot,
two-shot, multimem where the allocation has a multicast mapping): results against the sequential fp32 oracle
(every rank regenerates all ranks' inputs from seeds; bit-exact for one-/two-shot, 1 bf16 ulp for multimem),
then microseconds per call from CUDA-graph replays (max over ranks), next to NCCL's all_reduce of the same
tensor captured the same way. Rank 0 writes one JSON document."""
import json
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import torch
import torch.distributed as dist

from oracle import oracle
from vllm_omni_b200 import _lib, peer

HIDDEN = 2048
ROWS = [8, 64, 128, 256, 512]
REPS = 20


def inputs(world, rows, it=0):
    g = torch.Generator().manual_seed(1234 + rows + it)
    return [(torch.randn(rows, HIDDEN, generator=g) * 2).to(torch.bfloat16) for _ in range(world)]


def graph_us(fn, world):
    fn(); torch.cuda.synchronize(); dist.barrier()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(REPS):
            fn()
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize(); dist.barrier()
    ts = []
    for _ in range(7):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dist.barrier(); a.record(); g.replay(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3 / REPS)
    t = torch.tensor([sorted(ts)[len(ts) // 2]], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    doc = {"world": world, "hidden": HIDDEN, "transports": {}, "nccl_us": {}}
    for rows in ROWS:
        x = inputs(world, rows)[rank].cuda()
        doc["nccl_us"][str(rows)] = round(graph_us(lambda: dist.all_reduce(x), world), 2)
    for tr in ("symm", "ipc"):
        try:
            ar = peer.PeerAllReduce.create(max_numel=max(ROWS) * HIDDEN, transport=tr)
        except Exception as e:  # noqa: BLE001
            doc["transports"][tr] = {"unavailable": f"{type(e).__name__}: {str(e)[:400]}"}
            continue
        rec = {"multicast": ar.has_multicast, "algos": {}}
        algos = [_lib.AR_PUSH, _lib.AR_PUSH2, _lib.AR_ONESHOT, _lib.AR_TWOSHOT] + ([_lib.AR_MULTIMEM] if ar.has_multicast else [])
        for algo in algos:
            name = peer.ALGO_NAMES[algo]
            arec = {"exact": True, "max_abs_err": 0.0, "us": {}}
            try:
                for rows in ROWS:
                    if algo == _lib.AR_PUSH and 2 * rows * HIDDEN > ar.push_slot_bytes:
                        continue
                    if algo == _lib.AR_PUSH2 and 2 * rows * HIDDEN // world + 16 * world > ar.push_slot_bytes // 2:
                        continue
                    for it in range(2):
                        parts = inputs(world, rows, it)
                        xin = ar.input((rows, HIDDEN))
                        xin.copy_(parts[rank].cuda())
                        torch.cuda.synchronize(); dist.barrier()
                        out = ar.all_reduce(xin, algo=algo)
                        torch.cuda.synchronize(); dist.barrier()
                        ref = oracle.ref_all_reduce_bf16(parts)
                        got = out.cpu()
                        arec["exact"] &= bool(torch.equal(got.view(torch.int16), ref.view(torch.int16)))
                        arec["max_abs_err"] = max(arec["max_abs_err"], float((got.float() - ref.float()).abs().max()))
                    xin = ar.input((rows, HIDDEN))
                    arec["us"][str(rows)] = round(graph_us(lambda: ar.all_reduce(xin, algo=algo), world), 2)
                flags = torch.tensor([int(arec["exact"])], device="cuda")
                dist.all_reduce(flags, op=dist.ReduceOp.MIN)
                arec["exact_all_ranks"] = bool(flags.item())
            except Exception as e:  # noqa: BLE001
                arec["error"] = f"{type(e).__name__}: {str(e)[:400]}"
            rec["algos"][name] = arec
        rec["auto"] = {str(r): peer.ALGO_NAMES[peer.choose_algo(2 * r * HIDDEN, world, ar.has_multicast, ar.push_slot_bytes)] for r in ROWS}
        doc["transports"][tr] = rec
    if rank == 0:
        s = json.dumps(doc, indent=1)
        print(s, flush=True)
        if len(sys.argv) > 1:
            Path(sys.argv[1]).parent.mkdir(parents=True, exist_ok=True)
            Path(sys.argv[1]).write_text(s)
    dist.barrier()
    sys.stdout.flush()
    os._exit(0)


if __name__ == "__main__":
    main()
