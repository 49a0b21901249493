#!/usr/bin/env python
"""BASELINE.json configs[4]: split-K sweep of the long-context decode (cfg5: batch 16, 32k context) for the per-rank
shard of TP 1 / 2 / 4 / 8 (KV-head sharding; at TP 8 the KV head is replicated and the batch split inside the replica
group, see tp.ReplicaSplit). The decode has no collective, so every rank of a TP-N run does exactly the work of rank 0's
shard: it is measured here on one GPU (run it under torchrun to see the same numbers on N GPUs).

  python scripts/splitk_sweep.py [workload] [layers]      -> one JSON line per (tp, schedule)

Schedules: the fixed-split kernel (`paged_decode_stream_kernel` + `lse_reduce_kernel`) with splits 1 .. 64, and the
stream-K generations (no split count): v1 (per-warp chunks + reduce launch), v2, cooperative (default)."""
import json
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
os.environ.setdefault("B200ATTN_PLAN_SK2", "1")   # the step's plan launch also makes the second-generation plan (comparison arm)
import torch

from vllm_omni_b200 import ops, workloads

wl_name = sys.argv[1] if len(sys.argv) > 1 else "cfg5"
L = int(sys.argv[2]) if len(sys.argv) > 2 else 8
wl = workloads.WORKLOADS[wl_name]


def timed(fn, caches, reps=5, pre=None):
    """Median per-launch microseconds of a CUDA-graph replay of one launch per layer cache. `pre` (the step's plan
    launch of the planned schedules: once per STEP of a model's 48 layers, 18 us) runs eagerly before every replay,
    outside the timed region — how the vLLM backend runs it (metadata builder, then the captured decode)."""
    def one_pass():
        for kc, vc in caches:
            fn(kc, vc)
    if pre is not None:
        pre()
    one_pass()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        one_pass()
    for _ in range(10):
        if pre is not None:
            pre()
        g.replay()
    ts = []
    for _ in range(reps):
        if pre is not None:
            pre()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); g.replay(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) / len(caches) * 1e3)
    return sorted(ts)[len(ts) // 2]


for tp in (1, 2, 4, 8):
    inp = workloads.make_decode_inputs(wl, "cuda", num_layers=L, tp_size=tp, rank=0)
    sh = inp.shard
    B, hq, hkv, d = wl.batch, sh.num_q_heads, sh.num_kv_heads, wl.head_size
    q, bt, sl = inp.q, inp.block_table, inp.seq_lens
    note = ""
    if sh.kv_replicas > 1:   # the replica group splits the batch: this rank serves B / R sequences for R x its query heads
        per = B // sh.kv_replicas
        q = torch.randn(per, hq * sh.kv_replicas, d, device="cuda", dtype=torch.bfloat16)
        bt, sl, B, hq = bt[:per], sl[:per], per, hq * sh.kv_replicas
        note = f"KV head replicated on {sh.kv_replicas} ranks: batch split inside the replica group"
    caches = [ops.split_kv_cache(c) for c in inp.kv_caches]
    kv_bytes = int(sl.sum().item()) * 2 * hkv * d * (1 if wl.fp8 else 2)
    out = torch.empty(B, hq, d, dtype=torch.bfloat16, device="cuda")
    ws = ops.DecodeWorkspace("cuda", ops.decode_workspace_bytes(B, hq, d, 64))

    def run(variant, splits, planned=False):
        def fn(kc, vc):
            ops.paged_decode(q, kc, vc, bt, sl, inp.max_seq_len, inp.softmax_scale, out=out, num_splits=splits,
                             variant=variant, workspace=ws, kv_cache_dtype=wl.kv_cache_dtype, k_scale=inp.k_scale,
                             v_scale=inp.v_scale, planned=planned, early_kv_read=planned, early_q_read=planned)
        return fn

    rows = [(f"fixed-split {s}", ops.VARIANT_STREAM, s, False) for s in (1, 2, 4, 8, 16, 32, 64)]
    rows += [("stream-K v1 (reduce launch)", ops.VARIANT_STREAMK, 0, False),
             ("stream-K v2 (reduce launch)", ops.VARIANT_SK2, 0, True),
             ("cooperative stream-K (default)", ops.VARIANT_SK3, 0, True)]
    for name, variant, splits, planned in rows:
        if wl.fp8 and variant == ops.VARIANT_STREAM:
            continue
        if planned:
            for _ in range(26):   # let the SM shares settle (frozen after 24 plan launches)
                ops.decode_plan(sl, bt, B, hq, hkv, wl.block_size, ws)
                for kc, vc in caches:
                    run(variant, splits, True)(kc, vc)
        us = timed(run(variant, splits, planned), caches,
                   pre=(lambda: ops.decode_plan(sl, bt, B, hq, hkv, wl.block_size, ws)) if planned else None)
        print(json.dumps({"workload": wl_name, "tp": tp, "rank_shape": f"B={B} Hq={hq} Hkv={hkv}", "schedule": name,
                          "us_per_launch": round(us, 2), "GBps_per_gpu": round(kv_bytes / us / 1e3, 1),
                          "frac_of_8TBps": round(kv_bytes / us / 1e3 / 8000, 3), "kv_bytes_per_gpu": kv_bytes,
                          "note": note}), flush=True)
    del inp, caches
    torch.cuda.empty_cache()
