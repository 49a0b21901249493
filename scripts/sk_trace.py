#!/usr/bin/env python
"""Per-warp timeline of the stream-K decode kernel (b200attn_debug_set_trace) and CUDA-event
timing for several B200ATTN_SK_DYN / B200ATTN_PDL settings. Usage: sk_trace.py [workload] [layers]"""
import json
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch

from vllm_omni_b200 import _lib, ops, workloads

wl_name = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
L = int(sys.argv[2]) if len(sys.argv) > 2 else 8
wl = workloads.WORKLOADS[wl_name]
inp = workloads.make_decode_inputs(wl, "cuda", num_layers=L)
caches = [ops.split_kv_cache(c) for c in inp.kv_caches]
B, hq, d = wl.batch, inp.shard.num_q_heads, wl.head_size
out = torch.empty(B, hq, d, dtype=torch.bfloat16, device="cuda")
os.environ["B200ATTN_SK_DYN"] = "3"  # size the workspace for the dynamic schedule
ws = ops.DecodeWorkspace("cuda", ops.decode_workspace_bytes(B, hq, d, 1))
lib = _lib.load()
sms = ops.num_sms()
nw = sms * 12
trace = torch.zeros(nw * 3, dtype=torch.int64, device="cuda")


def partial(kc, vc):
    ops.paged_decode_partial(inp.q, kc, vc, inp.block_table, inp.seq_lens, inp.max_seq_len, inp.softmax_scale,
                             out, 1, ws, kv_cache_dtype=wl.kv_cache_dtype, k_scale=inp.k_scale,
                             v_scale=inp.v_scale, variant=ops.VARIANT_STREAMK)


def full(kc, vc):
    ops.paged_decode(inp.q, kc, vc, inp.block_table, inp.seq_lens, inp.max_seq_len, inp.softmax_scale, out=out,
                     kv_cache_dtype=wl.kv_cache_dtype, k_scale=inp.k_scale, v_scale=inp.v_scale, workspace=ws)


def timed(fn, reps=5):
    g = torch.cuda.CUDAGraph()
    for kc, vc in caches:
        fn(kc, vc)
    torch.cuda.synchronize()
    with torch.cuda.graph(g):
        for kc, vc in caches:
            fn(kc, vc)
    for _ in range(3):
        g.replay()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); g.replay(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) / L * 1e3)
    return sorted(ts)[len(ts) // 2]


ref = None
for dyn in ["0", "2", "3", "4"]:
    os.environ["B200ATTN_SK_DYN"] = dyn
    full(*caches[0])
    torch.cuda.synchronize()
    o = out.clone()
    if ref is None:
        ref = o
    err = (o.float() - ref.float()).abs().max().item()
    us_p = timed(partial)
    us_f = timed(full)
    # trace of one launch
    lib.b200attn_debug_set_trace(trace.data_ptr())
    partial(*caches[1 % L])
    torch.cuda.synchronize()
    lib.b200attn_debug_set_trace(None)
    t = trace.view(nw, 3).cpu().double()
    t0 = t[:, 0].min()
    entry, stat, end = (t[:, 0] - t0) / 1e3, (t[:, 1] - t0) / 1e3, (t[:, 2] - t0) / 1e3
    sm_end = end.view(sms, 12).max(dim=1).values
    print(json.dumps({
        "workload": wl_name, "dyn_shift": dyn, "partial_us": round(us_p, 2), "full_us": round(us_f, 2),
        "GBps_partial": round(inp.kv_bytes_per_layer / us_p / 1e3, 1), "max_abs_diff_vs_static": err,
        "entry_us_max": round(entry.max().item(), 2),
        "warp_end_us": [round(x, 2) for x in (end.min().item(), end.mean().item(), end.max().item())],
        "static_end_us": [round(x, 2) for x in (stat.min().item(), stat.mean().item(), stat.max().item())],
        "sm_end_us": [round(x, 2) for x in (sm_end.min().item(), sm_end.mean().item(), sm_end.max().item())],
    }), flush=True)
for pdl in ["0", "1"]:
    os.environ["B200ATTN_SK_DYN"] = "3"
    os.environ["B200ATTN_PDL"] = pdl
    print(json.dumps({"pdl": pdl, "partial_us": round(timed(partial), 2), "full_us": round(timed(full), 2)}), flush=True)
