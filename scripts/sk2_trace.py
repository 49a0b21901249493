#!/usr/bin/env python
"""Per-warp timeline of the stream-K v2 decode kernel (b200attn_debug_set_trace): for every warp the
%globaltimer stamps {entry, first stage landed, end of streaming, merges done} of ONE isolated launch
(device idle before it), plus CUDA-event timings of the launch modes. Usage: sk2_trace.py [workload] [layers]"""
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch

from vllm_omni_b200 import _lib, ops, workloads

wl_name = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
L = int(sys.argv[2]) if len(sys.argv) > 2 else 8
tp = int(sys.argv[3]) if len(sys.argv) > 3 else 1
wl = workloads.WORKLOADS[wl_name]
inp = workloads.make_decode_inputs(wl, "cuda", num_layers=L, tp_size=tp, rank=0)
caches = [ops.split_kv_cache(c) for c in inp.kv_caches]
B, hq, hkv, d = wl.batch, inp.shard.num_q_heads, inp.shard.num_kv_heads, wl.head_size
out = torch.empty(B, hq, d, dtype=torch.bfloat16, device="cuda")
ws = ops.DecodeWorkspace("cuda", ops.decode_workspace_bytes(B, hq, d, 1))
lib = _lib.load()
sms = ops.num_sms()
W = 12
nw = sms * W
trace = torch.zeros(nw * 4, dtype=torch.int64, device="cuda")
ops.decode_plan(inp.seq_lens, inp.block_table, B, hq, hkv, wl.block_size, ws)


def run(kc, vc, planned=True, early=True):
    ops.paged_decode(inp.q, kc, vc, inp.block_table, inp.seq_lens, inp.max_seq_len, inp.softmax_scale, out=out,
                     kv_cache_dtype=wl.kv_cache_dtype, k_scale=inp.k_scale, v_scale=inp.v_scale, workspace=ws,
                     variant=ops.VARIANT_SK2, planned=planned, early_kv_read=early)


for i in range(3):
    run(*caches[i % L])
torch.cuda.synchronize()
lib.b200attn_debug_set_trace(trace.data_ptr())
run(*caches[3 % L])
torch.cuda.synchronize()
lib.b200attn_debug_set_trace(None)
t = trace.view(nw, 4).cpu().double()
t0 = t[:, 0].min()
cols = [(t[:, i] - t0) / 1e3 for i in range(4)]


def mmm(x):
    return [round(x.min().item(), 2), round(x.mean().item(), 2), round(x.max().item(), 2)]


sm_end = cols[2].view(sms, W).max(dim=1).values
sm_first_end = cols[2].view(sms, W).min(dim=1).values
print(json.dumps({"workload": wl_name, "tp": tp, "kv_bytes": inp.kv_bytes_per_layer,
                  "entry_us": mmm(cols[0]), "first_landed_us": mmm(cols[1]), "stream_end_us": mmm(cols[2]),
                  "merge_end_us": mmm(cols[3]), "sm_stream_end_us": mmm(sm_end),
                  "spread_within_sm_us": mmm(sm_end - sm_first_end)}), flush=True)
