#!/usr/bin/env python
"""Per-warp timeline of the stream-K decode kernels (VARIANT=6 cooperative, default; VARIANT=5 second generation) (b200attn_debug_set_trace): for every warp the
%globaltimer stamps {entry, first stage landed, end of streaming, merges done} and the SM id of isolated
launches (device idle before each). Usage: sk_timeline.py [workload] [layers] [tp] [launches]"""
import json
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch

from vllm_omni_b200 import _lib, ops, workloads

wl_name = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
L = int(sys.argv[2]) if len(sys.argv) > 2 else 8
tp = int(sys.argv[3]) if len(sys.argv) > 3 else 1
launches = int(sys.argv[4]) if len(sys.argv) > 4 else 3
contig = os.environ.get("CONTIG", "0") == "1"
variant = int(os.environ.get("VARIANT", "6"))
wl = workloads.WORKLOADS[wl_name]
inp = workloads.make_decode_inputs(wl, "cuda", num_layers=L, tp_size=tp, rank=0, contiguous_blocks=contig)
caches = [ops.split_kv_cache(c) for c in inp.kv_caches]
B, hq, hkv, d = wl.batch, inp.shard.num_q_heads, inp.shard.num_kv_heads, wl.head_size
out = torch.empty(B, hq, d, dtype=torch.bfloat16, device="cuda")
ws = ops.DecodeWorkspace("cuda", ops.decode_workspace_bytes(B, hq, d, 1))
lib = _lib.load()
sms = ops.num_sms()
W = 12
nw = sms * W
trace = torch.zeros(nw * 8, dtype=torch.int64, device="cuda")
ops.decode_plan(inp.seq_lens, inp.block_table, B, hq, hkv, wl.block_size, ws)


def run(kc, vc, planned=True, early=True):
    ops.paged_decode(inp.q, kc, vc, inp.block_table, inp.seq_lens, inp.max_seq_len, inp.softmax_scale, out=out,
                     kv_cache_dtype=wl.kv_cache_dtype, k_scale=inp.k_scale, v_scale=inp.v_scale, workspace=ws,
                     variant=variant, planned=planned, early_kv_read=early, early_q_read=early)


def mmm(x):
    return [round(x.min().item(), 2), round(x.mean().item(), 2), round(x.max().item(), 2)]


calib_steps = int(os.environ.get("CALIB_STEPS", "12"))
for i in range(calib_steps):  # a "step" = plan + L layers; the SM weights settle over the first 8 plans
    ops.decode_plan(inp.seq_lens, inp.block_table, B, hq, hkv, wl.block_size, ws)
    for kc, vc in caches:
        run(kc, vc)
torch.cuda.synchronize()
# Sk2Ctrl (csrc/decode_common.cuh) at byte 256: claimed[2][512], finished[2], calib_updates, pad, weight[512], ...
ctrl = ws.buf[256:256 + 4 * (1024 + 4 + 512 * 3)].view(torch.int32)
w = ctrl[1028: 1028 + sms].view(torch.float32).cpu()
print(json.dumps({"calib_updates": int(ctrl[1026]), "weight_min_mean_max_x_sms":
                  [round(float(w.min()) * sms, 3), round(float(w.mean()) * sms, 3), round(float(w.max()) * sms, 3)]}), flush=True)
per_sm = []
for rep in range(launches):
    lib.b200attn_debug_set_trace(trace.data_ptr())
    run(*caches[(3 + rep) % L])
    torch.cuda.synchronize()
    lib.b200attn_debug_set_trace(None)
    t = trace.view(nw, 8).cpu().double()
    t0 = t[:, 0].min()
    cols = [(t[:, i] - t0) / 1e3 for i in range(4)]
    smid = t[:, 4].long()
    # per SM (by hardware id): last warp's end of streaming
    end_by_sm = torch.zeros(int(smid.max()) + 1)
    end_by_sm.scatter_reduce_(0, smid, cols[2].float(), reduce="amax")
    first_by_sm = torch.full((int(smid.max()) + 1,), 1e9)
    first_by_sm.scatter_reduce_(0, smid, cols[2].float(), reduce="amin")
    used = torch.zeros(int(smid.max()) + 1, dtype=torch.bool)
    used[smid] = True
    per_sm.append(end_by_sm)
    # does an SM that got its first tiles late also end late? (per-SM mean of its warps' first landing vs its end)
    fl_sum = torch.zeros(int(smid.max()) + 1).scatter_add_(0, smid, cols[1].float())
    fl_cnt = torch.zeros(int(smid.max()) + 1).scatter_add_(0, smid, torch.ones(nw))
    fl_by_sm = fl_sum / fl_cnt.clamp(min=1)
    cc = torch.corrcoef(torch.stack([fl_by_sm[used], end_by_sm[used]]))[0, 1]
    tiles_by_sm = torch.zeros(int(smid.max()) + 1).scatter_add_(0, smid, t[:, 5].float())
    rate = tiles_by_sm[used] / (end_by_sm[used] - fl_by_sm[used]).clamp(min=1e-3)   # tiles / us after the first landing
    rate = rate if float(rate.sum()) > 0 else torch.ones_like(rate)                  # (kernels that do not count tiles)
    print(json.dumps({"launch": rep, "corr_first_landed_vs_end_by_sm": round(float(cc), 3),
                      "first_landed_by_sm_us": mmm(fl_by_sm[used]), "tiles_by_sm": mmm(tiles_by_sm[used]),
                      "rate_after_first_landing_tiles_per_us": mmm(rate),
                      "rate_rel_std": round(float(rate.std() / rate.mean()), 4),
                      "end_std_us": round(float(end_by_sm[used].std()), 2)}), flush=True)
    print(json.dumps({"workload": wl_name, "variant": variant, "tp": tp, "launch": rep, "kv_bytes": inp.kv_bytes_per_layer,
                      "interleave": os.environ.get("B200ATTN_SK2_DEBUG_INTERLEAVE", "0"), "contiguous_blocks": contig,
                      "smid_max": int(smid.max()), "sms_used": int(used.sum()),
                      "entry_us": mmm(cols[0]), "first_landed_us": mmm(cols[1]), "stream_end_us": mmm(cols[2]),
                      "merge_end_us": mmm(cols[3]), "sm_stream_end_us": mmm(end_by_sm[used]),
                      "spread_within_sm_us": mmm((end_by_sm - first_by_sm)[used])}), flush=True)
if launches >= 2:
    a = torch.stack(per_sm)
    c = torch.corrcoef(a)
    print(json.dumps({"sm_end_correlation_between_launches": [[round(float(x), 3) for x in r] for r in c],
                      "slowest_sms_launch0": torch.topk(a[0], 8).indices.tolist(),
                      "slowest_sms_launch1": torch.topk(a[1], 8).indices.tolist(),
                      "fastest_sms_launch0": torch.topk(-a[0], 8).indices.tolist(),
                      "fastest_sms_launch1": torch.topk(-a[1], 8).indices.tolist(),
                      "sm_end_launch0_by_smid": [round(float(x), 1) for x in a[0]]}), flush=True)
