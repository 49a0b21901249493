#!/usr/bin/env python
"""Same-box GPU comparators: the decode kernels behind the reference's stock attention
backends — FlashInfer's TRT-LLM-gen paged decode (what vLLM's FLASHINFER backend calls on sm_100,
vllm/v1/attention/backends/flashinfer.py:1694-1822) and FlashAttention's paged varlen call (the
FLASH_ATTN backend, flash_attn.py:758-819) — run on the SAME cache memory, block table and queries
as this repo's kernel, timed the same way (L distinct layer caches in rotation, CUDA events), and
checked for agreement. Library kernels are comparators only; nothing here is on the product path.

  python scripts/compare_vllm_backends.py [workload] [layers]   -> one JSON line per implementation
"""
import json
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
os.environ.setdefault("B200ATTN_PLAN_SK2", "1")   # the step's plan launch also makes the second-generation plan (comparison arm)
import torch

from vllm_omni_b200 import ops, workloads

wl_name = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
L = int(sys.argv[2]) if len(sys.argv) > 2 else 8
wl = workloads.WORKLOADS[wl_name]
inp = workloads.make_decode_inputs(wl, "cuda", num_layers=L)
B, hq, hkv, d = wl.batch, wl.total_q_heads, wl.total_kv_heads, wl.head_size
caches = [ops.split_kv_cache(c) for c in inp.kv_caches]
kv_bytes = inp.kv_bytes_per_layer
ws = ops.DecodeWorkspace("cuda", ops.decode_workspace_bytes(B, hq, d, 1))
out_ours = torch.empty(B, hq, d, dtype=torch.bfloat16, device="cuda")


def make_ours(variant, planned=False, early=False, write=False):
    def fn(kc, vc, out):
        if write:
            ops.reshape_and_cache(inp.new_k, inp.new_v, kc, vc, inp.slot_mapping, wl.kv_cache_dtype, inp.k_scale,
                                  inp.v_scale)
        ops.paged_decode(inp.q, kc, vc, inp.block_table, inp.seq_lens, inp.max_seq_len, inp.softmax_scale, out=out,
                         kv_cache_dtype=wl.kv_cache_dtype, k_scale=inp.k_scale, v_scale=inp.v_scale, workspace=ws,
                         variant=variant, planned=planned, early_kv_read=early, early_q_read=early)
    return fn


ours = make_ours(ops.VARIANT_AUTO)


def plan_step():
    ops.decode_plan(inp.seq_lens, inp.block_table, B, hq, hkv, wl.block_size, ws)


def timed(fn, reps=7, pre=None):
    """Median per-launch microseconds over `reps` passes of the L caches (graph replay when capturable).
    `pre` runs once per pass, inside the timed graph (the step's plan launch of the planned modes)."""
    out = torch.empty(B, hq, d, dtype=torch.bfloat16, device="cuda")

    def one_pass():
        if pre is not None:
            pre()
        for kc, vc in caches:
            fn(kc, vc, out)

    one_pass()
    torch.cuda.synchronize()
    run, mode = one_pass, "eager"
    try:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            one_pass()
        run, mode = g.replay, "cuda-graph"
    except Exception:
        torch.cuda.synchronize()
    for _ in range(12):  # warm-up; also lets the stream-K SM weights settle (frozen after 8 plan launches)
        run()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); run(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) / L * 1e3)
    return sorted(ts)[len(ts) // 2], mode


def report(name, fn, note="", pre=None):
    try:
        out = torch.empty(B, hq, d, dtype=torch.bfloat16, device="cuda")
        fn(caches[0][0], caches[0][1], out)
        torch.cuda.synchronize()
        diff = (out.float() - out_ours.float()).abs().max().item()
        us, mode = timed(fn, pre=pre)
        print(json.dumps({"impl": name, "workload": wl_name, "us_per_launch": round(us, 2),
                          "GBps": round(kv_bytes / us / 1e3, 1), "frac_of_8TBps": round(kv_bytes / us / 1e3 / 8000, 3),
                          "max_abs_diff_vs_b200attn": diff, "launch_mode": mode, "note": note}), flush=True)
    except Exception as e:
        print(json.dumps({"impl": name, "workload": wl_name, "unavailable": f"{type(e).__name__}: {str(e)[:300]}"}),
              flush=True)


ours(caches[0][0], caches[0][1], out_ours)
torch.cuda.synchronize()
ops.decode_plan(inp.seq_lens, inp.block_table, B, hq, hkv, wl.block_size, ws)  # the step's plan (planned modes)
report("b200attn cooperative stream-K (default call: plan launch + decode, nothing planned ahead)", ours)
report("b200attn cooperative stream-K, planned once per step", make_ours(ops.VARIANT_SK3, planned=True), pre=plan_step,
       note=f"one plan launch per {L} layers inside the timed graph")
report("b200attn cooperative stream-K, planned + early KV / q reads (what the backend passes)",
       make_ours(ops.VARIANT_SK3, planned=True, early=True), pre=plan_step)
report("b200attn cooperative stream-K, planned + early, cache write + decode per layer",
       make_ours(ops.VARIANT_SK3, planned=True, early=True, write=True), note="two launches per layer", pre=plan_step)
report("b200attn stream-K v2 (round 2 first half: per-warp chunks + reduce launch), planned + early",
       make_ours(ops.VARIANT_SK2, planned=True, early=True), pre=plan_step)
report("b200attn stream-K v1 (round 1: decode + reduce launch)", make_ours(ops.VARIANT_STREAMK))
report("b200attn stream-K v1, cache write + decode + reduce per layer", make_ours(ops.VARIANT_STREAMK, write=True),
       note="three launches per layer")

fp8 = wl.fp8
if fp8:
    ks, vs = float(inp.k_scale.flatten()[0]), float(inp.v_scale.flatten()[0])

# ---- FlashInfer TRT-LLM-gen decode on the same memory (HND view of our cache)
try:
    import flashinfer

    fi_ws = torch.zeros(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")

    def fi(kc, vc, out):
        k = kc.permute(0, 2, 1, 3)  # [nb, Hkv, bs, d] view, no copy
        v = vc.permute(0, 2, 1, 3)
        if fp8:
            k, v = k.view(torch.float8_e4m3fn), v.view(torch.float8_e4m3fn)
        flashinfer.decode.trtllm_batch_decode_with_kv_cache(
            inp.q if not fp8 else inp.q, (k, v), fi_ws, inp.block_table, inp.seq_lens, inp.max_seq_len,
            bmm1_scale=inp.softmax_scale * (ks if fp8 else 1.0), bmm2_scale=(vs if fp8 else 1.0), out=out,
            kv_layout="HND")

    if fp8 and inp.k_scale.numel() > 1:
        print(json.dumps({"impl": "flashinfer trtllm-gen", "workload": wl_name,
                          "unavailable": "per-head KV scales are not expressible (scalar bmm scales only)"}))
    else:
        report("flashinfer trtllm-gen decode (vLLM FLASHINFER backend)", fi)
except Exception as e:
    print(json.dumps({"impl": "flashinfer", "unavailable": f"{type(e).__name__}: {e}"}))

# ---- FlashAttention paged varlen (vLLM FLASH_ATTN backend), every version that loads here
try:
    from vllm.vllm_flash_attn import flash_attn_varlen_func

    cu_q = torch.arange(B + 1, dtype=torch.int32, device="cuda")
    for ver in (2, 3, 4):
        def fa(kc, vc, out, ver=ver):
            flash_attn_varlen_func(inp.q, kc, vc, max_seqlen_q=1, cu_seqlens_q=cu_q, max_seqlen_k=inp.max_seq_len,
                                   seqused_k=inp.seq_lens, softmax_scale=inp.softmax_scale, causal=True,
                                   block_table=inp.block_table, out=out, fa_version=ver)
        if fp8:
            print(json.dumps({"impl": f"flash-attn v{ver}", "workload": wl_name, "unavailable": "fp8 KV not run"}))
        else:
            report(f"flash-attn v{ver} paged varlen (vLLM FLASH_ATTN backend)", fa)
except Exception as e:
    print(json.dumps({"impl": "flash-attn", "unavailable": f"{type(e).__name__}: {e}"}))
