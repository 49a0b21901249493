#!/usr/bin/env python
"""Prefill / extend kernel timing: causal varlen attention over the paged cache,
TFLOP/s counted as 4 * d * Hq * (number of attended (query, key) pairs); FlashAttention's paged varlen
call on the same memory as the comparator. One JSON line per case and implementation."""
import json
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch

from vllm_omni_b200 import layout, ops

HQ, HKV, D, BS = 32, 4, 128, 16
CASES = [  # (name, [(context incl. new tokens, new tokens)] per sequence)
    ("prefill 1 x 8k", [(8192, 8192)]),
    ("prefill 4 x 2k", [(2048, 2048)] * 4),
    ("chunked prefill: 2k new tokens on 30k context", [(32768, 2048)]),
    ("mixed extend 16 x (4k ctx, 64 new)", [(4096, 64)] * 16),
]


def pairs(S, n):
    return n * (S - n) + n * (n + 1) // 2


for fp8 in ((False,) if "--bf16-only" in sys.argv else (True,) if "--fp8-only" in sys.argv else (False, True)):
    for name, seqs in CASES:
        B = len(seqs)
        lens = torch.tensor([s for s, _ in seqs], dtype=torch.int32)
        qlens = [n for _, n in seqs]
        T = sum(qlens)
        pages = [(int(s) + BS - 1) // BS for s in lens]
        nb = sum(pages) + 8
        perm = torch.randperm(nb, generator=torch.Generator().manual_seed(2))[: sum(pages)].to(torch.int32)
        bt = torch.zeros(B, max(pages), dtype=torch.int32)
        o = 0
        for b, p in enumerate(pages):
            bt[b, :p] = perm[o:o + p]
            o += p
        dt = "fp8" if fp8 else "auto"
        kv = layout.alloc_kv_cache(nb, BS, HKV, D, dt, "cuda", zero=False)
        if fp8:
            kv.copy_(torch.randint(0, 120, kv.shape, dtype=torch.uint8, device="cuda"))
        else:
            kv.normal_()
        q = torch.randn(T, HQ, D, device="cuda").to(torch.bfloat16)
        qsl = torch.tensor([0] + list(torch.tensor(qlens).cumsum(0)), dtype=torch.int32).cuda()
        ks = vs = torch.ones(1, dtype=torch.float32, device="cuda") if fp8 else None
        bt_d, lens_d = bt.cuda(), lens.cuda()
        out = torch.empty_like(q)
        flops = 4.0 * D * HQ * sum(pairs(int(s), n) for s, n in seqs)

        staging = ops.PrefillStaging("cuda") if fp8 else None  # fp8 on the tcgen05 kernel: bf16 staging of the pages

        def ours():
            ops.paged_prefill(q, kv[0], kv[1], bt_d, lens_d, qsl, max(qlens), D ** -0.5, out=out, kv_cache_dtype=dt,
                              k_scale=ks, v_scale=vs, staging=staging, max_seq_len=int(lens.max()),
                              total_kv_tokens=int(lens.sum()))

        def fa():
            from vllm.vllm_flash_attn import flash_attn_varlen_func
            flash_attn_varlen_func(q, kv[0], kv[1], max_seqlen_q=max(qlens), cu_seqlens_q=qsl,
                                   max_seqlen_k=int(lens.max()), seqused_k=lens_d, softmax_scale=D ** -0.5,
                                   causal=True, block_table=bt_d, out=out2, fa_version=ver)

        def ours_umma():
            os.environ["B200ATTN_PREFILL_UMMA"] = "1"
            try:
                ours()
            finally:
                os.environ["B200ATTN_PREFILL_UMMA"] = "0"

        os.environ["B200ATTN_PREFILL_UMMA"] = "0"
        os.environ["B200ATTN_PREFILL_FP8_STAGED"] = "force"  # time the staged path on every case (default: a rule)
        impls = [("b200attn paged_prefill (mma.sync)", ours)]
        impls.append(("b200attn paged_prefill (tcgen05/TMEM" + (", fp8 pages staged as bf16 per call)" if fp8 else ")"),
                      ours_umma))
        if not fp8:
            out2 = torch.empty_like(q)
            for ver in (2, 4):
                impls.append((f"flash-attn v{ver} paged varlen", fa))
        ref = None
        for iname, fn in impls:
            if "flash-attn" in iname:
                ver = int(iname.split("v")[1][0])
            try:
                for _ in range(2):
                    fn()
                torch.cuda.synchronize()
                ts = []
                for _ in range(5):
                    a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record(); fn(); b_.record(); torch.cuda.synchronize()
                    ts.append(a.elapsed_time(b_))
                ms = sorted(ts)[2]
                res = out if fn in (ours, ours_umma) else out2
                if ref is None:
                    ref = res.float().clone()
                print(json.dumps({"case": name, "kv": "fp8" if fp8 else "bf16", "impl": iname, "ms": round(ms, 3),
                                  "TFLOPs": round(flops / ms / 1e9, 1), "tokens": T,
                                  "max_abs_diff_vs_ours": (res.float() - ref).abs().max().item()}), flush=True)
            except Exception as e:
                print(json.dumps({"case": name, "impl": iname, "unavailable": f"{type(e).__name__}: {str(e)[:200]}"}),
                      flush=True)
        del kv
