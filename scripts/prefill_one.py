#!/usr/bin/env python
"""One prefill shape, a few launches (for ncu): python scripts/prefill_one.py [ctx] [new_tokens] [batch]"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch

from vllm_omni_b200 import layout, ops

S, N, B = (int(sys.argv[1]) if len(sys.argv) > 1 else 8192, int(sys.argv[2]) if len(sys.argv) > 2 else 8192,
           int(sys.argv[3]) if len(sys.argv) > 3 else 1)
HQ, HKV, D, BS = 32, 4, 128, 16
pages = (S + BS - 1) // BS
nb = B * pages + 8
perm = torch.randperm(nb, generator=torch.Generator().manual_seed(2))[: B * pages].to(torch.int32)
bt = perm.view(B, pages).cuda()
kv = layout.alloc_kv_cache(nb, BS, HKV, D, "auto", "cuda", zero=False)
kv.normal_()
q = torch.randn(B * N, HQ, D, device="cuda").to(torch.bfloat16)
qsl = (torch.arange(B + 1, dtype=torch.int32) * N).cuda()
lens = torch.full((B,), S, dtype=torch.int32, device="cuda")
out = torch.empty_like(q)
for _ in range(4):
    ops.paged_prefill(q, kv[0], kv[1], bt, lens, qsl, N, D ** -0.5, out=out)
torch.cuda.synchronize()
print("done", float(out.float().abs().mean()))
