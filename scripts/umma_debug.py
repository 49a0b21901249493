#!/usr/bin/env python
"""Quick look at the tcgen05 prefill kernel against the oracle: error statistics and where the errors sit."""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
os.environ["B200ATTN_PREFILL_UMMA"] = sys.argv[1] if len(sys.argv) > 1 else "1"
import torch

from oracle import oracle
from tests.helpers import cache_to_dev, make_case
from vllm_omni_b200 import ops

for seq_lens, q_lens, hq, hkv in [([16], [16], 8, 1), ([130], [130], 8, 1), ([300, 129, 48], [44, 129, 17], 8, 1),
                                  ([1000], [1000], 32, 4), ([2049, 700], [513, 699], 28, 4)]:
    c = make_case(seq_lens, hq=hq, hkv=hkv, seed=sum(seq_lens), garbage_fill=True)
    g = torch.Generator().manual_seed(1)
    T = sum(q_lens)
    q = torch.randn(T, hq, 128, generator=g).to(torch.bfloat16)
    qsl = torch.tensor([0] + list(torch.tensor(q_lens).cumsum(0)), dtype=torch.int32)
    cache = cache_to_dev(c.kv_cache)
    out = torch.full((T, hq, 128), 7.0, dtype=torch.bfloat16, device="cuda")
    ops.paged_prefill(q.cuda(), cache[0], cache[1], c.block_table.cuda(), c.seq_lens.cuda(), qsl.cuda(), max(q_lens),
                      c.scale, out=out)
    torch.cuda.synchronize()
    ref = oracle.ref_paged_prefill(q, c.kv_cache[0], c.kv_cache[1], c.block_table, c.seq_lens, qsl, c.scale)
    o = out.float().cpu()
    err = (o - ref).abs()
    bad = err > 2e-2 * ref.abs().max()
    print(f"case {seq_lens} q{q_lens} {hq}/{hkv}: finite={bool(torch.isfinite(o).all())} max_err={err.max():.4f} "
          f"ref_max={ref.abs().max():.3f} bad_frac={bad.float().mean():.4f} "
          f"bad tokens={sorted(set(bad.nonzero()[:, 0].tolist()))[:12]} bad heads={sorted(set(bad.nonzero()[:, 1].tolist()))} "
          f"bad d={sorted(set(bad.nonzero()[:, 2].tolist()))[:16]}", flush=True)
    if bad.any() and T <= 16:
        print("out[0,0,:8]", o[0, 0, :8].tolist(), "ref", ref[0, 0, :8].tolist())
        print("out[5,3,:8]", o[5, 3, :8].tolist(), "ref", ref[5, 3, :8].tolist())
