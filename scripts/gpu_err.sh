#!/bin/bash
timeout 600 python - <<'PY'
import sys, torch
sys.path.insert(0, ".")
from tests.helpers import make_case, row_excess_rel_err, excess_rel_err, rel_l2, kernel_rel_l2, cache_to_dev
from oracle import c_oracle
from vllm_omni_b200 import ops
from vllm.vllm_flash_attn import flash_attn_varlen_func
for lens in ([5, 37, 130, 1, 700, 64, 33], [17, 49, 142, 13, 712, 76, 45], [4096, 2049, 800, 33]):
    c = make_case(lens, hq=16, hkv=2, seed=sum(lens))
    cache = cache_to_dev(c.kv_cache)
    ref, _ = c_oracle.paged_decode(c.q, c.kv_cache[0], c.kv_cache[1], c.block_table, c.seq_lens, c.scale)
    q, bt, sl = c.q.cuda(), c.block_table.cuda(), c.seq_lens.cuda()
    res = {}
    for name, v in (("sk3", ops.VARIANT_SK3), ("sk2", ops.VARIANT_SK2), ("v1", ops.VARIANT_STREAMK), ("fixed", ops.VARIANT_STREAM)):
        out = ops.paged_decode(q, cache[0], cache[1], bt, sl, c.max_seq_len, c.scale, variant=v)
        torch.cuda.synchronize()
        res[name] = out
    B = q.shape[0]
    cu_q = torch.arange(B + 1, dtype=torch.int32, device="cuda")
    out = torch.empty_like(q)
    flash_attn_varlen_func(q, cache[0], cache[1], max_seqlen_q=1, cu_seqlens_q=cu_q, max_seqlen_k=c.max_seq_len, seqused_k=sl,
                           softmax_scale=c.scale, causal=True, block_table=bt, out=out, fa_version=2)
    torch.cuda.synchronize()
    res["fa2"] = out
    print(lens)
    for k, o in res.items():
        print(f"  {k:6s} excess {excess_rel_err(o, ref):.2e} row {row_excess_rel_err(o, ref):.2e} rel_l2 {rel_l2(o, ref):.3e} kernel_l2 {kernel_rel_l2(o, ref):.2e}")
PY
