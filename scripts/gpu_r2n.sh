#!/bin/bash
mkdir -p gpurun_out/golden
timeout 600 python tests/golden/gen_vllm_gpu_golden.py gpurun_out/golden --fp8-only 2>&1 | tail -6
timeout 600 python - <<'PY'
import sys, numpy as np, torch
sys.path.insert(0, ".")
from tests.helpers import make_case, excess_rel_err, kernel_rel_l2, row_excess_rel_err, rel_l2
from oracle import c_oracle
from vllm_omni_b200 import layout, ops
for name in ("decode_fp8_32q4kv", "decode_fp8_16q2kv"):
    z = np.load(f"gpurun_out/golden/vllm_{name}.npz")
    hq, hkv = int(z["hq"]), int(z["hkv"])
    c = make_case(z["seq_lens"].tolist(), hq=hq, hkv=hkv, seed=int(z["seed"]))
    ks = torch.full((1,), float(z["k_scale"]), device="cuda"); vs = torch.full((1,), float(z["v_scale"]), device="cuda")
    kv = layout.alloc_kv_cache(c.num_blocks, 16, hkv, 128, "fp8", "cuda")
    ops.reshape_and_cache(c.hist_k.cuda(), c.hist_v.cuda(), kv[0], kv[1], c.slots.cuda(), "fp8", ks, vs)
    want = torch.from_numpy(z["cache_vllm_fp8"].copy())
    print(name, "cache bytes equal vLLM's:", bool(torch.equal(kv.cpu().view(torch.uint8).contiguous(), want)))
    out = ops.paged_decode(c.q.cuda(), kv[0], kv[1], c.block_table.cuda(), c.seq_lens.cuda(), c.max_seq_len, c.scale,
                           kv_cache_dtype="fp8", k_scale=ks, v_scale=vs)
    torch.cuda.synchronize()
    ref, _ = c_oracle.paged_decode(c.q, want[0], want[1], c.block_table, c.seq_lens, c.scale, True, ks.cpu(), vs.cpu())
    print(" ours vs oracle(dequantised cache): excess", excess_rel_err(out, ref), "row", row_excess_rel_err(out, ref))
    for key in ("out_trtllm_qbf16", "out_trtllm_qfp8"):
        if key in z.files:
            st = torch.from_numpy(z[key].copy()).view(torch.bfloat16)
            print(" ", key, "vs oracle: excess", excess_rel_err(st, ref), "rel_l2", rel_l2(st, ref), "| vs ours: max abs",
                  float((st.float() - out.cpu().float()).abs().max()), "of max", float(ref.abs().max()))
PY
timeout 900 python -m pytest tests/test_gpu_decode.py -x -q -m gpu --timeout 300 2>&1 | tail -4
timeout 300 python - <<'PY'
import sys, torch
sys.path.insert(0, ".")
from tests.helpers import make_case, row_excess_rel_err, excess_rel_err, cache_to_dev, to_dev
from oracle import c_oracle
from vllm_omni_b200 import ops, workloads
worst = 0.0
for name, fp8 in (("cfg2", False), ("cfg3", False), ("cfg4", True)):
    wl = workloads.WORKLOADS[name]
    c = make_case([wl.ctx, wl.ctx - 5, 777, 33], hq=wl.total_q_heads, hkv=wl.total_kv_heads, seed=7, fp8=fp8)
    cache = cache_to_dev(c.kv_cache)
    out = ops.paged_decode(c.q.cuda(), cache[0], cache[1], c.block_table.cuda(), c.seq_lens.cuda(), c.max_seq_len, c.scale,
                           kv_cache_dtype=c.kv_cache_dtype, k_scale=to_dev(c.k_scale), v_scale=to_dev(c.v_scale))
    torch.cuda.synchronize()
    ref, _ = c_oracle.paged_decode(c.q, c.kv_cache[0], c.kv_cache[1], c.block_table, c.seq_lens, c.scale, c.fp8, c.k_scale, c.v_scale)
    r, e = row_excess_rel_err(out, ref), excess_rel_err(out, ref)
    worst = max(worst, r)
    print(f"{name}: tensor-wide excess_rel_err {e:.3e}, per-row {r:.3e}")
print("worst per-row", worst)
PY
timeout 900 python scripts/splitk_sweep.py cfg5 8 > gpurun_out/sweep_cfg5.json 2> gpurun_out/sweep_cfg5.err; echo "sweep rc=$?"; cut -c1-220 gpurun_out/sweep_cfg5.json; tail -3 gpurun_out/sweep_cfg5.err
