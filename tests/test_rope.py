"""Rotary embedding: the oracle restatement against golden vectors produced by vLLM's own
RotaryEmbedding / MRotaryEmbedding.forward_native (tests/golden/gen_rope_golden.py) — bit-exact — and,
on the GPU, the fused RoPE + cache-write kernel against the oracle and the same golden vectors."""
from pathlib import Path

import numpy as np
import pytest
import torch

from oracle import oracle

GOLD = Path(__file__).resolve().parent / "golden"
NAMES = ["neox", "gptj", "partial", "mrope", "mrope_interleaved", "mrope_text_only"]


def bf16(a):
    return torch.from_numpy(np.ascontiguousarray(a)).view(torch.bfloat16)


def load(name):
    z = np.load(GOLD / f"rope_{name}.npz")
    pos = torch.from_numpy(z["positions"])
    sec = [int(x) for x in z["mrope_section"]]
    return dict(q=bf16(z["q"]), k=bf16(z["k"]), q_out=bf16(z["q_out"]), k_out=bf16(z["k_out"]), pos=pos,
                cs=bf16(z["cos_sin_cache"]), rot=int(z["rot_dim"]), neox=bool(z["is_neox"]),
                sec=sec if pos.ndim == 2 else None, inter=bool(z["mrope_interleaved"]))


@pytest.mark.parametrize("name", NAMES)
def test_oracle_rope_matches_vllm_golden_bit_exact(name):
    g = load(name)
    T = g["pos"].shape[-1]
    q, k = g["q"].view(T, -1, 128), g["k"].view(T, -1, 128)
    qo, ko = oracle.ref_rope(g["pos"], q, k, g["cs"], g["rot"], g["neox"], g["sec"], g["inter"])
    assert torch.equal(qo.reshape(T, -1).view(torch.int16), g["q_out"].view(torch.int16))
    assert torch.equal(ko.reshape(T, -1).view(torch.int16), g["k_out"].view(torch.int16))


@pytest.mark.gpu
@pytest.mark.parametrize("fp8", [False, True])
@pytest.mark.parametrize("name", NAMES)
def test_fused_rope_cache_write_golden(lib, name, fp8):
    """Kernel on the golden inputs: rotated q/k bit-exact vs vLLM's vectors; cache bytes bit-exact vs the
    oracle's cache write of the rotated key."""
    from vllm_omni_b200 import layout, ops
    g = load(name)
    T = g["pos"].shape[-1]
    hkv = g["k"].numel() // T // 128
    q, k = g["q"].view(T, -1, 128).clone().cuda(), g["k"].view(T, -1, 128).clone().cuda()
    v = torch.randn(T, hkv, 128, generator=torch.Generator().manual_seed(3)).to(torch.bfloat16)
    dt = "fp8" if fp8 else "auto"
    nb, bs = 5, 16
    slots = torch.tensor([3, 17, -1, 40, 41, 79, 0, 64, 65], dtype=torch.int64)[:T]
    ks = torch.tensor([0.5, 2.0][:hkv], dtype=torch.float32) if fp8 else None
    vs = torch.tensor([1.5, 0.25][:hkv], dtype=torch.float32) if fp8 else None
    kv = layout.alloc_kv_cache(nb, bs, hkv, 128, dt, "cuda")
    ops.rope_and_cache(g["pos"].cuda(), q, k, v.cuda(), g["cs"].cuda(), g["neox"], kv[0], kv[1], slots.cuda(), dt,
                       None if ks is None else ks.cuda(), None if vs is None else vs.cuda(), rot_dim=g["rot"],
                       mrope_section=g["sec"], mrope_interleaved=g["inter"])
    torch.cuda.synchronize()
    assert torch.equal(q.cpu().reshape(T, -1).view(torch.int16), g["q_out"].view(torch.int16))
    assert torch.equal(k.cpu().reshape(T, -1).view(torch.int16), g["k_out"].view(torch.int16))
    ref = layout.alloc_kv_cache(nb, bs, hkv, 128, dt, "cpu")
    oracle.ref_reshape_and_cache(g["k_out"].view(T, hkv, 128), v, ref[0], ref[1], slots, fp8, ks, vs)
    assert torch.equal(kv.cpu().view(torch.uint8 if fp8 else torch.int16), ref.view(torch.uint8 if fp8 else torch.int16))


@pytest.mark.gpu
@pytest.mark.parametrize("fp8", [False, True])
def test_fused_rope_cache_write_decode_sized_step_from_fused_qkv(lib, fp8):
    """cfg3-shaped decode step (64 tokens, 32q/4kv): q/k/v are slices of one fused qkv buffer (strided),
    slot_mapping shorter than the padded token count; equals unfused oracle RoPE + oracle cache write."""
    from vllm_omni_b200 import layout, ops
    T, pad, hq, hkv, d = 64, 8, 32, 4, 128
    g = torch.Generator().manual_seed(11)
    fused = torch.randn(T + pad, (hq + 2 * hkv) * d, generator=g).to(torch.bfloat16)
    pos = torch.randint(0, 8192, (T + pad,), generator=g)
    cs = oracle.make_cos_sin_cache(8192, d)
    qc, kc, vc = fused.split([hq * d, hkv * d, hkv * d], dim=-1)
    q_ref, k_ref = oracle.ref_rope(pos, qc.reshape(T + pad, hq, d), kc.reshape(T + pad, hkv, d), cs, d, True)
    nb, bs = 40, 16
    slots = torch.randperm(nb * bs, generator=g)[:T].to(torch.int64)
    slots[5] = -1
    dt = "fp8" if fp8 else "auto"
    ks = torch.rand(hkv, generator=g) + 0.5 if fp8 else None
    vs = torch.rand(hkv, generator=g) + 0.5 if fp8 else None
    ref = layout.alloc_kv_cache(nb, bs, hkv, d, dt, "cpu")
    oracle.ref_reshape_and_cache(k_ref[:T], vc.reshape(T + pad, hkv, d)[:T], ref[0], ref[1], slots, fp8, ks, vs)
    dev = fused.cuda()
    qd, kd, vd = dev.split([hq * d, hkv * d, hkv * d], dim=-1)
    kv = layout.alloc_kv_cache(nb, bs, hkv, d, dt, "cuda")
    ops.rope_and_cache(pos.cuda(), qd.view(T + pad, hq, d), kd.view(T + pad, hkv, d), vd.view(T + pad, hkv, d),
                       cs.cuda(), True, kv[0], kv[1], slots.cuda(), dt, None if ks is None else ks.cuda(),
                       None if vs is None else vs.cuda())
    torch.cuda.synchronize()
    out = dev.cpu()
    qo, ko, vo = out.split([hq * d, hkv * d, hkv * d], dim=-1)
    assert torch.equal(qo.reshape(T + pad, hq, d), q_ref) and torch.equal(ko.reshape(T + pad, hkv, d), k_ref)
    assert torch.equal(vo, vc)
    assert torch.equal(kv.cpu().view(torch.uint8 if fp8 else torch.int16), ref.view(torch.uint8 if fp8 else torch.int16))
