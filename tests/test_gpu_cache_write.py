"""GPU parity: KV-cache write through the C ABI vs the oracle — bit-exact (bf16 and e4m3)."""
from pathlib import Path

import numpy as np
import pytest
import torch

from oracle import oracle
from tests.helpers import cache_to_dev, make_case, to_dev
from vllm_omni_b200 import layout, ops

pytestmark = pytest.mark.gpu
GOLDEN = Path(__file__).parent / "golden"


def _bits(t: torch.Tensor) -> torch.Tensor:
    return (t.view(torch.int16) if t.dtype == torch.bfloat16 else t.view(torch.uint8)).cpu()


def _same_codes(dev: torch.Tensor, ref: torch.Tensor) -> bool:
    """Bit equality; for e4m3 the two NaN encodings (0x7F / 0xFF) are one value — the sign of a NaN
    is not defined by the conversion and never occurs in a real cache."""
    a, b = _bits(dev), _bits(ref)
    if a.dtype != torch.uint8:
        return torch.equal(a, b)
    nan = ((a & 0x7F) == 0x7F) & ((b & 0x7F) == 0x7F)
    return bool(torch.all((a == b) | nan))


def _write_both(key, value, slots, nb, bs, hkv, d, fp8, ks, vs, pad_rows=0, prefill_garbage=False):
    kind = "fp8_e4m3" if fp8 else "auto"
    ref = layout.alloc_kv_cache(nb, bs, hkv, d, kind, "cpu")
    dev = layout.alloc_kv_cache(nb, bs, hkv, d, kind, "cuda")
    if prefill_garbage:
        ref.view(torch.uint8 if fp8 else torch.int16).fill_(0x5A)
        dev.copy_(ref)
    oracle.ref_reshape_and_cache(key, value, ref[0], ref[1], slots, fp8, ks, vs)
    kd, vd = key.cuda(), value.cuda()
    if pad_rows:  # key/value padded past slot_mapping, as vLLM passes them (flash_attn.py:868-873)
        kd = torch.cat([kd, torch.full((pad_rows, hkv, d), 7.0, dtype=kd.dtype, device="cuda")])
        vd = torch.cat([vd, torch.full((pad_rows, hkv, d), 7.0, dtype=vd.dtype, device="cuda")])
    ops.reshape_and_cache(kd, vd, dev[0], dev[1], slots.cuda(), kind, to_dev(ks), to_dev(vs))
    torch.cuda.synchronize()
    return ref, dev


@pytest.mark.parametrize("fp8", [False, True])
@pytest.mark.parametrize("hkv,bs", [(4, 16), (1, 16), (2, 32), (8, 64)])
def test_cache_write_bit_exact(lib, fp8, hkv, bs):
    c = make_case([1, 17, 40, 0, 16, 33], hq=hkv, hkv=hkv, bs=bs, fp8=fp8, seed=hkv + bs)
    slots = c.slots.clone()
    slots[::5] = -1
    ref, dev = _write_both(c.hist_k, c.hist_v, slots, c.num_blocks, bs, hkv, 128, fp8, c.k_scale,
                           c.v_scale, pad_rows=3, prefill_garbage=True)
    assert torch.equal(_bits(dev), _bits(ref))


def test_cache_write_from_fused_qkv_slices(lib):
    """key/value are strided views of the qkv projection output (qwen2_old.py:146-156)."""
    T, hq, hkv, d = 37, 8, 2, 128
    g = torch.Generator().manual_seed(2)
    qkv = torch.randn(T, (hq + 2 * hkv) * d, generator=g).to(torch.bfloat16)
    q, k, v = qkv.split([hq * d, hkv * d, hkv * d], dim=-1)
    k, v = k.view(T, hkv, d), v.view(T, hkv, d)
    assert not k.is_contiguous()
    slots = torch.randperm(5 * 16, generator=g)[:T]
    ref = layout.alloc_kv_cache(5, 16, hkv, d, "auto", "cpu")
    oracle.ref_reshape_and_cache(k, v, ref[0], ref[1], slots)
    dev = layout.alloc_kv_cache(5, 16, hkv, d, "auto", "cuda")
    qkv_d = qkv.cuda()
    _, kd, vd = qkv_d.split([hq * d, hkv * d, hkv * d], dim=-1)
    ops.reshape_and_cache(kd.view(T, hkv, d), vd.view(T, hkv, d), dev[0], dev[1], slots.cuda())
    assert torch.equal(_bits(dev), _bits(ref))


@pytest.mark.parametrize("per_head", [False, True])
def test_fp8_rounding_known_answers_on_device(lib, per_head):
    """0, subnormals, +-448, out-of-range, ties, NaN with scales {1, 0.5, 3}."""
    hkv, d = 3, 128
    vals = torch.tensor([0.0, -0.0, 2.0 ** -9, -(2.0 ** -9), 2.0 ** -10, 3 * 2.0 ** -10, 2.0 ** -6, 1.0,
                         1.0625, 1.1875, 1.9375, 240.0, 448.0, -448.0, 449.0, 464.0, 480.0, 1e4, -1e4,
                         float("inf"), float("-inf"), float("nan"), 0.3, -3.0, 224.0, 225.0, 1344.0,
                         1345.0, 1e-3, 6.1e-5], dtype=torch.float32)
    key = vals.repeat((d + len(vals) - 1) // len(vals))[:d].view(1, 1, d).repeat(4, hkv, 1)
    key = key * torch.tensor([1.0, -1.0, 0.5, 2.0]).view(4, 1, 1)
    key = key.to(torch.bfloat16)
    sc = torch.tensor([1.0, 0.5, 3.0]) if per_head else torch.tensor([0.5])
    slots = torch.tensor([3, 18, 0, 31])
    ref, dev = _write_both(key, key.flip(0).contiguous(), slots, 2, 16, hkv, d, True, sc, sc * 1.5)
    assert _same_codes(dev, ref)
    # spot-check the semantics, not only the agreement: head 0 scale 1 (per_head) row of +vals
    if per_head:
        row = dev[0][0, 3, 0].cpu().numpy()
        want = oracle.e4m3_encode_satfinite(key[0, 0].to(torch.float32).numpy())
        nan = (want & 0x7F) == 0x7F
        assert np.array_equal(row[~nan], want[~nan]) and np.all((row[nan] & 0x7F) == 0x7F)
        assert row[12] == 0x7E and row[14] == 0x7E and row[17] == 0x7E and (row[21] & 0x7F) == 0x7F


@pytest.mark.parametrize("name", ["bf16", "fp8"])
def test_reference_reader_golden_through_device_write_and_gather(lib, name):
    """Device write -> device gather must reproduce what the reference's own reader extracted
    (tests/golden/kv_reader_*.npz, produced by OmniKVTransferManager._extract_kv_cache)."""
    z = np.load(GOLDEN / f"kv_reader_{name}.npz")
    fp8 = name == "fp8"
    nb, bs, hkv, d = (int(v) for v in z["shape"])
    key = torch.from_numpy(z["key"]).view(torch.bfloat16).cuda()
    value = torch.from_numpy(z["value"]).view(torch.bfloat16).cuda()
    ks = torch.from_numpy(z["k_scale"]).cuda() if fp8 else None
    vs = torch.from_numpy(z["v_scale"]).cuda() if fp8 else None
    kind = "fp8_e4m3" if fp8 else "auto"
    cache = layout.alloc_kv_cache(nb, bs, hkv, d, kind, "cuda")
    ops.reshape_and_cache(key, value, cache[0], cache[1], torch.from_numpy(z["slots"]).cuda(), kind, ks, vs)
    ids = torch.from_numpy(z["block_ids"]).cuda()
    S = int(z["seq_len"])
    gk, gv = ops.gather_kv(cache[0], ids, S), ops.gather_kv(cache[1], ids, S)
    view = (lambda t: t.cpu().numpy()) if fp8 else (lambda t: t.view(torch.int16).cpu().numpy())
    assert np.array_equal(view(gk), z["extracted_key"])
    assert np.array_equal(view(gv), z["extracted_value"])
    # and the reference's indexing expression itself on the device tensor
    k_ref = cache[0][ids.long()].flatten(0, 1)[:S].contiguous()
    assert torch.equal(gk, k_ref)


def test_cache_write_matches_vllm_kernel_when_available(lib):
    """Same inputs through vLLM's reshape_and_cache_flash (the kernel this replaces) — bit-exact."""
    try:
        from vllm import _custom_ops as vops
        torch.ops._C_cache_ops.reshape_and_cache_flash  # noqa: B018
    except Exception as e:  # pragma: no cover
        pytest.skip(f"vLLM _C not loadable here: {e}")
    T, hkv, d, nb, bs = 257, 4, 128, 40, 16
    g = torch.Generator().manual_seed(7)
    k = (torch.randn(T, hkv, d, generator=g) * 3).to(torch.bfloat16).cuda()
    v = (torch.randn(T, hkv, d, generator=g) * 3).to(torch.bfloat16).cuda()
    slots = torch.randperm(nb * bs, generator=g)[:T].cuda()
    slots[::9] = -1
    for kind, ks, vs in (("auto", torch.ones(1), torch.ones(1)),
                         ("fp8_e4m3", torch.tensor([0.02]), torch.tensor([0.03]))):
        ks, vs = ks.cuda(), vs.cuda()
        dt = torch.uint8 if kind != "auto" else torch.bfloat16
        theirs = torch.zeros(2, nb, bs, hkv, d, dtype=dt, device="cuda")   # vLLM NHD layout
        ours = torch.zeros_like(theirs)
        vops.reshape_and_cache_flash(k, v, theirs[0], theirs[1], slots, "fp8" if kind != "auto" else "auto", ks, vs)
        ops.reshape_and_cache(k, v, ours[0], ours[1], slots, kind, ks, vs)
        assert torch.equal(_bits(ours), _bits(theirs)), kind


def test_cache_write_prefill_sized(lib):
    """32k-token chunk (max_num_batched_tokens of the reference stage config): property check —
    every written slot reads back the token, every other slot is untouched."""
    T, hkv, d, bs = 32768, 4, 128, 16
    nb = T // bs + 7
    g = torch.Generator(device="cuda").manual_seed(3)
    k = torch.randn(T, hkv, d, generator=g, device="cuda").to(torch.bfloat16)
    v = torch.randn(T, hkv, d, generator=g, device="cuda").to(torch.bfloat16)
    slots = torch.randperm(nb * bs, generator=g, device="cuda")[:T]
    cache = layout.alloc_kv_cache(nb, bs, hkv, d, "auto", "cuda")
    ops.reshape_and_cache(k, v, cache[0], cache[1], slots)
    flat_k = cache[0].reshape(nb * bs, hkv, d)
    flat_v = cache[1].reshape(nb * bs, hkv, d)
    assert torch.equal(flat_k[slots], k) and torch.equal(flat_v[slots], v)
    mask = torch.ones(nb * bs, dtype=torch.bool, device="cuda")
    mask[slots] = False
    assert torch.count_nonzero(flat_k[mask]) == 0 and torch.count_nonzero(flat_v[mask]) == 0
