"""CPU: pins the oracle (both restatements) against golden vectors before anything trusts it."""
import json
import math
from pathlib import Path

import numpy as np
import pytest
import torch

from oracle import c_oracle, oracle
from tests.helpers import make_case
from vllm_omni_b200 import layout

GOLDEN = Path(__file__).parent / "golden"


def _kat():
    vs = json.loads((GOLDEN / "e4m3_kat.json").read_text())["vectors"]
    x = np.array([float.fromhex(v["value"]) if v["value"] not in ("inf", "-inf", "nan")
                  else float(v["value"]) for v in vs], dtype=np.float32)
    return x, np.array([v["code"] for v in vs], dtype=np.uint8)


def test_e4m3_known_answers_both_oracles():
    x, want = _kat()
    assert np.array_equal(oracle.e4m3_encode_satfinite(x), want)
    assert np.array_equal(c_oracle.e4m3_encode(x), want)


def test_e4m3_encode_matches_torch_cast_in_finite_range():
    # every bf16 bit pattern (the cache-write input type) with |x| <= 448, plus random fp32 quotients
    bits = torch.arange(0, 65536, dtype=torch.int32).to(torch.int16)
    x = bits.view(torch.bfloat16).to(torch.float32)
    x = x[torch.isfinite(x) & (x.abs() <= 448.0)]
    g = torch.Generator().manual_seed(0)
    r = torch.randn(200_000, generator=g) * 3.0 / torch.rand(200_000, generator=g).clamp_min(0.01)
    r = r[r.abs() <= 448.0]
    for v in (x, r, r * 1e-3):
        want = v.to(torch.float8_e4m3fn).view(torch.uint8).numpy()
        assert np.array_equal(oracle.e4m3_encode_satfinite(v.numpy()), want)
        assert np.array_equal(c_oracle.e4m3_encode(v.numpy()), want)


def test_e4m3_saturates_where_torch_gives_nan():
    x = np.array([465.0, 1e6, -1e6, np.inf], dtype=np.float32)
    assert list(oracle.e4m3_encode_satfinite(x)) == [0x7E, 0x7E, 0xFE, 0x7E]
    assert list(c_oracle.e4m3_encode(x)) == [0x7E, 0x7E, 0xFE, 0x7E]


def test_e4m3_decode_table_matches_torch_and_c():
    codes = np.arange(256, dtype=np.uint8)
    want = torch.from_numpy(codes).view(torch.float8_e4m3fn).to(torch.float32).numpy()
    for got in (oracle.e4m3_decode(codes), c_oracle.e4m3_decode(codes)):
        assert np.array_equal(np.isnan(got), np.isnan(want))
        assert np.array_equal(got[~np.isnan(want)], want[~np.isnan(want)])
    # round trip of every finite code
    fin = codes[~np.isnan(want)]
    assert np.array_equal(oracle.e4m3_encode_satfinite(want[~np.isnan(want)]) & 0x7F, fin & 0x7F)


@pytest.mark.parametrize("fp8", [False, True])
def test_cache_write_python_vs_c_and_slot_semantics(fp8):
    c = make_case([1, 17, 40, 0, 16], hq=8, hkv=2, fp8=fp8, seed=3)
    slots = c.slots.clone()
    slots[::7] = -1  # padding tokens are skipped
    a = layout.alloc_kv_cache(c.num_blocks, 16, 2, 128, c.kv_cache_dtype, "cpu")
    b = layout.alloc_kv_cache(c.num_blocks, 16, 2, 128, c.kv_cache_dtype, "cpu")
    oracle.ref_reshape_and_cache(c.hist_k, c.hist_v, a[0], a[1], slots, fp8, c.k_scale, c.v_scale)
    c_oracle.reshape_and_cache(c.hist_k, c.hist_v, b[0], b[1], slots, fp8, c.k_scale, c.v_scale)
    assert torch.equal(a.view(torch.uint8) if fp8 else a.view(torch.int16),
                       b.view(torch.uint8) if fp8 else b.view(torch.int16))
    # a skipped token leaves its slot zero; a written one holds the token
    blk, off = int(c.slots[0]) // 16, int(c.slots[0]) % 16
    assert torch.count_nonzero(a[0][blk, off]) == 0
    blk, off = int(c.slots[1]) // 16, int(c.slots[1]) % 16
    if not fp8:
        assert torch.equal(a[0][blk, off], c.hist_k[1])


@pytest.mark.parametrize("name", ["bf16", "fp8"])
def test_cache_write_layout_against_reference_reader_golden(name):
    """tests/golden/kv_reader_*.npz: what the reference's OmniKVTransferManager._extract_kv_cache
    returned on a cache the oracle wrote in the backend layout (gen_cpu_golden.py)."""
    z = np.load(GOLDEN / f"kv_reader_{name}.npz")
    fp8 = name == "fp8"
    nb, bs, hkv, d = (int(v) for v in z["shape"])
    key = torch.from_numpy(z["key"]).view(torch.bfloat16)
    value = torch.from_numpy(z["value"]).view(torch.bfloat16)
    slots = torch.from_numpy(z["slots"])
    ks = torch.from_numpy(z["k_scale"]) if fp8 else None
    vs = torch.from_numpy(z["v_scale"]) if fp8 else None
    for write in (oracle.ref_reshape_and_cache, c_oracle.reshape_and_cache):
        cache = layout.alloc_kv_cache(nb, bs, hkv, d, "fp8_e4m3" if fp8 else "auto", "cpu")
        write(key, value, cache[0], cache[1], slots, fp8, ks, vs)
        ids, S = z["block_ids"].tolist(), int(z["seq_len"])
        gk = oracle.ref_gather_kv(cache[0], ids, S)
        gv = oracle.ref_gather_kv(cache[1], ids, S)
        view = (lambda t: t.numpy()) if fp8 else (lambda t: t.view(torch.int16).numpy())
        assert np.array_equal(view(gk), z["extracted_key"])
        assert np.array_equal(view(gv), z["extracted_value"])
    if not fp8:  # bf16: the reader must hand back exactly the tokens that were written
        assert np.array_equal(z["extracted_key"], z["key"])


@pytest.mark.parametrize("fp8", [False, True])
@pytest.mark.parametrize("lens", [[1], [16], [17, 1, 33], [5, 0, 64, 31], [300, 129]])
def test_decode_python_vs_c_oracle(lens, fp8):
    c = make_case(lens, hq=8, hkv=2, fp8=fp8, seed=len(lens))
    o1, l1 = oracle.ref_paged_decode(c.q, c.kv_cache[0], c.kv_cache[1], c.block_table, c.seq_lens,
                                     c.scale, fp8, c.k_scale, c.v_scale)
    o2, l2 = c_oracle.paged_decode(c.q, c.kv_cache[0], c.kv_cache[1], c.block_table, c.seq_lens,
                                   c.scale, fp8, c.k_scale, c.v_scale)
    assert torch.allclose(o1, o2, atol=2e-6, rtol=1e-5)
    fin = torch.isfinite(l1)
    assert torch.equal(fin, torch.isfinite(l2))
    assert torch.allclose(l1[fin], l2[fin], atol=1e-5)
    for b, S in enumerate(lens):
        if S == 0:
            assert torch.count_nonzero(o1[b]) == 0 and torch.all(l1[b] == -math.inf)


def test_decode_hand_checkable_case():
    """d-dimensional one-hot keys make the softmax weights explicit."""
    hq, hkv, d, bs = 2, 1, 128, 16
    S = 3
    cache = layout.alloc_kv_cache(2, bs, hkv, d, "auto", "cpu")
    k = torch.zeros(S, hkv, d)
    v = torch.zeros(S, hkv, d)
    for j in range(S):
        k[j, 0, j] = 1.0
        v[j, 0, :] = float(j + 1)
    oracle.ref_reshape_and_cache(k.to(torch.bfloat16), v.to(torch.bfloat16), cache[0], cache[1],
                                 torch.tensor([16, 17, 18]))
    q = torch.zeros(1, hq, d)
    q[0, 0, 0] = 2.0      # head 0 looks at key 0
    q[0, 1, :3] = 1.0     # head 1 is uniform over the 3 keys
    bt = torch.tensor([[1, 0]], dtype=torch.int32)
    lens = torch.tensor([S], dtype=torch.int32)
    out, lse = oracle.ref_paged_decode(q.to(torch.bfloat16), cache[0], cache[1], bt, lens, 1.0)
    w0 = math.exp(2.0) / (math.exp(2.0) + 2.0)
    want0 = w0 * 1 + (1 - w0) / 2 * (2 + 3)
    assert abs(float(out[0, 0, 5]) - want0) < 1e-5
    assert abs(float(out[0, 1, 5]) - 2.0) < 1e-5
    assert abs(float(lse[0, 1]) - math.log(3 * math.e)) < 1e-5


def test_merge_attn_states_equals_unsplit():
    c = make_case([200], hq=8, hkv=2, seed=9)
    full, lse = oracle.ref_paged_decode(c.q, c.kv_cache[0], c.kv_cache[1], c.block_table, c.seq_lens, c.scale)
    # split the context at a block boundary by presenting two shorter sequences
    lens_a = torch.tensor([96], dtype=torch.int32)
    oa, la = oracle.ref_paged_decode(c.q, c.kv_cache[0], c.kv_cache[1], c.block_table, lens_a, c.scale)
    bt_b = c.block_table[:, 6:]
    ob, lb = oracle.ref_paged_decode(c.q, c.kv_cache[0], c.kv_cache[1], bt_b,
                                     torch.tensor([104], dtype=torch.int32), c.scale)
    mo, ml = oracle.ref_merge_attn_states(torch.stack([oa, ob]), torch.stack([la, lb]))
    assert torch.allclose(mo, full, atol=1e-5)
    assert torch.allclose(ml, lse, atol=1e-5)


def test_prefill_oracle_last_row_equals_decode():
    c = make_case([37, 64], hq=8, hkv=2, seed=4)
    # queries: last 5 tokens of seq 0, last 1 of seq 1
    g = torch.Generator().manual_seed(1)
    q = torch.randn(6, 8, 128, generator=g).to(torch.bfloat16)
    qsl = torch.tensor([0, 5, 6], dtype=torch.int32)
    out = oracle.ref_paged_prefill(q, c.kv_cache[0], c.kv_cache[1], c.block_table, c.seq_lens, qsl, c.scale)
    dec, _ = oracle.ref_paged_decode(torch.stack([q[4], q[5]]), c.kv_cache[0], c.kv_cache[1],
                                     c.block_table, c.seq_lens, c.scale)
    assert torch.allclose(out[4], dec[0], atol=1e-5) and torch.allclose(out[5], dec[1], atol=1e-5)
    # first of the 5 rows sees 33 tokens
    dec33, _ = oracle.ref_paged_decode(q[0:1], c.kv_cache[0], c.kv_cache[1], c.block_table[:1],
                                       torch.tensor([33], dtype=torch.int32), c.scale)
    assert torch.allclose(out[0], dec33[0], atol=1e-5)
