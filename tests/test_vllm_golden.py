"""CPU: the oracle pinned against outputs of the reference's own stock kernels.

The arithmetic of this path lives in vLLM, whose compiled kernels cannot run in the build
container. `tests/golden/gen_vllm_gpu_golden.py` ran them on a B200 box — vLLM's `reshape_and_cache_flash`
for the slot writes, the FlashAttention-2 paged varlen call of the FLASH_ATTN backend and FlashInfer's
TRT-LLM-gen decode of the FLASHINFER backend for the attention output — and the committed
`tests/golden/vllm_decode_*.npz` hold their complete inputs and outputs. Here both restatements of the
oracle (Python and C) are checked against them: slot writes bit-exact, attention output within the
north_star's 2e-3 (beyond the bf16 rounding of the stock kernel's own output)."""
from pathlib import Path

import numpy as np
import pytest
import torch

from oracle import c_oracle, oracle
from tests.helpers import excess_rel_err, kernel_rel_l2, make_case

GOLDEN = Path(__file__).resolve().parent / "golden"
FILES = ["vllm_decode_32q4kv.npz", "vllm_decode_28q4kv.npz", "vllm_decode_16q2kv.npz"]
STOCK = ["out_fa2", "out_trtllm"]


def bf16(a: np.ndarray) -> torch.Tensor:
    return torch.from_numpy(a.copy()).view(torch.bfloat16)


def load(name):
    p = GOLDEN / name
    assert p.exists(), f"{p} missing: run tests/golden/gen_vllm_gpu_golden.py on a GPU box with vLLM"
    z = np.load(p)
    lens = z["seq_lens"].tolist()
    c = make_case(lens, hq=int(z["hq"]), hkv=int(z["hkv"]), seed=int(z["seed"]))   # the generator's inputs, from the seed
    return z, c


@pytest.mark.parametrize("name", FILES)
def test_slot_writes_equal_vllm_reshape_and_cache_flash(name):
    z, c = load(name)
    assert torch.equal(bf16(z["q"]), c.q) and np.array_equal(z["block_table"], c.block_table.numpy())
    want = bf16(z["cache_vllm"])                       # [2, nb, bs, Hkv, d] written by vLLM's CUDA kernel
    nb, bs, hkv, d = want.shape[1:]
    for impl in (oracle.ref_reshape_and_cache, c_oracle.reshape_and_cache):
        kv = torch.zeros(2, nb, bs, hkv, d, dtype=torch.bfloat16)
        impl(c.hist_k, c.hist_v, kv[0], kv[1], c.slots)
        assert torch.equal(kv.view(torch.int16), want.view(torch.int16)), impl.__module__
    # and the cache make_case() builds (this repo's physical layout, same logical tensor) holds the same rows
    assert torch.equal(c.kv_cache.contiguous().view(torch.int16), want.view(torch.int16))


@pytest.mark.parametrize("name", FILES)
def test_decode_oracle_within_tolerance_of_vllm_stock_kernels(name):
    z, c = load(name)
    kv = bf16(z["cache_vllm"])
    ref_py, _ = oracle.ref_paged_decode(c.q, kv[0], kv[1], c.block_table, c.seq_lens, float(z["scale"]))
    ref_c, _ = c_oracle.paged_decode(c.q, kv[0], kv[1], c.block_table, c.seq_lens, float(z["scale"]))
    present = [k for k in STOCK if k in z.files]
    assert present, "no stock-kernel output in the fixture"
    for key in present:
        stock = bf16(z[key])                           # [B, Hq, d] bf16 from the stock kernel
        for ref in (ref_py, ref_c):
            e, l2 = excess_rel_err(stock, ref), kernel_rel_l2(stock, ref)
            assert e <= 2e-3 and l2 <= 2e-3, (name, key, e, l2)   # tolerance of north_star, stated here
    if len(present) == 2:                              # the two stock kernels agree with each other to bf16 rounding
        a, b = bf16(z[present[0]]).float(), bf16(z[present[1]]).float()
        assert float((a - b).abs().max()) <= 2 ** -7 * float(a.abs().max())


PREFILL_FILES = ["vllm_prefill_32q4kv.npz", "vllm_prefill_16q2kv.npz"]


@pytest.mark.parametrize("name", PREFILL_FILES)
def test_prefill_oracle_within_tolerance_of_vllm_flash_attn(name):
    """The causal varlen oracle (`ref_paged_prefill`) against vLLM's FLASH_ATTN call for max_query_len > 1
    (FlashAttention-2 paged varlen, flash_attn.py:758-819) on prefill / extend / single-token rows."""
    p = GOLDEN / name
    assert p.exists(), f"{p} missing: run tests/golden/gen_vllm_gpu_golden.py on a GPU box with vLLM"
    z = np.load(p)
    c = make_case(z["seq_lens"].tolist(), hq=int(z["hq"]), hkv=int(z["hkv"]), seed=int(z["seed"]))
    kv = bf16(z["cache_vllm"])
    assert torch.equal(c.kv_cache.contiguous().view(torch.int16), kv.view(torch.int16))   # same cache as vLLM wrote
    q = bf16(z["q"])
    qsl = torch.from_numpy(z["query_start_loc"].copy())
    ref = oracle.ref_paged_prefill(q, kv[0], kv[1], c.block_table, c.seq_lens, qsl, float(z["scale"]))
    stock = bf16(z["out_fa2"])
    e, l2 = excess_rel_err(stock, ref), kernel_rel_l2(stock, ref)
    assert e <= 2e-3 and l2 <= 2e-3, (name, e, l2)


# ------------------------------------------------------------------------------------------------
# fp8-e4m3 KV cache, per-tensor scales (what the stock fp8 decode can express): the cache bytes written by vLLM's fp8
# `reshape_and_cache_flash`, and FlashInfer's TRT-LLM-gen decode (bf16 query) reading them.
# ------------------------------------------------------------------------------------------------
FP8_FILES = ["vllm_decode_fp8_32q4kv.npz", "vllm_decode_fp8_16q2kv.npz"]


def load_fp8(name):
    p = GOLDEN / name
    assert p.exists(), f"{p} missing: run tests/golden/gen_vllm_gpu_golden.py --fp8-only on a GPU box with vLLM"
    z = np.load(p)
    c = make_case(z["seq_lens"].tolist(), hq=int(z["hq"]), hkv=int(z["hkv"]), seed=int(z["seed"]))
    ks = torch.full((1,), float(z["k_scale"]), dtype=torch.float32)
    vs = torch.full((1,), float(z["v_scale"]), dtype=torch.float32)
    return z, c, ks, vs


@pytest.mark.parametrize("name", FP8_FILES)
def test_fp8_slot_writes_equal_vllm_reshape_and_cache_flash(name):
    """e4m3_satfinite(x / scale) of both oracle restatements, byte for byte against vLLM's CUDA kernel."""
    z, c, ks, vs = load_fp8(name)
    want = torch.from_numpy(z["cache_vllm_fp8"].copy())       # [2, nb, bs, Hkv, d] uint8
    nb, bs, hkv, d = want.shape[1:]
    for impl in (oracle.ref_reshape_and_cache, c_oracle.reshape_and_cache):
        kv = torch.zeros(2, nb, bs, hkv, d, dtype=torch.uint8)
        impl(c.hist_k, c.hist_v, kv[0], kv[1], c.slots, True, ks, vs)
        assert torch.equal(kv, want), impl.__module__


@pytest.mark.parametrize("name", FP8_FILES)
def test_fp8_decode_oracle_within_tolerance_of_trtllm_gen(name):
    """The oracle fed the dequantised fp8 cache against the stock fp8 decode (bf16 query, scalar k/v scales folded
    into bmm1 / bmm2 exactly as vLLM's FLASHINFER backend passes them): the FP8 tolerance of this repo's tests —
    2e-3 against the dequantised cache — is one the stock kernel meets too."""
    z, c, ks, vs = load_fp8(name)
    assert "out_trtllm_qbf16" in z.files
    kv = torch.from_numpy(z["cache_vllm_fp8"].copy())
    stock = bf16(z["out_trtllm_qbf16"])
    for ref in (oracle.ref_paged_decode(c.q, kv[0], kv[1], c.block_table, c.seq_lens, float(z["scale"]), True, ks, vs)[0],
                c_oracle.paged_decode(c.q, kv[0], kv[1], c.block_table, c.seq_lens, float(z["scale"]), True, ks, vs)[0]):
        e, l2 = excess_rel_err(stock, ref), kernel_rel_l2(stock, ref)
        assert e <= 2e-3 and l2 <= 2e-3, (name, e, l2)
