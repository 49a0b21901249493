"""GPU parity: causal varlen attention over the paged cache (prefill / extend / mixed batches,
SURVEY §8f row 1) through b200attn_paged_prefill vs the fp32 oracle (`ref_paged_prefill`, the
causal restatement of vllm/v1/attention/backends/flash_attn.py:758-819). Same 2e-3 bound as decode;
fp8 caches are compared against the oracle fed the same quantised cache."""
import pytest
import torch

from oracle import oracle
from tests.helpers import cache_to_dev, excess_rel_err, kernel_rel_l2, make_case, to_dev
from vllm_omni_b200 import ops

pytestmark = pytest.mark.gpu
TOL = 2e-3


@pytest.fixture(autouse=True, params=["mma_sync", "umma"])
def prefill_kernel(request, monkeypatch):
    """Every case runs on both prefill kernels: the mma.sync one (paged_prefill.cu) and the tcgen05 / TMEM one
    (paged_prefill_umma.cu; fp8 caches reach it through the bf16 staging of b200attn_paged_prefill_fp8_staged).
    The library and ops.paged_prefill read the switch at every call."""
    monkeypatch.setenv("B200ATTN_PREFILL_UMMA", "1" if request.param == "umma" else "0")
    # fp8: always staged in the tcgen05 runs (the default stages only calls with enough new tokens per sequence)
    monkeypatch.setenv("B200ATTN_PREFILL_FP8_STAGED", "force")
    return request.param


def run_prefill(c, q_lens, hq, seed=0, q_std=1.0, total_kv_tokens=0):
    g = torch.Generator().manual_seed(seed)
    T = sum(q_lens)
    q = (torch.randn(T, hq, 128, generator=g) * q_std).to(torch.bfloat16)
    qsl = torch.tensor([0] + list(torch.tensor(q_lens).cumsum(0)), dtype=torch.int32)
    cache = cache_to_dev(c.kv_cache)
    out = torch.full((T, hq, 128), 7.0, dtype=torch.bfloat16, device="cuda")
    staging = ops.PrefillStaging("cuda") if c.fp8 else None
    ops.paged_prefill(q.cuda(), cache[0], cache[1], c.block_table.cuda(), c.seq_lens.cuda(), qsl.cuda(),
                      max(q_lens), c.scale, out=out, kv_cache_dtype=c.kv_cache_dtype, k_scale=to_dev(c.k_scale),
                      v_scale=to_dev(c.v_scale), staging=staging, max_seq_len=c.max_seq_len,
                      total_kv_tokens=total_kv_tokens)
    torch.cuda.synchronize()
    if c.fp8:  # the staged path ran exactly when the tcgen05 kernel is selected
        import os
        assert (staging.buf is not None) == (os.environ["B200ATTN_PREFILL_UMMA"] == "1")
    ref = oracle.ref_paged_prefill(q, c.kv_cache[0], c.kv_cache[1], c.block_table, c.seq_lens, qsl, c.scale,
                                   c.fp8, c.k_scale, c.v_scale)
    return out, ref


def check(out, ref):
    assert torch.isfinite(out.float()).all(), "non-finite output"
    e, l2 = excess_rel_err(out, ref), kernel_rel_l2(out, ref)
    assert e <= TOL, f"excess rel err {e:.3e}"
    assert l2 <= TOL, f"kernel rel L2 {l2:.3e}"


CASES = [
    # (seq_lens, q_lens): pure prefill, extend with context, single-token rows, ragged tails
    ([16], [16]),
    ([17], [17]),
    ([1], [1]),
    ([33, 64], [5, 1]),
    ([100, 257, 31], [100, 1, 31]),
    ([300, 129, 48], [44, 129, 17]),
    ([1000], [1000]),
    ([2049, 700], [513, 699]),
]


@pytest.mark.parametrize("fp8", [False, True])
@pytest.mark.parametrize("seq_lens,q_lens", CASES)
def test_prefill_vs_oracle(lib, seq_lens, q_lens, fp8):
    c = make_case(seq_lens, hq=8, hkv=1, seed=sum(seq_lens), fp8=fp8, garbage_fill=True)
    out, ref = run_prefill(c, q_lens, 8, seed=len(q_lens))
    check(out, ref)


@pytest.mark.parametrize("fp8", [False, True])
@pytest.mark.parametrize("hq,hkv", [(32, 4), (28, 4), (16, 2), (4, 1), (4, 4), (16, 1), (24, 2)])
def test_prefill_gqa_group_shapes(lib, hq, hkv, fp8):
    c = make_case([130, 77, 256], hq=hq, hkv=hkv, seed=hq, fp8=fp8)
    out, ref = run_prefill(c, [130, 20, 1], hq, seed=hkv)
    check(out, ref)


@pytest.mark.parametrize("fp8", [False, True])
@pytest.mark.parametrize("bs", [32, 64])
def test_prefill_larger_block_sizes(lib, bs, fp8):
    c = make_case([200, 65], hq=8, hkv=2, bs=bs, seed=bs, fp8=fp8)
    out, ref = run_prefill(c, [200, 3], 8)
    check(out, ref)


@pytest.mark.parametrize("per_head", [True, False])
def test_fp8_prefill_staging_ragged_batch(lib, per_head):
    """Staged fp8 pages: ragged lengths (a sequence of one token, lengths on and off page boundaries), garbage in the
    unused slots, the tight page budget (sum of the lengths known on the host), per-head and scalar scales — and a
    second step on the same staging buffer with other sequences."""
    staging = ops.PrefillStaging("cuda")
    for lens, q_lens, seed in (([1, 700, 16, 33, 2048, 95], [1, 300, 16, 2, 64, 95], 3), ([129, 40], [129, 7], 4)):
        c = make_case(lens, hq=16, hkv=2, seed=seed, fp8=True, per_head_scale=per_head, garbage_fill=True)
        g = torch.Generator().manual_seed(seed)
        q = torch.randn(sum(q_lens), 16, 128, generator=g).to(torch.bfloat16)
        qsl = torch.tensor([0] + list(torch.tensor(q_lens).cumsum(0)), dtype=torch.int32)
        cache = cache_to_dev(c.kv_cache)
        out = ops.paged_prefill(q.cuda(), cache[0], cache[1], c.block_table.cuda(), c.seq_lens.cuda(), qsl.cuda(),
                                max(q_lens), c.scale, kv_cache_dtype=c.kv_cache_dtype, k_scale=to_dev(c.k_scale),
                                v_scale=to_dev(c.v_scale), staging=staging, max_seq_len=max(lens),
                                total_kv_tokens=sum(lens))
        torch.cuda.synchronize()
        ref = oracle.ref_paged_prefill(q, c.kv_cache[0], c.kv_cache[1], c.block_table, c.seq_lens, qsl, c.scale,
                                       True, c.k_scale, c.v_scale)
        check(out, ref)


def test_fp8_prefill_staging_rule(lib, prefill_kernel, monkeypatch):
    """Default policy of ops.paged_prefill: stage when group x new tokens per sequence >= 256 (long prompts, chunked
    prefill), not for a few new tokens per sequence over long contexts."""
    if prefill_kernel != "umma":
        pytest.skip("one run is enough")
    monkeypatch.setenv("B200ATTN_PREFILL_FP8_STAGED", "1")
    for lens, q_lens, staged in (([300, 200], [300, 150], True), ([300, 200, 150], [3, 2, 1], False)):
        c = make_case(lens, hq=8, hkv=1, seed=7, fp8=True)
        staging = ops.PrefillStaging("cuda")
        g = torch.Generator().manual_seed(5)
        q = torch.randn(sum(q_lens), 8, 128, generator=g).to(torch.bfloat16)
        qsl = torch.tensor([0] + list(torch.tensor(q_lens).cumsum(0)), dtype=torch.int32)
        cache = cache_to_dev(c.kv_cache)
        out = ops.paged_prefill(q.cuda(), cache[0], cache[1], c.block_table.cuda(), c.seq_lens.cuda(), qsl.cuda(),
                                max(q_lens), c.scale, kv_cache_dtype=c.kv_cache_dtype, k_scale=to_dev(c.k_scale),
                                v_scale=to_dev(c.v_scale), staging=staging, max_seq_len=max(lens))
        torch.cuda.synchronize()
        assert (staging.buf is not None) == staged
        check(out, oracle.ref_paged_prefill(q, c.kv_cache[0], c.kv_cache[1], c.block_table, c.seq_lens, qsl, c.scale,
                                            True, c.k_scale, c.v_scale))


def test_fp8_prefill_staging_too_small_is_rejected(lib, prefill_kernel):
    from vllm_omni_b200 import _lib
    if prefill_kernel != "umma":
        pytest.skip("one run is enough")
    c = make_case([100], hq=8, hkv=1, seed=1, fp8=True)
    cache = cache_to_dev(c.kv_cache)
    q = torch.zeros(4, 8, 128, dtype=torch.bfloat16, device="cuda")
    qsl = torch.tensor([0, 4], dtype=torch.int32).cuda()
    need = ops.prefill_staging_bytes(1, 100, 0, 1, 128, 16)
    assert need >= 7 * 2 * 16 * 128 * 2
    small = torch.empty(need - 1024, dtype=torch.uint8, device="cuda")
    off = (-small.data_ptr()) % 1024
    rc = lib.b200attn_paged_prefill_fp8_staged(
        q.data_ptr(), q.stride(0), cache[0].data_ptr(), cache[1].data_ptr(), cache[0].stride(0), cache[0].stride(1),
        cache[0].stride(2), cache[0].shape[0], q.data_ptr(), q.stride(0), c.block_table.cuda().data_ptr(),
        c.block_table.stride(0), c.seq_lens.cuda().data_ptr(), qsl.data_ptr(), 1, 4, 4, 100, 0, 8, 1, 128, 16, 0.1,
        to_dev(c.k_scale).data_ptr(), to_dev(c.v_scale).data_ptr(), 1, small.data_ptr() + off, small.numel() - off,
        None)
    assert rc == -_lib_code("EWORKSPACE") and b"staging needs" in lib.b200attn_last_error()


def _lib_code(name):
    from vllm_omni_b200 import _lib
    return {v: k for k, v in _lib.ERRORS.items()}[name]


def test_prefill_last_row_equals_decode_kernel(lib):
    """The last query row of every sequence sees the whole context: it must agree with the decode
    kernel on the same cache (two different kernels, same math)."""
    c = make_case([333, 64, 1025], hq=16, hkv=2, seed=9)
    q_lens = [40, 64, 1]
    out, _ = run_prefill(c, q_lens, 16, seed=3)
    g = torch.Generator().manual_seed(3)
    q = torch.randn(sum(q_lens), 16, 128, generator=g).to(torch.bfloat16)
    last = torch.tensor(q_lens).cumsum(0) - 1
    cache = cache_to_dev(c.kv_cache)
    dec = ops.paged_decode(q[last].cuda(), cache[0], cache[1], c.block_table.cuda(), c.seq_lens.cuda(),
                           c.max_seq_len, c.scale)
    torch.cuda.synchronize()
    assert (out[last.cuda()].float() - dec.float()).abs().max().item() <= 2e-2 * dec.float().abs().max().item()


def test_prefill_padded_token_count_and_strided_q(lib):
    """q may be a slice of a wider fused-qkv tensor and carry padding rows past query_start_loc[-1]."""
    c = make_case([90, 40], hq=8, hkv=2, seed=5)
    g = torch.Generator().manual_seed(0)
    T, pad = 50, 14
    fused = torch.randn(T + pad, 8 * 128 + 2 * 2 * 128, generator=g).to(torch.bfloat16).cuda()
    q = fused[:, : 8 * 128].view(T + pad, 8, 128)
    qsl = torch.tensor([0, 30, 50], dtype=torch.int32)
    cache = cache_to_dev(c.kv_cache)
    out = torch.full((T + pad, 8, 128), 7.0, dtype=torch.bfloat16, device="cuda")
    ops.paged_prefill(q, cache[0], cache[1], c.block_table.cuda(), c.seq_lens.cuda(), qsl.cuda(), 30, c.scale, out=out)
    torch.cuda.synchronize()
    ref = oracle.ref_paged_prefill(q[:T].cpu(), c.kv_cache[0], c.kv_cache[1], c.block_table, c.seq_lens, qsl, c.scale)
    check(out[:T], ref)
    assert (out[T:].float() == 7.0).all(), "rows past query_start_loc[-1] must not be written"


def test_prefill_long_chunk_properties(lib):
    """An 8k-token chunk against an 8k context (cfg5-style prefix, one kv head group): sampled rows vs
    the oracle, and chunked == unchunked (two 4k chunks with the second extending the first)."""
    S = 8192
    c = make_case([S], hq=8, hkv=1, seed=11)
    out, _ = run_prefill(c, [S], 8, seed=1)
    g = torch.Generator().manual_seed(1)
    q = torch.randn(S, 8, 128, generator=g).to(torch.bfloat16)
    rows = [0, 15, 16, 4095, 4096, 8191]
    for r in rows:
        ref, _ = oracle.ref_paged_decode(q[r:r + 1], c.kv_cache[0], c.kv_cache[1], c.block_table,
                                         torch.tensor([r + 1], dtype=torch.int32), c.scale)
        assert excess_rel_err(out[r:r + 1], ref) <= TOL
    # second half as an extend chunk over the same cache
    cache = cache_to_dev(c.kv_cache)
    qsl = torch.tensor([0, S // 2], dtype=torch.int32).cuda()
    out2 = ops.paged_prefill(q[S // 2:].cuda(), cache[0], cache[1], c.block_table.cuda(), c.seq_lens.cuda(), qsl,
                             S // 2, c.scale)
    torch.cuda.synchronize()
    assert torch.equal(out2, out[S // 2:])


@pytest.mark.parametrize("name", ["vllm_prefill_32q4kv.npz", "vllm_prefill_16q2kv.npz"])
def test_prefill_against_vllm_flash_attn_fixtures(lib, name):
    """Committed outputs of vLLM's own FLASH_ATTN prefill call (FlashAttention-2 paged varlen,
    tests/golden/gen_vllm_gpu_golden.py): both prefill kernels must agree with the oracle to 2e-3 and with the stock
    kernel to a few bf16 roundings of the output scale."""
    import numpy as np
    from pathlib import Path
    z = np.load(Path(__file__).resolve().parent / "golden" / name)
    hq, hkv = int(z["hq"]), int(z["hkv"])
    c = make_case(z["seq_lens"].tolist(), hq=hq, hkv=hkv, seed=int(z["seed"]))
    q = torch.from_numpy(z["q"].copy()).view(torch.bfloat16)
    qsl = torch.from_numpy(z["query_start_loc"].copy())
    q_lens = (qsl[1:] - qsl[:-1]).tolist()
    cache = cache_to_dev(c.kv_cache)
    out = torch.full(q.shape, 7.0, dtype=torch.bfloat16, device="cuda")
    ops.paged_prefill(q.cuda(), cache[0], cache[1], c.block_table.cuda(), c.seq_lens.cuda(), qsl.cuda(), max(q_lens),
                      float(z["scale"]), out=out)
    torch.cuda.synchronize()
    ref = oracle.ref_paged_prefill(q, c.kv_cache[0], c.kv_cache[1], c.block_table, c.seq_lens, qsl, float(z["scale"]))
    check(out, ref)
    stock = torch.from_numpy(z["out_fa2"].copy()).view(torch.bfloat16).float()
    assert float((out.cpu().float() - stock).abs().max()) <= 2 ** -6 * float(stock.abs().max())
