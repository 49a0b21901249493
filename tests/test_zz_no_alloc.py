"""b200attn_paged_decode_static: decode with the split-KV partials in the library's own `static __device__` array
(no workspace argument, no allocation). No counterpart in the reference.
Collected last on purpose.

The GPU test calls the new entry point through the C ABI twice in a row (different batches, so the second call
finds the first one's partials in the array) and compares each result with the existing entry point:
bit for bit against b200attn_paged_decode on the same schedule (variant STREAM, same split count), and within the
decode tests' tolerance (2e-3, tests/test_gpu_decode.py) against the default schedule and the fp32 oracle.
"""
import re
from pathlib import Path

import pytest
import torch

from tests.helpers import cache_to_dev, excess_rel_err, make_case
from vllm_omni_b200 import _lib, ops

TOL = 2e-3
HEADER = Path(__file__).resolve().parent.parent / "include" / "b200attn.h"


def test_static_scratch_limits_match_the_header(lib):
    text = HEADER.read_text()
    assert int(re.search(r"#define B200ATTN_STATIC_SCRATCH_BYTES (\d+)", text).group(1)) == _lib.STATIC_SCRATCH_BYTES
    assert int(re.search(r"#define B200ATTN_STATIC_MAX_ROWS (\d+)", text).group(1)) == _lib.STATIC_MAX_ROWS
    assert lib.b200attn_static_scratch_bytes() == _lib.STATIC_SCRATCH_BYTES
    # partial O rows (128 f32) + their LSE behind the 64 KiB control head every decode workspace starts with
    assert 65536 + _lib.STATIC_MAX_ROWS * (128 + 1) * 4 + 256 <= _lib.STATIC_SCRATCH_BYTES


def _device_case(lens, hq, hkv, seed):
    c = make_case(lens, hq=hq, hkv=hkv, seed=seed)
    cache = cache_to_dev(c.kv_cache)
    return c, c.q.cuda(), cache, c.block_table.cuda(), c.seq_lens.cuda()


def _static(c, q, cache, bt, sl, splits):
    B, hq, d = q.shape
    out = torch.full((B, hq, d), 7.0, dtype=torch.bfloat16, device="cuda")
    lse = torch.full((B, hq), 123.0, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    ops.paged_decode_static(q, cache[0], cache[1], bt, sl, c.max_seq_len, c.scale, out=out, lse=lse, num_splits=splits)
    assert torch.cuda.memory_allocated() == before, "the call allocated device memory through torch"
    torch.cuda.synchronize()
    return out, lse


def _existing(c, q, cache, bt, sl, splits, variant):
    B, hq, d = q.shape
    out = torch.full((B, hq, d), 7.0, dtype=torch.bfloat16, device="cuda")
    lse = torch.full((B, hq), 123.0, dtype=torch.float32, device="cuda")
    ops.paged_decode(q, cache[0], cache[1], bt, sl, c.max_seq_len, c.scale, out=out, lse=lse, num_splits=splits,
                     variant=variant)
    torch.cuda.synchronize()
    return out, lse


def _compare(c, dev, splits):
    from oracle import c_oracle

    out, lse = _static(c, *dev, splits)
    same, same_lse = _existing(c, *dev, splits, ops.VARIANT_STREAM)
    assert torch.equal(out, same), "differs from b200attn_paged_decode on the same schedule"
    assert torch.equal(lse, same_lse)
    dflt, dflt_lse = _existing(c, *dev, 0, ops.VARIANT_AUTO)
    assert torch.isfinite(out.float()).all()
    # two bf16 outputs, each within TOL of the fp32 result and each rounded once: twice the bound between them
    e = excess_rel_err(out, dflt)
    assert e <= 2 * TOL, f"excess rel err {e:.3e} against the default schedule"
    fin = torch.isfinite(dflt_lse)
    assert torch.equal(torch.isfinite(lse), fin)
    assert torch.allclose(lse[fin], dflt_lse[fin], atol=2e-3, rtol=1e-4)
    ref, _ = c_oracle.paged_decode(c.q, c.kv_cache[0], c.kv_cache[1], c.block_table, c.seq_lens, c.scale, c.fp8,
                                   c.k_scale, c.v_scale)
    e = excess_rel_err(out, ref)
    assert e <= TOL, f"excess rel err {e:.3e} against the oracle"
    return out


@pytest.mark.gpu
def test_two_static_calls_in_a_row_match_the_existing_entry_point(lib):
    # first call: 4 sequences x 32 heads x 8 splits = 1024 partial rows (one sequence empty, one a single tile)
    c1, *d1 = _device_case([700, 0, 9, 1234], hq=32, hkv=4, seed=11)
    first = _compare(c1, d1, 8)
    # second call, straight after: another batch and split count over the rows the first one left behind
    c2, *d2 = _device_case([333, 2048, 65, 17, 512, 1], hq=16, hkv=2, seed=12)
    _compare(c2, d2, 5)
    # heuristic split count, then the first batch again: the same bits as before
    _compare(c2, d2, 0)
    again, _ = _static(c1, *d1, 8)
    assert torch.equal(again, first)


@pytest.mark.gpu
def test_static_call_rejects_what_does_not_fit(lib):
    c, *d = _device_case([64] * 8, hq=32, hkv=4, seed=13)
    splits = _lib.STATIC_MAX_ROWS // (8 * 32) + 1           # 63 splits x 256 rows > 16000
    with pytest.raises(_lib.B200AttnError, match="EWORKSPACE.*static scratch holds 16000 partial rows"):
        _static(c, *d, splits)
    out, _ = _static(c, *d, 1)                              # one range per sequence needs no scratch: no limit
    ref, _ = _existing(c, *d, 1, ops.VARIANT_STREAM)
    assert torch.equal(out, ref)
    q, cache, bt, sl = d
    args = ops._decode_args(q, cache[0], cache[1], out, None, bt, sl, c.max_seq_len, c.scale, "auto", None, None, 2,
                            ops.VARIANT_STREAM, None)
    rc = lib.b200attn_paged_decode_static(*args[:21], _lib.KV_FP8_E4M3, 2, args[30])
    assert rc == -2 and b"bf16 caches only" in lib.b200attn_last_error()
