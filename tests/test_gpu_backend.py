"""GPU: the vLLM-facing plugin classes (AttentionImpl.do_kv_cache_update / forward) end to end."""
import types

import pytest
import torch

from oracle import c_oracle
from tests.helpers import excess_rel_err, kernel_rel_l2, make_case, rel_l2
from vllm_omni_b200 import layout, ops

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def be():
    pytest.importorskip("vllm")
    from vllm_omni_b200 import backend
    return backend


def _layer(k_scale=None, v_scale=None):
    one = torch.ones(1, dtype=torch.float32, device="cuda")
    return types.SimpleNamespace(_k_scale=one if k_scale is None else k_scale.cuda(),
                                 _v_scale=one if v_scale is None else v_scale.cuda())


def _metadata(be, c, ws, splits):
    B = c.seq_lens.numel()
    qsl = torch.arange(B + 1, dtype=torch.int32, device="cuda")
    return be.B200AttentionMetadata(
        num_actual_tokens=B, max_query_len=1, query_start_loc=qsl, max_seq_len=c.max_seq_len,
        seq_lens=c.seq_lens.cuda(), block_table=c.block_table.cuda(), slot_mapping=None,
        num_decodes=B, num_decode_tokens=B, num_prefills=0, num_prefill_tokens=0,
        prefill_query_start_loc=None, prefill_max_query_len=0, decode_num_splits=splits, workspace=ws)


@pytest.mark.parametrize("fp8", [False])
def test_impl_write_then_decode_step(lib, be, fp8):
    """One decode step the way Attention.forward drives it (attention.py:509-528): cache update with the
    new token's K/V, then forward() writing into the preallocated output."""
    hq, hkv, d = 32, 4, 128
    c = make_case([257, 64, 1, 1000], hq=hq, hkv=hkv, seed=17, fp8=fp8)
    impl = be.B200AttentionImpl(hq, d, c.scale, hkv, None, None, c.kv_cache_dtype)
    layer = _layer(c.k_scale, c.v_scale)
    B = 4
    # device cache holds the history WITHOUT the last token of each sequence
    kv = layout.alloc_kv_cache(c.num_blocks, 16, hkv, d, c.kv_cache_dtype, "cuda")
    l64 = c.seq_lens.long()
    last = torch.cumsum(l64, 0) - 1
    keep = torch.ones(int(l64.sum()), dtype=torch.bool)
    keep[last] = False
    impl.do_kv_cache_update(layer, c.hist_k[keep].cuda(), c.hist_v[keep].cuda(), kv, c.slots[keep].cuda())
    # the step: padded key/value (B + 2 rows), slot_mapping of B rows
    pad = torch.zeros(2, hkv, d, dtype=torch.bfloat16)
    key = torch.cat([c.hist_k[last], pad]).cuda()
    value = torch.cat([c.hist_v[last], pad]).cuda()
    impl.do_kv_cache_update(layer, key, value, kv, c.slots[last].cuda())
    ws = ops.DecodeWorkspace("cuda", ops.decode_workspace_bytes(B, hq, d, 8))
    query = torch.cat([c.q, torch.zeros(2, hq, d, dtype=torch.bfloat16)]).cuda()
    output = torch.full((B + 2, hq * d), 9.0, dtype=torch.bfloat16, device="cuda")
    ret = impl.forward(layer, query, key, value, kv, _metadata(be, c, ws, 8), output)
    torch.cuda.synchronize()
    assert ret is output
    ref, _ = c_oracle.paged_decode(c.q, c.kv_cache[0], c.kv_cache[1], c.block_table, c.seq_lens, c.scale,
                                   fp8, c.k_scale, c.v_scale)
    got = output[:B].view(B, hq, d)
    assert excess_rel_err(got, ref) <= 2e-3 and kernel_rel_l2(got, ref) <= 2e-3
    assert torch.all(output[B:] == 9.0)  # padded rows are not touched
    # the cache now equals the oracle-written cache bit for bit
    assert torch.equal(kv.cpu().view(torch.int16), c.kv_cache.view(torch.int16))


def test_builder_builds_without_sync_and_sizes_workspace(lib, be):
    """B200AttentionMetadataBuilder against a minimal stand-in VllmConfig."""
    from vllm.v1.attention.backend import CommonAttentionMetadata
    mc = types.SimpleNamespace(get_num_attention_heads=lambda pc: 32, get_num_kv_heads=lambda pc: 4,
                               get_head_size=lambda: 128, max_model_len=4096)
    cfg = types.SimpleNamespace(model_config=mc, parallel_config=None,
                                scheduler_config=types.SimpleNamespace(max_num_seqs=64))
    spec = types.SimpleNamespace(block_size=16)
    b = be.B200AttentionMetadataBuilder(spec, ["l0"], cfg, torch.device("cuda"))
    assert b.workspace.nbytes >= ops.decode_workspace_bytes(
        64, 32, 128, ops.decode_num_splits(64, 32, 4, 4096))
    c = make_case([100, 200, 300], hq=32, hkv=4, seed=2)
    qsl_cpu = torch.arange(4, dtype=torch.int32)
    cam = CommonAttentionMetadata(
        query_start_loc=qsl_cpu.cuda(), query_start_loc_cpu=qsl_cpu, seq_lens=c.seq_lens.cuda(), num_reqs=3,
        num_actual_tokens=3, max_query_len=1, max_seq_len=300, block_table_tensor=c.block_table.cuda(),
        slot_mapping=torch.zeros(3, dtype=torch.int64, device="cuda"))
    md = b.build(0, cam)
    assert (md.num_decodes, md.num_prefills, md.num_decode_tokens) == (3, 0, 3)
    # 0 = the library's default schedule (stream-K), nothing for the host to choose
    assert md.decode_num_splits == 0 and md.workspace is b.workspace
    md2 = b.build_for_cudagraph_capture(cam)
    assert md2.num_decodes == 3


@pytest.mark.parametrize("fp8", [False, True])
def test_mixed_batch_decodes_then_prefills_through_builder_and_impl(lib, be, fp8):
    """A step with 2 decode requests followed by 2 prefill/extend requests (vLLM orders decodes first,
    reorder_batch_threshold = 1): metadata from the builder, cache update for all new tokens, one
    forward() call that routes rows to the decode and the prefill kernels."""
    from vllm.v1.attention.backend import CommonAttentionMetadata
    from oracle import oracle
    hq, hkv, d = 16, 2, 128
    seq_lens = [200, 33, 150, 64]      # context INCLUDING this step's new tokens
    q_lens = [1, 1, 70, 64]            # decode, decode, extend (80 cached + 70 new), pure prefill
    c = make_case(seq_lens, hq=hq, hkv=hkv, seed=23, fp8=fp8)
    mc = types.SimpleNamespace(get_num_attention_heads=lambda pc: hq, get_num_kv_heads=lambda pc: hkv,
                               get_head_size=lambda: d, max_model_len=1024)
    cfg = types.SimpleNamespace(model_config=mc, parallel_config=None,
                                scheduler_config=types.SimpleNamespace(max_num_seqs=8))
    b = be.B200AttentionMetadataBuilder(types.SimpleNamespace(block_size=16), ["l0"], cfg, torch.device("cuda"))
    T = sum(q_lens)
    qsl_cpu = torch.tensor([0] + list(torch.tensor(q_lens).cumsum(0)), dtype=torch.int32)
    # new tokens = the last q_len positions of every sequence
    l64 = c.seq_lens.long()
    starts = torch.cumsum(l64, 0) - l64
    new_idx = torch.cat([torch.arange(int(starts[i] + l64[i] - q_lens[i]), int(starts[i] + l64[i])) for i in range(4)])
    old = torch.ones(int(l64.sum()), dtype=torch.bool)
    old[new_idx] = False
    impl = be.B200AttentionImpl(hq, d, c.scale, hkv, None, None, c.kv_cache_dtype)
    layer = _layer(c.k_scale, c.v_scale)
    kv = layout.alloc_kv_cache(c.num_blocks, 16, hkv, d, c.kv_cache_dtype, "cuda")
    impl.do_kv_cache_update(layer, c.hist_k[old].cuda(), c.hist_v[old].cuda(), kv, c.slots[old].cuda())
    cam = CommonAttentionMetadata(
        query_start_loc=qsl_cpu.cuda(), query_start_loc_cpu=qsl_cpu, seq_lens=c.seq_lens.cuda(), num_reqs=4,
        num_actual_tokens=T, max_query_len=max(q_lens), max_seq_len=max(seq_lens),
        block_table_tensor=c.block_table.cuda(), slot_mapping=c.slots[new_idx].cuda())
    md = b.build(0, cam)
    assert (md.num_decodes, md.num_prefills, md.num_decode_tokens, md.num_prefill_tokens) == (2, 2, 2, T - 2)
    assert md.prefill_max_query_len == 70
    key, value = c.hist_k[new_idx].cuda(), c.hist_v[new_idx].cuda()
    impl.do_kv_cache_update(layer, key, value, kv, md.slot_mapping)
    g = torch.Generator().manual_seed(5)
    query = torch.randn(T, hq, d, generator=g).to(torch.bfloat16)
    output = torch.full((T, hq * d), 9.0, dtype=torch.bfloat16, device="cuda")
    impl.forward(layer, query.cuda(), key, value, kv, md, output)
    torch.cuda.synchronize()
    ref = oracle.ref_paged_prefill(query, c.kv_cache[0], c.kv_cache[1], c.block_table, c.seq_lens, qsl_cpu, c.scale,
                                   fp8, c.k_scale, c.v_scale)
    got = output.view(T, hq, d)
    assert excess_rel_err(got, ref) <= 2e-3 and kernel_rel_l2(got, ref) <= 2e-3
