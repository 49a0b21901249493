"""CPU: host-side mirror of the reference interface — TP rule, layout contract, workloads, backend classes."""
import pytest
import torch

from vllm_omni_b200 import layout, tp, workloads


# ---- TP rule (vllm_omni/model_executor/models/qwen2_5_omni/qwen2_old.py:89-106) ----
@pytest.mark.parametrize("tpn,hq,hkv,rep", [(1, 32, 4, 1), (2, 16, 2, 1), (4, 8, 1, 1), (8, 4, 1, 2)])
def test_shard_heads_qwen3_omni(tpn, hq, hkv, rep):
    shards = [tp.shard_heads(32, 4, tpn, r) for r in range(tpn)]
    assert all(s.num_q_heads == hq and s.num_kv_heads == hkv and s.kv_replicas == rep for s in shards)
    # query heads partition [0, 32); every KV head is held by exactly `rep` ranks
    assert sorted(h for s in shards for h in s.q_heads) == list(range(32))
    owners = [h for s in shards for h in s.kv_heads]
    assert sorted(owners) == sorted(list(range(4)) * rep)
    # a rank's query heads map onto its own KV heads under the global GQA grouping
    for s in shards:
        for h in s.q_heads:
            assert h // 8 in s.kv_heads


def test_shard_heads_rejects_what_the_reference_asserts_on():
    with pytest.raises(ValueError):
        tp.shard_heads(32, 4, 3, 0)      # q heads not divisible
    with pytest.raises(ValueError):
        tp.shard_heads(48, 4, 6, 0)      # 6 % 4 != 0 and 4 % 6 != 0
    with pytest.raises(ValueError):
        tp.shard_heads(32, 4, 2, 2)


def test_shard_scales_follow_kv_heads():
    sc = torch.arange(4, dtype=torch.float32)
    assert tp.shard_scales(sc, tp.shard_heads(32, 4, 2, 1)).tolist() == [2.0, 3.0]
    assert tp.shard_scales(sc, tp.shard_heads(32, 4, 8, 5)).tolist() == [2.0]
    one = torch.ones(1)
    assert tp.shard_scales(one, tp.shard_heads(32, 4, 8, 5)) is one


# ---- layout ----
def test_kv_cache_layout_contract():
    c = layout.alloc_kv_cache(10, 16, 4, 128, "auto", "cpu")
    assert tuple(c.shape) == (2, 10, 16, 4, 128) and c.dtype == torch.bfloat16
    # physical [nb, Hkv, 2, bs, d]: K page and V page of one (block, head) are adjacent
    assert c.stride() == (16 * 128, 4 * 2 * 16 * 128, 128, 2 * 16 * 128, 1)
    assert c[0][3].shape == (16, 4, 128)  # cache[kv][block_id] -> [bs, Hkv, d] (kv_utils.py:417-450)
    f = layout.alloc_kv_cache(10, 32, 1, 128, "fp8_e4m3", "cpu")
    assert f.dtype == torch.uint8 and f.shape[2] == 32
    with pytest.raises(ValueError):
        layout.kv_cache_shape(10, 8, 4, 128)


# ---- workloads ----
def test_named_workloads_match_baseline_table():
    w = workloads.WORKLOADS
    assert (w["cfg3"].total_q_heads, w["cfg3"].total_kv_heads, w["cfg3"].batch, w["cfg3"].ctx) == (32, 4, 64, 4096)
    assert w["cfg1"].kv_bytes_per_layer() == 524_288
    assert w["cfg2"].kv_bytes_per_layer() == 134_217_728
    assert w["cfg3"].kv_bytes_per_layer() == 536_870_912
    assert w["cfg4"].kv_bytes_per_layer() == 1_073_741_824
    # 16 x 32768 tokens x 2 x 4 x 128 x 2 B = 1 GiB (SURVEY.md §8 / BASELINE.md print 512 MiB for
    # this row, which is an arithmetic slip there; the formula is the same for every row)
    assert w["cfg5"].kv_bytes_per_layer() == 1_073_741_824
    # per-rank bytes under KV-head TP: 512 / 256 / 128 / 128 MiB
    per_rank = [w["cfg3"].kv_bytes_per_layer(tp.shard_heads(32, 4, n, 0).num_kv_heads) >> 20 for n in (1, 2, 4, 8)]
    assert per_rank == [512, 256, 128, 128]


def test_block_table_is_a_permutation_and_slots_are_unique():
    lens = workloads.make_seq_lens(8, 100, ragged=True)
    assert lens.min() >= 50 and lens.max() <= 100
    bt, nb = workloads.make_block_table(lens, 16)
    pages = (lens + 15) // 16
    used = torch.cat([bt[b, :pages[b]] for b in range(8)])
    assert used.unique().numel() == used.numel() and int(used.max()) < nb
    assert nb == int(pages.sum()) + 64
    seq_idx = torch.repeat_interleave(torch.arange(8), lens.long())
    pos = torch.cat([torch.arange(int(n)) for n in lens])
    slots = workloads.slots_of_positions(bt, seq_idx, pos, 16)
    assert slots.unique().numel() == slots.numel()
    assert torch.equal(workloads.make_block_table(lens, 16)[0], bt)  # seeded: reproducible


# ---- backend classes against vLLM's ABCs ----
@pytest.fixture(scope="module")
def backend():
    pytest.importorskip("vllm")
    from vllm_omni_b200 import backend as be
    return be


def test_backend_static_contract(backend):
    from vllm.v1.attention.backend import AttentionBackend, AttentionImpl, AttentionMetadataBuilder
    B = backend.B200AttentionBackend
    assert issubclass(B, AttentionBackend)
    assert issubclass(B.get_impl_cls(), AttentionImpl)
    assert issubclass(B.get_builder_cls(), AttentionMetadataBuilder)
    assert B.get_name() == "CUSTOM"
    assert B.forward_includes_kv_cache_update is False
    assert B.get_kv_cache_shape(7, 16, 4, 128) == (2, 7, 16, 4, 128)
    assert B.get_kv_cache_block_dim(16, 4, 128) == 1
    order = B.get_kv_cache_stride_order()
    assert sorted(order) == list(range(5)) and order == layout.KV_CACHE_STRIDE_ORDER
    assert sorted(B.get_kv_cache_stride_order(include_num_layers_dimension=True)) == list(range(6))
    assert B.supports_block_size(16) and B.supports_block_size(64) and not B.supports_block_size(8)
    assert B.supports_head_size(128) and not B.supports_head_size(64)
    assert B.supports_kv_cache_dtype("fp8_e4m3") and not B.supports_kv_cache_dtype("fp8_e5m2")
    assert B.supports_per_head_quant_scales()
    # the reference's builder patch probes these two attributes (vllm_omni/worker/gpu_model_runner.py:63-75)
    assert not hasattr(B.get_builder_cls(), "scheduler_metadata")
    assert not hasattr(B.get_builder_cls(), "max_num_splits")


def test_backend_validates_on_blackwell_only(backend):
    from vllm.platforms.interface import DeviceCapability
    B = backend.B200AttentionBackend
    args = dict(head_size=128, dtype=torch.bfloat16, kv_cache_dtype="fp8_e4m3", block_size=16,
                use_mla=False, has_sink=False, use_sparse=False, use_mm_prefix=False,
                use_per_head_quant_scales=True, attn_type="decoder")
    assert B.validate_configuration(device_capability=DeviceCapability(10, 0), **args) == []
    assert B.validate_configuration(device_capability=DeviceCapability(9, 0), **args) != []
    bad = dict(args, dtype=torch.float16)
    assert "dtype not supported" in B.validate_configuration(device_capability=DeviceCapability(10, 0), **bad)


def test_plugin_registers_custom_backend(backend):
    from vllm.v1.attention.backends.registry import AttentionBackendEnum
    from vllm_omni_b200 import plugin
    plugin.register()
    plugin.register()  # idempotent: the loaders may call it more than once per process
    assert AttentionBackendEnum.CUSTOM.get_path() == plugin.BACKEND_PATH
    assert AttentionBackendEnum.CUSTOM.get_class() is backend.B200AttentionBackend
    assert plugin.force_for_stage({})["attention_backend"] == "CUSTOM"
    assert plugin.force_for_stage({"attention_backend": "FLASH_ATTN"})["attention_backend"] == "FLASH_ATTN"


def test_impl_rejects_unsupported_features(backend):
    Impl = backend.B200AttentionImpl
    Impl(32, 128, 0.088, 4, None, None, "auto")
    Impl(32, 128, 0.088, 4, None, None, "fp8_e4m3", None, "decoder", None)
    for kw in (dict(alibi_slopes=[0.1] * 32), dict(sliding_window=128), dict(logits_soft_cap=30.0)):
        with pytest.raises(NotImplementedError):
            Impl(32, 128, 0.088, 4, kv_cache_dtype="auto", **kw)
    with pytest.raises(ValueError):
        Impl(32, 128, 0.088, 4, None, None, "fp8_e5m2")
    impl = Impl(32, 128, 0.088, 4, None, None, "auto")
    out = torch.ones(3, 32 * 128)
    assert impl.forward(None, None, None, None, None, None, out).abs().sum() == 0  # profiling run
    with pytest.raises(NotImplementedError):
        impl.forward(None, None, None, None, None, object(), out, output_scale=torch.ones(1))


# ---- peer-memory all-reduce: host arithmetic (no GPU) ----
def test_peer_allreduce_schedule_choice_and_region_layout(lib, monkeypatch):
    from vllm_omni_b200 import _lib, peer
    monkeypatch.delenv("B200ATTN_AR_MULTIMEM", raising=False)
    # two ranks: the same bytes cross NVLink either way, so never pay the extra barrier
    assert peer.choose_algo(8 << 20, 2, False) == _lib.AR_ONESHOT
    for world in (4, 8):
        assert peer.choose_algo(64 << 10, world, False) == _lib.AR_ONESHOT
        assert peer.choose_algo(512 << 10, world, False) == _lib.AR_TWOSHOT
        assert peer.choose_algo(512 << 10, world, True) == _lib.AR_TWOSHOT      # multimem is opt-in
    monkeypatch.setenv("B200ATTN_AR_MULTIMEM", "1")
    assert peer.choose_algo(512 << 10, 8, True) == _lib.AR_MULTIMEM
    assert peer.choose_algo(512 << 10, 8, False) == _lib.AR_TWOSHOT
    # PUSH (barrier-free) while the posted bytes stay latency-sized and fit the provisioned slot
    assert peer.choose_algo(512 << 10, 2, False, 512 << 10) == _lib.AR_PUSH
    assert peer.choose_algo(512 << 10, 2, False, 256 << 10) == _lib.AR_ONESHOT      # slot too small
    assert peer.choose_algo(128 << 10, 8, False, 1 << 20) == _lib.AR_PUSH
    assert peer.choose_algo(2 << 20, 8, False, 1 << 20) == _lib.AR_PUSH2            # 14 MiB of one-shot pushes: scatter-reduce instead
    assert peer.choose_algo(1 << 20, 4, False, 1 << 20) == _lib.AR_PUSH2
    assert peer.choose_algo(8 << 20, 8, False, 1 << 20) == _lib.AR_TWOSHOT          # a slice no longer fits half a slot
    assert peer.push_layout(128 * 2048, 1) == 0 and peer.push_layout(128 * 2048, 2) == 512 << 10
    assert peer.push_layout(1 << 22, 8) == peer.PUSH_MAX_BYTES
    assert peer.region_layout(128 * 2048, 2)[3] == peer.region_layout(128 * 2048)[3] + 3 * 2 * (512 << 10)
    off_in, off_scr, cap, total = peer.region_layout(512 * 2048)
    assert off_in == _lib.PEER_FLAG_BYTES and off_in % 256 == 0 and off_scr % 256 == 0
    assert cap >= 512 * 2048 * 2 + 16 and off_scr == off_in + cap and total == off_scr + cap


def test_peer_allreduce_api_rejects_bad_arguments_without_a_gpu(lib):
    import ctypes as C
    from vllm_omni_b200 import _lib
    arr = (C.c_void_p * 2)(C.c_void_p(4096), C.c_void_p(8192))
    f = lib.b200attn_peer_allreduce
    assert f(arr, arr, arr, None, 0, 65536, None, None, 64, 3, 0, 1, 0, 0, None) == -2      # world 3 unsupported
    assert f(arr, arr, arr, None, 0, 65536, None, None, 12, 2, 0, 1, 0, 0, None) == -1      # numel % 8
    assert f(arr, arr, arr, None, 0, 65536, None, None, 64, 2, 2, 1, 0, 0, None) == -1      # rank out of range
    assert f(arr, arr, arr, None, 0, 65536, None, None, 64, 2, 0, 3, 0, 0, None) == -1      # multimem without multicast ptrs
    assert f(arr, arr, arr, None, 0, 4096, None, None, 64, 2, 0, 1, 0, 0, None) == -1       # out aliases the input
    assert f(arr, arr, arr, None, 0, 65536, None, None, 0, 2, 0, 1, 0, 0, None) == 0        # empty: nothing to do
    assert f(arr, arr, arr, None, 0, 65536, None, None, 64, 2, 0, 4, 0, 0, None) == -1      # PUSH without receive buffers
    assert f(arr, arr, arr, arr, 64, 65536, None, None, 64, 2, 0, 4, 0, 0, None) == -1      # PUSH slot smaller than the message


def test_oracle_all_reduce_is_sequential_fp32_then_one_rounding():
    import torch
    from oracle import oracle
    a = torch.tensor([1.0, 256.0, -3.0]).to(torch.bfloat16)
    b = torch.tensor([0.00390625, 1.0, 3.0]).to(torch.bfloat16)       # 2^-8: lost only at the final rounding
    c = torch.tensor([0.00390625, 1.0, 0.0]).to(torch.bfloat16)
    out = oracle.ref_all_reduce_bf16([a, b, c])
    assert out.tolist() == [1.0078125, 258.0, 0.0]                    # 1 + 2^-7 is representable; a bf16 running sum would give 1.0


def test_fp8_prefill_staging_size_is_host_arithmetic(lib):
    """b200attn_prefill_staging_bytes: pages for batch x ceil(max_seq_len / bs), or for the tighter host bound of the
    summed lengths, plus the prefix and the staged block table (each part rounded to 1 KiB)."""
    f = lib.b200attn_prefill_staging_bytes
    page = 4 * 2 * 16 * 128 * 2                       # [Hkv][K|V][bs][d] bf16
    up = lambda n: (n + 1023) // 1024 * 1024          # noqa: E731
    dense = f(3, 1000, 0, 4, 128, 16)
    assert dense == 3 * 63 * page + up(4 * 4) + up(3 * 63 * 4)
    tight = f(3, 1000, 1100, 4, 128, 16)              # 1100 tokens in all: at most 1100 // 16 + 3 pages
    assert tight == up(71 * page) + up(4 * 4) + up(3 * 63 * 4) and tight < dense
    assert f(3, 1000, 10 ** 9, 4, 128, 16) == dense   # a useless bound changes nothing
    assert f(0, 1000, 0, 4, 128, 16) == 0 and f(3, 0, 0, 4, 128, 16) == 0


def test_fp8_prefill_staging_rule_is_host_arithmetic(monkeypatch):
    """ops._fp8_prefill_staged: group x new tokens per sequence >= 256 (break-even ~210, see its docstring)."""
    from vllm_omni_b200 import ops
    monkeypatch.delenv("B200ATTN_PREFILL_FP8_STAGED", raising=False)
    monkeypatch.delenv("B200ATTN_PREFILL_UMMA", raising=False)
    assert ops._fp8_prefill_staged(8192, 1, 8)            # one long prompt
    assert ops._fp8_prefill_staged(16 * 64, 16, 8)        # 64 new tokens on each of 16 contexts (measured 1.9x)
    assert not ops._fp8_prefill_staged(16 * 16, 16, 8)    # 16 new tokens each: the staging would cost what it saves
    assert not ops._fp8_prefill_staged(64 * 4, 64, 8)     # speculative decoding rows
    assert ops._fp8_prefill_staged(64 * 64, 64, 4) and not ops._fp8_prefill_staged(64 * 32, 64, 4)
    monkeypatch.setenv("B200ATTN_PREFILL_FP8_STAGED", "force")
    assert ops._fp8_prefill_staged(3, 3, 8)
    monkeypatch.setenv("B200ATTN_PREFILL_UMMA", "0")
    assert not ops._fp8_prefill_staged(8192, 1, 8)
    monkeypatch.setenv("B200ATTN_PREFILL_UMMA", "1")
    monkeypatch.setenv("B200ATTN_PREFILL_FP8_STAGED", "0")
    assert not ops._fp8_prefill_staged(8192, 1, 8)
