"""vLLM AttentionBackend / MetadataBuilder / AttentionImpl for the B200 paged-attention path.

This is layer 1 of the drop-in boundary: the class interface vLLM's `Attention`
layer calls for vllm-omni's thinker and talker stages
(vllm/model_executor/layers/attention/attention.py:437-529 -> `impl.do_kv_cache_update`,
`impl.forward`; caller of both: vllm_omni/worker/gpu_ar_model_runner.py:421-520).
All arithmetic happens in libb200attn.so through `ops`; there is no other code path.

Selection: `register_backend(AttentionBackendEnum.CUSTOM, ...)` (see `plugin.py`) plus
`attention_backend: CUSTOM` in the stage's engine_args — the same knob the reference forces for
ROCm in vllm_omni/engine/stage_init_utils.py:282-293.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import ClassVar

import torch
from vllm.v1.attention.backend import (
    AttentionBackend,
    AttentionCGSupport,
    AttentionImpl,
    AttentionMetadataBuilder,
    AttentionType,
    CommonAttentionMetadata,
    MultipleOf,
)
from vllm.v1.attention.backends.utils import split_decodes_and_prefills

from . import layout, ops

BACKEND_PATH = "vllm_omni_b200.backend.B200AttentionBackend"


class B200AttentionBackend(AttentionBackend):
    """Static capability + layout contract (vllm/v1/attention/backend.py:55-343)."""

    supported_dtypes: ClassVar[list[torch.dtype]] = [torch.bfloat16]
    supported_kv_cache_dtypes: ClassVar[list[str]] = ["auto", "bfloat16", "fp8", "fp8_e4m3"]
    # the cache write is a separate call (unified_kv_cache_update), as for FLASH_ATTN
    forward_includes_kv_cache_update: bool = False

    @staticmethod
    def get_name() -> str:
        # Attention.__init__ does AttentionBackendEnum[backend.get_name()] (attention.py:386)
        return "CUSTOM"

    @staticmethod
    def get_impl_cls() -> type["B200AttentionImpl"]:
        return B200AttentionImpl

    @staticmethod
    def get_builder_cls() -> type["B200AttentionMetadataBuilder"]:
        return B200AttentionMetadataBuilder

    @staticmethod
    def get_supported_kernel_block_sizes() -> list[int | MultipleOf]:
        # 16-token tiles, 16 * 2^k tokens per page: what the stream-K schedules (the only ones with an fp8 path)
        # index with shifts. Anything else would be refused at the first decode, so it is refused here, at start-up.
        return [16, 32, 64, 128]

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int, head_size: int,
                           cache_dtype_str: str = "auto") -> tuple[int, ...]:
        return layout.kv_cache_shape(num_blocks, block_size, num_kv_heads, head_size)

    @staticmethod
    def get_kv_cache_stride_order(include_num_layers_dimension: bool = False) -> tuple[int, ...]:
        if include_num_layers_dimension:
            # (num_blocks, num_kv_heads, num_layers, 2, block_size, head_size)
            return (2, 4, 0, 1, 3, 5)
        return layout.KV_CACHE_STRIDE_ORDER

    @classmethod
    def get_supported_head_sizes(cls) -> list[int]:
        return [128]

    @classmethod
    def supports_per_head_quant_scales(cls) -> bool:
        return True

    @classmethod
    def supports_compute_capability(cls, capability) -> bool:
        return capability.major == 10

    @classmethod
    def supports_attn_type(cls, attn_type: str) -> bool:
        return attn_type == AttentionType.DECODER


@dataclass
class B200AttentionMetadata:
    """Per-step metadata, shared by all layers of a KV-cache group (model: FlashAttentionMetadata,
    vllm/v1/attention/backends/flash_attn.py:222-256). The batch is ordered decodes first."""

    num_actual_tokens: int
    max_query_len: int
    query_start_loc: torch.Tensor
    max_seq_len: int
    seq_lens: torch.Tensor
    block_table: torch.Tensor
    slot_mapping: torch.Tensor
    num_decodes: int
    num_decode_tokens: int
    num_prefills: int
    num_prefill_tokens: int
    # prefill slice, rebased so that row 0 is the first prefill token
    prefill_query_start_loc: torch.Tensor | None
    prefill_max_query_len: int
    decode_num_splits: int
    workspace: ops.DecodeWorkspace
    causal: bool = True
    # the step's stream-K plan is in `workspace` (made once by the builder; every layer's decode skips its own)
    decode_planned: bool = False
    # fp8 caches: bf16 staging of the prefill sequences' pages for the tcgen05 prefill kernel (shared by the layers)
    prefill_staging: ops.PrefillStaging | None = None
    prefill_total_kv_tokens: int = 0  # host bound of sum(seq_lens[num_decodes:]), 0 = unknown


class B200AttentionMetadataBuilder(AttentionMetadataBuilder[B200AttentionMetadata]):
    """Builds B200AttentionMetadata once per step (vllm/v1/attention/backend.py:516-663).

    No host sync and no device allocation that a captured graph could outlive: the split-K
    workspace is allocated once here for the largest decode batch. Deliberately has neither a
    `scheduler_metadata` nor a `max_num_splits` attribute, so the reference's patch in
    vllm_omni/worker/gpu_model_runner.py:63-75 leaves it alone.
    """

    _cudagraph_support: ClassVar[AttentionCGSupport] = AttentionCGSupport.UNIFORM_SINGLE_TOKEN_DECODE
    reorder_batch_threshold: int = 1

    def __init__(self, kv_cache_spec, layer_names, vllm_config, device):
        super().__init__(kv_cache_spec, layer_names, vllm_config, device)
        mc, pc = vllm_config.model_config, vllm_config.parallel_config
        self.num_heads_q = mc.get_num_attention_heads(pc)
        self.num_heads_kv = mc.get_num_kv_heads(pc)
        self.head_size = mc.get_head_size()
        self.block_size = kv_cache_spec.block_size
        self.max_model_len = mc.max_model_len
        self.max_num_seqs = vllm_config.scheduler_config.max_num_seqs
        self._sms = ops.num_sms()
        self.prefill_staging = ops.PrefillStaging(device)  # nothing is allocated until an fp8 prefill needs it
        # the stream-K schedules' workspace is batch-independent apart from the plan's [batch + 1] prefix; the
        # fixed-split fallback (only taken when stream-K does not fit, i.e. never with this size) is not provisioned
        self.workspace = ops.DecodeWorkspace(device, ops.decode_workspace_bytes(
            max(1, self.max_num_seqs), self.num_heads_q, self.head_size, 1))
        # KV head replicated on several TP ranks (tp > total KV heads): optionally split the batch inside the replica
        # group instead (tp.ReplicaSplit, opt-in: B200ATTN_REPLICA_SPLIT=1); built once per worker from vLLM's TP
        # group and handed to this builder's layers, staging buffers sized for max_num_seqs here, outside any graph
        from . import tp as _tp
        try:
            rs = _tp.replica_split_for_vllm(vllm_config, self.max_num_seqs)
        except Exception:  # noqa: BLE001 - no distributed environment (single-process tests): nothing to split
            rs = None
        if rs is not None:
            ctx = vllm_config.compilation_config.static_forward_context
            for name in layer_names:
                impl = getattr(ctx.get(name), "impl", None)
                if isinstance(impl, B200AttentionImpl):
                    impl.enable_replica_split(rs, self.max_num_seqs)

    def build(self, common_prefix_len: int, common_attn_metadata: CommonAttentionMetadata,
              fast_build: bool = False) -> B200AttentionMetadata:
        cam = common_attn_metadata
        nd, np_, ndt, npt = split_decodes_and_prefills(cam, decode_threshold=1)
        # 0 = the library's default schedule (stream-K: KV-tile chunks per warp of a persistent grid, nothing
        # to choose on the host; falls back to fixed splits sized to the workspace)
        splits = 0
        planned = False
        if nd > 0:
            # one plan launch per step instead of one per layer (what FlashAttention's scheduler metadata /
            # FlashInfer's plan() are to the stock builders, flash_attn.py:388-575); no host sync
            ops.decode_plan(cam.seq_lens, cam.block_table_tensor, nd, self.num_heads_q, self.num_heads_kv,
                            self.block_size, self.workspace)
            planned = True
        pqsl, pmax, ptotal = None, 0, 0
        if np_ > 0:
            pqsl = cam.query_start_loc[nd:] - ndt
            cpu = cam.query_start_loc_cpu
            pmax = int((cpu[nd + 1:] - cpu[nd:-1]).max()) if nd > 0 else cam.max_query_len
            ub = getattr(cam, "seq_lens_cpu_upper_bound", None)  # host copy without a sync (newer vLLM only)
            ptotal = int(ub[nd:nd + np_].sum()) if ub is not None else 0
        return B200AttentionMetadata(
            num_actual_tokens=cam.num_actual_tokens, max_query_len=cam.max_query_len,
            query_start_loc=cam.query_start_loc, max_seq_len=cam.max_seq_len, seq_lens=cam.seq_lens,
            block_table=cam.block_table_tensor, slot_mapping=cam.slot_mapping, num_decodes=nd,
            num_decode_tokens=ndt, num_prefills=np_, num_prefill_tokens=npt,
            prefill_query_start_loc=pqsl, prefill_max_query_len=pmax, decode_num_splits=splits,
            workspace=self.workspace, causal=cam.causal, decode_planned=planned,
            prefill_staging=self.prefill_staging, prefill_total_kv_tokens=ptotal)


def _batch_sizes(max_num_seqs: int):
    b = 1
    while b < max_num_seqs:
        yield b
        b = b + 1 if b < 64 else b + 8
    yield max(1, max_num_seqs)


class B200AttentionImpl(AttentionImpl[B200AttentionMetadata]):
    """KV-cache write + paged attention through the C ABI (vllm/v1/attention/backend.py:763-841)."""

    can_return_lse_for_decode: bool = True
    # keep Q in bf16: Attention.forward then skips the fp8 query quantisation (attention.py:461-471)
    supports_quant_query_input: bool = False

    def __init__(self, num_heads: int, head_size: int, scale: float, num_kv_heads: int | None = None,
                 alibi_slopes: list[float] | None = None, sliding_window: int | None = None,
                 kv_cache_dtype: str = "auto", logits_soft_cap: float | None = None,
                 attn_type: str = AttentionType.DECODER,
                 kv_sharing_target_layer_name: str | None = None, **extra) -> None:
        self.num_heads = num_heads
        self.head_size = head_size
        self.scale = float(scale)
        self.num_kv_heads = num_kv_heads if num_kv_heads is not None else num_heads
        self.kv_cache_dtype = kv_cache_dtype
        self.kv_sharing_target_layer_name = kv_sharing_target_layer_name
        self.attn_type = attn_type
        unsupported = []
        if alibi_slopes is not None:
            unsupported.append("alibi_slopes")
        if sliding_window is not None:
            unsupported.append("sliding_window")
        if logits_soft_cap:
            unsupported.append("logits_soft_cap")
        if extra.get("sinks") is not None:
            unsupported.append("sinks")
        if attn_type != AttentionType.DECODER:
            unsupported.append(f"attn_type={attn_type}")
        if self.num_heads % self.num_kv_heads != 0:
            unsupported.append("num_heads not a multiple of num_kv_heads")
        if unsupported:
            raise NotImplementedError(f"B200AttentionImpl does not support: {', '.join(unsupported)}")
        self._kv_code = ops.kv_dtype_code(kv_cache_dtype)
        # optional: batch split inside a KV-head replica group (tp > total_kv_heads), see tp.ReplicaSplit
        self.replica_split = None
        self._split_ws: ops.DecodeWorkspace | None = None
        self._split_q = self._split_out = None
        self._q_rotated_by_update = False

    def enable_replica_split(self, replica_split, max_batch: int = 0) -> None:
        """Serve decode rows with the batch split over the ranks that replicate this KV head instead of
        every rank streaming the whole head (tp.ReplicaSplit; identical results, 1/R of the KV bytes).
        `max_batch` > 0 sizes the staging buffers and the decode workspace now (nothing is allocated in forward(),
        which may run under CUDA-graph capture); the live decode batch may then be anything up to it."""
        self.replica_split = replica_split
        self._split_q = self._split_out = None
        if max_batch > 0:
            rs = replica_split
            per = -(-max_batch // rs.R)
            ex = rs.exchange
            dev = ex.q_grp.device if ex is not None else torch.device("cuda", torch.cuda.current_device())
            if ex is None:
                self._split_q = torch.zeros(per, rs.group_q_heads, self.head_size, dtype=torch.bfloat16, device=dev)
            self._split_out = (ex.out_grp if ex is not None else
                               torch.zeros(per, rs.group_q_heads, self.head_size, dtype=torch.bfloat16, device=dev))
            self._split_ws = ops.DecodeWorkspace(dev, ops.decode_workspace_bytes(per, rs.group_q_heads, self.head_size, 1))

    # ------------------------------------------------------------------ cache write
    def do_kv_cache_update(self, layer, key: torch.Tensor, value: torch.Tensor,
                           kv_cache: torch.Tensor, slot_mapping: torch.Tensor) -> None:
        """`key`/`value` may be padded past slot_mapping (flash_attn.py:850-883): the token count is
        slot_mapping.shape[0]. Skipped when this layer shares another layer's cache."""
        if self.kv_sharing_target_layer_name is not None:
            return
        self._q_rotated_by_update = False   # q is final before this kernel starts (see forward: early_q_read)
        k_cache, v_cache = ops.split_kv_cache(kv_cache)
        ops.reshape_and_cache(key, value, k_cache, v_cache, slot_mapping, self._kv_code,
                              getattr(layer, "_k_scale", None), getattr(layer, "_v_scale", None))

    # ------------------------------------------------------------------ fused RoPE + cache write
    def fused_rope_kvcache_supported(self) -> bool:
        """vLLM's RopeKVCacheFusionPass may replace rotary_emb + unified_kv_cache_update by one call of
        do_rope_and_kv_cache_update (vllm/compilation/passes/fusion/rope_kvcache_fusion.py)."""
        return self.kv_sharing_target_layer_name is None

    def do_rope_and_kv_cache_update(self, layer, query: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
                                    positions: torch.Tensor, cos_sin_cache: torch.Tensor, is_neox: bool,
                                    kv_cache: torch.Tensor, layer_slot_mapping: torch.Tensor) -> None:
        """In-place RoPE on query/key + cache write of (rotated key, value) in ONE launch
        (vllm/v1/attention/backend.py:823-840). query/key/value may arrive flat ([T, H*d])."""
        self._q_rotated_by_update = True    # this launch rotates q in place: the decode may not read q ahead of it
        k_cache, v_cache = ops.split_kv_cache(kv_cache)
        q = query.view(query.shape[0], -1, self.head_size)
        k = key.view(key.shape[0], -1, self.head_size)
        v = value.view(value.shape[0], -1, self.head_size)
        cs = cos_sin_cache if cos_sin_cache.dtype == torch.bfloat16 else self._cos_sin_bf16(cos_sin_cache)
        ops.rope_and_cache(positions, q, k, v, cs, is_neox, k_cache, v_cache, layer_slot_mapping, self._kv_code,
                           getattr(layer, "_k_scale", None), getattr(layer, "_v_scale", None))

    def _cos_sin_bf16(self, cos_sin_cache: torch.Tensor) -> torch.Tensor:
        """vLLM casts the table to the activation dtype on every call (`_match_cos_sin_cache_dtype`);
        keep one bf16 copy per table instead."""
        key = (cos_sin_cache.data_ptr(), tuple(cos_sin_cache.shape))
        cache = self.__dict__.setdefault("_cs_cache", {})
        if key not in cache:
            cache[key] = cos_sin_cache.to(torch.bfloat16)
        return cache[key]

    # ------------------------------------------------------------------ attention
    def forward(self, layer, query: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
                kv_cache: torch.Tensor, attn_metadata: B200AttentionMetadata,
                output: torch.Tensor, output_scale: torch.Tensor | None = None,
                output_block_scale: torch.Tensor | None = None) -> torch.Tensor:
        if output_scale is not None or output_block_scale is not None:
            raise NotImplementedError("fused output quantization is not supported by B200AttentionImpl")
        if attn_metadata is None:  # profiling run
            return output.fill_(0)
        md = attn_metadata
        k_cache, v_cache = ops.split_kv_cache(kv_cache)
        q = query.view(-1, self.num_heads, self.head_size)
        out = output.view(-1, self.num_heads, self.head_size)
        k_scale = getattr(layer, "_k_scale", None)
        v_scale = getattr(layer, "_v_scale", None)
        self.last_scale_numel = None if k_scale is None else int(k_scale.numel())   # 1 = per tensor, Hkv = per head
        nd, ndt = md.num_decodes, md.num_decode_tokens
        if nd > 0 and self.replica_split is not None:
            rs = self.replica_split
            per, b0, b1 = rs.plan(nd)
            q_grp = rs.gather_q(q[:ndt], out=None if self._split_q is None else self._split_q[:per])
            if self._split_out is not None:      # preallocated (enable_replica_split(max_batch)): nothing allocated here
                out_grp = self._split_out[:per]
                if b1 - b0 < per:
                    out_grp[b1 - b0:].zero_()
            else:
                out_grp = torch.zeros_like(q_grp) if b1 - b0 < per else torch.empty_like(q_grp)
            if b1 > b0:
                if self._split_ws is None:  # sized for the group's head count, batch-independent (stream-K)
                    self._split_ws = ops.DecodeWorkspace(q.device, ops.decode_workspace_bytes(
                        per, rs.group_q_heads, self.head_size, 1))
                ops.paged_decode(q_grp[:b1 - b0], k_cache, v_cache, md.block_table[b0:b1], md.seq_lens[b0:b1],
                                 md.max_seq_len, self.scale, out=out_grp[:b1 - b0], kv_cache_dtype=self._kv_code,
                                 k_scale=k_scale, v_scale=v_scale, num_splits=md.decode_num_splits,
                                 workspace=self._split_ws)
            rs.scatter_out(out_grp, nd, out=out[:ndt])
        elif nd > 0:
            # early_kv_read: the only KV bytes this step's cache write touched in a decoded sequence are its
            # newest token's (vLLM writes slot_mapping = position seq_len - 1 for a decode), see b200attn.h
            ops.paged_decode(q[:ndt], k_cache, v_cache, md.block_table, md.seq_lens, md.max_seq_len,
                             self.scale, out=out[:ndt], kv_cache_dtype=self._kv_code,
                             k_scale=k_scale, v_scale=v_scale, num_splits=md.decode_num_splits,
                             workspace=md.workspace, planned=md.decode_planned and md.decode_num_splits == 0,
                             early_kv_read=True, early_q_read=not self._q_rotated_by_update)
        if md.num_prefills > 0:
            n = md.num_actual_tokens
            ops.paged_prefill(q[ndt:n], k_cache, v_cache, md.block_table[nd:], md.seq_lens[nd:],
                              md.prefill_query_start_loc, md.prefill_max_query_len, self.scale,
                              out=out[ndt:n], kv_cache_dtype=self._kv_code, k_scale=k_scale,
                              v_scale=v_scale, staging=md.prefill_staging, max_seq_len=md.max_seq_len,
                              total_kv_tokens=md.prefill_total_kv_tokens)
        return output
