"""Drives the plugin (B200AttentionImpl) from HOST buffers, one decode step over L layers.

This is the end-to-end shape of the hot path as a user of the reference-facing interface sees it:
per layer, the step's q/k/v arrive in pinned host memory, are copied to the device, go through
`impl.do_kv_cache_update` + `impl.forward` exactly as vLLM's `Attention.forward` calls them
(vllm/model_executor/layers/attention/attention.py:509-528), and the attention output is copied back
to pinned host memory. Copies run on their own streams, double-buffered, so layer l+1's H2D and layer
l-1's D2H overlap layer l's kernels.

`capture()` records one whole step — the H2D copies, every kernel and the D2H copies, on all three
streams — into a CUDA graph bound to the caller's pinned staging buffers, the way vLLM replays decode
steps (persistent input buffers + FULL cudagraphs, `qwen3_omni_moe.yaml:20`): `replay()` then costs one
launch on the host, and the copies are still part of every step.
"""
from __future__ import annotations

import types

import torch


class HostStepRunner:
    def __init__(self, impl, metadata, kv_caches: list[torch.Tensor], batch: int, num_q_heads: int,
                 num_kv_heads: int, head_size: int, slot_mapping: torch.Tensor,
                 k_scale: torch.Tensor | None = None, v_scale: torch.Tensor | None = None, after_layer=None,
                 plan_per_step: bool = False):
        dev = kv_caches[0].device
        self.impl, self.md, self.kv_caches, self.slot_mapping = impl, metadata, kv_caches, slot_mapping
        self.after_layer = after_layer  # e.g. the o_proj all-reduce of a tensor-parallel layer
        # one stream-K plan launch per step, as B200AttentionMetadataBuilder.build does for a scheduler step
        self.plan_per_step = plan_per_step and metadata.num_decodes > 0
        metadata.decode_planned = self.plan_per_step
        self.block_size = kv_caches[0].shape[-3] if kv_caches[0].ndim == 5 else None
        one = torch.ones(1, dtype=torch.float32, device=dev)
        self.layer = types.SimpleNamespace(_k_scale=one if k_scale is None else k_scale,
                                           _v_scale=one if v_scale is None else v_scale)
        B, hq, hkv, d = batch, num_q_heads, num_kv_heads, head_size
        self.width = (hq + 2 * hkv) * d
        self.split = [hq * d, hkv * d, hkv * d]
        self.shapes = (B, hq, hkv, d)
        # fused qkv rows, as the model's qkv_proj produces them
        self.dev_in = [torch.empty(B, self.width, dtype=torch.bfloat16, device=dev) for _ in range(2)]
        self.dev_out = [torch.empty(B, hq * d, dtype=torch.bfloat16, device=dev) for _ in range(2)]
        self.s_in, self.s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
        self.h2d_bytes_per_layer = B * self.width * 2
        self.d2h_bytes_per_layer = B * hq * d * 2

    def step(self, host_qkv: list[torch.Tensor], host_out: list[torch.Tensor]) -> None:
        """host_qkv[l]: pinned [B, (Hq+2Hkv)*d] bf16; host_out[l]: pinned [B, Hq*d] bf16 (written)."""
        B, hq, hkv, d = self.shapes
        cur = torch.cuda.current_stream()
        L = len(self.kv_caches)
        in_ready = [torch.cuda.Event() for _ in range(2)]
        in_free = [torch.cuda.Event() for _ in range(2)]
        out_ready = [torch.cuda.Event() for _ in range(2)]
        out_free = [torch.cuda.Event() for _ in range(2)]
        self.s_in.wait_stream(cur)
        self.s_out.wait_stream(cur)
        if self.plan_per_step:
            from . import ops
            md = self.md
            ops.decode_plan(md.seq_lens, md.block_table, md.num_decodes, hq, hkv, self.block_size, md.workspace)
        with torch.cuda.stream(self.s_in):
            self.dev_in[0].copy_(host_qkv[0], non_blocking=True)
            in_ready[0].record(self.s_in)
        for l in range(L):
            b = l & 1
            if l + 1 < L:  # prefetch the next layer's inputs
                with torch.cuda.stream(self.s_in):
                    if l >= 1:
                        self.s_in.wait_event(in_free[b ^ 1])
                    self.dev_in[b ^ 1].copy_(host_qkv[l + 1], non_blocking=True)
                    in_ready[b ^ 1].record(self.s_in)
            cur.wait_event(in_ready[b])
            if l >= 2:
                cur.wait_event(out_free[b])
            q, k, v = self.dev_in[b].split(self.split, dim=-1)
            k, v = k.view(B, hkv, d), v.view(B, hkv, d)
            self.impl.do_kv_cache_update(self.layer, k, v, self.kv_caches[l], self.slot_mapping)
            self.impl.forward(self.layer, q.view(B, hq, d), k, v, self.kv_caches[l], self.md, self.dev_out[b])
            if self.after_layer is not None:
                self.after_layer()
            in_free[b].record(cur)
            out_ready[b].record(cur)
            with torch.cuda.stream(self.s_out):
                self.s_out.wait_event(out_ready[b])
                host_out[l].copy_(self.dev_out[b], non_blocking=True)
                out_free[b].record(self.s_out)
        cur.wait_stream(self.s_out)
        cur.wait_stream(self.s_in)

    def capture(self, host_qkv: list[torch.Tensor], host_out: list[torch.Tensor]) -> torch.cuda.CUDAGraph:
        """One step over these pinned buffers as a CUDA graph (copy nodes included). Refill `host_qkv`
        in place between replays; results land in `host_out` when the replay's stream work completes."""
        for t in (*host_qkv, *host_out):
            if not t.is_pinned():
                raise ValueError("capture() needs pinned host buffers (their addresses are baked into the graph)")
        self.step(host_qkv, host_out)       # lazy initialisation (tensor maps, workspaces) outside the capture
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            self.step(host_qkv, host_out)
        return graph
