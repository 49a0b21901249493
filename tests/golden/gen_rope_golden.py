#!/usr/bin/env python
"""Golden vectors for the rotary embedding from the third-party module that owns this arithmetic on
the reference's path: vLLM's RotaryEmbedding / MRotaryEmbedding `forward_native` (CPU, bf16), as
called by the reference's attention layers (vllm_omni/.../qwen2_old.py:151-153; M-RoPE positions come
from vllm_omni/model_executor/layers/rotary_embedding/mrope.py). Run in the build container
(vLLM 0.22 installed): python tests/golden/gen_rope_golden.py -> tests/golden/rope_*.npz"""
from pathlib import Path

import numpy as np
import torch
from vllm.config import VllmConfig, set_current_vllm_config
from vllm.model_executor.layers.rotary_embedding import get_rope

OUT = Path(__file__).resolve().parent
T, HQ, HKV, D = 9, 4, 2, 128


def bits(t):
    return t.contiguous().view(torch.int16).numpy()


def case(name, rope_parameters, is_neox, positions):
    from vllm.config import DeviceConfig
    with set_current_vllm_config(VllmConfig(device_config=DeviceConfig(device="cpu"))):
        rope = get_rope(D, 4096, is_neox, rope_parameters, dtype=torch.bfloat16)
    g = torch.Generator().manual_seed(7)
    q = torch.randn(T, HQ * D, generator=g).to(torch.bfloat16)
    k = torch.randn(T, HKV * D, generator=g).to(torch.bfloat16)
    qo, ko = rope.forward_native(positions, q.clone(), k.clone())
    np.savez_compressed(
        OUT / f"rope_{name}.npz", q=bits(q), k=bits(k), q_out=bits(qo), k_out=bits(ko), positions=positions.numpy(),
        cos_sin_cache=bits(rope.cos_sin_cache.to(torch.bfloat16)), rot_dim=rope.rotary_dim, is_neox=int(is_neox),
        mrope_section=np.array(getattr(rope, "mrope_section", None) or [0, 0, 0]),
        mrope_interleaved=int(getattr(rope, "mrope_interleaved", False)), cls=type(rope).__name__)
    print(name, type(rope).__name__, rope.rotary_dim)


pos1 = torch.tensor([0, 1, 2, 17, 255, 1000, 4095, 3, 3], dtype=torch.int64)
pos3 = torch.stack([pos1, (pos1 * 7 + 3) % 4096, (pos1 * 13 + 5) % 4096])
case("neox", {"rope_type": "default", "rope_theta": 1e6}, True, pos1)
case("gptj", {"rope_type": "default", "rope_theta": 1e4}, False, pos1)
case("partial", {"rope_type": "default", "rope_theta": 1e4, "partial_rotary_factor": 0.5}, True, pos1)
case("mrope", {"rope_type": "default", "rope_theta": 1e6, "mrope_section": [16, 24, 24]}, True, pos3)
case("mrope_interleaved", {"rope_type": "default", "rope_theta": 1e6, "mrope_section": [24, 20, 20],
                           "mrope_interleaved": True}, True, pos3)
case("mrope_text_only", {"rope_type": "default", "rope_theta": 1e6, "mrope_section": [16, 24, 24]}, True, pos1)
