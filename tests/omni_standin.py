"""Test-harness model for the e2e run inside vllm-omni: the reference's own Qwen2 decoder
(vllm_omni/model_executor/models/qwen2_5_omni/qwen2_old.py — the attention caller SURVEY §8(a1/a2) cites), with the
one thing its thinker wrappers add for `GPUARModelRunner`: `forward()` accepts the runner's extra keyword arguments
(sampling_metadata, additional_information, ... gpu_ar_model_runner.py:511) and passes the decoder what it takes.
Worker, model runner, attention module and decoder layers are the reference's, unmodified."""
from vllm_omni.model_executor.models.qwen2_5_omni.qwen2_old import Qwen2ForCausalLM as _RefQwen2ForCausalLM


class ThinkerStandIn(_RefQwen2ForCausalLM):
    def forward(self, input_ids, positions, intermediate_tensors=None, inputs_embeds=None, **omni_kwargs):
        return super().forward(input_ids, positions, intermediate_tensors, inputs_embeds)
