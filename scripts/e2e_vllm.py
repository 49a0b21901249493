#!/usr/bin/env python
"""End-to-end smoke inside vLLM: a random-init Qwen3-style decoder served by
vLLM's own engine with `attention_backend=CUSTOM` (this repo's backend) or a stock backend, on the
same prompts; prints one JSON line with the greedy tokens and their logprobs so that two runs can
be compared. No checkpoint, no tokenizer, no network: config.json is written to a temp dir,
`load_format="dummy"` (vLLM's seeded dummy initialiser gives identical weights in every run).

  python scripts/e2e_vllm.py --backend CUSTOM [--kv-cache-dtype fp8] [--graphs] --out file.json
"""
import argparse
import json
import os
import sys
import tempfile
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))

ap = argparse.ArgumentParser()
ap.add_argument("--backend", default="CUSTOM")
ap.add_argument("--kv-cache-dtype", default="auto")
ap.add_argument("--graphs", action="store_true", help="CUDA-graph decode (no torch.compile) instead of eager")
ap.add_argument("--max-tokens", type=int, default=12)
ap.add_argument("--layers", type=int, default=4)
ap.add_argument("--omni", action="store_true",
                help="run inside vllm-omni: the reference's OmniEngineArgs (its plugin loader finds this repo's entry "
                     "point), GPUARWorker / GPUARModelRunner, and its own Qwen2 decoder (qwen2_old.py) as the model")
ap.add_argument("--kv-scales", default="", help="with --omni: 'per-head' = the decoder is built with a compressed-tensors "
                "KV-cache scheme (strategy attn_head) routed through the reference's ComponentQuantizationConfig, and "
                "its per-KV-head k/v scales are set to very different values per head")
ap.add_argument("--out", default="")
args = ap.parse_args()

os.environ.setdefault("VLLM_ENABLE_V1_MULTIPROCESSING", "0")  # engine core in this process
os.environ.setdefault("VLLM_USE_FLASHINFER_SAMPLER", "0")
os.environ.setdefault("VLLM_LOGGING_LEVEL", "WARNING")
os.environ.setdefault("TOKENIZERS_PARALLELISM", "false")

CONFIG = {
    "architectures": ["Qwen3ForCausalLM"], "model_type": "qwen3", "vocab_size": 2048, "hidden_size": 1024,
    "intermediate_size": 2048, "num_hidden_layers": args.layers, "num_attention_heads": 16,
    "num_key_value_heads": 2, "head_dim": 128, "max_position_embeddings": 4096, "rms_norm_eps": 1e-6,
    "rope_theta": 1000000.0, "tie_word_embeddings": False, "torch_dtype": "bfloat16", "hidden_act": "silu",
    "attention_bias": False, "bos_token_id": 1, "eos_token_id": 2,
}


OMNI_CONFIG = {  # the reference's own decoder for its Qwen2.5-Omni thinker (qwen2_old.py: bias in qkv, no q/k norm)
    "architectures": ["B200ThinkerStandIn"], "model_type": "qwen2", "vocab_size": 2048, "hidden_size": 2048,
    "intermediate_size": 2048, "num_hidden_layers": args.layers, "num_attention_heads": 16,
    "num_key_value_heads": 2, "max_position_embeddings": 4096, "rms_norm_eps": 1e-6,
    "rope_theta": 1000000.0, "tie_word_embeddings": False, "torch_dtype": "bfloat16", "hidden_act": "silu",
    "bos_token_id": 1, "eos_token_id": 2, "sliding_window": None, "use_sliding_window": False,
}


def run_omni():
    """The same prompts through the reference's own engine arguments, worker and AR model runner."""
    import torch
    from tests import ref_import
    stubbed = ref_import.setup()
    import vllm_omni  # noqa: F401  (applies the reference's patches to vLLM)
    from vllm import SamplingParams
    from vllm.v1.engine.llm_engine import LLMEngine
    from vllm_omni.engine.arg_utils import OmniEngineArgs
    from vllm_omni.engine.stage_init_utils import extract_stage_metadata  # noqa: F401  (where a backend is forced)

    # the reference's decoder with a forward() that takes the AR runner's keyword arguments (tests/omni_standin.py)
    from vllm_omni.model_executor.models.registry import OmniModelRegistry
    from vllm import ModelRegistry
    for reg in (OmniModelRegistry, ModelRegistry):   # (the reference mirrors its own models into vLLM's registry too)
        reg.register_model("B200ThinkerStandIn", "tests.omni_standin:ThinkerStandIn")
    d = tempfile.mkdtemp(prefix="tiny_qwen2_old_")
    cfg = dict(OMNI_CONFIG)
    per_head = args.kv_scales == "per-head"
    if per_head:
        # what an llm-compressor checkpoint with an fp8 KV cache and one scale per attention head declares
        cfg["quantization_config"] = {
            "quant_method": "compressed-tensors", "format": "dense", "config_groups": {}, "ignore": [],
            "kv_cache_scheme": {"num_bits": 8, "type": "float", "strategy": "attn_head", "symmetric": True,
                                "dynamic": False}}
    Path(d, "config.json").write_text(json.dumps(cfg))
    # stage engine_args as a stage YAML would give them (qwen3_omni_moe.yaml:14-27), with the backend knob
    stage_args = dict(model=d, load_format="dummy", skip_tokenizer_init=True, dtype="bfloat16", max_model_len=2048,
                      gpu_memory_utilization=0.15, kv_cache_dtype=args.kv_cache_dtype, seed=0, max_num_seqs=16,
                      enable_prefix_caching=False, model_stage="thinker", model_arch="B200ThinkerStandIn",
                      worker_type="ar", engine_output_type="text",
                      worker_cls="vllm_omni.worker.gpu_ar_worker.GPUARWorker")
    if args.backend == "CUSTOM":
        from vllm_omni_b200 import plugin
        stage_args = plugin.force_for_stage(stage_args)     # next to stage_init_utils.py:282-293's ROCm precedent
    else:
        stage_args["attention_backend"] = args.backend
    if args.graphs:
        stage_args["compilation_config"] = {"mode": 0, "cudagraph_mode": "FULL_DECODE_ONLY",
                                            "cudagraph_capture_sizes": [1, 2, 4, 8]}
    else:
        stage_args["enforce_eager"] = True
    # vLLM version skew, outside the attention path: the reference pins vLLM 0.19.0 (docker/Dockerfile.cuda:1), this
    # image has 0.22, whose GPUModelRunner._bookkeeping_sync no longer takes spec_decode_metadata
    import inspect
    from vllm.v1.worker.gpu_model_runner import GPUModelRunner
    from vllm_omni.worker.gpu_ar_model_runner import GPUARModelRunner
    n_params = len(inspect.signature(GPUModelRunner._bookkeeping_sync).parameters)
    skew = []
    if n_params == 6:
        base_sync = GPUModelRunner._bookkeeping_sync
        GPUARModelRunner._bookkeeping_sync = lambda self, *a: base_sync(self, *a[:5])
        skew.append("_bookkeeping_sync(spec_decode_metadata dropped)")
    engine_args = OmniEngineArgs(**stage_args)   # __post_init__ runs vllm_omni.plugins.load_omni_general_plugins()
    routed = None
    if per_head:
        # vllm_omni/quantization: per-component routing by longest layer-name prefix (component_config.py:63-82);
        # the decoder lives under "model.", so its attention layers resolve to the checkpoint's config and
        # everything else (lm_head) to "no quantisation"
        from vllm.usage.usage_lib import UsageContext
        from vllm_omni.quantization.component_config import ComponentQuantizationConfig
        vllm_config = engine_args.create_engine_config(UsageContext.LLM_CLASS)
        inner = vllm_config.quant_config
        assert inner is not None and getattr(inner, "kv_cache_scheme", None), "checkpoint KV-cache scheme was not parsed"
        vllm_config.quant_config = ComponentQuantizationConfig({"model": inner}, default_config=None)
        routed = {"attn": type(vllm_config.quant_config.resolve("model.layers.0.self_attn.attn")).__name__,
                  "lm_head": type(vllm_config.quant_config.resolve("lm_head")).__name__}
        engine = LLMEngine.from_vllm_config(vllm_config, UsageContext.LLM_CLASS)
    else:
        engine = LLMEngine.from_engine_args(engine_args)
    worker = engine.engine_core.engine_core.model_executor.driver_worker.worker
    mr = worker.model_runner
    gen = torch.Generator(device="cuda").manual_seed(1234)
    with torch.no_grad():
        for name, prm in sorted(mr.model.named_parameters()):
            if prm.ndim == 1 and "norm" in name:
                prm.fill_(1.0)
            elif prm.ndim == 1:
                prm.zero_()                                   # qkv bias
            else:
                prm.copy_((torch.randn(prm.shape, generator=gen, device="cuda") * 0.03).to(prm.dtype))
    scales = None
    if per_head:
        # the scales a checkpoint would carry, one per KV head and very different between the heads: reading the
        # wrong head's scale is a 16x error in that head's scores / values
        ks = torch.tensor([0.25, 4.0], device="cuda")
        vs = torch.tensor([4.0, 0.25], device="cuda")
        for layer in mr.compilation_config.static_forward_context.values():
            assert layer._k_scale.numel() == 2 and layer._v_scale.numel() == 2, "per-head scales did not reach the layer"
            with torch.no_grad():
                layer._k_scale.copy_(ks)
                layer._v_scale.copy_(vs)
        scales = {"k": ks.tolist(), "v": vs.tolist()}
    g = torch.Generator().manual_seed(0)
    lens = [5, 37, 130, 1, 700, 64, 33]
    sp = SamplingParams(temperature=0.0, max_tokens=args.max_tokens, logprobs=1, detokenize=False, ignore_eos=True)
    for i, n in enumerate(lens):
        engine.add_request(str(i), {"prompt_token_ids": torch.randint(3, OMNI_CONFIG["vocab_size"], (n,), generator=g).tolist()}, sp)
    done = {}
    while engine.has_unfinished_requests():
        for o in engine.step():
            if o.finished:
                done[int(o.request_id)] = o
    res = []
    for i in range(len(lens)):
        c = done[i].outputs[0]
        res.append({"tokens": list(c.token_ids), "logprobs": [next(iter(lp.values())).logprob for lp in c.logprobs]})
    layers = mr.compilation_config.static_forward_context
    attn = next(iter(layers.values()))
    builders = [type(b).__name__ for grp in getattr(mr, "attn_groups", []) for g_ in grp for b in getattr(g_, "metadata_builders", [])]
    from vllm_omni import plugins as omni_plugins
    line = json.dumps({"backend": args.backend, "impl": type(attn.impl).__name__, "kv_cache_dtype": args.kv_cache_dtype,
                       "graphs": args.graphs, "omni": True, "worker": type(worker).__name__,
                       "model_runner": type(mr).__name__, "model": type(mr.model).__module__ + "." + type(mr.model).__name__,
                       "model_config": type(engine.vllm_config.model_config).__name__,
                       "metadata_builders": sorted(set(builders)), "omni_plugins_loaded": bool(omni_plugins.omni_plugins_loaded),
                       "stubbed_imports": sorted(stubbed), "vllm_version_skew_adapters": skew,
                       "quant_routing": routed, "kv_scales": scales,
                       "impl_scale_numel": getattr(attn.impl, "last_scale_numel", None),
                       "results": res})
    if args.out:
        Path(args.out).write_text(line + "\n")
    print(line, flush=True)


def main():
    import torch
    from vllm import LLM, SamplingParams

    if args.omni:
        return run_omni()
    if args.backend == "CUSTOM":
        from vllm_omni_b200 import plugin
        plugin.register()
    d = tempfile.mkdtemp(prefix="tiny_qwen3_")
    Path(d, "config.json").write_text(json.dumps(CONFIG))
    kw = dict(model=d, load_format="dummy", skip_tokenizer_init=True, dtype="bfloat16", max_model_len=2048,
              gpu_memory_utilization=0.15, attention_backend=args.backend, kv_cache_dtype=args.kv_cache_dtype,
              seed=0, max_num_seqs=16, enable_prefix_caching=False)
    if args.graphs:
        kw["compilation_config"] = {"mode": 0, "cudagraph_mode": "FULL_DECODE_ONLY", "cudagraph_capture_sizes": [1, 2, 4, 8]}
    else:
        kw["enforce_eager"] = True
    llm = LLM(**kw)
    # vLLM's dummy initialiser draws weights from U(-1e-3, 1e-3): the logits come out uniform and
    # attention cannot change a token. Re-draw them (seeded, identical in every run) at a scale that
    # gives peaked logits and O(1) attention scores.
    mr = llm.llm_engine.model_executor.driver_worker.worker.model_runner
    gen = torch.Generator(device="cuda").manual_seed(1234)
    with torch.no_grad():
        for name, prm in sorted(mr.model.named_parameters()):
            if prm.ndim == 1:  # RMSNorm weights (incl. q_norm / k_norm)
                prm.fill_(1.0)
            else:
                prm.copy_((torch.randn(prm.shape, generator=gen, device="cuda") * 0.04).to(prm.dtype))
    g = torch.Generator().manual_seed(0)
    lens = [5, 37, 130, 1, 700, 64, 33]
    prompts = [{"prompt_token_ids": torch.randint(3, CONFIG["vocab_size"], (n,), generator=g).tolist()} for n in lens]
    sp = SamplingParams(temperature=0.0, max_tokens=args.max_tokens, logprobs=1, detokenize=False, ignore_eos=True)
    outs = llm.generate(prompts, sp, use_tqdm=False)
    res = []
    for o in outs:
        c = o.outputs[0]
        lps = [next(iter(lp.values())).logprob for lp in c.logprobs]
        res.append({"tokens": list(c.token_ids), "logprobs": lps})
    # which attention implementation actually ran
    impl = "?"
    try:
        from vllm.config import get_layers_from_vllm_config  # noqa: F401
        mr = llm.llm_engine.model_executor.driver_worker.worker.model_runner
        layers = mr.compilation_config.static_forward_context
        impl = type(next(iter(layers.values())).impl).__name__
    except Exception as e:  # pragma: no cover - vLLM internals differ between releases
        impl = f"unknown ({type(e).__name__})"
    line = json.dumps({"backend": args.backend, "impl": impl, "kv_cache_dtype": args.kv_cache_dtype,
                       "graphs": args.graphs, "results": res})
    if args.out:
        Path(args.out).write_text(line + "\n")
    print(line, flush=True)


if __name__ == "__main__":
    main()
