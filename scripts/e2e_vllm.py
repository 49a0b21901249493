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


def main():
    import torch
    from vllm import LLM, SamplingParams

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
