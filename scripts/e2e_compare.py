#!/usr/bin/env python
"""Compare two e2e_vllm.py outputs: greedy tokens and logprobs of the chosen tokens."""
import json
import sys

a, b = (json.loads(open(f).read()) for f in sys.argv[1:3])
tol = float(sys.argv[3]) if len(sys.argv) > 3 else 0.05
n_tok = n_same = 0
worst = 0.0
for ra, rb in zip(a["results"], b["results"]):
    for i, (ta, tb) in enumerate(zip(ra["tokens"], rb["tokens"])):
        n_tok += 1
        if ta != tb:
            break  # after the first divergence the continuations are different sequences
        n_same += 1
        worst = max(worst, abs(ra["logprobs"][i] - rb["logprobs"][i]))
lp = [x for r in a["results"] for x in r["logprobs"]]
print(json.dumps({"a": a["impl"], "b": b["impl"], "kv": a["kv_cache_dtype"], "graphs": a["graphs"],
                  "tokens_compared": n_tok, "tokens_equal_until_divergence": n_same,
                  "max_abs_logprob_diff": worst, "logprob_range": [min(lp), max(lp)]}))
ok = n_same >= 0.9 * n_tok and worst <= tol and max(lp) - min(lp) > 0.5
sys.exit(0 if ok else 1)
