#!/usr/bin/env python
"""bench.py — decode throughput of the B200 paged-attention hot path (BASELINE.json metric).

A "step" is one decode step of the hot path over one synthetic batch: ONE stream-K plan launch, then for
each of the model's L layers the KV-cache write of the new token (reshape_and_cache) followed by paged
decode attention (decode + LSE-reduce kernels) over that layer's own KV cache (+ the o_proj all-reduce
stand-in and the replica-group q/out exchange at N > 1). Inputs are resident in HBM when the timed region
starts; the L distinct layer caches (>= 24 GB at cfg3) are far larger than L2.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg3] [--impl reference]
  torchrun --nproc-per-node N ... bench.py --gpus N ...        (one rank per GPU, KV-head TP)

Prints ONE JSON line on rank 0 (see README / DESIGN.md for the fields).
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

import torch  # noqa: E402

METRIC, UNIT = "decode_tokens_per_s", "tokens/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg3")
    ap.add_argument("--layers", type=int, default=0, help="0 = the model's layer count")
    ap.add_argument("--ragged", action="store_true", help="seq_lens ~ U[ctx/2, ctx] instead of uniform")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: global batch = workload batch x N (per-GPU KV bytes fixed up to the "
                         "KV-head count); strong: global batch fixed (the other one is reported under `extras`)")
    ap.add_argument("--variant", type=int, default=0, help="0 auto, 1 TMA, 2 cp.async, 3 stream, 4 stream-K v1, 5 stream-K v2, 6 cooperative stream-K (default)")
    ap.add_argument("--splits", type=int, default=0, help="0 = heuristic")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the extra arms (ragged / strong scaling / "
                    "north_star collectives) reported under `extras`")
    ap.add_argument("--no-verify", action="store_true", help="skip the output check against a 1-rank run")
    ap.add_argument("--contiguous-blocks", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="eager launches instead of CUDA-graph replay")
    ap.add_argument("--replica-mode", default="split", choices=["split", "replicate"],
                    help="when tp > KV heads: 'replicate' = every rank of a replica group streams the whole KV "
                         "head for its own query heads (the reference rule); 'split' = the group splits the "
                         "batch and exchanges q/out (tp.ReplicaSplit: same results, 1/R of the KV bytes per rank)")
    ap.add_argument("--replica-exchange", default="peer", choices=["peer", "nccl"],
                    help="q/out exchange of --replica-mode split: this repo's peer-memory kernels or NCCL all_to_all")
    ap.add_argument("--allreduce", default="peer", choices=["peer", "nccl"],
                    help="the o_proj row-parallel sum after every layer (N > 1): 'peer' = this repo's peer-memory "
                         "kernel over NVLink (csrc/peer_allreduce.cu), 'nccl' = dist.all_reduce")
    ap.add_argument("--total-kv-heads", type=int, default=0, help="override the workload's KV-head count "
                    "(e.g. 1 with --gpus 2 exercises KV-head replication on two GPUs)")
    ap.add_argument("--cpu-budget-s", type=float, default=150.0, help="wall budget of the reference arm's steps")
    return ap.parse_args()


# --------------------------------------------------------------------------------------------
# helpers
# --------------------------------------------------------------------------------------------
def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, torch copy read+write)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def kernel_sources_sha() -> str:
    """sha256 over the decode kernel sources: a committed ncu capture is only quoted while it matches."""
    h = hashlib.sha256()
    for name in ("paged_decode_sk3.cu", "sk_math.cuh", "decode_common.cuh", "ptx.cuh"):
        h.update((ROOT / "vllm_omni_b200" / "csrc" / name).read_bytes())
    return h.hexdigest()[:16]


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled every 100 ms while the step runs back to back: started before the
    warm-up (nvidia-smi needs up to a second before its first line, longer than a short timed region), read by a
    thread, stopped right after the timed region. `count()` = lines so far."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines, self.thread = index, None, [], None
        self.result = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return self

        def pump():
            for line in self.proc.stdout:
                self.lines.append(line)

        self.thread = threading.Thread(target=pump, daemon=True)
        self.thread.start()
        return self

    def count(self) -> int:
        return len(self.lines)

    def __exit__(self, *exc):
        if self.proc is None:
            return
        time.sleep(0.12)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        if self.thread is not None:
            self.thread.join(timeout=2)
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in list(self.lines):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if sm:
            self.result = {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "power_w_max": max(pw),
                           "reasons": sorted(reasons), "samples": len(sm),
                           "window": "last warm-up steps + timed region (the same step replayed back to back)"}


T0 = time.perf_counter()


def log(msg: str) -> None:
    """Progress on stderr (rank 0): a hung run shows where it stopped."""
    if int(os.environ.get("RANK", "0")) == 0:
        print(f"[bench +{time.perf_counter() - T0:6.1f}s] {msg}", file=sys.stderr, flush=True)


def dist_setup(n_gpus: int):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        torch.cuda.set_device(0)
    if world != n_gpus:
        raise SystemExit(f"--gpus {n_gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {n_gpus}")
    return rank, world, local


def max_over_ranks(ms: float, world: int) -> float:
    if world == 1:
        return ms
    import torch.distributed as dist
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier(world: int):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()


# --------------------------------------------------------------------------------------------
# reference arm: the path's CPU implementation on the host cores. vLLM's genuine CPU attention kernels
# (cpu_attn, K12 in SURVEY.md) are absent from the installed CUDA wheel and the reference tree has no native
# sources, so this is a port (kind = "port") of the construction vLLM's CPU backend uses
# (vllm/v1/attention/backends/cpu_attn.py:400-455: gather the pages of every sequence, ONE
# scaled_dot_product_attention with enable_gqa) plus the cache write of the new tokens.
# --------------------------------------------------------------------------------------------
def cpu_layer_inputs(wl, ragged: bool):
    """One layer of the workload, ALL sequences, as CPU tensors (cache written once, outside the timed part)."""
    from oracle import oracle  # the checker's slot-write restatement fills the cache (set-up only)
    from vllm_omni_b200 import layout, workloads
    lens = workloads.make_seq_lens(wl.batch, wl.ctx, ragged)
    bt, nb = workloads.make_block_table(lens, wl.block_size)
    g = torch.Generator().manual_seed(workloads.SEED_KV)
    ks = vs = None
    if wl.fp8:
        ks, vs = workloads.default_scales(wl.total_kv_heads)
    cache = layout.alloc_kv_cache(nb, wl.block_size, wl.total_kv_heads, wl.head_size, wl.kv_cache_dtype, "cpu")
    l64 = lens.long()
    for b in range(wl.batch):  # per sequence: bounded temporaries
        n = int(l64[b])
        k = torch.randn(n, wl.total_kv_heads, wl.head_size, generator=g).to(torch.bfloat16)
        v = torch.randn(n, wl.total_kv_heads, wl.head_size, generator=g).to(torch.bfloat16)
        slots = workloads.slots_of_positions(bt, torch.full((n,), b), torch.arange(n), wl.block_size)
        oracle.ref_reshape_and_cache(k, v, cache[0], cache[1], slots, wl.fp8, ks, vs)
    q = torch.randn(wl.batch, wl.total_q_heads, wl.head_size, generator=g).to(torch.bfloat16)
    new_k = torch.randn(wl.batch, wl.total_kv_heads, wl.head_size, generator=g).to(torch.bfloat16)
    new_v = torch.randn(wl.batch, wl.total_kv_heads, wl.head_size, generator=g).to(torch.bfloat16)
    last = workloads.slots_of_positions(bt, torch.arange(wl.batch), l64 - 1, wl.block_size)
    return dict(q=q, cache=cache, bt=bt, lens=lens, ks=ks, vs=vs, new_k=new_k, new_v=new_v, last=last,
                uniform=bool((lens == lens[0]).all()))


def cpu_layer(wl, s) -> torch.Tensor:
    """One layer of the hot path on the host: slot write of the new tokens, page gather, one fp32 SDPA."""
    import torch.nn.functional as F
    kc, vc = s["cache"][0], s["cache"][1]          # logical [nb, bs, Hkv, d]
    nb, bs = kc.shape[0], kc.shape[1]
    # cache write (reshape_and_cache semantics; fp8: e4m3(x / scale))
    if wl.fp8:
        kq = (s["new_k"].float() / s["ks"].view(1, -1, 1)).clamp(-448, 448).to(torch.float8_e4m3fn).view(torch.uint8)
        vq = (s["new_v"].float() / s["vs"].view(1, -1, 1)).clamp(-448, 448).to(torch.float8_e4m3fn).view(torch.uint8)
    else:
        kq, vq = s["new_k"], s["new_v"]
    blk, off = s["last"] // bs, s["last"] % bs
    kc[blk, off] = kq
    vc[blk, off] = vq
    scale = wl.head_size ** -0.5

    def deq(x, sc):
        if not wl.fp8:
            return x.float()
        return x.view(torch.float8_e4m3fn).float() * sc.view(1, 1, -1, 1)

    q = s["q"].float().unsqueeze(2)                # [B, Hq, 1, d]
    if s["uniform"]:
        S = int(s["lens"][0])
        pages = (S + bs - 1) // bs
        idx = s["bt"][:, :pages].long()
        K = deq(kc[idx].flatten(1, 2)[:, :S], s["ks"]).transpose(1, 2)     # [B, Hkv, S, d]
        V = deq(vc[idx].flatten(1, 2)[:, :S], s["vs"]).transpose(1, 2)
        return F.scaled_dot_product_attention(q, K, V, scale=scale, enable_gqa=True).squeeze(2)
    outs = []
    for b in range(q.shape[0]):
        S = int(s["lens"][b])
        idx = s["bt"][b, :(S + bs - 1) // bs].long()
        K = deq(kc[idx].flatten(0, 1)[:S].unsqueeze(0), s["ks"]).transpose(1, 2)
        V = deq(vc[idx].flatten(0, 1)[:S].unsqueeze(0), s["vs"]).transpose(1, 2)
        outs.append(F.scaled_dot_product_attention(q[b:b + 1], K, V, scale=scale, enable_gqa=True).squeeze(2))
    return torch.cat(outs)


def time_cpu(wl, ragged: bool, steps: int, warmup: int, budget_s: float, L: int):
    """Times `steps` steps of `layers` layers each (all sequences); `layers` = L if the budget allows, else the
    largest count that fits (the sample is then a stated fraction of the step, and `value` is scaled by it)."""
    cores = len(os.sched_getaffinity(0))
    torch.set_num_threads(cores)
    s = cpu_layer_inputs(wl, ragged)
    cpu_layer(wl, s)                                   # warm the allocator / thread pool
    t0 = time.perf_counter()
    cpu_layer(wl, s)
    t_layer = time.perf_counter() - t0
    layers = max(1, min(L, int(budget_s / max(1e-9, (steps + warmup) * t_layer))))
    for _ in range(warmup):
        for _ in range(layers):
            cpu_layer(wl, s)
    ts = []
    for _ in range(steps):
        t0 = time.perf_counter()
        for _ in range(layers):
            cpu_layer(wl, s)
        ts.append(time.perf_counter() - t0)
    t_step = sum(ts) / len(ts)                         # seconds per timed step (= `layers` layers)
    tok_s = wl.batch / (t_step * L / layers)           # a token needs all L layers
    return dict(tok_s=tok_s, cores=cores, t_step=t_step, layers=layers, t_layer_first=t_layer)


def run_reference(args, wl):
    """`--impl reference`: rank 0 times the CPU port on the host cores; other ranks exit 0 without work."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    L = args.layers or wl.num_layers
    t0 = time.perf_counter()
    r = time_cpu(wl, args.ragged, max(1, args.steps), max(0, args.warmup), args.cpu_budget_s, L)
    frac = r["layers"] / L
    sample = (f"each timed step = all {wl.batch} sequences x {r['layers']} of {L} layers "
              f"(page gather + one torch fp32 SDPA with enable_gqa per layer + the new tokens' slot write, "
              f"{r['cores']} host threads)" + ("" if frac == 1.0 else f"; tokens/s scaled by {L}/{r['layers']} layers"))
    line = {
        "impl": "reference", "metric": METRIC, "value": r["tok_s"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["t_step"] * 1e3,
        "ms_per_step_note": "measured wall time of one timed step (the sample described in cpu_baseline.sample)",
        "sample_fraction_of_step": frac, "extrapolated": frac != 1.0,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": workload_config(wl, args, 1, wl.batch, args.ragged, args.replica_mode, L),
        "cpu_baseline": {"value": r["tok_s"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": sample},
        "e2e": {"value": r["tok_s"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": time.perf_counter() - t0,
    }
    print(json.dumps(line), flush=True)


def workload_config(wl, args, world, global_batch, ragged, replica_mode, L):
    return {
        "workload": (f"{wl.name}: {wl.model} decode, global batch {global_batch}, ctx {wl.ctx}, "
                     f"{'fp8-e4m3' if wl.fp8 else 'bf16'} paged KV (block {wl.block_size}), "
                     f"{wl.total_q_heads}q/{wl.total_kv_heads}kv heads d={wl.head_size}, L={L}, "
                     f"{'ragged' if ragged else 'uniform'} seq_lens, "
                     f"{'contiguous' if args.contiguous_blocks else 'random-permutation'} block table"),
        "global_batch": global_batch, "seq_len": wl.ctx, "parallelism": (f"tp{world} (KV-head sharding" + (
            "" if world <= wl.total_kv_heads else
            f"; {world // wl.total_kv_heads} ranks per KV head, " + (
                "batch split inside the replica group, q/out exchanged over peer memory"
                if replica_mode == "split" else "KV head replicated (reference rule)")) + ")"),
        "l2": "inputs larger than L2: every layer streams its own KV cache (no reuse between launches)",
    }


# --------------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------------
class Arm:
    """One measured configuration of the step on this rank: inputs, the step closure, its CUDA graph."""

    def __init__(self, args, wl, world, rank, groups, *, global_batch, ragged, replica_mode, allreduce, exchange, L):
        from vllm_omni_b200 import ops, workloads
        self.args, self.wl, self.world, self.rank, self.L = args, wl, world, rank, L
        self.global_batch, self.ragged, self.replica_mode = global_batch, ragged, replica_mode
        self.graphs = []
        wl_run = workloads.scaled(wl, batch=global_batch)
        self.wl_run = wl_run
        inp = workloads.make_decode_inputs(wl_run, "cuda", tp_size=world, rank=rank, ragged=ragged,
                                           num_layers=L, contiguous_blocks=args.contiguous_blocks)
        self.inp, sh = inp, inp.shard
        self.sh = sh
        B, hq, hkv, d = global_batch, sh.num_q_heads, sh.num_kv_heads, wl.head_size
        self.B, self.hq, self.hkv, self.d = B, hq, hkv, d
        self.caches = [ops.split_kv_cache(c) for c in inp.kv_caches]
        self.out = torch.empty(B, hq, d, dtype=torch.bfloat16, device="cuda")
        # KV-head replication (tp > KV heads): split the batch inside the replica group unless told otherwise
        self.rs, self.ex, self.ex_note = None, None, None
        self.dq, self.dout, self.dbt, self.dsl, self.dhq = inp.q, self.out, inp.block_table, inp.seq_lens, hq
        self.kv_bytes = inp.kv_bytes_per_layer
        self.q_grp = self.out_grp = None
        if sh.kv_replicas > 1 and replica_mode == "split":
            from vllm_omni_b200 import tp as tpmod
            my_group = groups[rank // sh.kv_replicas]
            self.ex_note = "nccl all_to_all"
            if exchange == "peer":
                from vllm_omni_b200 import peer
                try:
                    self.ex = peer.PeerExchange.create(my_group, B, hq, d, timeout_s=30.0)
                    self.ex.q.copy_(inp.q)                     # the qkv projection would write q here
                    self.ex_note = f"peer-memory gather/scatter kernels ({self.ex.mem.transport} mapping)"
                except Exception as exn:  # noqa: BLE001 - reported in the JSON line
                    self.ex = None
                    self.ex_note = f"nccl all_to_all (peer-memory setup failed: {type(exn).__name__}: {str(exn)[:200]})"
                    log(self.ex_note)
            self.rs = tpmod.ReplicaSplit(sh, my_group, exchange=self.ex)
            per, b0, b1 = self.rs.plan(B)
            self.dhq = self.rs.group_q_heads
            if self.ex is not None:
                self.q_grp, self.out_grp, self.out = self.ex.q_grp, self.ex.out_grp, self.ex.out
            else:
                self.q_grp = torch.zeros(per, self.dhq, d, dtype=torch.bfloat16, device="cuda")
                self.out_grp = torch.zeros(per, self.dhq, d, dtype=torch.bfloat16, device="cuda")
            self.dq, self.dout = self.q_grp[:b1 - b0], self.out_grp[:b1 - b0]
            self.dbt, self.dsl = inp.block_table[b0:b1], inp.seq_lens[b0:b1]
            self.kv_bytes = int(inp.seq_lens[b0:b1].sum().item()) * 2 * hkv * d * (1 if wl.fp8 else 2)
        self.dB = int(self.dq.shape[0])
        # schedule: stream-K (no split count) unless a split count or a fixed-split variant is asked for
        self.streamk = args.variant in (ops.VARIANT_STREAMK, ops.VARIANT_SK2, ops.VARIANT_SK3) or (
            args.variant == ops.VARIANT_AUTO and args.splits == 0)
        # default schedule = the library's own choice (B200ATTN_VARIANT_AUTO): the cooperative stream-K kernel, except for
        # fp8 caches (issue-bound: the leaner first-generation stream-K + its reduce launch, see api.cu)
        dflt = ops.VARIANT_STREAMK if (wl.fp8 and global_batch <= 2048) else ops.VARIANT_SK3
        self.variant = (args.variant or dflt) if self.streamk else (args.variant or ops.VARIANT_STREAM)
        self.planned = self.variant in (ops.VARIANT_SK2, ops.VARIANT_SK3)
        self.splits = 0 if self.streamk else (args.splits or ops.decode_num_splits(B, hq, hkv, inp.max_seq_len))
        self.ws = ops.DecodeWorkspace("cuda", ops.decode_workspace_bytes(max(self.dB, 1), self.dhq, d, max(self.splits, 1)))
        self.hidden = torch.zeros(B, wl.hidden_size, dtype=torch.bfloat16, device="cuda")  # o_proj output stand-in
        self.ar, self.ar_note, self.hidden_out = None, None, None
        if world > 1:
            self.ar_note = "nccl all_reduce"
            if allreduce == "peer":
                from vllm_omni_b200 import peer
                try:
                    self.ar = peer.PeerAllReduce.create(max_numel=self.hidden.numel(), timeout_s=30.0)
                    self.hidden = self.ar.input(self.hidden.shape)       # o_proj would write its partial sums here
                    self.hidden.zero_()
                    self.hidden_out = torch.empty(B, wl.hidden_size, dtype=torch.bfloat16, device="cuda")
                    algo = peer.ALGO_NAMES[peer.choose_algo(2 * self.hidden.numel(), world, self.ar.has_multicast,
                                                            self.ar.push_slot_bytes)]
                    self.ar_note = (f"peer-memory {algo} kernel ({self.ar.transport} mapping, "
                                    f"multicast={self.ar.has_multicast})")
                except Exception as exc:  # noqa: BLE001 - say so in the JSON line, never silently
                    self.ar = None
                    self.ar_note = f"nccl all_reduce (peer-memory setup failed: {type(exc).__name__}: {str(exc)[:200]})"
                    log(self.ar_note)
        # our kernels only (not NCCL / copies): plan + L x (cache write + decode [+ reduce] [+ all-reduce] [+ 2 exchanges]);
        # the cooperative stream-K kernel folds its partials itself (no reduce launch)
        reduce = (self.splits > 1) if not self.streamk else self.variant != ops.VARIANT_SK3
        per_layer = 2 + (1 if reduce else 0) + (1 if self.ar is not None else 0) + (2 if self.ex is not None else 0)
        self.launches_per_step = L * per_layer + (1 if self.planned and self.dB > 0 else 0)

    # ---- pieces of the step
    def o_proj_sum(self):
        """The only collective next to the path: o_proj's row-parallel sum of [B, hidden] bf16."""
        if self.ar is not None:
            self.ar.all_reduce(self.hidden, out=self.hidden_out)
        else:
            import torch.distributed as dist
            dist.all_reduce(self.hidden)

    def plan(self):
        from vllm_omni_b200 import ops
        if self.planned and self.dB > 0:
            ops.decode_plan(self.dsl, self.dbt, self.dB, self.dhq, self.hkv, self.wl.block_size, self.ws)

    def decode(self, kc, vc, variant=None, splits=None, out=None, planned=None):
        from vllm_omni_b200 import ops
        inp, wl = self.inp, self.wl
        if self.dB == 0:
            return
        v = self.variant if variant is None else variant
        pl = (self.planned if planned is None else planned) and v in (ops.VARIANT_SK2, ops.VARIANT_SK3)
        ops.paged_decode(self.dq, kc, vc, self.dbt, self.dsl, inp.max_seq_len, inp.softmax_scale,
                         out=self.dout if out is None else out, kv_cache_dtype=wl.kv_cache_dtype, k_scale=inp.k_scale,
                         v_scale=inp.v_scale, num_splits=self.splits if splits is None else splits, variant=v,
                         workspace=self.ws, planned=pl, early_kv_read=True, early_q_read=True)

    def layer(self, kc, vc, **kw):
        from vllm_omni_b200 import ops
        inp, wl = self.inp, self.wl
        ops.reshape_and_cache(inp.new_k, inp.new_v, kc, vc, inp.slot_mapping, wl.kv_cache_dtype, inp.k_scale, inp.v_scale)
        if self.rs is not None:
            self.rs.gather_q(self.ex.q if self.ex is not None else inp.q, out=self.q_grp)
        self.decode(kc, vc, **kw)
        if self.rs is not None:
            self.rs.scatter_out(self.out_grp, self.B, out=self.out)
        if self.world > 1:
            self.o_proj_sum()

    def step(self):
        self.plan()
        for kc, vc in self.caches:
            self.layer(kc, vc)

    def capture(self, fn):
        """CUDA-graph the launches of `fn` (how vLLM replays decode steps); None if capture fails."""
        if self.args.no_graph:
            return None
        try:
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                fn()
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                fn()
            self.graphs.append(g)
            return g
        except Exception as ex:  # report, fall back to eager launches
            print(f"[bench] graph capture failed ({type(ex).__name__}: {ex}); eager launches", file=sys.stderr)
            torch.cuda.synchronize()
            return None

    def time_step(self, steps: int, warmup: int, clk: ClockSampler | None = None):
        """(ms per step as max over ranks, launch mode). Warm-up >= 12 steps: the stream-K SM weights settle over
        the first 8 plan launches and are frozen afterwards."""
        world, rank = self.world, self.rank
        self.step()  # first call outside capture: lazy init (tensor maps, func attributes, NCCL)
        torch.cuda.synchronize()
        g_step = self.capture(self.step)
        run_step = g_step.replay if g_step is not None else self.step
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(max(warmup, 12)):
            run_step()
        if clk is not None:
            # keep the load on until rank 0's sampler has its first line (every rank runs the same number of extra
            # steps: the steps contain collectives), at most ~4 s
            t_lim = time.perf_counter() + 4.0
            while True:
                need = torch.tensor([1 if (rank == 0 and clk.proc is not None and clk.count() == 0
                                           and time.perf_counter() < t_lim) else 0], device="cuda")
                if world > 1:
                    import torch.distributed as dist
                    dist.broadcast(need, 0)
                if int(need.item()) == 0:
                    break
                for _ in range(8):
                    run_step()
        barrier(world)
        e0.record()
        for _ in range(steps):
            run_step()
        e1.record()
        torch.cuda.synchronize()
        ms_total = max_over_ranks(e0.elapsed_time(e1), world)
        barrier(world)
        return ms_total / steps, ("cuda-graph replay" if g_step is not None else "eager")

    def time_kernels(self, steps: int):
        """Average launch duration of (a) decode + reduce — what `roofline.achieved` is computed from — and (b) the
        decode kernel alone: CUDA events around a graph of ONLY those launches, one per layer cache back to back
        (plus the step's one plan launch), so the average excludes host launch gaps."""
        from vllm_omni_b200 import ops

        def full():
            self.plan()
            for kc, vc in self.caches:
                self.decode(kc, vc)

        def partial_only():
            self.plan()
            for kc, vc in self.caches:
                inp, wl = self.inp, self.wl
                ops.paged_decode_partial(self.dq, kc, vc, self.dbt, self.dsl, inp.max_seq_len, inp.softmax_scale,
                                         self.dout, max(self.splits, 1), self.ws, kv_cache_dtype=wl.kv_cache_dtype,
                                         k_scale=inp.k_scale, v_scale=inp.v_scale, variant=self.variant,
                                         planned=self.planned, early_kv_read=True, early_q_read=True)

        res = {}
        for name, fn in (("decode_plus_reduce", full), ("decode_only", partial_only)):
            if self.dB == 0:
                res[name] = (0.0, 0.0)
                continue
            g = self.capture(fn)
            run = g.replay if g is not None else fn
            for _ in range(3):
                run()
            ts = []
            for _ in range(steps):
                a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                run()
                b_.record()
                torch.cuda.synchronize()
                ts.append(a.elapsed_time(b_) / self.L)
            res[name] = (sum(ts) / len(ts), statistics.median(ts))
        return res

    def free(self):
        """Drop the graphs (they hold references to NCCL / peer kernels) and the big tensors."""
        self.graphs.clear()
        self.caches = None
        self.inp = None
        torch.cuda.synchronize()
        torch.cuda.empty_cache()


def verify_outputs(arm: Arm):
    """the concatenated per-rank outputs must equal the 1-GPU output bit for bit per head. Checked
    on layer 0 with the schedule-independent kernel (fixed split count 1: every (sequence, kv head) is computed by
    one CTA whatever the sharding), through the same cache write, replica-group exchange and head slicing as the
    timed step; the stream-K fast path (whose fp32 summation order follows the per-rank chunking) is then held to
    two bf16 roundings of that reference. Rank 0 recomputes the unsharded reference from the same seeds."""
    from vllm_omni_b200 import ops, workloads
    world, rank, wl = arm.world, arm.rank, arm.wl
    kc, vc = arm.caches[0]
    torch.cuda.synchronize()

    def run_layer(**kw):
        arm.layer(kc, vc, **kw) if world > 1 or arm.rs is not None else (arm.plan(), arm.layer(kc, vc, **kw))
        torch.cuda.synchronize()
        return arm.out.clone()

    arm.plan()
    det = run_layer(variant=ops.VARIANT_STREAM, splits=1)      # deterministic, schedule-independent
    fast = run_layer()                                          # the timed path
    if world > 1:
        import torch.distributed as dist
        parts_det = [torch.empty_like(det) for _ in range(world)] if rank == 0 else None
        parts_fast = [torch.empty_like(fast) for _ in range(world)] if rank == 0 else None
        dist.gather(det, parts_det, dst=0)
        dist.gather(fast, parts_fast, dst=0)
    else:
        parts_det, parts_fast = [det], [fast]
    res = None
    if rank == 0:
        full = workloads.make_decode_inputs(arm.wl_run, "cuda", tp_size=1, rank=0, ragged=arm.ragged, num_layers=1,
                                            contiguous_blocks=arm.args.contiguous_blocks)
        fk, fv = ops.split_kv_cache(full.kv_caches[0])
        ops.reshape_and_cache(full.new_k, full.new_v, fk, fv, full.slot_mapping, wl.kv_cache_dtype, full.k_scale,
                              full.v_scale)
        ref = ops.paged_decode(full.q, fk, fv, full.block_table, full.seq_lens, full.max_seq_len, full.softmax_scale,
                               kv_cache_dtype=wl.kv_cache_dtype, k_scale=full.k_scale, v_scale=full.v_scale,
                               num_splits=1, variant=ops.VARIANT_STREAM) if not wl.fp8 else None
        if ref is None:  # fp8 exists in the stream-K kernels only: the 1-rank stream-K run is the reference
            ref = ops.paged_decode(full.q, fk, fv, full.block_table, full.seq_lens, full.max_seq_len,
                                   full.softmax_scale, kv_cache_dtype=wl.kv_cache_dtype, k_scale=full.k_scale,
                                   v_scale=full.v_scale)
        torch.cuda.synchronize()
        got_det = torch.cat(parts_det, dim=1)
        got_fast = torch.cat(parts_fast, dim=1)
        bit_exact = bool(torch.equal(got_det, ref)) if not wl.fp8 else None
        tol = 2.0 ** -7 * float(ref.float().abs().max())
        diff_fast = float((got_fast.float() - ref.float()).abs().max())
        res = {"ranks": world, "rows": int(ref.shape[0]), "heads": int(ref.shape[1]), "layer": 0,
               "bit_exact": bit_exact, "bit_exact_kernel": "fixed-split (1 split) stream kernel, same write/exchange/slicing",
               "fast_path_max_abs_diff": diff_fast, "fast_path_tolerance": tol, "fast_path_ok": diff_fast <= tol}
        del full
    return res


def run_b200(args, wl):
    from vllm_omni_b200 import workloads
    rank, world, local = dist_setup(args.gpus)
    L = args.layers or wl.num_layers
    groups = None
    sh0 = workloads.shard_heads(wl.total_q_heads, wl.total_kv_heads, world, rank)
    if world > 1 and sh0.kv_replicas > 1:
        import torch.distributed as dist
        from vllm_omni_b200 import tp as tpmod
        groups = [dist.new_group(r) for r in tpmod.ReplicaSplit.groups(world, sh0.kv_replicas)]
    gb_primary = wl.batch * (world if args.scaling == "weak" else 1)

    def make(global_batch, ragged=args.ragged, replica_mode=args.replica_mode, allreduce=args.allreduce):
        return Arm(args, wl, world, rank, groups, global_batch=global_batch, ragged=ragged, replica_mode=replica_mode,
                   allreduce=allreduce, exchange=args.replica_exchange, L=L)

    arm = make(gb_primary)
    sh = arm.sh
    log(f"inputs ready: tp{world} rank{rank} Hq={arm.hq} Hkv={arm.hkv} replicas={sh.kv_replicas} L={L} B={arm.B}")
    with ClockSampler(local) as clk:
        ms_per_step, launch_mode = arm.time_step(args.steps, args.warmup, clk)
    log(f"timed region done: {ms_per_step:.3f} ms/step")
    value = arm.global_batch / (ms_per_step * 1e-3)

    verify = None
    if not args.no_verify:
        try:
            verify = verify_outputs(arm)
        except Exception as ex:  # report, never hide
            verify = {"error": f"{type(ex).__name__}: {str(ex)[:300]}"}
        log(f"verify: {verify}")

    # ---- roofline of the dominant kernels (decode + reduce), see Arm.time_kernels
    kt = arm.time_kernels(args.steps)
    full_avg_ms, full_med_ms = kt["decode_plus_reduce"]
    part_avg_ms, _ = kt["decode_only"]
    full_avg_ms = max_over_ranks(full_avg_ms, world)
    log(f"kernel-only timing done: decode+reduce {full_avg_ms * 1e3:.1f} us/launch, decode {part_avg_ms * 1e3:.1f} us")
    peak, peak_src = peaks()
    kv_bytes = arm.kv_bytes
    achieved = kv_bytes / (full_avg_ms * 1e-3) / 1e9 if full_avg_ms > 0 else 0.0
    traffic, traffic_note = None, "no ncu capture of this build committed"
    tj = ROOT / "profiles" / "traffic.json"
    if world == 1 and not args.ragged and tj.exists():
        tdoc = json.loads(tj.read_text())
        ent = tdoc.get(wl.name, {})
        if ent.get("kernel_sources_sha") == kernel_sources_sha() and ("sk3" in ent.get("kernel", "")) == (arm.variant == 6):
            traffic = ent.get("dram_bytes_per_launch")
            traffic_note = f"ncu --set full capture {ent.get('capture', '')} of these kernel sources"
        else:
            traffic_note = "committed ncu capture is of older kernel sources: not quoted"
    from vllm_omni_b200 import ops
    kname = {1: "paged_decode_tma_kernel + lse_reduce_kernel", 2: "paged_decode_cpasync_kernel + lse_reduce_kernel",
             3: "paged_decode_stream_kernel + lse_reduce_kernel", 4: "paged_decode_streamk_kernel + streamk_reduce_kernel",
             5: "paged_decode_sk2_kernel + sk2_reduce_kernel",
             6: "paged_decode_sk3_kernel (partials folded inside the kernel: no reduce launch)"}[arm.variant]
    roofline = {
        "bound": "hbm", "kernel": kname,
        "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
        "traffic_note": traffic_note, "peak_source": peak_src, "algorithmic_bytes_per_launch": kv_bytes,
        "kernel_avg_us": full_avg_ms * 1e3, "kernel_median_us": full_med_ms * 1e3,
        "kernel_avg_note": ("the complete decode of a layer (SURVEY 8d: t_kernel = partial + reduce; this kernel folds "
                            "its partials itself, there is no reduce launch), max over ranks" if arm.variant == 6 else
                            "decode + LSE-reduce launches (SURVEY 8d: t_kernel = partial + reduce), max over ranks"),
        "decode_only_us": part_avg_ms * 1e3,
        "achieved_decode_only": kv_bytes / (part_avg_ms * 1e-3) / 1e9 if part_avg_ms > 0 else 0.0,
        "kernel_share_of_step": full_avg_ms * L / ms_per_step, "frac_of_8TBps": achieved / 8000.0,
        "step_gbs_per_gpu": kv_bytes * L / (ms_per_step * 1e-3) / 1e9, "launch_mode": launch_mode,
        "num_splits": arm.splits, "schedule": "stream-K" if arm.streamk else "fixed-split",
    }

    # ---- end to end through the plugin with host buffers (H2D + D2H inside the timed region) ----
    e2e = None
    if not args.no_e2e:
        try:
            e2e = run_e2e(args, arm)
        except Exception as ex:  # report, never hide
            e2e = {"value": None, "unit": UNIT, "error": f"{type(ex).__name__}: {ex}"}
    log("e2e leg done")

    per_rank = {"q_heads": arm.hq, "kv_heads": arm.hkv, "kv_replicas": sh.kv_replicas,
                "replica_mode": (args.replica_mode if sh.kv_replicas > 1 else None),
                "decode_q_heads": arm.dhq, "decode_batch": arm.dB, "kv_bytes_per_layer": kv_bytes}
    ar_note, ex_note, launches = arm.ar_note, arm.ex_note, arm.launches_per_step
    primary_gbs = roofline["step_gbs_per_gpu"]
    arm.free()

    # ---- extra arms, each a complete measurement of the same step at another named configuration
    extras = {}
    if not args.no_extras:
        def measure(tag, **kw):
            a = None
            try:
                a = make(**kw)
                ms, _ = a.time_step(min(args.steps, 10), 3)
                kt2 = a.time_kernels(min(args.steps, 5))
                us = max_over_ranks(kt2["decode_plus_reduce"][0], world) * 1e3
                extras[tag] = {
                    "global_batch": a.global_batch, "ms_per_step": ms, "value": a.global_batch / (ms * 1e-3), "unit": UNIT,
                    "ragged": a.ragged, "replica_mode": a.replica_mode if a.sh.kv_replicas > 1 else None,
                    "collective": a.ar_note, "replica_exchange": a.ex_note, "kv_bytes_per_layer_per_gpu": a.kv_bytes,
                    "decode_plus_reduce_us": us, "achieved_gbs": a.kv_bytes / us / 1e3 if us > 0 else 0.0,
                    "step_gbs_per_gpu": a.kv_bytes * L / (ms * 1e-3) / 1e9}
            except Exception as ex:  # noqa: BLE001
                extras[tag] = {"error": f"{type(ex).__name__}: {str(ex)[:300]}"}
            finally:
                if a is not None:
                    a.free()
            log(f"extra arm {tag} done")

        if world == 1:
            measure("ragged" if not args.ragged else "uniform", global_batch=gb_primary, ragged=not args.ragged)
        else:
            other = wl.batch if args.scaling == "weak" else wl.batch * world
            measure("strong_scaling" if args.scaling == "weak" else "weak_scaling", global_batch=other)
            # the north_star's own arms: NCCL for the o_proj sum, KV heads replicated when tp > KV heads
            measure("north_star_nccl_allreduce", global_batch=gb_primary, allreduce="nccl",
                    replica_mode="replicate" if sh.kv_replicas > 1 else args.replica_mode)
            if sh.kv_replicas > 1:
                measure("replicate_kv_heads_peer_allreduce", global_batch=gb_primary, replica_mode="replicate")
        for e in extras.values():
            if "step_gbs_per_gpu" in e:
                e["per_gpu_gbs_vs_primary"] = e["step_gbs_per_gpu"] / primary_gbs if primary_gbs > 0 else None

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        r = time_cpu(wl, args.ragged, steps=1, warmup=0, budget_s=20.0, L=L)
        cpu = {"value": r["tok_s"], "unit": UNIT, "cores": r["cores"], "kind": "port",
               "sample": (f"1 step = all {wl.batch} sequences x {r['layers']} of {L} layers "
                          f"({r['t_step']:.2f} s measured; page gather + one torch fp32 SDPA with enable_gqa per layer "
                          f"+ slot write)" + ("" if r["layers"] == L else f", tokens/s scaled by {L}/{r['layers']} layers"))}
        log("cpu baseline done")

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 12), "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None,
            "dtype": "fp8_e4m3 KV / bf16 q, f32 accumulate" if wl.fp8 else "bf16",
            "data": "synthetic", "config": workload_config(wl, args, world, gb_primary, args.ragged, args.replica_mode, L),
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches * args.steps,
            "clocks": clk.result, "verify": verify, "per_rank": per_rank, "collective": ar_note,
            "replica_exchange": ex_note, "extras": extras,
        }
        print(json.dumps(line), flush=True)
    ok = verify is None or "error" in verify or (verify.get("bit_exact") in (True, None) and verify.get("fast_path_ok", True))
    if rank != 0:
        ok = True
    teardown(world)
    if not ok:
        raise SystemExit("bench.py: output verification against the 1-rank run FAILED (see `verify` in the JSON line)")


def teardown(world: int) -> None:
    """All CUDA graphs were dropped by Arm.free() (a graph keeps its NCCL / peer kernels alive, and destroying a
    communicator that live graphs still reference blocked forever in round 1). A watchdog still bounds the exit."""
    if world <= 1:
        return
    import torch.distributed as dist
    barrier(world)
    sys.stdout.flush()
    sys.stderr.flush()

    def bail():
        print("[bench] process-group teardown did not return in 20 s; exiting", file=sys.stderr, flush=True)
        os._exit(0)

    t = threading.Timer(20.0, bail)
    t.daemon = True
    t.start()
    dist.destroy_process_group()
    t.cancel()


def run_e2e(args, arm: Arm):
    from vllm_omni_b200 import backend as be
    from vllm_omni_b200.host_runner import HostStepRunner
    wl, inp, world = arm.wl, arm.inp, arm.world
    B, hq, hkv, d, L = arm.B, arm.hq, arm.hkv, arm.d, arm.L
    impl = be.B200AttentionImpl(hq, d, inp.softmax_scale, hkv, None, None, wl.kv_cache_dtype)
    if arm.rs is not None:
        impl.enable_replica_split(arm.rs, B)   # staging buffers sized once, outside the captured region
    qsl = torch.arange(B + 1, dtype=torch.int32, device="cuda")
    md = be.B200AttentionMetadata(
        num_actual_tokens=B, max_query_len=1, query_start_loc=qsl, max_seq_len=inp.max_seq_len,
        seq_lens=inp.seq_lens, block_table=inp.block_table, slot_mapping=inp.slot_mapping, num_decodes=B,
        num_decode_tokens=B, num_prefills=0, num_prefill_tokens=0, prefill_query_start_loc=None,
        prefill_max_query_len=0, decode_num_splits=arm.splits, workspace=arm.ws)
    runner = HostStepRunner(impl, md, inp.kv_caches, B, hq, hkv, d, inp.slot_mapping, inp.k_scale, inp.v_scale,
                            after_layer=arm.o_proj_sum if world > 1 else None,
                            plan_per_step=arm.planned and arm.rs is None)
    qkv_dev = torch.cat([inp.q.reshape(B, -1), inp.new_k.reshape(B, -1), inp.new_v.reshape(B, -1)], dim=1)
    host_qkv = [qkv_dev.cpu().pin_memory() for _ in range(L)]
    host_out = [torch.empty(B, hq * d, dtype=torch.bfloat16).pin_memory() for _ in range(L)]
    # one step = H2D of every layer's q/k/v + the layer kernels + D2H of every layer's output; replayed as a CUDA
    # graph with the copies as graph nodes (how vLLM replays decode), eager launches if capture fails / --no-graph
    run_step, mode = (lambda: runner.step(host_qkv, host_out)), "eager"
    if not args.no_graph:
        try:
            g = runner.capture(host_qkv, host_out)
            arm.graphs.append(g)
            run_step, mode = g.replay, "cuda-graph replay (H2D/D2H copy nodes inside)"
        except Exception as ex:  # noqa: BLE001
            print(f"[bench] e2e graph capture failed ({type(ex).__name__}: {ex}); eager", file=sys.stderr)
            torch.cuda.synchronize()
    for _ in range(3):
        run_step()
    barrier(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        run_step()
    e1.record()
    torch.cuda.synchronize()
    ms = max_over_ranks(e0.elapsed_time(e1), world) / args.steps
    return {"value": arm.global_batch / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms,
            "h2d_bytes_per_step": runner.h2d_bytes_per_layer * L, "d2h_bytes_per_step": runner.d2h_bytes_per_layer * L,
            "launch_mode": mode,
            "api": "B200AttentionImpl.do_kv_cache_update + forward via HostStepRunner (pinned host q/k/v in, output out; "
                   "one decode_plan per step as B200AttentionMetadataBuilder.build does)"}


def main():
    args = parse_args()
    from vllm_omni_b200 import workloads
    wl = workloads.WORKLOADS[args.workload]
    if args.total_kv_heads:
        wl = workloads.scaled(wl, total_kv_heads=args.total_kv_heads)
    if args.impl == "reference":
        run_reference(args, wl)
    else:
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a CUDA device (no CPU fallback for the product path)")
        run_b200(args, wl)


if __name__ == "__main__":
    main()
