#!/usr/bin/env python
"""bench.py — decode throughput of the B200 paged-attention hot path (BASELINE.json metric).

A "step" is one decode step of the hot path over one synthetic batch: for each of the model's L
layers, the KV-cache write of the new token (reshape_and_cache) followed by split-K paged decode
attention (partial + LSE-reduce kernels) over that layer's own KV cache. Inputs are resident in HBM
when the timed region starts; the L distinct layer caches (>= 24 GB at cfg3) are far larger than L2.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg3] [--impl reference]
  torchrun --nproc-per-node N ... bench.py --gpus N ...        (one rank per GPU, KV-head TP)

Prints ONE JSON line on rank 0 (see README / DESIGN.md for the fields).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
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
                         "KV-head count); strong: global batch fixed")
    ap.add_argument("--variant", type=int, default=0, help="0 auto, 1 TMA, 2 cp.async, 3 stream, 4 stream-K")
    ap.add_argument("--splits", type=int, default=0, help="0 = heuristic")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
                         "kernel over NVLink (one-/two-shot, csrc/peer_allreduce.cu), 'nccl' = dist.all_reduce")
    ap.add_argument("--total-kv-heads", type=int, default=0, help="override the workload's KV-head count "
                    "(e.g. 1 with --gpus 2 exercises KV-head replication on two GPUs)")
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


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled every 100 ms while the step runs back to back: started before the
    warm-up (nvidia-smi needs up to a second before its first line, longer than a short timed region), read by a
    thread, stopped right after the timed region. `count()` = lines so far."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines, self.thread = index, None, [], None

    def __enter__(self):
        import threading
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
        self.result = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
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
# reference arm: the path's CPU implementation (oracle port of vLLM's CPU attention construction)
# --------------------------------------------------------------------------------------------
def cpu_sample_inputs(wl, n_seqs: int):
    """One layer of the workload restricted to `n_seqs` sequences, as CPU tensors (oracle-written cache)."""
    from oracle import oracle
    from vllm_omni_b200 import layout, workloads
    lens = workloads.make_seq_lens(wl.batch, wl.ctx)[:n_seqs]
    bt, nb = workloads.make_block_table(lens, wl.block_size)
    g = torch.Generator().manual_seed(workloads.SEED_KV)
    total = int(lens.sum())
    k = torch.randn(total, wl.total_kv_heads, wl.head_size, generator=g).to(torch.bfloat16)
    v = torch.randn(total, wl.total_kv_heads, wl.head_size, generator=g).to(torch.bfloat16)
    l64 = lens.long()
    seq_idx = torch.repeat_interleave(torch.arange(n_seqs), l64)
    pos = torch.arange(total) - torch.repeat_interleave(torch.cumsum(l64, 0) - l64, l64)
    slots = workloads.slots_of_positions(bt, seq_idx, pos, wl.block_size)
    ks = vs = None
    if wl.fp8:
        ks, vs = workloads.default_scales(wl.total_kv_heads)
    cache = layout.alloc_kv_cache(nb, wl.block_size, wl.total_kv_heads, wl.head_size, wl.kv_cache_dtype, "cpu")
    oracle.ref_reshape_and_cache(k, v, cache[0], cache[1], slots, wl.fp8, ks, vs)
    q = torch.randn(n_seqs, wl.total_q_heads, wl.head_size, generator=g).to(torch.bfloat16)
    new_k, new_v = k[:n_seqs].clone(), v[:n_seqs].clone()
    last = workloads.slots_of_positions(bt, torch.arange(n_seqs), l64 - 1, wl.block_size)
    return dict(q=q, cache=cache, bt=bt, lens=lens, ks=ks, vs=vs, new_k=new_k, new_v=new_v, last=last)


def cpu_step(wl, s) -> None:
    """One layer of the hot path on the CPU: cache write of the new tokens + paged decode."""
    from oracle import oracle
    oracle.ref_reshape_and_cache(s["new_k"], s["new_v"], s["cache"][0], s["cache"][1], s["last"], wl.fp8,
                                 s["ks"], s["vs"])
    oracle.ref_paged_decode(s["q"], s["cache"][0], s["cache"][1], s["bt"], s["lens"], wl.head_size ** -0.5,
                            wl.fp8, s["ks"], s["vs"])


def time_cpu(wl, n_seqs: int, reps: int, warm: int = 1):
    cores = len(os.sched_getaffinity(0))
    torch.set_num_threads(cores)
    s = cpu_sample_inputs(wl, n_seqs)
    for _ in range(warm):
        cpu_step(wl, s)
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        cpu_step(wl, s)
        ts.append(time.perf_counter() - t0)
    t_layer = statistics.median(ts)              # seconds for n_seqs sequences of one layer
    tok_s = n_seqs / (t_layer * wl.num_layers)   # a token needs all L layers
    return tok_s, cores, t_layer


def run_reference(args, wl):
    """`--impl reference`: the path's CPU implementation on the host cores. The reference's genuine
    CPU kernels (vLLM cpu_attn, K12 in SURVEY.md) are absent from the installed CUDA wheel and the
    reference tree has no native sources, so this is the oracle port (kind = "port")."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n_seqs = min(wl.batch, 8)
    t0 = time.perf_counter()
    tok_s, cores, t_layer = time_cpu(wl, n_seqs, reps=max(1, args.steps), warm=max(1, min(args.warmup, 2)))
    sample = (f"{n_seqs} of {wl.batch} sequences x 1 of {wl.num_layers} layers per step "
              f"(torch fp32 SDPA over gathered pages, all host threads), scaled by L")
    line = {
        "impl": "reference", "metric": METRIC, "value": tok_s, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": t_layer * wl.num_layers * 1e3 * wl.batch / n_seqs,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": workload_config(wl, args, 1, wl.batch),
        "cpu_baseline": {"value": tok_s, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": tok_s, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": time.perf_counter() - t0,
    }
    print(json.dumps(line), flush=True)


def workload_config(wl, args, world, global_batch):
    return {
        "workload": (f"{wl.name}: {wl.model} decode, global batch {global_batch}, ctx {wl.ctx}, "
                     f"{'fp8-e4m3' if wl.fp8 else 'bf16'} paged KV (block {wl.block_size}), "
                     f"{wl.total_q_heads}q/{wl.total_kv_heads}kv heads d={wl.head_size}, L={args.layers or wl.num_layers}, "
                     f"{'ragged' if args.ragged else 'uniform'} seq_lens, "
                     f"{'contiguous' if args.contiguous_blocks else 'random-permutation'} block table"),
        "global_batch": global_batch, "seq_len": wl.ctx, "parallelism": (f"tp{world} (KV-head sharding" + (
            "" if world <= wl.total_kv_heads else
            f"; {world // wl.total_kv_heads} ranks per KV head, " + (
                "batch split inside the replica group, q/out exchanged by all-to-all"
                if args.replica_mode == "split" else "KV head replicated (reference rule)")) + ")"),
        "l2": "inputs larger than L2: every layer streams its own KV cache (no reuse between launches)",
    }


# --------------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------------
def run_b200(args, wl):
    from vllm_omni_b200 import ops, workloads
    rank, world, local = dist_setup(args.gpus)
    L = args.layers or wl.num_layers
    global_batch = wl.batch * (world if args.scaling == "weak" else 1)
    wl_run = workloads.scaled(wl, batch=global_batch)
    inp = workloads.make_decode_inputs(wl_run, "cuda", tp_size=world, rank=rank, ragged=args.ragged,
                                       num_layers=L, contiguous_blocks=args.contiguous_blocks)
    sh = inp.shard
    B, hq, hkv, d = global_batch, sh.num_q_heads, sh.num_kv_heads, wl.head_size
    caches = [ops.split_kv_cache(c) for c in inp.kv_caches]
    out = torch.empty(B, hq, d, dtype=torch.bfloat16, device="cuda")
    # KV-head replication (tp > KV heads): split the batch inside the replica group unless told otherwise
    rs, ex, ex_note = None, None, None
    dq, dout, dbt, dsl, dhq = inp.q, out, inp.block_table, inp.seq_lens, hq   # what the decode kernel sees
    kv_bytes = inp.kv_bytes_per_layer
    if sh.kv_replicas > 1 and args.replica_mode == "split":
        import torch.distributed as dist
        from vllm_omni_b200 import tp as tpmod
        groups = [dist.new_group(r) for r in tpmod.ReplicaSplit.groups(world, sh.kv_replicas)]
        my_group = groups[rank // sh.kv_replicas]
        ex, ex_note = None, "nccl all_to_all"
        if args.replica_exchange == "peer":
            from vllm_omni_b200 import peer
            try:
                ex = peer.PeerExchange.create(my_group, B, hq, d, timeout_s=30.0)
                ex.q.copy_(inp.q)                     # the qkv projection would write q here
                ex_note = f"peer-memory gather/scatter kernels ({ex.mem.transport} mapping)"
            except Exception as exn:  # noqa: BLE001 - reported in the JSON line
                ex, ex_note = None, f"nccl all_to_all (peer-memory setup failed: {type(exn).__name__}: {str(exn)[:200]})"
                log(ex_note)
        rs = tpmod.ReplicaSplit(sh, my_group, exchange=ex)
        per, b0, b1 = rs.plan(B)
        dhq = rs.group_q_heads
        if ex is not None:
            q_grp, out_grp, out = ex.q_grp, ex.out_grp, ex.out
        else:
            q_grp = torch.zeros(per, dhq, d, dtype=torch.bfloat16, device="cuda")
            out_grp = torch.zeros(per, dhq, d, dtype=torch.bfloat16, device="cuda")
        dq, dout, dbt, dsl = q_grp[:b1 - b0], out_grp[:b1 - b0], inp.block_table[b0:b1], inp.seq_lens[b0:b1]
        kv_bytes = int(inp.seq_lens[b0:b1].sum().item()) * 2 * hkv * d * (1 if wl.fp8 else 2)
    # schedule: stream-K (no split count) unless a split count or a fixed-split variant is asked for
    streamk = args.variant == ops.VARIANT_STREAMK or (args.variant == ops.VARIANT_AUTO and args.splits == 0)
    variant = ops.VARIANT_STREAMK if streamk else (args.variant or ops.VARIANT_STREAM)
    splits = 0 if streamk else (args.splits or ops.decode_num_splits(B, hq, hkv, inp.max_seq_len))
    ws = ops.DecodeWorkspace("cuda", ops.decode_workspace_bytes(B, dhq, d, max(splits, 1)))
    hidden = torch.zeros(B, wl.hidden_size, dtype=torch.bfloat16, device="cuda")  # o_proj output stand-in
    ar, ar_note, hidden_out = None, None, None
    if world > 1:
        import torch.distributed as dist
        ar_note = "nccl all_reduce"
        if args.allreduce == "peer":
            from vllm_omni_b200 import peer
            try:
                ar = peer.PeerAllReduce.create(max_numel=hidden.numel(), timeout_s=30.0)
                hidden = ar.input(hidden.shape)           # o_proj would write its partial sums here
                hidden.zero_()
                hidden_out = torch.empty(B, wl.hidden_size, dtype=torch.bfloat16, device="cuda")
                algo = peer.ALGO_NAMES[peer.choose_algo(2 * hidden.numel(), world, ar.has_multicast, ar.push_slot_bytes)]
                ar_note = f"peer-memory {algo} kernel ({ar.transport} mapping, multicast={ar.has_multicast})"
            except Exception as exc:  # noqa: BLE001 - say so in the JSON line, never silently
                ar = None
                ar_note = f"nccl all_reduce (peer-memory setup failed: {type(exc).__name__}: {str(exc)[:200]})"
                log(ar_note)

    def o_proj_sum():
        """The only collective next to the path: o_proj's row-parallel sum of [B, hidden] bf16."""
        if ar is not None:
            ar.all_reduce(hidden, out=hidden_out)
        else:
            dist.all_reduce(hidden)

    def step():
        for kc, vc in caches:
            ops.reshape_and_cache(inp.new_k, inp.new_v, kc, vc, inp.slot_mapping, wl.kv_cache_dtype,
                                  inp.k_scale, inp.v_scale)
            if rs is not None:
                rs.gather_q(ex.q if ex is not None else inp.q, out=q_grp)
            ops.paged_decode(dq, kc, vc, dbt, dsl, inp.max_seq_len,
                             inp.softmax_scale, out=dout, kv_cache_dtype=wl.kv_cache_dtype,
                             k_scale=inp.k_scale, v_scale=inp.v_scale, num_splits=splits,
                             variant=variant, workspace=ws)
            if rs is not None:
                rs.scatter_out(out_grp, B, out=out)
            if world > 1:
                o_proj_sum()

    # our kernels only (not NCCL / copies): cache write + decode (+ reduce) (+ peer all-reduce)
    launches_per_step = L * (2 + (1 if (splits > 1 or streamk) else 0) + (1 if ar is not None else 0)
                             + (2 if ex is not None else 0))

    def capture(fn):
        """CUDA-graph the launches of `fn` (how vLLM replays decode steps); None if capture fails."""
        if args.no_graph:
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
            return g
        except Exception as ex:  # report, fall back to eager launches
            print(f"[bench] graph capture failed ({type(ex).__name__}: {ex}); eager launches", file=sys.stderr)
            torch.cuda.synchronize()
            return None

    log(f"inputs ready: tp{world} rank{rank} Hq={hq} Hkv={hkv} replicas={sh.kv_replicas} L={L} B={B}")
    step()  # first call outside capture: lazy init (tensor maps, func attributes, NCCL)
    torch.cuda.synchronize()
    log("first step done")
    g_step = capture(step)
    log(f"step {'captured' if g_step is not None else 'NOT captured (eager)'}")
    run_step = g_step.replay if g_step is not None else step
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clk:
        for _ in range(max(args.warmup, 3)):
            run_step()
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
        for _ in range(args.steps):
            run_step()
        e1.record()
        torch.cuda.synchronize()
    ms_total = max_over_ranks(e0.elapsed_time(e1), world)
    barrier(world)
    log(f"timed region done: {ms_total / args.steps:.3f} ms/step")
    ms_per_step = ms_total / args.steps
    value = global_batch / (ms_per_step * 1e-3)

    # ---- roofline of the dominant kernel: CUDA events around a graph of ONLY its L launches
    # (one per layer cache, back to back), so the average excludes host launch gaps; the eager
    # fallback brackets every launch with its own event pair.
    def partial_only():
        for kc, vc in caches:
            ops.paged_decode_partial(dq, kc, vc, dbt, dsl, inp.max_seq_len,
                                     inp.softmax_scale, dout, max(splits, 1), ws,
                                     kv_cache_dtype=wl.kv_cache_dtype, k_scale=inp.k_scale,
                                     v_scale=inp.v_scale, variant=variant)

    g_part = capture(partial_only)
    kern_ms = []
    if g_part is not None:
        for _ in range(3):
            g_part.replay()
        for _ in range(args.steps):
            a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            g_part.replay()
            b_.record()
            torch.cuda.synchronize()
            kern_ms.append(a.elapsed_time(b_) / L)
    else:
        for _ in range(args.steps):
            for kc, vc in caches:
                a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                ops.paged_decode_partial(dq, kc, vc, dbt, dsl, inp.max_seq_len,
                                         inp.softmax_scale, dout, max(splits, 1), ws,
                                         kv_cache_dtype=wl.kv_cache_dtype, k_scale=inp.k_scale,
                                         v_scale=inp.v_scale, variant=variant)
                b_.record()
                kern_ms.append((a, b_))
        torch.cuda.synchronize()
        kern_ms = [a.elapsed_time(b_) for a, b_ in kern_ms]
    kern_avg_ms = sum(kern_ms) / len(kern_ms)
    log(f"kernel-only timing done: {kern_avg_ms * 1e3:.1f} us/launch")
    peak, peak_src = peaks()
    achieved = kv_bytes / (kern_avg_ms * 1e-3) / 1e9
    traffic = None  # DRAM read+write bytes per launch from the committed ncu --set full capture (N=1 shapes)
    tj = ROOT / "profiles" / "traffic.json"
    if world == 1 and not args.ragged and tj.exists():
        traffic = json.loads(tj.read_text()).get(wl.name, {}).get("dram_bytes_per_launch")
    roofline = {
        "bound": "hbm", "kernel": {1: "paged_decode_tma_kernel", 2: "paged_decode_cpasync_kernel",
                                   3: "paged_decode_stream_kernel", 4: "paged_decode_streamk_kernel"}[variant],
        "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
        "peak_source": peak_src, "algorithmic_bytes_per_launch": kv_bytes,
        "kernel_avg_us": kern_avg_ms * 1e3, "kernel_median_us": statistics.median(kern_ms) * 1e3,
        "kernel_share_of_step": kern_avg_ms * L / ms_per_step, "frac_of_8TBps": achieved / 8000.0,
        "step_gbs_per_gpu": kv_bytes * L / (ms_per_step * 1e-3) / 1e9, "launch_mode": "cuda-graph replay" if g_step is not None else "eager", "num_splits": splits, "schedule": "stream-K" if streamk else "fixed-split",
    }

    # ---- end to end through the plugin with host buffers (H2D + D2H inside the timed region) ----
    e2e = None
    if not args.no_e2e:
        try:
            e2e = run_e2e(args, wl, inp, B, hq, hkv, d, L, splits, ws, world, global_batch, rs,
                          o_proj_sum if world > 1 else None)
        except Exception as ex:  # report, never hide
            e2e = {"value": None, "unit": UNIT, "error": f"{type(ex).__name__}: {ex}"}

    log("e2e leg done")
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        n_seqs = min(wl.batch, 8)
        tok_s, cores, t_layer = time_cpu(wl, n_seqs, reps=3)
        cpu = {"value": tok_s, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"{n_seqs} of {wl.batch} sequences x 1 of {wl.num_layers} layers, 3 reps "
                         f"(median {t_layer * 1e3:.1f} ms), torch fp32 SDPA over gathered pages, scaled by L"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "fp8_e4m3 KV / bf16 q, f32 accumulate" if wl.fp8 else "bf16",
            "data": "synthetic", "config": workload_config(wl, args, world, global_batch),
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches_per_step * args.steps,
            "clocks": clk.result, "per_rank": {"q_heads": hq, "kv_heads": hkv, "kv_replicas": sh.kv_replicas,
                                               "replica_mode": (args.replica_mode if sh.kv_replicas > 1 else None),
                                               "decode_q_heads": dhq, "decode_batch": int(dq.shape[0]),
                                               "kv_bytes_per_layer": kv_bytes},
            "collective": ar_note, "replica_exchange": ex_note,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        # Leave without tearing NCCL down: destroying a communicator whose kernels are still referenced
        # by live CUDA graphs can block forever (seen on the 2-GPU box: the line above was printed and
        # the ranks never exited). Everything is synchronised and flushed, so exit hard on every rank.
        barrier(world)
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


def run_e2e(args, wl, inp, B, hq, hkv, d, L, splits, ws, world, global_batch, rs=None, after_layer=None):
    from vllm_omni_b200 import backend as be
    from vllm_omni_b200.host_runner import HostStepRunner
    impl = be.B200AttentionImpl(hq, d, inp.softmax_scale, hkv, None, None, wl.kv_cache_dtype)
    if rs is not None:
        impl.enable_replica_split(rs)
    qsl = torch.arange(B + 1, dtype=torch.int32, device="cuda")
    md = be.B200AttentionMetadata(
        num_actual_tokens=B, max_query_len=1, query_start_loc=qsl, max_seq_len=inp.max_seq_len,
        seq_lens=inp.seq_lens, block_table=inp.block_table, slot_mapping=inp.slot_mapping, num_decodes=B,
        num_decode_tokens=B, num_prefills=0, num_prefill_tokens=0, prefill_query_start_loc=None,
        prefill_max_query_len=0, decode_num_splits=splits, workspace=ws)
    runner = HostStepRunner(impl, md, inp.kv_caches, B, hq, hkv, d, inp.slot_mapping, inp.k_scale, inp.v_scale,
                            after_layer=after_layer)
    qkv_dev = torch.cat([inp.q.reshape(B, -1), inp.new_k.reshape(B, -1), inp.new_v.reshape(B, -1)], dim=1)
    host_qkv = [qkv_dev.cpu().pin_memory() for _ in range(L)]
    host_out = [torch.empty(B, hq * d, dtype=torch.bfloat16).pin_memory() for _ in range(L)]
    # one step = H2D of every layer's q/k/v + the layer kernels + D2H of every layer's output; replayed as a CUDA
    # graph with the copies as graph nodes (how vLLM replays decode), eager launches if capture fails / --no-graph
    run_step, mode = (lambda: runner.step(host_qkv, host_out)), "eager"
    if not args.no_graph:
        try:
            g = runner.capture(host_qkv, host_out)
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
    return {"value": global_batch / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms,
            "h2d_bytes_per_step": runner.h2d_bytes_per_layer * L, "d2h_bytes_per_step": runner.d2h_bytes_per_layer * L,
            "launch_mode": mode,
            "api": "B200AttentionImpl.do_kv_cache_update + forward via HostStepRunner (pinned host q/k/v in, output out)"}


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
