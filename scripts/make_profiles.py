#!/usr/bin/env python
"""Turn the scratch artefacts in gpurun_out/ into the tracked summaries under profiles/.

  python scripts/make_profiles.py r1          # prefix of the profiles/ files

Inputs (whatever exists): gpurun_out/r1_launches_*.csv (ncu --metrics gpu__time_duration.sum launch
lists), gpurun_out/r1_prof_*.ncu-rep (ncu --set full captures, read with `ncu -i ... --page raw --csv`),
gpurun_out/sk_trace_*.log, gpurun_out/bench_*.json, and the built library for the SASS census.
"""
import collections
import csv
import glob
import io
import json
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
OUT = ROOT / "profiles"
SRC = ROOT / "gpurun_out"
tag = sys.argv[1] if len(sys.argv) > 1 else "r1"
OUT.mkdir(exist_ok=True)

METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__bytes_read.sum.per_second", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__cycles_active.min", "sm__cycles_active.avg",
    "sm__cycles_active.max", "sm__cycles_elapsed.max", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__grid_size", "launch__block_size",
    "l1tex__data_bank_conflicts_pipe_lsu.sum", "smsp__inst_executed.sum",
]


def short(name: str) -> str:
    n = name.split("(")[0]
    return n.replace("void ", "").replace("b200::", "")


def launches():
    for f in sorted(glob.glob(str(SRC / f"{tag}_launches_*.csv"))):
        rows = list(csv.DictReader(l for l in open(f) if l.startswith('"')))
        agg = collections.OrderedDict()
        for r in rows:
            k = short(r["Kernel Name"])
            if "b200" not in r["Kernel Name"]:
                k = "(torch setup) " + k[-50:]
            agg.setdefault(k, []).append(float(r["Metric Value"]) / 1e3)
        ours = {k: v for k, v in agg.items() if not k.startswith("(torch")}
        # setup launches of our own kernels (history fill) are much longer than the step's: split them off
        lines = [f"# ncu launch list: {Path(f).name}", "",
                 "`ncu --metrics gpu__time_duration.sum --clock-control none` over a short bench run "
                 "(per-launch times are cold-cache and serialised; the SHARE of the step is what is compared "
                 "with bench.py's live `kernel_share_of_step`).", "",
                 "| kernel | launches | avg us | min us | max us | sum us |", "|---|---|---|---|---|---|"]
        for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
            lines.append(f"| `{k}` | {len(v)} | {sum(v)/len(v):.2f} | {min(v):.2f} | {max(v):.2f} | {sum(v):.1f} |")
        step = {}
        for k, v in ours.items():
            med = sorted(v)[len(v) // 2]
            step[k] = [x for x in v if x < 8 * med] if "cache_write" in k else v
        per_step = {k: v for k, v in step.items() if "_plan_kernel" in k}
        step = {k: v for k, v in step.items() if "_plan_kernel" not in k}
        tot = sum(sum(v) / len(v) for v in step.values())
        lines += ["", "Once per step: " + ", ".join(f"`{k}` {sum(v)/len(v):.2f} us" for k, v in per_step.items()), "",
                  "Per-layer step (one launch of each of our kernels, setup-sized cache writes excluded):", ""]
        for k, v in step.items():
            a = sum(v) / len(v)
            lines.append(f"- `{k}`: {a:.2f} us = {100*a/tot:.1f} % of the layer")
        (OUT / (Path(f).stem + ".md")).write_text("\n".join(lines) + "\n")
        print("wrote", OUT / (Path(f).stem + ".md"))


UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def kernel_sources_sha() -> str:
    """Same digest as bench.py: a capture is only quoted as `roofline.traffic` while the kernel sources match."""
    import hashlib
    h = hashlib.sha256()
    for name in ("paged_decode_sk3.cu", "sk_math.cuh", "decode_common.cuh", "ptx.cuh"):
        h.update((ROOT / "vllm_omni_b200" / "csrc" / name).read_bytes())
    return h.hexdigest()[:16]


def record_traffic(workload, kernel, hdr, units, r):
    """profiles/traffic.json: measured DRAM bytes per launch of the dominant kernel (bench.py's roofline.traffic)."""
    tot = 0.0
    for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        i = hdr.index(m)
        tot += float(r[i]) * UNIT[units[i]]
    path = OUT / "traffic.json"
    d = json.loads(path.read_text()) if path.exists() else {}
    d[workload] = {"kernel": kernel, "dram_bytes_per_launch": tot, "source": f"{tag}_prof_{workload}.ncu-rep",
                   "capture": f"{tag}_prof_{workload}.ncu-rep", "kernel_sources_sha": kernel_sources_sha()}
    path.write_text(json.dumps(d, indent=1, sort_keys=True) + "\n")


def ncu_reports():
    for f in sorted(glob.glob(str(SRC / f"{tag}_prof_*.ncu-rep"))):
        try:
            raw = subprocess.run(["ncu", "-i", f, "--page", "raw", "--csv"], capture_output=True, text=True,
                                 check=True).stdout
        except (OSError, subprocess.CalledProcessError) as e:
            print("ncu failed on", f, e)
            continue
        rows = list(csv.reader(io.StringIO(raw)))
        hdr, units = rows[0], rows[1]
        lines = [f"# ncu --set full: {Path(f).name}", "",
                 "`ncu --set full --clock-control none --import-source on`, one launch per kernel, "
                 "read with `ncu -i ... --page raw --csv`.", ""]
        for r in rows[2:]:
            name = short(r[hdr.index("Kernel Name")])
            if "paged_decode_sk3_kernel" in name or "paged_decode_streamk_kernel" in name:
                record_traffic(Path(f).stem.split("_prof_")[-1], name, hdr, units, r)
            lines += [f"## `{name}`", "", "| metric | value | unit |", "|---|---|---|"]
            for m in METRICS:
                if m in hdr:
                    i = hdr.index(m)
                    lines.append(f"| {m} | {r[i]} | {units[i]} |")
            lines.append("")
        (OUT / (Path(f).stem + ".md")).write_text("\n".join(lines))
        print("wrote", OUT / (Path(f).stem + ".md"))


def sass():
    lib = ROOT / "vllm_omni_b200" / "libb200attn.so"
    if not lib.exists():
        return
    out = subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True).stdout
    keys = ("UTMALDG", "UTMAPF", "HMMA", "LDSM", "MOVM", "SYNCS", "F2FP", "MUFU.EX2", "LDGSTS", "ATOMG", "UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTCATOMSWS", "MULTIMEM", "REDG",
            "ACQBULK", "LDG.E.128", "STG.E.128", "STG.E.64")
    cnt, name = collections.OrderedDict(), None
    for line in out.splitlines():
        if "Function :" in line:
            name = line.split("Function :")[1].strip()
            cnt[name] = collections.Counter()
            continue
        if name is None:
            continue
        for tok in line.split():
            if tok.startswith(keys):
                cnt[name][tok.rstrip(";")] += 1
    lines = ["SASS mnemonic census of libb200attn.so (cuobjdump -sass, sm_100a). UTMALDG = TMA tensor load,",
             "SYNCS = mbarrier ops, HMMA.16816 = mma.sync tensor-core MMA, LDSM = ldmatrix, MOVM = movmatrix,",
             "F2FP...E4M3.UNPACK = e4m3 -> f16 conversion, UTCHMMA = tcgen05.mma, UTCBAR = tcgen05.commit, LDTM / STTM =",
             "tcgen05.ld / st (TMEM), UTCATOMSWS = TMEM allocation.", ""]
    for fn, c in cnt.items():
        dem = subprocess.run(["c++filt", fn], capture_output=True, text=True).stdout.strip() or fn
        lines.append(short(dem))
        for k, v in sorted(c.items()):
            lines.append(f"    {v:5d}  {k}")
    (OUT / f"{tag}_sass_census.txt").write_text("\n".join(lines) + "\n")
    print("wrote", OUT / f"{tag}_sass_census.txt")


def copies():
    for pat in ("sk_trace_*.log", "bench_*.json", "prefill_*.json", "compare_*.json", "scale_*.json", "sweep_*.json"):
        for f in sorted(glob.glob(str(SRC / pat))):
            txt = open(f).read().strip()
            if txt:
                (OUT / f"{tag}_{Path(f).name}").write_text(txt + "\n")


if __name__ == "__main__":
    launches()
    ncu_reports()
    sass()
    copies()
