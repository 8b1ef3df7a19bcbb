"""
Turns the raw captures that a gpurun call brings back in gpurun_out/ into the tracked
summaries in profiles/ (run in the build container, needs `ncu` for reading the reports):

    python profiles/make_summaries.py [round tag, default r02]

Inputs (written by profiles/capture.sh on the GPU box):
  gpurun_out/<tag>_launch_list_raw.csv   ncu --metrics gpu__time_duration.sum launch list
  gpurun_out/<tag>_kin_cached.ncu-rep    ncu --set full capture of kinetics_lanes_kernel, cached lists
  gpurun_out/<tag>_kin_scan.ncu-rep      the same kernel with the list cache off (dense tiles scanned)
"""
from __future__ import annotations

import collections
import csv
import json
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
OUT = ROOT / "profiles"
G = ROOT / "gpurun_out"
TAG = sys.argv[1] if len(sys.argv) > 1 else "r02"
BENCH_CMD = "python bench.py --gpus 1 --steps 2 --warmup 1 --no-cpu-baseline --no-extra --e2e-steps 1 --scan-steps 2"
CELLS, P, S = 1 << 20, 64, 32


def launch_list():
    src = G / f"{TAG}_launch_list_raw.csv"
    rows = [r for r in csv.reader(open(src)) if len(r) > 10]
    hdr = rows[0]
    kn, mv, mu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.defaultdict(list)
    for r in rows[1:]:
        v = float(r[mv].replace(",", ""))
        ms = {"us": v / 1e3, "ms": v, "ns": v / 1e6, "s": v * 1e3, "second": v * 1e3}.get(r[mu], v)
        name = r[kn].split("(")[0].strip().replace("void ", "")
        agg[name].append(ms)
    tot = sum(sum(v) for v in agg.values())
    lines = [
        f"# ncu launch list of `{BENCH_CMD}` (1,048,576 cells, p=64, s=32, one B200)",
        "# ncu --metrics gpu__time_duration.sum --clock-control none "
        "-k regex:^(kinetics_|stencil_|correct_kernel|permeate_|scatter_|degrade_kernel|pixel_index|early_exit_verdict) -c 200",
        "# the command runs 5 steps with cached lists (1 builds them) and 4 steps of the scan path; per-launch",
        "# times are cold-cache and serialised: compare SHARES, not absolutes",
        "kernel,launches,total_ms,avg_ms,min_ms,max_ms,share",
    ]
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        lines.append(f"{k},{len(v)},{sum(v):.3f},{sum(v) / len(v):.4f},{min(v):.4f},{max(v):.4f},{sum(v) / tot:.4f}")
    (OUT / f"{TAG}_launch_list_summary.csv").write_text("\n".join(lines) + "\n")
    shutil.copy(src, OUT / f"{TAG}_launch_list_raw.csv")
    print("\n".join(lines))


def full_capture(name: str, what: str, cached: bool):
    rep = G / f"{TAG}_kin_{name}.ncu-rep"
    raw = subprocess.run(["ncu", "-i", str(rep), "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr = rows[0]
    keys = ["Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
            "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
            "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
            "smsp__issue_active.avg.pct_of_peak_sustained_active",
            "sm__warps_active.avg.pct_of_peak_sustained_active",
            "smsp__thread_inst_executed_per_inst_executed.ratio",
            "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_bytes.sum"]
    flags = "" if cached else " --no-list-cache"
    out = [f"# ncu --set full --clock-control none --import-source on -k regex:kinetics_lanes -s 2 -c 1  ({BENCH_CMD}{flags})",
           f"# 1,048,576 cells, p=64, s=32, one B200: {what}"]
    r = rows[2]
    d = {k: (r[hdr.index(k)], rows[1][hdr.index(k)]) for k in keys if k in hdr}
    out.append("")
    out += [f"{k} = {v} {u}" for k, (v, u) in d.items()]
    n_inst = float(d["smsp__inst_executed.sum"][0].replace(",", ""))
    out.append(f"warp instructions per cell = {n_inst / CELLS:.0f}")
    items = [(hdr[i], r[i]) for i in range(len(hdr))
             if "pcsamp_warps_issue_stalled" in hdr[i] and "not_issued" not in hdr[i]]
    tot = sum(float(v) for _, v in items) or 1.0
    out.append("warp stall sampling (all samples):")
    for k, v in sorted(items, key=lambda x: -float(x[1]))[:10]:
        out.append(f"  {100 * float(v) / tot:5.1f}%  " + k.replace("smsp__pcsamp_warps_issue_stalled_", ""))
    (OUT / f"{TAG}_ncu_kinetics_{name}_1M.txt").write_text("\n".join(out) + "\n")
    print("\n".join(out))
    split = subprocess.run([sys.executable, str(OUT / "phase_split.py"), str(rep), str(CELLS)],
                           capture_output=True, text=True).stdout
    (OUT / f"{TAG}_phase_split_{name}_1M.txt").write_text(
        f"# python profiles/phase_split.py gpurun_out/{rep.name} {CELLS}   ({what})\n" + split)

    def num(key):
        v, u = d[key]
        return float(v.replace(",", "")) * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}.get(u, 1.0)

    return {"cells": CELLS, "proteins": P, "signals": S, "cached": cached,
            "dram_bytes_per_launch": num("dram__bytes_read.sum") + num("dram__bytes_write.sum"),
            "source": f"profiles/{TAG}_ncu_kinetics_{name}_1M.txt (dram__bytes_read.sum + dram__bytes_write.sum, one launch)"}


def ptxas():
    sys.path.insert(0, str(ROOT))
    from magicsoup_b200 import build as B

    cmd = [B.nvcc(), *B.NVCC_FLAGS, "-Xptxas", "-v", "-o", "/tmp/ptxas_check.so", *[str(B.CSRC / s) for s in B.SOURCES]]
    err = subprocess.run(cmd, capture_output=True, text=True).stderr.splitlines()
    keep = []
    for i, ln in enumerate(err):
        if "Compiling entry function" in ln and ("kinetics_lanes_kernel" in ln or "kinetics_fused_kernel" in ln):
            keep += err[i:i + 4]
    (OUT / f"{TAG}_ptxas.txt").write_text("# nvcc " + " ".join(B.NVCC_FLAGS) + " -Xptxas -v\n" + "\n".join(keep) + "\n")


if __name__ == "__main__":
    launch_list()
    entries = [full_capture("cached", "steady-state step, lists from the cache", True),
               full_capture("scan", "list cache off: dense tiles scanned in the step", False)]
    (OUT / "kinetics_traffic.json").write_text(json.dumps({"entries": entries}, indent=1) + "\n")
    ptxas()
    for f in ("bench_default.json", "bench_2gpu.json"):
        if (G / f).exists():
            shutil.copy(G / f, OUT / f"{TAG}_{f}")
