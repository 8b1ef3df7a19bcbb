"""
Turns the raw captures that a gpurun call brings back in gpurun_out/ into the tracked
summaries in profiles/ (run in the build container, needs `ncu` for reading the report):

    python profiles/make_summaries.py

Inputs (names as written by the capture command documented in each summary header):
  gpurun_out/r01_launches.csv          ncu --metrics gpu__time_duration.sum launch list
  gpurun_out/r01_prof_kin_1M.ncu-rep   ncu --set full capture of kinetics_fused_kernel
  gpurun_out/r01_bench_default.json    plain `python bench.py` line
"""
from __future__ import annotations

import collections
import csv
import json
import shutil
import subprocess
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
OUT = ROOT / "profiles"
G = ROOT / "gpurun_out"
BENCH_CMD = "python bench.py --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 2"


def launch_list():
    rows = [r for r in csv.reader(open(G / "r01_launches.csv")) if len(r) > 10]
    hdr = rows[0]
    kn, mv, mu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.defaultdict(list)
    for r in rows[1:]:
        v = float(r[mv].replace(",", ""))
        u = r[mu]
        ms = {"us": v / 1e3, "ms": v, "ns": v / 1e6, "s": v * 1e3, "second": v * 1e3}.get(u, v)
        agg[r[kn].split("(")[0].strip()].append(ms)
    tot = sum(sum(v) for v in agg.values())
    lines = [
        f"# ncu launch list of `{BENCH_CMD}` (1,048,576 cells, p=64, s=32, one B200)",
        "# ncu --metrics gpu__time_duration.sum --clock-control none "
        "-k regex:^(kinetics_fused|stencil_|correct_|permeate_|scatter_kernel) -c 120",
        "# per-launch times are cold-cache and serialised: compare SHARES, not absolutes",
        "kernel,launches,total_ms,avg_ms,share",
    ]
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        lines.append(f"{k},{len(v)},{sum(v):.3f},{sum(v) / len(v):.4f},{sum(v) / tot:.4f}")
    (OUT / "r01_launch_list_summary.csv").write_text("\n".join(lines) + "\n")
    shutil.copy(G / "r01_launches.csv", OUT / "r01_launch_list_raw.csv")
    print("\n".join(lines))


def full_capture():
    raw = subprocess.run(["ncu", "-i", str(G / "r01_prof_kin_1M.ncu-rep"), "--page", "raw", "--csv"],
                         capture_output=True, text=True).stdout
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
    out = [f"# ncu --set full --clock-control none --import-source on -k regex:kinetics_fused -s 6 -c 2  ({BENCH_CMD})",
           "# 1,048,576 cells, p=64, s=32, one B200.  launch 1 = tier 1 (all cells), launch 2 = tier 2",
           "# (cells deferred because their tile overflows the small lists: none in this population)"]
    first = None
    for r in rows[2:]:
        d = {k: (r[hdr.index(k)], rows[1][hdr.index(k)]) for k in keys if k in hdr}
        first = first or d
        out.append("")
        out += [f"{k} = {v} {u}" for k, (v, u) in d.items()]
        items = [(hdr[i], r[i]) for i in range(len(hdr))
                 if "pcsamp_warps_issue_stalled" in hdr[i] and "not_issued" not in hdr[i]]
        tot = sum(float(v) for _, v in items) or 1.0
        out.append("warp stall sampling (all samples):")
        for k, v in sorted(items, key=lambda x: -float(x[1]))[:10]:
            out.append(f"  {100 * float(v) / tot:5.1f}%  " + k.replace("smsp__pcsamp_warps_issue_stalled_", ""))
    (OUT / "r01_ncu_kinetics_full_1M.txt").write_text("\n".join(out) + "\n")
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    num = lambda x: float(x.replace(",", ""))
    rd = num(first["dram__bytes_read.sum"][0]) * scale[first["dram__bytes_read.sum"][1]]
    wr = num(first["dram__bytes_write.sum"][0]) * scale[first["dram__bytes_write.sum"][1]]
    json.dump({"cells": 1048576, "proteins": 64, "signals": 32, "dram_bytes_per_launch": rd + wr,
               "dram_read_bytes": rd, "dram_write_bytes": wr,
               "source": "profiles/r01_ncu_kinetics_full_1M.txt (ncu --set full, one launch of kinetics_fused_kernel)"},
              open(OUT / "kinetics_traffic.json", "w"), indent=1)
    print("\n".join(out[:32]))
    print("DRAM bytes per cell:", (rd + wr) / 1048576)


if __name__ == "__main__":
    launch_list()
    full_capture()
    if (G / "r01_bench_default.json").exists():
        shutil.copy(G / "r01_bench_default.json", OUT / "r01_bench_default_1gpu.json")
