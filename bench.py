"""
bench.py -- cell-steps/s of one World.enzymatic_activity() + one World.diffuse_molecules()
("step") on synthetic populations (BASELINE.json metric), one process per GPU.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  N > 1: python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload at N = 1: BASELINE.json configs[3], "synthetic 1M cells on 1024x1024 map, 64
proteins/cell, 32 signals" (42,248 algorithmic bytes per cell-step, 44 GB of parameters
resident in HBM: far larger than L2, so no flush is needed between iterations).
N > 1: every rank owns a row slab of the map with its own 1M cells (weak scaling).

Prints ONE JSON line (rank 0).  `value` = cells x K / device time with all state resident
in HBM; `e2e` = same through the public World API with the per-step state (cell
molecules, molecule map, cell positions) coming from pinned host memory and the result
read back every step; `roofline` = algorithmic bytes of the fused kinetics kernel / its
CUDA-event duration against the measured HBM peak; `cpu_baseline` = the CPU oracle
(oracle/ms_oracle.c, a port) with all host threads on a bounded sample.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

PARAMS = ("N", "Nf", "Nb", "A", "Kmr", "Ke", "Kmf", "Kmb", "Vmax")
METRIC = "cell-steps/sec (enzymatic_activity+diffuse_molecules)"
FALLBACK_HBM_GBS = 6650.0  # /opt/skills/guides/B200_PROFILING.md fallback


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cells", type=int, default=1 << 20)
    ap.add_argument("--proteins", type=int, default=64)
    ap.add_argument("--molecules", type=int, default=16)
    ap.add_argument("--map", type=int, default=1024)
    ap.add_argument("--cpu-sample-cells", type=int, default=65536)
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-list-cache", action="store_true",
                    help="scan the dense parameter tiles in every step (Kinetics.list_cache = False)")
    ap.add_argument("--energies", default="modest", choices=["modest", "survey"],
                    help="molecule energies of the synthetic chemistry: U(10,50) kJ (default) or SURVEY §8(d)'s U(10,500) kJ")
    return ap.parse_args()


def hbm_peak() -> tuple[float, str]:
    f = ROOT / "MEASURED_PEAKS.json"
    if f.exists():
        try:
            return float(json.loads(f.read_text())["hbm_gbs"]), "measured"
        except Exception:
            pass
    return FALLBACK_HBM_GBS, "fallback"


def kinetics_bytes_per_cell(p: int, s: int) -> int:
    """every parameter once, X in/out, positions"""
    return 20 * p * s + 16 * p + 8 * s + 8


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region"""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.lines: list[str] = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def energy_range(args):
    return (10e3, 500e3) if args.energies == "survey" else (10e3, 50e3)


def build_world(args, device, seed, world_size=1):
    import torch
    import magicsoup_b200 as ms
    from magicsoup_b200.synthetic import make_population

    pop = make_population(args.cells, args.proteins, args.molecules, args.map, seed=seed, device=device,
                          energy_range=energy_range(args))
    mols = [ms.Molecule(f"b{i}", 10e3, diffusivity=pop.diffusivities[i], permeability=pop.permeabilities[i])
            for i in range(pop.m)]
    chem = ms.Chemistry(molecules=mols, reactions=[])
    if world_size > 1:
        # every rank owns one (map x map) row slab of a (world_size*map x map) torus and the
        # cells on it: halo rows, plane sums and early-exit flags go over NCCL every step
        from magicsoup_b200.distributed import SlabWorld

        world = SlabWorld(chemistry=chem, global_map_size=(args.map * world_size, args.map), device=str(device))
    else:
        world = ms.World(chemistry=chem, map_size=args.map, device=str(device))
    # install the population without the (c,p,s)-sized copies the append path would make
    world.molecule_map = pop.molecule_map
    world.cell_positions = pop.cell_positions
    world.cell_molecules = pop.cell_molecules
    world.cell_map[pop.cell_positions[:, 0].long(), pop.cell_positions[:, 1].long()] = True
    world.cell_lifetimes = torch.zeros(pop.c, dtype=torch.int32, device=device)
    world.cell_divisions = torch.zeros(pop.c, dtype=torch.int32, device=device)
    world.n_cells = pop.c
    world.cell_genomes = [""] * pop.c
    world.cell_labels = [""] * pop.c
    for k in PARAMS:
        setattr(world.kinetics, k, getattr(pop, k))
    world.kinetics.list_cache = not args.no_list_cache
    return world, pop


def cpu_leg(args, pop_params_host, X0_host, map_host, cm_host, pos_host, diffs, perms, full_cells):
    """times the CPU oracle on a bounded sample; returns (cell-steps/s extrapolated, description, cores)"""
    import numpy as np
    from oracle import oracle as O

    cores = os.cpu_count() or 1
    P = O.Params(*pop_params_host)
    n = P.c
    t0 = time.perf_counter()
    O.integrate_signals(P, X0_host, n_threads=cores)
    t_kin = time.perf_counter() - t0
    t0 = time.perf_counter()
    O.diffuse_molecules(cm_host, map_host, pos_host, diffs, perms)
    t_diff = time.perf_counter() - t0
    # kinetics scales with cells; diffusion was timed on the full map, permeation on the sample
    t_step = t_kin * (full_cells / n) + t_diff
    desc = (f"kinetics on the first {n} of {full_cells} cells ({t_kin:.2f} s, {cores} OpenMP threads, "
            f"scaled linearly) + diffusion of the full {map_host.shape[0]}x{map_host.shape[1]}x{map_host.shape[2]} "
            f"map and permeation of the sample ({t_diff:.2f} s, 1 thread)")
    return full_cells / t_step, desc, cores


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path = the oracle port"""
    import numpy as np
    import torch
    from magicsoup_b200.synthetic import make_population

    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n = min(args.cells, args.cpu_sample_cells // 4)  # ~2 s of CPU work per step
    pop = make_population(n, args.proteins, args.molecules, args.map, seed=0, device="cpu",
                          energy_range=energy_range(args))
    par = [t.numpy() for t in pop.params()]
    X0 = pop.X0().numpy()
    vals = []
    desc = ""
    cores = os.cpu_count() or 1
    for i in range(args.warmup + args.steps):
        v, desc, cores = cpu_leg(args, par, X0, pop.molecule_map.numpy(), pop.cell_molecules.numpy(),
                                 pop.cell_positions.numpy(), pop.diffusivities, pop.permeabilities, args.cells)
        if i >= args.warmup:
            vals.append(v)
    value = sum(vals) / len(vals)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "cell-steps/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * args.cells / value,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, 1),
        "cpu_baseline": {"value": value, "unit": "cell-steps/s", "cores": cores, "kind": "port", "sample": desc},
        "e2e": {"value": value, "unit": "cell-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def workload_config(args, n_gpus):
    return {
        "workload": f"synthetic {args.cells} cells/GPU on {args.map}x{args.map} map slab/GPU, {args.proteins} "
                    f"proteins/cell, {2 * args.molecules} signals (BASELINE.json configs[3])",
        "cells_per_gpu": args.cells, "proteins": args.proteins, "signals": 2 * args.molecules,
        "map": [args.map, args.map], "parallelism": f"cells+map-slab x{n_gpus}",
        "l2": "inputs (44 GB of parameters at the default size) exceed L2; no flush needed",
    }


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
        return

    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    if world_size != args.gpus and world_size > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world_size}")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world_size > 1:
        dist.init_process_group("nccl", device_id=device)

    def barrier():
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize()

    world, pop = build_world(args, device, seed=rank, world_size=world_size)
    c, p, m = pop.c, pop.p, pop.m
    s = 2 * m
    torch.cuda.synchronize()

    def step():
        world.enzymatic_activity()
        world.diffuse_molecules()

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    world._kernel_events = []
    t_start, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_start.record()
    for _ in range(args.steps):
        step()
    t_end.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms_total = t_start.elapsed_time(t_end)
    kin_ms = [a.elapsed_time(b) for a, b in world._kernel_events]
    n_kin_launches = len(kin_ms)
    world._kernel_events = None
    t = torch.tensor([ms_total], dtype=torch.float64, device=device)
    if world_size > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    value = c * world_size * args.steps / (ms_total / 1e3)

    # ---- end to end through the public API with host-resident per-step state
    host_cm = world.cell_molecules.cpu().pin_memory()
    host_map = world.molecule_map.cpu().pin_memory()
    host_pos = world.cell_positions.cpu().pin_memory()
    out_cm = torch.empty_like(host_cm).pin_memory()
    out_map = torch.empty_like(host_map).pin_memory()
    h2d = host_cm.numel() * 4 + host_map.numel() * 4 + host_pos.numel() * 4
    d2h = out_cm.numel() * 4 + out_map.numel() * 4

    def e2e_step():
        world.cell_molecules = host_cm.to(device, non_blocking=True)
        world.molecule_map = host_map.to(device, non_blocking=True)
        world.cell_positions = host_pos.to(device, non_blocking=True)
        step()
        out_cm.copy_(world.cell_molecules, non_blocking=True)
        out_map.copy_(world.molecule_map, non_blocking=True)
        torch.cuda.synchronize()

    for _ in range(2):
        e2e_step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.e2e_steps):
        e2e_step()
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=device)
    if world_size > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = c * world_size * args.e2e_steps / (float(t.item()) / 1e3)

    if rank != 0:
        if world_size > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = hbm_peak()
    kin_avg_ms = sum(kin_ms) / max(len(kin_ms), 1)
    bytes_per_launch = kinetics_bytes_per_cell(p, s) * c
    achieved = bytes_per_launch / (kin_avg_ms / 1e3) / 1e9
    traffic = None
    tf = ROOT / "profiles" / "kinetics_traffic.json"
    if tf.exists():
        try:
            tj = json.loads(tf.read_text())
            if tj.get("cells") == c and tj.get("proteins") == p and tj.get("signals") == s:
                traffic = tj.get("dram_bytes_per_launch")
        except Exception:
            pass
    line = {
        "metric": METRIC, "value": value, "unit": "cell-steps/s", "n_gpus": world_size, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, world_size),
        "roofline": {"bound": "hbm", "kernel": "kinetics_fused_kernel", "achieved": achieved, "peak": peak,
                     "peak_source": peak_src, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                     "algorithmic_bytes_per_launch": bytes_per_launch, "kernel_ms": kin_avg_ms,
                     "kernel_share_of_step": kin_avg_ms * n_kin_launches / ms_total},
        "e2e": {"value": e2e_value, "unit": "cell-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": args.e2e_steps},
        "gpu_launches": 4 * args.steps + n_kin_launches,
        "clocks": clocks,
    }
    if not args.no_cpu_baseline:
        n = min(c, args.cpu_sample_cells)
        par = [getattr(pop, k)[:n].cpu().numpy() for k in PARAMS]
        cm = host_cm[:n].numpy()
        pos = host_pos[:n].numpy()
        mp = host_map.numpy()
        X0 = torch.cat([host_cm[:n], host_map[:, host_pos[:n, 0].long(), host_pos[:n, 1].long()].T], 1).numpy()
        v, desc, cores = cpu_leg(args, par, X0, mp, cm, pos, pop.diffusivities, pop.permeabilities, c)
        line["cpu_baseline"] = {"value": v, "unit": "cell-steps/s", "cores": cores, "kind": "port", "sample": desc}
    print(json.dumps(line))
    if world_size > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
