"""
bench.py -- cell-steps/s of one World.enzymatic_activity() + one World.diffuse_molecules()
("step") on synthetic populations (BASELINE.json metric), one process per GPU.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                  [--config c4|c3|c5] [--scaling weak|strong]
  N > 1: python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workloads (BASELINE.json configs)
  c4 (default, configs[3])  1,048,576 cells on a 1024 x 1024 map, 64 proteins, 32 signals
       weak   (default)     every rank owns such a population on its own 1024 x 1024 row slab
       strong               ONE such population and map split over the N ranks
  c3 (configs[2])           16,384 cells saturating a 128 x 128 map, 35 proteins, 40 signals, 1 GPU
  c5 (configs[4])           16,777,216 cells on a 4096 x 4096 map over 8 GPUs (2 M cells + 512 rows each)
The 42 KB of parameters per cell are far larger than L2: no flush between iterations.

Prints ONE JSON line (rank 0).
  value     cells x K / device time, all state resident in HBM, the package's defaults: the kernel's
            sparse lists are cached per parameter row, so a steady-state step does not re-read the
            dense tiles (`roofline.scan_path` is the same kernel with the cache off)
  e2e       the same through the public World API with the per-step state (cell molecules, molecule
            map, cell positions) coming from pinned host memory and the result read back, every step
  roofline  kinetics kernel: SURVEY 8(d)'s algorithmic bytes per cell-step x cells / CUDA-event time,
            against the measured HBM peak; `moved_*` = the bytes a launch actually has to move
  cpu_baseline / --impl reference: the CPU oracle (oracle/ms_oracle.c, a port), all host threads,
            full-size steps
  extra     N = 1: the c3 workload, the cost of kill + divide of 1 % of the cells per step at c4, and the
            reference's own benchmark scripts (performance/check.py, run_simulation.py; bench_workloads.py);
            N > 1: strong scaling of c4 and a bit-exact parity check of the multi-GPU path against the
            single-domain oracle (the run fails if it mismatches)
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
CONFIGS = {
    #        cells    proteins molecules map   BASELINE.json
    "c4": (1 << 20, 64, 16, 1024, "configs[3]"),
    "c3": (1 << 14, 35, 20, 128, "configs[2]"),
    "c5": (1 << 24, 64, 16, 4096, "configs[4]"),
}
CPU_CHUNK = 65536  # cells per oracle pass of the CPU legs


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="c4", choices=sorted(CONFIGS))
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--cells", type=int, default=None, help="override the config's cell count (per GPU when weak)")
    ap.add_argument("--proteins", type=int, default=None)
    ap.add_argument("--molecules", type=int, default=None)
    ap.add_argument("--map", type=int, default=None)
    ap.add_argument("--e2e-steps", type=int, default=10)
    ap.add_argument("--scan-steps", type=int, default=5, help="steps timed with the list cache off (0 = skip)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true")
    ap.add_argument("--skew", type=float, default=0.0,
                    help="N > 1: also time a run where rank 0 holds SKEW x the cells of every other rank (extra.density_skew)")
    ap.add_argument("--no-list-cache", action="store_true",
                    help="scan the dense parameter tiles in every step (Kinetics.list_cache = False)")
    ap.add_argument("--energies", default="modest", choices=["modest", "survey"],
                    help="molecule energies of the synthetic chemistry: U(10,50) kJ (default) or SURVEY 8(d)'s U(10,500) kJ")
    args = ap.parse_args()
    cells, p, m, size, _ = CONFIGS[args.config]
    args.cells = args.cells or cells
    args.proteins = args.proteins or p
    args.molecules = args.molecules or m
    args.map = args.map or size
    return args


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


def energy_range(args):
    return (10e3, 500e3) if args.energies == "survey" else (10e3, 50e3)


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


# ---------------------------------------------------------------------------------- worlds
def build_world(cells, proteins, molecules, map_rows, map_cols, device, seed, world_size, erange, list_cache=True):
    """a World (one GPU) or the rank's SlabWorld of a (world_size * map_rows) x map_cols torus"""
    import torch
    import magicsoup_b200 as ms
    from magicsoup_b200.synthetic import make_population

    pop = make_population(cells, proteins, molecules, (map_rows, map_cols), seed=seed, device=device,
                          energy_range=erange)
    mols = [ms.Molecule(f"b{i}", 10e3, diffusivity=pop.diffusivities[i], permeability=pop.permeabilities[i])
            for i in range(pop.m)]
    chem = ms.Chemistry(molecules=mols, reactions=[])
    if world_size > 1:
        # every rank owns one row slab and the cells on it: halo rows, plane sums and early-exit
        # flags go over NCCL every step
        from magicsoup_b200.distributed import SlabWorld

        world = SlabWorld(chemistry=chem, global_map_size=(map_rows * world_size, map_cols), device=str(device))
        world.deferred_settle = True  # every rank runs the same loop: the early-exit check may trail the step
    else:
        world = ms.World(chemistry=chem, map_size=map_rows if map_rows == map_cols else (map_rows, map_cols),
                         device=str(device))
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
    world.kinetics.list_cache = list_cache
    return world, pop


def timed_steps(world, n_steps, barrier, device, world_size):
    """(total ms over n_steps as max over ranks, kinetics launch durations in ms)"""
    import torch
    import torch.distributed as dist

    world._kernel_events = []
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t0.record()
    for _ in range(n_steps):
        world.enzymatic_activity()
        world.diffuse_molecules()
    t1.record()
    barrier()
    kin_ms = [a.elapsed_time(b) for a, b in world._kernel_events]
    world._kernel_events = None
    t = torch.tensor([t0.elapsed_time(t1)], dtype=torch.float64, device=device)
    if world_size > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    return {"ms_total": ms_total, "ms_per_step": ms_total / n_steps, "kernel_ms": sum(kin_ms) / max(len(kin_ms), 1),
            "kernel_share_of_step": sum(kin_ms) / ms_total}


def measure(world, steps, warmup, barrier, device, world_size):
    for _ in range(warmup):
        world.enzymatic_activity()
        world.diffuse_molecules()
    return timed_steps(world, steps, barrier, device, world_size)


def lifecycle_rounds(world, device, step_ms, rounds=4, frac=0.01):
    """
    The reference's per-step bookkeeping (performance/run_simulation.py:56-100) at the bench size: kill
    `frac` of the cells, let as many divide onto the freed pixels, step once.  Host clock around
    each call with a device synchronize on both sides (these calls are host work + small launches).
    """
    import torch

    g = torch.Generator()
    g.manual_seed(5)
    out = []
    for _ in range(rounds):
        n = world.n_cells
        k = max(1, int(n * frac))
        kill = torch.randperm(n, generator=g)[:k].tolist()
        freed = world.cell_positions[kill].clone()
        parents = torch.randperm(n - k, generator=g)[:k].tolist()
        torch.cuda.synchronize(device)
        t0 = time.perf_counter()
        world.kill_cells(cell_idxs=kill)
        torch.cuda.synchronize(device)
        t1 = time.perf_counter()
        world.divide_cells_at(parents, freed)
        torch.cuda.synchronize(device)
        t2 = time.perf_counter()
        world.enzymatic_activity()
        world.diffuse_molecules()
        torch.cuda.synchronize(device)
        t3 = time.perf_counter()
        # genomes of 1 % of the cells changed (mutations): their lists are rebuilt from the dense tiles in the next step
        kin = world.kinetics
        stale = torch.randperm(world.n_cells, generator=g)[:k].to(device)
        kin._stale(kin._phys(stale))
        torch.cuda.synchronize(device)
        t4 = time.perf_counter()
        world.enzymatic_activity()
        world.diffuse_molecules()
        torch.cuda.synchronize(device)
        t5 = time.perf_counter()
        out.append({"kill_ms": (t1 - t0) * 1e3, "divide_ms": (t2 - t1) * 1e3, "step_after_ms": (t3 - t2) * 1e3,
                    "step_with_1pct_stale_lists_ms": (t5 - t4) * 1e3})
    k = max(1, int(world.n_cells * frac))
    return {"what": f"kill {k} random cells, divide {k} random cells onto the freed pixels, one step, then the lists of {k} cells "
                    f"marked stale (as after mutations) and one more step; {rounds} rounds "
                    "(the first one includes the one-off growth of the parameter buffers by half)",
            "rounds": out, "steady_step_ms": step_ms,
            "parameter_bytes_per_cell": kinetics_bytes_per_cell(int(world.kinetics._buf["N"].shape[1]),
                                                                int(world.kinetics._buf["N"].shape[2]))}


def genome_update_times(device, n_cells=10_000, genome_size=1000):
    """
    World.update_cells' work for `n_cells` random genomes (translate_genomes + set_cell_params,
    world.py:880-908; the reference quotes 5.95 s per 10,000 cells on a T4, performance/check.py:15):
    host translation (C++ / OpenMP), flattening, one kernel launch.  14 molecules / 28 signals like
    the reference's Wood-Ljungdahl chemistry.
    """
    import random

    import numpy as np
    import torch

    import magicsoup_b200 as ms

    random.seed(11)
    mols = [ms.Molecule(f"bench-g{i}", 10e3 + 15e3 * i, diffusivity=1.0, permeability=0.1) for i in range(14)]
    reactions = [([mols[i]], [mols[i + 1]]) for i in range(0, 12, 2)] + \
                [([mols[i], mols[i + 2]], [mols[i + 1], mols[i + 3]]) for i in range(0, 10, 2)]
    kin = ms.Kinetics(chemistry=ms.Chemistry(molecules=mols, reactions=reactions), device=str(device))
    gen = ms.Genetics()
    genomes = [ms.random_genome(genome_size) for _ in range(n_cells)]
    out = {"cells": n_cells, "genome_size": genome_size, "host_threads": len(os.sched_getaffinity(0))}
    for rep in range(2):  # the second round is the warm one
        t0 = time.perf_counter()
        prot_off, dom_off, doms = gen.translate_genomes_flat(genomes)
        t1 = time.perf_counter()
        counts = prot_off[1:] - prot_off[:-1]
        p_max = int(counts.max())
        if rep == 0:
            kin.increase_max_cells(by_n=n_cells)
        kin.increase_max_proteins(max_n=p_max)
        has = counts > 0
        keep = [i for i, h in enumerate(has.tolist()) if h]
        offsets = np.concatenate([prot_off[:1], prot_off[1:][has]])
        torch.cuda.synchronize(device)
        t2 = time.perf_counter()
        kin.set_cell_params_flat(keep, offsets, dom_off, doms)
        torch.cuda.synchronize(device)
        t3 = time.perf_counter()
        t4 = time.perf_counter()
        nested = gen.translate_genomes(genomes)  # the reference's interface: nested lists of tuples
        t5 = time.perf_counter()
        del nested
    out.update(translate_genomes_flat_ms=(t1 - t0) * 1e3, set_cell_params_ms=(t3 - t2) * 1e3, proteins_max=p_max,
               translate_genomes_nested_ms=(t5 - t4) * 1e3,
               reference_quote="5.95 s per 10,000 cells for update_cells on a T4 (performance/check.py:15)")
    return out


def roofline_block(c, p, s, res, scan, rec_words, traffic):
    """res / scan: timed_steps() of the default run (cached lists) and of the scan path"""
    peak, peak_src = hbm_peak()
    alg = kinetics_bytes_per_cell(p, s) * c
    moved = (4 * rec_words + 8 * s + 8) * c if rec_words else alg
    out = {
        "bound": "hbm", "kernel": "kinetics_lanes_kernel" if p <= 64 and s <= 64 else "kinetics_fused_kernel",
        "achieved": alg / (res["kernel_ms"] / 1e3) / 1e9, "peak": peak, "peak_source": peak_src, "unit": "GB/s",
        "frac": alg / (res["kernel_ms"] / 1e3) / 1e9 / peak, "traffic": traffic,
        "algorithmic_bytes_per_launch": alg, "kernel_ms": res["kernel_ms"],
        "kernel_share_of_step": res["kernel_share_of_step"],
        "definition": "achieved = SURVEY 8(d)'s algorithmic bytes (dense dtypes: 20ps+16p+8s+8 per cell-step) / kernel "
                      "time.  With the list cache on (default) a launch only has to move moved_bytes_per_launch "
                      "(one cached record + X in/out per cell); scan_path is the same kernel reading the dense "
                      "tiles in every step (list cache off)",
        "moved_bytes_per_launch": moved, "moved_achieved": moved / (res["kernel_ms"] / 1e3) / 1e9,
        "moved_frac": moved / (res["kernel_ms"] / 1e3) / 1e9 / peak,
    }
    if scan is not None:
        a = alg / (scan["kernel_ms"] / 1e3) / 1e9
        out["scan_path"] = {"kernel_ms": scan["kernel_ms"], "achieved": a, "frac": a / peak,
                            "ms_per_step": scan["ms_per_step"], "cell_steps_per_s": c / (scan["ms_per_step"] / 1e3)}
    return out


def traffic_for(c, p, s, cached: bool):
    try:
        tj = json.loads((ROOT / "profiles" / "kinetics_traffic.json").read_text())
        for e in tj.get("entries", []):
            if (e.get("cells"), e.get("proteins"), e.get("signals"), e.get("cached")) == (c, p, s, cached):
                return e.get("dram_bytes_per_launch")
    except Exception:
        pass
    return None


# ---------------------------------------------------------------------------------- CPU legs
def cpu_population(args):
    from magicsoup_b200.synthetic import make_population

    n = min(args.cells, CPU_CHUNK)
    n_passes = max(1, args.cells // n)
    pop = make_population(n, args.proteins, args.molecules, args.map, seed=0, device="cpu", energy_range=energy_range(args))
    return pop, n, n_passes


def cpu_step_time(pop, n_passes):
    """one full-size step of the CPU oracle: n_passes passes over a resident population + the map once"""
    from oracle import oracle as O

    cores = os.cpu_count() or 1
    P = O.Params(*[t.numpy() for t in pop.params()])
    X0 = pop.X0().numpy()
    cm, mp, pos = pop.cell_molecules.numpy(), pop.molecule_map.numpy(), pop.cell_positions.numpy()
    t0 = time.perf_counter()
    for _ in range(n_passes):
        O.integrate_signals(P, X0, n_threads=cores)
    t_kin = time.perf_counter() - t0
    t0 = time.perf_counter()
    mm = O.diffuse(mp, pop.diffusivities)
    for _ in range(n_passes):
        O.permeate(cm, mm, pos, pop.permeabilities)
    t_diff = time.perf_counter() - t0
    return t_kin, t_diff, cores


def cpu_sample_text(args, n, n_passes, t_kin, t_diff, cores):
    gb = kinetics_bytes_per_cell(args.proteins, 2 * args.molecules) * args.cells / 1e9
    return (f"one full-size step = {n_passes} oracle passes over a resident population of {n} cells "
            f"({n * n_passes} cell-steps of kinetics: {t_kin:.2f} s on {cores} OpenMP threads) + one diffusion of the "
            f"{args.molecules}x{args.map}x{args.map} map and the permeation of every pass ({t_diff:.2f} s, 1 thread); "
            f"nothing is extrapolated (distinct parameters for all {args.cells} cells would be {gb:.0f} GB of host memory)")


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path = the oracle port, full-size steps"""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    pop, n, n_passes = cpu_population(args)
    times, desc, cores = [], "", 1
    for i in range(args.warmup + args.steps):
        t_kin, t_diff, cores = cpu_step_time(pop, n_passes)
        if i >= args.warmup:
            times.append(t_kin + t_diff)
            desc = cpu_sample_text(args, n, n_passes, t_kin, t_diff, cores)
    step_s = sum(times) / len(times)
    value = n * n_passes / step_s
    cfg = workload_config(args, args.gpus)  # the same config as our arm at this N
    if args.gpus > 1 and not split_over_ranks(args, args.gpus):
        # weak scaling: the N-GPU workload is N such populations; a timed step is a bounded sample of it
        desc = f"bounded sample: 1/{args.gpus} of the {args.gpus}-GPU workload's cells per step (cell-steps/s does not depend on it); " + desc
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "cell-steps/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * step_s,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": cfg,
        "cpu_baseline": {"value": value, "unit": "cell-steps/s", "cores": cores, "kind": "port", "sample": desc},
        "e2e": {"value": value, "unit": "cell-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def split_over_ranks(args, n_gpus) -> bool:
    return n_gpus > 1 and (args.scaling == "strong" or args.config == "c5")


def workload_config(args, n_gpus):
    _, _, _, _, where = CONFIGS[args.config]
    s = 2 * args.molecules
    if split_over_ranks(args, n_gpus):
        what = (f"synthetic {args.cells} cells on a {args.map}x{args.map} map split over {n_gpus} GPUs "
                f"({args.cells // n_gpus} cells and {args.map // n_gpus} map rows each)")
    else:
        what = f"synthetic {args.cells} cells/GPU on {args.map}x{args.map} map slab/GPU"
    return {
        "workload": f"{what}, {args.proteins} proteins/cell, {s} signals (BASELINE.json {where})",
        "cells_per_gpu": args.cells // n_gpus if split_over_ranks(args, n_gpus) else args.cells,
        "proteins": args.proteins, "signals": s, "map": [args.map, args.map],
        "parallelism": f"cells+map-slab x{n_gpus}", "energies": args.energies, "list_cache": not args.no_list_cache,
        "deferred_settle": True,
        "l2": "parameters (42 KB per cell) exceed L2 by orders of magnitude; no flush needed",
    }


def per_rank_shape(args, world_size):
    """(cells, map rows, map cols) of one rank"""
    if split_over_ranks(args, world_size):
        if args.map % world_size or args.cells % world_size:
            raise SystemExit(f"{args.map} rows / {args.cells} cells do not split over {world_size} ranks")
        return args.cells // world_size, args.map // world_size, args.map
    if args.config == "c5":
        raise SystemExit("c5 (16M cells, 700 GB of parameters) needs --gpus 8")
    return args.cells, args.map, args.map


def pin_to_own_cpus(local_rank, local_world):
    """spread the ranks over the host cores (the e2e leg is a host-memory copy loop per rank)"""
    try:
        cpus = sorted(os.sched_getaffinity(0))
        per = max(1, len(cpus) // max(local_world, 1))
        mine = cpus[local_rank * per:(local_rank + 1) * per] or cpus
        os.sched_setaffinity(0, mine)
        return len(mine)
    except Exception:
        return None


def slab_parity_check(world_size):
    """bit-exact comparison of the N-rank SlabWorld with the single-domain CPU oracle (small problem)"""
    from tests import dist_worker

    t0 = time.perf_counter()
    try:
        dist_worker.run("nccl", steps=3, cells=20000, rows_per_rank=128, cols=128, seed=5, quiet=True)
        ok, err = True, ""
    except AssertionError as e:
        ok, err = False, str(e)[:300]
    return {"ranks": world_size, "ok": ok,
            "problem": f"20000 cells x 12 proteins x 8 signals on a {128 * world_size}x128 map: 3 steps "
                       "(activity, degradation, diffusion), kill + cells migrating across slab boundaries + 1 step, "
                       "then most of rank 0's cells killed + rebalance() of the slab boundaries + 1 step",
            "result": "bit-identical to the single-domain oracle" if ok else "MISMATCH: " + err,
            "seconds": round(time.perf_counter() - t0, 2)}


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
    n_cpus = None
    if world_size > 1:
        n_cpus = pin_to_own_cpus(local_rank, int(os.environ.get("LOCAL_WORLD_SIZE", world_size)))
        dist.init_process_group("nccl", device_id=device)

    def barrier():
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize()

    cells, rows, cols = per_rank_shape(args, world_size)
    erange = energy_range(args)
    world, pop = build_world(cells, args.proteins, args.molecules, rows, cols, device, seed=rank,
                             world_size=world_size, erange=erange, list_cache=not args.no_list_cache)
    c, p, m = pop.c, pop.p, pop.m
    s = 2 * m
    torch.cuda.synchronize()

    for _ in range(args.warmup):
        world.enzymatic_activity()
        world.diffuse_molecules()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    res = timed_steps(world, args.steps, barrier, device, world_size)
    clocks = sampler.stop() if rank == 0 else None
    peer_active = world.__dict__.get("_peer_ctx") is not None  # flags, plane sums and halos over peer memory (csrc/peer.cu)
    value = c * world_size * args.steps / (res["ms_total"] / 1e3)
    rec_words = int(world.kinetics._cache.shape[1]) if world.kinetics._cache is not None else 0

    # ---- the same kernel reading the dense tiles in every step (list cache off)
    scan = None
    if args.scan_steps > 0 and not args.no_list_cache:
        world.kinetics.list_cache = False
        scan = measure(world, args.scan_steps, 2, barrier, device, world_size)
        world.kinetics.list_cache = True
        world.enzymatic_activity()  # rebuilds the records
    roofline = roofline_block(c, p, s, res, scan, rec_words, traffic_for(c, p, s, rec_words > 0))

    # ---- end to end through the public API with host-resident per-step state
    host_cm = world.cell_molecules.cpu().pin_memory()
    host_map = world.molecule_map.cpu().pin_memory()
    host_pos = world.cell_positions.cpu().pin_memory()
    out_cm = torch.empty_like(host_cm).pin_memory()
    out_map = torch.empty_like(host_map).pin_memory()
    h2d = host_cm.numel() * 4 + host_map.numel() * 4 + host_pos.numel() * 4
    d2h = out_cm.numel() * 4 + out_map.numel() * 4

    # Every step copies its inputs from pinned host memory and reads its results back.  The copies run
    # on their own streams, two input slots deep: the upload of step t+1 and the read-back of step
    # t-1 overlap the kernels of step t (the host blocks once per step, on the early-exit flags).
    compute = torch.cuda.current_stream(device)
    s_in, s_out = torch.cuda.Stream(device), torch.cuda.Stream(device)
    slots = [dict(cm=torch.empty_like(host_cm, device=device), map=torch.empty_like(host_map, device=device),
                  pos=torch.empty_like(host_pos, device=device), ev_in=torch.cuda.Event(), ev_free=torch.cuda.Event())
             for _ in range(2)]
    for sl in slots:
        sl["ev_free"].record(compute)

    def upload(t):
        sl = slots[t % 2]
        with torch.cuda.stream(s_in):
            s_in.wait_event(sl["ev_free"])  # the slot's previous step has been computed and read back
            sl["cm"].copy_(host_cm, non_blocking=True)
            sl["map"].copy_(host_map, non_blocking=True)
            sl["pos"].copy_(host_pos, non_blocking=True)
            sl["ev_in"].record(s_in)

    keep = []

    def run(n):
        upload(0)
        for t in range(n):
            sl = slots[t % 2]
            if t + 1 < n:
                upload(t + 1)
            compute.wait_event(sl["ev_in"])
            world.cell_molecules, world.molecule_map, world.cell_positions = sl["cm"], sl["map"], sl["pos"]
            world.enzymatic_activity()
            world.diffuse_molecules()
            done = torch.cuda.Event()
            done.record(compute)
            res_cm = world.cell_molecules  # a new tensor of this step (world.py:625, 664-665)
            with torch.cuda.stream(s_out):
                s_out.wait_event(done)
                out_cm.copy_(res_cm, non_blocking=True)
                out_map.copy_(sl["map"], non_blocking=True)
                sl["ev_free"].record(s_out)
            keep.append(res_cm)
            del keep[:-3]
        torch.cuda.synchronize()

    run(3)
    barrier()
    t_wall = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    run(args.e2e_steps)
    e1.record()
    torch.cuda.synchronize()
    t_wall = (time.perf_counter() - t_wall) * 1e3
    barrier()
    t = torch.tensor([max(e0.elapsed_time(e1), t_wall)], dtype=torch.float64, device=device)
    if world_size > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = c * world_size * args.e2e_steps / (float(t.item()) / 1e3)
    del slots, keep
    del host_cm, host_map, host_pos, out_cm, out_map

    # ---- extras (outside the timed regions above)
    extra = {}
    if not args.no_extra and world_size > 1:
        extra["parity_check"] = slab_parity_check(world_size)
        if args.config == "c4" and args.scaling == "weak" and args.map % world_size == 0:
            del world, pop  # strong scaling: ONE c4 population and map split over the ranks
            torch.cuda.empty_cache()
            w2, _ = build_world(args.cells // world_size, p, m, args.map // world_size, args.map, device,
                                seed=100 + rank, world_size=world_size, erange=erange)
            r2 = measure(w2, max(args.steps, 10), max(args.warmup, 3), barrier, device, world_size)
            extra["c4_strong"] = {
                "workload": f"{args.cells} cells on a {args.map}x{args.map} map split over {world_size} GPUs "
                            f"({args.cells // world_size} cells and {args.map // world_size} rows each)",
                "scaling": "strong", "ms_per_step": r2["ms_per_step"], "kernel_ms": r2["kernel_ms"],
                "value": args.cells / (r2["ms_per_step"] / 1e3), "unit": "cell-steps/s"}
            del w2
    if args.skew > 1.0 and world_size > 1:
        # cells live with their pixels, so a crowded slab is a slow rank: rank 0 gets SKEW x the cells
        torch.cuda.empty_cache()
        base = int(args.cells / args.skew)
        n_mine = int(base * args.skew) if rank == 0 else base
        w4, _ = build_world(n_mine, p, m, args.map, args.map, device, seed=200 + rank, world_size=world_size, erange=erange)
        r4 = measure(w4, max(args.steps, 10), max(args.warmup, 3), barrier, device, world_size)
        total = int(base * args.skew) + base * (world_size - 1)
        w5, _ = build_world(total // world_size, p, m, args.map, args.map, device, seed=300 + rank, world_size=world_size,
                            erange=erange)
        r5 = measure(w5, max(args.steps, 10), max(args.warmup, 3), barrier, device, world_size)
        del w5
        torch.cuda.empty_cache()
        barrier()
        t0 = time.perf_counter()
        heights = w4.rebalance()  # slab boundaries follow the cells; map rows and cells move to their new owners
        barrier()
        t_rebalance = time.perf_counter() - t0
        r6 = measure(w4, max(args.steps, 10), max(args.warmup, 3), barrier, device, world_size)
        extra["density_skew"] = {
            "workload": f"rank 0: {int(base * args.skew)} cells, every other rank: {base} cells, {args.map}x{args.map} slab each",
            "ms_per_step": r4["ms_per_step"], "value": total / (r4["ms_per_step"] / 1e3),
            "balanced_ms_per_step": r5["ms_per_step"], "balanced_value": world_size * (total // world_size) / (r5["ms_per_step"] / 1e3),
            "rebalanced_ms_per_step": r6["ms_per_step"], "rebalanced_value": total / (r6["ms_per_step"] / 1e3),
            "rebalance_s": t_rebalance, "slab_heights_after": heights, "cells_per_rank_after": w4.global_n_cells(),
            "note": "cells live with their pixels, so the step takes as long as the fullest slab's cells; "
                    "SlabWorld.rebalance() moves the slab boundaries (and the rows and cells behind them) until every "
                    "rank carries the same share of the step's algorithmic bytes"}
        del w4
    if not args.no_extra and world_size == 1 and args.config == "c4":
        extra["lifecycle"] = lifecycle_rounds(world, device, res["ms_per_step"])
        del world, pop
        torch.cuda.empty_cache()
        cc, cp, cmol, cmap, where = CONFIGS["c3"]
        w3, _ = build_world(cc, cp, cmol, cmap, cmap, device, seed=3, world_size=1, erange=erange)
        r3 = measure(w3, 200, 20, barrier, device, 1)
        w3.kinetics.list_cache = False
        r3s = measure(w3, 100, 10, barrier, device, 1)
        b3 = kinetics_bytes_per_cell(cp, 2 * cmol) * cc
        extra["c3"] = {
            "workload": f"{cc} cells saturating a {cmap}x{cmap} map, {cp} proteins, {2 * cmol} signals (BASELINE.json {where})",
            "value": cc / (r3["ms_per_step"] / 1e3), "unit": "cell-steps/s", "ms_per_step": r3["ms_per_step"],
            "kernel_ms": r3["kernel_ms"], "roofline_frac": b3 / (r3["kernel_ms"] / 1e3) / 1e9 / hbm_peak()[0],
            "scan_path": {"ms_per_step": r3s["ms_per_step"], "kernel_ms": r3s["kernel_ms"],
                          "value": cc / (r3s["ms_per_step"] / 1e3),
                          "roofline_frac": b3 / (r3s["kernel_ms"] / 1e3) / 1e9 / hbm_peak()[0]}}
        del w3
        extra["genome_update"] = genome_update_times(device)
        from bench_workloads import reference_workloads

        extra["reference_workloads"] = reference_workloads(device)
    if "parity_check" in extra and not extra["parity_check"]["ok"]:
        if rank == 0:
            print(json.dumps({"error": "multi-GPU parity check failed", "parity_check": extra["parity_check"]}))
        dist.destroy_process_group()
        sys.exit(2)

    if rank != 0:
        if world_size > 1:
            dist.destroy_process_group()
        return

    line = {
        "metric": METRIC, "value": value, "unit": "cell-steps/s", "n_gpus": world_size, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": res["ms_per_step"], "higher_is_better": True,
        "scaling": "strong" if split_over_ranks(args, world_size) else "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": dict(workload_config(args, world_size),
                       peer_exchange=bool(world_size > 1 and peer_active)), "roofline": roofline,
        "e2e": {"value": e2e_value, "unit": "cell-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": args.e2e_steps, "host_cpus_per_rank": n_cpus,
                "how": "every step uploads its inputs (cell molecules, molecule map, positions) from pinned host "
                       "memory and reads its results back; copies on their own streams, so the upload of step t+1 and "
                       "the read-back of step t-1 overlap the kernels of step t; timed with CUDA events and the host "
                       "clock around all steps (the larger of the two)"},
        # per step: kinetics (lane kernel + the launch for deferred cells), early-exit verdict, write-back, stencil,
        # correction, permeation
        "gpu_launches": 7 * args.steps,
        "clocks": clocks,
    }
    if extra:
        line["extra"] = extra
    if not args.no_cpu_baseline:
        cpop, n, n_passes = cpu_population(args)
        t_kin, t_diff, cores = cpu_step_time(cpop, n_passes)
        line["cpu_baseline"] = {"value": n * n_passes / (t_kin + t_diff), "unit": "cell-steps/s", "cores": cores,
                                "kind": "port", "sample": cpu_sample_text(args, n, n_passes, t_kin, t_diff, cores)}
    print(json.dumps(line))
    if world_size > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
