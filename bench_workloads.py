"""
The reference's own benchmark workloads, timed through this package's World API on one B200
(bench.py puts the result under `extra.reference_workloads` of its N = 1 line):

* `check_times`: what performance/check.py:45-117 times -- spawn_cells, update_cells, divide_cells
  ("replicate"), one enzymatic_activity, and mutations (point mutations + neighbours +
  recombinations) on 10,000 cells with ~1000-bp random genomes and the Wood-Ljungdahl chemistry;
  the published numbers (docstring of check.py:11-25) are quoted beside ours.
* `simulation_loop`: the step of performance/run_simulation.py:56-100 (BASELINE.json configs[1]) --
  activity, kill (ATP < 1), replicate (ATP > 5), recombinate, mutate, degrade + diffuse + lifetimes --
  at ~1,000 and ~40,000 cells on a 256 x 256 map, per-step and per-stage times; the published
  per-step notes are run_simulation.py:18-20.

Host clock with a device synchronize on both sides of every timed call, as the numbers are host
work + launches; the reference's scripts read time.time() without synchronizing.
"""
from __future__ import annotations

import random
import time

PUBLISHED_CHECK = {  # performance/check.py:11-25, seconds, magicsoup v0.14.1
    "T4 (g4dn.xlarge)": {"spawn_cells": 6.64, "update_cells": 5.95, "replicate_cells": 0.28,
                         "enzymatic_activity": 0.16, "mutations": 0.46},
    "CPU (i5-10210U)": {"spawn_cells": 7.66, "update_cells": 7.17, "replicate_cells": 0.37,
                        "enzymatic_activity": 4.51, "mutations": 0.40},
}
PUBLISHED_LOOP = {"1k cells": 0.03, "40k cells": 0.30, "40k cells, activity": 0.25,
                  "source": "performance/run_simulation.py:20 (2023-12-19, 'EC2 GPU'), seconds per step"}


def _genomes(n: int, size: int, spread: float = 0.1) -> list[str]:
    import magicsoup_b200 as ms

    sizes = (size - int(size * spread), size, size + int(size * spread))
    return [ms.random_genome(random.choice(sizes)) for _ in range(n)]


def _stats(ts: list[float]) -> dict:
    mu = sum(ts) / len(ts)
    sd = (sum((t - mu) ** 2 for t in ts) / len(ts)) ** 0.5
    return {"mean_s": mu, "sd_s": sd, "runs": len(ts)}


class _Clock:
    def __init__(self, device):
        import torch

        self.sync = lambda: torch.cuda.synchronize(device)

    def __call__(self, fn) -> float:
        self.sync()
        t0 = time.perf_counter()
        fn()
        self.sync()
        return time.perf_counter() - t0


def check_times(device, n_cells: int = 10_000, genome_size: int = 1_000, repeats: int = 3) -> dict:
    import magicsoup_b200 as ms
    from magicsoup_b200.examples.wood_ljungdahl import CHEMISTRY

    random.seed(21)
    clock = _Clock(device)
    dev = str(device)
    out: dict = {k: [] for k in ("spawn_cells", "update_cells", "replicate_cells", "enzymatic_activity", "mutations")}

    def fresh(spawn: bool = True):
        world = ms.World(chemistry=CHEMISTRY, device=dev)
        genomes = _genomes(n_cells, genome_size)
        idxs = world.spawn_cells(genomes=genomes) if spawn else []
        return world, genomes, idxs

    for _ in range(repeats):
        world, genomes, _ = fresh(spawn=False)
        out["spawn_cells"].append(clock(lambda: world.spawn_cells(genomes=genomes)))
        pairs = [(g, i) for i, g in enumerate(world.cell_genomes)]
        out["update_cells"].append(clock(lambda: world.update_cells(genome_idx_pairs=pairs)))
        out["enzymatic_activity"].append(clock(world.enzymatic_activity))  # first call: lists are built here
        idxs = list(range(world.n_cells))
        out["replicate_cells"].append(clock(lambda: world.divide_cells(cell_idxs=idxs)))

    world, genomes, _ = fresh()

    def mutate():
        ms.point_mutations(seqs=genomes)
        pairs = world.get_neighbors(cell_idxs=list(range(world.n_cells)))
        ms.recombinations(seq_pairs=[(genomes[a], genomes[b]) for a, b in pairs])

    for _ in range(repeats):
        out["mutations"].append(clock(mutate))
    steady = [clock(world.enzymatic_activity) for _ in range(5)]
    res = {k: _stats(v) for k, v in out.items()}
    res["enzymatic_activity_steady"] = _stats(steady[1:])
    return {"what": f"{n_cells} cells, ~{genome_size}-bp random genomes, Wood-Ljungdahl chemistry (14 molecules), "
                    f"128 x 128 map; performance/check.py:45-117", "ours": res, "published": PUBLISHED_CHECK,
            "proteins_max": int(world.kinetics.N.shape[1])}


def simulation_loop(device, n_start: int, n_steps: int, map_size: int = 256, genome_size: int = 500) -> dict:
    import torch

    import magicsoup_b200 as ms
    from magicsoup_b200.examples.wood_ljungdahl import CHEMISTRY

    random.seed(22)
    torch.manual_seed(22)
    clock = _Clock(device)
    world = ms.World(chemistry=CHEMISTRY, map_size=map_size, mol_map_init="randn", device=str(device))
    atp = [m.name for m in CHEMISTRY.molecules].index("ATP")
    stages: dict = {k: [] for k in ("addCells", "activity", "kill", "replicate", "recombinateGenomes",
                                    "mutateGenomes", "wrapUp", "perStep")}
    cells = []

    def kill():
        idx = torch.argwhere(world.cell_molecules[:, atp] < 1.0).flatten().tolist()
        world.kill_cells(cell_idxs=idx)

    def replicate():
        idx = torch.argwhere(world.cell_molecules[:, atp] > 5.0).flatten().tolist()
        world.cell_molecules[idx, atp] -= 4.0
        world.divide_cells(cell_idxs=idx)

    def wrap_up():
        world.degrade_molecules()
        world.diffuse_molecules()
        world.increment_cell_lifetimes()

    for _ in range(n_steps):
        step = 0.0
        t = 0.0
        if world.n_cells < n_start:
            genomes = [ms.random_genome(genome_size) for _ in range(n_start - world.n_cells)]
            t = clock(lambda: world.spawn_cells(genomes=genomes))
        stages["addCells"].append(t)
        step += t
        for name, fn in (("activity", world.enzymatic_activity), ("kill", kill), ("replicate", replicate),
                         ("recombinateGenomes", world.recombinate_cells), ("mutateGenomes", world.mutate_cells),
                         ("wrapUp", wrap_up)):
            t = clock(fn)
            stages[name].append(t)
            step += t
        stages["perStep"].append(step)
        cells.append(world.n_cells)
    skip = 1 if n_steps > 2 else 0  # the first step carries the one-off spawn of all cells
    return {"cells_start": n_start, "cells_end": cells[-1], "map": map_size, "genome_size": genome_size, "steps": n_steps,
            "mean_s": {k: sum(v[skip:]) / max(len(v[skip:]), 1) for k, v in stages.items()},
            "first_step_s": stages["perStep"][0], "proteins_max": int(world.kinetics.N.shape[1])}


def readme_loop(device, n_steps: int = 1000) -> dict:
    """BASELINE.json configs[0]: the loop of the reference's README (README.md:49-114)"""
    import torch

    import magicsoup_b200 as ms

    random.seed(23)
    torch.manual_seed(23)
    mols = [ms.Molecule("rl-NADPH", 200e3), ms.Molecule("rl-NADP", 100e3), ms.Molecule("rl-formiat", 20e3),
            ms.Molecule("rl-CO2", 10e3)]
    chem = ms.Chemistry(molecules=mols, reactions=[([mols[3], mols[0]], [mols[2], mols[1]])])
    world = ms.World(chemistry=chem, device=str(device))
    world.spawn_cells(genomes=[ms.random_genome(s=500) for _ in range(100)])

    def sample(p):
        return torch.argwhere(torch.bernoulli(p)).flatten().tolist()

    def loop():
        for _ in range(n_steps):
            world.enzymatic_activity()
            x = world.cell_molecules[:, 2]
            sizes = torch.tensor([float(len(g)) for g in world.cell_genomes], device=x.device)
            too_big = sample(sizes ** 7 / (sizes ** 7 + 3000.0 ** 7))  # docs/tutorials.md:191-208: without it genomes grow without bound
            world.kill_cells(cell_idxs=sorted(set(sample(0.01 / (0.01 + x)) + too_big)))
            x = world.cell_molecules[:, 2]
            world.divide_cells(cell_idxs=sample(x ** 3 / (x ** 3 + 20.0 ** 3)))
            world.mutate_cells(p=1e-4)
            world.recombinate_cells(p=1e-6)
            world.diffuse_molecules()

    t = _Clock(device)(loop)
    return {"what": "README.md:49-114: CO2 + NADPH <-> formiat + NADP, 100 random 500-bp genomes, 128 x 128 map, "
                    "activity / kill (+ the size-dependent kill of docs/tutorials.md:191-208) / replicate / mutate / "
                    "recombinate / diffuse", "steps": n_steps, "total_s": t,
            "ms_per_step": t / n_steps * 1e3, "cells_end": world.n_cells}


def reference_workloads(device) -> dict:
    return {
        "readme_loop": readme_loop(device),
        "check": check_times(device),
        "simulation_loop": {
            "what": "the step of performance/run_simulation.py:56-100 (BASELINE.json configs[1]): Wood-Ljungdahl "
                    "chemistry, 500-bp random genomes, 256 x 256 map, cells kept at >= cells_start by spawning",
            "1k cells": simulation_loop(device, 1_000, 20),
            "40k cells": simulation_loop(device, 40_000, 8),
            "published": PUBLISHED_LOOP,
        },
    }


if __name__ == "__main__":
    import json

    import torch

    print(json.dumps(reference_workloads(torch.device("cuda", 0)), indent=1))
