"""
Many steps of an N-rank SlabWorld under lifecycle churn (kills, cells crossing slab boundaries,
rebalanced slabs) against ONE single-GPU World that holds the whole map and all cells, on real
kernels (NCCL; launched by tests/test_gpu_distributed.py):

    python -m torch.distributed.run --nproc-per-node 2 tests/dist_stress_worker.py [--blocking] [--steps 24]

Rank 0 drives the single-GPU World through the same operations, identified by global pixel, and
compares after every step: the map bit for bit, and every cell's molecules by the pixel it sits on.
The single-GPU World is itself held to the CPU oracle by tests/test_gpu_world.py.
"""
from __future__ import annotations

import argparse
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

import magicsoup_b200 as ms  # noqa: E402
from magicsoup_b200.distributed import SlabWorld  # noqa: E402
from magicsoup_b200.synthetic import make_population  # noqa: E402

PARAMS = ("N", "Nf", "Nb", "A", "Kmr", "Ke", "Kmf", "Kmb", "Vmax")


def gathered(world: SlabWorld):
    """(global map, {global pixel: cell molecule bytes}) on every rank"""
    H, W = world.global_map_size
    parts = [None] * world.world_size
    pos = world.global_positions().cpu().numpy() % np.array([[H, W]])
    dist.all_gather_object(parts, {"map": world.molecule_map.cpu().numpy(), "pos": pos,
                                   "cm": world.cell_molecules.cpu().numpy()}, group=world.group)
    mm = np.concatenate([p["map"] for p in parts], axis=1)
    cells = {(int(a), int(b)): c.tobytes() for p in parts for (a, b), c in zip(p["pos"], p["cm"])}
    assert len(cells) == sum(len(p["pos"]) for p in parts), "two cells on one pixel"
    return mm, cells


def run(steps: int, cells: int, rows_per_rank: int, cols: int, deferred: bool, seed: int = 3):
    rank, ws = dist.get_rank(), dist.get_world_size()
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    device = f"cuda:{torch.cuda.current_device()}"
    H = rows_per_rank * ws + 3  # slabs of unequal heights from the start
    pop = make_population(cells, 12, 4, (H, cols), seed=seed)
    mols = [ms.Molecule(f"st{i}", 10e3, half_life=1000 + 50 * i, diffusivity=pop.diffusivities[i],
                        permeability=pop.permeabilities[i]) for i in range(pop.m)]
    chem = ms.Chemistry(molecules=mols, reactions=[])
    world = SlabWorld(chemistry=chem, global_map_size=(H, cols), device=device)
    world.deferred_settle = deferred
    r0, r1 = world.row0, world.row0 + world.slab_rows
    world.molecule_map = pop.molecule_map[:, r0:r1].clone().to(device)
    mine = torch.nonzero((pop.cell_positions[:, 0] >= r0) & (pop.cell_positions[:, 0] < r1)).flatten()
    local = pop.cell_positions[mine].clone()
    local[:, 0] -= r0
    world.add_cells_from_params(local, cell_molecules=pop.cell_molecules[mine], **{k: getattr(pop, k)[mine] for k in PARAMS})
    ref = None
    if rank == 0:  # the whole problem on one GPU
        ref = ms.World(chemistry=chem, map_size=(H, cols), device=device)
        ref.molecule_map = pop.molecule_map.clone().to(device)
        ref.add_cells_from_params(pop.cell_positions, cell_molecules=pop.cell_molecules, **{k: getattr(pop, k) for k in PARAMS})

    def ref_indices(pixels: set) -> list[int]:
        return [i for i, (a, b) in enumerate(ref.cell_positions.tolist()) if (a, b) in pixels]

    rng = np.random.default_rng(seed)  # the same stream on every rank
    n_migrated = n_killed = 0
    for step in range(steps):
        for w in (world, ref):
            if w is None:
                continue
            w.enzymatic_activity()
            if step % 2:
                w.diffuse_molecules(degrade_first=True)
            else:
                w.degrade_molecules()
                w.diffuse_molecules()
        mm, cells_now = gathered(world)
        pixels = sorted(cells_now)
        if step % 3 == 1:  # kill every 9th cell (by pixel order), plus most of the first slab once
            doomed = {p for i, p in enumerate(pixels) if i % 9 == int(rng.integers(9))}
            if step == 7:
                doomed |= {p for p in pixels if p[0] < world.row_starts[1] and p[1] % 4 != 0}
            gp = (world.global_positions().cpu().numpy() % np.array([[H, cols]])).tolist()
            world.kill_cells([i for i, (a, b) in enumerate(gp) if (a, b) in doomed])
            if ref is not None:
                ref.kill_cells(ref_indices(doomed))
            n_killed += len(doomed)
        if step % 4 == 2:  # every cell in the first row of a slab steps up where the pixel above is free
            occupied = set(pixels)
            starts = set(world.row_starts[:-1])
            movers = {p for p in pixels if p[0] in starts and ((p[0] - 1) % H, p[1]) not in occupied}
            gp = (world.global_positions().cpu().numpy() % np.array([[H, cols]])).tolist()
            for i, (a, b) in enumerate(gp):
                if (a, b) in movers:
                    world.cell_map[int(world.cell_positions[i, 0]), b] = False
                    world.cell_positions[i, 0] -= 1  # row -1: the slab above (the map is a torus)
            world.migrate_cells()
            if ref is not None:
                for i in ref_indices(movers):
                    a, b = (int(v) for v in ref.cell_positions[i])
                    ref.cell_map[a, b] = False
                    ref.cell_positions[i, 0] = (a - 1) % H
                    ref.cell_map[(a - 1) % H, b] = True
            n_migrated += len(movers)
        if step % 6 == 5:
            world.rebalance(tolerance=0.02)
        # compare
        mm, cells_now = gathered(world)
        if ref is not None:
            want = {(int(a), int(b)): c.tobytes()
                    for (a, b), c in zip(ref.cell_positions.tolist(), ref.cell_molecules.cpu().numpy())}
            assert np.array_equal(mm, ref.molecule_map.cpu().numpy()), f"step {step}: maps differ"
            assert cells_now == want, f"step {step}: cells differ ({len(cells_now)} vs {len(want)})"
        assert int(world.cell_map.sum()) == world.n_cells == world.kinetics.N.shape[0]
    assert world.peer_status() == 0
    if rank == 0:
        print(f"dist_stress ok: ranks={ws} steps={steps} deferred={deferred} peer={world.__dict__.get('_peer_ctx') is not None} "
              f"cells {cells} -> {len(cells_now)} killed={n_killed} moved across slabs={n_migrated} heights={world.slab_heights}")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=24)
    ap.add_argument("--cells", type=int, default=3000)
    ap.add_argument("--rows-per-rank", type=int, default=48)
    ap.add_argument("--cols", type=int, default=64)
    ap.add_argument("--blocking", action="store_true")
    args = ap.parse_args()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    try:
        run(args.steps, args.cells, args.rows_per_rank, args.cols, deferred=not args.blocking)
    finally:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
