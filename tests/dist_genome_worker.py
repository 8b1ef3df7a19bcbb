"""
The genome-driven loop of the reference (performance/run_simulation.py:56-100: spawn, activity, kill,
divide, recombinate, mutate, move, degrade, diffuse) on an N-rank SlabWorld with real kernels (NCCL;
launched by tests/test_gpu_distributed.py).  Every rank spawns, divides and mutates its own cells;
the Moore-neighbourhood searches stay inside the slab.  Checked on every rank and step:

  * the step itself: enzymatic_activity + diffuse_molecules of all ranks together equal the CPU
    oracle on the gathered world, bit for bit (parameters as the ranks' stores show them);
  * the bookkeeping: one cell per pixel, all cells inside their slab, per-cell lists and tensors of
    one length, and -- at the end -- parameters equal to a fresh derivation from the genomes.
"""
from __future__ import annotations

import os
import random
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

import magicsoup_b200 as ms  # noqa: E402
from magicsoup_b200._hostlib import lib as hostlib  # noqa: E402
from magicsoup_b200.distributed import SlabWorld  # noqa: E402
from oracle import oracle as O  # noqa: E402
from tests.dist_worker import gather_params, gather_world  # noqa: E402

PARAMS = ("N", "Nf", "Nb", "A", "Kmr", "Ke", "Kmf", "Kmb", "Vmax")


def run(steps: int = 8, rows_per_rank: int = 20, cols: int = 24, genomes_per_rank: int = 120):
    rank, ws = dist.get_rank(), dist.get_world_size()
    device = f"cuda:{torch.cuda.current_device()}"
    random.seed(50)  # the same chemistry and token maps on every rank ...
    torch.manual_seed(50)
    mols = [ms.Molecule(f"dg{i}", e, diffusivity=d, permeability=q)
            for i, (e, d, q) in enumerate(zip((10e3, 25e3, 60e3, 150e3, 8e3), (1.0, 0.1, 0.01, 0.0, 0.5), (1.0, 0.1, 0.0, 0.5, 0.0)))]
    chem = ms.Chemistry(molecules=mols, reactions=[([mols[0]], [mols[1]]), ([mols[1], mols[2]], [mols[3]]),
                                                    ([mols[4], mols[4]], [mols[0], mols[2]])])
    H = rows_per_rank * ws + 1
    world = SlabWorld(chemistry=chem, global_map_size=(H, cols), device=device)
    world.genetics = ms.Genetics(p_catal_dom=0.1, p_transp_dom=0.1, p_reg_dom=0.1)
    world.deferred_settle = True
    g = torch.Generator().manual_seed(51)
    whole = torch.randn(len(mols), H, cols, generator=g).abs() * 3 + 5
    world.molecule_map = whole[:, world.row0:world.row0 + world.slab_rows].clone().to(device)
    random.seed(100 + rank)  # ... different genomes and choices per rank
    hostlib().seed(200 + rank)
    world.spawn_cells(genomes=[ms.random_genome(random.choice((300, 600))) for _ in range(genomes_per_rank)])
    assert world.n_cells == genomes_per_rank
    diffs, perms = [m.diffusivity for m in mols], [m.permeability for m in mols]
    for step in range(steps):
        # the step against the oracle on the gathered world
        mm0, cm0, pos0 = gather_world(world)
        par = gather_params(world)
        o_cm, o_mm = O.enzymatic_activity(O.Params(*[par[k] for k in PARAMS]), pos0, cm0, mm0)
        o_cm, o_mm = O.diffuse_molecules(o_cm, o_mm, pos0, diffs, perms)
        world.enzymatic_activity()
        world.diffuse_molecules()
        mm1, cm1, pos1 = gather_world(world)
        assert np.array_equal(pos1, pos0)
        assert np.array_equal(mm1, o_mm), f"step {step}: map differs from the oracle"
        assert np.array_equal(cm1, o_cm), f"step {step}: cell molecules differ from the oracle"
        # the reference's loop, every rank on its own cells
        world.kill_cells(cell_idxs=[i for i in range(world.n_cells) if i % 9 == step % 9])
        pairs = world.divide_cells(cell_idxs=[i for i in range(world.n_cells) if i % 4 == 0])
        assert all(world.cell_genomes[a] == world.cell_genomes[b] for a, b in pairs)
        world.recombinate_cells(p=1e-4)
        world.mutate_cells(p=1e-3)
        world.move_cells(cell_idxs=list(range(0, world.n_cells, 3)))
        world.degrade_molecules()
        world.increment_cell_lifetimes()
        n = world.n_cells
        assert len(world.cell_genomes) == len(world.cell_labels) == n == world.cell_molecules.shape[0] == world.kinetics.N.shape[0]
        xs, ys = world.cell_positions[:, 0].long(), world.cell_positions[:, 1].long()
        assert bool((xs >= 0).all()) and bool((xs < world.slab_rows).all()) and bool((ys >= 0).all()) and bool((ys < cols).all())
        assert int(world.cell_map.sum()) == n and bool(world.cell_map[xs, ys].all())
        assert len({(int(a), int(b)) for a, b in world.cell_positions.tolist()}) == n
    # parameters follow the genomes
    fresh = ms.Kinetics(chemistry=chem, device=device)
    for name in ("km_map", "vmax_map", "sign_map", "hill_map", "reaction_map", "transport_map", "effector_map"):
        setattr(fresh, name, getattr(world.kinetics, name))
    proteomes = world.genetics.translate_genomes(world.cell_genomes)
    fresh.increase_max_cells(by_n=world.n_cells)
    fresh.increase_max_proteins(max_n=int(world.kinetics.N.shape[1]))
    keep = [i for i, p in enumerate(proteomes) if p]
    fresh.set_cell_params(cell_idxs=keep, proteomes=[proteomes[i] for i in keep])
    for k in ("N", "Nf", "Nb", "A", "Vmax"):
        assert torch.equal(getattr(world.kinetics, k), getattr(fresh, k)), k
    counts = world.global_n_cells()
    assert world.peer_status() == 0
    if rank == 0:
        print(f"dist_genome ok: ranks={ws} steps={steps} cells per rank {counts} proteins {int(world.kinetics.N.shape[1])}")


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    try:
        run()
    finally:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
