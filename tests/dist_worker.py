"""
Worker of the multi-rank tests (spawned by tests/test_distributed_cpu.py with gloo on CPU,
or launched with torchrun + NCCL on GPUs by tests/test_gpu_distributed.py / by hand):

    python -m torch.distributed.run --nproc-per-node 2 tests/dist_worker.py --backend nccl

Every rank owns one row slab of a seeded global world (cells live with their pixels),
runs the step loop through magicsoup_b200.distributed.SlabWorld and rank 0 checks that the
gathered result is BIT-IDENTICAL to the CPU oracle run on the whole map.

With --backend gloo there is no GPU: the five kernel-launch hooks of World are replaced by
oracle calls (test-only subclass below), so the test covers the host-side exchange logic:
halo rows, integer plane sums, early-exit flags, cell migration and re-indexing.
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

from magicsoup_b200.distributed import SlabWorld  # noqa: E402
from magicsoup_b200.synthetic import make_population  # noqa: E402
from magicsoup_b200.world import Chemistry, Molecule  # noqa: E402
from oracle import oracle as O  # noqa: E402

PARAMS = ("N", "Nf", "Nb", "A", "Kmr", "Ke", "Kmf", "Kmb", "Vmax")


class OracleSlabWorld(SlabWorld):
    """SlabWorld whose kernel launches are CPU oracle calls (tests only)"""

    def _device_for_step(self):
        return torch.device("cpu")

    def _launch_activity(self, dev, pos, cm, mm, X, eq):
        kin = self.kinetics
        P = O.Params(*[getattr(kin, k).numpy() for k in PARAMS])
        xs, ys = pos[:, 0].long(), pos[:, 1].long()
        X0 = torch.cat([cm, mm[:, xs, ys].T], dim=1).numpy()
        out, flags = O.integrate_signals_iters(P, X0, eq)
        X.copy_(torch.from_numpy(out))
        return [int(f) for f in flags]

    def _launch_scatter(self, dev, X, pos, mm):
        m = self.n_molecules
        xs, ys = pos[:, 0].long(), pos[:, 1].long()
        mm[:, xs, ys] = X[:, m:].T
        return X[:, :m].clone()

    def _const_f32(self, key, values, dev):
        return torch.tensor(values, dtype=torch.float32)

    def _launch_degrade(self, dev, mm, cm):
        f = torch.tensor(self._mol_degrads, dtype=torch.float32)
        mm *= f[:, None, None]
        cm *= f[None, :]

    def _launch_stencil(self, dev, mm, top, bottom, decay=None):
        a = np.array([w[0] for w in self._diffusion], dtype=np.float32)
        b = np.array([w[1] for w in self._diffusion], dtype=np.float32)
        mmn = mm.numpy()
        if decay is not None:  # degradation folded in: every value the stencil reads is decayed first
            f = decay.numpy()
            mmn = mmn * f[:, None, None]
            top = None if top is None else top * decay[:, None]
            bottom = None if bottom is None else bottom * decay[:, None]
        out, sums = O.diffuse_stencil(mmn, None if top is None else top.numpy(),
                                      None if bottom is None else bottom.numpy(), a, b)
        return torch.from_numpy(out), torch.from_numpy(sums)

    def _launch_correct(self, dev, out, mm, sums, n_total):
        mm.copy_(torch.from_numpy(O.diffuse_correct(out.numpy(), sums.numpy(), n_total)))

    def _launch_permeate(self, dev, cm, mm, pos, decay=None):
        q = np.array(self._permeation, dtype=np.float32)
        cm2, mm2 = cm.numpy().copy(), mm.numpy()
        if decay is not None:
            cm2 = cm2 * decay.numpy()[None, :]
        import ctypes as C
        lib = O._lib(np.float32)
        m, H, W = mm2.shape
        posn = np.ascontiguousarray(pos.numpy(), dtype=np.int32)
        lib.ms_oracle_permeate(O._p(cm2), O._p(mm2), O._p(posn), O._p(q), C.c_int64(cm2.shape[0]),
                               C.c_int64(m), C.c_int64(H), C.c_int64(W))
        return torch.from_numpy(cm2)


def gather_world(world: SlabWorld):
    """all ranks' map slabs and cells on rank 0 (global coordinates, rank order)"""
    parts = [None] * world.world_size
    payload = {
        "map": world.molecule_map.cpu().numpy(),
        "cm": world.cell_molecules.cpu().numpy(),
        "pos": world.global_positions().cpu().numpy(),
    }
    dist.all_gather_object(parts, payload, group=world.group)
    mm = np.concatenate([p["map"] for p in parts], axis=1)
    cm = np.concatenate([p["cm"] for p in parts], axis=0)
    pos = np.concatenate([p["pos"] for p in parts], axis=0)
    return mm, cm, pos


def gather_params(world: SlabWorld):
    """the nine parameter tensors of all ranks' cells on every rank (rank order), padded to the widest protein axis"""
    parts = [None] * world.world_size
    mine = {k: v.cpu().numpy() for k, v in world.kinetics.cell_params(list(range(world.n_cells))).items()}
    dist.all_gather_object(parts, mine, group=world.group)
    p = max(q["N"].shape[1] for q in parts)
    out = {}
    for k in PARAMS:
        padded = []
        for q in parts:
            a = q[k]
            pad = [(0, 0), (0, p - a.shape[1])] + [(0, 0)] * (a.ndim - 2)
            padded.append(np.pad(a, pad))
        out[k] = np.concatenate(padded, axis=0)
    return out


def run(backend: str, steps: int, cells: int, rows_per_rank: int, cols: int, seed: int, quiet: bool = False,
        rows: int | None = None, deferred: bool = True) -> None:
    rank, ws = dist.get_rank(), dist.get_world_size()
    if backend == "nccl":
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
        device = f"cuda:{torch.cuda.current_device()}"
        cls = SlabWorld
    else:
        device = "cpu"
        cls = OracleSlabWorld
    H = rows if rows else rows_per_rank * ws  # `rows` need not be a multiple of the number of ranks
    # the same seeded global population on every rank; each keeps the cells of its slab
    pop = make_population(cells, 12, 4, (H, cols), seed=seed)
    mols = [Molecule(f"d{i}", 10e3, diffusivity=pop.diffusivities[i], permeability=pop.permeabilities[i])
            for i in range(pop.m)]
    world = cls(chemistry=Chemistry(molecules=mols, reactions=[]), global_map_size=(H, cols), device=device)
    world.deferred_settle = deferred  # all ranks run the same sequence of calls here (no effect on the gloo hooks)
    r0, r1 = world.row0, world.row0 + world.slab_rows
    world.molecule_map = pop.molecule_map[:, r0:r1].clone().to(device)
    mine = torch.nonzero((pop.cell_positions[:, 0] >= r0) & (pop.cell_positions[:, 0] < r1)).flatten()
    local_pos = pop.cell_positions[mine].clone()
    local_pos[:, 0] -= r0
    world.add_cells_from_params(local_pos, cell_molecules=pop.cell_molecules[mine],
                                **{k: getattr(pop, k)[mine] for k in PARAMS})
    counts = world.global_n_cells()
    assert sum(counts) == cells and world.global_cell_offset() == sum(counts[:rank])

    # the oracle on the whole map, cells in rank order (= the gathered order)
    order = torch.argsort(world.owner_of_rows(pop.cell_positions[:, 0].long()), stable=True)
    P = O.Params(*[getattr(pop, k)[order].numpy() for k in PARAMS])
    o_pos = pop.cell_positions[order].numpy()
    o_cm, o_mm = pop.cell_molecules[order].numpy(), pop.molecule_map.numpy()

    for step in range(steps):
        world.enzymatic_activity()
        if step % 2 == 0:
            world.degrade_molecules()
            world.diffuse_molecules()
        else:  # the same, with the degradation folded into the diffusion kernels
            world.diffuse_molecules(degrade_first=True)
        o_cm, o_mm = O.enzymatic_activity(P, o_pos, o_cm, o_mm)
        f = np.array(world._mol_degrads, dtype=np.float32)
        o_cm, o_mm = o_cm * f[None, :], o_mm * f[:, None, None]
        o_cm, o_mm = O.diffuse_molecules(o_cm, o_mm, o_pos, pop.diffusivities, pop.permeabilities)
        mm, cm, pos = gather_world(world)
        if rank == 0:
            assert np.array_equal(pos, o_pos), f"step {step}: positions differ"
            assert np.array_equal(mm, o_mm), f"step {step}: map differs from the single-domain oracle"
            assert np.array_equal(cm, o_cm), f"step {step}: cell molecules differ from the oracle"

    # re-indexing: kill some cells, move a cell across the slab boundary, step again
    kill = [i for i in range(world.n_cells) if i % 7 == 3]
    world.kill_cells(kill)
    # first local cell steps over the upper edge of its slab (the map is a torus) onto a free
    # pixel of the neighbour's last row
    last_rows = [None] * ws
    dist.all_gather_object(last_rows, world.cell_map[-1].cpu().tolist(), group=world.group)
    free_cols = [y for y, occ in enumerate(last_rows[(rank - 1) % ws]) if not occ]
    if world.n_cells > 0 and free_cols and ws > 1:
        x, y = int(world.cell_positions[0, 0]), int(world.cell_positions[0, 1])
        world.cell_map[x, y] = False
        world.cell_positions[0, 0] = -1
        world.cell_positions[0, 1] = free_cols[0]
    received = world.migrate_cells()
    if ws > 1:
        assert received == 1, received
    world.enzymatic_activity()
    world.diffuse_molecules()
    mm, cm, pos = gather_world(world)
    total = int(sum(world.global_n_cells()))
    if rank == 0:
        assert cm.shape[0] == total == pos.shape[0]
        assert len({(int(a), int(b)) for a, b in pos}) <= total
        assert np.isfinite(mm).all() and np.isfinite(cm).all() and (mm >= 0).all()
    assert world.kinetics.N.shape[0] == world.n_cells == world.cell_molecules.shape[0]

    # bulk migration: every cell in a slab's first row steps up onto the neighbour's last row where
    # that pixel is free; cells keep their molecules and parameters (checked by global position)
    def snapshot():
        parts = [None] * ws
        payload = {"pos": world.global_positions().cpu().numpy() % np.array([[H, cols]]),
                   "cm": world.cell_molecules.cpu().numpy(), "vmax": world.kinetics.cell_params(
                       list(range(world.n_cells)))["Vmax"].cpu().numpy()}
        dist.all_gather_object(parts, payload, group=world.group)
        return {(int(a), int(b)): (c.tobytes(), v.tobytes())
                for p_ in parts for (a, b), c, v in zip(p_["pos"], p_["cm"], p_["vmax"])}

    last_rows = [None] * ws
    dist.all_gather_object(last_rows, world.cell_map[-1].cpu().tolist(), group=world.group)
    above = last_rows[(rank - 1) % ws]
    movers = [i for i in range(world.n_cells)
              if int(world.cell_positions[i, 0]) == 0 and not above[int(world.cell_positions[i, 1])]]
    if ws > 1:
        for i in movers:
            world.cell_map[0, int(world.cell_positions[i, 1])] = False
            world.cell_positions[i, 0] = -1
    before = snapshot()
    n_moved = [None] * ws
    dist.all_gather_object(n_moved, len(movers) if ws > 1 else 0, group=world.group)
    received = world.migrate_cells()
    assert received == (n_moved[(rank + 1) % ws] if ws > 1 else 0), (received, n_moved)
    after = snapshot()
    assert before == after and len(after) == int(sum(world.global_n_cells()))
    assert bool((world.cell_positions[:, 0] >= 0).all()) and bool((world.cell_positions[:, 0] < world.slab_rows).all())
    world.enzymatic_activity()
    world.diffuse_molecules()

    # load balance: rank 0 loses most of its cells, the slab boundaries follow the cells; the world is
    # the same before and after, and the next step equals the single-domain oracle's step on it
    if rank == 0:
        world.kill_cells([i for i in range(world.n_cells) if i % 5 != 0])
    mm0, cm0, pos0 = gather_world(world)
    before = snapshot()
    heights0 = list(world.slab_heights)
    cost = world.row_costs()
    heights1 = world.rebalance(tolerance=0.01)
    assert heights1 == world.slab_heights and sum(heights1) == H and min(heights1) >= 1
    assert world.molecule_map.shape[1] == world.slab_rows == heights1[rank]
    mm1, cm1, pos1 = gather_world(world)
    assert before == snapshot()
    assert np.array_equal(mm0, mm1), "the map changed under rebalance()"
    assert int(world.cell_map.sum()) == world.n_cells
    xs, ys = world.cell_positions[:, 0].long(), world.cell_positions[:, 1].long()
    assert world.n_cells == 0 or bool(world.cell_map[xs, ys].all())
    if ws > 1 and cells >= 50:  # (a handful of cells leaves nothing to balance)
        def fullest(hs):
            at, worst = 0, 0.0
            for h in hs:
                worst = max(worst, float(cost[at:at + h].sum()))
                at += h
            return worst
        assert heights1 != heights0 and fullest(heights1) < fullest(heights0), (heights0, heights1)
    par = gather_params(world)
    o_cm, o_mm = O.enzymatic_activity(O.Params(*[par[k] for k in PARAMS]), pos1, cm1, mm1)
    o_cm, o_mm = O.diffuse_molecules(o_cm, o_mm, pos1, pop.diffusivities, pop.permeabilities)
    world.enzymatic_activity()
    world.diffuse_molecules()
    mm2, cm2, pos2 = gather_world(world)
    if rank == 0:
        assert np.array_equal(pos2, pos1)
        assert np.array_equal(mm2, o_mm), "map differs from the single-domain oracle after rebalance()"
        assert np.array_equal(cm2, o_cm), "cell molecules differ from the oracle after rebalance()"
    if rank == 0 and not quiet:
        print(f"dist_worker ok: backend={backend} ranks={ws} steps={steps} cells={cells} received={received}")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--backend", default="gloo")
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--cells", type=int, default=300)
    ap.add_argument("--rows-per-rank", type=int, default=16)
    ap.add_argument("--rows", type=int, default=None, help="total map rows (need not divide by the number of ranks)")
    ap.add_argument("--blocking", action="store_true", help="SlabWorld.deferred_settle = False (the multi-rank default)")
    ap.add_argument("--cols", type=int, default=24)
    ap.add_argument("--seed", type=int, default=5)
    args = ap.parse_args()
    dist.init_process_group(args.backend)
    try:
        run(args.backend, args.steps, args.cells, args.rows_per_rank, args.cols, args.seed, rows=args.rows,
            deferred=not args.blocking)
    finally:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
