"""
GPU tests of the parameter store's row map and of the list cache (ms_*_cached entry points,
include/magicsoup_b200.h): whatever the store does behind the nine public attributes -- spare
capacity, rows removed without moving data, cached sparse lists -- the kernels must return the
oracle's bits for the dense tensors the attributes show.
"""
import numpy as np
import pytest
import torch

from tests.helpers import make_population, bits_equal, describe_mismatch, dense_adversarial, PARAMS
from oracle import oracle as O

pytestmark = pytest.mark.gpu


class _Chem:
    def __init__(self, m):
        from magicsoup_b200.world import Molecule
        self.molecules = [Molecule(f"st{i}", 10e3) for i in range(m)]
        self.reactions = []


def _store(pop, device="cuda"):
    import magicsoup_b200 as ms
    kin = ms.Kinetics(chemistry=_Chem(pop.m), device=device)
    kin.increase_max_proteins(pop.p)
    kin.increase_max_cells(pop.c)
    kin.set_cell_params_from_tensors(list(range(pop.c)), **{k: getattr(pop, k) for k in PARAMS})
    return kin


def _oracle(kin, X):
    P = O.Params(*[kin.cell_params(list(range(kin._n)))[k].cpu().numpy() for k in PARAMS])
    return O.integrate_signals(P, X.cpu().numpy())


@pytest.mark.parametrize("shape", [(700, 64, 16, 32, 3), (333, 35, 14, 32, 4), (100, 22, 4, 16, 5)],
                         ids=["p64s32", "p35s28", "p22s8"])
def test_cached_lists_give_the_same_bits(shape):
    c, p, m, size, seed = shape
    pop = make_population(c, p, m, size, seed=seed)
    kin = _store(pop)
    X = pop.X0().cuda()
    ora = _oracle(kin, X)
    first = kin.integrate_signals(X)   # scans the tiles and stores the lists
    assert kin._cache is not None and int((kin._cache[: c, 0] >> 24 == 1).sum()) > 0.9 * c
    second = kin.integrate_signals(X)  # integrates from the stored lists
    assert bits_equal(first, ora), describe_mismatch(first, ora)
    assert bits_equal(second, ora), describe_mismatch(second, ora)
    kin.list_cache = False
    third = kin.integrate_signals(X)
    assert kin._cache is None and bits_equal(third, ora)


def test_overflowing_and_unpackable_cells_with_cache():
    pop = make_population(160, 32, 8, 16, seed=21)
    dense = dense_adversarial(160, 32, 16, seed=22)
    pick = torch.arange(160) % 5 == 0
    for k in PARAMS:
        getattr(pop, k)[pick] = getattr(dense, k)[pick]
    pop.Nf[7, 3, 2] = 300  # beyond the 8-bit packing
    pop.N[7, 3, 2] = -300
    kin = _store(pop)
    X = pop.X0().cuda()
    ora = _oracle(kin, X)
    for _ in range(3):
        out = kin.integrate_signals(X)
        assert bits_equal(out, ora), describe_mismatch(out, ora)
    status = (kin._cache[:160, 0] >> 24).cpu()
    assert set(status[pick].tolist()) == {2} and int((status == 1).sum()) == int((~pick).sum()) - 1


def test_row_map_and_cache_through_a_lifecycle():
    g = torch.Generator().manual_seed(5)
    pop = make_population(600, 64, 16, 32, seed=6)
    src = {k: getattr(pop, k) for k in PARAMS}
    kin = _store(pop)
    n = 600
    X = pop.X0().cuda()
    for step in range(6):
        ora = _oracle(kin, X)
        out = kin.integrate_signals(X)
        assert bits_equal(out, ora), f"step {step}: " + describe_mismatch(out, ora)
        # kill a fifth of the cells (order preserved), divide a few, mutate a few, add new ones
        keep = torch.rand(n, generator=g) > 0.2
        kin.remove_cell_params(keep=keep)
        X = out[keep.cuda()]
        n = int(keep.sum())
        assert kin._rows is not None and kin._n == n
        parents = torch.randint(0, n, (20,), generator=g).tolist()
        kin.increase_max_cells(by_n=20)
        kin.copy_cell_params(from_idxs=parents, to_idxs=list(range(n, n + 20)))
        X = torch.cat([X, X[parents]])
        n += 20
        mutated = torch.randint(0, n, (15,), generator=g).tolist()
        donors = torch.randint(0, 600, (15,), generator=g)
        kin.set_cell_params_from_tensors(mutated, **{k: v[donors] for k, v in src.items()})
        kin.unset_cell_params(cell_idxs=[0, n - 1])
        if step == 3:  # a user pokes a value through the attribute view (compacts, then invalidates)
            kin.Vmax[5] *= 0.5
            assert kin._rows is None
    ora = _oracle(kin, X)
    out = kin.integrate_signals(X)
    assert bits_equal(out, ora), describe_mismatch(out, ora)


def test_world_steps_with_store_match_oracle():
    import magicsoup_b200 as ms
    pop = make_population(500, 40, 8, 32, seed=9)
    mols = [ms.Molecule(f"sw{i}", 10e3, diffusivity=pop.diffusivities[i], permeability=pop.permeabilities[i])
            for i in range(pop.m)]
    w = ms.World(chemistry=ms.Chemistry(molecules=mols, reactions=[]), map_size=32, device="cuda")
    w.molecule_map = pop.molecule_map.clone().cuda()
    w.add_cells_from_params(pop.cell_positions, cell_molecules=pop.cell_molecules,
                            **{k: getattr(pop, k) for k in PARAMS})
    cm, mm = pop.cell_molecules.numpy(), pop.molecule_map.numpy()
    pos = pop.cell_positions.numpy()
    keep = np.ones(500, dtype=bool)
    for step in range(4):
        idx = np.nonzero(keep)[0]
        P = O.Params(*[getattr(pop, k)[idx].numpy() for k in PARAMS])
        cm, mm = O.enzymatic_activity(P, pos, cm, mm)
        cm, mm = O.diffuse_molecules(cm, mm, pos, pop.diffusivities, pop.permeabilities)
        w.enzymatic_activity()
        w.diffuse_molecules()
        assert bits_equal(w.cell_molecules, cm), f"step {step}: " + describe_mismatch(w.cell_molecules, cm)
        assert bits_equal(w.molecule_map, mm), f"step {step}"
        if step < 3:  # kill every 7th surviving cell: spill + row map edit, cached lists stay valid
            kill = list(range(step, len(idx), 7))
            mm = mm.copy()
            mm[:, pos[kill, 0], pos[kill, 1]] += cm[kill].T
            w.kill_cells(kill)
            sub = np.ones(len(idx), dtype=bool)
            sub[kill] = False
            keep[idx[kill]] = False
            cm, pos = cm[sub], pos[sub]
