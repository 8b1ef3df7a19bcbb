"""
CPU-only tests: the C-ABI library loads and exports every symbol include/magicsoup_b200.h
declares, the host-side logic (early-exit protocol, parameter store, cell bookkeeping)
and the no-CPU-fallback rule.  No kernel is launched here.
"""
import ctypes
import re
from pathlib import Path

import pytest
import torch

import magicsoup_b200 as ms
from magicsoup_b200 import _native as nat
from magicsoup_b200.kinetics import _settle_early_exit, _ALL_PARAMS
from magicsoup_b200.synthetic import make_population
from magicsoup_b200.world import diffusion_weights, permeation_factor

ROOT = Path(__file__).resolve().parents[1]


def test_library_exports_every_declared_symbol():
    header = (ROOT / "include" / "magicsoup_b200.h").read_text()
    declared = set(re.findall(r"\b(ms_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 18
    lib = nat.lib()
    for name in declared:
        assert hasattr(lib, name), name
    assert declared <= set(nat.exported_symbols())
    assert lib.ms_version() >= 100
    assert lib.ms_error_string(-3).decode().startswith("scratch")
    assert lib.ms_kinetics_scratch_bytes(1000, 64, 32) > 4000
    assert lib.ms_diffuse_scratch_bytes(16, 1024, 1024) >= 16 * 1024 * 1024 * 4


def test_no_cpu_fallback():
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    pop = make_population(4, 3, 2, 4, seed=0)
    kin = ms.Kinetics(chemistry=ms.Chemistry([ms.Molecule("a", 1.0), ms.Molecule("b", 2.0)], []))
    for k in _ALL_PARAMS:
        setattr(kin, k, getattr(pop, k))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        kin.integrate_signals(pop.X0())
    w = ms.World(chemistry=ms.Chemistry([ms.Molecule("a", 1.0)], []), map_size=4)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        w.diffuse_molecules()


def test_settle_early_exit_protocol():
    full = [1] * 24
    eq = [4, 4, 4]
    assert _settle_early_exit(full, eq, 3) is False and eq == [4, 4, 4]
    # nothing adjustable anywhere from iteration 2 of part 0 on: later iterations were no-ops
    f = list(full)
    f[2] = f[3] = f[14] = f[15] = 0
    eq = [4, 4, 4]
    assert _settle_early_exit(f, eq, 3) is False and eq == [4, 4, 4]
    # part 1, iteration 1: only low-impact proteins still adjustable -> the reference stops there
    f = list(full)
    f[4 + 1] = 0
    eq = [4, 4, 4]
    assert _settle_early_exit(f, eq, 3) is True and eq == [4, 1, 4]
    # after the repeat the same flags settle
    assert _settle_early_exit(f, eq, 3) is False
    # single-part calls only look at part 0
    f = [0] * 24
    f[12] = 1
    eq = [4, 0, 0]
    assert _settle_early_exit(f, eq, 1) is True and eq == [0, 0, 0]


def test_weights_match_reference_formulas():
    assert diffusion_weights(0.0) == (0.0, 1.0)
    a, b = diffusion_weights(0.5)
    assert abs(a - 0.1) < 1e-12 and abs(8 * a + b - 1.0) < 1e-12
    assert diffusion_weights(-3.0) == diffusion_weights(1.0)
    assert permeation_factor(0.0) == 0.0 and permeation_factor(1.0) == 0.5 and permeation_factor(-7.0) == 0.5


def _kin(m=2):
    return ms.Kinetics(chemistry=ms.Chemistry([ms.Molecule(f"hk{i}", 1.0) for i in range(m)], []))


def test_parameter_store_grow_copy_remove():
    kin = _kin()
    assert kin.N.shape == (0, 0, 4) and kin.N.dtype is torch.int32 and kin.Ke.dtype is torch.float32
    kin.increase_max_cells(by_n=3)
    kin.increase_max_proteins(max_n=2)
    kin.increase_max_proteins(max_n=1)  # never shrinks (kinetics.py:705-723)
    assert kin.N.shape == (3, 2, 4) and kin.Vmax.shape == (3, 2)
    pop = make_population(3, 2, 2, 4, seed=3)
    kin.set_cell_params_from_tensors([0, 1, 2], **{k: getattr(pop, k) for k in _ALL_PARAMS})
    kin.copy_cell_params(from_idxs=[0], to_idxs=[2])
    assert torch.equal(kin.Kmr[2], kin.Kmr[0]) and torch.equal(kin.N[2], pop.N[0])
    kin.unset_cell_params(cell_idxs=[1])
    assert kin.Ke[1].abs().sum() == 0 and kin.A[1].abs().sum() == 0
    kin.remove_cell_params(keep=torch.tensor([True, False, True]))
    assert kin.N.shape == (2, 2, 4) and torch.equal(kin.N[1], pop.N[0])
    # fewer proteins than the store: padded with zeros
    kin.increase_max_cells(by_n=1)
    small = make_population(1, 1, 2, 4, seed=4)
    kin.set_cell_params_from_tensors([2], **{k: getattr(small, k) for k in _ALL_PARAMS})
    assert torch.equal(kin.N[2, 0], small.N[0, 0]) and kin.N[2, 1].abs().sum() == 0


def test_world_bookkeeping_on_cpu():
    mols = [ms.Molecule("wa", 1.0, half_life=10), ms.Molecule("wb", 2.0, diffusivity=1.0, permeability=0.5)]
    w = ms.World(chemistry=ms.Chemistry(mols, []), map_size=8)
    assert w.molecule_map.shape == (2, 8, 8) and w.cell_map.dtype is torch.bool
    assert w.cell_positions.dtype is torch.int32 and w.cell_molecules.shape == (0, 2)
    pop = make_population(5, 3, 2, 8, seed=5)
    before = w.molecule_map.sum()
    idxs = w.add_cells_from_params(pop.cell_positions, **{k: getattr(pop, k) for k in _ALL_PARAMS})
    assert idxs == [0, 1, 2, 3, 4] and w.n_cells == 5 and int(w.cell_map.sum()) == 5
    assert abs(float(before - w.molecule_map.sum() - w.cell_molecules.sum())) < 1e-3
    with pytest.raises(ValueError):
        w.add_cells_from_params(pop.cell_positions[:1], **{k: getattr(pop, k)[:1] for k in _ALL_PARAMS})
    w.kill_cells([1, 3])
    assert w.n_cells == 3 and w.kinetics.N.shape[0] == 3 and len(w.cell_labels) == 3
    assert torch.equal(w.cell_positions, pop.cell_positions[[0, 2, 4]])
    free = torch.nonzero(~w.cell_map)[:1].int()
    assert w.divide_cells_at([0], free) == [(0, 3)]
    assert w.n_cells == 4 and int(w.cell_divisions[0]) == 1 and int(w.cell_divisions[3]) == 1
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="no CPU fallback"):
            w.degrade_molecules()  # a kernel launch like the other step methods
    w.increment_cell_lifetimes()
    assert int(w.cell_lifetimes[1]) == 1
    with pytest.raises(ValueError):
        ms.World(chemistry=ms.Chemistry(mols, []), map_size=4, mol_map_init="nope")


def test_synthetic_population_invariants():
    pop = make_population(64, 16, 6, 16, seed=9)
    assert pop.N.shape == (64, 16, 12) and pop.N.dtype is torch.int32
    assert (pop.Nf >= 0).all() and (pop.Nb >= 0).all()
    assert torch.equal(pop.N, pop.Nb - pop.Nf)
    assert ((pop.Kmr == 1.0) | (pop.A != 0)).all()
    assert (pop.Ke >= 1e-36).all() and (pop.Ke <= 1e36).all()
    assert len({(int(x), int(y)) for x, y in pop.cell_positions.tolist()}) == 64
    again = make_population(64, 16, 6, 16, seed=9)
    assert torch.equal(pop.N, again.N) and torch.equal(pop.Kmr, again.Kmr)


class _NaiveStore:
    """the reference's way (kinetics.py:627-723): torch.cat growth, boolean-mask removal"""

    def __init__(self, s):
        self.t = {k: (torch.zeros(0, 0, s, dtype=torch.int32) if k in ("N", "Nf", "Nb", "A") else
                      torch.zeros(0, 0, s) if k == "Kmr" else torch.zeros(0, 0)) for k in _ALL_PARAMS}

    def grow_c(self, n):
        for k, t in self.t.items():
            self.t[k] = torch.cat([t, torch.zeros(n, *t.shape[1:], dtype=t.dtype)], 0)

    def grow_p(self, p):
        for k, t in self.t.items():
            if p > t.shape[1]:
                shape = list(t.shape)
                shape[1] = p - t.shape[1]
                self.t[k] = torch.cat([t, torch.zeros(shape, dtype=t.dtype)], 1)


def test_parameter_store_matches_reference_semantics_under_random_lifecycle():
    g = torch.Generator().manual_seed(0)
    kin, ref = _kin(), _NaiveStore(4)
    pop = make_population(64, 3, 2, 8, seed=1)
    src = {k: getattr(pop, k) for k in _ALL_PARAMS}
    kin.increase_max_proteins(3)
    ref.grow_p(3)
    n = 0
    for step in range(60):
        op = int(torch.randint(0, 5, (1,), generator=g))
        if op == 0 or n < 4:  # new cells with parameters
            k_new = int(torch.randint(1, 6, (1,), generator=g))
            pick = torch.randint(0, 64, (k_new,), generator=g)
            kin.increase_max_cells(by_n=k_new)
            ref.grow_c(k_new)
            idxs = list(range(n, n + k_new))
            kin.set_cell_params_from_tensors(idxs, **{k: v[pick] for k, v in src.items()})
            for k in _ALL_PARAMS:
                ref.t[k][idxs] = src[k][pick]
            n += k_new
        elif op == 1:  # kill
            keep = torch.rand(n, generator=g) > 0.3
            kin.remove_cell_params(keep=keep)
            for k in _ALL_PARAMS:
                ref.t[k] = ref.t[k][keep]
            n = int(keep.sum())
        elif op == 2:  # divide: children appended, parameters copied
            k_new = int(torch.randint(1, 4, (1,), generator=g))
            parents = torch.randint(0, n, (k_new,), generator=g).tolist()
            kin.increase_max_cells(by_n=k_new)
            ref.grow_c(k_new)
            kids = list(range(n, n + k_new))
            kin.copy_cell_params(from_idxs=parents, to_idxs=kids)
            for k in _ALL_PARAMS:
                ref.t[k][kids] = ref.t[k][parents]
            n += k_new
        elif op == 3:
            idx = torch.randint(0, n, (2,), generator=g).tolist()
            kin.unset_cell_params(cell_idxs=idx)
            for k in _ALL_PARAMS:
                ref.t[k][idx] = 0
        else:  # peek without compacting
            idx = torch.randint(0, n, (3,), generator=g).tolist()
            got = kin.cell_params(idx)
            for k in _ALL_PARAMS:
                assert torch.equal(got[k], ref.t[k][idx])
        if step % 7 == 0 or step == 59:  # reading an attribute compacts in place
            for k in _ALL_PARAMS:
                assert torch.equal(getattr(kin, k), ref.t[k]), (step, k)
            assert kin._rows is None and kin._phys_end == n
    assert kin._n == n


def test_parameter_store_growth_is_amortised_and_removal_moves_nothing():
    kin = _kin()
    pop = make_population(200, 4, 2, 16, seed=2)
    kin.increase_max_cells(by_n=100)
    kin.set_cell_params_from_tensors(list(range(100)), **{k: getattr(pop, k)[:100] for k in _ALL_PARAMS})
    kin.increase_max_cells(by_n=10)  # grows the capacity by half
    ptr = kin._buf["N"].data_ptr()
    assert kin._buf["N"].shape[0] >= 150
    for _ in range(4):
        kin.increase_max_cells(by_n=10)  # fits the spare rows: no reallocation
        assert kin._buf["N"].data_ptr() == ptr
    keep = torch.ones(150, dtype=torch.bool)
    keep[::3] = False
    kin.remove_cell_params(keep=keep)
    assert kin._buf["N"].data_ptr() == ptr and kin._rows is not None and kin._n == 100
    assert torch.equal(kin._rows.long(), torch.nonzero(keep).flatten())
    assert kin.N.shape == (100, 4, 4) and kin._rows is None  # attribute access compacted the rows
    assert torch.equal(kin.N[:66], pop.N[:100][keep[:100]])


def test_parameter_store_notices_outside_writes_and_assignment():
    kin = _kin()
    pop = make_population(8, 3, 2, 4, seed=3)
    kin.increase_max_cells(by_n=8)
    kin.set_cell_params_from_tensors(list(range(8)), **{k: getattr(pop, k) for k in _ALL_PARAMS})
    assert not kin._written_from_outside()
    kin.N[2, 1, 0] = 7  # in place through the attribute view, as users do
    assert kin._written_from_outside() and int(kin._buf["N"][2, 1, 0]) == 7
    kin._stamp()
    kin.Vmax = torch.ones(8, 3)  # assignment adopts the tensor
    assert kin._managed() and not kin._written_from_outside() and float(kin.Vmax.sum()) == 24.0
    kin.Ke = torch.ones(5, 3)  # inconsistent shapes are allowed, as in the reference
    assert not kin._managed() and kin.Ke.shape == (5, 3)
    import pickle
    kin.Ke = torch.ones(8, 3)
    kin.remove_cell_params(keep=torch.tensor([True, False] * 4))
    back = pickle.loads(pickle.dumps(kin))
    assert back.N.shape == (4, 3, 4) and torch.equal(back.N, kin.N) and back._cache is None


def test_slab_height_rules():
    """distributed.py: slabs as even as the rows allow; boundaries that follow a per-row cost"""
    from magicsoup_b200.distributed import balanced_slab_heights, even_slab_heights

    assert even_slab_heights(10, 3) == [4, 3, 3] and even_slab_heights(8, 8) == [1] * 8
    assert even_slab_heights(1024, 8) == [128] * 8
    cost = torch.ones(12)
    assert balanced_slab_heights(cost, 3) == [4, 4, 4]
    cost = torch.tensor([9.0, 1, 1, 1, 1, 1, 1, 1, 1, 1])  # a crowded first row
    h = balanced_slab_heights(cost, 2)
    assert h == [1, 9]
    cost = torch.tensor([0.0, 0, 0, 0, 0, 0, 100.0, 0])  # everything on one row: every slab still gets a row
    h = balanced_slab_heights(cost, 4)
    assert sum(h) == 8 and min(h) >= 1 and len(h) == 4
    assert balanced_slab_heights(torch.zeros(9), 3) == [3, 3, 3]
    g = torch.Generator().manual_seed(0)
    for _ in range(50):
        H, ws = int(torch.randint(4, 200, (1,), generator=g)), int(torch.randint(1, 9, (1,), generator=g))
        if H < ws:
            continue
        cost = torch.rand(H, generator=g) ** 4 * 100
        h = balanced_slab_heights(cost, ws)
        assert sum(h) == H and min(h) >= 1 and len(h) == ws
        loads, at = [], 0
        for x in h:
            loads.append(float(cost[at:at + x].sum()))
            at += x
        # no slab exceeds the mean by more than the heaviest single row
        assert max(loads) <= float(cost.sum()) / ws + float(cost.max()) + 1e-9


def test_world_persistence_round_trip(tmp_path):
    """save / from_file and save_state / load_state (world.py:747-878), incl. the reference's file names"""
    import magicsoup_b200 as ms

    mols = [ms.Molecule(f"ps{i}", 10e3 * (i + 1)) for i in range(3)]
    w = ms.World(chemistry=ms.Chemistry(molecules=mols, reactions=[([mols[0]], [mols[1]])]), map_size=8)
    n, p, s = 5, 2, 6
    pos = torch.tensor([[0, 0], [1, 2], [3, 3], [7, 7], [4, 0]], dtype=torch.int32)
    g = torch.Generator().manual_seed(3)
    params = {k: torch.randint(-2, 3, (n, p, s), generator=g, dtype=torch.int32) for k in ("N", "Nf", "Nb", "A")}
    params.update(Kmr=torch.rand(n, p, s, generator=g), **{k: torch.rand(n, p, generator=g) for k in ("Ke", "Kmf", "Kmb", "Vmax")})
    w.add_cells_from_params(pos, **params)
    w.cell_genomes[:] = ["ACGT" * (i + 1) for i in range(n)]
    w.kill_cells([1])
    w.save(rundir=tmp_path / "run")
    w2 = ms.World.from_file(rundir=tmp_path / "run", device="cpu")
    assert w2.n_cells == 4 and w2.cell_genomes == w.cell_genomes and w2.cell_labels == w.cell_labels
    for key in ("cell_molecules", "cell_map", "molecule_map", "cell_lifetimes", "cell_positions", "cell_divisions"):
        assert torch.equal(getattr(w, key), getattr(w2, key)), key
    for k in ("N", "Nf", "Nb", "A", "Kmr", "Ke", "Kmf", "Kmb", "Vmax"):
        assert torch.equal(getattr(w.kinetics, k), getattr(w2.kinetics, k)), k
    w.save_state(statedir=tmp_path / "run" / "step=0")
    names = sorted(f.name for f in (tmp_path / "run" / "step=0").iterdir())
    assert names == ["cell_divisions.pt", "cell_lifetimes.pt", "cell_map.pt", "cell_molecules.pt", "cell_positions.pt",
                     "cells.fasta", "molecule_map.pt"]
    assert (tmp_path / "run" / "step=0" / "cells.fasta").read_text().startswith(f">0 {w.cell_labels[0]}\nACGT\n>1 ")
    w.cell_molecules += 1.0
    w.cell_genomes[0] = "TTTT"
    w.load_state(statedir=tmp_path / "run" / "step=0", ignore_cell_params=True)
    assert w.cell_genomes == w2.cell_genomes and w.cell_labels == w2.cell_labels and w.n_cells == 4
    assert torch.equal(w.cell_molecules, w2.cell_molecules) and w.cell_positions.dtype == torch.int32
