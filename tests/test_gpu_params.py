"""
GPU tests of Kinetics.set_cell_params on the device (ms_set_cell_params, reference
python/magicsoup/kinetics.py:521-625) and of the genome-driven World loop built on it
(spawn / divide / mutate / recombinate, reference performance/run_simulation.py:56-100).
"""
import random

import numpy as np
import pytest
import torch

import magicsoup_b200 as ms
from oracle import host_oracle as HO
from oracle import oracle as O
from tests.helpers import PARAMS, bits_equal, describe_mismatch
from tests.test_host_lib import _params_fixture, check_params_against_reference, unflatten

pytestmark = pytest.mark.gpu


def _chem(n_mols=5, tag="gp"):
    mols = [ms.Molecule(f"{tag}{i}", e, diffusivity=d, permeability=q)
            for i, (e, d, q) in enumerate(zip((10e3, 25e3, 60e3, 150e3, 8e3, 40e3)[:n_mols], (1.0, 0.1, 0.01, 0.0, 0.5, 1.0),
                                              (1.0, 0.1, 0.0, 0.5, 0.0, 0.2)))]
    reactions = [([mols[0]], [mols[1]]), ([mols[1], mols[2]], [mols[3]]), ([mols[4], mols[4]], [mols[0], mols[2]])]
    return ms.Chemistry(molecules=mols, reactions=reactions)


def test_set_cell_params_kernel_against_the_reference_fixture():
    z, tables = _params_fixture()
    c, p, s = z["N"].shape
    kin = ms.Kinetics(chemistry=_chem(), abs_temp=float(z["abs_temp"]), device="cuda")
    kin.mol_energies = torch.from_numpy(z["mol_energies"])
    kin.km_map.weights = torch.from_numpy(tables["km_weights"])      # the reference's tests overwrite the maps
    kin.vmax_map.weights = torch.from_numpy(tables["vmax_weights"])  # the same way (test_kinetics.py:103-109)
    kin.sign_map.signs = torch.from_numpy(tables["signs"])
    kin.hill_map.numbers = torch.from_numpy(tables["hill_numbers"])
    kin.reaction_map.M = torch.from_numpy(tables["reaction_M"])
    kin.transport_map.M = torch.from_numpy(tables["transport_M"])
    kin.effector_map.M = torch.from_numpy(tables["effector_M"])
    kin.increase_max_cells(by_n=c + 3)
    kin.increase_max_proteins(max_n=p)
    proteomes = unflatten(z["prot_off"], z["dom_off"], z["doms"])
    order = list(range(c + 2, 2, -1))  # cells c+2 .. 3 get proteomes 0 .. c-1: rows need not be in order
    kin.set_cell_params(cell_idxs=order, proteomes=proteomes)
    got = {k: getattr(kin, k)[order].cpu().numpy() for k in PARAMS}
    check_params_against_reference(got, z)
    mine = HO.set_cell_params(z["prot_off"], z["dom_off"], z["doms"], tables, z["mol_energies"], float(z["abs_temp"]), p, s)
    for k in ("N", "Nf", "Nb", "A"):
        assert np.array_equal(got[k], mine[k])
    for k in ("Vmax", "Kmr", "Ke", "Kmf", "Kmb"):
        assert np.allclose(got[k], mine[k], rtol=2e-5, atol=0, equal_nan=True), k
    assert float(kin.N[:3].abs().sum()) == 0  # untouched cells stay unset
    # setting again after unsetting reproduces the same bits (tests/slow/test_kinetics.py:18-78)
    kin.unset_cell_params(cell_idxs=order)
    kin.set_cell_params(cell_idxs=order, proteomes=proteomes)
    for k in PARAMS:
        again = getattr(kin, k)[order].cpu().numpy()
        assert np.array_equal(again, got[k], equal_nan=k not in ("N", "Nf", "Nb", "A")), k
    with pytest.raises(ValueError):
        kin.set_cell_params(cell_idxs=[0], proteomes=[proteomes[0] * 50])
    bad = [([((1, 1, 1, 1, 9999), 0, 21)], 0, 30, True)]
    with pytest.raises(IndexError):
        kin.set_cell_params(cell_idxs=[0], proteomes=[bad])


def test_genome_driven_world_loop_matches_oracle():
    random.seed(23)
    torch.manual_seed(23)
    chem = _chem(tag="gw")
    w = ms.World(chemistry=chem, map_size=32, device="cuda")
    w.genetics = ms.Genetics(p_catal_dom=0.1, p_transp_dom=0.1, p_reg_dom=0.1)
    idxs = w.spawn_cells(genomes=[ms.random_genome(random.choice((300, 500, 1000))) for _ in range(300)])
    assert idxs == list(range(300)) and w.n_cells == 300 and int(w.cell_map.sum()) == 300
    assert w.kinetics.N.shape[0] == 300 and w.kinetics.N.shape[2] == 10
    n_active = int((w.kinetics.N.abs().sum((1, 2)) > 0).sum())
    assert n_active > 200  # most random genomes encode a working protein
    for step in range(6):
        # kinetics against the CPU oracle on the parameters the store shows
        P = O.Params(*[getattr(w.kinetics, k).cpu().numpy() for k in PARAMS])
        pos = w.cell_positions.cpu().numpy()
        cm, mm = O.enzymatic_activity(P, pos, w.cell_molecules.cpu().numpy(), w.molecule_map.cpu().numpy())
        w.enzymatic_activity()
        assert bits_equal(w.cell_molecules, cm), f"step {step}: " + describe_mismatch(w.cell_molecules, cm)
        assert bits_equal(w.molecule_map, mm), f"step {step}"
        # the reference's loop (performance/run_simulation.py:56-100)
        w.kill_cells(cell_idxs=[i for i in range(w.n_cells) if i % 11 == step])
        pairs = w.divide_cells(cell_idxs=[i for i in range(w.n_cells) if i % 3 == 0])
        assert len(pairs) > 0 and all(w.cell_genomes[a] == w.cell_genomes[b] for a, b in pairs)
        w.recombinate_cells(p=1e-4)
        w.mutate_cells(p=1e-3)
        w.move_cells(cell_idxs=list(range(0, w.n_cells, 5)))
        w.degrade_molecules()
        w.diffuse_molecules()
        w.increment_cell_lifetimes()
        n = w.n_cells
        assert len(w.cell_genomes) == len(w.cell_labels) == n == w.cell_molecules.shape[0] == w.kinetics.N.shape[0]
        assert w.cell_positions.shape == (n, 2) and int(w.cell_map.sum()) == n
        xs, ys = w.cell_positions[:, 0].long(), w.cell_positions[:, 1].long()
        assert bool(w.cell_map[xs, ys].all()) and len({(int(a), int(b)) for a, b in w.cell_positions.tolist()}) == n
        assert torch.isfinite(w.molecule_map).all() and (w.molecule_map >= 0).all() and (w.cell_molecules >= 0).all()
    # parameters follow the genomes: re-deriving them from scratch gives the same tensors
    fresh = ms.Kinetics(chemistry=chem, device="cuda")
    for name in ("km_map", "vmax_map", "sign_map", "hill_map", "reaction_map", "transport_map", "effector_map"):
        setattr(fresh, name, getattr(w.kinetics, name))
    proteomes = w.genetics.translate_genomes(w.cell_genomes)
    fresh.increase_max_cells(by_n=w.n_cells)
    fresh.increase_max_proteins(max_n=int(w.kinetics.N.shape[1]))
    keep = [i for i, p in enumerate(proteomes) if p]
    fresh.set_cell_params(cell_idxs=keep, proteomes=[proteomes[i] for i in keep])
    # (as in the reference, protein slots added later by increase_max_proteins hold zeros where a
    # fresh derivation writes the no-op values Kmr = 0^0 = 1, Ke = 1 ...: compare each cell's own
    # proteins in all nine tensors, the padding only where it matters to the kinetics)
    n_prots = torch.tensor([len(p) for p in proteomes], device="cuda")
    own = torch.arange(fresh.N.shape[1], device="cuda")[None, :] < n_prots[:, None]
    for k in PARAMS:
        a, b = getattr(w.kinetics, k), getattr(fresh, k)
        assert a.shape == b.shape, (k, tuple(a.shape), tuple(b.shape))
        differ = (a != b) & ~(torch.isnan(a) & torch.isnan(b)) if a.is_floating_point() else a != b
        if k in ("Kmr", "Ke", "Kmf", "Kmb"):
            differ &= own[:, :, None] if differ.dim() == 3 else own
        assert not bool(differ.any()), (k, int(differ.sum()), differ.nonzero()[:5].tolist())
    w.reposition_cells(cell_idxs=list(range(10)))
    assert int(w.cell_map.sum()) == w.n_cells
    nb = w.get_neighbors(cell_idxs=list(range(w.n_cells)))
    assert all(a < b for a, b in nb) and len(nb) == len(set(nb))


def test_config0_readme_loop_1000_steps():
    """
    BASELINE.json configs[0], the reference's README example (README.md:49-114): CO2 + NADPH <->
    formiat + NADP, 100 random 500-bp genomes on a 128 x 128 map, 1000 steps of activity / kill /
    replicate / mutate / recombinate / diffuse.  enzymatic_activity and diffuse_molecules are compared
    bit for bit with the CPU oracle in each of the first 100 steps and in every 10th step after.
    Without a penalty on genome size this loop lets genomes (and with them the protein axis) grow
    without bound -- at step 250 the largest genome has 1.4 M bases and 35,000 proteins -- "a technical
    problem" the reference's docs answer with a size-dependent kill (docs/tutorials.md:191-208,
    524-540); that selection is part of the loop here.
    """
    random.seed(31)
    torch.manual_seed(31)
    mols = [ms.Molecule("rd-NADPH", 200e3), ms.Molecule("rd-NADP", 100e3), ms.Molecule("rd-formiat", 20e3),
            ms.Molecule("rd-CO2", 10e3)]
    chem = ms.Chemistry(molecules=mols, reactions=[([mols[3], mols[0]], [mols[2], mols[1]])])
    w = ms.World(chemistry=chem, device="cuda")
    w.spawn_cells(genomes=[ms.random_genome(s=500) for _ in range(100)])
    assert w.n_cells == 100 and tuple(w.molecule_map.shape) == (4, 128, 128)
    diffs, perms = [m.diffusivity for m in mols], [m.permeability for m in mols]

    def sample(p):
        return torch.argwhere(torch.bernoulli(p)).flatten().tolist()

    checked = 0
    for step in range(1000):
        check = step < 100 or step % 10 == 0
        if check:
            P = O.Params(*[getattr(w.kinetics, k).cpu().numpy() for k in PARAMS])
            cm, mm = O.enzymatic_activity(P, w.cell_positions.cpu().numpy(), w.cell_molecules.cpu().numpy(),
                                          w.molecule_map.cpu().numpy())
        w.enzymatic_activity()
        if check:
            assert bits_equal(w.cell_molecules, cm), f"step {step}: " + describe_mismatch(w.cell_molecules, cm)
            assert bits_equal(w.molecule_map, mm), f"step {step}"
        x = w.cell_molecules[:, 2]
        sizes = torch.tensor([float(len(g)) for g in w.cell_genomes], device="cuda")
        w.kill_cells(cell_idxs=sorted(set(sample(0.01 / (0.01 + x)) + sample(sizes ** 7 / (sizes ** 7 + 3000.0 ** 7)))))
        x = w.cell_molecules[:, 2]
        w.divide_cells(cell_idxs=sample(x ** 3 / (x ** 3 + 20.0 ** 3)))
        w.mutate_cells(p=1e-4)
        w.recombinate_cells(p=1e-6)
        if check:
            cm, mm = O.diffuse_molecules(w.cell_molecules.cpu().numpy(), w.molecule_map.cpu().numpy(),
                                         w.cell_positions.cpu().numpy(), diffs, perms)
        w.diffuse_molecules()
        if check:
            assert bits_equal(w.cell_molecules, cm), f"step {step} (diffusion)"
            assert bits_equal(w.molecule_map, mm), f"step {step} (diffusion)"
            checked += 1
        assert len(w.cell_genomes) == w.n_cells == w.cell_molecules.shape[0]
    assert checked == 190 and w.n_cells > 0
    assert torch.isfinite(w.molecule_map).all() and (w.molecule_map >= 0).all() and (w.cell_molecules >= 0).all()
    assert int(w.cell_map.sum()) == w.n_cells


def test_save_state_load_state_rederives_parameters(tmp_path):
    """world.py:795-878: a state holds tensors + genomes; loading it re-translates the genomes"""
    random.seed(41)
    torch.manual_seed(41)
    chem = _chem(tag="gs")
    w = ms.World(chemistry=chem, map_size=24, device="cuda")
    w.genetics = ms.Genetics(p_catal_dom=0.1, p_transp_dom=0.1, p_reg_dom=0.1)
    w.spawn_cells(genomes=[ms.random_genome(random.choice((200, 600))) for _ in range(120)])
    w.enzymatic_activity()
    w.kill_cells(cell_idxs=list(range(0, w.n_cells, 7)))
    w.divide_cells(cell_idxs=list(range(0, w.n_cells, 4)))
    w.save_state(statedir=tmp_path / "s0")
    w.save(rundir=tmp_path)
    want = {k: getattr(w.kinetics, k).clone() for k in PARAMS}
    state = {k: getattr(w, k).clone() for k in ("cell_molecules", "cell_map", "molecule_map", "cell_positions",
                                                  "cell_lifetimes", "cell_divisions")}
    genomes, labels = list(w.cell_genomes), list(w.cell_labels)
    n_prots = torch.tensor([len(p) for p in w.genetics.translate_genomes(genomes)], device="cuda")
    # go on, then come back
    w.enzymatic_activity()
    w.diffuse_molecules()
    w.kill_cells(cell_idxs=list(range(0, w.n_cells, 3)))
    w.load_state(statedir=tmp_path / "s0")
    assert w.cell_genomes == genomes and w.cell_labels == labels and w.n_cells == len(genomes)
    for k, v in state.items():
        assert torch.equal(getattr(w, k), v), k
    own = torch.arange(want["N"].shape[1], device="cuda")[None, :] < n_prots[:, None]
    for k in PARAMS:
        a, b = getattr(w.kinetics, k), want[k]
        assert a.shape == b.shape, k
        differ = (a != b) & ~(torch.isnan(a) & torch.isnan(b)) if a.is_floating_point() else a != b
        if k in ("Kmr", "Ke", "Kmf", "Kmb"):  # padding proteins hold zeros or the no-op values, see above
            differ &= own[:, :, None] if differ.dim() == 3 else own
        assert not bool(differ.any()), k
    w.enzymatic_activity()  # and the loaded world steps
    # the pickled world, on the CPU and back on the GPU
    w2 = ms.World.from_file(rundir=tmp_path, device="cpu")
    assert w2.device == "cpu" and w2.molecule_map.device.type == "cpu" and w2.kinetics.N.device.type == "cpu"
    assert w2.cell_genomes == genomes and torch.equal(w2.kinetics.N, want["N"].cpu())
    w3 = ms.World.from_file(rundir=tmp_path, device="cuda")
    w3.enzymatic_activity()
    assert torch.isfinite(w3.cell_molecules).all()
