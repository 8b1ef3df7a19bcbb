"""
Generates the golden fixtures in this directory from the REAL reference
(/root/reference, magicsoup 1.0.5, imported read-only with a stubbed Rust `_lib`).
Runs only in the build container; the fixtures it writes are committed.

    python tests/golden/make_golden.py

Fixtures
  ref_tests_kinetics.npz   every call the reference's own hot-path tests
                           (tests/fast/test_kinetics.py:955-2234) make into
                           Kinetics.integrate_signals and its helpers: parameter
                           tensors, arguments and the reference's return values.
  ref_synthetic_<tag>.npz  seeded synthetic populations (magicsoup_b200.synthetic)
                           pushed through the reference stage by stage and end to
                           end, incl. the reference's own summation-order noise
                           floor (same inputs, proteins permuted).  Tags c1, wl, two
                           (small) and c4, c4e, c3, s82 (the benchmarked tile shapes).
  ref_drift_1000[_c4].npz  population statistics of a 1000-step reference trajectory (and of
                           the reference with permuted proteins) every 100 steps.
  ref_set_cell_params.npz  Kinetics.set_cell_params on random proteomes: token tables, flattened
                           domain tokens, the nine tensors the reference derives from them.
  ref_world.npz            World.diffuse_molecules / enzymatic_activity on small
                           maps incl. the 5x5 known-answer map of
                           tests/fast/test_world.py:8-51.
"""
from __future__ import annotations

import importlib.util
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

from oracle import ref_loader  # noqa: E402
from magicsoup_b200.synthetic import make_population  # noqa: E402

OUT = Path(__file__).resolve().parent
PARAMS = ("N", "Nf", "Nb", "A", "Kmr", "Ke", "Kmf", "Kmb", "Vmax")
HOT_TESTS = (
    "test_simple_mm_kinetic",
    "test_mm_kinetic_with_proportions",
    "test_mm_kinetic_with_multiple_substrates",
    "test_mm_kinetic_with_cofactors",
    "test_mm_kinetic_with_allosteric_action",
    "test_reduce_velocity_to_avoid_negative_concentrations",
    "test_reduce_velocity_in_multiple_proteins",
    "test_multiply_signals",
    "test_get_quotient",
    "test_zeros_dont_stop_reactions",
    "test_equilibrium_is_quickly_reached",
    "test_get_negative_adjusted_nv",
    "test_get_equilibrium_adjusted_x",
)
MAX_RECORDS_PER_TEST_AND_METHOD = 12


def _np(t):
    if isinstance(t, torch.Tensor):
        return t.detach().cpu().numpy().copy()
    return np.asarray(t)


def record_reference_tests(ms) -> dict[str, np.ndarray]:
    Kinetics = ms.kinetics.Kinetics
    spec = importlib.util.spec_from_file_location(
        "ref_test_kinetics", ref_loader.REFERENCE_ROOT / "tests" / "fast" / "test_kinetics.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)

    store: dict[str, np.ndarray] = {}
    state = {"test": "", "counts": {}, "depth": 0}
    methods = {
        "integrate_signals": ("X",),
        "_integrate_signals_part": ("adj_vmax", "X0"),
        "_get_velocities": ("X", "Vmax"),
        "_multiply_signals": ("X", "N"),
        "_get_quotient": ("X",),
        "_get_negative_adjusted_nv": ("NV", "X"),
        "_get_equilibrium_adjusted_x": ("X0", "X1", "NV", "V"),
    }
    originals = {k: getattr(Kinetics, k) for k in methods}

    def wrap(name, argnames):
        orig = originals[name]

        def wrapped(self, *args, **kwargs):
            bound = dict(zip(argnames, args))
            bound.update(kwargs)
            snap_args = {k: _np(v) for k, v in bound.items()}  # before the call (in-place ops)
            snap_par = {k: _np(getattr(self, k)) for k in PARAMS}
            depth = state["depth"]
            state["depth"] += 1
            try:
                out = orig(self, *args, **kwargs)
            finally:
                state["depth"] -= 1
            key = (state["test"], name)
            n = state["counts"].get(key, 0)
            if n < MAX_RECORDS_PER_TEST_AND_METHOD:
                state["counts"][key] = n + 1
                pre = f"{state['test']}|{name}|{n}|"
                for k, v in snap_args.items():
                    store[pre + "arg_" + k] = v
                for k, v in snap_par.items():
                    store[pre + "par_" + k] = v
                outs = out if isinstance(out, tuple) else (out,)
                for i, o in enumerate(outs):
                    store[pre + f"out_{i}"] = _np(o)
                eq_mocked = type(self)._get_equilibrium_adjusted_x is not wrapped_eq[0]
                store[pre + "meta"] = np.array([depth, int(eq_mocked)], dtype=np.int64)
            return out

        return wrapped

    wrapped_eq = [None]
    for name, argnames in methods.items():
        w = wrap(name, argnames)
        if name == "_get_equilibrium_adjusted_x":
            wrapped_eq[0] = w
        setattr(Kinetics, name, w)
    try:
        for t in HOT_TESTS:
            state["test"] = t
            torch.manual_seed(0)
            getattr(mod, t)()  # the reference's own asserts run too: the record is of a PASSING test
            print("recorded", t, {k[1]: v for k, v in state["counts"].items() if k[0] == t})
    finally:
        for name, f in originals.items():
            setattr(Kinetics, name, f)
    return store


def _ref_kinetics(ms, pop, perm=None):
    chem = ms.Chemistry(molecules=[ms.Molecule(f"syn{pop.m}m{i}", 10e3) for i in range(pop.m)], reactions=[])
    kin = ms.kinetics.Kinetics(chemistry=chem)
    for k in PARAMS:
        t = getattr(pop, k)
        setattr(kin, k, (t if perm is None else t[:, perm]).clone())
    return kin


def _mismatch(a: np.ndarray, b: np.ndarray, tol=1e-4) -> float:
    rel = np.abs(a - b) / np.maximum(np.abs(b), 1e-6)
    return float((rel > tol).mean())


def synthetic_fixture(ms, tag, c, p, m, map_size, seed, n_steps, energy_range=(10e3, 50e3)):
    pop = make_population(c, p, m, map_size, seed=seed, energy_range=energy_range)
    kin = _ref_kinetics(ms, pop)
    X0 = pop.X0()
    d: dict[str, np.ndarray] = {k: _np(getattr(pop, k)) for k in PARAMS}
    d["X0"] = _np(X0)
    # stage by stage, part 1 (trim 0.7), teacher forced on the reference's own intermediates
    adj = (kin.Vmax * 0.7).clamp(0.0)
    V = kin._get_velocities(X=X0, Vmax=adj)
    NV = torch.einsum("cps,cp->cps", kin.N, V)
    NV_adj = kin._get_negative_adjusted_nv(NV=NV, X=X0)
    X1 = X0 + NV_adj.sum(1)
    X1[X1 < 0.0] = 0.0
    X1_adj = kin._get_equilibrium_adjusted_x(X0=X0, X1=X1.clone(), NV=NV_adj, V=V)
    d.update(adj_vmax=_np(adj), V=_np(V), NV_adj=_np(NV_adj), X1_pre=_np(X1), X1_adj=_np(X1_adj))
    d["Q_X1"] = _np(kin._get_quotient(X=X1))
    xx, inv = kin._multiply_signals(X=X0, N=kin.Nf)
    d["xx_f"], d["inv_f"] = _np(xx), _np(inv)
    # end to end
    X = X0.clone()
    traj = []
    for _ in range(n_steps):
        X = kin.integrate_signals(X=X)
        traj.append(_np(X))
    d["X_step1"] = traj[0]
    d["X_traj_last"] = traj[-1]
    d["traj_mean"] = np.stack([t.mean(0) for t in traj])
    d["traj_zero_frac"] = np.array([(t == 0).mean() for t in traj])
    # the reference's own noise floor: identical inputs, proteins permuted
    g = torch.Generator().manual_seed(seed + 1)
    perm = torch.randperm(p, generator=g)
    kin2 = _ref_kinetics(ms, pop, perm)
    Xp = X0.clone()
    trajp = []
    for _ in range(n_steps):
        Xp = kin2.integrate_signals(X=Xp)
        trajp.append(_np(Xp))
    d["noise_floor_step1"] = np.array(_mismatch(trajp[0], traj[0]))
    d["noise_floor_last"] = np.array(_mismatch(trajp[-1], traj[-1]))
    d["traj_mean_perm"] = np.stack([t.mean(0) for t in trajp])
    d["n_steps"] = np.array(n_steps)
    d["cell_positions"] = _np(pop.cell_positions)
    d["cell_molecules"] = _np(pop.cell_molecules)
    d["molecule_map"] = _np(pop.molecule_map)
    np.savez_compressed(OUT / f"ref_synthetic_{tag}.npz", **d)
    print(f"synthetic {tag}: c={c} p={p} s={2*m} noise floor step1={d['noise_floor_step1']:.4f} "
          f"last={d['noise_floor_last']:.4f}")


def bench_shape_fixtures(ms):
    """
    The shapes BASELINE.json's configs are benchmarked on (round 2): C4's tile (p=64, s=32) with
    the bench's default energies and with SURVEY.md 8(d)'s U(10,500) kJ, C3 (a saturated map,
    p=35, s=40) and the co2_fixing chemistry's size (s=82).  `python make_golden.py bench` writes
    only these.
    """
    synthetic_fixture(ms, "c4", c=1024, p=64, m=16, map_size=64, seed=31, n_steps=10)
    synthetic_fixture(ms, "c4e", c=512, p=64, m=16, map_size=64, seed=32, n_steps=10, energy_range=(10e3, 500e3))
    synthetic_fixture(ms, "c3", c=1024, p=35, m=20, map_size=32, seed=33, n_steps=10)  # 32 x 32 map, every pixel occupied
    synthetic_fixture(ms, "s82", c=256, p=40, m=41, map_size=16, seed=34, n_steps=10)  # 16 x 16 map, every pixel occupied


def drift_fixture(ms, c=96, p=20, m=6, n_steps=1000, seed=21, name="ref_drift_1000.npz"):
    """
    north_star: "drift over 1000 steps".  Population statistics of the reference's own
    trajectory (and of the same run with its proteins permuted = the reference's own drift)
    every 100 steps; parameters and X0 are regenerated from the seed by the test.
    """
    pop = make_population(c, p, m, 16, seed=seed)
    g = torch.Generator().manual_seed(seed + 1)
    perm = torch.randperm(p, generator=g)
    out = {"shape": np.array([c, p, m, 16, seed, n_steps])}
    for tag, pr in (("ref", None), ("perm", perm)):
        kin = _ref_kinetics(ms, pop, pr)
        X = pop.X0().clone()
        means, zeros, q99, tot = [], [], [], []
        for step in range(1, n_steps + 1):
            X = kin.integrate_signals(X=X)
            if step % 100 == 0:
                x = _np(X)
                means.append(x.mean(0))
                zeros.append((x == 0).mean())
                q99.append(np.quantile(x, 0.99, axis=0))
                tot.append(x.sum(dtype=np.float64))
        out[f"{tag}_mean"] = np.stack(means)
        out[f"{tag}_zero_frac"] = np.array(zeros)
        out[f"{tag}_q99"] = np.stack(q99)
        out[f"{tag}_total"] = np.array(tot)
        out[f"{tag}_final"] = _np(X)
    np.savez_compressed(OUT / name, **out)
    print("drift fixture: total mass ref", out["ref_total"][-1], "perm", out["perm_total"][-1])


def world_fixture(ms):
    d: dict[str, np.ndarray] = {}
    # (1) the 5x5 known-answer map of tests/fast/test_world.py:8-51
    ms.Molecule._instances = {}
    mols = [ms.Molecule("g5m0", 10, diffusivity=0.0), ms.Molecule("g5m1", 10, diffusivity=0.5)]
    world = ms.World(chemistry=ms.Chemistry(molecules=mols, reactions=[]), map_size=5)
    l0 = torch.zeros(5, 5)
    l0[2, 2] = 1.0
    l1 = torch.zeros(5, 5)
    l1[0, 0] = l1[2, 3] = l1[3, 3] = 1.0
    world.molecule_map = torch.stack([l0, l1])
    d["k5_in"] = _np(world.molecule_map)
    world.diffuse_molecules()
    d["k5_out"] = _np(world.molecule_map)
    d["k5_diffusivities"] = np.array([0.0, 0.5])

    # (2) random maps, all diffusivity classes, with cells -> diffusion + permeation
    for tag, size, m, c, seed in (("w24", 24, 5, 150, 3), ("w64", 64, 6, 1000, 4)):
        diffs = [1.0, 0.1, 0.01, 0.0, 0.5, 2.0][:m]
        perms = [1.0, 0.1, 0.0, 0.5, 3.0, 0.01][:m]
        mols = [ms.Molecule(f"{tag}m{i}", 10e3, diffusivity=diffs[i], permeability=perms[i]) for i in range(m)]
        world = ms.World(chemistry=ms.Chemistry(molecules=mols, reactions=[]), map_size=size)
        g = torch.Generator().manual_seed(seed)
        world.molecule_map = (torch.randn(m, size, size, generator=g) + 10.0).abs()
        pix = torch.randperm(size * size, generator=g)[:c]
        world.cell_positions = torch.stack([pix // size, pix % size], 1).int()
        world.cell_molecules = torch.rand(c, m, generator=g) * 20.0
        world.n_cells = c
        d[f"{tag}_map_in"] = _np(world.molecule_map)
        d[f"{tag}_cm_in"] = _np(world.cell_molecules)
        d[f"{tag}_pos"] = _np(world.cell_positions)
        d[f"{tag}_diffusivities"] = np.array(diffs)
        d[f"{tag}_permeabilities"] = np.array(perms)
        world.diffuse_molecules()
        d[f"{tag}_map_out1"] = _np(world.molecule_map)
        d[f"{tag}_cm_out1"] = _np(world.cell_molecules)
        for _ in range(9):
            world.diffuse_molecules()
        d[f"{tag}_map_out10"] = _np(world.molecule_map)
        d[f"{tag}_cm_out10"] = _np(world.cell_molecules)

    # (3) enzymatic_activity through World (gather / integrate / scatter, world.py:610-625)
    pop = make_population(200, 10, 4, 16, seed=7)
    mols = [ms.Molecule(f"eam{i}", 10e3) for i in range(4)]
    world = ms.World(chemistry=ms.Chemistry(molecules=mols, reactions=[]), map_size=16)
    for k in PARAMS:
        setattr(world.kinetics, k, getattr(pop, k).clone())
        d["ea_" + k] = _np(getattr(pop, k))
    world.molecule_map = pop.molecule_map.clone()
    world.cell_positions = pop.cell_positions.clone()
    world.cell_molecules = pop.cell_molecules.clone()
    world.n_cells = pop.c
    d["ea_map_in"], d["ea_cm_in"], d["ea_pos"] = _np(pop.molecule_map), _np(pop.cell_molecules), _np(pop.cell_positions)
    world.enzymatic_activity()
    d["ea_map_out"], d["ea_cm_out"] = _np(world.molecule_map), _np(world.cell_molecules)
    np.savez_compressed(OUT / "ref_world.npz", **d)
    print("world fixture written")


def set_cell_params_fixture(ms):
    """
    Kinetics.set_cell_params of the reference (kinetics.py:521-625) on random proteomes: the
    reference's token tables, the flattened domain tokens and the nine tensors it derives.
    Proteomes come from magicsoup_b200's own genome translation (the reference's Rust extension
    cannot be built here; tests/test_host_lib.py pins that translation to the reference's vectors).
    """
    import random
    from magicsoup_b200 import Genetics, random_genome
    from magicsoup_b200._hostlib import lib as hostlib

    random.seed(17)
    ms.Molecule._instances = {}
    mols = [ms.Molecule(f"scp{i}", e) for i, e in enumerate((10e3, 25e3, 60e3, 150e3, 8e3))]
    reactions = [([mols[0]], [mols[1]]), ([mols[1], mols[2]], [mols[3]]), ([mols[4], mols[4]], [mols[0], mols[2]])]
    chem = ms.Chemistry(molecules=mols, reactions=reactions)
    kin = ms.kinetics.Kinetics(chemistry=chem)
    gen = Genetics(p_catal_dom=0.1, p_transp_dom=0.1, p_reg_dom=0.1)  # many domains per protein
    genomes = [random_genome(random.choice((300, 1000, 2000))) for _ in range(80)]
    proteomes = [p for p in gen.translate_genomes(genomes) if len(p) > 0]
    n, p_max = len(proteomes), max(len(p) for p in proteomes)
    kin.increase_max_cells(by_n=n)
    kin.increase_max_proteins(max_n=p_max)
    kin.set_cell_params(cell_idxs=list(range(n)), proteomes=proteomes)
    prot_off, dom_off, doms = hostlib().flatten_proteomes(proteomes)
    d = {k: _np(getattr(kin, k)) for k in PARAMS}
    d.update(prot_off=prot_off, dom_off=dom_off, doms=doms, abs_temp=np.array(kin.abs_temp),
             mol_energies=_np(kin.mol_energies), km_weights=_np(kin.km_map.weights), vmax_weights=_np(kin.vmax_map.weights),
             signs=_np(kin.sign_map.signs), hill_numbers=_np(kin.hill_map.numbers), reaction_M=_np(kin.reaction_map.M),
             transport_M=_np(kin.transport_map.M), effector_M=_np(kin.effector_map.M))
    np.savez_compressed(OUT / "ref_set_cell_params.npz", **d)
    print(f"set_cell_params fixture: {n} cells, up to {p_max} proteins, {doms.size // 5} domains")


def main():
    ms = ref_loader.load()
    torch.manual_seed(0)
    if sys.argv[1:] == ["bench"]:
        bench_shape_fixtures(ms)
        return
    if sys.argv[1:] == ["drift_c4"]:  # the benchmark's tile shape (p=64, s=32) over 1000 steps
        drift_fixture(ms, c=256, p=64, m=16, n_steps=1000, seed=22, name="ref_drift_1000_c4.npz")
        return
    if sys.argv[1:] == ["params"]:
        set_cell_params_fixture(ms)
        return
    store = record_reference_tests(ms)
    np.savez_compressed(OUT / "ref_tests_kinetics.npz", **store)
    print("ref_tests_kinetics.npz:", len(store), "arrays")
    synthetic_fixture(ms, "c1", c=100, p=22, m=4, map_size=16, seed=11, n_steps=20)
    synthetic_fixture(ms, "wl", c=256, p=35, m=14, map_size=32, seed=12, n_steps=20)
    synthetic_fixture(ms, "two", c=2, p=6, m=3, map_size=4, seed=13, n_steps=30)
    bench_shape_fixtures(ms)
    drift_fixture(ms)
    drift_fixture(ms, c=256, p=64, m=16, n_steps=1000, seed=22, name="ref_drift_1000_c4.npz")
    world_fixture(ms)
    set_cell_params_fixture(ms)


if __name__ == "__main__":
    main()
