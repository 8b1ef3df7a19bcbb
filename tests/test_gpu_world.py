"""
GPU parity tests of World.diffuse_molecules / World.enzymatic_activity (C ABI:
ms_diffuse_molecules, ms_permeate, ms_enzymatic_activity, ms_scatter_signals).
Known answers: /root/reference/tests/fast/test_world.py:8-51 (5x5 map, abs 1e-7);
reference outputs recorded in tests/golden/ref_world.npz; bit parity with the oracle on
seeded maps incl. the shared-memory tiled path; the reference's mass-integrity
properties (tests/slow/test_world.py:7-16, 58-88).
"""
import pickle

import numpy as np
import pytest
import torch

from tests.replay import GOLDEN, mismatch_fraction
from tests.helpers import make_population, oracle_params, bits_equal, describe_mismatch, PARAMS
from oracle import oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ms():
    import magicsoup_b200
    from magicsoup_b200 import _native
    _native.require_cuda()
    return magicsoup_b200


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLDEN / "ref_world.npz")


def _world(ms, m, size, diffs=None, perms=None):
    diffs = diffs if diffs is not None else [0.1] * m
    perms = perms if perms is not None else [0.0] * m
    mols = [ms.Molecule(f"w{i}", 10e3, diffusivity=float(diffs[i]), permeability=float(perms[i])) for i in range(m)]
    return ms.World(chemistry=ms.Chemistry(molecules=mols, reactions=[]), map_size=size, device="cuda")


def _world_from_pop(ms, pop):
    H = pop.molecule_map.shape[1]
    w = _world(ms, pop.m, H, pop.diffusivities, pop.permeabilities)
    w.molecule_map = pop.molecule_map.clone().cuda()
    w.add_cells_from_params(pop.cell_positions, cell_molecules=pop.cell_molecules,
                            **{k: getattr(pop, k) for k in PARAMS})
    return w


def test_diffuse_known_answer_5x5(ms, gold):
    w = _world(ms, 2, 5, diffs=[0.0, 0.5])
    w.molecule_map = torch.from_numpy(gold["k5_in"])  # CPU tensor assigned, as the reference's test does
    w.diffuse_molecules()
    exp1 = torch.tensor([[0.2, 0.1, 0.0, 0.0, 0.1], [0.1, 0.1, 0.1, 0.1, 0.2], [0.0, 0.0, 0.2, 0.3, 0.2],
                         [0.0, 0.0, 0.2, 0.3, 0.2], [0.1, 0.1, 0.1, 0.1, 0.2]])
    out = w.molecule_map.cpu()
    assert out.shape == (2, 5, 5)
    assert ((out[0] - torch.from_numpy(gold["k5_in"][0])).abs() < 1e-7).all()
    assert ((out[1] - exp1).abs() < 1e-7).all()
    assert (np.abs(out.numpy() - gold["k5_out"]) < 1e-7).all()


@pytest.mark.parametrize("tag", ["w24", "w64"])
def test_diffuse_and_permeate_vs_reference(ms, gold, tag):
    diffs, perms = gold[f"{tag}_diffusivities"], gold[f"{tag}_permeabilities"]
    m, size = gold[f"{tag}_map_in"].shape[:2]
    w = _world(ms, m, size, diffs, perms)
    w.molecule_map = torch.from_numpy(gold[f"{tag}_map_in"]).cuda()
    c = gold[f"{tag}_pos"].shape[0]
    w.cell_positions = torch.from_numpy(gold[f"{tag}_pos"]).cuda()
    w.cell_molecules = torch.from_numpy(gold[f"{tag}_cm_in"]).cuda()
    w.n_cells = c
    cm, mm = gold[f"{tag}_cm_in"], gold[f"{tag}_map_in"]
    for step in range(10):
        before = w.cell_molecules
        w.diffuse_molecules()
        assert w.cell_molecules.data_ptr() != before.data_ptr()  # rebinds like world.py:665
        cm, mm = O.diffuse_molecules(cm, mm, gold[f"{tag}_pos"], diffs, perms)
        assert bits_equal(w.molecule_map, mm), describe_mismatch(w.molecule_map, mm, f"map step {step}")
        assert bits_equal(w.cell_molecules, cm), describe_mismatch(w.cell_molecules, cm, f"cm step {step}")
        if step == 0:
            assert np.abs(w.molecule_map.cpu().numpy() - gold[f"{tag}_map_out1"]).max() <= 1e-5
            assert np.abs(w.cell_molecules.cpu().numpy() - gold[f"{tag}_cm_out1"]).max() <= 1e-5
    assert np.abs(w.molecule_map.cpu().numpy() - gold[f"{tag}_map_out10"]).max() <= 5e-5
    assert np.abs(w.cell_molecules.cpu().numpy() - gold[f"{tag}_cm_out10"]).max() <= 5e-5


@pytest.mark.parametrize("shape", [(3, 128, 128), (5, 64, 256), (2, 96, 128), (4, 33, 47), (1, 1, 1), (2, 2, 3)],
                         ids=lambda s: "x".join(map(str, s)))
def test_diffusion_bit_exact_tiled_and_generic(ms, shape):
    m, H, W = shape
    diffs = [1.0, 0.1, 0.01, 0.0, 0.5][:m]
    g = torch.Generator().manual_seed(H * 1000 + W)
    mm = (torch.randn(m, H, W, generator=g) + 10.0).abs()
    mm[0, 0, 0] = 0.0
    from magicsoup_b200 import _native as nat
    from magicsoup_b200.world import diffusion_weights
    a = torch.tensor([diffusion_weights(d)[0] for d in diffs], dtype=torch.float32, device="cuda")
    b = torch.tensor([diffusion_weights(d)[1] for d in diffs], dtype=torch.float32, device="cuda")
    dm = mm.cuda().contiguous()
    n = int(nat.lib().ms_diffuse_scratch_bytes(m, H, W))
    scratch = torch.empty(n, dtype=torch.uint8, device="cuda")
    ref = mm.numpy()
    for step in range(4):
        nat.check(nat.lib().ms_diffuse_molecules(nat.ptr(dm), nat.ptr(a), nat.ptr(b), m, H, W, nat.ptr(scratch), n,
                                                 nat.stream_ptr(dm.device)))
        ref = O.diffuse(ref, diffs)
        assert bits_equal(dm, ref), describe_mismatch(dm, ref, f"step {step}")
    assert (dm >= 0).all()


def test_diffusion_mass_integrity(ms):
    # tests/slow/test_world.py:7-16: 100 steps on 128^2, 14 molecules
    diffs = [1.0, 0.1, 0.01, 0.0, 0.5, 0.1, 0.1] * 2
    w = _world(ms, 14, 128, diffs=diffs)
    exp = w.molecule_map.sum(dim=[1, 2], dtype=torch.float64)
    for step_i in range(100):
        w.diffuse_molecules()
        res = w.molecule_map.sum(dim=[1, 2], dtype=torch.float64)
        assert (res.sum() - exp.sum()).abs() < 10.0, step_i
        assert torch.all(torch.abs(res - exp) < 1.0), step_i
    assert w.molecule_map.dtype is torch.float32 and (w.molecule_map >= 0).all()


def test_enzymatic_activity_vs_reference_fixture(ms, gold):
    m, size = gold["ea_map_in"].shape[:2]
    w = _world(ms, m, size)
    w.molecule_map = torch.from_numpy(gold["ea_map_in"]).cuda()
    w.add_cells_from_params(torch.from_numpy(gold["ea_pos"]), cell_molecules=torch.from_numpy(gold["ea_cm_in"]),
                            **{k: torch.from_numpy(gold["ea_" + k]) for k in PARAMS})
    w.enzymatic_activity()
    assert mismatch_fraction(w.cell_molecules.cpu().numpy(), gold["ea_cm_out"]) <= 0.03
    pos = gold["ea_pos"]
    occ = np.zeros((size, size), dtype=bool)
    occ[pos[:, 0], pos[:, 1]] = True
    out = w.molecule_map.cpu().numpy()
    assert (out[:, ~occ] == gold["ea_map_in"][:, ~occ]).all()
    assert mismatch_fraction(out[:, occ], gold["ea_map_out"][:, occ]) <= 0.03


@pytest.mark.parametrize("shape", [(200, 10, 4, 16, 7), (1500, 35, 14, 64, 8), (4096, 64, 16, 64, 9)],
                         ids=["c200", "c1500", "saturated64"])
def test_full_step_bit_exact(ms, shape):
    c, p, m, size, seed = shape
    pop = make_population(c, p, m, size, seed=seed)
    w = _world_from_pop(ms, pop)
    w._mol_degrads = [0.99 - 0.03 * i for i in range(m)]  # the reference's tests poke this list too (test_world.py:68)
    P = oracle_params(pop)
    cm, mm, pos = pop.cell_molecules.numpy(), pop.molecule_map.numpy(), pop.cell_positions.numpy()
    for step in range(4):
        w.enzymatic_activity()
        cm, mm = O.enzymatic_activity(P, pos, cm, mm)
        assert bits_equal(w.cell_molecules, cm), describe_mismatch(w.cell_molecules, cm, f"ea cm {step}")
        assert bits_equal(w.molecule_map, mm), describe_mismatch(w.molecule_map, mm, f"ea map {step}")
        if step % 2 == 0:
            w.degrade_molecules()  # world.py:667-678: one launch, in place
            w.diffuse_molecules()
        else:  # the same with the degradation folded into the diffusion kernels
            w.diffuse_molecules(degrade_first=True)
        f = np.array(w._mol_degrads, dtype=np.float32)
        cm, mm = cm * f[None, :], mm * f[:, None, None]
        cm, mm = O.diffuse_molecules(cm, mm, pos, pop.diffusivities, pop.permeabilities)
        assert bits_equal(w.cell_molecules, cm), describe_mismatch(w.cell_molecules, cm, f"diff cm {step}")
        assert bits_equal(w.molecule_map, mm), describe_mismatch(w.molecule_map, mm, f"diff map {step}")


def test_full_loop_stability(ms):
    # tests/slow/test_world.py:58-88 shape: 1000 cells, 100 x (degrade, diffuse, activity)
    pop = make_population(1000, 35, 14, 128, seed=61)
    w = _world_from_pop(ms, pop)
    for _ in range(100):
        w.degrade_molecules()
        w.diffuse_molecules()
        w.enzymatic_activity()
        w.increment_cell_lifetimes()
    assert w.molecule_map.min() >= 0.0 and w.cell_molecules.min() >= 0.0
    assert 0.0 < w.molecule_map.mean() < 50.0
    assert w.molecule_map.max() < 500.0 and w.cell_molecules.max() < 500.0
    assert torch.isfinite(w.cell_molecules).all()
    assert w.cell_molecules.dtype is torch.float32 and w.cell_lifetimes.dtype is torch.int32
    assert int(w.cell_lifetimes[0]) == 100


def test_lifecycle_reindexing_keeps_kernels_consistent(ms):
    pop = make_population(300, 12, 4, 32, seed=71)
    w = _world_from_pop(ms, pop)
    kill = [0, 17, 150, 299]
    keep = [i for i in range(300) if i not in kill]
    total = w.molecule_map.sum(dtype=torch.float64) + w.cell_molecules.sum(dtype=torch.float64)
    w.kill_cells(kill)
    assert w.n_cells == 296 and w.kinetics.N.shape[0] == 296
    after = w.molecule_map.sum(dtype=torch.float64) + w.cell_molecules.sum(dtype=torch.float64)
    assert abs(float(total - after)) < 0.1  # spilled onto the map (tests/fast/test_world.py:254)
    assert torch.equal(w.kinetics.N.cpu(), pop.N[keep])
    free = torch.nonzero(~w.cell_map)[:3].int()
    pairs = w.divide_cells_at([1, 2, 3], free)
    assert pairs == [(1, 296), (2, 297), (3, 298)]
    assert torch.equal(w.kinetics.Kmr[296], w.kinetics.Kmr[1])
    assert torch.equal(w.cell_molecules[296], w.cell_molecules[1])
    # the step still matches the oracle on the re-indexed state
    P = O.Params(*[getattr(w.kinetics, k).cpu().numpy() for k in PARAMS])
    cm, mm = O.enzymatic_activity(P, w.cell_positions.cpu().numpy(), w.cell_molecules.cpu().numpy(),
                                  w.molecule_map.cpu().numpy())
    w.enzymatic_activity()
    assert bits_equal(w.cell_molecules, cm) and bits_equal(w.molecule_map, mm)


def test_world_pickles_without_device_handles(ms):
    pop = make_population(50, 8, 4, 16, seed=81)
    w = _world_from_pop(ms, pop)
    w.enzymatic_activity()
    w.diffuse_molecules()
    w2 = pickle.loads(pickle.dumps(w))
    assert w2._scratch == {} and w2.kinetics._scratch == {}
    w.enzymatic_activity()
    w2.enzymatic_activity()
    assert torch.equal(w.cell_molecules, w2.cell_molecules)


def test_position_outside_the_map_raises_like_the_reference(ms):
    # the reference's advanced indexing raises IndexError (world.py:621); here the kernels never touch
    # memory outside the map and the step reports it
    pop = make_population(64, 8, 4, 16, seed=71)
    w = _world_from_pop(ms, pop)
    before = w.molecule_map.clone()
    w.cell_positions[5, 0] = 16
    cm_before = w.cell_molecules.clone()
    w.enzymatic_activity()  # returns before the flags are read ...
    w.diffuse_molecules()
    with pytest.raises(IndexError):  # ... the error surfaces at the next touch of the state (like an asynchronous CUDA error)
        w.molecule_map
    assert torch.equal(w.molecule_map, before) and torch.equal(w.cell_molecules, cm_before)  # nothing was committed
    w.deferred_settle = False
    with pytest.raises(IndexError):  # blocking mode: inside the call, as the reference on the CPU
        w.enzymatic_activity()
    assert torch.equal(w.molecule_map, before)
    w.deferred_settle = True
    w.cell_positions[5, 0] = -3
    w.diffuse_molecules()  # permeation skips the cell, the map stays finite
    assert torch.isfinite(w.molecule_map).all() and torch.isfinite(w.cell_molecules).all()


@pytest.mark.parametrize("size", [24, 128], ids=["generic-stencil", "tiled-stencil"])
def test_non_finite_value_poisons_its_plane_like_torch_sum(ms, size):
    # world.py:640-648: total_before / total_after are float32 torch.sum results, so one NaN or Inf
    # in a plane makes the correction NaN and with it the whole plane; other planes are untouched
    m = 3
    mols = [ms.Molecule(f"nf{size}_{i}", 10e3, diffusivity=[1.0, 0.1, 0.5][i]) for i in range(m)]
    w = ms.World(chemistry=ms.Chemistry(molecules=mols, reactions=[]), map_size=size, device="cuda")
    g = torch.Generator().manual_seed(3)
    mm = (torch.randn(m, size, size, generator=g) + 10.0).abs()
    clean = O.diffuse(mm.numpy(), [1.0, 0.1, 0.5])
    mm[0, 3, 5] = float("nan")
    mm[2, 7, 1] = float("inf")
    w.molecule_map = mm.cuda()
    w.diffuse_molecules()
    out = w.molecule_map.cpu()
    assert torch.isnan(out[0]).all() and torch.isnan(out[2]).all()
    assert bits_equal(out[1], clean[1])


@pytest.mark.parametrize("cells,seed", [(1, 1), (1, 2), (1, 13), (2, 3), (3, 5), (40, 6), (700, 7)])
def test_deferred_early_exit_check_equals_the_blocking_one(ms, cells, seed):
    """
    World.deferred_settle: the rest of the step is enqueued behind a device-side verdict on the
    early-exit flags and the host reads the flags at the next touch of the state.  Same bits as
    blocking on the flags after the launch -- also where the verdict is "repeat" (1-3 cells: the
    reference's global early exit fires, nothing is committed and the step is redone), with
    degradation and both diffusion variants queued in between.
    """
    pop = make_population(cells, 10, 4, 12 if cells < 100 else 32, seed=seed)
    worlds = []
    for deferred in (True, False):
        w = _world_from_pop(ms, pop)
        w.deferred_settle = deferred
        worlds.append(w)
    a, b = worlds
    redone = 0
    for step in range(6):
        for w in (a, b):
            w.enzymatic_activity()
            if step % 3 == 0:
                w.degrade_molecules()
                w.diffuse_molecules()
            elif step % 3 == 1:
                w.diffuse_molecules(degrade_first=True)
            else:
                w.diffuse_molecules()
                w.diffuse_molecules()
            w.increment_cell_lifetimes()
        assert a._pending is not None and b._pending is None
        gate = int(a._pending["gate"][:4].view(torch.int32)[0])
        redone += gate == 0
        assert bits_equal(a.cell_molecules, b.cell_molecules.cpu().numpy()), f"step {step}"
        assert a._pending is None
        assert bits_equal(a.molecule_map, b.molecule_map.cpu().numpy()), f"step {step}"
    if cells == 1:  # (seeds chosen with the oracle: the first launch has to be repeated in part 3, 1 and 2)
        assert redone > 0  # the replay path was exercised
    # parameters touched while a step is pending: the step is finished with the old ones first
    a.enzymatic_activity()
    b.enzymatic_activity()
    for w in (a, b):
        w.kinetics.Vmax = w.kinetics.Vmax * 0.5
    assert a._pending is None
    assert bits_equal(a.cell_molecules, b.cell_molecules.cpu().numpy())
    a.enzymatic_activity()
    b.enzymatic_activity()
    a.kill_cells([0])
    b.kill_cells([0])
    assert bits_equal(a.cell_molecules, b.cell_molecules.cpu().numpy()) and bits_equal(a.molecule_map, b.molecule_map.cpu().numpy())
