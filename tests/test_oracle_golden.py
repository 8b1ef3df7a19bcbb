"""
Pins the CPU oracle (oracle/ms_oracle.c) against the REAL reference:
(1) every call the reference's own hot-path tests make (recorded, see
    tests/golden/make_golden.py) -- /root/reference/tests/fast/test_kinetics.py:955-2234;
(2) seeded synthetic populations pushed through the reference stage by stage
    (teacher forced) and end to end, judged against the reference's own
    summation-order noise floor;
(3) World.diffuse_molecules / enzymatic_activity fixtures incl. the 5x5 known-answer
    map of /root/reference/tests/fast/test_world.py:8-51.
CPU only.
"""
import numpy as np
import pytest

from oracle import oracle as O
from tests.replay import (load_records, full_params, rel_err, mismatch_fraction, GOLDEN, SYNTHETIC_TAGS,
                          velocity_bound)

_TOL = 1e-4  # the reference tests' own absolute tolerance (test_kinetics.py:15)


def _ids(recs):
    return [r.id for r in recs]


_MS = load_records("_multiply_signals")
_Q = load_records("_get_quotient")
_V = load_records("_get_velocities")
_NEG = load_records("_get_negative_adjusted_nv")
_EQ = load_records("_get_equilibrium_adjusted_x")
_PART = load_records("_integrate_signals_part")
_INT = load_records("integrate_signals")


def test_record_counts():
    assert len(_MS) >= 30 and len(_Q) >= 10 and len(_V) >= 15
    assert len(_NEG) >= 15 and len(_EQ) >= 10 and len(_PART) >= 15 and len(_INT) >= 5
    assert {r.test for r in _PART} >= {"test_simple_mm_kinetic", "test_mm_kinetic_with_cofactors"}


@pytest.mark.parametrize("rec", _MS, ids=_ids(_MS))
def test_multiply_signals(rec):
    xx, inv = O.multiply_signals(rec.args["X"], rec.args["N"])
    ref_xx, ref_inv = rec.outs
    assert (inv == ref_inv).all()
    # torch.pow vs repeated multiplication: a few ulp; saturation exact
    assert rel_err(xx, ref_xx, floor=1e-30).max() <= 2e-6
    sat = (ref_xx == np.float32(1e36)) | (ref_xx == 0)
    assert (xx[sat] == ref_xx[sat]).all()


@pytest.mark.parametrize("rec", _Q, ids=_ids(_Q))
def test_get_quotient(rec):
    Q = O.get_quotient(rec.args["X"], rec.params["Nf"], rec.params["Nb"])
    ref = rec.outs[0]
    assert rel_err(Q, ref, floor=1e-37).max() <= 4e-6
    sat = (ref == np.float32(1e36)) | (ref == np.float32(1e-36)) | (ref == 1.0)
    assert (Q[sat] == ref[sat]).all()


@pytest.mark.parametrize("rec", _V, ids=_ids(_V))
def test_get_velocities(rec):
    X = rec.args["X"]
    c, s = X.shape
    p = rec.params["N"].shape[1]
    P = O.Params(**full_params(rec, c, p, s))
    V = O.get_velocities(P, X, rec.args["Vmax"])
    ref = rec.outs[0]
    assert np.abs(V - ref).max() <= 2e-5 * max(1.0, np.abs(ref).max())


@pytest.mark.parametrize("rec", _NEG, ids=_ids(_NEG))
def test_get_negative_adjusted_nv(rec):
    out = O.get_negative_adjusted_nv(rec.args["NV"], rec.args["X"])
    ref = rec.outs[0]
    assert np.abs(out - ref).max() <= 2e-5 * max(1.0, np.abs(ref).max())


@pytest.mark.parametrize("rec", _EQ, ids=_ids(_EQ))
def test_get_equilibrium_adjusted_x(rec):
    out, _ = O.get_equilibrium_adjusted_x(
        rec.args["X0"], rec.args["X1"], rec.args["NV"], rec.args["V"],
        rec.params["Nf"], rec.params["Nb"], rec.params["Ke"])
    assert np.abs(out - rec.outs[0]).max() <= _TOL


@pytest.mark.parametrize("rec", _PART, ids=_ids(_PART))
def test_integrate_signals_part(rec):
    X0 = rec.args["X0"]
    c, s = X0.shape
    p = rec.params["N"].shape[1]
    P = O.Params(**full_params(rec, c, p, s))
    X1, _, _ = O.integrate_signals_part(P, rec.args["adj_vmax"], X0,
                                        max_eq_iters=0 if rec.eq_mocked else 4)
    assert np.abs(X1 - rec.outs[0]).max() <= _TOL


@pytest.mark.parametrize("rec", _INT, ids=_ids(_INT))
def test_integrate_signals(rec):
    X = rec.args["X"]
    c, s = X.shape
    p = rec.params["N"].shape[1]
    P = O.Params(**full_params(rec, c, p, s))
    out = O.integrate_signals(P, X)
    assert np.abs(out - rec.outs[0]).max() <= _TOL


# ---------------------------------------------------------------- synthetic populations

@pytest.fixture(scope="module", params=list(SYNTHETIC_TAGS))
def syn(request):
    z = np.load(GOLDEN / f"ref_synthetic_{request.param}.npz")
    P = O.Params(*[z[k] for k in ("N", "Nf", "Nb", "A", "Kmr", "Ke", "Kmf", "Kmb", "Vmax")])
    return request.param, z, P


def test_stage_velocities(syn):
    _, z, P = syn
    V = O.get_velocities(P, z["X0"], z["adj_vmax"])
    ref = z["V"]
    # P1: rel <= 2e-5 (f32-vs-f64 noise of the reference itself is 1.4e-5) + 4 ulp of Vmax'
    assert (np.abs(V - ref) <= velocity_bound(ref, z["adj_vmax"])).all()
    xx, inv = O.multiply_signals(z["X0"], z["Nf"])
    assert (inv == z["inv_f"]).all()
    assert rel_err(xx, z["xx_f"], floor=1e-30).max() <= 4e-6


def test_stage_negative_adjust_and_x1(syn):
    _, z, P = syn
    NV = z["N"].astype(np.float32) * z["V"][:, :, None]  # teacher forced on the reference's V
    NV_adj = O.get_negative_adjusted_nv(NV, z["X0"])
    ref = z["NV_adj"]
    assert (np.abs(NV_adj - ref) <= 2e-5 * np.maximum(np.abs(ref), 1e-3)).all()
    X1 = np.maximum(z["X0"] + NV_adj.sum(1), 0)
    assert (np.abs(X1 - z["X1_pre"]) <= 1e-5 * np.maximum(1.0, np.abs(z["X0"]))).all()


def test_stage_quotient(syn):
    _, z, P = syn
    Q = O.get_quotient(z["X1_pre"], z["Nf"], z["Nb"])
    ref = z["Q_X1"]
    assert rel_err(Q, ref, floor=1e-37).max() <= 1e-5
    sat = (ref >= np.float32(1e36)) | (ref <= np.float32(1e-36))
    assert (Q[sat] == ref[sat]).all()


def test_stage_equilibrium_adjust_teacher_forced(syn):
    tag, z, P = syn
    out, done = O.get_equilibrium_adjusted_x(z["X0"], z["X1_pre"], z["NV_adj"], z["V"],
                                             z["Nf"], z["Nb"], z["Ke"])
    ref = z["X1_adj"]
    bad = np.abs(out - ref) > 1e-5 * np.maximum(1.0, np.abs(z["X0"]))
    # identical inputs: decisions may differ only where Q/Ke sits on a threshold
    assert bad.any(axis=1).mean() <= 0.02, bad.any(axis=1).mean()


def test_end_to_end_within_reference_noise_floor(syn):
    tag, z, P = syn
    out = O.integrate_signals(P, z["X0"])
    ref = z["X_step1"]
    assert np.isfinite(out).all() and (out >= 0).all()
    mm = mismatch_fraction(out, ref)
    floor = float(z["noise_floor_step1"])
    # P2: not worse than the reference against itself with its proteins permuted
    assert mm <= max(floor * 1.25, 0.01) or tag == "two", (mm, floor)
    if tag == "two":  # 2 cells: the global early exit fires; 12 numbers only
        assert mm <= max(floor, 0.5)


def test_trajectory_statistics(syn):
    tag, z, P = syn
    n = int(z["n_steps"])
    X = z["X0"]
    means = []
    for _ in range(n):
        X = O.integrate_signals(P, X)
        means.append(X.mean(0))
    means = np.stack(means)
    assert np.isfinite(X).all() and (X >= 0).all()
    if tag == "two":
        return
    # P3: per-signal population means drift no more than 2x the reference's own
    # permuted-order drift (or 1e-2 relative)
    ref = z["traj_mean"]
    own = np.abs(z["traj_mean_perm"] - ref)
    bound = np.maximum(2 * own.max(), 1e-2 * np.abs(ref).max())
    assert np.abs(means - ref).max() <= bound


# ---------------------------------------------------------------- world fixtures

@pytest.fixture(scope="module")
def world():
    return np.load(GOLDEN / "ref_world.npz")


def test_diffuse_known_answer_5x5(world):
    # /root/reference/tests/fast/test_world.py:8-51 : abs 1e-7 against exact expectations
    out = O.diffuse(world["k5_in"], world["k5_diffusivities"])
    exp1 = np.array([[0.2, 0.1, 0.0, 0.0, 0.1], [0.1, 0.1, 0.1, 0.1, 0.2], [0.0, 0.0, 0.2, 0.3, 0.2],
                     [0.0, 0.0, 0.2, 0.3, 0.2], [0.1, 0.1, 0.1, 0.1, 0.2]], dtype=np.float32)
    assert (np.abs(out[0] - world["k5_in"][0]) < 1e-7).all()
    assert (np.abs(out[1] - exp1) < 1e-7).all()
    assert (np.abs(out - world["k5_out"]) < 1e-7).all()


@pytest.mark.parametrize("tag", ["w24", "w64"])
def test_diffuse_and_permeate_vs_reference(world, tag):
    cm, mm = world[f"{tag}_cm_in"], world[f"{tag}_map_in"]
    pos = world[f"{tag}_pos"]
    diffs, perms = world[f"{tag}_diffusivities"], world[f"{tag}_permeabilities"]
    cm1, mm1 = O.diffuse_molecules(cm, mm, pos, diffs, perms)
    # P4: <= 1e-5 abs on |N(10,1)| maps for one step
    assert np.abs(mm1 - world[f"{tag}_map_out1"]).max() <= 1e-5
    assert np.abs(cm1 - world[f"{tag}_cm_out1"]).max() <= 1e-5
    for _ in range(9):
        cm1, mm1 = O.diffuse_molecules(cm1, mm1, pos, diffs, perms)
    assert np.abs(mm1 - world[f"{tag}_map_out10"]).max() <= 5e-5
    assert np.abs(cm1 - world[f"{tag}_cm_out10"]).max() <= 5e-5
    # mass conservation at least as good as the reference's (tests/slow/test_world.py:7-16)
    m0 = mm.sum(dtype=np.float64) + cm.sum(dtype=np.float64)
    m1 = mm1.sum(dtype=np.float64) + cm1.sum(dtype=np.float64)
    mr = world[f"{tag}_map_out10"].sum(dtype=np.float64) + world[f"{tag}_cm_out10"].sum(dtype=np.float64)
    # (reference's own bound: < 1.0 per molecule over 100 steps; its drift here is ~1e-2)
    assert abs(m1 - m0) <= max(3 * abs(mr - m0), 0.1)


def test_enzymatic_activity_vs_reference(world):
    P = O.Params(*[world["ea_" + k] for k in ("N", "Nf", "Nb", "A", "Kmr", "Ke", "Kmf", "Kmb", "Vmax")])
    cm, mm = O.enzymatic_activity(P, world["ea_pos"], world["ea_cm_in"], world["ea_map_in"])
    ref_cm, ref_mm = world["ea_cm_out"], world["ea_map_out"]
    assert mismatch_fraction(cm, ref_cm) <= 0.03
    # pixels without a cell are untouched
    occupied = np.zeros(mm.shape[1:], dtype=bool)
    occupied[world["ea_pos"][:, 0], world["ea_pos"][:, 1]] = True
    assert (mm[:, ~occupied] == world["ea_map_in"][:, ~occupied]).all()
    assert mismatch_fraction(mm[:, occupied], ref_mm[:, occupied]) <= 0.03


# ---------------------------------------------------------------- drift over 1000 steps

DRIFT_FIXTURES = ["ref_drift_1000.npz", "ref_drift_1000_c4.npz"]  # 96 x 20 x 12, and the benchmark's tile shape 256 x 64 x 32


def drift_population(name: str = "ref_drift_1000.npz"):
    from magicsoup_b200.synthetic import make_population

    z = np.load(GOLDEN / name)
    c, p, m, size, seed, n_steps = [int(v) for v in z["shape"]]
    pop = make_population(c, p, m, size, seed=seed)
    return z, pop, n_steps


def check_drift(z, stats):
    """
    north_star: "drift over 1000 steps".  Element-wise agreement with the reference is not
    attainable after many steps (the reference drifts from ITSELF when its
    proteins are merely permuted), so the bound is on population statistics every 100
    steps: within max(1e-2 relative, 2x the reference's own permuted-order drift).
    """
    for key, scale_key in (("mean", "ref_mean"), ("q99", "ref_q99")):
        ref, perm, got = z[f"ref_{key}"], z[f"perm_{key}"], stats[key]
        own = np.abs(perm - ref).max()
        bound = max(2 * own, 1e-2 * np.abs(z[scale_key]).max())
        assert np.abs(got - ref).max() <= bound, (key, np.abs(got - ref).max(), bound)
    own = np.abs(z["perm_total"] - z["ref_total"]).max()
    assert np.abs(stats["total"] - z["ref_total"]).max() <= max(2 * own, 1e-3 * z["ref_total"].max())
    assert np.abs(stats["zero_frac"] - z["ref_zero_frac"]).max() <= max(
        2 * np.abs(z["perm_zero_frac"] - z["ref_zero_frac"]).max(), 0.02)


@pytest.mark.parametrize("name", DRIFT_FIXTURES)
def test_drift_over_1000_steps(name):
    z, pop, n_steps = drift_population(name)
    P = O.Params(*[t.numpy() for t in pop.params()])
    X = pop.X0().numpy()
    stats = {"mean": [], "q99": [], "total": [], "zero_frac": []}
    for step in range(1, n_steps + 1):
        X = O.integrate_signals(P, X)
        if step % 100 == 0:
            stats["mean"].append(X.mean(0))
            stats["q99"].append(np.quantile(X, 0.99, axis=0))
            stats["total"].append(X.sum(dtype=np.float64))
            stats["zero_frac"].append((X == 0).mean())
    assert np.isfinite(X).all() and (X >= 0).all()
    check_drift(z, {k: np.array(v) for k, v in stats.items()})
