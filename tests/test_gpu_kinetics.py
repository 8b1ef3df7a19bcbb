"""
GPU parity tests of the kinetics kernels, through the C ABI (magicsoup_b200.Kinetics is a
thin ctypes caller).  Three layers:
  (1) replay of every call the reference's own hot-path tests make, against the
      reference's recorded results (tests/golden/ref_tests_kinetics.npz);
  (2) bit-exact parity with the CPU oracle (oracle/ms_oracle.c, pinned to the reference
      by tests/test_oracle_golden.py) on seeded populations of many shapes, incl. the
      sparse / dense paths, the global early exit and ragged / empty inputs;
  (3) the reference's robustness properties (tests/slow/test_kinetics.py:81-178).
"""
import numpy as np
import pytest
import torch

from tests.replay import (load_records, full_params, rel_err, mismatch_fraction, GOLDEN, SYNTHETIC_TAGS,
                          velocity_bound)
from tests.helpers import (make_population, oracle_params, install, bits_equal, describe_mismatch,
                           dense_adversarial, PARAMS)
from oracle import oracle as O

pytestmark = pytest.mark.gpu
_TOL = 1e-4


@pytest.fixture(scope="module")
def ms():
    import magicsoup_b200
    from magicsoup_b200 import _native
    _native.require_cuda()
    return magicsoup_b200


class _Chem:
    def __init__(self, m):
        from magicsoup_b200.world import Molecule
        self.molecules = [Molecule(f"t{i}", 10e3) for i in range(m)]
        self.reactions = []


def _kinetics(ms, s, cls=None):
    return (cls or ms.Kinetics)(chemistry=_Chem(s // 2), device="cuda")


def _mocked_cls(ms):
    # the reference tests' _MockedKinetics (tests/fast/test_kinetics.py:87-96)
    class Mocked(ms.Kinetics):
        def integrate_signals(self, X):
            return self._integrate_signals_part(adj_vmax=self.Vmax, X0=X)

        def _get_equilibrium_adjusted_x(self, *_, **kwargs):
            return kwargs["X1"]

    return Mocked


def _set(kin, params):
    for k, v in params.items():
        setattr(kin, k, torch.from_numpy(np.ascontiguousarray(v)))  # CPU tensors, like the reference's tests


# ------------------------------------------------------------------ (1) recorded reference calls
_MS = load_records("_multiply_signals")
_Q = load_records("_get_quotient")
_V = load_records("_get_velocities")
_NEG = load_records("_get_negative_adjusted_nv")
_EQ = load_records("_get_equilibrium_adjusted_x")
_PART = load_records("_integrate_signals_part")
_INT = load_records("integrate_signals")


def _ids(recs):
    return [r.id for r in recs]


@pytest.mark.parametrize("rec", _MS, ids=_ids(_MS))
def test_ref_multiply_signals(ms, rec):
    kin = _kinetics(ms, rec.args["X"].shape[1])
    xx, inv = kin._multiply_signals(X=torch.from_numpy(rec.args["X"]), N=torch.from_numpy(rec.args["N"]))
    assert inv.dtype is torch.bool and xx.dtype is torch.float32
    assert (inv.cpu().numpy() == rec.outs[1]).all()
    xx = xx.cpu().numpy()
    assert rel_err(xx, rec.outs[0], floor=1e-30).max() <= 2e-6
    sat = (rec.outs[0] == np.float32(1e36)) | (rec.outs[0] == 0)
    assert (xx[sat] == rec.outs[0][sat]).all()
    oxx, oinv = O.multiply_signals(rec.args["X"], rec.args["N"])
    assert bits_equal(xx, oxx), describe_mismatch(xx, oxx)


@pytest.mark.parametrize("rec", _Q, ids=_ids(_Q))
def test_ref_get_quotient(ms, rec):
    kin = _kinetics(ms, rec.args["X"].shape[1])
    _set(kin, {"Nf": rec.params["Nf"], "Nb": rec.params["Nb"]})
    Q = kin._get_quotient(X=torch.from_numpy(rec.args["X"])).cpu().numpy()
    assert rel_err(Q, rec.outs[0], floor=1e-37).max() <= 4e-6
    assert bits_equal(Q, O.get_quotient(rec.args["X"], rec.params["Nf"], rec.params["Nb"]))


@pytest.mark.parametrize("rec", _V, ids=_ids(_V))
def test_ref_get_velocities(ms, rec):
    X = rec.args["X"]
    c, s = X.shape
    p = rec.params["N"].shape[1]
    par = full_params(rec, c, p, s)
    kin = _kinetics(ms, s)
    _set(kin, par)
    V = kin._get_velocities(X=torch.from_numpy(X), Vmax=torch.from_numpy(rec.args["Vmax"])).cpu().numpy()
    assert np.abs(V - rec.outs[0]).max() <= 2e-5 * max(1.0, np.abs(rec.outs[0]).max())
    assert bits_equal(V, O.get_velocities(O.Params(**par), X, rec.args["Vmax"]))


@pytest.mark.parametrize("rec", _NEG, ids=_ids(_NEG))
def test_ref_get_negative_adjusted_nv(ms, rec):
    kin = _kinetics(ms, rec.args["X"].shape[1])
    out = kin._get_negative_adjusted_nv(NV=torch.from_numpy(rec.args["NV"]), X=torch.from_numpy(rec.args["X"]))
    out = out.cpu().numpy()
    assert np.abs(out - rec.outs[0]).max() <= 2e-5 * max(1.0, np.abs(rec.outs[0]).max())
    assert bits_equal(out, O.get_negative_adjusted_nv(rec.args["NV"], rec.args["X"]))


@pytest.mark.parametrize("rec", _EQ, ids=_ids(_EQ))
def test_ref_get_equilibrium_adjusted_x(ms, rec):
    kin = _kinetics(ms, rec.args["X0"].shape[1])
    _set(kin, {k: rec.params[k] for k in ("Nf", "Nb", "Ke")})
    out = kin._get_equilibrium_adjusted_x(**{k: torch.from_numpy(v) for k, v in rec.args.items()}).cpu().numpy()
    assert np.abs(out - rec.outs[0]).max() <= _TOL
    ora, _ = O.get_equilibrium_adjusted_x(rec.args["X0"], rec.args["X1"], rec.args["NV"], rec.args["V"],
                                          rec.params["Nf"], rec.params["Nb"], rec.params["Ke"])
    assert bits_equal(out, ora), describe_mismatch(out, ora)


@pytest.mark.parametrize("rec", _PART, ids=_ids(_PART))
def test_ref_integrate_signals_part(ms, rec):
    X0 = rec.args["X0"]
    c, s = X0.shape
    p = rec.params["N"].shape[1]
    par = full_params(rec, c, p, s)
    kin = _kinetics(ms, s, cls=_mocked_cls(ms) if rec.eq_mocked else None)
    _set(kin, par)
    X1 = kin._integrate_signals_part(adj_vmax=torch.from_numpy(rec.args["adj_vmax"]), X0=torch.from_numpy(X0))
    X1 = X1.cpu().numpy()
    assert np.abs(X1 - rec.outs[0]).max() <= _TOL
    ora, _, _ = O.integrate_signals_part(O.Params(**par), rec.args["adj_vmax"], X0,
                                         max_eq_iters=0 if rec.eq_mocked else 4)
    assert bits_equal(X1, ora), describe_mismatch(X1, ora)


@pytest.mark.parametrize("rec", _INT, ids=_ids(_INT))
def test_ref_integrate_signals(ms, rec):
    X = rec.args["X"]
    c, s = X.shape
    p = rec.params["N"].shape[1]
    par = full_params(rec, c, p, s)
    kin = _kinetics(ms, s)
    _set(kin, par)
    Xin = torch.from_numpy(X.copy())
    out = kin.integrate_signals(X=Xin)
    assert out.data_ptr() != Xin.data_ptr() and out.dtype is torch.float32
    out = out.cpu().numpy()
    assert np.abs(out - rec.outs[0]).max() <= _TOL
    ora = O.integrate_signals(O.Params(**par), X)
    assert bits_equal(out, ora), describe_mismatch(out, ora)


# ------------------------------------------------------------------ (2) bit parity with the oracle
_SHAPES = [
    # (c, p, m, map, seed)         path exercised
    (100, 22, 4, 16, 11),      # C1-like, s=8 (lanes beyond s idle)
    (256, 35, 14, 32, 12),     # C2/WL-like, s=28
    (300, 35, 20, 32, 13),     # C3-like, s=40: two signals per lane
    (64, 40, 41, 16, 14),      # co2_fixing-sized chemistry, s=82: four signals per lane
    (2048, 64, 16, 64, 15),    # C4 tile shape, s=32
    (50, 7, 3, 8, 16),         # ragged: s=6, p=7
    (40, 300, 4, 8, 17),       # p > 256: dense path for every cell
    (30, 12, 70, 8, 18),       # s=140 > 128: dense path for every cell
    (3, 40000, 4, 4, 20),      # a protein axis grown by runaway genomes (p > 32768), s=8
    (1, 5, 2, 4, 19),          # a single cell: the global early exit fires
    (2, 6, 3, 4, 13),          # two cells (the reference's unit-test regime)
]


@pytest.mark.parametrize("shape", _SHAPES, ids=[f"c{c}p{p}s{2*m}" for c, p, m, _, _ in _SHAPES])
def test_integrate_signals_bit_exact(ms, shape):
    c, p, m, size, seed = shape
    pop = make_population(c, p, m, size, seed=seed)
    kin = install(_kinetics(ms, 2 * m), pop)
    P = oracle_params(pop)
    X = pop.X0()
    Xn = X.numpy()
    for step in range(3):
        out = kin.integrate_signals(X=X.cuda())
        ora, done = O.integrate_signals(P, Xn, return_eq_done=True)
        assert bits_equal(out, ora), f"step {step} eq_done {done}: " + describe_mismatch(out, ora)
        X, Xn = out.cpu(), ora


_PATH_SWITCHES = [
    {},                                                        # default: lane path, staged scan
    {"MS_KIN_NO_STAGED_SCAN": "1"},                            # lane path, register scan
    {"MS_KIN_NO_LANES": "1"},                                  # batch mode, staged scan two steps ahead
    {"MS_KIN_NO_LANES": "1", "MS_KIN_STAGED_DEPTH1": "1"},     # batch mode, staged scan one step ahead
    {"MS_KIN_NO_LANES": "1", "MS_KIN_NO_STAGED_SCAN": "1"},    # batch mode, register scan
    {"MS_KIN_NO_LANES": "1", "MS_KIN_NO_BATCH": "1"},          # warp-per-cell path on variable-length lists
]


@pytest.mark.parametrize("shape", [_SHAPES[0], _SHAPES[1], _SHAPES[4]],
                         ids=[f"c{c}p{p}s{2*m}" for c, p, m, _, _ in [_SHAPES[0], _SHAPES[1], _SHAPES[4]]])
def test_kernel_paths_agree_bit_for_bit(ms, shape, monkeypatch):
    """every scan / integration path of the fused kernel gives the oracle's bits (the switches are
    read once per process; ms_kinetics_reload_switches re-reads them)"""
    from magicsoup_b200 import _native
    c, p, m, size, seed = shape
    pop = make_population(c, p, m, size, seed=seed)
    kin = install(_kinetics(ms, 2 * m), pop)
    X = pop.X0()
    ora = O.integrate_signals(oracle_params(pop), X.numpy())
    for env in _PATH_SWITCHES:
        for k in ("MS_KIN_STAGED_DEPTH1", "MS_KIN_NO_STAGED_SCAN", "MS_KIN_NO_BATCH", "MS_KIN_NO_LANES"):
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        _native.lib().ms_kinetics_reload_switches()
        try:
            out = kin.integrate_signals(X=X.cuda())
        finally:
            for k in env:
                monkeypatch.delenv(k, raising=False)
            _native.lib().ms_kinetics_reload_switches()
        assert bits_equal(out, ora), f"{env}: " + describe_mismatch(out, ora)


@pytest.mark.parametrize("batch", ["32", "96", "512"])
@pytest.mark.parametrize("shape", [_SHAPES[0], _SHAPES[1], _SHAPES[2], _SHAPES[4]],
                         ids=[f"c{c}p{p}s{2*m}" for c, p, m, _, _ in [_SHAPES[0], _SHAPES[1], _SHAPES[2], _SHAPES[4]]])
def test_phased_lane_kernel_gives_the_same_bits(ms, shape, batch, monkeypatch):
    """MS_KIN_PHASED: the lane path with the list cache off as one CTA per SM that alternates between
    a build phase and an integration phase (a measured variant); rounds smaller and
    larger than the population, through the row map as well"""
    from magicsoup_b200 import _native
    c, p, m, size, seed = shape
    pop = make_population(c, p, m, size, seed=seed)
    kin = install(_kinetics(ms, 2 * m), pop)
    kin.list_cache = False
    P = oracle_params(pop)
    X = pop.X0()
    monkeypatch.setenv("MS_KIN_PHASED", "1")
    monkeypatch.setenv("MS_KIN_PHASE_BATCH", batch)
    _native.lib().ms_kinetics_reload_switches()
    try:
        out = kin.integrate_signals(X=X.cuda())
        ora = O.integrate_signals(P, X.numpy())
        assert bits_equal(out, ora), describe_mismatch(out, ora)
        keep = torch.ones(c, dtype=torch.bool)
        keep[::3] = False
        kin.remove_cell_params(keep=keep.cuda())  # parameter rows behind a row map
        out = kin.integrate_signals(X=X[keep].cuda())
        sub = O.Params(*[t[keep.numpy()] for t in (P.N, P.Nf, P.Nb, P.A, P.Kmr, P.Ke, P.Kmf, P.Kmb, P.Vmax)])
        ora = O.integrate_signals(sub, X[keep].numpy())
        assert bits_equal(out, ora), describe_mismatch(out, ora)
    finally:
        monkeypatch.delenv("MS_KIN_PHASED", raising=False)
        monkeypatch.delenv("MS_KIN_PHASE_BATCH", raising=False)
        _native.lib().ms_kinetics_reload_switches()


@pytest.mark.parametrize("shape", _SHAPES[:6], ids=[f"c{c}p{p}s{2*m}" for c, p, m, _, _ in _SHAPES[:6]])
def test_stages_bit_exact(ms, shape):
    c, p, m, size, seed = shape
    pop = make_population(c, p, m, size, seed=seed)
    kin = install(_kinetics(ms, 2 * m), pop)
    P = oracle_params(pop)
    X0 = pop.X0()
    adj = (pop.Vmax * 0.7).clamp(0.0)
    V = kin._get_velocities(X=X0, Vmax=adj)
    oV = O.get_velocities(P, X0.numpy(), adj.numpy())
    assert bits_equal(V, oV), describe_mismatch(V, oV, "V")
    NV = pop.N.float() * torch.from_numpy(oV)[:, :, None]
    NVa = kin._get_negative_adjusted_nv(NV=NV, X=X0)
    oNVa = O.get_negative_adjusted_nv(NV.numpy(), X0.numpy())
    assert bits_equal(NVa, oNVa), describe_mismatch(NVa, oNVa, "NV_adj")
    X1, _, _ = O.integrate_signals_part(P, adj.numpy(), X0.numpy(), max_eq_iters=0)
    Q = kin._get_quotient(X=torch.from_numpy(X1))
    assert bits_equal(Q, O.get_quotient(X1, P.Nf, P.Nb))
    X2 = kin._get_equilibrium_adjusted_x(X0=X0, X1=torch.from_numpy(X1), NV=torch.from_numpy(oNVa),
                                         V=torch.from_numpy(oV))
    oX2, _ = O.get_equilibrium_adjusted_x(X0.numpy(), X1, oNVa, oV, P.Nf, P.Nb, P.Ke)
    assert bits_equal(X2, oX2), describe_mismatch(X2, oX2, "X1_adj")
    # fused single part == staged composition == oracle
    Xp = kin._integrate_signals_part(adj_vmax=adj, X0=X0)
    oXp, _, _ = O.integrate_signals_part(P, adj.numpy(), X0.numpy())
    assert bits_equal(Xp, oXp), describe_mismatch(Xp, oXp, "part")


def test_dense_path_matches_oracle(ms):
    # every tile overflows the sparse lists (dense random N, A), incl. exponents up to 8
    pop = dense_adversarial(60, 40, 28, seed=5)
    kin = install(_kinetics(ms, 28), pop)
    P = oracle_params(pop)
    X = torch.cat([pop.cell_molecules, pop.cell_molecules.flip(1)], 1).contiguous()
    Xn = X.numpy()
    for _ in range(3):
        out = kin.integrate_signals(X=X)
        ora = O.integrate_signals(P, Xn)
        assert bits_equal(out, ora), describe_mismatch(out, ora)
        X, Xn = out.cpu(), ora


def test_mixed_sparse_and_dense_cells(ms):
    pop = make_population(128, 32, 8, 16, seed=21)
    dense = dense_adversarial(128, 32, 16, seed=22)
    pick = torch.arange(128) % 5 == 0  # every 5th cell gets a dense tile / large coefficients
    for k in PARAMS:
        getattr(pop, k)[pick] = getattr(dense, k)[pick]
    pop.Nf[7, 3, 2] = 300  # beyond the 8-bit packing: dense path for that cell
    pop.N[7, 3, 2] = -300
    kin = install(_kinetics(ms, 16), pop)
    X = pop.X0()
    out = kin.integrate_signals(X=X)
    ora = O.integrate_signals(oracle_params(pop), X.numpy())
    assert bits_equal(out, ora), describe_mismatch(out, ora)


def test_negative_stoichiometry_entries(ms):
    # Nf/Nb < 0 are invalid but the reference handles them (pow(0, k<0) = inf, kinetics.py:905-915)
    pop = make_population(32, 10, 4, 8, seed=31)
    pop.Nf[3, 1, 2] = -2
    pop.Nb[5, 0, 1] = -1
    kin = install(_kinetics(ms, 8), pop)
    X = pop.X0()
    out = kin.integrate_signals(X=X)
    ora = O.integrate_signals(oracle_params(pop), X.numpy())
    assert bits_equal(out, ora), describe_mismatch(out, ora)


def test_nan_parameters_propagate_like_reference(ms):
    pop = make_population(16, 8, 4, 8, seed=33)
    pop.Vmax[2, 0] = float("nan")
    kin = install(_kinetics(ms, 8), pop)
    X = pop.X0()
    out = kin.integrate_signals(X=X).cpu().numpy()
    ora = O.integrate_signals(oracle_params(pop), X.numpy())
    assert np.isnan(out[2]).all() and np.isnan(ora[2]).all()  # 0 * NaN reaches every signal of the cell
    assert bits_equal(out, ora), describe_mismatch(out, ora)


def test_empty_and_unset_cells(ms):
    kin = _kinetics(ms, 8)
    out = kin.integrate_signals(X=torch.zeros(0, 8))
    assert out.shape == (0, 8)
    pop = make_population(20, 6, 4, 8, seed=34)
    kin = install(kin, pop)
    kin.unset_cell_params(cell_idxs=[0, 5, 19])  # all-zero parameters incl. Ke = 0 (kinetics.py:627-644)
    X = pop.X0()
    out = kin.integrate_signals(X=X)
    P = O.Params(*[getattr(kin, k).cpu().numpy() for k in PARAMS])
    ora = O.integrate_signals(P, X.numpy())
    assert bits_equal(out, ora), describe_mismatch(out, ora)
    assert bits_equal(out[[0, 5, 19]], X[[0, 5, 19]])  # cells without proteins are unchanged


@pytest.mark.parametrize("tag", [t for t in SYNTHETIC_TAGS if t != "two"])
def test_golden_synthetic_vs_reference(ms, tag):
    # the reference's own outputs for seeded populations (incl. the benchmarked tile shapes c4,
    # c4e, c3, s82): within its summation-order noise floor, and bit-identical to the oracle
    z = np.load(GOLDEN / f"ref_synthetic_{tag}.npz")
    kin = _kinetics(ms, z["X0"].shape[1])
    _set(kin, {k: z[k] for k in PARAMS})
    out = kin.integrate_signals(X=torch.from_numpy(z["X0"])).cpu().numpy()
    mm = mismatch_fraction(out, z["X_step1"])
    assert mm <= max(1.25 * float(z["noise_floor_step1"]), 0.01), (tag, mm)
    P = O.Params(*[z[k] for k in PARAMS])
    ora = O.integrate_signals(P, z["X0"])
    assert bits_equal(out, ora), describe_mismatch(out, ora)
    V = kin._get_velocities(X=torch.from_numpy(z["X0"]), Vmax=torch.from_numpy(z["adj_vmax"])).cpu().numpy()
    assert (np.abs(V - z["V"]) <= velocity_bound(z["V"], z["adj_vmax"])).all()


# ------------------------------------------------------------------ (3) robustness properties
def test_random_kinetics_stay_zero(ms):
    # tests/slow/test_kinetics.py:81-127 (fewer steps)
    pop = dense_adversarial(100, 100, 28, seed=41)
    pop.Vmax = torch.randn(100, 100).abs() * 100
    kin = install(_kinetics(ms, 28), pop)
    X = torch.zeros(100, 28, device="cuda")
    for _ in range(50):
        X = kin.integrate_signals(X=X)
        assert X.min() == 0.0 and X.max() == 0.0
    assert kin.N.dtype is torch.int32 and kin.Kmr.dtype is torch.float32 and X.dtype is torch.float32


def test_random_kinetics_dont_explode(ms):
    # tests/slow/test_kinetics.py:130-178 (fewer steps)
    pop = dense_adversarial(100, 100, 28, seed=42)
    kin = install(_kinetics(ms, 28), pop)
    g = torch.Generator().manual_seed(1)
    X = (torch.randn(100, 28, generator=g).abs().clamp(max=1.0) * 100).cuda()
    for _ in range(100):
        X = kin.integrate_signals(X=X)
        assert not torch.any(X < 0.0)
        assert not torch.any(X.isnan())
        assert torch.all(X.isfinite())
        assert torch.all(X < 10_000)


def test_large_population_invariants(ms):
    # C4 tile shape at 1/16 of its cell count; oracle comparison on a subsample (cells are
    # independent while every equilibrium iteration is applied)
    pop = make_population(65536, 64, 16, 256, seed=51, device="cuda")
    kin = ms.Kinetics(chemistry=_Chem(16), device="cuda")
    for k in PARAMS:
        setattr(kin, k, getattr(pop, k))
    X = pop.X0()
    out = kin.integrate_signals(X=X)
    assert torch.isfinite(out).all() and (out >= 0).all()
    idx = torch.arange(0, 65536, 97, device="cuda")
    sub = O.Params(*[t[idx].cpu().numpy() for t in pop.params()])
    ora, done = O.integrate_signals(sub, X[idx].cpu().numpy(), return_eq_done=True)
    assert done == [4, 4, 4]
    assert bits_equal(out[idx], ora), describe_mismatch(out[idx], ora)
    # idempotent launch: same inputs, same bits
    assert torch.equal(out, kin.integrate_signals(X=X))


@pytest.mark.parametrize("name", ["ref_drift_1000.npz", "ref_drift_1000_c4.npz"])
def test_drift_over_1000_steps_gpu(ms, name):
    # 1000 steps on the device: bit-identical to the oracle every 100 steps, and within the
    # reference's own drift bounds (tests/test_oracle_golden.py::check_drift)
    from tests.test_oracle_golden import drift_population, check_drift

    z, pop, n_steps = drift_population(name)
    kin = install(_kinetics(ms, pop.s), pop)
    P = oracle_params(pop)
    X = pop.X0().cuda()
    Xo = pop.X0().numpy()
    stats = {"mean": [], "q99": [], "total": [], "zero_frac": []}
    for step in range(1, n_steps + 1):
        X = kin.integrate_signals(X=X)
        Xo = O.integrate_signals(P, Xo)
        if step % 100 == 0:
            x = X.cpu().numpy()
            assert bits_equal(x, Xo), f"step {step}: " + describe_mismatch(x, Xo)
            stats["mean"].append(x.mean(0))
            stats["q99"].append(np.quantile(x, 0.99, axis=0))
            stats["total"].append(x.sum(dtype=np.float64))
            stats["zero_frac"].append((x == 0).mean())
    check_drift(z, {k: np.array(v) for k, v in stats.items()})
