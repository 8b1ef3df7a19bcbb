"""
Seeded synthetic cell populations for parity tests and benchmarks.

Mirrors what the reference's `Kinetics.set_cell_params` produces from random genomes
(/root/reference/python/magicsoup/kinetics.py:571-625; invariants in SURVEY.md A.11)
without needing genomes: the reference's own `spawn_cells`/`divide_cells` are
quadratic in the cell count and unusable at 1M cells.

Everything is generated with torch on the requested device, in cell chunks so that a
1M x 64 x 32 population (44 GB of parameters) can be built directly in HBM.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch

GAS_CONSTANT = 8.31446261815324  # reference: python/magicsoup/constants.py
_EPS = 1e-36
_MAX = 1e36


@dataclass
class Population:
    """parameter tensors shaped like the reference's Kinetics attributes + world state"""

    N: torch.Tensor  # (c,p,s) int32
    Nf: torch.Tensor
    Nb: torch.Tensor
    A: torch.Tensor
    Kmr: torch.Tensor  # (c,p,s) f32
    Ke: torch.Tensor  # (c,p) f32
    Kmf: torch.Tensor
    Kmb: torch.Tensor
    Vmax: torch.Tensor
    cell_positions: torch.Tensor  # (c,2) int32
    cell_molecules: torch.Tensor  # (c,m) f32
    molecule_map: torch.Tensor  # (m,H,W) f32
    diffusivities: list[float]
    permeabilities: list[float]

    @property
    def c(self) -> int:
        return int(self.N.shape[0])

    @property
    def p(self) -> int:
        return int(self.N.shape[1])

    @property
    def s(self) -> int:
        return int(self.N.shape[2])

    @property
    def m(self) -> int:
        return self.s // 2

    def params(self) -> tuple[torch.Tensor, ...]:
        return (self.N, self.Nf, self.Nb, self.A, self.Kmr, self.Ke, self.Kmf, self.Kmb, self.Vmax)

    def X0(self) -> torch.Tensor:
        """world.py:619-621: [intracellular | extracellular at the cell's pixel]"""
        xs = self.cell_positions[:, 0].long()
        ys = self.cell_positions[:, 1].long()
        return torch.cat([self.cell_molecules, self.molecule_map[:, xs, ys].T], dim=1).contiguous()


def _lognormal_in_range(shape, lo: float, hi: float, gen, device) -> torch.Tensor:
    """log-normal centred in [lo,hi] (log space), clipped to the range (kinetics.py:53-66)"""
    mu = 0.5 * (math.log(lo) + math.log(hi))
    sd = (math.log(hi) - math.log(lo)) / 6.0
    x = torch.randn(shape, generator=gen, device=device) * sd + mu
    return x.clamp(math.log(lo), math.log(hi)).exp()


def _fill_chunk(c: int, p: int, m: int, gen, device, energies: torch.Tensor, abs_temp: float):
    s = 2 * m
    dev = device
    # active proteins per cell: Poisson(p/2) clipped to [1, p]
    n_act = torch.poisson(torch.full((c,), p / 2.0, device=dev), generator=gen).clamp(1, p)
    active = torch.arange(p, device=dev)[None, :] < n_act[:, None]  # (c,p)

    N = torch.zeros(c, p, s, dtype=torch.int32, device=dev)
    Nf = torch.zeros_like(N)
    Nb = torch.zeros_like(N)
    A = torch.zeros_like(N)
    Ka_sum = torch.zeros(c, p, s, device=dev)
    Ka_cnt = torch.zeros(c, p, s, device=dev)

    def rnd(*shape):
        return torch.rand(*shape, generator=gen, device=dev)

    def rint(lo, hi, *shape):
        return torch.randint(lo, hi, shape, generator=gen, device=dev)

    # 1-3 work domains per protein, 1.38 on average (1.35-1.41 for random genomes)
    u = rnd(c, p)
    n_dom = 1 + (u > 0.70).int() + (u > 0.92).int()
    for d in range(3):
        has = active & (n_dom > d)
        kind = torch.where(rnd(c, p) < 0.6, 1, 2)  # catalytic more likely
        sign = torch.where(rnd(c, p) < 0.5, 1, -1).int()
        Nd = torch.zeros(c, p, s, dtype=torch.int32, device=dev)
        # catalytic: <=4 intracellular signals with coefficients in {-2..2}
        for _ in range(2):
            sig = rint(0, m, c, p)
            coef = rint(-2, 3, c, p).int()
            use = (has & (kind == 1)).int()
            Nd.scatter_add_(2, sig[..., None], (coef * use)[..., None])
        # make sure a catalytic domain has both sides: substrate at sig_a, product at sig_b
        sig_a = rint(0, m, c, p)
        sig_b = (sig_a + rint(1, m, c, p)) % m
        use = (has & (kind == 1)).int()
        Nd.scatter_add_(2, sig_a[..., None], (-use)[..., None])
        Nd.scatter_add_(2, sig_b[..., None], use[..., None])
        Nd.clamp_(-3, 3)
        # transporter: -1 at i, +1 at i+m
        ti = rint(0, m, c, p)
        use = (has & (kind == 2)).int()
        Nd.scatter_add_(2, ti[..., None], (-use)[..., None])
        Nd.scatter_add_(2, (ti + m)[..., None], use[..., None])
        Nd = Nd * sign[..., None]
        N += Nd
        Nf += torch.where(Nd < 0, -Nd, 0)
        Nb += torch.where(Nd > 0, Nd, 0)

    # 0-2 regulatory domains per protein
    n_reg = torch.where(rnd(c, p) < 0.5, 0, torch.where(rnd(c, p) < 0.7, 1, 2))
    hill_p = torch.tensor([0.52, 0.26, 0.13, 0.06, 0.03], device=dev)
    for d in range(2):
        has = active & (n_reg > d)
        sig = rint(0, s, c, p)
        hill = (torch.multinomial(hill_p, c * p, replacement=True, generator=gen).view(c, p) + 1).int()
        sign = torch.where(rnd(c, p) < 0.5, 1, -1).int()
        A.scatter_add_(2, sig[..., None], (hill * sign * has.int())[..., None])
        ka = _lognormal_in_range((c, p), 1e-2, 100.0, gen, dev)
        Ka_sum.scatter_add_(2, sig[..., None], (ka * has)[..., None])
        Ka_cnt.scatter_add_(2, sig[..., None], has.float()[..., None])

    Ka = torch.where(Ka_cnt > 0, Ka_sum / Ka_cnt.clamp(min=1), 0.0)
    Kmr = torch.pow(Ka, A)  # kinetics.py:592 ; 0^0 = 1 where not regulating

    Vmax = _lognormal_in_range((c, p), 1e-3, 100.0, gen, dev) * active
    Km = _lognormal_in_range((c, p), 1e-2, 100.0, gen, dev) * active
    E = torch.einsum("cps,s->cp", N.float(), energies)
    Ke = torch.exp(-E / abs_temp / GAS_CONSTANT).clamp(_EPS, _MAX)
    is_fwd = Ke >= 1.0
    Kmf = torch.where(is_fwd, Km, Km / Ke).clamp(_EPS, _MAX)
    Kmb = torch.where(is_fwd, Km * Ke, Km).clamp(_EPS, _MAX)
    return N, Nf, Nb, A, Kmr, Ke, Kmf, Kmb, Vmax


def make_population(
    n_cells: int,
    n_proteins: int,
    n_molecules: int,
    map_size: int | tuple[int, int],
    seed: int = 0,
    device: str | torch.device = "cpu",
    abs_temp: float = 310.0,
    chunk: int = 32768,
    energy_range: tuple[float, float] = (10e3, 50e3),
) -> Population:
    """
    SURVEY.md §8(d) direct-parameter synthetic population.  `energy_range`: the molecule
    energies are U(lo, hi) J/mol.  The default (10e3, 50e3) keeps Ke = exp(-dE/RT) inside
    float32 for most proteins, so the equilibrium decisions are exercised near their
    thresholds; SURVEY.md §8(d)'s (10e3, 500e3) saturates most Ke (pass it explicitly:
    `bench.py --energies survey` and the `c4e` golden fixture use it).
    """
    H, W = (map_size, map_size) if isinstance(map_size, int) else map_size
    if n_cells > H * W:
        raise ValueError("more cells than pixels")
    device = torch.device(device)
    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    m, p, c, s = n_molecules, n_proteins, n_cells, 2 * n_molecules

    e_lo, e_hi = energy_range
    e_int = torch.rand(m, generator=gen, device=device) * (e_hi - e_lo) + e_lo
    energies = torch.cat([e_int, e_int])

    names = ("N", "Nf", "Nb", "A", "Kmr", "Ke", "Kmf", "Kmb", "Vmax")
    out: dict[str, torch.Tensor] = {}
    for k in names:
        shape = (c, p, s) if k in ("N", "Nf", "Nb", "A", "Kmr") else (c, p)
        dt = torch.int32 if k in ("N", "Nf", "Nb", "A") else torch.float32
        out[k] = torch.empty(shape, dtype=dt, device=device)
    for a in range(0, c, chunk):
        b = min(c, a + chunk)
        vals = _fill_chunk(b - a, p, m, gen, device, energies, abs_temp)
        for k, v in zip(names, vals):
            out[k][a:b] = v
        del vals

    molecule_map = (torch.randn(m, H, W, generator=gen, device=device) + 10.0).abs()
    perm = torch.randperm(H * W, generator=gen, device=device)[:c]
    pos = torch.stack([perm // W, perm % W], dim=1).int()
    xs, ys = pos[:, 0].long(), pos[:, 1].long()
    pickup = molecule_map[:, xs, ys] * 0.5  # world.py:335-338
    cell_molecules = pickup.T.contiguous()
    molecule_map[:, xs, ys] -= pickup

    diffs = [(1.0, 0.1, 0.01, 0.0)[i % 4] for i in range(m)]
    perms = [(1.0, 0.1, 0.0)[i % 3] for i in range(m)]
    return Population(
        cell_positions=pos, cell_molecules=cell_molecules, molecule_map=molecule_map,
        diffusivities=diffs, permeabilities=perms, **out,
    )
