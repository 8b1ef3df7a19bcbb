"""
`World` with the reference's interface (mRcSchwering/magic-soup 1.0.5,
python/magicsoup/world.py:36-1004).  The per-step world math runs as CUDA kernels:

  enzymatic_activity()        world.py:610-625   fused kinetics kernel + write-back
  diffuse_molecules()         world.py:627-665   tiled stencil, mass correction, permeation
  degrade_molecules()         world.py:667-678   one launch, or folded into the diffusion
  increment_cell_lifetimes()  world.py:680-687

and the bookkeeping around it keeps every tensor the kernels read index-consistent:

  spawn_cells / divide_cells / update_cells / kill_cells / move_cells / reposition_cells /
  mutate_cells / recombinate_cells / get_neighbors          world.py:229-608, 689-745
  add_cells_from_params / divide_cells_at                   the same with ready-made parameter
                                                            tensors / caller-chosen pixels
  save / from_file / save_state / load_state                world.py:747-878

String and map searches run in `magicsoup_b200._lib` (C++, the interface of the reference's Rust
extension), `Kinetics.set_cell_params` as one kernel.

State attributes keep the reference's names, shapes and dtypes (world.py:192-204):
cell_map (H,W) bool, cell_positions (c,2) i32, cell_lifetimes/cell_divisions (c,) i32,
cell_molecules (c,m) f32, molecule_map (m,H,W) f32.  As in the reference,
enzymatic_activity/diffuse_molecules update molecule_map in place and REBIND
cell_molecules to a new tensor.  cell_molecules, molecule_map and cell_positions are properties:
reading one first finishes a step whose early-exit check is still pending (`_settle`).

No CPU fallback: the step methods raise without a CUDA device.
"""
from __future__ import annotations

import ctypes as C
import io
import math
import pickle
import random
from dataclasses import dataclass
from pathlib import Path
from typing import Any

import numpy as np
import torch

from magicsoup_b200 import _native as nat
from magicsoup_b200._hostlib import lib as _hostlib
from magicsoup_b200.genetics import Genetics
from magicsoup_b200.mutations import point_mutations, recombinations
from magicsoup_b200.util import randstr
from magicsoup_b200.kinetics import Kinetics, _settle_early_exit, _ALL_PARAMS, _STAGE_METHODS


@dataclass(frozen=True)
class Molecule:
    """the attributes of the reference's Molecule that the world math reads (containers.py:91-132)"""

    name: str
    energy: float
    half_life: int = 100_000
    diffusivity: float = 0.1
    permeability: float = 0.0


@dataclass
class Chemistry:
    """molecules (ordered) + reactions, as containers.py's Chemistry"""

    molecules: list
    reactions: list


def permeation_factor(mol_perm_rate: float) -> float:
    """world.py:935-946 _get_permeate"""
    rate = min(abs(mol_perm_rate), 1.0)
    if rate == 0.0:
        return 0.0
    return 1 / (1 / rate + 1)


def diffusion_weights(mol_diff_rate: float) -> tuple[float, float]:
    """world.py:948-964 _get_diffuse: (neighbour weight a, centre weight b) of the 3x3 kernel"""
    rate = min(abs(mol_diff_rate), 1.0)
    if rate == 0.0:
        return 0.0, 1.0
    d = 1 / rate
    a = 1 / (d + 8)
    b = d * a
    b = b + 1.0 - (8 * a + b)
    return a, b


_KERNEL_HOOKS = ("_launch_activity", "_launch_scatter", "_launch_stencil", "_launch_correct", "_launch_permeate",
                 "_launch_degrade", "_device_for_step")


class _StepGate:
    """`with`: the library's committing kernels launched by this thread obey the gate word (ms_step_gate)"""

    def __init__(self, gate: torch.Tensor | None):
        self.gate = gate

    def __enter__(self):
        if self.gate is not None:
            nat.lib().ms_step_gate(nat.ptr(self.gate))

    def __exit__(self, *exc):
        if self.gate is not None:
            nat.lib().ms_step_gate(None)
        return False


_STATE_TENSORS = {"cell_molecules": torch.float32, "cell_map": torch.bool, "molecule_map": torch.float32,
                  "cell_lifetimes": torch.int32, "cell_positions": torch.int32, "cell_divisions": torch.int32}


class _TensorRehomingUnpickler(pickle.Unpickler):
    """pickle.load that puts every tensor on `device` (torch pickles storages through one loader function)"""

    def __init__(self, fh, device: str | None):
        super().__init__(fh)
        self._device = device

    def find_class(self, module, name):
        if self._device is not None and (module, name) == ("torch.storage", "_load_from_bytes"):
            return lambda raw: torch.load(io.BytesIO(raw), map_location=self._device, weights_only=False)
        return super().find_class(module, name)


def _remove_items(items: list, dead_sorted: list[int]):
    """drop the (ascending, unique) indices from `items` in place"""
    if len(dead_sorted) < 16:
        for i in reversed(dead_sorted):
            del items[i]
        return
    from magicsoup_b200._hostlib import lib as hostlib

    hostlib().remove_items(items, dead_sorted)


class World:
    """see module docstring; constructor signature of world.py:147-157"""

    def __init__(
        self,
        chemistry: Any,
        map_size: int = 128,
        abs_temp: float = 310.0,
        mol_map_init: str = "randn",
        start_codons: tuple[str, ...] = ("TTG", "GTG", "ATG"),
        stop_codons: tuple[str, ...] = ("TGA", "TAG", "TAA"),
        device: str = "cpu",
        batch_size: int | None = None,
    ):
        self.device = device
        self.batch_size = batch_size
        self.map_size = map_size  # the reference's maps are square; a (rows, cols) pair is accepted too
        self.abs_temp = abs_temp
        self.chemistry = chemistry
        self.start_codons = start_codons
        self.stop_codons = stop_codons
        self.kinetics = Kinetics(chemistry=chemistry, abs_temp=abs_temp, device=device)
        self.genetics = Genetics(start_codons=start_codons, stop_codons=stop_codons)

        molecules = list(chemistry.molecules)
        self.n_molecules = len(molecules)
        self._int_mol_idxs = list(range(self.n_molecules))
        self._ext_mol_idxs = list(range(self.n_molecules, self.n_molecules * 2))
        self._mol_degrads = [math.exp(-math.log(2) / d.half_life) for d in molecules]
        self._diffusion = [diffusion_weights(d.diffusivity) for d in molecules]
        self._permeation = [permeation_factor(d.permeability) for d in molecules]

        self._pending = None  # a step enqueued behind a kinetics launch whose flags are not read yet (_settle)
        self.n_cells = 0
        self.cell_genomes: list[str] = []
        self.cell_labels: list[str] = []
        n_rows, n_cols = (map_size, map_size) if isinstance(map_size, int) else map_size
        self.cell_map = torch.zeros(n_rows, n_cols, dtype=torch.bool, device=device)
        self.cell_positions = torch.zeros(0, 2, dtype=torch.int32, device=device)
        self.cell_lifetimes = torch.zeros(0, dtype=torch.int32, device=device)
        self.cell_divisions = torch.zeros(0, dtype=torch.int32, device=device)
        self.cell_molecules = torch.zeros(0, self.n_molecules, dtype=torch.float32, device=device)
        self.molecule_map = self._get_molecule_map(n=self.n_molecules, size=(n_rows, n_cols), init=mol_map_init)
        self._scratch: dict[str, torch.Tensor] = {}
        # when set to a list, (start, end) CUDA event pairs around each kinetics kernel
        # launch are appended to it (used by bench.py for the roofline of that kernel)
        self._kernel_events: list | None = None

    def __getstate__(self):
        self._settle()
        state = dict(self.__dict__)
        state["_scratch"] = {}
        state["_kernel_events"] = None
        state["_pending"] = None
        return state

    # ------------------------------------------------------------------ plumbing
    def _cuda(self) -> torch.device:
        nat.require_cuda()
        dev = torch.device(self.device)
        if dev.type != "cuda" or dev.index is None:
            dev = torch.device("cuda", torch.cuda.current_device())
        return dev

    def _buffer(self, key: str, nbytes: int, dev: torch.device) -> torch.Tensor:
        buf = self._scratch.get(key)
        if buf is None or buf.numel() < nbytes or buf.device != dev:
            buf = torch.empty(max(nbytes, 256), dtype=torch.uint8, device=dev)
            self._scratch[key] = buf
        return buf

    def _const_f32(self, key: str, values: list[float], dev: torch.device) -> torch.Tensor:
        """per-molecule factors as float32 device vectors, refreshed when the lists change"""
        cached = self._scratch.get(key)
        stamp = self._scratch.get(key + "_src")
        if cached is None or stamp != tuple(values) or cached.device != dev:
            cached = torch.tensor(values, dtype=torch.float32, device=dev)
            self._scratch[key] = cached
            self._scratch[key + "_src"] = tuple(values)  # type: ignore[assignment]
        return cached

    def _state_on_device(self, dev: torch.device):
        """normalise user-assigned state tensors (tests assign CPU tensors, world.py H5)"""
        mm = self._st_molecule_map
        if mm.device != dev or mm.dtype != torch.float32 or not mm.is_contiguous():
            mm = mm.to(device=dev, dtype=torch.float32).contiguous()
            self._st_molecule_map = mm
        cm = self._st_cell_molecules.to(device=dev, dtype=torch.float32).contiguous()
        pos = self._st_cell_positions.to(device=dev, dtype=torch.int32).contiguous()
        return mm, cm, pos

    # ------------------------------------------------------------------ deferred early-exit check
    # The reference's adjustment loop ends on a GLOBAL condition (kinetics.py:846-847), which the
    # host settles from 24 flag words after the kinetics launch (Kinetics._settle_early_exit); with
    # realistic populations the answer is always "the launch stands".  Blocking on the flags right
    # after the launch leaves the GPU idle while the host then enqueues the write-back and the whole
    # diffusion (nine launches and, on several GPUs, three collectives: ~0.45 ms per step).  Instead
    # the verdict is also computed on the device (ms_early_exit_verdict) and everything that commits
    # results until the flags are read is launched behind that gate (ms_step_gate): the host returns
    # from enzymatic_activity() without a synchronisation, enqueues the rest of the step while the
    # kinetics kernel runs, and reads the flags at the next touch of the state (_settle: the next
    # enzymatic_activity(), any access to cell_molecules / molecule_map / cell_positions or to the
    # kinetics parameters, any lifecycle call).  If the verdict was "repeat" nothing was committed,
    # and the step is redone from the untouched state the blocking way.  An out-of-map cell position
    # raises its IndexError at that point, like an asynchronous CUDA error.
    deferred_settle = True  # False: block on the flags inside enzymatic_activity()

    def _can_defer(self) -> bool:
        cls = type(self)
        return bool(self.deferred_settle) and all(getattr(cls, h) is getattr(World, h) for h in _KERNEL_HOOKS)

    def _reduce_flags_device(self, flags: torch.Tensor) -> torch.Tensor:
        """the flags of all ranks, still on the device (SlabWorld: MAX over ranks, no host round trip)"""
        return flags

    def _gate(self):
        """context: launches inside commit only if the pending verdict allows it"""
        return _StepGate(self._pending["gate"] if self._pending is not None else None)

    def _settle(self):
        pend = self._pending
        if pend is None:
            return
        self._pending = None
        pend["kin"]._guard = None
        flags = pend["flags"].tolist()  # the step's one synchronisation
        eq = [4, 4, 4]
        try:
            redo = _settle_early_exit(flags, eq, 3)
        except IndexError:
            self._st_cell_molecules = pend["cm"]  # nothing was committed
            raise
        if not redo:
            return
        # the reference would have left an adjustment loop early: the gate stayed shut, state is as before the step
        self._st_cell_molecules = pend["cm"]
        self._activity_blocking(eq)
        for op in pend["replay"]:
            op()

    def _activity_blocking(self, eq: list[int]):
        """enzymatic_activity with the flags read after every launch (eq is lowered until the launch stands)"""
        dev = self._device_for_step()
        mm, cm, pos = self._state_on_device(dev)
        c, s = self.n_cells, 2 * self.n_molecules
        X = torch.empty(c, s, dtype=torch.float32, device=dev)
        while True:
            flags = self._launch_activity(dev, pos, cm, mm, X, eq) if c > 0 else [0] * nat.N_FLAGS
            if not _settle_early_exit(self._reduce_flags(flags), eq, 3):
                break
        if c > 0:
            self._st_cell_molecules = self._launch_scatter(dev, X, pos, mm)

    # ------------------------------------------------------------------ step methods
    # The step methods are written against a few hooks so that the multi-GPU SlabWorld
    # (magicsoup_b200/distributed.py) only adds the exchange steps: _reduce_flags,
    # _exchange_halos, _reduce_sums, _n_pixels_total.
    def _reduce_flags(self, flags) -> list[int]:
        """the 24 early-exit flags as host integers (a device tensor is read back here: one sync)"""
        return flags.tolist() if isinstance(flags, torch.Tensor) else flags

    def _exchange_halos(self, mm: torch.Tensor):
        return None, None  # one device holds the whole map: rows wrap around inside it

    def _reduce_sums(self, sums: torch.Tensor) -> torch.Tensor:
        return sums

    def _n_pixels_total(self, H: int, W: int) -> int:
        return H * W

    def _launch_activity(self, dev, pos, cm, mm, X, eq):
        """ms_enzymatic_activity with eq iterations -> the 24 early-exit flags (int32, on the device)"""
        kin = self.kinetics
        m = self.n_molecules
        _, H, W = mm.shape
        c, s = self.n_cells, 2 * m
        st, cache, keep, p = kin._launch_args(dev, c, s)
        scratch, nbytes = kin._kin_scratch(dev, c, p, s)
        flags = kin._flags(dev)
        if self._kernel_events is not None:
            ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
            ev[0].record(torch.cuda.current_stream(dev))
        if cache is None:
            rc = nat.lib().ms_enzymatic_activity(
                st, nat.ptr(pos), nat.ptr(cm), nat.ptr(mm), m, H, W, nat.ptr(X), nat.eq_array(eq),
                nat.ptr(flags), nat.ptr(scratch), nbytes, nat.stream_ptr(dev))
        else:
            rc = nat.lib().ms_enzymatic_activity_cached(
                st, C.byref(cache), nat.ptr(pos), nat.ptr(cm), nat.ptr(mm), m, H, W, nat.ptr(X),
                nat.eq_array(eq), nat.ptr(flags), nat.ptr(scratch), nbytes, nat.stream_ptr(dev))
        nat.check(rc)
        if self._kernel_events is not None:
            ev[1].record(torch.cuda.current_stream(dev))
            self._kernel_events.append(ev)
        return flags[: nat.N_FLAGS * 4].view(torch.int32)

    def _pixel_index(self, dev, pos, H: int, W: int) -> torch.Tensor | None:
        """
        pixel -> cell index for the pixel-ordered kernels, or None on sparsely populated maps (there
        the per-cell kernels touch less memory).  Rebuilt when cell_positions change (tensor identity
        and version) -- one small launch.
        """
        c = self.n_cells
        if 4 * c < H * W or self.n_molecules > 32:
            return None
        # the stamp holds the tensor itself, so its address cannot be recycled for another one
        stamp = (pos, pos._version, c, H, W)
        old = self._scratch.get("pixel_index_stamp")
        if old is None or old[0] is not pos or old[1:] != stamp[1:]:
            buf = self._buffer("pixel_index", 4 * H * W, dev)
            nat.check(nat.lib().ms_build_pixel_index(nat.ptr(pos), c, H, W, nat.ptr(buf), nat.stream_ptr(dev)))
            self._scratch["pixel_index_stamp"] = stamp  # type: ignore[assignment]
        return self._scratch["pixel_index"]

    def _launch_scatter(self, dev, X, pos, mm) -> torch.Tensor:
        m = self.n_molecules
        _, H, W = mm.shape
        new_cm = torch.empty(self.n_cells, m, dtype=torch.float32, device=dev)
        pix = self._pixel_index(dev, pos, H, W)
        if pix is not None:
            nat.check(nat.lib().ms_scatter_signals_by_pixel(nat.ptr(X), nat.ptr(pix), m, H, W, nat.ptr(new_cm),
                                                            nat.ptr(mm), nat.stream_ptr(dev)))
            return new_cm
        nat.check(nat.lib().ms_scatter_signals(nat.ptr(X), nat.ptr(pos), self.n_cells, m, H, W, nat.ptr(new_cm),
                                               nat.ptr(mm), nat.stream_ptr(dev)))
        return new_cm

    def _launch_stencil(self, dev, mm, top, bottom, decay=None) -> tuple[torch.Tensor, torch.Tensor]:
        """ms_diffuse_stencil -> (stencil output in scratch, int64 plane sums + flags (5*m)); decay: degradation folded in"""
        m, H, W = mm.shape
        a = self._const_f32("diff_a", [w[0] for w in self._diffusion], dev)
        b = self._const_f32("diff_b", [w[1] for w in self._diffusion], dev)
        out = self._buffer("diffuse", m * H * W * 4, dev)
        sums = self._buffer("diffuse_sums", 5 * m * 8, dev)[: 5 * m * 8].view(torch.int64)
        if decay is None:
            rc = nat.lib().ms_diffuse_stencil(nat.ptr(mm), nat.ptr(top), nat.ptr(bottom), nat.ptr(a), nat.ptr(b),
                                              m, H, W, nat.ptr(out), nat.ptr(sums), nat.stream_ptr(dev))
        else:
            rc = nat.lib().ms_degrade_diffuse_stencil(nat.ptr(mm), nat.ptr(top), nat.ptr(bottom), nat.ptr(a), nat.ptr(b),
                                                      nat.ptr(decay), m, H, W, nat.ptr(out), nat.ptr(sums),
                                                      nat.stream_ptr(dev))
        nat.check(rc)
        return out, sums

    def _launch_correct(self, dev, out, mm, sums, n_total: int):
        m, H, W = mm.shape
        nat.check(nat.lib().ms_diffuse_correct(nat.ptr(out), nat.ptr(mm), nat.ptr(sums), m, H, W, n_total,
                                               nat.stream_ptr(dev)))

    def _launch_permeate(self, dev, cm, mm, pos, decay=None) -> torch.Tensor:
        m, H, W = mm.shape
        perm = self._const_f32("perm", list(self._permeation), dev)
        new_cm = torch.empty_like(cm)
        pix = self._pixel_index(dev, pos, H, W)
        if pix is not None:
            nat.check(nat.lib().ms_permeate_by_pixel(nat.ptr(cm), nat.ptr(new_cm), nat.ptr(mm), nat.ptr(pix), nat.ptr(perm),
                                                     nat.ptr(decay), m, H, W, nat.stream_ptr(dev)))
            return new_cm
        if decay is None:
            rc = nat.lib().ms_permeate(nat.ptr(cm), nat.ptr(new_cm), nat.ptr(mm), nat.ptr(pos), nat.ptr(perm),
                                       self.n_cells, m, H, W, nat.stream_ptr(dev))
        else:
            rc = nat.lib().ms_degrade_permeate(nat.ptr(cm), nat.ptr(new_cm), nat.ptr(mm), nat.ptr(pos), nat.ptr(perm),
                                               nat.ptr(decay), self.n_cells, m, H, W, nat.stream_ptr(dev))
        nat.check(rc)
        return new_cm

    def _launch_degrade(self, dev, mm, cm):
        m, H, W = mm.shape
        decay = self._const_f32("decay", list(self._mol_degrads), dev)
        nat.check(nat.lib().ms_degrade_molecules(nat.ptr(mm), nat.ptr(cm), nat.ptr(decay), cm.shape[0], m, H, W,
                                                 nat.stream_ptr(dev)))

    def _device_for_step(self) -> torch.device:
        return self._cuda()

    def enzymatic_activity(self):
        """world.py:610-625"""
        self._settle()
        if self._all_ranks_empty():
            return
        dev = self._device_for_step()
        kin = self.kinetics
        m = self.n_molecules
        c, s = self.n_cells, 2 * m
        if kin._overrides("integrate_signals") or any(kin._overrides(k) for k in _STAGE_METHODS):
            mm, cm, pos = self._state_on_device(dev)
            xs, ys = pos[:, 0].long(), pos[:, 1].long()
            X = kin.integrate_signals(X=torch.cat([cm, mm[:, xs, ys].T], dim=1))
            if c > 0:
                self._st_cell_molecules = self._launch_scatter(dev, X, pos, mm)
            return
        if not self._can_defer():
            self._activity_blocking([4, 4, 4])
            return
        mm, cm, pos = self._state_on_device(dev)
        X = torch.empty(c, s, dtype=torch.float32, device=dev)
        eq = [4, 4, 4]
        if c > 0:
            flags = self._launch_activity(dev, pos, cm, mm, X, eq)
        else:  # a rank without cells still takes part in the other ranks' decision
            flags = kin._flags(dev)[: nat.N_FLAGS * 4].view(torch.int32)
            flags.zero_()
        flags = self._reduce_flags_device(flags)
        gate = self._buffer("gate", 256, dev)
        nat.check(nat.lib().ms_early_exit_verdict(nat.ptr(flags), nat.eq_array(eq), 3, nat.ptr(gate), nat.stream_ptr(dev)))
        self._pending = {"flags": flags, "gate": gate, "cm": cm, "kin": kin, "replay": []}
        kin._guard = self._settle
        if c > 0:
            with self._gate():
                self._st_cell_molecules = self._launch_scatter(dev, X, pos, mm)

    def _all_ranks_empty(self) -> bool:
        return self.n_cells == 0

    @torch.no_grad()
    def diffuse_molecules(self, degrade_first: bool = False):
        """
        world.py:627-665.  degrade_first=True folds a preceding degrade_molecules() (world.py:667-678)
        into the stencil and permeation reads: same bits as calling both, one pass over the map less.
        """
        dev = self._device_for_step()
        mm, cm, pos = self._state_on_device(dev)
        m, H, W = mm.shape
        if m == 0:
            return
        decay = self._const_f32("decay", list(self._mol_degrads), dev) if degrade_first else None
        if self._pending is not None:  # not committed unless the pending verdict allows it; redone otherwise
            self._pending["replay"].append(lambda: self.diffuse_molecules(degrade_first=degrade_first))
        top, bottom = self._exchange_halos(mm)
        out, sums = self._launch_stencil(dev, mm, top, bottom, decay)
        sums = self._reduce_sums(sums)
        with self._gate():
            self._launch_correct(dev, out, mm, sums, self._n_pixels_total(H, W))
            if self.n_cells > 0:
                self._st_cell_molecules = self._launch_permeate(dev, cm, mm, pos, decay)

    def degrade_molecules(self):
        """world.py:667-678: per-molecule decay of the map and of the cell molecules, in place, one launch"""
        dev = self._device_for_step()
        mm, cm, _ = self._state_on_device(dev)
        if self._pending is not None:
            self._pending["replay"].append(self.degrade_molecules)
            cm = cm.clone() if cm is self._pending["cm"] else cm  # the step's input stays as it was
        self._st_cell_molecules = cm
        with self._gate():
            self._launch_degrade(dev, mm, cm)

    def increment_cell_lifetimes(self):
        """world.py:680-687"""
        self.cell_lifetimes += 1

    # ------------------------------------------------------------------ cell bookkeeping
    def add_cells_from_params(self, positions: torch.Tensor, cell_molecules: torch.Tensor | None = None,
                              **params: torch.Tensor) -> list[int]:
        """
        Append cells at the given free pixels with ready-made kinetic parameters (what
        spawn_cells + set_cell_params produce, world.py:287-341).  Without `cell_molecules`
        each new cell takes half of its pixel's molecules (world.py:335-338).
        """
        self._settle()
        n = int(positions.shape[0])
        if n == 0:
            return []
        dev = self.device
        pos = positions.to(device=dev, dtype=torch.int32)
        xs, ys = pos[:, 0].long(), pos[:, 1].long()
        if bool(self.cell_map[xs, ys].any()):
            raise ValueError("some positions are already occupied")
        new_idxs = list(range(self.n_cells, self.n_cells + n))
        self.n_cells += n
        self.cell_genomes.extend("" for _ in range(n))
        self.cell_labels.extend("".join(random.choices("abcdefghijklmnopqrstuvwxyz", k=12)) for _ in range(n))
        self.cell_lifetimes = self._expand_c(self.cell_lifetimes, n)
        self.cell_positions = self._expand_c(self.cell_positions, n)
        self.cell_divisions = self._expand_c(self.cell_divisions, n)
        self.cell_molecules = self._expand_c(self.cell_molecules, n)
        self.kinetics.increase_max_cells(by_n=n)
        self.cell_map[xs, ys] = True
        self.cell_positions[new_idxs] = pos
        if cell_molecules is None:
            pickup = self.molecule_map[:, xs, ys] * 0.5
            self.cell_molecules[new_idxs, :] += pickup.T
            self.molecule_map[:, xs, ys] -= pickup
        else:
            self.cell_molecules[new_idxs, :] = cell_molecules.to(device=dev, dtype=torch.float32)
        self.kinetics.set_cell_params_from_tensors(new_idxs, **params)
        return new_idxs

    def kill_cells(self, cell_idxs: list[int] | None = None):
        """world.py:495-540: spill molecules, compact every per-cell tensor order-preservingly"""
        self._settle()
        if cell_idxs is None:
            cell_idxs = list(range(self.n_cells))
        if len(cell_idxs) == 0:
            return
        dead = sorted(set(cell_idxs))
        idx = torch.as_tensor(dead, dtype=torch.long, device=self.cell_positions.device)  # one upload
        xs = self.cell_positions[idx, 0].long()
        ys = self.cell_positions[idx, 1].long()
        self.cell_map[xs, ys] = False
        self.molecule_map[:, xs, ys] += self.cell_molecules[idx, :].T
        keep = torch.ones(self.cell_lifetimes.size(0), dtype=torch.bool, device=idx.device)
        keep[idx] = False
        self.cell_lifetimes = self.cell_lifetimes[keep]
        self.cell_positions = self.cell_positions[keep]
        self.cell_divisions = self.cell_divisions[keep]
        self.cell_molecules = self.cell_molecules[keep]
        self.kinetics.remove_cell_params(keep=keep)
        _remove_items(self.cell_genomes, dead)
        _remove_items(self.cell_labels, dead)
        self.n_cells -= len(dead)

    def divide_cells_at(self, parent_idxs: list[int], child_positions: torch.Tensor) -> list[tuple[int, int]]:
        """
        world.py:446-475 after the reference's Rust helper has picked the children's pixels:
        children are appended, inherit the parents' parameters and half of their molecules.
        """
        self._settle()
        n = len(parent_idxs)
        if n == 0:
            return []
        dev = self.cell_positions.device
        child_idxs = list(range(self.n_cells, self.n_cells + n))
        parents = torch.as_tensor(parent_idxs, dtype=torch.long, device=dev)  # one upload, reused below
        children = torch.arange(self.n_cells, self.n_cells + n, dtype=torch.long, device=dev)
        both = torch.cat([parents, children])
        self.n_cells += n
        genomes, labels = self.cell_genomes, self.cell_labels
        genomes.extend([genomes[d] for d in parent_idxs])
        labels.extend([labels[d] for d in parent_idxs])
        self.cell_lifetimes = self._expand_c(self.cell_lifetimes, n)
        self.cell_positions = self._expand_c(self.cell_positions, n)
        self.cell_divisions = self._expand_c(self.cell_divisions, n)
        self.cell_molecules = self._expand_c(self.cell_molecules, n)
        self.kinetics.increase_max_cells(by_n=n)
        self.kinetics.copy_cell_params(from_idxs=parents, to_idxs=children)
        pos = child_positions.to(device=dev, dtype=torch.int32)
        self.cell_map[pos[:, 0].long(), pos[:, 1].long()] = True
        self.cell_positions[children] = pos
        self.cell_molecules[children] = self.cell_molecules[parents]
        self.cell_molecules[both] *= 0.5
        self.cell_divisions[children] = self.cell_divisions[parents]
        self.cell_divisions[both] += 1
        self.cell_lifetimes[both] = 0
        return list(zip(parent_idxs, child_idxs))

    # ------------------------------------------------------------------ genome-driven lifecycle
    # (world.py:229-608, 689-745, 880-920; string and map searches in magicsoup_b200._lib)
    def _positions_array(self):
        """cell positions as an (n, 2) host array for the map searches of `_lib`"""
        return self.cell_positions.cpu().numpy()

    def _map_shape(self):
        """the map as `_lib`'s searches take it: an int (square torus, the reference's maps) or (rows, cols, wrap_rows)"""
        if isinstance(self.map_size, int):
            return self.map_size
        rows, cols = self.map_size
        return (int(rows), int(cols), True)

    def get_neighbors(self, cell_idxs: list[int], nghbr_idxs: list[int] | None = None) -> list[tuple[int, int]]:
        """world.py:229-285: unique pairs (lower index first) of cells in each other's Moore neighbourhood"""
        if len(cell_idxs) == 0:
            return []
        from_idxs = list(set(cell_idxs))
        to_idxs = list(set(cell_idxs)) if nghbr_idxs is None else list(set(nghbr_idxs))
        if len(to_idxs) == 0:
            return []
        return _hostlib().get_neighbors(from_idxs, to_idxs, self._positions_array(), self._map_shape())

    def spawn_cells(self, genomes: list[str]) -> list[int]:
        """world.py:287-341: new cells on random free pixels, each takes half of its pixel's molecules"""
        self._settle()
        n = len(genomes)
        if n == 0:
            return []
        free = self._find_free_random_positions(n_cells=n)
        if free.size(0) == 0:
            return []
        if free.size(0) < n:
            n = int(free.size(0))
            genomes = random.sample(list(genomes), k=n)
        pos = free[:n]
        new_idxs = list(range(self.n_cells, self.n_cells + n))
        self.n_cells += n
        self.cell_genomes.extend(genomes)
        self.cell_labels.extend(randstr(n=12) for _ in range(n))
        self.cell_lifetimes = self._expand_c(self.cell_lifetimes, n)
        self.cell_positions = self._expand_c(self.cell_positions, n)
        self.cell_divisions = self._expand_c(self.cell_divisions, n)
        self.cell_molecules = self._expand_c(self.cell_molecules, n)
        self.kinetics.increase_max_cells(by_n=n)
        xs, ys = pos[:, 0].long(), pos[:, 1].long()
        self.cell_map[xs, ys] = True
        self.cell_positions[new_idxs] = pos.to(self.cell_positions.device)
        pickup = self.molecule_map[:, xs, ys] * 0.5
        self.cell_molecules[new_idxs, :] += pickup.T.to(self.cell_molecules.device)
        self.molecule_map[:, xs, ys] -= pickup
        self._update_cell_params(genomes=list(genomes), idxs=new_idxs)
        return new_idxs

    def divide_cells(self, cell_idxs: list[int]) -> list[tuple[int, int]]:
        """world.py:406-475: every cell with a free Moore pixel puts a clone there; (ancestor, new) pairs"""
        self._settle()
        if len(cell_idxs) == 0:
            return []
        parents, children, child_pos = _hostlib().divide_cells_if_possible(
            list(set(cell_idxs)), self._positions_array(), self.n_cells, self._map_shape())
        if len(children) == 0:
            return []
        pos = torch.tensor(child_pos, dtype=torch.int32, device=self.cell_positions.device)
        return self.divide_cells_at(list(parents), pos)

    def update_cells(self, genome_idx_pairs: list[tuple[str, int]]):
        """world.py:476-492: new genomes for existing cells, parameters re-derived"""
        self._settle()
        if len(genome_idx_pairs) == 0:
            return
        last = {}
        for genome, idx in genome_idx_pairs:  # a cell named twice keeps its last genome, and the matching parameters
            self.cell_genomes[idx] = genome
            last[idx] = genome
        self._update_cell_params(genomes=list(last.values()), idxs=list(last.keys()))

    def move_cells(self, cell_idxs: list[int] | None = None):
        """world.py:543-574: every cell steps onto a random free pixel of its Moore neighbourhood, if any"""
        self._settle()
        if cell_idxs is None:
            cell_idxs = list(range(self.n_cells))
        if len(cell_idxs) == 0:
            return
        new_pos, moved = _hostlib().move_cells(sorted(set(cell_idxs)), self._positions_array(), self._map_shape())
        if len(moved) == 0:
            return
        old = self.cell_positions[moved].long()
        self.cell_map[old[:, 0], old[:, 1]] = False
        new = torch.tensor(new_pos, dtype=torch.int32, device=self.cell_positions.device)
        self.cell_map[new[:, 0].long(), new[:, 1].long()] = True
        self.cell_positions[moved] = new

    def reposition_cells(self, cell_idxs: list[int] | None = None):
        """world.py:576-608: the cells are taken off the map and put on random free pixels"""
        self._settle()
        if cell_idxs is None:
            cell_idxs = list(range(self.n_cells))
        if len(cell_idxs) == 0:
            return
        cell_idxs = list(set(cell_idxs))
        old = self.cell_positions[cell_idxs].long()
        self.cell_map[old[:, 0], old[:, 1]] = False
        new = self._find_free_random_positions(n_cells=len(cell_idxs))
        self.cell_map[new[:, 0].long(), new[:, 1].long()] = True
        self.cell_positions[cell_idxs] = new.to(self.cell_positions.device)

    def mutate_cells(self, cell_idxs: list[int] | None = None, p: float = 1e-6, p_indel: float = 0.4, p_del: float = 0.66):
        """world.py:689-714: point mutations in the genomes, parameters of the changed cells re-derived"""
        self._settle()
        if cell_idxs is None:
            mutated = point_mutations(seqs=self.cell_genomes, p=p, p_indel=p_indel, p_del=p_del)
            self.update_cells(genome_idx_pairs=mutated)
        else:
            seqs = [self.cell_genomes[d] for d in cell_idxs]
            mutated = point_mutations(seqs=seqs, p=p, p_indel=p_indel, p_del=p_del)
            self.update_cells(genome_idx_pairs=[(g, cell_idxs[i]) for g, i in mutated])

    def recombinate_cells(self, cell_idxs: list[int] | None = None, p: float = 1e-7):
        """world.py:716-745: recombinations between the genomes of neighbouring cells"""
        self._settle()
        if cell_idxs is None:
            cell_idxs = list(range(self.n_cells))
        if len(cell_idxs) == 0:
            return
        if type(self).get_neighbors is World.get_neighbors and isinstance(self.cell_genomes, list):
            hits = _hostlib().neighbor_recombinations(self.cell_genomes, list(set(cell_idxs)), self._positions_array(),
                                                     self._map_shape(), p)
            changed = [pair for g0, g1, a, b in hits for pair in ((g0, a), (g1, b))]
        else:
            nghbrs = self.get_neighbors(cell_idxs=cell_idxs)
            pairs = [(self.cell_genomes[a], self.cell_genomes[b]) for a, b in nghbrs]
            changed = []
            for g0, g1, idx in recombinations(seq_pairs=pairs, p=p):
                a, b = nghbrs[idx]
                changed += [(g0, a), (g1, b)]
        self.update_cells(genome_idx_pairs=changed)

    def _update_cell_params(self, genomes: list[str], idxs: list[int]):
        """world.py:880-908: translate, unset cells without proteins, grow the protein axis, set the rest"""
        if len(genomes) == 0:
            return
        if type(self.genetics).translate_genomes is Genetics.translate_genomes and \
                type(self.kinetics).set_cell_params is Kinetics.set_cell_params:
            # flat arrays from the translation straight into the kernel: no nested Python lists
            prot_off, dom_off, doms = self.genetics.translate_genomes_flat(genomes)
            counts = prot_off[1:] - prot_off[:-1]
            has = counts > 0
            unset_idxs = [i for i, h in zip(idxs, has.tolist()) if not h]
            set_idxs = [i for i, h in zip(idxs, has.tolist()) if h]
            if unset_idxs:
                self.kinetics.unset_cell_params(cell_idxs=unset_idxs)
            if not set_idxs:
                return
            self.kinetics.increase_max_proteins(max_n=int(counts.max()))
            offsets = np.concatenate([prot_off[:1], prot_off[1:][has]])  # cells without proteins own no entries
            self.kinetics.set_cell_params_flat(set_idxs, offsets, dom_off, doms)
            return
        proteomes = self.genetics.translate_genomes(genomes=genomes)
        set_idxs, set_proteomes, unset_idxs = [], [], []
        for idx, proteome in zip(idxs, proteomes):
            if len(proteome) > 0:
                set_idxs.append(idx)
                set_proteomes.append(proteome)
            else:
                unset_idxs.append(idx)
        if unset_idxs:
            self.kinetics.unset_cell_params(cell_idxs=unset_idxs)
        if not set_idxs:
            return
        self.kinetics.increase_max_proteins(max_n=max(len(p) for p in set_proteomes))
        self.kinetics.set_cell_params(cell_idxs=set_idxs, proteomes=set_proteomes)

    def _find_free_random_positions(self, n_cells: int) -> torch.Tensor:
        """world.py:910-920: up to n_cells distinct free pixels, int32 (n,2)"""
        pxls = torch.nonzero(~self.cell_map).int()
        n = min(n_cells, int(pxls.size(0)))
        return pxls[random.sample(range(int(pxls.size(0))), k=n)]

    # ------------------------------------------------------------------ helpers
    # ------------------------------------------------------------------ persistence
    # (world.py:747-878; same file names and formats, so states written by either package load in the other)
    def save(self, rundir, name: str = "world.pkl"):
        """the whole object as one pickle (chemistry, genetics, parameters, state)"""
        rundir = Path(rundir)
        rundir.mkdir(parents=True, exist_ok=True)
        with open(rundir / name, "wb") as fh:
            pickle.dump(self, fh)

    @classmethod
    def from_file(cls, rundir, name: str = "world.pkl", device: str | None = None) -> "World":
        """a world written by save(); `device` re-homes all tensors (e.g. "cpu" for a world saved on a GPU)"""
        with open(Path(rundir) / name, "rb") as fh:
            world = _TensorRehomingUnpickler(fh, device).load()
        if device is not None:
            world.device = device
            world.kinetics.device = device
        return world

    def save_state(self, statedir):
        """what changes during a simulation: six tensors and the genomes with their labels as FASTA"""
        self._settle()
        statedir = Path(statedir)
        statedir.mkdir(parents=True, exist_ok=True)
        for key in _STATE_TENSORS:
            torch.save(getattr(self, key), statedir / f"{key}.pt")
        with open(statedir / "cells.fasta", "w", encoding="utf-8") as fh:
            fh.write("\n".join(f">{i} {label}\n{genome}"
                               for i, (genome, label) in enumerate(zip(self.cell_genomes, self.cell_labels))))

    def load_state(self, statedir, ignore_cell_params: bool = False):
        """a state written by save_state(); parameters are re-derived from the genomes unless told otherwise"""
        self._settle()
        statedir = Path(statedir)
        if not ignore_cell_params:
            self.kill_cells(cell_idxs=list(range(self.n_cells)))
        for key, dtype in _STATE_TENSORS.items():
            setattr(self, key, torch.load(statedir / f"{key}.pt", map_location=self.device).to(dtype))
        text = (statedir / "cells.fasta").read_text(encoding="utf-8")
        self.cell_genomes, self.cell_labels = [], []
        for record in (r.strip() for r in text.split(">")):
            if not record:
                continue
            head, _, seq = record.partition("\n")
            words = head.split()
            self.cell_genomes.append(seq.split("\n")[0])
            self.cell_labels.append(words[1].strip() if len(words) > 1 else "")
        self.n_cells = len(self.cell_genomes)
        self.kinetics.increase_max_cells(by_n=self.n_cells)
        if not ignore_cell_params:
            self.update_cells(genome_idx_pairs=[(g, i) for i, g in enumerate(self.cell_genomes)])

    def _get_molecule_map(self, n: int, size, init: str) -> torch.Tensor:
        """world.py:922-933"""
        rows, cols = (size, size) if isinstance(size, int) else size
        if init == "zeros":
            return torch.zeros(n, rows, cols, device=self.device, dtype=torch.float32)
        if init == "randn":
            return (torch.randn(n, rows, cols, dtype=torch.float32, device=self.device) + 10.0).abs()
        raise ValueError(f"Didnt recognize mol_map_init={init}. Should be one of: 'zeros', 'randn'.")

    def _expand_c(self, t: torch.Tensor, n: int) -> torch.Tensor:
        zeros = torch.zeros(n, *t.shape[1:], dtype=t.dtype, device=t.device)
        return torch.cat([t, zeros], dim=0)

    def __repr__(self) -> str:
        return f"{type(self).__name__}(map_size:{self.map_size!r},abs_temp:{self.abs_temp!r},device:{self.device!r})"


def _state_property(name: str) -> property:
    """
    cell_molecules, molecule_map, cell_positions: plain tensor attributes in the reference
    (world.py:161-204); here reading or assigning one first finishes a step that is still behind
    its early-exit check (World._settle), so users always see committed state.
    """
    key = "_st_" + name

    def getter(self) -> torch.Tensor:
        if self._pending is not None:
            self._settle()
        return self.__dict__[key]

    def setter(self, value):
        if self.__dict__.get("_pending") is not None:
            self._settle()
        self.__dict__[key] = value

    return property(getter, setter, doc=f"World.{name} (world.py:161-204)")


for _name in ("cell_molecules", "molecule_map", "cell_positions"):
    setattr(World, _name, _state_property(_name))
