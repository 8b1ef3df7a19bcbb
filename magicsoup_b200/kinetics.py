"""
`Kinetics` with the reference's interface (mRcSchwering/magic-soup 1.0.5,
python/magicsoup/kinetics.py:292-992) whose numeric core runs as hand-written sm_100a
CUDA kernels through the C ABI of include/magicsoup_b200.h.

Mirrored from the reference (same names, argument meaning, dtypes, error behaviour):
  parameter tensors  Ke, Kmf, Kmb, Vmax (c,p) f32; Kmr (c,p,s) f32; N, Nf, Nb, A (c,p,s) i32
                     (kinetics.py:399-409)
  integrate_signals(X)                                  kinetics.py:725
  _integrate_signals_part / _get_velocities / _multiply_signals / _get_quotient /
  _get_negative_adjusted_nv / _get_equilibrium_adjusted_x   kinetics.py:753-918
  unset_cell_params / copy_cell_params / remove_cell_params /
  increase_max_cells / increase_max_proteins            kinetics.py:627-723

  set_cell_params(cell_idxs, proteomes)                 kinetics.py:521-625, one kernel launch
  token maps km_map / vmax_map / sign_map / hill_map / reaction_map / transport_map /
  effector_map (kinetics.py:16-289; magicsoup_b200/tokens.py)

The parameter tensors are attributes that users and tests overwrite (`kinetics.N = tensor`),
possibly with CPU tensors: every call reads the current attributes and normalises them (device,
dtype, contiguity) at the boundary.  Behind them sits a capacity store with a row map and the
kernel's list cache (see "parameter store" below).

There is no CPU fallback: the numeric methods raise without a CUDA device.
`get_proteome` (parameters -> domain descriptions for the reference's `Cell` container,
kinetics.py:462-519) is not part of this package; `_lib.get_proteome` has the Rust function's interface.
"""
from __future__ import annotations

import ctypes as C
from typing import Any, Sequence

import torch

from magicsoup_b200 import _native as nat
from magicsoup_b200 import tokens as tk

_EPS = 1e-36
_MAX = 1e36
_MIN = -1e36
_TRIMS = (0.7, 0.2, 0.1)  # kinetics.py:746
_INT_PARAMS = ("N", "Nf", "Nb", "A")
_CPS_PARAMS = ("N", "Nf", "Nb", "A", "Kmr")
_ALL_PARAMS = ("N", "Nf", "Nb", "A", "Kmr", "Ke", "Kmf", "Kmb", "Vmax")
_STAGE_METHODS = ("_integrate_signals_part", "_get_velocities", "_multiply_signals", "_get_quotient",
                  "_get_negative_adjusted_nv", "_get_equilibrium_adjusted_x")


def _settle_early_exit(flags: list[int], eq: list[int], n_parts: int) -> bool:
    """
    Also raises IndexError if the kernel met a cell position outside the map (flag word 24), as the
    reference's advanced indexing does (world.py:621).

    kinetics.py:846-847: the reference leaves a part's adjustment loop at the first
    iteration where NO cell has an adjustable protein with |V| > 0.1.  `flags` come from
    a launch that applied eq[q] iterations in part q.  Returns True if eq was lowered
    (the launch must be repeated), False if the launch already equals the reference.
    """
    if len(flags) > nat.FLAG_BAD_POSITION and flags[nat.FLAG_BAD_POSITION]:
        raise IndexError("cell_positions holds a pixel outside the molecule map")
    for q in range(n_parts):
        for k in range(eq[q]):
            if flags[q * 4 + k] == 0:
                if flags[12 + q * 4 + k] != 0:
                    eq[q] = k  # some low-impact protein kept adjusting: not what the reference does
                    return True
                break  # nothing adjustable anywhere: later iterations were no-ops
    return False


class Kinetics:
    """see module docstring; constructor signature of kinetics.py:380-389"""

    def __init__(
        self,
        chemistry: Any,
        abs_temp: float = 310.0,
        km_range: tuple[float, float] = (1e-2, 100.0),
        vmax_range: tuple[float, float] = (1e-3, 100.0),
        device: str = "cpu",
        scalar_enc_size: int = 64 - 3,
        vector_enc_size: int = 4096 - 3 * 64,
    ):
        self.abs_temp = abs_temp
        self.device = device
        self.km_range = km_range
        self.vmax_range = vmax_range
        self.scalar_enc_size = scalar_enc_size
        self.vector_enc_size = vector_enc_size
        molecules = list(chemistry.molecules)
        self.mol_names = [d.name for d in molecules]
        self.mol_energies = torch.tensor([d.energy for d in molecules] * 2, dtype=torch.float32, device=device)
        n_signals = 2 * len(molecules)
        # token maps (kinetics.py:418-449), drawn like the reference's; set_cell_params reads them
        mol_2_idx = {d: i for i, d in enumerate(molecules)}
        self.km_map = tk.LogNormWeightMap(max_token=scalar_enc_size, weight_range=km_range, device=device)
        self.vmax_map = tk.LogNormWeightMap(max_token=scalar_enc_size, weight_range=vmax_range, device=device)
        self.sign_map = tk.SignMap(max_token=scalar_enc_size, device=device)
        self.hill_map = tk.HillMap(max_token=scalar_enc_size, device=device)
        self.reaction_map = tk.VectorMap(
            max_token=vector_enc_size, n_signals=n_signals, device=device,
            vectors=tk.reaction_vectors(mol_2_idx, list(getattr(chemistry, "reactions", [])), n_signals))
        self.transport_map = tk.VectorMap(max_token=vector_enc_size, n_signals=n_signals, device=device,
                                          vectors=tk.transport_vectors(len(molecules)))
        self.effector_map = tk.VectorMap(max_token=vector_enc_size, n_signals=n_signals, device=device,
                                         vectors=tk.effector_vectors(len(molecules)))
        z_i = lambda: self._zeros_i32_tensor(0, 0, n_signals)  # noqa: E731
        z_f = lambda: self._zeros_f32_tensor(0, 0)  # noqa: E731
        self._buf: dict[str, torch.Tensor] = {
            "N": z_i(), "Nf": z_i(), "Nb": z_i(), "A": z_i(), "Kmr": self._zeros_f32_tensor(0, 0, n_signals),
            "Ke": z_f(), "Kmf": z_f(), "Kmb": z_f(), "Vmax": z_f()}
        self._len: dict[str, int] = {k: 0 for k in _ALL_PARAMS}
        self._rows: torch.Tensor | None = None   # int32 row map (logical cell -> buffer row); None = identity
        self._phys_end = 0                       # buffer rows in use
        self._cache: torch.Tensor | None = None  # (buffer rows, record words) int32 list-cache records
        self.list_cache = True                   # keep the kernel's sparse lists between steps (p, s <= 64)
        self._stamps: dict[str, tuple[int, int]] = {}
        self._stamp()
        self._scratch: dict[str, torch.Tensor] = {}

    # ------------------------------------------------------------------ pickling
    # A World that has enqueued a step behind a kinetics launch whose early-exit flags it has not read
    # yet (World._settle) hangs itself in here: whoever reads or writes the parameters, or uses the
    # flag buffer, lets it finish that step first.
    _guard = None

    def _touch(self):
        guard = self._guard
        if guard is not None:
            guard()

    def __getstate__(self):
        self._touch()
        self._materialize()
        state = dict(self.__dict__)
        state.pop("_guard", None)
        state["_scratch"] = {}  # device scratch is rebuilt lazily (world.py:747-793 pickles the object)
        state["_cache"] = None
        state["_buf"] = {k: self._view(k).clone() for k in _ALL_PARAMS}  # without the spare capacity
        return state

    def __setstate__(self, state):
        self.__dict__.update(state)
        self._stamp()

    # ------------------------------------------------------------------ device plumbing
    def _cuda(self) -> torch.device:
        nat.require_cuda()
        dev = torch.device(self.device)
        if dev.type != "cuda" or dev.index is None:
            dev = torch.device("cuda", torch.cuda.current_device())
        return dev

    def _f32(self, t: Any, dev: torch.device) -> torch.Tensor:
        if not isinstance(t, torch.Tensor):
            t = torch.as_tensor(t)
        return t.to(device=dev, dtype=torch.float32).contiguous()

    def _i32(self, t: Any, dev: torch.device) -> torch.Tensor:
        if not isinstance(t, torch.Tensor):
            t = torch.as_tensor(t)
        return t.to(device=dev, dtype=torch.int32).contiguous()

    def _buffer(self, key: str, nbytes: int, dev: torch.device) -> torch.Tensor:
        buf = self._scratch.get(key)
        if buf is None or buf.numel() < nbytes or buf.device != dev:
            buf = torch.empty(max(nbytes, 256), dtype=torch.uint8, device=dev)
            self._scratch[key] = buf
        return buf

    def _params(self, dev: torch.device, c: int, s: int, need: Sequence[str] = _ALL_PARAMS):
        """
        Normalised dense views of the current attributes + the C struct.  Tensors a call does
        not read may be left at their initial (0,0[,s]) shape, as the reference's tests do.
        """
        p = None
        for k in need:
            t = self._view(k)
            if t.dim() >= 2 and t.shape[0] == c:
                p = int(t.shape[1])
                break
        if p is None:
            p = int(self._view(need[0]).shape[1])
        keep = {}
        for k in _ALL_PARAMS:
            t = self._view(k)
            want = (c, p, s) if k in _CPS_PARAMS else (c, p)
            if k in need:
                if tuple(t.shape) != want:
                    raise ValueError(f"Kinetics.{k} has shape {tuple(t.shape)}, expected {want}")
                keep[k] = self._i32(t, dev) if k in _INT_PARAMS else self._f32(t, dev)
            else:
                keep[k] = None
        st = nat.KineticsParams(**{k: nat.ptr(v) for k, v in keep.items()}, c=c, p=p, s=s)
        return st, keep, p

    def _launch_args(self, dev: torch.device, c: int, s: int):
        """
        Arguments of the fused three-part launches (integrate_signals, World.enzymatic_activity):
        (params struct, cache struct or None, tensors to keep alive, p).  With the store in its
        managed state on `dev` the kernels read the capacity buffers in place through the row
        map, and the list cache is used; anything else (user-assigned CPU tensors, odd dtypes)
        goes through _params' normalised copies without a cache.
        """
        b = self._buf
        # the same store as at the last launch (same tensors at the same versions, same row map, same
        # cache): the same arguments (this runs in every step; building them costs ~30 us of Python)
        memo = self._scratch.get("launch_args")
        rec_now = int(nat.lib().ms_kinetics_cache_record_words(int(b["N"].shape[1]), s)) \
            if self.list_cache and b["N"].dim() == 3 else 0  # (depends on the library's diagnostic switches)
        key = (dev, c, s, rec_now, id(self._rows), id(self._cache), self._len["N"],
               tuple(t.data_ptr() for t in b.values()))
        if memo is not None and memo[0] == key and not self._written_from_outside():
            return memo[1]
        if not (self._managed() and b["N"].device == dev and self._len["N"] == c and b["N"].shape[2] == s
                and all(t.is_contiguous() for t in b.values())):
            st, keep, p = self._params(dev, c, s)
            return st, None, keep, p
        p = int(b["N"].shape[1])
        st = nat.KineticsParams(**{k: nat.ptr(b[k]) for k in _ALL_PARAMS}, c=c, p=p, s=s)
        rec_words = rec_now
        if rec_words == 0:
            self._cache = None
        else:
            rows = int(b["N"].shape[0])
            if self._cache is None or tuple(self._cache.shape) != (rows, rec_words) or self._cache.device != dev:
                self._cache = torch.zeros(rows, rec_words, dtype=torch.int32, device=dev)
                self._stamp()
            elif self._written_from_outside():
                self._stale(None)  # someone wrote into an attribute view: every record may be stale
                self._stamp()
        if self._cache is None and self._rows is None:
            value = (st, None, b, p)
        else:
            cs = nat.KineticsCache(records=nat.ptr(self._cache), rec_words=rec_words, row_map=nat.ptr(self._rows))
            value = (st, cs, (dict(b), self._cache, self._rows), p)  # holds the tensors: ids in the key stay theirs
        key = (dev, c, s, rec_now, id(self._rows), id(self._cache), self._len["N"],
               tuple(t.data_ptr() for t in b.values()))  # (the cache may just have been created)
        self._scratch["launch_args"] = (key, value)
        return value

    def _kin_scratch(self, dev: torch.device, c: int, p: int, s: int):
        n = int(nat.lib().ms_kinetics_scratch_bytes(c, p, s))
        return self._buffer("kin", n, dev), n

    def _flags(self, dev: torch.device) -> torch.Tensor:
        return self._buffer("flags", nat.N_FLAGS * 4, dev)

    def _read_flags(self, flags: torch.Tensor) -> list[int]:
        return flags[: nat.N_FLAGS * 4].view(torch.int32).tolist()  # device -> host sync

    def _overrides(self, name: str) -> bool:
        return getattr(type(self), name) is not getattr(Kinetics, name)

    # ------------------------------------------------------------------ hot path
    def integrate_signals(self, X: torch.Tensor) -> torch.Tensor:
        """
        kinetics.py:725-751.  X (c,s) >= 0 -> new (c,s) tensor.  One fused kernel launch
        (plus, only if the reference's global early exit would have fired, a repeat).
        Subclasses that override a stage method get the staged composition instead.
        """
        self._touch()
        if any(self._overrides(m) for m in _STAGE_METHODS):
            for trim in _TRIMS:
                X = self._integrate_signals_part(adj_vmax=(self.Vmax * trim).clamp(0.0), X0=X)
            return X
        dev = self._cuda()
        Xd = self._f32(X, dev)
        c, s = Xd.shape
        out = torch.empty_like(Xd)
        if c == 0 or s == 0:
            return out
        st, cache, keep, p = self._launch_args(dev, c, s)
        scratch, nbytes = self._kin_scratch(dev, c, p, s)
        flags = self._flags(dev)
        eq = [4, 4, 4]
        while True:
            if cache is None:
                rc = nat.lib().ms_integrate_signals(
                    st, nat.ptr(Xd), nat.ptr(out), nat.eq_array(eq), nat.ptr(flags), nat.ptr(scratch),
                    nbytes, nat.stream_ptr(dev))
            else:
                rc = nat.lib().ms_integrate_signals_cached(
                    st, C.byref(cache), nat.ptr(Xd), nat.ptr(out), nat.eq_array(eq), nat.ptr(flags),
                    nat.ptr(scratch), nbytes, nat.stream_ptr(dev))
            nat.check(rc)
            if not _settle_early_exit(self._read_flags(flags), eq, 3):
                return out

    def _integrate_signals_part(self, adj_vmax: torch.Tensor, X0: torch.Tensor) -> torch.Tensor:
        """kinetics.py:753-769"""
        dev = self._cuda()
        X0d = self._f32(X0, dev)
        c, s = X0d.shape
        if c == 0 or s == 0:
            return X0d.clone()
        staged = any(self._overrides(m) for m in _STAGE_METHODS if m != "_integrate_signals_part")
        if staged:
            # composition through the (possibly overridden) stage methods
            V = self._get_velocities(X=X0d, Vmax=adj_vmax)
            NV = self._i32(self.N, dev).float() * V[:, :, None]  # kinetics.py:759, one rounding
            NV_adj = self._get_negative_adjusted_nv(NV=NV, X=X0d)
            X1 = self._sum_over_proteins(X0d, NV_adj)
            return self._get_equilibrium_adjusted_x(X0=X0d, X1=X1, NV=NV_adj, V=V)
        X1, _ = self._part_kernel(adj_vmax, X0d, dev)
        return X1

    def _part_kernel(self, adj_vmax, X0d, dev, want_v: bool = False):
        self._touch()
        c, s = X0d.shape
        st, keep, p = self._params(dev, c, s)
        adj = self._f32(adj_vmax, dev)
        if tuple(adj.shape) != (c, p):
            raise ValueError(f"adj_vmax has shape {tuple(adj.shape)}, expected {(c, p)}")
        scratch, nbytes = self._kin_scratch(dev, c, p, s)
        flags = self._flags(dev)
        X1 = torch.empty_like(X0d)
        V = torch.empty(c, p, dtype=torch.float32, device=dev) if want_v else None
        eq = [4, 0, 0]
        while True:
            nat.check(nat.lib().ms_integrate_signals_part(
                st, nat.ptr(adj), nat.ptr(X0d), nat.ptr(X1), eq[0], nat.ptr(flags), nat.ptr(V),
                nat.ptr(scratch), nbytes, nat.stream_ptr(dev)))
            if not _settle_early_exit(self._read_flags(flags), eq, 1):
                return X1, V

    def _sum_over_proteins(self, X0: torch.Tensor, NV: torch.Tensor) -> torch.Tensor:
        """X1 = max(X0 + sum_p NV, 0) in spec order (kinetics.py:763-764)"""
        dev = X0.device
        c, p, s = NV.shape
        X1 = torch.empty_like(X0)
        F = torch.empty(c, p, dtype=torch.float32, device=dev)
        nat.check(nat.lib().ms_add_protein_rates(nat.ptr(X0), nat.ptr(NV), c, p, s, nat.ptr(X1), nat.ptr(F),
                                                 nat.stream_ptr(dev)))
        return X1

    def _get_velocities(self, X: torch.Tensor, Vmax: torch.Tensor) -> torch.Tensor:
        """kinetics.py:771-806"""
        dev = self._cuda()
        Xd = self._f32(X, dev)
        c, s = Xd.shape
        st, keep, p = self._params(dev, c, s, need=("Nf", "Nb", "A", "Kmr", "Kmf", "Kmb"))
        vmax = self._f32(Vmax, dev)
        V = torch.empty(c, p, dtype=torch.float32, device=dev)
        nat.check(nat.lib().ms_get_velocities(st, nat.ptr(Xd), nat.ptr(vmax), nat.ptr(V), nat.stream_ptr(dev)))
        return V

    def _multiply_signals(self, X: torch.Tensor, N: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor]:
        """kinetics.py:894-918 -> (xx (c,p) f32, involved (c,p) bool)"""
        dev = self._cuda()
        Xd, Nd = self._f32(X, dev), self._i32(N, dev)
        c, p, s = Nd.shape
        xx = torch.empty(c, p, dtype=torch.float32, device=dev)
        inv = torch.empty(c, p, dtype=torch.uint8, device=dev)
        nat.check(nat.lib().ms_multiply_signals(nat.ptr(Xd), nat.ptr(Nd), c, p, s, nat.ptr(xx), nat.ptr(inv),
                                                nat.stream_ptr(dev)))
        return xx, inv.bool()

    def _get_quotient(self, X: torch.Tensor) -> torch.Tensor:
        """kinetics.py:881-892"""
        dev = self._cuda()
        Xd = self._f32(X, dev)
        Nf, Nb = self._i32(self.Nf, dev), self._i32(self.Nb, dev)
        c, p, s = Nf.shape
        Q = torch.empty(c, p, dtype=torch.float32, device=dev)
        nat.check(nat.lib().ms_get_quotient(nat.ptr(Xd), nat.ptr(Nf), nat.ptr(Nb), c, p, s, nat.ptr(Q),
                                            nat.stream_ptr(dev)))
        return Q

    def _get_negative_adjusted_nv(self, NV: torch.Tensor, X: torch.Tensor) -> torch.Tensor:
        """kinetics.py:861-879 (the reference's test passes an int32 NV; result is float32)"""
        dev = self._cuda()
        NVd, Xd = self._f32(NV, dev), self._f32(X, dev)
        c, p, s = NVd.shape
        out = torch.empty_like(NVd)
        fsig = torch.empty(c, s, dtype=torch.float32, device=dev)
        nat.check(nat.lib().ms_get_negative_adjusted_nv(nat.ptr(NVd), nat.ptr(Xd), c, p, s, nat.ptr(out),
                                                        nat.ptr(fsig), nat.stream_ptr(dev)))
        return out

    def _get_equilibrium_adjusted_x(self, X0: torch.Tensor, X1: torch.Tensor, NV: torch.Tensor,
                                    V: torch.Tensor) -> torch.Tensor:
        """kinetics.py:808-859 on dense NV (c,p,s)"""
        self._touch()
        dev = self._cuda()
        X0d, X1d, NVd, Vd = (self._f32(t, dev) for t in (X0, X1, NV, V))
        c, p, s = NVd.shape
        Nf, Nb, Ke = self._i32(self.Nf, dev), self._i32(self.Nb, dev), self._f32(self.Ke, dev)
        out = torch.empty_like(X1d)
        F = torch.empty(c, p, dtype=torch.float32, device=dev)
        flags = self._flags(dev)
        eq = [4, 0, 0]
        while True:
            nat.check(nat.lib().ms_get_equilibrium_adjusted_x(
                nat.ptr(X0d), nat.ptr(X1d), nat.ptr(NVd), nat.ptr(Vd), nat.ptr(Nf), nat.ptr(Nb), nat.ptr(Ke),
                c, p, s, nat.ptr(out), nat.ptr(F), eq[0], nat.ptr(flags), nat.stream_ptr(dev)))
            if not _settle_early_exit(self._read_flags(flags), eq, 1):
                return out

    # ------------------------------------------------------------------ parameter store
    # The nine public attributes (kinetics.py:399-409) are views of capacity buffers.  The
    # reference re-allocates all of them on every change of the cell count (torch.cat in
    # kinetics.py:688-703, boolean-mask copies in :667-686: 42 KB per cell at p=64, s=32, i.e. two
    # 44 GB copies per kill or divide call at 1M cells).  Here
    #   * growth is amortised (capacity grows by half, new rows are zeroed);
    #   * removing cells only edits an int32 row map (logical cell -> buffer row, increasing); the
    #     kernels read parameter tiles through it, nothing is moved;
    #   * reading an attribute compacts the buffers in place first (chunked, no second copy), so
    #     users and tests always see the reference's dense (c,p[,s]) tensors;
    #   * assigning an attribute adopts the tensor as the new buffer, as in the reference.
    # A per-row cache of the kernel's sparse lists rides along (see _launch_args).

    def _managed(self) -> bool:
        """all nine tensors describe the same cells in same-sized buffers on one device"""
        b = self._buf
        n, rows = self._len["N"], b["N"].shape[0]
        if b["N"].dim() != 3:
            return False
        p, s_ = b["N"].shape[1], b["N"].shape[2]
        dev = b["N"].device
        for k in _ALL_PARAMS:
            t = b[k]
            want = (rows, p, s_) if k in _CPS_PARAMS else (rows, p)
            dt = torch.int32 if k in _INT_PARAMS else torch.float32
            if tuple(t.shape) != want or self._len[k] != n or t.device != dev or t.dtype != dt:
                return False
        return True

    @property
    def _n(self) -> int:
        return self._len["N"]

    def _view(self, k: str) -> torch.Tensor:
        self._touch()
        self._materialize()
        t = self._buf[k]
        return t if t.dim() == 0 or t.shape[0] == self._len[k] else t[: self._len[k]]

    def _adopt(self, k: str, value: Any):
        """`kinetics.N = tensor` (kinetics.py's attributes are plain tensors that users replace)"""
        self._touch()
        self._materialize()
        if not isinstance(value, torch.Tensor):
            value = torch.as_tensor(value)
        self._buf[k] = value
        self._len[k] = int(value.shape[0]) if value.dim() > 0 else 0
        self._phys_end = self._len["N"]
        self._cache = None
        self._stamp()

    def _stamp(self):
        self._stamps = {k: (t.data_ptr(), t._version) for k, t in self._buf.items()}

    def _written_from_outside(self) -> bool:
        return any(self._stamps.get(k) != (t.data_ptr(), t._version) for k, t in self._buf.items())

    def _phys(self, idxs) -> torch.Tensor:
        """buffer rows of logical cells"""
        dev = self._buf["N"].device
        idx = torch.as_tensor(idxs, dtype=torch.long, device=dev)
        if idx.dtype == torch.bool:
            idx = torch.nonzero(idx).flatten()
        return idx if self._rows is None else self._rows[idx].long()

    def _materialize(self):
        """compact the buffers in place so that row i holds cell i (the row map is increasing)"""
        if self._rows is None:
            return
        rows, n = self._rows.long(), self._len["N"]
        self._rows = None
        self._phys_end = n
        if n == 0:
            return
        moved = torch.nonzero(rows != torch.arange(n, device=rows.device)).flatten()
        if moved.numel() == 0:
            return
        start = int(moved[0])
        bufs = [self._buf[k] for k in _ALL_PARAMS] + ([self._cache] if self._cache is not None else [])
        row_bytes = max(t[0].numel() * t.element_size() for t in bufs)
        chunk = max(1, (256 << 20) // max(row_bytes, 1))
        for a in range(start, n, chunk):
            r = rows[a:a + chunk]
            for t in bufs:  # rows[i] >= i: writing rows [a, a+chunk) never destroys an unread row
                t[a:a + r.numel()] = t[r]
        self._stamp()

    def _reserve(self, new_rows: int):
        end = self._phys_end
        for k in _ALL_PARAMS:
            old = self._buf[k]
            new = torch.empty((new_rows,) + tuple(old.shape[1:]), dtype=old.dtype, device=old.device)
            new[:end] = old[:end]
            self._buf[k] = new
            del old
        if self._cache is not None:
            new = torch.zeros((new_rows, self._cache.shape[1]), dtype=torch.int32, device=self._cache.device)
            new[:end] = self._cache[:end]
            self._cache = new

    def _stale(self, phys: torch.Tensor | None):
        """the list-cache records of these buffer rows no longer describe their parameters"""
        if self._cache is not None:
            if phys is None:
                self._cache[:, 0] = 0
            else:
                self._cache[phys.to(self._cache.device), 0] = 0

    def cell_params(self, cell_idxs) -> dict[str, torch.Tensor]:
        """the nine parameter slices of the given cells (copies), without compacting the store"""
        self._touch()
        if not self._managed():
            return {k: self._view(k)[cell_idxs] for k in _ALL_PARAMS}
        phys = self._phys(cell_idxs)
        return {k: self._buf[k][phys] for k in _ALL_PARAMS}

    def set_cell_params_from_tensors(self, cell_idxs: list[int] | torch.Tensor, **tensors: torch.Tensor):
        """
        Install ready-made parameters for `cell_idxs` (what set_cell_params derives from
        proteomes, kinetics.py:571-625).  Keyword names are the attribute names.
        """
        self._touch()
        p_new = int(tensors["N"].shape[1])
        self.increase_max_proteins(max_n=p_new)
        managed = self._managed()
        phys = self._phys(cell_idxs) if managed else cell_idxs
        for k in _ALL_PARAMS:
            dst = self._buf[k] if managed else self._view(k)
            p = int(dst.shape[1])
            src = tensors[k].to(device=dst.device, dtype=dst.dtype)
            if p_new < p:  # pad like unset proteins
                pad_shape = list(src.shape)
                pad_shape[1] = p - p_new
                src = torch.cat([src, torch.zeros(pad_shape, dtype=src.dtype, device=src.device)], dim=1)
            dst[phys] = src
        if managed:
            self._stale(phys)
        self._stamp()

    def set_cell_params(self, cell_idxs: list[int], proteomes: list):
        """
        kinetics.py:521-625: translate proteomes (Genetics.translate_genomes) into the nine parameter
        tensors of the given cells.  One kernel launch (ms_set_cell_params) writes the cells' rows
        from the flattened domain tokens and the token maps; nothing of size (c,p,d,s) is built.
        """
        if len(cell_idxs) == 0:
            return
        if len(cell_idxs) != len(proteomes):
            raise ValueError("cell_idxs and proteomes differ in length")
        from magicsoup_b200._hostlib import lib as hostlib

        prot_off, dom_off, doms = hostlib().flatten_proteomes(list(proteomes))
        self.set_cell_params_flat(cell_idxs, prot_off, dom_off, doms)

    def set_cell_params_flat(self, cell_idxs: list[int], prot_off, dom_off, doms):
        """
        set_cell_params on flattened proteomes (int32 numpy arrays: protein offsets per cell, domain
        offsets per protein, 5 tokens per domain; `_lib.flatten_proteomes` / `translate_genomes_flat`).
        No counterpart in the reference, which pads nested lists into (c,p,d) tensors (kinetics.py:920-970).
        """
        if len(cell_idxs) == 0:
            return
        if len(prot_off) != len(cell_idxs) + 1:
            raise ValueError("cell_idxs and proteomes differ in length")
        self._touch()
        dev = self._cuda()
        if not self._managed() or self._buf["N"].device != dev or not all(t.is_contiguous() for t in self._buf.values()):
            self._materialize()
            for k in _ALL_PARAMS:  # user-assigned or CPU tensors: adopt normalised device copies
                t = self._view(k)
                self._adopt(k, (self._i32(t, dev) if k in _INT_PARAMS else self._f32(t, dev)).clone())
            if not self._managed():
                raise ValueError("the parameter tensors do not describe the same cells; cannot set cell parameters")
        p, s = int(self._buf["N"].shape[1]), int(self._buf["N"].shape[2])
        n_max = int((prot_off[1:] - prot_off[:-1]).max()) if len(prot_off) > 1 else 0
        if n_max > p:
            raise ValueError(f"a proteome has {n_max} proteins, the parameter tensors hold {p}: call increase_max_proteins first")
        doms2 = doms.reshape(-1, 5)
        tables = {"vmax_weights": self._f32(self.vmax_map.weights, dev), "km_weights": self._f32(self.km_map.weights, dev),
                  "signs": self._i32(self.sign_map.signs, dev), "hill_numbers": self._i32(self.hill_map.numbers, dev),
                  "reaction_M": self._i32(self.reaction_map.M, dev), "transport_M": self._i32(self.transport_map.M, dev),
                  "effector_M": self._i32(self.effector_map.M, dev)}
        if doms2.shape[0]:
            n_scalar = min(int(tables[k].shape[0]) for k in ("vmax_weights", "km_weights", "signs", "hill_numbers"))
            n_vector = min(int(tables[k].shape[0]) for k in ("reaction_M", "transport_M", "effector_M"))
            if doms2.min() < 0 or doms2[:, 1:4].max() >= n_scalar or doms2[:, 4].max() >= n_vector:
                raise IndexError("a domain token lies outside the token maps")
        for k in ("reaction_M", "transport_M", "effector_M"):
            if tables[k].dim() != 2 or int(tables[k].shape[1]) != s:
                raise ValueError(f"{k} has shape {tuple(tables[k].shape)}, expected (tokens, {s})")
        rows = self._phys(cell_idxs).to(torch.int32).contiguous()
        d_prot, d_dom, d_doms = (torch.from_numpy(a).to(dev) for a in (prot_off, dom_off, doms))
        energies = self._f32(self.mol_energies, dev)
        tt = nat.TokenTables(**{k: nat.ptr(v) for k, v in tables.items()})
        out = nat.KineticsParamsOut(**{k: nat.ptr(self._buf[k]) for k in _ALL_PARAMS}, p=p, s=s)
        nat.check(nat.lib().ms_set_cell_params(
            C.byref(tt), nat.ptr(energies), float(self.abs_temp), nat.ptr(rows), nat.ptr(d_prot), nat.ptr(d_dom),
            nat.ptr(d_doms), len(cell_idxs), C.byref(out), nat.stream_ptr(dev)))
        self._stale(rows.long())  # the kernel wrote the rows behind torch's back: no version stamp moved

    def unset_cell_params(self, cell_idxs: list[int]):
        """kinetics.py:627-644"""
        self._touch()
        managed = self._managed()
        phys = self._phys(cell_idxs) if managed else cell_idxs
        for k in _ALL_PARAMS:
            (self._buf[k] if managed else self._view(k))[phys] = 0
        if managed:
            self._stale(phys)
        self._stamp()

    def copy_cell_params(self, from_idxs: list[int], to_idxs: list[int]):
        """kinetics.py:646-665"""
        self._touch()
        managed = self._managed()
        src = self._phys(from_idxs) if managed else from_idxs
        dst = self._phys(to_idxs) if managed else to_idxs
        for k in _ALL_PARAMS:
            t = self._buf[k] if managed else self._view(k)
            t[dst] = t[src]
        if managed and self._cache is not None:
            self._cache[dst] = self._cache[src]  # same parameters, same lists
        self._stamp()

    def remove_cell_params(self, keep: torch.Tensor):
        """kinetics.py:667-686: order-preserving removal; only the row map is edited"""
        self._touch()
        if not self._managed():
            for k in _ALL_PARAMS:
                self._adopt(k, self._view(k)[keep])
            return
        dev = self._buf["N"].device
        keep = torch.as_tensor(keep, dtype=torch.bool, device=dev)
        n = self._len["N"]
        rows = self._rows if self._rows is not None else torch.arange(n, dtype=torch.int32, device=dev)
        self._rows = rows[keep].contiguous()
        for k in _ALL_PARAMS:
            self._len[k] = int(self._rows.numel())

    def increase_max_cells(self, by_n: int):
        """kinetics.py:688-703: by_n new all-zero cells at the end"""
        if by_n <= 0:
            return
        self._touch()
        if not self._managed():
            for k in _ALL_PARAMS:
                self._adopt(k, self._expand(self._view(k), dim=0, n=by_n))
            return
        n, end, cap = self._len["N"], self._phys_end, int(self._buf["N"].shape[0])
        if end + by_n > cap:
            if self._rows is not None and (end - n) >= by_n and 4 * (end - n) >= cap:
                self._materialize()  # enough freed rows inside: compact instead of growing
                end = n
            if end + by_n > cap:
                self._reserve(max(end + by_n, cap + cap // 2 + 16) if cap > 0 else by_n)
        for k in _ALL_PARAMS:
            self._buf[k][end:end + by_n].zero_()
            self._len[k] = n + by_n
        if self._cache is not None:
            self._cache[end:end + by_n].zero_()
        if self._rows is not None:
            new = torch.arange(end, end + by_n, dtype=torch.int32, device=self._rows.device)
            self._rows = torch.cat([self._rows, new])
        self._phys_end = end + by_n
        self._stamp()

    def increase_max_proteins(self, max_n: int):
        """kinetics.py:705-723: the protein dimension only ever grows"""
        n_prots = int(self._buf["N"].size(1))
        if max_n <= n_prots:
            return
        self._touch()
        if not self._managed():
            for k in _ALL_PARAMS:
                self._adopt(k, self._expand(self._view(k), dim=1, n=max_n - n_prots))
            return
        end = self._phys_end
        for k in _ALL_PARAMS:
            old = self._buf[k]
            shape = list(old.shape)
            shape[1] = max_n
            new = torch.zeros(shape, dtype=old.dtype, device=old.device)
            new[:end, :n_prots] = old[:end]
            self._buf[k] = new
            del old
        self._cache = None  # the record size depends on p
        self._stamp()

    def _expand(self, t: torch.Tensor, dim: int, n: int) -> torch.Tensor:
        shape = list(t.shape)
        shape[dim] = n
        return torch.cat([t, torch.zeros(shape, dtype=t.dtype, device=t.device)], dim=dim)

    def _zeros_i32_tensor(self, *args) -> torch.Tensor:
        return torch.zeros(*args, device=self.device, dtype=torch.int32)

    def _zeros_f32_tensor(self, *args) -> torch.Tensor:
        return torch.zeros(*args, device=self.device, dtype=torch.float32)


def _param_property(name: str) -> property:
    def getter(self) -> torch.Tensor:
        return self._view(name)

    def setter(self, value):
        self._adopt(name, value)

    return property(getter, setter, doc=f"Kinetics.{name} (kinetics.py:399-409): dense tensor of the current cells")


for _k in _ALL_PARAMS:
    setattr(Kinetics, _k, _param_property(_k))
