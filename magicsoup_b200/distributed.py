"""
Multi-GPU world: one process per GPU (`torch.distributed`, NCCL over NVLink on the B200
box, gloo in the CPU tests).  The reference is single-device; this is the
sharding north_star asks for:

  * the molecule map (m, H, W) is split into `world_size` row slabs (as even as H allows by
    default; any heights >= 1 that add up to H can be given, and `rebalance()` moves the slab
    boundaries so that every rank carries the same share of the step's memory traffic);
  * every cell lives on the rank that owns its pixel's slab, together with its slice of the
    nine kinetics parameter tensors, so the gather/scatter of `enzymatic_activity`
    (world.py:619-625) and the permeation (world.py:651-665) stay local;
  * `diffuse_molecules` exchanges ONE edge row per neighbour (m*W floats each way, circular:
    the slab above rank 0 is the last rank's) and all-reduces the 4*m exact integer plane sums
    of the mass correction (world.py:640-648); because those sums are integers the result is
    bit-identical to the single-device result on the whole map;
  * `enzymatic_activity` all-reduces (max) the 24 early-exit flags so that every rank takes the
    same decision about the reference's GLOBAL early exit (kinetics.py:846-847).
  * With NCCL (i.e. GPUs on one NVLink domain) those three small exchanges do not go through
    collective calls: every rank's buffer is mapped into every process
    (torch.distributed._symmetric_memory) and short kernels store a rank's contribution into the
    peers' buffers, publish a sequence number and reduce what has arrived (csrc/peer.cu,
    `peer_exchange`); NCCL remains the fallback, and what moves the cells (`migrate_cells`).

Cell indices are local to the rank; the global index of local cell i is
`global_cell_offset() + i` (ranks in order).  `kill_cells`, `add_cells_from_params` and
`divide_cells_at` act on the local partition and keep all per-cell tensors and the parameter
slices consistent; `migrate_cells` moves cells (with their parameters) to the rank that owns
their new pixel, e.g. after a division or move across a slab boundary.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Any, Sequence

import torch
import torch.distributed as dist

from magicsoup_b200 import _native as nat
from magicsoup_b200.kinetics import _ALL_PARAMS
from magicsoup_b200.world import World, _KERNEL_HOOKS


def even_slab_heights(n_rows: int, world_size: int) -> list[int]:
    """row slabs as even as n_rows allows (the first n_rows % world_size slabs get one row more)"""
    base, extra = divmod(n_rows, world_size)
    return [base + (1 if r < extra else 0) for r in range(world_size)]


def balanced_slab_heights(cost_per_row: torch.Tensor, world_size: int) -> list[int]:
    """
    Slab heights (>= 1 row each) whose summed per-row cost is as equal as whole rows allow:
    boundary k sits at the first row where the running cost reaches k / world_size of the total.
    """
    H = int(cost_per_row.numel())
    cum = torch.cumsum(cost_per_row.to(torch.float64).cpu(), 0)
    total = float(cum[-1])
    if total <= 0:
        return even_slab_heights(H, world_size)
    bounds = [0]
    for k in range(1, world_size):
        b = int(torch.searchsorted(cum, torch.tensor(total * k / world_size, dtype=torch.float64)).item()) + 1
        b = max(b, bounds[-1] + 1)            # at least one row per slab ...
        b = min(b, H - (world_size - k))      # ... also for the slabs that follow
        bounds.append(b)
    bounds.append(H)
    return [bounds[k + 1] - bounds[k] for k in range(world_size)]


class SlabWorld(World):
    """`World` of one row slab; constructor takes the GLOBAL map size (rows, cols)"""

    def __init__(
        self,
        chemistry: Any,
        global_map_size: int | tuple[int, int],
        group: dist.ProcessGroup | None = None,
        abs_temp: float = 310.0,
        mol_map_init: str = "randn",
        device: str = "cpu",
        slab_heights: Sequence[int] | None = None,
    ):
        if not dist.is_initialized():
            raise RuntimeError("SlabWorld needs an initialised torch.distributed process group")
        self.group = group
        self.rank = dist.get_rank(group)
        self.world_size = dist.get_world_size(group)
        H, W = (global_map_size, global_map_size) if isinstance(global_map_size, int) else global_map_size
        if H < self.world_size:
            raise ValueError(f"{H} map rows cannot be split over {self.world_size} ranks")
        self.global_map_size = (H, W)
        self._set_slabs(list(slab_heights) if slab_heights is not None else even_slab_heights(H, self.world_size))
        super().__init__(chemistry=chemistry, map_size=(self.slab_rows, W), abs_temp=abs_temp,
                         mol_map_init=mol_map_init, device=device)

    def __getstate__(self):
        state = super().__getstate__()
        state["group"] = None
        state.pop("_peer_ctx", None)  # mappings of other processes' memory do not travel
        return state

    def _settle(self):
        pending = self._pending is not None
        super()._settle()
        if pending and self.peer_status():
            raise RuntimeError("an exchange over peer memory gave up waiting for a rank (see csrc/peer.cu)")

    def _map_shape(self):
        # the slab's rows end where other ranks begin: the Moore-neighbourhood searches of `_lib`
        # (divide_cells, move_cells, get_neighbors, recombinate_cells) stay inside the slab; columns wrap
        return (int(self.slab_rows), int(self.global_map_size[1]), self.world_size == 1)

    def _set_slabs(self, heights: list[int]):
        H = self.global_map_size[0]
        if len(heights) != self.world_size or any(int(h) < 1 for h in heights) or sum(heights) != H:
            raise ValueError(f"slab heights {heights} must be {self.world_size} positive numbers adding up to {H}")
        self.slab_heights = [int(h) for h in heights]
        self.row_starts = [sum(self.slab_heights[:r]) for r in range(self.world_size + 1)]
        self.slab_rows = self.slab_heights[self.rank]
        self.row0 = self.row_starts[self.rank]  # first global row of this slab

    # ------------------------------------------------------------------ global bookkeeping
    def global_n_cells(self) -> list[int]:
        """number of cells on every rank"""
        counts = torch.zeros(self.world_size, dtype=torch.int64, device=self._comm_device())
        counts[self.rank] = self.n_cells
        dist.all_reduce(counts, op=dist.ReduceOp.SUM, group=self.group)
        return [int(x) for x in counts.tolist()]

    def global_cell_offset(self) -> int:
        """global index of this rank's local cell 0 (ranks in order)"""
        return sum(self.global_n_cells()[: self.rank])

    def global_positions(self) -> torch.Tensor:
        """(c,2) pixel coordinates on the whole map"""
        pos = self.cell_positions.clone()
        pos[:, 0] += self.row0
        return pos

    def owner_of_rows(self, rows: torch.Tensor) -> torch.Tensor:
        """rank that owns each global row"""
        inner = torch.tensor(self.row_starts[1:-1], dtype=rows.dtype, device=rows.device)
        return torch.bucketize(rows, inner, right=True)

    def _comm_device(self) -> torch.device:
        backend = dist.get_backend(self.group)
        if backend == "nccl":
            return self._cuda()
        return torch.device("cpu")

    # ------------------------------------------------------------------ exchange steps
    def _all_ranks_empty(self) -> bool:
        # A rank without cells still takes part in the flag reduction of the other ranks' cells
        # (asking all ranks for their counts would cost one more collective and sync per step).
        return False

    def _reduce_flags(self, flags) -> list[int]:
        """MAX over ranks; flags from the kernel are reduced where they are (one read-back)"""
        if self.world_size == 1:
            return super()._reduce_flags(flags)
        comm = self._comm_device()
        if isinstance(flags, torch.Tensor) and flags.device == comm:
            t = flags
        else:
            t = torch.as_tensor(flags, dtype=torch.int32).to(comm)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        return t.tolist()

    # ------------------------------------------------------------------ exchanges over peer memory
    peer_exchange = True  # False (or MS_NO_PEER=1): NCCL collectives for the flags, the plane sums and the halos

    def _peer(self):
        """
        The rank's exchange buffer, mapped by all ranks (csrc/peer.cu), or None: one rank, not NCCL,
        switched off, more than 16 ranks, or symmetric memory not available here (then NCCL is used).
        Created collectively at the first step.
        """
        ctx = self.__dict__.get("_peer_ctx", False)
        if ctx is not False:
            if ctx is not None and ctx["shape"] != (self.n_molecules, self.global_map_size[1]):
                raise RuntimeError("the map's planes or columns changed under a SlabWorld with peer exchange")
            return ctx
        ctx = None
        if (self.peer_exchange and not os.environ.get("MS_NO_PEER") and 1 < self.world_size <= 16
                and dist.get_backend(self.group) == "nccl"):
            try:
                import torch.distributed._symmetric_memory as symm_mem

                dev = self._cuda()
                m, W = self.n_molecules, self.global_map_size[1]
                nbytes = int(nat.lib().ms_peer_buffer_bytes(self.world_size, m, W))
                buf = symm_mem.empty(nbytes, dtype=torch.uint8, device=dev)
                buf.zero_()
                handle = symm_mem.rendezvous(buf, self.group if self.group is not None else dist.group.WORLD)
                torch.cuda.synchronize(dev)
                dist.barrier(group=self.group)  # every buffer is zero before anybody stores into it
                ptrs = (C.c_uint64 * self.world_size)(*[int(p) for p in handle.buffer_ptrs])
                ctx = {"buf": buf, "handle": handle, "ptrs": ptrs, "seq_reduce": 0, "seq_halo": 0, "shape": (m, W),
                       "status": torch.zeros(1, dtype=torch.int32, device=dev)}
            except Exception as err:  # no symmetric memory on this system: the collectives do the same job
                import warnings

                warnings.warn(f"SlabWorld: peer exchange not available ({err!r}); using NCCL collectives")
                ctx = None
        # all ranks must agree (a rank that fell back would wait in a collective for ranks that did not)
        ok = torch.tensor([1 if ctx is not None else 0], dtype=torch.int32, device=self._comm_device())
        if self.world_size > 1:
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)
        if int(ok.item()) == 0:
            ctx = None
        self.__dict__["_peer_ctx"] = ctx
        return ctx

    def _peer_allreduce(self, ctx, data: torch.Tensor, op: int):
        ctx["seq_reduce"] += 1
        dev = data.device
        nat.check(nat.lib().ms_peer_allreduce(ctx["ptrs"], self.world_size, self.rank, ctx["seq_reduce"], nat.ptr(data),
                                              int(data.numel()), op, nat.ptr(ctx["status"]), nat.stream_ptr(dev)))

    def peer_status(self) -> int:
        """non-zero if an exchange over peer memory gave up waiting for a rank (synchronises)"""
        ctx = self.__dict__.get("_peer_ctx")
        return int(ctx["status"].item()) if ctx else 0

    def _reduce_flags_device(self, flags: torch.Tensor) -> torch.Tensor:
        """MAX over ranks in place, enqueued like any other step of the launch sequence (no host round trip)"""
        if self.world_size > 1:
            ctx = self._peer()
            if ctx is not None:
                self._peer_allreduce(ctx, flags, 1)
            else:
                dist.all_reduce(flags, op=dist.ReduceOp.MAX, group=self.group)
        return flags

    # On several ranks the deferred early-exit check (World._settle) is opt-in: when the verdict is
    # "repeat" (tiny or fully equilibrated populations) the step is redone with collectives at the
    # rank's next touch of its state, so every rank has to touch its state at the same points of
    # the program -- true for a plain SPMD loop (bench.py sets it), not for a program where, say,
    # only rank 0 reads cell_molecules for logging between two steps.
    deferred_settle = False

    def _can_defer(self) -> bool:
        cls = type(self)
        hooks_own = all(getattr(cls, h) is getattr(World, h) for h in _KERNEL_HOOKS)
        if self.world_size == 1:  # one rank: nothing collective in a repeat, World's default applies
            wanted = self.__dict__.get("deferred_settle", World.deferred_settle)
            return bool(wanted) and hooks_own
        # the collectives of the deferred sequence run on the step's device: NCCL only
        return bool(self.deferred_settle) and hooks_own and dist.get_backend(self.group) == "nccl"

    def _reduce_sums(self, sums: torch.Tensor) -> torch.Tensor:
        if self.world_size > 1:
            ctx = self._peer() if sums.is_cuda else None
            if ctx is not None:
                self._peer_allreduce(ctx, sums, 0)
            elif sums.device != self._comm_device():
                moved = sums.to(self._comm_device())
                dist.all_reduce(moved, op=dist.ReduceOp.SUM, group=self.group)
                sums.copy_(moved)
            else:
                dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=self.group)
        return sums

    def _n_pixels_total(self, H: int, W: int) -> int:
        return self.global_map_size[0] * self.global_map_size[1]

    def _exchange_halos(self, mm: torch.Tensor):
        """
        One edge row per neighbour: my first row becomes the BOTTOM halo of the rank above,
        my last row the TOP halo of the rank below (circular).  Returns (top, bottom) halos
        of shape (m, W) on mm's device.
        """
        if self.world_size == 1:
            return None, None
        ctx = self._peer() if mm.is_cuda else None
        if ctx is not None:
            ctx["seq_halo"] += 1
            m, H, W = mm.shape
            nat.check(nat.lib().ms_peer_halo(ctx["ptrs"], self.world_size, self.rank, ctx["seq_halo"], nat.ptr(mm), m, H, W,
                                             nat.ptr(ctx["status"]), nat.stream_ptr(mm.device)))
            halos = []
            for bottom in (0, 1):
                off = int(nat.lib().ms_peer_halo_offset(self.world_size, m, W, ctx["seq_halo"] & 1, bottom))
                halos.append(ctx["buf"][off:off + m * W * 4].view(torch.float32).view(m, W))
            return halos[0], halos[1]
        comm = self._comm_device()
        first = mm[:, 0, :].contiguous().to(comm)
        last = mm[:, -1, :].contiguous().to(comm)
        top = torch.empty_like(first)
        bottom = torch.empty_like(first)
        prev = (self.rank - 1) % self.world_size
        nxt = (self.rank + 1) % self.world_size
        if self.group is not None:
            prev, nxt = dist.get_global_rank(self.group, prev), dist.get_global_rank(self.group, nxt)
        ops = [
            dist.P2POp(dist.isend, last, nxt, group=self.group),
            dist.P2POp(dist.irecv, top, prev, group=self.group),
            dist.P2POp(dist.isend, first, prev, group=self.group),
            dist.P2POp(dist.irecv, bottom, nxt, group=self.group),
        ]
        for req in dist.batch_isend_irecv(ops):
            req.wait()
        return top.to(mm.device), bottom.to(mm.device)

    # ------------------------------------------------------------------ moving cells between ranks
    def export_cells(self, cell_idxs: list[int]) -> dict[str, torch.Tensor]:
        """per-cell state + parameter slices of the given local cells, positions global"""
        self._settle()
        idx = torch.as_tensor(cell_idxs, dtype=torch.long, device=self.cell_positions.device)
        out = {
            "positions": self.global_positions()[idx],
            "molecules": self.cell_molecules[idx],
            "lifetimes": self.cell_lifetimes[idx],
            "divisions": self.cell_divisions[idx],
        }
        out.update(self.kinetics.cell_params(idx))
        return out

    def import_cells(self, payload: dict[str, torch.Tensor]) -> list[int]:
        """append cells exported by another rank; their pixels must lie in this slab"""
        self._settle()
        n = int(payload["positions"].shape[0])
        if n == 0:
            return []
        pos = payload["positions"].clone()
        pos[:, 0] -= self.row0
        if bool(((pos[:, 0] < 0) | (pos[:, 0] >= self.slab_rows)).any()):
            raise ValueError("imported cells do not live in this rank's slab")
        new = self.add_cells_from_params(pos, cell_molecules=payload["molecules"],
                                         **{k: payload[k] for k in _ALL_PARAMS})
        self.cell_lifetimes[new] = payload["lifetimes"].to(self.cell_lifetimes.device)
        self.cell_divisions[new] = payload["divisions"].to(self.cell_divisions.device)
        return new

    def migrate_cells(self) -> int:
        """
        Move every local cell whose (local) row left the slab -- after the caller changed
        `cell_positions`, e.g. a division or move across the slab boundary -- to the rank
        that owns its pixel.  Collective: every rank must call it.  Returns the number of
        cells received.  Local indices are compacted exactly like `kill_cells` does
        (world.py:527-540); received cells are appended (sender ranks in order).

        One small all-to-all of per-destination counts, then the per-cell tensors and parameter
        slices of the leaving cells go rank to rank as packed device buffers (`batch_isend_irecv`,
        only between pairs that exchange cells): no pickling, no per-cell host work.
        """
        self._settle()
        ws, H = self.world_size, self.global_map_size[0]
        comm = self._comm_device()
        gpos = self.global_positions()
        gpos[:, 0] = gpos[:, 0] % H  # the map is a torus
        owner = self.owner_of_rows(gpos[:, 0].long())
        leaving = torch.nonzero(owner != self.rank).flatten()
        order = leaving[torch.argsort(owner[leaving], stable=True)]  # grouped by destination, cell order kept
        # per destination: (cells to send, my protein dimension); the protein dimension only ever grows
        # per rank (kinetics.py:705-723), so the receiver sizes its buffers by the sender's
        p_mine = int(self.kinetics._buf["N"].shape[1])
        send_meta = torch.stack([torch.bincount(owner[order], minlength=ws).to(torch.int64).cpu(),
                                 torch.full((ws,), p_mine, dtype=torch.int64)], dim=1).contiguous()
        recv_meta = send_meta.clone()
        if ws > 1:
            sc = send_meta.to(comm)
            rc = torch.empty_like(sc)
            dist.all_to_all_single(rc, sc, group=self.group)
            recv_meta = rc.cpu()
        send_counts = send_meta[:, 0].tolist()
        recv_counts, recv_p = recv_meta[:, 0].tolist(), recv_meta[:, 1].tolist()

        payload = self.export_cells(order) if order.numel() else None
        if payload is not None:
            payload["positions"][:, 0] = payload["positions"][:, 0] % H
        names = ["positions", "molecules", "lifetimes", "divisions", *_ALL_PARAMS]
        ops, inbox, keep_alive = [], {}, []
        offset = 0
        for dst in range(ws):
            n = send_counts[dst]
            if n and dst != self.rank:
                for k in names:
                    buf = payload[k][offset:offset + n].contiguous().to(comm)
                    keep_alive.append(buf)
                    ops.append(dist.P2POp(dist.isend, buf, self._global_rank(dst), group=self.group))
            offset += n
        for src in range(ws):
            n = recv_counts[src]
            if n and src != self.rank:
                inbox[src] = {k: torch.empty((n,) + shape, dtype=dt, device=comm)
                              for k, (shape, dt) in self._payload_templates(recv_p[src]).items()}
                for k in names:
                    ops.append(dist.P2POp(dist.irecv, inbox[src][k], self._global_rank(src), group=self.group))
        if ops:
            for req in dist.batch_isend_irecv(ops):
                req.wait()
        if order.numel():
            self._drop_cells(order)
        received = 0
        for src in sorted(inbox):
            received += len(self.import_cells(inbox[src]))
        return received

    # ------------------------------------------------------------------ load balance
    def row_costs(self) -> torch.Tensor:
        """
        Algorithmic bytes per step and global map row: B_kin = 20ps + 16p + 8s + 8 per
        cell on the row plus the row's share of the diffusion, 16 m W + (16 m + 8) per cell.  (H,) float64
        on the host, the same on every rank.  Collective.
        """
        self._settle()
        H, W = self.global_map_size
        comm = self._comm_device()
        hist = torch.zeros(H, dtype=torch.int64, device=comm)
        if self.n_cells > 0:
            rows = (self.cell_positions[:, 0].long() + self.row0) % H
            hist += torch.bincount(rows, minlength=H).to(comm)
        p_max = torch.tensor([int(self.kinetics._buf["N"].shape[1]) if self.kinetics._buf["N"].dim() == 3 else 0],
                             dtype=torch.int64, device=comm)
        if self.world_size > 1:
            dist.all_reduce(hist, op=dist.ReduceOp.SUM, group=self.group)
            dist.all_reduce(p_max, op=dist.ReduceOp.MAX, group=self.group)
        m, p = self.n_molecules, int(p_max.item())
        s = 2 * m
        per_cell = 20 * p * s + 16 * p + 8 * s + 8 + 16 * m + 8
        return hist.cpu().to(torch.float64) * per_cell + 16.0 * m * W

    def rebalance(self, tolerance: float = 0.05) -> list[int]:
        """
        Move the slab boundaries so that every rank carries the same share of the step's memory
        traffic (`row_costs`), then move map rows and cells to their new owners.  Nothing happens
        while the fullest slab is within `tolerance` of the mean.  Collective; returns the slab
        heights in effect afterwards.  The step's results do not depend on where the boundaries are.
        """
        cost = self.row_costs()
        loads = [float(cost[self.row_starts[r]:self.row_starts[r + 1]].sum()) for r in range(self.world_size)]
        mean = sum(loads) / self.world_size
        if self.world_size == 1 or mean <= 0 or max(loads) <= (1.0 + tolerance) * mean:
            return list(self.slab_heights)
        heights = balanced_slab_heights(cost, self.world_size)
        new_loads = []
        at = 0
        for h in heights:
            new_loads.append(float(cost[at:at + h].sum()))
            at += h
        if max(new_loads) >= max(loads):  # whole rows cannot do better
            return list(self.slab_heights)
        self.set_slab_heights(heights)
        return list(self.slab_heights)

    def set_slab_heights(self, heights: Sequence[int]):
        """re-cut the map into slabs of the given heights; map rows and cells move to their new owners. Collective."""
        self._settle()
        old_starts = list(self.row_starts)
        old_r0 = self.row0
        mm_old = self.molecule_map
        self._set_slabs([int(h) for h in heights])
        if self.row_starts == old_starts:
            return
        ws, (H, W), m = self.world_size, self.global_map_size, self.n_molecules
        comm = self._comm_device()
        new_r0, new_r1 = self.row0, self.row0 + self.slab_rows
        mm_new = torch.empty(m, self.slab_rows, W, dtype=mm_old.dtype, device=mm_old.device)
        ops, landing, keep_alive = [], [], []
        for q in range(ws):
            # rows I held that rank q holds now / rows rank q held that I hold now
            a, b = max(old_starts[self.rank], self.row_starts[q]), min(old_starts[self.rank + 1], self.row_starts[q + 1])
            if a < b:
                block = mm_old[:, a - old_r0:b - old_r0, :]
                if q == self.rank:
                    mm_new[:, a - new_r0:b - new_r0, :] = block
                else:
                    buf = block.contiguous().to(comm)
                    keep_alive.append(buf)
                    ops.append(dist.P2POp(dist.isend, buf, self._global_rank(q), group=self.group))
            a, b = max(new_r0, old_starts[q]), min(new_r1, old_starts[q + 1])
            if a < b and q != self.rank:
                buf = torch.empty(m, b - a, W, dtype=mm_old.dtype, device=comm)
                landing.append((a - new_r0, b - new_r0, buf))
                ops.append(dist.P2POp(dist.irecv, buf, self._global_rank(q), group=self.group))
        if ops:
            for req in dist.batch_isend_irecv(ops):
                req.wait()
        for a, b, buf in landing:
            mm_new[:, a:b, :] = buf.to(mm_new.device)
        self.molecule_map = mm_new
        self.map_size = (self.slab_rows, W)
        # cells: local rows relative to the new first row; the ones outside move to their owners
        self.cell_positions = self.cell_positions.clone()
        self.cell_positions[:, 0] += old_r0 - new_r0
        rows = self.cell_positions[:, 0].long()
        stay = (rows >= 0) & (rows < self.slab_rows)
        self.cell_map = torch.zeros(self.slab_rows, W, dtype=torch.bool, device=self.cell_map.device)
        self.cell_map[rows[stay], self.cell_positions[stay, 1].long()] = True
        self.migrate_cells()

    def _global_rank(self, r: int) -> int:
        return dist.get_global_rank(self.group, r) if self.group is not None else r

    def _payload_templates(self, p: int) -> dict[str, tuple[tuple[int, ...], torch.dtype]]:
        """trailing shape and dtype of every per-cell tensor that travels with a cell (p: the sender's proteins)"""
        s = 2 * self.n_molecules
        out = {"positions": ((2,), torch.int32), "molecules": ((self.n_molecules,), torch.float32),
               "lifetimes": ((), torch.int32), "divisions": ((), torch.int32)}
        for k in _ALL_PARAMS:
            out[k] = ((p, s) if k in ("N", "Nf", "Nb", "A", "Kmr") else (p,),
                      torch.int32 if k in ("N", "Nf", "Nb", "A") else torch.float32)
        return out

    def _drop_cells(self, cell_idxs):
        """kill_cells' compaction (world.py:527-540) without the spill and cell_map update"""
        self._settle()
        idx = torch.as_tensor(cell_idxs, dtype=torch.long, device=self.cell_lifetimes.device)
        keep = torch.ones(self.cell_lifetimes.size(0), dtype=torch.bool, device=self.cell_lifetimes.device)
        keep[idx] = False
        self.cell_lifetimes = self.cell_lifetimes[keep]
        self.cell_positions = self.cell_positions[keep]
        self.cell_divisions = self.cell_divisions[keep]
        self.cell_molecules = self.cell_molecules[keep]
        self.kinetics.remove_cell_params(keep=keep)
        gone = set(idx.tolist())
        self.cell_genomes = [g for i, g in enumerate(self.cell_genomes) if i not in gone]
        self.cell_labels = [g for i, g in enumerate(self.cell_labels) if i not in gone]
        self.n_cells -= len(gone)
