"""
Multi-GPU parity (needs >= 2 CUDA devices; skipped on a single-GPU box): N ranks over NCCL,
each owning a row slab and its cells, must reproduce the single-domain CPU oracle bit for
bit (tests/dist_worker.py).  Run by hand with
    gpurun --gpus 2 -- 'python -m pytest tests/test_gpu_distributed.py -m gpu -q'
"""
import os
import socket
import subprocess
import sys
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parents[1]


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("blocking,peer", [(False, True), (True, True), (False, False)],
                         ids=["deferred-peer", "blocking-peer", "deferred-nccl"])
@pytest.mark.parametrize("shape", [(300, 16, 24), (20000, 128, 128), (5, 4, 8)],
                         ids=["generic-stencil", "tiled-stencil", "five-cells"])
def test_nccl_slab_world_matches_oracle(shape, blocking, peer):
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    cells, rows, cols = shape
    nproc = 2 if cells < 100 else min(n, 4)  # five cells: the global early exit fires and steps are repeated
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           str(ROOT / "tests" / "dist_worker.py"), "--backend", "nccl", "--cells", str(cells),
           "--rows-per-rank", str(rows), "--cols", str(cols)]
    if blocking:  # SlabWorld's multi-rank default; otherwise the early-exit check trails the step (opt-in)
        cmd.append("--blocking")
    env = dict(os.environ)
    if not peer:  # flags, plane sums and halos through NCCL collectives instead of the peer-memory kernels (csrc/peer.cu)
        env["MS_NO_PEER"] = "1"
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=str(ROOT), env=env)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "dist_worker ok" in res.stdout


@pytest.mark.parametrize("blocking", [False, True], ids=["deferred", "blocking"])
def test_slab_world_equals_one_gpu_under_lifecycle_churn(blocking):
    """24 steps with kills, cells crossing slab boundaries and rebalanced slabs: the N-rank SlabWorld
    (peer-memory exchanges, deferred or blocking early-exit check) against one World holding everything"""
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={min(n, 4)}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           str(ROOT / "tests" / "dist_stress_worker.py")] + (["--blocking"] if blocking else [])
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=str(ROOT), env=dict(os.environ))
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "dist_stress ok" in res.stdout


def test_genome_driven_loop_on_slabs():
    """spawn / kill / divide / recombinate / mutate / move on every rank's own cells (tests/dist_genome_worker.py):
    each step of all ranks together equals the CPU oracle on the gathered world, bookkeeping stays consistent"""
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={min(n, 4)}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), str(ROOT / "tests" / "dist_genome_worker.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=str(ROOT), env=dict(os.environ))
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "dist_genome ok" in res.stdout
