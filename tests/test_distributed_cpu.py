"""
World-size-2 (and 3) runs of the multi-GPU host logic on CPU with the gloo backend: halo
rows, all-reduced integer plane sums and early-exit flags, cell migration.  The kernels
are replaced by the CPU oracle inside the workers (tests/dist_worker.py); rank 0 asserts
bit-identity with the single-domain oracle.
"""
import os
import socket
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("nproc,cells,rows", [(2, 300, None), (3, 200, None), (1, 100, None), (3, 250, 50), (2, 200, 31),
                                              (2, 5, 8)])  # five cells: the global early exit fires across ranks
def test_slab_world_matches_single_domain_oracle(nproc, cells, rows):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           str(ROOT / "tests" / "dist_worker.py"), "--backend", "gloo", "--cells", str(cells)]
    if rows:  # slabs of unequal heights
        cmd += ["--rows", str(rows)]
    env = dict(os.environ, OMP_NUM_THREADS="2")
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=str(ROOT))
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "dist_worker ok" in res.stdout
