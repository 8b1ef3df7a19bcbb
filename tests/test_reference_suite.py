"""
The reference's OWN test files, run unmodified against the reference's Python package with
`magicsoup_b200._lib` standing in for its Rust extension `magicsoup._lib` (rust/lib.rs:18-205;
no Rust toolchain exists here, so without a stand-in 23 of these tests cannot even import).
This pins the C++ module to everything the reference's tests ask of the Rust one: genome
translation, mutations, neighbour / division / movement searches, `get_proteome`.

Needs `/root/reference` (this container); skipped where it is absent (the GPU box).  The fast
suite (89 tests, ~10 s) always runs; the slow one (11 tests, ~2 min) with MS_REFERENCE_SLOW=1
(last run: profiles/r02_reference_tests_with_our_lib.log).
"""
import os
import re
import subprocess
import sys
from pathlib import Path

import pytest

REFERENCE = Path("/root/reference")
ROOT = Path(__file__).resolve().parents[1]

PLUGIN = f"""
import sys
sys.path.insert(0, {str(ROOT)!r})
sys.path.insert(0, {str(REFERENCE / "python")!r})
import magicsoup_b200._hostlib as hostlib
sys.modules["magicsoup._lib"] = hostlib.lib()   # before the reference package is imported
"""


def _run(suite: str, tmp_path: Path, timeout: int) -> tuple[int, str]:
    (tmp_path / "ms_b200_lib_plugin.py").write_text(PLUGIN)
    env = dict(os.environ, PYTHONPATH=f"{tmp_path}:{REFERENCE}")
    cmd = [sys.executable, "-m", "pytest", str(REFERENCE / "tests" / suite), "-p", "ms_b200_lib_plugin",
           "-p", "no:cacheprovider", "-q", f"--rootdir={tmp_path}"]
    res = subprocess.run(cmd, cwd=tmp_path, env=env, capture_output=True, text=True, timeout=timeout)
    return res.returncode, res.stdout[-4000:] + res.stderr[-2000:]


@pytest.mark.skipif(not (REFERENCE / "tests" / "fast").is_dir(), reason="the reference is not on this machine")
def test_reference_fast_suite_passes_with_our_lib(tmp_path):
    rc, out = _run("fast", tmp_path, timeout=900)
    assert rc == 0, out
    m = re.search(r"(\d+) passed", out)
    assert m and int(m.group(1)) >= 89 and "failed" not in out, out


@pytest.mark.skipif(not (REFERENCE / "tests" / "slow").is_dir() or not os.environ.get("MS_REFERENCE_SLOW"),
                    reason="set MS_REFERENCE_SLOW=1 (about two minutes); needs the reference")
def test_reference_slow_suite_passes_with_our_lib(tmp_path):
    rc, out = _run("slow", tmp_path, timeout=3000)
    assert rc == 0, out
    m = re.search(r"(\d+) passed", out)
    assert m and int(m.group(1)) >= 11 and "failed" not in out, out
