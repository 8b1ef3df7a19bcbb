"""
bench.py's output contract (the driver parses ONE JSON line): the CPU arm (`--impl reference`) on a
tiny workload here, our arm on a small workload on the GPU.
"""
import json
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
             "vs_baseline", "dtype", "data", "config", "e2e", "gpu_launches"}


def _run(*args, timeout=600):
    res = subprocess.run([sys.executable, str(ROOT / "bench.py"), *args], capture_output=True, text=True,
                         timeout=timeout, cwd=str(ROOT))
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    lines = [ln for ln in res.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1, res.stdout[-2000:]
    return json.loads(lines[0])


def test_reference_arm_line():
    d = _run("--impl", "reference", "--steps", "1", "--warmup", "0", "--cells", "512", "--map", "32")
    assert d["impl"] == "reference" and BASE_KEYS <= set(d) and d["gpu_launches"] == 0
    assert d["unit"] == "cell-steps/s" and d["value"] > 0 and d["higher_is_better"] is True and d["vs_baseline"] is None
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and "oracle" in cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["config"]["proteins"] == 64 and d["config"]["signals"] == 32 and "workload" in d["config"]


@pytest.mark.gpu
def test_our_arm_line():
    d = _run("--steps", "4", "--warmup", "3", "--cells", "65536", "--map", "256", "--e2e-steps", "3", "--scan-steps", "2",
             "--no-extra")
    assert BASE_KEYS | {"roofline", "cpu_baseline", "clocks"} <= set(d) and "impl" not in d
    assert d["n_gpus"] == 1 and d["steps"] == 4 and d["warmup"] == 3 and d["dtype"] == "f32" and d["data"] == "synthetic"
    assert d["gpu_launches"] == 7 * 4 and d["value"] > 1e6
    assert abs(d["value"] - 65536 / (d["ms_per_step"] / 1e3)) < 1e-6 * d["value"]
    r = d["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    assert r["kernel_ms"] <= d["ms_per_step"] and 0 < r["scan_path"]["frac"] < 1.0
    e = d["e2e"]
    assert e["h2d_bytes_per_step"] == 65536 * 16 * 4 + 16 * 256 * 256 * 4 + 65536 * 2 * 4
    assert e["d2h_bytes_per_step"] == 65536 * 16 * 4 + 16 * 256 * 256 * 4 and 0 < e["value"] <= 1.05 * d["value"]
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1
    assert set(d["clocks"]) >= {"sm_mhz", "sm_max_mhz", "reasons"}
