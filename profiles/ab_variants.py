"""
A/B builds of the kinetics kernel's tuning switches (MS_TUNE_* in csrc/kinetics.cu).

    python profiles/ab_variants.py name1:-DMS_TUNE_X=1,-DMS_TUNE_Y=4 name2: ...

builds variants/<name>.so (git-ignored, travels with gpurun) and prints the shell loop that times
each of them on the GPU box with the same bench command (the library file is swapped in place; the
source-hash stamp stays valid because the sources do not change).
"""
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from magicsoup_b200 import build as B  # noqa: E402

out = ROOT / "variants"
out.mkdir(exist_ok=True)
names = []
for spec in sys.argv[1:]:
    name, _, flags = spec.partition(":")
    cmd = [B.nvcc(), *B.NVCC_FLAGS, *[f for f in flags.split(",") if f], "-o", str(out / f"{name}.so"),
           *[str(B.CSRC / s) for s in B.SOURCES]]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.exit(res.stderr)
    names.append(name)
    print("built", name, flags)
print("for v in " + " ".join(names) + "; do cp variants/$v.so magicsoup_b200/libmagicsoup_b200.so; echo $v; "
      "python bench.py --steps 8 --warmup 5 --cells 262144 --map 512 --no-cpu-baseline --e2e-steps 1 "
      "| grep -o '\"kernel_ms\": [0-9.]*'; done")
