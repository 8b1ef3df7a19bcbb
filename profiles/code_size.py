"""
SASS size of the kinetics kernels, split by the source file / function the line info attributes each
instruction to (run in the build container: needs cuobjdump and nvdisasm, no GPU):

    python profiles/code_size.py > profiles/r02_code_size.txt

The number that matters: the code one warp of kinetics_lanes_kernel<4> runs through when
it scans, builds and integrates a cell, against the 32 KB instruction cache level of an SM.
"""
from __future__ import annotations

import bisect
import collections
import re
import subprocess
import sys
import tempfile
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
CSRC = ROOT / "magicsoup_b200" / "csrc"
KERNELS = ["_ZN2ms21kinetics_lanes_kernelILi4EEEvNS_7KinArgsE", "_ZN2ms28kinetics_lanes_phased_kernelILi4EEEvNS_7KinArgsE",
           "_ZN2ms21kinetics_fused_kernelILi4EEEvNS_7KinArgsEPf"]


def functions_of(path: Path):
    out = []
    for i, line in enumerate(path.read_text().splitlines(), 1):
        m = re.match(r"^(?:template\s*<[^>]*>\s*)?(?:__device__|__global__|static|inline|MS_)[^;{]*?\b([A-Za-z_][A-Za-z_0-9]*)\s*\(", line)
        if m and not line.startswith(" "):
            out.append((i, m.group(1)))
    return out


def main():
    with tempfile.TemporaryDirectory() as tmp:
        subprocess.run(["cuobjdump", "-xelf", "all", str(ROOT / "magicsoup_b200" / "libmagicsoup_b200.so")], cwd=tmp,
                       check=True, capture_output=True)
        sass = subprocess.run(["nvdisasm", "-g", str(Path(tmp) / "kinetics.sm_100a.cubin")], check=True,
                              capture_output=True, text=True).stdout
    funcs = {p.name: functions_of(p) for p in CSRC.glob("*.cu*")}
    for kern in KERNELS:
        start = sass.find(f".text.{kern}:")
        if start < 0:
            continue
        end = sass.find("//--------------------- .text.", start + 10)
        cur, by = None, collections.Counter()
        for line in sass[start:end if end > 0 else None].splitlines():
            m = re.search(r'//## File "([^"]+)", line (\d+)', line)
            if m:
                cur = (Path(m.group(1)).name, int(m.group(2)))
            elif re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", line):
                by[cur] += 1
        total = sum(by.values())
        agg = collections.Counter()
        for (f, ln), n in by.items():
            names = funcs.get(f)
            if names:
                k = bisect.bisect_right([x[0] for x in names], ln) - 1
                agg[(f, names[k][1] if k >= 0 else "?")] += n
            else:
                agg[(f, "(library inline)")] += n
        print(f"{kern}: {total} SASS instructions = {total * 16 / 1024:.1f} KB")
        for (f, name), n in agg.most_common(14):
            print(f"  {n:6d}  {n * 16 / 1024:5.1f} KB  {f}: {name}")
        print("  (the division / reciprocal slow paths the compiler appends to the kernel are attributed to the\n"
              "   last source line before them, i.e. to the function defined just above the kernel)\n")


if __name__ == "__main__":
    sys.exit(main())
