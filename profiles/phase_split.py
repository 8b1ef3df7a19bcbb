"""
Per-phase instruction split of the kinetics kernel from an ncu report captured with
`--set full --import-source on` (the library is built with -lineinfo):

    python profiles/phase_split.py gpurun_out/<name>.ncu-rep [cells]

Aggregates "Instructions Executed" per CUDA source line (ncu --page source --print-source
cuda,sass) into the phases of csrc/kinetics.cu by the line ranges of its functions, and
prints a SASS opcode histogram weighted by executed instructions.
"""
import csv
import io
import re
import subprocess
import sys
from collections import Counter, defaultdict
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]


def function_ranges(src: Path):
    """(name, first line, last line) of the top-level device functions, by brace matching"""
    lines = src.read_text().splitlines()
    out = []
    i = 0
    pat = re.compile(r"^(?:template.*\n)?(?:__device__|__global__|static|template|extern).*?(\w+)\s*\(")
    while i < len(lines):
        ln = lines[i]
        if (ln.startswith("__device__") or ln.startswith("__global__") or ln.startswith("template") or
                ln.startswith("__host__")) and not ln.rstrip().endswith(";"):
            # find the opening brace
            j = i
            while j < len(lines) and "{" not in lines[j]:
                j += 1
            head = " ".join(lines[i:j + 1])
            m = re.search(r"(\w+)\s*\([^;]*$", head.split("{")[0])
            name = m.group(1) if m else f"line{i+1}"
            depth = 0
            k = j
            while k < len(lines):
                depth += lines[k].count("{") - lines[k].count("}")
                if depth == 0:
                    break
                k += 1
            out.append((name, i + 1, k + 1))
            i = k + 1
        else:
            i += 1
    return out


def main():
    rep = sys.argv[1]
    cells = int(sys.argv[2]) if len(sys.argv) > 2 else None
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source=cuda,sass"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    cur_file = None
    header = None
    per_line = defaultdict(lambda: [0, 0, 0, 0])  # inst, thread inst, samples, barrier samples
    opcodes = Counter()
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            cur_file = Path(r[1]).name
            continue
        if r[0] == "Line No":
            header = r
            iI = header.index("Instructions Executed")
            iT = header.index("Thread Instructions Executed")
            iS = header.index("# Samples")
            iB = header.index("stall_barrier")
            continue
        if header is None or len(r) < len(header):
            continue
        if r[0] != "":
            try:
                per_line[(cur_file, int(r[0]))][0] += int(r[iI])
                per_line[(cur_file, int(r[0]))][1] += int(r[iT])
                per_line[(cur_file, int(r[0]))][2] += int(r[iS])
                per_line[(cur_file, int(r[0]))][3] += int(r[iB])
            except ValueError:
                pass
        elif r[2] not in ("", "-", "..."):
            try:
                n = int(r[iI])
            except ValueError:
                continue
            toks = r[3].split()
            op = toks[1] if toks and toks[0].startswith("@") and len(toks) > 1 else (toks[0] if toks else "?")
            opcodes[op.split(".")[0]] += n
    total = sum(v[0] for v in per_line.values())
    total_s = sum(v[2] for v in per_line.values())
    print(f"total warp instructions {total}" + (f" = {total / cells:.0f} per cell" if cells else ""))
    csrc = ROOT / "magicsoup_b200" / "csrc"
    fr = {f.name: function_ranges(f) for f in list(csrc.glob("*.cu")) + list(csrc.glob("*.cuh"))}
    by_fn = defaultdict(lambda: [0, 0, 0, 0])
    for (f, ln), v in per_line.items():
        name = f
        if f in fr:
            name = next((n for n, a, b in fr[f] if a <= ln <= b), f + ":other")
        for i in range(4):
            by_fn[name][i] += v[i]
    print(f"{'function':28s} {'inst%':>6s} {'inst/cell':>9s} {'lanes':>6s} {'samples%':>8s} {'barrier%':>8s}")
    for name, v in sorted(by_fn.items(), key=lambda kv: -kv[1][0]):
        per_cell = f"{v[0] / cells:9.0f}" if cells else " " * 9
        print(f"{name:28s} {100 * v[0] / total:6.1f} {per_cell} {v[1] / max(v[0], 1):6.1f} "
              f"{100 * v[2] / max(total_s, 1):8.1f} {100 * v[3] / max(v[2], 1):8.1f}")
    print("\nhottest source lines")
    for (f, ln), v in sorted(per_line.items(), key=lambda kv: -kv[1][0])[:40]:
        print(f"{f}:{ln:5d} {100 * v[0] / total:5.1f}% inst {v[1] / max(v[0], 1):5.1f} lanes {100 * v[2] / max(total_s, 1):5.1f}% samples")
    print("\nSASS opcode histogram (executed warp instructions)")
    tot_op = sum(opcodes.values())
    for op, n in opcodes.most_common(30):
        print(f"{op:10s} {100 * n / tot_op:5.1f}%")


if __name__ == "__main__":
    main()
