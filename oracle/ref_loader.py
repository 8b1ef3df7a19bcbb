"""
Loads the REAL reference (python/magicsoup from /root/reference, read-only) so that
the oracle can be validated against it and golden vectors can be generated.

TEST INFRASTRUCTURE.  Works only in the build container: /root/reference does not
exist on the GPU box, so nothing under `-m gpu`, smoke() or bench.py calls this.
The reference imports its Rust extension `magicsoup._lib` at module top
(kinetics.py:7, world.py:14); Rust cannot be built here (no cargo), and the hot
path never calls it, so a stub module is injected instead.
"""
from __future__ import annotations

import sys
import types
import warnings
from pathlib import Path

REFERENCE_ROOT = Path("/root/reference")


def available() -> bool:
    return (REFERENCE_ROOT / "python" / "magicsoup" / "kinetics.py").exists()


def load():
    """returns the reference's `magicsoup` package (imports it on first call)"""
    if "magicsoup" in sys.modules and getattr(sys.modules["magicsoup"], "_is_reference", False):
        return sys.modules["magicsoup"]
    if not available():
        raise RuntimeError("reference not present (only in the build container)")
    stub = types.ModuleType("magicsoup._lib")

    def _missing(name):
        def f(*_a, **_k):
            raise NotImplementedError(f"magicsoup._lib.{name}: Rust extension not built here")
        return f

    for name in ("translate_genomes", "get_coding_regions", "extract_domains",
                 "reverse_complement", "point_mutations", "recombinations", "get_neighbors",
                 "divide_cells_if_possible", "move_cells", "dist_1d", "free_moores_nghbhd",
                 "get_proteome"):
        setattr(stub, name, _missing(name))
    sys.modules["magicsoup._lib"] = stub
    sys.path.insert(0, str(REFERENCE_ROOT / "python"))
    try:
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            import magicsoup  # noqa
    finally:
        sys.path.remove(str(REFERENCE_ROOT / "python"))
    magicsoup._is_reference = True
    return magicsoup
