"""
magicsoup_b200: B200-native (sm_100a) implementation of magicsoup's per-step world math behind the
reference's own `Kinetics` / `World` API (`Kinetics.integrate_signals`, `World.enzymatic_activity`,
`World.diffuse_molecules`), with the genome side around it (`Genetics`, mutations, `_lib` in C++,
`set_cell_params` as a kernel) and `distributed.SlabWorld` for several GPUs.
"""
from magicsoup_b200.kinetics import Kinetics  # noqa: F401
from magicsoup_b200.world import World, Molecule, Chemistry  # noqa: F401

__version__ = "0.1.0"
from magicsoup_b200.genetics import Genetics  # noqa: F401,E402
from magicsoup_b200.mutations import point_mutations, recombinations  # noqa: F401,E402
from magicsoup_b200.util import random_genome, randstr, variants, codons  # noqa: F401,E402
