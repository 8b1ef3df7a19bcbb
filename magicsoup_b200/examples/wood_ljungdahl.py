"""
The reference's Wood-Ljungdahl example chemistry (python/magicsoup/examples/wood_ljungdahl.py:17-63):
14 molecule species, 6 reactions.  It is the chemistry of the reference's own benchmarks
(performance/check.py:34, performance/run_simulation.py:33), so bench.py's `reference_workloads`
and the tests of BASELINE.json configs[1] use it.  Same names, energies (J/mol), order and
stoichiometry; everything else about a molecule is the default (diffusivity 0.1, permeability 0),
except CO2, which diffuses and permeates freely.
"""
from magicsoup_b200.world import Chemistry, Molecule

_KJ = 1e3
# name -> (energy in kJ/mol, extra attributes)
_SPECIES = {
    "NADPH": (200.0, {}),
    "NADP": (100.0, {}),
    "ATP": (100.0, {}),
    "ADP": (70.0, {}),
    "methyl-FH4": (360.0, {}),
    "methylen-FH4": (300.0, {}),
    "formyl-FH4": (240.0, {}),
    "FH4": (200.0, {}),
    "formiat": (20.0, {}),
    "CO2": (10.0, {"diffusivity": 1.0, "permeability": 1.0}),
    "Ni-ACS": (200.0, {}),
    "methyl-Ni-ACS": (300.0, {}),
    "HS-CoA": (200.0, {}),
    "acetyl-CoA": (260.0, {}),
}
# substrates -> products, by name: the methyl branch (formate dehydrogenase, formate-FH4 ligase,
# cyclohydrolase, reductase) and the carbonyl branch (methyltransferase, CO dehydrogenase / ACS)
_STEPS = (
    ("CO2 NADPH", "formiat NADP"),
    ("formiat FH4 ATP", "formyl-FH4 ADP"),
    ("formyl-FH4 NADPH", "methylen-FH4 NADP"),
    ("methylen-FH4 NADPH", "methyl-FH4 NADP"),
    ("methyl-FH4 Ni-ACS", "FH4 methyl-Ni-ACS"),
    ("methyl-Ni-ACS CO2 HS-CoA", "Ni-ACS acetyl-CoA"),
)

_BY_NAME = {name: Molecule(name, kj * _KJ, **extra) for name, (kj, extra) in _SPECIES.items()}
MOLECULES = list(_BY_NAME.values())
REACTIONS = [([_BY_NAME[n] for n in lhs.split()], [_BY_NAME[n] for n in rhs.split()]) for lhs, rhs in _STEPS]
CHEMISTRY = Chemistry(molecules=MOLECULES, reactions=REACTIONS)
