"""constants of the reference (python/magicsoup/constants.py:1-10)"""
from itertools import product

CODON_SIZE = 3  # nucleotides
GAS_CONSTANT = 8.31446261815324  # J/(K*mol)
ALL_NTS = tuple("TCGA")
ALL_CODONS = set("".join(d) for d in product(ALL_NTS, ALL_NTS, ALL_NTS))

# ((domain type, i0, i1, i2, i3), start, end) and (domains, cds_start, cds_end, is_fwd)
DomainSpecType = tuple[tuple[int, int, int, int, int], int, int]
ProteinSpecType = tuple[list[DomainSpecType], int, int, bool]
