"""
The names `magicsoup.constants` exports (python/magicsoup/constants.py:1-10), for code that imports
them from here: codon size, the gas constant of Ke = exp(-dE / (R T)), the nucleotide alphabet in
the order of the 2-bit codes used by `_lib`, and the shapes of the translation results.
"""
CODON_SIZE = 3  # nucleotides per codon
GAS_CONSTANT = 8.31446261815324  # J / (K mol)
ALL_NTS = ("T", "C", "G", "A")
ALL_CODONS = {a + b + c for a in ALL_NTS for b in ALL_NTS for c in ALL_NTS}

# what translate_genomes returns per genome: a list of proteins, each
#   (domains, cds_start, cds_end, is_fwd)   with domains = [((type, i0, i1, i2, i3), start, end), ...]
_DomainTokens = tuple[int, int, int, int, int]
DomainSpecType = tuple[_DomainTokens, int, int]
ProteinSpecType = tuple[list[DomainSpecType], int, int, bool]
