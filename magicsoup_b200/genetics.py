"""
`Genetics` with the reference's interface (python/magicsoup/genetics.py:18-178): codon -> token
maps and genome translation through the native extension.
"""
from __future__ import annotations

import random
import warnings

from magicsoup_b200._hostlib import lib as _lib
from magicsoup_b200.constants import CODON_SIZE, ProteinSpecType
from magicsoup_b200.util import codons


def _get_n(p: float, s: int, name: str) -> int:
    n = int(p * s)
    if n == 0 and p > 0.0:
        warnings.warn(f"There will be no {name}. Increase dom_type_size to accomodate low probabilities of having {name}.")
    return n


class Genetics:
    """constructor signature and attributes of genetics.py:58-122"""

    def __init__(
        self,
        start_codons: tuple[str, ...] = ("TTG", "GTG", "ATG"),
        stop_codons: tuple[str, ...] = ("TGA", "TAG", "TAA"),
        p_catal_dom: float = 0.01,
        p_transp_dom: float = 0.01,
        p_reg_dom: float = 0.01,
        n_dom_type_codons: int = 2,
    ):
        if any(len(d) != CODON_SIZE for d in start_codons):
            raise ValueError(f"Not all start codons are of length {CODON_SIZE}")
        if any(len(d) != CODON_SIZE for d in stop_codons):
            raise ValueError(f"Not all stop codons are of length {CODON_SIZE}")
        both = set(start_codons) & set(stop_codons)
        if both:
            raise ValueError(f"Overlapping start and stop codons: {','.join(sorted(both))}")
        if p_catal_dom + p_transp_dom + p_reg_dom > 1.0:
            raise ValueError("p_catal_dom, p_transp_dom, p_reg_dom together must not be greater 1.0")
        self.start_codons = list(start_codons)
        self.stop_codons = list(stop_codons)
        # a domain = type codons + 3 one-codon tokens + 1 two-codon token (genetics.py:91-98)
        self.dom_size = (n_dom_type_codons + 5) * CODON_SIZE
        self.dom_type_size = n_dom_type_codons * CODON_SIZE

        sets = codons(n=n_dom_type_codons, excl_codons=self.start_codons)
        random.shuffle(sets)
        counts = (_get_n(p_catal_dom, len(sets), "catalytic domains"),
                  _get_n(p_transp_dom, len(sets), "transporter domains"),
                  _get_n(p_reg_dom, len(sets), "allosteric domains"))
        self.domain_types: dict[int, list[str]] = {}  # 1 = catalytic, 2 = transporter, 3 = regulatory
        at = 0
        for kind, n in zip((1, 2, 3), counts):
            self.domain_types[kind] = sets[at:at + n]
            at += n
        self.domain_map = {d: k for k, v in self.domain_types.items() for d in v}
        self.one_codon_map = {d: i + 1 for i, d in enumerate(self._get_single_codons())}
        self.two_codon_map = {d: i + 1 for i, d in enumerate(self._get_double_codons())}
        self.idx_2_one_codon = {v: k for k, v in self.one_codon_map.items()}
        self.idx_2_two_codon = {v: k for k, v in self.two_codon_map.items()}

    def translate_genomes(self, genomes: list[str]) -> list[list[ProteinSpecType]]:
        """proteomes of all genomes, forward and reverse complement (genetics.py:124-168)"""
        if len(genomes) < 1:
            return []
        return _lib().translate_genomes(genomes, self.start_codons, self.stop_codons, self.domain_map,
                                        self.one_codon_map, self.two_codon_map, self.dom_size, self.dom_type_size)

    def translate_genomes_flat(self, genomes: list[str]):
        """
        translate_genomes as three int32 arrays (protein offsets per genome, domain offsets per
        protein, (type, i0, i1, i2, i3) per domain) instead of nested lists: what
        Kinetics.set_cell_params_flat takes.  No counterpart in the reference.
        """
        return _lib().translate_genomes_flat(list(genomes), self.start_codons, self.stop_codons, self.domain_map,
                                             self.one_codon_map, self.two_codon_map, self.dom_size, self.dom_type_size)

    def _get_single_codons(self) -> list[str]:
        return [d for d in codons(n=1) if d not in self.stop_codons]  # no premature stop

    def _get_double_codons(self) -> list[str]:
        return [d for d in codons(n=2) if d[:CODON_SIZE] not in self.stop_codons]  # the second may be a stop
