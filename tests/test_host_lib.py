"""
CPU tests of magicsoup_b200._lib (C++ / pybind11, csrc/lib_host.cpp), the counterpart of the
reference's Rust extension (rust/lib.rs:18-176), and of the Genetics / util / mutations modules on
top of it.  Golden vectors are the ones the reference's own tests hold
(/root/reference/tests/fast/test_genetics.py:10-201, test_util.py:95-125, test_world.py:650-732);
random inputs are checked against the plain restatements in oracle/host_oracle.py.
"""
import random
from collections import Counter

import pytest

import magicsoup_b200 as ms
from magicsoup_b200 import mutations, util
from magicsoup_b200._hostlib import lib
from magicsoup_b200.genetics import Genetics
from oracle import host_oracle as HO

STARTS, STOPS = ["TTG", "GTG", "ATG"], ["TGA", "TAG", "TAA"]

# (genome, [(start, stop)]), min CDS size 18 -- test_genetics.py:10-59
_CDS = [
    ("""TACCGGATA GCAGCTTTT CTTGGAATA GCCAAGGGT CGCCTTTAT ACCTATCTA CAACTACTA CTCGGTTGG
        TAACAAAGG TTAAAACGC CAAACGAGT ATCGGCCAA TCCTGTCAC TGTGAGAAG TTTCAATTA TAGATTCCT
        GGGGCGATT GGCGATGGT""", [(68, 122)]),
    ("""AACATATCC ACCATCCCT TAAGGGGCG ATGAATTAC GAAAGCGGG CGTACTACT TCTGGGGAT ACGATTAGT
        GTACTCGGT TCTCTTAAC GACTACCCT GTGTTACGT TATTGAAAG AGCAAATTG CGAGCTCCC CGTGACACT
        TGTGCGGCG CTATACACC CCTGCAGTT ATTTAAGGG CTTAGGCGA GAAGTTCCG CCTGCTAAG GAGTCCCTG
        TTGGGTGAA GTAACGCAC AGCCAGGCC TTGGCAGGA CGTTTCCGT TCTCGT""",
     [(27, 114), (70, 229), (110, 140), (123, 177), (136, 229), (143, 185), (145, 229)]),
    ("TTGAAAGA GCAAATTT GA", [(0, 18)]),
    ("GTGTGCTCG AAAGAGAAC GCAAATTCG TAACCTAG", [(0, 30), (2, 35)]),
]


def test_reverse_complement():
    assert lib().reverse_complement("ACTGG") == "CCAGT"


@pytest.mark.parametrize("seq, exp", _CDS)
def test_get_coding_regions_golden(seq, exp):
    seq = "".join(seq.split())
    res = lib().get_coding_regions(seq, 18, STARTS, STOPS, False)
    assert sorted((a, b) for a, b, _ in res) == sorted(exp)
    assert all(f is False for _, _, f in res)
    assert res == HO.get_coding_regions(seq, 18, STARTS, STOPS, False)  # same order as the reference's algorithm


def test_extract_domains_golden():
    # test_genetics.py:122-201
    dom_type_map = {"AAA": 1, "GGG": 2, "CCC": 3}
    two = {"ACTGAT": 1, "CTGTAT": 2, "CCGCGA": 3, "GGAATC": 4, "TGTCGA": 5}
    one = {"ACT": 1, "CTG": 2, "CCG": 3, "GGA": 4, "TGT": 5}
    genome = ("AGACAAAAACTGTGTACTCCGCGATAGACTAGACG" "AGACTATAGCTAGAAGCCCCTGTACTCCGTGTCGATAGACG"
              "AGACTAGGGCCGGGACTGCCGCGACTAGAAGCTAGACTAACG" "AAACCGGGATGTCTGTAT" "CCCCCGGGACTGCCGCGAGGGACTCTGCCGGGAATC")
    cdss = [(0, 35, True), (35, 76, False), (76, 118, True), (118, 136, False), (136, 172, True)]
    res = lib().extract_domains(genome, cdss, 18, 3, dom_type_map, one, two)
    assert [(r[1], r[2], r[3]) for r in res] == [(0, 35, True), (76, 118, True), (118, 136, False), (136, 172, True)]
    assert [len(r[0]) for r in res] == [1, 1, 1, 2]
    assert res[0][0][0] == ((1, 2, 5, 1, 3), 6, 24)
    assert res[1][0][0] == ((2, 3, 4, 2, 3), 6, 24)
    assert res[2][0][0] == ((1, 3, 4, 5, 2), 0, 18)
    assert res[3][0][0] == ((3, 3, 4, 2, 3), 0, 18)
    assert res[3][0][1] == ((2, 1, 2, 3, 4), 18, 36)
    assert res == HO.extract_domains(genome, cdss, 18, 3, dom_type_map, one, two)


def test_translate_genomes_matches_the_plain_restatement():
    random.seed(5)
    gen = Genetics()
    genomes = [util.random_genome(s=random.choice([0, 10, 200, 500, 1000, 3000])) for _ in range(200)]
    res = gen.translate_genomes(genomes)
    assert len(res) == 200
    n_prot = 0
    for g, prots in zip(genomes, res):
        exp = HO.translate_genome(g, gen.start_codons, gen.stop_codons, gen.domain_map, gen.one_codon_map,
                                  gen.two_codon_map, gen.dom_size, gen.dom_type_size)
        assert prots == exp
        n_prot += len(prots)
    assert n_prot > 100
    assert gen.translate_genomes([]) == []


@pytest.mark.parametrize("n_type_codons, probs", [(2, 0.05), (1, 0.3), (3, 0.1)])
def test_translation_paths_agree(n_type_codons, probs):
    """table-driven walk (two type codons) vs the look-ups by key (any other size) vs the plain
    restatement, and the flat arrays vs flatten_proteomes of the nested lists"""
    import numpy as np

    random.seed(8 + n_type_codons)
    gen = Genetics(p_catal_dom=probs, p_transp_dom=probs, p_reg_dom=probs, n_dom_type_codons=n_type_codons)
    genomes = [util.random_genome(s=random.choice([0, 2, 30, 200, 700])) for _ in range(150)]
    genomes += ["", "ATG", "ATGNNNNNNTGA" * 20, genomes[5].lower()]  # characters that are no nucleotides
    res = gen.translate_genomes(genomes)
    for g, prots in zip(genomes, res):
        exp = HO.translate_genome(g, gen.start_codons, gen.stop_codons, gen.domain_map, gen.one_codon_map,
                                  gen.two_codon_map, gen.dom_size, gen.dom_type_size)
        assert prots == exp
    assert sum(len(p) for p in res) > 50
    nested = lib().flatten_proteomes(res)
    flat = gen.translate_genomes_flat(genomes)
    for a, b in zip(nested, flat):
        assert a.dtype == b.dtype == np.int32 and np.array_equal(a, b)
    assert len(flat[0]) == len(genomes) + 1 and flat[0][-1] == len(flat[1]) - 1 and flat[1][-1] * 5 == len(flat[2])


def test_genetics_maps_and_domain_frequencies():
    # genetics.py:104-122 and the reference's statistical check (test_genetics.py:204-246)
    random.seed(1)
    gen = Genetics(p_catal_dom=0.1, p_transp_dom=0.1, p_reg_dom=0.1)
    assert len(gen.one_codon_map) == 61 and len(gen.two_codon_map) == 3904
    assert set(gen.domain_map.values()) == {1, 2, 3} and gen.dom_size == 21 and gen.dom_type_size == 6
    assert not any(c[:3] in gen.start_codons or c[3:] in gen.start_codons for c in gen.domain_map)
    kinds = Counter()
    for prots in gen.translate_genomes([util.random_genome(1000) for _ in range(300)]):
        for doms, start, stop, _ in prots:
            assert stop - start >= gen.dom_size and any(d[0][0] != 3 for d in doms)
            kinds.update(d[0][0] for d in doms)
    tot = sum(kinds.values())
    assert all(0.2 < kinds[k] / tot < 0.5 for k in (1, 2, 3)), kinds
    with pytest.warns(UserWarning):
        Genetics(n_dom_type_codons=1)
    with pytest.raises(ValueError):
        Genetics(start_codons=("TTG",), stop_codons=("TTG",))


@pytest.mark.parametrize("a, b, exp", [(0, 1, 1), (0, 2, 2), (0, 3, 2), (0, 4, 1), (0, 0, 0)])
def test_dist_1d(a, b, exp):
    assert util.dist_1d(a=a, b=b, m=5) == exp  # test_util.py:95-107


@pytest.mark.parametrize("x, y, exp", [
    (2, 2, [(1, 1), (1, 2), (1, 3), (3, 1), (3, 2), (3, 3), (2, 1), (2, 3)]),
    (0, 0, [(4, 4), (4, 0), (4, 1), (1, 4), (1, 0), (1, 1), (0, 4), (0, 1)]),
    (4, 4, [(3, 3), (3, 4), (3, 0), (0, 3), (0, 4), (0, 0), (4, 3), (4, 0)]),
    (0, 4, [(4, 3), (4, 4), (4, 0), (1, 3), (1, 4), (1, 0), (0, 3), (0, 0)]),
    (4, 0, [(3, 4), (4, 4), (0, 4), (3, 1), (4, 1), (0, 1), (3, 0), (0, 0)]),
])
def test_free_moores_nghbhd(x, y, exp):
    res = util.free_moores_nghbhd(x=x, y=y, positions=[], map_size=5)  # test_util.py:110-125
    assert res == HO.moores_nghbhd(x, y, 5) and set(res) == set(exp)
    assert set(util.free_moores_nghbhd(x=x, y=y, positions=[res[0]], map_size=5)) == set(exp) - {res[0]}
    assert util.free_moores_nghbhd(x=x, y=y, positions=res, map_size=5) == []


def test_get_neighbors_golden_and_random():
    # test_world.py:650-732
    pos = [(1, 1), (1, 2), (0, 1), (2, 1), (8, 8), (0, 8), (2, 0), (8, 6), (8, 7)]
    nghbrs = [{(0, 2), (0, 1), (0, 6), (0, 3)}, {(1, 2), (0, 1), (1, 3)}, {(0, 2), (1, 2)}, {(0, 3), (1, 3), (3, 6)},
              {(4, 8), (4, 5)}, {(5, 8), (4, 5)}, {(0, 6), (3, 6)}, {(7, 8)}, {(7, 8), (4, 8), (5, 8)}]
    everyone = list(range(9))
    for i, exp in enumerate(nghbrs):
        res = lib().get_neighbors([i], everyone, pos, 9)
        assert len(res) == len(exp) and set(res) == exp
    res = lib().get_neighbors([1, 4, 8], everyone, pos, 9)
    assert len(res) == len(set(res)) and set(res) == nghbrs[1] | nghbrs[4] | nghbrs[8]
    res = lib().get_neighbors([0, 3], [6, 2], pos, 9)
    assert set(res) == {d for d in nghbrs[0] | nghbrs[3] if d[0] in (6, 2) or d[1] in (6, 2)}
    rnd = random.Random(3)
    for size in (2, 3, 7, 40):
        cells = rnd.sample([(x, y) for x in range(size) for y in range(size)], k=min(size * size, 60))
        frm = rnd.sample(range(len(cells)), k=max(1, len(cells) // 2))
        to = rnd.sample(range(len(cells)), k=max(1, len(cells) * 2 // 3))
        assert lib().get_neighbors(frm, to, cells, size) == HO.get_neighbors(frm, to, cells, size)


def test_divide_and_move_keep_the_map_consistent():
    rnd = random.Random(7)
    lib().seed(11)
    size = 12
    cells = rnd.sample([(x, y) for x in range(size) for y in range(size)], k=90)
    movers = sorted(rnd.sample(range(90), k=60))
    parents, children, child_pos = lib().divide_cells_if_possible(movers, cells, 90, size)
    assert children == list(range(90, 90 + len(parents))) and len(set(child_pos)) == len(child_pos)
    assert not set(child_pos) & set(cells) and parents == [i for i in movers if i in set(parents)]
    for p, c in zip(parents, child_pos):
        assert c in HO.moores_nghbhd(*cells[p], size)
    # a cell that is not divided had no free pixel left when its turn came
    taken = set(cells)
    for i in movers:
        if i in parents:
            taken.add(child_pos[parents.index(i)])
        else:
            assert all(q in taken for q in HO.moores_nghbhd(*cells[i], size))
    new_pos, moved = lib().move_cells(movers, cells, size)
    after = list(cells)
    for i, q in zip(moved, new_pos):
        assert q in HO.moores_nghbhd(*cells[i], size)
        after[i] = q
    assert len(set(after)) == 90 and moved == [i for i in movers if i in set(moved)]
    # a full map: nobody divides, nobody moves
    full = [(x, y) for x in range(4) for y in range(4)]
    assert lib().divide_cells_if_possible(list(range(16)), full, 16, 4) == ([], [], [])
    assert lib().move_cells(list(range(16)), full, 4) == ([], [])
    # the choice among free pixels is uniform (reference: slice.choose, rust/world.rs:90)
    seen = Counter(lib().divide_cells_if_possible([0], [(5, 5)], 1, 11)[2][0] for _ in range(1600))
    assert len(seen) == 8 and min(seen.values()) > 120


def test_point_mutations_and_recombinations_statistics():
    # tests/fast/test_mutations.py, tests/slow/test_mutations.py: statistical, the generator is random
    lib().seed(3)
    seqs = [util.random_genome(1000) for _ in range(400)]
    res = mutations.point_mutations(seqs, p=1e-3)
    assert 200 < len(res) < 320  # P(no mutation) = exp(-1)
    idxs = [i for _, i in res]
    assert idxs == sorted(set(idxs)) and all(set(s) <= set("ACTG") for s, _ in res)
    diffs = Counter(len(s) - 1000 for s, _ in res)
    assert diffs[0] > 0 and sum(v for k, v in diffs.items() if k < 0) > sum(v for k, v in diffs.items() if k > 0)
    assert mutations.point_mutations(seqs, p=0.0) == [] and mutations.point_mutations([], p=0.5) == []
    only_sub = mutations.point_mutations(seqs, p=1e-2, p_indel=0.0)
    assert len(only_sub) == 400 and all(len(s) == 1000 for s, _ in only_sub)
    n_changed = sum(a != b for a, b in zip(only_sub[0][0], seqs[only_sub[0][1]]))
    assert 1 <= n_changed <= 30
    everything = mutations.point_mutations(["ACGT" * 5], p=10.0, p_indel=1.0, p_del=1.0)
    assert everything == [("", 0)]  # more mutations than nucleotides: capped at one per nucleotide, all deleted
    pairs = [(util.random_genome(500), util.random_genome(700)) for _ in range(300)]
    rec = mutations.recombinations(pairs, p=1e-3)
    assert 150 < len(rec) < 260  # P(no break) = exp(-1.2)
    for a, b, i in rec:
        assert len(a) + len(b) == 1200 and Counter(a + b) == Counter(pairs[i][0] + pairs[i][1])
    assert mutations.recombinations(pairs, p=0.0) == [] and mutations.recombinations([], p=0.1) == []


def test_package_surface():
    assert ms.Genetics is Genetics and ms.random_genome is util.random_genome
    assert ms.point_mutations is mutations.point_mutations and ms.recombinations is mutations.recombinations


def _params_fixture():
    import numpy as np
    from pathlib import Path

    z = np.load(Path(__file__).resolve().parent / "golden" / "ref_set_cell_params.npz")
    tables = {k: z[k] for k in ("vmax_weights", "km_weights", "signs", "hill_numbers", "reaction_M", "transport_M", "effector_M")}
    return z, tables


def unflatten(prot_off, dom_off, doms):
    """flattened tokens back into the nested lists translate_genomes returns (positions are not kept)"""
    doms = doms.reshape(-1, 5)
    return [[([(tuple(int(x) for x in doms[d]), 0, 0) for d in range(dom_off[q], dom_off[q + 1])], 0, 0, True)
             for q in range(prot_off[k], prot_off[k + 1])] for k in range(len(prot_off) - 1)]


PARAM_TOLERANCES = {"Vmax": 2e-6, "Kmr": 2e-5, "Ke": 2e-4, "Kmf": 2e-4, "Kmb": 2e-4}


def check_params_against_reference(got, z):
    """
    integers exact; floats relative: means of a few float32 2e-6, powf with exponents up to 25 2e-5,
    exp of an energy sum of magnitude 1e2 (float32 rounding of the sum enters the exponent) 2e-4
    """
    import numpy as np

    for k in ("N", "Nf", "Nb", "A"):
        assert np.array_equal(got[k], z[k]), k
    for k, tol in PARAM_TOLERANCES.items():
        a, b = np.asarray(got[k], dtype=np.float64), z[k].astype(np.float64)
        assert np.array_equal(np.isnan(a), np.isnan(b)) and np.array_equal(np.isinf(a), np.isinf(b)), k
        fin = np.isfinite(b)
        sat = fin & ((b == np.float32(1e36)) | (b == np.float32(1e-36)))
        assert np.array_equal(a[sat], b[sat]), k  # clamped values are exact
        assert (np.abs(a[fin] - b[fin]) <= tol * np.abs(b[fin])).all(), (k, np.abs(a[fin] / b[fin] - 1).max())


def test_set_cell_params_restatement_matches_the_reference():
    # pins oracle/host_oracle.py::set_cell_params to the reference's Kinetics.set_cell_params
    # (tests/golden/make_golden.py::set_cell_params_fixture); the CUDA kernel is held to both in
    # tests/test_gpu_params.py
    z, tables = _params_fixture()
    c, p, s = z["N"].shape
    got = HO.set_cell_params(z["prot_off"], z["dom_off"], z["doms"], tables, z["mol_energies"], float(z["abs_temp"]), p, s)
    check_params_against_reference(got, z)
    assert unflatten(z["prot_off"], z["dom_off"], z["doms"])[0][0][0][0][0] == tuple(z["doms"][:5])
    assert lib().flatten_proteomes(unflatten(z["prot_off"], z["dom_off"], z["doms"]))[2].tolist() == z["doms"].tolist()


def test_remove_items_in_place():
    """the list edit behind World.kill_cells (world.py:527-540 pops genome and label per index)"""
    import random as _r

    rng = _r.Random(4)
    for n, k in ((0, 0), (1, 1), (40, 17), (5000, 300), (5000, 5000)):
        items = [f"g{i}" for i in range(n)]
        same = items
        dead = sorted(rng.sample(range(n), k))
        want = [v for i, v in enumerate(items) if i not in set(dead)]
        lib().remove_items(items, dead)
        assert items == want and items is same
        items.append("tail")  # the list stays a working list
        assert items[-1] == "tail" and len(items) == n - k + 1
    with pytest.raises(ValueError):
        lib().remove_items([1, 2, 3], [1, 1])
    with pytest.raises(IndexError):
        lib().remove_items([1, 2, 3], [3])


def test_lazy_mutation_entry_points():
    """point_mutations / recombinations read only the sequences that are hit; the indexed and the
    fused neighbour variants give what the reference's two-step composition gives"""
    import numpy as np

    random.seed(12)
    genomes = [util.random_genome(s=random.choice([0, 50, 400])) for _ in range(400)]
    assert lib().point_mutations(genomes, 0.0, 0.4, 0.66) == []
    assert lib().recombinations([(genomes[0], genomes[1])], 0.0) == []
    with pytest.raises(TypeError):
        lib().point_mutations(["ACGT", 5], 0.5, 0.4, 0.66)
    with pytest.raises(TypeError):
        lib().recombinations([("ACGT",)], 0.5)
    # tuples as well as lists, generators of pairs
    res = lib().recombinations([[genomes[1], genomes[2]], (genomes[3], genomes[4])], 0.05)
    for a, b, i in res:
        src = (genomes[1] + genomes[2], genomes[3] + genomes[4])[i]
        assert sorted(a + b) == sorted(src)
    # same seed, same pairs: the indexed variant is the same computation
    pairs = [(random.randrange(400), random.randrange(400)) for _ in range(300)]
    lib().seed(77)
    by_string = lib().recombinations([(genomes[a], genomes[b]) for a, b in pairs], 0.01)
    lib().seed(77)
    by_index = lib().recombinations_indexed(genomes, pairs, 0.01)
    assert by_string == by_index and len(by_index) > 20
    with pytest.raises(IndexError):
        lib().recombinations_indexed(genomes, [(0, 400)], 0.01)
    # fused: neighbours among the given cells, then recombinations between them
    size = 30
    cells = random.sample([(x, y) for x in range(size) for y in range(size)], 400)
    pos = np.array(cells, dtype=np.int32)
    idxs = list(range(0, 400, 2))
    nghbrs = lib().get_neighbors(idxs, idxs, pos, size)
    assert nghbrs == lib().get_neighbors(idxs, idxs, cells, size) == HO.get_neighbors(idxs, idxs, cells, size)
    lib().seed(78)
    want = [(a, b) + nghbrs[i] for a, b, i in lib().recombinations_indexed(genomes, nghbrs, 0.02)]
    lib().seed(78)
    got = lib().neighbor_recombinations(genomes, idxs, pos, size, 0.02)
    assert got == want and len(got) > 3


def test_map_searches_on_a_row_slab():
    """(rows, cols, wrap_rows=False): a row slab of a map split over several GPUs -- columns wrap,
    the first and the last row end at another rank (magicsoup_b200/distributed.py)"""
    rng = random.Random(9)
    rows, cols = 7, 11
    pixels = rng.sample([(x, y) for x in range(rows) for y in range(cols)], 40)
    idxs = list(range(len(pixels)))

    def near(a, b, wrap_rows):
        dx = abs(a[0] - b[0])
        if wrap_rows:
            dx = min(dx, rows - dx)
        dy = abs(a[1] - b[1])
        return dx <= 1 and min(dy, cols - dy) <= 1

    for wrap in (False, True):
        got = lib().get_neighbors(idxs, idxs, pixels, (rows, cols, wrap))
        want = {(i, j) for i in idxs for j in idxs if i < j and near(pixels[i], pixels[j], wrap)}
        assert set(got) == want and len(got) == len(want)
    assert set(lib().get_neighbors(idxs, idxs, pixels, (rows, cols))) == \
        set(lib().get_neighbors(idxs, idxs, pixels, (rows, cols, True)))
    # children and moves stay inside the slab and on free pixels
    lib().seed(5)
    parents, children, child_pos = lib().divide_cells_if_possible(idxs, pixels, len(pixels), (rows, cols, False))
    assert children == list(range(len(pixels), len(pixels) + len(parents))) and len(parents) > 20
    taken = set(pixels)
    for par, pos in zip(parents, child_pos):
        pos = tuple(pos)
        assert 0 <= pos[0] < rows and 0 <= pos[1] < cols and pos not in taken and near(pixels[par], pos, False)
        taken.add(pos)
    new_pos, moved = lib().move_cells(idxs, pixels, (rows, cols, False))
    now = list(pixels)
    for i, pos in zip(moved, new_pos):
        pos = tuple(pos)
        assert 0 <= pos[0] < rows and near(pixels[i], pos, False)
        now[i] = pos
    assert len(set(now)) == len(now)
    # a cell in a one-row slab can only go sideways
    assert lib().divide_cells_if_possible([0], [(0, 0)], 1, (1, 3, False))[2] in ([(0, 1)], [(0, 2)])
    with pytest.raises(ValueError):
        lib().get_neighbors([0], [0], [(0, 0)], (0, 4))
    with pytest.raises(IndexError):
        lib().get_neighbors([0], [0], [(5, 0)], (4, 4, False))
