"""
Plain-Python restatements of the reference's native host helpers (rust/genetics.rs, rust/world.rs,
rust/util.rs), written as the slow, obvious algorithms.  TEST INFRASTRUCTURE ONLY: tests/ use them
to check magicsoup_b200._lib (C++) on random inputs; the product never imports this module.
"""
from __future__ import annotations

CODON = 3


def get_coding_regions(seq, min_cds_size, start_codons, stop_codons, is_fwd):
    """rust/genetics.rs:13-50"""
    out = []
    n = len(seq)
    if n < min_cds_size:
        return out
    starts = [[], [], []]
    for i in range(n - CODON + 1):
        frame, j = i % CODON, i + CODON
        codon = seq[i:j]
        if codon in start_codons:
            starts[frame].append(i)
        elif codon in stop_codons:
            while starts[frame]:
                d = starts[frame].pop()
                if j - d >= min_cds_size:
                    out.append((d, j, is_fwd))
    return out


def extract_domains(genome, cdss, dom_size, dom_type_size, dom_type_map, one_codon_map, two_codon_map):
    """rust/genetics.rs:58-123"""
    out = []
    for cds_start, cds_stop, is_fwd in cdss:
        n, i, useful, doms = cds_stop - cds_start, 0, False, []
        while i + dom_size <= n:
            d0 = cds_start + i
            t = dom_type_map.get(genome[d0:d0 + dom_type_size])
            if t is None:
                i += CODON
                continue
            useful |= t != 3
            spec = genome[d0 + dom_type_size:d0 + dom_size]
            doms.append(((t, one_codon_map[spec[0:3]], one_codon_map[spec[3:6]], one_codon_map[spec[6:9]],
                          two_codon_map[spec[9:15]]), i, i + dom_size))
            i += dom_size
        if useful:
            out.append((doms, cds_start, cds_stop, is_fwd))
    return out


def reverse_complement(seq):
    """rust/genetics.rs:126-137"""
    m = {"A": "T", "C": "G", "T": "A", "G": "C"}
    return "".join(m[c] for c in reversed(seq) if c in m)


def translate_genome(genome, start_codons, stop_codons, domain_map, one_codon_map, two_codon_map, dom_size, dom_type_size):
    """rust/genetics.rs:141-176"""
    fwd = extract_domains(genome, get_coding_regions(genome, dom_size, start_codons, stop_codons, True),
                          dom_size, dom_type_size, domain_map, one_codon_map, two_codon_map)
    bwd_seq = reverse_complement(genome)
    bwd = extract_domains(bwd_seq, get_coding_regions(bwd_seq, dom_size, start_codons, stop_codons, False),
                          dom_size, dom_type_size, domain_map, one_codon_map, two_codon_map)
    return fwd + bwd


def dist_1d(a, b, m):
    """rust/util.rs:15-25"""
    hi, lo = max(a, b), min(a, b)
    return min(hi - lo, m - hi + lo)


def moores_nghbhd(x, y, m):
    """rust/util.rs:27-43"""
    e, w = (0 if x + 1 == m else x + 1), (m - 1 if x == 0 else x - 1)
    s, n = (0 if y + 1 == m else y + 1), (m - 1 if y == 0 else y - 1)
    return [(w, n), (w, y), (w, s), (x, n), (x, s), (e, n), (e, y), (e, s)]


def get_neighbors(from_idxs, to_idxs, positions, map_size):
    """rust/world.rs:9-56: pairs in from/to order, first occurrence kept"""
    out = []
    for f in from_idxs:
        x0, y0 = positions[f]
        for t in to_idxs:
            if t == f:
                continue
            x1, y1 = positions[t]
            if dist_1d(x0, x1, map_size) <= 1 and dist_1d(y0, y1, map_size) <= 1:
                pair = (min(f, t), max(f, t))
                if pair not in out:
                    out.append(pair)
    return out


def set_cell_params(prot_off, dom_off, doms, tables, mol_energies, abs_temp, p, s):
    """
    Kinetics.set_cell_params (python/magicsoup/kinetics.py:521-625) as plain loops over the
    flattened domain tokens; returns the nine (c,p[,s]) arrays.  Means are sum / count in
    float32 in domain order, the energy sum in float64 (as the CUDA kernel does).
    """
    import numpy as np

    f32 = np.float32
    c = len(prot_off) - 1
    doms = np.asarray(doms).reshape(-1, 5)
    out = {k: np.zeros((c, p, s), dtype=np.int32) for k in ("N", "Nf", "Nb", "A")}
    out["Kmr"] = np.ones((c, p, s), dtype=f32)
    out.update(Ke=np.ones((c, p), f32), Kmf=np.full((c, p), 1e-36, f32), Kmb=np.full((c, p), 1e-36, f32),
               Vmax=np.zeros((c, p), f32))
    R = f32(8.31446261815324)
    with np.errstate(all="ignore"):
        for k in range(c):
            for j, q in enumerate(range(prot_off[k], prot_off[k + 1])):
                vs, ks = [], []
                N = np.zeros(s, np.int32); Nf = np.zeros(s, np.int32); Nb = np.zeros(s, np.int32)
                Af = np.zeros(s, f32); rsum = np.zeros(s, f32); rcnt = np.zeros(s, np.int32)
                for dt, i0, i1, i2, i3 in doms[dom_off[q]:dom_off[q + 1]]:
                    reg = dt == 3
                    vm = tables["vmax_weights"][0 if reg else i0]
                    km = tables["km_weights"][i1]
                    if not np.isnan(vm):
                        vs.append(vm)
                    if not reg and not np.isnan(km):
                        ks.append(km)
                    sign = int(tables["signs"][i2])
                    nd = (tables["reaction_M"][i3 if dt == 1 else 0] + tables["transport_M"][i3 if dt == 2 else 0]) * sign
                    N += nd; Nf += np.where(nd < 0, -nd, 0); Nb += np.where(nd > 0, nd, 0)
                    eff = tables["effector_M"][i3 if reg else 0].astype(f32)
                    hill = int(tables["hill_numbers"][i0 if reg else 0])
                    Af = Af + eff * f32(sign * hill)
                    kv = eff * (km if reg else f32(np.nan))
                    use = (kv != 0) & ~np.isnan(kv)
                    rsum = np.where(use, rsum + kv, rsum).astype(f32); rcnt += use
                A = Af.astype(np.int32)
                kmr = np.where(rcnt > 0, rsum / np.maximum(rcnt, 1).astype(f32), f32(0)).astype(f32)
                vmax = f32(0) if not vs else f32(sum(vs[1:], start=f32(vs[0]))) / f32(len(vs))
                acc = f32(0)
                for x in ks:
                    acc = f32(acc + x)
                kmn = f32(0) if not ks else acc / f32(len(ks))
                E = f32(np.dot(N.astype(np.float64), mol_energies.astype(np.float64)))
                ke = np.exp((-E / f32(abs_temp)) / R).astype(f32)
                ke = np.clip(ke, f32(1e-36), f32(1e36))
                fwd = ke >= 1
                out["N"][k, j], out["Nf"][k, j], out["Nb"][k, j], out["A"][k, j] = N, Nf, Nb, A
                out["Kmr"][k, j] = np.power(kmr, A.astype(f32)).astype(f32)
                out["Vmax"][k, j], out["Ke"][k, j] = vmax, ke
                out["Kmf"][k, j] = np.clip(kmn if fwd else kmn / ke, f32(1e-36), f32(1e36))
                out["Kmb"][k, j] = np.clip(kmn * ke if fwd else kmn, f32(1e-36), f32(1e36))
    return out
