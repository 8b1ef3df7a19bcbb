// magicsoup_b200._lib -- host-side helpers with the interface of the reference's Rust extension
// `magicsoup._lib` (mRcSchwering/magic-soup 1.0.5, rust/lib.rs:18-176): genome translation, point
// mutations and recombinations on strings, and the neighbourhood / division / movement searches on
// the cell map.  C++17 + pybind11, OpenMP over genomes, the GIL released while working.
//
// Not a translation of the Rust sources: the map searches use an occupancy grid (hash of pixel ->
// cells), so a query costs O(1) instead of the reference's linear scans over all positions
// (rust/world.rs:9-146, rust/util.rs:45-64: O(n^2) per call), and coding regions are found with
// per-frame stacks over a 2-bit codon code instead of string compares.  Same results for the same
// inputs wherever the reference is deterministic; where it draws random numbers (choice of a free
// pixel, mutations) the distributions are the same and `seed()` makes them reproducible, which
// the reference's thread_rng is not.
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <algorithm>
#include <array>
#include <cstdint>
#include <cmath>
#include <cstring>
#include <mutex>
#include <random>
#include <string>
#include <tuple>
#include <unordered_map>
#include <unordered_set>
#include <vector>

namespace py = pybind11;

namespace {

constexpr size_t CODON = 3;

using DomainSpec = std::tuple<std::tuple<int, int, int, int, int>, size_t, size_t>;
using ProteinSpec = std::tuple<std::vector<DomainSpec>, size_t, size_t, bool>;
using Cds = std::tuple<size_t, size_t, bool>;
using Pos = std::pair<uint16_t, uint16_t>;

// ---------------------------------------------------------------------------- random numbers
std::mutex g_rng_mutex;
std::mt19937_64 g_rng{std::random_device{}()};

uint64_t next_seed() {
  std::lock_guard<std::mutex> lock(g_rng_mutex);
  return g_rng();
}

// One small generator per sequence (xoshiro256**, seeded through splitmix64 from the call's seed
// and the sequence's index): the streams are independent of the thread schedule, and setting one
// up costs a few instructions where a Mersenne twister initialises 312 words.
struct SeqRng {
  using result_type = uint64_t;
  uint64_t s[4];
  SeqRng(uint64_t seed, uint64_t index) {
    uint64_t x = seed ^ (0x9e3779b97f4a7c15ull * (index + 1));
    for (auto &w : s) {
      x += 0x9e3779b97f4a7c15ull;
      uint64_t z = x;
      z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
      z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
      w = z ^ (z >> 31);
    }
  }
  static constexpr uint64_t min() { return 0; }
  static constexpr uint64_t max() { return ~0ull; }
  static uint64_t rotl(uint64_t v, int k) { return (v << k) | (v >> (64 - k)); }
  uint64_t operator()() {
    const uint64_t r = rotl(s[1] * 5, 7) * 9, t = s[1] << 17;
    s[2] ^= s[0];
    s[3] ^= s[1];
    s[1] ^= s[2];
    s[0] ^= s[3];
    s[2] ^= t;
    s[3] = rotl(s[3], 45);
    return r;
  }
  double uniform() { return (double)((*this)() >> 11) * (1.0 / 9007199254740992.0); }  // [0, 1)
};

// ---------------------------------------------------------------------------- genetics
// nucleotide -> 2 bits (anything else: 4, never part of a known codon)
inline unsigned nt_code(char c) {
  switch (c) {
    case 'T': return 0;
    case 'C': return 1;
    case 'G': return 2;
    case 'A': return 3;
    default: return 4;
  }
}

// codon membership table over the 64 codon codes
std::array<bool, 64> codon_table(const std::vector<std::string> &codons) {
  std::array<bool, 64> t{};
  for (const auto &c : codons) {
    if (c.size() != CODON) continue;
    const unsigned a = nt_code(c[0]), b = nt_code(c[1]), d = nt_code(c[2]);
    if (a < 4 && b < 4 && d < 4) t[a * 16 + b * 4 + d] = true;
  }
  return t;
}

// rust/genetics.rs:13-50: every start codon opens a CDS in its frame, the first in-frame stop
// codon closes all CDSs open in that frame (most recent first); CDSs shorter than min_size are
// dropped, unstopped ones too.  One pass, a rolling 6-bit codon code.
std::vector<Cds> coding_regions(const std::string &seq, size_t min_size, const std::array<bool, 64> &starts,
                                const std::array<bool, 64> &stops, bool is_fwd) {
  std::vector<Cds> out;
  const size_t n = seq.size();
  if (n < min_size || n < CODON) return out;
  std::array<std::vector<size_t>, 3> open;
  unsigned code = 0, bad = 0;  // bad: how many of the last 3 characters are not nucleotides
  for (size_t j = 0; j < n; j++) {
    const unsigned c = nt_code(seq[j]);
    code = ((code << 2) | (c & 3u)) & 63u;
    bad = ((bad << 1) | (c > 3 ? 1u : 0u)) & 7u;
    if (j + 1 < CODON || bad) continue;
    const size_t i = j + 1 - CODON, frame = i % CODON;
    if (starts[code]) {
      open[frame].push_back(i);
    } else if (stops[code]) {
      auto &st = open[frame];
      while (!st.empty()) {
        const size_t d = st.back();
        st.pop_back();
        if (j + 1 - d >= min_size) out.emplace_back(d, j + 1, is_fwd);
      }
    }
  }
  return out;
}

struct CodonMaps {
  std::unordered_map<std::string, int> dom_type, one, two;
  // dense tables over codon codes, usable when the domain type is two codons long (the
  // reference's default) and every key is made of nucleotides; -1 = not in the map
  bool dense = false;
  std::vector<int> type_tab, one_tab, two_tab;
};

// 6-bit code of a codon string, -1 if it holds anything but T, C, G, A
inline int codon_code(const char *c) {
  const unsigned a = nt_code(c[0]), b = nt_code(c[1]), d = nt_code(c[2]);
  return (a | b | d) > 3 ? -1 : (int)(a * 16 + b * 4 + d);
}

void make_dense_tables(CodonMaps &m) {
  m.type_tab.assign(4096, -1);
  m.one_tab.assign(64, -1);
  m.two_tab.assign(4096, -1);
  const auto two_code = [](const std::string &k) {
    if (k.size() != 2 * CODON) return -1;
    const int a = codon_code(k.data()), b = codon_code(k.data() + CODON);
    return a < 0 || b < 0 ? -1 : a * 64 + b;
  };
  for (const auto &[k, v] : m.dom_type) {
    const int c = two_code(k);
    if (c < 0 || v < 0) return;
    m.type_tab[c] = v;
  }
  for (const auto &[k, v] : m.one) {
    const int c = k.size() == CODON ? codon_code(k.data()) : -1;
    if (c < 0 || v < 0) return;
    m.one_tab[c] = v;
  }
  for (const auto &[k, v] : m.two) {
    const int c = two_code(k);
    if (c < 0 || v < 0) return;
    m.two_tab[c] = v;
  }
  m.dense = true;
}

// rust/genetics.rs:58-123: walk each CDS; a known domain-type prefix consumes a whole domain
// (type, 3 one-codon tokens, 1 two-codon token), anything else one codon; proteins without a
// catalytic or transporter domain are dropped.  Map look-ups by substring (any key length).
std::vector<ProteinSpec> domains_of_by_key(const std::string &genome, const std::vector<Cds> &cdss, size_t dom_size,
                                           size_t dom_type_size, const CodonMaps &maps) {
  std::vector<ProteinSpec> out;
  out.reserve(cdss.size());
  for (const auto &[cds_start, cds_stop, is_fwd] : cdss) {
    const size_t n = cds_stop - cds_start;
    std::vector<DomainSpec> doms;
    bool useful = false;
    size_t i = 0;
    while (i + dom_size <= n) {
      const size_t d0 = cds_start + i;
      auto it = maps.dom_type.find(genome.substr(d0, dom_type_size));
      if (it == maps.dom_type.end()) {
        i += CODON;
        continue;
      }
      const int type = it->second;
      if (type != 3) useful = true;  // 1 = catalytic, 2 = transporter, 3 = regulatory
      const size_t t0 = d0 + dom_type_size;
      const auto tok1 = [&](size_t k) -> int {
        auto f = maps.one.find(genome.substr(t0 + k * CODON, CODON));
        if (f == maps.one.end()) throw std::runtime_error("Incomplete one_codon_map");
        return f->second;
      };
      auto f2 = maps.two.find(genome.substr(t0 + 3 * CODON, 2 * CODON));
      if (f2 == maps.two.end()) throw std::runtime_error("Incomplete two_codon_map");
      doms.emplace_back(std::make_tuple(type, tok1(0), tok1(1), tok1(2), f2->second), i, i + dom_size);
      i += dom_size;
    }
    if (useful) out.emplace_back(std::move(doms), cds_start, cds_stop, is_fwd);
  }
  return out;
}

// the same walk over per-position codon codes and dense tables: CDSs nest (every start codon
// before a stop opens one), so a genome's codons are visited many times and a table look-up per
// visit instead of a substring + hash is what makes translation cheap
std::vector<ProteinSpec> domains_of(const std::string &genome, const std::vector<Cds> &cdss, size_t dom_size,
                                    size_t dom_type_size, const CodonMaps &maps) {
  if (!maps.dense || dom_type_size != 2 * CODON || dom_size != dom_type_size + 5 * CODON)
    return domains_of_by_key(genome, cdss, dom_size, dom_type_size, maps);
  std::vector<ProteinSpec> out;
  if (cdss.empty()) return out;
  const size_t len = genome.size();
  std::vector<int8_t> code(len, -1);  // codon starting at each position
  for (size_t i = 0; i + CODON <= len; i++) code[i] = (int8_t)codon_code(genome.data() + i);
  out.reserve(cdss.size());
  for (const auto &[cds_start, cds_stop, is_fwd] : cdss) {
    const size_t n = cds_stop - cds_start;
    std::vector<DomainSpec> doms;
    bool useful = false;
    size_t i = 0;
    while (i + dom_size <= n) {
      const size_t d0 = cds_start + i;
      const int c0 = code[d0], c1 = code[d0 + CODON];
      const int type = (c0 | c1) < 0 ? -1 : maps.type_tab[c0 * 64 + c1];
      if (type < 0) {
        i += CODON;
        continue;
      }
      if (type != 3) useful = true;
      const size_t t0 = d0 + dom_type_size;
      int tok[3];
      for (size_t k = 0; k < 3; k++) {
        const int c = code[t0 + k * CODON];
        tok[k] = c < 0 ? -1 : maps.one_tab[c];
        if (tok[k] < 0) throw std::runtime_error("Incomplete one_codon_map");
      }
      const int e0 = code[t0 + 3 * CODON], e1 = code[t0 + 4 * CODON];
      const int tok2 = (e0 | e1) < 0 ? -1 : maps.two_tab[e0 * 64 + e1];
      if (tok2 < 0) throw std::runtime_error("Incomplete two_codon_map");
      doms.emplace_back(std::make_tuple(type, tok[0], tok[1], tok[2], tok2), i, i + dom_size);
      i += dom_size;
    }
    if (useful) out.emplace_back(std::move(doms), cds_start, cds_stop, is_fwd);
  }
  return out;
}

std::string reverse_complement(const std::string &seq) {
  std::string out;
  out.reserve(seq.size());
  for (auto it = seq.rbegin(); it != seq.rend(); ++it) {
    switch (*it) {
      case 'A': out.push_back('T'); break;
      case 'C': out.push_back('G'); break;
      case 'T': out.push_back('A'); break;
      case 'G': out.push_back('C'); break;
      default: break;  // dropped, like the reference
    }
  }
  return out;
}

std::vector<std::vector<ProteinSpec>> translate_genomes(const std::vector<std::string> &genomes,
                                                        const std::vector<std::string> &start_codons,
                                                        const std::vector<std::string> &stop_codons,
                                                        const CodonMaps &maps, size_t dom_size,
                                                        size_t dom_type_size) {
  const auto starts = codon_table(start_codons), stops = codon_table(stop_codons);
  std::vector<std::vector<ProteinSpec>> out(genomes.size());
  std::string error;
#pragma omp parallel for schedule(dynamic, 16)
  for (long long g = 0; g < (long long)genomes.size(); g++) {
    try {
      const std::string &fwd = genomes[g];
      auto prots = domains_of(fwd, coding_regions(fwd, dom_size, starts, stops, true), dom_size, dom_type_size, maps);
      const std::string bwd = reverse_complement(fwd);
      auto more = domains_of(bwd, coding_regions(bwd, dom_size, starts, stops, false), dom_size, dom_type_size, maps);
      prots.insert(prots.end(), std::make_move_iterator(more.begin()), std::make_move_iterator(more.end()));
      out[g] = std::move(prots);
    } catch (const std::exception &e) {
#pragma omp critical
      error = e.what();
    }
  }
  if (!error.empty()) throw std::runtime_error(error);
  return out;
}

// ---------------------------------------------------------------------------- mutations
const char NTS[4] = {'A', 'C', 'T', 'G'};

// k distinct indices out of n, ascending
template <class Rng>
std::vector<size_t> sample_sorted(Rng &rng, size_t n, size_t k) {
  std::vector<size_t> idx;
  if (k * 4 >= n) {  // dense: partial shuffle
    std::vector<size_t> all(n);
    for (size_t i = 0; i < n; i++) all[i] = i;
    for (size_t i = 0; i < k; i++) std::swap(all[i], all[i + std::uniform_int_distribution<size_t>(0, n - 1 - i)(rng)]);
    idx.assign(all.begin(), all.begin() + k);
  } else {  // sparse: rejection
    std::unordered_set<size_t> seen;
    std::uniform_int_distribution<size_t> pick(0, n - 1);
    while (seen.size() < k) seen.insert(pick(rng));
    idx.assign(seen.begin(), seen.end());
  }
  std::sort(idx.begin(), idx.end());
  return idx;
}

// number of events ~ Poisson(lambda), at most cap.  Small rates (the usual case: 1e-6 per base)
// by inversion of the distribution function from one uniform number.
template <class Rng>
size_t poisson_count(Rng &rng, double lambda, size_t cap) {
  if (!(lambda > 0.0)) return 0;
  if (lambda < 16.0) {
    const double u = std::generate_canonical<double, 53>(rng);
    if (u < 1.0 - lambda) return 0;  // 1 - lambda <= exp(-lambda): the same answer without the exponential
    double pk = std::exp(-lambda), cdf = pk;
    size_t n = 0;
    while (u >= cdf && n < cap && pk > 0.0) {
      n++;
      pk *= lambda / (double)n;
      cdf += pk;
    }
    return std::min(n, cap);
  }
  const size_t n = (size_t)std::poisson_distribution<long long>(lambda)(rng);
  return std::min(n, cap);
}

// rust/mutations.rs:11-72: Poisson(p * len) positions; each an indel with p_indel (deletion with
// p_del, else insertion of a random nucleotide before it), else a substitution.
// The number of events depends on the length alone, so the callers below draw it for every
// sequence first and copy only the sequences that are hit out of their Python strings: at the
// reference's default rates (1e-6 per base) that is a handful of genomes out of thousands.
std::string mutate_one(const std::string &src, SeqRng &rng, size_t n_muts, double p_indel, double p_del) {
  const auto where = sample_sorted(rng, src.size(), std::min(n_muts, src.size()));
  const double q_indel = std::clamp(p_indel, 0.0, 1.0), q_del = std::clamp(p_del, 0.0, 1.0);
  std::string out;
  out.reserve(src.size() + n_muts);
  size_t next = 0;
  for (size_t i = 0; i < src.size(); i++) {
    if (next < where.size() && where[next] == i) {
      next++;
      if (rng.uniform() < q_indel) {
        if (rng.uniform() < q_del) continue;  // deletion
        out.push_back(NTS[rng() >> 62]);      // insertion before position i
        out.push_back(src[i]);
      } else {
        out.push_back(NTS[rng() >> 62]);      // substitution (may draw the same nucleotide, like the reference)
      }
    } else {
      out.push_back(src[i]);
    }
  }
  return out;
}

// rust/mutations.rs:78-154: Poisson(p * (n0 + n1)) strand breaks over both sequences, the
// fragments are shuffled and re-joined into two sequences at a random cut of the fragment list
std::pair<std::string, std::string> recombine_one(const std::string &a, const std::string &b, SeqRng &rng,
                                                  size_t n_breaks) {
  const size_t n0 = a.size(), n_both = a.size() + b.size();
  const auto where = sample_sorted(rng, n_both, std::min(n_breaks, n_both));
  std::vector<std::string> parts;
  parts.reserve(where.size() + 2);
  size_t i = 0, k = 0;
  for (; k < where.size() && where[k] < n0; k++) {
    parts.push_back(a.substr(i, where[k] - i));
    i = where[k];
  }
  parts.push_back(a.substr(i));
  i = 0;
  for (; k < where.size(); k++) {
    parts.push_back(b.substr(i, where[k] - n0 - i));
    i = where[k] - n0;
  }
  parts.push_back(b.substr(i));
  std::shuffle(parts.begin(), parts.end(), rng);
  const size_t cut = (size_t)(rng.uniform() * (double)parts.size());
  std::string x, y;
  for (size_t q = 0; q < parts.size(); q++) (q < cut ? x : y) += parts[q];
  return {std::move(x), std::move(y)};
}

// a Python str as (length in code points) / (UTF-8 copy); genomes are ASCII, so both agree
size_t str_len(PyObject *o) {
  if (!PyUnicode_Check(o)) throw py::type_error("expected str");
  return (size_t)PyUnicode_GET_LENGTH(o);
}
std::string str_copy(PyObject *o) {
  Py_ssize_t n = 0;
  const char *d = PyUnicode_AsUTF8AndSize(o, &n);
  if (!d) throw py::error_already_set();
  return std::string(d, (size_t)n);
}

std::vector<std::pair<std::string, size_t>> point_mutations(const py::sequence &seqs, double p, double p_indel,
                                                            double p_del) {
  const py::object fast = py::reinterpret_steal<py::object>(PySequence_Fast(seqs.ptr(), "seqs must be a sequence of str"));
  if (!fast) throw py::error_already_set();
  const size_t n = (size_t)PySequence_Fast_GET_SIZE(fast.ptr());
  const uint64_t seed = next_seed();
  std::vector<size_t> hit, hit_len;
  std::vector<std::string> src;
  for (size_t s = 0; s < n; s++) {
    PyObject *o = PySequence_Fast_GET_ITEM(fast.ptr(), (Py_ssize_t)s);
    const size_t len = str_len(o);
    if (len == 0) continue;
    SeqRng rng(seed, s);
    if (poisson_count(rng, p * (double)len, len) == 0) continue;
    hit.push_back(s);
    hit_len.push_back(len);
    src.push_back(str_copy(o));
  }
  std::vector<std::pair<std::string, size_t>> res(hit.size());
  {
    py::gil_scoped_release release;
#pragma omp parallel for schedule(dynamic, 16) if (hit.size() > 64)
    for (long long h = 0; h < (long long)hit.size(); h++) {
      SeqRng rng(seed, hit[h]);
      const size_t n_muts = poisson_count(rng, p * (double)hit_len[h], hit_len[h]);  // the same draw again
      res[h] = {mutate_one(src[h], rng, n_muts, p_indel, p_del), hit[h]};
    }
  }
  return res;
}

// `pair_at(s)` yields the two Python strings of pair s
template <class PairAt>
std::vector<std::tuple<std::string, std::string, size_t>> recombine_pairs(size_t n_pairs, double p, PairAt pair_at) {
  const uint64_t seed = next_seed();
  std::vector<size_t> hit, hit_len;
  std::vector<std::pair<std::string, std::string>> src;
  for (size_t s = 0; s < n_pairs; s++) {
    const std::pair<PyObject *, PyObject *> ab = pair_at(s);
    const size_t n_both = str_len(ab.first) + str_len(ab.second);
    if (n_both == 0) continue;
    SeqRng rng(seed, s);
    if (poisson_count(rng, p * (double)n_both, n_both) == 0) continue;
    hit.push_back(s);
    hit_len.push_back(n_both);
    src.emplace_back(str_copy(ab.first), str_copy(ab.second));
  }
  std::vector<std::tuple<std::string, std::string, size_t>> res(hit.size());
  {
    py::gil_scoped_release release;
#pragma omp parallel for schedule(dynamic, 16) if (hit.size() > 64)
    for (long long h = 0; h < (long long)hit.size(); h++) {
      SeqRng rng(seed, hit[h]);
      const size_t n_breaks = poisson_count(rng, p * (double)hit_len[h], hit_len[h]);
      auto xy = recombine_one(src[h].first, src[h].second, rng, n_breaks);
      res[h] = std::make_tuple(std::move(xy.first), std::move(xy.second), hit[h]);
    }
  }
  return res;
}

std::vector<std::tuple<std::string, std::string, size_t>> recombinations(const py::sequence &seq_pairs, double p) {
  const py::object fast = py::reinterpret_steal<py::object>(PySequence_Fast(seq_pairs.ptr(), "seq_pairs must be a sequence"));
  if (!fast) throw py::error_already_set();
  const size_t n = (size_t)PySequence_Fast_GET_SIZE(fast.ptr());
  return recombine_pairs(n, p, [&](size_t s) {
    PyObject *pair = PySequence_Fast_GET_ITEM(fast.ptr(), (Py_ssize_t)s);
    if (PyTuple_Check(pair) && PyTuple_GET_SIZE(pair) == 2)
      return std::pair<PyObject *, PyObject *>(PyTuple_GET_ITEM(pair, 0), PyTuple_GET_ITEM(pair, 1));
    if (PyList_Check(pair) && PyList_GET_SIZE(pair) == 2)
      return std::pair<PyObject *, PyObject *>(PyList_GET_ITEM(pair, 0), PyList_GET_ITEM(pair, 1));
    throw py::type_error("seq_pairs must hold pairs of str");
  });
}

// recombinations((genomes[a], genomes[b]) for a, b in pairs) without building the pairs of strings.
// The lengths of all genomes are read once; only the pairs that are hit touch their strings.
std::vector<std::tuple<std::string, std::string, size_t>> recombinations_indexed(
    const py::list &genomes, const std::vector<std::pair<size_t, size_t>> &pairs, double p) {
  const size_t n_genomes = (size_t)py::len(genomes);
  for (const auto &ab : pairs)
    if (ab.first >= n_genomes || ab.second >= n_genomes) throw py::index_error("genome index out of range");
  std::vector<size_t> len(n_genomes);
  for (size_t g = 0; g < n_genomes; g++) len[g] = str_len(PyList_GET_ITEM(genomes.ptr(), (Py_ssize_t)g));
  const uint64_t seed = next_seed();
  std::vector<size_t> hit;
  {
    const size_t chunk = 8192, n_chunks = (pairs.size() + chunk - 1) / chunk;
    std::vector<std::vector<size_t>> parts(n_chunks);
#pragma omp parallel for schedule(dynamic, 1) if (n_chunks > 2)
    for (long long c = 0; c < (long long)n_chunks; c++) {
      const size_t s1 = std::min(pairs.size(), ((size_t)c + 1) * chunk);
      for (size_t s = (size_t)c * chunk; s < s1; s++) {
        const size_t n_both = len[pairs[s].first] + len[pairs[s].second];
        if (n_both == 0) continue;
        SeqRng rng(seed, s);
        if (poisson_count(rng, p * (double)n_both, n_both) > 0) parts[(size_t)c].push_back(s);
      }
    }
    for (const auto &part : parts) hit.insert(hit.end(), part.begin(), part.end());
  }
  std::vector<std::pair<std::string, std::string>> src;
  src.reserve(hit.size());
  for (const size_t s : hit)
    src.emplace_back(str_copy(PyList_GET_ITEM(genomes.ptr(), (Py_ssize_t)pairs[s].first)),
                     str_copy(PyList_GET_ITEM(genomes.ptr(), (Py_ssize_t)pairs[s].second)));
  std::vector<std::tuple<std::string, std::string, size_t>> res(hit.size());
  {
    py::gil_scoped_release release;
#pragma omp parallel for schedule(dynamic, 16) if (hit.size() > 64)
    for (long long h = 0; h < (long long)hit.size(); h++) {
      const size_t s = hit[h], n_both = len[pairs[s].first] + len[pairs[s].second];
      SeqRng rng(seed, s);
      const size_t n_breaks = poisson_count(rng, p * (double)n_both, n_both);  // the same draw again
      auto xy = recombine_one(src[h].first, src[h].second, rng, n_breaks);
      res[h] = std::make_tuple(std::move(xy.first), std::move(xy.second), s);
    }
  }
  return res;
}

// ---------------------------------------------------------------------------- world
inline uint32_t pixel_key(uint16_t x, uint16_t y) { return ((uint32_t)x << 16) | y; }

uint16_t dist_1d(uint16_t a, uint16_t b, uint16_t m) {
  const uint16_t hi = std::max(a, b), lo = std::min(a, b);
  const uint16_t d0 = hi - lo, d1 = (uint16_t)(m - hi + lo);
  return std::min(d0, d1);
}

// The map the searches run on.  The reference's maps are square tori (`map_size`); a row slab of a
// map that is split over several GPUs (magicsoup_b200/distributed.py) is `rows x cols`, wraps
// around in its columns only, and what lies beyond its first and last row belongs to other ranks.
struct MapShape {
  uint16_t rows, cols;
  bool wrap_rows;
};

struct Moore {
  std::array<Pos, 8> p;
  int n;
  const Pos *begin() const { return p.data(); }
  const Pos *end() const { return p.data() + n; }
};

// rust/util.rs:27-43: W-N, W, W-S, N, S, E-N, E, E-S with "east" = x + 1, "south" = y + 1
Moore moore(uint16_t x, uint16_t y, const MapShape &m) {
  const bool has_e = m.wrap_rows || x + 1 < m.rows, has_w = m.wrap_rows || x > 0;
  const uint16_t e = (uint16_t)(x + 1 >= m.rows ? 0 : x + 1), w = (uint16_t)(x == 0 ? m.rows - 1 : x - 1);
  const uint16_t s = (uint16_t)(y + 1 >= m.cols ? 0 : y + 1), n = (uint16_t)(y == 0 ? m.cols - 1 : y - 1);
  Moore out{};
  out.n = 0;
  if (has_w) {
    out.p[out.n++] = {w, n};
    out.p[out.n++] = {w, y};
    out.p[out.n++] = {w, s};
  }
  out.p[out.n++] = {x, n};
  out.p[out.n++] = {x, s};
  if (has_e) {
    out.p[out.n++] = {e, n};
    out.p[out.n++] = {e, y};
    out.p[out.n++] = {e, s};
  }
  return out;
}

std::vector<Pos> free_moores_nghbhd(uint16_t x, uint16_t y, const std::vector<Pos> &positions, uint16_t map_size_) {
  const MapShape map_size{map_size_, map_size_, true};
  std::unordered_set<uint32_t> occ;
  occ.reserve(positions.size() * 2);
  for (const auto &p : positions) occ.insert(pixel_key(p.first, p.second));
  std::vector<Pos> out;
  for (const auto &p : moore(x, y, map_size))
    if (!occ.count(pixel_key(p.first, p.second))) out.push_back(p);
  return out;
}

// One value per pixel of a map: a dense array where that is small next to the number of
// entries it serves (4 Mi pixels, or 16 per entry), a hash map on huge, sparsely populated maps.
template <class T>
struct PixelTable {
  size_t rows, m;  // m: columns (the row stride)
  T fill;
  bool dense;
  std::vector<T> flat;
  std::unordered_map<uint64_t, T> sparse;
  PixelTable(const MapShape &shape, size_t n_entries, T fill_value) : rows(shape.rows), m(shape.cols), fill(fill_value) {
    const size_t n_pixels = rows * m;
    dense = n_pixels <= ((size_t)1 << 22) || n_pixels <= 16 * n_entries;
    if (dense) flat.assign(n_pixels, fill);
    else sparse.reserve(2 * n_entries);
  }
  void check(const Pos &p) const {
    if (p.first >= rows || p.second >= m) throw py::index_error("position outside the map");
  }
  T &at(const Pos &p) {  // created with the fill value if absent
    const size_t k = (size_t)p.first * m + p.second;
    if (dense) return flat[k];
    return sparse.try_emplace(k, fill).first->second;
  }
  T get(const Pos &p) const {
    const size_t k = (size_t)p.first * m + p.second;
    if (dense) return flat[k];
    const auto it = sparse.find(k);
    return it == sparse.end() ? fill : it->second;
  }
};

// first occurrences of a list of indices, order kept
std::vector<size_t> unique_in_order(const std::vector<size_t> &idxs, size_t n) {
  std::vector<char> seen(n, 0);
  std::vector<size_t> out;
  out.reserve(idxs.size());
  for (const size_t i : idxs) {
    if (i >= n) throw py::index_error("cell index out of range");
    if (!seen[i]) {
      seen[i] = 1;
      out.push_back(i);
    }
  }
  return out;
}

// rust/world.rs:9-56: unique pairs (smaller index first) of cells from `from_idxs` and `to_idxs` at
// torus distance <= 1 in both coordinates, in the reference's order (from_idxs order, then
// to_idxs order, first occurrence kept).  The reference compares every `from` with every `to`;
// here the `to` cells are chained per pixel of an occupancy grid and a `from` cell looks at its
// 3 x 3 block.  A pair is found twice exactly when both cells are in both lists: the second find
// is recognised by the ranks of the two cells in from_idxs, no set of seen pairs is kept.
std::vector<std::pair<size_t, size_t>> get_neighbors(const std::vector<size_t> &from_raw,
                                                     const std::vector<size_t> &to_raw,
                                                     const std::vector<Pos> &positions, const MapShape &map_size) {
  const size_t n = positions.size();
  const std::vector<size_t> from_idxs = unique_in_order(from_raw, n), to_idxs = unique_in_order(to_raw, n);
  constexpr int64_t NONE = -1;
  std::vector<int64_t> from_rank(n, NONE), to_rank(n, NONE);
  for (size_t r = 0; r < from_idxs.size(); r++) from_rank[from_idxs[r]] = (int64_t)r;
  for (size_t r = 0; r < to_idxs.size(); r++) to_rank[to_idxs[r]] = (int64_t)r;
  // pixel -> chain of ranks in to_idxs (ascending rank along the chain)
  const size_t m = map_size.cols;
  PixelTable<int64_t> head(map_size, to_idxs.size(), NONE);
  std::vector<int64_t> next(to_idxs.size(), NONE);
  for (size_t r = to_idxs.size(); r-- > 0;) {
    const Pos &q = positions[to_idxs[r]];
    head.check(q);
    int64_t &h = head.at(q);
    next[r] = h;
    h = (int64_t)r;
  }
  for (const size_t from : from_idxs) head.check(positions[from]);
  // the from cells in chunks, one output list per chunk (joined in order: the result does not depend on the threads)
  const size_t n_from = from_idxs.size(), chunk = 2048, n_chunks = (n_from + chunk - 1) / chunk;
  std::vector<std::vector<std::pair<size_t, size_t>>> parts(n_chunks);
#pragma omp parallel for schedule(dynamic, 1) if (n_chunks > 2)
  for (long long c = 0; c < (long long)n_chunks; c++) {
    auto &out = parts[(size_t)c];
    std::vector<int64_t> ranks;
    const size_t f1 = std::min(n_from, ((size_t)c + 1) * chunk);
    for (size_t f = (size_t)c * chunk; f < f1; f++) {
      const size_t from = from_idxs[f];
      const Pos &p0 = positions[from];
      std::array<uint32_t, 9> pixels;  // the 3 x 3 block, without duplicates on tiny maps
      size_t n_px = 0;
      pixels[n_px++] = pixel_key(p0.first, p0.second);
      for (const auto &q : moore(p0.first, p0.second, map_size)) pixels[n_px++] = pixel_key(q.first, q.second);
      if (m < 3 || map_size.rows < 3) {
        std::sort(pixels.begin(), pixels.begin() + n_px);
        n_px = (size_t)(std::unique(pixels.begin(), pixels.begin() + n_px) - pixels.begin());
      }
      ranks.clear();
      for (size_t k = 0; k < n_px; k++)
        for (int64_t r = head.get({(uint16_t)(pixels[k] >> 16), (uint16_t)(pixels[k] & 0xffffu)}); r != NONE; r = next[r])
          ranks.push_back(r);
      std::sort(ranks.begin(), ranks.end());
      for (const int64_t r : ranks) {
        const size_t to = to_idxs[(size_t)r];
        if (to == from) continue;
        // found before, from the other side: `to` came earlier in from_idxs and `from` is a `to` cell
        if (from_rank[to] != NONE && from_rank[to] < from_rank[from] && to_rank[from] != NONE) continue;
        out.emplace_back(std::min(from, to), std::max(from, to));
      }
    }
  }
  size_t total = 0;
  for (const auto &part : parts) total += part.size();
  std::vector<std::pair<size_t, size_t>> out;
  out.reserve(total);
  for (const auto &part : parts) out.insert(out.end(), part.begin(), part.end());
  return out;
}

// cells per pixel of a square map
struct Occupancy {
  PixelTable<int32_t> count;
  Occupancy(const std::vector<Pos> &positions, const MapShape &map_size) : count(map_size, positions.size(), 0) {
    for (const auto &p : positions) {
      count.check(p);
      count.at(p)++;
    }
  }
  int32_t &at(const Pos &p) { return count.at(p); }
  int32_t get(const Pos &p) const { return count.get(p); }
};

// rust/world.rs:59-99: in the order of cell_idxs, every cell picks one free Moore pixel at random
// (free of all current cells and of the children placed before it); children get indices n_cells..
std::tuple<std::vector<size_t>, std::vector<size_t>, std::vector<Pos>> divide_cells_if_possible(
    const std::vector<size_t> &cell_idxs, const std::vector<Pos> &positions, size_t n_cells, const MapShape &map_size) {
  Occupancy occ(positions, map_size);
  SeqRng rng(next_seed(), 0);
  std::vector<size_t> parents, children;
  std::vector<Pos> child_pos;
  size_t next = n_cells;
  for (const size_t idx : cell_idxs) {
    const Pos &p0 = positions.at(idx);
    std::array<Pos, 8> opts;
    size_t n_opts = 0;
    for (const auto &p : moore(p0.first, p0.second, map_size))
      if (occ.get(p) == 0) opts[n_opts++] = p;
    if (n_opts == 0) continue;
    const Pos pick = opts[(size_t)(rng.uniform() * (double)n_opts)];
    occ.at(pick)++;
    parents.push_back(idx);
    children.push_back(next++);
    child_pos.push_back(pick);
  }
  return {parents, children, child_pos};
}

// rust/world.rs:102-146: in the order of cell_idxs, every cell moves to one free Moore pixel at
// random: free of the cells that do not move and of the moving cells' current pixels
std::tuple<std::vector<Pos>, std::vector<size_t>> move_cells(const std::vector<size_t> &cell_idxs,
                                                             const std::vector<Pos> &positions,
                                                             const MapShape &map_size) {
  Occupancy occ(positions, map_size);
  SeqRng rng(next_seed(), 0);
  std::vector<Pos> new_pos;
  std::vector<size_t> moved;
  for (const size_t idx : cell_idxs) {
    const Pos p0 = positions.at(idx);
    std::array<Pos, 8> opts;
    size_t n_opts = 0;
    for (const auto &p : moore(p0.first, p0.second, map_size))
      if (occ.get(p) == 0) opts[n_opts++] = p;
    if (n_opts == 0) continue;
    const Pos pick = opts[(size_t)(rng.uniform() * (double)n_opts)];
    occ.at(p0)--;
    occ.at(pick)++;
    new_pos.push_back(pick);
    moved.push_back(idx);
  }
  return {new_pos, moved};
}

// `map_size` as the reference passes it (an int: a square torus) or (rows, cols[, wrap_rows]) for a row slab
MapShape shape_from(const py::object &obj) {
  const auto dim = [](long long v) {
    if (v < 1 || v > 0xffff) throw py::value_error("map sizes must be in 1..65535");
    return (uint16_t)v;
  };
  if (py::isinstance<py::int_>(obj)) {
    const uint16_t m = dim(obj.cast<long long>());
    return MapShape{m, m, true};
  }
  const auto seq = obj.cast<py::sequence>();
  if (py::len(seq) != 2 && py::len(seq) != 3) throw py::value_error("map_size: int or (rows, cols[, wrap_rows])");
  return MapShape{dim(seq[0].cast<long long>()), dim(seq[1].cast<long long>()), py::len(seq) == 3 ? seq[2].cast<bool>() : true};
}

// positions as a list of (x, y) pairs (the reference's interface) or as an (n, 2) integer array
std::vector<Pos> positions_from(const py::object &obj) {
  if (py::isinstance<py::array>(obj)) {
    const auto arr = py::array_t<int64_t, py::array::c_style | py::array::forcecast>::ensure(obj);
    if (!arr || arr.ndim() != 2 || arr.shape(1) != 2) throw py::value_error("positions must have shape (n, 2)");
    const auto v = arr.unchecked<2>();
    std::vector<Pos> out((size_t)arr.shape(0));
    for (py::ssize_t i = 0; i < arr.shape(0); i++) {
      if (v(i, 0) < 0 || v(i, 0) > 0xffff || v(i, 1) < 0 || v(i, 1) > 0xffff) throw py::index_error("position outside the map");
      out[(size_t)i] = {(uint16_t)v(i, 0), (uint16_t)v(i, 1)};
    }
    return out;
  }
  return obj.cast<std::vector<Pos>>();
}

// rust/kinetics.rs:17-201: keyword arguments for the reference's Protein / domain containers
// (containers.py) from a proteome's index tuples and the values Kinetics looked up for them.
// Per protein i and domain j: vmaxs[i][j], kms[i][j], hills[i][j], signs[i][j] and the three
// signal vectors reacts / trnspts / effectors [i][j][:].
py::list get_proteome(const std::vector<ProteinSpec> &proteome, const std::vector<std::vector<float>> &vmaxs,
                      const std::vector<std::vector<float>> &kms, const std::vector<std::vector<int>> &hills,
                      const std::vector<std::vector<int>> &signs,
                      const std::vector<std::vector<std::vector<int>>> &reacts,
                      const std::vector<std::vector<std::vector<int>>> &trnspts,
                      const std::vector<std::vector<std::vector<int>>> &effectors,
                      const std::vector<std::string> &molecules) {
  const size_t n_mols = molecules.size();
  const auto first_nonzero = [](const std::vector<int> &v, const char *what) {
    for (size_t i = 0; i < v.size(); i++)
      if (v[i] != 0) return i;
    throw std::runtime_error(what);
  };
  py::list out;
  for (size_t i = 0; i < proteome.size(); i++) {
    const auto &[doms, cds_start, cds_end, is_fwd] = proteome[i];
    py::list domains;
    for (size_t j = 0; j < doms.size(); j++) {
      const auto &[idxs, start, end] = doms[j];
      const int kind = std::get<0>(idxs);  // 1 = catalytic, 2 = transporter, 3 = regulatory
      const int sign = signs.at(i).at(j);
      py::dict spec;
      spec["start"] = start;
      spec["end"] = end;
      if (kind == 1) {
        py::list lhs, rhs;  // a molecule appears |n| times on its side
        const auto &vec = reacts.at(i).at(j);
        for (size_t k = 0; k < vec.size(); k++) {
          const int n = vec[k] * sign;
          if (n == 0) continue;
          const py::str name(molecules.at(k));
          for (int r = 0; r < std::abs(vec[k]); r++) (n > 0 ? rhs : lhs).append(name);
        }
        spec["km"] = kms.at(i).at(j);
        spec["vmax"] = vmaxs.at(i).at(j);
        spec["reaction"] = py::make_tuple(lhs, rhs);
      } else if (kind == 2) {
        const auto &vec = trnspts.at(i).at(j);
        const size_t k = first_nonzero(vec, "No transporter molecule identified");
        spec["km"] = kms.at(i).at(j);
        spec["vmax"] = vmaxs.at(i).at(j);
        spec["is_exporter"] = vec[k] * sign < 0;
        spec["molecule"] = molecules.at(k);
      } else if (kind == 3) {
        const auto &vec = effectors.at(i).at(j);
        const size_t k = first_nonzero(vec, "No effector molecule identified");
        spec["km"] = kms.at(i).at(j);
        spec["hill"] = hills.at(i).at(j);
        spec["is_transmembrane"] = k >= n_mols;  // the second half of the signals is extracellular
        spec["is_inhibiting"] = vec[k] * sign < 0;
        spec["effector"] = molecules.at(k >= n_mols ? k - n_mols : k);
      }
      py::dict dom;
      dom["spec"] = spec;
      dom["type"] = kind == 1 ? "C" : kind == 2 ? "T" : kind == 3 ? "R" : " ";
      domains.append(dom);
    }
    py::dict prot;
    prot["cds_start"] = cds_start;
    prot["cds_end"] = cds_end;
    prot["is_fwd"] = is_fwd;
    prot["domains"] = domains;
    out.append(prot);
  }
  return out;
}

CodonMaps make_maps(const std::unordered_map<std::string, int> &dom_type, const std::unordered_map<std::string, int> &one,
                    const std::unordered_map<std::string, int> &two) {
  CodonMaps m{dom_type, one, two};
  make_dense_tables(m);
  return m;
}

}  // namespace

PYBIND11_MODULE(_lib, m) {
  m.doc() = "host-side helpers with the interface of magicsoup._lib (C++17, pybind11)";
  m.def("seed", [](uint64_t s) {
    std::lock_guard<std::mutex> lock(g_rng_mutex);
    g_rng.seed(s);
  }, py::arg("seed"), "seed the generator behind the random choices (the reference's is not seedable)");
  m.def("dist_1d", &dist_1d, py::arg("a"), py::arg("b"), py::arg("m"));
  m.def("free_moores_nghbhd", &free_moores_nghbhd, py::arg("x"), py::arg("y"), py::arg("positions"), py::arg("map_size"));
  m.def("point_mutations", &point_mutations, py::arg("seqs"), py::arg("p"), py::arg("p_indel"), py::arg("p_del"));
  m.def("recombinations", &recombinations, py::arg("seq_pairs"), py::arg("p"));
  m.def("recombinations_indexed", &recombinations_indexed, py::arg("genomes"), py::arg("pairs"), py::arg("p"),
        "no counterpart in the reference: recombinations() over index pairs into one list of genomes");
  m.def("get_coding_regions",
        [](const std::string &seq, size_t min_cds_size, const std::vector<std::string> &start_codons,
           const std::vector<std::string> &stop_codons, bool is_fwd) {
          return coding_regions(seq, min_cds_size, codon_table(start_codons), codon_table(stop_codons), is_fwd);
        },
        py::arg("seq"), py::arg("min_cds_size"), py::arg("start_codons"), py::arg("stop_codons"), py::arg("is_fwd"));
  m.def("extract_domains",
        [](const std::string &genome, const std::vector<Cds> &cdss, size_t dom_size, size_t dom_type_size,
           const std::unordered_map<std::string, int> &dom_type_map, const std::unordered_map<std::string, int> &one_codon_map,
           const std::unordered_map<std::string, int> &two_codon_map) {
          return domains_of(genome, cdss, dom_size, dom_type_size, make_maps(dom_type_map, one_codon_map, two_codon_map));
        },
        py::arg("genome"), py::arg("cdss"), py::arg("dom_size"), py::arg("dom_type_size"), py::arg("dom_type_map"),
        py::arg("one_codon_map"), py::arg("two_codon_map"));
  m.def("reverse_complement", &reverse_complement, py::arg("seq"));
  m.def("translate_genomes",
        [](const std::vector<std::string> &genomes, const std::vector<std::string> &start_codons,
           const std::vector<std::string> &stop_codons, const std::unordered_map<std::string, int> &domain_map,
           const std::unordered_map<std::string, int> &one_codon_map,
           const std::unordered_map<std::string, int> &two_codon_map, size_t dom_size, size_t dom_type_size) {
          const CodonMaps maps = make_maps(domain_map, one_codon_map, two_codon_map);
          py::gil_scoped_release release;
          return translate_genomes(genomes, start_codons, stop_codons, maps, dom_size, dom_type_size);
        },
        py::arg("genomes"), py::arg("start_codons"), py::arg("stop_codons"), py::arg("domain_map"),
        py::arg("one_codon_map"), py::arg("two_codon_map"), py::arg("dom_size"), py::arg("dom_type_size"));
  m.def("translate_genomes_flat",
        [](const std::vector<std::string> &genomes, const std::vector<std::string> &start_codons,
           const std::vector<std::string> &stop_codons, const std::unordered_map<std::string, int> &domain_map,
           const std::unordered_map<std::string, int> &one_codon_map,
           const std::unordered_map<std::string, int> &two_codon_map, size_t dom_size, size_t dom_type_size) {
          // translate_genomes + flatten_proteomes without the nested Python lists in between
          // (half a million objects per 10,000 genomes): what World._update_cell_params needs
          const CodonMaps maps = make_maps(domain_map, one_codon_map, two_codon_map);
          std::vector<int32_t> prot_off{0}, dom_off{0}, doms;
          {
            py::gil_scoped_release release;
            const auto all = translate_genomes(genomes, start_codons, stop_codons, maps, dom_size, dom_type_size);
            for (const auto &proteome : all) {
              for (const auto &protein : proteome) {
                for (const auto &dom : std::get<0>(protein)) {
                  const auto &t = std::get<0>(dom);
                  doms.insert(doms.end(), {(int32_t)std::get<0>(t), (int32_t)std::get<1>(t), (int32_t)std::get<2>(t),
                                           (int32_t)std::get<3>(t), (int32_t)std::get<4>(t)});
                }
                dom_off.push_back((int32_t)(doms.size() / 5));
              }
              prot_off.push_back((int32_t)(dom_off.size() - 1));
            }
          }
          auto arr = [](const std::vector<int32_t> &v) { return py::array_t<int32_t>(v.size(), v.data()); };
          return py::make_tuple(arr(prot_off), arr(dom_off), arr(doms));
        },
        py::arg("genomes"), py::arg("start_codons"), py::arg("stop_codons"), py::arg("domain_map"),
        py::arg("one_codon_map"), py::arg("two_codon_map"), py::arg("dom_size"), py::arg("dom_type_size"),
        "no counterpart in the reference: the three int32 arrays of flatten_proteomes(translate_genomes(...))");
  m.def("flatten_proteomes",
        [](const py::list &proteomes) {
          // nested (proteome -> protein -> domain) lists of translate_genomes -> three int32 arrays:
          // protein offsets per proteome, domain offsets per protein, (type, i0, i1, i2, i3) per domain
          std::vector<int32_t> prot_off{0}, dom_off{0}, doms;
          for (const auto &proteome : proteomes) {
            for (const auto &protein : py::reinterpret_borrow<py::sequence>(proteome)) {
              const auto prot = py::reinterpret_borrow<py::sequence>(protein);
              for (const auto &dom : py::reinterpret_borrow<py::sequence>(prot[0])) {
                const auto spec = py::reinterpret_borrow<py::sequence>(py::reinterpret_borrow<py::sequence>(dom)[0]);
                for (int q = 0; q < 5; q++) doms.push_back(spec[q].cast<int32_t>());
              }
              dom_off.push_back((int32_t)(doms.size() / 5));
            }
            prot_off.push_back((int32_t)(dom_off.size() - 1));
          }
          auto arr = [](const std::vector<int32_t> &v) { return py::array_t<int32_t>(v.size(), v.data()); };
          return py::make_tuple(arr(prot_off), arr(dom_off), arr(doms));
        },
        py::arg("proteomes"), "no counterpart in the reference (which pads nested lists into (c,p,d) tensors, kinetics.py:920-970)");
  m.def("remove_items",
        [](py::list items, const std::vector<int64_t> &dead_sorted) {
          // order-preserving removal from a Python list in place (World.kill_cells' genome and
          // label lists, world.py:527-540): the kept pointers slide down in one pass and only the
          // removed objects are touched; a pop per index moves the tail every time, and building
          // a new list touches the reference count of every kept object (a cache miss each)
          PyListObject *lst = (PyListObject *)items.ptr();
          const int64_t n = (int64_t)Py_SIZE(lst);
          int64_t prev = -1;
          for (int64_t i : dead_sorted) {
            if (i < 0 || i >= n) throw py::index_error("index out of range");
            if (i <= prev) throw py::value_error("indices must be strictly ascending");
            prev = i;
          }
          std::vector<PyObject *> removed;
          removed.reserve(dead_sorted.size());
          PyObject **it = lst->ob_item;
          int64_t o = 0, a = 0;
          for (int64_t i : dead_sorted) {
            if (i > a) std::memmove(it + o, it + a, (size_t)(i - a) * sizeof(PyObject *));
            o += i - a;
            removed.push_back(it[i]);
            a = i + 1;
          }
          if (n > a) std::memmove(it + o, it + a, (size_t)(n - a) * sizeof(PyObject *));
          o += n - a;
          Py_SET_SIZE(lst, o);
          for (PyObject *r : removed) Py_DECREF(r);  // after the list is consistent again
        },
        py::arg("items"), py::arg("dead_sorted"));
  m.def("get_proteome", &get_proteome, py::arg("proteome"), py::arg("vmaxs"), py::arg("kms"), py::arg("hills"),
        py::arg("signs"), py::arg("reacts"), py::arg("trnspts"), py::arg("effectors"), py::arg("molecules"));
  m.def("get_neighbors",
        [](const std::vector<size_t> &from_idxs, const std::vector<size_t> &to_idxs, const py::object &positions,
           const py::object &map_size) {
          const auto pos = positions_from(positions);
          const MapShape shape = shape_from(map_size);
          py::gil_scoped_release release;
          return get_neighbors(from_idxs, to_idxs, pos, shape);
        },
        py::arg("from_idxs"), py::arg("to_idxs"), py::arg("positions"), py::arg("map_size"));
  m.def("neighbor_recombinations",
        [](const py::list &genomes, const std::vector<size_t> &cell_idxs, const py::object &positions,
           const py::object &map_size, double p) {
          // World.recombinate_cells in one call (world.py:716-745): neighbour pairs among cell_idxs,
          // recombinations between their genomes; only the pairs that were hit come back, as
          // (genome a, genome b, a, b).  No list of pairs crosses the language boundary.
          const auto pos = positions_from(positions);
          if ((size_t)py::len(genomes) != pos.size()) throw py::value_error("one genome per position expected");
          const auto pairs = get_neighbors(cell_idxs, cell_idxs, pos, shape_from(map_size));
          const auto hits = recombinations_indexed(genomes, pairs, p);
          std::vector<std::tuple<std::string, std::string, size_t, size_t>> out;
          out.reserve(hits.size());
          for (const auto &[x, y, s] : hits) out.emplace_back(x, y, pairs[s].first, pairs[s].second);
          return out;
        },
        py::arg("genomes"), py::arg("cell_idxs"), py::arg("positions"), py::arg("map_size"), py::arg("p"),
        "no counterpart in the reference");
  m.def("divide_cells_if_possible",
        [](const std::vector<size_t> &cell_idxs, const py::object &positions, size_t n_cells, const py::object &map_size) {
          const auto pos = positions_from(positions);
          const MapShape shape = shape_from(map_size);
          py::gil_scoped_release release;
          return divide_cells_if_possible(cell_idxs, pos, n_cells, shape);
        },
        py::arg("cell_idxs"), py::arg("positions"), py::arg("n_cells"), py::arg("map_size"));
  m.def("move_cells",
        [](const std::vector<size_t> &cell_idxs, const py::object &positions, const py::object &map_size) {
          const auto pos = positions_from(positions);
          const MapShape shape = shape_from(map_size);
          py::gil_scoped_release release;
          return move_cells(cell_idxs, pos, shape);
        },
        py::arg("cell_idxs"), py::arg("positions"), py::arg("map_size"));
}
