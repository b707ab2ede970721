"""Known-answer and cross-implementation tests that pin the CPU oracle as far as it can be pinned
(the reference ships no tests; anchors = hand-computable definitions + the seqan3
minimiser_hash unit-test vector + closed forms of the IBF constants).  CPU only."""
import math
import random
import numpy as np
import oracle
from tests import pyref


def test_kmer_hash_hand_computed():
    # "ACGT", k=4, sigma=5: 0*125 + 1*25 + 2*5 + 4 = 39; reverse complement of ACGT is ACGT.
    v = oracle.kmer_values("ACGT", k=4, seed=0)
    assert list(v) == [39]
    # N is rank 3 and is hashed, not skipped: "ANA" k=3 -> 0*25+3*5+0 = 15 ; rc = T N T = 4*25+3*5+4 = 119
    assert list(oracle.kmer_values("ANA", k=3, seed=0)) == [15]
    # lower case and U
    assert list(oracle.kmer_values("acgu", k=4, seed=0)) == [39]


def test_seqan3_minimiser_hash_unit_vector():
    # seqan3 test/unit/search/views/minimiser_hash_test.cpp (recalled): dna4, ungapped 4, window 8, seed 0
    got = oracle.minimisers("ACGGCGACGTTTAG", k=4, w=8, seed=0, sigma=4)
    assert list(got) == [26, 97, 27, 6, 1]
    assert list(oracle.minimisers("A" * 20, k=4, w=8, seed=0, sigma=4)) == [0, 0, 0]
    assert pyref.minimisers("ACGGCGACGTTTAG", 4, 8, 0, 4) == [26, 97, 27, 6, 1]
    assert pyref.minimisers("A" * 20, 4, 8, 0, 4) == [0, 0, 0]


def test_default_seed_applied():
    v0 = oracle.kmer_values("ACGTACGTACGTACGTACGTAGG", k=19, seed=0)
    v1 = oracle.kmer_values("ACGTACGTACGTACGTACGTAGG", k=19)
    f = pyref.kmer_values("ACGTACGTACGTACGTACGTAGG", 19, 0)
    assert list(v0) == f
    assert list(v1) == pyref.kmer_values("ACGTACGTACGTACGTACGTAGG", 19)
    assert any(a != b for a, b in zip(v0, v1))


def _rand_seq(rng, n, alphabet="ACGT"):
    return "".join(rng.choice(alphabet) for _ in range(n))


def test_minimisers_vs_python_restatement_random_and_tie_heavy():
    rng = random.Random(1234)
    cases = []
    for n in (0, 1, 18, 19, 20, 40, 41, 42, 60, 100, 333, 1000):
        cases.append(_rand_seq(rng, n))
    cases += ["A" * 200, "AC" * 150, "ACG" * 90, "N" * 120, "A" * 50 + "N" * 30 + "ACGT" * 30,
              _rand_seq(rng, 300, "AC"), _rand_seq(rng, 300, "ACGTN"), ("ACGTTGCA" * 40)[:311]]
    for _ in range(40):                       # repeats with 2 % noise
        unit = _rand_seq(rng, rng.randint(1, 12))
        s = list((unit * 80)[:rng.randint(100, 500)])
        for i in range(len(s)):
            if rng.random() < 0.02:
                s[i] = rng.choice("ACGTN")
        cases.append("".join(s))
    for s in cases:
        for (k, w) in ((19, 41), (19, 19), (5, 12), (26, 60)):
            assert list(oracle.minimisers(s, k, w)) == pyref.minimisers(s, k, w), (len(s), k, w)


def test_ibf_constants_closed_forms():
    # seeds equal floor(2^64 / c) for documented constants; multiplier = floor(2^64 / phi)
    two64 = 2.0 ** 64
    approx = [two64 / (math.e / 2), two64 / math.sqrt(2), two64 / math.sqrt(3), two64 / (math.sqrt(5) / 2), 2.0 ** 63 / (3 * math.pi / 5)]
    for i, a in enumerate(approx):
        assert abs(oracle.lib().orc_ibf_seed(i) - a) / a < 1e-12
        assert oracle.lib().orc_ibf_seed(i) == pyref.IBF_SEEDS[i]
    assert abs(11400714819323198485 - two64 / ((1 + math.sqrt(5)) / 2)) / two64 < 1e-12


def test_bin_size_formula():
    # ceil(3n / 0.242630...) (reference src/utils.cpp:75-90)
    n = 448_500
    d = math.log(1 - math.exp(math.log(0.01) / 3))
    assert oracle.lib().orc_bin_size_in_bits(n) == math.ceil(-(n * 3) / d)
    assert oracle.lib().orc_bin_size_in_bits(10 ** 12) == 2 ** 32 - 3     # capped (include/index_arguments.hpp:20)


def test_ibf_build_query_and_rows():
    rng = random.Random(5)
    refs = [np.frombuffer(_rand_seq(rng, 3000).encode(), np.uint8) for _ in range(3)]
    idx = oracle.Index.build(refs, rec_bin=[0, 1, 2], bin_cat=[0, 1, 1], categories=["host", "microbial"], k=19, w=41, threads=2)
    assert (idx.bins, idx.bin_words, idx.technical_bins, idx.hash_funs, idx.ncat) == (3, 1, 64, 3, 2)
    assert idx.hash_shift == 64 - idx.bin_size.bit_length()
    rows = idx.rows()
    for b, r in enumerate(refs):
        mins = oracle.minimisers(r)
        masks = idx.query(mins)
        assert np.all((masks[:, 0] >> np.uint64(b)) & np.uint64(1) == 1)      # no false negatives
        for h in map(int, mins[:20]):
            for i in range(3):
                row = pyref.ibf_row(h, i, idx.bin_size)
                assert (int(rows[row]) >> b) & 1
    assert int(np.bitwise_or.reduce(rows)) < (1 << 3)                           # only real bins ever set


def test_index_file_roundtrip(tmp_path):
    rng = random.Random(6)
    refs = [np.frombuffer(_rand_seq(rng, 5000).encode(), np.uint8) for _ in range(2)]
    idx = oracle.Index.build(refs, [0, 1], [0, 1], ["microbial", "human"], threads=1)
    p = tmp_path / "t.idx"
    idx.save(p)
    back = oracle.Index.load(p)
    assert (back.k, back.w, back.bins, back.bin_size, back.hash_shift, back.categories) == (19, 41, 2, idx.bin_size, idx.hash_shift, ["microbial", "human"])
    assert back.host_index == 1
    assert np.array_equal(back.rows(), idx.rows())
    raw = p.read_bytes()
    assert raw[0] == 41 and raw[1] == 19 and np.frombuffer(raw[2:10], np.float64)[0] == 0.01


def test_model_probability_properties():
    # Model::prob (classify_stats.hpp:370-389): x == 1 forces p_pos = 1; x = 0 is dominated by p_err + p_neg
    assert oracle.model_prob(0.0) < 1e-3
    p1 = oracle.model_prob(1.0)
    assert 0.99 < p1 <= 1.0
    assert 0.9 < oracle.model_prob(0.5) <= 1.0
    pos, neg = oracle.kde_defaults()
    assert np.all(np.diff(pos) >= 0) and np.all(np.diff(neg) >= 0)             # KDEParams ctor sorts (:217)
    # independent float32 restatement of the KDE sum at one point
    x = np.float32(0.1)
    def kde(ds, h):
        tot = np.float32(0)
        for xi in ds:
            z = np.float32((x - xi) / np.float32(h))
            tot = np.float32(tot + np.float32(math.exp(-float(z) ** 2 / 2) / math.sqrt(2 * 3.141592653589793238463)))
        return np.float32(tot / np.float32(np.float32(h) * np.float32(len(ds))))
    p_err = np.float32(math.exp(np.float32(np.log(np.float32(300))) - np.float32(300) * x))
    pp, pn = kde(pos, 0.1), kde(neg, 0.001)
    want = float(np.float32(pp / np.float32(np.float32(p_err + pp) + pn)))
    assert abs(oracle.model_prob(0.1) - want) <= 2e-6 * want


def test_compression_ratio_matches_python_zlib():
    import zlib
    rng = random.Random(9)
    s = _rand_seq(rng, 5000).encode()
    c = zlib.compressobj(-1, zlib.DEFLATED, 15 | 16, 8, zlib.Z_DEFAULT_STRATEGY)
    n = len(c.compress(s) + c.flush())
    assert abs(oracle.compression_ratio(s) - np.float32(n / 5000)) < 1e-7
