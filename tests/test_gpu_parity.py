"""GPU parity tests proper: the CUDA path, called through the C-ABI (ctypes), against the CPU oracle on the same
seeded inputs.  Bit-exact for num_hashes / counts / unique counts / confidence / call; probabilities <= 1e-6 relative
(north_star).  Run on the B200 box: pytest -m gpu."""
import random
import numpy as np
import pytest
import oracle
from tests import common

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def small():
    idx, refs = common.build_oracle_index([300_000, 300_000], [0, 1], ["host", "microbial"], seed=11)
    gpu = common.gpu_from_oracle(idx)
    import charon_b200
    gpu.set_model(charon_b200.DehostOptions(), host_index=idx.host_index)
    yield idx, refs, gpu
    gpu.close()


def test_filter_layout_roundtrip(small):
    idx, refs, gpu = small
    assert gpu.info().row_bits == 2 and gpu.info().single_pass == 1
    assert np.array_equal(gpu.debug_dump_filter(), idx.rows())


def test_membership_query_matches_oracle(small):
    idx, refs, gpu = small
    rng = np.random.default_rng(3)
    vals = np.concatenate([oracle.minimisers(refs[0][:50_000]), oracle.minimisers(refs[1][:50_000]),
                           rng.integers(0, 2**63, 5000, dtype=np.uint64)])
    assert np.array_equal(gpu.debug_query(vals), idx.query(vals))


def _tie_heavy_cases():
    rng = random.Random(77)
    rs = lambda n, a="ACGT": "".join(rng.choice(a) for _ in range(n))
    cases = ["", "A", "ACGT" * 4 + "AC", "ACGT" * 4 + "ACG", rs(19), rs(20), rs(40), rs(41), rs(42), rs(64), rs(1000), rs(2984), rs(2985),
             rs(2986), rs(3100), rs(9000), "A" * 5000, "AC" * 3000, "ACG" * 1500, "N" * 4000, rs(3000) + "N" * 3000 + rs(3000),
             rs(3000, "AC"), rs(7000, "ACGTN"), ("ACGTTGCA" * 900)[:6111], rs(2940) + "A" * 200 + rs(3000), "acgtn" * 700 + "RYKMSWBDHVU" * 30]
    for _ in range(12):
        unit = rs(rng.randint(1, 12))
        s = list((unit * 4000)[:rng.randint(2500, 9500)])
        for i in range(len(s)):
            if rng.random() < 0.02:
                s[i] = rng.choice("ACGTN")
        cases.append("".join(s))
    return cases


def test_minimiser_stream_matches_oracle_including_ties(small):
    idx, refs, gpu = small
    for s in _tie_heavy_cases():
        want = np.sort(oracle.minimisers(s))
        got = gpu.debug_minimisers(s)
        assert np.array_equal(got, want), (len(s), s[:40], len(got), len(want))


def test_long_read_tiling(small):
    idx, refs, gpu = small
    s = refs[0][1000:151_000]
    assert np.array_equal(gpu.debug_minimisers(s), np.sort(oracle.minimisers(s)))


def test_mixed_lengths_dispatch_between_warp_and_cta_kernels(small):
    """Reads shorter than 65 536 bases run on the warp-per-read kernel, longer ones on the CTA-per-read kernel of the
    same batch; both must agree with the oracle (incl. a read exactly at the threshold and a low-complexity long read)."""
    idx, refs, gpu = small
    rng = np.random.default_rng(21)
    lens = [100, 8000, 65_535, 65_536, 65_537, 120_000, 30, 5000, 250_000, 713 + 40, 714 + 40, 1426 + 40]
    pieces = []
    for ln in lens:
        st = int(rng.integers(0, 300_000 - ln)) if ln < 300_000 else 0
        pieces.append(refs[int(rng.integers(0, 2))][st:st + ln].copy())
    pieces.append(np.frombuffer((b"ACGTTGCA" * 20000)[:100_001], np.uint8))
    bases, seg_off = common.ragged_batch(rng, pieces)
    n = len(pieces)
    mq, cr = common.scalars(n)
    want = idx.classify_batch(bases, seg_off, mean_q=mq, compression=cr)
    got = gpu.classify_batch(bases, seg_off, mq, cr, np.diff(seg_off).astype(np.uint32))
    common.assert_parity(got, want)


def test_counts_and_calls_dehost_config1_like(small):
    idx, refs, gpu = small
    n, L = 2000, 5000
    bases, seg_off, src = common.fixed_length_batch(5, refs, [0.45, 0.45], n, L)
    rng = np.random.default_rng(1)
    mq, cr = common.scalars(n, rng)
    length = np.full(n, L, np.uint32)
    want = idx.classify_batch(bases, seg_off, mean_q=mq, compression=cr)
    got = gpu.classify_batch(bases, seg_off, mq, cr, length)
    common.assert_parity(got, want)
    assert (want["call"] != 255).sum() > n // 2            # the workload really classifies
    only_counts = gpu.count_batch(bases, seg_off)
    common.assert_parity(only_counts, want)
    ev = gpu.eval_batch(want["num_hashes"], want["counts"], want["unique_counts"], mq, cr, length)
    assert np.array_equal(ev["call"], want["call"]) and np.array_equal(ev["confidence"], want["confidence"])


def test_ragged_empty_short_and_paired(small):
    idx, refs, gpu = small
    rng = np.random.default_rng(9)
    pieces = []
    for ln in [0, 1, 18, 19, 20, 41, 42, 139, 140, 141, 500, 2985, 2986, 6000, 12_345, 0, 33]:
        st = int(rng.integers(0, 200_000))
        pieces.append(refs[int(rng.integers(0, 2))][st:st + ln].copy())
    bases, seg_off = common.ragged_batch(rng, pieces)
    n = len(pieces)
    mq, cr = common.scalars(n)
    length = np.diff(seg_off).astype(np.uint32)
    want = idx.classify_batch(bases, seg_off, mean_q=mq, compression=cr)
    got = gpu.classify_batch(bases, seg_off, mq, cr, length)
    common.assert_parity(got, want)
    assert want["num_hashes"][0] == 0 and np.isnan(got["prob"][0]).all()
    # paired: two segments per read, minimisers of mate 1 then mate 2 into one entry (src/dehost_main.cpp:458-465)
    pieces2 = pieces[:16]
    bases2, seg_off2 = common.ragged_batch(rng, pieces2)
    read_seg = np.arange(0, 17, 2, dtype=np.uint32)
    mq2, cr2 = common.scalars(8)
    len2 = np.add.reduceat(np.diff(seg_off2), read_seg[:-1]).astype(np.uint32)
    want2 = idx.classify_batch(bases2, seg_off2, read_seg=read_seg, mean_q=mq2, compression=cr2)
    got2 = gpu.classify_batch(bases2, seg_off2, mq2, cr2, len2, read_seg=read_seg)
    common.assert_parity(got2, want2)


def test_empty_batch(small):
    idx, refs, gpu = small
    res = gpu.count_batch(np.zeros(0, np.uint8), np.zeros(1, np.uint64))
    assert res["num_hashes"].shape == (0,)


@pytest.mark.parametrize("nbins,bin_cat,cats", [
    (3, [0, 1, 1], ["host", "microbial"]),                       # multi-bin category -> two-pass best-bin path, 4-bit rows
    (7, [0, 1, 2, 0, 1, 2, 2], ["a", "host", "c"]),              # 8-bit rows
    (20, [i % 5 for i in range(20)], [f"c{i}" for i in range(5)]),  # 32-bit rows
    (40, [i % 40 for i in range(40)], [f"c{i}" for i in range(40)]),  # 64-bit rows, one bin per category
    (70, [i % 9 for i in range(70)], [f"c{i}" for i in range(9)]),   # 128-bit rows, two mask words
    (255, list(range(255)), [f"c{i}" for i in range(255)]),          # 256-bit rows (config 4 shape)
    (200, [i % 100 for i in range(200)], [f"c{i}" for i in range(100)]),
])
def test_many_bins_classify(nbins, bin_cat, cats):
    import charon_b200
    idx, refs = common.build_oracle_index([20_000] * nbins, bin_cat, cats, seed=100 + nbins)
    gpu = common.gpu_from_oracle(idx)
    try:
        assert np.array_equal(gpu.debug_dump_filter(), idx.rows())
        gpu.set_model(charon_b200.ClassifyOptions(), host_index=0)
        n, L = 300, 3000
        fr = [0.9 / nbins] * nbins
        bases, seg_off, src = common.fixed_length_batch(6, refs, fr, n, L)
        mq, cr = common.scalars(n)
        opts = oracle.default_options(dehost=False)
        want = idx.classify_batch(bases, seg_off, mean_q=mq, compression=cr, opts=opts)
        got = gpu.classify_batch(bases, seg_off, mq, cr, np.full(n, L, np.uint32))
        # classify uses beta densities through lgamma/log/exp in float: best-effort 1e-4; integers exact
        common.assert_parity(got, want, prob_rtol=2e-4)
    finally:
        gpu.close()


@pytest.mark.parametrize("k,w", [(19, 19), (5, 12), (13, 40), (14, 14), (26, 60), (11, 255)])
def test_other_k_w(k, w):
    idx, refs = common.build_oracle_index([50_000, 50_000], [0, 1], ["human", "other"], seed=55, k=k, w=w)
    gpu = common.gpu_from_oracle(idx)
    try:
        for s in [refs[0][:7000], refs[1][100:400], np.frombuffer(b"A" * 3000, np.uint8), np.frombuffer(b"ACGTN" * 900, np.uint8)]:
            assert np.array_equal(gpu.debug_minimisers(s), np.sort(oracle.minimisers(s, k, w))), (k, w, len(s))
        bases, seg_off, _ = common.fixed_length_batch(2, refs, [0.5, 0.4], 200, 1500)
        common.assert_parity(gpu.count_batch(bases, seg_off), idx.classify_batch(bases, seg_off))
    finally:
        gpu.close()


def test_sd_vector_payload_decoded_on_device(tmp_path, small):
    """The .idx payload (sdsl sd_vector low/high words) is decoded on the GPU to the same filter bits."""
    idx, refs, gpu = small
    rows = idx.rows()
    pos = np.flatnonzero(np.unpackbits(rows.view(np.uint8), bitorder="little")).astype(np.uint64)
    n, m = idx.technical_bins * idx.bin_size, len(pos)
    logm, logn = int(m).bit_length(), int(n).bit_length()
    if logm == logn:
        logm -= 1
    wl = logn - logm
    lowvals = pos & np.uint64((1 << wl) - 1)
    low_bits = m * wl
    lowbits = ((lowvals[:, None] >> np.arange(wl, dtype=np.uint64)) & np.uint64(1)).astype(np.uint8).reshape(-1)
    low = np.packbits(np.pad(lowbits, (0, (-len(lowbits)) % 64)), bitorder="little").view(np.uint64)
    high_bits = m + (1 << logm)
    hb = np.zeros(high_bits + (-high_bits) % 64, np.uint8)
    hb[(pos >> np.uint64(wl)) + np.arange(m, dtype=np.uint64)] = 1
    high = np.packbits(hb, bitorder="little").view(np.uint64)
    import charon_b200
    g2 = charon_b200.GpuClassifier(k=idx.k, w=idx.w, bins=idx.bins, bin_size=idx.bin_size, hash_funs=3, bin_to_cat=idx.bin_to_cat,
                                   ncat=idx.ncat, sd=(low, low_bits, high, high_bits, wl))
    try:
        assert np.array_equal(g2.debug_dump_filter(), rows)
    finally:
        g2.close()
    bad = high.copy()
    z = int(np.flatnonzero(hb[:high_bits] == 0)[0])                 # an extra set bit must be rejected, not ignored
    bad[z // 64] |= np.uint64(1) << np.uint64(z % 64)
    with pytest.raises(charon_b200.CharonGpuError):
        charon_b200.GpuClassifier(k=idx.k, w=idx.w, bins=idx.bins, bin_size=idx.bin_size, hash_funs=3, bin_to_cat=idx.bin_to_cat,
                                  ncat=idx.ncat, sd=(low, low_bits, bad, high_bits, wl))


@pytest.mark.parametrize("nbins,lens", [(2, [400_000, 150_000]), (5, [30_000, 70_000, 10, 200_000, 18]), (70, [5_000] * 70)])
def test_index_built_on_gpu_equals_oracle_index(nbins, lens):
    """cg_build_index (device-side `charon index` data path) produces the same bin_size, distinct-hash counts and
    filter bits as the oracle's restatement of src/index_main.cpp:118-263."""
    import charon_b200
    cats = [f"c{i}" for i in range(min(nbins, 3))]
    bin_cat = [i % len(cats) for i in range(nbins)]
    idx, refs = common.build_oracle_index(lens, bin_cat, cats, seed=900 + nbins)
    gpu = charon_b200.GpuClassifier.from_references(refs, list(range(nbins)), bin_cat, len(cats))
    try:
        assert gpu.bin_size == idx.bin_size
        for b, r in enumerate(refs):
            assert int(gpu.hashes_per_bin[b]) == len(np.unique(oracle.minimisers(r))), b
        assert np.array_equal(gpu.debug_dump_filter(), idx.rows())
    finally:
        gpu.close()


@pytest.mark.parametrize("which", [1, 2])
def test_other_kernels_still_match(small, which):
    """The any-(k,w) kernel (1) and the CTA-per-read specialised kernel (2: used for long reads) stay parity-green on the
    default (19, 41) when forced for every read."""
    idx, refs, gpu = small
    gpu.force_generic(which)
    try:
        for s in _tie_heavy_cases():
            assert np.array_equal(gpu.debug_minimisers(s), np.sort(oracle.minimisers(s))), len(s)
        bases, seg_off, _ = common.fixed_length_batch(15, refs, [0.45, 0.45], 500, 5000)
        common.assert_parity(gpu.count_batch(bases, seg_off), idx.classify_batch(bases, seg_off))
    finally:
        gpu.force_generic(False)


def test_equal_values_exactly_one_window_apart(small):
    """Two equal minima exactly W k-mers apart never share a window, yet the tracked minimiser must be re-emitted
    (seqan3 recomputes when the old one falls out).  Periodic sequences of period 23 create exactly that."""
    idx, refs, gpu = small
    rng = random.Random(5)
    for period in (23, 46, 22, 24, 41):
        unit = "".join(rng.choice("ACGT") for _ in range(period))
        for n in (200, 3000, 7000):
            s = (unit * (n // period + 2))[:n]
            assert np.array_equal(gpu.debug_minimisers(s), np.sort(oracle.minimisers(s))), (period, n)


def test_duplicate_values_inside_one_window(small):
    """The fast path keeps ONE tie test (the k-mer that has just left a window carries that window's minimum, count_kernel_v2.cuh
    p2_block) and compares run starts on full values.  Stress it with equal canonical values at every distance below, at and
    above the window length: a 19-mer repeated d bases later, and a 19-mer followed by its reverse complement (same canonical
    value on the other strand), planted at random offsets so that they also meet block and tile boundaries."""
    idx, refs, gpu = small
    rng = random.Random(11)
    comp = {"A": "T", "C": "G", "G": "C", "T": "A"}
    rnd = lambda n: "".join(rng.choice("ACGT") for _ in range(n))
    for mode in (0, 2):   # warp-per-read and CTA-per-read kernels
        gpu.force_generic(mode)
        try:
            for d in list(range(1, 26)) + [40, 41, 45, 46, 64]:
                parts = [rnd(rng.randrange(0, 60))]
                for rep in range(40):
                    kmer = rnd(19)
                    twin = kmer if rep % 2 == 0 else "".join(comp[c] for c in reversed(kmer))
                    if d >= 19:
                        parts += [kmer, rnd(d - 19), twin]
                    else:   # overlapping copies: a periodic stretch of period d holds the 19-mer again d bases later
                        parts += [(kmer[:d] * (40 // d + 2))[:19 + d]]
                    parts.append(rnd(rng.randrange(0, 90)))
                s = "".join(parts)
                assert np.array_equal(gpu.debug_minimisers(s), np.sort(oracle.minimisers(s))), (mode, d, len(s))
        finally:
            gpu.force_generic(0)


def test_tie_run_crossing_a_tile_boundary_regression(small):
    """Golden regression (tests/golden/tie_tile_boundary_809.txt): a run of equal window minima crosses the 713-window
    warp tile while an EARLIER run's tracked element also sits past the tile end — only the run alive at the last window
    may hand its position to the next tile.  Also shifted so the same structure meets the 2921-window CTA tile."""
    import os
    idx, refs, gpu = small
    t = open(os.path.join(os.path.dirname(__file__), "golden", "tie_tile_boundary_809.txt")).read().strip()
    rng = random.Random(3)
    for mode in (0, 2, 1):
        gpu.force_generic(mode)
        try:
            for pad in (0, 2921 - 713, 1, 5, 712, 2208 + 713):
                s = "".join(rng.choice("ACGT") for _ in range(pad)) + t
                assert np.array_equal(gpu.debug_minimisers(s), np.sort(oracle.minimisers(s))), (mode, pad)
        finally:
            gpu.force_generic(0)


@pytest.mark.gpu
def test_two_slot_pipeline_matches_the_oracle():
    """cg_submit_batch / cg_wait_batch: two different batches in flight at once, then reuse of both slots."""
    import charon_b200
    idx, refs = common.build_oracle_index([150_000, 110_000], [0, 1], ["host", "microbial"], seed=61)
    gpu = common.gpu_from_oracle(idx)
    gpu.set_model(charon_b200.DehostOptions(), host_index=idx.host_index)
    batches = []
    for i, (n, L) in enumerate([(300, 3000), (150, 5000), (64, 700), (500, 1500)]):
        bases, seg_off, _ = common.fixed_length_batch(70 + i, refs, [0.45, 0.45], n, L)
        mq, cr = common.scalars(n, np.random.default_rng(i))
        batches.append((bases, seg_off, mq, cr, np.full(n, L, np.uint32)))
    want = [idx.classify_batch(b, s, mean_q=mq, compression=cr) for b, s, mq, cr, _ in batches]
    r0 = gpu.submit(0, *batches[0])
    r1 = gpu.submit(1, *batches[1])
    with pytest.raises(charon_b200.CharonGpuError, match="in flight"):
        gpu.submit(1, *batches[2])
    gpu.wait(0)
    r2 = gpu.submit(0, *batches[2])
    gpu.wait(1)
    r3 = gpu.submit(1, batches[3][0], batches[3][1])        # counts only on the second use of slot 1
    gpu.wait(0); gpu.wait(1)
    for got, w in zip((r0, r1, r2), want[:3]):
        common.assert_parity(got, w)
    common.assert_parity(r3, want[3])
    # the synchronous call still works afterwards
    common.assert_parity(gpu.classify_batch(*batches[0]), want[0])
    gpu.close()


def test_elias_fano_encode_decode_round_trip(small):
    """cg_export_sd_vector (device encoder, `charon index` output) -> CG_FILTER_SD_VECTOR (device decoder, `.idx` input):
    the filter must come back bit for bit, and the parts must satisfy sdsl::sd_vector's invariants."""
    import charon_b200
    idx, refs, gpu = small
    size, wl, low, low_bits, high, high_bits = gpu.export_sd_vector()
    ones = int(np.unpackbits(idx.rows().view(np.uint8)).sum())
    assert size == 64 * idx.bin_size and low_bits == ones * wl
    assert int(np.unpackbits(high.view(np.uint8)).sum()) == ones
    logm = max(ones, 1).bit_length(); logn = size.bit_length()
    if logm == logn:
        logm -= 1
    assert wl == logn - logm and high_bits == ones + (1 << logm)
    again = charon_b200.GpuClassifier(k=idx.k, w=idx.w, bins=idx.bins, bin_size=idx.bin_size, hash_funs=idx.hash_funs, bin_to_cat=idx.bin_to_cat,
                                      ncat=idx.ncat, sd=(low, low_bits, high, high_bits, wl))
    assert np.array_equal(again.debug_dump_filter(), idx.rows())
    again.close()


def test_gather_microbenchmark_runs(small):
    idx, refs, gpu = small
    for mode in (0, 1):
        assert gpu.bench_gather(1_000_000, mode=mode, reps=2) > 0.0


def test_full_size_batch_properties():
    """BASELINE config 1 shape at a size the oracle cannot replay in seconds (2 x 200 Mbp index built on the GPU, 131 072 reads
    of 8 kb): size-independent properties of the path + an oracle spot check on the oracle's own full-size index."""
    import charon_b200
    from tools import synth
    refs = [synth.reference(1000 + b, 200_000_000) for b in range(2)]
    gpu = charon_b200.GpuClassifier.from_references(refs, [0, 1], [0, 1], 2)
    gpu.set_model(charon_b200.DehostOptions(), host_index=0)
    n, L = 131_072, 8000
    bases, src = synth.reads(42, refs, [0.45, 0.45], n, L, 0.02)
    seg_off = np.arange(n + 1, dtype=np.uint64) * np.uint64(L)
    mq, cr, ln = np.full(n, 20.0, np.float32), np.full(n, 0.3, np.float32), np.full(n, L, np.uint32)
    whole = gpu.classify_batch(bases, seg_off, mq, cr, ln)
    # (1) plausibility of the whole batch: emission density of iid sequence, reads classified to their source
    dens = whole["num_hashes"].astype(np.float64).mean() / L
    assert 0.085 < dens < 0.095
    assert ((whole["call"] == src) | (src == 2))[whole["call"] != 255].mean() > 0.999
    assert (whole["call"] != 255).mean() > 0.85
    # (2) a batch is the concatenation of its parts: any split gives the same rows
    cut = 50_001
    a = gpu.classify_batch(bases[:cut * L], seg_off[:cut + 1], mq[:cut], cr[:cut], ln[:cut])
    b = gpu.classify_batch(bases[cut * L:], seg_off[cut:] - seg_off[cut], mq[cut:], cr[cut:], ln[cut:])
    for key in ("num_hashes", "counts", "unique_counts", "call", "confidence", "prob"):
        assert np.array_equal(np.concatenate([a[key], b[key]]), whole[key]), key
    # (3) order of reads does not matter
    perm = np.random.default_rng(3).permutation(n)
    shuffled = bases.reshape(n, L)[perm].reshape(-1)
    p = gpu.classify_batch(shuffled, seg_off, mq, cr, ln)
    for key in ("num_hashes", "counts", "unique_counts", "call", "confidence"):
        assert np.array_equal(p[key], whole[key][perm]), key
    # (4) paired reads: mates are hashed separately and pooled -> totals add (one bin per category: counts are plain sums)
    m = 20_000
    read_seg = np.arange(0, 2 * m + 1, 2, dtype=np.uint32)
    pair = gpu.count_batch(bases[:2 * m * L], seg_off[:2 * m + 1], read_seg=read_seg)
    for key in ("num_hashes", "counts", "unique_counts"):
        w = whole[key][:2 * m]
        assert np.array_equal(pair[key], w[0::2] + w[1::2]), key
    # (5) oracle spot check at full index size (the oracle builds its own 2 x 200 Mbp index)
    idx = oracle.Index.build(refs, rec_bin=[0, 1], bin_cat=[0, 1], categories=["host", "microbial"])
    assert idx.bin_size == gpu.bin_size
    k = 2000
    want = idx.classify_batch(bases[:k * L], seg_off[:k + 1], mean_q=mq[:k], compression=cr[:k])
    common.assert_parity({key: whole[key][:k] for key in ("num_hashes", "counts", "unique_counts", "call", "confidence", "prob")}, want)
    gpu.close()


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["1", "0"])
def test_both_gather_carriers(monkeypatch, mode):
    """The deferred filter gathers travel either in registers (plain loads) or as cp.async into shared memory; the library picks
    by filter size (only filters beyond L2 take the second), so the small test filters force each in turn (CHARON_ASYNC_GATHER)."""
    import charon_b200
    monkeypatch.setenv("CHARON_ASYNC_GATHER", mode)
    idx, refs = common.build_oracle_index([300_000, 300_000], [0, 1], ["host", "microbial"], seed=11)
    gpu = common.gpu_from_oracle(idx)
    try:
        gpu.set_model(charon_b200.DehostOptions(), host_index=idx.host_index)
        rng = np.random.default_rng(5)
        # long enough reads for several tiles with more than DQA full groups in flight, plus ragged / short / paired ones
        bases, seg_off, _ = common.fixed_length_batch(31, refs, [0.45, 0.45], 1500, 9001)
        mq, cr = common.scalars(1500, rng)
        want = idx.classify_batch(bases, seg_off, mean_q=mq, compression=cr)
        for pack in (1, 0):
            gpu.set_host_pack(pack)
            common.assert_parity(gpu.classify_batch(bases, seg_off, mq, cr, np.full(1500, 9001, np.uint32)), want)
        pieces = [refs[int(rng.integers(0, 2))][s:s + ln].copy() for ln in [0, 5, 41, 800, 2985, 2986, 30_000, 70_000, 1426 + 40, 19, 64, 12_000]
                  for s in [int(rng.integers(0, 200_000))]]
        pieces.append(np.frombuffer(b"ACGTTGCA" * 3000, np.uint8))       # tie tiles: the exact path runs between deferred groups
        pieces.append(np.frombuffer(b"A" * 9000, np.uint8))
        b2, so2 = common.ragged_batch(rng, pieces)
        read_seg = np.arange(0, len(pieces) + 1, 2, dtype=np.uint32)
        want2 = idx.classify_batch(b2, so2, read_seg=read_seg)
        got2 = gpu.count_batch(b2, so2, read_seg=read_seg)
        common.assert_parity(got2, want2)
        common.assert_parity(gpu.count_batch(b2, so2), idx.classify_batch(b2, so2))
    finally:
        gpu.close()
