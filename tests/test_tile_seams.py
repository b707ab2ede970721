"""Tile seams of the minimiser kernels.

Equal canonical values a < b less than a window apart, both the strict minimum of their windows, swept over every position
around the end of the first (and second) kernel tile, for every distance d in 1..22, as a repeated 19-mer and as a
reverse-complement twin (tests/seam_cases.py).  seqan3 emits once more when `a` leaves the window; the GPU
kernels must do so whether the departure happens inside the tile that saw both, in its last window, or in the next tile.
"""
import os
import random
import numpy as np
import pytest
import oracle
from tests import common, pyref, seam_cases

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
WARP_TILE = 31 * 23      # windows per tile, count_kernel_v3 (V3Cfg::TW)
CTA_TILE = 127 * 23      # count_kernel_v2 (V2Cfg::TW) and the generic kernel ((NT - 1) * W)


def test_seam_generator_plants_what_it_promises():
    """CPU: the planted pair is equal, is the strict minimum of every window it touches, and costs the value-only emission
    rule exactly one emission against the oracle (and against the independent pure-Python statement)."""
    cases = seam_cases.sweep(WARP_TILE, 1500, seed=3, positions=[700, 712, 713, 720], ds=[3, 7, 8, 18, 19, 22], cores_per_kind=1)
    assert len(cases) >= 40
    for d, twin, pa, s in cases:
        v = oracle.kmer_values(s)
        assert v[pa] == v[pa + d]
        lo, hi = pa - 22, pa + d + 23
        assert (np.delete(v[lo:hi], [pa - lo, pa + d - lo]) > v[pa]).all()
        wm = np.lib.stride_tricks.sliding_window_view(v, 23).min(axis=1)
        assert len(oracle.minimisers(s)) == 1 + int((wm[1:] != wm[:-1]).sum()) + 1
    for d, twin, pa, s in random.Random(1).sample(cases, 4):
        assert list(oracle.minimisers(s)) == pyref.minimisers(s, 19, 41)


def test_golden_period4_seam_sequence_on_the_oracle():
    """The 941-base input of VERDICT r1 (period-4 repeat at base 718: equal values at k-mers 718 and 722 inside the last
    window of the first 713-window tile): 86 emissions."""
    s = open(os.path.join(GOLDEN, "seam_period4_941.txt")).read().strip()
    v = oracle.kmer_values(s)
    assert len(s) == 941 and v[718] == v[722]
    assert len(oracle.minimisers(s)) == 86 and pyref.minimisers(s, 19, 41) == list(oracle.minimisers(s))


@pytest.fixture(scope="module")
def seam_batches():
    """(name, tile, sequences) — generated once per module (about 25 s of pure Python)."""
    warp = seam_cases.sweep(WARP_TILE, 1900, seed=5)
    warp2 = seam_cases.sweep(2 * (WARP_TILE - 1) + 1, 2400, seed=6, ds=[1, 2, 3, 4, 7, 11, 18, 19, 22])   # the second seam: tiles overlap by a window
    cta = seam_cases.sweep(CTA_TILE, 4200, seed=7)
    return [("warp-tile seam 1", warp), ("warp-tile seam 2", warp2), ("cta-tile seam 1", cta)]


def _batch(cases):
    pieces = [np.frombuffer(s.encode(), np.uint8) for _, _, _, s in cases]
    return common.ragged_batch(None, pieces)


@pytest.mark.gpu
@pytest.mark.parametrize("carrier", ["0", "1"])
def test_equal_minima_across_tile_seams(monkeypatch, seam_batches, carrier):
    """num_hashes / counts / unique counts of every seam sequence, all kernels (0 warp-per-read, 2 CTA-per-read, 1 generic),
    both transports, both gather carriers, against the oracle."""
    import charon_b200
    monkeypatch.setenv("CHARON_ASYNC_GATHER", carrier)
    idx, refs = common.build_oracle_index([200_000, 200_000], [0, 1], ["host", "microbial"], seed=17)
    gpu = common.gpu_from_oracle(idx)
    try:
        for name, cases in seam_batches:
            bases, seg_off = _batch(cases)
            want = idx.classify_batch(bases, seg_off)
            for mode in (0, 2, 1):
                gpu.force_generic(mode)
                for pack in ((1, 0) if mode != 1 else (0,)):
                    gpu.set_host_pack(pack)
                    got = gpu.count_batch(bases, seg_off)
                    bad = np.flatnonzero(got["num_hashes"] != want["num_hashes"])
                    assert bad.size == 0, (name, f"kernel mode {mode} pack {pack} carrier {carrier}: {bad.size} of {len(cases)} sequences differ; first: "
                                           f"d={cases[bad[0]][0]} twin={cases[bad[0]][1]} pos_a={cases[bad[0]][2]} "
                                           f"gpu={got['num_hashes'][bad[0]]} oracle={want['num_hashes'][bad[0]]}")
                    common.assert_parity(got, want)
    finally:
        gpu.force_generic(0)
        gpu.close()


@pytest.mark.gpu
def test_golden_period4_seam_sequence_on_the_gpu():
    import charon_b200
    s = open(os.path.join(GOLDEN, "seam_period4_941.txt")).read().strip()
    idx, refs = common.build_oracle_index([50_000, 50_000], [0, 1], ["host", "microbial"], seed=19)
    gpu = common.gpu_from_oracle(idx)
    try:
        want = np.sort(oracle.minimisers(s))
        rng = random.Random(4)
        for mode in (0, 2, 1):
            gpu.force_generic(mode)
            for pad in (0, 1, 2921 - 713, 712):     # also moved onto the CTA tile's seam and the second warp seam
                t = "".join(rng.choices("ACGT", k=pad)) + s
                assert np.array_equal(gpu.debug_minimisers(t), np.sort(oracle.minimisers(t))), (mode, pad)
        gpu.force_generic(0)
        assert len(gpu.debug_minimisers(s)) == 86 == len(want)
    finally:
        gpu.force_generic(0)
        gpu.close()


@pytest.mark.gpu
def test_two_hundred_thousand_random_reads_match_the_oracle():
    """>= 2e5 random 8 kb reads (the 2 000-read spot check let a 3e-5-per-read miscount through): every integer
    output and call of every read against the oracle, both transports."""
    import charon_b200
    idx, refs = common.build_oracle_index([5_000_000, 5_000_000], [0, 1], ["host", "microbial"], seed=23)
    gpu = common.gpu_from_oracle(idx)
    try:
        gpu.set_model(charon_b200.DehostOptions(), host_index=idx.host_index)
        n, L = 200_000, 8000
        bases, seg_off, _ = common.fixed_length_batch(77, refs, [0.45, 0.45], n, L)
        mq, cr = common.scalars(n)
        want = idx.classify_batch(bases, seg_off, mean_q=mq, compression=cr)
        for pack in (1, 0):
            gpu.set_host_pack(pack)
            got = gpu.classify_batch(bases, seg_off, mq, cr, np.full(n, L, np.uint32))
            common.assert_parity(got, want)
    finally:
        gpu.close()


@pytest.mark.gpu
def test_full_size_batch_properties(monkeypatch):
    """BASELINE's step size (131 072 reads x 8 kb = 1.05 G bases), where the oracle would take minutes: properties that need no
    oracle.  The outputs of a read may depend on nothing but the read: the same rows come back when the batch is permuted,
    cut in two, sent as ASCII or as two bits per base, and gathered by either carrier; and the first 4 096 reads agree with
    the oracle."""
    import charon_b200
    from charon_b200 import api as A
    idx, refs = common.build_oracle_index([20_000_000, 20_000_000], [0, 1], ["host", "microbial"], seed=29)
    n, L = 131_072, 8000
    bases, seg_off, _ = common.fixed_length_batch(78, refs, [0.45, 0.45], n, L)
    mq, cr = common.scalars(n)
    ln = np.full(n, L, np.uint32)
    keys = ("num_hashes", "counts", "unique_counts", "call", "confidence")
    results = {}
    for carrier in ("0", "1"):
        monkeypatch.setenv("CHARON_ASYNC_GATHER", carrier)
        gpu = common.gpu_from_oracle(idx)
        try:
            gpu.set_model(charon_b200.DehostOptions(), host_index=idx.host_index)
            gpu.set_host_pack(0)
            whole = gpu.classify_batch(bases, seg_off, mq, cr, ln)
            results[carrier] = whole
            if carrier == "0":
                continue
            packed = gpu.classify_batch(A.pack_batch(bases, seg_off, A.BASES_2BIT, align=64), None, mq, cr, ln)
            for k in keys:
                assert np.array_equal(packed[k], whole[k]), ("two-bit transport", k)
            assert np.array_equal(packed["prob"], whole["prob"], equal_nan=True)
            # cut in two
            h = 50_001
            a = gpu.classify_batch(bases[: h * L], seg_off[: h + 1], mq[:h], cr[:h], ln[:h])
            b = gpu.classify_batch(bases[h * L:], seg_off[h:] - seg_off[h], mq[h:], cr[h:], ln[h:])
            for k in keys:
                assert np.array_equal(np.concatenate([a[k], b[k]]), whole[k]), ("split", k)
            # permuted
            perm = np.random.default_rng(5).permutation(n)
            pb = np.ascontiguousarray(bases.reshape(n, L)[perm]).reshape(-1)
            p = gpu.classify_batch(pb, seg_off, mq, cr, ln)
            for k in keys:
                assert np.array_equal(p[k], whole[k][perm]), ("permutation", k)
            assert np.array_equal(p["prob"], whole["prob"][perm], equal_nan=True)
            # and the oracle on a slice
            m = 4096
            want = idx.classify_batch(bases[: m * L], seg_off[: m + 1], mean_q=mq[:m], compression=cr[:m])
            common.assert_parity({k: v[:m] for k, v in whole.items()}, want)
        finally:
            gpu.close()
    for k in keys:
        assert np.array_equal(results["0"][k], results["1"][k]), ("gather carriers", k)
    assert (results["1"]["call"] != 255).mean() > 0.8
