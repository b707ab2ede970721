"""Packed transport of the bases (include/charon_gpu.h cg_set_host_pack / cg_pack_bases): the library may send 4-bit dna5
ranks over PCIe instead of ASCII.  CPU: the packer against its scalar definition.  GPU: every integer must be the same
whichever transport carried the batch, and equal to the oracle's."""
import numpy as np
import pytest
import oracle
from tests import common


def _rank_reference(b):
    u = b & 0xDF
    r = np.full(b.shape, 3, np.uint8)
    r[u == ord("A")] = 0
    r[u == ord("C")] = 1
    r[u == ord("G")] = 2
    r[(u == ord("T")) | (u == ord("U"))] = 4
    if r.size & 1:
        r = np.concatenate([r, np.zeros(1, np.uint8)])
    return (r[0::2] | (r[1::2] << 4)).astype(np.uint8)


@pytest.mark.parametrize("n", [0, 1, 2, 31, 32, 63, 64, 65, 4097, 262_144, 262_145, 3 * 262_144 + 77, 40_000_001])
def test_packer_matches_scalar_definition(n):
    from charon_b200 import api
    rng = np.random.default_rng(n)
    b = rng.choice(np.frombuffer(b"ACGTacgtNnUuRYKMSWBDHV-*.!@`[{\x00\x01\x7f\xc1\xe1\xff", np.uint8), n) if n < 1_000_000 else \
        rng.choice(np.frombuffer(b"ACGTN", np.uint8), n, p=[0.24, 0.25, 0.25, 0.24, 0.02])
    if 256 <= n < 1_000_000:
        b[:256] = np.arange(256, dtype=np.uint8)       # every byte value once
    got = api.pack_bases(b)
    assert got.shape == ((n + 1) // 2,)
    assert np.array_equal(got, _rank_reference(b))


def test_packer_is_deterministic_across_pool_sizes():
    from charon_b200 import api
    rng = np.random.default_rng(5)
    b = rng.choice(np.frombuffer(b"ACGTN", np.uint8), 5_000_003)
    a = api.pack_bases(b)
    for _ in range(3):
        assert np.array_equal(api.pack_bases(b), a)


@pytest.fixture(scope="module")
def small():
    idx, refs = common.build_oracle_index([300_000, 300_000], [0, 1], ["host", "microbial"], seed=11)
    gpu = common.gpu_from_oracle(idx)
    import charon_b200
    gpu.set_model(charon_b200.DehostOptions(), host_index=idx.host_index)
    yield idx, refs, gpu
    gpu.close()


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [1, 0])
def test_minimiser_stream_same_on_both_transports(small, mode):
    from tests.test_gpu_parity import _tie_heavy_cases
    idx, refs, gpu = small
    gpu.set_host_pack(mode)
    for s in _tie_heavy_cases() + ["acgtnACGTNuU" * 300 + "RYKMSWBDHV" * 50 + "ACGGTCA" * 500]:
        want = np.sort(oracle.minimisers(s))
        got = gpu.debug_minimisers(s)
        assert np.array_equal(got, want), (mode, len(s), s[:40], len(got), len(want))
        if len(s):
            assert gpu.info().last_packed == mode
    gpu.set_host_pack(-1)


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [1, 0])
def test_ragged_odd_offsets_long_reads_and_pairs(small, mode):
    """Segments start at odd base indices (nibble-unaligned tiles), lengths straddle every tile / vector boundary, reads beyond
    65 536 bases go to the CTA-per-read kernel, shorter than one window take the single-window path."""
    idx, refs, gpu = small
    gpu.set_host_pack(mode)
    rng = np.random.default_rng(31 + mode)
    lens = [0, 1, 7, 18, 19, 20, 40, 41, 42, 63, 64, 65, 713 + 40, 714 + 40, 735, 736, 737, 1426 + 40, 2985, 2986, 8001, 12_345, 65_535, 65_536,
            65_537, 120_001, 33, 5]
    pieces = []
    for ln in lens:
        st = int(rng.integers(0, 300_000 - ln))
        pieces.append(refs[int(rng.integers(0, 2))][st:st + ln].copy())
    pieces.append(np.frombuffer((b"ACGTTGCA" * 20000)[:100_001], np.uint8))
    pieces.append(np.frombuffer(b"acgtn" * 700 + b"RYKMSWBDHVU" * 30, np.uint8))
    bases, seg_off = common.ragged_batch(rng, pieces)
    n = len(pieces)
    mq, cr = common.scalars(n)
    want = idx.classify_batch(bases, seg_off, mean_q=mq, compression=cr)
    got = gpu.classify_batch(bases, seg_off, mq, cr, np.diff(seg_off).astype(np.uint32))
    common.assert_parity(got, want)
    assert gpu.info().last_packed == mode
    # paired reads: two segments per read
    pieces2 = pieces[:28]
    bases2, seg_off2 = common.ragged_batch(rng, pieces2)
    read_seg = np.arange(0, 29, 2, dtype=np.uint32)
    mq2, cr2 = common.scalars(14)
    len2 = np.add.reduceat(np.diff(seg_off2), read_seg[:-1]).astype(np.uint32)
    want2 = idx.classify_batch(bases2, seg_off2, read_seg=read_seg, mean_q=mq2, compression=cr2)
    got2 = gpu.classify_batch(bases2, seg_off2, mq2, cr2, len2, read_seg=read_seg)
    common.assert_parity(got2, want2)
    gpu.set_host_pack(-1)


@pytest.mark.gpu
def test_pipeline_with_packing_and_transport_switches(small):
    """Two slots in flight with packed batches, then the same batches as ASCII: identical integers; the library's own packing
    (two bits per base since round 2) quarters the copied bases."""
    idx, refs, gpu = small
    n, L = 3000, 4001            # odd read length: every other read starts on a high nibble
    batches = []
    for i in range(4):
        bases, seg_off, _ = common.fixed_length_batch(100 + i, refs, [0.45, 0.45], n, L)
        mq, cr = common.scalars(n, np.random.default_rng(i))
        batches.append((bases, seg_off, mq, cr, np.full(n, L, np.uint32)))
    want = [idx.classify_batch(b[0], b[1], mean_q=b[2], compression=b[3]) for b in batches]
    sent = {}
    for mode in (1, 0, 1):
        gpu.set_host_pack(mode)
        gpu.submit(0, *batches[0])
        for i in range(1, len(batches) + 1):
            if i < len(batches):
                gpu.submit(i & 1, *batches[i])
            got = gpu.wait((i - 1) & 1)
            common.assert_parity(got, want[i - 1])
        gpu.select_slot(0)
        sent[mode] = gpu.info().last_h2d_bytes
    gpu.set_host_pack(-1)
    assert sent[0] - sent[1] == n * L - (n * L + 3) // 4


@pytest.mark.gpu
def test_many_bins_and_other_kernels_on_packed_input():
    """Wide rows (MODE 1-3 kernels) and the forced CTA-per-read kernel decode the packed transport too; a (k, w) other than
    (19, 41) silently stays on ASCII."""
    for nbins, bin_cat, ncat, force in ((70, [i % 9 for i in range(70)], 9, 0), (3, [0, 1, 1], 2, 0), (2, [0, 1], 2, 2), (20, [i % 5 for i in range(20)], 5, 0)):
        idx, refs = common.build_oracle_index([20_000] * nbins, bin_cat, [f"c{i}" for i in range(ncat)], seed=100 + nbins)
        gpu = common.gpu_from_oracle(idx)
        try:
            gpu.set_host_pack(1)
            if force:
                gpu.force_generic(force)
            bases, seg_off, _ = common.fixed_length_batch(7, refs, [0.9 / nbins] * nbins, 300, 3001)
            mq, cr = common.scalars(300)
            want = idx.classify_batch(bases, seg_off, mean_q=mq, compression=cr, opts=oracle.default_options(dehost=False))
            common.assert_parity(gpu.count_batch(bases, seg_off), want)
            assert gpu.info().last_packed == 1
        finally:
            gpu.close()
    idx, refs = common.build_oracle_index([50_000, 50_000], [0, 1], ["human", "other"], seed=3, k=13, w=40)
    gpu = common.gpu_from_oracle(idx)
    gpu.set_host_pack(1)
    bases, seg_off, _ = common.fixed_length_batch(7, refs, [0.5, 0.5], 200, 2000)
    common.assert_parity(gpu.count_batch(bases, seg_off), idx.classify_batch(bases, seg_off))
    assert gpu.info().last_packed == 0
    gpu.close()


@pytest.mark.gpu
def test_random_ragged_batches_agree_across_transports(small):
    """Seeded random batches (lengths from one base to several tiles, random alphabets incl. N runs and lower case, random
    pairing): ASCII and packed transport give the same integers as each other and as the oracle."""
    idx, refs, gpu = small
    rng = np.random.default_rng(2024)
    alphabets = [b"ACGT", b"ACGTN", b"acgtACGT", b"AC", b"ACGTRYKM"]
    for trial in range(6):
        pieces = []
        for _ in range(int(rng.integers(20, 80))):
            ln = int(rng.choice([rng.integers(0, 64), rng.integers(64, 800), rng.integers(800, 6000)]))
            kind = rng.integers(0, 4)
            if kind == 0:
                st = int(rng.integers(0, 300_000 - ln))
                p = refs[int(rng.integers(0, 2))][st:st + ln].copy()
            elif kind == 1:
                p = rng.choice(np.frombuffer(alphabets[int(rng.integers(0, len(alphabets)))], np.uint8), ln)
            elif kind == 2:
                unit = rng.choice(np.frombuffer(b"ACGT", np.uint8), int(rng.integers(1, 30)))
                p = np.tile(unit, ln // len(unit) + 1)[:ln]
            else:
                st = int(rng.integers(0, 300_000 - ln))
                p = refs[0][st:st + ln].copy()
                if ln > 40:
                    a = int(rng.integers(0, ln - 20))
                    p[a:a + int(rng.integers(1, 20))] = ord("N")
            pieces.append(np.ascontiguousarray(p, np.uint8))
        paired = bool(trial & 1)
        if paired and len(pieces) & 1:
            pieces.pop()
        bases, seg_off = common.ragged_batch(rng, pieces)
        read_seg = np.arange(0, len(pieces) + 1, 2, dtype=np.uint32) if paired else None
        n = len(pieces) // 2 if paired else len(pieces)
        mq, cr = common.scalars(n, rng)
        seglen = np.diff(seg_off)
        length = (np.add.reduceat(seglen, read_seg[:-1]) if paired else seglen).astype(np.uint32)
        want = idx.classify_batch(bases, seg_off, read_seg=read_seg, mean_q=mq, compression=cr)
        for mode in (1, 0):
            gpu.set_host_pack(mode)
            got = gpu.classify_batch(bases, seg_off, mq, cr, length, read_seg=read_seg)
            common.assert_parity(got, want)
            assert gpu.info().last_packed == (mode if len(bases) else 0)
    gpu.set_host_pack(-1)
