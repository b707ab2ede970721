"""Base transports of the C-ABI (include/charon_gpu.h enum cg_bases_format): ASCII, 4-bit dna5 ranks, 2 bits per base with
an exception list for everything that is not A/C/G/T/U.  CPU: the 2-bit writer against its scalar definition.  GPU: every
integer output must be the same whichever format carried the batch — padded segments, paired reads, N runs, the generic
any-(k, w) kernel (which unpacks on the device) — and equal to the oracle's."""
import ctypes as C
import numpy as np
import pytest
import oracle
from tests import common

ALPHABET = np.frombuffer(b"ACGTacgtNnUuRYKMSWBDHV-*.!@`[{\x00\x01\x7f\xc1\xe1\xff", np.uint8)


def _two_bit_reference(b):
    u = b & 0xDF
    c = np.zeros(b.shape, np.uint8)
    c[u == ord("C")] = 1
    c[u == ord("G")] = 2
    c[(u == ord("T")) | (u == ord("U"))] = 3
    exc = np.flatnonzero(~((u == ord("A")) | (u == ord("C")) | (u == ord("G")) | (u == ord("T")) | (u == ord("U"))))
    c = np.concatenate([c, np.zeros((-len(c)) % 4, np.uint8)])
    return (c[0::4] | (c[1::4] << 2) | (c[2::4] << 4) | (c[3::4] << 6)).astype(np.uint8), exc.astype(np.uint64)


def _pack2(b, pos0=0, cap=None):
    from charon_b200 import api
    L = api.load_library()
    out = np.full((len(b) + 3) // 4 + 8, 0xEE, np.uint8)
    cap = len(b) if cap is None else cap
    exc = np.zeros(max(cap, 1), np.uint64)
    k = C.c_uint64(0)
    assert L.cg_pack_bases_2bit(b.ctypes.data, len(b), out.ctypes.data, pos0, exc.ctypes.data, cap, C.byref(k)) == 0
    return out, exc, int(k.value)


@pytest.mark.parametrize("n", [0, 1, 3, 4, 5, 31, 32, 33, 63, 64, 65, 127, 128, 129, 4099, 1_000_003])
def test_two_bit_writer_matches_scalar_definition(n):
    rng = np.random.default_rng(n + 1)
    b = rng.choice(ALPHABET, n) if n < 100_000 else rng.choice(np.frombuffer(b"ACGTN", np.uint8), n, p=[0.24, 0.25, 0.25, 0.25, 0.01])
    if 256 <= n < 100_000:
        b[:256] = np.arange(256, dtype=np.uint8)
    want, want_exc = _two_bit_reference(b)
    got, exc, k = _pack2(b, pos0=4 * 7)
    assert np.array_equal(got[:len(want)], want)
    assert (got[len(want):] == 0xEE).all()                       # nothing written past (n + 3) / 4 bytes
    assert k == len(want_exc) and np.array_equal(exc[:k], want_exc + np.uint64(28))


def test_two_bit_writer_reports_overflow_and_rejects_misaligned_runs():
    from charon_b200 import api
    b = np.frombuffer(b"ACGTNNNNACGTNRYK" * 20, np.uint8)
    _, want_exc = _two_bit_reference(b)
    got, exc, k = _pack2(b, cap=5)
    assert k == len(want_exc) and np.array_equal(exc[:5], want_exc[:5])
    L = api.load_library()
    out = np.zeros(100, np.uint8)
    assert L.cg_pack_bases_2bit(b.ctypes.data, len(b), out.ctypes.data, 2, None, 0, None) != 0


def test_pack_batch_pads_segments_and_keeps_lengths():
    from charon_b200 import api
    rng = np.random.default_rng(4)
    pieces = [rng.choice(np.frombuffer(b"ACGTN", np.uint8), n) for n in (0, 1, 63, 64, 65, 1000, 7)]
    bases, seg_off = common.ragged_batch(rng, pieces)
    for fmt in (api.BASES_2BIT, api.BASES_RANK4):
        pb = api.pack_batch(bases, seg_off, fmt)
        assert np.array_equal(pb.seg_len, np.diff(seg_off)) and (pb.seg_off % 64 == 0).all()
        assert (np.diff(pb.seg_off) >= pb.seg_len).all()
    pb = api.pack_batch(bases, seg_off, api.BASES_2BIT)
    n_per_piece = [int((p == ord("N")).sum()) for p in pieces]
    assert len(pb.exc_pos) == sum(n_per_piece) and (np.diff(pb.exc_pos.astype(np.int64)) > 0).all()


@pytest.fixture(scope="module")
def small():
    idx, refs = common.build_oracle_index([300_000, 300_000], [0, 1], ["host", "microbial"], seed=11)
    gpu = common.gpu_from_oracle(idx)
    import charon_b200
    gpu.set_model(charon_b200.DehostOptions(), host_index=idx.host_index)
    yield idx, refs, gpu
    gpu.close()


def _mixed_batch(refs, rng):
    lens = [0, 1, 18, 19, 20, 41, 42, 63, 64, 65, 139, 500, 754, 755, 2985, 2986, 6000, 12_345, 0, 33, 70_000, 8000, 8000, 65_536]
    pieces = []
    for ln in lens:
        st = int(rng.integers(0, 200_000))
        p = refs[int(rng.integers(0, 2))][st:st + ln].copy()
        pieces.append(p)
    # bases that are not A/C/G/T: single N, N runs across tile and vector boundaries, lower case, IUPAC, U
    pieces[16][100] = ord("N"); pieces[16][713:760] = ord("N"); pieces[16][3000:3003] = np.frombuffer(b"ryk", np.uint8)
    pieces[17][::997] = ord("n"); pieces[17][5000:5100] = ord("u")
    pieces[20][60_000:60_900] = ord("N"); pieces[20][1] = ord("N"); pieces[20][-1] = ord("N")
    pieces[21][:] = ord("N")
    pieces[10][5] = ord("N")
    pieces[4][0] = ord("N")
    pieces.append(np.frombuffer(b"acgtnACGTNuU" * 300 + b"RYKMSWBDHV" * 50 + b"ACGGTCA" * 500, np.uint8).copy())
    return pieces


@pytest.mark.gpu
@pytest.mark.parametrize("carrier", ["0", "1"])
def test_every_format_gives_the_oracles_integers(monkeypatch, carrier):
    import charon_b200
    from charon_b200 import api
    monkeypatch.setenv("CHARON_ASYNC_GATHER", carrier)
    idx, refs = common.build_oracle_index([300_000, 300_000], [0, 1], ["host", "microbial"], seed=11)
    gpu = common.gpu_from_oracle(idx)
    try:
        gpu.set_model(charon_b200.DehostOptions(), host_index=idx.host_index)
        rng = np.random.default_rng(8)
        pieces = _mixed_batch(refs, rng)
        bases, seg_off = common.ragged_batch(rng, pieces)
        n = len(pieces)
        mq, cr = common.scalars(n, rng)
        length = np.diff(seg_off).astype(np.uint32)
        want = idx.classify_batch(bases, seg_off, mean_q=mq, compression=cr)
        for mode in (0, 2, 1):           # warp-per-read + CTA-per-read dispatch, CTA-per-read only, generic (unpacks on the device)
            gpu.force_generic(mode)
            for fmt in (api.BASES_2BIT, api.BASES_RANK4):
                for align in (64, 4, 128):
                    pb = api.pack_batch(bases, seg_off, fmt, align=align)
                    got = gpu.classify_batch(pb, None, mq, cr, length)
                    common.assert_parity(got, want)
            gpu.set_host_pack(0)
            common.assert_parity(gpu.classify_batch(bases, seg_off, mq, cr, length), want)
            gpu.set_host_pack(-1)
        gpu.force_generic(0)
        # paired reads: two (padded) segments per read
        m = (n // 2) * 2
        b2, so2 = common.ragged_batch(rng, pieces[:m])
        read_seg = np.arange(0, m + 1, 2, dtype=np.uint32)
        want2 = idx.classify_batch(b2, so2, read_seg=read_seg)
        for fmt in (api.BASES_2BIT, api.BASES_RANK4):
            pb = api.pack_batch(b2, so2, fmt, read_seg=read_seg)
            common.assert_parity(gpu.count_batch(pb, None), want2)
        # two slots in flight with packed input
        pb = api.pack_batch(bases, seg_off, api.BASES_2BIT)
        r0 = gpu.submit(0, pb, None, mq, cr, length)
        r1 = gpu.submit(1, pb, None)
        gpu.wait(0); gpu.wait(1)
        common.assert_parity(r0, want)
        common.assert_parity(r1, want)
    finally:
        gpu.force_generic(0)
        gpu.close()


@pytest.mark.gpu
def test_two_bit_batch_of_many_reads(small):
    """A few thousand 8 kb reads with sprinkled N: the hot shape of the transport (64-base aligned reads, exception list)."""
    from charon_b200 import api
    idx, refs, gpu = small
    n, L = 3000, 8000
    bases, seg_off, _ = common.fixed_length_batch(91, refs, [0.45, 0.45], n, L)
    rng = np.random.default_rng(2)
    bases = bases.copy()
    bases[rng.integers(0, len(bases), 5000)] = ord("N")
    mq, cr = common.scalars(n, rng)
    want = idx.classify_batch(bases, seg_off, mean_q=mq, compression=cr)
    pb = api.pack_batch(bases, seg_off, api.BASES_2BIT)
    assert len(pb.exc_pos) > 4000
    common.assert_parity(gpu.classify_batch(pb, None, mq, cr, np.full(n, L, np.uint32)), want)


@pytest.mark.gpu
def test_bad_batches_are_rejected(small):
    """validate_batch: offsets that are not monotone, exception lists out of order or with the wrong format."""
    import charon_b200
    from charon_b200 import api
    idx, refs, gpu = small
    b = np.frombuffer(b"ACGT" * 100, np.uint8)
    for so in ([1, 400], [0, 300, 200, 400]):
        with pytest.raises(charon_b200.CharonGpuError):
            gpu.count_batch(b, np.array(so, np.uint64))
    with pytest.raises(charon_b200.CharonGpuError):
        gpu.count_batch(b, np.array([0, 200, 400], np.uint64), read_seg=np.array([0, 2, 1, 2], np.uint32)[:3])
    pb = api.pack_batch(b, np.array([0, 400], np.uint64), api.BASES_2BIT)
    pb.exc_pos = np.array([7, 3], np.uint64)
    with pytest.raises(charon_b200.CharonGpuError):
        gpu.count_batch(pb, None)
    pb.exc_pos = np.array([3, 7], np.uint64); pb.fmt = api.BASES_RANK4
    with pytest.raises(charon_b200.CharonGpuError):
        gpu.count_batch(pb, None)
    pb.fmt = api.BASES_2BIT; pb.seg_len = np.array([500], np.uint64)
    with pytest.raises(charon_b200.CharonGpuError):
        gpu.count_batch(pb, None)


def test_two_bit_writer_avx2_path_in_a_subprocess():
    """The AVX-512 path hides the AVX2 one on this CPU: run the scalar-definition check again with CHARON_NO_AVX512=1."""
    import os, subprocess, sys
    code = ("import numpy as np; from tests.test_base_formats import _pack2, _two_bit_reference, ALPHABET\n"
            "rng = np.random.default_rng(3)\n"
            "for n in (0, 5, 31, 32, 33, 64, 100, 4099, 70001):\n"
            "    b = rng.choice(ALPHABET, n)\n"
            "    want, we = _two_bit_reference(b); got, exc, k = _pack2(b)\n"
            "    assert np.array_equal(got[:len(want)], want) and k == len(we) and np.array_equal(exc[:k], we), n\n"
            "print('ok')\n")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.check_output([sys.executable, "-c", code], cwd=root, env=dict(os.environ, CHARON_NO_AVX512="1"))
    assert out.strip().endswith(b"ok")
