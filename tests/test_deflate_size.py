"""The compression-ratio scalar (reference src/utils.cpp:114-124: gzip::compress = zlib level 6, gzip wrapper, memLevel 8).

charon_b200/csrc/deflate_size.cuh restates zlib's deflate_slow + trees.c far enough to give the SIZE of the stream; its
sequential core is compiled for the host here (tests/deflate_size_host.cpp, test infrastructure) and pinned against the real
zlib of this box (python's zlib module = the system libz; zlib is a third-party dependency of the reference, absent from
/root/reference, versions 1.2.11 .. 1.3.1 behave alike at level 6).  The GPU kernel built from the same header is compared
with zlib in the gpu test below through the C-ABI."""
import os
import random
import subprocess
import zlib

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def gzip_size(s: bytes) -> int:
    c = zlib.compressobj(-1, zlib.DEFLATED, 31, 8, 0)
    return len(c.compress(s) + c.flush())


def sequences(seed, n_random=250):
    rng = random.Random(seed)
    rs = lambda n, letters="ACGT": "".join(rng.choice(letters) for _ in range(n))
    seqs = [rs(n) for n in [0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 19, 50, 100, 261, 262, 263, 264, 500, 1000, 3000, 5000, 8000, 8000, 12000, 20000]]
    seqs += ["A" * 5000, "AC" * 3000, "ACG" * 2500, "ACGTTGCA" * 1000, "A" * 300, "A" * 258 + "C" + "A" * 259, "N" * 1000, "T" * 3, "GGGG"]
    for n in [500, 4000, 9000]:            # a period-37 repeat with point changes: long matches at many distances
        unit = rs(37)
        s = list((unit * (n // 37 + 1))[:n])
        for i in range(0, n, 53):
            s[i] = rng.choice("ACGTN")
        seqs.append("".join(s))
    base = rs(3000)
    seqs += [base + base, base + base[:1500] + "N" + base[1500:], base[:100] + rs(4200) + base[:100]]      # TOO_FAR: a 3-letter match 4 kb back
    # match-length and window limits: runs around MAX_MATCH = 258, periods 1..40 across the 32 506-letter distance limit and the
    # 4096-letter TOO_FAR rule, Fibonacci-like letter frequencies (deep Huffman trees)
    seqs += ["A" * n for n in (257, 258, 259, 260, 516, 517, 1000)] + ["AC" * 129 + "G", "ACGT" * 65 + "T"]
    for period in (1, 2, 3, 5, 7, 11, 19, 40):
        unit = rs(period)
        seqs.append((unit * (33000 // period + 1))[:33000])
    seqs.append(rs(5000) + "N" * 3 + rs(4090) + "NNN" + rs(100))          # a 3-letter match about 4096 back
    seqs.append("".join(c * k for c, k in zip("ACGTN" * 4, [1, 1, 2, 3, 5, 8, 13, 21, 34, 55, 89, 144, 233, 377, 610, 987, 1597, 2584, 4181, 6765])))
    for _ in range(n_random):
        n = rng.choice([30, 200, 900, 2500, 6000, 8000, 15000])
        if rng.random() < 0.7:
            seqs.append("".join(rng.choice("ACGT") if rng.random() > 0.01 else "N" for _ in range(n)))
        else:
            seqs.append("".join(rng.choice("AC" if rng.random() < 0.9 else "GT") for _ in range(n)))        # skewed: dynamic trees win
    return seqs


LONG = [33000, 40000, 65273, 65274, 65280, 70000, 98000, 131072, 200000]   # several blocks; beyond 65273 letters zlib slides its window (once, twice, ... five times)


@pytest.fixture(scope="module")
def host_exe(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("dsz") / "deflate_size_host")
    gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    subprocess.check_call([gxx, "-O2", "-std=c++17", os.path.join(ROOT, "tests", "deflate_size_host.cpp"), "-o", exe])
    return exe


def test_restated_deflate_size_equals_zlib_on_the_host(host_exe):
    rng = random.Random(5)
    seqs = sequences(1) + ["".join(rng.choice("ACGT") for _ in range(n)) for n in LONG]
    seqs += ["A" * 100_000, ("ACGTTGCATT" * 9000)[:88_000], "".join(rng.choice("AC") for _ in range(150_000))]     # long and repetitive: chain ranks far beyond 255
    out = subprocess.run([host_exe], input=("\n".join(seqs) + "\n").encode(), capture_output=True, check=True).stdout.decode().split()
    assert len(out) == len(seqs)
    bad = [(len(s), int(o), gzip_size(s.encode())) for s, o in zip(seqs, out) if int(o) != gzip_size(s.encode())]
    assert not bad, f"{len(bad)} of {len(seqs)} sizes differ from zlib {zlib.ZLIB_RUNTIME_VERSION}; first (length, got, zlib): {bad[:5]}"
    assert gzip_size(b"") == 20 and int(out[0]) == 20


def _batch_of(seqs):
    from tests import common
    return common.ragged_batch(None, [np.frombuffer(s.encode(), np.uint8) for s in seqs])


@pytest.mark.gpu
def test_gpu_gzip_size_equals_zlib():
    """cg_compression_batch against the box's zlib: sizes bit-exact, ratios as the reference forms them, both transports."""
    import charon_b200
    from tests import common
    idx, refs = common.build_oracle_index([50_000, 50_000], [0, 1], ["host", "microbial"], seed=41)
    gpu = common.gpu_from_oracle(idx)
    try:
        rng = random.Random(9)
        seqs = sequences(2, n_random=400) + ["".join(rng.choice("ACGT") for _ in range(n)) for n in LONG]
        seqs += ["A" * 100_000, ("ACGTTGCATT" * 9000)[:88_000]]
        seqs += ["acgtnACGTN" * 300, "ACGURYKM" * 500]                                                            # folding to dna5
        want = np.array([gzip_size(s.upper().translate(str.maketrans("URYKMSWBDHV", "TNNNNNNNNNN")).encode()) for s in seqs], np.uint32)
        bases, seg_off = _batch_of(seqs)
        for pack in (1, 0):
            gpu.set_host_pack(pack)
            got = gpu.compression_batch(bases, seg_off)
            bad = np.flatnonzero(got["size"] != want)
            assert bad.size == 0, f"pack {pack}: {bad.size} of {len(seqs)} sizes differ; first: read {bad[0]} length {len(seqs[bad[0]])} gpu {got['size'][bad[0]]} zlib {want[bad[0]]}"
            lens = np.array([len(s) for s in seqs], np.float64)
            with np.errstate(divide="ignore", invalid="ignore"):
                ratio = (want.astype(np.float64) / lens).astype(np.float32)
            assert np.array_equal(got["ratio"], ratio, equal_nan=True)
        # paired reads: the two mates are compressed as one string (src/dehost_main.cpp:449-452)
        pairs = [s for s in seqs if len(s) <= 20_000][30:90]       # a pair longer than the kernel's limit would be flagged, not compressed
        bases2, seg_off2 = _batch_of(pairs)
        read_seg = np.arange(0, len(pairs) + 1, 2, dtype=np.uint32)
        got2 = gpu.compression_batch(bases2, seg_off2, read_seg=read_seg)
        want2 = np.array([gzip_size((pairs[i] + pairs[i + 1]).encode()) for i in range(0, len(pairs), 2)], np.uint32)
        assert np.array_equal(got2["size"], want2)
    finally:
        gpu.close()


@pytest.mark.gpu
def test_classify_with_ratios_computed_on_the_device():
    """cg_classify_batch / cg_submit_batch with cg_read_scalars.compression == NULL: same calls as with zlib's ratios passed in."""
    import charon_b200
    from tests import common
    idx, refs = common.build_oracle_index([300_000, 300_000], [0, 1], ["host", "microbial"], seed=43)
    gpu = common.gpu_from_oracle(idx)
    try:
        import oracle
        gpu.set_model(charon_b200.ClassifyOptions(), host_index=0)                  # min_compression 0.15 (classify_arguments.hpp)
        n, L = 3000, 3000
        bases, seg_off, _ = common.fixed_length_batch(8, refs, [0.45, 0.45], n, L)
        bases[: 40 * L] = np.frombuffer(b"A" * (40 * L), np.uint8)                  # low-complexity reads: the compression gate fires
        texts = [bytes(bases[i * L:(i + 1) * L]) for i in range(n)]
        cr = np.array([np.float32(gzip_size(t) / float(L)) for t in texts], np.float32)
        mq = np.full(n, 20.0, np.float32)
        ln = np.full(n, L, np.uint32)
        want = idx.classify_batch(bases, seg_off, mean_q=mq, compression=cr, opts=oracle.default_options(dehost=False))
        assert (cr < 0.15).sum() == 40 and (want["call"][40:] != 255).any()
        got = gpu.classify_batch(bases, seg_off, mq, None, ln)
        common.assert_parity(got, want, prob_rtol=2e-4)
        assert gpu.info().last_gzip_kernel_ms > 0
        assert np.array_equal(gpu.fetch_compression(0, n), cr)
    finally:
        gpu.close()


@pytest.mark.gpu
def test_gpu_gzip_size_on_twenty_thousand_reads():
    """20 000 synthetic 8 kb reads (2 % substitutions, insertions and deletions each) against zlib: every size."""
    import charon_b200
    from charon_b200 import api as A
    from tests import common
    refs = [common.synth.reference(71, 3_000_000), common.synth.reference(72, 3_000_000)]
    gpu = charon_b200.GpuClassifier.from_references(refs, [0, 1], [0, 1], 2, device=0)
    try:
        n, L = 20_000, 8000
        bases, seg_off, _ = common.fixed_length_batch(73, refs, [0.45, 0.45], n, L)
        got = gpu.compression_batch(A.pack_batch(bases, seg_off, A.BASES_2BIT, align=64), None)
        want = np.array([gzip_size(bytes(bases[i * L:(i + 1) * L])) for i in range(n)], np.uint32)
        bad = np.flatnonzero(got["size"] != want)
        assert bad.size == 0, f"{bad.size} of {n} sizes differ; first: read {bad[0]} gpu {got['size'][bad[0]]} zlib {want[bad[0]]}"
        assert np.array_equal(got["ratio"], (want.astype(np.float64) / L).astype(np.float32))
    finally:
        gpu.close()


@pytest.mark.gpu
def test_two_slots_with_device_ratios():
    """cg_submit_batch on both slots with cg_read_scalars.compression == NULL, different batches in flight: each slot's calls equal
    the oracle fed zlib's ratios, and cg_fetch_compression returns each slot's own ratios after its wait."""
    import charon_b200
    import oracle
    from tests import common
    idx, refs = common.build_oracle_index([300_000, 300_000], [0, 1], ["host", "microbial"], seed=47)
    gpu = common.gpu_from_oracle(idx)
    try:
        gpu.set_model(charon_b200.ClassifyOptions(), host_index=0)
        opts = oracle.default_options(dehost=False)
        batches = []
        for s, (n, L) in enumerate([(1500, 2500), (900, 4000), (2000, 1200)]):
            bases, seg_off, _ = common.fixed_length_batch(80 + s, refs, [0.45, 0.45], n, L)
            bases[: 20 * L] = np.frombuffer(((b"AC" if s else b"A") * (20 * L))[: 20 * L], np.uint8)                 # low-complexity reads
            cr = np.array([np.float32(gzip_size(bytes(bases[i * L:(i + 1) * L])) / float(L)) for i in range(n)], np.float32)
            mq = np.full(n, 20.0, np.float32)
            ln = np.full(n, L, np.uint32)
            want = idx.classify_batch(bases, seg_off, mean_q=mq, compression=cr, opts=opts)
            batches.append((bases, seg_off, mq, ln, cr, want, n))
        res = {}
        for i, (bases, seg_off, mq, ln, cr, want, n) in enumerate(batches):
            slot = i & 1
            if i >= 2:
                got = gpu.wait(slot)
                j = i - 2
                common.assert_parity(got, batches[j][5], prob_rtol=2e-4)
                assert np.array_equal(gpu.fetch_compression(slot, batches[j][6]), batches[j][4])
            res[slot] = gpu.submit(slot, bases, seg_off, mq, None, ln)
        for j in (1, 2):
            slot = j & 1
            got = gpu.wait(slot)
            common.assert_parity(got, batches[j][5], prob_rtol=2e-4)
            assert np.array_equal(gpu.fetch_compression(slot, batches[j][6]), batches[j][4])
    finally:
        gpu.close()


def structured_strings(count, seed=20260501):
    """Strings assembled from the building blocks that steer zlib's decisions — random stretches, copies of earlier stretches
    at every distance class (short, beyond TOO_FAR, beyond the 32 506-letter limit), runs, short periods, point changes — in
    five alphabets from one to five letters."""
    rng = random.Random(seed)
    seqs = []
    for _ in range(count):
        letters = rng.choice(["A", "AC", "ACG", "ACGT", "ACGTN"])
        target = rng.choice([40, 300, 2000, 9000, 9000, 40000])
        parts, total = [], 0
        while total < target:
            kind = rng.random()
            if kind < 0.35 or not parts:
                p = "".join(rng.choice(letters) for _ in range(rng.choice([1, 2, 3, 5, 17, 60, 300, 1500])))
            elif kind < 0.65:                                  # copy of an earlier stretch, possibly with a point change
                src = "".join(parts)
                ln = rng.choice([3, 4, 5, 8, 16, 40, 128, 129, 258, 259, 600])
                st = rng.randrange(0, max(1, len(src) - ln))
                p = list(src[st:st + ln])
                if p and rng.random() < 0.4:
                    p[rng.randrange(len(p))] = rng.choice(letters)
                p = "".join(p)
            elif kind < 0.8:
                p = rng.choice(letters) * rng.choice([2, 3, 7, 30, 257, 258, 259, 700])
            else:
                unit = "".join(rng.choice(letters) for _ in range(rng.choice([2, 3, 4, 7, 13])))
                p = unit * rng.choice([2, 5, 40, 200])
            parts.append(p)
            total += len(p)
        seqs.append("".join(parts))
    return seqs


def test_restated_deflate_size_on_structured_random_strings(host_exe):
    """CPU, seeded: 1 500 structured strings, every size against zlib."""
    seqs = structured_strings(1500)
    out = subprocess.run([host_exe], input=("\n".join(seqs) + "\n").encode(), capture_output=True, check=True).stdout.decode().split()
    assert len(out) == len(seqs)
    bad = [(len(s), int(o), gzip_size(s.encode())) for s, o in zip(seqs, out) if int(o) != gzip_size(s.encode())]
    assert not bad, f"{len(bad)} of {len(seqs)} sizes differ from zlib; first (length, got, zlib): {bad[:5]}"


@pytest.mark.gpu
def test_gpu_gzip_size_on_structured_random_strings():
    """The same family of strings (another seed) through cg_compression_batch, both transports."""
    import charon_b200
    from tests import common
    idx, refs = common.build_oracle_index([50_000, 50_000], [0, 1], ["host", "microbial"], seed=49)
    gpu = common.gpu_from_oracle(idx)
    try:
        seqs = structured_strings(1200, seed=77)
        want = np.array([gzip_size(s.encode()) for s in seqs], np.uint32)
        bases, seg_off = _batch_of(seqs)
        for pack in (1, 0):
            gpu.set_host_pack(pack)
            got = gpu.compression_batch(bases, seg_off)
            bad = np.flatnonzero(got["size"] != want)
            assert bad.size == 0, f"pack {pack}: {bad.size} of {len(seqs)} differ; first: length {len(seqs[bad[0]])} gpu {got['size'][bad[0]]} zlib {want[bad[0]]}"
    finally:
        gpu.close()
