"""Structural self-check of the sdsl::select_support_mcl tables that `charon index` appends to the `.idx` file
(host/index_writer.hpp).  The real sdsl-lite is not on this image, so the byte layout itself stays
unverified — but whatever reads these tables the way sdsl does (superblock start every 4096 occurrences; a "long" superblock
lists all its positions, otherwise one offset every 64 occurrences relative to the superblock start, then a scan) must get
select(i) right: that is re-derived here from the written bytes and compared with a direct scan, for set and unset bits,
dense and sparse vectors, superblocks that end exactly on a boundary, and the unset padding of the last word."""
import os
import struct
import subprocess
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def dumper(tmp_path_factory):
    d = tmp_path_factory.mktemp("sel")
    exe = str(d / "dump")
    gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    subprocess.check_call([gxx, "-O1", "-std=c++14", os.path.join(ROOT, "tests", "select_support_dump.cpp"), "-o", exe])
    return exe, d


class _Reader:
    def __init__(self, data):
        self.b, self.o = data, 0

    def take(self, fmt):
        v = struct.unpack_from("<" + fmt, self.b, self.o)
        self.o += struct.calcsize("<" + fmt)
        return v[0]

    def int_vector(self):
        width, growth, bits = self.take("B"), self.take("f"), self.take("Q")
        assert growth == 1.5 and (width == 0 or bits % width == 0)
        nwords = (bits + 63) // 64
        words = np.frombuffer(self.b, np.uint64, nwords, self.o)
        self.o += 8 * nwords
        if not width:
            return []
        bitsarr = np.unpackbits(words.view(np.uint8), bitorder="little")[:bits].reshape(-1, width)
        return [int(x) for x in (bitsarr.astype(object) * (2 ** np.arange(width, dtype=object))).sum(axis=1)]


def _check_one(r, positions, n_bits):
    """One select support: header, superblock table, then per superblock either all positions or one offset per 64."""
    cnt = r.take("Q")
    logn, logn2, logn4 = r.take("B"), r.take("B"), r.take("B")
    assert cnt == len(positions)
    assert logn == max(n_bits, 1).bit_length() and logn2 == (logn * logn) & 0xFF and logn4 == (logn2 * logn2) & 0xFF
    if cnt == 0:
        return
    sb = (cnt + 4095) >> 12
    superblock = r.int_vector()
    assert superblock == [positions[k << 12] for k in range(sb)]
    member = set(positions)
    for k in range(sb):
        blk = positions[k << 12:(k + 1) << 12]
        flag = r.take("B")
        table = r.int_vector()
        if flag == 0:                                   # long: every position spelled out
            assert len(table) == 4096 and table[:len(blk)] == blk
        else:                                           # miniblocks: offset of every 64th occurrence, then a scan
            assert len(table) == 64
            for i in range(0, len(blk), 61):            # select of a sample of occurrences
                p = superblock[k] + table[i >> 6]
                left = i & 63
                while left:
                    p += 1
                    if p in member:
                        left -= 1
                assert p == blk[i]
        span = blk[-1] - blk[0]
        assert (flag == 0) == (span > max(logn4, 1))   # sdsl: a superblock longer than log^4(n) bits is stored "long"


@pytest.mark.parametrize("n_bits,density,seed", [(1, 1.0, 1), (64, 0.5, 2), (65, 0.5, 3), (5000, 0.9, 4), (300_000, 0.5, 5), (300_000, 0.02, 6),
                                                 (4096 * 3, 1.0, 7), (2_000_000, 0.004, 8), (70_000, 0.0, 9), (8192 + 17, 0.999, 10)])
def test_select_tables_answer_select_queries(dumper, n_bits, density, seed):
    exe, d = dumper
    rng = np.random.default_rng(seed)
    bits = (rng.random(n_bits) < density).astype(np.uint8)
    words = np.packbits(np.concatenate([bits, np.zeros((-n_bits) % 64, np.uint8)]), bitorder="little").view(np.uint64)
    out = str(d / f"s_{n_bits}_{seed}.bin")
    text = f"{n_bits}\n" + "\n".join(f"{int(w):x}" for w in words) + "\n"
    subprocess.run([exe, out], input=text.encode(), check=True)
    r = _Reader(open(out, "rb").read())
    _check_one(r, [int(x) for x in np.flatnonzero(bits == 1)], n_bits)
    _check_one(r, [int(x) for x in np.flatnonzero(bits == 0)], n_bits)      # the padding of the last word must not count as zeros
    assert r.o == len(r.b)
