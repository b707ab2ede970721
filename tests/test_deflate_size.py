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
    for _ in range(n_random):
        n = rng.choice([30, 200, 900, 2500, 6000, 8000, 15000])
        if rng.random() < 0.7:
            seqs.append("".join(rng.choice("ACGT") if rng.random() > 0.01 else "N" for _ in range(n)))
        else:
            seqs.append("".join(rng.choice("AC" if rng.random() < 0.9 else "GT") for _ in range(n)))        # skewed: dynamic trees win
    return seqs


LONG = [33000, 40000, 65273]       # more than 16383 symbols: several blocks; 65273 = the longest input without a window slide


@pytest.fixture(scope="module")
def host_exe(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("dsz") / "deflate_size_host")
    gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    subprocess.check_call([gxx, "-O2", "-std=c++17", os.path.join(ROOT, "tests", "deflate_size_host.cpp"), "-o", exe])
    return exe


def test_restated_deflate_size_equals_zlib_on_the_host(host_exe):
    rng = random.Random(5)
    seqs = sequences(1) + ["".join(rng.choice("ACGT") for _ in range(n)) for n in LONG]
    out = subprocess.run([host_exe], input=("\n".join(seqs) + "\n").encode(), capture_output=True, check=True).stdout.decode().split()
    assert len(out) == len(seqs)
    bad = [(len(s), int(o), gzip_size(s.encode())) for s, o in zip(seqs, out) if int(o) != gzip_size(s.encode())]
    assert not bad, f"{len(bad)} of {len(seqs)} sizes differ from zlib {zlib.ZLIB_RUNTIME_VERSION}; first (length, got, zlib): {bad[:5]}"
    assert gzip_size(b"") == 20 and int(out[0]) == 20
