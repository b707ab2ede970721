#!/usr/bin/env python3
"""tools/gz_probe.py [reads=131072] [len=8000] — time of the compression-ratio kernel on one batch, next to zlib on one host core."""
import os, sys, time, zlib
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import common
import charon_b200
from charon_b200 import api as A

n = int(sys.argv[1]) if len(sys.argv) > 1 else 131072
L = int(sys.argv[2]) if len(sys.argv) > 2 else 8000
refs = [common.synth.reference(1, 5_000_000), common.synth.reference(2, 5_000_000)]
gpu = charon_b200.GpuClassifier.from_references(refs, [0, 1], [0, 1], 2, device=0)
bases, seg_off, _ = common.fixed_length_batch(3, refs, [0.45, 0.45], n, L)
pb = A.pack_batch(bases, seg_off, A.BASES_2BIT, align=64)
for rep in range(3):
    got = gpu.compression_batch(pb, None)
    print(f"gpu: {n} reads x {L}: gzip kernel {gpu.info().last_gzip_kernel_ms:.2f} ms  ({n / gpu.info().last_gzip_kernel_ms / 1e3:.3f} M reads/s)", flush=True)
m = 400
t0 = time.time()
want = []
for i in range(m):
    c = zlib.compressobj(-1, zlib.DEFLATED, 31, 8, 0)
    want.append(len(c.compress(bytes(bases[i * L:(i + 1) * L])) + c.flush()))
dt = time.time() - t0
print(f"zlib, one core: {dt / m * 1e6:.0f} us per read ({m / dt:.0f} reads/s); sizes equal: {np.array_equal(got['size'][:m], np.array(want, np.uint32))}; mean ratio {got['ratio'].mean():.4f}")
gpu.close()
