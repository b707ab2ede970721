#!/usr/bin/env python3
"""Single-process probe of the row-range partitioned filter on two GPUs (for ncu's NVLink counters): part 0 on device 0,
part 1 on device 1, one batch counted on part 0 — half of its row requests travel over NVLink.
  python tools/partition_probe.py [reads] [read_len] [ref_bp]"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tests import common

n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
L = int(sys.argv[2]) if len(sys.argv) > 2 else 5000
bp = int(sys.argv[3]) if len(sys.argv) > 3 else 20_000_000
idx, refs = common.build_oracle_index([bp, bp], [0, 1], ["host", "microbial"], seed=5)
parts = [common.gpu_from_oracle(idx, device=r, part=(r, 2)) for r in range(2)]
parts[0].attach_peer(1, parts[1]); parts[1].attach_peer(0, parts[0])
whole = common.gpu_from_oracle(idx, device=0)
bases, seg_off, _ = common.fixed_length_batch(3, refs, [0.45, 0.45], n, L)
want = whole.count_batch(bases, seg_off)
for rep in range(3):
    t0 = time.time()
    got = parts[0].count_batch(bases, seg_off)
    dt = time.time() - t0
assert all(np.array_equal(got[k], want[k]) for k in want)
i = parts[0].info()
print(f"partitioned == replicated on {n} reads; count kernel {i.last_count_kernel_ms:.3f} ms partitioned vs {whole.info().last_count_kernel_ms:.3f} ms replicated, "
      f"{int(got['num_hashes'].sum())} minimisers, filter part {i.filter_bytes} B")
