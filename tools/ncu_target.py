#!/usr/bin/env python3
"""tools/ncu_target.py [workload] [launches] — the smallest program that runs whole-batch launches of the count kernel of a
bench.py workload (index built on the GPU, one resident 2-bit batch, cg_run_staged): the command ncu profiles.
  python tools/ncu_target.py human 3 && ncu --set full --clock-control none --import-source on -k regex:count_kernel_v3 -s 1 -c 1 -o prof python tools/ncu_target.py human 3
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    import charon_b200
    from charon_b200 import api as A
    name = sys.argv[1] if len(sys.argv) > 1 else "human"
    launches = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    wl = bench.WORKLOADS[name]
    R, L = wl["reads_per_step"], wl["read_len"]
    ncat = len(bench.categories(wl))
    records, rec_bin = bench.make_refs(wl)
    gpu = charon_b200.GpuClassifier.from_references(records, rec_bin, bench.bin_categories(wl), ncat, device=0)
    gpu.set_model(charon_b200.DehostOptions() if wl["mode"] == "dehost" else charon_b200.ClassifyOptions(), host_index=0)
    bases, seg_off, _ = bench.make_reads(wl, records, rec_bin, R, first_index=0)
    pb = A.pack_batch(bases, seg_off, A.BASES_2BIT, align=64)
    gpu.stage(pb, None, np.full(R, 20.0, np.float32), np.full(R, 0.3, np.float32), np.full(R, L, np.uint32))
    for _ in range(launches):
        gpu.run_staged(True)
    gpu.sync()
    res = gpu.fetch_staged(True)
    print(name, "reads", R, "minimisers", int(res["num_hashes"].astype(np.int64).sum()), "count kernel ms", gpu.info().last_count_kernel_ms)
    gpu.close()


if __name__ == "__main__":
    main()
