#!/usr/bin/env python3
"""tools/e2e_probe.py — why does the end-to-end arm lose a third against the resident arm?  (human-scale workload)

  T1  kernels alone on a resident batch (cg_run_staged loop)
  T2  the same loop while an UNRELATED 0.27 GB pinned->device copy runs again and again on another stream (torch)
  T3  cg_submit_batch / cg_wait_batch on two slots (what bench.py's e2e arm times)
  T4  the copy loop of T2 alone

Prints one JSON line.  Environment: CHARON_ASYNC_GATHER=0|1 picks the gather carrier of the count kernel.
"""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    import torch
    import charon_b200
    from charon_b200 import api as A
    from charon_b200.api import pinned_array
    name = sys.argv[1] if len(sys.argv) > 1 else "human"
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    wl = bench.WORKLOADS[name]
    R, L = wl["reads_per_step"], wl["read_len"]
    ncat = len(bench.categories(wl))
    records, rec_bin = bench.make_refs(wl)
    gpu = charon_b200.GpuClassifier.from_references(records, rec_bin, bench.bin_categories(wl), ncat, device=0)
    gpu.set_model(charon_b200.DehostOptions() if wl["mode"] == "dehost" else charon_b200.ClassifyOptions(), host_index=0)
    bases = pinned_array((R * L,), np.uint8)
    _, seg_off, _ = bench.make_reads(wl, records, rec_bin, R, first_index=0, out=bases)
    Lp = (L + 63) // 64 * 64
    p2 = pinned_array((R * Lp // 4 + 64,), np.uint8)
    pb = A.pack_batch(bases, seg_off, A.BASES_2BIT, align=64, out=p2)
    p2_off = pinned_array((R + 1,), np.uint64); p2_off[:] = pb.seg_off
    p2_len = pinned_array((R,), np.uint64); p2_len[:] = pb.seg_len
    mq = pinned_array((R,), np.float32); mq[:] = 20.0
    cr = pinned_array((R,), np.float32); cr[:] = 0.3
    ln = pinned_array((R,), np.uint32); ln[:] = L
    assert len(pb.exc_pos) == 0

    def outputs():
        o = dict(num_hashes=pinned_array((R,), np.uint32), counts=pinned_array((R, ncat), np.uint32), unique_counts=pinned_array((R, ncat), np.uint32),
                 prob=pinned_array((R, ncat), np.float64), call=pinned_array((R,), np.uint8), confidence=pinned_array((R,), np.uint8))
        return o, A.CountsOut(o["num_hashes"].ctypes.data, o["counts"].ctypes.data, o["unique_counts"].ctypes.data), \
            A.CallsOut(o["prob"].ctypes.data, o["call"].ctypes.data, o["confidence"].ctypes.data)

    slots = [outputs(), outputs()]
    batch = A.Batch(p2.ctypes.data, p2_off.ctypes.data, None, R, R, A.BASES_2BIT, p2_len.ctypes.data, None, 0)
    sc = A.ReadScalars(mq.ctypes.data, cr.ctypes.data, ln.ctypes.data)
    Lb = gpu._L
    out = {"workload": name, "steps": steps, "carrier_env": os.environ.get("CHARON_ASYNC_GATHER", "auto")}

    gpu._check(Lb.cg_stage_batch(gpu._h, C.byref(batch), C.byref(sc)), "cg_stage_batch")
    gpu._staged_n = R

    def kernel_loop(n):
        gpu.sync()
        gpu.timer_start()
        for _ in range(n):
            gpu.run_staged(True)
        ms, cms, ems, _ = gpu.timer_stop()
        return ms / n, cms / n

    kernel_loop(3)
    out["T1_kernels_alone_ms"], out["T1_count_ms"] = kernel_loop(steps)

    # an unrelated copy stream
    nbytes = R * Lp // 4
    src = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    dst = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    side = torch.cuda.Stream()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def copies(n):
        with torch.cuda.stream(side):
            e0.record()
            for _ in range(n):
                dst.copy_(src, non_blocking=True)
            e1.record()

    copies(2); torch.cuda.synchronize()
    copies(steps); torch.cuda.synchronize()
    out["T4_copy_alone_ms"] = e0.elapsed_time(e1) / steps
    copies(steps + 2)
    out["T2_kernels_next_to_copies_ms"], out["T2_count_ms"] = kernel_loop(steps)
    torch.cuda.synchronize()
    out["T2_copy_ms"] = e0.elapsed_time(e1) / (steps + 2)

    def e2e_steps(n):
        for i in range(n):
            sl = i & 1
            if i >= 2:
                gpu._check(Lb.cg_wait_batch(gpu._h, sl), "cg_wait_batch")
            ts = time.time()
            gpu._check(Lb.cg_submit_batch(gpu._h, sl, C.byref(batch), C.byref(sc), C.byref(slots[sl][1]), C.byref(slots[sl][2])), "cg_submit_batch")
            if os.environ.get("CHARON_TRACE_SLOTS"):
                sys.stderr.write(f"  submit(slot {sl}) took {(time.time() - ts) * 1e3:.3f} ms of host time\n")
        gpu._check(Lb.cg_wait_batch(gpu._h, 0), "cg_wait_batch")
        gpu._check(Lb.cg_wait_batch(gpu._h, 1), "cg_wait_batch")

    e2e_steps(3)
    gpu.sync()
    t0 = time.time()
    gpu.timer_start()
    e2e_steps(steps)
    ms, _, _, _ = gpu.timer_stop()
    out["T3_e2e_ms"] = ms / steps
    out["T3_e2e_wall_ms"] = (time.time() - t0) * 1e3 / steps
    info = gpu.info()
    out["T3_last"] = {"h2d": info.last_h2d_ms, "count": info.last_count_kernel_ms, "eval": info.last_eval_kernel_ms}
    print(json.dumps(out))
    gpu.close()


if __name__ == "__main__":
    main()
