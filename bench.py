#!/usr/bin/env python3
"""bench.py — reads/s of Charon's read-classification hot path on B200 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload bacterial|smoke|human] [--impl b200|reference]

One "step" = one pass of the hot path (minimiser hashing -> IBF membership -> per-category counting -> KDE model ->
call) over one batch of synthetic reads.  Default workload = BASELINE.json configs[1] ("bacterial-scale: 2 categories
~200 Mbp each, 8 kb reads, index replicated on 1xB200"); 8 steps x 131072 reads = the config's 1M reads.
N > 1 (under torchrun): one rank per GPU, filter replicated, reads sharded, NO data-path collective; torch.distributed
is used only for the barrier and the max-over-ranks of the device-measured time ("scaling": "weak").

value : whole-job reads/s with the batch already resident in HBM (CUDA events on the launching stream).
e2e   : the same through the C-ABI calls cg_submit_batch / cg_wait_batch with pinned ASCII HOST buffers (H2D + kernels + D2H
        inside the timed region; with >= 8 host threads per GPU the library packs the bases to 4 bits on the way, also inside it).
--impl reference : the reference's CPU implementation of the path = the oracle port (the real charon cannot be built
        here: its CMake downloads seqan3 & co.), all host threads, on a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (ref bp per category [host, microbial], record bp, read length, reads per step, description)
    "smoke": dict(ref_bp=[5_000_000, 5_000_000], rec_bp=5_000_000, read_len=5000, reads_per_step=10_000,
                  desc="config0: 2 categories x 5 Mbp, 5 kb reads"),
    "bacterial": dict(ref_bp=[200_000_000, 200_000_000], rec_bp=5_000_000, read_len=8000, reads_per_step=131_072,
                      desc="config1 bacterial-scale: 2 categories x 200 Mbp, 8 kb reads, filter replicated per GPU"),
    "human": dict(ref_bp=[3_100_000_000, 1_500_000_000], rec_bp=50_000_000, read_len=8000, reads_per_step=131_072,
                  desc="config2 human-scale: 3.1 Gbp host + 1.5 Gbp microbial, 8 kb reads, filter replicated per GPU"),
}
CATEGORIES = ["host", "microbial"]
ALGO_BYTES_PER_MINIMISER = 96   # 3 hash functions x one 32 B sector
ALGO_BYTES_PER_BASE = 1


def make_refs(wl):
    """Seeded iid ACGT references, cut into chromosome-like records (all records of a category share one bin)."""
    from tools import synth
    records, rec_bin = [], []
    for b, bp in enumerate(wl["ref_bp"]):
        full = synth.reference(1000 + b, bp)
        for s in range(0, bp, wl["rec_bp"]):
            records.append(full[s:s + wl["rec_bp"]])
            rec_bin.append(b)
    return records, rec_bin


def make_reads(wl, records, rec_bin, n_reads, first_index, out=None):
    from tools import synth
    per_bin = [np.concatenate([r for r, b in zip(records, rec_bin) if b == bb]) for bb in range(len(wl["ref_bp"]))]
    bases, src = synth.reads(42, per_bin, [0.45, 0.45], n_reads, wl["read_len"], 0.02, first_index=first_index, out=out)
    seg_off = np.arange(n_reads + 1, dtype=np.uint64) * np.uint64(wl["read_len"])
    return bases, seg_off, src


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index):
        self.rows, self.proc, self.idx = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.idx}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for t, r in self.rows if t0 <= t <= t1 + 0.2 and len(r) >= 8] or [r for _, r in self.rows if len(r) >= 8]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm = sorted(float(r[1]) for r in rows)
        reasons = [n for i, n in ((4, "hw_slowdown"), (5, "hw_thermal_slowdown"), (6, "sw_thermal_slowdown"), (7, "sw_power_cap"))
                   if any(r[i].lower().startswith("active") for r in rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][2]), "power_w_max": max(float(r[3]) for r in rows),
                "samples": len(rows), "reasons": reasons}


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def cpu_baseline(wl, records, rec_bin, seconds_target=15.0, threads=0):
    """The oracle port (reference loop body restated) on the host cores, bounded sample of the same workload."""
    import oracle
    # every core this process may run on: torchrun exports OMP_NUM_THREADS=1 to its children, which must not shrink the CPU arm
    try:
        host_cores = len(os.sched_getaffinity(0))
    except AttributeError:
        host_cores = os.cpu_count() or 1
    threads = threads or max(host_cores, oracle.lib().orc_max_threads())
    t0 = time.time()
    idx = oracle.Index.build(records, rec_bin, [0, 1], CATEGORIES, threads=threads)
    build_s = time.time() - t0
    probe = 256 * max(1, threads // 8)
    bases, seg_off, _ = make_reads(wl, records, rec_bin, probe, first_index=10_000_000)
    r = idx.classify_batch(bases, seg_off, threads=threads)
    rate = probe / max(r["seconds"], 1e-6)
    n = int(min(max(rate * seconds_target, probe), 200_000))
    bases, seg_off, _ = make_reads(wl, records, rec_bin, n, first_index=20_000_000)
    r = idx.classify_batch(bases, seg_off, threads=threads)
    hot = n / r["seconds"]
    n_gz = max(probe, n // 4)
    rg = idx.classify_batch(bases[:n_gz * wl["read_len"]], seg_off[:n_gz + 1], with_gzip=True, threads=threads)
    return idx, dict(value=hot, unit="reads/s", cores=threads, kind="port",
                     sample=f"{n} reads x {wl['read_len']} b of the same workload, hot path only (minimiser+IBF+count+model+call, plain "
                            f"bit-matrix filter = upper bound on the real sd_vector reference); with the per-read gzip ratio of the "
                            f"reference loop: {n_gz / rg['seconds']:.1f} reads/s on {n_gz} reads",
                     with_gzip_value=n_gz / rg["seconds"], index_build_s=round(build_s, 2), seconds=round(r["seconds"], 2)), (bases, seg_off, r)


def parity_mismatch(g, o):
    """None when every integer output agrees, else a description of the first differing read."""
    for key in ("num_hashes", "counts", "unique_counts", "confidence", "call"):
        a, b = np.asarray(g[key]), np.asarray(o[key])
        if not np.array_equal(a, b):
            r = int(np.argwhere(a != b)[0][0])
            return (f"first difference in `{key}` at read {r}: gpu num_hashes={int(g['num_hashes'][r])} counts={g['counts'][r].tolist()} "
                    f"unique={g['unique_counts'][r].tolist()} call={int(g['call'][r])} | oracle num_hashes={int(o['num_hashes'][r])} "
                    f"counts={o['counts'][r].tolist()} unique={o['unique_counts'][r].tolist()} call={int(o['call'][r])}; "
                    f"{int((np.asarray(g['num_hashes']) != np.asarray(o['num_hashes'])).sum())} reads differ in num_hashes")
    return None


def run_reference(args, wl):
    rank, local_rank, world = dist_env()
    if rank != 0:
        return
    records, rec_bin = make_refs(wl)
    per_step_s = []
    idx, cb, _ = cpu_baseline(wl, records, rec_bin, seconds_target=max(3.0, 40.0 / max(1, args.steps + args.warmup)))
    # K timed steps, each a bounded sample sized from the probe
    n = int(max(256, min(cb["value"] * 4.0, 100_000)))
    bases, seg_off, _ = make_reads(wl, records, rec_bin, n, first_index=30_000_000)
    for i in range(args.warmup + args.steps):
        r = idx.classify_batch(bases, seg_off, threads=cb["cores"])
        if i >= args.warmup:
            per_step_s.append(r["seconds"])
    total = sum(per_step_s)
    val = n * len(per_step_s) / total
    line = {"impl": "reference", "metric": "reads_per_sec", "value": val, "unit": "reads/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1000 * total / len(per_step_s), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": {"workload": wl["desc"], "reads_per_step": n, "read_len": wl["read_len"], "k": 19, "w": 41, "note": "CPU oracle port, bounded sample per step"},
            "cpu_baseline": {"value": val, "unit": "reads/s", "cores": cb["cores"], "kind": "port",
                             "sample": f"{n} reads/step x {args.steps} steps; hot path only; with gzip ratio: {cb['with_gzip_value']:.1f} reads/s"},
            "gbases_per_sec": val * wl["read_len"] / 1e9,
            "e2e": {"value": val, "unit": "reads/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def run_b200(args, wl):
    rank, local_rank, world = dist_env()
    import charon_b200
    from charon_b200.api import pinned_array
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    R, L = args.reads_per_step or wl["reads_per_step"], wl["read_len"]
    records, rec_bin = make_refs(wl)
    t0 = time.time()
    gpu = charon_b200.GpuClassifier.from_references(records, rec_bin, [0, 1], 2, device=local_rank)   # index built on the GPU
    build_s = time.time() - t0
    partitioned = bool(args.partitioned and world > 1)
    if partitioned:
        # BASELINE config 4: the filter is split by row range over the ranks; rows owned by a peer are fetched through its
        # NVLink peer mapping inside the count kernel.  Reads stay sharded, still no data-path collective.
        import torch
        from charon_b200 import multi
        gpu.partition(rank, world)
        multi.exchange_partitions(gpu, dist, rank, world, device=torch.device("cuda", local_rank))
    gpu.set_model(charon_b200.DehostOptions(), host_index=0)
    # transport of the bases over PCIe: the library packs them to 4-bit ranks on host threads when it has at least 8 of them
    # (cg_set_host_pack); the ranks of one box share its cores
    try:
        host_cores = len(os.sched_getaffinity(0))
    except AttributeError:
        host_cores = os.cpu_count() or 1
    pack_threads = max(1, min(16, host_cores // world))
    pack_mode = {"auto": -1, "off": 0, "on": 1}[args.host_pack]
    gpu.set_host_pack(pack_mode, pack_threads)
    # this rank's batch (weak scaling: R reads per GPU per step), in pinned host memory
    bases = pinned_array((R * L,), np.uint8)
    _, seg_off, _ = make_reads(wl, records, rec_bin, R, first_index=rank * R, out=bases)
    seg_pin = pinned_array((R + 1,), np.uint64); seg_pin[:] = seg_off
    mq = pinned_array((R,), np.float32); mq[:] = 20.0
    cr = pinned_array((R,), np.float32); cr[:] = 0.3
    ln = pinned_array((R,), np.uint32); ln[:] = L
    out = dict(num_hashes=pinned_array((R,), np.uint32), counts=pinned_array((R, 2), np.uint32), unique_counts=pinned_array((R, 2), np.uint32),
               prob=pinned_array((R, 2), np.float64), call=pinned_array((R,), np.uint8), confidence=pinned_array((R,), np.uint8))
    from charon_b200 import api as A
    import ctypes as C
    batch = A.Batch(bases.ctypes.data, seg_pin.ctypes.data, None, R, R)
    sc = A.ReadScalars(mq.ctypes.data, cr.ctypes.data, ln.ctypes.data)
    co = A.CountsOut(out["num_hashes"].ctypes.data, out["counts"].ctypes.data, out["unique_counts"].ctypes.data)
    ca = A.CallsOut(out["prob"].ctypes.data, out["call"].ctypes.data, out["confidence"].ctypes.data)
    Lb = gpu._L

    def barrier():
        gpu.sync()
        if dist is not None:
            dist.barrier()

    def max_over_ranks(x):
        if dist is None:
            return x
        import torch
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- kernel-only arm: inputs resident in HBM --------------------------------------------------------------
    gpu._check(Lb.cg_stage_batch(gpu._h, C.byref(batch), C.byref(sc)), "cg_stage_batch")
    gpu._staged_n = R
    for _ in range(args.warmup):
        gpu.run_staged(True)
    barrier()
    sampler = ClockSampler(local_rank); sampler.start(); time.sleep(0.25)
    w0 = time.time()
    gpu.timer_start()
    for _ in range(args.steps):
        gpu.run_staged(True)
    total_ms, count_ms, eval_ms, n_runs = gpu.timer_stop()
    w1 = time.time()
    barrier()
    clocks = sampler.stop(w0, w1)
    res = gpu.fetch_staged(True)
    resident_packed = bool(gpu.info().last_packed)
    minimisers = int(res["num_hashes"].astype(np.int64).sum())
    total_ms_max = max_over_ranks(total_ms)
    # the same kernels on the other resident format (ASCII, one byte per base), for comparison only
    other_value = None
    if resident_packed:
        gpu.set_host_pack(0, 0)
        gpu._check(Lb.cg_stage_batch(gpu._h, C.byref(batch), C.byref(sc)), "cg_stage_batch")
        for _ in range(args.warmup):
            gpu.run_staged(True)
        barrier()
        gpu.timer_start()
        for _ in range(args.steps):
            gpu.run_staged(True)
        o_ms, o_count_ms, _, o_runs = gpu.timer_stop()
        barrier()
        res_o = gpu.fetch_staged(True)
        assert np.array_equal(res_o["num_hashes"], res["num_hashes"]) and np.array_equal(res_o["counts"], res["counts"]) and np.array_equal(res_o["call"], res["call"])
        other_value = {"reads_per_sec": R * args.steps * world / (max_over_ranks(o_ms) / 1e3), "count_kernel_ms": o_count_ms / max(o_runs, 1)}
        gpu.set_host_pack(pack_mode, 0)
    # ---- end-to-end arm: host buffers through the C-ABI (H2D + kernels + D2H of every step inside the timed region) -------
    # cg_submit_batch / cg_wait_batch on two slots: the PCIe copy of step i+1 overlaps the kernels of step i
    out2 = dict(num_hashes=pinned_array((R,), np.uint32), counts=pinned_array((R, 2), np.uint32), unique_counts=pinned_array((R, 2), np.uint32),
                prob=pinned_array((R, 2), np.float64), call=pinned_array((R,), np.uint8), confidence=pinned_array((R,), np.uint8))
    co2 = A.CountsOut(out2["num_hashes"].ctypes.data, out2["counts"].ctypes.data, out2["unique_counts"].ctypes.data)
    ca2 = A.CallsOut(out2["prob"].ctypes.data, out2["call"].ctypes.data, out2["confidence"].ctypes.data)
    slots = [(co, ca), (co2, ca2)]

    def e2e_steps(n):
        for i in range(n):
            sl = i & 1
            if i >= 2:
                gpu._check(Lb.cg_wait_batch(gpu._h, sl), "cg_wait_batch")
            gpu._check(Lb.cg_submit_batch(gpu._h, sl, C.byref(batch), C.byref(sc), C.byref(slots[sl][0]), C.byref(slots[sl][1])), "cg_submit_batch")
        gpu._check(Lb.cg_wait_batch(gpu._h, 0), "cg_wait_batch")
        gpu._check(Lb.cg_wait_batch(gpu._h, 1), "cg_wait_batch")

    e2e_steps(max(2, min(args.warmup, 3)))
    barrier()
    gpu.timer_start()
    e2e_steps(args.steps)
    e2e_ms, _, _, _ = gpu.timer_stop()
    barrier()
    e2e_ms_max = max_over_ranks(e2e_ms)
    e2e_info = gpu.info()
    assert np.array_equal(out["num_hashes"], res["num_hashes"]) and np.array_equal(out["call"], res["call"])
    assert args.steps < 2 or (np.array_equal(out2["num_hashes"], res["num_hashes"]) and np.array_equal(out2["call"], res["call"]))
    # ---- serial variant of the same call (cg_classify_batch: copy, kernels, copy back, one after the other) -----------------
    gpu._check(Lb.cg_classify_batch(gpu._h, C.byref(batch), C.byref(sc), C.byref(co), C.byref(ca)), "cg_classify_batch")
    barrier()
    gpu.timer_start()
    for _ in range(args.steps):
        gpu._check(Lb.cg_classify_batch(gpu._h, C.byref(batch), C.byref(sc), C.byref(co), C.byref(ca)), "cg_classify_batch")
    serial_ms, _, _, _ = gpu.timer_stop()
    barrier()
    serial_ms_max = max_over_ranks(serial_ms)
    # ---- what the memory system gives a kernel that ONLY gathers, on this very filter -------------------------
    g_query_ms = gpu.bench_gather(minimisers, mode=0, reps=5)
    g_sector_ms = gpu.bench_gather(3 * minimisers, mode=1, reps=5)
    info = gpu.info()
    if rank != 0:
        return
    reads_total = R * args.steps * world
    value = reads_total / (total_ms_max / 1e3)
    e2e = reads_total / (e2e_ms_max / 1e3)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    algo_bytes = ALGO_BYTES_PER_MINIMISER * minimisers + ALGO_BYTES_PER_BASE * R * L
    avg_count_ms = count_ms / max(n_runs, 1)
    achieved = algo_bytes / (avg_count_ms / 1e3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "count_kernel_traffic.json")
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get(args.workload)
        except Exception:
            traffic = None
    line = {"metric": "reads_per_sec", "value": value, "unit": "reads/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64",
            "data": "synthetic",
            "config": {"workload": wl["desc"] + (" -- filter PARTITIONED by row range across the GPUs (NVLink peer gathers)" if partitioned else ""),
                       "filter_partitioned": partitioned, "reads_per_step_per_gpu": R, "read_len": L, "k": 19, "w": 41, "hash_funs": 3,
                       "filter_rows": int(info.bin_size), "filter_bytes_hbm": int(info.filter_bytes), "row_bits": int(info.row_bits),
                       "minimisers_per_step_per_gpu": minimisers, "classified_frac": float((res["call"] != 255).mean()),
                       "l2_note": "step input (0.5 or 1 B/base x reads = 0.5 / 1 GB) is larger than the 126 MB L2 and is re-read every step; no flush needed",
                       "hbm_resident_bases": "4-bit dna5 ranks (the library's transport format)" if resident_packed else "ASCII, 1 byte per base",
                       "value_on_ascii_resident_bases": other_value,
                       "index_build_on_gpu_s": round(build_s, 2)},
            "gbases_per_sec": value * L / 1e9,
            "e2e": {"value": e2e, "unit": "reads/s", "h2d_bytes_per_step": int(e2e_info.last_h2d_bytes) * world,
                    "transport": "ASCII host buffers packed to 4-bit ranks inside the call, chunk copies overlapped" if e2e_info.last_packed else "ASCII copied as is",
                    "host_pack_threads_per_gpu": pack_threads if e2e_info.last_packed else 0, "host_pack_ms_per_step": float(e2e_info.last_pack_ms),
                    "d2h_bytes_per_step": int(R * (4 + 8 + 8 + 16 + 2)) * world, "gbases_per_sec": e2e * L / 1e9,
                    "api": "cg_submit_batch/cg_wait_batch, two slots (copy of step i+1 overlaps kernels of step i)",
                    "serial_cg_classify_batch_value": reads_total / (serial_ms_max / 1e3)},
            "gpu_launches": int(3 * args.steps),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": "cgk::count_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs (burst copy)" if peaks else "fallback 6650",
                         "algorithmic_bytes_per_launch": algo_bytes, "avg_launch_ms": avg_count_ms,
                         "count_kernel_share_of_step": count_ms / total_ms,
                         "gather_microbench": {
                             "what": "same filter, same 3-row addressing, nothing else in the kernel (cg_bench_gather mode 0) / one random 32 B sector per load (mode 1)",
                             "query_addressing_GBps": ALGO_BYTES_PER_MINIMISER * minimisers / (g_query_ms / 1e3) / 1e9,
                             "pure_sector_GBps": 32 * 3 * minimisers / (g_sector_ms / 1e3) / 1e9,
                             "count_kernel_gather_GBps": ALGO_BYTES_PER_MINIMISER * minimisers / (avg_count_ms / 1e3) / 1e9,
                             "frac_of_achievable_gather": (g_query_ms / avg_count_ms)}, "eval_kernels_ms_per_step": eval_ms / max(n_runs, 1),
                         "note": "96 B/minimiser + 1 B/base (SURVEY 8d) regardless of L2 hits; compacted filter of this workload is "
                                 f"{info.filter_bytes / 1e6:.0f} MB"}}
    if world == 1 and not args.no_cpu_baseline:
        idx, cb, (sb, so, sr) = cpu_baseline(wl, records, rec_bin)
        # parity GATE at full index size: the oracle's own index + sample vs the GPU, every integer output of every read
        n = len(so) - 1
        g = gpu.classify_batch(sb, so, np.full(n, 20.0, np.float32), np.full(n, 0.3, np.float32), np.full(n, L, np.uint32))
        bad = parity_mismatch(g, sr)
        cb["parity_on_sample"] = bad is None
        cb["parity_sample_reads"] = n
        line["cpu_baseline"] = cb
        if bad is not None:
            sys.stderr.write(f"bench.py: PARITY FAILURE against the oracle on the cpu_baseline sample ({n} reads): {bad}\n"
                             "bench.py: no bench line is printed for a kernel that is not bit-exact\n")
            sys.stderr.flush()
            gpu.close()
            sys.exit(3)
    emit(line)
    gpu.close()
    if dist is not None:
        dist.destroy_process_group()


_REAL_STDOUT = None


def emit(line: dict):
    """The ONE JSON line goes to the process's original stdout; everything else (NCCL banners, library chatter) was
    re-routed to stderr in main()."""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="bacterial", choices=sorted(WORKLOADS))
    ap.add_argument("--reads-per-step", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--host-pack", default="auto", choices=["auto", "on", "off"], help="transport of the bases over PCIe: 4-bit ranks packed on host threads (auto: when >= 8 threads per GPU) or ASCII")
    ap.add_argument("--partitioned", action="store_true", help="N>1: split the filter by row range across the GPUs instead of replicating it")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    wl = WORKLOADS[args.workload]
    rank, local_rank, world = dist_env()
    if args.impl == "reference":
        run_reference(args, wl)
        return
    if world != args.gpus and world == 1 and args.gpus > 1:
        # launched without torchrun: re-exec under torch.distributed.run
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}", "--master-addr", "127.0.0.1",
               "--master-port", "29511"] + sys.argv
        os.dup2(_REAL_STDOUT, 1)   # the child sets up its own re-routing
        os.execv(sys.executable, cmd)
    run_b200(args, wl)


if __name__ == "__main__":
    main()
