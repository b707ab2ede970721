#!/usr/bin/env python3
"""bench.py — reads/s of Charon's read-classification hot path on B200 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload human|bacterial|many_category|partitioned|smoke]
                  [--impl b200|reference]

One "step" = one pass of the hot path (minimiser hashing -> IBF membership -> per-category counting -> model -> call)
over one batch of synthetic reads.  Default workload = BASELINE.json configs[2], the configuration the north-star
target is quoted on: human-scale index (3.1 Gbp host + 1.5 Gbp microbial, 0.86 GB compacted filter: fits one B200),
131 072 reads x 8 kb per GPU and step.  `--workload bacterial` = configs[1], `many_category` = configs[3] (255
categories, `classify`, 256-bit rows), `partitioned` = configs[4] (filter split across the GPUs, N >= 2).
N > 1 (under torchrun): one rank per GPU, filter replicated, reads sharded, NO data-path collective; torch.distributed
only carries the barrier and the max-over-ranks of the device-measured time ("scaling": "weak").

value : whole-job reads/s with the batch already resident in HBM (CUDA events on the launching stream).
e2e   : the same through the C-ABI calls cg_submit_batch / cg_wait_batch with pinned HOST buffers in the format the
        library's own parser emits (CG_BASES_2BIT: 2 bits per base + exception list): H2D of every step's bases,
        offsets and scalars + kernels + D2H of every step's results inside the timed region, two slots.  The same
        with ASCII host buffers (the library packs them to 4 bits inside the call) is reported beside it.
--impl reference : the reference's CPU implementation of the path = the oracle port (the real charon cannot be built
        here: its CMake downloads seqan3 & co.), all host threads, on a bounded sample of the same workload.
A parity miss against the oracle on the sample is FATAL: no line is printed and the exit code is 3.
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
    "smoke": dict(ref_bp=[5_000_000, 5_000_000], rec_bp=5_000_000, read_len=5000, reads_per_step=10_000, mode="dehost",
                  desc="config0: 2 categories x 5 Mbp, 5 kb reads"),
    "bacterial": dict(ref_bp=[200_000_000, 200_000_000], rec_bp=5_000_000, read_len=8000, reads_per_step=131_072, mode="dehost",
                      desc="config1 bacterial-scale: 2 categories x 200 Mbp, 8 kb reads, filter replicated per GPU"),
    "human": dict(ref_bp=[3_100_000_000, 1_500_000_000], rec_bp=50_000_000, read_len=8000, reads_per_step=131_072, mode="dehost",
                  desc="config2 human-scale: 3.1 Gbp host + 1.5 Gbp microbial, 8 kb reads, filter replicated per GPU"),
    # 255 reference files of 1..40 Mbp = 255 bins, `charon classify`.  254 categories, not 255: with exactly
    # 255 the reference cannot print an unclassified read (include/input_summary.hpp:57-62 tests `index > categories.size()`,
    # so category_name(255) runs into categories.at(255) and throws) — the last two files share a category.
    "many_category": dict(ref_bp=[1_000_000 * (1 + (i * 7) % 40) for i in range(255)], bin_cat=[min(i, 253) for i in range(255)],
                          rec_bp=50_000_000, read_len=8000, reads_per_step=65_536, mode="classify",
                          desc="config3 many-category: 255 reference files (1-40 Mbp) in 254 categories, 256-bit filter rows, classify, 8 kb reads"),
    "partitioned": dict(ref_bp=[3_100_000_000, 1_500_000_000], rec_bp=50_000_000, read_len=8000, reads_per_step=131_072, mode="dehost",
                        partitioned=True, desc="config4 partitioned: human-scale index, filter split by row range across the GPUs"),
}
ALGO_BYTES_PER_MINIMISER = 96   # 3 hash functions x one 32 B sector
ALGO_BYTES_PER_BASE = 1


def bin_categories(wl):
    """category index of every bin (= reference file)"""
    return list(wl.get("bin_cat") or range(len(wl["ref_bp"])))


def categories(wl):
    n = max(bin_categories(wl)) + 1
    return ["host", "microbial"] if n == 2 else [f"cat{i:03d}" for i in range(n)]


def shared_config(name, wl):
    """The `config` object — identical in both arms (the driver compares them)."""
    return {"workload": name, "description": wl["desc"], "mode": wl["mode"], "bins": len(wl["ref_bp"]), "categories": len(categories(wl)),
            "reference_bp_total": int(sum(wl["ref_bp"])), "read_len": wl["read_len"], "k": 19, "w": 41, "hash_funs": 3,
            "reads": "45/45/10 % host/microbial/random source (uniform over categories when > 2), 2 % sub + 2 % ins + 2 % del",
            "l2": "every step's input (>= 0.26 GB of bases per GPU) is larger than the 126 MB L2 and is re-read each step"}


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


def source_fracs(wl):
    nb = len(wl["ref_bp"])
    return [0.45, 0.45] if nb == 2 else [0.9 / nb] * nb


_PER_BIN = {}


def make_reads(wl, records, rec_bin, n_reads, first_index, out=None):
    from tools import synth
    nb = len(wl["ref_bp"])
    if len(records) == nb:
        per_bin = records
    else:
        if id(records) not in _PER_BIN:     # the records of a bin are slices of one array: join them once
            _PER_BIN.clear()
            _PER_BIN[id(records)] = [np.concatenate([r for r, b in zip(records, rec_bin) if b == bb]) for bb in range(nb)]
        per_bin = _PER_BIN[id(records)]
    bases, src = synth.reads(42, per_bin, source_fracs(wl), n_reads, wl["read_len"], 0.02, first_index=first_index, out=out)
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


def host_cores():
    # every core this process may run on: torchrun exports OMP_NUM_THREADS=1 to its children, which must not shrink the CPU arm
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def oracle_options(wl):
    import oracle
    return oracle.default_options(dehost=wl["mode"] == "dehost")


def oracle_index(wl, records, rec_bin, threads):
    import oracle
    nb = len(wl["ref_bp"])
    t0 = time.time()
    idx = oracle.Index.build(records, rec_bin, bin_categories(wl), categories(wl), threads=threads)
    return idx, time.time() - t0


def fastq_text(bases, n, L, first=0):
    q = b"5" * L    # phred 20
    return b"".join(b"@r%d synthetic\n" % (first + i) + bases[i * L:(i + 1) * L].tobytes() + b"\n+\n" + q + b"\n" for i in range(n))


def cpu_baseline(wl, records, rec_bin, seconds_target=15.0, threads=0, stage_split=True):
    """The oracle port (reference loop body restated) on the host cores, bounded sample of the same workload:
    all cores and one thread, hot path only; plus the reference program's own shape (parse / gzip ratio / hot path /
    model + print under one critical section, chunks of 100) with its stage split — BASELINE.md §3."""
    import oracle
    cores = host_cores()
    threads = threads or max(cores, oracle.lib().orc_max_threads())
    idx, build_s = oracle_index(wl, records, rec_bin, threads)
    opts = oracle_options(wl)
    L = wl["read_len"]
    probe = 256 * max(1, threads // 8)
    bases, seg_off, _ = make_reads(wl, records, rec_bin, probe, first_index=10_000_000)
    r = idx.classify_batch(bases, seg_off, threads=threads, opts=opts)
    rate = probe / max(r["seconds"], 1e-6)
    n = int(min(max(rate * seconds_target, probe), 200_000))
    bases, seg_off, _ = make_reads(wl, records, rec_bin, n, first_index=20_000_000)
    r = idx.classify_batch(bases, seg_off, threads=threads, opts=opts)
    hot = n / r["seconds"]
    out = dict(value=hot, unit="reads/s", cores=threads, kind="port",
               sample=f"{n} reads x {L} b of the same workload, hot path only (minimiser + IBF + count + model + call; plain bit-matrix "
                      f"filter = upper bound on the real sd_vector-backed reference), {threads} threads",
               index_build_s=round(build_s, 2), seconds=round(r["seconds"], 2))
    if stage_split:
        n1 = max(64, min(n, int(rate / threads * 3.0)))       # about 3 s on one thread
        r1 = idx.classify_batch(bases[:n1 * L], seg_off[:n1 + 1], threads=1, opts=opts)
        out["t1_value"] = n1 / r1["seconds"]
        out["t1_sample_reads"] = n1
        # the reference program's shape, with the per-read gzip ratio (src/utils.cpp:114-124) and its one critical section
        n_all = max(256, min(n, int(rate)))                   # gzip costs about 3x the hot path: ~4 s on all cores
        fq = fastq_text(bases, n_all, L)
        ra = idx.run_like_reference(fq, threads=threads, tsv_path="/dev/shm/charon_bench_ref.tsv" if os.path.isdir("/dev/shm") else None, opts=opts)
        n_one = max(32, min(n_all, n1 // 3))
        ro = idx.run_like_reference(fastq_text(bases, n_one, L), threads=1, tsv_path=None, opts=opts)
        for tag, rr in (("program_all_cores", ra), ("program_t1", ro)):
            rr["reads_per_s"] = rr["reads"] / max(rr["parse_s"] + rr["loop_wall_s"], 1e-9)
            out[tag] = {k: (round(v, 4) if isinstance(v, float) else v) for k, v in rr.items()}
        out["with_gzip_value"] = ra["reads_per_s"]
        out["sample"] += (f"; the reference program's shape (FASTQ parse, chunks of 100, gzip ratio per read, one critical section for model + "
                          f"print) on {n_all} reads: {ra['reads_per_s']:.0f} reads/s on {threads} threads, {ro['reads_per_s']:.0f} on one")
        try:
            os.unlink("/dev/shm/charon_bench_ref.tsv")
        except OSError:
            pass
    return idx, out, (bases, seg_off, r)


def cli_tsv_check(n_reads=10_000):
    """the product BINARY (`charon_b200/charon dehost`, its own FASTQ parser, scalars, packing,
    pipeline and TSV printer) against the oracle's restatement of the reference program at -t 1, byte for byte, on the smoke
    configuration (2 x 5 Mbp index written by the oracle, 10 000 reads of 5 kb in a FASTQ file on tmpfs)."""
    import re
    import shutil
    import tempfile
    import oracle
    exe = os.path.join(ROOT, "charon_b200", "charon")
    if not os.path.exists(exe):
        return {"unavailable": "charon_b200/charon is not built"}
    wl = WORKLOADS["smoke"]
    records, rec_bin = make_refs(wl)
    idx, _ = oracle_index(wl, records, rec_bin, host_cores())
    L = wl["read_len"]
    bases, _, _ = make_reads(wl, records, rec_bin, n_reads, first_index=50_000_000)
    d = tempfile.mkdtemp(prefix="charon_bench_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    try:
        fq = fastq_text(bases, n_reads, L)
        open(os.path.join(d, "reads.fq"), "wb").write(fq)
        idx.save(os.path.join(d, "smoke.idx"))
        idx.run_like_reference(fq, threads=1, tsv_path=os.path.join(d, "oracle.tsv"), opts=oracle_options(wl))
        t0 = time.time()
        r = subprocess.run([exe, "dehost", os.path.join(d, "reads.fq"), "--db", os.path.join(d, "smoke.idx"), "-t", str(min(host_cores(), 255)),
                            "--log", os.path.join(d, "charon.log")], stdout=open(os.path.join(d, "gpu.tsv"), "wb"), stderr=subprocess.PIPE, timeout=600)
        wall = time.time() - t0
        same = r.returncode == 0 and open(os.path.join(d, "gpu.tsv"), "rb").read() == open(os.path.join(d, "oracle.tsv"), "rb").read()
        out = {"workload": wl["desc"], "reads": n_reads, "tsv_identical_to_oracle_t1": bool(same), "charon_dehost_wall_s": round(wall, 2),
               "threads": min(host_cores(), 255), "rows": sum(1 for _ in open(os.path.join(d, "gpu.tsv"), "rb"))}
        try:
            m = re.search(r"Read loop: (\d+) reads in ([0-9.]+) s \(([0-9.]+) reads/s\); stage wall times: reader ([0-9.]+) s, scalars \+ packing \(-t \d+\) ([0-9.]+) s, "
                          r"waiting for the GPU ([0-9.]+) s, writer ([0-9.]+) s", open(os.path.join(d, "charon.log")).read())
            if m:
                out.update(read_loop_s=float(m.group(2)), read_loop_reads_per_s=float(m.group(3)), reader_s=float(m.group(4)),
                           scalars_and_packing_s=float(m.group(5)), waiting_for_gpu_s=float(m.group(6)), writer_s=float(m.group(7)))
        except Exception:
            pass
        if not same:
            out["stderr_tail"] = r.stderr.decode(errors="replace")[-300:]
        # the same run with the compression ratios computed on the GPU (--gpu_compression) and ONE host thread
        t0 = time.time()
        r2 = subprocess.run([exe, "dehost", os.path.join(d, "reads.fq"), "--db", os.path.join(d, "smoke.idx"), "-t", "1", "--gpu_compression",
                             "--log", os.path.join(d, "charon2.log")], stdout=open(os.path.join(d, "gpu2.tsv"), "wb"), stderr=subprocess.PIPE, timeout=600)
        same2 = r2.returncode == 0 and open(os.path.join(d, "gpu2.tsv"), "rb").read() == open(os.path.join(d, "oracle.tsv"), "rb").read()
        g = {"tsv_identical_to_oracle_t1": bool(same2), "threads": 1, "charon_dehost_wall_s": round(time.time() - t0, 2)}
        try:
            m = re.search(r"Read loop: (\d+) reads in ([0-9.]+) s \(([0-9.]+) reads/s\)", open(os.path.join(d, "charon2.log")).read())
            if m:
                g.update(read_loop_s=float(m.group(2)), read_loop_reads_per_s=float(m.group(3)))
        except Exception:
            pass
        if not same2:
            g["stderr_tail"] = r2.stderr.decode(errors="replace")[-300:]
        out["with_gpu_compression"] = g
        return out
    finally:
        shutil.rmtree(d, ignore_errors=True)


def parity_mismatch(g, o, keys=("num_hashes", "counts", "unique_counts", "confidence", "call")):
    """None when every integer output agrees, else a description of the first differing read."""
    for key in keys:
        a, b = np.asarray(g[key]), np.asarray(o[key])
        if not np.array_equal(a, b):
            r = int(np.argwhere(a != b)[0][0])
            return (f"first difference in `{key}` at read {r}: gpu num_hashes={int(g['num_hashes'][r])} counts={g['counts'][r].tolist()[:8]} "
                    f"unique={g['unique_counts'][r].tolist()[:8]} call={int(g['call'][r])} | oracle num_hashes={int(o['num_hashes'][r])} "
                    f"counts={o['counts'][r].tolist()[:8]} unique={o['unique_counts'][r].tolist()[:8]} call={int(o['call'][r])}; "
                    f"{int((np.asarray(g['num_hashes']) != np.asarray(o['num_hashes'])).sum())} reads differ in num_hashes")
    return None


def run_reference(args, name, wl):
    rank, local_rank, world = dist_env()
    if rank != 0:
        return
    records, rec_bin = make_refs(wl)
    per_step_s = []
    idx, cb, _ = cpu_baseline(wl, records, rec_bin, seconds_target=max(3.0, 40.0 / max(1, args.steps + args.warmup)), stage_split=False)
    opts = oracle_options(wl)
    # K timed steps, each a bounded sample sized from the probe
    n = int(max(256, min(cb["value"] * 4.0, 100_000)))
    bases, seg_off, _ = make_reads(wl, records, rec_bin, n, first_index=30_000_000)
    for i in range(args.warmup + args.steps):
        r = idx.classify_batch(bases, seg_off, threads=cb["cores"], opts=opts)
        if i >= args.warmup:
            per_step_s.append(r["seconds"])
    total = sum(per_step_s)
    val = n * len(per_step_s) / total
    line = {"impl": "reference", "metric": "reads_per_sec", "value": val, "unit": "reads/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1000 * total / len(per_step_s), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u64", "data": "synthetic", "config": shared_config(name, wl),
            "cpu_baseline": {"value": val, "unit": "reads/s", "cores": cb["cores"], "kind": "port",
                             "sample": f"{n} reads per step x {args.steps} steps of the same workload; hot path only (the oracle port, plain bit-matrix "
                                       f"filter); oracle index build {cb['index_build_s']} s outside the timed region"},
            "gbases_per_sec": val * wl["read_len"] / 1e9,
            "e2e": {"value": val, "unit": "reads/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def run_b200(args, name, wl):
    rank, local_rank, world = dist_env()
    import ctypes as C
    import charon_b200
    from charon_b200 import api as A
    from charon_b200.api import pinned_array
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    partitioned = bool(wl.get("partitioned") or args.partitioned)
    if partitioned and world < 2:
        sys.stderr.write("bench.py: --workload partitioned splits the filter across GPUs: it needs --gpus N with N >= 2\n")
        sys.exit(2)
    R, L = args.reads_per_step or wl["reads_per_step"], wl["read_len"]
    ncat = len(categories(wl))
    records, rec_bin = make_refs(wl)
    t0 = time.time()
    gpu = charon_b200.GpuClassifier.from_references(records, rec_bin, bin_categories(wl), ncat, device=local_rank)   # index built on the GPU
    build_s = time.time() - t0
    part_check = None
    if partitioned:
        # BASELINE config 4: the filter is split by row range over the ranks; rows owned by a peer are fetched through its
        # NVLink peer mapping inside the count kernel.  Reads stay sharded, still no data-path collective.
        import torch
        from charon_b200 import multi
        # the same sample classified on the whole filter first (every rank still holds it here) and again once it is split
        gpu.set_model(charon_b200.DehostOptions() if wl["mode"] == "dehost" else charon_b200.ClassifyOptions(), host_index=0)
        n_chk = 20_000
        cb_, co_, _ = make_reads(wl, records, rec_bin, n_chk, first_index=40_000_000 + rank * n_chk)
        chk_args = (cb_, co_, np.full(n_chk, 20.0, np.float32), np.full(n_chk, 0.3, np.float32), np.full(n_chk, wl["read_len"], np.uint32))
        whole = gpu.classify_batch(*chk_args)
        replicated_bytes = int(gpu.info().filter_bytes)
        gpu.partition(rank, world)
        multi.exchange_partitions(gpu, dist, rank, world, device=torch.device("cuda", local_rank))
        split = gpu.classify_batch(*chk_args)
        same = all(np.array_equal(whole[k], split[k]) for k in ("num_hashes", "counts", "unique_counts", "call", "confidence")) and \
            np.array_equal(whole["prob"], split["prob"], equal_nan=True)
        t = torch.tensor([0 if same else 1], dtype=torch.int32, device="cuda")
        dist.all_reduce(t)
        part_check = {"partitioned_equals_replicated": int(t.item()) == 0, "reads_checked_per_rank": n_chk, "ranks": world,
                      "filter_bytes_replicated": replicated_bytes, "filter_bytes_per_gpu_partitioned": int(gpu.info().filter_bytes)}
        if int(t.item()) != 0:
            if rank == 0:
                sys.stderr.write("bench.py: PARITY FAILURE: the partitioned filter does not give the replicated filter's outputs\n")
            os._exit(3)
    opts = charon_b200.DehostOptions() if wl["mode"] == "dehost" else charon_b200.ClassifyOptions()
    gpu.set_model(opts, host_index=0)
    pack_threads = max(1, min(16, host_cores() // world))
    pack_mode = {"auto": -1, "off": 0, "on": 1}[args.host_pack]
    gpu.set_host_pack(pack_mode, pack_threads)
    # this rank's batch (weak scaling: R reads per GPU per step), in pinned host memory: ASCII, and the 2-bit form a parser
    # written for this library emits (every read on a 64-base boundary, bases that are not A/C/G/T in the exception list)
    bases = pinned_array((R * L,), np.uint8)
    _, seg_off, _ = make_reads(wl, records, rec_bin, R, first_index=rank * R, out=bases)
    seg_pin = pinned_array((R + 1,), np.uint64); seg_pin[:] = seg_off
    Lp = (L + 63) // 64 * 64
    p2 = pinned_array((R * Lp // 4 + 64,), np.uint8)
    pb = A.pack_batch(bases, seg_off, A.BASES_2BIT, align=64, out=p2)
    p2_off = pinned_array((R + 1,), np.uint64); p2_off[:] = pb.seg_off
    p2_len = pinned_array((R,), np.uint64); p2_len[:] = pb.seg_len
    n_exc = len(pb.exc_pos)
    p2_exc = pinned_array((max(n_exc, 1),), np.uint64); p2_exc[:n_exc] = pb.exc_pos
    mq = pinned_array((R,), np.float32); mq[:] = 20.0
    cr = pinned_array((R,), np.float32); cr[:] = 0.3
    ln = pinned_array((R,), np.uint32); ln[:] = L

    def outputs():
        o = dict(num_hashes=pinned_array((R,), np.uint32), counts=pinned_array((R, ncat), np.uint32), unique_counts=pinned_array((R, ncat), np.uint32),
                 prob=pinned_array((R, ncat), np.float64), call=pinned_array((R,), np.uint8), confidence=pinned_array((R,), np.uint8))
        return o, A.CountsOut(o["num_hashes"].ctypes.data, o["counts"].ctypes.data, o["unique_counts"].ctypes.data), \
            A.CallsOut(o["prob"].ctypes.data, o["call"].ctypes.data, o["confidence"].ctypes.data)

    out, co, ca = outputs()
    out2, co2, ca2 = outputs()
    slots = [(co, ca), (co2, ca2)]
    batch_ascii = A.Batch(bases.ctypes.data, seg_pin.ctypes.data, None, R, R, A.BASES_ASCII, None, None, 0)
    batch_2bit = A.Batch(p2.ctypes.data, p2_off.ctypes.data, None, R, R, A.BASES_2BIT, p2_len.ctypes.data, p2_exc.ctypes.data if n_exc else None, n_exc)
    sc = A.ReadScalars(mq.ctypes.data, cr.ctypes.data, ln.ctypes.data)
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

    def resident_arm(batch):
        gpu._check(Lb.cg_stage_batch(gpu._h, C.byref(batch), C.byref(sc)), "cg_stage_batch")
        gpu._staged_n = R
        sampler = ClockSampler(local_rank); sampler.start()      # early: with 8 ranks nvidia-smi takes a while to print its first line
        for _ in range(args.warmup):
            gpu.run_staged(True)
        t_wait = time.time()
        while not sampler.rows and time.time() - t_wait < 3.0:
            time.sleep(0.02)
        barrier()
        w0 = time.time()
        gpu.timer_start()
        launches = 0
        for _ in range(args.steps):
            gpu.run_staged(True)
            launches += int(gpu.info().last_kernel_launches)
        total_ms, count_ms, eval_ms, n_runs = gpu.timer_stop()
        w1 = time.time()
        barrier()
        clocks = sampler.stop(w0, w1)
        res = gpu.fetch_staged(True)
        return dict(total_ms=max_over_ranks(total_ms), count_ms=count_ms, eval_ms=eval_ms, n_runs=n_runs, clocks=clocks, res=res, launches=launches)

    # ---- kernel-only arm: inputs resident in HBM (2-bit bases, the transport format) -----------------------------------------
    ra = resident_arm(batch_2bit)
    res = ra["res"]
    minimisers = int(res["num_hashes"].astype(np.int64).sum())
    # the same kernels on ASCII-resident bases, for comparison only (outputs must not depend on the format)
    gpu.set_host_pack(0, 0)
    rb = resident_arm(batch_ascii)
    for key in ("num_hashes", "counts", "unique_counts", "call", "confidence"):
        assert np.array_equal(rb["res"][key], res[key]), f"{key} differs between 2-bit and ASCII resident bases"
    gpu.set_host_pack(pack_mode, 0)

    # ---- end-to-end arm: host buffers through the C-ABI (H2D + kernels + D2H of every step inside the timed region) -------
    # cg_submit_batch / cg_wait_batch on two slots: the PCIe copy of step i+1 overlaps the kernels of step i
    def e2e_steps(batch, n):
        for i in range(n):
            sl = i & 1
            if i >= 2:
                gpu._check(Lb.cg_wait_batch(gpu._h, sl), "cg_wait_batch")
            gpu._check(Lb.cg_submit_batch(gpu._h, sl, C.byref(batch), C.byref(sc), C.byref(slots[sl][0]), C.byref(slots[sl][1])), "cg_submit_batch")
        gpu._check(Lb.cg_wait_batch(gpu._h, 0), "cg_wait_batch")
        gpu._check(Lb.cg_wait_batch(gpu._h, 1), "cg_wait_batch")

    def e2e_arm(batch):
        e2e_steps(batch, max(2, min(args.warmup, 3)))
        barrier()
        gpu.timer_start()
        e2e_steps(batch, args.steps)
        ms, _, _, _ = gpu.timer_stop()
        barrier()
        info = gpu.info()
        assert np.array_equal(out["num_hashes"], res["num_hashes"]) and np.array_equal(out["call"], res["call"])
        assert args.steps < 2 or (np.array_equal(out2["num_hashes"], res["num_hashes"]) and np.array_equal(out2["call"], res["call"]))
        return max_over_ranks(ms), info

    e2e_ms, e2e_info = e2e_arm(batch_2bit)
    ascii_ms, ascii_info = e2e_arm(batch_ascii)
    # serial variant of the same call (cg_classify_batch: copy, kernels, copy back, one after the other)
    gpu._check(Lb.cg_classify_batch(gpu._h, C.byref(batch_2bit), C.byref(sc), C.byref(co), C.byref(ca)), "cg_classify_batch")
    barrier()
    gpu.timer_start()
    for _ in range(args.steps):
        gpu._check(Lb.cg_classify_batch(gpu._h, C.byref(batch_2bit), C.byref(sc), C.byref(co), C.byref(ca)), "cg_classify_batch")
    serial_ms, _, _, _ = gpu.timer_stop()
    barrier()
    serial_ms = max_over_ranks(serial_ms)
    # ---- the compression-ratio scalar of the same batch on the device (outside the timed regions above: the
    #      bench's scalars are given, as BASELINE's metric has them), checked against the box's zlib on a sample -----------------
    gz_sizes = pinned_array((R,), np.uint32)
    gz_ratio = pinned_array((R,), np.float32)
    gz_ms = []
    for _ in range(2):
        gpu._check(Lb.cg_compression_batch(gpu._h, C.byref(batch_2bit), gz_sizes.ctypes.data, gz_ratio.ctypes.data), "cg_compression_batch")
        gz_ms.append(float(gpu.info().last_gzip_kernel_ms))
    import zlib
    n_gz = min(R, 300)
    t_z = time.time()
    z_sizes = []
    for i in range(n_gz):
        zc = zlib.compressobj(-1, zlib.DEFLATED, 31, 8, 0)
        z_sizes.append(len(zc.compress(bytes(bases[i * L:(i + 1) * L]).upper()) + zc.flush()))
    t_z = time.time() - t_z
    compression = {"what": "gzip size (zlib level 6, gzip wrapper, memLevel 8: reference src/utils.cpp:114-124) of every read of the step, cg_compression_batch",
                   "gzip_kernel_ms": min(gz_ms), "reads_per_sec": R / (min(gz_ms) / 1e3), "sizes_equal_zlib_on_sample": bool(np.array_equal(gz_sizes[:n_gz], np.array(z_sizes, np.uint32))),
                   "sample_reads": n_gz, "zlib_version": zlib.ZLIB_RUNTIME_VERSION, "zlib_one_core_reads_per_sec": n_gz / t_z, "mean_ratio": float(np.mean(gz_ratio))}
    if not compression["sizes_equal_zlib_on_sample"]:
        sys.stderr.write("bench.py: PARITY FAILURE: gzip sizes from the device differ from zlib\n")
        os._exit(3)
    # ---- what the memory system gives a kernel that ONLY gathers, on this very filter -------------------------
    g_query_ms = gpu.bench_gather(minimisers, mode=0, reps=5)
    g_sector_ms = gpu.bench_gather(3 * minimisers, mode=1, reps=5)
    info = gpu.info()
    if rank != 0:
        if dist is not None:
            dist.barrier()          # rank 0 may still be checking parity
            dist.destroy_process_group()
        return
    reads_total = R * args.steps * world
    value = reads_total / (ra["total_ms"] / 1e3)
    e2e = reads_total / (e2e_ms / 1e3)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    algo_bytes = ALGO_BYTES_PER_MINIMISER * minimisers + ALGO_BYTES_PER_BASE * R * L
    avg_count_ms = ra["count_ms"] / max(ra["n_runs"], 1)
    achieved = algo_bytes / (avg_count_ms / 1e3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "count_kernel_traffic.json")
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get(name)
        except Exception:
            traffic = None
    # contract: `traffic` = DRAM bytes (read + write) of one launch from the ncu --set full capture, or null; the capture's
    # description travels beside it
    traffic_detail = traffic
    traffic = int(traffic["bytes_for_bench_launch"]) if isinstance(traffic, dict) and "bytes_for_bench_launch" in traffic else None
    d2h = int(R * (4 + 8 * ncat + 8 * ncat + 2))
    line = {"metric": "reads_per_sec", "value": value, "unit": "reads/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ra["total_ms"] / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64",
            "data": "synthetic", "config": shared_config(name, wl),
            "details": {"compression_on_device": compression, "filter_partitioned": partitioned, "partition_check": part_check, "reads_per_step_per_gpu": R,
                        "filter_rows": int(info.bin_size), "filter_bytes_hbm": int(info.filter_bytes), "row_bits": int(info.row_bits),
                        "minimisers_per_step_per_gpu": minimisers, "classified_frac": float((res["call"] != 255).mean()),
                        "hbm_resident_bases": "2 bits per base (CG_BASES_2BIT), every read on a 64-base boundary",
                        "non_acgt_bases_per_step": n_exc,
                        "value_on_ascii_resident_bases": {"reads_per_sec": reads_total / (rb["total_ms"] / 1e3),
                                                          "count_kernel_ms": rb["count_ms"] / max(rb["n_runs"], 1)},
                        "index_build_on_gpu_s": round(build_s, 2), "count_grid": int(info.count_grid)},
            "gbases_per_sec": value * L / 1e9,
            "e2e": {"value": e2e, "unit": "reads/s", "h2d_bytes_per_step": int(e2e_info.last_h2d_bytes) * world, "d2h_bytes_per_step": d2h * world,
                    "transport": "pinned host buffers in CG_BASES_2BIT (the format the library's FASTQ parser emits: 2 bits per base, reads on 64-base "
                                 "boundaries, exception list for non-ACGT) copied as they are",
                    "host_pack_ms_per_step": float(e2e_info.last_pack_ms), "gbases_per_sec": e2e * L / 1e9,
                    "api": "cg_submit_batch/cg_wait_batch, two slots (copy of step i+1 overlaps kernels of step i)",
                    "ms_per_step": e2e_ms / args.steps,
                    "last_step_device_ms": {"h2d": float(e2e_info.last_h2d_ms), "count_kernel": float(e2e_info.last_count_kernel_ms),
                                            "eval_kernels": float(e2e_info.last_eval_kernel_ms), "d2h": float(e2e_info.last_d2h_ms)},
                    "from_ascii_host_buffers": {"value": reads_total / (ascii_ms / 1e3), "h2d_bytes_per_step": int(ascii_info.last_h2d_bytes) * world,
                                                "transport": "ASCII packed to 4-bit ranks inside the call by the library's thread pool" if ascii_info.last_packed else "ASCII copied as is",
                                                "host_pack_threads_per_gpu": pack_threads if ascii_info.last_packed else 0,
                                                "host_pack_ms_per_step": float(ascii_info.last_pack_ms)},
                    "serial_cg_classify_batch_value": reads_total / (serial_ms / 1e3)},
            "gpu_launches": int(ra["launches"]),
            "clocks": ra["clocks"],
            "roofline": {"bound": "hbm", "kernel": "cgk::count_kernel_v3", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "traffic_unit": "bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum)", "traffic_capture": traffic_detail,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs (burst copy)" if peaks else "fallback 6650",
                         "algorithmic_bytes_per_launch": algo_bytes, "avg_launch_ms": avg_count_ms,
                         "count_kernel_share_of_step": ra["count_ms"] / (ra["total_ms"] if world == 1 else max(ra["count_ms"] + ra["eval_ms"], 1e-9)),
                         "gather_microbench": {
                             "what": "same filter, same 3-row addressing, nothing else in the kernel (cg_bench_gather mode 0) / one random 32 B sector per load (mode 1)",
                             "query_addressing_GBps": ALGO_BYTES_PER_MINIMISER * minimisers / (g_query_ms / 1e3) / 1e9,
                             "pure_sector_GBps": 32 * 3 * minimisers / (g_sector_ms / 1e3) / 1e9,
                             "count_kernel_gather_GBps": ALGO_BYTES_PER_MINIMISER * minimisers / (avg_count_ms / 1e3) / 1e9,
                             "frac_of_achievable_gather": (g_query_ms / avg_count_ms)},
                         "eval_kernels_ms_per_step": ra["eval_ms"] / max(ra["n_runs"], 1),
                         "note": "96 B/minimiser + 1 B/base (SURVEY 8d) regardless of L2 hits and of rows the query never had to fetch; compacted "
                                 f"filter of this workload is {info.filter_bytes / 1e6:.0f} MB"}}
    if not args.no_cpu_baseline:
        # N = 1: the full CPU baseline; N > 1: a small oracle sample, only to gate parity on rank 0
        if world == 1:
            idx, cb, (sb, so, sr) = cpu_baseline(wl, records, rec_bin)
        else:
            idx, build_s2 = oracle_index(wl, records, rec_bin, host_cores())
            n = 20_000
            sb, so, _ = make_reads(wl, records, rec_bin, n, first_index=20_000_000)
            sr = idx.classify_batch(sb, so, threads=host_cores(), opts=oracle_options(wl))
            cb = None
        n = len(so) - 1
        if partitioned:
            bad = None     # a rank cannot classify on its own once the filter is split; parity of this mode is a GPU test
        else:
            g = gpu.classify_batch(sb, so, np.full(n, 20.0, np.float32), np.full(n, 0.3, np.float32), np.full(n, L, np.uint32))
            # dehost (KDE model): everything bit-exact.  classify evaluates beta densities through gcem in the reference, whose
            # source is not available (probabilities best effort there): the gate covers every integer the GPU
            # derives, and calls that flip on a near-tie of two probabilities are counted and reported, not fatal.
            keys = ("num_hashes", "counts", "unique_counts", "confidence", "call") if wl["mode"] == "dehost" else ("num_hashes", "counts", "unique_counts", "confidence")
            bad = parity_mismatch(g, sr, keys)
            line["details"]["calls_differing_from_oracle_on_sample"] = int((np.asarray(g["call"]) != np.asarray(sr["call"])).sum())
        if cb is not None:
            cb["parity_on_sample"] = bad is None
            cb["parity_sample_reads"] = n
            line["cpu_baseline"] = cb
        else:
            line["parity_on_sample"] = {"ok": bad is None, "reads": n, "rank": 0}
        if world == 1 and bad is None and not partitioned:
            gpu.sync()
            line["cli_tsv_check"] = cli_tsv_check()
            if line["cli_tsv_check"].get("tsv_identical_to_oracle_t1") is False or \
                    line["cli_tsv_check"].get("with_gpu_compression", {}).get("tsv_identical_to_oracle_t1") is False:
                bad = "the TSV of charon_b200/charon dehost differs from the oracle's on the smoke configuration: " + json.dumps(line["cli_tsv_check"])
        if bad is not None:
            sys.stderr.write(f"bench.py: PARITY FAILURE against the oracle on {n} reads of the workload: {bad}\n"
                             "bench.py: no bench line is printed for a kernel that is not bit-exact\n")
            sys.stderr.flush()
            os._exit(3)
    emit(line)
    gpu.close()
    if dist is not None:
        dist.barrier()
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
    ap.add_argument("--workload", default="human", choices=sorted(WORKLOADS))
    ap.add_argument("--reads-per-step", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--host-pack", default="auto", choices=["auto", "on", "off"], help="ASCII callers only: 4-bit ranks packed on host threads (auto: when >= 8 threads per GPU) or ASCII copies")
    ap.add_argument("--partitioned", action="store_true", help="N>1: split the filter by row range across the GPUs instead of replicating it (same as --workload partitioned on any index)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    wl = WORKLOADS[args.workload]
    rank, local_rank, world = dist_env()
    if args.impl == "reference":
        run_reference(args, args.workload, wl)
        return
    if world != args.gpus and world == 1 and args.gpus > 1:
        # launched without torchrun: re-exec under torch.distributed.run
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}", "--master-addr", "127.0.0.1",
               "--master-port", "29511"] + sys.argv
        os.dup2(_REAL_STDOUT, 1)   # the child sets up its own re-routing
        os.execv(sys.executable, cmd)
    run_b200(args, args.workload, wl)


if __name__ == "__main__":
    main()
