"""ctypes binding of libcharon_gpu.so (include/charon_gpu.h) and a thin host-side mirror of the reference's
classification objects for Python callers (tests, bench).

The product path has NO CPU fallback: if the CUDA library is missing or no B200 is visible, construction raises.
Nothing in this package imports the oracle.
"""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CHARON_GPU_LIB") or os.path.join(_HERE, "libcharon_gpu.so")   # env override: kernel-variant experiments

u8, u16, u32, u64, i32, f32, f64, vp = C.c_uint8, C.c_uint16, C.c_uint32, C.c_uint64, C.c_int32, C.c_float, C.c_double, C.c_void_p


class BuildDesc(C.Structure):
    _fields_ = [("kmer_size", u8), ("window_size", u8), ("hash_funs", u32), ("max_fpr", f64), ("max_bits", u64), ("bases", vp),
                ("rec_off", vp), ("rec_bin", vp), ("n_records", u32), ("bins", u32), ("ncat", u16), ("bin_to_cat", vp), ("force_max_hashes", u64)]


class IndexDesc(C.Structure):
    _fields_ = [("kmer_size", u8), ("window_size", u8), ("bins", u64), ("technical_bins", u64), ("bin_size", u64), ("hash_shift", u64),
                ("bin_words", u64), ("hash_funs", u64), ("filter_format", i32), ("rows", vp), ("sd_low", vp), ("sd_low_bits", u64),
                ("sd_high", vp), ("sd_high_bits", u64), ("sd_wl", u8), ("ncat", u16), ("bin_to_cat", vp)]


class CategoryModel(C.Structure):
    _fields_ = [("dist", i32), ("pos", vp), ("npos", u32), ("h_pos", f32), ("neg", vp), ("nneg", u32), ("h_neg", f32),
                ("beta_pos", f32 * 2), ("beta_neg", f32 * 2), ("gamma_pos", f32 * 3), ("gamma_neg", f32 * 3)]


class Thresholds(C.Structure):
    _fields_ = [("min_quality", f32), ("min_length", u32), ("min_compression", f32), ("confidence_threshold", C.c_int8),
                ("confidence_probability_threshold", f32), ("min_hits", u8), ("host_unique_prop_lo_threshold", f32),
                ("min_proportion_difference", f32), ("min_prob_difference", f32), ("host_index", u8)]


class Batch(C.Structure):
    _fields_ = [("bases", vp), ("seg_off", vp), ("read_seg", vp), ("n_reads", u32), ("n_segs", u32),
                ("bases_format", i32), ("seg_len", vp), ("exc_pos", vp), ("n_exc", u64)]


BASES_ASCII, BASES_RANK4, BASES_2BIT = 0, 1, 2


class CountsOut(C.Structure):
    _fields_ = [("num_hashes", vp), ("counts", vp), ("unique_counts", vp)]


class ReadScalars(C.Structure):
    _fields_ = [("mean_quality", vp), ("compression", vp), ("length", vp)]


class CallsOut(C.Structure):
    _fields_ = [("probabilities", vp), ("call", vp), ("confidence", vp)]


class Info(C.Structure):
    _fields_ = [("bin_size", u64), ("filter_bytes", u64), ("row_bits", u32), ("single_pass", u32), ("sm_count", u32), ("count_grid", u32),
                ("count_block", u32), ("count_smem_bytes", u32), ("last_minimisers", u64), ("last_bases", u64),
                ("last_kernel_launches", u32), ("last_count_kernel_ms", f32), ("last_eval_kernel_ms", f32), ("last_h2d_ms", f32),
                ("last_d2h_ms", f32), ("last_h2d_bytes", u64), ("last_pack_ms", f32), ("last_packed", u32), ("last_gzip_kernel_ms", f32)]


EVAL_STATIC_MAX_READS, EVAL_STATIC_MAX_PAIRS = 16384, 65536    # include/charon_gpu.h

EXPORTS = ["cg_create", "cg_build_index", "cg_destroy", "cg_last_error", "cg_set_stream", "cg_sync", "cg_set_model", "cg_count_batch", "cg_eval_batch",
           "cg_classify_batch", "cg_stage_batch", "cg_run_staged", "cg_fetch_staged", "cg_get_info", "cg_host_alloc", "cg_host_free", "cg_timer_start", "cg_timer_stop", "cg_debug_minimisers",
           "cg_debug_query", "cg_debug_dump_filter", "cg_debug_force_generic", "cg_create_partitioned", "cg_partition_from_replicated",
           "cg_partition_export", "cg_partition_attach_ipc", "cg_partition_attach_peer", "cg_bench_gather", "cg_submit_batch", "cg_wait_batch", "cg_export_sd_vector", "cg_select_slot", "cg_set_host_pack", "cg_pack_bases", "cg_pack_bases_2bit", "cg_eval_batch_static", "cg_compression_batch", "cg_fetch_compression"]

_lib = None


def load_library():
    """dlopen libcharon_gpu.so; raises (never falls back) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: build it with `make -C charon_b200/csrc` (or __graft_entry__.build()); "
                           "charon_b200 has no CPU fallback")
    L = C.CDLL(LIB_PATH)
    P = C.POINTER
    L.cg_create.argtypes = [i32, P(IndexDesc), C.c_size_t, P(vp)]
    L.cg_build_index.argtypes = [i32, P(BuildDesc), C.c_size_t, vp, P(vp)]
    L.cg_create_partitioned.argtypes = [i32, P(IndexDesc), u32, u32, P(vp)]
    L.cg_partition_from_replicated.argtypes = [vp, u32, u32]
    L.cg_partition_export.argtypes = [vp, vp]
    L.cg_partition_attach_ipc.argtypes = [vp, u32, vp]
    L.cg_partition_attach_peer.argtypes = [vp, u32, vp]
    L.cg_submit_batch.argtypes = [vp, i32, P(Batch), P(ReadScalars), P(CountsOut), P(CallsOut)]
    L.cg_wait_batch.argtypes = [vp, i32]
    L.cg_select_slot.argtypes = [vp, i32]
    L.cg_set_host_pack.argtypes = [vp, i32, i32]
    L.cg_pack_bases.argtypes = [vp, u64, vp]
    L.cg_pack_bases_2bit.argtypes = [vp, u64, vp, u64, vp, u64, P(u64)]
    L.cg_export_sd_vector.argtypes = [vp, P(u64), P(C.c_uint8), P(u64), P(u64), vp, vp]
    L.cg_bench_gather.argtypes = [vp, u64, i32, i32, P(f32)]
    L.cg_destroy.argtypes = [vp]; L.cg_destroy.restype = None
    L.cg_last_error.argtypes = [vp]; L.cg_last_error.restype = C.c_char_p
    L.cg_set_stream.argtypes = [vp, vp]
    L.cg_sync.argtypes = [vp]
    L.cg_set_model.argtypes = [vp, P(CategoryModel), P(Thresholds), i32]
    L.cg_count_batch.argtypes = [vp, P(Batch), P(CountsOut)]
    L.cg_eval_batch.argtypes = [vp, u32, vp, vp, vp, P(ReadScalars), P(CallsOut)]
    L.cg_compression_batch.argtypes = [vp, P(Batch), vp, vp]
    L.cg_fetch_compression.argtypes = [vp, i32, vp]
    L.cg_eval_batch_static.argtypes = [vp, u32, vp, vp, vp, P(ReadScalars), P(CallsOut)]
    L.cg_classify_batch.argtypes = [vp, P(Batch), P(ReadScalars), P(CountsOut), P(CallsOut)]
    L.cg_stage_batch.argtypes = [vp, P(Batch), P(ReadScalars)]
    L.cg_run_staged.argtypes = [vp, i32]
    L.cg_fetch_staged.argtypes = [vp, P(CountsOut), P(CallsOut)]
    L.cg_get_info.argtypes = [vp, P(Info)]
    L.cg_host_alloc.argtypes = [C.c_size_t, P(vp)]
    L.cg_host_free.argtypes = [vp]; L.cg_host_free.restype = None
    L.cg_timer_start.argtypes = [vp]
    L.cg_timer_stop.argtypes = [vp, P(f32), P(f32), P(f32), P(u32)]
    L.cg_debug_minimisers.argtypes = [vp, vp, u64, vp, u64, P(u64)]
    L.cg_debug_query.argtypes = [vp, vp, u64, vp]
    L.cg_debug_dump_filter.argtypes = [vp, vp]
    L.cg_debug_force_generic.argtypes = [vp, i32]
    _lib = L
    return L


class CharonGpuError(RuntimeError):
    pass


def pinned_array(shape, dtype):
    """numpy array backed by cudaHostAlloc'd (pinned) memory; freed when the array is garbage collected."""
    L = load_library()
    dt = np.dtype(dtype)
    n = int(np.prod(shape)) * dt.itemsize
    p = vp()
    rc = L.cg_host_alloc(n, C.byref(p))
    if rc != 0:
        raise CharonGpuError(f"cg_host_alloc failed ({rc}): {L.cg_last_error(None).decode()}")
    buf = (C.c_char * max(n, 1)).from_address(p.value)
    arr = np.frombuffer(buf, dtype=dt, count=int(np.prod(shape))).reshape(shape)

    class _Owner:
        def __init__(self, ptr):
            self.ptr = ptr

        def __del__(self):
            try:
                L.cg_host_free(self.ptr)
            except Exception:
                pass
    _PINNED[id(buf)] = _Owner(p)          # keep the owner alive as long as the process (bench buffers)
    return arr


_PINNED = {}


def pack_bases(bases) -> np.ndarray:
    """4-bit dna5 ranks of an ASCII base array, two per byte (cg_pack_bases: the transport the library uses on PCIe)."""
    L = load_library()
    b = np.ascontiguousarray(np.frombuffer(bases.encode() if isinstance(bases, str) else bases, np.uint8) if not isinstance(bases, np.ndarray) else bases, np.uint8)
    out = np.zeros((b.size + 1) // 2, np.uint8)
    rc = L.cg_pack_bases(b.ctypes.data_as(C.c_void_p), b.size, out.ctypes.data_as(C.c_void_p))
    if rc:
        raise RuntimeError(f"cg_pack_bases failed ({rc})")
    return out


class PackedBatch:
    """A batch whose bases are already in one of the packed transports (what a parser written for this library emits).
    `bases` packed bytes, `seg_off` / `seg_len` in bases (segments may be padded apart), `exc_pos` for BASES_2BIT."""

    def __init__(self, fmt, bases, seg_off, seg_len=None, exc_pos=None, read_seg=None):
        self.fmt, self.bases, self.seg_off, self.seg_len, self.exc_pos, self.read_seg = fmt, bases, seg_off, seg_len, exc_pos, read_seg


def pack_batch(bases, seg_off, fmt=BASES_2BIT, align=64, read_seg=None, out=None) -> PackedBatch:
    """Pack an ASCII batch (concatenated bases + offsets) read by read into `fmt`, every segment starting on a multiple of
    `align` bases — what host/fastx.hpp does while it parses.  `out`: optional preallocated uint8 buffer (e.g. pinned)."""
    L = load_library()
    b = np.ascontiguousarray(bases, np.uint8)
    so = np.ascontiguousarray(seg_off, np.uint64)
    n = len(so) - 1
    lens = np.diff(so).astype(np.uint64)
    per = 4 if fmt == BASES_2BIT else 2
    assert fmt in (BASES_2BIT, BASES_RANK4) and align % per == 0
    padded = (lens + np.uint64(align - 1)) // np.uint64(align) * np.uint64(align)
    new_off = np.zeros(n + 1, np.uint64)
    np.cumsum(padded, out=new_off[1:])
    total = int(new_off[-1])
    nbytes = (total + per - 1) // per + 16
    buf = out if out is not None else np.zeros(nbytes, np.uint8)
    assert buf.size >= nbytes
    exc = np.zeros(max(16, total // 64), np.uint64)
    n_exc = 0
    k = u64(0)
    for s in range(n):
        lo, ln, dst = int(so[s]), int(lens[s]), int(new_off[s])
        if ln == 0:
            continue
        src = b[lo:lo + ln]
        if fmt == BASES_2BIT:
            while True:
                cap = exc.size - n_exc
                rc = L.cg_pack_bases_2bit(src.ctypes.data, ln, buf.ctypes.data + dst // 4, dst, exc.ctypes.data + 8 * n_exc, cap, C.byref(k))
                if rc:
                    raise CharonGpuError(f"cg_pack_bases_2bit failed ({rc}): {L.cg_last_error(None).decode()}")
                if k.value <= cap:
                    break
                exc = np.concatenate([exc, np.zeros(max(exc.size, int(k.value)), np.uint64)])
            n_exc += int(k.value)
        else:
            rc = L.cg_pack_bases(src.ctypes.data, ln, buf.ctypes.data + dst // 2)
            if rc:
                raise CharonGpuError(f"cg_pack_bases failed ({rc})")
    return PackedBatch(fmt, buf, new_off, lens, exc[:n_exc].copy() if fmt == BASES_2BIT else None, read_seg)


def _ptr(a):
    return a.ctypes.data if a is not None else None


class DehostOptions:
    """Thresholds of DehostArguments (reference include/dehost_arguments.hpp:22-43)."""
    mode = 0

    def __init__(self, **kw):
        self.dist = "kde"; self.min_quality = 15.0; self.min_length = 140; self.min_compression = 0.0
        self.confidence_threshold = 7; self.confidence_probability_threshold = 0.0; self.host_unique_prop_lo_threshold = 0.05
        self.min_proportion_difference = 0.04; self.min_prob_difference = 0.0; self.min_hits = 0
        self.__dict__.update(kw)


class ClassifyOptions(DehostOptions):
    """Thresholds of ClassifyArguments (reference include/classify_arguments.hpp:16-39)."""
    mode = 1

    def __init__(self, **kw):
        super().__init__()
        self.dist = "beta"; self.min_quality = 10.0; self.min_compression = 0.15; self.confidence_threshold = 2
        self.min_proportion_difference = 0.0
        self.__dict__.update(kw)


class GpuClassifier:
    """One handle = one GPU holding the compacted filter.  Mirrors what Index::agent() + Result give the
    reference's loop (src/dehost_main.cpp:321,331): `count_batch`, `classify_batch`."""

    @classmethod
    def from_references(cls, records, rec_bin, bin_to_cat, ncat, k=19, w=41, hash_funs=3, max_fpr=0.01, device=0, mem_budget=0):
        """`charon index` data path on the GPU (cg_build_index): records = list of uint8 arrays, rec_bin[i] = bin."""
        L = load_library()
        self = cls.__new__(cls)
        self._L = L
        cat = np.ascontiguousarray(np.concatenate(records)) if len(records) > 1 else np.ascontiguousarray(records[0], dtype=np.uint8)
        off = np.zeros(len(records) + 1, dtype=np.uint64); off[1:] = np.cumsum([len(r) for r in records])
        rb = np.ascontiguousarray(rec_bin, dtype=np.uint32); b2c = np.ascontiguousarray(bin_to_cat, dtype=np.uint8)
        d = BuildDesc(k, w, hash_funs, max_fpr, 0, cat.ctypes.data, off.ctypes.data, rb.ctypes.data, len(records), len(b2c), ncat, b2c.ctypes.data, 0)
        hpb = np.zeros(len(b2c), np.uint64)
        h = vp()
        rc = L.cg_build_index(device, C.byref(d), mem_budget, hpb.ctypes.data, C.byref(h))
        if rc != 0:
            raise CharonGpuError(f"cg_build_index failed ({rc}): {L.cg_last_error(None).decode()}")
        self._h = h
        self._keep = []; self._model_keep = None
        self.hashes_per_bin = hpb
        self.k, self.w, self.bins, self.ncat = k, w, len(b2c), ncat
        self.bin_words = (self.bins + 63) // 64
        self.bin_size = int(self.info().bin_size)
        return self

    def __init__(self, *, k, w, bins, bin_size, hash_funs, bin_to_cat, ncat, rows=None, sd=None, device=0, mem_budget=0, part=None):
        """part=(rank, world): store only this rank's row range of the filter (cg_create_partitioned); the other parts must
        then be attached (attach_peer / attach_ipc) before the first batch."""
        L = load_library()
        self._L = L
        self.k, self.w, self.bins, self.bin_size, self.ncat = k, w, bins, bin_size, ncat
        self.bin_words = (bins + 63) // 64
        d = IndexDesc()
        d.kmer_size, d.window_size, d.bins, d.bin_words = k, w, bins, self.bin_words
        d.technical_bins, d.bin_size, d.hash_funs = self.bin_words * 64, bin_size, hash_funs
        d.hash_shift = 64 - int(bin_size).bit_length()
        self._keep = []
        if rows is not None:
            rows = np.ascontiguousarray(rows, dtype=np.uint64)
            assert rows.size == bin_size * self.bin_words
            d.filter_format, d.rows = 0, rows.ctypes.data
            self._keep.append(rows)
        else:
            low, low_bits, high, high_bits, wl = sd
            low = np.ascontiguousarray(low, dtype=np.uint64); high = np.ascontiguousarray(high, dtype=np.uint64)
            d.filter_format, d.sd_low, d.sd_low_bits, d.sd_high, d.sd_high_bits, d.sd_wl = 1, low.ctypes.data, low_bits, high.ctypes.data, high_bits, wl
            self._keep += [low, high]
        b2c = np.ascontiguousarray(bin_to_cat, dtype=np.uint8)
        d.ncat, d.bin_to_cat = ncat, b2c.ctypes.data
        self._keep.append(b2c)
        h = vp()
        if part is not None:
            rc = L.cg_create_partitioned(device, C.byref(d), int(part[0]), int(part[1]), C.byref(h))
        else:
            rc = L.cg_create(device, C.byref(d), mem_budget, C.byref(h))
        if rc != 0:
            raise CharonGpuError(f"cg_create failed ({rc}): {L.cg_last_error(None).decode()}")
        self._h = h
        self._model_keep = None

    # ---- row-range partitioned filter -----------------------------------------------------------
    def partition(self, rank, world):
        """Keep only rows of part `rank` of `world` of a replicated filter (cg_partition_from_replicated)."""
        self._check(self._L.cg_partition_from_replicated(self._h, rank, world), "cg_partition_from_replicated")

    def export_partition(self) -> bytes:
        buf = C.create_string_buffer(64)
        self._check(self._L.cg_partition_export(self._h, buf), "cg_partition_export")
        return buf.raw

    def attach_ipc(self, peer_part: int, handle: bytes):
        self._check(self._L.cg_partition_attach_ipc(self._h, peer_part, C.create_string_buffer(handle, 64)), "cg_partition_attach_ipc")

    def attach_peer(self, peer_part: int, peer: "GpuClassifier"):
        self._check(self._L.cg_partition_attach_peer(self._h, peer_part, peer._h), "cg_partition_attach_peer")

    # ---- plumbing -----------------------------------------------------------------------------------------
    def _check(self, rc, what):
        if rc != 0:
            raise CharonGpuError(f"{what} failed ({rc}): {self._L.cg_last_error(self._h).decode()}")

    def close(self):
        if getattr(self, "_h", None):
            self._L.cg_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def info(self) -> Info:
        i = Info()
        self._check(self._L.cg_get_info(self._h, C.byref(i)), "cg_get_info")
        return i

    def set_stream(self, cuda_stream_ptr):
        self._check(self._L.cg_set_stream(self._h, cuda_stream_ptr), "cg_set_stream")

    def sync(self):
        self._check(self._L.cg_sync(self._h), "cg_sync")

    # ---- model ----------------------------------------------------------------------------------------------
    def set_model(self, opts, host_index=0, kde_pos=None, kde_neg=None, h_pos=0.1, h_neg=0.001):
        """opts: DehostOptions / ClassifyOptions.  kde_pos/kde_neg: datasets in summation order (default: the
        reference's built-in data, sorted ascending as KDEParams' ctor does, classify_stats.hpp:214-218)."""
        from .kde_defaults import default_pos_sorted, default_neg_sorted
        pos = np.ascontiguousarray(kde_pos if kde_pos is not None else default_pos_sorted(), dtype=np.float32)
        neg = np.ascontiguousarray(kde_neg if kde_neg is not None else default_neg_sorted(), dtype=np.float32)
        dist = {"kde": 0, "beta": 1, "gamma": 2}[opts.dist]
        arr = (CategoryModel * self.ncat)()
        for m in arr:
            m.dist = dist
            m.pos, m.npos, m.h_pos = pos.ctypes.data, len(pos), h_pos
            m.neg, m.nneg, m.h_neg = neg.ctypes.data, len(neg), h_neg
            m.beta_pos[:] = [6, 4]; m.beta_neg[:] = [6, 40]                      # classify_stats.hpp:268-269
            m.gamma_pos[:] = [25, 0, 0.02]; m.gamma_neg[:] = [10, 0, 0.005]      # classify_stats.hpp:266-267
        t = Thresholds(opts.min_quality, opts.min_length, opts.min_compression, opts.confidence_threshold,
                       opts.confidence_probability_threshold, opts.min_hits, opts.host_unique_prop_lo_threshold,
                       opts.min_proportion_difference, opts.min_prob_difference, host_index)
        self._model_keep = (pos, neg)
        self._check(self._L.cg_set_model(self._h, arr, C.byref(t), opts.mode), "cg_set_model")

    # ---- batches --------------------------------------------------------------------------------------------
    def _batch(self, bases, seg_off, read_seg):
        if isinstance(bases, PackedBatch):
            pb = bases
            so = np.ascontiguousarray(pb.seg_off, np.uint64)
            sl = np.ascontiguousarray(pb.seg_len, np.uint64) if pb.seg_len is not None else None
            ex = np.ascontiguousarray(pb.exc_pos, np.uint64) if pb.exc_pos is not None and len(pb.exc_pos) else None
            rs = np.ascontiguousarray(pb.read_seg if pb.read_seg is not None else read_seg, dtype=np.uint32) if (pb.read_seg is not None or read_seg is not None) else None
            n_segs = len(so) - 1
            n_reads = len(rs) - 1 if rs is not None else n_segs
            b = Batch(pb.bases.ctypes.data, so.ctypes.data, _ptr(rs), n_reads, n_segs, pb.fmt, _ptr(sl), _ptr(ex), len(ex) if ex is not None else 0)
            return b, (pb.bases, so, sl, ex, rs), n_reads
        bases = np.ascontiguousarray(bases, dtype=np.uint8)
        seg_off = np.ascontiguousarray(seg_off, dtype=np.uint64)
        rs = np.ascontiguousarray(read_seg, dtype=np.uint32) if read_seg is not None else None
        n_segs = len(seg_off) - 1
        n_reads = len(rs) - 1 if rs is not None else n_segs
        b = Batch(bases.ctypes.data, seg_off.ctypes.data, _ptr(rs), n_reads, n_segs)
        return b, (bases, seg_off, rs), n_reads

    def _outs(self, n, with_calls):
        res = dict(num_hashes=np.zeros(n, np.uint32), counts=np.zeros((n, self.ncat), np.uint32), unique_counts=np.zeros((n, self.ncat), np.uint32))
        co = CountsOut(res["num_hashes"].ctypes.data, res["counts"].ctypes.data, res["unique_counts"].ctypes.data)
        ca = None
        if with_calls:
            res.update(prob=np.zeros((n, self.ncat), np.float64), call=np.zeros(n, np.uint8), confidence=np.zeros(n, np.uint8))
            ca = CallsOut(res["prob"].ctypes.data, res["call"].ctypes.data, res["confidence"].ctypes.data)
        return res, co, ca

    @staticmethod
    def _scalars(n, mean_q, compression, length):
        mq = np.ascontiguousarray(mean_q, dtype=np.float32)
        cr = np.ascontiguousarray(compression, dtype=np.float32) if compression is not None else None    # None: computed on the device
        ln = np.ascontiguousarray(length, dtype=np.uint32)
        assert len(mq) == n and (cr is None or len(cr) == n) and len(ln) == n
        return ReadScalars(mq.ctypes.data, cr.ctypes.data if cr is not None else None, ln.ctypes.data), (mq, cr, ln)

    def count_batch(self, bases, seg_off, read_seg=None):
        b, keep, n = self._batch(bases, seg_off, read_seg)
        res, co, _ = self._outs(n, False)
        self._check(self._L.cg_count_batch(self._h, C.byref(b), C.byref(co)), "cg_count_batch")
        return res

    def classify_batch(self, bases, seg_off, mean_q, compression, length, read_seg=None):
        b, keep, n = self._batch(bases, seg_off, read_seg)
        sc, keep2 = self._scalars(n, mean_q, compression, length)
        res, co, ca = self._outs(n, True)
        self._check(self._L.cg_classify_batch(self._h, C.byref(b), C.byref(sc), C.byref(co), C.byref(ca)), "cg_classify_batch")
        return res

    # two-slot pipeline ------------------------------------------------------------------------------------------
    def submit(self, slot, bases, seg_off, mean_q=None, compression=None, length=None, read_seg=None):
        """cg_submit_batch: returns at once; the result dict is filled when wait(slot) returns."""
        b, keep, n = self._batch(bases, seg_off, read_seg)
        with_calls = mean_q is not None
        sc, keep2 = self._scalars(n, mean_q, compression, length) if with_calls else (None, None)
        res, co, ca = self._outs(n, with_calls)
        self._check(self._L.cg_submit_batch(self._h, slot, C.byref(b), C.byref(sc) if with_calls else None, C.byref(co),
                                            C.byref(ca) if with_calls else None), "cg_submit_batch")
        if not hasattr(self, "_inflight"):
            self._inflight = {}
        self._inflight[slot] = (b, keep, sc, keep2, co, ca, res)   # host buffers must outlive the wait
        return res

    def select_slot(self, slot):
        self._check(self._L.cg_select_slot(self._h, slot), "cg_select_slot")

    def set_host_pack(self, mode=-1, threads=0):
        """Transport of the bases (cg_set_host_pack): -1 automatic, 0 ASCII copies, 1 packed 4-bit ranks; threads > 0 resizes the pool."""
        self._check(self._L.cg_set_host_pack(self._h, mode, threads), "cg_set_host_pack")

    def wait(self, slot):
        self._check(self._L.cg_wait_batch(self._h, slot), "cg_wait_batch")
        return self._inflight.pop(slot)[-1] if getattr(self, "_inflight", {}).get(slot) else None

    def compression_batch(self, bases, seg_off, read_seg=None):
        """cg_compression_batch: gzip size (zlib level 6, gzip wrapper) and compression ratio of every read, computed on the GPU."""
        b, keep, n = self._batch(bases, seg_off, read_seg)
        res = dict(size=np.zeros(n, np.uint32), ratio=np.zeros(n, np.float32))
        self._check(self._L.cg_compression_batch(self._h, C.byref(b), res["size"].ctypes.data, res["ratio"].ctypes.data), "cg_compression_batch")
        return res

    def fetch_compression(self, slot, n):
        out = np.zeros(n, np.float32)
        self._check(self._L.cg_fetch_compression(self._h, slot, out.ctypes.data), "cg_fetch_compression")
        return out

    def eval_batch(self, num_hashes, counts, unique_counts, mean_q, compression, length, static=False):
        n = len(num_hashes)
        nh = np.ascontiguousarray(num_hashes, np.uint32); ct = np.ascontiguousarray(counts, np.uint32); uq = np.ascontiguousarray(unique_counts, np.uint32)
        sc, keep2 = self._scalars(n, mean_q, compression, length)
        res = dict(prob=np.zeros((n, self.ncat), np.float64), call=np.zeros(n, np.uint8), confidence=np.zeros(n, np.uint8))
        ca = CallsOut(res["prob"].ctypes.data, res["call"].ctypes.data, res["confidence"].ctypes.data)
        entry = self._L.cg_eval_batch_static if static else self._L.cg_eval_batch
        self._check(entry(self._h, n, nh.ctypes.data, ct.ctypes.data, uq.ctypes.data, C.byref(sc), C.byref(ca)), "cg_eval_batch_static" if static else "cg_eval_batch")
        return res

    def eval_batch_static(self, num_hashes, counts, unique_counts, mean_q, compression, length):
        """cg_eval_batch_static: the same results from the library's static device scratch, no allocation inside the call
        (at most EVAL_STATIC_MAX_READS reads and EVAL_STATIC_MAX_PAIRS read-category pairs; not in the reference)."""
        return self.eval_batch(num_hashes, counts, unique_counts, mean_q, compression, length, static=True)

    # staged path (device-resident inputs) -----------------------------------------------------------------------
    def stage(self, bases, seg_off, mean_q=None, compression=None, length=None, read_seg=None):
        b, keep, n = self._batch(bases, seg_off, read_seg)
        sc = None
        if mean_q is not None:
            sc, keep2 = self._scalars(n, mean_q, compression, length)
        self._check(self._L.cg_stage_batch(self._h, C.byref(b), C.byref(sc) if sc is not None else None), "cg_stage_batch")
        self.sync()
        self._staged_n = n

    def run_staged(self, with_eval=True):
        self._check(self._L.cg_run_staged(self._h, 1 if with_eval else 0), "cg_run_staged")

    def fetch_staged(self, with_calls=True):
        res, co, ca = self._outs(self._staged_n, with_calls)
        self._check(self._L.cg_fetch_staged(self._h, C.byref(co), C.byref(ca) if ca is not None else None), "cg_fetch_staged")
        return res

    def bench_gather(self, n_values, mode=0, reps=5) -> float:
        """Average ms of the random-gather microbenchmark (cg_bench_gather) on this filter."""
        ms = f32(0)
        self._check(self._L.cg_bench_gather(self._h, int(n_values), mode, reps, C.byref(ms)), "cg_bench_gather")
        return float(ms.value)

    def export_sd_vector(self):
        """(size_bits, wl, low words, low_bits, high words, high_bits) of the filter as sdsl::sd_vector parts (cg_export_sd_vector)."""
        size, wl, lb, hb = u64(0), C.c_uint8(0), u64(0), u64(0)
        self._check(self._L.cg_export_sd_vector(self._h, C.byref(size), C.byref(wl), C.byref(lb), C.byref(hb), None, None), "cg_export_sd_vector")
        low = np.zeros((lb.value + 63) // 64 + 1, np.uint64); high = np.zeros((hb.value + 63) // 64, np.uint64)
        self._check(self._L.cg_export_sd_vector(self._h, C.byref(size), C.byref(wl), C.byref(lb), C.byref(hb), low.ctypes.data, high.ctypes.data), "cg_export_sd_vector")
        return size.value, wl.value, low, lb.value, high, hb.value

    def timer_start(self):
        self._check(self._L.cg_timer_start(self._h), "cg_timer_start")

    def timer_stop(self):
        """-> (total_ms, count_kernel_ms_sum, eval_kernel_ms_sum, n_runs) measured with CUDA events on the handle's stream"""
        t, c, e, n = f32(0), f32(0), f32(0), u32(0)
        self._check(self._L.cg_timer_stop(self._h, C.byref(t), C.byref(c), C.byref(e), C.byref(n)), "cg_timer_stop")
        return t.value, c.value, e.value, n.value

    # test hooks ---------------------------------------------------------------------------------------------------
    def debug_minimisers(self, seq) -> np.ndarray:
        a = np.frombuffer(seq.encode() if isinstance(seq, str) else bytes(seq), dtype=np.uint8) if not isinstance(seq, np.ndarray) else np.ascontiguousarray(seq, np.uint8)
        cap = max(len(a), 1)
        out = np.zeros(cap, np.uint64); n = u64(0)
        self._check(self._L.cg_debug_minimisers(self._h, _ptr(a) if len(a) else None, len(a), out.ctypes.data, cap, C.byref(n)), "cg_debug_minimisers")
        return out[:n.value].copy()

    def force_generic(self, on=True):
        """0/False: default dispatch (warp-per-read + CTA-per-read for long reads); 1/True: generic any-(k,w) kernel;
        2: CTA-per-read specialised kernel for every read."""
        self._check(self._L.cg_debug_force_generic(self._h, int(on)), "cg_debug_force_generic")

    def debug_query(self, values) -> np.ndarray:
        v = np.ascontiguousarray(values, np.uint64)
        out = np.zeros((len(v), self.bin_words), np.uint64)
        self._check(self._L.cg_debug_query(self._h, v.ctypes.data, len(v), out.ctypes.data), "cg_debug_query")
        return out

    def debug_dump_filter(self) -> np.ndarray:
        out = np.zeros(self.bin_size * self.bin_words, np.uint64)
        self._check(self._L.cg_debug_dump_filter(self._h, out.ctypes.data), "cg_debug_dump_filter")
        return out
