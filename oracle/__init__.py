"""ctypes binding of oracle/liboracle.so — the CPU ORACLE (test infrastructure; parity unpinned, see oracle.hpp).

Only tests/, bench.py's cpu_baseline / --impl reference leg and __graft_entry__.smoke() may import this.
The product package (charon_b200) never does.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")
SEED = 0x8F3F73B5CF1C9ADE


class Options(C.Structure):
    _fields_ = [("dehost", C.c_int32), ("dist", C.c_int32), ("lo_hi_threshold", C.c_float), ("num_reads_to_fit", C.c_uint32),
                ("min_quality", C.c_float), ("min_length", C.c_uint32), ("min_compression", C.c_float),
                ("confidence_threshold", C.c_int32), ("confidence_probability_threshold", C.c_float),
                ("host_unique_prop_lo_threshold", C.c_float), ("min_proportion_difference", C.c_float),
                ("min_prob_difference", C.c_float), ("min_hits", C.c_uint32), ("run_extract", C.c_int32)]


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


_lib_cache = None


def lib():
    global _lib_cache
    if _lib_cache is not None:
        return _lib_cache
    if not os.path.exists(_SO):
        build()
    L = C.CDLL(_SO)
    vp, u64, u32, i32, f32, f64 = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int32, C.c_float, C.c_double
    L.orc_default_options.argtypes = [i32, C.POINTER(Options)]
    L.orc_minimisers.argtypes = [vp, u64, u32, u32, u64, u32, vp, u64]; L.orc_minimisers.restype = u64
    L.orc_kmer_values.argtypes = [vp, u64, u32, u64, vp, u64]; L.orc_kmer_values.restype = u64
    L.orc_ibf_seed.argtypes = [u32]; L.orc_ibf_seed.restype = u64
    L.orc_compression_ratio.argtypes = [vp, u64]; L.orc_compression_ratio.restype = f32
    L.orc_model_prob.argtypes = [C.POINTER(Options), f32]; L.orc_model_prob.restype = f64
    L.orc_bin_size_in_bits.argtypes = [u64]; L.orc_bin_size_in_bits.restype = u64
    L.orc_kde_defaults.argtypes = [vp, vp]
    L.orc_index_build.argtypes = [vp, vp, vp, u32, u32, vp, C.c_char_p, u32, u32, i32]; L.orc_index_build.restype = vp
    L.orc_index_load.argtypes = [C.c_char_p]; L.orc_index_load.restype = vp
    L.orc_index_save.argtypes = [vp, C.c_char_p]; L.orc_index_save.restype = i32
    L.orc_index_free.argtypes = [vp]
    L.orc_index_header.argtypes = [vp, vp]
    L.orc_index_rows.argtypes = [vp]; L.orc_index_rows.restype = vp
    L.orc_index_bin_to_cat.argtypes = [vp]; L.orc_index_bin_to_cat.restype = vp
    L.orc_index_category_name.argtypes = [vp, u32, C.c_char_p, u32]; L.orc_index_category_name.restype = i32
    L.orc_index_host_category.argtypes = [vp]; L.orc_index_host_category.restype = i32
    L.orc_index_query.argtypes = [vp, vp, u64, vp]
    L.orc_classify_batch.argtypes = [vp, C.POINTER(Options), vp, vp, vp, u32, vp, vp, i32, i32, vp, vp, vp, vp, vp, vp, vp]
    L.orc_classify_batch.restype = f64
    L.orc_run_tsv.argtypes = [vp, C.POINTER(Options), C.c_char_p, vp, vp, vp, u32, vp, vp]; L.orc_run_tsv.restype = vp
    L.orc_run_like_reference.argtypes = [vp, C.POINTER(Options), vp, u64, i32, u32, C.c_char_p, vp]; L.orc_run_like_reference.restype = C.c_longlong
    L.orc_free.argtypes = [vp]
    L.orc_max_threads.restype = i32
    _lib_cache = L
    return L


def default_options(dehost=True) -> Options:
    o = Options()
    lib().orc_default_options(1 if dehost else 0, C.byref(o))
    return o


def _bytes(seq):
    if isinstance(seq, str):
        seq = seq.encode()
    return np.frombuffer(bytes(seq), dtype=np.uint8) if not isinstance(seq, np.ndarray) else np.ascontiguousarray(seq, dtype=np.uint8)


def minimisers(seq, k=19, w=41, seed=SEED, sigma=5) -> np.ndarray:
    a = _bytes(seq)
    out = np.empty(max(len(a), 1), dtype=np.uint64)
    n = lib().orc_minimisers(a.ctypes.data, len(a), k, w, seed, sigma, out.ctypes.data, len(out))
    return out[:n].copy()


def kmer_values(seq, k=19, seed=SEED) -> np.ndarray:
    a = _bytes(seq)
    out = np.empty(max(len(a), 1), dtype=np.uint64)
    n = lib().orc_kmer_values(a.ctypes.data, len(a), k, seed, out.ctypes.data, len(out))
    return out[:n].copy()


def compression_ratio(seq) -> float:
    a = _bytes(seq)
    return float(lib().orc_compression_ratio(a.ctypes.data, len(a)))


def model_prob(x: float, opts: Options = None) -> float:
    o = opts or default_options(True)
    return float(lib().orc_model_prob(C.byref(o), C.c_float(x)))


def kde_defaults():
    pos = np.empty(500, dtype=np.float32); neg = np.empty(500, dtype=np.float32)
    lib().orc_kde_defaults(pos.ctypes.data, neg.ctypes.data)
    return pos, neg


class Index:
    """An in-memory Charon index held by the oracle (orc::Index)."""

    def __init__(self, handle):
        if not handle:
            raise RuntimeError("oracle index handle is NULL")
        self._h = handle
        hdr = np.zeros(9, dtype=np.uint64)
        lib().orc_index_header(self._h, hdr.ctypes.data)
        (self.k, self.w, self.bins, self.technical_bins, self.bin_size, self.hash_shift, self.bin_words, self.hash_funs,
         self.ncat) = [int(x) for x in hdr]
        buf = C.create_string_buffer(4096)
        self.categories = []
        for i in range(self.ncat):
            lib().orc_index_category_name(self._h, i, buf, 4096)
            self.categories.append(buf.value.decode())
        p = lib().orc_index_bin_to_cat(self._h)
        self.bin_to_cat = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(self.bins,)).copy()
        self.host_index = lib().orc_index_host_category(self._h)

    @classmethod
    def build(cls, records, rec_bin, bin_cat, categories, k=19, w=41, threads=0):
        """records: list of uint8 arrays (one per reference record); rec_bin[i]: bin of record i;
        bin_cat[b]: category index of bin b; categories: names in index order."""
        cat = np.ascontiguousarray(np.concatenate(records))
        off = np.zeros(len(records) + 1, dtype=np.uint64); off[1:] = np.cumsum([len(r) for r in records])
        rb = np.asarray(rec_bin, dtype=np.uint32); bc = np.asarray(bin_cat, dtype=np.uint32)
        threads = threads or lib().orc_max_threads()
        h = lib().orc_index_build(cat.ctypes.data, off.ctypes.data, rb.ctypes.data, len(records), len(bc), bc.ctypes.data,
                                  "\n".join(categories).encode(), k, w, threads)
        return cls(h)

    @classmethod
    def load(cls, path):
        return cls(lib().orc_index_load(str(path).encode()))

    def save(self, path):
        if lib().orc_index_save(self._h, str(path).encode()) != 0:
            raise RuntimeError("index save failed")

    def rows(self) -> np.ndarray:
        """The plain bit matrix: bin_size * bin_words uint64 (view into oracle memory)."""
        p = lib().orc_index_rows(self._h)
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint64)), shape=(self.bin_size * self.bin_words,))

    def query(self, values) -> np.ndarray:
        v = np.ascontiguousarray(values, dtype=np.uint64)
        out = np.empty((len(v), self.bin_words), dtype=np.uint64)
        lib().orc_index_query(self._h, v.ctypes.data, len(v), out.ctypes.data)
        return out

    def classify_batch(self, bases, seg_off, read_seg=None, mean_q=None, compression=None, opts=None, with_gzip=False, threads=0):
        """Loop body of dehost_reads/classify_reads for a batch.  Returns dict of arrays + 'seconds'."""
        o = opts or default_options(True)
        bases = np.ascontiguousarray(bases, dtype=np.uint8); seg_off = np.ascontiguousarray(seg_off, dtype=np.uint64)
        n = (len(read_seg) - 1) if read_seg is not None else (len(seg_off) - 1)
        rs = np.ascontiguousarray(read_seg, dtype=np.uint32) if read_seg is not None else None
        mq = np.ascontiguousarray(mean_q, dtype=np.float32) if mean_q is not None else np.full(n, 20.0, np.float32)
        cr = np.ascontiguousarray(compression, dtype=np.float32) if compression is not None else np.full(n, 0.3, np.float32)
        res = dict(num_hashes=np.zeros(n, np.uint32), counts=np.zeros((n, self.ncat), np.uint32),
                   unique_counts=np.zeros((n, self.ncat), np.uint32), prob=np.zeros((n, self.ncat), np.float64),
                   call=np.zeros(n, np.uint8), confidence=np.zeros(n, np.uint8), compression=np.zeros(n, np.float32))
        threads = threads or lib().orc_max_threads()
        sec = lib().orc_classify_batch(self._h, C.byref(o), bases.ctypes.data, seg_off.ctypes.data, rs.ctypes.data if rs is not None else None,
                                       n, mq.ctypes.data, cr.ctypes.data, 1 if with_gzip else 0, threads, res["num_hashes"].ctypes.data,
                                       res["counts"].ctypes.data, res["unique_counts"].ctypes.data, res["prob"].ctypes.data,
                                       res["call"].ctypes.data, res["confidence"].ctypes.data, res["compression"].ctypes.data)
        res["seconds"] = sec
        res["mean_q"] = mq
        return res

    def run_tsv(self, ids, bases, seg_off, mean_q, compression, read_seg=None, opts=None) -> str:
        o = opts or default_options(True)
        bases = np.ascontiguousarray(bases, dtype=np.uint8); seg_off = np.ascontiguousarray(seg_off, dtype=np.uint64)
        rs = np.ascontiguousarray(read_seg, dtype=np.uint32) if read_seg is not None else None
        mq = np.ascontiguousarray(mean_q, dtype=np.float32); cr = np.ascontiguousarray(compression, dtype=np.float32)
        p = lib().orc_run_tsv(self._h, C.byref(o), "\n".join(ids).encode(), bases.ctypes.data, seg_off.ctypes.data,
                              rs.ctypes.data if rs is not None else None, len(ids), mq.ctypes.data, cr.ctypes.data)
        s = C.string_at(p).decode()
        lib().orc_free(p)
        return s

    def run_like_reference(self, fastq: bytes, threads=1, chunk=100, tsv_path=None, opts=None):
        """`charon dehost|classify -t threads` restated over an in-memory FASTQ (orc_run_like_reference): parse, chunks of
        `chunk` reads, OpenMP over a chunk, one critical section for model + call + print.  Returns a dict of stage times."""
        o = opts or default_options(True)
        buf = np.frombuffer(fastq, np.uint8)
        st = np.zeros(5, np.float64)
        n = lib().orc_run_like_reference(self._h, C.byref(o), buf.ctypes.data, len(buf), threads, chunk,
                                         tsv_path.encode() if tsv_path else None, st.ctypes.data)
        if n < 0:
            raise RuntimeError(f"orc_run_like_reference failed ({n})")
        return dict(reads=int(n), threads=threads, parse_s=float(st[0]), loop_wall_s=float(st[1]), gzip_cpu_s=float(st[2]),
                    hot_path_cpu_s=float(st[3]), model_print_cpu_s=float(st[4]))

    def __del__(self):
        try:
            lib().orc_index_free(self._h)
        except Exception:
            pass
