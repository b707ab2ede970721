"""ctypes binding of tools/libcharon_synth.so: seeded synthetic references and reads."""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libcharon_synth.so")


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


def _lib():
    if not os.path.exists(_SO):
        build()
    lib = C.CDLL(_SO)
    lib.synth_reference.argtypes = [C.c_uint64, C.c_uint64, C.c_void_p]
    lib.synth_reads.argtypes = [C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_uint32,
                                C.c_double, C.c_void_p, C.c_void_p, C.c_uint64]
    return lib


def reference(seed: int, n: int) -> np.ndarray:
    """n iid uniform ACGT bases as a uint8 array of ASCII codes."""
    out = np.empty(n, dtype=np.uint8)
    _lib().synth_reference(seed, n, out.ctypes.data)
    return out


def reads(seed: int, refs, fracs, n_reads: int, length: int, err: float = 0.02, first_index: int = 0, out=None):
    """n_reads reads of `length` bases.  refs: list of uint8 arrays; fracs: source probability per ref (the
    remainder is pure random sequence).  Returns (bases[n_reads*length] uint8, src[n_reads] uint8)."""
    cat = np.ascontiguousarray(np.concatenate(refs)) if len(refs) > 1 else np.ascontiguousarray(refs[0])
    off = np.zeros(len(refs) + 1, dtype=np.uint64)
    off[1:] = np.cumsum([len(r) for r in refs])
    fr = np.asarray(fracs, dtype=np.float64)
    if out is None:
        out = np.empty(n_reads * length, dtype=np.uint8)
    src = np.empty(n_reads, dtype=np.uint8)
    _lib().synth_reads(seed, cat.ctypes.data, off.ctypes.data, len(refs), fr.ctypes.data, n_reads, length, err,
                       out.ctypes.data, src.ctypes.data, first_index)
    return out, src
