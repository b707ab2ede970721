"""Shared builders for the parity tests: synthetic references/reads (tools/synth), the oracle index, and the GPU
classifier constructed from the very same filter bits."""
import numpy as np
import oracle
from tools import synth


def build_oracle_index(ref_lens, bin_cat, categories, seed=1, k=19, w=41, threads=0):
    refs = [synth.reference(seed + i, n) for i, n in enumerate(ref_lens)]
    idx = oracle.Index.build(refs, rec_bin=list(range(len(refs))), bin_cat=bin_cat, categories=categories, k=k, w=w, threads=threads)
    return idx, refs


def gpu_from_oracle(idx, device=0, **kw):
    import charon_b200
    return charon_b200.GpuClassifier(k=idx.k, w=idx.w, bins=idx.bins, bin_size=idx.bin_size, hash_funs=idx.hash_funs,
                                     bin_to_cat=idx.bin_to_cat, ncat=idx.ncat, rows=idx.rows(), device=device, **kw)


def fixed_length_batch(seed, refs, fracs, n_reads, length, err=0.02):
    bases, src = synth.reads(seed, refs, fracs, n_reads, length, err)
    seg_off = (np.arange(n_reads + 1, dtype=np.uint64) * np.uint64(length))
    return bases, seg_off, src


def ragged_batch(rng, pieces):
    """pieces: list of uint8 arrays -> (bases, seg_off)"""
    seg_off = np.zeros(len(pieces) + 1, dtype=np.uint64)
    seg_off[1:] = np.cumsum([len(p) for p in pieces])
    bases = np.concatenate(pieces) if pieces and seg_off[-1] > 0 else np.zeros(0, np.uint8)
    return np.ascontiguousarray(bases, dtype=np.uint8), seg_off


def scalars(n, rng=None):
    if rng is None:
        return np.full(n, 20.0, np.float32), np.full(n, 0.3, np.float32)
    return rng.uniform(8, 30, n).astype(np.float32), rng.uniform(0.05, 0.5, n).astype(np.float32)


def assert_parity(gpu_res, orc_res, prob_rtol=1e-6):
    for key in ("num_hashes", "counts", "unique_counts"):
        assert np.array_equal(gpu_res[key], orc_res[key]), key + " differs: first at " + str(np.argwhere(gpu_res[key] != orc_res[key])[:5])
    if "call" in gpu_res:
        assert np.array_equal(gpu_res["confidence"], orc_res["confidence"]), "confidence differs"
        assert np.array_equal(gpu_res["call"], orc_res["call"]), "call differs"
        a, b = gpu_res["prob"], orc_res["prob"]
        assert np.array_equal(np.isnan(a), np.isnan(b))
        ok = ~np.isnan(a)
        denom = np.maximum(np.abs(b[ok]), 1e-300)
        rel = np.abs(a[ok] - b[ok]) / denom
        assert rel.size == 0 or rel.max() <= prob_rtol, f"probabilities differ: max rel {rel.max()}"
