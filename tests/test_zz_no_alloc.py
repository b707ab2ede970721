"""cg_eval_batch_static — the model/call kernels on the library's static device scratch (no cudaMalloc inside the call).

The entry point has no counterpart in the reference (include/charon_gpu.h says so).
It must return what cg_eval_batch returns: calls and confidences bit-exact, probabilities identical (the same
kernels run on the same inputs; the tolerance of tests/common.assert_parity, 1e-6 relative, is the bound written here).
Two calls in a row with DIFFERENT inputs, the second smaller than the first, so that anything the shared array kept from
the first call would show.  Collected last on purpose (file name), see DESIGN.md.
"""
import numpy as np
import pytest
from tests import common


def _inputs(rng, n, ncat):
    nh = rng.integers(0, 700, n).astype(np.uint32)
    nh[: min(3, n)] = 0                                       # reads without minimisers: NaN proportions
    counts = (rng.random((n, ncat)) * (nh[:, None] + 1)).astype(np.uint32)
    uniq = (rng.random((n, ncat)) * (counts + 1)).astype(np.uint32)
    mq, cr = common.scalars(n, rng)
    length = rng.integers(100, 20_000, n).astype(np.uint32)
    return nh, counts, uniq, mq, cr, length


def _same(a, b):
    assert np.array_equal(a["call"], b["call"]), "call differs"
    assert np.array_equal(a["confidence"], b["confidence"]), "confidence differs"
    pa, pb = a["prob"], b["prob"]
    assert np.array_equal(np.isnan(pa), np.isnan(pb))
    ok = ~np.isnan(pa)
    rel = np.abs(pa[ok] - pb[ok]) / np.maximum(np.abs(pb[ok]), 1e-300)
    assert rel.size == 0 or rel.max() <= 1e-6, f"probabilities differ: max rel {rel.max()}"


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["dehost", "classify"])
def test_static_scratch_eval_matches_eval_batch_twice_in_a_row(mode):
    import charon_b200
    from charon_b200 import api
    if mode == "dehost":
        idx, refs = common.build_oracle_index([300_000, 300_000], [0, 1], ["host", "microbial"], seed=31)
    else:
        idx, refs = common.build_oracle_index([120_000] * 5, [0, 1, 2, 3, 3], ["c0", "c1", "c2", "c3"], seed=32)
    gpu = common.gpu_from_oracle(idx)
    try:
        if mode == "dehost":
            gpu.set_model(charon_b200.DehostOptions(), host_index=idx.host_index)
        else:
            gpu.set_model(charon_b200.ClassifyOptions())
        rng = np.random.default_rng(2024)
        ncat = gpu.ncat
        # set-up call (n_reads == 0): resolves the device address of the array, computes nothing
        gpu.eval_batch_static(*[x[:0] for x in _inputs(rng, 4, ncat)])
        first_n = min(api.EVAL_STATIC_MAX_READS, api.EVAL_STATIC_MAX_PAIRS // ncat)          # the largest batch the array holds
        for n in (first_n, 1237):
            args = _inputs(rng, n, ncat)
            want = gpu.eval_batch(*args)
            got = gpu.eval_batch_static(*args)
            _same(got, want)
            assert (want["call"] != 255).any() and (want["call"] == 255).any()              # both outcomes occur
        # real counts through the ordinary path, evaluated by both entry points
        bases, seg_off, _ = common.fixed_length_batch(3, refs[:2], [0.45, 0.45], 1500, 4000)
        mq, cr = common.scalars(1500)
        length = np.full(1500, 4000, np.uint32)
        c = gpu.count_batch(bases, seg_off)
        _same(gpu.eval_batch_static(c["num_hashes"], c["counts"], c["unique_counts"], mq, cr, length),
              gpu.eval_batch(c["num_hashes"], c["counts"], c["unique_counts"], mq, cr, length))
        # beyond the stated maximum: refused with the limits in the message, nothing computed
        too_many = _inputs(rng, first_n + 1, ncat)
        with pytest.raises(charon_b200.CharonGpuError, match="exceed the static scratch"):
            gpu.eval_batch_static(*too_many)
    finally:
        gpu.close()
