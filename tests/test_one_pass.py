"""One-pass counting (CountParams::amb_mode): where only a few bins belong to categories that own several bins, the
warp-per-read kernel decides uniqueness during the counting pass and parks only the minimisers that hit such a bin
(ReadEntry::get_counts, reference include/read_entry.hpp:96-141, needs the best bin of every category first).  Results must
not depend on the shortcut: parked and resolved, parked beyond capacity (the read is redone in two passes) and never
parked, all against the oracle; CHARON_ONE_PASS forces the mode on (even where every bin is shared) or off."""
import numpy as np
import pytest
import oracle
from tests import common

pytestmark = pytest.mark.gpu

CASES = [
    # bins, bin -> category
    (20, list(range(18)) + [18, 18]),                  # 32-bit rows: 18 single-bin categories + one with two bins
    (40, [0, 0, 0] + list(range(1, 38))),              # 64-bit rows: one category of three bins
    (100, [i // 2 for i in range(8)] + list(range(4, 96))),   # 128-bit rows: four categories of two bins
    (255, [min(i, 253) for i in range(255)]),          # 256-bit rows: the many_category workload of bench.py
    (70, [i % 9 for i in range(70)]),                  # every bin shared: only parks when forced
]


@pytest.mark.parametrize("force", [None, "1", "0"])
@pytest.mark.parametrize("nbins,bin_cat", CASES)
def test_one_pass_counting_matches_the_oracle(monkeypatch, nbins, bin_cat, force):
    import charon_b200
    if force is None:
        monkeypatch.delenv("CHARON_ONE_PASS", raising=False)
    else:
        monkeypatch.setenv("CHARON_ONE_PASS", force)
    cats = [f"c{i}" for i in range(max(bin_cat) + 1)]
    idx, refs = common.build_oracle_index([20_000] * nbins, bin_cat, cats, seed=300 + nbins)
    gpu = common.gpu_from_oracle(idx)
    try:
        gpu.set_model(charon_b200.ClassifyOptions(), host_index=0)
        rng = np.random.default_rng(nbins)
        multi = [b for b in range(nbins) if bin_cat.count(bin_cat[b]) > 1]
        pieces = []
        for i in range(400):
            # half of the reads come from bins of multi-bin categories (every minimiser of such a read is parked: short ones
            # fit, long ones overflow and are redone), the rest from anywhere; lengths from below one window to several tiles
            b = int(rng.choice(multi)) if i % 2 else int(rng.integers(0, nbins))
            ln = int(rng.choice([30, 60, 150, 400, 900, 1500, 2500, 4000, 9000]))
            st = int(rng.integers(0, 20_000 - min(ln, 19_999)))
            piece = refs[b][st:st + ln].copy()
            if i % 7 == 0 and len(piece) > 200:                                     # a chimera: two sources in one read
                piece[len(piece) // 2:] = refs[int(rng.choice(multi))][: len(piece) - len(piece) // 2]
            pieces.append(piece)
        pieces.append(refs[multi[0]][:19_000].copy())                               # long enough for the CTA-per-read kernel
        bases, seg_off = common.ragged_batch(rng, pieces)
        n = len(pieces)
        mq, cr = common.scalars(n)
        opts = oracle.default_options(dehost=False)
        want = idx.classify_batch(bases, seg_off, mean_q=mq, compression=cr, opts=opts)
        for pack in (1, 0):
            gpu.set_host_pack(pack)
            got = gpu.classify_batch(bases, seg_off, mq, cr, np.diff(seg_off).astype(np.uint32))
            common.assert_parity(got, want, prob_rtol=2e-4)
        # paired reads: two segments per read share one entry (and one set of parked masks)
        read_seg = np.arange(0, 401, 2, dtype=np.uint32)
        bases2, seg_off2 = common.ragged_batch(rng, pieces[:400])
        want2 = idx.classify_batch(bases2, seg_off2, read_seg=read_seg, opts=opts)
        got2 = gpu.count_batch(bases2, seg_off2, read_seg=read_seg)
        common.assert_parity(got2, want2)
        assert (want["unique_counts"].sum(axis=0) > 0).sum() >= min(len(cats), 5)    # the cases really produce unique hits
    finally:
        gpu.close()
