"""Filter partitioned by row range across GPUs (SURVEY 8e / BASELINE config 4): every part answers bulk_contains with rows
fetched from its peers' partitions, so results must equal the replicated filter's — and the oracle's — bit for bit.
On a one-GPU box all parts live on device 0 (peer mapping degenerates to a plain pointer); as soon as two or more GPUs are
visible the parts are spread over them (part r on device r mod n_gpus: NVLink peer access) — there is no fallback to one
device then, and `test_parts_really_live_on_different_devices` fails if the spreading did not happen."""
import numpy as np
import pytest

import charon_b200
from tests import common


def _parts(idx, world, devices):
    parts = [common.gpu_from_oracle(idx, device=devices[r], part=(r, world)) for r in range(world)]
    for r, g in enumerate(parts):
        for q, peer in enumerate(parts):
            if q != r:
                g.attach_peer(q, peer)
    return parts


def _n_gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


def _devices(world):
    n = max(1, _n_gpus())
    return [r % n for r in range(world)]


@pytest.mark.gpu
def test_parts_really_live_on_different_devices():
    """With >= 2 GPUs visible the partition tests below must exercise real peer mappings."""
    n = _n_gpus()
    devs = _devices(4)
    assert len(set(devs)) == min(4, max(1, n))
    if n >= 2:
        idx, refs = common.build_oracle_index([60_000, 50_000], [0, 1], ["host", "microbial"], seed=77)
        parts = _parts(idx, 2, [0, 1])
        bases, seg_off, _ = common.fixed_length_batch(2, refs, [0.45, 0.45], 200, 2000)
        want = idx.classify_batch(bases, seg_off)
        for g in parts:
            common.assert_parity(g.count_batch(bases, seg_off), want)
        for g in parts:          # only now: a part's memory is what its peers read
            g.close()


@pytest.mark.gpu
@pytest.mark.parametrize("world,bins", [(2, 2), (3, 2), (4, 70), (8, 5)])
def test_partitioned_filter_equals_oracle(world, bins):
    lens = [40_000 + 7_000 * (b % 5) for b in range(bins)]
    ncat = min(bins, 3)
    bin_cat = [b % ncat for b in range(bins)]
    idx, refs = common.build_oracle_index(lens, bin_cat, ["host", "microbial", "other"][:ncat], seed=40 + bins)
    parts = _parts(idx, world, _devices(world))
    n, L = 300, 3000
    bases, seg_off, _ = common.fixed_length_batch(9, refs[:4], [0.2] * min(4, len(refs)), n, L)
    want = idx.classify_batch(bases, seg_off)
    for g in parts:
        info = g.info()
        assert info.filter_bytes < (idx.bin_size * info.row_bits) // 8 // world + 4096     # holds only its share
        got = g.count_batch(bases, seg_off)
        common.assert_parity(got, want)
    # membership masks straight from the partitioned query
    vals = np.random.default_rng(1).integers(0, 2**63, 5000, dtype=np.uint64)
    assert np.array_equal(parts[-1].debug_query(vals), idx.query(vals))
    for g in parts:
        g.close()


@pytest.mark.gpu
def test_unattached_part_is_an_error():
    idx, refs = common.build_oracle_index([30_000, 30_000], [0, 1], ["host", "microbial"], seed=51)
    g = common.gpu_from_oracle(idx, device=0, part=(0, 2))
    bases, seg_off, _ = common.fixed_length_batch(3, refs, [0.5, 0.5], 8, 500)
    with pytest.raises(charon_b200.CharonGpuError, match="not attached"):
        g.count_batch(bases, seg_off)
    g.close()


@pytest.mark.gpu
def test_repartition_of_a_replicated_filter_and_dehost_calls():
    idx, refs = common.build_oracle_index([120_000, 90_000], [0, 1], ["host", "microbial"], seed=52)
    world = 2
    devs = _devices(world)
    parts = [common.gpu_from_oracle(idx, device=devs[r]) for r in range(world)]
    for r, g in enumerate(parts):
        g.partition(r, world)
    for r, g in enumerate(parts):
        g.attach_peer(1 - r, parts[1 - r])
        g.set_model(charon_b200.DehostOptions(), host_index=idx.host_index)
    n, L = 400, 4000
    bases, seg_off, _ = common.fixed_length_batch(5, refs, [0.45, 0.45], n, L)
    mq, cr = common.scalars(n)
    want = idx.classify_batch(bases, seg_off, mean_q=mq, compression=cr)
    for g in parts:
        got = g.classify_batch(bases, seg_off, mq, cr, np.full(n, L, np.uint32))
        common.assert_parity(got, want)
    for g in parts:
        g.close()
