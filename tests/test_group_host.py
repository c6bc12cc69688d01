"""CPU: host-side arithmetic of the multi-GPU handle that needs no device — LPT chromosome assignment (gg_shard_assign)
against the numpy restatement the gloo tests use — and the loud failure of gg_group_create without a GPU."""
import ctypes

import numpy as np
import pytest


def test_shard_assign_matches_the_numpy_lpt():
    from genogrove_b200 import shard, shard_assign, synth
    rng = np.random.default_rng(1)
    for n_chrom, world in ((24, 8), (24, 2), (1, 4), (7, 7), (300, 8), (5, 1)):
        counts = rng.integers(0, 10_000_000, n_chrom).astype(np.uint64)
        counts[rng.integers(0, n_chrom)] = 0
        own = shard_assign(counts, world)
        assert np.array_equal(own, shard.assign_chromosomes(counts, world))
        load = np.bincount(own, weights=counts.astype(np.float64), minlength=world)
        # LPT guarantee: no rank carries more than the average plus the largest chromosome
        assert load.max() <= counts.sum() / world + counts.max() + 1
    # hg38-shaped: 24 chromosomes on 8 GPUs stay within 10 % of the ideal share
    counts = synth._split_by_length(200_000_000, synth.HG38_LENGTHS[:24]).astype(np.uint64)
    own = shard_assign(counts, 8)
    load = np.bincount(own, weights=counts.astype(np.float64), minlength=8)
    assert load.max() / (counts.sum() / 8) < 1.10


def test_group_create_fails_loudly_without_a_gpu():
    from genogrove_b200 import _lib, DeviceGroup, GGError
    L = _lib.lib()
    if L.gg_device_count() > 0:
        pytest.skip("a CUDA device is present")
    h = ctypes.c_void_p()
    devs = (ctypes.c_int * 2)(0, 1)
    assert L.gg_group_create(ctypes.byref(h), devs, 2) == 2  # GG_ERR_CUDA: there is no CPU path
    with pytest.raises(GGError):
        DeviceGroup([0])
