"""CPU: host-side logic — synthetic workloads, sharding helpers, the interval/key mirror types."""
import numpy as np
import pytest

from genogrove_b200 import Interval, Key, QueryResult, shard, synth


def test_interval_mirror_semantics():
    # interval.hpp:70-74, 87-105
    with pytest.raises(ValueError):
        Interval(5, 4)
    assert Interval(1, 5) < Interval(1, 6) < Interval(2, 0 + 2)
    assert Interval(3, 9) > Interval(3, 8)
    assert Interval(0, 0) == Interval(0, 0)
    assert Interval().get_start() == 2**64 - 1 and Interval().get_end() == 2**64 - 1
    assert Interval(10, 20).to_string() == "[10,20]"
    r = QueryResult(Interval(1, 2))
    with pytest.raises(ValueError):
        r.add_key(None)  # query_result.hpp:108-111
    k = Key(Interval(1, 2), "x")
    r.add_key(k)
    assert r.get_keys() == [k] and r.get_query() == Interval(1, 2) and k.get_data() == "x"


def test_synth_is_deterministic_and_valid():
    a, b = synth.cfg2(5000, 3000), synth.cfg2(5000, 3000)
    for x, y in ((a.start, b.start), (a.end, b.end), (a.chrom, b.chrom), (a.q_start, b.q_start)):
        assert np.array_equal(x, y)
    for w in (synth.cfg1(2000, 500), a, synth.cfg3(300, 2000), synth.cfg4(5000, 20000)):
        assert (w.start <= w.end).all() and (w.q_start <= w.q_end).all()
        assert w.chrom.max() < w.n_chrom and w.q_chrom.max() < w.n_chrom
        assert w.end.max() < 2**32
        L = synth.HG38_LENGTHS[w.chrom.astype(np.int64)]
        assert (w.end.astype(np.int64) < L).all()
    w = synth.cfg4(5000, 20000)
    key = w.q_chrom.astype(np.int64) * 2**32 + w.q_start.astype(np.int64)
    assert (np.diff(key) >= 0).all(), "cfg4 queries are coordinate-sorted per chromosome"
    assert ((w.q_end - w.q_start) == 149).all()
    w3 = synth.cfg3(300, 100)
    pairs = np.stack([w3.chrom.astype(np.uint64), w3.start, w3.end], 1)
    assert len(np.unique(pairs, axis=0)) < len(pairs), "cfg3 must contain identical (start,end) features"


def test_query_shard_partitions_the_batch():
    for nq in (0, 1, 7, 100, 1_000_003):
        for world in (1, 2, 3, 8):
            spans = [shard.query_shard(nq, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == nq
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def test_lpt_assignment_balances_hg38():
    counts = synth._split_by_length(200_000_000, synth.HG38_LENGTHS)
    owner = shard.assign_chromosomes(counts, 8)
    load = np.bincount(owner, weights=counts, minlength=8)
    assert load.min() > 0 and load.max() / load.mean() < 1.08
    local = shard.localize_chromosomes(owner, 3)
    mine = np.nonzero(owner == 3)[0]
    assert list(local[mine]) == list(range(len(mine))) and (np.delete(local, mine) == 0xFFFFFFFF).all()


def test_route_and_merge_keep_query_order():
    rng = np.random.default_rng(0)
    owner = np.array([0, 1, 0, 1, 1])
    qc = rng.integers(0, 7, 1000).astype(np.uint32)  # ids 5, 6 are unknown chromosomes
    parts = []
    for r in range(2):
        qi = shard.route_queries(qc, owner, r)
        assert (np.diff(qi) > 0).all() and (owner[qc[qi]] == r).all()
        cnt = (qi % 3).astype(np.uint64)
        off = np.concatenate([[0], np.cumsum(cnt)]).astype(np.uint64)
        hits = np.repeat(qi, cnt.astype(np.int64)).astype(np.uint32)
        parts.append((qi, off, hits))
    off, hits = shard.merge_routed_results(len(qc), parts)
    for q in range(len(qc)):
        exp = (q % 3) if qc[q] < 5 else 0
        assert off[q + 1] - off[q] == exp
        assert (hits[int(off[q]):int(off[q + 1])] == q).all()
    assert shard.global_offset_base([5, 7, 9], 2) == 12 and shard.gather_hit_totals(11) == [11]


def test_sorted_slice_is_a_partition_of_one_sorted_batch():
    """synth.hg38_sorted_slice: contiguous parts of ONE coordinate-sorted batch (what bench.py --mode cfg5 hands the ranks):
    the parts tile the batch, stay sorted across part boundaries, and follow the per-chromosome apportionment."""
    from genogrove_b200 import synth
    q = 200_003
    cuts = [0, 1, 777, 25_000, 25_001, 130_000, q]
    parts = [synth.hg38_sorted_slice(q, a, b, 150, 5) for a, b in zip(cuts[:-1], cuts[1:])]
    c = np.concatenate([p[0] for p in parts])
    s = np.concatenate([p[1] for p in parts]).astype(np.int64)
    e = np.concatenate([p[2] for p in parts]).astype(np.int64)
    assert len(c) == q and np.array_equal(e - s, np.full(q, 149))
    assert np.array_equal(np.bincount(c, minlength=24), synth._split_by_length(q, synth.HG38_LENGTHS))
    assert np.all(np.diff(c.astype(np.int64)) >= 0)
    same = np.diff(c.astype(np.int64)) == 0
    assert np.all(np.diff(s)[same] >= 0)
    assert np.all(s >= 0) and np.all(e < synth.HG38_LENGTHS[c])
    again = synth.hg38_sorted_slice(q, 777, 25_000, 150, 5)
    assert np.array_equal(again[1], parts[2][1])  # deterministic
