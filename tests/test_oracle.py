"""CPU: pin the oracle (oracle/gg_oracle.c) against the reference's golden vectors (tests/golden, generated from
the unmodified reference) and, when oracle/_ref is present, against the reference itself on random inputs."""
import numpy as np
import pytest

import oracle as O
from helpers import ChainReplay, load_golden

G = load_golden()


def test_golden_file_shape():
    assert len(G["cases"]) >= 100 and sum(len(c["queries"]) for c in G["cases"]) >= 2000


@pytest.mark.parametrize("case", G["cases"], ids=[c["name"] for c in G["cases"]])
def test_oracle_matches_reference_golden(case):
    rp = ChainReplay(case["order"])
    for op in case["ops"]:
        rp.apply(op)
    if not case["queries"]:
        return
    assert rp.query_ids(case["queries"]) == case["expected"]


def _pure(case, kind):
    return len(case["ops"]) == 1 and case["ops"][0]["op"] == kind and case["queries"]


@pytest.mark.parametrize("case", [c for c in G["cases"] if _pure(c, "regular") or _pure(c, "sorted")],
                         ids=lambda c: c["name"])
def test_c_chain_emulation_matches_golden(case):
    """orc_chain_regular / orc_chain_sorted_append (the C restatements of the insert paths' leaf-chain order)."""
    items = np.array(case["ops"][0]["items"], dtype=np.uint64).reshape(-1, 3)
    s, e, ids = items[:, 0], items[:, 1], items[:, 2].astype(np.int64)
    perm = O.chain_regular(s, e) if case["ops"][0]["op"] == "regular" else O.chain_sorted_append(s, e, case["order"])
    q = np.array(case["queries"], dtype=np.uint64).reshape(-1, 3)
    qc = np.where(q[:, 0] == case["ops"][0]["chrom"], 0, O.NO_CHROM).astype(np.uint32)
    off, ranks = O.flat_intersect_batch(s[perm], e[perm], [0, len(s)], qc, q[:, 1], q[:, 2])
    got = [[int(ids[perm[int(r)]]) for r in ranks[int(off[k]):int(off[k + 1])]] for k in range(len(q))]
    assert got == case["expected"]


@pytest.mark.parametrize("case", [c for c in G["cases"] if _pure(c, "bulk_sorted")], ids=lambda c: c["name"])
def test_c_bptree_restatement_matches_golden(case):
    """orc_tree_build_bottom_up + orc_tree_search_overlaps: same node shapes, same descent + leaf walk."""
    items = np.array(case["ops"][0]["items"], dtype=np.uint64).reshape(-1, 3)
    t = O.OracleTree(items[:, 0], items[:, 1], case["order"])
    for (c, qs, qe), exp in zip(case["queries"], case["expected"]):
        got = [int(items[int(r), 2]) for r in t.intersect(qs, qe)] if c == case["ops"][0]["chrom"] else []
        assert got == exp


def test_interval_arithmetic_known_answers():
    L = O.oracle_lib()
    for row in G["interval_arithmetic"]:
        a, b = row["a"], row["b"]
        assert L.orc_overlaps(*a, *b) == row["overlaps"]
        assert L.orc_overlaps(*b, *a) == row["overlaps_rev"]
        assert L.orc_less(*a, *b) == row["less"]
        assert L.orc_greater(*a, *b) == row["greater"]


def test_distribute_evenly_known_answers():
    for row in G["distribute_evenly"]:
        assert list(O.distribute_evenly(row["total"], row["max"])) == row["result"]
    assert O.distribute_evenly(3_000_000_000, 100) == (30_000_000, 100, 0)  # bulk_distribution_test.cpp:66-73


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built (no reference checkout on this machine)")
@pytest.mark.parametrize("order", [3, 4, 32])
def test_oracle_vs_live_reference_random(order):
    rng = np.random.default_rng(order)
    n, nq = 1500, 400
    s = (rng.integers(0, 20000, n) // 16 * 16).astype(np.uint64)
    ln = rng.integers(0, 300, n).astype(np.uint64)
    ln[rng.random(n) < 0.05] = 5000
    e = s + ln
    qs = rng.integers(0, 21000, nq).astype(np.uint64)
    qe = qs + rng.integers(0, 1500, nq).astype(np.uint64)
    zc = np.zeros(nq, np.uint32)
    # regular inserts
    g = O.RefGrove(order)
    for i in range(n):
        g.insert(0, int(s[i]), int(e[i]), i)
    perm = O.chain_regular(s, e)
    off, ranks = O.flat_intersect_batch(s[perm], e[perm], [0, n], zc, qs, qe)
    roff, rids, _ = g.intersect_batch(zc, qs, qe)
    assert np.array_equal(off, roff) and np.array_equal(perm[ranks], rids)
    # sorted single appends
    p0 = O.chain_stable_sort(s, e)
    ss, se = s[p0], e[p0]
    g = O.RefGrove(order)
    for i in range(n):
        g.insert(0, int(ss[i]), int(se[i]), i, O.RefGrove.SORTED)
    perm = O.chain_sorted_append(ss, se, order)
    off, ranks = O.flat_intersect_batch(ss[perm], se[perm], [0, n], zc, qs, qe)
    roff, rids, _ = g.intersect_batch(zc, qs, qe)
    assert np.array_equal(off, roff) and np.array_equal(perm[ranks], rids)
    # sorted bulk: tree restatement + flat oracle
    g = O.RefGrove(order)
    g.insert_bulk(0, ss, se, np.arange(n, dtype=np.uint32), presorted=True)
    roff, rids, _ = g.intersect_batch(zc, qs, qe, threads=2)
    off, ranks = O.flat_intersect_batch(ss, se, [0, n], zc, qs, qe)
    assert np.array_equal(off, roff) and np.array_equal(ranks, rids)
    t = O.OracleTree(ss, se, order)
    for k in range(0, nq, 7):
        assert np.array_equal(t.intersect(int(qs[k]), int(qe[k])).astype(np.uint32), rids[int(roff[k]):int(roff[k + 1])])


def test_invalid_inputs_match_reference_errors():
    assert not O.OracleTree([1], [2], 2)._t  # order < 3 builds nothing (grove.hpp:158-162)
    if O.ref_available():
        with pytest.raises(ValueError):
            O.RefGrove(2)
        g = O.RefGrove(3)
        with pytest.raises(ValueError):
            g.insert(0, 5, 4, 0)
