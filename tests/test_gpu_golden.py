"""GPU: the reference's own hot-path tests (restated as golden cases from the unmodified reference,
tests/golden/reference_cases.json) run through the Grove mirror -> C ABI -> CUDA kernels.  Expected id lists are
the reference's, order included."""
import numpy as np
import pytest

from helpers import load_golden

pytestmark = pytest.mark.gpu
G = load_golden()


def _replay(case):
    from genogrove_b200 import Grove, Interval
    g = Grove(case["order"])
    by_id = {}
    name = lambda c: f"chr{c}"  # noqa: E731
    for op in case["ops"]:
        idx = name(op["chrom"])
        if op["op"] in ("regular", "sorted"):
            for s, e, i in op["items"]:
                by_id[i] = g.insert_data(idx, Interval(s, e), i, sorted=op["op"] == "sorted")
        elif op["op"] in ("bulk", "bulk_sorted"):
            keys = g.insert_data_bulk(idx, [(Interval(s, e), i) for s, e, i in op["items"]], presorted=op["op"] == "bulk_sorted")
            for k in keys:
                by_id[k.get_data()] = k
        elif op["op"] == "remove":
            for i in op["ids"]:
                assert g.remove_key(idx, by_id[i])
    return g, name


@pytest.mark.parametrize("case", [c for c in G["cases"] if c["queries"]], ids=lambda c: c["name"])
def test_grove_mirror_matches_reference_golden(case):
    from genogrove_b200 import Interval
    g, name = _replay(case)
    # one intersect(query, index) per query, as the reference's tests do
    for (c, qs, qe), exp in list(zip(case["queries"], case["expected"]))[:40]:
        res = g.intersect(Interval(qs, qe), name(c))
        assert [k.get_data() for k in res.get_keys()] == exp
        assert res.get_query() == Interval(qs, qe)
    # and all queries in one batch call
    q = np.array(case["queries"], dtype=np.uint64).reshape(-1, 3)
    cids = np.array([g.chrom_id(name(int(c))) for c in q[:, 0]], dtype=np.uint32)
    off, hits = g.intersect_batch(cids, q[:, 1], q[:, 2])
    for k, exp in enumerate(case["expected"]):
        got = [key.get_data() for key in g.keys_at(int(cids[k]), hits[int(off[k]):int(off[k + 1])])] if exp else []
        assert got == exp and int(off[k + 1] - off[k]) == len(exp)


def test_reference_api_error_behaviour():
    from genogrove_b200 import Grove, Interval
    with pytest.raises(ValueError):
        Grove(2)  # grove.hpp:158-162
    g = Grove()
    assert g.get_order() == 3
    with pytest.raises(ValueError):
        Interval(3, 2)
    assert g.intersect(Interval(0, 10), "chr1").get_keys() == []  # empty grove, unknown index
    k = g.insert_data("chr1", Interval(10, 20), 1)
    assert g.intersect(Interval(10, 20), "chrX").get_keys() == []  # precondition_test.cpp:105-111
    assert g.intersect(Interval(20, 30), "chr1").get_keys() == [k]
    assert g.remove_key("chr1", k) and not g.remove_key("chr1", k) and not g.remove_key("chr1", None)
    assert g.intersect(Interval(10, 20), "chr1").get_keys() == []


def test_intersect_all_indices():
    from genogrove_b200 import Grove, Interval
    g = Grove(4)
    k1 = [g.insert_data("chr1", Interval(i * 10, i * 10 + 5), i, sorted=True) for i in range(5)]
    k2 = [g.insert_data("chr2", Interval(i * 10, i * 10 + 5), 100 + i, sorted=True) for i in range(5)]
    g.remove_key("chr1", k1[2])  # grove_remove_test.cpp:606-625
    res = g.intersect(Interval(20, 25))
    assert k1[2] not in res.get_keys() and k2[2] in res.get_keys()
    assert sorted(k.get_data() for k in g.intersect(Interval(0, 60)).get_keys()) == [0, 1, 3, 4, 100, 101, 102, 103, 104]
