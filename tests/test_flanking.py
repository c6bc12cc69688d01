"""Flanking (nearest predecessor / successor): the oracle pinned against the unmodified reference's golden answers
(CPU), and the CUDA path (C ABI -> flanking_kernel) against the golden cases and the oracle (GPU).
Reference: grove::flanking -> detail::search_flanking, include/genogrove/structure/grove/query_engine.hpp:238-311;
cases restate tests/structure/grove_flanking_test.cpp:20-117."""
import json
import os

import numpy as np
import pytest

import oracle as O
from helpers import ChainReplay, chrom_offsets, expected_chain, random_index, random_queries

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "flanking_cases.json")
with open(GOLD) as f:
    F = json.load(f)
NONE = O.NO_RANK


def _oracle_ids(case):
    """Replay the ops into chains with the oracle's insert-path rules, answer with the flat flanking oracle."""
    rp = ChainReplay(case["order"])
    for op in case["ops"]:
        rp.apply(op)
    n_chrom, chrom, start, end, ids = rp.arrays()
    coff = chrom_offsets(chrom, n_chrom)
    q = np.array(case["queries"], dtype=np.uint64).reshape(-1, 3)
    pred, succ = O.flat_flanking_batch(start, end, coff, q[:, 0].astype(np.uint32), q[:, 1], q[:, 2])

    def to_id(c, r):
        return None if int(r) == NONE else int(ids[int(coff[int(c)]) + int(r)])

    return [to_id(c, r) for c, r in zip(q[:, 0], pred)], [to_id(c, r) for c, r in zip(q[:, 0], succ)]


def test_flanking_golden_shape():
    assert len(F["cases"]) >= 20 and sum(len(c["queries"]) for c in F["cases"]) >= 900


@pytest.mark.parametrize("case", F["cases"], ids=[c["name"] for c in F["cases"]])
def test_flat_flanking_oracle_matches_reference_golden(case):
    pred, succ = _oracle_ids(case)
    assert pred == case["pred"]
    assert succ == case["succ"]


@pytest.mark.parametrize("case", [c for c in F["cases"] if len(c["ops"]) == 1 and c["ops"][0]["op"] == "bulk_sorted"],
                         ids=lambda c: c["name"])
def test_tree_flanking_restatement_matches_golden(case):
    """orc_tree_search_flanking: the reference's pruned depth-first search on the reference's node shapes."""
    items = np.array(case["ops"][0]["items"], dtype=np.uint64).reshape(-1, 3)
    t = O.OracleTree(items[:, 0], items[:, 1], case["order"])
    for (c, qs, qe), ep, es in zip(case["queries"], case["pred"], case["succ"]):
        if c != case["ops"][0]["chrom"]:
            continue
        p, s = t.flanking(qs, qe)
        assert (None if p is None else int(items[p, 2])) == ep
        assert (None if s is None else int(items[s, 2])) == es


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built")
def test_flat_flanking_oracle_matches_reference_random():
    rng = np.random.default_rng(99)
    for trial in range(6):
        n = int(rng.integers(1, 500))
        chrom, s, e = random_index(rng, n, 1, span=3000, max_len=300, dup=trial % 2 == 0, nested=trial % 3 == 0)
        perm = O.chain_stable_sort(s, e)
        g = O.RefGrove(int(rng.choice([3, 5, 32])))
        g.insert_bulk(0, s[perm], e[perm], perm.astype(np.uint32), presorted=True)
        qc, qs, qe = random_queries(rng, 400, 1, span=3000, max_len=200)
        rp, rs = g.flanking_batch(qc, qs, qe)
        fp, fs = O.flat_flanking_batch(s[perm], e[perm], [0, n], qc, qs, qe)
        assert np.array_equal(rp, np.where(fp == NONE, NONE, perm[np.minimum(fp, n - 1)]))
        assert np.array_equal(rs, np.where(fs == NONE, NONE, perm[np.minimum(fs, n - 1)]))


# ----------------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
@pytest.mark.parametrize("case", F["cases"], ids=[c["name"] for c in F["cases"]])
def test_gpu_grove_flanking_matches_reference_golden(case):
    from genogrove_b200 import Grove, Interval
    g = Grove(case["order"])
    for op in case["ops"]:
        idx = f"chr{op['chrom']}"
        if op["op"] in ("regular", "sorted"):
            for s, e, i in op["items"]:
                g.insert_data(idx, Interval(s, e), i, sorted=op["op"] == "sorted")
        else:
            g.insert_data_bulk(idx, [(Interval(s, e), i) for s, e, i in op["items"]], presorted=True)
    for (c, qs, qe), ep, es in list(zip(case["queries"], case["pred"], case["succ"])):
        r = g.flanking(Interval(qs, qe), f"chr{c}")
        assert (None if r.get_predecessor() is None else r.get_predecessor().get_data()) == ep
        assert (None if r.get_successor() is None else r.get_successor().get_data()) == es


@pytest.mark.gpu
@pytest.mark.parametrize("n,n_chrom,q,kw", [
    (0, 1, 100, {}),
    (1, 1, 100, {}),
    (33, 2, 500, {}),
    (5000, 3, 20000, {}),
    (5000, 3, 20000, {"dup": True}),
    (20000, 5, 50001, {"nested": True}),
    (3000, 2, 10000, {"wide": True}),
    (200000, 24, 300000, {"dup": True, "nested": True}),
])
def test_gpu_flanking_vs_oracle(n, n_chrom, q, kw):
    from genogrove_b200 import DeviceIndex
    rng = np.random.default_rng(n * 31 + q)
    span = 100_000 if n < 100000 else 3_000_000
    chrom, s, e = random_index(rng, n, n_chrom, span=span, **kw)
    qc, qs, qe = random_queries(rng, q, n_chrom, span=span, wide=kw.get("wide", False), extra_chrom=1)
    qs[: q // 50] = 0  # start 0: no predecessor possible
    qe[: q // 50] = rng.integers(0, 50, q // 50)
    idx = DeviceIndex(0)
    perm = idx.build(chrom, s, e, n_chrom)
    assert np.array_equal(perm, expected_chain(chrom, s, e))
    coff = chrom_offsets(chrom, n_chrom)
    ep, es = O.flat_flanking_batch(s[perm], e[perm], coff, qc, qs, qe)
    for order in ("random", "sorted"):
        if order == "sorted":
            o = np.lexsort((qs, qc))
            qc, qs, qe, ep, es = qc[o], qs[o], qe[o], ep[o], es[o]
        pred, succ = idx.flanking_batch(qc, qs, qe)
        assert np.array_equal(pred, ep), order
        assert np.array_equal(succ, es), order


@pytest.mark.gpu
def test_gpu_flanking_u32_index_u64_queries_beyond_domain_and_errors():
    from genogrove_b200 import DeviceIndex, GGError
    idx = DeviceIndex(0)
    s = np.array([10, 20, 20, 4_000_000_000], np.uint64)
    e = np.array([15, 30, 30, 4_294_967_295], np.uint64)
    idx.build(np.zeros(4, np.uint32), s, e, 1)
    assert idx.stats().coord_bytes == 4
    qs = np.array([2**33, 16, 0, 4_294_967_295], np.uint64)
    qe = np.array([2**34, 2**35, 9, 4_294_967_295], np.uint64)
    pred, succ = idx.flanking_batch(np.zeros(4, np.uint32), qs, qe)
    assert pred.tolist() == [3, 0, NONE, 1] and succ.tolist() == [NONE, NONE, 0, NONE]
    with pytest.raises(GGError) as ei:
        idx.flanking_batch(np.zeros(1, np.uint32), np.array([5], np.uint64), np.array([4], np.uint64))
    assert ei.value.status == 4
    p, su = idx.flanking_batch(np.zeros(0, np.uint32), np.zeros(0, np.uint64), np.zeros(0, np.uint64))
    assert len(p) == 0 and len(su) == 0


@pytest.mark.gpu
def test_gpu_flanking_device_buffers_u32():
    import torch
    from genogrove_b200 import DeviceIndex, synth
    w = synth.cfg2(n=100_000, q=400_000, seed=3)
    idx = DeviceIndex(0)
    perm = idx.build(w.chrom, w.start, w.end, w.n_chrom)
    coff = chrom_offsets(w.chrom, w.n_chrom)
    ep, es = O.flat_flanking_batch(w.start[perm], w.end[perm], coff, w.q_chrom, w.q_start, w.q_end)
    dc = torch.from_numpy(w.q_chrom.astype(np.int32)).cuda()
    ds = torch.from_numpy(w.q_start.astype(np.uint32).view(np.int32)).cuda()
    de = torch.from_numpy(w.q_end.astype(np.uint32).view(np.int32)).cuda()
    dp = torch.empty(len(w.q_chrom), dtype=torch.int32, device="cuda")
    dsu = torch.empty_like(dp)
    idx.flanking_device(dc.data_ptr(), ds.data_ptr(), de.data_ptr(), 4, len(w.q_chrom), dp.data_ptr(), dsu.data_ptr(),
                        stream=torch.cuda.current_stream().cuda_stream)
    assert np.array_equal(dp.cpu().numpy().view(np.uint32), ep)
    assert np.array_equal(dsu.cpu().numpy().view(np.uint32), es)
