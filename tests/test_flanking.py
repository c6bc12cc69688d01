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


# ------------------------------------------------------------ genomic_coordinate keys, strand-compatible predicate
# grove<genomic_coordinate, T>::flanking(q, index, pred) with the predicate the reference documents
# (include/genogrove/structure/grove/grove_flanking.ipp:47-54).  Golden answers: tests/golden/flanking_strand_cases.json,
# written by tests/golden/make_golden_flanking_strand.py from the unmodified reference (identical over 4 tree orders x 3
# insert paths).
GOLD_S = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "flanking_strand_cases.json")
with open(GOLD_S) as f:
    FS = json.load(f)


def _gc_arrays(case):
    it = case["items"]  # [strand, start, end, id], id = chain rank
    st = np.frombuffer("".join(t[0] for t in it).encode(), dtype=np.uint8)
    s = np.array([t[1] for t in it], dtype=np.uint64)
    e = np.array([t[2] for t in it], dtype=np.uint64)
    q = case["queries"]
    qc = np.array([t[0] for t in q], dtype=np.uint32)
    qst = np.frombuffer("".join(t[1] for t in q).encode(), dtype=np.uint8)
    qs = np.array([t[2] for t in q], dtype=np.uint64)
    qe = np.array([t[3] for t in q], dtype=np.uint64)
    exp_p = np.array([NONE if v is None else v for v in case["pred"]], dtype=np.uint32)
    exp_s = np.array([NONE if v is None else v for v in case["succ"]], dtype=np.uint32)
    return st, s, e, qc, qst, qs, qe, exp_p, exp_s


def test_flanking_strand_golden_shape():
    assert len(FS["cases"]) >= 7 and sum(len(c["queries"]) for c in FS["cases"]) >= 900
    assert len(FS["default_predicate_shape_dependence"]) >= 1


@pytest.mark.parametrize("case", FS["cases"], ids=[c["name"] for c in FS["cases"]])
def test_flat_stranded_flanking_oracle_matches_reference_golden(case):
    st, s, e, qc, qst, qs, qe, exp_p, exp_s = _gc_arrays(case)
    pred, succ = O.flat_flanking_batch_stranded(s, e, st, [0, len(s)], qc, qs, qe, qst)
    assert np.array_equal(pred, exp_p)
    assert np.array_equal(succ, exp_s)


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built")
def test_default_predicate_on_genomic_coordinate_keys_depends_on_tree_shape():
    """Why the default predicate has no device form on genomic_coordinate keys: the unmodified reference gives different
    answers for the SAME keys depending on how they were inserted (the pruning rule of query_engine.hpp:176-199 assumes
    non-overlapping candidates lie wholly before / after the query, which strand-incompatible overlaps violate)."""
    rec = FS["default_predicate_shape_dependence"][0]
    items, (c, st, qs, qe) = rec["items"], rec["query"]
    a, b = O.RefGcGrove(rec["order"]), O.RefGcGrove(rec["order"])
    rng = np.random.default_rng(5)
    seen = set()
    for trial in range(40):  # some insertion order reproduces a second answer
        g = O.RefGcGrove(rec["order"])
        for j in rng.permutation(len(items)):
            g.insert(0, *items[j])
        seen.add(g.flanking(c, st, qs, qe))
        g.close()
    b.insert_bulk(0, "".join(t[0] for t in items), [t[1] for t in items], [t[2] for t in items], [t[3] for t in items])
    seen.add(b.flanking(c, st, qs, qe))
    assert len(seen) >= 2, seen
    # ... while the strand-compatible predicate gives one answer whatever the shape
    for j in rng.permutation(len(items)):
        a.insert(0, *items[j])
    assert a.flanking(c, st, qs, qe, strand_predicate=True) == b.flanking(c, st, qs, qe, strand_predicate=True)


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built")
def test_flat_stranded_flanking_oracle_matches_reference_random():
    rng = np.random.default_rng(123)
    sym = np.frombuffer(b"*.+-", dtype=np.uint8)
    for trial in range(5):
        n = int(rng.integers(1, 400))
        s = (rng.integers(0, 500, n) * 4).astype(np.uint64)
        e = s + (rng.integers(0, 12, n) * 4).astype(np.uint64) + np.where(rng.random(n) < 0.05, 800, 0).astype(np.uint64)
        st = sym[rng.integers(0, 4, n)]
        keys = sorted({(int(a), int(b), int(c)) for a, b, c in zip(s, e, st)}, key=lambda k: (k[0], k[1], O.STRAND_ORDER[chr(k[2])]))
        s, e, st = (np.array([k[i] for k in keys], dtype=dt) for i, dt in ((0, np.uint64), (1, np.uint64), (2, np.uint8)))
        g = O.RefGcGrove(int(rng.choice([3, 6, 32])))
        for j in rng.permutation(len(s)):
            g.insert(0, chr(st[j]), int(s[j]), int(e[j]), int(j))
        qs = rng.integers(0, 2100, 300).astype(np.uint64)
        qe = qs + rng.integers(0, 200, 300).astype(np.uint64)
        qst = sym[rng.integers(0, 4, 300)]
        fp, fsu = O.flat_flanking_batch_stranded(s, e, st, [0, len(s)], np.zeros(300, np.uint32), qs, qe, qst)
        for i in range(300):
            assert g.flanking(0, chr(qst[i]), int(qs[i]), int(qe[i]), strand_predicate=True) == (int(fp[i]), int(fsu[i]))
        g.close()


@pytest.mark.gpu
@pytest.mark.parametrize("case", FS["cases"], ids=[c["name"] for c in FS["cases"]])
def test_gpu_stranded_flanking_matches_reference_golden(case):
    from genogrove_b200 import DeviceIndex
    st, s, e, qc, qst, qs, qe, exp_p, exp_s = _gc_arrays(case)
    idx = DeviceIndex(0)
    idx.upload_sorted(np.zeros(len(s), np.uint32), s, e, 1)
    idx.set_strands(st)
    pred, succ = idx.flanking_batch_stranded(qc, qs, qe, qst)
    assert np.array_equal(pred, exp_p)
    assert np.array_equal(succ, exp_s)


@pytest.mark.gpu
@pytest.mark.parametrize("wide", [False, True])
@pytest.mark.parametrize("skew", ["uniform", "one_strand_runs"])
def test_gpu_stranded_flanking_matches_oracle_random(wide, skew):
    """Several chromosomes (one empty, one tiny), duplicated end values, long runs of a single strand (the per-node count
    search instead of the mask of the node at hand), u64 layout, unknown chromosomes; and the error contract."""
    import torch
    from genogrove_b200 import DeviceIndex, GGError, TIE_EXPLICIT
    rng = np.random.default_rng(77 + wide)
    n_chrom = 5
    n = 30_000
    chrom = rng.integers(0, n_chrom, n).astype(np.uint32)
    chrom[chrom == 3] = 2
    chrom[:3] = 3
    s = (rng.integers(0, 400_000, n) // 8 * 8).astype(np.uint64)
    e = s + (rng.integers(0, 40, n) * 8).astype(np.uint64)
    long_ = rng.random(n) < 0.01
    e[long_] += np.uint64(100_000)
    sym = np.frombuffer(b"*.+-", dtype=np.uint8)
    st = sym[rng.integers(0, 4, n)]
    if skew == "one_strand_runs":  # thousands of consecutive '+' keys: a '-' query has to jump over them
        st = np.where((s > 100_000) & (s < 300_000), np.uint8(ord("+")), st).astype(np.uint8)
    if wide:
        s, e = s + np.uint64(2**34), e + np.uint64(2**34)
    so = np.array([O.STRAND_ORDER[chr(c)] for c in st], dtype=np.uint64)
    idx = DeviceIndex(0)
    perm = idx.build(chrom, s, e, n_chrom, tie_mode=TIE_EXPLICIT, tie_key=(so << np.uint64(32)) | np.arange(n, dtype=np.uint64))
    idx.set_strands(st[perm])
    coff = chrom_offsets(chrom, n_chrom)
    nq = 20_000
    qc = rng.integers(0, n_chrom + 1, nq).astype(np.uint32)
    qs = rng.integers(0, 520_000, nq).astype(np.uint64)
    qs[::50] = 0
    qe = qs + rng.integers(0, 3000, nq).astype(np.uint64)
    if wide:
        qs, qe = qs + np.uint64(2**34), qe + np.uint64(2**34)
        qs[::50] = 0
    qst = sym[rng.integers(0, 4, nq)]
    ep, es = O.flat_flanking_batch_stranded(s[perm], e[perm], st[perm], coff, qc, qs, qe, qst)
    pred, succ = idx.flanking_batch_stranded(qc, qs, qe, qst)
    assert np.array_equal(pred, ep) and np.array_equal(succ, es)
    assert (pred != NONE).sum() > nq // 3 and (succ != NONE).sum() > nq // 3
    if not wide:  # device-buffer form, u32 coordinates
        dev = torch.device("cuda:0")
        t = [torch.from_numpy(a).to(dev) for a in (qc.view(np.int32), qs.astype(np.uint32).view(np.int32),
                                                   qe.astype(np.uint32).view(np.int32), qst)]
        d_p = torch.empty(nq, dtype=torch.int32, device=dev)
        d_s = torch.empty(nq, dtype=torch.int32, device=dev)
        idx.flanking_stranded_device(t[0].data_ptr(), t[1].data_ptr(), t[2].data_ptr(), t[3].data_ptr(), 4, nq, d_p.data_ptr(),
                                     d_s.data_ptr(), torch.cuda.current_stream().cuda_stream)
        assert np.array_equal(d_p.cpu().numpy().view(np.uint32), ep) and np.array_equal(d_s.cpu().numpy().view(np.uint32), es)
    bad = qst.copy()
    bad[7] = ord("x")
    with pytest.raises(GGError) as ei:
        idx.flanking_batch_stranded(qc, qs, qe, bad)
    assert ei.value.status == 1  # GG_ERR_INVALID_ARG: the reference's std::invalid_argument for a bad strand
    plain = DeviceIndex(0)
    plain.build(chrom, s, e, n_chrom)
    with pytest.raises(GGError):
        plain.flanking_batch_stranded(qc, qs, qe, qst)  # no strands attached
