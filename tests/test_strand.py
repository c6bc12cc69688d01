"""Strand-aware intersect (genomic_coordinate keys): the oracle pinned against the unmodified reference's golden answers
(CPU), and the CUDA path (gg_index_set_strands + gg_intersect_*stranded*) against the golden cases and the oracle (GPU).
Reference: include/genogrove/data_type/genomic_coordinate.hpp:97-110 (order), :156-160 (overlaps);
cases restate tests/structure/genomic_coordinate_grove_test.cpp:150-266."""
import json
import os

import numpy as np
import pytest

import oracle as O
from helpers import chrom_offsets, random_index, random_queries

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "strand_cases.json")
with open(GOLD) as f:
    S = json.load(f)
SYM = np.frombuffer(b"+-.*", dtype=np.uint8)


def _case_arrays(case):
    it = case["items"]
    st = np.frombuffer("".join(t[0] for t in it).encode(), dtype=np.uint8)
    s = np.array([t[1] for t in it], dtype=np.uint64)
    e = np.array([t[2] for t in it], dtype=np.uint64)
    ids = np.array([t[3] for t in it], dtype=np.int64)
    return st, s, e, ids


def _case_queries(case, n_chrom):
    q = case["queries"]
    qc = np.array([c if c < n_chrom else O.NO_CHROM for c, _, _, _ in q], dtype=np.uint32)
    qst = np.frombuffer("".join(t[1] for t in q).encode(), dtype=np.uint8)
    return qc, np.array([t[2] for t in q], np.uint64), np.array([t[3] for t in q], np.uint64), qst


def test_strand_golden_shape():
    assert len(S["cases"]) >= 14 and sum(len(c["queries"]) for c in S["cases"]) >= 1200


@pytest.mark.parametrize("case", S["cases"], ids=[c["name"] for c in S["cases"]])
def test_flat_strand_oracle_matches_reference_golden(case):
    st, s, e, ids = _case_arrays(case)
    perm = O.gc_chain(s, e, st, "regular" if case["path"] == "regular" else "stable")
    c = case["chrom"]
    coff = np.array([0] * (c + 1) + [len(s)], dtype=np.uint64)  # chromosomes below c are empty
    qc, qs, qe, qst = _case_queries(case, c + 1)
    off, ranks = O.flat_intersect_batch_stranded(s[perm], e[perm], st[perm], coff, qc, qs, qe, qst)
    got = [[int(ids[perm[int(r)]]) for r in ranks[int(off[k]):int(off[k + 1])]] for k in range(len(qc))]
    assert got == case["expected"]


def test_strand_arithmetic_known_answers():
    """genomic_coordinate.hpp:34-38 documented examples + the strand order, on the C restatement."""
    L = O.oracle_lib()
    ov = lambda a, b: L.orc_gc_overlaps(ord(a[0]), a[1], a[2], ord(b[0]), b[1], b[2])  # noqa: E731
    assert ov(("+", 100, 200), ("+", 150, 250)) == 1
    assert ov(("+", 100, 200), ("-", 150, 250)) == 0
    assert ov(("+", 100, 200), ("*", 150, 250)) == 1
    assert ov(("+", 100, 200), ("+", 300, 400)) == 0
    assert ov((".", 0, 10), (".", 10, 20)) == 1 and ov((".", 0, 10), ("+", 10, 20)) == 0
    assert [L.orc_strand_order(ord(c)) for c in "*.+-"] == [0, 1, 2, 3] and L.orc_strand_order(ord("x")) == -1
    assert L.orc_gc_less(ord("*"), 5, 9, ord("."), 5, 9) == 1 and L.orc_gc_less(ord("-"), 5, 9, ord("+"), 5, 9) == 0
    assert L.orc_gc_less(ord("-"), 5, 8, ord("*"), 5, 9) == 1


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built")
def test_strand_arithmetic_matches_reference():
    L, R = O.oracle_lib(), O.ref_lib()
    rng = np.random.default_rng(3)
    for _ in range(2000):
        a, b = (int(x) for x in rng.integers(0, 30, 2))
        c, d = (int(x) for x in rng.integers(0, 30, 2))
        sa, sb = (bytes([SYM[int(x)]]) for x in rng.integers(0, 4, 2))
        ka, kb = (sa, min(a, b), max(a, b)), (sb, min(c, d), max(c, d))
        assert L.orc_gc_less(ka[0][0], ka[1], ka[2], kb[0][0], kb[1], kb[2]) == R.ref_gc_less(*ka, *kb)
        assert L.orc_gc_overlaps(ka[0][0], ka[1], ka[2], kb[0][0], kb[1], kb[2]) == R.ref_gc_overlaps(*ka, *kb)


# ----------------------------------------------------------------------------------------------- GPU
def _build_stranded(idx, chrom, s, e, st, n_chrom, regular=False):
    """Chain order (start, end, strand order[, reverse input for the regular path]) through GG_TIE_EXPLICIT."""
    from genogrove_b200 import TIE_EXPLICIT
    n = len(s)
    so = np.array([O.STRAND_ORDER[chr(c)] for c in st], dtype=np.uint64)
    seq = np.arange(n, dtype=np.uint64)
    tie = (so << np.uint64(32)) | (np.uint64(n - 1) - seq if regular and n else seq)
    perm = idx.build(chrom, s, e, n_chrom, tie_mode=TIE_EXPLICIT, tie_key=tie)
    idx.set_strands(st[perm])
    return perm


@pytest.mark.gpu
@pytest.mark.parametrize("case", S["cases"], ids=[c["name"] for c in S["cases"]])
def test_gpu_stranded_matches_reference_golden(case):
    from genogrove_b200 import DeviceIndex
    st, s, e, ids = _case_arrays(case)
    c = case["chrom"]
    idx = DeviceIndex(0)
    perm = _build_stranded(idx, np.full(len(s), c, np.uint32), s, e, st, c + 1, regular=case["path"] == "regular")
    qc, qs, qe, qst = _case_queries(case, c + 1)
    off, ranks = idx.intersect_batch_stranded(qc, qs, qe, qst)
    got = [[int(ids[perm[int(r)]]) for r in ranks[int(off[k]):int(off[k + 1])]] for k in range(len(qc))]
    assert got == case["expected"]


@pytest.mark.gpu
@pytest.mark.parametrize("n,n_chrom,q,kw", [
    (0, 1, 50, {}),
    (1, 1, 50, {}),
    (40, 2, 700, {"dup": True}),
    (6000, 3, 30000, {"dup": True}),
    (30000, 4, 60001, {"nested": True}),
    (4000, 2, 9000, {"wide": True}),
    (200000, 24, 250000, {"dup": True, "nested": True}),
])
def test_gpu_stranded_vs_oracle(n, n_chrom, q, kw):
    from genogrove_b200 import DeviceIndex
    rng = np.random.default_rng(n * 17 + q)
    span = 100_000 if n < 100000 else 3_000_000
    chrom, s, e = random_index(rng, n, n_chrom, span=span, **kw)
    st = SYM[rng.integers(0, 4, n)]
    qc, qs, qe = random_queries(rng, q, n_chrom, span=span, wide=kw.get("wide", False), extra_chrom=1)
    qst = SYM[rng.integers(0, 4, q)]
    idx = DeviceIndex(0)
    idx.set_option("chunk_queries", 100_000)  # several chunks: chained offsets
    perm = _build_stranded(idx, chrom, s, e, st, n_chrom)
    coff = chrom_offsets(chrom, n_chrom)
    eo, eh = O.flat_intersect_batch_stranded(s[perm], e[perm], st[perm], coff, qc, qs, qe, qst)
    off, hits = idx.intersect_batch_stranded(qc, qs, qe, qst)
    assert np.array_equal(off, eo)
    assert np.array_equal(hits, eh)
    # the wildcard query on every key == the interval path
    o2, h2 = idx.intersect_batch_stranded(qc, qs, qe, np.full(q, ord("*"), np.uint8))
    o3, h3 = idx.intersect_batch(qc, qs, qe)
    assert np.array_equal(o2, o3) and np.array_equal(h2, h3)


@pytest.mark.gpu
def test_gpu_stranded_errors_and_device_buffers():
    import torch
    from genogrove_b200 import DeviceIndex, GGError
    idx = DeviceIndex(0)
    s = np.array([10, 10, 50], np.uint64)
    e = np.array([20, 20, 60], np.uint64)
    idx.build(np.zeros(3, np.uint32), s, e, 1)
    z = np.zeros(1, np.uint32)
    with pytest.raises(GGError) as ei:  # no strands attached yet
        idx.intersect_batch_stranded(z, [0], [5], b"+")
    assert ei.value.status == 5
    with pytest.raises(GGError) as ei:
        idx.set_strands(b"+x-")
    assert ei.value.status == 1
    with pytest.raises(GGError):
        idx.set_strands(b"+-")  # wrong length
    idx.set_strands(b"+-.")
    with pytest.raises(GGError) as ei:  # genomic_coordinate ctor: invalid strand (genomic_coordinate.hpp:78-81)
        idx.intersect_batch_stranded(z, [0], [5], b"?")
    assert ei.value.status == 1
    with pytest.raises(GGError) as ei:
        idx.intersect_batch_stranded(z, [9], [5], b"+")
    assert ei.value.status == 4
    off, hits = idx.intersect_batch_stranded(np.zeros(4, np.uint32), [0, 0, 0, 0], [100, 100, 100, 100], b"+-.*")
    assert off.tolist() == [0, 1, 2, 3, 6] and hits.tolist() == [0, 1, 2, 0, 1, 2]
    # device buffers, u32 coordinates, capacity contract
    dc = torch.zeros(4, dtype=torch.int32, device="cuda")
    ds = torch.zeros(4, dtype=torch.int32, device="cuda")
    de = torch.full((4,), 100, dtype=torch.int32, device="cuda")
    dst = torch.tensor(list(b"+-.*"), dtype=torch.uint8, device="cuda")
    doff = torch.zeros(5, dtype=torch.int64, device="cuda")
    dh = torch.zeros(8, dtype=torch.int32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    total = idx.stranded_device(dc.data_ptr(), ds.data_ptr(), de.data_ptr(), dst.data_ptr(), 4, 4, doff.data_ptr(), dh.data_ptr(), 8, stream)
    assert total == 6 and doff.cpu().tolist() == [0, 1, 2, 3, 6] and dh.cpu().tolist()[:6] == [0, 1, 2, 0, 1, 2]
    with pytest.raises(GGError) as ei:
        idx.stranded_device(dc.data_ptr(), ds.data_ptr(), de.data_ptr(), dst.data_ptr(), 4, 4, doff.data_ptr(), dh.data_ptr(), 4, stream)
    assert ei.value.status == 7 and idx.last_total == 6
    idx.build(np.zeros(3, np.uint32), s, e, 1)  # a rebuild drops the strands
    with pytest.raises(GGError):
        idx.intersect_batch_stranded(z, [0], [5], b"+")


@pytest.mark.gpu
def test_set_strands_failure_leaves_no_strands():
    """A device allocation that fails inside gg_index_set_strands must fail the call and leave the handle without
    strand records (the stranded entry points then refuse to run) — ADVICE r1: the error used to be swallowed."""
    import genogrove_b200 as gg
    rng = np.random.default_rng(3)
    n = 5000
    start = np.sort(rng.integers(0, 100_000, n)).astype(np.uint64)
    end = start + rng.integers(1, 500, n).astype(np.uint64)
    idx = gg.DeviceIndex(0)
    idx.build(np.zeros(n, np.uint32), start, end, 1)
    strands = np.frombuffer(b"+-.*", dtype=np.uint8)[rng.integers(0, 4, n)]
    q = (np.zeros(10, np.uint32), start[:10], end[:10], strands[:10])
    idx.set_strands(strands)
    idx.intersect_batch_stranded(*q)
    for k in (1, 2, 3, 5):  # the k-th allocation of the call fails
        idx.set_option("inject_alloc_failure", k)
        with pytest.raises(gg.GGError) as ei:
            idx.set_strands(strands)
        assert ei.value.status in (2, 3)
        idx.set_option("inject_alloc_failure", 0)
        with pytest.raises(gg.GGError) as ei:
            idx.intersect_batch_stranded(*q)
        assert ei.value.status == 5  # GG_ERR_NOT_BUILT: no strands attached
    idx.set_strands(strands)
    off, hits = idx.intersect_batch_stranded(*q)
    assert off[-1] >= 10
