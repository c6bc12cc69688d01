"""BASELINE.json's configurations at FULL size on the GPU, checked through size-independent properties.

The oracle cannot answer 10^7-10^8 queries in seconds, so each full-size run is verified by four properties that
together pin the exact reference result (hits = { r : start[r] <= q.end and end[r] >= q.start }
in ascending chain rank):
  1. the build permutation is a stable lexicographic sort of (chrom, start, end)  -> the chain is the reference's;
  2. offsets are a non-decreasing CSR whose per-query counts equal ub(S, q.end) - ub(E, q.start - 1) computed
     independently with torch.searchsorted (library ops, none of this repo's kernels);
  3. every reported rank really overlaps its query, and ranks are strictly ascending inside a query
     (valid + no duplicates + the right count = exactly the reference's set, in the reference's order);
  4. a random sample of queries equals the CPU oracle's hit lists verbatim.
cfg1 is small enough to be compared with the oracle in full.
"""
import numpy as np
import pytest

from helpers import oracle_csr

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gg():
    import genogrove_b200 as g
    return g


def _chain(chrom, start, end, perm, n_chrom):
    """Sorted SoA of the grove (chain order) + per-chromosome begin offsets, and check property 1."""
    c, s, e = chrom[perm], start[perm], end[perm]
    # lexicographic non-decreasing, ties in input order (stable)
    same_c = c[1:] == c[:-1]
    same_s = same_c & (s[1:] == s[:-1])
    same_e = same_s & (e[1:] == e[:-1])
    assert (c[1:] >= c[:-1]).all()
    assert (s[1:][same_c] >= s[:-1][same_c]).all()
    assert (e[1:][same_s] >= e[:-1][same_s]).all()
    assert (perm[1:][same_e] > perm[:-1][same_e]).all()
    assert np.array_equal(np.sort(perm), np.arange(len(perm), dtype=perm.dtype))
    cbegin = np.concatenate([[0], np.cumsum(np.bincount(c, minlength=n_chrom))]).astype(np.int64)
    return s, e, cbegin


def _verify(torch, s_chain, e_chain, cbegin, n_chrom, qc, qs, qe, d_off, d_hits, sample=4000, seed=5):
    """Properties 2-4.  d_off (int64[Q+1]) and d_hits (int32[H]) are device tensors; the rest numpy."""
    dev = d_off.device
    nq = len(qc)
    t_qc = torch.from_numpy(qc.astype(np.int64)).to(dev)
    t_qs = torch.from_numpy(qs.astype(np.int64)).to(dev)
    t_qe = torch.from_numpy(qe.astype(np.int64)).to(dev)
    t_s = torch.from_numpy(s_chain.astype(np.int64)).to(dev)
    t_e = torch.from_numpy(e_chain.astype(np.int64)).to(dev)
    t_cb = torch.from_numpy(cbegin).to(dev)
    # 2. independent counts
    expect = torch.zeros(nq, dtype=torch.int64, device=dev)
    for c in range(n_chrom):
        m = (t_qc == c).nonzero().squeeze(1)
        if m.numel() == 0:
            continue
        a, b = int(cbegin[c]), int(cbegin[c + 1])
        if a == b:
            continue
        sc = t_s[a:b]
        ec_sorted, _ = torch.sort(t_e[a:b])
        hi = torch.searchsorted(sc, t_qe[m], right=True)
        lo = torch.searchsorted(ec_sorted, t_qs[m] - 1, right=True)
        expect[m] = hi - lo
        del sc, ec_sorted, hi, lo, m
    counts = d_off[1:] - d_off[:-1]
    assert int(d_off[0]) == 0
    assert bool((counts >= 0).all())
    assert bool(torch.equal(counts, expect)), "per-query hit counts differ from the independent searchsorted count"
    total = int(d_off[-1])
    assert total == int(expect.sum())
    del expect
    # 3. validity and strict ascent of every hit (in slabs of <= ~400M hits: cfg5's share has 1.9e10 hit pairs)
    width = t_cb[1:] - t_cb[:-1]
    q_lo = 0
    while q_lo < nq and total:
        q_hi = int(torch.searchsorted(d_off, d_off[q_lo:q_lo + 1] + 400_000_000, right=True)[0]) - 1
        q_hi = min(max(q_hi, q_lo + 1), nq)
        h0, h1 = int(d_off[q_lo]), int(d_off[q_hi])
        if h1 > h0:
            qid = q_lo + torch.repeat_interleave(torch.arange(q_hi - q_lo, device=dev), counts[q_lo:q_hi])
            h = d_hits[h0:h1].to(torch.int64)
            cq = t_qc[qid]
            assert bool((h < width[cq]).all())
            g = t_cb[cq] + h
            assert bool((t_s[g] <= t_qe[qid]).all()) and bool((t_e[g] >= t_qs[qid]).all()), "a reported rank does not overlap"
            same_q = qid[1:] == qid[:-1]
            assert bool((h[1:][same_q] > h[:-1][same_q]).all()), "ranks not strictly ascending inside a query"
            del qid, h, g, cq, same_q
        q_lo = q_hi
    # 4. verbatim sample against the CPU oracle
    rng = np.random.default_rng(seed)
    pick = np.sort(rng.choice(nq, size=min(sample, nq), replace=False))
    off = d_off.cpu().numpy().view(np.uint64)
    starts, ends = off[pick], off[pick + 1]
    got = [d_hits[int(a):int(b)].cpu().numpy().view(np.uint32) for a, b in zip(starts, ends)]
    # oracle on the chain itself (already sorted: chrom ids recovered from cbegin)
    chain_chrom = np.repeat(np.arange(n_chrom, dtype=np.uint32), np.diff(cbegin))
    o_off, o_hits = oracle_csr(chain_chrom, s_chain, e_chain, n_chrom, qc[pick], qs[pick], qe[pick])
    for i in range(len(pick)):
        assert np.array_equal(got[i], o_hits[int(o_off[i]):int(o_off[i + 1])]), f"query {pick[i]} differs from the oracle"
    return total


def _run_device(torch, idx, qc, qs, qe, nbytes=4):
    dev = torch.device("cuda:0")
    dt, tt = (np.uint32, np.int32) if nbytes == 4 else (np.uint64, np.int64)
    d_c = torch.from_numpy(qc.view(np.int32)).to(dev)
    d_s = torch.from_numpy(qs.astype(dt).view(tt)).to(dev)
    d_e = torch.from_numpy(qe.astype(dt).view(tt)).to(dev)
    nq = len(qc)
    d_off = torch.zeros(nq + 1, dtype=torch.int64, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    total = idx.count_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), nbytes, nq, d_off.data_ptr(), st, sync_total=True)
    d_hits = torch.empty(max(total, 1), dtype=torch.int32, device=dev)
    d_off.zero_()
    got = idx.fused_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), nbytes, nq, d_off.data_ptr(), d_hits.data_ptr(),
                           total, st, sync_total=True)
    assert got == total
    return d_off, d_hits


def test_cfg1_full_oracle(gg):
    """cfg1: 100k intervals on one chromosome, 10k queries of 1 kb — the reference's own CPU-runnable case."""
    from genogrove_b200 import synth
    w = synth.cfg1()
    idx = gg.DeviceIndex(0)
    perm = idx.build(w.chrom, w.start, w.end, w.n_chrom)
    off, hits = idx.intersect_batch(w.q_chrom, w.q_start, w.q_end)
    s, e, cbegin = _chain(w.chrom, w.start, w.end, perm, w.n_chrom)
    o_off, o_hits = oracle_csr(np.zeros(len(s), np.uint32), s, e, 1, w.q_chrom, w.q_start, w.q_end)
    assert np.array_equal(off, o_off) and np.array_equal(hits, o_hits)


@pytest.mark.parametrize("sorted_queries", [False, True])
def test_cfg2_full_size_host_path(gg, sorted_queries):
    """cfg2: 1M-interval hg38-shaped grove, 10M random 1 kb queries, through the host-buffer C ABI call."""
    torch = pytest.importorskip("torch")
    from genogrove_b200 import synth
    w = synth.cfg2(sorted_queries=sorted_queries)
    idx = gg.DeviceIndex(0)
    perm = idx.build(w.chrom, w.start, w.end, w.n_chrom)
    s, e, cbegin = _chain(w.chrom, w.start, w.end, perm, w.n_chrom)
    off, hits = idx.intersect_batch(w.q_chrom, w.q_start, w.q_end, copy_hits=False)
    assert len(off) == len(w.q_chrom) + 1
    dev = torch.device("cuda:0")
    d_off = torch.from_numpy(off.view(np.int64)).to(dev)
    d_hits = torch.from_numpy(np.ascontiguousarray(hits).view(np.int32)).to(dev)
    total = _verify(torch, s, e, cbegin, w.n_chrom, w.q_chrom, w.q_start, w.q_end, d_off, d_hits)
    assert total > len(w.q_chrom)  # ~1.7 hits per query


def test_cfg3_gencode_shaped_nested_duplicates(gg):
    """cfg3: GENCODE-shaped nested gene/transcript/exon features (many equal (start, end): the tie rule; long genes:
    max-end pruning), 50M queries.  u64 query coordinates, the facade's type."""
    torch = pytest.importorskip("torch")
    from genogrove_b200 import synth
    w = synth.cfg3()
    idx = gg.DeviceIndex(0)
    perm = idx.build(w.chrom, w.start, w.end, w.n_chrom)
    s, e, cbegin = _chain(w.chrom, w.start, w.end, perm, w.n_chrom)
    d_off, d_hits = _run_device(torch, idx, w.q_chrom, w.q_start, w.q_end, nbytes=8)
    _verify(torch, s, e, cbegin, w.n_chrom, w.q_chrom, w.q_start, w.q_end, d_off, d_hits, sample=2000)


@pytest.mark.parametrize("sorted_queries", [True, False])
def test_cfg4_full_size(gg, sorted_queries):
    """cfg4: 10M-interval grove x 100M 150 bp read-shaped queries (the bench workload), one GPU's share."""
    torch = pytest.importorskip("torch")
    from genogrove_b200 import synth
    w = synth.cfg4(sorted_queries=sorted_queries)
    idx = gg.DeviceIndex(0)
    perm = idx.build(w.chrom, w.start, w.end, w.n_chrom)
    s, e, cbegin = _chain(w.chrom, w.start, w.end, perm, w.n_chrom)
    d_off, d_hits = _run_device(torch, idx, w.q_chrom, w.q_start, w.q_end)
    total = _verify(torch, s, e, cbegin, w.n_chrom, w.q_chrom, w.q_start, w.q_end, d_off, d_hits)
    assert total == 744081723  # the bench's hit-pair count for this seed


def test_cfg5_one_gpu_share_of_chromosome_sharded_build(gg):
    """cfg5: 200M intervals / 1B queries chromosome-sharded over 8 GPUs: this is rank 0's share at full per-GPU
    size (LPT assignment of chromosomes by interval count, local chromosome ids, ~25M intervals, ~125M queries)."""
    torch = pytest.importorskip("torch")
    from genogrove_b200 import shard, synth
    n_total, q_total, world = 200_000_000, 1_000_000_000, 8
    per = synth._split_by_length(n_total, synth.HG38_LENGTHS)
    owner = shard.assign_chromosomes(per, world)
    mine = np.flatnonzero(owner == 0)
    assert len(mine) >= 1
    # this rank's intervals and queries only, generated per chromosome with the cfg2 recipe
    rng = np.random.default_rng(77)
    qper = synth._split_by_length(q_total, synth.HG38_LENGTHS)
    chrom, start, end, qc, qs, qe = [], [], [], [], [], []
    for local, c in enumerate(mine):
        L = int(synth.HG38_LENGTHS[c])
        n = int(per[c])
        ln = np.exp(rng.uniform(np.log(100), np.log(10000), n)).astype(np.int64)
        st = (rng.random(n) * (L - ln - 1)).astype(np.int64)
        chrom.append(np.full(n, local, np.uint32)), start.append(st.astype(np.uint64)), end.append((st + ln - 1).astype(np.uint64))
        m = int(qper[c])
        q0 = np.sort(rng.integers(0, L - 150, m, dtype=np.int64)).astype(np.uint64)
        qc.append(np.full(m, local, np.uint32)), qs.append(q0), qe.append(q0 + np.uint64(149))
    chrom, start, end = np.concatenate(chrom), np.concatenate(start), np.concatenate(end)
    sh = rng.permutation(len(chrom))  # unsorted input, as cfg5 states
    chrom, start, end = chrom[sh], start[sh], end[sh]
    qc, qs, qe = np.concatenate(qc), np.concatenate(qs), np.concatenate(qe)
    assert 20_000_000 < len(chrom) < 30_000_000 and 100_000_000 < len(qc) < 150_000_000
    idx = gg.DeviceIndex(0)
    perm = idx.build(chrom, start, end, len(mine))
    s, e, cbegin = _chain(chrom, start, end, perm, len(mine))
    d_off, d_hits = _run_device(torch, idx, qc, qs, qe)
    _verify(torch, s, e, cbegin, len(mine), qc, qs, qe, d_off, d_hits, sample=2000)


@pytest.mark.parametrize("cfg", ["cfg4", "cfg3"])
def test_full_size_groves_against_the_live_reference(gg, cfg):
    """The CUDA path against the UNMODIFIED reference itself (oracle/_ref/libgg_ref.so: grove<interval, uint32_t>::intersect
    per query, order 32) on the full-size cfg4 / cfg3 groves, id for id and in order, for 2 M / 1 M queries taken from the
    middle of the real batches — not against the flat restatement the other tests use."""
    import oracle as O
    if not O.ref_available():
        pytest.skip("oracle/_ref/libgg_ref.so not built")
    from genogrove_b200 import synth
    w = synth.cfg4(q=20_000_000) if cfg == "cfg4" else synth.cfg3(q=5_000_000)
    nq = 2_000_000 if cfg == "cfg4" else 1_000_000
    lo = (len(w.q_chrom) - nq) // 2
    qc, qs, qe = w.q_chrom[lo:lo + nq], w.q_start[lo:lo + nq], w.q_end[lo:lo + nq]
    ref = O.RefGrove(32)
    ids = np.arange(len(w.chrom), dtype=np.uint32)
    for c in range(w.n_chrom):
        m = w.chrom == c
        # cfg3 holds many equal (start, end) keys (cfg4 possibly a few): the reference's sorted-bulk path keeps equal keys in
        # input order, which is what the stable device sort does; its unsorted-bulk path runs an unstable std::sort
        order = np.lexsort((w.end[m], w.start[m]))
        ref.insert_bulk(c, w.start[m][order], w.end[m][order], ids[m][order], presorted=True)
    r_off, r_ids, _ = ref.intersect_batch(qc, qs, qe, threads=ref.hardware_threads())
    idx = gg.DeviceIndex(0)
    perm = idx.build(w.chrom, w.start, w.end, w.n_chrom)  # stable: equal keys in input order, like sorted-bulk above
    cbegin = np.concatenate([[0], np.cumsum(np.bincount(w.chrom, minlength=w.n_chrom))]).astype(np.int64)
    for form in ("csr", "compact"):
        if form == "csr":
            off, hits = idx.intersect_batch(qc, qs, qe)
        else:
            off, hits = idx.intersect_batch_compact(qc, qs.astype(np.uint32), qe.astype(np.uint32))
        assert np.array_equal(off, r_off)
        counts = np.diff(off).astype(np.int64)
        got_ids = perm[np.repeat(cbegin[qc], counts) + hits]
        assert np.array_equal(got_ids, r_ids), f"{cfg} {form}: hit ids differ from the reference's"
    ref.close()


@pytest.mark.parametrize("dense_emit", [-1, 1, 0])
def test_uniformly_dense_tiles_between_stage_and_dense_threshold(gg, dense_emit):
    """~36 hits per query on every query (a 50 M-interval hg38-shaped grove has this density): warp tiles of ~1150 hits exceed
    the emit stage but stay below the 2048-hit threshold of the dense kernel.  The shape test (dense_tile, emit_scan.cuh)
    sends them to the segment-flush kernel; a tile with one long list among short ones keeps the direct walk."""
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(131)
    n, span = 2_500_000, 150_000_000
    ln = np.exp(rng.uniform(np.log(100), np.log(10000), n)).astype(np.int64)
    start = (rng.random(n) * (span - ln - 1)).astype(np.int64)
    end = start + ln - 1
    end[::100_000] = start[::100_000] + 40_000_000   # a few huge intervals: skewed tiles among the uniform ones
    chrom = np.zeros(n, np.uint32)
    idx = gg.DeviceIndex(0)
    idx.set_option("dense_emit", dense_emit)
    perm = idx.build(chrom, start.astype(np.uint64), end.astype(np.uint64), 1)
    s, e, cbegin = _chain(chrom, start.astype(np.uint64), end.astype(np.uint64), perm, 1)
    nq = 2_000_000
    qs = np.sort(rng.integers(0, span - 150, nq, dtype=np.int64)).astype(np.uint64)
    qs[::7] = rng.integers(0, span - 150, len(qs[::7]), dtype=np.int64).astype(np.uint64)  # and some out-of-order ones
    qc, qe = np.zeros(nq, np.uint32), qs + np.uint64(149)
    for _ in range(2):  # the second pass runs with the density the first one reported (dense_emit = -1)
        d_off, d_hits = _run_device(torch, idx, qc, qs, qe)
        total = _verify(torch, s, e, cbegin, 1, qc, qs, qe, d_off, d_hits, sample=1500)
    assert total > 30 * nq
