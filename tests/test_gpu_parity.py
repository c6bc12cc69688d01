"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle, bit-exact (offsets and the
ordered rank list per query).  Marked gpu: run on the B200 box."""
import numpy as np
import pytest

import oracle as O
from helpers import chrom_offsets, expected_chain, oracle_csr, random_index, random_queries

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gg():
    import genogrove_b200 as g
    return g


def _check(idx, chrom, start, end, n_chrom, q, perm=None, two_phase=False):
    qc, qs, qe = q
    exp_off, exp_hits = oracle_csr(chrom, start, end, n_chrom, qc, qs, qe, perm)
    off, hits = (idx.intersect_two_phase if two_phase else idx.intersect_batch)(qc, qs, qe)
    assert np.array_equal(off, exp_off)
    assert np.array_equal(hits, exp_hits)
    return int(exp_off[-1])


@pytest.mark.parametrize("n", [0, 1, 2, 31, 32, 33, 63, 64, 65, 1000, 1023, 1024, 1025, 4096, 33000, 100_003])
def test_build_sizes_single_chrom(gg, oracle, n):
    rng = np.random.default_rng(n + 1)
    chrom, start, end = random_index(rng, n, 1, span=50 * max(n, 10))
    idx = gg.DeviceIndex(0)
    perm = idx.build(chrom, start, end, 1)
    assert np.array_equal(perm, expected_chain(chrom, start, end))
    _check(idx, chrom, start, end, 1, random_queries(rng, 5000, 1, span=50 * max(n, 10)))
    st = idx.stats()
    assert st.n_intervals == n and st.coord_bytes == 4 and st.fanout == 32


@pytest.mark.parametrize("dup,nested", [(False, False), (True, False), (False, True), (True, True)])
@pytest.mark.parametrize("n_chrom", [1, 3, 24, 300])
def test_multi_chrom_shapes(gg, oracle, n_chrom, dup, nested):
    rng = np.random.default_rng(1000 + n_chrom + 2 * dup + 4 * nested)
    n = 60_000
    chrom, start, end = random_index(rng, n, n_chrom, span=200_000, dup=dup, nested=nested)
    if n_chrom >= 3:
        chrom[chrom == 1] = 0  # an empty chromosome in the middle
    idx = gg.DeviceIndex(0)
    perm = idx.build(chrom, start, end, n_chrom)
    assert np.array_equal(perm, expected_chain(chrom, start, end))
    q = random_queries(rng, 40_000, n_chrom, span=200_000, extra_chrom=2)
    q[0][::97] = gg.NO_CHROM  # unknown index -> empty result
    hits = _check(idx, chrom, start, end, n_chrom, q)
    assert hits > 0
    counts = idx.chrom_counts()
    assert np.array_equal(counts, np.bincount(chrom, minlength=n_chrom))


@pytest.mark.parametrize("first", ["upload_sorted", "build"])
def test_update_chrom_splices_one_chromosome(gg, oracle, first):
    """gg_index_update_chrom: replacing, emptying, growing and appending single chromosomes leaves exactly the snapshot a
    full gg_index_upload_sorted of the same chains gives, and uploads only the changed chromosome's records."""
    rng = np.random.default_rng(321)
    n_chrom = 6
    chrom, start, end = random_index(rng, 50_000, n_chrom, span=300_000, dup=True, nested=True)
    order = expected_chain(chrom, start, end)
    chains = {c: (start[order][chrom[order] == c].copy(), end[order][chrom[order] == c].copy()) for c in range(n_chrom)}
    idx = gg.DeviceIndex(0)
    if first == "build":
        idx.build(chrom, start, end, n_chrom)
    else:
        idx.upload_sorted(chrom[order], start[order], end[order], n_chrom)

    def flat():
        cs = sorted(chains)
        c = np.concatenate([np.full(len(chains[k][0]), k, np.uint32) for k in cs])
        return c, np.concatenate([chains[k][0] for k in cs]), np.concatenate([chains[k][1] for k in cs])

    def new_chain(m, wide=False):
        s = np.sort(rng.integers(0, 300_000, m).astype(np.uint64))
        if wide:
            s = s + np.uint64(2**33)
        return s, s + rng.integers(0, 3000, m).astype(np.uint64)

    steps = [(2, new_chain(20_000)), (0, new_chain(0)), (5, new_chain(7)), (6, new_chain(1000)), (8 - 1, new_chain(33)),
             (3, new_chain(4000, wide=True)), (3, new_chain(10))]
    for c, (s_c, e_c) in steps:
        idx.update_chrom(c, s_c, e_c)
        chains[c] = (s_c, e_c)
        st = idx.stats()
        fc, fs, fe = flat()
        assert st.h2d_bytes == 16 * len(s_c) and st.n_intervals == len(fc) and st.n_chrom == max(chains) + 1
        assert st.coord_bytes == (8 if fe.max() >= 2**32 else 4)
        nc = st.n_chrom
        q = random_queries(rng, 20_000, nc, span=300_000, extra_chrom=1, wide=bool(st.coord_bytes == 8))
        coff = np.concatenate([[0], np.cumsum(np.bincount(fc, minlength=nc))]).astype(np.uint64)
        exp_off, exp_hits = oracle.flat_intersect_batch(fs, fe, coff, *q)
        off, hits = idx.intersect_batch(*q)
        assert np.array_equal(off, exp_off) and np.array_equal(hits, exp_hits), (c, len(s_c))
    idx.set_option("keep_build_workspace", 0)
    with pytest.raises(Exception):
        idx.update_chrom(1, *new_chain(5))


@pytest.mark.parametrize("legacy", [0, 1])
@pytest.mark.parametrize("wide", [False, True])
def test_sort_forms_large(gg, oracle, legacy, wide):
    """Both radix-sort forms (onesweep passes with a decoupled look-back over ~1000 tiles; the three-kernel passes) give the
    stable (chrom, start, end) permutation on an input with heavy duplicates, constant high digits and a ragged last tile."""
    rng = np.random.default_rng(77 + legacy + 2 * wide)
    n = 3_000_017
    chrom, start, end = random_index(rng, n, 7, span=3_000_000, dup=True, wide=wide)
    idx = gg.DeviceIndex(0)
    idx.set_option("sort_legacy", legacy)
    for tie_mode in (0, 1):
        perm = idx.build(chrom, start, end, 7, tie_mode=tie_mode)
        assert np.array_equal(perm, expected_chain(chrom, start, end, tie_mode))
    _check(idx, chrom, start, end, 7, random_queries(rng, 20_000, 7, span=3_000_000, wide=wide), perm=perm)


@pytest.mark.parametrize("tie_mode", [0, 1, 2])
def test_tie_modes(gg, oracle, tie_mode):
    rng = np.random.default_rng(5 + tie_mode)
    n = 20_000
    chrom, start, end = random_index(rng, n, 4, span=20_000, dup=True)
    tie = rng.integers(0, 50, n).astype(np.uint64) if tie_mode == 2 else None
    idx = gg.DeviceIndex(0)
    perm = idx.build(chrom, start, end, 4, tie_mode=tie_mode, tie_key=tie)
    exp = expected_chain(chrom, start, end, tie_mode, tie)
    assert np.array_equal(perm, exp)
    _check(idx, chrom, start, end, 4, random_queries(rng, 10_000, 4, span=20_000), perm=exp)


def test_wide_coordinates_u64_layout(gg, oracle):
    rng = np.random.default_rng(11)
    chrom, start, end = random_index(rng, 30_000, 5, wide=True, nested=True)
    # coordinates up to SIZE_MAX (tests/data_type/interval_test.cpp:428-451)
    start[:3] = np.array([2**64 - 10, 2**64 - 1, 0], dtype=np.uint64)
    end[:3] = np.array([2**64 - 1, 2**64 - 1, 2**64 - 1], dtype=np.uint64)
    idx = gg.DeviceIndex(0)
    perm = idx.build(chrom, start, end, 5)
    assert idx.stats().coord_bytes == 8
    assert np.array_equal(perm, expected_chain(chrom, start, end))
    qc, qs, qe = random_queries(rng, 20_000, 5, wide=True)
    qs[:4] = np.array([2**64 - 1, 0, 2**64 - 5, 5], dtype=np.uint64)
    qe[:4] = np.array([2**64 - 1, 2**64 - 1, 2**64 - 2, 5], dtype=np.uint64)
    _check(idx, chrom, start, end, 5, (qc, qs, qe))


def test_u32_boundary_values(gg, oracle):
    # index stays in the u32 layout with coordinates right at 2^32-1; queries beyond it (u64) must clamp
    chrom = np.zeros(6, np.uint32)
    start = np.array([0, 0, 10, 2**32 - 2, 2**32 - 1, 2**32 - 1], np.uint64)
    end = np.array([0, 2**32 - 1, 20, 2**32 - 1, 2**32 - 1, 2**32 - 1], np.uint64)
    idx = gg.DeviceIndex(0)
    idx.build(chrom, start, end, 1)
    assert idx.stats().coord_bytes == 4
    qs = np.array([0, 2**32 - 1, 2**32, 2**32 - 1, 0, 21, 2**40], np.uint64)
    qe = np.array([0, 2**32 - 1, 2**33, 2**64 - 1, 2**64 - 1, 2**32 - 3, 2**41], np.uint64)
    _check(idx, chrom, start, end, 1, (np.zeros(7, np.uint32), qs, qe))


def test_upload_sorted_chain_and_errors(gg, oracle):
    rng = np.random.default_rng(21)
    chrom, start, end = random_index(rng, 10_000, 3, dup=True)
    # a host-defined chain: regular-insert order (reverse insertion among equal keys)
    perm = expected_chain(chrom, start, end, 1)
    idx = gg.DeviceIndex(0)
    idx.upload_sorted(chrom[perm], start[perm], end[perm], 3)
    q = random_queries(rng, 5000, 3)
    _check(idx, chrom, start, end, 3, q, perm=perm)
    _check(idx, chrom, start, end, 3, q, perm=perm, two_phase=True)
    with pytest.raises(gg.GGError) as ei:
        idx.upload_sorted(chrom, start, end, 3)  # not ordered
    assert ei.value.status == 6
    with pytest.raises(gg.GGError) as ei:
        idx.build(np.array([0], np.uint32), np.array([5], np.uint64), np.array([4], np.uint64), 1)
    assert ei.value.status == 4  # start > end, like interval's std::invalid_argument
    with pytest.raises(gg.GGError) as ei:
        idx.build(np.array([7], np.uint32), np.array([1], np.uint64), np.array([4], np.uint64), 3)
    assert ei.value.status == 1
    with pytest.raises(gg.GGError) as ei:
        idx.intersect_batch(np.zeros(1, np.uint32), np.zeros(1, np.uint64), np.zeros(1, np.uint64))
    assert ei.value.status == 5  # failed builds leave the handle without a snapshot


def test_invalid_query_is_rejected(gg, oracle):
    idx = gg.DeviceIndex(0)
    idx.build(np.zeros(3, np.uint32), np.array([1, 5, 9], np.uint64), np.array([3, 7, 11], np.uint64), 1)
    with pytest.raises(gg.GGError) as ei:
        idx.intersect_batch(np.zeros(2, np.uint32), np.array([1, 8], np.uint64), np.array([2, 6], np.uint64))
    assert ei.value.status == 4
    off, hits = idx.intersect_batch(np.zeros(0, np.uint32), np.zeros(0, np.uint64), np.zeros(0, np.uint64))
    assert off.tolist() == [0] and len(hits) == 0


@pytest.mark.parametrize("opt", [("chunk_queries", 4096), ("chunk_queries", 20_000), ("stage_bytes", 0),
                                 ("stage_bytes", 16384), ("stage_bytes", 196608), ("count_threads", 128),
                                 ("count_threads", 1024), ("count_mode", 1), ("count_mode", 0), ("count_mode", 2), ("fused_kernel", 1), ("fused_kernel", 0),
                                 ("emit_walk", 16), ("emit_tma", 0), ("emit_tma", 1), ("emit_tma", 2), ("host_mode", 1),
                                 ("host_mode", 0), ("fused_regs40", 1), ("fused_ctas_per_sm", 1)])
def test_options_do_not_change_results(gg, oracle, opt):
    rng = np.random.default_rng(31)
    chrom, start, end = random_index(rng, 150_000, 24, span=400_000, nested=True)
    idx = gg.DeviceIndex(0)
    if opt[0] in ("stage_bytes", "count_threads", "count_mode"):
        idx.set_option("host_mode", 1)  # these knobs belong to the count -> scan -> emit kernels
    if opt[0] in ("stage_bytes", "count_threads"):
        idx.set_option("count_mode", 1)  # ... and these to the tiered tree kernel
    if opt[0] in ("fused_regs40", "fused_ctas_per_sm"):
        idx.set_option("fused_kernel", 1)  # knobs of the one-kernel form of the single-call path
    idx.set_option(*opt)
    idx.build(chrom, start, end, 24)
    q = random_queries(rng, 100_001, 24, span=400_000)
    _check(idx, chrom, start, end, 24, q)
    if opt[0] == "chunk_queries":
        assert idx.batch_stats().n_chunks == -(-100_001 // (opt[1] // 4096 * 4096))
    # u32 query entry point on the same handle
    exp_off, exp_hits = oracle_csr(chrom, start, end, 24, *q)
    off, hits = idx.intersect_batch(q[0], q[1].astype(np.uint32), q[2].astype(np.uint32))
    assert np.array_equal(off, exp_off) and np.array_equal(hits, exp_hits)


@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("wide", [False, True])
def test_sorted_batches_take_the_bracket_paths(gg, oracle, mode, wide):
    """Coordinate-sorted batches (the BAM-like case): root directory (mode 0) and TMA window / warp tiers (mode 1)."""
    rng = np.random.default_rng(61 + mode)
    chrom, start, end = random_index(rng, 200_000, 3, span=3_000_000, max_len=4000, nested=True, wide=wide)
    idx = gg.DeviceIndex(0)
    idx.set_option("count_mode", mode)
    idx.set_option("host_mode", 1)
    idx.build(chrom, start, end, 3)
    qc, qs, qe = random_queries(rng, 150_000, 3, span=3_000_000, max_len=300, wide=wide)
    order = np.lexsort((qs, qc))
    qc, qs, qe = qc[order], qs[order], qe[order]
    qs[:300] = 0  # q.start == 0 edge (no end can be < 0)
    qe[:300] = np.sort(qe[:300])
    _check(idx, chrom, start, end, 3, (qc, qs, qe))
    # dense queries on a tiny span: every tile fits one window; and a stride that breaks windows
    sub = slice(0, 150_000, 7)
    _check(idx, chrom, start, end, 3, (qc[sub], qs[sub], qe[sub]))


def test_device_buffer_api(gg, oracle):
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(41)
    chrom, start, end = random_index(rng, 80_000, 24, span=300_000, nested=True)
    idx = gg.DeviceIndex(0)
    idx.set_option("timing", 1)
    idx.build(chrom, start, end, 24)
    qc, qs, qe = random_queries(rng, 200_000, 24, span=300_000)
    exp_off, exp_hits = oracle_csr(chrom, start, end, 24, qc, qs, qe)
    dev = torch.device("cuda:0")
    for dt, nb in ((np.uint32, 4), (np.uint64, 8)):
        d_c = torch.from_numpy(qc.view(np.int32)).to(dev)
        d_s = torch.from_numpy(qs.astype(dt).view(np.int32 if nb == 4 else np.int64)).to(dev)
        d_e = torch.from_numpy(qe.astype(dt).view(np.int32 if nb == 4 else np.int64)).to(dev)
        d_off = torch.empty(len(qc) + 1, dtype=torch.int64, device=dev)
        stream = torch.cuda.current_stream().cuda_stream
        total = idx.count_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), nb, len(qc), d_off.data_ptr(),
                                 stream, sync_total=True)
        assert total == int(exp_off[-1])
        d_hits = torch.empty(max(total, 1), dtype=torch.int32, device=dev)
        idx.emit_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), nb, len(qc), d_off.data_ptr(),
                        d_hits.data_ptr(), stream)
        torch.cuda.synchronize()
        assert np.array_equal(d_off.cpu().numpy().view(np.uint64), exp_off)
        assert np.array_equal(d_hits.cpu().numpy().view(np.uint32)[:total], exp_hits)
        bs = idx.batch_stats()
        assert bs.count_ms > 0 and bs.emit_ms > 0 and bs.kernel_launches == 4


def test_fused_device_api_and_capacity(gg, oracle):
    """gg_intersect_fused_device: one kernel; exact capacity, generous capacity, and a too-small buffer
    (GG_ERR_CAPACITY with complete offsets and nothing written at or beyond the capacity)."""
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(43)
    chrom, start, end = random_index(rng, 80_000, 24, span=300_000, nested=True)
    idx = gg.DeviceIndex(0)
    idx.build(chrom, start, end, 24)
    qc, qs, qe = random_queries(rng, 200_003, 24, span=300_000, extra_chrom=1)
    order = np.lexsort((qs, qc))  # half sorted, half shuffled
    h = len(qc) // 2
    qc[:h], qs[:h], qe[:h] = qc[order][:h], qs[order][:h], qe[order][:h]
    exp_off, exp_hits = oracle_csr(chrom, start, end, 24, qc, qs, qe)
    total = int(exp_off[-1])
    dev = torch.device("cuda:0")
    stream = torch.cuda.current_stream().cuda_stream
    for dt, nb in ((np.uint32, 4), (np.uint64, 8)):
        d_c = torch.from_numpy(qc.view(np.int32)).to(dev)
        d_s = torch.from_numpy(qs.astype(dt).view(np.int32 if nb == 4 else np.int64)).to(dev)
        d_e = torch.from_numpy(qe.astype(dt).view(np.int32 if nb == 4 else np.int64)).to(dev)
        for cap in (total, total + 12345):
            d_off = torch.zeros(len(qc) + 1, dtype=torch.int64, device=dev)
            d_hits = torch.full((cap + 64,), -1, dtype=torch.int32, device=dev)
            got = idx.fused_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), nb, len(qc), d_off.data_ptr(),
                                   d_hits.data_ptr(), cap, stream, sync_total=True)
            assert got == total
            assert idx.batch_stats().kernel_launches in (1, 3, 4)  # one look-back kernel (small batches), or count, block scan, emit_scan (+ emit_dense)
            assert np.array_equal(d_off.cpu().numpy().view(np.uint64), exp_off)
            h_hits = d_hits.cpu().numpy()
            assert np.array_equal(h_hits.view(np.uint32)[:total], exp_hits)
            assert (h_hits[total:] == -1).all()
        # too small: offsets complete, ranks below the capacity either correct or untouched, nothing beyond it
        cap = total // 3
        d_off = torch.zeros(len(qc) + 1, dtype=torch.int64, device=dev)
        d_hits = torch.full((total + 64,), -1, dtype=torch.int32, device=dev)
        with pytest.raises(gg.GGError) as ei:
            idx.fused_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), nb, len(qc), d_off.data_ptr(),
                             d_hits.data_ptr(), cap, stream, sync_total=True)
        assert ei.value.status == 7 and idx.last_total == total
        assert np.array_equal(d_off.cpu().numpy().view(np.uint64), exp_off)
        h_hits = d_hits.cpu().numpy()
        assert (h_hits[cap:] == -1).all()
        written = h_hits[:cap] != -1
        assert written.sum() > 0.9 * cap
        assert np.array_equal(h_hits[:cap][written].view(np.uint32), exp_hits[:cap][written])
    # enqueue-only form on a side stream, two batches back to back on one handle
    s2 = torch.cuda.Stream()
    with torch.cuda.stream(s2):
        d_off = torch.zeros(len(qc) + 1, dtype=torch.int64, device=dev)
        d_hits = torch.empty(total, dtype=torch.int32, device=dev)
        for _ in range(2):
            idx.fused_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 8, len(qc), d_off.data_ptr(),
                             d_hits.data_ptr(), total, s2.cuda_stream)
    s2.synchronize()
    assert np.array_equal(d_off.cpu().numpy().view(np.uint64), exp_off)
    assert np.array_equal(d_hits.cpu().numpy().view(np.uint32), exp_hits)


@pytest.mark.parametrize("path", ["batch8", "batch16", "one_kernel", "tma_bulk", "tma_vec"])
def test_single_call_paths_sorted_unaligned_and_ragged(gg, oracle, path):
    """The single-call device path (count, 4 consecutive queries per thread -> block scan -> emit_scan, with 8- and
    16-rank walk steps) and its one-kernel look-back form on: coordinate-sorted batches (finger search from the
    previous query of the thread), runs of identical queries, chromosome changes inside a thread's group, a
    descending run, pointers that are not 16-byte aligned (scalar load path) and batch sizes that leave ragged
    last tiles and blocks."""
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(71 + len(path))
    chrom, start, end = random_index(rng, 120_000, 5, span=2_000_000, max_len=5000, nested=True, dup=True)
    idx = gg.DeviceIndex(0)
    idx.set_option("fused_kernel", 1 if path == "one_kernel" else 0)
    idx.set_option("emit_walk", 16 if path == "batch16" else 8)
    # pass 2: the TMA kernel (ring of bulk loads in, bulk or vector stores out; emit_tma.cuh) or emit_scan (emit_scan.cuh)
    idx.set_option("emit_tma", {"tma_bulk": 1, "tma_vec": 2}.get(path, 0))
    idx.build(chrom, start, end, 5)
    qc, qs, qe = random_queries(rng, 90_007, 5, span=2_000_000, max_len=400, extra_chrom=1)
    order = np.lexsort((qs, qc))
    qc, qs, qe = qc[order], qs[order], qe[order]
    qs[1000:1400] = qs[1000]
    qe[1000:1400] = qe[1000]          # identical queries
    qc[5000:5003] = [4, 0, 4]         # chromosome flips inside a thread's consecutive group
    qs[20_000:20_050] = 0             # start 0: no end can lie before it
    qe[20_000:20_050] = np.sort(qe[20_000:20_050])
    sl = slice(30_000, 30_400)        # a descending run (finger must not be used)
    qs[sl], qe[sl] = qs[sl][::-1].copy(), qe[sl][::-1].copy()
    bad = qs > qe
    qe[bad] = qs[bad]
    exp_off, exp_hits = oracle_csr(chrom, start, end, 5, qc, qs, qe)
    dev = torch.device("cuda:0")
    stream = torch.cuda.current_stream().cuda_stream
    for nb, dt in ((4, np.uint32), (8, np.uint64)):
        for shift in (0, 1):          # shift 1: every pointer 4 (or 8) bytes past a 16-byte boundary
            for nq in (len(qc), 1, 127, 128, 129, 1023, 1024, 4099, 0):
                off_np = exp_off[:nq + 1]
                total = int(off_np[-1])
                d_c = torch.zeros(nq + 4, dtype=torch.int32, device=dev)
                d_s = torch.zeros(nq + 4, dtype=torch.int32 if nb == 4 else torch.int64, device=dev)
                d_e = torch.zeros_like(d_s)
                d_c[shift:shift + nq] = torch.from_numpy(qc[:nq].view(np.int32)).to(dev)
                d_s[shift:shift + nq] = torch.from_numpy(qs[:nq].astype(dt).view(np.int32 if nb == 4 else np.int64)).to(dev)
                d_e[shift:shift + nq] = torch.from_numpy(qe[:nq].astype(dt).view(np.int32 if nb == 4 else np.int64)).to(dev)
                d_off = torch.zeros(nq + 4, dtype=torch.int64, device=dev)
                d_hits = torch.full((total + 8,), -1, dtype=torch.int32, device=dev)
                got = idx.fused_device(d_c[shift:].data_ptr(), d_s[shift:].data_ptr(), d_e[shift:].data_ptr(), nb, nq,
                                       d_off[shift:].data_ptr(), d_hits.data_ptr(), total, stream, sync_total=True)
                assert got == total
                assert np.array_equal(d_off.cpu().numpy().view(np.uint64)[shift:shift + nq + 1], off_np)
                h = d_hits.cpu().numpy()
                assert np.array_equal(h.view(np.uint32)[:total], exp_hits[:total]) and (h[total:] == -1).all()


@pytest.mark.parametrize("shape", ["bed", "nested", "dense"])
def test_compact_result_form_decodes_to_the_same_csr(gg, oracle, shape):
    """gg_intersect_batch_compact: {offset, hi slot} per query + 1/2/4-byte back-deltas per hit decode (gg_compact_expand,
    gg_compact_rank / _count) to exactly the CSR of gg_intersect_batch.  bed: short intervals, 1-byte deltas; nested: long
    intervals push deltas past 255 / 65535, the chunk is re-run one width up; dense: lists longer than the stage."""
    import ctypes as C
    rng = np.random.default_rng({"bed": 91, "nested": 92, "dense": 93}[shape])
    if shape == "bed":
        chrom, start, end = random_index(rng, 200_000, 24, span=2_000_000, max_len=3000)
    elif shape == "nested":
        chrom, start, end = random_index(rng, 300_000, 3, span=3_000_000, max_len=400, nested=True)
        end[:3] = start[:3] + np.uint64(2_900_000)   # spans ~100 k ranks: needs 4-byte deltas
    else:
        n = 40_000
        start = np.sort(rng.integers(0, 100_000, n)).astype(np.uint64)
        end = start + rng.integers(1, 3000, n).astype(np.uint64)
        chrom = np.zeros(n, np.uint32)
    n_chrom = int(chrom.max()) + 1
    idx = gg.DeviceIndex(0)
    idx.set_option("chunk_queries", 8192)
    idx.build(chrom, start, end, n_chrom)
    span = int(end.max())
    qc, qs, qe = random_queries(rng, 50_003, n_chrom, span=span, max_len=300, extra_chrom=1)
    order = np.lexsort((qs, qc))
    qc, qs, qe = qc[order], qs[order], qe[order]
    exp_off, exp_hits = oracle_csr(chrom, start, end, n_chrom, qc, qs, qe)
    for rep in range(2):  # second call: warm buffers, sticky delta width
        off, hits = idx.intersect_batch_compact(qc, qs.astype(np.uint32), qe.astype(np.uint32), threads=1 + rep)
        assert np.array_equal(off, exp_off) and np.array_equal(hits, exp_hits)
    res = idx.last_compact
    widths = {res.chunks[k].delta_bytes for k in range(res.n_chunks)}
    assert widths <= {1, 2, 4} and res.n_chunks == -(-50_003 // 8192)
    if shape == "bed":
        assert widths == {1}
    if shape == "nested":
        assert max(widths) >= 2  # (some chunk met a delta past 255 and was re-run wider)
    # element-wise decode helpers on a sample
    L = idx._L
    for q in rng.integers(0, len(qc), 300):
        n = L.gg_compact_count(C.byref(res), int(q))
        assert n == int(exp_off[q + 1] - exp_off[q])
        got = [L.gg_compact_rank(C.byref(res), C.c_void_p(qc.ctypes.data), int(q), j) for j in range(n)]
        assert got == exp_hits[int(exp_off[q]):int(exp_off[q + 1])].tolist()
    # empty batch and invalid interval
    off, hits = idx.intersect_batch_compact(qc[:0], qs[:0].astype(np.uint32), qe[:0].astype(np.uint32))
    assert off.tolist() == [0] and len(hits) == 0
    bad_s = qs.astype(np.uint32).copy()
    bad_s[17] = np.uint32(qe[17] + 5)
    with pytest.raises(gg.GGError) as ei:
        idx.intersect_batch_compact(qc, bad_s, qe.astype(np.uint32))
    assert ei.value.status == 4


def test_tma_emit_many_blocks_per_warp(gg, oracle):
    """intersect_emit_tma_kernel with a grid pinned to one CTA per SM: every warp owns many 128-query blocks, so the TMA
    ring wraps and the block hop of its tile sequence runs (the default grid gives small batches one block per warp)."""
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(97)
    chrom, start, end = random_index(rng, 150_000, 4, span=2_000_000, max_len=3000, nested=True)
    exp = None
    for nbytes in (4, 8):
        for mode, cps in ((0, 0), (1, 1), (2, 1), (1, 0)):
            idx = gg.DeviceIndex(0)
            idx.set_option("fused_kernel", 0)
            idx.set_option("emit_tma", mode)
            idx.set_option("emit_ctas_per_sm", cps)
            idx.build(chrom, start, end, 4)
            qc, qs, qe = random_queries(np.random.default_rng(98), 1_300_021, 4, span=2_000_000, max_len=300)
            order = np.lexsort((qs, qc))
            qc, qs, qe = qc[order], qs[order], qe[order]
            dev = torch.device("cuda:0")
            dt, tt = (np.uint32, torch.int32) if nbytes == 4 else (np.uint64, torch.int64)
            d_c = torch.from_numpy(qc.view(np.int32)).to(dev)
            d_s = torch.from_numpy(qs.astype(dt).view(np.int32 if nbytes == 4 else np.int64)).to(dev)
            d_e = torch.from_numpy(qe.astype(dt).view(np.int32 if nbytes == 4 else np.int64)).to(dev)
            nq = len(qc)
            d_off = torch.zeros(nq + 1, dtype=torch.int64, device=dev)
            st = torch.cuda.current_stream().cuda_stream
            tot = idx.count_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), nbytes, nq, d_off.data_ptr(), st, sync_total=True)
            d_hits = torch.full((tot + 8,), -1, dtype=torch.int32, device=dev)
            d_off.zero_()
            got = idx.fused_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), nbytes, nq, d_off.data_ptr(), d_hits.data_ptr(), tot, st, sync_total=True)
            assert got == tot
            if exp is None:  # the emit_scan result, itself checked against the oracle on a prefix
                exp = (d_off.cpu().numpy(), d_hits.cpu().numpy())
                o_off, o_hits = oracle_csr(chrom, start, end, 4, qc[:20_000], qs[:20_000], qe[:20_000])
                assert np.array_equal(exp[0][:20_001].view(np.uint64), o_off)
                assert np.array_equal(exp[1][:len(o_hits)].view(np.uint32), o_hits)
            else:
                assert np.array_equal(d_off.cpu().numpy(), exp[0]) and np.array_equal(d_hits.cpu().numpy(), exp[1])


def test_fused_host_pipeline_regrows_on_dense_chunks(gg, oracle):
    """Host-buffer call, fused mode: chunks whose hit count exceeds the running estimate are re-run exactly;
    very long hit lists (> 512 per warp tile) take the direct-to-global walk."""
    rng = np.random.default_rng(47)
    n = 40_000
    start = np.sort(rng.integers(0, 100_000, n)).astype(np.uint64)
    end = start + rng.integers(1, 3000, n).astype(np.uint64)   # ~600 hits per query
    chrom = np.zeros(n, np.uint32)
    idx = gg.DeviceIndex(0)
    idx.set_option("chunk_queries", 8192)
    idx.build(chrom, start, end, 1)
    nq = 30_000
    qs = rng.integers(0, 200_000, nq).astype(np.uint64)  # second half of the span: no hits at all
    qs[:10_000] = np.sort(qs[:10_000])
    q = (np.zeros(nq, np.uint32), qs, qs + rng.integers(0, 50, nq).astype(np.uint64))
    hits = _check(idx, chrom, start, end, 1, q)
    assert hits > 50 * nq
    assert idx.batch_stats().n_chunks == 4
    _check(idx, chrom, start, end, 1, q)  # second call on warm buffers


def test_many_contigs_offset_table_outside_shared_memory(gg, oracle):
    # 5000 contigs x 2 levels: the level-offset table (40 KB) is read through L1 instead of shared memory
    rng = np.random.default_rng(53)
    n_chrom = 5000
    chrom, start, end = random_index(rng, 200_000, n_chrom, span=50_000)
    chrom[:2000] = 7  # one contig deep enough for a second level
    idx = gg.DeviceIndex(0)
    idx.build(chrom, start, end, n_chrom)
    _check(idx, chrom, start, end, n_chrom, random_queries(rng, 60_000, n_chrom, span=50_000))


@pytest.mark.parametrize("host_mode", [0, 1])
def test_scaffold_level_assembly_two_phase(gg, oracle, host_mode):
    # 60 000 contigs x 2 levels = 480 KB of level offsets: more than any CTA's shared memory.  The two-phase API
    # (gg_intersect_count / _emit, host_mode 1) must read the table through L1 like the single-call path does.
    rng = np.random.default_rng(59)
    n_chrom = 60_000
    chrom, start, end = random_index(rng, 300_000, n_chrom, span=20_000)
    chrom[:3000] = 11  # one contig deep enough for a second level
    idx = gg.DeviceIndex(0)
    idx.set_option("host_mode", host_mode)
    idx.build(chrom, start, end, n_chrom)
    q = random_queries(rng, 50_000, n_chrom, span=20_000)
    _check(idx, chrom, start, end, n_chrom, q)
    _check(idx, chrom, start, end, n_chrom, q, two_phase=True)
    idx.set_option("count_mode", 1)  # the TMA-window count kernel keeps the table in shared memory: falls back, same results
    _check(idx, chrom, start, end, n_chrom, q, two_phase=True)


def test_long_nested_intervals_use_max_end_pruning(gg, oracle):
    # one_spanning / nested shapes of tests/structure/routing_test.cpp:52-109 at scale: a few huge intervals
    # far to the left of dense short ones
    rng = np.random.default_rng(51)
    n = 300_000
    start = np.sort(rng.integers(0, 30_000_000, n)).astype(np.uint64)
    end = start + rng.integers(1, 200, n).astype(np.uint64)
    end[::50_000] = start[::50_000] + np.uint64(25_000_000)
    chrom = np.zeros(n, np.uint32)
    idx = gg.DeviceIndex(0)
    idx.build(chrom, start, end, 1)
    qs = rng.integers(0, 30_000_000, 50_000).astype(np.uint64)
    _check(idx, chrom, start, end, 1, (np.zeros(50_000, np.uint32), qs, qs + np.uint64(100)))


@pytest.mark.parametrize("budget", [0, 1, 3, 12])
@pytest.mark.parametrize("wide", [False, True])
def test_warp_cooperative_walk_on_skewed_lists(gg, oracle, budget, wide):
    """Dense-batch emit with the warp-cooperative walk (emit_scan.cuh, coop_walk): GENCODE-like nesting — a few very long
    intervals over many short ones, heavy duplicates, several chromosomes with ragged sizes (256-slot blocks straddle
    chromosome boundaries), queries at position 0 (pads hold end = 0), lists far larger than the stage (global form) —
    for per-lane budgets from "always cooperative" to the default, and budget 0 (per-lane only) as the control."""
    rng = np.random.default_rng(900 + budget + 10 * wide)
    n_chrom = 5
    sizes = [70_001, 33, 9_000, 1, 41_237]
    chrom = np.concatenate([np.full(m, c, np.uint32) for c, m in enumerate(sizes)])
    n = len(chrom)
    start = rng.integers(0, 2_000_000, n).astype(np.uint64)
    ln = rng.integers(0, 300, n).astype(np.uint64)
    nest = rng.random(n) < 0.01
    ln[nest] = rng.integers(100_000, 1_900_000, int(nest.sum())).astype(np.uint64)   # genes / transcripts
    ln[rng.random(n) < 0.0005] = 2_000_000                                            # spans everything after it
    start = (start // 16) * 16                                                        # duplicates
    if wide:
        start = start + np.uint64(2**36)
    end = start + ln
    sh = rng.permutation(n)
    chrom, start, end = chrom[sh], start[sh], end[sh]
    idx = gg.DeviceIndex(0)
    idx.set_option("dense_emit", 1)
    idx.set_option("coop_budget", budget)
    idx.set_option("fused_kernel", 0)
    perm = idx.build(chrom, start, end, n_chrom)
    assert np.array_equal(perm, expected_chain(chrom, start, end))
    nq = 60_000
    qc = rng.integers(0, n_chrom + 1, nq).astype(np.uint32)
    qs = rng.integers(0, 2_100_000, nq).astype(np.uint64)
    qs[::17] = 0
    qe = qs + rng.integers(0, 1000, nq).astype(np.uint64)
    qe[::1001] = qs[::1001] + np.uint64(1_500_000)   # a few enormous lists among short ones: tiles beyond the stage
    if wide:
        qs, qe = qs + np.uint64(2**36), qe + np.uint64(2**36)
        qs[::17] = 0
    _check(idx, chrom, start, end, n_chrom, (qc, qs, qe))
    o = np.lexsort((qs, qc))
    _check(idx, chrom, start, end, n_chrom, (qc[o], qs[o], qe[o]))


@pytest.mark.parametrize("wide", [False, True])
def test_emit_shared_candidate_list_per_tile(gg, oracle, wide):
    """emit_scan with one shared candidate list per sorted 32-query tile (option emit_tile_list): fixed-length sorted reads
    (every tile qualifies), equal starts, chromosome changes and unsorted stretches inside the batch (those tiles take the
    walk), nested long intervals, tiles whose list exceeds the 64-candidate cap, the u64 layout."""
    rng = np.random.default_rng(707 + wide)
    n_chrom = 3
    chrom, start, end = random_index(rng, 120_000, n_chrom, span=4_000_000, max_len=900, dup=True, nested=True, wide=wide)
    idx = gg.DeviceIndex(0)
    idx.set_option("fused_kernel", 0)
    idx.set_option("dense_emit", 0)
    idx.set_option("emit_tile_list", 1)
    idx.build(chrom, start, end, n_chrom)
    nq = 300_007
    off = np.uint64(2**40) if wide else np.uint64(0)
    qc = np.sort(rng.integers(0, n_chrom + 1, nq)).astype(np.uint32)
    qs = rng.integers(0, 4_100_000, nq).astype(np.uint64)
    qs[5000:5600] = 1234
    o = np.lexsort((qs, qc))
    qc, qs = qc[o], qs[o]
    qs[100_000:101_000] = rng.permutation(qs[100_000:101_000])   # an unsorted stretch
    qs = qs + off
    qs[::97] = 0                                                   # position 0 (pads hold end = 0), breaks the order too
    qe = qs + np.uint64(149)
    qe[200_000:200_064] += np.uint64(40_000)                       # lists beyond the cap
    total = _check(idx, chrom, start, end, n_chrom, (qc, qs, qe))
    assert total > nq


@pytest.mark.parametrize("sorted_batch", [True, False])
def test_emit_vector_copy_out(gg, oracle, sorted_batch):
    """emit_scan with the phase-aligned stage and 16-byte copy-out (option emit_vec_out): every alignment of a tile's first
    hit (0, 4, 8, 12 bytes), tiles of 0 / 1 / 2 / 3 hits (head or tail only), tiles near the stage size."""
    rng = np.random.default_rng(404 + sorted_batch)
    n_chrom = 3
    chrom, start, end = random_index(rng, 90_000, n_chrom, span=900_000, max_len=400, dup=True, nested=True)
    idx = gg.DeviceIndex(0)
    idx.set_option("fused_kernel", 0)
    idx.set_option("dense_emit", 0)
    idx.set_option("emit_vec_out", 1)
    idx.build(chrom, start, end, n_chrom)
    nq = 200_003
    qc, qs, qe = random_queries(rng, nq, n_chrom, span=900_000, max_len=60, extra_chrom=1)
    qe[::7] = qs[::7]                                    # point queries: many empty / single-hit lists
    qe[::501] = qs[::501] + np.uint64(30_000)            # a few long lists: tiles near and above the 512-hit stage
    if sorted_batch:
        o = np.lexsort((qs, qc))
        qc, qs, qe = qc[o], qs[o], qe[o]
    total = _check(idx, chrom, start, end, n_chrom, (qc, qs, qe))
    assert total > nq


@pytest.mark.parametrize("delta_min", [1, 512, 0])
@pytest.mark.parametrize("wide", [False, True])
def test_sorted_tile_candidate_list(gg, oracle, delta_min, wide):
    """Dense-batch emit, sorted tiles (emit_scan.cuh, emit_tile_delta): one candidate list per 32-query tile.  Fixed-length
    reads sorted by start on a dense index (every tile qualifies), with equal starts, chromosome changes inside tiles
    (those tiles fall back), jumps in the query stream larger than the stage holds (fall back), nested long intervals,
    the u64 layout, and a hit buffer that is too small.  delta_min 1 forces the path for every sorted tile, 0 is the control."""
    import torch
    from genogrove_b200 import GGError
    rng = np.random.default_rng(606 + wide)
    n_chrom = 4
    sizes = [60_000, 257, 30_011, 12_000]
    chrom = np.concatenate([np.full(m, c, np.uint32) for c, m in enumerate(sizes)])
    n = len(chrom)
    start = rng.integers(0, 300_000, n).astype(np.uint64)
    ln = np.exp(rng.uniform(np.log(20), np.log(4000), n)).astype(np.uint64)
    ln[rng.random(n) < 0.002] = 250_000
    start[rng.random(n) < 0.2] //= np.uint64(64)
    start[rng.random(n) < 0.2] *= np.uint64(1)
    off = np.uint64(2**35) if wide else np.uint64(0)
    start, end = start + off, start + ln + off
    sh = rng.permutation(n)
    chrom, start, end = chrom[sh], start[sh], end[sh]
    idx = gg.DeviceIndex(0)
    idx.set_option("dense_emit", 1)
    idx.set_option("fused_kernel", 0)
    idx.set_option("delta_min_hits", delta_min)
    idx.build(chrom, start, end, n_chrom)
    nq = 50_001
    qc = np.sort(rng.integers(0, n_chrom, nq)).astype(np.uint32)
    qs = rng.integers(0, 310_000, nq).astype(np.uint64)
    qs[rng.random(nq) < 0.3] //= np.uint64(16)          # runs of equal starts, and a cluster near 0
    qs[1000:1400] = 5                                   # 400 identical queries
    o = np.lexsort((qs, qc))
    qc, qs = qc[o], qs[o]
    qs[20_000:20_032] += np.uint64(200_000)             # a jump the stage cannot hold (still sorted inside its tile?)
    qs = qs + off
    qe = qs + np.uint64(149)
    exp_total = _check(idx, chrom, start, end, n_chrom, (qc, qs, qe))
    assert exp_total / nq > 30
    # the same batch with a hit buffer that is too small: offsets complete, nothing beyond the capacity
    if not wide:
        dev = torch.device("cuda:0")
        d = [torch.from_numpy(a).to(dev) for a in (qc.view(np.int32), qs.astype(np.uint32).view(np.int32), qe.astype(np.uint32).view(np.int32))]
        d_off = torch.empty(nq + 1, dtype=torch.int64, device=dev)
        cap = exp_total // 2
        d_hits = torch.full((cap + 64,), -1, dtype=torch.int32, device=dev)
        with pytest.raises(GGError) as ei:
            idx.fused_device(d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), 4, nq, d_off.data_ptr(), d_hits.data_ptr(), cap,
                             torch.cuda.current_stream().cuda_stream, sync_total=True)
        assert ei.value.status == 7 and idx.last_total == exp_total
        got = d_hits.cpu().numpy()
        assert (got[cap:] == -1).all()
        exp_off, exp_hits = oracle_csr(chrom, start, end, n_chrom, qc, qs, qe)
        assert np.array_equal(d_off.cpu().numpy().view(np.uint64), exp_off)
        last_full = (int(np.searchsorted(exp_off, cap, side="right")) - 1) // 128 * 128  # every complete block below the capacity is exact
        upto = int(exp_off[last_full]) if last_full > 0 else 0
        assert upto > cap // 2 and np.array_equal(got[:upto].view(np.uint32), exp_hits[:upto])


def test_two_phase_device_emit_rejects_foreign_offsets(gg, oracle):
    """gg_intersect_emit_device gets the offsets back from the caller: offsets that are not the ones the count pass wrote
    must neither corrupt the per-warp stage nor pass silently — the kernel skips the affected tiles, and with the
    "debug_checks" option the call synchronises and fails with GG_ERR_INVALID_ARG."""
    import torch
    from genogrove_b200 import GGError
    rng = np.random.default_rng(808)
    chrom, start, end = random_index(rng, 60_000, 2, span=500_000, max_len=800, nested=True)
    idx = gg.DeviceIndex(0)
    idx.build(chrom, start, end, 2)
    nq = 100_000
    qc, qs, qe = random_queries(rng, nq, 2, span=500_000, max_len=600)
    exp_off, exp_hits = oracle_csr(chrom, start, end, 2, qc, qs, qe)
    dev = torch.device("cuda:0")
    d = [torch.from_numpy(a).to(dev) for a in (qc.view(np.int32), qs.astype(np.uint32).view(np.int32), qe.astype(np.uint32).view(np.int32))]
    d_off = torch.empty(nq + 1, dtype=torch.int64, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    total = idx.count_device(d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), 4, nq, d_off.data_ptr(), st, sync_total=True)
    assert total == len(exp_hits)
    d_hits = torch.full((total,), -1, dtype=torch.int32, device=dev)
    idx.set_option("debug_checks", 1)
    idx.emit_device(d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), 4, nq, d_off.data_ptr(), d_hits.data_ptr(), st)  # untouched: fine
    assert np.array_equal(d_hits.cpu().numpy().view(np.uint32), exp_hits)
    bad = d_off.clone()
    bad[50_001:50_020] += 40_000       # lists that claim to be far longer than the stage
    bad[70_000] -= 3                   # and one that shrinks / goes backwards
    d_hits.fill_(-1)
    with pytest.raises(GGError) as ei:
        idx.emit_device(d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), 4, nq, bad.data_ptr(), d_hits.data_ptr(), st)
    assert ei.value.status == 1
    got = d_hits.cpu().numpy().view(np.uint32)
    lo, hi = int(exp_off[49_920]), int(exp_off[50_176])   # the 256-query tile around the first tampering is left alone ...
    assert np.array_equal(got[:lo], exp_hits[:lo])       # ... everything before it is exact
    assert (got[lo:hi] == 0xFFFFFFFF).any()
    idx.set_option("debug_checks", 0)
    idx.count_device(d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), 4, nq, d_off.data_ptr(), st, sync_total=True)
    idx.emit_device(d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), 4, nq, bad.data_ptr(), d_hits.data_ptr(), st)  # asynchronous form: no crash
    torch.cuda.synchronize()


def test_smoke_entry(gg, oracle):
    import __graft_entry__ as ge
    ge.smoke()


@pytest.mark.parametrize("n,span,max_len,q", [(3000, 4000, 3000, 5000), (40000, 200000, 30000, 20001), (600, 2000, 1500, 2000),
                                              (20000, 400000, 1500, 20000), (30000, 300000, 900, 30001)])
def test_dense_tiles_windowed_emit(n, span, max_len, q):
    """Warp tiles with far more hits than the 512-entry stage (hundreds to thousands of hits per query): written by
    intersect_emit_dense_kernel in windows.  Interval and stranded paths, sorted and random batch order, and the
    capacity contract (offsets complete, ranks beyond the buffer dropped)."""
    import torch
    from genogrove_b200 import DeviceIndex, GGError, TIE_EXPLICIT
    rng = np.random.default_rng(n + q)
    chrom, s, e = random_index(rng, n, 2, span=span, max_len=max_len, dup=n == 600)
    st = np.frombuffer(b"+-.*", dtype=np.uint8)[rng.integers(0, 4, n)]
    qc, qs, qe = random_queries(rng, q, 2, span=span, max_len=max(2, max_len // 3), extra_chrom=1)
    qst = np.frombuffer(b"+-.*", dtype=np.uint8)[rng.integers(0, 4, q)]
    idx = DeviceIndex(0)
    idx.set_option("dense_emit", 1 if n != 600 else -1)  # forced on; the last case relies on the density seen so far
    so = np.array([O.STRAND_ORDER[chr(c)] for c in st], dtype=np.uint64)
    perm = idx.build(chrom, s, e, 2, tie_mode=TIE_EXPLICIT, tie_key=(so << np.uint64(32)) | np.arange(n, dtype=np.uint64))
    idx.set_strands(st[perm])
    coff = chrom_offsets(chrom, 2)
    for order in ("random", "sorted"):
        if order == "sorted":
            o = np.lexsort((qs, qc))
            qc, qs, qe, qst = qc[o], qs[o], qe[o], qst[o]
        eo, eh = O.flat_intersect_batch(s[perm], e[perm], coff, qc, qs, qe)
        assert len(eh) / q > 16  # most 32-query tiles exceed the 512-entry stage; the densest reach the dense kernel's 2048
        off, hits = idx.intersect_batch(qc, qs, qe)
        assert np.array_equal(off, eo) and np.array_equal(hits, eh)
        so_, sh_ = O.flat_intersect_batch_stranded(s[perm], e[perm], st[perm], coff, qc, qs, qe, qst)
        off, hits = idx.intersect_batch_stranded(qc, qs, qe, qst)
        assert np.array_equal(off, so_) and np.array_equal(hits, sh_)
    # device call with a buffer that is too small: offsets complete, the prefix that fits is exact
    dev = torch.device("cuda:0")
    d_c = torch.from_numpy(qc.view(np.int32)).to(dev)
    d_s = torch.from_numpy(qs.astype(np.uint32).view(np.int32)).to(dev)
    d_e = torch.from_numpy(qe.astype(np.uint32).view(np.int32)).to(dev)
    d_off = torch.empty(q + 1, dtype=torch.int64, device=dev)
    cap = len(eh) // 3
    d_hits = torch.full((cap + 64,), -1, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    with pytest.raises(GGError) as ei:
        idx.fused_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, q, d_off.data_ptr(), d_hits.data_ptr(), cap, stream,
                         sync_total=True)
    assert ei.value.status == 7 and idx.last_total == len(eh)
    assert np.array_equal(d_off.cpu().numpy().view(np.uint64), eo)
    got = d_hits.cpu().numpy()
    assert (got[cap:] == -1).all()  # nothing written beyond the capacity
    # every complete 32-query tile below the capacity is exact
    last_full = int(np.searchsorted(eo, cap, side="right")) - 1
    last_full = (last_full // 128) * 128
    upto = int(eo[last_full]) if last_full > 0 else 0
    assert np.array_equal(got[:upto].view(np.uint32), eh[:upto])
