"""gg_intersect_fused_device_static — the single call for callers that cannot allocate device memory.

Who that is: a service that gave every byte of its HBM to its own pool when it started (a cudaMalloc issued by this
library would fail there), a process whose free GPU memory is already held by another framework's caching allocator,
and a latency-bound query loop that cannot afford cudaMalloc's implicit device synchronisation.  For them the library
keeps the scratch of the one-kernel path (tile ticket, look-back words, error / total mailbox) in a fixed array on the
device, which caps a call at STATIC_MAX_QUERIES queries and makes overlapping calls on one GPU race.

The check: through the C ABI, the static entry point returns exactly what gg_intersect_fused_device returns (offsets
and ordered ranks, bit for bit — integer work, no tolerance) on two calls in a row with different batches, the second
smaller than the first so that anything the first call left in the shared array would show.  The file is named to be
collected last: under `pytest -x` a failure here hides no other test.  It builds its own handle and needs nothing from
the tests before it, so it passes alone in a fresh process as well as at the end of the whole suite.
"""
import numpy as np
import pytest

from helpers import random_index, random_queries

pytestmark = pytest.mark.gpu


def _device_queries(torch, dev, qc, qs, qe, nbytes):
    dt = np.uint32 if nbytes == 4 else np.uint64
    view = np.int32 if nbytes == 4 else np.int64
    return (torch.from_numpy(qc.view(np.int32)).to(dev),
            torch.from_numpy(qs.astype(dt).view(view)).to(dev),
            torch.from_numpy(qe.astype(dt).view(view)).to(dev))


def _run(torch, dev, call, d_q, nbytes, nq, cap, stream):
    d_off = torch.full((nq + 1,), -1, dtype=torch.int64, device=dev)
    d_hits = torch.full((cap + 64,), -1, dtype=torch.int32, device=dev)
    total = call(d_q[0].data_ptr(), d_q[1].data_ptr(), d_q[2].data_ptr(), nbytes, nq, d_off.data_ptr(), d_hits.data_ptr(),
                 cap, stream, sync_total=True)
    return total, d_off.cpu().numpy().view(np.uint64), d_hits.cpu().numpy()


def test_static_scratch_call_matches_regular_call_twice_in_a_row():
    torch = pytest.importorskip("torch")
    import genogrove_b200 as gg

    rng = np.random.default_rng(2026)
    n_chrom = 24
    chrom, start, end = random_index(rng, 120_000, n_chrom, span=400_000, nested=True)
    idx = gg.DeviceIndex(0)
    idx.build(chrom, start, end, n_chrom)
    dev = torch.device("cuda:0")
    stream = torch.cuda.current_stream().cuda_stream

    # two different batches, the larger one first; half of each coordinate-sorted, half shuffled; one unknown chromosome id
    batches = []
    for nq in (300_007, 70_001):
        qc, qs, qe = random_queries(rng, nq, n_chrom, span=400_000, extra_chrom=1)
        order = np.lexsort((qs, qc))
        h = nq // 2
        qc[:h], qs[:h], qe[:h] = qc[order][:h], qs[order][:h], qe[order][:h]
        batches.append((qc, qs, qe))

    for nbytes in (4, 8):
        for qc, qs, qe in batches:  # consecutive calls on the shared array: no clean-up of ours in between
            nq = len(qc)
            d_q = _device_queries(torch, dev, qc, qs, qe, nbytes)
            # the regular call, twice: with no hit buffer to learn the total (status 7 carries it), then for the result
            idx.set_option("fused_kernel", 1)  # the same kernel, scratch on the handle
            with pytest.raises(gg.GGError) as sized:
                _run(torch, dev, idx.fused_device, d_q, nbytes, nq, 0, stream)
            assert sized.value.status == 7
            exp_total = idx.last_total
            exp_total2, exp_off, exp_hits = _run(torch, dev, idx.fused_device, d_q, nbytes, nq, exp_total, stream)
            assert exp_total2 == exp_total and exp_total > nq
            idx.set_option("fused_kernel", -1)

            total, off, hits = _run(torch, dev, idx.fused_device_static, d_q, nbytes, nq, exp_total, stream)
            assert total == exp_total
            assert np.array_equal(off, exp_off)
            assert np.array_equal(hits[:total], exp_hits[:total])
            assert (hits[total:] == -1).all()  # nothing past the last hit
            assert idx.batch_stats().kernel_launches == 1

    # status codes follow the regular call: a too-small hit buffer reports the needed size, offsets stay complete
    qc, qs, qe = batches[1]
    d_q = _device_queries(torch, dev, qc, qs, qe, 4)
    with pytest.raises(gg.GGError) as ei:
        _run(torch, dev, idx.fused_device_static, d_q, 4, len(qc), exp_total // 2, stream)
    assert ei.value.status == 7 and idx.last_total == exp_total
    # ... and the call after a failed one is clean again
    total, off, hits = _run(torch, dev, idx.fused_device_static, d_q, 4, len(qc), exp_total, stream)
    assert total == exp_total and np.array_equal(off, exp_off) and np.array_equal(hits[:total], exp_hits[:total])

    # an empty batch: offsets[0] = 0, nothing else
    total, off, _ = _run(torch, dev, idx.fused_device_static, d_q, 4, 0, 0, stream)
    assert total == 0 and off[0] == 0


def test_static_scratch_call_rejects_batches_over_the_limit():
    torch = pytest.importorskip("torch")
    import genogrove_b200 as gg
    from genogrove_b200 import _lib

    rng = np.random.default_rng(7)
    chrom, start, end = random_index(rng, 1000, 1)
    idx = gg.DeviceIndex(0)
    idx.build(chrom, start, end, 1)
    dev = torch.device("cuda:0")
    nq = _lib.STATIC_MAX_QUERIES + 1
    d_c = torch.zeros(nq, dtype=torch.int32, device=dev)
    d_s = torch.zeros(nq, dtype=torch.int32, device=dev)
    d_e = torch.zeros(nq, dtype=torch.int32, device=dev)
    d_off = torch.full((nq + 1,), -1, dtype=torch.int64, device=dev)
    d_hits = torch.full((16,), -1, dtype=torch.int32, device=dev)
    with pytest.raises(gg.GGError) as ei:
        idx.fused_device_static(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, d_off.data_ptr(), d_hits.data_ptr(), 16,
                                sync_total=True)
    assert ei.value.status == 7 and "GG_STATIC_MAX_QUERIES" in str(ei.value)
    torch.cuda.synchronize()
    assert (d_off.cpu().numpy() == -1).all()  # rejected before anything was launched
    # exactly at the limit the call runs
    nq = _lib.STATIC_MAX_QUERIES
    cap = 64 * 1024 * 1024
    d_hits = torch.empty(cap, dtype=torch.int32, device=dev)
    total = idx.fused_device_static(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, d_off.data_ptr(), d_hits.data_ptr(),
                                    cap, sync_total=True)
    ref_off = torch.full((nq + 1,), -1, dtype=torch.int64, device=dev)
    idx.set_option("fused_kernel", 1)
    ref_total = idx.fused_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, ref_off.data_ptr(), d_hits.data_ptr(), cap,
                                 sync_total=True)
    assert total == ref_total
    assert torch.equal(d_off[: nq + 1], ref_off)
