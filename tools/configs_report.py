#!/usr/bin/env python
"""Measured throughput of the batch intersect path on every BASELINE.json configuration that fits one GPU
(cfg1-4 at full size; cfg5 as rank 0's chromosome share of the 8-GPU job), next to the unmodified reference
(oracle/_ref, all host threads) on a bounded sample of the same workload.  One JSON line per configuration.
Development / reporting aid: the headline metric stays bench.py's (cfg4)."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from genogrove_b200 import DeviceIndex, shard, synth  # noqa: E402

dev = torch.device("cuda:0")
want = sys.argv[1:] or ["cfg1", "cfg2", "cfg2s", "cfg3", "cfg4", "cfg4r", "cfg5"]


def cfg5_share():
    n_total, q_total, world = 200_000_000, 1_000_000_000, 8
    per = synth._split_by_length(n_total, synth.HG38_LENGTHS)
    owner = shard.assign_chromosomes(per, world)
    mine = np.flatnonzero(owner == 0)
    rng = np.random.default_rng(77)
    qper = synth._split_by_length(q_total, synth.HG38_LENGTHS)
    chrom, start, end, qc, qs, qe = [], [], [], [], [], []
    for local, c in enumerate(mine):
        L, n, m = int(synth.HG38_LENGTHS[c]), int(per[c]), int(qper[c])
        ln = np.exp(rng.uniform(np.log(100), np.log(10000), n)).astype(np.int64)
        st = (rng.random(n) * (L - ln - 1)).astype(np.int64)
        chrom.append(np.full(n, local, np.uint32)), start.append(st.astype(np.uint64)), end.append((st + ln - 1).astype(np.uint64))
        q0 = np.sort(rng.integers(0, L - 150, m, dtype=np.int64)).astype(np.uint64)
        qc.append(np.full(m, local, np.uint32)), qs.append(q0), qe.append(q0 + np.uint64(149))
    chrom, start, end = np.concatenate(chrom), np.concatenate(start), np.concatenate(end)
    sh = rng.permutation(len(chrom))
    return synth.Workload("cfg5 (rank 0 of 8: %d chromosomes)" % len(mine), len(mine), chrom[sh], start[sh], end[sh],
                          np.concatenate(qc), np.concatenate(qs), np.concatenate(qe))


MAKE = {
    "cfg1": lambda: synth.cfg1(),
    "cfg2": lambda: synth.cfg2(sorted_queries=False),
    "cfg2s": lambda: synth.cfg2(sorted_queries=True),
    "cfg3": lambda: synth.cfg3(),
    "cfg4": lambda: synth.cfg4(sorted_queries=True),
    "cfg4r": lambda: synth.cfg4(sorted_queries=False),
    "cfg5": cfg5_share,
}

for name in want:
    w = MAKE[name]()
    nq, n = len(w.q_chrom), len(w.chrom)
    idx = DeviceIndex(0)
    for kv in os.environ.get("GG_OPTS", "").split(","):  # e.g. GG_OPTS=fused_kernel=1,dense_emit=0
        if kv:
            idx.set_option(kv.split("=")[0], int(kv.split("=")[1]))
    idx.build(w.chrom, w.start, w.end, w.n_chrom, want_perm=False)
    idx.build(w.chrom, w.start, w.end, w.n_chrom, want_perm=False)
    ist = idx.stats()
    d_c = torch.from_numpy(w.q_chrom.view(np.int32)).to(dev)
    d_s = torch.from_numpy(w.q_start.astype(np.uint32).view(np.int32)).to(dev)
    d_e = torch.from_numpy(w.q_end.astype(np.uint32).view(np.int32)).to(dev)
    d_off = torch.empty(nq + 1, dtype=torch.int64, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    total = idx.count_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, d_off.data_ptr(), st, sync_total=True)
    d_hits = torch.empty(max(total, 1), dtype=torch.int32, device=dev)
    steps = 10 if nq >= 1_000_000 else 50
    for _ in range(3):
        idx.fused_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, d_off.data_ptr(), d_hits.data_ptr(), d_hits.numel(), st, sync_total=True)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        idx.fused_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, d_off.data_ptr(), d_hits.data_ptr(), d_hits.numel(), st, sync_total=True)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / steps
    row = {"config": w.name, "intervals": n, "queries": nq, "hits": total, "hits_per_query": total / max(nq, 1),
           "ms_per_step": ms, "queries_per_sec": nq / (ms / 1e3), "hit_pairs_per_sec": total / (ms / 1e3),
           "path_bytes": 12 * nq + 8 * (nq + 1) + 4 * total, "path_gbs": (12 * nq + 8 * (nq + 1) + 4 * total) / (ms / 1e3) / 1e9,
           "build_total_ms": ist.total_ms, "build_sort_ms": ist.sort_ms, "build_levels_ms": ist.levels_ms, "index_mb": ist.device_bytes / 1e6}
    del d_hits, d_off, d_c, d_s, d_e
    torch.cuda.empty_cache()
    # reference on a bounded sample (middle of the batch), all host threads
    try:
        import oracle as O
        if O.ref_available() and os.environ.get("NO_REF") != "1":
            g = O.RefGrove(32)
            t0 = time.time()
            ids = np.arange(n, dtype=np.uint32)
            for c in range(w.n_chrom):
                m = w.chrom == c
                g.insert_bulk(c, w.start[m], w.end[m], ids[m], presorted=False)
            S = min(nq, 2_000_000)
            lo = (nq - S) // 2
            off, _, secs = g.intersect_batch(w.q_chrom[lo:lo + S], w.q_start[lo:lo + S], w.q_end[lo:lo + S],
                                             threads=g.hardware_threads(), want_ids=False)
            row["reference_cpu"] = {"queries_per_sec": S / secs, "sample": S, "threads": g.hardware_threads(), "order": 32,
                                    "build_s": time.time() - t0 - secs}
            g.close()
    except Exception as ex:  # noqa: BLE001
        row["reference_cpu"] = {"error": repr(ex)}
    print(json.dumps(row), flush=True)
    del idx
