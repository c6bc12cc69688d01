#!/usr/bin/env python
"""GPU tuning sweep (development aid, not part of the product): device time of gg_intersect_fused_device (the
single-call path: count -> block scan -> emit_scan, or the one-kernel look-back form) and of the two-phase API's
kernels on the cfg4 workload; every variant's offsets and hits are compared on the device with the result of the
one-query-per-thread two-phase kernels."""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from genogrove_b200 import DeviceIndex, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--intervals", type=int, default=10_000_000)
ap.add_argument("--queries", type=int, default=100_000_000)
ap.add_argument("--variants", default="0:8,0:16,1:8", help="one_kernel:walk[:ctas_per_sm],... (single-call device path)")
ap.add_argument("--two-phase", default="", help="count_mode,... (times the count / scan / emit kernels of the two-phase API)")
ap.add_argument("--shuffled", action="store_true")
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--qbytes", type=int, default=4)
args = ap.parse_args()

dev = torch.device("cuda:0")
c, s, e = synth.hg38_intervals(args.intervals, 42)
qc, qs, qe = synth.hg38_queries(args.queries, 150, 43, not args.shuffled)
nq = len(qc)
idx = DeviceIndex(0)
idx.build(c, s, e, 24, want_perm=False)
idx.set_option("count_mode", 0)  # the reference result comes from the one-query-per-thread kernels
nb = args.qbytes
dt, tt = (np.uint32, np.int32) if nb == 4 else (np.uint64, np.int64)
d_c = torch.from_numpy(qc.view(np.int32)).to(dev)
d_s = torch.from_numpy(qs.astype(dt).view(tt)).to(dev)
d_e = torch.from_numpy(qe.astype(dt).view(tt)).to(dev)
d_off = torch.empty(nq + 1, dtype=torch.int64, device=dev)
st = torch.cuda.current_stream().cuda_stream
total = idx.count_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), nb, nq, d_off.data_ptr(), st, sync_total=True)
d_hits = torch.empty(max(total, 1), dtype=torch.int32, device=dev)
idx.emit_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), nb, nq, d_off.data_ptr(), d_hits.data_ptr(), st)
torch.cuda.synchronize()
ref_off, ref_hits = d_off.clone(), d_hits.clone()
idx.set_option("timing", 1)
print(f"nq={nq} hits={total} shuffled={args.shuffled} qbytes={nb}", flush=True)
for v in [x for x in args.two_phase.split(",") if x]:
    mode = int(v)
    idx.set_option("count_mode", mode)
    cm, sm, em = [], [], []
    for r in range(args.reps + 1):
        d_off.zero_()
        d_hits.fill_(-1)
        idx.count_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), nb, nq, d_off.data_ptr(), st, sync_total=True)
        idx.emit_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), nb, nq, d_off.data_ptr(), d_hits.data_ptr(), st)
        torch.cuda.synchronize()
        b = idx.batch_stats()
        if r:
            cm.append(b.count_ms), sm.append(b.scan_ms), em.append(b.emit_ms)
    ok = bool(torch.equal(d_off, ref_off)) and bool(torch.equal(d_hits, ref_hits))
    tot = np.median(cm) + np.median(sm) + np.median(em)
    print(f"two-phase count_mode={mode}: count {np.median(cm):6.3f} scan {np.median(sm):6.3f} emit {np.median(em):6.3f} "
          f"= {tot:6.3f} ms  {nq / tot / 1e6:8.2f} Gq/s  parity={'ok' if ok else 'MISMATCH'}", flush=True)
for v in [x for x in args.variants.split(",") if x]:
    f = [int(x) for x in v.split(":")] + [0, 0]
    one_kernel, walk, ctas = f[0], f[1] or 8, f[2]
    idx.set_option("fused_kernel", one_kernel)
    idx.set_option("emit_walk", walk)
    idx.set_option("fused_ctas_per_sm", ctas)
    idx.set_option("emit_ctas_per_sm", ctas)
    ms, parts = [], []
    for r in range(args.reps + 1):
        d_off.zero_()
        d_hits.fill_(-1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        idx.fused_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), nb, nq, d_off.data_ptr(), d_hits.data_ptr(),
                         d_hits.numel(), st)
        e1.record()
        torch.cuda.synchronize()
        if r:
            b = idx.batch_stats()
            ms.append(e0.elapsed_time(e1))
            parts.append((b.count_ms, b.scan_ms, b.emit_ms))
    ok = bool(torch.equal(d_off, ref_off)) and bool(torch.equal(d_hits, ref_hits))
    pm = np.median(np.array(parts), axis=0)
    print(f"single-call one_kernel={one_kernel} walk={walk} ctas_per_sm={ctas}: {np.median(ms):7.3f} ms (min {min(ms):7.3f}; count "
          f"{pm[0]:.3f} scan {pm[1]:.3f} emit {pm[2]:.3f})  {nq / np.median(ms) / 1e6:8.2f} Gq/s  parity={'ok' if ok else 'MISMATCH'}",
          flush=True)
