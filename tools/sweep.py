#!/usr/bin/env python
"""GPU tuning sweep (development aid, not part of the product): kernel times of the device-buffer path for
combinations of the library's tuning options on the cfg4 workload."""
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
ap.add_argument("--threads", default="128,256,512,1024")
ap.add_argument("--stage", default="0,16384,98304")
ap.add_argument("--shuffled", action="store_true")
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--mode", type=int, default=0)
args = ap.parse_args()

dev = torch.device("cuda:0")
c, s, e = synth.hg38_intervals(args.intervals, 42)
qc, qs, qe = synth.hg38_queries(args.queries, 150, 43, not args.shuffled)
nq = len(qc)
idx = DeviceIndex(0)
idx.set_option("count_mode", args.mode)
idx.build(c, s, e, 24, want_perm=False)
d_c = torch.from_numpy(qc.view(np.int32)).to(dev)
d_s = torch.from_numpy(qs.astype(np.uint32).view(np.int32)).to(dev)
d_e = torch.from_numpy(qe.astype(np.uint32).view(np.int32)).to(dev)
d_off = torch.empty(nq + 1, dtype=torch.int64, device=dev)
st = torch.cuda.current_stream().cuda_stream
total = idx.count_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, d_off.data_ptr(), st, sync_total=True)
d_hits = torch.empty(max(total, 1), dtype=torch.int32, device=dev)
idx.set_option("timing", 1)
print(f"nq={nq} hits={total} shuffled={args.shuffled}")
for stage in [int(x) for x in args.stage.split(",")]:
    idx.set_option("stage_bytes", stage)
    for th in [int(x) for x in args.threads.split(",")]:
        idx.set_option("count_threads", th)
        cm, em = [], []
        for r in range(args.reps + 1):
            idx.count_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, d_off.data_ptr(), st, sync_total=True)
            idx.emit_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, d_off.data_ptr(), d_hits.data_ptr(), st)
            torch.cuda.synchronize()
            b = idx.batch_stats()
            if r:
                cm.append(b.count_ms), em.append(b.emit_ms)
        print(f"stage_bytes={stage:6d} count_threads={th:4d} staged_levels={idx.stats().staged_levels} "
              f"count+scan={np.median(cm):7.3f} ms emit={np.median(em):7.3f} ms")
