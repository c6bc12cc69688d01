#!/usr/bin/env python
"""A/B aid (development, not product): kernel times of the single-call batch path (count_group -> block scan ->
emit_scan) on cfg4 for the library named by GG_B200_LIB (default: the in-tree build) and a few option settings."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from genogrove_b200 import DeviceIndex, synth, _lib  # noqa: E402

nq = int(os.environ.get("AB_QUERIES", 100_000_000))
dev = torch.device("cuda:0")
c, s, e = synth.hg38_intervals(10_000_000, 42)
idx = DeviceIndex(0)
idx.build(c, s, e, 24, want_perm=False)
st = torch.cuda.current_stream().cuda_stream
print("lib:", _lib.LIB_PATH)
for shuffled in (False, True):
    qc, qs, qe = synth.hg38_queries(nq, 150, 43, not shuffled)
    d_c = torch.from_numpy(qc.view(np.int32)).to(dev)
    d_s = torch.from_numpy(qs.astype(np.uint32).view(np.int32)).to(dev)
    d_e = torch.from_numpy(qe.astype(np.uint32).view(np.int32)).to(dev)
    d_off = torch.empty(nq + 1, dtype=torch.int64, device=dev)
    total = idx.count_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, d_off.data_ptr(), st, sync_total=True)
    d_hits = torch.empty(max(total, 1), dtype=torch.int32, device=dev)
    idx.set_option("timing", 1)
    for opts in ({}, {"count_ctas_per_sm": 4}, {"count_ctas_per_sm": 6}, {"emit_walk": 16}, {"emit_ctas_per_sm": 6}):
        for k in ("count_ctas_per_sm", "emit_ctas_per_sm"):
            idx.set_option(k, 0)
        idx.set_option("emit_walk", 8)
        for k, v in opts.items():
            idx.set_option(k, v)
        cm, sm, em = [], [], []
        for r in range(8):
            idx.fused_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, d_off.data_ptr(), d_hits.data_ptr(),
                             d_hits.numel(), st, sync_total=True)
            b = idx.batch_stats()
            if r >= 2:
                cm.append(b.count_ms), sm.append(b.scan_ms), em.append(b.emit_ms)
        print(f"shuffled={int(shuffled)} {str(opts):32s} count={np.median(cm):6.3f} scan={np.median(sm):6.3f} "
              f"emit={np.median(em):6.3f} sum={np.median(cm) + np.median(sm) + np.median(em):6.3f} ms")
    idx.set_option("timing", 0)
    del d_hits, d_c, d_s, d_e, d_off
    torch.cuda.empty_cache()
