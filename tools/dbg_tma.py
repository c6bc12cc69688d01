"""Debug: compare the TMA emit (emit_tma 1/2) with emit_scan (emit_tma 0) on the device path at sizes where a warp owns several blocks."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from genogrove_b200 import DeviceIndex, synth

dev = torch.device("cuda:0")
c, s, e = synth.hg38_intervals(1_000_000, 42)
idx = DeviceIndex(0)
idx.build(c, s, e, 24, want_perm=False)
idx.set_option("fused_kernel", 0)
for nq, sorted_q, nb in ((300_000, True, 4), (2_000_000, True, 4), (2_000_000, False, 4), (2_000_000, False, 8), (10_000_000, False, 8)):
    qc, qs, qe = synth.hg38_queries(nq, 1000, 7, sorted_q)
    dt, tt = (np.uint32, np.int32) if nb == 4 else (np.uint64, np.int64)
    d_c = torch.from_numpy(qc.view(np.int32)).to(dev)
    d_s = torch.from_numpy(qs.astype(dt).view(tt)).to(dev)
    d_e = torch.from_numpy(qe.astype(dt).view(tt)).to(dev)
    st = torch.cuda.current_stream().cuda_stream
    res = {}
    for mode, cps in ((0, 0), (1, 0), (2, 0), (1, 1), (2, 1)):
        idx.set_option("emit_tma", mode)
        idx.set_option("emit_ctas_per_sm", cps)
        d_off = torch.zeros(nq + 1, dtype=torch.int64, device=dev)
        tot = idx.count_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), nb, nq, d_off.data_ptr(), st, sync_total=True)
        d_hits = torch.full((tot + 8,), -1, dtype=torch.int32, device=dev)
        d_off.zero_()
        got = idx.fused_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), nb, nq, d_off.data_ptr(), d_hits.data_ptr(), tot, st, sync_total=True)
        res[(mode, cps)] = (d_off.cpu().numpy(), d_hits.cpu().numpy(), got, tot)
    ref = res[(0, 0)]
    for k, v in res.items():
        okoff = np.array_equal(v[0], ref[0]); okh = np.array_equal(v[1], ref[1])
        msg = f"nq={nq} sorted={sorted_q} nb={nb} mode={k}: total {v[2]}/{v[3]} offsets_equal={okoff} hits_equal={okh}"
        if not okoff:
            bad = np.nonzero(v[0] != ref[0])[0]
            msg += f" first bad q={bad[0]} n_bad={len(bad)} got={v[0][bad[0]]} exp={ref[0][bad[0]]} tile={bad[0]//32} blk={bad[0]//128}"
        elif not okh:
            bad = np.nonzero(v[1] != ref[1])[0]
            q = np.searchsorted(ref[0], bad[0], side='right') - 1
            msg += f" first bad hit={bad[0]} n_bad={len(bad)} got={v[1][bad[0]]} exp={ref[1][bad[0]]} query={q} tile={q//32}"
        print(msg, flush=True)
