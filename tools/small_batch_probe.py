"""Small-batch latency probe (development aid): the single-call device path at batch sizes from 1 k to 10 M queries on the
cfg2 index, three-kernel form (count_group -> block scan -> emit_scan) against the one-kernel look-back form."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from genogrove_b200 import DeviceIndex, synth

w = synth.cfg2(sorted_queries=os.environ.get("SORTED", "1") == "1")
dev = torch.device("cuda:0")
idx = DeviceIndex(0)
idx.build(w.chrom, w.start, w.end, w.n_chrom, want_perm=False)
st = torch.cuda.current_stream().cuda_stream
for nq in (1_000, 10_000, 100_000, 1_000_000, 4_000_000, 10_000_000):
    d_c = torch.from_numpy(w.q_chrom[:nq].view(np.int32)).to(dev)
    d_s = torch.from_numpy(w.q_start[:nq].astype(np.uint32).view(np.int32)).to(dev)
    d_e = torch.from_numpy(w.q_end[:nq].astype(np.uint32).view(np.int32)).to(dev)
    d_off = torch.empty(nq + 1, dtype=torch.int64, device=dev)
    d_hits = torch.empty(4 * nq + 1024, dtype=torch.int32, device=dev)
    row = []
    for mode in (0, 1):
        idx.set_option("fused_kernel", mode)
        reps = 200 if nq <= 1_000_000 else 30
        for _ in range(5):
            idx.fused_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, d_off.data_ptr(), d_hits.data_ptr(), d_hits.numel(), st, sync_total=True)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            idx.fused_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, d_off.data_ptr(), d_hits.data_ptr(), d_hits.numel(), st, sync_total=True)
        b.record(); torch.cuda.synchronize()
        row.append(a.elapsed_time(b) / reps * 1e3)
    print(f"nq={nq:9d}  three-kernel {row[0]:9.1f} us   one-kernel {row[1]:9.1f} us", flush=True)
