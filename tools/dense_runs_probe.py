"""What one owner GPU sees in the chromosome-sharded cfg5 job (development aid): its share of the index (25.7 M intervals on 3
chromosomes) and a received batch made of R coordinate-sorted runs (one per source rank), each spanning its chromosomes.
usage: dense_runs_probe.py total_queries runs"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from genogrove_b200 import DeviceIndex, synth

nq, runs = int(sys.argv[1]), int(sys.argv[2])
rng = np.random.default_rng(5)
L = synth.HG38_LENGTHS[:3]
chrom, start, end = [], [], []
for c in range(3):
    n = 8_584_000
    ln = np.exp(rng.uniform(np.log(100), np.log(10000), n)).astype(np.int64)
    st = (rng.random(n) * (int(L[c]) - ln - 1)).astype(np.int64)
    chrom.append(np.full(n, c, np.uint32)); start.append(st.astype(np.uint64)); end.append((st + ln - 1).astype(np.uint64))
chrom, start, end = np.concatenate(chrom), np.concatenate(start), np.concatenate(end)
qc, qs = [], []
for r in range(runs):
    for c in range(3):
        m = int(nq / runs * L[c] / L.sum())
        qc.append(np.full(m, c, np.uint32)); qs.append(np.sort(rng.integers(0, int(L[c]) - 150, m, dtype=np.int64)).astype(np.uint64))
qc, qs = np.concatenate(qc), np.concatenate(qs)
qe = qs + np.uint64(149)
nq = len(qc)
dev = torch.device("cuda:0")
idx = DeviceIndex(0)
idx.set_option("dense_emit", 1)
for kv in os.environ.get("GG_OPTS", "").split(","):
    if kv:
        idx.set_option(kv.split("=")[0], int(kv.split("=")[1]))
idx.build(chrom, start, end, 3, want_perm=False)
d_c = torch.from_numpy(qc.view(np.int32)).to(dev)
d_s = torch.from_numpy(qs.astype(np.uint32).view(np.int32)).to(dev)
d_e = torch.from_numpy(qe.astype(np.uint32).view(np.int32)).to(dev)
d_off = torch.empty(nq + 1, dtype=torch.int64, device=dev)
st_ = torch.cuda.current_stream().cuda_stream
total = idx.count_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, d_off.data_ptr(), st_, sync_total=True)
d_hits = torch.empty(total, dtype=torch.int32, device=dev)
idx.set_option("timing", 1)
for _ in range(2):
    idx.fused_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, d_off.data_ptr(), d_hits.data_ptr(), total, st_, sync_total=True)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(3):
    idx.fused_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, d_off.data_ptr(), d_hits.data_ptr(), total, st_, sync_total=True)
b.record(); torch.cuda.synchronize()
ms = a.elapsed_time(b) / 3
bs = idx.batch_stats()
print(f"runs={runs} nq={nq} hits/query={total / nq:.1f} {ms:.2f} ms = {ms / (nq / 1e6):.3f} ms per M queries, {total / ms / 1e6:.0f} G hit-pairs/s; count {bs.count_ms:.2f} scan {bs.scan_ms:.2f} emit {bs.emit_ms:.2f} ms")
