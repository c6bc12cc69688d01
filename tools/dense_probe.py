"""Dense-index probe (development aid): a cfg5-density workload (25.7M intervals on 3 chromosomes, ~150 hits per
query) at a reduced query count, through the single-call device path; for ncu captures of intersect_emit_dense_kernel."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from genogrove_b200 import DeviceIndex, synth

nq = int(sys.argv[1]) if len(sys.argv) > 1 else 16_000_000
rng = np.random.default_rng(5)
L = synth.HG38_LENGTHS[:3]
chrom, start, end, qc, qs, qe = [], [], [], [], [], []
for c in range(3):
    n = int(os.environ.get("N_PER", 8_500_000))  # 2_000_000: a uniform ~20 hits/query batch
    ln = np.exp(rng.uniform(np.log(100), np.log(10000), n)).astype(np.int64)
    st = (rng.random(n) * (int(L[c]) - ln - 1)).astype(np.int64)
    chrom.append(np.full(n, c, np.uint32)); start.append(st.astype(np.uint64)); end.append((st + ln - 1).astype(np.uint64))
    m = nq // 3
    q0 = np.sort(rng.integers(0, int(L[c]) - 150, m, dtype=np.int64)).astype(np.uint64)
    qc.append(np.full(m, c, np.uint32)); qs.append(q0); qe.append(q0 + np.uint64(149))
chrom, start, end = np.concatenate(chrom), np.concatenate(start), np.concatenate(end)
qc, qs, qe = np.concatenate(qc), np.concatenate(qs), np.concatenate(qe)
nq = len(qc)
dev = torch.device("cuda:0")
idx = DeviceIndex(0)
idx.set_option("dense_emit", int(os.environ.get("DENSE", "1")))
for kv in os.environ.get("GG_OPTS", "").split(","):  # e.g. GG_OPTS=delta_min_hits=0
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
for _ in range(3):
    idx.fused_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, d_off.data_ptr(), d_hits.data_ptr(), total, st_, sync_total=True)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(5):
    idx.fused_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, d_off.data_ptr(), d_hits.data_ptr(), total, st_, sync_total=True)
b.record(); torch.cuda.synchronize()
ms = a.elapsed_time(b) / 5
print(f"dense_emit={os.environ.get('DENSE', '1')} nq={nq} hits={total} ({total / nq:.1f}/query) {ms:.3f} ms  {total / ms / 1e6:.1f} G hit-pairs/s  hit bytes {4 * total / 1e9:.2f} GB")
