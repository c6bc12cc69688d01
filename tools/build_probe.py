"""Times gg_index_build twice on the cfg4 index (second call = warm allocations) and prints the library's stats."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from genogrove_b200 import DeviceIndex, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
c, s, e = synth.hg38_intervals(n, 42)
idx = DeviceIndex(0)
for kv in os.environ.get("GG_OPTS", "").split(","):  # e.g. GG_OPTS=sort_legacy=1
    if kv:
        idx.set_option(kv.split("=")[0], int(kv.split("=")[1]))
for k in range(3):
    t0 = time.perf_counter()
    idx.build(c, s, e, 24, want_perm=False)
    dt = (time.perf_counter() - t0) * 1e3
    st = idx.stats()
    print(f"build {k}: wall {dt:.1f} ms, total {st.total_ms:.1f} ms, sort {st.sort_ms:.2f} ms, levels {st.levels_ms:.2f} ms, device {st.device_bytes/1e6:.1f} MB")
