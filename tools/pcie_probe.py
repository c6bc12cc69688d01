#!/usr/bin/env python
"""Host <-> device copy ceiling of the box: N processes (one per GPU), pinned buffers, bare cudaMemcpyAsync.

    python tools/pcie_probe.py                      # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/pcie_probe.py

For each pattern — H2D alone, D2H alone, both at once (two streams) — every rank moves the byte counts of one e2e step of
bench.py (defaults: the round-1 call, 2.0 GB up / 3.78 GB down; --compact: 1.2 GB up / 1.55 GB down) after a barrier; the
time is the max over ranks.  Prints one JSON line: per-GPU and aggregate GB/s per pattern.  This is the denominator the
e2e number of bench.py is read against ("nothing in the repo measures that box ceiling")."""
import argparse
import json
import os
import time

import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--h2d-mb", type=float, default=2000.0)
    ap.add_argument("--d2h-mb", type=float, default=3780.0)
    ap.add_argument("--compact", action="store_true")
    ap.add_argument("--reps", type=int, default=5)
    a = ap.parse_args()
    if a.compact:
        a.h2d_mb, a.d2h_mb = 1200.0, 1550.0
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    up, dn = int(a.h2d_mb * 1e6), int(a.d2h_mb * 1e6)
    up_h, dn_h = torch.empty(up, dtype=torch.uint8).pin_memory(), torch.empty(dn, dtype=torch.uint8).pin_memory()
    up_d, dn_d = torch.empty(up, dtype=torch.uint8, device=dev), torch.empty(dn, dtype=torch.uint8, device=dev)
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def run(do_up, do_dn):
        best = None
        for _ in range(a.reps):
            barrier()
            t0 = time.perf_counter()
            if do_up:
                with torch.cuda.stream(s1):
                    up_d.copy_(up_h, non_blocking=True)
            if do_dn:
                with torch.cuda.stream(s2):
                    dn_h.copy_(dn_d, non_blocking=True)
            s1.synchronize()
            s2.synchronize()
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        t = torch.tensor([best], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    t_up, t_dn, t_both = run(True, False), run(False, True), run(True, True)
    if rank == 0:
        print(json.dumps({"n_gpus": world, "h2d_bytes": up, "d2h_bytes": dn,
                          "h2d_alone_gbs_per_gpu": up / t_up / 1e9, "d2h_alone_gbs_per_gpu": dn / t_dn / 1e9,
                          "both_ms": t_both * 1e3, "both_h2d_gbs_per_gpu": up / t_both / 1e9, "both_d2h_gbs_per_gpu": dn / t_both / 1e9,
                          "both_aggregate_gbs": (up + dn) * world / t_both / 1e9,
                          "queries_per_sec_ceiling_100M_per_gpu": 1e8 * world / t_both}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
