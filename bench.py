#!/usr/bin/env python
"""bench.py — overlap queries/s of the batched grove intersect path on N B200s (one process per GPU).

Workload (BASELINE.json configs[3], the configuration the metric is quoted on): a 10M-interval hg38-shaped
grove (24 chromosomes) and, PER GPU, 100M read-shaped (150 bp) queries, coordinate-sorted per chromosome
(BAM-like); index replicated on every GPU, query batch sharded (weak scaling: per-GPU work is fixed).
A step = one pass of the hot path over the rank's 100M-query batch through the single-call device entry point
(gg_intersect_fused_device): count (4 consecutive queries per thread) -> scan of the per-block totals -> emit_scan
(offsets + hits), genogrove_b200/csrc/{intersect,emit_scan}.cuh.  --one-kernel times the one-kernel look-back form
(fused.cuh), --two-phase the count / scan / emit kernels of the two-phase API.

  value : whole-job queries/s with queries and results resident in HBM (device-buffer C ABI).
  e2e   : the same metric through the host-buffer C ABI call (gg_intersect_batch, u64 coordinates as the
          reference's interval uses), pinned host buffers, H2D + D2H inside the timed region.
  --impl reference : the reference's own CPU intersect loop (oracle/_ref = unmodified genogrove) on this
          box's host cores, all threads, on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "overlap_queries_per_sec"
UNIT = "queries/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--intervals", type=int, default=10_000_000)
    ap.add_argument("--queries", type=int, default=100_000_000, help="queries per GPU")
    ap.add_argument("--shuffled", action="store_true", help="shuffled instead of coordinate-sorted queries")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--cpu-sample", type=int, default=32_000_000, help="queries in the CPU baseline sample")
    ap.add_argument("--cpu-order", type=int, default=32, help="B+ tree order of the reference grove (3 = library default)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the flanking / stranded side measurements")
    ap.add_argument("--two-phase", action="store_true",
                    help="time the count -> scan -> emit kernels (two-phase API) instead of the single-call path")
    ap.add_argument("--one-kernel", action="store_true", help="single-call path as ONE kernel (decoupled look-back scan)")
    ap.add_argument("--regs40", action="store_true", help="40-register build of the one-kernel form")
    ap.add_argument("--ctas-per-sm", type=int, default=0)
    ap.add_argument("--emit-tma", type=int, default=-1, help="pass 2 kernel: 1 TMA ring + bulk store [default], 2 TMA ring + vector stores, 0 emit_scan")
    ap.add_argument("--emit-ctas-per-sm", type=int, default=0)
    return ap.parse_args()


def workload_config(args, world):
    return {
        "workload": "cfg4: 10M-interval hg38-shaped grove (24 chromosomes) x 100M 150bp read-shaped queries per GPU, "
                    + ("shuffled" if args.shuffled else "coordinate-sorted per chromosome (BAM-like)"),
        "intervals": args.intervals,
        "queries_per_gpu": args.queries,
        "query_order": "shuffled" if args.shuffled else "sorted",
        "parallelism": f"index replicated, query batch sharded x{world}",
        "l2_policy": "inputs+outputs per step (>4 GB) exceed the 126 MB L2; no flush needed",
    }


# ----------------------------------------------------------------------------------------- clocks sampler
class ClockSampler:
    """Polls NVML (SM clock + clock event reasons) every few ms from a thread while the timed region runs."""
    REASONS = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}

    def __init__(self, gpu_index: int, period_s: float = 0.004):
        self.samples = []
        self.period = period_s
        self.ok = False
        self._stop = threading.Event()
        try:
            import pynvml as N
            N.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = gpu_index
            if vis:
                ids = [v.strip() for v in vis.split(",") if v.strip()]
                if gpu_index < len(ids) and ids[gpu_index].isdigit():
                    phys = int(ids[gpu_index])
            self.N, self.h = N, N.nvmlDeviceGetHandleByIndex(phys)
            self.max_sm = float(N.nvmlDeviceGetMaxClockInfo(self.h, N.NVML_CLOCK_SM))
            self.ok = True
        except Exception as ex:  # noqa: BLE001
            self.err = repr(ex)

    def _poll(self):
        N = self.N
        while not self._stop.is_set():
            try:
                clk = float(N.nvmlDeviceGetClockInfo(self.h, N.NVML_CLOCK_SM))
                try:
                    rs = int(N.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                except Exception:  # noqa: BLE001
                    rs = int(N.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                self.samples.append((time.time(), clk, rs))
            except Exception:  # noqa: BLE001
                pass
            time.sleep(self.period)

    def start(self):
        if self.ok:
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()

    def stop(self, t0, t1):
        if not self.ok:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable: " + getattr(self, "err", "?")]}
        self._stop.set()
        self.t.join(timeout=2)
        inside = [(c, r) for ts, c, r in self.samples if t0 <= ts <= t1] or [(c, r) for _, c, r in self.samples]
        reasons = sorted(nm for nm, bit in self.REASONS.items() if any(r & bit for _, r in inside))
        return {"sm_mhz": float(np.median([c for c, _ in inside])) if inside else None, "sm_max_mhz": self.max_sm,
                "reasons": reasons, "samples": len(inside)}


# ----------------------------------------------------------------------------------------- reference arm
def cpu_reference_run(args, w, steps, warmup, threads=None):
    """Times the unmodified reference (oracle/_ref) on a bounded sample; returns (q/s, info dict)."""
    import oracle as O
    if not O.ref_available():
        return None, {"unavailable": "oracle/_ref/libgg_ref.so not built"}
    g = O.RefGrove(args.cpu_order)
    T = threads or g.hardware_threads()
    t0 = time.time()
    ids = np.arange(len(w.chrom), dtype=np.uint32)
    for c in range(w.n_chrom):
        m = w.chrom == c
        g.insert_bulk(c, w.start[m], w.end[m], ids[m], presorted=False)
    build_s = time.time() - t0
    nq = len(w.q_chrom)
    S = min(args.cpu_sample, nq)
    lo = (nq - S) // 2
    qc, qs, qe = w.q_chrom[lo:lo + S], w.q_start[lo:lo + S], w.q_end[lo:lo + S]
    secs, hits = [], 0
    for i in range(warmup + steps):
        off, _, s = g.intersect_batch(qc, qs, qe, threads=T, want_ids=False)
        hits = int(off[-1])
        if i >= warmup:
            secs.append(s)
    ms = 1e3 * float(np.mean(secs))
    info = {"value": S / (ms / 1e3), "unit": UNIT, "cores": T, "kind": "reference",
            "sample": f"{S} contiguous queries from the middle of the batch x {steps} steps, reference grove order "
                      f"{args.cpu_order}, bulk build {build_s:.1f}s, {hits / S:.2f} hits/query",
            "ms_per_step": ms, "hit_pairs_per_sec": hits / (ms / 1e3)}
    return info["value"], info


def run_reference_arm(args, rank, world):
    if rank != 0:
        return
    from genogrove_b200 import synth
    w = synth.cfg4(args.intervals, min(args.queries, max(args.cpu_sample * 2, 1_000_000)), sorted_queries=not args.shuffled)
    steps, warmup = max(1, min(args.steps, 5)), max(0, min(args.warmup, 1))
    v, info = cpu_reference_run(args, w, steps, warmup)
    if v is None:
        print(json.dumps({"impl": "reference", "unavailable": info["unavailable"]}))
        return
    out = {"metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps, "warmup": warmup,
           "ms_per_step": info["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "u64", "data": "synthetic", "impl": "reference", "config": workload_config(args, world),
           "cpu_baseline": {k: info[k] for k in ("value", "unit", "cores", "kind", "sample")},
           "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "hit_pairs_per_sec": info["hit_pairs_per_sec"], "gpu_launches": 0}
    print(json.dumps(out))


# ----------------------------------------------------------------------------------------- B200 arm
def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    from genogrove_b200 import DeviceIndex, PinnedArray, synth, shard

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    # ---- synthetic inputs: same grove on every rank, a different 100M-query shard per rank
    c, s, e = synth.hg38_intervals(args.intervals, 42)
    qc, qs, qe = synth.hg38_queries(args.queries, 150, 43 + rank, not args.shuffled)
    nq = len(qc)

    idx = DeviceIndex(local_rank)
    idx.build(c, s, e, 24, want_perm=False)
    idx.build(c, s, e, 24, want_perm=False)  # second build: the reported build times exclude first-call allocation / context cost
    ist = idx.stats()

    d_c = torch.from_numpy(qc.view(np.int32)).to(dev)
    d_s = torch.from_numpy(qs.astype(np.uint32).view(np.int32)).to(dev)
    d_e = torch.from_numpy(qe.astype(np.uint32).view(np.int32)).to(dev)
    d_off = torch.empty(nq + 1, dtype=torch.int64, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    # sizing pass (untimed): a real caller sizes the hit buffer from the previous batch
    total = idx.count_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, d_off.data_ptr(), stream, sync_total=True)
    d_hits = torch.empty(max(total, 1), dtype=torch.int32, device=dev)
    d_tot = torch.zeros(1, dtype=torch.int64, device=dev)
    d_all = torch.zeros(world, dtype=torch.int64, device=dev)
    fused = not args.two_phase
    one_kernel = fused and args.one_kernel
    idx.set_option("fused_kernel", 1 if one_kernel else 0)
    if args.regs40:
        idx.set_option("fused_regs40", 1)
    if args.ctas_per_sm:
        idx.set_option("fused_ctas_per_sm", args.ctas_per_sm)
    if args.emit_tma >= 0:
        idx.set_option("emit_tma", args.emit_tma)
    if args.emit_ctas_per_sm:
        idx.set_option("emit_ctas_per_sm", args.emit_ctas_per_sm)
    idx.set_option("timing", 1)  # library CUDA events around its kernels, on the launching stream
    kms = {"count": [], "scan": [], "emit": [], "fused": []}

    def step(record=False):
        if fused:
            # ONE kernel: count -> look-back scan -> emit; the host reads the hit total (8 B) after it
            t = idx.fused_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, d_off.data_ptr(),
                                 d_hits.data_ptr(), d_hits.numel(), stream, sync_total=True)
        else:
            # pass 1 + scan; the host needs the hit total to size the output (a real caller allocates here)
            t = idx.count_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, d_off.data_ptr(), stream, sync_total=True)
            assert t <= d_hits.numel()
            idx.emit_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, d_off.data_ptr(), d_hits.data_ptr(), stream)
        if world > 1:  # the path's only collective: gather hit totals -> global CSR base
            d_tot.copy_(d_off[nq:nq + 1])
            dist.all_gather_into_tensor(d_all, d_tot)
        if record:
            torch.cuda.current_stream().synchronize()
            bs = idx.batch_stats()
            if one_kernel:
                kms["fused"].append(bs.emit_ms)
            else:
                kms["count"].append(bs.count_ms), kms["scan"].append(bs.scan_ms), kms["emit"].append(bs.emit_ms)
        return t

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    time.sleep(0.05)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_wall0 = time.time()
    e0.record()
    for k in range(args.steps):
        step(record=True)
    e1.record()
    barrier()
    t_wall1 = time.time()
    clocks = sampler.stop(t_wall0, t_wall1)
    ms_total = e0.elapsed_time(e1)
    t_ms = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
    ms_per_step = float(t_ms.item()) / args.steps
    idx.set_option("timing", 0)
    kernel_ms = {k: float(np.mean(v)) for k, v in kms.items() if v}

    totals = torch.tensor([total, nq], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(totals, op=dist.ReduceOp.SUM)
    job_hits, job_q = int(totals[0].item()), int(totals[1].item())
    value = job_q / (ms_per_step / 1e3)

    # ---- roofline of the dominant kernel (algorithmic bytes)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_kind = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    bytes_path = 12 * nq + 8 * (nq + 1) + 4 * total  # B_q = read (chrom,start,end) u32, write offsets u64, ranks u32
    if one_kernel:
        dom, dom_ms, dom_bytes = "intersect_fused_kernel", kernel_ms["fused"], bytes_path
        launches_per_step = 1
    elif fused:
        bytes_count = 12 * nq + 8 * nq                   # read (chrom,start,end) u32, write count u32 + hi u32
        bytes_emit = 4 * total + 8 * (nq + 1) + 16 * nq  # write ranks + offsets, read count, hi, chrom, start
        if kernel_ms["emit"] >= kernel_ms["count"]:
            dom, dom_ms, dom_bytes = "intersect_emit_scan_kernel", kernel_ms["emit"], bytes_emit
        else:
            dom, dom_ms, dom_bytes = "intersect_count_group_kernel", kernel_ms["count"], bytes_count
        launches_per_step = 3  # count_group, block scan, emit_scan (the dense-tile kernel only runs for batches averaging >= 12 hits/query)
    else:
        bytes_count = 12 * nq + 8 * nq                   # read (chrom,start,end) u32, write count u32 + hi u32
        bytes_emit = 4 * total + 8 * (nq + 1) + 12 * nq  # write ranks, read offsets, chrom/start/hi
        if kernel_ms["emit"] >= kernel_ms["count"]:
            dom, dom_ms, dom_bytes = "intersect_emit_kernel", kernel_ms["emit"], bytes_emit
        else:
            dom, dom_ms, dom_bytes = "intersect_count_dir_kernel", kernel_ms["count"], bytes_count
        launches_per_step = 4
    traffic = None
    try:  # per-launch DRAM bytes of that kernel from the committed ncu --set full capture of this same command
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        key = dom + ("" if not args.shuffled else ":shuffled")
        if key in tj and tj[key].get("queries") == nq:
            traffic = tj[key]["dram_bytes_per_launch"]
    except Exception:
        pass
    achieved = dom_bytes / (dom_ms / 1e3) / 1e9
    roofline = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_source": peak_kind,
                "algorithmic_bytes_per_launch": dom_bytes, "kernel_ms": kernel_ms,
                "path_achieved_gbs": bytes_path / (ms_per_step / 1e3) / 1e9,
                "path_frac": bytes_path / (ms_per_step / 1e3) / 1e9 / peak}

    # ---- side measurements (not the headline): the two other query kinds on the same index and batch, device-resident
    extras = None
    if not args.no_extras:
        ex_steps = 10
        d_pred = torch.empty(nq, dtype=torch.int32, device=dev)
        d_succ = torch.empty(nq, dtype=torch.int32, device=dev)

        def timed(fn, n):
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(n):
                fn()
            b.record()
            torch.cuda.synchronize()
            return a.elapsed_time(b) / n

        fl_ms = timed(lambda: idx.flanking_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, d_pred.data_ptr(),
                                                  d_succ.data_ptr(), stream, sync=False), ex_steps)
        del d_pred, d_succ
        rs = np.random.default_rng(7)
        idx.set_strands(np.frombuffer(b"+-.", dtype=np.uint8)[rs.integers(0, 3, ist.n_intervals)])
        d_str = torch.from_numpy(np.frombuffer(b"+-", dtype=np.uint8)[rs.integers(0, 2, nq)]).to(dev)
        st_total = [0]

        def stranded_step():
            st_total[0] = idx.stranded_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), d_str.data_ptr(), 4, nq,
                                              d_off.data_ptr(), d_hits.data_ptr(), d_hits.numel(), stream, sync_total=True)

        st_ms = timed(stranded_step, ex_steps)
        del d_str
        fl_bytes = 20 * nq  # read (chrom,start,end) u32, write pred + succ u32
        st_bytes = 13 * nq + 8 * (nq + 1) + 4 * st_total[0]
        extras = {"flanking": {"queries_per_sec": nq * world / (fl_ms / 1e3), "ms_per_step": fl_ms, "kernel": "flanking_kernel",
                               "algorithmic_bytes": fl_bytes, "achieved_gbs": fl_bytes / (fl_ms / 1e3) / 1e9,
                               "frac": fl_bytes / (fl_ms / 1e3) / 1e9 / peak},
                  "stranded": {"queries_per_sec": nq * world / (st_ms / 1e3), "ms_per_step": st_ms,
                               "kernels": "strand_count_group_kernel + exclusive_scan_kernel (block totals) + intersect_emit_scan_kernel<kStrand>",
                               "hits_per_query": st_total[0] / nq, "algorithmic_bytes": st_bytes,
                               "achieved_gbs": st_bytes / (st_ms / 1e3) / 1e9, "frac": st_bytes / (st_ms / 1e3) / 1e9 / peak,
                               "strands": "index: uniform over + - . ; queries: uniform over + -"},
                  "note": "per-rank device time, same index and query batch as the headline; steps=%d" % ex_steps}

    # ---- e2e: host-buffer C ABI call (u64 coordinates, pinned buffers), H2D + D2H inside the timed region
    e2e = None
    if not args.no_e2e:
        p_c, p_s, p_e = PinnedArray(nq, np.uint32), PinnedArray(nq, np.uint64), PinnedArray(nq, np.uint64)
        p_off = PinnedArray(nq + 1, np.uint64)
        p_c.array[:], p_s.array[:], p_e.array[:] = qc, qs, qe
        del d_hits
        torch.cuda.empty_cache()

        def e2e_step():
            off, hits = idx.intersect_batch(p_c.array, p_s.array, p_e.array, offsets_out=p_off.array, copy_hits=False)
            return int(off[-1])

        e2e_step()  # warm-up: grows the pinned result buffer and the pipeline slots
        e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.e2e_steps):
            h = e2e_step()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / args.e2e_steps
        assert h == total
        t_e = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t_e, op=dist.ReduceOp.MAX)
        e2e = {"value": job_q / float(t_e.item()), "unit": UNIT, "h2d_bytes_per_step": int(nq * 20),
               "d2h_bytes_per_step": int((nq + 1) * 8 + total * 4), "ms_per_step": float(t_e.item()) * 1e3,
               "steps": args.e2e_steps, "api": "gg_intersect_batch (u64 coordinates, pinned host buffers)",
               "timer": "host wall clock around the blocking C ABI call, max over ranks"}

    # ---- CPU baseline beside it (rank 0, N=1 only): the unmodified reference on a bounded sample
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        w = synth.Workload("cfg4", 24, c, s, e, qc, qs, qe)
        v, info = cpu_reference_run(args, w, steps=2, warmup=1)
        cpu = {k: info[k] for k in ("value", "unit", "cores", "kind", "sample")} if v else info
        v1, info1 = cpu_reference_run(args, w, steps=1, warmup=0, threads=1) if v else (None, None)
        if v1:
            cpu["single_thread_value"] = v1

    if rank == 0:
        out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
               "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
               "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
               "config": workload_config(args, world), "hit_pairs_per_sec": job_hits / (ms_per_step / 1e3),
               "hits_per_query": job_hits / job_q, "gpu_launches": launches_per_step * args.steps, "clocks": clocks,
               "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "extras": extras,
               "index": {"levels": ist.n_levels, "fanout": ist.fanout, "staged_levels": ist.staged_levels,
                         "coord_bytes": ist.coord_bytes, "device_mb": ist.device_bytes / 1e6,
                         "build_sort_ms": ist.sort_ms, "build_levels_ms": ist.levels_ms, "build_total_ms": ist.total_ms}}
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
