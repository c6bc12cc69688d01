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
  e2e   : the same metric through the host-buffer C ABI call with pinned host buffers, H2D + D2H inside the timed
          region: gg_intersect_batch_compact (u32 coordinates up, 8 B per query + a 1-byte back-delta per hit down — every
          hit is delivered; e2e.full_csr is gg_intersect_batch_u32 with u64 offsets + u32 ranks, e2e.u64_api the round-1
          call with u64 coordinates); e2e.copy_ceiling = the same bytes moved by bare concurrent cudaMemcpyAsync.
  strong / chrom_shard (N > 1): the same 100M-query batch cut into N slices (BASELINE cfg4 as written), and the grove
          sharded by chromosome with queries routed to the owner over NVLink (gg_group_*, rank form, NCCL).
  --mode cfg5 : BASELINE configs[4] — 200M-interval chromosome-sharded build + 1B routed queries over N GPUs.
  --impl reference : the reference's own CPU intersect loop (oracle/_ref = unmodified genogrove) on this
          box's host cores, all threads, on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import ctypes
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
    ap.add_argument("--mode", default="cfg4", choices=["cfg4", "cfg5"],
                    help="cfg5: chromosome-sharded 200M-interval build + 1B-query batch over --gpus ranks (BASELINE configs[4])")
    ap.add_argument("--cfg5-intervals", type=int, default=200_000_000)
    ap.add_argument("--cfg5-queries", type=int, default=1_000_000_000)
    ap.add_argument("--cfg5-chunk", type=int, default=64_000_000, help="queries per rank and routed call")
    ap.add_argument("--cfg5-layout", default="genome_wide_chunks", choices=["contiguous", "genome_wide_chunks"],
                    help="how the 1B-query coordinate-sorted batch reaches the ranks: contiguous = rank r reads the r-th contiguous "
                         "part of the sorted batch and submits it in consecutive chunks (a reader of a sorted BAM) [default]; "
                         "genome_wide_chunks = every chunk is its own sorted batch spanning the whole genome (round-2 early runs)")
    ap.add_argument("--no-multi", action="store_true", help="skip the strong-scaling / chromosome-sharded side measurements (N > 1)")
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
        "scaling_note": "headline: weak (100M queries per GPU); 'strong' holds the fixed 100M-query batch cut into N slices",
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


def timed_ms(torch, fn, n, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


def other_configs(torch, dev, peak, steps=10):
    """BASELINE.json's other single-GPU configurations, device-resident, at full size (N = 1 only): configs[1] (1M intervals x
    10M 1-kb queries, random and coordinate-sorted) and configs[2] (GENCODE-shaped nested features x 50M queries).  Same
    single call and same timing rules as the headline (CUDA events, 3 warm-up calls, inputs and outputs far larger than L2);
    a total-hits check against the two-phase count pass guards each number."""
    from genogrove_b200 import DeviceIndex, synth
    out = {}
    for name, make in (("cfg2_random", lambda: synth.cfg2(sorted_queries=False)), ("cfg2_sorted", lambda: synth.cfg2(sorted_queries=True)),
                       ("cfg3_gencode", lambda: synth.cfg3())):
        w = make()
        nq = len(w.q_chrom)
        idx = DeviceIndex(dev.index)
        idx.build(w.chrom, w.start, w.end, w.n_chrom, want_perm=False)
        d_c = torch.from_numpy(w.q_chrom.view(np.int32)).to(dev)
        d_s = torch.from_numpy(w.q_start.astype(np.uint32).view(np.int32)).to(dev)
        d_e = torch.from_numpy(w.q_end.astype(np.uint32).view(np.int32)).to(dev)
        d_off = torch.empty(nq + 1, dtype=torch.int64, device=dev)
        st = torch.cuda.current_stream().cuda_stream
        total = idx.count_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, d_off.data_ptr(), st, sync_total=True)
        d_hits = torch.empty(max(total, 1), dtype=torch.int32, device=dev)

        def step():
            t = idx.fused_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, d_off.data_ptr(), d_hits.data_ptr(),
                                 d_hits.numel(), st, sync_total=True)
            assert t == total, (name, t, total)

        ms = timed_ms(torch, step, steps)
        b = 12 * nq + 8 * (nq + 1) + 4 * total
        out[name] = {"workload": w.name, "intervals": int(len(w.chrom)), "queries": nq, "hits": int(total),
                     "hits_per_query": total / nq, "ms_per_step": ms, "queries_per_sec": nq / (ms / 1e3),
                     "hit_pairs_per_sec": total / (ms / 1e3), "algorithmic_bytes": b, "path_frac": b / (ms / 1e3) / 1e9 / peak,
                     "kernel_launches": int(idx.batch_stats().kernel_launches)}
        del d_c, d_s, d_e, d_off, d_hits, w
        idx.close()
        torch.cuda.empty_cache()
    return out


def copy_ceiling(torch, dev, h2d_bytes, d2h_bytes, barrier, dist, reps=3):
    """Bare concurrent pinned copies of the e2e call's byte counts on two streams (all ranks at once): the box's ceiling for
    that result form.  Returns per-rank GB/s both ways and the queries/s it would allow (filled in by the caller's ratio)."""
    try:
        up_h = torch.empty(h2d_bytes, dtype=torch.uint8).pin_memory()
        dn_h = torch.empty(d2h_bytes, dtype=torch.uint8).pin_memory()
        up_d = torch.empty(h2d_bytes, dtype=torch.uint8, device=dev)
        dn_d = torch.empty(d2h_bytes, dtype=torch.uint8, device=dev)
    except Exception as ex:  # noqa: BLE001
        return {"error": repr(ex)}
    s_up, s_dn = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    best = None
    for _ in range(reps):
        barrier()
        t0 = time.perf_counter()
        with torch.cuda.stream(s_up):
            up_d.copy_(up_h, non_blocking=True)
        with torch.cuda.stream(s_dn):
            dn_h.copy_(dn_d, non_blocking=True)
        s_up.synchronize()
        s_dn.synchronize()
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    t = torch.tensor([best], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dt = float(t.item())
    world = dist.get_world_size() if dist is not None else 1
    nq = h2d_bytes // 12
    return {"ms": dt * 1e3, "h2d_gbs_per_gpu": h2d_bytes / dt / 1e9, "d2h_gbs_per_gpu": d2h_bytes / dt / 1e9,
            "aggregate_gbs": (h2d_bytes + d2h_bytes) * world / dt / 1e9, "queries_per_sec": nq * world / dt,
            "note": "concurrent cudaMemcpyAsync H2D + D2H of exactly the e2e call's bytes, pinned buffers, max over ranks"}


def make_group(torch, dist, dev, rank, world):
    """gg_group in the rank form: the NCCL unique id travels over torch.distributed (plumbing), the communicator is the library's."""
    from genogrove_b200 import DeviceGroup
    idt = torch.zeros(128, dtype=torch.uint8, device=dev)
    if rank == 0:
        idt.copy_(torch.frombuffer(bytearray(DeviceGroup.unique_id()), dtype=torch.uint8))
    if world > 1:
        dist.broadcast(idt, 0)
    return DeviceGroup.from_rank(dev.index, rank, world, bytes(idt.cpu().numpy().tobytes()))


def _device_copy(torch, dev, ptr, n, dtype):
    """A torch tensor holding a copy of n elements of library-owned device memory (plain cudaMemcpy, device to device)."""
    import ctypes as C
    t = torch.empty(max(int(n), 1), dtype=dtype, device=dev)
    if n:
        rt = C.CDLL("libcudart.so.12")
        rc = rt.cudaMemcpy(C.c_void_p(t.data_ptr()), C.c_void_p(ptr), C.c_size_t(int(n) * t.element_size()), 3)
        if rc != 0:
            raise RuntimeError(f"cudaMemcpy D2D failed: {rc}")
    return t[:int(n)]


def _csr_checksum(torch, off, hits, ids):
    """Two wrapping int64 sums over a CSR whose queries carry job-wide ids: sum(count_q * w(id_q)) and
    sum((sum of (rank + 1) over q's hits) * w(id_q)) with w an odd multiplicative hash — independent of where a query was
    answered and in which order the queries are listed, so the routed result can be compared with the replicated one."""
    w = ids * -7046029254386353131 + 1442695040888963407  # (wraps; 0x9E3779B97F4A7C15 as int64)
    cnt = off[1:] - off[:-1]
    cs = torch.zeros(hits.numel() + 1, dtype=torch.int64, device=hits.device)
    torch.cumsum(hits.to(torch.int64) + 1, 0, out=cs[1:])
    seg = cs[off[1:]] - cs[off[:-1]]
    return torch.stack([(cnt * w).sum(), (seg * w).sum(), cnt.sum()])


def chrom_shard_side(args, torch, dist, dev, rank, world, c, s, e, d_c, d_s, d_e, nq, idx, d_off, d_hits, stream, steps=5):
    """The cfg4 grove sharded by chromosome over the N ranks (gg_group_*, GG_SHARD_CHROM): every rank passes its slice of the
    interval input (records routed to their owners), then its own 100M-query batch is routed to the owners and answered there
    (hit lists stay on the owner: return_to_source = 0), and — second number — returned to the source in original order."""
    from genogrove_b200 import SHARD_CHROM
    grp = make_group(torch, dist, dev, rank, world)
    lo, hi = rank * len(c) // world, (rank + 1) * len(c) // world
    grp.build(SHARD_CHROM, c[lo:hi], s[lo:hi], e[lo:hi], 24, want_perm=False)
    grp.build(SHARD_CHROM, c[lo:hi], s[lo:hi], e[lo:hi], 24, want_perm=False)  # warm buffers; report the second
    build_ms = torch.tensor([grp.stats().build_ms], dtype=torch.float64, device=dev)
    dist.all_reduce(build_ms, op=dist.ReduceOp.MAX)
    out = {"build_ms": float(build_ms.item()), "intervals": int(len(c)),
           "owners": grp.owners(24).tolist()}
    # the replicated index's answer to this rank's batch: what the routed results are checked against below
    ref_total = idx.fused_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, d_off.data_ptr(), d_hits.data_ptr(),
                                 d_hits.numel(), stream, sync_total=True)
    torch.cuda.synchronize()
    ref_sum = _csr_checksum(torch, d_off, d_hits[:ref_total], torch.arange(nq, dtype=torch.int64, device=dev) + rank * nq)
    dist.all_reduce(ref_sum)  # (int64 sums wrap the same way on both sides)
    verified = {}
    for back in (False, True):
        def step():
            return grp.intersect_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, nq, return_to_source=back)
        for _ in range(2):
            step()
        dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(steps):
            res = step()
        torch.cuda.synchronize()
        dt = torch.tensor([(time.perf_counter() - t0) / steps], dtype=torch.float64, device=dev)
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        st = grp.stats()
        # ---- parity of the routed result with the replicated index (untimed)
        r_off = _device_copy(torch, dev, res.d_offsets, res.n_queries + 1, torch.int64)
        r_hits = _device_copy(torch, dev, res.d_hits, res.n_hits, torch.int32)
        if back:  # this rank's own CSR in original query order: bit for bit
            same = bool(res.n_queries == nq and res.n_hits == ref_total and torch.equal(r_off, d_off) and
                        torch.equal(r_hits, d_hits[:ref_total]))
            flag = torch.tensor([1 if same else 0], device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            verified["returned_to_source"] = {"equal_to_replicated_csr": bool(flag.item()), "check": "torch.equal on offsets and hits, every rank"}
        else:     # hit lists stay on the owner: job-wide checksums keyed by (source rank, query index)
            origin = _device_copy(torch, dev, res.d_origin, res.n_queries, torch.int64)
            ids = (origin >> 40) * nq + (origin & ((1 << 40) - 1))
            got = _csr_checksum(torch, r_off, r_hits, ids)
            dist.all_reduce(got)
            verified["results_on_owner"] = {"equal_to_replicated_checksums": bool(torch.equal(got, ref_sum)),
                                            "hits_job": int(got[2].item()), "hits_job_replicated": int(ref_sum[2].item()),
                                            "check": "sum(count_q * w(id_q)), sum(sum_hits_q(rank + 1) * w(id_q)), sum(count_q) over all ranks, int64 wrapping"}
            del origin, ids
        del r_off, r_hits
        per = torch.tensor([st.n_answered, st.n_hits, st.bytes_sent, st.route_ms, st.exchange_ms, st.intersect_ms, st.return_ms],
                           dtype=torch.float64, device=dev)
        allp = [torch.zeros_like(per) for _ in range(world)]
        dist.all_gather(allp, per)
        allp = torch.stack(allp).cpu().numpy()
        hits_job = float(allp[:, 1].sum())
        out["results_on_owner" if not back else "returned_to_source"] = {
            "ms_per_step": float(dt.item()) * 1e3, "queries_per_sec": nq * world / float(dt.item()),
            "hit_pairs_per_sec": hits_job / float(dt.item()),
            "answered_imbalance_max_over_mean": float(allp[:, 0].max() / max(allp[:, 0].mean(), 1.0)),
            "nvlink_bytes_sent_per_step": float(allp[:, 2].sum()),
            "phase_ms_max_over_ranks": {"route": float(allp[:, 3].max()), "exchange": float(allp[:, 4].max()),
                                        "intersect": float(allp[:, 5].max()), "return": float(allp[:, 6].max())},
            "timer": "host wall clock around the synchronising C ABI call, max over ranks (the call reads count matrices and hit totals back)"}
        del res
    out["verified"] = verified
    if not all(v.get("equal_to_replicated_csr", v.get("equal_to_replicated_checksums")) for v in verified.values()):
        raise RuntimeError(f"chromosome-sharded result differs from the replicated index: {verified}")
    out["note"] = ("gg_group rank form over NCCL: LPT chromosome owners, device-side stable partition by owner, all-to-all-v of "
                   "(chrom, start, end) as grouped ncclSend/ncclRecv, gg_intersect_fused_device on the owner")
    grp.close()
    return out


def run_cfg5(args, rank, local_rank, world):
    """BASELINE configs[4]: bulk build of 200M intervals + 1B-query batch, chromosome-sharded over the N ranks.
    Every rank holds 1/N of the (unsorted) interval input and 1/N of the query batch — as if each had read its slice of the
    files — and everything else happens behind the C ABI (gg_group_*, rank form): records are routed to the owner of their
    chromosome over NVLink and each rank sorts / builds what it owns; queries are partitioned by owner on the device,
    exchanged, answered on the owner, and the hit lists stay there (1B queries x ~140 hits do not fit anywhere else)."""
    import torch
    import torch.distributed as dist
    from genogrove_b200 import SHARD_CHROM, synth

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29517")
    dist.init_process_group("nccl", device_id=dev, rank=rank, world_size=world)
    grp = make_group(torch, dist, dev, rank, world)
    N, Q, chunk = args.cfg5_intervals, args.cfg5_queries, args.cfg5_chunk
    c, s, e = synth.hg38_intervals_part(N, rank, world, 42)
    dist.barrier()
    t0 = time.perf_counter()
    grp.build(SHARD_CHROM, c, s, e, 24, want_perm=False)
    t_build = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    dist.all_reduce(t_build, op=dist.ReduceOp.MAX)
    dist.barrier()
    t0 = time.perf_counter()
    grp.build(SHARD_CHROM, c, s, e, 24, want_perm=False)  # second build: warm allocations (both reported)
    t_build2 = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    dist.all_reduce(t_build2, op=dist.ReduceOp.MAX)
    ist = IndexStats_of(grp)
    owned = torch.tensor([float(grp.n_owned), ist.sort_ms, ist.levels_ms, ist.total_ms], dtype=torch.float64, device=dev)
    owned_all = [torch.zeros_like(owned) for _ in range(world)]
    dist.all_gather(owned_all, owned)
    owned_all = torch.stack(owned_all).cpu().numpy()
    del c, s, e

    lo, hi = rank * Q // world, (rank + 1) * Q // world
    per_rank_max = -(-Q // world)
    n_chunks = -(-per_rank_max // chunk)
    times, answered, hits, sent = [], 0, 0, 0
    phase = np.zeros(4)
    for k in range(n_chunks):
        a, b = min(lo + k * chunk, hi), min(lo + (k + 1) * chunk, hi)
        m = b - a
        if not m:
            qc, qs, qe = (np.zeros(0, np.uint32),) * 3
        elif args.cfg5_layout == "contiguous":
            qc, qs, qe = synth.hg38_sorted_slice(Q, a, b, 150, 1000)
        else:
            qc, qs, qe = synth.hg38_queries(m, 150, 1000 + rank * 4096 + k, True)
        d_c = torch.from_numpy(np.ascontiguousarray(qc).view(np.int32)).to(dev)
        d_s = torch.from_numpy(qs.astype(np.uint32).view(np.int32)).to(dev)
        d_e = torch.from_numpy(qe.astype(np.uint32).view(np.int32)).to(dev)
        reps = 2 if k == 0 else 1  # first chunk once untimed: buffer growth
        for r_ in range(reps):
            torch.cuda.synchronize()
            dist.barrier()
            t0 = time.perf_counter()
            res = grp.intersect_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, m, return_to_source=False)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
        times.append(dt)
        st = grp.stats()
        answered += int(res.n_queries)
        hits += int(res.n_hits)
        sent += int(st.bytes_sent)
        phase += np.array([st.route_ms, st.exchange_ms, st.intersect_ms, st.return_ms])
        del d_c, d_s, d_e
    t = torch.tensor(times, dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    job_s = float(t.sum().item())
    tot = torch.tensor([float(answered), float(hits), float(sent)] + phase.tolist(), dtype=torch.float64, device=dev)
    tot_all = [torch.zeros_like(tot) for _ in range(world)]
    dist.all_gather(tot_all, tot)
    tot_all = torch.stack(tot_all).cpu().numpy()
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        H = float(tot_all[:, 1].sum())
        bytes_path = 12 * Q + 8 * (Q + world) + 4 * H
        out = {"metric": METRIC, "value": Q / job_s, "unit": UNIT, "n_gpus": world, "steps": 1, "warmup": 1,
               "ms_per_step": job_s * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u32",
               "data": "synthetic", "hit_pairs_per_sec": H / job_s, "hits_per_query": H / Q,
               "config": {"workload": "cfg5: bulk build of %d hg38-shaped intervals (unsorted input, 1/N per rank) + %d 150bp queries, "
                                      "chromosome-sharded over %d B200 (LPT on interval count), hit lists stay on the owner" % (N, Q, world),
                          "intervals": N, "queries": Q, "chunk_queries_per_rank": chunk, "chunks": n_chunks,
                          "query_layout": ("ONE coordinate-sorted 1B-query batch; rank r holds its r-th contiguous part and submits it in "
                                           "consecutive chunks" if args.cfg5_layout == "contiguous" else
                                           "every chunk of every rank is its own coordinate-sorted batch spanning the whole genome"),
                          "parallelism": f"chromosome-sharded x{world} (gg_group rank form, NCCL)"},
               "build": {"seconds_first": float(t_build.item()), "seconds": float(t_build2.item()),
                         "intervals_per_sec": N / float(t_build2.item()),
                         "algorithmic_bytes": 32.3 * N, "frac_of_hbm_peak": 32.3 * N / float(t_build2.item()) / 1e9 / (peak * world),
                         "per_rank_owned_intervals": owned_all[:, 0].tolist(),
                         "per_rank_device_ms": {"sort": owned_all[:, 1].tolist(), "levels": owned_all[:, 2].tolist(),
                                                "build_total": owned_all[:, 3].tolist()},
                         "note": "wall time of gg_group_build, max over ranks: H2D of the rank's records, routing over NVLink, "
                                 "device sort + level construction of the owned chromosomes"},
               "per_rank": {"answered_queries": tot_all[:, 0].tolist(), "hits": tot_all[:, 1].tolist(),
                            "imbalance_answered_max_over_mean": float(tot_all[:, 0].max() / tot_all[:, 0].mean()),
                            "imbalance_hits_max_over_mean": float(tot_all[:, 1].max() / max(tot_all[:, 1].mean(), 1.0))},
               "nvlink_bytes_sent_total": float(tot_all[:, 2].sum()),
               "nvlink_bytes_per_query": float(tot_all[:, 2].sum()) / Q,
               "phase_ms_sum_over_chunks_max_over_ranks": {"route": float(tot_all[:, 3].max()), "exchange": float(tot_all[:, 4].max()),
                                                           "intersect": float(tot_all[:, 5].max()), "return": float(tot_all[:, 6].max())},
               "roofline": {"bound": "hbm", "kernel": "whole path (route + exchange + count + scan + emit)", "achieved": bytes_path / job_s / 1e9,
                            "peak": peak * world, "unit": "GB/s", "frac": bytes_path / job_s / 1e9 / (peak * world), "traffic": None,
                            "algorithmic_bytes": bytes_path},
               "timer": "host wall clock around each synchronising gg_group_intersect_device call (queries resident on the source GPU), "
                        "max over ranks per chunk, summed over chunks",
               "gpu_launches": None}
        print(json.dumps(out))
    grp.close()
    dist.barrier()
    dist.destroy_process_group()


def IndexStats_of(grp):
    """gg_index_stats of the group's local index (ctypes)."""
    from genogrove_b200._lib import IndexStats
    st = IndexStats()
    L = grp._L
    L.gg_index_get_stats(L.gg_group_index(grp._h, 0), ctypes.byref(st))
    return st


# ----------------------------------------------------------------------------------------- B200 arm
def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return
    if args.mode == "cfg5":
        run_cfg5(args, rank, local_rank, world)
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
    # achieved / frac: SURVEY §8(d)'s ALGORITHMIC bytes of the path (12 Q read + 8 (Q + 1) + 4 H written, every datum once)
    # over the dominant kernel's launch time.  What that kernel itself moves — including the cnt / hi scratch round trip
    # and the second read of chrom / start that the two-pass design adds — is reported separately as kernel_bytes.
    achieved = bytes_path / (dom_ms / 1e3) / 1e9
    roofline = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_source": peak_kind,
                "algorithmic_bytes_per_launch": bytes_path, "kernel_bytes_per_launch": dom_bytes,
                "kernel_bytes_gbs": dom_bytes / (dom_ms / 1e3) / 1e9, "kernel_ms": kernel_ms,
                "path_achieved_gbs": bytes_path / (ms_per_step / 1e3) / 1e9,
                "path_frac": bytes_path / (ms_per_step / 1e3) / 1e9 / peak,
                "build": {"algorithmic_bytes": 32.3 * ist.n_intervals, "sort_ms": ist.sort_ms, "levels_ms": ist.levels_ms,
                          "device_ms": ist.sort_ms + ist.levels_ms,
                          "achieved_gbs": 32.3 * ist.n_intervals / ((ist.sort_ms + ist.levels_ms) / 1e3) / 1e9,
                          "frac": 32.3 * ist.n_intervals / ((ist.sort_ms + ist.levels_ms) / 1e3) / 1e9 / peak,
                          "note": "gg_index_build device time (sorts + level construction) over SURVEY 8(d)'s 32.3 B per interval"}}

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
        fls_ms = None
        try:  # flanking with the strand-compatible predicate on the same stranded index (a side number: never fatal)
            d_pred = torch.empty(nq, dtype=torch.int32, device=dev)
            d_succ = torch.empty(nq, dtype=torch.int32, device=dev)
            fls_ms = timed(lambda: idx.flanking_stranded_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), d_str.data_ptr(), 4, nq,
                                                                d_pred.data_ptr(), d_succ.data_ptr(), stream, sync=False), ex_steps)
            del d_pred, d_succ
        except Exception as ex:  # noqa: BLE001
            fls_ms = repr(ex)
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
                  "flanking_stranded": ({"queries_per_sec": nq * world / (fls_ms / 1e3), "ms_per_step": fls_ms, "kernel": "flanking_strand_kernel",
                                         "algorithmic_bytes": 21 * nq, "achieved_gbs": 21 * nq / (fls_ms / 1e3) / 1e9,
                                         "frac": 21 * nq / (fls_ms / 1e3) / 1e9 / peak,
                                         "predicate": "strand-compatible (grove_flanking.ipp:47-54)"} if isinstance(fls_ms, float) else {"error": fls_ms}),
                  "note": "per-rank device time, same index and query batch as the headline; steps=%d" % ex_steps}

    # ---- side measurements that need the headline buffers: the shuffled batch, strong scaling, the chromosome-sharded grove
    shuffled = None
    if not args.no_extras and not args.shuffled:
        perm_t = torch.randperm(nq, device=dev)
        s_c, s_s, s_e = d_c[perm_t].contiguous(), d_s[perm_t].contiguous(), d_e[perm_t].contiguous()
        del perm_t

        def shuf_step():
            idx.fused_device(s_c.data_ptr(), s_s.data_ptr(), s_e.data_ptr(), 4, nq, d_off.data_ptr(), d_hits.data_ptr(),
                             d_hits.numel(), stream, sync_total=True)

        sh_ms = timed_ms(torch, shuf_step, 5)
        shuffled = {"ms_per_step": sh_ms, "queries_per_sec": nq * world / (sh_ms / 1e3),
                    "path_frac": bytes_path / (sh_ms / 1e3) / 1e9 / peak,
                    "note": "the same 100M queries per GPU in random order (SURVEY 8(d) lists both orders); per-rank device time"}
        del s_c, s_s, s_e

    strong = None
    if world > 1 and not args.no_multi:
        # BASELINE cfg4 as written: ONE 100M-query batch (rank 0's) cut into N contiguous slices, index replicated
        gc_, gs_, ge_ = synth.hg38_queries(args.queries, 150, 43, not args.shuffled)
        lo, hi = shard.query_shard(len(gc_), rank, world)
        m = hi - lo
        t_c = torch.from_numpy(gc_[lo:hi].view(np.int32)).to(dev)
        t_s = torch.from_numpy(gs_[lo:hi].astype(np.uint32).view(np.int32)).to(dev)
        t_e = torch.from_numpy(ge_[lo:hi].astype(np.uint32).view(np.int32)).to(dev)
        del gc_, gs_, ge_

        def strong_step():
            idx.fused_device(t_c.data_ptr(), t_s.data_ptr(), t_e.data_ptr(), 4, m, d_off.data_ptr(), d_hits.data_ptr(),
                             d_hits.numel(), stream, sync_total=True)
            d_tot.copy_(d_off[m:m + 1])
            dist.all_gather_into_tensor(d_all, d_tot)  # global CSR base of the slice

        for _ in range(3):
            strong_step()
        barrier()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record()
        for _ in range(args.steps):
            strong_step()
        a1.record()
        barrier()
        t_st = torch.tensor([a0.elapsed_time(a1) / args.steps], dtype=torch.float64, device=dev)
        dist.all_reduce(t_st, op=dist.ReduceOp.MAX)
        strong = {"scaling": "strong", "queries_total": int(args.queries), "queries_per_gpu": m,
                  "ms_per_step": float(t_st.item()), "queries_per_sec": args.queries / (float(t_st.item()) / 1e3),
                  "note": "one 100M-query cfg4 batch cut into N contiguous slices, index replicated; device time, max over ranks"}
        del t_c, t_s, t_e

    chrom_shard = None
    if world > 1 and not args.no_multi:
        chrom_shard = chrom_shard_side(args, torch, dist, dev, rank, world, c, s, e, d_c, d_s, d_e, nq, idx, d_off, d_hits, stream)

    # ---- e2e: host-buffer C ABI calls with pinned buffers, H2D + D2H inside the timed region
    e2e = None
    if not args.no_e2e:
        del d_hits, d_off
        torch.cuda.empty_cache()
        p_c, p_s32, p_e32 = PinnedArray(nq, np.uint32), PinnedArray(nq, np.uint32), PinnedArray(nq, np.uint32)
        p_c.array[:], p_s32.array[:], p_e32.array[:] = qc, qs, qe

        def host_timed(fn):
            fn()  # warm-up: grows the pinned result buffers and the pipeline slots
            fn()
            barrier()
            t0 = time.perf_counter()
            for _ in range(args.e2e_steps):
                h = fn()
            torch.cuda.synchronize()
            dt = (time.perf_counter() - t0) / args.e2e_steps
            assert h == total, (h, total)
            t_e = torch.tensor([dt], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(t_e, op=dist.ReduceOp.MAX)
            return float(t_e.item())

        cres = {}

        def compact_step():
            cres["r"] = idx.intersect_batch_compact(p_c.array, p_s32.array, p_e32.array, expand=False)
            return int(cres["r"].n_hits)

        dt_c = host_timed(compact_step)
        r = cres["r"]
        d2h_c = int(nq * 8 + r.delta_bytes_total)
        widths = sorted({int(r.chunks[k].delta_bytes) for k in range(r.n_chunks)})
        # the same bytes moved by bare concurrent copies on two streams: what the link allows for this result form
        ceiling = copy_ceiling(torch, dev, nq * 12, d2h_c, barrier, dist if world > 1 else None)
        # host-side decode into the plain CSR, for callers that want it (not needed to read hits: gg_compact_rank)
        off_x, hits_x = np.empty(nq + 1, np.uint64), np.empty(total, np.uint32)
        t0 = time.perf_counter()
        st = idx._L.gg_compact_expand(ctypes.byref(r), ctypes.c_void_p(p_c.array.ctypes.data), ctypes.c_void_p(off_x.ctypes.data),
                                      ctypes.c_void_p(hits_x.ctypes.data), 0)
        expand_s = time.perf_counter() - t0
        assert st == 0 and int(off_x[-1]) == total
        del off_x, hits_x

        p_off = PinnedArray(nq + 1, np.uint64)

        def full_step():
            off, _ = idx.intersect_batch(p_c.array, p_s32.array, p_e32.array, offsets_out=p_off.array, copy_hits=False)
            return int(off[-1])

        dt_f = host_timed(full_step)
        e2e = {"value": job_q / dt_c, "unit": UNIT, "h2d_bytes_per_step": int(nq * 12), "d2h_bytes_per_step": d2h_c,
               "ms_per_step": dt_c * 1e3, "steps": args.e2e_steps,
               "api": "gg_intersect_batch_compact (u32 coordinates; per query {offset, hi slot}, per hit a back-delta of "
                      + "/".join(str(w) for w in widths) + " byte(s); pinned host buffers)",
               "timer": "host wall clock around the blocking C ABI call, max over ranks",
               "copy_ceiling": ceiling,
               "frac_of_copy_ceiling": (job_q / dt_c) / ceiling["queries_per_sec"] if ceiling else None,
               "host_expand_to_csr_ms": expand_s * 1e3,
               "full_csr": {"value": job_q / dt_f, "ms_per_step": dt_f * 1e3, "h2d_bytes_per_step": int(nq * 12),
                            "d2h_bytes_per_step": int((nq + 1) * 8 + total * 4),
                            "api": "gg_intersect_batch_u32 (u64 offsets + u32 ranks)"}}
        if world == 1:  # the round-1 e2e call (u64 coordinates up), kept for comparison
            p_s, p_e = PinnedArray(nq, np.uint64), PinnedArray(nq, np.uint64)
            p_s.array[:], p_e.array[:] = qs, qe

            def u64_step():
                off, _ = idx.intersect_batch(p_c.array, p_s.array, p_e.array, offsets_out=p_off.array, copy_hits=False)
                return int(off[-1])

            dt_u = host_timed(u64_step)
            e2e["u64_api"] = {"value": job_q / dt_u, "ms_per_step": dt_u * 1e3, "h2d_bytes_per_step": int(nq * 20),
                              "d2h_bytes_per_step": int((nq + 1) * 8 + total * 4), "api": "gg_intersect_batch (u64 coordinates)"}
            del p_s, p_e

    # ---- CPU baseline beside it (rank 0, N=1 only): the unmodified reference on a bounded sample
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        w = synth.Workload("cfg4", 24, c, s, e, qc, qs, qe)
        v, info = cpu_reference_run(args, w, steps=2, warmup=1)
        cpu = {k: info[k] for k in ("value", "unit", "cores", "kind", "sample")} if v else info
        v1, info1 = cpu_reference_run(args, w, steps=1, warmup=0, threads=1) if v else (None, None)
        if v1:
            cpu["single_thread_value"] = v1

    # ---- BASELINE configs[1] and [2] (parity cases in tests/test_gpu_configs.py; timed here so the driver's record has them)
    if world == 1 and not args.no_extras and not args.shuffled and extras is not None:
        del idx
        torch.cuda.empty_cache()
        extras["configs"] = other_configs(torch, dev, peak)

    if rank == 0:
        out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
               "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
               "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
               "config": workload_config(args, world), "hit_pairs_per_sec": job_hits / (ms_per_step / 1e3),
               "hits_per_query": job_hits / job_q, "gpu_launches": launches_per_step * args.steps, "clocks": clocks,
               "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "extras": extras, "shuffled": shuffled,
               "strong": strong, "chrom_shard": chrom_shard,
               "index": {"levels": ist.n_levels, "fanout": ist.fanout, "staged_levels": ist.staged_levels,
                         "coord_bytes": ist.coord_bytes, "device_mb": ist.device_bytes / 1e6,
                         "build_sort_ms": ist.sort_ms, "build_levels_ms": ist.levels_ms, "build_total_ms": ist.total_ms}}
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
