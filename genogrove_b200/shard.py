"""Multi-GPU sharding of the intersect path: one process per GPU, torch.distributed for plumbing.

The path shards without any data-path exchange:

* **replicate index, shard queries** (default): every rank holds the whole device index; rank r answers the
  contiguous query slice ``query_shard(nq, r, world)`` into a local CSR.  The only collective is the final
  gather of the per-rank hit totals (8 bytes per rank), which turns local CSR offsets into global ones.
* **shard by chromosome** (large genomes): chromosomes are bin-packed onto ranks by interval count (LPT);
  a rank builds only its chromosomes, queries are routed to the owner of their chromosome and answered
  locally; results carry the original query index.  Same final gather.

Everything here is host logic (numpy + torch.distributed); it runs with the gloo backend on CPU in tests and
with NCCL over NVLink on the GPU box.
"""
from __future__ import annotations

from typing import List, Tuple

import numpy as np

from .synth import lpt_assign


def query_shard(nq: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous slice [lo, hi) of a batch of nq queries owned by `rank` (sizes differ by at most 1)."""
    base, extra = divmod(nq, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def assign_chromosomes(counts, world: int) -> np.ndarray:
    """owner[c] = rank holding chromosome c (longest-processing-time packing on interval counts)."""
    return lpt_assign(counts, world)


def route_queries(q_chrom: np.ndarray, owner: np.ndarray, rank: int) -> np.ndarray:
    """Indices (ascending, i.e. original order kept) of the queries whose chromosome `rank` owns.
    Unknown chromosomes (id >= len(owner), incl. NO_CHROM) have no owner and yield no hits anywhere."""
    q_chrom = np.asarray(q_chrom)
    known = q_chrom < len(owner)
    own = np.zeros(len(q_chrom), dtype=bool)
    own[known] = owner[q_chrom[known]] == rank
    return np.nonzero(own)[0]


def localize_chromosomes(owner: np.ndarray, rank: int) -> np.ndarray:
    """Map global chrom id -> local id on `rank` (dense, ascending), or 0xFFFFFFFF when not owned."""
    local = np.full(len(owner), 0xFFFFFFFF, dtype=np.uint32)
    mine = np.nonzero(owner == rank)[0]
    local[mine] = np.arange(len(mine), dtype=np.uint32)
    return local


def gather_hit_totals(local_total: int, device=None) -> List[int]:
    """The path's one collective: all-gather of per-rank hit totals (u64 each).  NCCL when `device` is a CUDA
    device, gloo otherwise.  Returns the list of totals indexed by rank."""
    import torch
    import torch.distributed as dist

    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size() == 1:
        return [int(local_total)]
    t = torch.tensor([int(local_total)], dtype=torch.int64, device=device if device is not None else "cpu")
    out = torch.empty(dist.get_world_size(), dtype=torch.int64, device=t.device)
    dist.all_gather_into_tensor(out, t)
    return [int(x) for x in out.cpu().tolist()]


def global_offset_base(totals: List[int], rank: int) -> int:
    """Exclusive prefix of the gathered totals: what rank adds to its local CSR offsets."""
    return int(sum(totals[:rank]))


def merge_routed_results(nq: int, parts) -> Tuple[np.ndarray, np.ndarray]:
    """Chromosome-sharded mode: combine per-rank (query_index, offsets, hits) into one CSR in original query
    order.  Each query was answered by at most one rank.  Vectorised (no per-query loop): the destination of hit k of a
    part is offsets[query of k] + (k - first hit of that query), both obtained with one np.repeat per part — the host-side
    twin of gather_lists_kernel (genogrove_b200/csrc/gg_group.cu), which does this on the source GPU inside
    gg_group_intersect_batch; this function only serves host-side tests and tools."""
    counts = np.zeros(nq, dtype=np.uint64)
    for qidx, off, _ in parts:
        counts[np.asarray(qidx, dtype=np.int64)] = np.diff(np.asarray(off, dtype=np.uint64))
    offsets = np.zeros(nq + 1, dtype=np.uint64)
    np.cumsum(counts, out=offsets[1:])
    hits = np.zeros(int(offsets[-1]), dtype=np.uint32)
    for qidx, off, h in parts:
        qidx = np.asarray(qidx, dtype=np.int64)
        off = np.asarray(off, dtype=np.int64)
        n = int(off[-1] - off[0]) if len(off) else 0
        if n == 0:
            continue
        per = np.diff(off)
        src_first = np.repeat(off[:-1], per)                      # first hit (in the part) of the query each hit belongs to
        dst_first = np.repeat(offsets[qidx].astype(np.int64), per)
        k = np.arange(int(off[0]), int(off[-1]), dtype=np.int64)
        hits[dst_first + (k - src_first)] = np.asarray(h)[k]
    return offsets, hits
