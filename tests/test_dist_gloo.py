"""CPU: the N>1 host logic with torch.distributed (gloo, world_size 2): query sharding + the path's one collective
(gather of per-rank hit totals -> global CSR offsets), and chromosome-sharded routing.  The per-rank "answer" is
produced by the oracle here (checker role only) — on the GPU box the same logic runs over NCCL in bench.py."""
import os
import socket

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

from helpers import expected_chain, chrom_offsets


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, mode, out_dir):
    import oracle as O
    from genogrove_b200 import shard, synth
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    w = synth.cfg2(20_000, 30_001, seed=5)
    perm = expected_chain(w.chrom, w.start, w.end)
    coff = chrom_offsets(w.chrom, w.n_chrom)
    s, e = w.start[perm], w.end[perm]
    if mode == "query_shard":
        lo, hi = shard.query_shard(len(w.q_chrom), rank, world)
        off, hits = O.flat_intersect_batch(s, e, coff, w.q_chrom[lo:hi], w.q_start[lo:hi], w.q_end[lo:hi])
        totals = shard.gather_hit_totals(int(off[-1]))
        base = shard.global_offset_base(totals, rank)
        np.savez(os.path.join(out_dir, f"r{rank}.npz"), lo=lo, hi=hi, off=off + np.uint64(base), hits=hits,
                 totals=np.array(totals))
    else:
        owner = shard.assign_chromosomes(np.diff(coff.astype(np.int64)), world)
        local = shard.localize_chromosomes(owner, rank)
        mine = np.nonzero(owner == rank)[0]
        # this rank's index: only its chromosomes, renumbered densely
        keep = np.isin(w.chrom[perm], mine)
        l_coff = np.concatenate([[0], np.cumsum([coff[c + 1] - coff[c] for c in mine])]).astype(np.uint64)
        qi = shard.route_queries(w.q_chrom, owner, rank)
        off, hits = O.flat_intersect_batch(s[keep], e[keep], l_coff, local[w.q_chrom[qi]], w.q_start[qi], w.q_end[qi])
        totals = shard.gather_hit_totals(int(off[-1]))
        np.savez(os.path.join(out_dir, f"r{rank}.npz"), qi=qi, off=off, hits=hits, totals=np.array(totals))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["query_shard", "chrom_shard"])
def test_world_size_2_matches_single_process(tmp_path, mode):
    import oracle as O
    from genogrove_b200 import shard, synth
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), mode, str(tmp_path)), nprocs=world, join=True)
    w = synth.cfg2(20_000, 30_001, seed=5)
    perm = expected_chain(w.chrom, w.start, w.end)
    coff = chrom_offsets(w.chrom, w.n_chrom)
    exp_off, exp_hits = O.flat_intersect_batch(w.start[perm], w.end[perm], coff, w.q_chrom, w.q_start, w.q_end)
    parts = [np.load(tmp_path / f"r{r}.npz") for r in range(world)]
    assert all(list(p["totals"]) == list(parts[0]["totals"]) for p in parts)
    assert int(sum(parts[0]["totals"])) == int(exp_off[-1])
    if mode == "query_shard":
        off = np.concatenate([p["off"][:-1] for p in parts] + [parts[-1]["off"][-1:]])
        hits = np.concatenate([p["hits"] for p in parts])
    else:
        off, hits = shard.merge_routed_results(len(w.q_chrom), [(p["qi"], p["off"], p["hits"]) for p in parts])
    assert np.array_equal(off, exp_off) and np.array_equal(hits, exp_hits)
