"""GPU tests of the multi-GPU handle (gg_group_*, genogrove_b200/csrc/gg_group.cu): a grove over several GPUs must return
exactly what one gg_index returns for the same intervals and queries — same offsets, same ranks, same order.

On a one-GPU box the single-process form is exercised with the SAME device listed several times (each entry is its own rank
with its own gg_index, streams and buffers; the all-to-all then runs over device-to-device copies instead of NVLink peer
copies): the routing kernels, the partition, the count matrices, the return path and the re-ordering are all the real ones.
With two or more GPUs the peer-copy form and the one-process-per-GPU form (NCCL, launched with torchrun) run too."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from helpers import expected_chain, oracle_csr, random_index, random_queries

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def gg():
    import genogrove_b200 as g
    return g


def _n_gpus():
    import genogrove_b200._lib as L
    return L.lib().gg_device_count()


def _devices(world):
    n = max(_n_gpus(), 1)
    return [i % n for i in range(world)]


@pytest.mark.parametrize("world", [1, 2, 3, 8])
@pytest.mark.parametrize("mode", ["replicate", "chrom"])
def test_group_matches_single_index(gg, oracle, world, mode):
    rng = np.random.default_rng(100 + world + (7 if mode == "chrom" else 0))
    n_chrom = 11
    chrom, start, end = random_index(rng, 120_000, n_chrom, span=600_000, max_len=2500, dup=True, nested=True)
    chrom[chrom == 4] = 3  # an empty chromosome
    grp = gg.DeviceGroup(_devices(world))
    assert grp.world == world
    perm = grp.build(gg.SHARD_CHROM if mode == "chrom" else gg.SHARD_REPLICATE, chrom, start, end, n_chrom)
    assert np.array_equal(perm, expected_chain(chrom, start, end).astype(np.uint64))
    if mode == "chrom":
        owner = grp.owners(n_chrom)
        counts = np.bincount(chrom, minlength=n_chrom).astype(np.uint64)
        assert np.array_equal(owner, gg.shard_assign(counts, world))
        assert owner.max() < world
    for nq, wide in ((70_001, False), (70_001, True), (1, False), (0, False), (5, True)):
        qc, qs, qe = random_queries(rng, nq, n_chrom, span=600_000, max_len=900, extra_chrom=2)
        if nq > 10:
            qc[::53] = gg.NO_CHROM
        exp_off, exp_hits = oracle_csr(chrom, start, end, n_chrom, qc, qs, qe)
        a, b = (qs, qe) if wide else (qs.astype(np.uint32), qe.astype(np.uint32))
        off, hits, base = grp.intersect_batch(qc, a, b)
        assert base == 0
        assert np.array_equal(off, exp_off), (world, mode, nq, wide)
        assert np.array_equal(hits, exp_hits)
        st = grp.stats()
        assert st.n_queries == nq and st.n_hits == len(exp_hits)
        if mode == "chrom" and world > 1 and nq > 1000:
            assert st.n_answered == nq and st.bytes_sent == st.bytes_received and st.bytes_sent > 0
    grp.close()


def test_group_tie_modes_and_explicit_keys(gg, oracle):
    """The chromosome-sharded build keeps the reference's chain order among equal keys: every GG_TIE_* mode gives the same
    permutation as one gg_index_build over all chromosomes."""
    rng = np.random.default_rng(211)
    n_chrom = 5
    chrom, start, end = random_index(rng, 40_000, n_chrom, span=20_000, max_len=300, dup=True)
    tie = rng.integers(0, 4, len(chrom)).astype(np.uint64)
    for tie_mode in (0, 1, 2):
        grp = gg.DeviceGroup(_devices(3))
        perm = grp.build(gg.SHARD_CHROM, chrom, start, end, n_chrom, tie_mode=tie_mode, tie_key=tie if tie_mode == 2 else None)
        exp = expected_chain(chrom, start, end, tie_mode, tie)
        assert np.array_equal(perm, exp.astype(np.uint64))
        qc, qs, qe = random_queries(rng, 20_000, n_chrom, span=20_000, max_len=200)
        exp_off, exp_hits = oracle_csr(chrom, start, end, n_chrom, qc, qs, qe, exp)
        off, hits, _ = grp.intersect_batch(qc, qs, qe)
        assert np.array_equal(off, exp_off) and np.array_equal(hits, exp_hits)
        grp.close()


def test_group_device_call_results_stay_on_owner(gg, oracle):
    """gg_group_intersect_device with return_to_source = 0 (a group of one GPU): the result lists the answered queries with
    their origin ids; with return_to_source = 1 it is the slice's CSR in original order."""
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(223)
    n_chrom = 6
    chrom, start, end = random_index(rng, 80_000, n_chrom, span=400_000, max_len=2000)
    grp = gg.DeviceGroup([0])
    grp.build(gg.SHARD_CHROM, chrom, start, end, n_chrom, want_perm=False)
    qc, qs, qe = random_queries(rng, 33_333, n_chrom, span=400_000, max_len=500, extra_chrom=1)
    exp_off, exp_hits = oracle_csr(chrom, start, end, n_chrom, qc, qs, qe)
    dev = torch.device("cuda:0")
    d_c = torch.from_numpy(qc.view(np.int32)).to(dev)
    d_s = torch.from_numpy(qs.astype(np.uint32).view(np.int32)).to(dev)
    d_e = torch.from_numpy(qe.astype(np.uint32).view(np.int32)).to(dev)
    torch.cuda.synchronize()
    for back in (True, False):
        res = grp.intersect_device(d_c.data_ptr(), d_s.data_ptr(), d_e.data_ptr(), 4, len(qc), return_to_source=back)
        assert res.n_queries == len(qc) and res.n_hits == len(exp_hits)
        off = np.empty(len(qc) + 1, np.uint64)
        hits = np.empty(len(exp_hits), np.uint32)
        import ctypes as C
        cudart = C.CDLL("libcudart.so.12")
        assert cudart.cudaMemcpy(C.c_void_p(off.ctypes.data), C.c_void_p(res.d_offsets), C.c_size_t(off.nbytes), 2) == 0
        assert cudart.cudaMemcpy(C.c_void_p(hits.ctypes.data), C.c_void_p(res.d_hits), C.c_size_t(hits.nbytes), 2) == 0
        assert np.array_equal(off, exp_off) and np.array_equal(hits, exp_hits)
        if not back:
            origin = np.empty(len(qc), np.uint64)
            assert cudart.cudaMemcpy(C.c_void_p(origin.ctypes.data), C.c_void_p(res.d_origin), C.c_size_t(origin.nbytes), 2) == 0
            assert np.array_equal(origin, np.arange(len(qc), dtype=np.uint64))  # rank 0, arrival order = original order
    grp.close()


_RANK_SCRIPT = r'''
import os, sys, json
sys.path.insert(0, os.environ["GG_ROOT"]); sys.path.insert(0, os.path.join(os.environ["GG_ROOT"], "tests"))
import numpy as np, torch, torch.distributed as dist
import genogrove_b200 as gg
from helpers import oracle_csr, random_index, random_queries, expected_chain
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
if rank == 0:
    idt.copy_(torch.frombuffer(bytearray(gg.DeviceGroup.unique_id()), dtype=torch.uint8))
dist.broadcast(idt, 0)
grp = gg.DeviceGroup.from_rank(rank, rank, world, bytes(idt.cpu().numpy().tobytes()))
rng = np.random.default_rng(5)
n_chrom = 9
chrom, start, end = random_index(rng, 150_000, n_chrom, span=700_000, max_len=2500, dup=True, nested=True)
qc, qs, qe = random_queries(rng, 200_003, n_chrom, span=700_000, max_len=800, extra_chrom=1)
exp_off, exp_hits = oracle_csr(chrom, start, end, n_chrom, qc, qs, qe)
ok = True
# every rank holds a slice of the INPUT (as if each had read a part of the BED file) and a slice of the batch
lo_i, hi_i = rank * len(chrom) // world, (rank + 1) * len(chrom) // world
lo, hi = rank * len(qc) // world, (rank + 1) * len(qc) // world
for mode in (gg.SHARD_CHROM, gg.SHARD_REPLICATE):
    if mode == gg.SHARD_CHROM:
        perm = grp.build(mode, chrom[lo_i:hi_i], start[lo_i:hi_i], end[lo_i:hi_i], n_chrom, owned_capacity=len(chrom))
        owner = grp.owners(n_chrom)
        chain = expected_chain(chrom, start, end)
        mine = np.concatenate([chain[chrom[chain] == c] for c in range(n_chrom) if owner[c] == rank] or [np.zeros(0, np.int64)])
        ok &= bool(np.array_equal(perm, mine.astype(np.uint64)))
    else:
        grp.build(mode, chrom, start, end, n_chrom, want_perm=False)
    off, hits, base = grp.intersect_batch(qc[lo:hi], qs[lo:hi].astype(np.uint32), qe[lo:hi].astype(np.uint32))
    ok &= bool(base == int(exp_off[lo]))
    ok &= bool(np.array_equal(off + np.uint64(base), exp_off[lo:hi + 1]))
    ok &= bool(np.array_equal(hits, exp_hits[int(exp_off[lo]):int(exp_off[hi])]))
    st = grp.stats()
    if mode == gg.SHARD_CHROM:
        ok &= bool(st.bytes_sent > 0)
# ---- a failure on ONE rank ends the call on EVERY rank with that rank's status: nobody is left waiting in a collective
def status_of(fn):
    try:
        fn()
        return 0, ""
    except gg.GGError as e:
        return e.status, str(e)
bad = chrom[lo_i:hi_i].copy()
if rank == world - 1:
    bad[len(bad) // 2] = n_chrom + 5                       # one bad record in one rank's part of the input
st, msg = status_of(lambda: grp.build(gg.SHARD_CHROM, bad, start[lo_i:hi_i], end[lo_i:hi_i], n_chrom, want_perm=False))
ok &= st == 1 and (("rank %d failed" % (world - 1)) in msg if rank == 0 else "chrom id" in msg)
# rank_to_input too small on one rank: known only after the count matrix, reported before the exchange
st, msg = status_of(lambda: grp.build(gg.SHARD_CHROM, chrom[lo_i:hi_i], start[lo_i:hi_i], end[lo_i:hi_i], n_chrom,
                                      owned_capacity=1 if rank == world - 1 else len(chrom)))
ok &= st == 7
for mode in (gg.SHARD_CHROM, gg.SHARD_REPLICATE):
    grp.build(mode, chrom[lo_i:hi_i] if mode == gg.SHARD_CHROM else chrom, start[lo_i:hi_i] if mode == gg.SHARD_CHROM else start,
              end[lo_i:hi_i] if mode == gg.SHARD_CHROM else end, n_chrom, want_perm=False)
    big = np.tile(np.arange(lo, hi), 3)                    # a larger slice than any before: the upload buffers must grow
    if rank == world - 1:
        grp.set_option("inject_alloc_failure", 1)          # ... and on one rank the first allocation fails
    ok &= status_of(lambda: grp.intersect_batch(qc[big], qs[big].astype(np.uint32), qe[big].astype(np.uint32)))[0] == 3
    grp.set_option("inject_alloc_failure", 0)
    # the group is usable afterwards
    off, hits, base = grp.intersect_batch(qc[lo:hi], qs[lo:hi].astype(np.uint32), qe[lo:hi].astype(np.uint32))
    ok &= bool(np.array_equal(off + np.uint64(base), exp_off[lo:hi + 1]))
    ok &= bool(np.array_equal(hits, exp_hits[int(exp_off[lo]):int(exp_off[hi])]))
t = torch.tensor([1 if ok else 0], device="cuda")
dist.all_reduce(t, op=dist.ReduceOp.MIN)
if rank == 0:
    print(json.dumps({"ok": bool(t.item()), "world": world}))
grp.close()
dist.destroy_process_group()
'''


@pytest.mark.parametrize("world", [2])
def test_group_rank_form_over_nccl(gg, world, tmp_path):
    """One process per GPU (torchrun), NCCL communicator owned by the handle: chromosome-sharded build from per-rank input
    slices (records routed to their owners over NVLink), routed queries, results returned in original order."""
    if _n_gpus() < world:
        pytest.skip(f"needs {world} GPUs")
    script = tmp_path / "rank_script.py"
    script.write_text(_RANK_SCRIPT)
    env = dict(os.environ, GG_ROOT=ROOT)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", "29541", str(script)],
                       capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
    assert json.loads(line) == {"ok": True, "world": world}


def test_group_two_gpus_peer_copies(gg, oracle):
    if _n_gpus() < 2:
        pytest.skip("needs 2 GPUs")
    rng = np.random.default_rng(301)
    n_chrom = 24
    chrom, start, end = random_index(rng, 400_000, n_chrom, span=3_000_000, max_len=3000)
    grp = gg.DeviceGroup(list(range(min(_n_gpus(), 8))))
    grp.build(gg.SHARD_CHROM, chrom, start, end, n_chrom, want_perm=False)
    qc, qs, qe = random_queries(rng, 1_000_003, n_chrom, span=3_000_000, max_len=300)
    exp_off, exp_hits = oracle_csr(chrom, start, end, n_chrom, qc, qs, qe)
    off, hits, _ = grp.intersect_batch(qc, qs.astype(np.uint32), qe.astype(np.uint32))
    assert np.array_equal(off, exp_off) and np.array_equal(hits, exp_hits)
    grp.close()
