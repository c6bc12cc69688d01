"""DeviceGroup — Python owner of one ``gg_group`` handle: a grove over several B200s (C ABI: gg_group_*, include/gg_b200.h).

Two forms, as in the C ABI: ``DeviceGroup(devices=[0, 1, ...])`` drives several GPUs from this process (peer copies over
NVLink); ``DeviceGroup.from_rank(device, rank, world, unique_id)`` is one rank of a torchrun / MPI job (NCCL communicator
owned by the handle; the 128-byte id comes from ``DeviceGroup.unique_id()`` on rank 0 and is distributed by the launcher's
own plumbing, e.g. ``torch.distributed.broadcast``).  Nothing here computes on the CPU.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import GGError
from .index import _arr, _ptr

SHARD_REPLICATE, SHARD_CHROM = 0, 1


class GroupStats(C.Structure):
    _fields_ = [("n_queries", C.c_uint64), ("n_answered", C.c_uint64), ("n_hits", C.c_uint64), ("bytes_sent", C.c_uint64),
                ("bytes_received", C.c_uint64), ("route_ms", C.c_double), ("exchange_ms", C.c_double),
                ("intersect_ms", C.c_double), ("return_ms", C.c_double), ("build_ms", C.c_double)]


class GroupDeviceResult(C.Structure):
    _fields_ = [("n_queries", C.c_uint64), ("n_hits", C.c_uint64), ("hit_base", C.c_uint64), ("d_offsets", C.c_void_p),
                ("d_hits", C.c_void_p), ("d_origin", C.c_void_p)]


def shard_assign(counts, world: int) -> np.ndarray:
    """owner[c] by longest-processing-time packing (gg_shard_assign; host arithmetic, no GPU needed)."""
    L = _lib.lib()
    c = _arr(counts, np.uint64)
    owner = np.zeros(len(c), dtype=np.uint32)
    st = L.gg_shard_assign(_ptr(c), len(c), int(world), _ptr(owner))
    if st != 0:
        raise GGError(st, "gg_shard_assign")
    return owner


class DeviceGroup:
    def __init__(self, devices=(0,), _handle=None):
        self._L = _lib.lib()
        self._h = C.c_void_p()
        if _handle is not None:
            self._h = _handle
            return
        devs = (C.c_int * len(devices))(*[int(d) for d in devices])
        st = self._L.gg_group_create(C.byref(self._h), devs, len(devices))
        if st != 0:
            self._h = None
            raise GGError(st, "gg_group_create: no usable CUDA device (genogrove_b200 has no CPU fallback)")

    @staticmethod
    def unique_id() -> bytes:
        buf = C.create_string_buffer(128)
        st = _lib.lib().gg_group_unique_id(buf)
        if st != 0:
            raise GGError(st, "gg_group_unique_id: libnccl.so.2 not loadable")
        return buf.raw

    @classmethod
    def from_rank(cls, device: int, rank: int, world: int, unique_id: bytes) -> "DeviceGroup":
        L = _lib.lib()
        h = C.c_void_p()
        st = L.gg_group_create_rank(C.byref(h), int(device), int(rank), int(world), C.c_char_p(unique_id))
        if st != 0:
            raise GGError(st, "gg_group_create_rank")
        return cls(_handle=h)

    def _check(self, st: int):
        if st != 0:
            raise GGError(st, (self._L.gg_group_last_error(self._h) or b"").decode())

    def close(self):
        if getattr(self, "_h", None):
            self._L.gg_group_destroy(self._h)
            self._h = None

    def __del__(self):
        self.close()

    @property
    def world(self) -> int:
        return int(self._L.gg_group_world(self._h))

    def set_option(self, key: str, value: int, local: int = 0):
        st = self._L.gg_set_option(self._L.gg_group_index(self._h, local), key.encode(), int(value))
        if st != 0:
            raise GGError(st, f"gg_set_option({key})")

    def stats(self) -> GroupStats:
        s = GroupStats()
        self._check(self._L.gg_group_get_stats(self._h, C.byref(s)))
        return s

    def owners(self, n_chrom: int) -> np.ndarray:
        o = np.zeros(n_chrom, dtype=np.uint32)
        self._check(self._L.gg_group_owners(self._h, _ptr(o)))
        return o

    def build(self, shard_mode: int, chrom, start, end, n_chrom: int, tie_mode: int = 0, tie_key=None,
              want_perm: bool = True, owned_capacity: int | None = None):
        """gg_group_build.  Returns rank_to_input (u64; single-process form: n entries, chromosome-major; rank form: the
        entries of the chromosomes this rank owns) or None."""
        c, s, e = _arr(chrom, np.uint32), _arr(start, np.uint64), _arr(end, np.uint64)
        t = _arr(tie_key, np.uint64) if tie_key is not None else None
        cap = int(owned_capacity if owned_capacity is not None else len(c))
        perm = np.empty(cap, dtype=np.uint64) if want_perm else None
        n_owned = C.c_size_t(0)
        self._check(self._L.gg_group_build(self._h, int(shard_mode), _ptr(c), _ptr(s), _ptr(e), _ptr(t), int(tie_mode), len(c),
                                           int(n_chrom), _ptr(perm), cap, C.byref(n_owned)))
        self.n_owned = int(n_owned.value)
        return perm[:self.n_owned] if want_perm else None

    def intersect_batch(self, q_chrom, q_start, q_end):
        """gg_group_intersect_batch: (offsets u64[Q+1], hits u32[H], hit_base).  u32 coordinate arrays take the 4-byte path."""
        qc = _arr(q_chrom, np.uint32)
        a, b = np.asarray(q_start), np.asarray(q_end)
        u32 = a.dtype == np.uint32 and b.dtype == np.uint32
        a, b = _arr(a, np.uint32 if u32 else np.uint64), _arr(b, np.uint32 if u32 else np.uint64)
        nq = len(qc)
        offsets = np.empty(nq + 1, dtype=np.uint64)
        hp = C.POINTER(C.c_uint32)()
        nh, hb = C.c_uint64(0), C.c_uint64(0)
        self._check(self._L.gg_group_intersect_batch(self._h, _ptr(qc), _ptr(a), _ptr(b), 4 if u32 else 8, nq, _ptr(offsets),
                                                     C.byref(hp), C.byref(nh), C.byref(hb)))
        n_hits = int(nh.value)
        hits = np.ctypeslib.as_array(hp, shape=(n_hits,)).copy() if n_hits else np.zeros(0, dtype=np.uint32)
        return offsets, hits, int(hb.value)

    def intersect_device(self, d_chrom: int, d_start: int, d_end: int, coord_bytes: int, nq: int,
                         return_to_source: bool = True) -> GroupDeviceResult:
        res = GroupDeviceResult()
        self._check(self._L.gg_group_intersect_device(self._h, d_chrom, d_start, d_end, int(coord_bytes), int(nq),
                                                      1 if return_to_source else 0, C.byref(res)))
        return res
