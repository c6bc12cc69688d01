"""DeviceIndex — thin Python owner of one ``gg_index`` handle (C ABI, include/gg_b200.h).

Host arrays are numpy; the device-buffer entry points take raw device pointers (e.g. ``tensor.data_ptr()``)
so torch is only plumbing for memory and streams.  Nothing here computes on the CPU.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import GGError, IndexStats, BatchStats, NO_CHROM, NO_RANK, TIE_EXPLICIT, TIE_INPUT_ORDER, TIE_REVERSE_INPUT  # noqa: F401


def _ptr(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


def _arr(a, dtype):
    a = np.asarray(a)
    if a.dtype != dtype or not a.flags.c_contiguous:
        a = np.ascontiguousarray(a, dtype=dtype)
    return a


class PinnedArray:
    """numpy view over cudaHostAlloc'ed memory (gg_host_alloc); freed with the object."""

    def __init__(self, shape, dtype):
        self._L = _lib.lib()
        dtype = np.dtype(dtype)
        n = int(np.prod(shape))
        self._p = C.c_void_p()
        st = self._L.gg_host_alloc(C.byref(self._p), max(n * dtype.itemsize, 16))
        if st != 0:
            raise GGError(st, "gg_host_alloc")
        buf = (C.c_char * max(n * dtype.itemsize, 16)).from_address(self._p.value)
        self.array = np.frombuffer(buf, dtype=dtype, count=n).reshape(shape)

    def __del__(self):
        p = getattr(self, "_p", None)
        if p is not None and p.value:
            self.array = None
            self._L.gg_host_free(p)
            self._p = None


class DeviceIndex:
    def __init__(self, device: int = 0):
        self._L = _lib.lib()
        self._h = C.c_void_p()
        st = self._L.gg_index_create(C.byref(self._h), device)
        if st != 0:
            self._h = None
            raise GGError(st, "gg_index_create: no usable CUDA device (genogrove_b200 has no CPU fallback)")
        self.device = device

    # ------------------------------------------------------------------ plumbing
    def _check(self, st: int):
        if st != 0:
            raise GGError(st, (self._L.gg_last_error(self._h) or b"").decode())

    def close(self):
        if getattr(self, "_h", None):
            self._L.gg_index_destroy(self._h)
            self._h = None

    def __del__(self):
        self.close()

    def set_option(self, key: str, value: int):
        self._check(self._L.gg_set_option(self._h, key.encode(), int(value)))

    # ------------------------------------------------------------------ build
    def build(self, chrom, start, end, n_chrom: int, tie_mode: int = TIE_INPUT_ORDER, tie_key=None,
              want_perm: bool = True):
        """Device sort + level construction.  Returns rank_to_input (u32[n]) or None."""
        c, s, e = _arr(chrom, np.uint32), _arr(start, np.uint64), _arr(end, np.uint64)
        n = len(c)
        t = _arr(tie_key, np.uint64) if tie_key is not None else None
        perm = np.empty(n, dtype=np.uint32) if want_perm else None
        self._check(self._L.gg_index_build(self._h, _ptr(c), _ptr(s), _ptr(e), _ptr(t), tie_mode, n, n_chrom, _ptr(perm)))
        return perm

    def upload_sorted(self, chrom, start, end, n_chrom: int):
        c, s, e = _arr(chrom, np.uint32), _arr(start, np.uint64), _arr(end, np.uint64)
        self._check(self._L.gg_index_upload_sorted(self._h, _ptr(c), _ptr(s), _ptr(e), len(c), n_chrom))

    def update_chrom(self, chrom: int, start, end):
        """Replace one chromosome's leaf chain in the snapshot (chain order); only these records are uploaded."""
        s, e = _arr(start, np.uint64), _arr(end, np.uint64)
        self._check(self._L.gg_index_update_chrom(self._h, int(chrom), _ptr(s), _ptr(e), len(s)))

    def stats(self) -> IndexStats:
        st = IndexStats()
        self._check(self._L.gg_index_get_stats(self._h, C.byref(st)))
        return st

    def chrom_counts(self) -> np.ndarray:
        out = np.zeros(self.stats().n_chrom, dtype=np.uint64)
        self._check(self._L.gg_index_chrom_counts(self._h, _ptr(out)))
        return out

    def batch_stats(self) -> BatchStats:
        st = BatchStats()
        self._check(self._L.gg_last_batch_stats(self._h, C.byref(st)))
        return st

    # ------------------------------------------------------------------ host-buffer intersect
    def intersect_batch(self, q_chrom, q_start, q_end, offsets_out=None, copy_hits: bool = True):
        """count -> scan -> emit through the pipelined host-buffer call.  Coordinates u64 (or u32 arrays,
        which take the u32 entry point).  Returns (offsets u64[Q+1], hits u32[H])."""
        qc = _arr(q_chrom, np.uint32)
        a, b = np.asarray(q_start), np.asarray(q_end)
        u32 = a.dtype == np.uint32 and b.dtype == np.uint32
        a, b = _arr(a, np.uint32 if u32 else np.uint64), _arr(b, np.uint32 if u32 else np.uint64)
        nq = len(qc)
        offsets = offsets_out if offsets_out is not None else np.empty(nq + 1, dtype=np.uint64)
        hp = C.POINTER(C.c_uint32)()
        nh = C.c_uint64(0)
        fn = self._L.gg_intersect_batch_u32 if u32 else self._L.gg_intersect_batch
        self._check(fn(self._h, _ptr(qc), _ptr(a), _ptr(b), nq, _ptr(offsets), C.byref(hp), C.byref(nh)))
        n_hits = int(nh.value)
        if n_hits == 0:
            return offsets, np.zeros(0, dtype=np.uint32)
        view = np.ctypeslib.as_array(hp, shape=(n_hits,))
        return offsets, (view.copy() if copy_hits else view)

    def intersect_batch_compact(self, q_chrom, q_start, q_end, expand: bool = True, threads: int = 0):
        """gg_intersect_batch_compact (u32 coordinates): the PCIe-lean result form — 8 B per query + a 1/2/4-byte
        back-delta per hit.  expand=True decodes it on the host (gg_compact_expand) into the plain (offsets, hits) CSR;
        expand=False returns the raw ``CompactResult`` (library-owned pinned arrays, valid until the next call)."""
        qc, a, b = _arr(q_chrom, np.uint32), _arr(q_start, np.uint32), _arr(q_end, np.uint32)
        nq = len(qc)
        res = _lib.CompactResult()
        self._check(self._L.gg_intersect_batch_compact(self._h, _ptr(qc), _ptr(a), _ptr(b), nq, C.byref(res)))
        self.last_compact = res
        if not expand:
            return res
        offsets = np.empty(nq + 1, dtype=np.uint64)
        hits = np.empty(int(res.n_hits), dtype=np.uint32)
        st = self._L.gg_compact_expand(C.byref(res), _ptr(qc), _ptr(offsets), _ptr(hits) if len(hits) else None, threads)
        if st != 0:
            raise GGError(st, "gg_compact_expand")
        return offsets, hits

    def intersect_two_phase(self, q_chrom, q_start, q_end):
        """gg_intersect_count + gg_intersect_emit with a caller-owned hit buffer."""
        qc, a, b = _arr(q_chrom, np.uint32), _arr(q_start, np.uint64), _arr(q_end, np.uint64)
        nq = len(qc)
        offsets = np.empty(nq + 1, dtype=np.uint64)
        self._check(self._L.gg_intersect_count(self._h, _ptr(qc), _ptr(a), _ptr(b), nq, _ptr(offsets)))
        hits = np.empty(int(offsets[-1]), dtype=np.uint32)
        self._check(self._L.gg_intersect_emit(self._h, _ptr(hits), len(hits)))
        return offsets, hits

    # ------------------------------------------------------------------ device-buffer intersect
    def count_device(self, d_chrom: int, d_start: int, d_end: int, coord_bytes: int, nq: int, d_offsets: int,
                     stream: int = 0, sync_total: bool = False):
        total = C.c_uint64(0)
        self._check(self._L.gg_intersect_count_device(self._h, d_chrom, d_start, d_end, coord_bytes, nq, d_offsets,
                                                      C.byref(total) if sync_total else None, stream or None))
        return int(total.value) if sync_total else None

    def emit_device(self, d_chrom: int, d_start: int, d_end: int, coord_bytes: int, nq: int, d_offsets: int,
                    d_hits: int, stream: int = 0):
        self._check(self._L.gg_intersect_emit_device(self._h, d_chrom, d_start, d_end, coord_bytes, nq, d_offsets,
                                                     d_hits, stream or None))

    def fused_device(self, d_chrom: int, d_start: int, d_end: int, coord_bytes: int, nq: int, d_offsets: int,
                     d_hits: int, capacity: int, stream: int = 0, sync_total: bool = False):
        """Single-call path (count -> block-total scan -> emit_scan; one look-back kernel with option fused_kernel).  Raises GG_ERR_CAPACITY (status 7) when sync_total
        and the batch produced more than `capacity` hits; ``last_total`` then holds the needed size."""
        total = C.c_uint64(0)
        st = self._L.gg_intersect_fused_device(self._h, d_chrom, d_start, d_end, coord_bytes, nq, d_offsets, d_hits,
                                               int(capacity), C.byref(total) if sync_total else None, stream or None)
        self.last_total = int(total.value) if sync_total else None
        self._check(st)
        return self.last_total

    def fused_device_static(self, d_chrom: int, d_start: int, d_end: int, coord_bytes: int, nq: int, d_offsets: int,
                            d_hits: int, capacity: int, stream: int = 0, sync_total: bool = False):
        """gg_intersect_fused_device_static: same contract as :meth:`fused_device`, for callers that may not allocate
        device memory while querying (their HBM is fully handed to a pool, or cudaMalloc's stall is unacceptable).  The
        library's scratch is a fixed array on the device, so at most ``STATIC_MAX_QUERIES`` queries per call (status 7
        beyond that) and calls that overlap on one GPU race — keep them on one stream."""
        total = C.c_uint64(0)
        st = self._L.gg_intersect_fused_device_static(self._h, d_chrom, d_start, d_end, coord_bytes, nq, d_offsets, d_hits,
                                                      int(capacity), C.byref(total) if sync_total else None, stream or None)
        self.last_total = int(total.value) if sync_total else None
        self._check(st)
        return self.last_total

    # ------------------------------------------------------------------ flanking (nearest predecessor / successor)
    def flanking_batch(self, q_chrom, q_start, q_end):
        """gg_flanking_batch: (pred u32[Q], succ u32[Q]) per-chromosome chain ranks, NO_RANK when absent."""
        qc, a, b = _arr(q_chrom, np.uint32), _arr(q_start, np.uint64), _arr(q_end, np.uint64)
        nq = len(qc)
        pred = np.empty(nq, dtype=np.uint32)
        succ = np.empty(nq, dtype=np.uint32)
        self._check(self._L.gg_flanking_batch(self._h, _ptr(qc), _ptr(a), _ptr(b), nq, _ptr(pred), _ptr(succ)))
        return pred, succ

    def flanking_device(self, d_chrom: int, d_start: int, d_end: int, coord_bytes: int, nq: int, d_pred: int,
                        d_succ: int, stream: int = 0, sync: bool = True):
        self._check(self._L.gg_flanking_device(self._h, d_chrom, d_start, d_end, coord_bytes, nq, d_pred, d_succ,
                                               1 if sync else 0, stream or None))

    def flanking_batch_stranded(self, q_chrom, q_start, q_end, q_strand):
        """gg_flanking_batch_stranded: flanking on genomic_coordinate keys with the strand-compatible predicate."""
        qc, a, b = _arr(q_chrom, np.uint32), _arr(q_start, np.uint64), _arr(q_end, np.uint64)
        qst = _arr(np.frombuffer(q_strand, dtype=np.uint8) if isinstance(q_strand, (bytes, bytearray)) else q_strand, np.uint8)
        nq = len(qc)
        pred = np.empty(nq, dtype=np.uint32)
        succ = np.empty(nq, dtype=np.uint32)
        self._check(self._L.gg_flanking_batch_stranded(self._h, _ptr(qc), _ptr(a), _ptr(b), _ptr(qst), nq, _ptr(pred), _ptr(succ)))
        return pred, succ

    def flanking_stranded_device(self, d_chrom: int, d_start: int, d_end: int, d_strand: int, coord_bytes: int, nq: int,
                                 d_pred: int, d_succ: int, stream: int = 0, sync: bool = True):
        self._check(self._L.gg_flanking_stranded_device(self._h, d_chrom, d_start, d_end, d_strand, coord_bytes, nq, d_pred,
                                                        d_succ, 1 if sync else 0, stream or None))

    # ------------------------------------------------------------------ strand-aware intersect (genomic_coordinate)
    def set_strands(self, strands):
        """One strand byte (b'+-.*') per indexed interval in chain order (chromosomes concatenated in id order)."""
        a = _arr(np.frombuffer(strands, dtype=np.uint8) if isinstance(strands, (bytes, bytearray)) else strands, np.uint8)
        self._check(self._L.gg_index_set_strands(self._h, _ptr(a), len(a)))

    def intersect_batch_stranded(self, q_chrom, q_start, q_end, q_strand):
        """gg_intersect_batch_stranded: (offsets u64[Q+1], hits u32[H])."""
        qc, a, b = _arr(q_chrom, np.uint32), _arr(q_start, np.uint64), _arr(q_end, np.uint64)
        st = _arr(np.frombuffer(q_strand, dtype=np.uint8) if isinstance(q_strand, (bytes, bytearray)) else q_strand, np.uint8)
        nq = len(qc)
        offsets = np.empty(nq + 1, dtype=np.uint64)
        hp = C.POINTER(C.c_uint32)()
        nh = C.c_uint64(0)
        self._check(self._L.gg_intersect_batch_stranded(self._h, _ptr(qc), _ptr(a), _ptr(b), _ptr(st), nq, _ptr(offsets),
                                                        C.byref(hp), C.byref(nh)))
        n_hits = int(nh.value)
        if n_hits == 0:
            return offsets, np.zeros(0, dtype=np.uint32)
        return offsets, np.ctypeslib.as_array(hp, shape=(n_hits,)).copy()

    def stranded_device(self, d_chrom: int, d_start: int, d_end: int, d_strand: int, coord_bytes: int, nq: int,
                        d_offsets: int, d_hits: int, capacity: int, stream: int = 0, sync_total: bool = True):
        total = C.c_uint64(0)
        st = self._L.gg_intersect_stranded_device(self._h, d_chrom, d_start, d_end, d_strand, coord_bytes, nq, d_offsets,
                                                  d_hits, int(capacity), C.byref(total) if sync_total else None, stream or None)
        self.last_total = int(total.value) if sync_total else None
        self._check(st)
        return self.last_total
