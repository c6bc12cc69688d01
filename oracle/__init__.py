"""CPU oracle bindings — TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import this package, and only as the checker / the timed CPU baseline.  The product
(``genogrove_b200``) never imports it.

Two libraries (built by ``oracle/Makefile``):

* ``liboracle.so``       — ``oracle/gg_oracle.c``: plain-C restatement of the reference's intersect path
                           (array B+ tree with the reference's node shapes + flat filter).
* ``_ref/libgg_ref.so``  — the UNMODIFIED reference ``grove<interval, uint32_t>`` behind
                           ``oracle/ref_shim.cpp`` (compiled from /root/reference where it lies).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
NO_CHROM = 0xFFFFFFFF

_u64p = C.POINTER(C.c_uint64)
_u32p = C.POINTER(C.c_uint32)


def _p64(a):
    return a.ctypes.data_as(_u64p) if a is not None else None


def _p32(a):
    return a.ctypes.data_as(_u32p) if a is not None else None


def _u64(a):
    return np.ascontiguousarray(a, dtype=np.uint64)


def _u32(a):
    return np.ascontiguousarray(a, dtype=np.uint32)


def build(verbose: bool = False) -> None:
    """Compile liboracle.so and (when /root/reference exists) _ref/libgg_ref.so."""
    r = subprocess.run(["make", "-C", _HERE], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout, r.stderr)
    if r.returncode != 0:
        raise RuntimeError("oracle build failed")


_oracle = None
_ref = None


def oracle_lib():
    global _oracle
    if _oracle is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        L = C.CDLL(path)
        for name in ("orc_overlaps", "orc_less", "orc_greater"):
            getattr(L, name).argtypes = [C.c_uint64] * 4
            getattr(L, name).restype = C.c_int
        L.orc_distribute_evenly.argtypes = [C.c_uint64, C.c_uint64, _u64p]
        L.orc_tree_build_bottom_up.argtypes = [_u64p, _u64p, C.c_size_t, C.c_int]
        L.orc_tree_build_bottom_up.restype = C.c_void_p
        L.orc_tree_free.argtypes = [C.c_void_p]
        L.orc_tree_search_overlaps.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, _u64p, C.c_size_t]
        L.orc_tree_search_overlaps.restype = C.c_size_t
        L.orc_chain_regular.argtypes = [_u64p, _u64p, C.c_size_t, _u32p]
        L.orc_chain_sorted_append.argtypes = [_u64p, _u64p, C.c_size_t, C.c_int, _u32p]
        L.orc_chain_stable_sort.argtypes = [_u64p, _u64p, C.c_size_t, _u32p]
        L.orc_flat_intersect_batch.argtypes = [_u64p, _u64p, _u64p, C.c_uint32, _u32p, _u64p, _u64p,
                                               C.c_size_t, _u64p, _u64p, _u32p]
        L.orc_tree_search_flanking.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, _u64p]
        L.orc_flat_flanking_batch.argtypes = [_u64p, _u64p, _u64p, C.c_uint32, _u32p, _u64p, _u64p, C.c_size_t,
                                              _u32p, _u32p]
        L.orc_strand_order.argtypes = [C.c_int]
        L.orc_gc_less.argtypes = [C.c_int, C.c_uint64, C.c_uint64, C.c_int, C.c_uint64, C.c_uint64]
        L.orc_gc_overlaps.argtypes = [C.c_int, C.c_uint64, C.c_uint64, C.c_int, C.c_uint64, C.c_uint64]
        L.orc_flat_intersect_batch_stranded.argtypes = [_u64p, _u64p, C.c_void_p, _u64p, C.c_uint32, _u32p, _u64p, _u64p,
                                                        C.c_void_p, C.c_size_t, _u64p, _u64p, _u32p]
        _oracle = L
    return _oracle


def ref_available() -> bool:
    return os.path.exists(os.path.join(_HERE, "_ref", "libgg_ref.so"))


def ref_lib():
    global _ref
    if _ref is None:
        path = os.path.join(_HERE, "_ref", "libgg_ref.so")
        if not os.path.exists(path):
            raise FileNotFoundError("oracle/_ref/libgg_ref.so not built (needs /root/reference)")
        L = C.CDLL(path)
        L.ref_create.argtypes = [C.c_int]
        L.ref_create.restype = C.c_void_p
        L.ref_destroy.argtypes = [C.c_void_p]
        L.ref_insert.argtypes = [C.c_void_p, C.c_uint32, C.c_uint64, C.c_uint64, C.c_uint32, C.c_int]
        L.ref_insert_bulk.argtypes = [C.c_void_p, C.c_uint32, _u64p, _u64p, _u32p, C.c_size_t, C.c_int]
        L.ref_remove.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32]
        L.ref_intersect.argtypes = [C.c_void_p, C.c_uint32, C.c_uint64, C.c_uint64, _u32p, C.c_size_t]
        L.ref_intersect.restype = C.c_size_t
        L.ref_intersect_batch.argtypes = [C.c_void_p, _u32p, _u64p, _u64p, C.c_size_t, C.c_int, _u64p,
                                          _u64p, _u32p, _u64p]
        L.ref_intersect_batch.restype = C.c_double
        for name in ("ref_overlaps", "ref_less", "ref_greater"):
            getattr(L, name).argtypes = [C.c_uint64] * 4
            getattr(L, name).restype = C.c_int
        L.ref_distribute_evenly.argtypes = [C.c_uint64, C.c_uint64, _u64p]
        L.ref_hardware_threads.restype = C.c_int
        L.ref_gc_create.argtypes = [C.c_int]
        L.ref_gc_create.restype = C.c_void_p
        L.ref_gc_destroy.argtypes = [C.c_void_p]
        L.ref_gc_insert.argtypes = [C.c_void_p, C.c_uint32, C.c_char, C.c_uint64, C.c_uint64, C.c_uint32, C.c_int]
        L.ref_gc_insert_bulk.argtypes = [C.c_void_p, C.c_uint32, C.c_char_p, _u64p, _u64p, _u32p, C.c_size_t, C.c_int]
        L.ref_gc_intersect.argtypes = [C.c_void_p, C.c_uint32, C.c_char, C.c_uint64, C.c_uint64, _u32p, C.c_size_t]
        L.ref_gc_intersect.restype = C.c_size_t
        L.ref_gc_flanking.argtypes = [C.c_void_p, C.c_uint32, C.c_char, C.c_uint64, C.c_uint64, C.c_int, _u32p]
        L.ref_gc_flanking.restype = C.c_int
        for name in ("ref_gc_less", "ref_gc_overlaps"):
            getattr(L, name).argtypes = [C.c_char, C.c_uint64, C.c_uint64, C.c_char, C.c_uint64, C.c_uint64]
            getattr(L, name).restype = C.c_int
        L.ref_flanking_batch.argtypes = [C.c_void_p, _u32p, _u64p, _u64p, C.c_size_t, _u32p, _u32p]
        _ref = L
    return _ref


# ----------------------------------------------------------------------------- C restatement


def distribute_evenly(total: int, max_per_group: int, lib=None):
    out = np.zeros(3, dtype=np.uint64)
    (lib or oracle_lib()).orc_distribute_evenly(total, max_per_group, _p64(out))
    return tuple(int(x) for x in out)


class OracleTree:
    """Array B+ tree with the reference's shapes (build_tree_bottom_up) + search_overlaps."""

    def __init__(self, start, end, order: int = 3):
        self.start = _u64(start)
        self.end = _u64(end)
        self.n = len(self.start)
        self._L = oracle_lib()
        self._t = self._L.orc_tree_build_bottom_up(_p64(self.start), _p64(self.end), self.n, order)

    def intersect(self, qs: int, qe: int) -> np.ndarray:
        if not self._t:
            return np.zeros(0, dtype=np.uint64)
        cap = 64
        while True:
            out = np.zeros(cap, dtype=np.uint64)
            k = self._L.orc_tree_search_overlaps(self._t, qs, qe, _p64(out), cap)
            if k <= cap:
                return out[:k]
            cap = k

    def flanking(self, qs: int, qe: int):
        """(predecessor rank, successor rank) by the restated search_flanking; None when absent."""
        out = np.zeros(2, dtype=np.uint64)
        self._L.orc_tree_search_flanking(self._t, qs, qe, _p64(out))
        none = 0xFFFFFFFFFFFFFFFF
        return tuple(None if int(v) == none else int(v) for v in out)

    def __del__(self):
        if getattr(self, "_t", None):
            self._L.orc_tree_free(self._t)
            self._t = None


def chain_regular(start, end) -> np.ndarray:
    s, e = _u64(start), _u64(end)
    perm = np.zeros(len(s), dtype=np.uint32)
    oracle_lib().orc_chain_regular(_p64(s), _p64(e), len(s), _p32(perm))
    return perm


def chain_sorted_append(start, end, order: int) -> np.ndarray:
    s, e = _u64(start), _u64(end)
    perm = np.zeros(len(s), dtype=np.uint32)
    oracle_lib().orc_chain_sorted_append(_p64(s), _p64(e), len(s), order, _p32(perm))
    return perm


def chain_stable_sort(start, end) -> np.ndarray:
    s, e = _u64(start), _u64(end)
    perm = np.zeros(len(s), dtype=np.uint32)
    oracle_lib().orc_chain_stable_sort(_p64(s), _p64(e), len(s), _p32(perm))
    return perm


def flat_intersect_batch(start, end, coff, qchrom, qs, qe):
    """Flat oracle over concatenated per-chromosome chains.  Returns (offsets u64[Q+1], ranks u32[H])
    with per-chromosome chain ranks, ascending per query."""
    L = oracle_lib()
    s, e, co = _u64(start), _u64(end), _u64(coff)
    qc, a, b = _u32(qchrom), _u64(qs), _u64(qe)
    nq = len(qc)
    counts = np.zeros(nq, dtype=np.uint64)
    L.orc_flat_intersect_batch(_p64(s), _p64(e), _p64(co), len(co) - 1, _p32(qc), _p64(a), _p64(b), nq,
                               _p64(counts), None, None)
    offsets = np.zeros(nq + 1, dtype=np.uint64)
    np.cumsum(counts, out=offsets[1:])
    out = np.zeros(max(int(offsets[-1]), 1), dtype=np.uint32)
    L.orc_flat_intersect_batch(_p64(s), _p64(e), _p64(co), len(co) - 1, _p32(qc), _p64(a), _p64(b), nq,
                               None, _p64(offsets), _p32(out))
    return offsets, out[: int(offsets[-1])]


NO_RANK = 0xFFFFFFFF


def flat_flanking_batch(start, end, coff, qchrom, qs, qe):
    """Flat flanking oracle over concatenated per-chromosome chains: (pred u32[Q], succ u32[Q]) per-chromosome
    chain ranks, NO_RANK when absent."""
    L = oracle_lib()
    s, e, co = _u64(start), _u64(end), _u64(coff)
    qc, a, b = _u32(qchrom), _u64(qs), _u64(qe)
    nq = len(qc)
    pred = np.zeros(nq, dtype=np.uint32)
    succ = np.zeros(nq, dtype=np.uint32)
    L.orc_flat_flanking_batch(_p64(s), _p64(e), _p64(co), len(co) - 1, _p32(qc), _p64(a), _p64(b), nq,
                              _p32(pred), _p32(succ))
    return pred, succ


STRAND_ORDER = {"*": 0, ".": 1, "+": 2, "-": 3}


def _u8(a):
    if isinstance(a, (bytes, bytearray, str)):
        a = np.frombuffer(a.encode() if isinstance(a, str) else bytes(a), dtype=np.uint8)
    return np.ascontiguousarray(a, dtype=np.uint8)


def flat_intersect_batch_stranded(start, end, strand, coff, qchrom, qs, qe, qstrand):
    """Flat oracle for genomic_coordinate keys: chains (start, end, strand byte) concatenated per chromosome.
    Returns (offsets u64[Q+1], per-chromosome ranks u32[H])."""
    L = oracle_lib()
    s, e, co, st = _u64(start), _u64(end), _u64(coff), _u8(strand)
    qc, a, b, qst = _u32(qchrom), _u64(qs), _u64(qe), _u8(qstrand)
    nq = len(qc)
    counts = np.zeros(nq, dtype=np.uint64)
    args = (_p64(s), _p64(e), st.ctypes.data, _p64(co), len(co) - 1, _p32(qc), _p64(a), _p64(b), qst.ctypes.data, nq)
    L.orc_flat_intersect_batch_stranded(*args, _p64(counts), None, None)
    offsets = np.zeros(nq + 1, dtype=np.uint64)
    np.cumsum(counts, out=offsets[1:])
    out = np.zeros(max(int(offsets[-1]), 1), dtype=np.uint32)
    L.orc_flat_intersect_batch_stranded(*args, None, _p64(offsets), _p32(out))
    return offsets, out[: int(offsets[-1])]


def flat_flanking_batch_stranded(start, end, strand, coff, qchrom, qs, qe, qstrand):
    """Flat restatement of grove<genomic_coordinate>::flanking(q, index, strand-compatible predicate)
    (include/genogrove/structure/grove/grove_flanking.ipp:47-82 -> detail::search_flanking,
    include/genogrove/structure/grove/query_engine.hpp:238-311; leaf rule :251-281) over chains (start, end, strand byte)
    concatenated per chromosome: a candidate must be strand-compatible (equal strands, or either '*') and not overlap
    the query; predecessor = FIRST key in the chain among the candidates with the largest end < q.start, successor =
    first candidate in the chain whose start > q.end.  Returns (pred u32[Q], succ u32[Q]) per-chromosome ranks, NO_RANK
    when absent.  Plain numpy, one vectorised pass per query (test infrastructure: small cases only)."""
    s, e, co, st = _u64(start), _u64(end), _u64(coff), _u8(strand)
    qc, a, b, qst = _u32(qchrom), _u64(qs), _u64(qe), _u8(qstrand)
    nq, n_chrom = len(qc), len(co) - 1
    pred = np.full(nq, NO_RANK, dtype=np.uint32)
    succ = np.full(nq, NO_RANK, dtype=np.uint32)
    star = np.uint8(ord("*"))
    for i in range(nq):
        if qc[i] >= n_chrom:
            continue
        lo, hi = int(co[qc[i]]), int(co[qc[i] + 1])
        cs, ce, cst = s[lo:hi], e[lo:hi], st[lo:hi]
        ok = np.ones(hi - lo, dtype=bool) if qst[i] == star else (cst == qst[i]) | (cst == star)
        before = ok & (ce < a[i])
        if before.any():
            pred[i] = int(np.flatnonzero(before & (ce == ce[before].max()))[0])
        after = np.flatnonzero(ok & (cs > b[i]))
        if len(after):
            succ[i] = int(after[0])
    return pred, succ


def gc_chain(start, end, strand, mode: str = "stable"):
    """Leaf-chain order of genomic_coordinate keys: ascending (start, end, strand order); equal keys in input order
    (mode 'stable': sorted bulk) or reverse input order (mode 'regular': one insert_data per record)."""
    s, e, st = _u64(start), _u64(end), _u8(strand)
    so = np.array([STRAND_ORDER[chr(c)] for c in st], dtype=np.int64) if len(st) else np.zeros(0, np.int64)
    idx = np.arange(len(s), dtype=np.int64)
    return np.lexsort(((-idx if mode == "regular" else idx), so, e, s)).astype(np.uint32)


class RefGcGrove:
    """The unmodified reference grove<genomic_coordinate, uint32_t> (payload = id)."""

    def __init__(self, order: int = 3):
        self._L = ref_lib()
        self._g = self._L.ref_gc_create(order)
        if not self._g:
            raise ValueError("grove order must be >= 3")

    def insert(self, chrom: int, strand: str, start: int, end: int, id_: int, sorted_: bool = False) -> None:
        if self._L.ref_gc_insert(self._g, chrom, strand.encode(), start, end, id_, 1 if sorted_ else 0) != 0:
            raise ValueError("genomic_coordinate: invalid strand or start > end")

    def insert_bulk(self, chrom: int, strand, start, end, ids, presorted: bool = True) -> None:
        st, s, e, i = _u8(strand), _u64(start), _u64(end), _u32(ids)
        if self._L.ref_gc_insert_bulk(self._g, chrom, st.tobytes(), _p64(s), _p64(e), _p32(i), len(s), int(presorted)) != 0:
            raise ValueError("genomic_coordinate: invalid strand or start > end")

    def intersect(self, chrom: int, strand: str, qs: int, qe: int) -> np.ndarray:
        cap = 64
        while True:
            out = np.zeros(cap, dtype=np.uint32)
            k = self._L.ref_gc_intersect(self._g, chrom, strand.encode(), qs, qe, _p32(out), cap)
            if k == 2**64 - 1:
                raise ValueError("genomic_coordinate: invalid strand or start > end")
            if k <= cap:
                return out[:k]
            cap = k

    def flanking(self, chrom: int, strand: str, qs: int, qe: int, strand_predicate: bool = False):
        """grove.flanking(q, chrom[, strand-compatible predicate]) -> (pred id, succ id), NO_RANK (0xFFFFFFFF) when absent."""
        out = np.zeros(2, dtype=np.uint32)
        if self._L.ref_gc_flanking(self._g, chrom, strand.encode(), qs, qe, 1 if strand_predicate else 0, _p32(out)) != 0:
            raise ValueError("genomic_coordinate: invalid strand or start > end")
        return int(out[0]), int(out[1])

    def close(self):
        if getattr(self, "_g", None):
            self._L.ref_gc_destroy(self._g)
            self._g = None

    def __del__(self):
        self.close()


# ----------------------------------------------------------------------------- the real reference


class RefGrove:
    """The unmodified reference grove<interval, uint32_t> (payload = id)."""

    REGULAR, SORTED = 0, 1

    def __init__(self, order: int = 3):
        self._L = ref_lib()
        self._g = self._L.ref_create(order)
        if not self._g:
            raise ValueError("grove order must be >= 3")

    def insert(self, chrom: int, start: int, end: int, id_: int, mode: int = 0) -> None:
        if self._L.ref_insert(self._g, chrom, start, end, id_, mode) != 0:
            raise ValueError("interval: start must be <= end")

    def insert_bulk(self, chrom: int, start, end, ids, presorted: bool = False) -> None:
        s, e, i = _u64(start), _u64(end), _u32(ids)
        if self._L.ref_insert_bulk(self._g, chrom, _p64(s), _p64(e), _p32(i), len(s), int(presorted)) != 0:
            raise ValueError("interval: start must be <= end")

    def remove(self, chrom: int, id_: int) -> bool:
        return bool(self._L.ref_remove(self._g, chrom, id_))

    def intersect(self, chrom: int, qs: int, qe: int) -> np.ndarray:
        cap = 64
        while True:
            out = np.zeros(cap, dtype=np.uint32)
            k = self._L.ref_intersect(self._g, chrom, qs, qe, _p32(out), cap)
            if k <= cap:
                return out[:k]
            cap = k

    def intersect_batch(self, qchrom, qs, qe, threads: int = 1, want_ids: bool = True):
        """Returns (offsets, ids, seconds_of_query_loop)."""
        qc, a, b = _u32(qchrom), _u64(qs), _u64(qe)
        nq = len(qc)
        counts = np.zeros(nq, dtype=np.uint64)
        chk = C.c_uint64(0)
        secs = self._L.ref_intersect_batch(self._g, _p32(qc), _p64(a), _p64(b), nq, threads, _p64(counts),
                                           None, None, C.byref(chk))
        offsets = np.zeros(nq + 1, dtype=np.uint64)
        np.cumsum(counts, out=offsets[1:])
        if not want_ids:
            return offsets, None, secs
        ids = np.zeros(max(int(offsets[-1]), 1), dtype=np.uint32)
        self._L.ref_intersect_batch(self._g, _p32(qc), _p64(a), _p64(b), nq, threads, None, _p64(offsets),
                                    _p32(ids), C.byref(chk))
        return offsets, ids[: int(offsets[-1])], secs

    def flanking_batch(self, qchrom, qs, qe):
        """grove.flanking(q, chrom) per query: (pred ids, succ ids), NO_RANK when absent."""
        qc, a, b = _u32(qchrom), _u64(qs), _u64(qe)
        pred = np.zeros(len(qc), dtype=np.uint32)
        succ = np.zeros(len(qc), dtype=np.uint32)
        self._L.ref_flanking_batch(self._g, _p32(qc), _p64(a), _p64(b), len(qc), _p32(pred), _p32(succ))
        return pred, succ

    def hardware_threads(self) -> int:
        return int(self._L.ref_hardware_threads())

    def close(self):
        if getattr(self, "_g", None):
            self._L.ref_destroy(self._g)
            self._g = None

    def __del__(self):
        self.close()
