"""Host-side mirror of genogrove's ``grove<interval, T>`` for the intersect path (Python harness flavour;
the C++20 drop-in facade is include/genogrove_b200/grove.hpp).

Same names, argument meaning and error behaviour as the reference for this path:

* ``Interval(start, end)`` raises ``ValueError`` when ``start > end``       (interval.hpp:70-74, std::invalid_argument)
* ``Grove(order)`` raises ``ValueError`` when ``order < 3``; default 3        (grove.hpp:158-167)
* ``insert_data(index, interval, data)``                                      (grove_insert.ipp:36-43)
* ``insert_data(index, interval, data, sorted=True)``                         (grove_insert.ipp:19-26)
* ``insert_data_bulk(index, pairs, presorted=False)``                         (grove_insert.ipp:68-127,160-172)
* ``intersect(query, index)`` / ``intersect(query)`` -> ``QueryResult``       (grove_query.ipp:16-43)
* ``remove_key(index, key)``                                                  (grove_remove.ipp:24-62)
* unknown index -> empty result, not an error                                 (grove_query.ipp:37-39)

The host owns the keys (stable identity, like the reference's ``std::deque<key>``) and, per index, the
LEAF-CHAIN ORDER of key objects; the device holds a derived, flattened snapshot of that order
(DeviceIndex) which is rebuilt lazily after any mutation (dirty flag).  All overlap computation happens
in the CUDA library — there is no CPU search path here.

Leaf-chain order per insert path:
  regular insert      -> lower_bound insertion: equal keys in reverse insertion order
  sorted bulk         -> input order
  unsorted bulk       -> device radix sort, stable (equal keys in input order); the reference uses an
                         UNSTABLE std::sort there, so the order of equal keys is unspecified by it
  sorted single append-> lower_bound inside the rightmost leaf only, leaf splits at order-1
"""
from __future__ import annotations

import bisect
from typing import Any, Iterable, List, Optional, Sequence, Tuple

import numpy as np

from .index import DeviceIndex, NO_CHROM, NO_RANK, TIE_INPUT_ORDER


class Interval:
    """Closed interval [start, end] of two unsigned 64-bit coordinates (interval.hpp:54-221)."""

    __slots__ = ("start", "end")
    INVALID_POSITION = 2**64 - 1

    def __init__(self, start: int = INVALID_POSITION, end: int = INVALID_POSITION):
        if start > end:
            raise ValueError("interval: start must be <= end")
        self.start, self.end = int(start), int(end)

    def get_start(self) -> int:
        return self.start

    def get_end(self) -> int:
        return self.end

    def __lt__(self, o):  # interval.hpp:87-92
        return (self.start, self.end) < (o.start, o.end)

    def __gt__(self, o):  # interval.hpp:100-105
        return (self.start, self.end) > (o.start, o.end)

    def __eq__(self, o):
        return isinstance(o, Interval) and (self.start, self.end) == (o.start, o.end)

    def __hash__(self):
        return hash((self.start, self.end))

    def __repr__(self):
        return f"[{self.start},{self.end}]"

    def to_string(self) -> str:
        return repr(self)


class Key:
    """key<interval, T>: value + payload; object identity is what query results hand back (key.hpp:79-347)."""

    __slots__ = ("value", "data")

    def __init__(self, value: Interval, data: Any = None):
        self.value, self.data = value, data

    def get_value(self) -> Interval:
        return self.value

    def get_data(self) -> Any:
        return self.data


class QueryResult:
    """query copy + borrowed key references in leaf-chain order (query_result.hpp:59-118)."""

    def __init__(self, query: Interval):
        self.query = query
        self.keys: List[Key] = []

    def get_query(self) -> Interval:
        return self.query

    def get_keys(self) -> List[Key]:
        return self.keys

    def add_key(self, key: Key):
        if key is None:
            raise ValueError("query_result::add_key: key must not be nullptr")
        self.keys.append(key)


class FlankingQueryResult:
    """flanking_query_result<key_t, data_t>: borrowed nearest predecessor / successor, None when absent
    (data_type/flanking_query_result.hpp)."""

    __slots__ = ("predecessor", "successor")

    def __init__(self, predecessor: Optional[Key] = None, successor: Optional[Key] = None):
        self.predecessor, self.successor = predecessor, successor

    def get_predecessor(self) -> Optional[Key]:
        return self.predecessor

    def get_successor(self) -> Optional[Key]:
        return self.successor


class _Chain:
    """Leaf chain of one index (chromosome) on the host: key objects in chain order."""

    __slots__ = ("keys", "sort_keys", "leaf_begin", "leaf_known")

    def __init__(self):
        self.keys: List[Key] = []
        self.sort_keys: List[Tuple[int, int]] = []  # (start, end) parallel to keys, for bisect
        self.leaf_begin = 0       # first rank of the rightmost leaf (sorted-append emulation)
        self.leaf_known = True    # False once a path ran whose leaf boundaries we do not track


class Grove:
    def __init__(self, order: int = 3, device: int = 0):
        if order < 3:
            raise ValueError("grove order must be >= 3")
        self.order = order
        self._chains: dict[str, _Chain] = {}   # insertion-ordered index name -> chain
        self._dev = DeviceIndex(device)
        self._dirty = True
        self._names: List[str] = []

    def get_order(self) -> int:
        return self.order

    # ------------------------------------------------------------------ inserts
    def _chain(self, index: str) -> _Chain:
        ch = self._chains.get(index)
        if ch is None:
            ch = self._chains[index] = _Chain()
        return ch

    def insert_data(self, index: str, key_value: Interval, data: Any = None, sorted: bool = False) -> Key:
        ch = self._chain(index)
        key = Key(key_value, data)
        sk = (key_value.start, key_value.end)
        if not sorted:
            pos = bisect.bisect_left(ch.sort_keys, sk)        # lower_bound over the whole chain
            ch.leaf_known = False
        else:
            lo = ch.leaf_begin if ch.leaf_known else 0
            pos = bisect.bisect_left(ch.sort_keys, sk, lo)    # lower_bound inside the rightmost leaf
        ch.keys.insert(pos, key)
        ch.sort_keys.insert(pos, sk)
        if sorted and ch.leaf_known and len(ch.keys) - ch.leaf_begin == self.order:
            ch.leaf_begin = len(ch.keys) - 1                  # asymmetric split, mid = order - 1
        self._dirty = True
        return key

    def insert_data_bulk(self, index: str, data: Sequence[Tuple[Interval, Any]], presorted: bool = False) -> List[Key]:
        """insert_data(index, vec, bulk) / insert_data(index, vec, sorted, bulk)."""
        if len(data) == 0:
            return []
        ch = self._chain(index)
        keys = [Key(iv, d) for iv, d in data]
        if not presorted:
            # device radix sort of this batch (stable); one chromosome, so chrom ids are all zero
            s = np.fromiter((k.value.start for k in keys), dtype=np.uint64, count=len(keys))
            e = np.fromiter((k.value.end for k in keys), dtype=np.uint64, count=len(keys))
            perm = self._dev.build(np.zeros(len(keys), np.uint32), s, e, 1, TIE_INPUT_ORDER)
            keys_sorted = [keys[i] for i in perm]
        else:
            keys_sorted = keys
        if not ch.keys:
            # bottom-up build: the rightmost leaf holds the last group of distribute_evenly(n, order-1)
            n, cap = len(keys_sorted), self.order - 1
            groups = n // cap + (1 if n % cap else 0)
            ch.keys = list(keys_sorted)
            ch.sort_keys = [(k.value.start, k.value.end) for k in ch.keys]
            ch.leaf_begin = n - n // groups
            ch.leaf_known = True
        else:
            for k in keys_sorted:  # rightmost append loop (grove_insert.ipp:106-126)
                sk = (k.value.start, k.value.end)
                lo = ch.leaf_begin if ch.leaf_known else 0
                pos = bisect.bisect_left(ch.sort_keys, sk, lo)
                ch.keys.insert(pos, k)
                ch.sort_keys.insert(pos, sk)
                if ch.leaf_known and len(ch.keys) - ch.leaf_begin == self.order:
                    ch.leaf_begin = len(ch.keys) - 1
        self._dirty = True
        return keys if presorted else keys_sorted

    def remove_key(self, index: str, key: Optional[Key]) -> bool:
        if key is None:
            return False
        ch = self._chains.get(index)
        if ch is None:
            return False
        for i, k in enumerate(ch.keys):
            if k is key:
                del ch.keys[i]
                del ch.sort_keys[i]
                ch.leaf_known = False
                if not ch.keys:
                    del self._chains[index]
                self._dirty = True
                return True
        return False

    def size(self) -> int:
        return sum(len(c.keys) for c in self._chains.values())

    # ------------------------------------------------------------------ device snapshot
    def chrom_id(self, index: str) -> int:
        """u32 id of an index name in the current snapshot, NO_CHROM when absent (get_root -> nullptr)."""
        self._sync()
        try:
            return self._names.index(index)
        except ValueError:
            return NO_CHROM

    def _sync(self):
        if not self._dirty:
            return
        self._names = list(self._chains.keys())
        n = self.size()
        chrom = np.empty(n, np.uint32)
        start = np.empty(n, np.uint64)
        end = np.empty(n, np.uint64)
        p = 0
        for cid, name in enumerate(self._names):
            ch = self._chains[name]
            m = len(ch.keys)
            chrom[p:p + m] = cid
            if m:
                arr = np.array(ch.sort_keys, dtype=np.uint64).reshape(m, 2)
                start[p:p + m] = arr[:, 0]
                end[p:p + m] = arr[:, 1]
            p += m
        self._dev.upload_sorted(chrom, start, end, max(len(self._names), 1))
        self._dirty = False

    # ------------------------------------------------------------------ queries
    def intersect(self, query: Interval, index: Optional[str] = None) -> QueryResult:
        """One query = a batch of one (all indices: one query per index, in map iteration order)."""
        self._sync()
        res = QueryResult(query)
        if index is not None:
            cid = self.chrom_id(index)
            if cid == NO_CHROM:
                return res
            cids = [cid]
        else:
            cids = list(range(len(self._names)))
        if not cids:
            return res
        k = len(cids)
        off, hits = self._dev.intersect_batch(np.array(cids, np.uint32), np.full(k, query.start, np.uint64),
                                              np.full(k, query.end, np.uint64))
        for j, cid in enumerate(cids):
            keys = self._chains[self._names[cid]].keys
            for r in hits[int(off[j]):int(off[j + 1])]:
                res.add_key(keys[int(r)])
        return res

    def intersect_batch(self, chrom_ids, starts, ends):
        """The new batch entry point: CSR (offsets u64[Q+1], per-chromosome chain ranks u32[H])."""
        self._sync()
        return self._dev.intersect_batch(chrom_ids, starts, ends)

    def flanking(self, query: Interval, index: str) -> FlankingQueryResult:
        """grove::flanking(query, index) with the default predicate (grove_flanking.ipp): a flanking batch of one."""
        self._sync()
        cid = self.chrom_id(index)
        if cid == NO_CHROM:
            return FlankingQueryResult()
        pred, succ = self._dev.flanking_batch(np.array([cid], np.uint32), np.array([query.start], np.uint64),
                                              np.array([query.end], np.uint64))
        keys = self._chains[self._names[cid]].keys
        return FlankingQueryResult(None if int(pred[0]) == NO_RANK else keys[int(pred[0])],
                                   None if int(succ[0]) == NO_RANK else keys[int(succ[0])])

    def flanking_batch(self, chrom_ids, starts, ends):
        """Batch form: (pred u32[Q], succ u32[Q]) per-chromosome chain ranks, NO_RANK when absent."""
        self._sync()
        return self._dev.flanking_batch(chrom_ids, starts, ends)

    def keys_at(self, chrom_id: int, ranks: Iterable[int]) -> List[Key]:
        keys = self._chains[self._names[chrom_id]].keys
        return [keys[int(r)] for r in ranks]

    @property
    def device_index(self) -> DeviceIndex:
        self._sync()
        return self._dev
