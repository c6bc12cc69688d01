"""Shared helpers for the parity tests (CPU side: the oracle is the checker)."""
from __future__ import annotations

import numpy as np

import oracle as O


def expected_chain(chrom, start, end, tie_mode=0, tie_key=None):
    """rank -> input index, chromosomes concatenated in id order; ties per gg_tie_mode."""
    n = len(chrom)
    idx = np.arange(n, dtype=np.int64)
    if tie_mode == 0:
        last = idx
        keys = (last, end, start, chrom)
    elif tie_mode == 1:
        keys = (-idx, end, start, chrom)
    else:
        keys = (idx, np.asarray(tie_key), end, start, chrom)
    return np.lexsort(keys).astype(np.uint32)


def chrom_offsets(chrom, n_chrom):
    return np.concatenate([[0], np.cumsum(np.bincount(chrom, minlength=n_chrom))]).astype(np.uint64)


def oracle_csr(chrom, start, end, n_chrom, q_chrom, q_start, q_end, perm=None):
    """Expected CSR from the flat oracle for the chain given by perm (default: stable (chrom,start,end))."""
    chrom, start, end = np.asarray(chrom, np.uint32), np.asarray(start, np.uint64), np.asarray(end, np.uint64)
    if perm is None:
        perm = expected_chain(chrom, start, end)
    coff = chrom_offsets(chrom, n_chrom)
    return O.flat_intersect_batch(start[perm], end[perm], coff, q_chrom, q_start, q_end)


def random_index(rng, n, n_chrom=1, span=100_000, max_len=500, dup=False, nested=False, wide=False):
    chrom = rng.integers(0, n_chrom, n).astype(np.uint32)
    start = rng.integers(0, span, n).astype(np.uint64)
    ln = rng.integers(0, max_len, n).astype(np.uint64)
    if nested and n:
        big = rng.random(n) < 0.03
        ln[big] = rng.integers(span // 20, span // 2, int(big.sum())).astype(np.uint64)
    if dup:
        start = (start // 64) * 64
        ln = (ln // 128) * 128
    if wide:
        start = start + np.uint64(2**40)
    return chrom, start, start + ln


def random_queries(rng, q, n_chrom=1, span=100_000, max_len=1500, wide=False, extra_chrom=0):
    qc = rng.integers(0, n_chrom + extra_chrom, q).astype(np.uint32)
    qs = rng.integers(0, span + 1000, q).astype(np.uint64)
    qe = qs + rng.integers(0, max_len, q).astype(np.uint64)
    if wide:
        qs, qe = qs + np.uint64(2**40), qe + np.uint64(2**40)
    return qc, qs, qe


# ----------------------------------------------------------------------------- golden-case replay (CPU oracle side)
import bisect
import json
import os

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_cases.json")


def load_golden():
    with open(GOLDEN) as f:
        return json.load(f)


class ChainReplay:
    """Replays a golden case's ops into per-chromosome leaf chains with the ORACLE's rules (oracle/gg_oracle.c:
    orc_chain_regular / orc_chain_sorted_append / bottom-up build) and answers queries with the flat oracle."""

    def __init__(self, order):
        self.order = order
        self.chains = {}      # chrom -> list of (start, end, id)
        self.leaf_begin = {}  # chrom -> first rank of the rightmost leaf, or None when unknown

    def _append_sorted(self, c, item):
        ch = self.chains.setdefault(c, [])
        lb = self.leaf_begin.get(c, 0)
        lo = lb if lb is not None else 0
        keys = [(s, e) for s, e, _ in ch]
        pos = bisect.bisect_left(keys, (item[0], item[1]), lo)
        ch.insert(pos, tuple(item))
        if lb is not None and len(ch) - lb == self.order:
            self.leaf_begin[c] = len(ch) - 1
        elif c not in self.leaf_begin:
            self.leaf_begin[c] = 0

    def apply(self, op):
        c = op["chrom"]
        if op["op"] == "regular":
            ch = self.chains.setdefault(c, [])
            for it in op["items"]:
                keys = [(s, e) for s, e, _ in ch]
                ch.insert(bisect.bisect_left(keys, (it[0], it[1])), tuple(it))
            self.leaf_begin[c] = None
        elif op["op"] == "sorted":
            for it in op["items"]:
                self._append_sorted(c, it)
        elif op["op"] in ("bulk", "bulk_sorted"):
            items = [tuple(x) for x in op["items"]]
            if op["op"] == "bulk":
                items = sorted(items, key=lambda t: (t[0], t[1]))  # stable; golden bulk cases have distinct keys
            if not self.chains.get(c):
                n, cap = len(items), self.order - 1
                groups = n // cap + (1 if n % cap else 0)
                self.chains[c] = list(items)
                self.leaf_begin[c] = n - n // groups
            else:
                for it in items:
                    self._append_sorted(c, it)
        elif op["op"] == "remove":
            ch = self.chains[c]
            for i in op["ids"]:
                ch[:] = [t for t in ch if t[2] != i]
            self.leaf_begin[c] = None
            if not ch:
                del self.chains[c]

    def arrays(self):
        n_chrom = (max(self.chains) + 1) if self.chains else 1
        chrom, start, end, ids = [], [], [], []
        for c in range(n_chrom):
            for s, e, i in self.chains.get(c, []):
                chrom.append(c), start.append(s), end.append(e), ids.append(i)
        return (n_chrom, np.array(chrom, np.uint32), np.array(start, np.uint64), np.array(end, np.uint64),
                np.array(ids, np.int64))

    def query_ids(self, queries):
        n_chrom, chrom, start, end, ids = self.arrays()
        coff = chrom_offsets(chrom, n_chrom)
        q = np.array(queries, dtype=np.uint64).reshape(-1, 3)
        off, ranks = O.flat_intersect_batch(start, end, coff, q[:, 0].astype(np.uint32), q[:, 1], q[:, 2])
        out = []
        for k in range(len(q)):
            c = int(q[k, 0])
            base = int(coff[c]) if c < n_chrom else 0
            out.append([int(ids[base + int(r)]) for r in ranks[int(off[k]):int(off[k + 1])]])
        return out
