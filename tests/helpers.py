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
