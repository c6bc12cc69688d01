"""Synthetic workloads of BASELINE.json / SURVEY §8(d): hg38-shaped interval sets and query batches.

All generators are deterministic (numpy PCG64, fixed seeds), return plain numpy arrays
(chrom u32, start u64, end u64; intervals are closed [start, end]) and never touch the reference checkout.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

# hg38 primary assembly lengths, chr1..chr22, chrX, chrY
HG38_LENGTHS = np.array([
    248956422, 242193529, 198295559, 190214555, 181538259, 170805979, 159345973, 145138636,
    138394717, 133797422, 135086622, 133275309, 114364328, 107043718, 101991189, 90338345,
    83257441, 80373285, 58617616, 64444167, 46709983, 50818468, 156040895, 57227415], dtype=np.int64)


@dataclass
class Workload:
    name: str
    n_chrom: int
    chrom: np.ndarray      # u32[N]
    start: np.ndarray      # u64[N]
    end: np.ndarray        # u64[N]
    q_chrom: np.ndarray    # u32[Q]
    q_start: np.ndarray    # u64[Q]
    q_end: np.ndarray      # u64[Q]


def _split_by_length(total: int, lengths: np.ndarray) -> np.ndarray:
    """Apportion `total` items to chromosomes proportionally to length (largest remainder)."""
    w = lengths / lengths.sum()
    base = np.floor(w * total).astype(np.int64)
    rem = total - int(base.sum())
    order = np.argsort(-(w * total - base), kind="stable")
    base[order[:rem]] += 1
    return base


def hg38_intervals(n: int, seed: int = 42, min_len: int = 100, max_len: int = 10000, n_chrom: int = 24):
    """cfg2/4/5 recipe: n intervals spread over chromosomes proportionally to length, uniform starts,
    log-uniform lengths in [min_len, max_len].  Unsorted (input order is random)."""
    rng = np.random.default_rng(seed)
    lengths = HG38_LENGTHS[:n_chrom]
    per = _split_by_length(n, lengths)
    chrom = np.repeat(np.arange(n_chrom, dtype=np.uint32), per)
    L = np.repeat(lengths, per)
    ln = np.exp(rng.uniform(np.log(min_len), np.log(max_len), n)).astype(np.int64)
    start = (rng.random(n) * (L - ln - 1)).astype(np.int64)
    end = start + ln - 1
    shuffle = rng.permutation(n)
    return chrom[shuffle], start[shuffle].astype(np.uint64), end[shuffle].astype(np.uint64)


def hg38_intervals_part(n_total: int, part: int, parts: int, seed: int = 42, min_len: int = 100, max_len: int = 10000,
                        n_chrom: int = 24):
    """Part `part` of `parts` of an unsorted n_total-interval input (cfg5: every rank reads its slice of the file): the same
    recipe as hg38_intervals, drawn record by record — chromosome proportional to length, uniform start, log-uniform length
    — from a generator seeded per part, so the parts are independent, reproducible and any chromosome occurs in any part."""
    lo, hi = part * n_total // parts, (part + 1) * n_total // parts
    n = hi - lo
    rng = np.random.default_rng([seed, part, parts])
    lengths = HG38_LENGTHS[:n_chrom]
    cdf = np.cumsum(lengths / lengths.sum())
    chrom = np.minimum(np.searchsorted(cdf, rng.random(n), side="right"), n_chrom - 1).astype(np.uint32)
    ln = np.exp(rng.uniform(np.log(min_len), np.log(max_len), n)).astype(np.int64)
    start = (rng.random(n) * (lengths[chrom] - ln - 1)).astype(np.int64)
    return chrom, start.astype(np.uint64), (start + ln - 1).astype(np.uint64)


def hg38_queries(q: int, length: int, seed: int = 43, sorted_by_coord: bool = True, n_chrom: int = 24):
    """q queries of fixed length, chromosome proportional to length, uniform start.  sorted_by_coord:
    coordinate-sorted inside each chromosome, chromosomes in id order (BAM-like); else shuffled."""
    rng = np.random.default_rng(seed)
    lengths = HG38_LENGTHS[:n_chrom]
    per = _split_by_length(q, lengths)
    chrom = np.repeat(np.arange(n_chrom, dtype=np.uint32), per)
    start = np.empty(q, dtype=np.uint64)
    p = 0
    for c in range(n_chrom):
        m = int(per[c])
        s = rng.integers(0, int(lengths[c]) - length, m, dtype=np.int64)
        if sorted_by_coord:
            s.sort()
        start[p:p + m] = s
        p += m
    end = start + np.uint64(length - 1)
    if not sorted_by_coord:
        sh = rng.permutation(q)
        chrom, start, end = chrom[sh], start[sh], end[sh]
    return chrom, start, end


def hg38_sorted_slice(q_total: int, lo: int, hi: int, length: int, seed: int = 43, n_chrom: int = 24):
    """Records [lo, hi) of ONE coordinate-sorted batch of q_total fixed-length queries (chromosomes in id order, queries
    apportioned to chromosomes by length, uniform starts) — what a reader of a sorted BAM sees in a contiguous part of the
    file, without generating the rest: inside chromosome c the records [a, b) of its m_c sorted uniform starts are drawn as
    b - a sorted uniform starts on the sub-range [a / m_c, b / m_c) of the chromosome (the order statistics' expected
    span).  Deterministic in (seed, chromosome, a)."""
    lengths = HG38_LENGTHS[:n_chrom]
    per = _split_by_length(q_total, lengths)
    first = np.concatenate([[0], np.cumsum(per)])
    chrom, start = [], []
    for c in range(n_chrom):
        a, b = max(lo, int(first[c])) - int(first[c]), min(hi, int(first[c + 1])) - int(first[c])
        if b <= a:
            continue
        m, span = int(per[c]), int(lengths[c]) - length
        p0, p1 = a * span // m, max(b * span // m, a * span // m + 1)
        rng = np.random.default_rng([seed, c, a])
        s = rng.integers(p0, p1, b - a, dtype=np.int64)
        s.sort()
        chrom.append(np.full(b - a, c, np.uint32))
        start.append(s.astype(np.uint64))
    if not chrom:
        z = np.zeros(0, np.uint64)
        return np.zeros(0, np.uint32), z, z
    chrom, start = np.concatenate(chrom), np.concatenate(start)
    return chrom, start, start + np.uint64(length - 1)


def cfg1(n: int = 100_000, q: int = 10_000, seed: int = 42) -> Workload:
    """Single chromosome of 248,956,422 bp; lengths U[200,5200); 1 kb queries (reference runs as-is)."""
    rng = np.random.default_rng(seed)
    L = int(HG38_LENGTHS[0])
    start = rng.integers(0, L, n, dtype=np.int64)
    end = start + rng.integers(200, 5200, n, dtype=np.int64) - 1
    qs = rng.integers(0, L, q, dtype=np.int64)
    qe = qs + 999
    z = np.zeros
    return Workload("cfg1_100k_x_10k", 1, z(n, np.uint32), start.astype(np.uint64), end.astype(np.uint64),
                    z(q, np.uint32), qs.astype(np.uint64), qe.astype(np.uint64))


def cfg2(n: int = 1_000_000, q: int = 10_000_000, seed: int = 42, sorted_queries: bool = False) -> Workload:
    c, s, e = hg38_intervals(n, seed)
    qc, qs, qe = hg38_queries(q, 1000, seed + 1, sorted_queries)
    return Workload("cfg2_1M_x_10M_1kb", 24, c, s, e, qc, qs, qe)


def cfg3(n_genes: int = 60_000, q: int = 50_000_000, seed: int = 42) -> Workload:
    """GENCODE-shaped nested features: genes (log-normal length, median ~25 kb, tail to ~2.4 Mb), each with
    1-20 transcripts nested inside, each with 2-30 exons of 50-500 bp; exons are reused between
    transcripts so many features share identical (start, end).  Queries: 60 % inside genes, 40 % uniform,
    lengths 1-1000."""
    rng = np.random.default_rng(seed)
    lengths = HG38_LENGTHS
    per = _split_by_length(n_genes, lengths)
    g_chrom = np.repeat(np.arange(24, dtype=np.uint32), per)
    gL = np.repeat(lengths, per)
    g_len = np.minimum(np.exp(rng.normal(np.log(25_000), 1.3, n_genes)), 2_400_000).astype(np.int64) + 500
    g_start = (rng.random(n_genes) * (gL - g_len - 1)).astype(np.int64)
    n_tx = rng.integers(1, 21, n_genes)
    # exon pool per gene: up to 30 exons at random offsets inside the gene
    n_pool = rng.integers(2, 31, n_genes)
    feats_c, feats_s, feats_e = [g_chrom], [g_start], [g_start + g_len - 1]
    pool_gene = np.repeat(np.arange(n_genes), n_pool)
    pool_len = rng.integers(50, 501, len(pool_gene))
    pool_off = (rng.random(len(pool_gene)) * np.maximum(g_len[pool_gene] - pool_len, 1)).astype(np.int64)
    pool_s = g_start[pool_gene] + pool_off
    pool_e = np.minimum(pool_s + pool_len - 1, g_start[pool_gene] + g_len[pool_gene] - 1)
    pool_first = np.concatenate([[0], np.cumsum(n_pool)[:-1]])
    # transcripts: span of a random subset of the pool; exons = that subset (duplicates across transcripts)
    tx_gene = np.repeat(np.arange(n_genes), n_tx)
    n_t = len(tx_gene)
    take = np.minimum(rng.integers(2, 31, n_t), n_pool[tx_gene])
    ex_tx = np.repeat(np.arange(n_t), take)
    ex_pick = pool_first[tx_gene[ex_tx]] + (rng.random(len(ex_tx)) * n_pool[tx_gene[ex_tx]]).astype(np.int64)
    ex_s, ex_e = pool_s[ex_pick], pool_e[ex_pick]
    tx_s = np.full(n_t, np.iinfo(np.int64).max)
    tx_e = np.zeros(n_t, np.int64)
    np.minimum.at(tx_s, ex_tx, ex_s)
    np.maximum.at(tx_e, ex_tx, ex_e)
    feats_c += [g_chrom[tx_gene], g_chrom[tx_gene[ex_tx]]]
    feats_s += [tx_s, ex_s]
    feats_e += [tx_e, ex_e]
    chrom = np.concatenate(feats_c).astype(np.uint32)
    start = np.concatenate(feats_s).astype(np.uint64)
    end = np.concatenate(feats_e).astype(np.uint64)
    sh = rng.permutation(len(chrom))
    chrom, start, end = chrom[sh], start[sh], end[sh]
    # queries
    q_in = int(q * 0.6)
    gi = rng.integers(0, n_genes, q_in)
    qlen = rng.integers(1, 1001, q)
    qs_in = g_start[gi] + (rng.random(q_in) * g_len[gi]).astype(np.int64)
    uc, us, _ = hg38_queries(q - q_in, 1001, seed + 7, sorted_by_coord=False)
    qc = np.concatenate([g_chrom[gi], uc]).astype(np.uint32)
    qs = np.concatenate([qs_in, us.astype(np.int64)])
    qe = qs + qlen - 1
    sh = rng.permutation(q)
    return Workload("cfg3_gencode_nested", 24, chrom, start, end, qc[sh], qs[sh].astype(np.uint64), qe[sh].astype(np.uint64))


def cfg4(n: int = 10_000_000, q: int = 100_000_000, seed: int = 42, sorted_queries: bool = True) -> Workload:
    """10M-interval grove x 100M read-shaped (150 bp) queries — the configuration the metric is quoted on."""
    c, s, e = hg38_intervals(n, seed)
    qc, qs, qe = hg38_queries(q, 150, seed + 1, sorted_queries)
    return Workload("cfg4_10M_x_100M_150bp" + ("_sorted" if sorted_queries else "_shuffled"), 24, c, s, e, qc, qs, qe)


def lpt_assign(counts, n_bins: int):
    """Longest-processing-time bin packing of chromosomes onto GPUs by interval count (cfg5).
    Returns owner[c] in [0, n_bins)."""
    counts = np.asarray(counts, dtype=np.int64)
    owner = np.zeros(len(counts), dtype=np.int64)
    load = np.zeros(n_bins, dtype=np.int64)
    for c in np.argsort(-counts, kind="stable"):
        b = int(np.argmin(load))
        owner[c] = b
        load[b] += counts[c]
    return owner
