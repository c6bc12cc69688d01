/*
 * gg_oracle.c — CPU ORACLE.  TEST INFRASTRUCTURE ONLY: nothing in the product path may import, link
 * or call this file (only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg do, as the
 * checker).  Parity status: PINNED — checked in tests/test_oracle.py against the reference's own
 * known-answer cases (tests/golden/*.json, generated from the unmodified reference through
 * oracle/ref_shim.cpp by tests/golden/make_golden.py) and, when oracle/_ref/libgg_ref.so is present,
 * against the reference itself on random inputs.
 *
 * A plain-C restatement of genogrove's grove<interval,T>::intersect hot path.  All citations are
 * relative to the reference checkout (/root/reference):
 *
 *   interval order / overlap / aggregate      include/genogrove/data_type/interval.hpp:87-105,140-142,155-157
 *   detail::distribute_evenly                 include/genogrove/structure/grove/grove.hpp:100-106
 *   build_tree_bottom_up                      include/genogrove/structure/grove/grove_insert.ipp:545-669
 *   detail::search_overlaps                   include/genogrove/structure/grove/query_engine.hpp:76-134
 *   node::insert_key_ptr (lower_bound)        include/genogrove/structure/grove/node.hpp:247-254
 *   insert_sorted / asymmetric leaf split     grove_insert.ipp:501-531, 302-322 (mid = order - 1)
 *   grove::intersect(q, index)                include/genogrove/structure/grove/grove_query.ipp:33-43
 *
 * The tree here is array-based (no pointers) but keeps the reference's node SHAPES (leaf fill,
 * children per parent, separators = tracked child bounding boxes) and its exact search control flow,
 * so it answers with the same hit list in the same order.  Hits are reported as chain ranks (position
 * of the key in the leaf chain).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* ---------------------------------------------------------------- interval arithmetic */

/* interval.hpp:140-142  max(a.start,b.start) <= min(a.end,b.end); adjacency overlaps */
int orc_overlaps(uint64_t as, uint64_t ae, uint64_t bs, uint64_t be) {
    uint64_t lo = as > bs ? as : bs;
    uint64_t hi = ae < be ? ae : be;
    return lo <= hi;
}
/* interval.hpp:87-92 */
int orc_less(uint64_t as, uint64_t ae, uint64_t bs, uint64_t be) {
    if (as == bs) return ae < be;
    return as < bs;
}
/* interval.hpp:100-105 */
int orc_greater(uint64_t as, uint64_t ae, uint64_t bs, uint64_t be) {
    if (as == bs) return ae > be;
    return as > bs;
}

/* grove.hpp:100-106 — out3 = {num_groups, base_per_group, extra_groups} */
void orc_distribute_evenly(uint64_t total, uint64_t max_per_group, uint64_t* out3) {
    uint64_t groups = total / max_per_group + (total % max_per_group != 0 ? 1 : 0);
    out3[0] = groups;
    out3[1] = total / groups;
    out3[2] = total % groups;
}

/* ---------------------------------------------------------------- B+ tree restatement */

typedef struct {
    size_t n_nodes;
    size_t* first;   /* first child (level>0) or first chain rank (level 0) of node j; size n_nodes+1 */
    uint64_t* rs;    /* tracked full subtree range of node j: min start */
    uint64_t* re;    /*                                       max end   */
} orc_level;

typedef struct orc_tree {
    size_t n;
    int order;
    int n_levels;           /* level 0 = leaves, level n_levels-1 = root (single node) */
    orc_level* lv;
    const uint64_t* start;  /* chain order (borrowed copies) */
    const uint64_t* end;
    uint64_t* own_start;
    uint64_t* own_end;
} orc_tree;

static void level_alloc(orc_level* L, size_t n_nodes) {
    L->n_nodes = n_nodes;
    L->first = (size_t*)malloc((n_nodes + 1) * sizeof(size_t));
    L->rs = (uint64_t*)malloc(n_nodes * sizeof(uint64_t));
    L->re = (uint64_t*)malloc(n_nodes * sizeof(uint64_t));
}

void orc_tree_free(orc_tree* t) {
    if (!t) return;
    for (int l = 0; l < t->n_levels; ++l) {
        free(t->lv[l].first);
        free(t->lv[l].rs);
        free(t->lv[l].re);
    }
    free(t->lv);
    free(t->own_start);
    free(t->own_end);
    free(t);
}

/* grove_insert.ipp:545-669.  Chain order == input order (the caller passes sorted data, as
 * insert_data(idx, vec, sorted, bulk) does on an empty index, grove_insert.ipp:77-101). */
orc_tree* orc_tree_build_bottom_up(const uint64_t* start, const uint64_t* end, size_t n, int order) {
    if (order < 3 || n == 0) return NULL; /* grove.hpp:158-162; empty data -> no root (:553-555) */
    orc_tree* t = (orc_tree*)calloc(1, sizeof(orc_tree));
    t->n = n;
    t->order = order;
    t->own_start = (uint64_t*)malloc(n * sizeof(uint64_t));
    t->own_end = (uint64_t*)malloc(n * sizeof(uint64_t));
    memcpy(t->own_start, start, n * sizeof(uint64_t));
    memcpy(t->own_end, end, n * sizeof(uint64_t));
    t->start = t->own_start;
    t->end = t->own_end;
    t->lv = (orc_level*)calloc(64, sizeof(orc_level));

    /* Step 1 (:563-583): leaves filled evenly, at most order-1 keys each */
    uint64_t d[3];
    orc_distribute_evenly(n, (uint64_t)(order - 1), d);
    orc_level* L = &t->lv[0];
    level_alloc(L, d[0]);
    size_t pos = 0;
    for (size_t j = 0; j < d[0]; ++j) {
        size_t cnt = d[1] + (j < d[2] ? 1 : 0);
        L->first[j] = pos;
        /* leaf range = calc_keys_aggregate (:597-600; node.hpp:290-299) */
        uint64_t mn = t->start[pos], mx = t->end[pos];
        for (size_t k = 1; k < cnt; ++k) {
            if (t->start[pos + k] < mn) mn = t->start[pos + k];
            if (t->end[pos + k] > mx) mx = t->end[pos + k];
        }
        L->rs[j] = mn;
        L->re[j] = mx;
        pos += cnt;
    }
    L->first[d[0]] = pos;
    t->n_levels = 1;

    /* Step 3 (:606-665): parents get at most `order` children, evenly; parent range = fold of children */
    while (t->lv[t->n_levels - 1].n_nodes > 1) {
        orc_level* C = &t->lv[t->n_levels - 1];
        orc_level* P = &t->lv[t->n_levels];
        orc_distribute_evenly(C->n_nodes, (uint64_t)order, d);
        level_alloc(P, d[0]);
        size_t child = 0;
        for (size_t j = 0; j < d[0]; ++j) {
            size_t cnt = d[1] + (j < d[2] ? 1 : 0);
            P->first[j] = child;
            uint64_t mn = C->rs[child], mx = C->re[child];
            for (size_t k = 1; k < cnt; ++k) {
                if (C->rs[child + k] < mn) mn = C->rs[child + k];
                if (C->re[child + k] > mx) mx = C->re[child + k];
            }
            P->rs[j] = mn;
            P->re[j] = mx;
            child += cnt;
        }
        P->first[d[0]] = child;
        t->n_levels++;
    }
    return t;
}

/* query_engine.hpp:76-134.  Writes chain ranks of hits (ascending) to out[0..cap), returns #hits. */
size_t orc_tree_search_overlaps(const orc_tree* t, uint64_t qs, uint64_t qe, uint64_t* out, size_t cap) {
    if (!t) return 0; /* nullptr root -> nothing (:79-81) */
    size_t node = 0;
    /* internal descent (:113-133): separators of a node = ranges of children 0..nc-2 */
    for (int l = t->n_levels - 1; l > 0; --l) {
        const orc_level* P = &t->lv[l];
        const orc_level* C = &t->lv[l - 1];
        size_t cb = P->first[node], nc = P->first[node + 1] - cb;
        size_t nkeys = nc - 1;
        if (nkeys > 0 && qe < C->rs[cb]) return 0; /* early-out (:118-122) */
        size_t i = 0;
        while (i < nkeys && orc_greater(qs, qe, C->rs[cb + i], C->re[cb + i]) &&
               !orc_overlaps(C->rs[cb + i], C->re[cb + i], qs, qe)) {
            i++; /* (:124-128) */
        }
        node = cb + i;
    }
    /* leaf walk (:83-112) */
    const orc_level* L = &t->lv[0];
    size_t hits = 0;
    size_t leaf = node;
    for (;;) {
        for (size_t r = L->first[leaf]; r < L->first[leaf + 1]; ++r) {
            if (orc_overlaps(t->start[r], t->end[r], qs, qe)) {
                if (hits < cap) out[hits] = r;
                hits++;
            }
        }
        if (leaf + 1 >= L->n_nodes) break;                 /* next == nullptr (:94) */
        if (t->start[L->first[leaf + 1]] > qe) break;      /* next.keys[0].start > q.end (:103) */
        leaf++;
    }
    return hits;
}

/* ---------------------------------------------------------------- leaf-chain order per insert path */

static int key_lt(const uint64_t* s, const uint64_t* e, uint32_t a, uint32_t b) {
    return orc_less(s[a], e[a], s[b], e[b]);
}

/* Regular insert (grove_insert.ipp:182-274): routing on subtree max + lower_bound inside the leaf
 * (node.hpp:247-254) is equivalent to inserting at lower_bound of ONE flat sorted vector, so equal keys
 * end up in REVERSE insertion order.  perm[rank] = input index. */
void orc_chain_regular(const uint64_t* s, const uint64_t* e, size_t n, uint32_t* perm) {
    size_t len = 0;
    for (size_t i = 0; i < n; ++i) {
        size_t lo = 0, hi = len;
        while (lo < hi) { /* first element NOT less than key i */
            size_t mid = lo + (hi - lo) / 2;
            if (key_lt(s, e, perm[mid], (uint32_t)i)) lo = mid + 1; else hi = mid;
        }
        memmove(perm + lo + 1, perm + lo, (len - lo) * sizeof(uint32_t));
        perm[lo] = (uint32_t)i;
        len++;
    }
}

/* Sorted single-key append (grove_insert.ipp:501-531): the key goes into the RIGHTMOST LEAF ONLY at its
 * lower_bound there; when the leaf reaches `order` keys it splits at mid = order-1 (:316), i.e. the old
 * leaf keeps order-1 keys and the new rightmost leaf starts with the single last key. */
void orc_chain_sorted_append(const uint64_t* s, const uint64_t* e, size_t n, int order, uint32_t* perm) {
    size_t leaf_begin = 0, len = 0; /* rightmost leaf = perm[leaf_begin, len) */
    for (size_t i = 0; i < n; ++i) {
        size_t lo = leaf_begin, hi = len;
        while (lo < hi) {
            size_t mid = lo + (hi - lo) / 2;
            if (key_lt(s, e, perm[mid], (uint32_t)i)) lo = mid + 1; else hi = mid;
        }
        memmove(perm + lo + 1, perm + lo, (len - lo) * sizeof(uint32_t));
        perm[lo] = (uint32_t)i;
        len++;
        if (len - leaf_begin == (size_t)order) leaf_begin = len - 1;
    }
}

/* Stable merge sort by (start,end): the order the sorted-bulk path sees when the caller sorted stably,
 * and the device sort's default tie rule (input order).  The reference's unsorted-bulk path uses an
 * UNSTABLE std::sort (grove_insert.ipp:167-168): equal keys there are only pinned by running it. */
static void msort(const uint64_t* s, const uint64_t* e, uint32_t* a, uint32_t* tmp, size_t n) {
    if (n < 2) return;
    size_t h = n / 2;
    msort(s, e, a, tmp, h);
    msort(s, e, a + h, tmp, n - h);
    size_t i = 0, j = h, k = 0;
    while (i < h && j < n) tmp[k++] = key_lt(s, e, a[j], a[i]) ? a[j++] : a[i++];
    while (i < h) tmp[k++] = a[i++];
    while (j < n) tmp[k++] = a[j++];
    memcpy(a, tmp, n * sizeof(uint32_t));
}
void orc_chain_stable_sort(const uint64_t* s, const uint64_t* e, size_t n, uint32_t* perm) {
    uint32_t* tmp = (uint32_t*)malloc((n ? n : 1) * sizeof(uint32_t));
    for (size_t i = 0; i < n; ++i) perm[i] = (uint32_t)i;
    msort(s, e, perm, tmp, n);
    free(tmp);
}

/* ---------------------------------------------------------------- flat oracle (batch, multi-chromosome) */

/* The leaf walk of search_overlaps emits exactly { r : start[r] <= q.end && end[r] >= q.start } in
 * ascending chain rank (verified against the reference in tests/test_oracle.py).
 * Chains of all chromosomes are concatenated; chromosome c owns ranks [coff[c], coff[c+1]).
 * pmax (prefix max of end, per chromosome) only bounds the scan from the left; it does not change the
 * answer.  Pass 1: out == NULL, fills counts.  Pass 2: writes per-chromosome ranks at offsets[i].
 * qchrom == 0xFFFFFFFF or >= n_chrom -> zero hits (unknown index -> empty, grove_query.ipp:37-39). */
void orc_flat_intersect_batch(const uint64_t* start, const uint64_t* end, const uint64_t* coff,
                              uint32_t n_chrom, const uint32_t* qchrom, const uint64_t* qs,
                              const uint64_t* qe, size_t nq, uint64_t* counts,
                              const uint64_t* offsets, uint32_t* out) {
    uint64_t n = coff[n_chrom];
    uint64_t* pmax = (uint64_t*)malloc((n ? n : 1) * sizeof(uint64_t));
    for (uint32_t c = 0; c < n_chrom; ++c) {
        uint64_t m = 0;
        for (uint64_t r = coff[c]; r < coff[c + 1]; ++r) {
            if (end[r] > m) m = end[r];
            pmax[r] = m;
        }
    }
    for (size_t i = 0; i < nq; ++i) {
        uint64_t cnt = 0;
        if (qchrom[i] < n_chrom && qs[i] <= qe[i]) {
            uint64_t b = coff[qchrom[i]], t = coff[qchrom[i] + 1];
            uint64_t lo = b, hi = t; /* hi = first rank with start > q.end */
            while (lo < hi) {
                uint64_t mid = lo + (hi - lo) / 2;
                if (start[mid] <= qe[i]) lo = mid + 1; else hi = mid;
            }
            uint64_t ub = lo;
            lo = b; hi = ub; /* first rank whose prefix max end reaches q.start */
            while (lo < hi) {
                uint64_t mid = lo + (hi - lo) / 2;
                if (pmax[mid] < qs[i]) lo = mid + 1; else hi = mid;
            }
            uint32_t* dst = out ? out + offsets[i] : NULL;
            for (uint64_t r = lo; r < ub; ++r) {
                if (end[r] >= qs[i]) {
                    if (dst) dst[cnt] = (uint32_t)(r - b);
                    cnt++;
                }
            }
        }
        if (counts) counts[i] = cnt;
    }
    free(pmax);
}

/* ---------------------------------------------------------------- flanking (nearest predecessor / successor) */

/* detail::flanking_could_descend, interval branch  (query_engine.hpp:176-199).  have_* / best_* = the
 * state's current predecessor end and successor start. */
static int flank_could_descend(uint64_t as, uint64_t ae, uint64_t qs, uint64_t qe, int have_pred,
                               uint64_t pred_end, int have_succ, uint64_t succ_start) {
    int could_pred = as < qs;
    if (could_pred && have_pred) could_pred = ae > pred_end;
    int could_succ = ae > qe;
    if (could_succ && have_succ) could_succ = as < succ_start;
    return could_pred || could_succ;
}

typedef struct {
    int have_pred, have_succ;
    uint64_t pred, succ; /* chain ranks */
} flank_state;

/* detail::search_flanking (query_engine.hpp:238-311) on the array tree; default predicate (always
 * compatible, grove_flanking.ipp).  In the reference children 0..k-1 are pruned with the node's own
 * separator keys and the last child with subtree_range(child) (:296-307): both are the child's full
 * bounding box, which is what rs/re hold for every child here. */
static void flank_rec(const orc_tree* t, int l, size_t node, uint64_t qs, uint64_t qe, flank_state* st) {
    if (l == 0) {
        const orc_level* L = &t->lv[0];
        for (size_t r = L->first[node]; r < L->first[node + 1]; ++r) {
            uint64_t ks = t->start[r], ke = t->end[r];
            if (orc_overlaps(ks, ke, qs, qe)) continue;            /* :254 */
            if (orc_less(ks, ke, qs, qe)) {                        /* :256-272: largest END wins, strictly */
                if (!st->have_pred || ke > t->end[st->pred]) {
                    st->have_pred = 1;
                    st->pred = r;
                }
            } else if (orc_greater(ks, ke, qs, qe)) {              /* :273-281: smallest key wins, strictly */
                if (!st->have_succ || orc_less(ks, ke, t->start[st->succ], t->end[st->succ])) {
                    st->have_succ = 1;
                    st->succ = r;
                }
            }
        }
        return;
    }
    const orc_level* P = &t->lv[l];
    const orc_level* C = &t->lv[l - 1];
    for (size_t c = P->first[node]; c < P->first[node + 1]; ++c) { /* :287-309 */
        if (flank_could_descend(C->rs[c], C->re[c], qs, qe, st->have_pred,
                                st->have_pred ? t->end[st->pred] : 0, st->have_succ,
                                st->have_succ ? t->start[st->succ] : 0))
            flank_rec(t, l - 1, c, qs, qe, st);
    }
}

#define ORC_NO_RANK 0xFFFFFFFFFFFFFFFFull
/* out2[0] = predecessor chain rank, out2[1] = successor chain rank (ORC_NO_RANK when absent). */
void orc_tree_search_flanking(const orc_tree* t, uint64_t qs, uint64_t qe, uint64_t* out2) {
    flank_state st = {0, 0, 0, 0};
    if (t) flank_rec(t, t->n_levels - 1, 0, qs, qe, &st); /* null root -> empty result (grove_flanking.ipp) */
    out2[0] = st.have_pred ? st.pred : ORC_NO_RANK;
    out2[1] = st.have_succ ? st.succ : ORC_NO_RANK;
}

/* Flat statement of the same answer over concatenated chains (what the device computes):
 *   successor   = first chain rank with start > q.end
 *   predecessor = first chain rank among those whose end is the largest end < q.start
 * pred/succ receive per-chromosome ranks, 0xFFFFFFFF when absent (also for an unknown chromosome). */
typedef struct { uint64_t end; uint64_t rank; } end_rank;
static int cmp_end_rank(const void* a, const void* b) {
    const end_rank* x = (const end_rank*)a;
    const end_rank* y = (const end_rank*)b;
    if (x->end != y->end) return x->end < y->end ? -1 : 1;
    return x->rank < y->rank ? -1 : (x->rank > y->rank ? 1 : 0);
}
void orc_flat_flanking_batch(const uint64_t* start, const uint64_t* end, const uint64_t* coff,
                             uint32_t n_chrom, const uint32_t* qchrom, const uint64_t* qs,
                             const uint64_t* qe, size_t nq, uint32_t* pred, uint32_t* succ) {
    uint64_t n = coff[n_chrom];
    end_rank* er = (end_rank*)malloc((n ? n : 1) * sizeof(end_rank));
    for (uint32_t c = 0; c < n_chrom; ++c) {
        for (uint64_t r = coff[c]; r < coff[c + 1]; ++r) {
            er[r].end = end[r];
            er[r].rank = r - coff[c];
        }
        qsort(er + coff[c], coff[c + 1] - coff[c], sizeof(end_rank), cmp_end_rank);
    }
    for (size_t i = 0; i < nq; ++i) {
        pred[i] = succ[i] = 0xFFFFFFFFu;
        if (qchrom[i] >= n_chrom || qs[i] > qe[i]) continue;
        uint64_t b = coff[qchrom[i]], t = coff[qchrom[i] + 1];
        uint64_t lo = b, hi = t;
        while (lo < hi) { /* first rank with start > q.end */
            uint64_t mid = lo + (hi - lo) / 2;
            if (start[mid] <= qe[i]) lo = mid + 1; else hi = mid;
        }
        if (lo < t) succ[i] = (uint32_t)(lo - b);
        lo = b; hi = t;
        while (lo < hi) { /* number of ends < q.start */
            uint64_t mid = lo + (hi - lo) / 2;
            if (er[mid].end < qs[i]) lo = mid + 1; else hi = mid;
        }
        if (lo > b) {
            uint64_t v = er[lo - 1].end, l2 = b, h2 = lo;
            while (l2 < h2) { /* first (end == v): smallest chain rank with that end */
                uint64_t mid = l2 + (h2 - l2) / 2;
                if (er[mid].end < v) l2 = mid + 1; else h2 = mid;
            }
            pred[i] = (uint32_t)er[l2].rank;
        }
    }
    free(er);
}

/* ---------------------------------------------------------------- genomic_coordinate (strand-aware keys) */

/* strand order '*' < '.' < '+' < '-' (genomic_coordinate.hpp:97-110); -1 for an invalid strand byte (:78-81) */
int orc_strand_order(int s) {
    switch (s) {
        case '*': return 0;
        case '.': return 1;
        case '+': return 2;
        case '-': return 3;
        default: return -1;
    }
}
/* genomic_coordinate::operator< (:97-110) */
int orc_gc_less(int sa, uint64_t as, uint64_t ae, int sb, uint64_t bs, uint64_t be) {
    if (as != bs) return as < bs;
    if (ae != be) return ae < be;
    return orc_strand_order(sa) < orc_strand_order(sb);
}
/* genomic_coordinate::overlaps (:156-160) */
int orc_gc_overlaps(int sa, uint64_t as, uint64_t ae, int sb, uint64_t bs, uint64_t be) {
    if (as > be || bs > ae) return 0;
    if (sa == '*' || sb == '*') return 1;
    return sa == sb;
}

/* Flat statement of grove<genomic_coordinate,T>::intersect over concatenated chains (chain order = (start, end,
 * strand order), ties by insert path): search_overlaps (query_engine.hpp:76-134) skips a child only when it lies
 * before the query or its aggregate strand cannot match, then walks leaves to the first one starting past q.end,
 * so the hits are the spatial hits filtered by strand compatibility, ascending in chain rank.
 * Pass 1: out == NULL, fills counts.  Pass 2: writes per-chromosome ranks at offsets[i]. */
void orc_flat_intersect_batch_stranded(const uint64_t* start, const uint64_t* end, const uint8_t* strand,
                                       const uint64_t* coff, uint32_t n_chrom, const uint32_t* qchrom,
                                       const uint64_t* qs, const uint64_t* qe, const uint8_t* qstrand, size_t nq,
                                       uint64_t* counts, const uint64_t* offsets, uint32_t* out) {
    /* pmax (prefix max of end) only bounds the scan from the left, as in orc_flat_intersect_batch */
    uint64_t n = coff[n_chrom];
    uint64_t* pmax = (uint64_t*)malloc((n ? n : 1) * sizeof(uint64_t));
    for (uint32_t c = 0; c < n_chrom; ++c) {
        uint64_t m = 0;
        for (uint64_t r = coff[c]; r < coff[c + 1]; ++r) {
            if (end[r] > m) m = end[r];
            pmax[r] = m;
        }
    }
    for (size_t i = 0; i < nq; ++i) {
        uint64_t cnt = 0;
        if (qchrom[i] < n_chrom && qs[i] <= qe[i] && orc_strand_order(qstrand[i]) >= 0) {
            uint64_t b = coff[qchrom[i]], t = coff[qchrom[i] + 1];
            uint64_t lo = b, hi = t;
            while (lo < hi) { /* first rank with start > q.end */
                uint64_t mid = lo + (hi - lo) / 2;
                if (start[mid] <= qe[i]) lo = mid + 1; else hi = mid;
            }
            uint64_t ub = lo;
            lo = b; hi = ub;
            while (lo < hi) { /* first rank whose prefix max end reaches q.start */
                uint64_t mid = lo + (hi - lo) / 2;
                if (pmax[mid] < qs[i]) lo = mid + 1; else hi = mid;
            }
            uint32_t* dst = out ? out + offsets[i] : NULL;
            for (uint64_t r = lo; r < ub; ++r) {
                if (orc_gc_overlaps(strand[r], start[r], end[r], qstrand[i], qs[i], qe[i])) {
                    if (dst) dst[cnt] = (uint32_t)(r - b);
                    cnt++;
                }
            }
        }
        if (counts) counts[i] = cnt;
    }
    free(pmax);
}
