#!/usr/bin/env python
"""Generate tests/golden/flanking_strand_cases.json by RUNNING THE UNMODIFIED REFERENCE's
grove<genomic_coordinate, uint32_t>::flanking(query, index, pred) (oracle/_ref, built from /root/reference by
oracle/Makefile) with the strand-compatible predicate of the reference's own documentation
(include/genogrove/structure/grove/grove_flanking.ipp:47-54).  Run in the build container only; the JSON is committed.

Keys are distinct in (start, end, strand), so payload ids identify chain ranks whatever the insert path; many keys share
(start, end) across strands, share end values, and nest.  Every case is inserted three ways (regular inserts in random
order, sorted single appends, sorted bulk) and at several tree orders: the answers must agree — that is the claim that
this predicate's result does not depend on the tree's shape.

"default_predicate_shape_dependence" records the opposite for the DEFAULT predicate on genomic_coordinate keys: pairs of
groves holding the same keys (regular inserts vs sorted bulk) whose flanking answers differ.

    python tests/golden/make_golden_flanking_strand.py
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle as O  # noqa: E402

STR = "*.+-"
NONE = 0xFFFFFFFF


def make_keys(rng, n, span, step, dense):
    keys = set()
    while len(keys) < n:
        s = int(rng.integers(0, span // step)) * step
        ln = int(rng.integers(0, 8)) * step
        if rng.random() < 0.08:
            ln += int(rng.integers(span // 8, span // 2))
        if dense and rng.random() < 0.3:
            ln = 3 * step
        keys.add((STR[int(rng.integers(0, 4))], s, s + ln))
    keys = sorted(keys, key=lambda k: (k[1], k[2], O.STRAND_ORDER[k[0]]))
    return [[st, s, e, i] for i, (st, s, e) in enumerate(keys)]  # id = chain rank


def build(order, path, items, rng):
    g = O.RefGcGrove(order)
    if path == "regular":
        for j in rng.permutation(len(items)):
            st, s, e, i = items[j]
            g.insert(0, st, s, e, i)
    elif path == "sorted":
        for st, s, e, i in items:
            g.insert(0, st, s, e, i, sorted_=True)
    else:
        g.insert_bulk(0, "".join(t[0] for t in items), [t[1] for t in items], [t[2] for t in items], [t[3] for t in items])
    return g


def main():
    rng = np.random.default_rng(2026)
    cases, shape = [], []
    for ci, (n, span, step, dense) in enumerate([(1, 100, 5, False), (2, 100, 5, False), (7, 200, 5, True), (40, 400, 5, True),
                                                 (150, 3000, 10, False), (400, 2000, 4, True), (900, 20000, 8, False)]):
        items = make_keys(rng, n, span, step, dense)
        queries = []
        for _ in range(140):
            qs = int(rng.integers(0, span + 50))
            queries.append([0 if rng.random() > 0.05 else 1, STR[int(rng.integers(0, 4))], qs, qs + int(rng.integers(0, span // 6 + 1))])
        queries.append([0, "+", 0, 0])
        queries.append([0, "*", span * 2, span * 2 + 5])
        answers = None
        for order in (3, 4, 8, 32):
            for path in ("regular", "sorted", "bulk"):
                g = build(order, path, items, rng)
                got = [list(g.flanking(c, st, s, e, strand_predicate=True)) for c, st, s, e in queries]
                if answers is None:
                    answers = got
                assert got == answers, ("the strand predicate's answer depends on the tree shape", ci, order, path)
                if path == "regular" and len(shape) < 12 and n >= 40:
                    g2 = build(order, "bulk", items, rng)
                    for c, st, s, e in queries:
                        a, b = g.flanking(c, st, s, e), g2.flanking(c, st, s, e)
                        if a != b and len(shape) < 12:
                            shape.append({"order": order, "items": items, "query": [c, st, s, e],
                                          "regular_inserts": [int(x) for x in a], "sorted_bulk": [int(x) for x in b]})
                            break
                    g2.close()
                g.close()
        cases.append({"name": f"gc_flank_{ci}_n{n}", "items": items, "queries": queries,
                      "pred": [None if a[0] == NONE else int(a[0]) for a in answers],
                      "succ": [None if a[1] == NONE else int(a[1]) for a in answers]})
    assert shape, "expected at least one shape-dependent answer of the default predicate"
    out = {"generator": "tests/golden/make_golden_flanking_strand.py",
           "reference": "grove<genomic_coordinate,uint32_t>::flanking(q, idx, strand-compatible predicate); every case agreed over "
                        "orders 3/4/8/32 x regular / sorted-append / sorted-bulk inserts",
           "cases": cases, "default_predicate_shape_dependence": shape}
    path = os.path.join(ROOT, "tests", "golden", "flanking_strand_cases.json")
    with open(path, "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print(path, len(cases), "cases,", sum(len(c["queries"]) for c in cases), "queries,", len(shape), "shape-dependence records,",
          os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
