#!/usr/bin/env python
"""Generate tests/golden/strand_cases.json by RUNNING THE UNMODIFIED REFERENCE's
grove<genomic_coordinate, uint32_t>::intersect (oracle/_ref, built from /root/reference by oracle/Makefile).
Run in the build container only; the JSON is committed and is what the CPU and GPU suites compare against.

Cases restate /root/reference/tests/structure/genomic_coordinate_grove_test.cpp:150-266 (strand-specific, wildcard and
bulk strand-order tests) plus seeded random groves (regular inserts and sorted bulk) with duplicated keys.
items: [strand, start, end, id]; queries: [chrom, strand, start, end]; expected: id lists in the reference's order.

    python tests/golden/make_golden_strand.py
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle as O  # noqa: E402

cases = []


def run_case(name, source, order, path, chrom, items, queries):
    g = O.RefGcGrove(order)
    if path == "regular":
        for st, s, e, i in items:
            g.insert(chrom, st, s, e, i)
    else:  # sorted bulk: items must already be in (start, end, strand) order
        g.insert_bulk(chrom, "".join(t[0] for t in items), [t[1] for t in items], [t[2] for t in items], [t[3] for t in items])
    expected = [[int(x) for x in g.intersect(c, st, s, e)] for c, st, s, e in queries]
    cases.append({"name": name, "source": source, "order": order, "path": path, "chrom": chrom,
                  "items": [[st, int(s), int(e), int(i)] for st, s, e, i in items],
                  "queries": [[int(c), st, int(s), int(e)] for c, st, s, e in queries], "expected": expected})
    g.close()


T = "tests/structure/genomic_coordinate_grove_test.cpp"
run_case("unstranded_query_strict", T + ":150-177", 3, "regular", 0,
         [["+", 10, 30, 10], ["-", 20, 40, 20], [".", 30, 50, 30]], [(0, ".", 25, 45), (0, "+", 25, 45), (0, "-", 25, 45), (0, "*", 25, 45)])
run_case("stranded_query_matching_strand", T + ":179-218", 3, "regular", 0,
         [["+", 10, 30, 10], ["+", 40, 60, 70], ["+", 80, 100, 90], ["-", 10, 30, 20], ["-", 50, 90, 50], ["-", 100, 120, 70],
          [".", 10, 30, 30], [".", 25, 85, 40]],
         [(0, "+", 15, 85), (0, "-", 15, 55), (0, ".", 15, 55), (0, "*", 0, 200), (1, "+", 15, 85)])
run_case("wildcard_query_all_strands", T + ":220-244", 3, "regular", 0,
         [["+", 10, 30, 10], ["+", 40, 60, 40], ["-", 20, 50, 20], ["-", 55, 75, 55], [".", 15, 45, 15]], [(0, "*", 25, 65)])
run_case("bulk_strand_sorting", T + ":246-266", 5, "bulk_sorted", 0,
         [["*", 10, 20, 10], [".", 10, 20, 20], ["+", 10, 20, 30], ["-", 10, 20, 40]],
         [(0, "*", 10, 20), (0, "+", 10, 20), (0, "-", 0, 9), (0, ".", 20, 30)])

rng = np.random.default_rng(20260412)
SYM = "+-.*"
for k in range(10):
    n = int(rng.integers(30, 400))
    span = int(rng.integers(300, 6000))
    order = int(rng.choice([3, 4, 9, 32]))
    s = rng.integers(0, span, n)
    ln = rng.integers(0, max(2, span // 8), n)
    if k % 2 == 0:  # duplicated (start, end) with different and equal strands
        s = (s // 30) * 30
        ln = (ln // 60) * 60
    if k % 5 == 1:
        big = rng.random(n) < 0.04
        ln[big] += span // 2
    st = [SYM[int(x)] for x in rng.integers(0, 4 if k % 3 else 3, n)]
    items = [[st[i], int(s[i]), int(s[i] + ln[i]), i] for i in range(n)]
    path = "regular" if k % 2 == 0 else "bulk_sorted"
    if path == "bulk_sorted":
        items.sort(key=lambda t: (t[1], t[2], O.STRAND_ORDER[t[0]]))
    nq = 120
    qs = rng.integers(0, span + 100, nq)
    qe = qs + rng.integers(0, max(2, span // 10), nq)
    chrom = k % 3
    queries = [(chrom, SYM[int(rng.integers(0, 4))], int(a), int(b)) for a, b in zip(qs, qe)]
    queries += [(chrom, "*", 0, span * 3), (chrom, "+", 0, 0), (chrom + 1, "*", 0, span)]
    run_case(f"random_{k}_{path}", "seeded random", order, path, chrom, items, queries)

out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "strand_cases.json")
with open(out, "w") as f:
    json.dump({"generator": "tests/golden/make_golden_strand.py", "cases": cases}, f, separators=(",", ":"))
print(f"wrote {len(cases)} cases ({sum(len(c['queries']) for c in cases)} queries) to {out}")
