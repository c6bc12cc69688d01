#!/usr/bin/env python
"""Generate tests/golden/flanking_cases.json by RUNNING THE UNMODIFIED REFERENCE's grove::flanking
(oracle/_ref, built from /root/reference by oracle/Makefile).  Run in the build container only; the JSON is
committed and is what the CPU and GPU suites compare against.

Cases restate /root/reference/tests/structure/grove_flanking_test.cpp:20-117 (interval keys) plus seeded random
groves through three insert paths (regular, sorted append, sorted bulk) with duplicated and nested keys, which pin
the tie rule ("first key found in chain order") that no reference test pins.

    python tests/golden/make_golden_flanking.py
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle as O  # noqa: E402

NONE = 0xFFFFFFFF
cases = []


def run_case(name, source, order, ops, queries):
    """ops as in make_golden.py: {op: regular|sorted|bulk_sorted, chrom, items: [[s, e, id], ...]}."""
    g = O.RefGrove(order)
    for op in ops:
        if op["op"] in ("regular", "sorted"):
            for s, e, i in op["items"]:
                g.insert(op["chrom"], s, e, i, 1 if op["op"] == "sorted" else 0)
        else:
            it = np.array(op["items"], dtype=np.uint64).reshape(-1, 3)
            g.insert_bulk(op["chrom"], it[:, 0], it[:, 1], it[:, 2].astype(np.uint32), presorted=True)
    q = np.array(queries, dtype=np.uint64).reshape(-1, 3)
    pred, succ = g.flanking_batch(q[:, 0].astype(np.uint32), q[:, 1], q[:, 2])
    cases.append({"name": name, "source": source, "order": order, "ops": ops,
                  "queries": [[int(c), int(s), int(e)] for c, s, e in queries],
                  "pred": [None if int(p) == NONE else int(p) for p in pred],
                  "succ": [None if int(p) == NONE else int(p) for p in succ]})
    g.close()


T = "tests/structure/grove_flanking_test.cpp"
run_case("empty_grove", T + ":20-25", 8, [], [(0, 100, 200)])
run_case("missing_index", T + ":27-33", 8, [{"op": "sorted", "chrom": 0, "items": [[50, 60, 1]]}], [(1, 100, 200), (NONE, 100, 200)])
run_case("only_successor", T + ":35-41", 8, [{"op": "sorted", "chrom": 0, "items": [[500, 600, 1]]}], [(0, 100, 200)])
run_case("only_predecessor", T + ":43-49", 8, [{"op": "sorted", "chrom": 0, "items": [[100, 200, 1]]}], [(0, 500, 600)])
run_case("bracketed", T + ":51-58", 8, [{"op": "sorted", "chrom": 0, "items": [[100, 200, 1], [500, 600, 2]]}], [(0, 300, 400)])
run_case("overlapping_skipped", T + ":60-68", 8,
         [{"op": "sorted", "chrom": 0, "items": [[50, 60, 2], [100, 350, 1], [500, 600, 3]]}], [(0, 300, 400)])
run_case("abutting", T + ":70-80", 8, [{"op": "sorted", "chrom": 0, "items": [[50, 99, 1], [201, 300, 2]]}], [(0, 100, 200)])
run_case("closest_pred_among_many", T + ":82-90", 8,
         [{"op": "sorted", "chrom": 0, "items": [[0, 10, 1], [20, 30, 2], [40, 50, 3]]}], [(0, 100, 200)])
run_case("nested_pick_largest_end", T + ":92-102", 8,
         [{"op": "sorted", "chrom": 0, "items": [[50, 100, 1], [80, 90, 2]]}], [(0, 200, 300)])
run_case("multi_leaf", T + ":104-117", 4,
         [{"op": "sorted", "chrom": 0, "items": [[i * 100, i * 100 + 20, i] for i in range(16)]}],
         [(0, 440, 460), (0, 0, 5), (0, 1600, 1700), (0, 21, 99), (0, 0, 1520)])

rng = np.random.default_rng(20260411)
for k in range(12):
    n = int(rng.integers(20, 300))
    span = int(rng.integers(200, 5000))
    order = int(rng.choice([3, 4, 8, 32]))
    s = rng.integers(0, span, n)
    ln = rng.integers(0, max(2, span // 10), n)
    if k % 3 == 0:  # heavy duplication of (start, end) and of end values
        s = (s // 25) * 25
        ln = (ln // 40) * 40
    if k % 4 == 1:  # a few long intervals: nesting
        big = rng.random(n) < 0.05
        ln[big] += span // 2
    e = s + ln
    path = ("regular", "sorted", "bulk_sorted")[k % 3]
    items = [[int(s[i]), int(e[i]), i] for i in range(n)]
    if path != "regular":
        items.sort(key=lambda t: (t[0], t[1]))
    nq = 80
    qs = rng.integers(0, span + 100, nq)
    qe = qs + rng.integers(0, max(2, span // 20), nq)
    chrom = int(k % 2)
    queries = [(chrom, int(a), int(b)) for a, b in zip(qs, qe)] + [(chrom, 0, 0), (chrom, 0, span * 3), (1 - chrom, 5, 6)]
    run_case(f"random_{k}_{path}", "seeded random (tie rule: first key found in chain order)", order,
             [{"op": path, "chrom": chrom, "items": items}], queries)

out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "flanking_cases.json")
with open(out, "w") as f:
    json.dump({"generator": "tests/golden/make_golden_flanking.py", "cases": cases}, f, separators=(",", ":"))
print(f"wrote {len(cases)} cases to {out}")
