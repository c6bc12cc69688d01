#!/usr/bin/env python
"""Generate tests/golden/reference_cases.json by RUNNING THE UNMODIFIED REFERENCE (oracle/_ref, built from
/root/reference by oracle/Makefile).  Run in the build container only (the GPU box has no reference
checkout); the JSON is committed and is what the CPU and GPU suites compare against.

Each case restates one of the reference's own hot-path tests (file:line under /root/reference/tests) as:
ops (inserts/removes through a named insert path) + queries; the expected id lists come from the
reference's grove<interval,uint32_t>::intersect — order included.  Where the reference test shuffles with
std::mt19937 (not reproducible from Python) a numpy permutation with a fixed seed is used instead.

    python tests/golden/make_golden.py
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle as O  # noqa: E402

NO_CHROM = 0xFFFFFFFF
U64 = 2**64 - 1
cases = []


def run_case(name, source, order, ops, queries):
    """ops: list of dicts {op: regular|sorted|bulk|bulk_sorted|remove, chrom, items:[[s,e,id],...]} (remove: ids)."""
    g = O.RefGrove(order)
    for op in ops:
        if op["op"] in ("regular", "sorted"):
            for s, e, i in op["items"]:
                g.insert(op["chrom"], s, e, i, 1 if op["op"] == "sorted" else 0)
        elif op["op"] in ("bulk", "bulk_sorted"):
            it = np.array(op["items"], dtype=np.uint64).reshape(-1, 3)
            g.insert_bulk(op["chrom"], it[:, 0], it[:, 1], it[:, 2].astype(np.uint32), presorted=op["op"] == "bulk_sorted")
        elif op["op"] == "remove":
            for i in op["ids"]:
                assert g.remove(op["chrom"], i)
    expected = [[int(x) for x in g.intersect(c, s, e)] for c, s, e in queries]
    cases.append({"name": name, "source": source, "order": order, "ops": ops,
                  "queries": [[int(c), int(s), int(e)] for c, s, e in queries], "expected": expected})
    g.close()


def typed_data(count):  # tests/structure/interval_grove_test.cpp:33-43
    return [[i * 100, i * 100 + 50, i] for i in range(count)]


# ---- typed suite: regular / sorted / bulk insert then overlapping + gap queries
#      tests/structure/key_type_grove_test.hpp:386-539 with traits interval_grove_test.cpp:33-86
for count in (1, 2, 10, 100, 1000):
    for order in (3, 5, 10):
        data = typed_data(count)
        if count < 2:
            qs = [(0, data[0][0] + 10, data[0][1] - 10), (0, data[0][1] + 10, data[0][1] + 30)]
        else:
            hi = min(3, count - 1)
            qs = [(0, data[1][0] + 20, data[hi][0] + 20), (0, data[0][1] + 10, data[1][0] - 10)]
        for path in ("regular", "sorted", "bulk", "bulk_sorted"):
            run_case(f"typed_{path}_n{count}_o{order}", "tests/structure/key_type_grove_test.hpp:386-539",
                     order, [{"op": path, "chrom": 0, "items": data}], qs)

# ---- deep tree, shuffled regular inserts, self-queries return exactly one (interval_grove_test.cpp:105-129)
rng = np.random.default_rng(42)
pos = rng.permutation(500)
items = [[int(p) * 100, int(p) * 100 + 50, int(p)] for p in pos]
run_case("deep_tree_stays_globally_sorted", "tests/structure/interval_grove_test.cpp:105-129", 3,
         [{"op": "regular", "chrom": 0, "items": items}], [(0, int(p) * 100, int(p) * 100 + 50) for p in pos])


# ---- routing shapes x orders (tests/structure/routing_test.cpp:52-109,172-182)
def disjoint(n):
    return [[i * 100, i * 100 + 50, i] for i in range(n)]


def equal_starts(n):
    return [[(i // 3) * 100, (i // 3) * 100 + 10 + (i % 3) * 40, i] for i in range(n)]


def nested(n):
    return [[i * 100, i * 100 + (5000 if i % 3 == 0 else 40), i] for i in range(n)]


def duplicates(n):
    out = []
    for i in range(n // 2):
        out += [[i * 100, i * 100 + 50, 2 * i], [i * 100, i * 100 + 50, 2 * i + 1]]
    return out


def touching(n):
    return [[i * 10, i * 10 + 9, i] for i in range(n)]


def one_spanning(n):
    return disjoint(n - 1) + [[0, n * 100, n]]


shapes = {"disjoint": disjoint, "equal_starts": equal_starts, "nested": nested, "duplicates": duplicates,
          "touching": touching, "one_spanning": one_spanning}
for si, (sname, fn) in enumerate(shapes.items()):
    data = fn(300)
    for order in (3, 4, 10, 100):
        sh = np.random.default_rng(si + 1).permutation(len(data))
        shuffled = [data[i] for i in sh]
        qs = [(0, s, e) for s, e, _ in data[::7]] + [(0, 0, 40000), (0, 12345, 12400), (0, 29950, 31000)]
        run_case(f"routing_{sname}_o{order}", "tests/structure/routing_test.cpp:52-182", order,
                 [{"op": "regular", "chrom": 0, "items": shuffled}], qs)
        run_case(f"routing_{sname}_bulk_o{order}", "tests/structure/routing_test.cpp:52-109 (bulk path)", order,
                 [{"op": "bulk", "chrom": 0, "items": shuffled}], qs[:20] if sname != "duplicates" else [])

# ---- preconditions / edge cases (tests/structure/precondition_test.cpp:105-111,192-261)
run_case("missing_index_empty", "tests/structure/precondition_test.cpp:105-111", 10,
         [{"op": "regular", "chrom": 0, "items": [[10, 20, 1]]}], [(5, 10, 20), (0, 10, 20)])
run_case("single_element", "tests/structure/precondition_test.cpp:192-207", 10,
         [{"op": "regular", "chrom": 0, "items": [[10, 20, 42]]}], [(0, 15, 15), (0, 30, 40)])
run_case("duplicate_key", "tests/structure/precondition_test.cpp:209-216", 10,
         [{"op": "regular", "chrom": 0, "items": [[10, 20, 1], [10, 20, 2]]}], [(0, 10, 20)])
run_case("bulk_append_non_empty_cascades", "tests/structure/precondition_test.cpp:218-248", 3,
         [{"op": "sorted", "chrom": 0, "items": [[0, 5, 0]]},
          {"op": "bulk_sorted", "chrom": 0, "items": [[i * 10, i * 10 + 5, i] for i in range(1, 500)]}],
         [(0, 0, 5000), (0, 1234, 1300)])
run_case("keys_equal_to_order", "tests/structure/precondition_test.cpp:250-261", 5,
         [{"op": "regular", "chrom": 0, "items": [[i * 100, i * 100 + 50, i] for i in range(4)]}], [(0, 0, 350)])

# ---- eager grove of grove_view_test (tests/structure/grove_view_test.cpp:67-115)
run_case("grove_view_eager_two_indices", "tests/structure/grove_view_test.cpp:67-115", 4,
         [{"op": "sorted", "chrom": 0, "items": [[i * 10, i * 10 + 5, i] for i in range(60)]},
          {"op": "sorted", "chrom": 1, "items": [[i * 10, i * 10 + 5, 1000 + i] for i in range(40)]}],
         [(c, s, e) for c in (0, 1) for s, e in ((0, 5), (95, 130), (200, 260), (595, 595), (1000, 2000), (5000, 6000))])

# ---- after removal (tests/structure/grove_remove_test.cpp:573-625; insert_intervals = [i*10, i*10+5]... gap 10)
ins = [[i * 20, i * 20 + 10, i] for i in range(10)]
run_case("intersect_after_removal", "tests/structure/grove_remove_test.cpp:573-604", 4,
         [{"op": "regular", "chrom": 0, "items": ins}, {"op": "remove", "chrom": 0, "ids": [2, 5, 8]}],
         [(0, s, e) for s, e, _ in ins] + [(0, 0, 1000)])
run_case("stress_remove_reverse", "tests/structure/grove_remove_test.cpp:627-", 5,
         [{"op": "regular", "chrom": 0, "items": [[i * 20, i * 20 + 10, i] for i in range(50)]},
          {"op": "remove", "chrom": 0, "ids": list(range(49, 9, -1))}], [(0, 0, 2000), (0, 95, 130)])

# ---- hit ORDER pinned by the reference (tests/structure/heterogeneous_grove_test.cpp:98-121,151-177)
run_case("ordered_two_keys", "tests/structure/heterogeneous_grove_test.cpp:98-121", 3,
         [{"op": "regular", "chrom": 0, "items": [[10, 20, 0], [30, 40, 42]]}], [(0, 5, 45)])
run_case("ordered_inserted_descending", "tests/structure/heterogeneous_grove_test.cpp:151-177 (reversed input)", 3,
         [{"op": "regular", "chrom": 0, "items": [[30, 40, 100], [10, 20, 42]]}], [(0, 5, 45)])

# ---- CLI fixture (tests/cli/intersect-test.cpp:86-118; BED end is exclusive -> interval(start, end-1),
#      cli/src/handlers/bed.cpp:20-23).  chr1 = 0, chr2 = 1
run_case("cli_bed_fixture", "tests/cli/data/{target,query}.bed + tests/cli/intersect-test.cpp:86-118", 3,
         [{"op": "regular", "chrom": 0, "items": [[100, 499, 0], [600, 899, 1]]},
          {"op": "regular", "chrom": 1, "items": [[200, 399, 2]]}],
         [(0, 150, 249), (0, 850, 949), (1, 500, 599)])

# ---- order among EQUAL keys per insert path (no reference test pins it; only running the reference does)
eq = [[100, 200, i] for i in range(6)]
for order in (3, 4, 5):
    for path in ("regular", "sorted", "bulk_sorted"):
        run_case(f"equal_keys_{path}_o{order}", "SURVEY §8a row 14 (pinned by running the reference)", order,
                 [{"op": path, "chrom": 0, "items": eq}], [(0, 150, 150)])
mixed = [[(i // 4) * 50, (i // 4) * 50 + 30, i] for i in range(40)]
for order in (3, 4, 7):
    for path in ("regular", "sorted", "bulk_sorted"):
        run_case(f"runs_of_equal_keys_{path}_o{order}", "SURVEY §8a row 14", order,
                 [{"op": path, "chrom": 0, "items": mixed}], [(0, 0, 1000), (0, 120, 130)])

# ---- benchmark data shape (benchmarks/data/generate_intervals.cpp:64-75: [i*10, i*10+5])
bd = [[i * 10, i * 10 + 5, i] for i in range(10000)]
sh = np.random.default_rng(42).permutation(10000)
run_case("bench_10000_unsorted_regular", "benchmarks/data/10000_intervals_unsorted.txt (shape)", 3,
         [{"op": "regular", "chrom": 0, "items": [bd[i] for i in sh]}],
         [(0, 0, 100), (0, 49995, 50010), (0, 99990, 200000), (0, 6, 9)])
run_case("bench_10000_sorted_append", "benchmarks/data/10000_intervals_sorted.txt (shape)", 20,
         [{"op": "sorted", "chrom": 0, "items": bd}], [(0, 0, 100), (0, 49995, 50010), (0, 99990, 200000), (0, 6, 9)])

# ---- interval arithmetic known answers (tests/data_type/interval_test.cpp:121-176,374-451)
L = U64 - 1
arith = []
L_ = O.ref_lib()
for a, b in [((10, 30), (20, 40)), ((10, 20), (20, 30)), ((10, 20), (21, 30)), ((10, 50), (20, 30)), ((10, 30), (10, 30)),
             ((10, 25), (15, 35)), ((15, 35), (20, 40)), ((10, 25), (20, 40)), ((10, 25), (30, 50)), ((15, 15), (10, 20)),
             ((5, 5), (10, 20)), ((10, 10), (10, 20)), ((10, 10), (10, 10)), ((10, 10), (11, 11)), ((L - 100, L), (L - 50, L)),
             ((0, 100), (L - 100, L)), ((0, 0), (0, 0)), ((U64, U64), (U64, U64)), ((0, U64), (5, 5))]:
    arith.append({"a": list(a), "b": list(b), "overlaps": L_.ref_overlaps(*a, *b), "overlaps_rev": L_.ref_overlaps(*b, *a),
                  "less": L_.ref_less(*a, *b), "greater": L_.ref_greater(*a, *b)})

# ---- distribute_evenly (tests/structure/bulk_distribution_test.cpp:50-84)
dist = []
for total, mx in [(1, 4), (4, 4), (5, 4), (7, 2), (9, 2), (100, 3), (1000, 31), (3_000_000_000, 100), (2**31 + 5, 2), (10**12 + 7, 99)]:
    buf = np.zeros(3, np.uint64)
    L_.ref_distribute_evenly(total, mx, buf.ctypes.data_as(O._u64p))
    dist.append({"total": total, "max": mx, "result": [int(x) for x in buf]})

out = {"generator": "tests/golden/make_golden.py", "reference": "genogrove v0.26.1 (unmodified, oracle/_ref)",
       "cases": cases, "interval_arithmetic": arith, "distribute_evenly": dist}
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_cases.json")
with open(path, "w") as f:
    json.dump(out, f, separators=(",", ":"))
print(f"wrote {path}: {len(cases)} cases, {sum(len(c['queries']) for c in cases)} queries, {os.path.getsize(path) / 1e6:.2f} MB")
