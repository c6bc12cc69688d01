// grove_tests.cpp — host-side C++20 tests of the drop-in facade (include/genogrove_b200/grove.hpp).
//
// The scenarios restate the reference's own hot-path tests (tests/structure/key_type_grove_test.hpp:386-539,
// interval_grove_test.cpp:105-129, routing_test.cpp:52-182, precondition_test.cpp:105-111,192-261,
// grove_remove_test.cpp:573-625, heterogeneous_grove_test.cpp:98-177, tests/data_type/interval_test.cpp:121-176,
// 374-451) as templates over an `api` bundle, so the SAME test text runs against
//   * the facade            (facade_tests.bin: self-checked against brute force on the host), and
//   * the unmodified reference (ref_parity.bin, only built where /root/reference exists): every scenario's
//     trace of hit lists must be identical, payload for payload and in the same order; plus the
//     reference-side binding of INTEGRATION.md (device snapshot taken by walking the reference's real leaf
//     chain) checked pointer-for-pointer against the reference's own intersect.
// All overlap computation on the facade side runs on the GPU (no CPU path exists); needs a CUDA device.
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <numeric>
#include <random>
#include <set>
#include <sstream>
#include <string>
#include <utility>
#include <vector>

#include "genogrove_b200/grove.hpp"

#ifdef WITH_REFERENCE
#include <genogrove/data_type/genomic_coordinate.hpp>
#include <genogrove/data_type/interval.hpp>
#include <genogrove/structure/grove/grove.hpp>
#endif

// ------------------------------------------------------------------------------------------- mini harness
static int g_checks = 0, g_failures = 0;
#define CHECK(...)                                                                       \
    do {                                                                                 \
        ++g_checks;                                                                      \
        if (!(__VA_ARGS__)) {                                                                 \
            ++g_failures;                                                                \
            std::fprintf(stderr, "FAIL %s:%d: %s\n", __FILE__, __LINE__, #__VA_ARGS__);       \
        }                                                                                \
    } while (0)
#define CHECK_THROWS(expr, exc)                                                          \
    do {                                                                                 \
        ++g_checks;                                                                      \
        bool thrown_ = false;                                                            \
        try { (void)(expr); } catch (const exc&) { thrown_ = true; } catch (...) {}      \
        if (!thrown_) {                                                                  \
            ++g_failures;                                                                \
            std::fprintf(stderr, "FAIL %s:%d: %s did not throw %s\n", __FILE__, __LINE__, #expr, #exc); \
        }                                                                                \
    } while (0)

// ------------------------------------------------------------------------------------------- api bundles
struct facade_api {
    static constexpr const char* name = "facade";
    using interval = genogrove_b200::data_type::interval;
    template <typename D>
    using grove = genogrove_b200::structure::grove<interval, D>;
    template <typename D>
    using key = genogrove_b200::data_type::key<interval, D>;
    template <typename D>
    using query_result = genogrove_b200::data_type::query_result<interval, D>;
    static constexpr auto sorted = genogrove_b200::structure::sorted;
    static constexpr auto bulk = genogrove_b200::structure::bulk;
    using genomic_coordinate = genogrove_b200::data_type::genomic_coordinate;
    template <typename D>
    using gc_grove = genogrove_b200::structure::grove<genomic_coordinate, D>;
};
#ifdef WITH_REFERENCE
struct reference_api {
    static constexpr const char* name = "reference";
    using interval = genogrove::data_type::interval;
    template <typename D>
    using grove = genogrove::structure::grove<interval, D>;
    template <typename D>
    using key = genogrove::data_type::key<interval, D>;
    template <typename D>
    using query_result = genogrove::data_type::query_result<interval, D>;
    static constexpr auto sorted = genogrove::structure::sorted;
    static constexpr auto bulk = genogrove::structure::bulk;
    using genomic_coordinate = genogrove::data_type::genomic_coordinate;
    template <typename D>
    using gc_grove = genogrove::structure::grove<genomic_coordinate, D>;
};
#endif

using trace_t = std::vector<std::string>;
using items_t = std::vector<std::pair<std::pair<std::size_t, std::size_t>, int>>;  // ((start, end), payload)

template <typename A>
std::string hits_line(const char* tag, const typename A::template query_result<int>& r) {
    std::ostringstream os;
    os << tag << " q=" << r.get_query().to_string() << " ->";
    for (auto* k : r.get_keys()) os << ' ' << k->get_data();
    return os.str();
}

template <typename A>
std::vector<int> payloads(const typename A::template query_result<int>& r) {
    std::vector<int> v;
    for (auto* k : r.get_keys()) v.push_back(k->get_data());
    return v;
}

// brute force on the test's own copy of the data: the reference tests' verify_query_results (order-independent)
static std::multiset<int> expected_set(const items_t& items, std::size_t qs, std::size_t qe) {
    std::multiset<int> s;
    for (const auto& [iv, id] : items)
        if (std::max(iv.first, qs) <= std::min(iv.second, qe)) s.insert(id);
    return s;
}

template <typename A>
void check_result(const items_t& items, const typename A::template query_result<int>& r) {
    const auto q = r.get_query();
    auto p = payloads<A>(r);
    CHECK(std::multiset<int>(p.begin(), p.end()) == expected_set(items, q.get_start(), q.get_end()));
    // hits arrive in leaf-chain order: ascending (start, end)
    for (std::size_t i = 1; i < r.get_keys().size(); ++i)
        CHECK(!(r.get_keys()[i]->get_value() < r.get_keys()[i - 1]->get_value()));
}

enum class mode { regular, sorted_single, sorted_bulk, unsorted_bulk };

template <typename A>
void load(typename A::template grove<int>& g, const char* index, const items_t& items, mode m) {
    using interval = typename A::interval;
    if (m == mode::regular) {
        for (const auto& [iv, id] : items) g.insert_data(index, interval(iv.first, iv.second), id);
    } else if (m == mode::sorted_single) {
        for (const auto& [iv, id] : items) g.insert_data(index, interval(iv.first, iv.second), id, A::sorted);
    } else {
        std::vector<std::pair<interval, int>> v;
        for (const auto& [iv, id] : items) v.push_back({interval(iv.first, iv.second), id});
        if (m == mode::sorted_bulk) g.insert_data(index, v, A::sorted, A::bulk);
        else g.insert_data(index, v, A::bulk);
    }
}

static items_t sorted_copy(items_t v) {
    std::stable_sort(v.begin(), v.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
    return v;
}

// ------------------------------------------------------------------------------------------- data shapes
static items_t shape_disjoint(int n) {
    items_t d;
    for (int i = 0; i < n; ++i) d.push_back({{std::size_t(i) * 100, std::size_t(i) * 100 + 50}, i});
    return d;
}
static items_t shape_equal_starts(int n) {
    items_t d;
    for (int i = 0; i < n; ++i) {
        std::size_t s = std::size_t(i / 3) * 100;
        d.push_back({{s, s + 10 + std::size_t(i % 3) * 40}, i});
    }
    return d;
}
static items_t shape_nested(int n) {
    items_t d;
    for (int i = 0; i < n; ++i) {
        std::size_t s = std::size_t(i) * 100;
        d.push_back({{s, (i % 3 == 0) ? s + 5000 : s + 40}, i});
    }
    return d;
}
static items_t shape_duplicates(int n) {
    items_t d;
    for (int i = 0; i < n / 2; ++i) {
        d.push_back({{std::size_t(i) * 100, std::size_t(i) * 100 + 50}, 2 * i});
        d.push_back({{std::size_t(i) * 100, std::size_t(i) * 100 + 50}, 2 * i + 1});
    }
    return d;
}
static items_t shape_touching(int n) {
    items_t d;
    for (int i = 0; i < n; ++i) d.push_back({{std::size_t(i) * 10, std::size_t(i) * 10 + 9}, i});
    return d;
}
static items_t shape_one_spanning(int n) {
    items_t d = shape_disjoint(n - 1);
    d.push_back({{0, std::size_t(n) * 100}, n});
    return d;
}

// ------------------------------------------------------------------------------------------- scenarios
// key_type_grove_test.hpp:386-539 with the interval traits of interval_grove_test.cpp:33-86
template <typename A>
void scenario_typed(trace_t& t) {
    using interval = typename A::interval;
    for (int order : {3, 4, 10, 100})
        for (int n : {1, 2, 5, 50, 400})
            for (mode m : {mode::regular, mode::sorted_single, mode::sorted_bulk, mode::unsorted_bulk}) {
                typename A::template grove<int> g(order);
                items_t items = shape_disjoint(n);
                load<A>(g, "index1", items, m);
                interval overlapping = n < 2 ? interval(10, 40) : interval(120, std::size_t(std::min(3, n - 1)) * 100 + 20);
                auto r1 = g.intersect(overlapping, "index1");
                check_result<A>(items, r1);
                t.push_back(hits_line<A>("typed", r1));
                auto r2 = g.intersect(interval(60, n > 1 ? 90 : 80), "index1");  // the gap after the first interval
                CHECK(r2.get_keys().empty());
                auto r3 = g.intersect(overlapping, "no_such_index");
                CHECK(r3.get_keys().empty());
                auto r4 = g.intersect(overlapping);  // all indices
                t.push_back(hits_line<A>("typed_all", r4));
                CHECK(payloads<A>(r4) == payloads<A>(r1));
            }
}

// interval_grove_test.cpp:105-129 (#517): 500 shuffled keys at order 3, every key found by its own coordinates
template <typename A>
void scenario_deep_unsorted(trace_t& t) {
    using interval = typename A::interval;
    typename A::template grove<int> g(3);
    std::vector<int> pos(500);
    std::iota(pos.begin(), pos.end(), 0);
    std::shuffle(pos.begin(), pos.end(), std::mt19937{42});
    for (int p : pos) g.insert_data("chr1", interval(std::size_t(p) * 100, std::size_t(p) * 100 + 50), p);
    for (int p : pos) {
        auto r = g.intersect(interval(std::size_t(p) * 100, std::size_t(p) * 100 + 50), "chr1");
        CHECK(r.get_keys().size() == 1u);
        if (r.get_keys().size() == 1u) CHECK(r.get_keys()[0]->get_data() == p);
    }
    t.push_back(hits_line<A>("deep", g.intersect(interval(0, 100000), "chr1")));
}

// routing_test.cpp:52-182: six shapes x orders {3,4,10,100} x insert paths; the chain must equal the sorted keys
template <typename A>
void scenario_routing(trace_t& t) {
    using interval = typename A::interval;
    struct shape { const char* name; items_t (*make)(int); bool distinct; };
    const shape shapes[] = {{"disjoint", shape_disjoint, true},   {"equal_starts", shape_equal_starts, true},
                            {"nested", shape_nested, true},       {"duplicates", shape_duplicates, false},
                            {"touching", shape_touching, true},   {"one_spanning", shape_one_spanning, true}};
    unsigned seed = 1;
    for (const auto& sh : shapes)
        for (int order : {3, 4, 10, 100})
            for (mode m : {mode::regular, mode::sorted_single, mode::sorted_bulk, mode::unsorted_bulk}) {
                if (m == mode::unsorted_bulk && !sh.distinct) continue;  // the reference's std::sort leaves equal keys unordered
                items_t items = sh.make(120);
                items_t feed = items;
                if (m == mode::regular || m == mode::unsorted_bulk) std::shuffle(feed.begin(), feed.end(), std::mt19937{seed++});
                else feed = sorted_copy(feed);
                typename A::template grove<int> g(order);
                load<A>(g, "chr1", feed, m);
                // whole-index query == the leaf chain
                auto all = g.intersect(interval(0, 1u << 30), "chr1");
                CHECK(all.get_keys().size() == items.size());
                check_result<A>(items, all);
                t.push_back(hits_line<A>(sh.name, all));
                std::mt19937 rng(seed++);
                for (int k = 0; k < 12; ++k) {
                    const auto& it = items[rng() % items.size()];
                    auto r = g.intersect(interval(it.first.first, it.first.second), "chr1");
                    check_result<A>(items, r);
                    CHECK(!r.get_keys().empty());
                    t.push_back(hits_line<A>(sh.name, r));
                    std::size_t a = rng() % 13000, b = a + rng() % 700;
                    auto r2 = g.intersect(interval(a, b), "chr1");
                    check_result<A>(items, r2);
                    t.push_back(hits_line<A>(sh.name, r2));
                }
            }
}

// precondition_test.cpp:105-111,192-261 and the constructor / interval / query_result preconditions
template <typename A>
void scenario_edge_cases(trace_t& t) {
    using interval = typename A::interval;
    CHECK_THROWS(typename A::template grove<int>(2), std::invalid_argument);
    CHECK_THROWS(typename A::template grove<int>(0), std::invalid_argument);
    CHECK_THROWS(interval(10, 5), std::invalid_argument);
    {
        typename A::template query_result<int> qr{interval(0, 1)};
        CHECK_THROWS(qr.add_key(nullptr), std::invalid_argument);
    }
    {
        typename A::template grove<int> g;  // default order 3
        CHECK(g.get_order() == 3);
        auto r = g.intersect(interval(0, 100), "chr1");  // empty grove, unknown index
        CHECK(r.get_keys().empty());
        auto r2 = g.intersect(interval(0, 100));
        CHECK(r2.get_keys().empty());
    }
    {
        typename A::template grove<int> g(10);
        g.insert_data("chr1", interval{10, 20}, 42);
        auto r = g.intersect(interval{15, 15}, "chr1");
        CHECK(r.get_keys().size() == 1u && r.get_keys()[0]->get_data() == 42);
        CHECK(g.intersect(interval{30, 40}, "chr1").get_keys().empty());
        CHECK(g.intersect(interval{15, 15}, "chr2").get_keys().empty());
        // adjacency overlaps (closed intervals): [20,25] touches [10,20]
        CHECK(g.intersect(interval{20, 25}, "chr1").get_keys().size() == 1u);
        CHECK(g.intersect(interval{21, 25}, "chr1").get_keys().empty());
        CHECK(g.intersect(interval{0, 10}, "chr1").get_keys().size() == 1u);
        CHECK(g.intersect(interval{0, 9}, "chr1").get_keys().empty());
    }
    {
        typename A::template grove<int> g(10);
        g.insert_data("chr1", interval{10, 20}, 1);
        g.insert_data("chr1", interval{10, 20}, 2);
        auto r = g.intersect(interval{10, 20}, "chr1");
        CHECK(r.get_keys().size() == 2u);
        t.push_back(hits_line<A>("dup", r));
    }
    {   // #490: bulk sorted-append into a non-empty index, order 3, 500 keys
        typename A::template grove<int> g(3);
        g.insert_data("chr1", interval{0, 5}, 0, A::sorted);
        std::vector<std::pair<interval, int>> batch;
        for (int i = 1; i < 500; ++i) batch.emplace_back(interval{std::size_t(i) * 10, std::size_t(i) * 10 + 5}, i);
        g.insert_data("chr1", batch, A::sorted, A::bulk);
        auto r = g.intersect(interval{0, 5000}, "chr1");
        CHECK(r.get_keys().size() == 500u);
        t.push_back(hits_line<A>("append", r));
    }
    {
        typename A::template grove<int> g(5);
        for (int i = 0; i < 4; ++i) g.insert_data("chr1", interval{std::size_t(i) * 100, std::size_t(i) * 100 + 50}, i);
        CHECK(g.intersect(interval{0, 350}, "chr1").get_keys().size() == 4u);
    }
    {   // interval_test.cpp:374-451: zero-length points and coordinates near SIZE_MAX (u64 device layout)
        typename A::template grove<int> g(4);
        const std::size_t M = interval::INVALID_POSITION;
        g.insert_data("big", interval{M - 10, M - 1}, 1);
        g.insert_data("big", interval{M - 100, M - 50}, 2);
        g.insert_data("big", interval{5, 5}, 3);
        g.insert_data("big", interval{0, M}, 4);
        for (auto q : {interval{M - 5, M}, interval{M - 60, M - 55}, interval{5, 5}, interval{6, 6}, interval{0, 0},
                       interval{M, M}, interval{M - 49, M - 11}}) {
            auto r = g.intersect(q, "big");
            t.push_back(hits_line<A>("big", r));
        }
        CHECK(g.intersect(interval{5, 5}, "big").get_keys().size() == 2u);
        CHECK(g.intersect(interval{M - 5, M}, "big").get_keys().size() == 2u);
        CHECK(g.intersect(interval{M - 49, M - 11}, "big").get_keys().size() == 1u);
    }
}

// grove_remove_test.cpp:573-625: intersect after removal, single index and all indices
template <typename A>
void scenario_removal(trace_t& t) {
    using interval = typename A::interval;
    for (int order : {3, 5, 32}) {
        typename A::template grove<int> g(order);
        std::vector<typename A::template key<int>*> keys;
        items_t items = shape_nested(200);
        for (const auto& [iv, id] : items) keys.push_back(g.insert_data("chr1", interval(iv.first, iv.second), id));
        std::vector<typename A::template key<int>*> keys2;
        for (int i = 0; i < 40; ++i) keys2.push_back(g.insert_data("chr2", interval(std::size_t(i) * 50, std::size_t(i) * 50 + 60), 1000 + i));
        items_t live;
        for (std::size_t i = 0; i < keys.size(); ++i) {
            if (i % 3 == 1) CHECK(g.remove_key("chr1", keys[i]));
            else live.push_back(items[i]);
        }
        CHECK(!g.remove_key("chr1", keys[1]));        // already removed
        CHECK(!g.remove_key("chrX", keys[0]));        // unknown index
        CHECK(!g.remove_key("chr1", nullptr));
        auto r = g.intersect(interval(0, 1u << 30), "chr1");
        CHECK(r.get_keys().size() == live.size());
        check_result<A>(live, r);
        t.push_back(hits_line<A>("removed", r));
        for (std::size_t a : {0u, 150u, 5000u, 9990u, 19900u}) {
            auto rr = g.intersect(interval(a, a + 120), "chr1");
            check_result<A>(live, rr);
            t.push_back(hits_line<A>("removed", rr));
        }
        t.push_back(hits_line<A>("removed_all_indices", g.intersect(interval(100, 400))));
        // re-insert after removal, then query again
        g.insert_data("chr1", interval(150, 170), 7777);
        live.push_back({{150, 170}, 7777});
        auto r2 = g.intersect(interval(100, 200), "chr1");
        check_result<A>(live, r2);
        t.push_back(hits_line<A>("reinserted", r2));
        // empty an index completely: it disappears
        for (auto* k : keys2) CHECK(g.remove_key("chr2", k));
        CHECK(g.intersect(interval(0, 1u << 30), "chr2").get_keys().empty());
        t.push_back(hits_line<A>("emptied", g.intersect(interval(0, 500))));
    }
}

// grove_query.ipp:16-24: all-index queries append per-index hit lists in the host map's iteration order
template <typename A>
void scenario_multi_index(trace_t& t) {
    using interval = typename A::interval;
    typename A::template grove<int> g(6);
    const char* names[] = {"chr1", "chr10", "chr2", "chrX", "chrM", "scaffold_17", "chr3", "chrY", "chr21"};
    int id = 0;
    items_t all;
    for (const char* nm : names)
        for (int i = 0; i < 25; ++i) {
            std::size_t s = std::size_t(i) * 40 + (id % 7);
            g.insert_data(nm, interval(s, s + 55), id);
            all.push_back({{s, s + 55}, id});
            ++id;
        }
    for (std::size_t a : {0u, 100u, 333u, 990u, 2000u}) {
        auto r = g.intersect(interval(a, a + 30));
        auto p = payloads<A>(r);
        CHECK(std::multiset<int>(p.begin(), p.end()) == expected_set(all, a, a + 30));
        t.push_back(hits_line<A>("multi", r));
    }
}

// heterogeneous_grove_test.cpp:98-177: hits come back in ascending (start, end) order
template <typename A>
void scenario_order_pinned(trace_t& t) {
    using interval = typename A::interval;
    typename A::template grove<int> g(4);
    const std::pair<std::size_t, std::size_t> ivs[] = {{500, 600}, {100, 200}, {300, 400}, {100, 150}, {250, 700}, {100, 300}};
    int id = 0;
    for (auto [s, e] : ivs) g.insert_data("chr1", interval(s, e), id++);
    auto r = g.intersect(interval(0, 1000), "chr1");
    CHECK((payloads<A>(r) == std::vector<int>{3, 1, 5, 4, 2, 0}));
    t.push_back(hits_line<A>("ordered", r));
}

// order among EQUAL keys per insert path
template <typename A>
void scenario_tie_order(trace_t& t) {
    using interval = typename A::interval;
    {
        typename A::template grove<int> g(4);
        for (int i = 0; i < 6; ++i) g.insert_data("c", interval(10, 20), i);
        auto r = g.intersect(interval(10, 20), "c");
        CHECK((payloads<A>(r) == std::vector<int>{5, 4, 3, 2, 1, 0}));  // lower_bound insertion: reverse insertion order
        t.push_back(hits_line<A>("tie_regular", r));
    }
    {
        typename A::template grove<int> g(4);
        std::vector<std::pair<interval, int>> v;
        for (int i = 0; i < 6; ++i) v.push_back({interval(10, 20), i});
        g.insert_data("c", v, A::sorted, A::bulk);
        auto r = g.intersect(interval(10, 20), "c");
        CHECK((payloads<A>(r) == std::vector<int>{0, 1, 2, 3, 4, 5}));
        t.push_back(hits_line<A>("tie_sorted_bulk", r));
    }
    for (int order : {3, 4, 5, 8}) {
        typename A::template grove<int> g(order);
        for (int i = 0; i < 11; ++i) g.insert_data("c", interval(10, 20), i, A::sorted);
        auto r = g.intersect(interval(10, 20), "c");
        t.push_back(hits_line<A>("tie_sorted_append", r));
    }
    {   // six equal keys at order 4 (SURVEY probe: 3 2 1 5 4 0)
        typename A::template grove<int> g(4);
        for (int i = 0; i < 6; ++i) g.insert_data("c", interval(10, 20), i, A::sorted);
        auto r = g.intersect(interval(10, 20), "c");
        CHECK((payloads<A>(r) == std::vector<int>{3, 2, 1, 5, 4, 0}));
        t.push_back(hits_line<A>("tie_sorted_append6", r));
    }
    {   // sorted bulk on an empty index followed by sorted appends of keys equal to the current maximum
        typename A::template grove<int> g(5);
        std::vector<std::pair<interval, int>> v;
        for (int i = 0; i < 10; ++i) v.push_back({interval(std::size_t(i / 2) * 10, std::size_t(i / 2) * 10 + 5), i});
        g.insert_data("c", v, A::sorted, A::bulk);
        for (int i = 10; i < 16; ++i) g.insert_data("c", interval(40, 45), i, A::sorted);
        std::vector<std::pair<interval, int>> w;
        for (int i = 16; i < 24; ++i) w.push_back({interval(40 + std::size_t(i / 20) * 10, 45 + std::size_t(i / 20) * 10), i});
        g.insert_data("c", w, A::sorted, A::bulk);
        t.push_back(hits_line<A>("tie_mixed_sorted", g.intersect(interval(0, 100), "c")));
    }
    {   // regular inserts of duplicates on top of a bulk-built chain
        typename A::template grove<int> g(3);
        std::vector<std::pair<interval, int>> v;
        for (int i = 0; i < 12; ++i) v.push_back({interval(std::size_t(i / 3) * 10, std::size_t(i / 3) * 10 + 5), i});
        g.insert_data("c", v, A::sorted, A::bulk);
        for (int i = 12; i < 20; ++i) g.insert_data("c", interval(std::size_t(i % 4) * 10, std::size_t(i % 4) * 10 + 5), i);
        t.push_back(hits_line<A>("tie_regular_on_bulk", g.intersect(interval(0, 100), "c")));
    }
}

// randomised stress: duplicates + nested + several indices, regular and bulk paths
template <typename A>
void scenario_random(trace_t& t) {
    using interval = typename A::interval;
    std::mt19937_64 rng(20260101);
    for (int round = 0; round < 3; ++round) {
        const int order = round == 0 ? 3 : (round == 1 ? 16 : 64);
        typename A::template grove<int> g(order);
        const char* names[] = {"chr1", "chr2", "chr3"};
        std::vector<items_t> per(3);
        int id = 0;
        for (int c = 0; c < 3; ++c) {
            const int n = 3000 + 1500 * c;
            for (int i = 0; i < n; ++i) {
                std::size_t s = rng() % 200000;
                std::size_t len = (rng() % 50 == 0) ? 20000 + rng() % 50000 : rng() % 400;
                if (rng() % 10 == 0 && !per[c].empty()) {  // exact duplicate of an earlier key
                    auto prev = per[c][rng() % per[c].size()].first;
                    s = prev.first;
                    len = prev.second - prev.first;
                }
                per[c].push_back({{s, s + len}, id++});
            }
            if (c == 0) load<A>(g, names[c], per[c], mode::regular);
            else if (c == 1) load<A>(g, names[c], sorted_copy(per[c]), mode::sorted_bulk);
            else load<A>(g, names[c], sorted_copy(per[c]), mode::sorted_single);
        }
        for (int k = 0; k < 600; ++k) {
            const int c = int(rng() % 3);
            std::size_t a = rng() % 210000, b = a + (rng() % 20 == 0 ? rng() % 30000 : rng() % 300);
            auto r = g.intersect(interval(a, b), names[c]);
            check_result<A>(per[c], r);
            t.push_back(hits_line<A>("rand", r));
        }
    }
}

// flanking (nearest predecessor / successor): grove_flanking_test.cpp:20-117 restated + random groves with duplicated
// and nested keys through three insert paths; the trace carries the payloads, so ties must resolve the same way
template <typename A>
void scenario_flanking(trace_t& t) {
    using interval = typename A::interval;
    auto line = [&](const char* tag, auto& g, std::size_t a, std::size_t b, const char* index) {
        auto nn = g.flanking(interval(a, b), index);
        std::ostringstream os;
        os << tag << " q=[" << a << ',' << b << "] pred=";
        if (nn.get_predecessor()) os << nn.get_predecessor()->get_data(); else os << "null";
        os << " succ=";
        if (nn.get_successor()) os << nn.get_successor()->get_data(); else os << "null";
        t.push_back(os.str());
        return nn;
    };
    {
        typename A::template grove<int> g(8);
        auto nn = line("flank_empty", g, 100, 200, "chr1");
        CHECK(nn.get_predecessor() == nullptr && nn.get_successor() == nullptr);
        auto* far_pred = g.insert_data("chr1", interval(50, 60), 2, A::sorted);
        g.insert_data("chr1", interval(100, 350), 1, A::sorted);
        auto* far_succ = g.insert_data("chr1", interval(500, 600), 3, A::sorted);
        nn = line("flank_skip_overlap", g, 300, 400, "chr1");
        CHECK(nn.get_predecessor() == far_pred && nn.get_successor() == far_succ);
        nn = line("flank_missing_index", g, 300, 400, "chr2");
        CHECK(nn.get_predecessor() == nullptr && nn.get_successor() == nullptr);
    }
    {
        typename A::template grove<int> g(8);
        auto* outer = g.insert_data("chr1", interval(50, 100), 1, A::sorted);
        g.insert_data("chr1", interval(80, 90), 2, A::sorted);
        auto nn = line("flank_nested", g, 200, 300, "chr1");
        CHECK(nn.get_predecessor() == outer);
        auto* lhs = g.insert_data("chr1", interval(120, 199), 3, A::sorted);
        auto* rhs = g.insert_data("chr1", interval(301, 400), 4, A::sorted);
        nn = line("flank_abutting", g, 200, 300, "chr1");
        CHECK(nn.get_predecessor() == lhs && nn.get_successor() == rhs);
    }
    {
        typename A::template grove<int> g(4);
        for (std::size_t i = 0; i < 16; ++i) g.insert_data("chr1", interval(i * 100, i * 100 + 20), int(i), A::sorted);
        auto nn = line("flank_multi_leaf", g, 440, 460, "chr1");
        CHECK(nn.get_predecessor() && nn.get_predecessor()->get_value().get_end() == 420u);
        CHECK(nn.get_successor() && nn.get_successor()->get_value().get_start() == 500u);
    }
    std::mt19937_64 rng(777);
    for (int round = 0; round < 6; ++round) {
        const int order = round % 3 == 0 ? 3 : (round % 3 == 1 ? 8 : 40);
        typename A::template grove<int> g(order);
        items_t items;
        const int n = 400 + 300 * round;
        for (int i = 0; i < n; ++i) {
            std::size_t s = (rng() % 30000) / 16 * 16;
            std::size_t len = (rng() % 40 == 0) ? 5000 + rng() % 9000 : (rng() % 300) / 32 * 32;
            items.push_back({{s, s + len}, i});
        }
        const mode m = round % 3 == 0 ? mode::regular : (round % 3 == 1 ? mode::sorted_bulk : mode::sorted_single);
        load<A>(g, "c", m == mode::regular ? items : sorted_copy(items), m);
        for (int k = 0; k < 250; ++k) {
            std::size_t a = rng() % 46000, b = a + rng() % 500;
            auto nn = line("flank_rand", g, a, b, "c");
            // brute force on the test's own data: largest end below a, smallest start above b
            std::size_t best_end = 0, best_start = 0;
            bool hp = false, hs = false;
            for (const auto& [iv, id] : items) {
                if (iv.second < a && (!hp || iv.second > best_end)) { hp = true; best_end = iv.second; }
                if (iv.first > b && (!hs || iv.first < best_start)) { hs = true; best_start = iv.first; }
            }
            CHECK((nn.get_predecessor() != nullptr) == hp && (!hp || nn.get_predecessor()->get_value().get_end() == best_end));
            CHECK((nn.get_successor() != nullptr) == hs && (!hs || nn.get_successor()->get_value().get_start() == best_start));
        }
    }
}

// genomic_coordinate keys (strand-aware overlap): genomic_coordinate_grove_test.cpp:150-266 restated + random groves
// through regular / sorted / sorted-bulk / unsorted-bulk inserts; traces carry payloads in result order
template <typename A>
void scenario_stranded(trace_t& t) {
    using gc = typename A::genomic_coordinate;
    auto line = [&](const char* tag, auto& g, char st, std::size_t a, std::size_t b, const char* index) {
        auto r = g.intersect(gc(st, a, b), index);
        std::ostringstream os;
        os << tag << " q=" << r.get_query().to_string() << " ->";
        for (auto* k : r.get_keys()) os << ' ' << k->get_data();
        t.push_back(os.str());
        std::vector<int> v;
        for (auto* k : r.get_keys()) v.push_back(k->get_data());
        return v;
    };
    {
        typename A::template gc_grove<int> g(3);
        g.insert_data("chr1", gc('+', 10, 30), 10);
        g.insert_data("chr1", gc('+', 40, 60), 70);
        g.insert_data("chr1", gc('+', 80, 100), 90);
        g.insert_data("chr1", gc('-', 10, 30), 20);
        g.insert_data("chr1", gc('-', 50, 90), 50);
        g.insert_data("chr1", gc('-', 100, 120), 71);
        g.insert_data("chr1", gc('.', 10, 30), 30);
        g.insert_data("chr1", gc('.', 25, 85), 40);
        CHECK(line("gc_plus", g, '+', 15, 85, "chr1") == std::vector<int>({10, 70, 90}));
        CHECK(line("gc_minus", g, '-', 15, 55, "chr1") == std::vector<int>({20, 50}));
        CHECK(line("gc_dot", g, '.', 15, 55, "chr1") == std::vector<int>({30, 40}));
        CHECK(line("gc_wild", g, '*', 0, 200, "chr1").size() == 8u);
        CHECK(line("gc_missing", g, '*', 0, 200, "chr2").empty());
        CHECK_THROWS(gc('x', 1, 2), std::invalid_argument);
        CHECK_THROWS(gc('+', 3, 2), std::invalid_argument);
    }
    {
        typename A::template gc_grove<int> g(5);
        std::vector<std::pair<gc, int>> data = {{gc('*', 10, 20), 10}, {gc('.', 10, 20), 20}, {gc('+', 10, 20), 30}, {gc('-', 10, 20), 40}};
        g.insert_data("chr1", data, A::sorted, A::bulk);
        CHECK(line("gc_bulk_order", g, '*', 10, 20, "chr1") == std::vector<int>({10, 20, 30, 40}));
        CHECK(line("gc_bulk_plus", g, '+', 10, 20, "chr1") == std::vector<int>({10, 30}));  // '*' keys match any query strand
    }
    std::mt19937_64 rng(4242);
    const char sym[4] = {'+', '-', '.', '*'};
    for (int round = 0; round < 8; ++round) {
        const int order = round % 4 == 0 ? 3 : (round % 4 == 1 ? 6 : (round % 4 == 2 ? 32 : 4));
        typename A::template gc_grove<int> g(order);
        std::vector<std::pair<gc, int>> items;
        const int n = 300 + 250 * round;
        for (int i = 0; i < n; ++i) {
            std::size_t s = (rng() % 20000) / 8 * 8;
            std::size_t len = (rng() % 50 == 0) ? 4000 + rng() % 6000 : (rng() % 256) / 32 * 32;
            // distinct full keys in the unsorted-bulk rounds (the reference's std::sort leaves equal keys in no defined order)
            if (round % 4 == 3) s = static_cast<std::size_t>(i) * 8 + (rng() % 2) * 20000;
            items.push_back({gc(sym[rng() % (round % 2 ? 4 : 3)], s, s + len), i});
        }
        auto sorted_items = items;
        std::stable_sort(sorted_items.begin(), sorted_items.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
        if (round % 4 == 0) {
            for (const auto& [k, d] : items) g.insert_data("c", k, d);
        } else if (round % 4 == 1) {
            for (const auto& [k, d] : sorted_items) g.insert_data("c", k, d, A::sorted);
        } else if (round % 4 == 2) {
            g.insert_data("c", sorted_items, A::sorted, A::bulk);
        } else {
            g.insert_data("c", items, A::bulk);
        }
        for (int k = 0; k < 200; ++k) {
            const std::size_t a = rng() % 42000, b = a + (rng() % 25 == 0 ? rng() % 8000 : rng() % 400);
            const char st = sym[rng() % 4];
            auto got = line("gc_rand", g, st, a, b, "c");
            std::multiset<int> want;
            for (const auto& [kk, d] : items)
                if (gc::overlaps(kk, gc(st, a, b))) want.insert(d);
            CHECK(std::multiset<int>(got.begin(), got.end()) == want);
        }
    }
}

template <typename A>
trace_t run_scenarios() {
    trace_t t;
    scenario_typed<A>(t);
    scenario_deep_unsorted<A>(t);
    scenario_routing<A>(t);
    scenario_edge_cases<A>(t);
    scenario_removal<A>(t);
    scenario_multi_index<A>(t);
    scenario_order_pinned<A>(t);
    scenario_tie_order<A>(t);
    scenario_random<A>(t);
    scenario_flanking<A>(t);
    scenario_stranded<A>(t);
    return t;
}

// ------------------------------------------------------------------------------------------- facade-only
// interval_test.cpp:121-176: overlap known answers on the facade's own interval type
static void test_interval_arithmetic() {
    using interval = facade_api::interval;
    CHECK(interval::overlaps(interval(0, 10), interval(10, 20)));    // adjacent counts as overlap
    CHECK(!interval::overlaps(interval(0, 10), interval(11, 20)));
    CHECK(interval::overlaps(interval(0, 100), interval(40, 50)));   // contained
    CHECK(interval::overlaps(interval(5, 5), interval(5, 5)));       // identical points
    CHECK(!interval::overlaps(interval(5, 5), interval(6, 6)));
    CHECK(interval(1, 5) < interval(2, 3));
    CHECK(interval(1, 5) < interval(1, 6));
    CHECK(interval(2, 3) > interval(1, 9));
    CHECK(!(interval(1, 5) < interval(1, 5)) && !(interval(1, 5) > interval(1, 5)) && interval(1, 5) == interval(1, 5));
    CHECK(interval::aggregate(interval(5, 10), interval(1, 7)) == interval(1, 10));
    CHECK(interval(3, 9).to_string() == "[3,9]");
    interval d;
    CHECK(d.get_start() == interval::INVALID_POSITION && d.get_end() == interval::INVALID_POSITION);
    interval s(1, 2);
    CHECK_THROWS(s.set_range(9, 3), std::invalid_argument);
    s.set_range(4, 8);
    CHECK(s.get_start() == 4 && s.get_end() == 8);
}

// the new batch entry point agrees with per-query intersect, and maps ranks back to the same key pointers
static void test_batch_entry_point() {
    using interval = facade_api::interval;
    namespace gst = genogrove_b200::structure;
    gst::grove<interval, int> g(16);
    std::mt19937_64 rng(7);
    const char* names[] = {"chr1", "chr2"};
    for (int c = 0; c < 2; ++c) {
        std::vector<std::pair<interval, int>> v;
        for (int i = 0; i < 50000; ++i) {
            std::size_t s = rng() % 5000000;
            v.push_back({interval(s, s + rng() % 3000), c * 1000000 + i});
        }
        g.insert_data(names[c], v, gst::bulk);
    }
    const std::size_t nq = 20000;
    std::vector<std::uint32_t> qc(nq);
    std::vector<std::uint64_t> qs(nq), qe(nq);
    const std::uint32_t ids[3] = {g.index_id("chr1"), g.index_id("chr2"), g.index_id("nope")};
    CHECK(ids[2] == gst::grove<interval, int>::npos_index);
    for (std::size_t i = 0; i < nq; ++i) {
        qc[i] = ids[rng() % 3];
        qs[i] = rng() % 5000000;
        qe[i] = qs[i] + rng() % 2000;
    }
    auto br = g.intersect_batch(qc.data(), qs.data(), qe.data(), nq);
    CHECK(br.size() == nq);
    std::vector<std::vector<gst::grove<interval, int>::key_ptr>> got(nq);
    for (std::size_t i = 0; i < nq; ++i) got[i] = br.get_keys(i);  // before the next query invalidates the view
    std::uint64_t total = 0;
    for (std::size_t i = 0; i < nq; i += 7) {
        const char* nm = qc[i] == ids[0] ? "chr1" : (qc[i] == ids[1] ? "chr2" : "nope");
        auto r = g.intersect(interval(qs[i], qe[i]), nm);
        CHECK(r.get_keys() == got[i]);
        total += r.get_keys().size();
    }
    CHECK(total > 0);
    std::vector<std::uint64_t> bad_s{10}, bad_e{5};
    std::vector<std::uint32_t> bad_c{ids[0]};
    CHECK_THROWS(g.intersect_batch(bad_c.data(), bad_s.data(), bad_e.data(), 1), std::invalid_argument);
    // span overload
    std::vector<interval> qv{interval(0, 100000), interval(2000000, 2000500)};
    auto b2 = g.intersect_batch(std::span<const interval>(qv), "chr1");
    auto b2k0 = b2.get_keys(0), b2k1 = b2.get_keys(1);
    CHECK(g.intersect(qv[0], "chr1").get_keys() == b2k0);
    CHECK(g.intersect(qv[1], "chr1").get_keys() == b2k1);
    // move semantics: a moved-to grove answers like the source did
    gst::grove<interval, int> h(std::move(g));
    CHECK(h.intersect(qv[0], "chr1").get_keys() == b2k0);
}

#ifdef WITH_REFERENCE
// ------------------------------------------------------------------------------------------- reference-side binding
// what a genogrove maintainer adds — flatten the reference grove's REAL leaf chain into SoA
// arrays, hand them to gg_index_upload_sorted, answer batches with gg_intersect_batch and map ranks to key*.
namespace binding {
namespace gdt = genogrove::data_type;
namespace gst = genogrove::structure;

template <typename D>
struct device_snapshot {
    gg_index* handle = nullptr;
    std::vector<std::string> names;                                    // chrom id -> index name
    std::vector<std::vector<gdt::key<gdt::interval, D>*>> rank_to_key;  // per chrom id

    explicit device_snapshot(const gst::grove<gdt::interval, D>& g) {
        if (gg_index_create(&handle, 0) != GG_OK) throw std::runtime_error("no CUDA device");
        std::vector<std::uint32_t> c;
        std::vector<std::uint64_t> s, e;
        for (const auto& [name, root] : g.get_root_nodes()) names.push_back(name);
        std::sort(names.begin(), names.end());
        for (std::uint32_t id = 0; id < names.size(); ++id) {
            auto* n = g.get_root_nodes().at(names[id]);
            while (!n->get_is_leaf()) n = n->get_children().front();  // leftmost leaf
            rank_to_key.emplace_back();
            for (; n != nullptr; n = n->get_next())
                for (auto* k : n->get_keys()) {
                    rank_to_key.back().push_back(k);
                    c.push_back(id);
                    s.push_back(k->get_value().get_start());
                    e.push_back(k->get_value().get_end());
                }
        }
        if (gg_index_upload_sorted(handle, c.data(), s.data(), e.data(), c.size(), std::max<std::uint32_t>(1, names.size())) != GG_OK)
            throw std::runtime_error(gg_last_error(handle));
    }
    ~device_snapshot() { gg_index_destroy(handle); }
    std::uint32_t id_of(std::string_view name) const {
        auto it = std::lower_bound(names.begin(), names.end(), name);
        return (it != names.end() && *it == name) ? std::uint32_t(it - names.begin()) : GG_NO_CHROM;
    }
};

static void test_reference_binding() {
    std::mt19937_64 rng(99);
    for (int order : {3, 7, 50}) {
        gst::grove<gdt::interval, int> g(order);
        const char* names[] = {"chr1", "chr2", "chrX"};
        int id = 0;
        std::vector<gdt::key<gdt::interval, int>*> inserted;
        for (int c = 0; c < 3; ++c) {
            if (c == 1) {
                std::vector<std::pair<gdt::interval, int>> v;
                for (int i = 0; i < 4000; ++i) {
                    std::size_t s = rng() % 300000;
                    v.push_back({gdt::interval(s, s + rng() % 2000), id++});
                }
                g.insert_data(names[c], v, gst::bulk);
            } else {
                for (int i = 0; i < 4000; ++i) {
                    std::size_t s = rng() % 300000;
                    std::size_t len = (i % 40 == 0) ? 50000 : rng() % 500;
                    if (i % 9 == 0) s = (s / 1000) * 1000, len = 100;  // duplicates
                    inserted.push_back(g.insert_data(names[c], gdt::interval(s, s + len), id++));
                }
            }
        }
        for (std::size_t i = 0; i < inserted.size(); i += 5) g.remove_key(i < 4000 ? "chr1" : "chrX", inserted[i]);
        device_snapshot<int> snap(g);
        const std::size_t nq = 5000;
        std::vector<std::uint32_t> qc(nq);
        std::vector<std::uint64_t> qs(nq), qe(nq), off(nq + 1);
        std::vector<std::string> qname(nq);
        for (std::size_t i = 0; i < nq; ++i) {
            qname[i] = (rng() % 10 == 0) ? "unknown" : names[rng() % 3];
            qc[i] = snap.id_of(qname[i]);
            qs[i] = rng() % 310000;
            qe[i] = qs[i] + (rng() % 25 == 0 ? rng() % 60000 : rng() % 400);
        }
        const std::uint32_t* hits = nullptr;
        std::uint64_t n_hits = 0;
        CHECK(gg_intersect_batch(snap.handle, qc.data(), qs.data(), qe.data(), nq, off.data(), &hits, &n_hits) == GG_OK);
        std::uint64_t ref_total = 0;
        bool all_equal = true;
        for (std::size_t i = 0; i < nq; ++i) {
            auto r = g.intersect(gdt::interval(qs[i], qe[i]), qname[i]);
            const auto& ref_keys = r.get_keys();
            ref_total += ref_keys.size();
            if (off[i + 1] - off[i] != ref_keys.size()) { all_equal = false; continue; }
            for (std::size_t j = 0; j < ref_keys.size(); ++j)
                if (snap.rank_to_key[qc[i]][hits[off[i] + j]] != ref_keys[j]) all_equal = false;  // same key*, same position
        }
        CHECK(all_equal);
        CHECK(ref_total == n_hits);
        CHECK(ref_total > 0);
    }
}
}  // namespace binding
#endif

int main() {
    if (gg_device_count() < 1) {
#ifdef WITH_REFERENCE
        // no GPU: only check that the scenarios' own expectations hold on the unmodified reference
        trace_t ref_only = run_scenarios<reference_api>();
        std::printf("no CUDA device: reference-only run, %zu trace lines, %d checks, %d failures\n", ref_only.size(), g_checks, g_failures);
        return g_failures == 0 ? 77 : 1;
#else
        std::fprintf(stderr, "no CUDA device: the facade has no CPU path, nothing to test\n");
        return 77;
#endif
    }
    test_interval_arithmetic();
    trace_t facade = run_scenarios<facade_api>();
    test_batch_entry_point();
    std::printf("facade scenarios: %zu trace lines\n", facade.size());
#ifdef WITH_REFERENCE
    const int before = g_failures;
    trace_t ref = run_scenarios<reference_api>();  // also validates the scenarios' own expectations on the reference
    if (g_failures != before) std::fprintf(stderr, "note: %d scenario expectation(s) do not hold on the reference itself\n", g_failures - before);
    CHECK(ref.size() == facade.size());
    std::size_t diff = 0;
    for (std::size_t i = 0; i < std::min(ref.size(), facade.size()); ++i)
        if (ref[i] != facade[i]) {
            if (diff++ < 10) std::fprintf(stderr, "trace %zu differs\n  reference: %.300s\n  facade   : %.300s\n", i, ref[i].c_str(), facade[i].c_str());
        }
    CHECK(diff == 0);
    std::printf("reference parity: %zu trace lines compared, %zu differ\n", ref.size(), diff);
    binding::test_reference_binding();
#endif
    std::printf("%d checks, %d failures\n", g_checks, g_failures);
    return g_failures == 0 ? 0 : 1;
}
