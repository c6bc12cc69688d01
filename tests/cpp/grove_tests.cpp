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
#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <numeric>
#include <random>
#include <set>
#include <sstream>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

#include "genogrove_b200/grove.hpp"
#include "genogrove_b200/gg_loader.hpp"

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
        // flanking with the strand-compatible predicate (grove_flanking.ipp:47-82): nearest same-strand neighbours
        for (int k = 0; k < 80; ++k) {
            const std::size_t a = rng() % 42000, b = a + rng() % 300;
            const gc q(sym[rng() % 4], a, b);
            auto nn = g.flanking(q, "c", genogrove_b200::structure::strand_compatible);
            std::ostringstream os;
            os << "gc_flank q=" << q.to_string() << " pred=";
            if (nn.get_predecessor()) os << nn.get_predecessor()->get_data(); else os << "null";
            os << " succ=";
            if (nn.get_successor()) os << nn.get_successor()->get_data(); else os << "null";
            t.push_back(os.str());
            // brute force on the test's own data: largest compatible end below a, smallest compatible start above b
            std::size_t best_end = 0, best_start = 0;
            bool hp = false, hs = false;
            for (const auto& [kk, d] : items) {
                if (!genogrove_b200::structure::strand_compatible(kk, q)) continue;
                if (kk.get_end() < a && (!hp || kk.get_end() > best_end)) { hp = true; best_end = kk.get_end(); }
                if (kk.get_start() > b && (!hs || kk.get_start() < best_start)) { hs = true; best_start = kk.get_start(); }
            }
            CHECK((nn.get_predecessor() != nullptr) == hp && (!hp || nn.get_predecessor()->get_value().get_end() == best_end));
            CHECK((nn.get_successor() != nullptr) == hs && (!hs || nn.get_successor()->get_value().get_start() == best_start));
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
    {   // keep(): a result that outlives the next query on the grove
        auto kept = g.intersect_batch(qc.data(), qs.data(), qe.data(), nq);
        kept.keep();
        CHECK(kept.owns_ranks());
        (void)g.intersect_batch(qc.data(), qe.data(), qe.data(), nq);  // overwrites the library's pinned buffer
        bool ok = true;
        for (std::size_t i = 0; i < nq; i += 3) ok = ok && kept.get_keys(i) == got[i];
        CHECK(ok);
        // chained on the temporary: the result is moved out, its ranks are its own
        auto chained = g.intersect_batch(qc.data(), qs.data(), qe.data(), nq).keep();
        CHECK(chained.owns_ranks());
        // copies of a kept result own their ranks too (copy construction, copy assignment, move)
        auto copy = chained;
        CHECK(copy.owns_ranks() && copy.ranks().data() != chained.ranks().data());
        decltype(chained) assigned;
        assigned = chained;
        CHECK(assigned.owns_ranks() && assigned.ranks().data() != chained.ranks().data());
        auto moved = std::move(chained);
        CHECK(moved.owns_ranks());
        (void)g.intersect_batch(qc.data(), qe.data(), qe.data(), nq);  // overwrites the library's pinned buffer again
        bool ok2 = true;
        for (std::size_t i = 0; i < nq; i += 3)
            ok2 = ok2 && copy.get_keys(i) == got[i] && assigned.get_keys(i) == got[i] && moved.get_keys(i) == got[i];
        CHECK(ok2);
        // a copy of a plain view stays a view of the library's buffer
        auto view = g.intersect_batch(qc.data(), qs.data(), qe.data(), nq);
        auto view2 = view;
        CHECK(!view2.owns_ranks() && view2.ranks().data() == view.ranks().data());
    }
    {   // 32-bit coordinates: the u32 call and the compact result form give the same keys in the same order
        std::vector<std::uint32_t> qs32(qs.begin(), qs.end()), qe32(qe.begin(), qe.end());
        auto b32 = g.intersect_batch(qc.data(), qs32.data(), qe32.data(), nq);
        CHECK(b32.size() == nq && b32.total_hits() == br.total_hits());
        bool ok = true;
        for (std::size_t i = 0; i < nq; ++i) ok = ok && b32.get_keys(i) == got[i];
        CHECK(ok);
        auto cr = g.intersect_batch_compact(qc.data(), qs32.data(), qe32.data(), nq);
        CHECK(cr.size() == nq && cr.total_hits() == br.total_hits());
        bool okc = true;
        for (std::size_t i = 0; i < nq; ++i) okc = okc && cr.count(i) == got[i].size() && cr.get_keys(i) == got[i];
        CHECK(okc);
        auto ex = cr.expand();
        CHECK(ex.owns_ranks() && ex.total_hits() == br.total_hits());
        (void)g.intersect_batch(qc.data(), qe.data(), qe.data(), nq);  // the expanded CSR owns its ranks
        bool oke = true;
        for (std::size_t i = 0; i < nq; ++i) oke = oke && ex.get_keys(i) == got[i];
        CHECK(oke);
    }
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

// Cost model of the snapshot ("a dirty flag per chromosome"): a multi-index bulk load is ONE device
// build, a mutation of one index re-uploads that index only (gg_index_update_chrom), index ids never move.
static void test_snapshot_cost_model() {
    using interval = facade_api::interval;
    namespace gst = genogrove_b200::structure;
    using grove_t = gst::grove<interval, int>;
    std::mt19937_64 rng(99);
    const int n_idx = 24;
    std::vector<std::pair<std::string, std::vector<std::pair<interval, int>>>> load(n_idx);
    for (int c = 0; c < n_idx; ++c) {
        load[c].first = "chr" + std::to_string(c + 1);
        const int m = 2000 + 300 * c;
        for (int i = 0; i < m; ++i) {
            const std::size_t s = (rng() % 400000) / 16 * 16;  // duplicates on purpose
            load[c].second.push_back({interval(s, s + (rng() % 64) * 64), c * 100000 + i});
        }
    }
    grove_t g(8), per_index(8);
    auto keys = g.insert_data(load, gst::bulk);
    CHECK(keys.size() == static_cast<std::size_t>(n_idx) && g.device_build_count() == 1);
    for (auto& [name, v] : load) per_index.insert_data(name, v, gst::bulk);
    CHECK(per_index.device_build_count() == static_cast<std::size_t>(n_idx));
    auto same = [&](grove_t& a, grove_t& b, const interval& q, const std::string& name) {
        auto ra = a.intersect(q, name), rb = b.intersect(q, name);
        if (ra.get_keys().size() != rb.get_keys().size()) return false;
        for (std::size_t i = 0; i < ra.get_keys().size(); ++i) {
            const auto *ka = ra.get_keys()[i], *kb = rb.get_keys()[i];
            if (!(ka->get_value() == kb->get_value()) || ka->get_data() != kb->get_data()) return false;
        }
        return true;
    };
    auto sweep = [&](grove_t& a, grove_t& b) {
        bool ok = true;
        for (int t = 0; t < 300; ++t) {
            const std::size_t s = rng() % 400000;
            ok = ok && same(a, b, interval(s, s + rng() % 5000), load[rng() % n_idx].first);
        }
        return ok;
    };
    CHECK(sweep(g, per_index));
    CHECK(g.device_build_count() == 1 && g.full_upload_count() == 0 && g.partial_update_count() == 0);  // the build's snapshot served the queries
    CHECK(per_index.full_upload_count() == 1);
    // sorted key order came back per index
    for (int c = 0; c < n_idx; ++c) {
        CHECK(keys[c].size() == load[c].second.size());
        for (std::size_t i = 1; i < keys[c].size(); ++i) CHECK(!(keys[c][i]->get_value() < keys[c][i - 1]->get_value()));
    }

    // a mutation of chr2 between two chr1 batches: only chr2 goes up again, ids stay put
    const std::uint32_t id1 = g.index_id("chr1"), id2 = g.index_id("chr2"), id24 = g.index_id("chr24");
    CHECK(id1 == 0 && id2 == 1 && id24 == 23);
    std::vector<std::uint32_t> qc(500, id1);
    std::vector<std::uint64_t> qs(500), qe(500);
    for (std::size_t i = 0; i < qs.size(); ++i) {
        qs[i] = rng() % 400000;
        qe[i] = qs[i] + rng() % 3000;
    }
    std::vector<std::vector<grove_t::key_ptr>> before(qs.size());
    {
        auto br = g.intersect_batch(qc.data(), qs.data(), qe.data(), qs.size());
        for (std::size_t i = 0; i < qs.size(); ++i) before[i] = br.get_keys(i);
    }
    for (int i = 0; i < 50; ++i) {
        const std::size_t s = (rng() % 400000) / 16 * 16;
        g.insert_data("chr2", interval(s, s + 640), 7000000 + i);
        per_index.insert_data("chr2", interval(s, s + 640), 7000000 + i);
    }
    {
        auto br = g.intersect_batch(qc.data(), qs.data(), qe.data(), qs.size());
        bool ok = true;
        for (std::size_t i = 0; i < qs.size(); ++i) ok = ok && br.get_keys(i) == before[i];
        CHECK(ok);
    }
    CHECK(g.partial_update_count() == 1 && g.full_upload_count() == 0);
    gg_index_stats st{};
    CHECK(gg_index_get_stats(g.device_handle(), &st) == GG_OK);
    CHECK(st.h2d_bytes == 16 * g.chain("chr2").size());
    CHECK(g.index_id("chr1") == id1 && g.index_id("chr2") == id2 && g.index_id("chr24") == id24);
    CHECK(sweep(g, per_index));

    // a new index is appended (its id is the next one), an emptied index retires its id; nobody else moves
    g.insert_data("chrM", interval(100, 200), 1);
    per_index.insert_data("chrM", interval(100, 200), 1);
    CHECK(g.index_id("chrM") == 24 && g.index_id("chr24") == id24);
    CHECK(g.intersect(interval(150, 160), "chrM").get_keys().size() == 1);
    auto victim = g.intersect(interval(0, 1000000), "chr3").get_keys();
    auto victim2 = per_index.intersect(interval(0, 1000000), "chr3").get_keys();
    CHECK(victim.size() == load[2].second.size());
    for (auto* k : victim) CHECK(g.remove_key("chr3", k));
    for (auto* k : victim2) CHECK(per_index.remove_key("chr3", k));
    CHECK(g.index_id("chr3") == grove_t::npos_index && g.index_id("chr4") == 3 && g.index_id("chrM") == 24);
    CHECK(g.intersect(interval(0, 1000000), "chr3").get_keys().empty());
    CHECK(sweep(g, per_index));
    g.insert_data("chr3", interval(5, 6), 2);  // the name comes back under a NEW id
    CHECK(g.index_id("chr3") == 25 && g.intersect(interval(0, 10), "chr3").get_keys().size() == 1);
    CHECK(g.full_upload_count() == 0);
    // many indices changed at once -> one whole upload instead of a rebuild per index
    for (int c = 5; c < 15; ++c) g.insert_data(load[c].first, interval(1, 2), -c);
    CHECK(g.intersect(interval(0, 3), "chr6").get_keys().size() >= 1);
    CHECK(g.full_upload_count() == 1);

    // the handle dropped its build workspace (no chains on the device any more): the next mutation falls back to one
    // whole upload instead of failing
    CHECK(gg_set_option(g.device_handle(), "keep_build_workspace", 0) == GG_OK);
    const std::size_t full_before = g.full_upload_count(), partial_before = g.partial_update_count();
    g.insert_data("chr1", interval(77, 88), 4242);
    per_index.insert_data("chr1", interval(77, 88), 4242);
    CHECK(sweep(g, per_index));
    CHECK(g.full_upload_count() == full_before + 1 && g.partial_update_count() == partial_before);
    CHECK(gg_set_option(g.device_handle(), "keep_build_workspace", 1) == GG_OK);
    g.insert_data("chr2", interval(78, 89), 4243);  // (that upload ran without a kept workspace: one more whole upload ...)
    per_index.insert_data("chr2", interval(78, 89), 4243);
    CHECK(sweep(g, per_index));
    CHECK(g.full_upload_count() == full_before + 2 && g.partial_update_count() == partial_before);
    g.insert_data("chr2", interval(79, 90), 4244);  // (... which leaves the chains on the device: single-index updates again)
    per_index.insert_data("chr2", interval(79, 90), 4244);
    CHECK(sweep(g, per_index));
    CHECK(g.full_upload_count() == full_before + 2 && g.partial_update_count() == partial_before + 1);

    // genomic_coordinate keys: the multi-index build leaves a usable stranded snapshot too
    using gcoord = facade_api::genomic_coordinate;
    gst::grove<gcoord, int> sg(5), sg_ref(5);
    std::vector<std::pair<std::string, std::vector<std::pair<gcoord, int>>>> sload(3);
    const char strands[4] = {'+', '-', '.', '*'};
    for (int c = 0; c < 3; ++c) {
        sload[c].first = "s" + std::to_string(c);
        for (int i = 0; i < 3000; ++i) {
            const std::size_t s = (rng() % 100000) / 8 * 8;
            sload[c].second.push_back({gcoord(strands[rng() % 4], s, s + (rng() % 16) * 32), c * 10000 + i});
        }
    }
    sg.insert_data(sload, gst::bulk);
    for (auto& [name, v] : sload) sg_ref.insert_data(name, v, gst::bulk);
    bool ok = true;
    for (int t = 0; t < 400; ++t) {
        const std::size_t s = rng() % 100000;
        const gcoord q(strands[rng() % 4], s, s + rng() % 2000);
        const std::string& name = sload[rng() % 3].first;
        auto ra = sg.intersect(q, name), rb = sg_ref.intersect(q, name);
        ok = ok && ra.get_keys().size() == rb.get_keys().size();
        for (std::size_t i = 0; ok && i < ra.get_keys().size(); ++i)
            ok = ra.get_keys()[i]->get_value() == rb.get_keys()[i]->get_value() && ra.get_keys()[i]->get_data() == rb.get_keys()[i]->get_data();
    }
    CHECK(ok);
    CHECK(sg.device_build_count() == 1 && sg.full_upload_count() == 0);
}

// ------------------------------------------------------------------------------------------- .gg fast path
// a serialized grove (format 0.3) goes to the device through genogrove_b200::io::load_gg, which
// reads the leaf chain out of the blocks without building a pointer tree.  A "script" is a list of text lines:
//   Q <index|*> <strand> <start> <end> | <payload> ...   query and the reference's answer, in the reference's order
//   A <index> <strand> <start> <end> <payload>            sorted single append applied AFTER loading
//   I <index> <strand> <start> <end> <payload>            regular insert applied AFTER loading
//   R <index> <payload>                                   remove_key of the key inserted with that payload
//   S|U <index> <n> n x (<strand> <start> <end> <payload>) sorted / unsorted bulk insert of a run
//   T <index> L<keys>|N<children> ...                     the reference's real tree in pre-order (shape check)
// The reference writes the stream and the script (live in ref_parity.bin; committed under tests/golden/gg/ by
// `ref_parity.bin --write-gg-fixtures DIR`); replay() runs the script against the loaded facade grove.
namespace ggfast {
namespace fio = genogrove_b200::io;

template <typename K>
K make_value(char strand, std::size_t s, std::size_t e) {
    if constexpr (requires { K('+', s, e); }) return K(strand, s, e);
    else return K(s, e);
}
template <typename K>
char strand_of(const K& k) {
    if constexpr (requires { k.get_strand(); }) return k.get_strand();
    else return '.';
}
template <typename D>
std::string show(const D& d) {
    if constexpr (std::is_same_v<D, std::string>) return d;
    else return std::to_string(d);
}
template <typename D>
D parse_payload(const std::string& s) {
    if constexpr (std::is_same_v<D, std::string>) return s;
    else return static_cast<D>(std::stoll(s));
}

template <typename G, typename K>
std::string answer_line(G& g, const std::string& index, const K& q) {
    std::ostringstream os;
    os << "Q " << index << ' ' << strand_of(q) << ' ' << q.get_start() << ' ' << q.get_end() << " |";
    auto r = index == "*" ? g.intersect(q) : g.intersect(q, index);
    for (auto* k : r.get_keys()) os << ' ' << show(k->get_data());
    return os.str();
}

/// run a script against the facade grove restored from `stream`; returns the number of Q lines checked
/// run a script against a facade grove: restored from `stream`, or empty (order given) when the stream is empty
/// host_only: a detached grove (no device) — only the lines that need no overlap computed are checked: tree shapes,
/// and queries wide enough to return the whole chain, answered from chain()
template <typename K, typename D>
std::size_t replay(const std::string& stream, const std::vector<std::string>& script, fio::gg_stream_info* info = nullptr, int order = 3,
                   bool host_only = false) {
    using grove_t = genogrove_b200::structure::grove<K, D>;
    std::istringstream is(stream, std::ios::binary);
    grove_t g = host_only ? grove_t(order, genogrove_b200::structure::detached)
                          : (stream.empty() ? grove_t(order) : fio::load_gg<K, D>(is, 0, info));
    std::unordered_map<std::string, typename grove_t::key_ptr> by_payload;  // R lines name keys by payload
    std::size_t checked = 0, bad = 0;
    for (const std::string& line : script) {
        std::istringstream ls(line);
        std::string op, index;
        ls >> op >> index;
        if (op == "T") {  // the reference's real tree, pre-order: 'L<keys>' / 'N<children>'
            std::ostringstream os;
            os << "T " << index;
            for (const std::uint32_t t : g.shape_tokens(index)) {
                if ((t & 0x80000000u) != 0) os << " L" << (t & 0x7FFFFFFFu);
                else os << " N" << t;
            }
            if (os.str() != line && bad++ < 5) std::fprintf(stderr, "tree shape differs\n  reference: %.300s\n  facade   : %.300s\n", line.c_str(), os.str().c_str());
            ++checked;
            continue;
        }
        if (op == "R") {
            std::string payload;
            ls >> payload;
            auto it = by_payload.find(index + "/" + payload);
            CHECK(it != by_payload.end());
            if (it != by_payload.end()) {
                CHECK(g.remove_key(index, it->second));
                by_payload.erase(it);
            }
            continue;
        }
        if (op == "S" || op == "U") {  // sorted / unsorted bulk run: count, then count x (strand start end payload)
            std::size_t count;
            ls >> count;
            std::vector<std::pair<K, D>> run;
            std::vector<std::string> names;
            for (std::size_t i = 0; i < count; ++i) {
                char strand;
                std::size_t s, e;
                std::string payload;
                ls >> strand >> s >> e >> payload;
                run.push_back({make_value<K>(strand, s, e), parse_payload<D>(payload)});
            }
            auto keys = op == "S" ? g.insert_data(index, run, genogrove_b200::structure::sorted, genogrove_b200::structure::bulk)
                                  : g.insert_data(index, run, genogrove_b200::structure::bulk);
            for (auto* k : keys) by_payload[index + "/" + show(k->get_data())] = k;
            continue;
        }
        char strand;
        std::size_t s, e;
        ls >> strand >> s >> e;
        if (op == "Q" && host_only) {
            if (s != 0 || e < (1u << 30)) continue;
            std::ostringstream os;
            os << "Q " << index << ' ' << strand << ' ' << s << ' ' << e << " |";
            for (auto* k : g.chain(index)) os << ' ' << show(k->get_data());
            if (os.str() != line && bad++ < 5) std::fprintf(stderr, "chain order differs\n  reference: %.300s\n  facade   : %.300s\n", line.c_str(), os.str().c_str());
            ++checked;
        } else if (op == "Q") {
            const std::string got = answer_line(g, index, make_value<K>(strand, s, e));
            if (got != line && bad++ < 5) std::fprintf(stderr, "hit order differs\n  reference: %.300s\n  facade   : %.300s\n", line.c_str(), got.c_str());
            ++checked;
        } else {
            std::string payload;
            ls >> payload;
            auto* k = op == "A" ? g.insert_data(index, make_value<K>(strand, s, e), parse_payload<D>(payload), genogrove_b200::structure::sorted)
                                : g.insert_data(index, make_value<K>(strand, s, e), parse_payload<D>(payload));
            by_payload[index + "/" + payload] = k;
        }
    }
    CHECK(bad == 0);
    CHECK(checked > 0);
    return checked;
}

static std::string slurp(const std::string& path) {
    std::FILE* f = std::fopen(path.c_str(), "rb");
    if (f == nullptr) return {};
    std::string out;
    char buf[1 << 16];
    for (std::size_t n; (n = std::fread(buf, 1, sizeof(buf), f)) > 0;) out.append(buf, n);
    std::fclose(f);
    return out;
}
static std::vector<std::string> lines_of(const std::string& text) {
    std::vector<std::string> v;
    std::istringstream is(text);
    for (std::string l; std::getline(is, l);)
        if (!l.empty()) v.push_back(l);
    return v;
}

/// committed fixtures: streams written by the reference itself + its answers (tests/golden/gg/)
static void test_fixtures(const std::string& dir) {
    namespace fdt = genogrove_b200::data_type;
    struct fx { const char* stem; int kind; };
    std::size_t total = 0;
    for (const fx& f : {fx{"interval_int_o3", 0}, fx{"interval_int_o32", 0}, fx{"interval_str_o5", 1}, fx{"gcoord_int_o4", 2}, fx{"gcoord_str_o16", 3}}) {
        const std::string stream = slurp(dir + "/" + f.stem + ".gg"), script = slurp(dir + "/" + f.stem + ".script");
        CHECK(!stream.empty() && !script.empty());
        if (stream.empty() || script.empty()) {
            std::fprintf(stderr, "missing fixture %s/%s.{gg,script}\n", dir.c_str(), f.stem);
            continue;
        }
        const auto lines = lines_of(script);
        if (f.kind == 0) total += replay<fdt::interval, int>(stream, lines);
        if (f.kind == 1) total += replay<fdt::interval, std::string>(stream, lines);
        if (f.kind == 2) total += replay<fdt::genomic_coordinate, int>(stream, lines);
        if (f.kind == 3) total += replay<fdt::genomic_coordinate, std::string>(stream, lines);
    }
    std::printf("gg fast path: %zu fixture queries identical to the reference's answers\n", total);
    // mixed insert histories from an EMPTY grove (row f1): hit order AND tree shape after every step
    std::size_t mixed = 0;
    for (const char* stem : {"mixed_o3", "mixed_o4", "mixed_o5", "mixed_o8"}) {
        const std::string script = slurp(dir + "/" + stem + ".script");
        CHECK(!script.empty());
        if (script.empty()) continue;
        mixed += replay<fdt::interval, int>(std::string(), lines_of(script), nullptr, std::atoi(stem + 7));
    }
    std::printf("mixed histories: %zu fixture lines (hit order + tree shape) identical to the reference\n", mixed);
    // malformed input is reported the way the reference reports it: std::runtime_error
    const std::string good = slurp(dir + "/interval_int_o3.gg");
    if (!good.empty()) {
        auto load = [](const std::string& bytes) {
            std::istringstream is(bytes, std::ios::binary);
            return fio::load_gg<fdt::interval, int>(is).size();
        };
        CHECK_THROWS(load("GGB\x02" + good.substr(4)), std::runtime_error);       // wrong format version in the magic
        CHECK_THROWS(load(good.substr(0, good.size() / 2)), std::runtime_error);  // truncated
        CHECK_THROWS(load(std::string()), std::runtime_error);
        std::string flipped = good;
        flipped[flipped.size() - 9] = static_cast<char>(flipped[flipped.size() - 9] ^ 0x5a);  // inside the last deflate stream
        CHECK_THROWS(load(flipped), std::runtime_error);
        // the CLI's 12-byte wrapper (io/gg_format.hpp:24-40) in front of the same stream is skipped
        const std::string wrapped = std::string("GROV\x00\x03\x00\x01\x00\x01\x00\x00", 12) + good;
        fio::gg_stream_info info;
        std::istringstream is(wrapped, std::ios::binary);
        auto g = fio::load_gg<fdt::interval, int>(is, 0, &info);
        CHECK(info.cli_wrapper && info.payload_type == 1);
        CHECK(g.size() == info.leaf_keys && g.size() > 0);
        CHECK(g.get_order() == 3);
    }
}
}  // namespace ggfast

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
// ------------------------------------------------------------------------------------------- .gg fast path, reference side
// The reference builds groves through every insert path, serializes them, restores them with ITS deserialize()
// and answers the script; the facade restores the same bytes through load_gg (no pointer tree) and must give the
// same payload sequences — including after sorted appends of tied keys, which depend on the key count of the
// rightmost leaf in the file.
namespace ggfast {
namespace rdt = genogrove::data_type;
namespace rst = genogrove::structure;

template <typename RK, typename D>
std::vector<std::string> write_script(int order, std::uint64_t seed, int n, std::string& stream_out) {
    std::mt19937_64 rng(seed);
    rst::grove<RK, D> g(order);
    int next_id = 0;
    auto payload = [&]() -> D {
        const int v = next_id++;
        if constexpr (std::is_same_v<D, std::string>) return (v % 17 == 0 ? std::string() : "p" + std::to_string(v) + std::string(v % 5, 'x'));
        else return static_cast<D>(v);
    };
    const char sym[4] = {'+', '-', '.', '*'};
    auto value = [&](bool dup) {
        std::size_t s = rng() % 300000, len = rng() % 40 == 0 ? 50000 : rng() % 500;
        if (dup) s = s / 1000 * 1000, len = 100;
        return make_value<RK>(sym[rng() % 4], s, s + len);
    };
    // chr1: regular inserts (ties resolved by insertion history) and removals (leaf merges / borrows)
    std::vector<rdt::key<RK, D>*> ins;
    for (int i = 0; i < n; ++i) ins.push_back(g.insert_data("chr1", value(i % 7 == 0), payload()));
    for (int i = 0; i < n; i += 5) g.remove_key("chr1", ins[static_cast<std::size_t>(i)]);
    // chr2: unsorted bulk — the reference's std::sort leaves equal keys in an order only the stream records
    std::vector<std::pair<RK, D>> v;
    for (int i = 0; i < n; ++i) v.push_back({value(i % 11 == 0), payload()});
    g.insert_data("chr2", v, rst::bulk);
    // chrX: sorted bulk, then sorted single appends of the maximum (rightmost leaf fills and splits)
    v.clear();
    for (int i = 0; i < n; ++i) v.push_back({value(i % 9 == 0), payload()});
    std::stable_sort(v.begin(), v.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
    g.insert_data("chrX", v, rst::sorted, rst::bulk);
    const RK top = v.back().first;
    for (int j = 0; j < order / 2 + 1; ++j) g.insert_data("chrX", top, payload(), rst::sorted);
    // scaffold_7: a root that is still a leaf; "gone": created and emptied again
    g.insert_data("scaffold_7", make_value<RK>('+', 5, 9), payload(), rst::sorted);
    g.insert_data("scaffold_7", make_value<RK>('-', 7, 70), payload(), rst::sorted);
    auto* lone = g.insert_data("gone", make_value<RK>('.', 1, 2), payload());
    g.remove_key("gone", lone);

    std::ostringstream os(std::ios::binary);
    g.serialize(os);
    stream_out = os.str();
    std::istringstream is(stream_out, std::ios::binary);
    auto r = rst::grove<RK, D>::deserialize(is);

    std::vector<std::string> script;
    const char* names[] = {"chr1", "chr2", "chrX", "scaffold_7", "gone", "never", "*"};
    auto ask = [&](int count) {
        for (int k = 0; k < count; ++k) {
            const std::size_t a = rng() % 310000, b = a + (rng() % 25 == 0 ? rng() % 60000 : rng() % 400);
            script.push_back(answer_line(r, names[rng() % 7], make_value<RK>(sym[rng() % 4], k % 50 == 0 ? 0 : a, k % 50 == 0 ? 400000 : b)));
        }
    };
    ask(150);
    // sorted appends of tied keys on the restored grove: more than one leaf's worth, on the index built by
    // regular inserts + removals too (its rightmost-leaf key count is only known from the stream)
    auto apply = [&](const char* op, const char* index, const RK& k) {
        const D d = payload();
        std::ostringstream ls;
        ls << op << ' ' << index << ' ' << strand_of(k) << ' ' << k.get_start() << ' ' << k.get_end() << ' ' << (show(d).empty() ? std::string("-") : show(d));
        if (op[0] == 'A') r.insert_data(index, k, show(d).empty() ? parse_payload<D>("-") : d, rst::sorted);
        else r.insert_data(index, k, show(d).empty() ? parse_payload<D>("-") : d);
        script.push_back(ls.str());
    };
    const RK far1 = make_value<RK>('+', 400000, 400100), far2 = make_value<RK>('+', 400000, 400200);
    for (int j = 0; j < 2 * order + 1; ++j) {
        apply("A", "chr1", far1);
        apply("A", "chrX", j < order ? far1 : far2);
        if (j % 3 == 0) apply("A", "chr2", far1);
    }
    apply("A", "fresh", far1);
    ask(120);
    script.push_back(answer_line(r, "chr1", make_value<RK>('*', 399999, 400001)));
    script.push_back(answer_line(r, "chrX", make_value<RK>('*', 399999, 400001)));
    script.push_back(answer_line(r, "chr2", make_value<RK>('*', 399999, 400001)));
    // regular inserts after that land in front of every equal key
    for (int j = 0; j < order + 2; ++j) {
        apply("I", "chr1", far1);
        apply("I", "chr2", value(true));
    }
    ask(80);
    script.push_back(answer_line(r, "chr1", make_value<RK>('*', 399999, 400001)));
    return script;
}

/// the reference's real tree of one index in pre-order
template <typename G>
std::string tree_line(const G& g, const std::string& index) {
    std::ostringstream os;
    os << "T " << index;
    auto it = g.get_root_nodes().find(index);
    if (it == g.get_root_nodes().end()) return os.str();
    auto walk = [&](auto* n, auto&& self) -> void {
        if (n->get_is_leaf()) {
            os << " L" << n->get_keys().size();
            return;
        }
        os << " N" << n->get_children().size();
        for (auto* c : n->get_children()) self(c, self);
    };
    walk(it->second, walk);
    return os.str();
}

/// Random MIXED history on an empty reference grove — regular inserts, sorted appends of the running maximum,
/// sorted bulk runs, removals, all over a handful of coordinates so that most keys are tied — recording after every
/// few steps the whole chain (one wide query) and the real tree's shape.
static std::vector<std::string> write_mixed_script(int order, std::uint64_t seed, int steps) {
    std::mt19937_64 rng(seed);
    rst::grove<rdt::interval, int> g(order);
    std::vector<std::string> script;
    const char* names[] = {"a", "b"};
    struct live { std::vector<std::pair<rdt::key<rdt::interval, int>*, int>> keys; std::size_t top = 0; };
    live state[2];
    int next_id = 0;
    auto line = [&](const char* op, const char* index, std::size_t s, std::size_t e, int id) {
        std::ostringstream os;
        os << op << ' ' << index << " . " << s << ' ' << e << ' ' << id;
        script.push_back(os.str());
    };
    for (int step = 0; step < steps; ++step) {
        const int c = static_cast<int>(rng() % 2);
        live& st = state[c];
        const unsigned dice = static_cast<unsigned>(rng() % 100);
        if (dice < 35 || st.keys.empty()) {  // regular insert; coordinates on a coarse grid -> many ties, some above the top
            const std::size_t s = (rng() % (st.top / 10 + 3)) * 10;
            const int id = next_id++;
            st.keys.push_back({g.insert_data(names[c], rdt::interval(s, s + 5), id), id});
            st.top = std::max(st.top, s);
            line("I", names[c], s, s + 5, id);
        } else if (dice < 65) {  // sorted append: the running maximum again, or the next grid point
            const std::size_t s = rng() % 3 == 0 ? st.top + 10 : st.top;
            const int id = next_id++;
            st.keys.push_back({g.insert_data(names[c], rdt::interval(s, s + 5), id, rst::sorted), id});
            st.top = s;
            line("A", names[c], s, s + 5, id);
        } else if (dice < 72) {  // sorted bulk run onto the tail
            const std::size_t count = 1 + rng() % (2 * static_cast<std::size_t>(order));
            std::vector<std::pair<rdt::interval, int>> run;
            std::ostringstream os;
            os << "S " << names[c] << ' ' << count;
            std::size_t s = st.top;
            for (std::size_t i = 0; i < count; ++i) {
                if (rng() % 3 == 0) s += 10;
                run.push_back({rdt::interval(s, s + 5), next_id});
                os << " . " << s << ' ' << s + 5 << ' ' << next_id++;
            }
            auto keys = g.insert_data(names[c], run, rst::sorted, rst::bulk);
            for (std::size_t i = 0; i < keys.size(); ++i) st.keys.push_back({keys[i], run[i].second});
            st.top = s;
            script.push_back(os.str());
        } else {  // removal of a random live key (sometimes draining most of the index)
            const int burst = dice > 97 ? static_cast<int>(st.keys.size() * 3 / 4) : 1;
            for (int b = 0; b < burst && !st.keys.empty(); ++b) {
                const std::size_t at = rng() % st.keys.size();
                g.remove_key(names[c], st.keys[at].first);
                std::ostringstream os;
                os << "R " << names[c] << ' ' << st.keys[at].second;
                script.push_back(os.str());
                st.keys[at] = st.keys.back();
                st.keys.pop_back();
            }
            if (st.keys.empty()) st.top = 0;
        }
        if (step % 3 == 2 || step + 1 == steps) {
            script.push_back(answer_line(g, names[c], rdt::interval(0, 1u << 30)));
            script.push_back(tree_line(g, names[c]));
        }
    }
    return script;
}

template <typename RK, typename FK, typename D>
void live_case(int order, std::uint64_t seed, int n) {
    std::string stream;
    const auto script = write_script<RK, D>(order, seed, n, stream);
    replay<FK, D>(stream, script);
}

/// row f1: mixed histories, many seeds, small orders (deep trees, constant splits / borrows / merges)
static void test_mixed_live(bool host_only) {
    namespace fdt = genogrove_b200::data_type;
    std::size_t lines = 0;
    for (int order : {3, 4, 5, 6, 7, 8, 16})
        for (std::uint64_t seed = 1; seed <= 6; ++seed)
            lines += replay<fdt::interval, int>(std::string(), write_mixed_script(order, 1000 * static_cast<std::uint64_t>(order) + seed, 1500), nullptr, order, host_only);
    std::printf("mixed histories%s: %zu lines (%s + tree shape) checked against the live reference\n", host_only ? " (host bookkeeping only)" : "", lines,
                host_only ? "chain order" : "hit order");
}

static void test_live() {
    namespace fdt = genogrove_b200::data_type;
    test_mixed_live(false);
    for (int order : {3, 4, 7, 50}) {
        live_case<rdt::interval, fdt::interval, int>(order, 100 + order, 700);
        live_case<rdt::genomic_coordinate, fdt::genomic_coordinate, int>(order, 200 + order, 500);
    }
    live_case<rdt::interval, fdt::interval, std::string>(5, 31, 400);
    live_case<rdt::genomic_coordinate, fdt::genomic_coordinate, std::string>(16, 32, 400);

    // load cost: reference deserialize() + leaf-chain walk + upload against load_gg + upload
    using clock = std::chrono::steady_clock;
    auto ms = [](auto a, auto b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
    for (int order : {8, 128}) {
        const std::size_t n = 1000000;
        std::mt19937_64 rng(5);
        std::vector<std::pair<rdt::interval, int>> v;
        v.reserve(n);
        std::size_t pos = 0;
        for (std::size_t i = 0; i < n; ++i) {
            pos += rng() % 200;
            v.push_back({rdt::interval(pos, pos + rng() % 3000), static_cast<int>(i)});
        }
        rst::grove<rdt::interval, int> g(order);
        g.insert_data("chr1", v, rst::sorted, rst::bulk);
        std::ostringstream os(std::ios::binary);
        g.serialize(os);
        const std::string stream = os.str();
        const auto t0 = clock::now();
        std::istringstream is(stream, std::ios::binary);
        auto restored = rst::grove<rdt::interval, int>::deserialize(is);
        const auto t1 = clock::now();
        binding::device_snapshot<int> snap(restored);
        const auto t2 = clock::now();
        fio::gg_stream_info info;
        std::istringstream is2(stream, std::ios::binary);
        auto fast = fio::load_gg<fdt::interval, int>(is2, 0, &info);
        const auto t3 = clock::now();
        gg_index* h = fast.device_handle();
        const auto t4 = clock::now();
        CHECK(h != nullptr && fast.size() == n);
        // same device content: one batch over both handles
        const std::size_t nq = 20000;
        std::vector<std::uint32_t> qc(nq, 0);
        std::vector<std::uint64_t> qs(nq), qe(nq), off_a(nq + 1), off_b(nq + 1);
        for (std::size_t i = 0; i < nq; ++i) qs[i] = rng() % (pos + 1000), qe[i] = qs[i] + rng() % 5000;
        const std::uint32_t *ha = nullptr, *hb = nullptr;
        std::uint64_t na = 0, nb = 0;
        CHECK(gg_intersect_batch(snap.handle, qc.data(), qs.data(), qe.data(), nq, off_a.data(), &ha, &na) == GG_OK);
        std::vector<std::uint32_t> keep(ha, ha + na);
        CHECK(gg_intersect_batch(h, qc.data(), qs.data(), qe.data(), nq, off_b.data(), &hb, &nb) == GG_OK);
        CHECK(na == nb && off_a == off_b && std::equal(keep.begin(), keep.end(), hb));
        bool payloads_equal = true;
        const auto& chain = fast.chain("chr1");
        for (std::size_t i = 0; i < n; i += 997) payloads_equal = payloads_equal && chain[i]->get_data() == snap.rank_to_key[0][i]->get_data();
        CHECK(payloads_equal);
        std::printf("gg load, 1M keys order %d (%zu blocks, %.1f MB deflated): reference deserialize %.0f ms + chain walk/upload %.0f ms;"
                    " load_gg %.0f ms (read %.0f, inflate+parse %.0f on %u threads, adopt %.0f) + upload %.0f ms\n",
                    order, static_cast<std::size_t>(info.ext_block_begin), stream.size() / 1e6, ms(t0, t1), ms(t1, t2), ms(t2, t3), info.read_ms,
                    info.parse_ms, info.threads, info.adopt_ms, ms(t3, t4));
    }
}

static int write_fixtures(const std::string& dir) {
    auto put = [&](const char* stem, const std::string& stream, const std::vector<std::string>& script) {
        std::FILE* f = std::fopen((dir + "/" + stem + ".gg").c_str(), "wb");
        if (f == nullptr) return false;
        std::fwrite(stream.data(), 1, stream.size(), f);
        std::fclose(f);
        f = std::fopen((dir + "/" + stem + ".script").c_str(), "w");
        if (f == nullptr) return false;
        for (const auto& l : script) std::fprintf(f, "%s\n", l.c_str());
        std::fclose(f);
        std::printf("%s: %zu stream bytes, %zu script lines\n", stem, stream.size(), script.size());
        return true;
    };
    std::string stream;
    bool ok = true;
    for (int order : {3, 4, 5, 8}) {
        const std::string stem = "mixed_o" + std::to_string(order);
        std::FILE* f = std::fopen((dir + "/" + stem + ".script").c_str(), "w");
        if (f == nullptr) return 1;
        const auto script = write_mixed_script(order, 77 + static_cast<std::uint64_t>(order), 400);
        for (const auto& l : script) std::fprintf(f, "%s\n", l.c_str());
        std::fclose(f);
        std::printf("%s: %zu script lines\n", stem.c_str(), script.size());
    }
    auto s1 = write_script<rdt::interval, int>(3, 11, 120, stream);
    ok = ok && put("interval_int_o3", stream, s1);
    auto s2 = write_script<rdt::interval, int>(32, 12, 400, stream);
    ok = ok && put("interval_int_o32", stream, s2);
    auto s3 = write_script<rdt::interval, std::string>(5, 13, 150, stream);
    ok = ok && put("interval_str_o5", stream, s3);
    auto s4 = write_script<rdt::genomic_coordinate, int>(4, 14, 150, stream);
    ok = ok && put("gcoord_int_o4", stream, s4);
    auto s5 = write_script<rdt::genomic_coordinate, std::string>(16, 15, 300, stream);
    ok = ok && put("gcoord_str_o16", stream, s5);
    return ok ? 0 : 1;
}
}  // namespace ggfast
#endif

static std::string golden_gg_dir() {
    if (const char* env = std::getenv("GG_GOLDEN_DIR")) return env;
    char buf[4096];
    const ssize_t n = readlink("/proc/self/exe", buf, sizeof(buf) - 1);
    std::string exe = n > 0 ? std::string(buf, static_cast<std::size_t>(n)) : std::string("tests/cpp/x");
    return exe.substr(0, exe.rfind('/')) + "/../golden/gg";
}

int main(int argc, char** argv) {
#ifdef WITH_REFERENCE
    if (argc == 3 && std::string(argv[1]) == "--write-gg-fixtures") return ggfast::write_fixtures(argv[2]);
#endif
    (void)argc;
    (void)argv;
    if (gg_device_count() < 1) {
#ifdef WITH_REFERENCE
        // no GPU: only check that the scenarios' own expectations hold on the unmodified reference
        trace_t ref_only = run_scenarios<reference_api>();
        ggfast::test_mixed_live(/*host_only=*/true);
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
    test_snapshot_cost_model();
    ggfast::test_fixtures(golden_gg_dir());
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
    ggfast::test_live();
#endif
    std::printf("%d checks, %d failures\n", g_checks, g_failures);
    return g_failures == 0 ? 0 : 1;
}
