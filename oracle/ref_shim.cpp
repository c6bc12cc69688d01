// ref_shim.cpp — TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// A thin extern "C" driver around the UNMODIFIED reference library: it #includes the reference's
// own headers from /root/reference/include (never copied into this repo) and instantiates
// genogrove::structure::grove<interval, uint32_t>, payload = caller-supplied id.  Built by
// oracle/Makefile into oracle/_ref/libgg_ref.so (git-ignored).  Used for
//   * pinning oracle/gg_oracle.c (the C restatement) and generating tests/golden/*.json,
//   * the timed CPU baseline ("kind": "reference") in bench.py.
//
// Reference entry points driven here:
//   grove::insert_data(idx, key, data)                 include/genogrove/structure/grove/grove_insert.ipp:36-43
//   grove::insert_data(idx, key, data, sorted)         grove_insert.ipp:19-26
//   grove::insert_data(idx, vec, sorted, bulk)         grove_insert.ipp:68-127
//   grove::insert_data(idx, vec, bulk)                 grove_insert.ipp:160-172
//   grove::intersect(q, idx) / intersect(q)            grove_query.ipp:16-43
//   grove::remove_key(idx, key*)                       grove_remove.ipp:24-62
//   grove<genomic_coordinate, uint32_t> (ref_gc_*)     same entry points; genomic_coordinate.hpp:97-110, 156-160
//   grove::flanking(q, idx)                            grove_flanking.ipp -> detail::search_flanking query_engine.hpp:238-311
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <string>
#include <thread>
#include <utility>
#include <vector>

#include <genogrove/data_type/genomic_coordinate.hpp>
#include <genogrove/data_type/interval.hpp>
#include <genogrove/structure/grove/grove.hpp>

namespace gdt = genogrove::data_type;
namespace gst = genogrove::structure;

using ref_grove_t = gst::grove<gdt::interval, uint32_t>;
using ref_key_t = gdt::key<gdt::interval, uint32_t>;

struct ref_grove {
    ref_grove_t g;
    std::vector<std::string> names;  // chrom id -> index name
    std::vector<ref_key_t*> by_id;   // payload id -> key* (for remove)
    explicit ref_grove(int order) : g(order) {}
    const std::string& name(uint32_t c) {
        if (c >= names.size()) {
            size_t old = names.size();
            names.resize(c + 1);
            for (size_t i = old; i <= c; ++i) names[i] = "chr" + std::to_string(i);
        }
        return names[c];
    }
    void remember(uint32_t id, ref_key_t* k) {
        if (id >= by_id.size()) by_id.resize(static_cast<size_t>(id) + 1, nullptr);
        by_id[id] = k;
    }
};

extern "C" {

// mode for ref_insert
enum { REF_REGULAR = 0, REF_SORTED = 1 };

ref_grove* ref_create(int order) {
    try {
        return new ref_grove(order);
    } catch (...) {
        return nullptr;  // order < 3 -> std::invalid_argument (grove.hpp:158-162)
    }
}

void ref_destroy(ref_grove* r) { delete r; }

// returns 0 ok, 1 = invalid interval (start > end -> std::invalid_argument, interval.hpp:70-74)
int ref_insert(ref_grove* r, uint32_t chrom, uint64_t start, uint64_t end, uint32_t id, int mode) {
    try {
        gdt::interval iv(start, end);
        ref_key_t* k = (mode == REF_SORTED) ? r->g.insert_data(r->name(chrom), iv, id, gst::sorted)
                                            : r->g.insert_data(r->name(chrom), iv, id);
        r->remember(id, k);
        return 0;
    } catch (const std::invalid_argument&) {
        return 1;
    }
}

// bulk insert into ONE chromosome. presorted != 0 -> (sorted, bulk) path, else the sorting bulk path.
int ref_insert_bulk(ref_grove* r, uint32_t chrom, const uint64_t* start, const uint64_t* end,
                    const uint32_t* id, size_t n, int presorted) {
    try {
        std::vector<std::pair<gdt::interval, uint32_t>> v;
        v.reserve(n);
        for (size_t i = 0; i < n; ++i) v.emplace_back(gdt::interval(start[i], end[i]), id[i]);
        std::vector<ref_key_t*> keys = presorted ? r->g.insert_data(r->name(chrom), v, gst::sorted, gst::bulk)
                                                 : r->g.insert_data(r->name(chrom), std::move(v), gst::bulk);
        for (ref_key_t* k : keys) r->remember(k->get_data(), k);
        return 0;
    } catch (const std::invalid_argument&) {
        return 1;
    }
}

// returns 1 if removed, 0 otherwise
int ref_remove(ref_grove* r, uint32_t chrom, uint32_t id) {
    if (id >= r->by_id.size() || r->by_id[id] == nullptr) return 0;
    bool ok = r->g.remove_key(r->name(chrom), r->by_id[id]);
    if (ok) r->by_id[id] = nullptr;
    return ok ? 1 : 0;
}

// One query against one chromosome (chrom == 0xFFFFFFFF -> all indices, grove_query.ipp:16-24).
// Writes up to cap ids, returns the number of hits (may exceed cap).
size_t ref_intersect(ref_grove* r, uint32_t chrom, uint64_t qs, uint64_t qe, uint32_t* out, size_t cap) {
    gdt::interval q(qs, qe);
    auto res = (chrom == 0xFFFFFFFFu) ? r->g.intersect(q) : r->g.intersect(q, r->name(chrom));
    const auto& keys = res.get_keys();
    for (size_t i = 0; i < keys.size() && i < cap; ++i) out[i] = keys[i]->get_data();
    return keys.size();
}

// Batch: one grove.intersect(q, chrom) per query (the reference's loop, cli/src/intersect.cpp:39-58),
// query-parallel over `threads` contiguous shards of one shared read-only grove.
// Pass 1 (out_ids == nullptr): counts[i] = hits of query i.  Pass 2: ids written at offsets[i].
// Returns elapsed seconds of the query loop only.
double ref_intersect_batch(ref_grove* r, const uint32_t* qchrom, const uint64_t* qs, const uint64_t* qe,
                           size_t nq, int threads, uint64_t* counts, const uint64_t* offsets,
                           uint32_t* out_ids, uint64_t* checksum) {
    if (threads < 1) threads = 1;
    for (size_t i = 0; i < nq; ++i)
        if (qchrom[i] != 0xFFFFFFFFu) r->name(qchrom[i]);  // materialise names before threads start
    std::vector<uint64_t> sums(threads, 0);
    auto work = [&](int t) {
        size_t lo = nq * t / threads, hi = nq * (t + 1) / threads;
        uint64_t acc = 0;
        for (size_t i = lo; i < hi; ++i) {
            gdt::interval q(qs[i], qe[i]);
            static const std::string missing = "\x01<no such index>";
            auto res = r->g.intersect(q, qchrom[i] == 0xFFFFFFFFu ? missing : r->names[qchrom[i]]);
            const auto& keys = res.get_keys();
            if (counts) counts[i] = keys.size();
            acc += keys.size();
            if (out_ids) {
                uint32_t* dst = out_ids + offsets[i];
                for (size_t k = 0; k < keys.size(); ++k) dst[k] = keys[k]->get_data();
            } else {
                for (auto* k : keys) acc += k->get_data();  // consume results
            }
        }
        sums[t] = acc;
    };
    auto t0 = std::chrono::steady_clock::now();
    if (threads == 1) {
        work(0);
    } else {
        std::vector<std::thread> th;
        for (int t = 0; t < threads; ++t) th.emplace_back(work, t);
        for (auto& x : th) x.join();
    }
    auto t1 = std::chrono::steady_clock::now();
    uint64_t s = 0;
    for (auto v : sums) s += v;
    if (checksum) *checksum = s;
    return std::chrono::duration<double>(t1 - t0).count();
}

// Batch of grove.flanking(q, chrom): payload id of the nearest predecessor / successor, 0xFFFFFFFF when absent
// (null pointer in flanking_query_result, also for an unknown index name).
void ref_flanking_batch(ref_grove* r, const uint32_t* qchrom, const uint64_t* qs, const uint64_t* qe, size_t nq,
                        uint32_t* pred_ids, uint32_t* succ_ids) {
    static const std::string missing = "\x01<no such index>";
    for (size_t i = 0; i < nq; ++i) {
        gdt::interval q(qs[i], qe[i]);
        auto res = r->g.flanking(q, qchrom[i] == 0xFFFFFFFFu ? missing : r->name(qchrom[i]));
        pred_ids[i] = res.get_predecessor() ? res.get_predecessor()->get_data() : 0xFFFFFFFFu;
        succ_ids[i] = res.get_successor() ? res.get_successor()->get_data() : 0xFFFFFFFFu;
    }
}

// ---- grove<genomic_coordinate, uint32_t>: the strand-aware key type (one chromosome namespace as above)
using gc_grove_t = gst::grove<gdt::genomic_coordinate, uint32_t>;
struct ref_gc_grove {
    gc_grove_t g;
    explicit ref_gc_grove(int order) : g(order) {}
};
ref_gc_grove* ref_gc_create(int order) {
    try {
        return new ref_gc_grove(order);
    } catch (...) {
        return nullptr;
    }
}
void ref_gc_destroy(ref_gc_grove* r) { delete r; }
// mode: 0 regular insert, 1 sorted insert; returns 1 on std::invalid_argument (bad strand / start > end)
int ref_gc_insert(ref_gc_grove* r, uint32_t chrom, char strand, uint64_t start, uint64_t end, uint32_t id, int mode) {
    try {
        gdt::genomic_coordinate k(strand, start, end);
        const std::string name = "chr" + std::to_string(chrom);
        if (mode == REF_SORTED) r->g.insert_data(name, k, id, gst::sorted);
        else r->g.insert_data(name, k, id);
        return 0;
    } catch (const std::invalid_argument&) {
        return 1;
    }
}
int ref_gc_insert_bulk(ref_gc_grove* r, uint32_t chrom, const char* strand, const uint64_t* start, const uint64_t* end,
                       const uint32_t* id, size_t n, int presorted) {
    try {
        std::vector<std::pair<gdt::genomic_coordinate, uint32_t>> v;
        v.reserve(n);
        for (size_t i = 0; i < n; ++i) v.emplace_back(gdt::genomic_coordinate(strand[i], start[i], end[i]), id[i]);
        const std::string name = "chr" + std::to_string(chrom);
        if (presorted) r->g.insert_data(name, v, gst::sorted, gst::bulk);
        else r->g.insert_data(name, std::move(v), gst::bulk);
        return 0;
    } catch (const std::invalid_argument&) {
        return 1;
    }
}
// returns #hits (may exceed cap), or SIZE_MAX on std::invalid_argument from the query's constructor
size_t ref_gc_intersect(ref_gc_grove* r, uint32_t chrom, char strand, uint64_t qs, uint64_t qe, uint32_t* out, size_t cap) {
    try {
        gdt::genomic_coordinate q(strand, qs, qe);
        auto res = r->g.intersect(q, "chr" + std::to_string(chrom));
        const auto& keys = res.get_keys();
        for (size_t i = 0; i < keys.size() && i < cap; ++i) out[i] = keys[i]->get_data();
        return keys.size();
    } catch (const std::invalid_argument&) {
        return static_cast<size_t>(-1);
    }
}
// grove<genomic_coordinate>::flanking(q, idx[, pred]) -> payload ids {predecessor, successor}, 0xFFFFFFFF when absent.
// pred_mode 0: the default predicate (always compatible) — grove_flanking.ipp:33-37; 1: the strand-compatible predicate of
// the reference's own documentation (grove_flanking.ipp:47-54: equal strands, or either is '*').
// returns 1 on std::invalid_argument from the query's constructor
int ref_gc_flanking(ref_gc_grove* r, uint32_t chrom, char strand, uint64_t qs, uint64_t qe, int pred_mode, uint32_t* out2) {
    try {
        gdt::genomic_coordinate q(strand, qs, qe);
        const std::string name = "chr" + std::to_string(chrom);
        auto compatible = [](const gdt::genomic_coordinate& candidate, const gdt::genomic_coordinate& query) {
            return query.get_strand() == '*' || candidate.get_strand() == '*' || candidate.get_strand() == query.get_strand();
        };
        auto res = pred_mode == 1 ? r->g.flanking(q, name, compatible) : r->g.flanking(q, name);
        out2[0] = res.get_predecessor() ? res.get_predecessor()->get_data() : 0xFFFFFFFFu;
        out2[1] = res.get_successor() ? res.get_successor()->get_data() : 0xFFFFFFFFu;
        return 0;
    } catch (const std::invalid_argument&) {
        return 1;
    }
}
int ref_gc_less(char sa, uint64_t as, uint64_t ae, char sb, uint64_t bs, uint64_t be) {
    return gdt::genomic_coordinate(sa, as, ae) < gdt::genomic_coordinate(sb, bs, be) ? 1 : 0;
}
int ref_gc_overlaps(char sa, uint64_t as, uint64_t ae, char sb, uint64_t bs, uint64_t be) {
    return gdt::genomic_coordinate::overlaps(gdt::genomic_coordinate(sa, as, ae), gdt::genomic_coordinate(sb, bs, be)) ? 1 : 0;
}

// known-answer helpers for interval arithmetic (interval.hpp:87-105,140-157)
int ref_overlaps(uint64_t as, uint64_t ae, uint64_t bs, uint64_t be) {
    return gdt::interval::overlaps(gdt::interval(as, ae), gdt::interval(bs, be)) ? 1 : 0;
}
int ref_less(uint64_t as, uint64_t ae, uint64_t bs, uint64_t be) {
    return gdt::interval(as, ae) < gdt::interval(bs, be) ? 1 : 0;
}
int ref_greater(uint64_t as, uint64_t ae, uint64_t bs, uint64_t be) {
    return gdt::interval(as, ae) > gdt::interval(bs, be) ? 1 : 0;
}
void ref_distribute_evenly(uint64_t total, uint64_t max_per_group, uint64_t* out3) {
    auto d = gst::detail::distribute_evenly(total, max_per_group);
    out3[0] = d.num_groups;
    out3[1] = d.base_per_group;
    out3[2] = d.extra_groups;
}
int ref_hardware_threads() { return static_cast<int>(std::thread::hardware_concurrency()); }

}  // extern "C"
