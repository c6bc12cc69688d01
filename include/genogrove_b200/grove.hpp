// genogrove_b200/grove.hpp — C++20 host facade of the B200 intersect path (header-only, links libgg_b200.so).
//
// Mirrors the part of genogrove's public API that sits on the grove<interval, T>::intersect hot path — same
// names, argument meaning and error behaviour — so that code written against the reference compiles against
// this header after swapping the namespace aliases:
//
//     namespace gdt = genogrove_b200::data_type;   // reference: genogrove::data_type
//     namespace gst = genogrove_b200::structure;   // reference: genogrove::structure
//
// Reference interfaces mirrored (paths relative to the reference checkout):
//   data_type::interval            include/genogrove/data_type/interval.hpp:54-221
//   data_type::key<K, D>           include/genogrove/data_type/key.hpp:79-347
//   data_type::query_result<K, D>  include/genogrove/data_type/query_result.hpp:59-118
//   structure::sorted / bulk tags  include/genogrove/structure/grove/grove.hpp:56-67
//   structure::grove<K, D>         include/genogrove/structure/grove/grove.hpp:143-336
//     grove(int order)             grove.hpp:158-167   (order < 3 -> std::invalid_argument, default 3)
//     insert_data(...) x4, insert  include/genogrove/structure/grove/grove_insert.ipp:19-43,68-127,160-196
//     intersect(q) / (q, index)    include/genogrove/structure/grove/grove_query.ipp:16-43
//     remove_key / compact         include/genogrove/structure/grove/grove_remove.ipp:24-98
//   plus the NEW batch entry point grove::intersect_batch (millions of queries per call -> CSR).
//
// What lives where.  The host owns the keys (std::deque: stable key* identity, like grove.hpp:327) and, per
// index name, the LEAF-CHAIN ORDER of key pointers — the order search_overlaps appends hits in
// (query_engine.hpp:83-112).  The device holds a derived snapshot of that order (flattened B+ tree, built and
// searched by the hand-written sm_100a kernels behind include/gg_b200.h); it is refreshed lazily after any
// mutation.  Every overlap is computed on the GPU: there is no CPU search path in this header, and a missing
// device surfaces as genogrove_b200::device_error from the constructor.
//
// Leaf-chain order per insert path (the only freedom is the order of keys with EQUAL (start, end)):
//   insert_data(idx, k, d)               lower_bound insertion  -> equal keys in reverse insertion order
//   insert_data(idx, vec, sorted, bulk)  input order (empty index: bottom-up build; else rightmost append)
//   insert_data(idx, vec, bulk)          device radix sort, stable: equal keys keep input order (the reference
//                                        runs an UNSTABLE std::sort here, grove_insert.ipp:167-168, so it does
//                                        not define that order itself)
//   insert_data(idx, k, d, sorted)       lower_bound inside the RIGHTMOST LEAF only (grove_insert.ipp:501-531), so
//                                        the place among equal keys depends on that leaf's key count.  The count is
//                                        exact for every history: closed form while an index has only seen bulk /
//                                        sorted / adopted input, and from a counts-only replay of the reference's
//                                        split / borrow / merge rules (chain_shape.hpp) once a regular insert or a
//                                        removal touched it
#ifndef GENOGROVE_B200_GROVE_HPP
#define GENOGROVE_B200_GROVE_HPP

#include <algorithm>
#include <concepts>
#include <cstddef>
#include <cstdint>
#include <deque>
#include <functional>
#include <limits>
#include <memory>
#include <numeric>
#include <ranges>
#include <span>
#include <stdexcept>
#include <string>
#include <string_view>
#include <type_traits>
#include <unordered_map>
#include <utility>
#include <vector>

#include "../gg_b200.h"
#include "chain_shape.hpp"

namespace genogrove_b200 {

/// Any failure of the CUDA library other than an invalid interval (which maps to std::invalid_argument).
class device_error : public std::runtime_error {
  public:
    device_error(int status, const std::string& what) : std::runtime_error(what), status_(status) {}
    [[nodiscard]] int status() const noexcept { return status_; }

  private:
    int status_;
};

namespace data_type {

/// Closed interval [start, end] over size_t coordinates; adjacency counts as overlap.
class interval {
  public:
    static constexpr std::size_t INVALID_POSITION = std::numeric_limits<std::size_t>::max();

    constexpr interval() : start(INVALID_POSITION), end(INVALID_POSITION) {}
    constexpr interval(std::size_t s, std::size_t e) : start(s), end(e) {
        if (s > e) throw std::invalid_argument("interval: start must be <= end");
    }

    constexpr bool operator<(const interval& o) const { return start != o.start ? start < o.start : end < o.end; }
    constexpr bool operator>(const interval& o) const { return start != o.start ? start > o.start : end > o.end; }
    constexpr bool operator==(const interval& o) const { return start == o.start && end == o.end; }

    [[nodiscard]] static constexpr bool overlaps(const interval& a, const interval& b) {
        return (a.start > b.start ? a.start : b.start) <= (a.end < b.end ? a.end : b.end);
    }
    [[nodiscard]] static constexpr interval aggregate(const interval& a, const interval& b) {
        return interval(a.start < b.start ? a.start : b.start, a.end > b.end ? a.end : b.end);
    }

    [[nodiscard]] std::string to_string() const { return "[" + std::to_string(start) + "," + std::to_string(end) + "]"; }
    constexpr std::size_t get_start() const noexcept { return start; }
    constexpr std::size_t get_end() const noexcept { return end; }
    constexpr void set_range(std::size_t s, std::size_t e) {
        if (s > e) throw std::invalid_argument("interval: start must be <= end");
        start = s;
        end = e;
    }

  private:
    std::size_t start;
    std::size_t end;
};

/// Stranded closed interval: strand in {'+', '-', '.', '*'}, ordered by (start, end, strand) with '*' < '.' < '+' < '-';
/// overlaps = coordinates overlap AND (equal strands OR either one is the wildcard '*').  Same members and semantics
/// as genogrove's data_type/genomic_coordinate.hpp:62-260.
class genomic_coordinate {
  public:
    constexpr genomic_coordinate() : strand('.'), start(0), end(0) {}
    constexpr genomic_coordinate(char strand_, std::size_t s, std::size_t e) : strand(strand_), start(s), end(e) {
        if (!valid_strand(strand_)) throw std::invalid_argument("genomic_coordinate: strand must be one of '+', '-', '.', '*'");
        if (s > e) throw std::invalid_argument("genomic_coordinate: start must be <= end");
    }
    static constexpr bool is_interval = true;

    /// rank of a strand in the sort order: '*' 0, '.' 1, '+' 2, '-' 3
    [[nodiscard]] static constexpr unsigned strand_rank(char c) noexcept { return c == '*' ? 0u : c == '.' ? 1u : c == '+' ? 2u : 3u; }
    constexpr bool operator<(const genomic_coordinate& o) const {
        if (start != o.start) return start < o.start;
        if (end != o.end) return end < o.end;
        return strand_rank(strand) < strand_rank(o.strand);
    }
    constexpr bool operator>(const genomic_coordinate& o) const { return o < *this; }
    constexpr bool operator==(const genomic_coordinate& o) const { return strand == o.strand && start == o.start && end == o.end; }

    [[nodiscard]] static constexpr bool overlaps(const genomic_coordinate& a, const genomic_coordinate& b) {
        if (a.start > b.end || b.start > a.end) return false;
        if (a.strand == '*' || b.strand == '*') return true;
        return a.strand == b.strand;
    }
    [[nodiscard]] static constexpr genomic_coordinate aggregate(const genomic_coordinate& a, const genomic_coordinate& b) {
        return genomic_coordinate(a.strand == b.strand ? a.strand : '*', a.start < b.start ? a.start : b.start,
                                  a.end > b.end ? a.end : b.end);
    }
    [[nodiscard]] std::string to_string() const {
        return std::string(1, strand) + ":" + std::to_string(start) + "-" + std::to_string(end);
    }
    constexpr char get_strand() const noexcept { return strand; }
    constexpr std::size_t get_start() const noexcept { return start; }
    constexpr std::size_t get_end() const noexcept { return end; }
    constexpr void set_strand(char c) {
        if (!valid_strand(c)) throw std::invalid_argument("genomic_coordinate: strand must be one of '+', '-', '.', '*'");
        strand = c;
    }
    constexpr void set_range(std::size_t s, std::size_t e) {
        if (s > e) throw std::invalid_argument("genomic_coordinate: start must be <= end");
        start = s;
        end = e;
    }

  private:
    static constexpr bool valid_strand(char c) noexcept { return c == '+' || c == '-' || c == '.' || c == '*'; }
    char strand;
    std::size_t start;
    std::size_t end;
};

namespace detail {
struct no_data {};
}  // namespace detail

/// Value + optional payload.  The ADDRESS of a key owned by a grove is its identity in query results.
template <typename key_t, typename data_t = void>
class key {
    using stored_t = std::conditional_t<std::is_void_v<data_t>, detail::no_data, data_t>;

  public:
    key() requires std::default_initializable<key_t> && std::default_initializable<stored_t> : value{}, data{} {}
    explicit key(key_t kvalue) requires std::default_initializable<stored_t> : value(std::move(kvalue)), data{} {}
    template <typename D = data_t>
    key(key_t kvalue, D&& dvalue) requires(!std::is_void_v<data_t>)
        : value(std::move(kvalue)), data(std::forward<D>(dvalue)) {}

    [[nodiscard]] const key_t& get_value() const noexcept { return value; }
    void set_value(key_t v) { value = std::move(v); }

    template <typename D = data_t>
    [[nodiscard]] const D& get_data() const noexcept requires(!std::is_void_v<data_t>) { return data; }
    template <typename D = data_t>
    [[nodiscard]] D& get_data() noexcept requires(!std::is_void_v<data_t>) { return data; }
    template <typename D = data_t>
    void set_data(D d) requires(!std::is_void_v<data_t>) { data = std::move(d); }
    [[nodiscard]] constexpr bool has_data() const noexcept { return !std::is_void_v<data_t>; }

    [[nodiscard]] std::string to_string() const { return value.to_string(); }
    bool operator<(const key& o) const { return value < o.value; }
    bool operator>(const key& o) const { return value > o.value; }

  private:
    key_t value;
    [[no_unique_address]] stored_t data;
};

/// Copy of the query + borrowed key pointers in leaf-chain order; does not own the keys.
template <typename key_t, typename data_t = void>
class query_result {
  public:
    explicit query_result(key_t q) : query(std::move(q)), keys{} {}
    [[nodiscard]] const key_t& get_query() const noexcept { return query; }
    [[nodiscard]] const std::vector<key<key_t, data_t>*>& get_keys() const { return keys; }
    void add_key(key<key_t, data_t>* k) {
        if (k == nullptr) throw std::invalid_argument("query_result::add_key: key must not be nullptr");
        keys.push_back(k);
    }
    void reserve(std::size_t n) { keys.reserve(n); }

  private:
    key_t query;
    std::vector<key<key_t, data_t>*> keys;
};

/// Nearest non-overlapping predecessor / successor of a flanking query; borrowed pointers, nullptr when absent
/// (same members as genogrove's data_type/flanking_query_result.hpp).
template <typename key_t, typename data_t = void>
class flanking_query_result {
  public:
    flanking_query_result() = default;
    [[nodiscard]] key<key_t, data_t>* get_predecessor() const noexcept { return predecessor; }
    [[nodiscard]] key<key_t, data_t>* get_successor() const noexcept { return successor; }
    void set_predecessor(key<key_t, data_t>* k) noexcept { predecessor = k; }
    void set_successor(key<key_t, data_t>* k) noexcept { successor = k; }

  private:
    key<key_t, data_t>* predecessor = nullptr;
    key<key_t, data_t>* successor = nullptr;
};

}  // namespace data_type

namespace structure {

namespace gdt = genogrove_b200::data_type;

struct string_hash {
    using is_transparent = void;
    std::size_t operator()(std::string_view sv) const noexcept { return std::hash<std::string_view>{}(sv); }
};

struct sorted_t {};
struct bulk_t {};
inline constexpr sorted_t sorted{};
inline constexpr bulk_t bulk{};
/// The flanking predicate the reference documents for strand-aware keys (grove_flanking.ipp:47-54): candidate and query
/// have equal strands, or either one is '*'.  Passed by tag (not as a callable) so that it can run on the device.
struct strand_compatible_t {
    template <typename K>
    constexpr bool operator()(const K& candidate, const K& query) const noexcept {
        return query.get_strand() == '*' || candidate.get_strand() == '*' || candidate.get_strand() == query.get_strand();
    }
};
inline constexpr strand_compatible_t strand_compatible{};
/// Tag for a grove WITHOUT a device handle: host bookkeeping only (keys, leaf chains, tree shape — chain(),
/// shape_tokens()), so that host-side ordering logic can be tested where no GPU exists.  Anything that needs an
/// overlap computed — intersect, flanking, the device-sorted unsorted bulk — throws device_error: there is no CPU
/// search path.
struct detached_t {};
inline constexpr detached_t detached{};

template <typename key_type, typename data_type>
class compact_batch_result;

/// Result of grove::intersect_batch: CSR over the batch.  Query q's hits are ranks()[offsets()[q] ..
/// offsets()[q+1]) — leaf-chain ranks inside the query's index, ascending — and key(q, j) maps them back to the
/// grove's key pointers.  Valid until the grove is mutated, queried again or destroyed (the rank array is the
/// library's pinned result buffer).
template <typename key_type, typename data_type>
class batch_result {
  public:
    using key_ptr = gdt::key<key_type, data_type>*;

    [[nodiscard]] std::size_t size() const noexcept { return offsets_.empty() ? 0 : offsets_.size() - 1; }
    [[nodiscard]] std::uint64_t total_hits() const noexcept { return n_hits_; }
    [[nodiscard]] const std::vector<std::uint64_t>& offsets() const noexcept { return offsets_; }
    [[nodiscard]] std::span<const std::uint32_t> ranks() const noexcept { return {hits_, static_cast<std::size_t>(n_hits_)}; }
    [[nodiscard]] std::span<const std::uint32_t> ranks(std::size_t q) const {
        return {hits_ + offsets_[q], static_cast<std::size_t>(offsets_[q + 1] - offsets_[q])};
    }
    [[nodiscard]] std::size_t count(std::size_t q) const { return static_cast<std::size_t>(offsets_[q + 1] - offsets_[q]); }
    [[nodiscard]] key_ptr key(std::size_t q, std::size_t j) const {
        return (*chains_[q_index_[q]])[hits_[offsets_[q] + j]];
    }
    /// the reference-shaped view of one query of the batch
    [[nodiscard]] std::vector<key_ptr> get_keys(std::size_t q) const {
        std::vector<key_ptr> out;
        out.reserve(count(q));
        for (std::size_t j = 0; j < count(q); ++j) out.push_back(key(q, j));
        return out;
    }
    /// Copy the rank array out of the library's pinned buffer into this object: afterwards the result survives later
    /// queries on the grove (it still refers to the grove's keys, so not its mutation or destruction).  Costs one memcpy of
    /// 4 bytes per hit; results that are consumed before the next query do not need it.
    batch_result& keep() & {
        if (hits_ != nullptr && hits_ != owned_.data()) {
            owned_.assign(hits_, hits_ + n_hits_);
            hits_ = owned_.data();
        }
        return *this;
    }
    /// `auto r = g.intersect_batch(...).keep();` — on a temporary the kept result is handed over by value
    [[nodiscard]] batch_result keep() && {
        keep();
        return std::move(*this);
    }
    [[nodiscard]] bool owns_ranks() const noexcept { return hits_ == owned_.data() && hits_ != nullptr; }

    batch_result() = default;
    ~batch_result() = default;
    /// A kept result's rank pointer refers to its own vector: copies and moves re-seat it (a copy of a view stays a view).
    batch_result(const batch_result& o)
        : offsets_(o.offsets_), hits_(o.hits_), n_hits_(o.n_hits_), owned_(o.owned_), q_index_(o.q_index_), chains_(o.chains_) {
        if (o.owns_ranks()) hits_ = owned_.data();
    }
    batch_result(batch_result&& o) noexcept
        : offsets_(std::move(o.offsets_)), hits_(o.hits_), n_hits_(o.n_hits_), q_index_(std::move(o.q_index_)),
          chains_(std::move(o.chains_)) {
        const bool owned = o.owns_ranks();
        owned_ = std::move(o.owned_);
        if (owned) hits_ = owned_.data();
        o.hits_ = nullptr;
        o.n_hits_ = 0;
    }
    batch_result& operator=(const batch_result& o) {
        if (this != &o) {
            batch_result tmp(o);
            *this = std::move(tmp);
        }
        return *this;
    }
    batch_result& operator=(batch_result&& o) noexcept {
        if (this != &o) {
            const bool owned = o.owns_ranks();
            offsets_ = std::move(o.offsets_);
            hits_ = o.hits_;
            n_hits_ = o.n_hits_;
            owned_ = std::move(o.owned_);
            q_index_ = std::move(o.q_index_);
            chains_ = std::move(o.chains_);
            if (owned) hits_ = owned_.data();
            o.hits_ = nullptr;
            o.n_hits_ = 0;
        }
        return *this;
    }

  private:
    template <typename, typename, typename>
    friend class grove;
    template <typename, typename>
    friend class compact_batch_result;  // expand() fills one
    std::vector<std::uint64_t> offsets_;
    const std::uint32_t* hits_ = nullptr;
    std::uint64_t n_hits_ = 0;
    std::vector<std::uint32_t> owned_;                         // keep(): the ranks, once copied out of the pinned buffer
    std::vector<std::uint32_t> q_index_;                       // copy of the per-query index ids
    std::vector<const std::vector<key_ptr>*> chains_;          // index id -> chain (rank -> key*)
};

/// Result of grove::intersect_batch_compact: the same hits as intersect_batch in the form that crosses PCIe cheapest
/// (gg_intersect_batch_compact: 8 bytes per query + 1-4 bytes per hit instead of 8 + 4) and is read in place — the host
/// call is copy-bound, so on BED/BAM-shaped batches this form is 2-3x the plain CSR end to end.  count / rank / key decode
/// on the fly; expand() materialises the plain CSR for callers that need arrays.  A view of the library's pinned
/// buffers: valid until the grove is mutated, queried again or destroyed.
template <typename key_type, typename data_type>
class compact_batch_result {
  public:
    using key_ptr = gdt::key<key_type, data_type>*;
    [[nodiscard]] std::size_t size() const noexcept { return static_cast<std::size_t>(res_.n_queries); }
    [[nodiscard]] std::uint64_t total_hits() const noexcept { return res_.n_hits; }
    [[nodiscard]] std::size_t count(std::size_t q) const { return static_cast<std::size_t>(gg_compact_count(&res_, q)); }
    /// j-th hit of query q as a leaf-chain rank inside the query's index, ascending in j
    [[nodiscard]] std::uint32_t rank(std::size_t q, std::size_t j) const { return gg_compact_rank(&res_, q_index_.data(), q, j); }
    [[nodiscard]] key_ptr key(std::size_t q, std::size_t j) const { return (*chains_[q_index_[q]])[rank(q, j)]; }
    [[nodiscard]] std::vector<key_ptr> get_keys(std::size_t q) const {
        const std::size_t n = count(q);
        std::vector<key_ptr> out;
        out.reserve(n);
        for (std::size_t j = 0; j < n; ++j) out.push_back(key(q, j));
        return out;
    }
    /// the C-ABI view (chunk table, per-query records, delta bytes) for callers that decode in their own loops
    [[nodiscard]] const gg_compact_result& raw() const noexcept { return res_; }
    /// The plain CSR with its own rank array (survives later queries, like batch_result::keep()).  n_threads <= 0: all cores.
    [[nodiscard]] batch_result<key_type, data_type> expand(int n_threads = 0) const {
        batch_result<key_type, data_type> r;
        r.offsets_.resize(size() + 1);
        r.owned_.resize(static_cast<std::size_t>(res_.n_hits) + 1);
        const gg_status st = gg_compact_expand(&res_, q_index_.data(), r.offsets_.data(), r.owned_.data(), n_threads);
        if (st != GG_OK) throw device_error(st, "gg_compact_expand");
        r.hits_ = r.owned_.data();
        r.n_hits_ = res_.n_hits;
        r.q_index_ = q_index_;
        r.chains_ = chains_;
        return r;
    }

  private:
    template <typename, typename, typename>
    friend class grove;
    gg_compact_result res_{};
    std::vector<std::uint32_t> q_index_;
    std::vector<const std::vector<key_ptr>*> chains_;
};

template <typename key_type, typename data_type = void, typename edge_data_type = void>
class grove {
    static_assert(std::is_same_v<key_type, gdt::interval> || std::is_same_v<key_type, gdt::genomic_coordinate>,
                  "genogrove_b200::structure::grove accelerates grove<interval, T> and grove<genomic_coordinate, T> (SURVEY.md §8)");
    /// genomic_coordinate keys: one strand byte per chain rank rides along the device snapshot and every query
    /// carries its strand (gg_index_set_strands / gg_intersect_batch_stranded)
    static constexpr bool stranded = std::is_same_v<key_type, gdt::genomic_coordinate>;

  public:
    using key_t = gdt::key<key_type, data_type>;
    using key_ptr = key_t*;
    static constexpr std::uint32_t npos_index = GG_NO_CHROM;

    explicit grove(int order, int device = 0) : order(order) {
        if (order < 3) throw std::invalid_argument("grove order must be >= 3");
        gg_index* h = nullptr;
        gg_status st = gg_index_create(&h, device);
        if (st != GG_OK) throw device_error(st, "gg_index_create: no usable CUDA device (there is no CPU fallback)");
        dev.reset(h);
    }
    grove(int order, detached_t) : order(order) {
        if (order < 3) throw std::invalid_argument("grove order must be >= 3");
    }
    grove() : grove(3) {}
    ~grove() = default;
    grove(const grove&) = delete;
    grove& operator=(const grove&) = delete;
    grove(grove&&) noexcept = default;
    grove& operator=(grove&&) noexcept = default;

    [[nodiscard]] int get_order() const { return order; }
    /// number of keys indexed in the tree (the reference's indexed_vertex_count, grove_graph.ipp:204)
    [[nodiscard]] std::size_t indexed_vertex_count() const noexcept { return leaf_key_count; }
    [[nodiscard]] std::size_t size() const noexcept { return leaf_key_count; }
    [[nodiscard]] std::size_t key_storage_size() const noexcept { return key_storage.size(); }
    [[nodiscard]] gg_index* device_handle() { sync(); return dev.get(); }

    // ------------------------------------------------------------------------------------------ inserts
    template <typename D = data_type>
    key_ptr insert_data(std::string_view index, key_type key_value, D data_value, sorted_t)
        requires(!std::is_void_v<data_type>) {
        return insert_sorted(index, allocate_key(key_t(std::move(key_value), std::move(data_value))));
    }
    template <typename D = data_type>
    key_ptr insert_data(std::string_view index, key_type key_value, D data_value) requires(!std::is_void_v<data_type>) {
        return insert(index, key_t(std::move(key_value), std::move(data_value)));
    }
    key_ptr insert_data(std::string_view index, key_type key_value, sorted_t) requires(std::is_void_v<data_type>) {
        return insert_sorted(index, allocate_key(key_t(std::move(key_value))));
    }
    key_ptr insert_data(std::string_view index, key_type key_value) requires(std::is_void_v<data_type>) {
        return insert(index, key_t(std::move(key_value)));
    }

    /// regular insert: the chain position is lower_bound over the whole index (node.hpp:247-254 reached by
    /// top-down routing), so a key lands BEFORE existing equal keys.  Positions are resolved lazily, in one
    /// merge, before the next query.
    key_ptr insert(std::string_view index, const key_t& k) {
        chain_state& ch = chain_for(index);
        key_ptr p = allocate_key(k);
        ch.pending.push_back(p);
        ++leaf_key_count;
        touch(ch);
        return p;
    }

    template <typename Container>
    std::vector<key_ptr> insert_data(std::string_view index, const Container& data, sorted_t, bulk_t)
        requires(!std::is_void_v<data_type> && std::ranges::forward_range<Container> && std::ranges::sized_range<Container>) {
        std::vector<key_ptr> inserted;
        if (std::ranges::empty(data)) return inserted;
        inserted.reserve(std::ranges::size(data));
        for (const auto& [key_value, data_value] : data) inserted.push_back(allocate_key(key_t(key_value, data_value)));
        append_sorted_run(index, inserted, /*device_is_current=*/false);
        return inserted;
    }

    /// unsorted bulk: the sort runs on the device (hand-written LSD radix sort of (start, end), stable), then the
    /// sorted-bulk path.  Returns the key pointers in SORTED order, like the reference does.
    template <typename Container>
    std::vector<key_ptr> insert_data(std::string_view index, Container data, bulk_t)
        requires(!std::is_void_v<data_type> && std::ranges::random_access_range<Container> && std::ranges::sized_range<Container>) {
        const std::size_t n = std::ranges::size(data);
        if (n == 0) return {};
        need_device();
        std::vector<std::uint64_t> s(n), e(n);
        for (std::size_t i = 0; i < n; ++i) {
            s[i] = data[i].first.get_start();
            e[i] = data[i].first.get_end();
        }
        const std::vector<std::uint32_t> zeros(n, 0u);
        std::vector<std::uint32_t> perm(n);
        const bool was_empty = indices.empty() && next_id == 0;
        if constexpr (stranded) {  // third sort key: the strand's rank (genomic_coordinate.hpp:97-110), then input order
            std::vector<std::uint64_t> tie(n);
            for (std::size_t i = 0; i < n; ++i) tie[i] = key_type::strand_rank(data[i].first.get_strand());
            check(gg_index_build(dev.get(), zeros.data(), s.data(), e.data(), tie.data(), GG_TIE_EXPLICIT, n, 1, perm.data()));
        } else {
            check(gg_index_build(dev.get(), zeros.data(), s.data(), e.data(), nullptr, GG_TIE_INPUT_ORDER, n, 1, perm.data()));
        }
        ++device_builds;
        dirty = true;  // the handle now holds a snapshot of this batch alone
        snapshot_valid = false;
        std::vector<key_ptr> inserted;
        inserted.reserve(n);
        for (std::size_t r = 0; r < n; ++r) inserted.push_back(allocate_key(key_t(data[perm[r]].first, data[perm[r]].second)));
        append_sorted_run(index, inserted, /*device_is_current=*/was_empty && !stranded);
        return inserted;
    }

    /// NEW: unsorted bulk load of SEVERAL indices at once — what a 24-chromosome annotation load is.  On an empty grove
    /// this is ONE device build (gg_index_build over all indices: radix sort on (index, start, end), levels, max-end
    /// augmentation) whose snapshot is the grove's: no second upload follows.  The per-index call above sorts each index on
    /// the device and then has to re-upload everything (24 sorts + 1 upload).  Returns, per entry, the key pointers in
    /// sorted order.  On a grove that already holds keys it is the per-index loop.
    template <typename Container>
    std::vector<std::vector<key_ptr>> insert_data(const std::vector<std::pair<std::string, Container>>& per_index, bulk_t)
        requires(!std::is_void_v<data_type> && std::ranges::random_access_range<Container> && std::ranges::sized_range<Container>) {
        std::vector<std::vector<key_ptr>> out(per_index.size());
        bool one_build = indices.empty() && next_id == 0 && !per_index.empty();
        std::size_t n = 0;
        for (std::size_t a = 0; a < per_index.size() && one_build; ++a) {
            if (std::ranges::size(per_index[a].second) == 0) one_build = false;
            for (std::size_t b = 0; b < a; ++b)
                if (per_index[a].first == per_index[b].first) one_build = false;
            n += std::ranges::size(per_index[a].second);
        }
        if (!one_build) {
            for (std::size_t a = 0; a < per_index.size(); ++a) out[a] = insert_data(per_index[a].first, per_index[a].second, bulk);
            return out;
        }
        need_device();
        std::vector<std::uint32_t> c(n), perm(n);
        std::vector<std::uint64_t> s(n), e(n), tie;
        if constexpr (stranded) tie.resize(n);
        std::size_t p = 0;
        for (std::size_t a = 0; a < per_index.size(); ++a)
            for (const auto& [key_value, data_value] : per_index[a].second) {
                c[p] = static_cast<std::uint32_t>(a);
                s[p] = key_value.get_start();
                e[p] = key_value.get_end();
                if constexpr (stranded) tie[p] = key_type::strand_rank(key_value.get_strand());
                ++p;
            }
        check(gg_index_build(dev.get(), c.data(), s.data(), e.data(), stranded ? tie.data() : nullptr,
                             stranded ? GG_TIE_EXPLICIT : GG_TIE_INPUT_ORDER, n, static_cast<std::uint32_t>(per_index.size()), perm.data()));
        ++device_builds;
        std::size_t first = 0;  // of this index in the concatenated input, and in perm (grouped by id in id order)
        for (std::size_t a = 0; a < per_index.size(); ++a) {
            const Container& data = per_index[a].second;
            const std::size_t m = std::ranges::size(data);
            out[a].reserve(m);
            for (std::size_t r = 0; r < m; ++r) {
                const auto& rec = data[perm[first + r] - first];
                out[a].push_back(allocate_key(key_t(rec.first, rec.second)));
            }
            append_sorted_run(per_index[a].first, out[a], /*device_is_current=*/false);
            first += m;
        }
        if constexpr (stranded) {
            std::vector<std::uint8_t> st;
            st.reserve(n);
            for (const auto& keys : out)
                for (const key_ptr k : keys) st.push_back(static_cast<std::uint8_t>(k->get_value().get_strand()));
            check(gg_index_set_strands(dev.get(), st.data(), st.size()));
        }
        for (auto& [name, ch] : indices) ch.dirty = false;  // the build's snapshot IS this grove
        device_n_chrom = next_id;
        snapshot_valid = true;
        dirty = false;
        return out;
    }

    // ------------------------------------------------------------------------------------------ removal
    bool remove_key(std::string_view index_name, key_ptr key_to_remove) {
        if (key_to_remove == nullptr) return false;
        auto it = indices.find(index_name);
        if (it == indices.end()) return false;
        chain_state& ch = it->second;
        flush_pending(ch);
        const key_type& v = key_to_remove->get_value();
        auto lo = std::lower_bound(ch.chain.begin(), ch.chain.end(), v,
                                   [](const key_ptr a, const key_type& b) { return a->get_value() < b; });
        auto pos = ch.chain.end();
        for (auto p = lo; p != ch.chain.end() && !((*p)->get_value() > v); ++p)
            if (*p == key_to_remove) { pos = p; break; }
        if (pos == ch.chain.end()) {  // chains that violate the sorted-append precondition are not globally sorted
            pos = std::find(ch.chain.begin(), ch.chain.end(), key_to_remove);
            if (pos == ch.chain.end()) return false;
        }
        materialise(ch);
        ch.shape.remove_at(static_cast<std::size_t>(pos - ch.chain.begin()));
        ch.chain.erase(pos);
        ch.leaf_begin = ch.chain.size() - ch.shape.rightmost_leaf_keys();
        --leaf_key_count;
        touch(ch);
        if (ch.chain.empty()) {  // the reference erases an emptied root (grove_remove.ipp:44-50)
            emptied_ids.push_back(ch.id);  // its id is retired: the snapshot keeps it as an empty chromosome
            chain_by_id[ch.id] = &no_chain();
            indices.erase(it);
        }
        return true;
    }

    /// The reference re-packs its key deque here and invalidates key pointers.  Nothing to reclaim on this side
    /// beyond dropping unlinked keys' slots lazily; pointers handed out earlier stay valid.
    void compact() {}

    // ------------------------------------------------------------------------------------------ queries
    [[nodiscard]] gdt::query_result<key_type, data_type> intersect(const key_type& query) {
        gdt::query_result<key_type, data_type> result{query};
        sync();
        if (indices.empty()) return result;
        // one query per index, in the host map's iteration order — the order the reference visits its roots in
        scratch_c.clear();
        for (const auto& [name, ch] : indices) scratch_c.push_back(ch.id);
        scratch_s.assign(scratch_c.size(), query.get_start());
        scratch_e.assign(scratch_c.size(), query.get_end());
        scratch_off.resize(scratch_c.size() + 1);
        const std::uint32_t* hits = nullptr;
        std::uint64_t n_hits = 0;
        if constexpr (stranded) {
            const std::vector<std::uint8_t> qst(scratch_c.size(), static_cast<std::uint8_t>(query.get_strand()));
            check(gg_intersect_batch_stranded(dev.get(), scratch_c.data(), scratch_s.data(), scratch_e.data(), qst.data(),
                                              scratch_c.size(), scratch_off.data(), &hits, &n_hits));
        } else {
            check(gg_intersect_batch(dev.get(), scratch_c.data(), scratch_s.data(), scratch_e.data(), scratch_c.size(),
                                     scratch_off.data(), &hits, &n_hits));
        }
        result.reserve(n_hits);
        for (std::size_t j = 0; j < scratch_c.size(); ++j) {
            const std::vector<key_ptr>& chain = *chain_by_id[scratch_c[j]];
            for (std::uint64_t h = scratch_off[j]; h < scratch_off[j + 1]; ++h) result.add_key(chain[hits[h]]);
        }
        return result;
    }

    [[nodiscard]] gdt::query_result<key_type, data_type> intersect(const key_type& query, std::string_view index) {
        gdt::query_result<key_type, data_type> result{query};
        auto it = indices.find(index);
        if (it == indices.end()) return result;  // unknown index: empty result, not an error
        sync();
        const std::uint32_t c = it->second.id;
        const std::uint64_t s = query.get_start(), e = query.get_end();
        std::uint64_t off[2] = {0, 0};
        const std::uint32_t* hits = nullptr;
        std::uint64_t n_hits = 0;
        if constexpr (stranded) {
            const std::uint8_t qst = static_cast<std::uint8_t>(query.get_strand());
            check(gg_intersect_batch_stranded(dev.get(), &c, &s, &e, &qst, 1, off, &hits, &n_hits));
        } else {
            check(gg_intersect_batch(dev.get(), &c, &s, &e, 1, off, &hits, &n_hits));
        }
        result.reserve(n_hits);
        const std::vector<key_ptr>& chain = it->second.chain;
        for (std::uint64_t h = 0; h < n_hits; ++h) result.add_key(chain[hits[h]]);
        return result;
    }

    /// Id of an index name in the device snapshot, npos_index (== GG_NO_CHROM, "no hits") when absent.  An id is
    /// assigned when the index is created and never changes or gets reused while the grove lives: creating, filling or
    /// emptying OTHER indices does not move it (ids cached before such a change stay valid).
    [[nodiscard]] std::uint32_t index_id(std::string_view index) {
        auto it = indices.find(index);
        if (it == indices.end()) return npos_index;
        sync();
        return it->second.id;
    }

    /// NEW: millions of queries in one call.  q_index[i] = index_id(name) of query i (npos_index -> no hits),
    /// closed coordinates like interval; start > end in any query -> std::invalid_argument.
    [[nodiscard]] batch_result<key_type, data_type> intersect_batch(const std::uint32_t* q_index, const std::uint64_t* q_start,
                                                                    const std::uint64_t* q_end, std::size_t nq)
        requires(!stranded) {
        sync();
        batch_result<key_type, data_type> r;
        r.offsets_.resize(nq + 1);
        check(gg_intersect_batch(dev.get(), q_index, q_start, q_end, nq, r.offsets_.data(), &r.hits_, &r.n_hits_));
        r.q_index_.assign(q_index, q_index + nq);
        r.chains_ = chain_by_id;
        return r;
    }
    /// The same with 32-bit coordinates (BED / SAM / hg38-sized; needs a grove whose coordinates all fit 32 bits, which the
    /// device snapshot then stores as u32): 12 instead of 20 bytes per query over PCIe (gg_intersect_batch_u32).
    [[nodiscard]] batch_result<key_type, data_type> intersect_batch(const std::uint32_t* q_index, const std::uint32_t* q_start,
                                                                    const std::uint32_t* q_end, std::size_t nq)
        requires(!stranded) {
        sync();
        batch_result<key_type, data_type> r;
        r.offsets_.resize(nq + 1);
        check(gg_intersect_batch_u32(dev.get(), q_index, q_start, q_end, nq, r.offsets_.data(), &r.hits_, &r.n_hits_));
        r.q_index_.assign(q_index, q_index + nq);
        r.chains_ = chain_by_id;
        return r;
    }
    /// ... and with the result left in the compact form (gg_intersect_batch_compact), read in place: the fastest way to
    /// put a host-resident batch through the device (bench.py's `e2e`).  Same hits, same order.
    [[nodiscard]] compact_batch_result<key_type, data_type> intersect_batch_compact(const std::uint32_t* q_index,
                                                                                    const std::uint32_t* q_start,
                                                                                    const std::uint32_t* q_end, std::size_t nq)
        requires(!stranded) {
        sync();
        compact_batch_result<key_type, data_type> r;
        check(gg_intersect_batch_compact(dev.get(), q_index, q_start, q_end, nq, &r.res_));
        r.q_index_.assign(q_index, q_index + nq);
        r.chains_ = chain_by_id;
        return r;
    }
    /// genomic_coordinate keys: the same plus one strand byte ('+', '-', '.', '*') per query
    [[nodiscard]] batch_result<key_type, data_type> intersect_batch(const std::uint32_t* q_index, const std::uint64_t* q_start,
                                                                    const std::uint64_t* q_end, const std::uint8_t* q_strand,
                                                                    std::size_t nq)
        requires(stranded) {
        sync();
        batch_result<key_type, data_type> r;
        r.offsets_.resize(nq + 1);
        check(gg_intersect_batch_stranded(dev.get(), q_index, q_start, q_end, q_strand, nq, r.offsets_.data(), &r.hits_, &r.n_hits_));
        r.q_index_.assign(q_index, q_index + nq);
        r.chains_ = chain_by_id;
        return r;
    }
    [[nodiscard]] batch_result<key_type, data_type> intersect_batch(std::span<const key_type> queries, std::string_view index) {
        const std::uint32_t c = index_id(index);
        std::vector<std::uint32_t> qc(queries.size(), c);
        std::vector<std::uint64_t> qs(queries.size()), qe(queries.size());
        for (std::size_t i = 0; i < queries.size(); ++i) {
            qs[i] = queries[i].get_start();
            qe[i] = queries[i].get_end();
        }
        if constexpr (stranded) {
            std::vector<std::uint8_t> qst(queries.size());
            for (std::size_t i = 0; i < queries.size(); ++i) qst[i] = static_cast<std::uint8_t>(queries[i].get_strand());
            return intersect_batch(qc.data(), qs.data(), qe.data(), qst.data(), queries.size());
        } else {
            return intersect_batch(qc.data(), qs.data(), qe.data(), queries.size());
        }
    }

    /// grove::flanking(query, index) with the default (always compatible) predicate, grove_flanking.ipp: nearest
    /// non-overlapping predecessor (largest end below query.start, first such key in the chain) and successor (first
    /// key starting past query.end).  One gg_flanking_batch call of one query.
    [[nodiscard]] gdt::flanking_query_result<key_type, data_type> flanking(const key_type& query, std::string_view index)
        requires(!stranded) {
        gdt::flanking_query_result<key_type, data_type> result{};
        auto it = indices.find(index);
        if (it == indices.end()) return result;  // null root: empty pair
        sync();
        const std::uint32_t c = it->second.id;
        const std::uint64_t s = query.get_start(), e = query.get_end();
        std::uint32_t pred = GG_NO_RANK, succ = GG_NO_RANK;
        check(gg_flanking_batch(dev.get(), &c, &s, &e, 1, &pred, &succ));
        if (pred != GG_NO_RANK) result.set_predecessor(it->second.chain[pred]);
        if (succ != GG_NO_RANK) result.set_successor(it->second.chain[succ]);
        return result;
    }

    /// NEW: batch form.  pred[i] / succ[i] = chain rank in index q_index[i] (see chain()), GG_NO_RANK when absent.
    void flanking_batch(const std::uint32_t* q_index, const std::uint64_t* q_start, const std::uint64_t* q_end, std::size_t nq,
                        std::uint32_t* pred, std::uint32_t* succ)
        requires(!stranded) {
        sync();
        check(gg_flanking_batch(dev.get(), q_index, q_start, q_end, nq, pred, succ));
    }

    /// grove<genomic_coordinate, T>::flanking(query, index, pred) for pred = strand_compatible (grove_flanking.ipp:47-82):
    /// nearest same-strand (or wildcard) neighbours on either side.  One gg_flanking_batch_stranded call of one query.
    /// The default predicate is not offered on these keys: the reference's own answer depends on its tree's shape there
    /// (see gg_b200.h), so there is nothing well-defined to reproduce.
    [[nodiscard]] gdt::flanking_query_result<key_type, data_type> flanking(const key_type& query, std::string_view index, strand_compatible_t)
        requires(stranded) {
        gdt::flanking_query_result<key_type, data_type> result{};
        auto it = indices.find(index);
        if (it == indices.end()) return result;
        sync();
        const std::uint32_t c = it->second.id;
        const std::uint64_t s = query.get_start(), e = query.get_end();
        const std::uint8_t qst = static_cast<std::uint8_t>(query.get_strand());
        std::uint32_t pred = GG_NO_RANK, succ = GG_NO_RANK;
        check(gg_flanking_batch_stranded(dev.get(), &c, &s, &e, &qst, 1, &pred, &succ));
        if (pred != GG_NO_RANK) result.set_predecessor(it->second.chain[pred]);
        if (succ != GG_NO_RANK) result.set_successor(it->second.chain[succ]);
        return result;
    }
    void flanking_batch(const std::uint32_t* q_index, const std::uint64_t* q_start, const std::uint64_t* q_end,
                        const std::uint8_t* q_strand, std::size_t nq, std::uint32_t* pred, std::uint32_t* succ, strand_compatible_t)
        requires(stranded) {
        sync();
        check(gg_flanking_batch_stranded(dev.get(), q_index, q_start, q_end, q_strand, nq, pred, succ));
    }

    /// Take over one index whose keys are already in LEAF-CHAIN order — what a serialized grove's leaf blocks
    /// hold (gg_loader.hpp) or what a caller flattened from a reference grove.  make_key() is called n times and
    /// yields the keys in chain order; `shape` is the tree in pre-order tokens (leaf: chain_shape::leaf_bit | key
    /// count, internal node: child count; bottom_up_shape() gives the tokens of a freshly bulk-built tree), which
    /// later sorted appends, regular inserts and removals need to place tied keys exactly where the reference
    /// does.  The index must not exist yet; n == 0 creates nothing.
    template <typename MakeKey>
    std::vector<key_ptr> adopt_chain(std::string_view index, std::size_t n, std::span<const std::uint32_t> shape, MakeKey&& make_key) {
        if (indices.find(index) != indices.end()) throw std::invalid_argument("adopt_chain: index already exists");
        std::size_t leaf_keys = 0, last_leaf = 0;
        for (const std::uint32_t t : shape)
            if ((t & detail::chain_shape::leaf_bit) != 0) leaf_keys += (last_leaf = t & ~detail::chain_shape::leaf_bit);
        if (leaf_keys != n) throw std::invalid_argument("adopt_chain: the shape's leaves do not hold n keys");
        std::vector<key_ptr> adopted;
        if (n == 0) return adopted;
        adopted.reserve(n);
        for (std::size_t i = 0; i < n; ++i) adopted.push_back(allocate_key(make_key()));
        chain_state& ch = chain_for(index);
        ch.chain = adopted;
        ch.lazy_tokens.assign(shape.begin(), shape.end());
        ch.leaf_begin = n - last_leaf;
        leaf_key_count += n;
        touch(ch);
        return adopted;
    }
    /// pre-order tokens of the tree a sorted bulk insert of n keys builds on an empty index
    [[nodiscard]] std::vector<std::uint32_t> bottom_up_shape(std::size_t n) const {
        detail::chain_shape s(order);
        s.build_bottom_up(n);
        return s.tokens();
    }
    /// pre-order tokens of one index's tree as this side tracks it (tests compare them with the reference's real tree)
    [[nodiscard]] std::vector<std::uint32_t> shape_tokens(std::string_view index) {
        auto it = indices.find(index);
        if (it == indices.end()) return {};
        flush_pending(it->second);
        materialise(it->second);
        return it->second.shape.tokens();
    }

    /// Leaf chain of one index as the device sees it (rank -> key*); empty when the index does not exist.
    [[nodiscard]] const std::vector<key_ptr>& chain(std::string_view index) {
        static const std::vector<key_ptr> none;
        auto it = indices.find(index);
        if (it == indices.end()) return none;
        flush_pending(it->second);
        return it->second.chain;
    }

  private:
    struct chain_state {
        explicit chain_state(int order) : shape(order) {}
        std::vector<key_ptr> chain;    // leaf-chain order
        std::vector<key_ptr> pending;  // regular inserts not yet merged into the chain, in insertion order
        std::size_t leaf_begin = 0;    // first rank of the rightmost leaf: where a sorted append starts its lower_bound
        // Tree shape.  While an index has only seen bulk / sorted / adopted input its shape has a closed description
        // (tokens or a bottom-up build, then lazy_appends sorted appends) and leaf_begin follows from the append rule
        // alone; the first regular insert or removal replays that description into `shape`, which is kept from then on.
        detail::chain_shape shape;
        bool shape_live = false;
        std::size_t lazy_bottom_up = 0;
        std::vector<std::uint32_t> lazy_tokens;
        std::size_t lazy_appends = 0;
        std::uint32_t id = 0;          // chromosome id in the device snapshot: assigned at creation, never renumbered
        bool dirty = true;             // the snapshot does not hold this chain as it is now
    };
    struct handle_deleter {
        void operator()(gg_index* h) const noexcept { gg_index_destroy(h); }
    };

    void need_device() const {
        if (!dev) throw device_error(GG_ERR_NOT_BUILT, "genogrove_b200::grove: detached grove has no device (there is no CPU search path)");
    }

    void check(gg_status st) const {
        if (st == GG_OK) return;
        const std::string msg = gg_last_error(dev.get());
        if (st == GG_ERR_INVALID_INTERVAL) throw std::invalid_argument(msg.empty() ? "interval: start must be <= end" : msg);
        throw device_error(st, "gg_b200: " + (msg.empty() ? std::string("status ") + std::to_string(st) : msg));
    }

    key_ptr allocate_key(key_t k) {
        key_storage.push_back(std::move(k));
        return &key_storage.back();
    }

    chain_state& chain_for(std::string_view index) {
        auto it = indices.find(index);
        if (it != indices.end()) return it->second;
        dirty = true;
        chain_state& ch = indices.emplace(std::string(index), chain_state{order}).first->second;
        ch.id = next_id++;
        chain_by_id.push_back(&ch.chain);  // (map nodes do not move: the pointer stays valid until the index is erased)
        return ch;
    }

    void touch(chain_state& ch) {
        ch.dirty = true;
        dirty = true;
    }
    static const std::vector<key_ptr>& no_chain() {
        static const std::vector<key_ptr> none;
        return none;
    }

    bool all_pending_empty() const {
        for (const auto& [n, ch] : indices)
            if (!ch.pending.empty()) return false;
        return true;
    }

    static bool value_less(const key_ptr a, const key_ptr b) { return a->get_value() < b->get_value(); }

    /// Resolve the regular inserts since the last flush.  Sequential lower_bound insertion puts every new key in
    /// front of all equal keys present at that moment, so the final order is: pending keys by (start, end) with
    /// LATER insertions first among equals, merged into the chain with pending keys ahead of equal chain keys.
    void flush_pending(chain_state& ch) {
        if (ch.pending.empty()) return;
        std::vector<key_ptr> p(ch.pending.rbegin(), ch.pending.rend());
        std::stable_sort(p.begin(), p.end(), value_less);
        // replay the inserts on the tree shape, in insertion order: each landed at its lower_bound rank over the keys
        // present at that moment = chain keys below it + earlier pending keys below it (Fenwick tree over value order)
        materialise(ch);
        std::vector<std::size_t> fenwick(p.size() + 1, 0);
        for (const key_ptr k : ch.pending) {
            const std::size_t in_chain = static_cast<std::size_t>(std::lower_bound(ch.chain.begin(), ch.chain.end(), k, value_less) - ch.chain.begin());
            const std::size_t slot = static_cast<std::size_t>(std::lower_bound(p.begin(), p.end(), k, value_less) - p.begin());
            std::size_t earlier_below = 0;
            for (std::size_t i = slot; i > 0; i -= i & (~i + 1)) earlier_below += fenwick[i];
            ch.shape.insert_at(in_chain + earlier_below);
            for (std::size_t i = slot + 1; i <= p.size(); i += i & (~i + 1)) ++fenwick[i];
        }
        std::vector<key_ptr> merged;
        merged.reserve(ch.chain.size() + p.size());
        // std::merge is stable with the FIRST range winning ties: pending first
        std::merge(p.begin(), p.end(), ch.chain.begin(), ch.chain.end(), std::back_inserter(merged), value_less);
        ch.chain.swap(merged);
        ch.pending.clear();
        ch.leaf_begin = ch.chain.size() - ch.shape.rightmost_leaf_keys();
    }

    /// make ch.shape the explicit twin of the index's tree (see chain_state)
    void materialise(chain_state& ch) {
        if (ch.shape_live) return;
        if (!ch.lazy_tokens.empty()) ch.shape.build_from_tokens(ch.lazy_tokens);
        else ch.shape.build_bottom_up(ch.lazy_bottom_up);
        for (std::size_t i = 0; i < ch.lazy_appends; ++i) ch.shape.append_sorted();
        if (ch.shape.size() != ch.chain.size() || ch.chain.size() - ch.shape.rightmost_leaf_keys() != ch.leaf_begin)
            throw std::logic_error("genogrove_b200::grove: tree shape out of step with the leaf chain");
        ch.lazy_tokens.clear();
        ch.lazy_tokens.shrink_to_fit();
        ch.lazy_bottom_up = ch.lazy_appends = 0;
        ch.shape_live = true;
    }

    /// one key was appended through the sorted path: the rightmost leaf grew, and splits leaving ONE key right at `order`
    void note_sorted_append(chain_state& ch) {
        if (ch.shape_live) {
            ch.shape.append_sorted();
            ch.leaf_begin = ch.chain.size() - ch.shape.rightmost_leaf_keys();
        } else {
            ++ch.lazy_appends;
            if (ch.chain.size() - ch.leaf_begin == static_cast<std::size_t>(order)) ch.leaf_begin = ch.chain.size() - 1;
        }
    }

    /// position of a sorted append: lower_bound inside the rightmost leaf (node::insert_key_ptr on it)
    std::size_t sorted_position(const chain_state& ch, const key_ptr k) const {
        return static_cast<std::size_t>(std::lower_bound(ch.chain.begin() + static_cast<std::ptrdiff_t>(ch.leaf_begin), ch.chain.end(), k, value_less) -
                                        ch.chain.begin());
    }

    key_ptr insert_sorted(std::string_view index, key_ptr k) {
        chain_state& ch = chain_for(index);
        flush_pending(ch);
        ch.chain.insert(ch.chain.begin() + static_cast<std::ptrdiff_t>(sorted_position(ch, k)), k);
        note_sorted_append(ch);
        ++leaf_key_count;
        touch(ch);
        return k;
    }

    /// sorted-bulk: bottom-up build on an empty index (chain = input order, rightmost leaf = last group of the
    /// even distribution, grove.hpp:100-106), else the rightmost-append loop (grove_insert.ipp:106-126).
    void append_sorted_run(std::string_view index, const std::vector<key_ptr>& run, bool device_is_current) {
        const bool fresh_grove = indices.empty() && next_id == 0;
        chain_state& ch = chain_for(index);
        flush_pending(ch);
        if (ch.chain.empty()) {
            const std::size_t n = run.size(), cap = static_cast<std::size_t>(order - 1);
            const std::size_t groups = n / cap + (n % cap != 0 ? 1 : 0);
            ch.chain = run;
            ch.leaf_begin = n - n / groups;
            ch.shape.clear();
            ch.shape_live = false;
            ch.lazy_tokens.clear();
            ch.lazy_bottom_up = n;
            ch.lazy_appends = 0;
        } else {
            for (key_ptr k : run) {
                ch.chain.insert(ch.chain.begin() + static_cast<std::ptrdiff_t>(sorted_position(ch, k)), k);
                note_sorted_append(ch);
            }
        }
        leaf_key_count += run.size();
        touch(ch);
        if (device_is_current && fresh_grove) {  // the sort that ordered this batch already left its snapshot on the device
            ch.dirty = false;
            dirty = false;
            snapshot_valid = true;
            device_n_chrom = 1;
        }
    }

    /// Bring the device snapshot up to date.  Every index has its own dirty flag: when few of them changed (and they hold
    /// at most half the keys) only those chains are flattened and uploaded, one gg_index_update_chrom each — the device
    /// splices them in and rebuilds its levels at HBM speed; otherwise all chains go up in one gg_index_upload_sorted.
    void sync() {
        need_device();
        if (!dirty) return;
        std::size_t total = 0, changed = 0, n_changed = 0;
        std::vector<chain_state*> by_id(next_id, nullptr);
        for (auto& [name, ch] : indices) {
            flush_pending(ch);
            by_id[ch.id] = &ch;
            total += ch.chain.size();
            if (ch.dirty) {
                changed += ch.chain.size();
                ++n_changed;
            }
        }
        auto flatten = [](const chain_state& ch, std::vector<std::uint64_t>& s, std::vector<std::uint64_t>& e) {
            for (const key_ptr k : ch.chain) {
                s.push_back(k->get_value().get_start());
                e.push_back(k->get_value().get_end());
            }
        };
        const std::size_t new_ids = next_id - std::min(next_id, device_n_chrom);
        bool partial = snapshot_valid && n_changed + emptied_ids.size() + new_ids <= max_partial_updates && 2 * changed <= total;
        if (partial) {
            // GG_ERR_NOT_BUILT: the handle no longer holds the chains (its build workspace was dropped) -> whole upload below
            auto update = [&](std::uint32_t id, const std::uint64_t* s, const std::uint64_t* e, std::size_t n) {
                const gg_status st = gg_index_update_chrom(dev.get(), id, s, e, n);
                if (st == GG_ERR_NOT_BUILT) return false;
                check(st);
                return true;
            };
            std::vector<std::uint64_t> s, e;
            for (const std::uint32_t id : emptied_ids)
                if (partial && id < device_n_chrom) partial = update(id, nullptr, nullptr, 0);
            for (std::uint32_t id = 0; partial && id < next_id; ++id) {
                chain_state* ch = by_id[id];
                if (id < device_n_chrom && (ch == nullptr || !ch->dirty)) continue;
                s.clear(), e.clear();
                if (ch != nullptr) flatten(*ch, s, e);  // (an id created and retired between two syncs is appended empty)
                partial = update(id, s.data(), e.data(), s.size());
                if (partial) ++partial_updates;
            }
        }
        if (!partial) {
            std::vector<std::uint32_t> c;
            std::vector<std::uint64_t> s, e;
            c.reserve(total), s.reserve(total), e.reserve(total);
            for (std::uint32_t id = 0; id < next_id; ++id) {
                if (by_id[id] == nullptr) continue;
                flatten(*by_id[id], s, e);
                c.resize(s.size(), id);
            }
            check(gg_index_upload_sorted(dev.get(), c.data(), s.data(), e.data(), s.size(), std::max<std::uint32_t>(next_id, 1u)));
            ++full_uploads;
        }
        if constexpr (stranded) {
            std::vector<std::uint8_t> st;
            st.reserve(total);
            for (std::uint32_t id = 0; id < next_id; ++id)
                if (by_id[id] != nullptr)
                    for (const key_ptr k : by_id[id]->chain) st.push_back(static_cast<std::uint8_t>(k->get_value().get_strand()));
            check(gg_index_set_strands(dev.get(), st.data(), st.size()));
        }
        for (auto& [name, ch] : indices) ch.dirty = false;
        emptied_ids.clear();
        device_n_chrom = next_id;
        snapshot_valid = true;
        dirty = false;
    }

  public:
    /// how the snapshot was refreshed so far (tests / cost accounting): whole-grove uploads and single-index updates
    [[nodiscard]] std::size_t full_upload_count() const noexcept { return full_uploads; }
    [[nodiscard]] std::size_t partial_update_count() const noexcept { return partial_updates; }
    [[nodiscard]] std::size_t device_build_count() const noexcept { return device_builds; }
    /// more changed indices than this between two queries -> one whole upload instead (each update rebuilds the levels)
    std::size_t max_partial_updates = 4;

  private:
    int order;
    std::unique_ptr<gg_index, handle_deleter> dev;
    /// index name -> chain; same key type, hash and insert/erase history as the reference's root_nodes map
    /// (grove.hpp:321), hence the same iteration order for intersect(query) across all indices
    std::unordered_map<std::string, chain_state, string_hash, std::equal_to<>> indices;
    std::vector<const std::vector<key_ptr>*> chain_by_id;  // index id -> chain (retired ids: an empty chain)
    std::deque<key_t> key_storage;
    std::size_t leaf_key_count = 0;
    bool dirty = true;             // some chain differs from the snapshot
    bool snapshot_valid = false;   // the handle holds a snapshot of THIS grove (not of a lone bulk batch)
    std::uint32_t next_id = 0;     // ids handed out so far
    std::uint32_t device_n_chrom = 0;
    std::vector<std::uint32_t> emptied_ids;
    std::size_t full_uploads = 0, partial_updates = 0, device_builds = 0;
    std::vector<std::uint32_t> scratch_c;
    std::vector<std::uint64_t> scratch_s, scratch_e, scratch_off;
};

}  // namespace structure
}  // namespace genogrove_b200

#endif  // GENOGROVE_B200_GROVE_HPP
