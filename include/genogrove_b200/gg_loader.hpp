// genogrove_b200/gg_loader.hpp — `.gg` (grove stream format 0.3) -> device snapshot, without a pointer tree.
//
// The reference restores a serialized grove by rebuilding every node object, linking child / next pointers and
// re-seating the key deque (grove_serialize.ipp:111-161,617-646; node.hpp:575-642); a caller that wants the
// device snapshot then has to walk that tree's leaf chain again.  The device index only needs
// the LEAF CHAIN ORDER of each index, and the stream already holds it: every node is one length-prefixed,
// independently deflated block, a leaf block carries its keys and the block id of the next leaf, an internal
// block carries its children's block ids (node.hpp:527-571).  load_gg() therefore
//   1. reads the plain directory (magic, order, index names + root block ids, block counts),
//   2. inflates and parses the node blocks in parallel (blocks are independent; internal keys, per-key edge
//      records and external-key blocks are skipped, nothing is allocated per node),
//   3. per index walks the child block ids depth-first from the root, which yields the leaves in chain order
//      (checked against the leaves' `next` ids) and the tree's shape as pre-order counts, and hands the keys in
//      that order to grove::adopt_chain together with the shape (what later sorted appends, inserts and removals
//      need to place tied keys exactly where the reference puts them, chain_shape.hpp),
// and the grove's ordinary sync() uploads the chains with gg_index_upload_sorted.  Index names enter the grove in
// directory order, the order the reference's deserialize() inserts them into its root map, so intersect(query)
// over all indices visits them in the same order on both sides.
//
// Stream layout (restated from the reference's writer, grove_serialize.ipp:267-345; all fields little-endian, raw):
//   optional CLI wrapper  "GROV" + 8 bytes (io/gg_format.hpp:24-40; byte 9 = payload type)     -> skipped
//   "GGB\x03" | int order | u32 n_indices | n x { u32 len, name, u32 root_block } |
//   u32 n_blocks | u32 ext_block_begin | u64 leaf_keys | u64 external_keys |
//   n_blocks x { u64 clen, zlib stream }            blocks [0, ext_block_begin) are nodes, DFS pre-order per index
//   node block: u32 (bit 31 = leaf | n_keys), n_keys x { key, payload }, then
//               leaf: u32 next_block (0xFFFFFFFF = none), per-key edge lists (not needed here)
//               internal: u32 n_children (= n_keys + 1), n_children x u32 child_block
//   key:     interval = u64 start, u64 end (src/data_type/interval.cpp:18-24);
//            genomic_coordinate = char strand, u64 start, u64 end (src/data_type/genomic_coordinate.cpp:15-22)
//   payload: trivially copyable = raw bytes; std::string = u32 length + bytes (serialization_traits.hpp:118-160);
//            anything else: specialise genogrove_b200::io::gg_payload<T>.
// Graph edges and external keys are outside the intersect path and are not restored.
// Needs zlib (-lz).  Malformed input -> std::runtime_error, as the reference reports it.
#ifndef GENOGROVE_B200_GG_LOADER_HPP
#define GENOGROVE_B200_GG_LOADER_HPP

#include <zlib.h>

#include <chrono>
#include <cstring>
#include <istream>
#include <iterator>
#include <thread>

#include "grove.hpp"

namespace genogrove_b200::io {

/// How a payload type is laid out in a block.  read() consumes exactly the bytes the reference's
/// serializer<T>::write produced and must throw std::runtime_error when fewer than that remain.
template <typename T>
struct gg_payload {
    static_assert(std::is_trivially_copyable_v<T>, "specialise genogrove_b200::io::gg_payload<T> for this payload type");
    static T read(const unsigned char*& cur, const unsigned char* end) {
        if (static_cast<std::size_t>(end - cur) < sizeof(T)) throw std::runtime_error("gg_loader: block truncated inside a payload");
        T v;
        std::memcpy(&v, cur, sizeof(T));
        cur += sizeof(T);
        return v;
    }
    static void skip(const unsigned char*& cur, const unsigned char* end) {
        if (static_cast<std::size_t>(end - cur) < sizeof(T)) throw std::runtime_error("gg_loader: block truncated inside a payload");
        cur += sizeof(T);
    }
};
template <>
struct gg_payload<std::string> {
    static std::uint32_t length(const unsigned char*& cur, const unsigned char* end) {
        std::uint32_t len;
        if (static_cast<std::size_t>(end - cur) < sizeof(len)) throw std::runtime_error("gg_loader: block truncated inside a string length");
        std::memcpy(&len, cur, sizeof(len));
        cur += sizeof(len);
        if (static_cast<std::size_t>(end - cur) < len) throw std::runtime_error("gg_loader: string length exceeds the block");
        return len;
    }
    static std::string read(const unsigned char*& cur, const unsigned char* end) {
        const std::uint32_t len = length(cur, end);
        std::string s(reinterpret_cast<const char*>(cur), len);
        cur += len;
        return s;
    }
    static void skip(const unsigned char*& cur, const unsigned char* end) { cur += length(cur, end); }
};

/// What the directory said and where the load time went.
struct gg_stream_info {
    int order = 0;
    bool cli_wrapper = false;          ///< stream began with the CLI's "GROV" header
    std::uint8_t payload_type = 0;     ///< its payload byte (1 = BED, 2 = GFF), 0 without wrapper
    std::uint32_t num_blocks = 0, ext_block_begin = 0;
    std::uint64_t leaf_keys = 0, external_keys = 0;
    std::vector<std::pair<std::string, std::uint32_t>> index_roots;
    std::size_t compressed_bytes = 0, inflated_bytes = 0;
    unsigned threads = 0;
    double read_ms = 0, parse_ms = 0, adopt_ms = 0;
};

namespace detail {

inline constexpr std::uint32_t no_block = 0xFFFFFFFFu;

template <typename T>
T take(const unsigned char*& cur, const unsigned char* end, const char* what) {
    if (static_cast<std::size_t>(end - cur) < sizeof(T)) throw std::runtime_error(std::string("gg_loader: truncated ") + what);
    T v;
    std::memcpy(&v, cur, sizeof(T));
    cur += sizeof(T);
    return v;
}

/// one inflate state reused for every block of a worker (inflateReset per block, like zlib_streambuf.hpp:108-171)
class inflater {
  public:
    inflater() {
        std::memset(&z, 0, sizeof(z));
        if (inflateInit(&z) != Z_OK) throw std::runtime_error("gg_loader: inflateInit failed");
    }
    ~inflater() { inflateEnd(&z); }
    inflater(const inflater&) = delete;
    inflater& operator=(const inflater&) = delete;
    /// inflate [src, src+n) into buf (grown as needed); returns the inflated size
    std::size_t run(const unsigned char* src, std::size_t n, std::vector<unsigned char>& buf) {
        if (inflateReset(&z) != Z_OK) throw std::runtime_error("gg_loader: inflateReset failed");
        if (buf.size() < 4096) buf.resize(4096);
        z.next_in = const_cast<unsigned char*>(src);
        z.avail_in = static_cast<uInt>(n);
        std::size_t produced = 0;
        for (;;) {
            z.next_out = buf.data() + produced;
            z.avail_out = static_cast<uInt>(buf.size() - produced);
            const int rc = inflate(&z, Z_NO_FLUSH);
            produced = buf.size() - z.avail_out;
            if (rc == Z_STREAM_END) break;
            if (rc != Z_OK && rc != Z_BUF_ERROR) throw std::runtime_error("gg_loader: corrupt compressed block");
            if (z.avail_out == 0) {
                buf.resize(buf.size() * 2);
            } else if (z.avail_in == 0) {
                throw std::runtime_error("gg_loader: compressed block ends early");
            }
        }
        return produced;
    }

  private:
    z_stream z;
};

/// per node block: what the chain assembly needs
struct block_meta {
    std::uint32_t next = no_block;  ///< leaf: next leaf's block id
    std::uint32_t n_keys = 0;
    bool is_leaf = false;
    std::uint64_t begin = 0;        ///< leaf: position of its first key in the owning worker's key arrays;
                                    ///< internal: position of its first child id in the worker's child array
};

}  // namespace detail

/// Restore a grove from a format-0.3 stream straight into leaf-chain order (see the header comment).
/// `threads` = 0 picks min(hardware threads, 16).
template <typename key_type, typename data_type = void>
[[nodiscard]] structure::grove<key_type, data_type> load_gg(std::istream& is, int device = 0, gg_stream_info* info_out = nullptr,
                                                            unsigned threads = 0) {
    using clock = std::chrono::steady_clock;
    using grove_t = structure::grove<key_type, data_type>;
    using key_t = typename grove_t::key_t;
    constexpr bool stranded = std::is_same_v<key_type, genogrove_b200::data_type::genomic_coordinate>;
    constexpr bool has_payload = !std::is_void_v<data_type>;
    using stored_t = std::conditional_t<has_payload, data_type, char>;
    const auto t0 = clock::now();

    // ---- 1. the whole stream into memory (the directory is a few bytes; the blocks are parsed from here)
    std::vector<unsigned char> file;
    {
        std::string chunk(1 << 20, '\0');
        while (is) {
            is.read(chunk.data(), static_cast<std::streamsize>(chunk.size()));
            const std::streamsize got = is.gcount();
            if (got <= 0) break;
            file.insert(file.end(), chunk.data(), chunk.data() + got);
        }
    }
    const unsigned char* cur = file.data();
    const unsigned char* const end = file.data() + file.size();
    gg_stream_info info;
    if (file.size() >= 12 && std::memcmp(cur, "GROV", 4) == 0) {  // CLI wrapper, io/gg_format.hpp
        if (cur[4] != 0 || cur[5] != 3) throw std::runtime_error("gg_loader: unsupported .gg format version (this loader reads 0.3)");
        info.cli_wrapper = true;
        info.payload_type = cur[9];
        cur += 12;
    }
    if (end - cur < 4 || std::memcmp(cur, "GGB\x03", 4) != 0)
        throw std::runtime_error("Failed to deserialize grove: bad magic (not a format 0.3 grove stream)");
    cur += 4;
    info.order = detail::take<int>(cur, end, "order");
    if (info.order < 3) throw std::runtime_error("Failed to deserialize grove: invalid order");
    const auto n_indices = detail::take<std::uint32_t>(cur, end, "index count");
    for (std::uint32_t i = 0; i < n_indices; ++i) {
        const auto len = detail::take<std::uint32_t>(cur, end, "index name length");
        if (static_cast<std::size_t>(end - cur) < len) throw std::runtime_error("gg_loader: truncated index name");
        std::string name(reinterpret_cast<const char*>(cur), len);
        cur += len;
        info.index_roots.emplace_back(std::move(name), detail::take<std::uint32_t>(cur, end, "root block id"));
    }
    info.num_blocks = detail::take<std::uint32_t>(cur, end, "block count");
    info.ext_block_begin = detail::take<std::uint32_t>(cur, end, "external block begin");
    info.leaf_keys = detail::take<std::uint64_t>(cur, end, "leaf key count");
    info.external_keys = detail::take<std::uint64_t>(cur, end, "external key count");
    if (info.ext_block_begin > info.num_blocks) throw std::runtime_error("gg_loader: external block begin exceeds the block count");
    const std::uint32_t n_nodes = info.ext_block_begin;
    // block extents: one cheap sequential pass over the length prefixes (node blocks only)
    std::vector<std::pair<const unsigned char*, std::size_t>> extent(n_nodes);
    for (std::uint32_t b = 0; b < n_nodes; ++b) {
        const auto clen = detail::take<std::uint64_t>(cur, end, "block length");
        if (clen > static_cast<std::uint64_t>(end - cur))
            throw std::runtime_error("Failed to deserialize grove: compressed block length exceeds remaining stream");
        extent[b] = {cur, static_cast<std::size_t>(clen)};
        cur += clen;
        info.compressed_bytes += clen;
    }
    const auto t1 = clock::now();

    // ---- 2. inflate + parse, blocks split into contiguous ranges over the workers
    if (threads == 0) threads = std::min(16u, std::max(1u, std::thread::hardware_concurrency()));
    threads = std::max(1u, std::min<unsigned>(threads, n_nodes / 256 + 1));
    info.threads = threads;
    struct worker {
        std::vector<std::uint64_t> start, stop;
        std::vector<char> strand;
        std::vector<stored_t> payload;
        std::vector<std::uint32_t> children;
        std::size_t inflated = 0;
        std::string error;
    };
    std::vector<worker> work(threads);
    std::vector<detail::block_meta> meta(n_nodes);
    std::vector<std::uint32_t> owner(n_nodes);
    const std::uint32_t order_u = static_cast<std::uint32_t>(info.order);
    auto parse_range = [&](unsigned w, std::uint32_t b0, std::uint32_t b1) {
        worker& me = work[w];
        try {
            detail::inflater z;
            std::vector<unsigned char> raw;
            for (std::uint32_t b = b0; b < b1; ++b) {
                const std::size_t n = z.run(extent[b].first, extent[b].second, raw);
                me.inflated += n;
                const unsigned char* p = raw.data();
                const unsigned char* const pe = raw.data() + n;
                const auto packed = detail::take<std::uint32_t>(p, pe, "node header");
                detail::block_meta& m = meta[b];
                m.is_leaf = (packed & 0x80000000u) != 0;
                m.n_keys = packed & 0x7FFFFFFFu;
                owner[b] = w;
                if (m.n_keys >= order_u) throw std::runtime_error("Failed to deserialize node block: num_keys exceeds order");
                m.begin = m.is_leaf ? me.start.size() : me.children.size();
                for (std::uint32_t k = 0; k < m.n_keys; ++k) {
                    char st = '.';
                    if constexpr (stranded) st = detail::take<char>(p, pe, "key strand");
                    const auto s = detail::take<std::uint64_t>(p, pe, "key start");
                    const auto e = detail::take<std::uint64_t>(p, pe, "key end");
                    if (m.is_leaf) {
                        me.start.push_back(s);
                        me.stop.push_back(e);
                        if constexpr (stranded) me.strand.push_back(st);
                        if constexpr (has_payload) me.payload.push_back(gg_payload<data_type>::read(p, pe));
                    } else {
                        if constexpr (has_payload) gg_payload<data_type>::skip(p, pe);
                    }
                }
                if (m.is_leaf) {
                    m.next = detail::take<std::uint32_t>(p, pe, "next leaf id");  // edge records follow; not needed
                } else {
                    const auto n_children = detail::take<std::uint32_t>(p, pe, "child count");
                    if (n_children != m.n_keys + 1)
                        throw std::runtime_error("Failed to deserialize node block: child count does not match key count");
                    for (std::uint32_t c = 0; c < n_children; ++c) me.children.push_back(detail::take<std::uint32_t>(p, pe, "child id"));
                }
            }
        } catch (const std::exception& ex) {
            me.error = ex.what();
        }
    };
    {
        std::vector<std::thread> pool;
        for (unsigned w = 0; w < threads; ++w) {
            const std::uint32_t b0 = static_cast<std::uint32_t>(std::uint64_t(n_nodes) * w / threads);
            const std::uint32_t b1 = static_cast<std::uint32_t>(std::uint64_t(n_nodes) * (w + 1) / threads);
            if (w + 1 == threads) parse_range(w, b0, b1);
            else pool.emplace_back(parse_range, w, b0, b1);
        }
        for (auto& t : pool) t.join();
    }
    std::uint64_t parsed_leaf_keys = 0;
    for (const worker& w : work) {
        if (!w.error.empty()) throw std::runtime_error(w.error);
        parsed_leaf_keys += w.start.size();
        info.inflated_bytes += w.inflated;
    }
    if (parsed_leaf_keys != info.leaf_keys) throw std::runtime_error("Failed to deserialize grove: leaf key count does not match the directory");
    const auto t2 = clock::now();

    // ---- 3. per index: depth-first over the child ids = leaves in chain order + the tree shape in pre-order
    grove_t g(info.order, device);
    std::vector<std::uint32_t> leaves, shape, stack;
    std::vector<bool> visited(n_nodes, false);
    std::uint64_t adopted = 0;
    for (const auto& [name, root] : info.index_roots) {
        leaves.clear();
        shape.clear();
        stack.assign(1, root);
        std::uint64_t n = 0;
        while (!stack.empty()) {
            const std::uint32_t b = stack.back();
            stack.pop_back();
            if (b >= n_nodes) throw std::runtime_error("Failed to deserialize grove: invalid child block id");
            if (visited[b]) throw std::runtime_error("Failed to deserialize grove: node block referenced twice");
            visited[b] = true;
            const detail::block_meta& m = meta[b];
            if (m.is_leaf) {
                if (!leaves.empty() && meta[leaves.back()].next != b)
                    throw std::runtime_error("Failed to deserialize grove: leaf chain does not follow the tree");
                leaves.push_back(b);
                shape.push_back(structure::detail::chain_shape::leaf_bit | m.n_keys);
                n += m.n_keys;
            } else {
                shape.push_back(m.n_keys + 1);
                const std::uint32_t* kids = work[owner[b]].children.data() + m.begin;
                for (std::uint32_t c = m.n_keys + 1; c-- > 0;) stack.push_back(kids[c]);
            }
        }
        if (leaves.empty() || meta[leaves.back()].next != detail::no_block)
            throw std::runtime_error("Failed to deserialize grove: leaf chain does not end with the tree");
        adopted += n;
        if (adopted > info.leaf_keys) throw std::runtime_error("Failed to deserialize grove: leaf chains hold more keys than the directory");
        // generator over (leaf, slot) in chain order
        std::size_t li = 0;
        std::uint32_t slot = 0;
        auto next_key = [&]() -> key_t {
            while (slot >= meta[leaves[li]].n_keys) {
                ++li;
                slot = 0;
            }
            const detail::block_meta& m = meta[leaves[li]];
            worker& w = work[owner[leaves[li]]];
            const std::size_t at = static_cast<std::size_t>(m.begin) + slot++;
            key_type value = [&] {
                if constexpr (stranded) return key_type(w.strand[at], w.start[at], w.stop[at]);
                else return key_type(w.start[at], w.stop[at]);
            }();
            if constexpr (has_payload) return key_t(std::move(value), std::move(w.payload[at]));
            else return key_t(std::move(value));
        };
        g.adopt_chain(name, static_cast<std::size_t>(n), shape, next_key);
    }
    if (adopted != info.leaf_keys) throw std::runtime_error("Failed to deserialize grove: leaf chains do not cover every leaf key");
    const auto t3 = clock::now();
    auto ms = [](auto a, auto b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
    info.read_ms = ms(t0, t1);
    info.parse_ms = ms(t1, t2);
    info.adopt_ms = ms(t2, t3);
    if (info_out != nullptr) *info_out = std::move(info);
    return g;
}

}  // namespace genogrove_b200::io

#endif  // GENOGROVE_B200_GG_LOADER_HPP
