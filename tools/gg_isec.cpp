// gg_isec — batched `genogrove intersect` for BED files on the B200 path.
//
// Mirrors the reference CLI's target-file mode (cli/src/intersect.cpp:39-58,196-221 with the BED handlers
// cli/src/handlers/bed.cpp:11-38, cli/include/handlers/bed.hpp:50-62): every target record is inserted as
// interval(start, end - 1) under its chromosome name (BED is 0-based half-open, the grove is closed), every query
// record asks intersect(interval(start, end - 1), chrom), and each hit prints "chrom\tstart\tend\n" of the TARGET
// record, hits in leaf-chain order, queries in file order.  The difference is the shape of the work: the reference
// calls grove.intersect once per query line; this reads the whole query file into structure-of-arrays buffers and
// makes ONE grove.intersect_batch_compact call (u32 coordinates up, compact hits down; the u64 intersect_batch call when a
// coordinate needs more than 32 bits), then prints from the result in place.
//
//   gg_isec -q query.{bed,gff,gtf,sam} -t target.{bed,gff,gtf,sam} [-o out] [-k order] [-s]   (-s: target is sorted -> sorted inserts)
//
// Text SAM (by extension .sam; read-shaped queries, BASELINE configs[3]) follows the reference reader's coordinate rule
// (src/io/bam_reader.cpp:180-211,261-282, include/genogrove/io/bam_reader.hpp:244-248,283-287): start = POS - 1 (0-based),
// end = start + the reference bases the CIGAR consumes (M, D, N, =, X; bam_reader.hpp:112-116), half-open; unmapped
// records (FLAG & 4) are skipped as by the reader's default options; a record becomes the closed
// interval(start, end - 1) ONLY when it consumes reference (start < end): pure soft-clip ("100S") and hard-clip-only
// secondary records yield [pos, pos) and ask nothing (tests/io/bamfile-test.cpp:676-765).  A SAM target prints
// RNAME, start, end (0-based half-open) per hit.
//
// GFF/GTF files (by extension .gff / .gff3 / .gtf) are accepted on either side, as in the reference CLI: 1-based
// inclusive [start, end] becomes interval(start - 1, end - 1) (cli/src/handlers/gff.cpp:22-30,
// cli/include/handlers/gff.hpp:52-57), so BED and GFF records meet in one 0-based closed space and cross-type queries
// work; a hit prints the target record's own seqid / start / end (gff.hpp:61-63, bed.hpp:60-62).
// BED parsing follows io/bed_reader.cpp:395-431 for the three mandatory columns: tab/space separated, digits only,
// start < end, else the run fails with the reader's message.  Lines that are empty or start with '#', "track" or
// "browser" are skipped.
#include <charconv>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <stdexcept>
#include <string>
#include <string_view>
#include <vector>

#include "genogrove_b200/grove.hpp"

namespace gdt = genogrove_b200::data_type;
namespace gst = genogrove_b200::structure;

struct bed3 {  // one target / query record: the printed fields + its 0-based closed interval
    std::string chrom;
    std::size_t start = 0, end = 0;  // as written in the file (what a hit prints)
    std::size_t lo = 0, hi = 0;      // canonical closed interval: BED [start, end - 1], GFF [start - 1, end - 1]
};

static bool is_gff(const std::string& path) {
    auto ends = [&](const char* suf) {
        const std::size_t n = std::strlen(suf);
        return path.size() >= n && path.compare(path.size() - n, n, suf) == 0;
    };
    return ends(".gff") || ends(".gff3") || ends(".gtf");
}

static bool skip_line(std::string_view l) {
    return l.empty() || l[0] == '#' || l.substr(0, 5) == "track" || l.substr(0, 7) == "browser";
}

static std::string_view next_field(std::string_view l, std::size_t& pos) {
    while (pos < l.size() && (l[pos] == '\t' || l[pos] == ' ')) ++pos;
    const std::size_t b = pos;
    while (pos < l.size() && l[pos] != '\t' && l[pos] != ' ') ++pos;
    return l.substr(b, pos - b);
}

static std::size_t parse_coord(std::string_view f, std::size_t line_num) {
    std::size_t v = 0;
    if (f.empty() || f.find_first_not_of("0123456789") != std::string_view::npos)
        throw std::runtime_error("Invalid coordinate format at line " + std::to_string(line_num));
    auto [p, ec] = std::from_chars(f.data(), f.data() + f.size(), v);
    if (ec != std::errc{}) throw std::runtime_error("Coordinate out of range at line " + std::to_string(line_num));
    return v;
}

template <typename F>
static void for_each_bed(const std::string& path, F&& cb) {
    std::ifstream in(path);
    if (!in) throw std::runtime_error("Failed to open file: " + path);
    std::string line;
    std::size_t line_num = 0;
    while (std::getline(in, line)) {
        ++line_num;
        if (!line.empty() && line.back() == '\r') line.pop_back();
        if (skip_line(line)) continue;
        std::size_t pos = 0;
        const std::string_view l(line);
        const std::string_view c = next_field(l, pos), s = next_field(l, pos), e = next_field(l, pos);
        if (c.empty() || s.empty() || e.empty()) throw std::runtime_error("Invalid line format at line " + std::to_string(line_num));
        bed3 r{std::string(c), parse_coord(s, line_num), parse_coord(e, line_num), 0, 0};
        if (r.start >= r.end)
            throw std::runtime_error("Start coordinate is greater than or equal to the end coordinate at line " + std::to_string(line_num));
        r.lo = r.start;
        r.hi = r.end - 1;
        cb(std::move(r));
    }
}

// GFF3 / GTF: 9 tab-separated columns, seqid = 1, start = 4, end = 5 (1-based inclusive, start <= end, start >= 1)
template <typename F>
static void for_each_gff(const std::string& path, F&& cb) {
    std::ifstream in(path);
    if (!in) throw std::runtime_error("Failed to open file: " + path);
    std::string line;
    std::size_t line_num = 0;
    while (std::getline(in, line)) {
        ++line_num;
        if (!line.empty() && line.back() == '\r') line.pop_back();
        if (line.empty() || line[0] == '#') continue;
        std::string_view l(line), col[5];
        std::size_t pos = 0;
        int k = 0;
        for (; k < 5 && pos <= l.size(); ++k) {
            const std::size_t t = l.find('\t', pos);
            col[k] = l.substr(pos, t == std::string_view::npos ? std::string_view::npos : t - pos);
            if (t == std::string_view::npos) { ++k; break; }
            pos = t + 1;
        }
        if (k < 5 || col[0].empty()) throw std::runtime_error("Invalid line format at line " + std::to_string(line_num));
        bed3 r{std::string(col[0]), parse_coord(col[3], line_num), parse_coord(col[4], line_num), 0, 0};
        if (r.start < 1 || r.start > r.end)
            throw std::runtime_error("Invalid GFF coordinates at line " + std::to_string(line_num));
        r.lo = r.start - 1;
        r.hi = r.end - 1;
        cb(std::move(r));
    }
}

static bool is_sam(const std::string& path) { return path.size() >= 4 && path.compare(path.size() - 4, 4, ".sam") == 0; }

// reference bases consumed by a CIGAR string ("*" = none); ops other than MIDNSHP=X are an error like htslib's parser
static std::size_t cigar_ref_len(std::string_view cigar, std::size_t line_num) {
    if (cigar == "*") return 0;
    std::size_t total = 0, len = 0;
    bool have = false;
    for (const char ch : cigar) {
        if (ch >= '0' && ch <= '9') {
            len = len * 10 + static_cast<std::size_t>(ch - '0');
            have = true;
            continue;
        }
        if (!have || std::strchr("MIDNSHP=X", ch) == nullptr) throw std::runtime_error("Invalid CIGAR at line " + std::to_string(line_num));
        if (ch == 'M' || ch == 'D' || ch == 'N' || ch == '=' || ch == 'X') total += len;
        len = 0;
        have = false;
    }
    if (have) throw std::runtime_error("Invalid CIGAR at line " + std::to_string(line_num));
    return total;
}

// text SAM: QNAME FLAG RNAME POS MAPQ CIGAR ... (tab separated); '@' header lines skipped
template <typename F>
static void for_each_sam(const std::string& path, F&& cb) {
    std::ifstream in(path);
    if (!in) throw std::runtime_error("Failed to open file: " + path);
    std::string line;
    std::size_t line_num = 0;
    while (std::getline(in, line)) {
        ++line_num;
        if (!line.empty() && line.back() == '\r') line.pop_back();
        if (line.empty() || line[0] == '@') continue;
        std::string_view l(line), col[6];
        std::size_t pos = 0;
        int k = 0;
        for (; k < 6 && pos <= l.size(); ++k) {
            const std::size_t t = l.find('\t', pos);
            col[k] = l.substr(pos, t == std::string_view::npos ? std::string_view::npos : t - pos);
            if (t == std::string_view::npos) { ++k; break; }
            pos = t + 1;
        }
        if (k < 6) throw std::runtime_error("Invalid line format at line " + std::to_string(line_num));
        const std::size_t flag = parse_coord(col[1], line_num);
        if (flag & 4u) continue;  // unmapped: start == end == 0, skipped by the reader's default options
        const std::size_t pos1 = parse_coord(col[3], line_num);
        if (pos1 == 0 || col[2] == "*") continue;  // no position
        bed3 r{std::string(col[2]), pos1 - 1, pos1 - 1 + cigar_ref_len(col[5], line_num), 0, 0};
        if (!(r.start < r.end)) continue;  // consumes_reference() == false: no interval to ask with
        r.lo = r.start;
        r.hi = r.end - 1;
        cb(std::move(r));
    }
}

template <typename F>
static void for_each_record(const std::string& path, F&& cb) {
    if (is_gff(path)) for_each_gff(path, cb);
    else if (is_sam(path)) for_each_sam(path, cb);
    else for_each_bed(path, cb);
}

int main(int argc, char** argv) {
    std::string query, target, out_path = "stdout";
    int order = 3;
    bool sorted = false;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        auto need = [&](const char* what) -> std::string {
            if (i + 1 >= argc) throw std::runtime_error(std::string("Error: missing value for ") + what);
            return argv[++i];
        };
        try {
            if (a == "-q" || a == "--queryfile") query = need("-q");
            else if (a == "-t" || a == "--targetfile") target = need("-t");
            else if (a == "-o" || a == "--outputfile") out_path = need("-o");
            else if (a == "-k" || a == "--order") order = std::stoi(need("-k"));
            else if (a == "-s" || a == "--sorted") sorted = true;
            else if (a == "-h" || a == "--help") {
                std::puts("gg_isec -q query.bed -t target.bed [-o out] [-k order] [-s]");
                return 0;
            } else if (query.empty()) query = a;
            else if (target.empty()) target = a;
        } catch (const std::exception& ex) {
            std::fprintf(stderr, "%s\n", ex.what());
            return 2;
        }
    }
    try {
        if (query.empty()) throw std::runtime_error("Error: queryfile is required");
        if (target.empty()) throw std::runtime_error("Error: a target file (-t) is required");
        if (order < 3) throw std::runtime_error("Error: order must be at least 3");

        gst::grove<gdt::interval, bed3> grove(order);
        for_each_record(target, [&](bed3 r) {
            const gdt::interval iv(r.lo, r.hi);
            std::string chrom = r.chrom;
            if (sorted) grove.insert_data(chrom, iv, std::move(r), gst::sorted);
            else grove.insert_data(chrom, iv, std::move(r));
        });

        // the whole query file -> SoA buffers -> one batch call
        std::vector<std::uint32_t> qc;
        std::vector<std::uint64_t> qs, qe;
        for_each_record(query, [&](bed3 r) {
            qc.push_back(grove.index_id(r.chrom));  // npos_index for a chromosome the target lacks: no hits
            qs.push_back(r.lo);
            qe.push_back(r.hi);
        });
        std::ofstream file;
        if (out_path != "stdout") {
            file.open(out_path);
            if (!file.is_open()) throw std::runtime_error("Error: could not open output file: " + out_path);
        }
        std::ostream& out = out_path == "stdout" ? std::cout : file;
        std::string buf;
        buf.reserve(1 << 20);
        auto print_hits = [&](const auto& res) {
        for (std::size_t q = 0; q < res.size(); ++q) {
            for (std::size_t j = 0; j < res.count(q); ++j) {
                const bed3& t = res.key(q, j)->get_data();
                buf += t.chrom;
                buf += '\t';
                buf += std::to_string(t.start);
                buf += '\t';
                buf += std::to_string(t.end);
                buf += '\n';
                if (buf.size() > (1 << 20) - 256) {
                    out << buf;
                    buf.clear();
                }
            }
        }
        };
        // BED / GFF / SAM coordinates fit 32 bits on every real genome: then the batch goes up as u32 columns and the hits
        // come back in the compact form, read in place (12 + ~15 bytes per query over PCIe instead of 20 + 38 — the call
        // is copy-bound).  Anything wider takes the u64 call; both run the same kernels and print the same lines.
        bool narrow = true;
        for (std::size_t i = 0; i < qe.size() && narrow; ++i) narrow = qe[i] <= 0xFFFFFFFFull;
        gg_index_stats ist{};
        if (narrow && (gg_index_get_stats(grove.device_handle(), &ist) != GG_OK || ist.coord_bytes != 4)) narrow = false;
        if (narrow) {
            std::vector<std::uint32_t> qs32(qs.begin(), qs.end()), qe32(qe.begin(), qe.end());
            print_hits(grove.intersect_batch_compact(qc.data(), qs32.data(), qe32.data(), qc.size()));
        } else {
            print_hits(grove.intersect_batch(qc.data(), qs.data(), qe.data(), qc.size()));
        }
        out << buf;
        out.flush();
    } catch (const std::exception& ex) {
        std::fprintf(stderr, "%s\n", ex.what());
        return 1;
    }
    return 0;
}
