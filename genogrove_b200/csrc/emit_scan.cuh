// emit_scan.cuh — pass 2 of the batch path: offsets AND hits in one kernel.
//
// Replaces the leaf walk of detail::search_overlaps (include/genogrove/structure/grove/query_engine.hpp:83-112)
// and the growth of query_result's vector (data_type/query_result.hpp:108-113) for a whole batch.
// Input: cnt / hi from intersect_count_group_kernel and block_prefix[b] = CSR offset of the first query of block b
// (blocks of blk_q = 32 * QPT consecutive queries; the exclusive scan of that kernel's block totals, chunk base
// included; block_prefix[n_blocks] = base + total).  A warp owns whole blocks: it carries the running offset in a
// register, turns 32 counts into offsets with one warp scan (the per-query scan kernel of the two-phase API and its
// 12 B/query of traffic disappear), stores offsets[q], walks into its shared-memory stage and copies the stage out
// coalesced.  Warps never wait for each other: no barrier, no look-back, no atomics on the output.
// hits[] is chunk-local (index = offset - block_prefix[0]); ranks at or beyond `capacity` are not written and bit 1
// of err_flag reports the overflow (offsets are complete either way).
#pragma once
#include "intersect.cuh"
#include "strand.cuh"

namespace gg {

constexpr uint32_t kEmitScanSlack = 8;  // spare slots in front of each warp's stage (see emit_walk_staged)
// stage entries per warp of intersect_emit_scan_kernel: the dense-batch variant doubles the stage (4 KB per warp, 6 CTAs per
// SM instead of 8) so that tiles of 16-32 hits per query still leave through coalesced stores (GENCODE-shaped cfg3: 88 % of
// the tiles exceed 512 hits, 0.5 % exceed 1024)
__host__ __device__ constexpr uint32_t emit_stage_entries(bool strand, bool dense) { return dense && !strand ? 2u * kEmitStage : kEmitStage; }

// kDense: intersect_emit_dense_kernel follows and writes the tiles with >= kDenseMin hits (skipped here).
// kStrand (genomic_coordinate keys, strand.cuh): cnt_in holds the strand-compatible counts and the walk also tests
// the strand byte of every rank against qstrand[q]; strands = padded level-0 strand bytes, slot_base = tab[0][0].
struct StrandArgs {
    const uint8_t* strands;
    const uint8_t* qstrand;
    uint32_t slot_base;
};

// Warp tiles whose hit lists do not fit the stage (dense indexes, long nested intervals: cfg5 has ~150 hits per query,
// GENCODE-shaped cfg3 ~20).  All 32 lanes keep walking their own queries concurrently; each lane owns a private
// 16-entry segment of the warp's stage, fills it back to front, and when no lane can take another 8-rank group the
// warp flushes all segments: a lane's segment is a contiguous run of its output list, so it goes out as 64-byte
// pieces written by 16 threads each (whole 32-byte sectors instead of one 4-byte store per hit and lane — the direct
// walk costs 8x the L2 write transactions).  Runs in its own kernel (intersect_emit_dense_kernel) so the common
// path keeps its register budget.
constexpr uint32_t kDenseSeg = kEmitStage / 32;  // stage entries per lane (16)
// per-warp stage of intersect_emit_dense_kernel: the interval build also holds the shared candidate list of a sorted tile
// (ranks + ends; emit_tile_delta) and is as large as the stage of the dense-batch intersect_emit_scan_kernel
__host__ __device__ constexpr uint32_t dense_stage_entries(bool strand) { return strand ? kEmitStage : 2u * kEmitStage; }
// Tiles between the stage size and this many hits are walked straight to global memory by intersect_emit_scan_kernel
// (skewed lists: a few long ones among short ones — GENCODE-shaped data — gain nothing from the segment flush);
// from 64 hits per query on average the 8x saving in L2 write transactions wins (cfg5: 201 -> 89 ms).
constexpr uint32_t kDenseMin = 2048;
// Between the stage size and kDenseMin the choice depends on the SHAPE of the tile: 32 lists of similar length (a uniformly
// dense index: 16-64 hits per query) still want the segment flush — the direct walk writes them at 1e10 hits/s, measured on
// a 50 M-interval grove — while one long list among short ones does not.  Both kernels evaluate this same predicate on the
// tile's 32 counts, so exactly one of them writes the tile.
__device__ __forceinline__ bool dense_tile(uint64_t wtotal, uint32_t cmax, uint32_t stage_entries) {
    if (wtotal <= stage_entries || wtotal >= (1ull << 31)) return false;
    return wtotal >= kDenseMin || static_cast<uint64_t>(cmax) * 4u <= wtotal;
}

template <typename C, bool kStrand>
__device__ __forceinline__ void emit_dense_tile(const IndexView<C>& ix, const uint32_t* __restrict__ tabp, uint32_t stride,
                                                int n_levels, uint32_t chrom, C qs, uint32_t hi, uint32_t c, uint64_t o_lo,
                                                uint32_t stage_addr, uint32_t* __restrict__ out, uint64_t room,
                                                const uint8_t* __restrict__ strands, uint32_t q_strand) {
    static_assert(kDenseSeg == 16, "flush assumes two 16-entry segments per warp-wide store");
    const uint32_t lane = lane_id();
    const C* __restrict__ leaf_end = ix.M + (c != 0 ? tabp[chrom] : 0u);
    const uint32_t q4 = q_strand * 0x01010101u;
    const bool wild = !kStrand || q_strand == '*';
    const uint32_t seg_addr = stage_addr + lane * (kDenseSeg * 4u);
    uint32_t left = c, r_hi = hi;  // hits still to write (the next one goes to out[o_lo + left - 1])
    C ea[8];
    while (__any_sync(0xffffffffu, left != 0)) {
        uint32_t n = 0;  // entries in this lane's segment: slots [kDenseSeg - n, kDenseSeg)
        while (left != 0 && n + 8 <= kDenseSeg && r_hi != 0) {
            if ((r_hi & (kFanout - 1)) == 0 && n_levels > 1) {
                const uint32_t span = emit_skip_span<C>(ix, tabp, stride, n_levels, chrom, qs, r_hi);
                if (span) {
                    r_hi -= span;
                    continue;
                }
            }
            const uint32_t g8 = (r_hi - 1) & ~7u, top = r_hi - g8;
            load8(leaf_end + g8, ea);
            uint32_t m_lo = 0xFFFFFFFFu, m_hi = 0xFFFFFFFFu;
            if (kStrand && !wild) {
                const uint2 sw = *reinterpret_cast<const uint2*>(strands + g8);
                m_lo = strand_compat4(sw.x, q4);
                m_hi = strand_compat4(sw.y, q4);
            }
            uint32_t m = group_hit_mask<C>(ea, qs, top, m_lo, m_hi);
            while (m != 0 && left != 0) {  // hits of the group, highest rank first
                const uint32_t k = 31u - __clz(m);
                m ^= 1u << k;
                --left;
                ++n;
                sts32(seg_addr + (kDenseSeg - n) * 4u, g8 + k);
            }
            r_hi = g8;
        }
        if (r_hi == 0) left = 0;  // (cannot happen with counts that match the index; keeps the loop finite)
        __syncwarp();
        // flush: segment s (lane s) holds n_s entries for out[o_lo_s + left_s ...]; 16 threads per segment, two segments
        // per step (positions are tile-local and fit 32 bits: the tile has fewer than 2^31 hits)
        const uint32_t dst0 = static_cast<uint32_t>(o_lo) + left;
        const uint32_t e = lane & 15u, half = lane >> 4;
        const uint32_t room32 = room > 0xFFFFFFFFull ? 0xFFFFFFFFu : static_cast<uint32_t>(room);
        if (__any_sync(0xffffffffu, n != 0)) {
#pragma unroll 4
            for (uint32_t j = 0; j < 16; ++j) {
                const uint32_t sgm = 2 * j + half;
                const uint32_t n_s = __shfl_sync(0xffffffffu, n, sgm);
                const uint32_t pos = __shfl_sync(0xffffffffu, dst0, sgm) + e;
                if (e < n_s && pos < room32) __stcs(out + pos, lds32(stage_addr + (sgm * kDenseSeg + (kDenseSeg - n_s) + e) * 4u));
            }
        }
        __syncwarp();
    }
}

__device__ __forceinline__ uint32_t max_of8(const uint32_t (&e)[8]) {  // 4 three-input max instructions
    return max(__vimax3_u32(__vimax3_u32(e[0], e[1], e[2]), e[3], e[4]), __vimax3_u32(e[5], e[6], e[7]));
}
__device__ __forceinline__ uint64_t max_of8(const uint64_t (&e)[8]) {
    return max(max(max(e[0], e[1]), max(e[2], e[3])), max(max(e[4], e[5]), max(e[6], e[7])));
}

// ---- warp-cooperative walk (skewed hit lists: GENCODE-shaped annotations, one long list among short ones) ------------
// One lane = one query lets the longest walk of a 32-query tile set the pace (cfg3, ncu: 6 of 32 threads active per
// instruction).  Here the whole warp finishes ONE query: a coalesced load of 32 group maxima (M8: max end of every aligned
// group of 8 level-0 slots) covers 256 slots per step, a ballot says which groups can hold a hit, and those are tested four
// at a time, 8 lanes per group; the in-order write position of every hit falls out of the ballot.  At 1024-slot boundaries
// 32 entries of the flat per-1024 maxima are tested at once, so a nested query jumps straight to the ancestor that
// contains it.  Works on absolute level-0 slots (chromosome segments are padded to whole 32-slot nodes, pads hold end = 0
// and sit above the query's hi or below its chromosome: the walk takes the topmost `rem` hits and exactly count(q) slots of
// [cb, hi) pass the test, so it never reaches another chromosome).
template <typename C>
struct CoopArgs {
    const C* leaf;          // ends in chain order by absolute level-0 slot
    const C* m8;            // max end per 8 slots
    const C* lv1;           // max end per 1024 slots (flat hierarchy level 1), valid when n_lv > 1
    const uint32_t* cslot;  // [n_chrom + 1] first slot of each chromosome
    uint32_t n_groups;      // entries of m8
    int n_lv;
    int budget;             // per-lane walk steps before a query is handed to the warp (0: per-lane walks only)
    uint32_t delta_min;     // tiles with at least this many hits may take the sorted-tile path (emit_tile_delta); 0 = never
};

// dst[j] = max(src[8 j .. 8 j + 8)); src is padded to whole 32-slot nodes, so every group is complete
template <typename C>
__global__ void group8_max_kernel(const C* __restrict__ src, C* __restrict__ dst, uint32_t n_dst) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n_dst) return;
    C e[8];
    load8(src + static_cast<size_t>(j) * 8, e);
    dst[j] = max_of8(e);
}

__device__ __forceinline__ uint32_t strip_top_bit(uint32_t f) { return f != 0 ? f & ~(0x80000000u >> __clz(f)) : 0u; }

// All 32 lanes call this with the same arguments.  Hits are slots < abs_hi with end >= qs, highest first; value stored =
// slot - cb.  kShared: the next hit goes to shared address sp - 4, downwards.  Else: to dst[rem - 1] down to dst[0],
// positions at or beyond `room` dropped (hit buffer too small; the caller has flagged it).
__device__ __forceinline__ void sts_coord(uint32_t a, uint32_t v) { sts32(a, v); }
__device__ __forceinline__ void sts_coord(uint32_t a, uint64_t v) {
    sts32(a, static_cast<uint32_t>(v));
    sts32(a + 4u, static_cast<uint32_t>(v >> 32));
}
__device__ __forceinline__ void lds_coord(uint32_t a, uint32_t& v) { v = lds32(a); }
__device__ __forceinline__ void lds_coord(uint32_t a, uint64_t& v) {
    const uint2 w = lds64(a);
    v = (static_cast<uint64_t>(w.y) << 32) | w.x;
}

// ends_base != 0 (kShared only): the end of the hit stored at list index i also goes to shared address ends_base + i * sizeof(C)
template <typename C, bool kShared>
__device__ __forceinline__ void coop_walk(const CoopArgs<C>& ca, C qs, uint32_t abs_hi, uint32_t cb, uint32_t sp, uint32_t rem,
                                          uint32_t* __restrict__ dst, uint64_t room, uint32_t ends_base = 0) {
    const uint32_t lane = lane_id();
    const uint32_t sub = 3u - (lane >> 3);  // lanes 24-31 take the highest flagged group, 16-23 the next, ...
    uint32_t r_hi = abs_hi;
    while (rem != 0 && r_hi != 0) {
        if ((r_hi & 1023u) == 0 && ca.n_lv > 1) {  // 32 x 1024 slots below r_hi at once
            const uint32_t i1 = r_hi >> 10;
            const bool f = lane < i1 && !(ca.lv1[i1 - 1 - lane] < qs);
            const uint32_t m = __ballot_sync(0xffffffffu, f);
            if (m == 0) {
                r_hi = i1 > 32 ? (i1 - 32) << 10 : 0u;
                continue;
            }
            r_hi = (i1 - static_cast<uint32_t>(__ffs(m) - 1)) << 10;
        }
        const uint32_t blk = (r_hi - 1) >> 8;  // the 256-slot block holding slot r_hi - 1
        const uint32_t g = blk * 32 + lane;
        uint32_t flagged = __ballot_sync(0xffffffffu, g < ca.n_groups && g * 8 < r_hi && !(ca.m8[g] < qs));
        while (flagged != 0 && rem != 0) {
            uint32_t f = flagged;
            if (sub >= 1) f = strip_top_bit(f);
            if (sub >= 2) f = strip_top_bit(f);
            if (sub >= 3) f = strip_top_bit(f);
            const uint32_t slot = (blk * 32 + (31u - __clz(f | 1u))) * 8 + (lane & 7u);
            const C ev = (f != 0 && slot < r_hi) ? ca.leaf[slot] : C(0);
            const bool hit = f != 0 && slot < r_hi && !(ev < qs);
            const uint32_t hm = __ballot_sync(0xffffffffu, hit);
            const uint32_t k = __popc((hm >> 1) >> lane);  // hits in higher slots of this step
            if (hit && k < rem) {
                if (kShared) {
                    sts32(sp - 4u * (k + 1), slot - cb);
                    if (ends_base != 0) sts_coord(ends_base + (rem - 1 - k) * static_cast<uint32_t>(sizeof(C)), ev);
                } else if (rem - 1 - k < room) {
                    dst[rem - 1 - k] = slot - cb;
                }
            }
            const uint32_t nh = min(static_cast<uint32_t>(__popc(hm)), rem);
            sp -= 4u * nh;
            rem -= nh;
            flagged = strip_top_bit(strip_top_bit(strip_top_bit(strip_top_bit(flagged))));
        }
        r_hi = blk << 8;
    }
}

// ---- sorted dense tiles: one candidate list per 32-query tile ------------------------------------------------------
// On a dense index (cfg5: 150 hits per 150 bp query) the per-lane forms spend ~4 warp-instructions per hit testing ~650
// ranks per query and flushing per-lane segments.  In a coordinate-sorted batch the 32 queries of a tile ask almost the
// same question: with q.start and hi non-decreasing along the tile (one chromosome), every hit of query j is either a hit
// of query 0 or a rank in [hi_0, hi_j).  So the warp builds ONE candidate list U = hits(query 0) ++ [hi_0, hi_31) in shared
// memory, ranks ascending with their ends beside them — one cooperative walk for the whole tile — and each query's list is
// the prefix U[0, c_0 + hi_j - hi_0) filtered by end >= q_j.start: 32 candidates per step, one compare, a ballot for the
// write positions, one coalesced store.  ~0.5 warp-instructions per hit; no per-rank test per query.
// Both pass-2 kernels of the dense variant evaluate tile_delta_ok on the same values, so exactly one of them writes a tile.
template <typename C>
__host__ __device__ constexpr uint32_t delta_cap(uint32_t stage_entries) {  // candidates that fit the warp's stage (even)
    return (stage_entries * 4u / (4u + static_cast<uint32_t>(sizeof(C)))) & ~1u;
}
template <typename C>
__device__ __forceinline__ bool tile_delta_ok(uint32_t c, uint32_t hi, C qs, uint32_t chrom, uint64_t wtotal, uint32_t cap,
                                              uint32_t delta_min) {
    if (delta_min == 0 || wtotal < delta_min || wtotal >= (1ull << 31)) return false;
    const uint32_t hi_prev = __shfl_up_sync(0xffffffffu, hi, 1), ch0 = __shfl_sync(0xffffffffu, chrom, 0);
    const C qs_prev = __shfl_up_sync(0xffffffffu, qs, 1);
    const bool ok = c != 0 && chrom == ch0 && (lane_id() == 0 || (hi >= hi_prev && !(qs < qs_prev)));
    if (!__all_sync(0xffffffffu, ok)) return false;
    const uint32_t c0 = __shfl_sync(0xffffffffu, c, 0), hi0 = __shfl_sync(0xffffffffu, hi, 0), hi31 = __shfl_sync(0xffffffffu, hi, 31);
    return static_cast<uint64_t>(c0) + (hi31 - hi0) <= cap;
}

template <typename C>
__device__ __forceinline__ C warp_min(C v) {
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
        const C o = __shfl_xor_sync(0xffffffffu, v, d);
        v = o < v ? o : v;
    }
    return v;
}

// The tile's list is kept INCREMENTALLY (second version; the first filtered the whole candidate list for every query, 23
// instructions per 32 candidates): going from query j-1 to j the list gains the ranks [hi_{j-1}, hi_j) at its end and loses
// the members whose end fell below q_j.start — known from the smallest end among the members, so most queries lose nothing
// and their list goes out as a plain shared -> global copy.
// The list outlives the tile: a warp of the dense kernel takes up to 32 consecutive tiles, and when the next sorted tile
// continues where this one stopped (same chromosome, q.start and hi not below the last query's) its first query is just
// one more incremental step — the cooperative walk that builds a list from nothing (~650 ranks on cfg5) then runs once per
// 1024 queries instead of once per 32.
template <typename C>
struct DeltaCarry {
    bool valid = false;
    uint32_t chrom = 0, len = 0, prev_hi = 0;
    C last_qs = 0, min_end = 0;
};

template <typename C>
__device__ __forceinline__ void emit_tile_delta(const CoopArgs<C>& ca, uint32_t c, uint32_t hi, C qs, uint32_t chrom, uint64_t excl,
                                                uint64_t wbase, uint32_t stage_addr, uint32_t cap, uint32_t* __restrict__ hits,
                                                uint64_t capacity, DeltaCarry<C>& carry) {
    constexpr uint32_t kSz = static_cast<uint32_t>(sizeof(C));
    const uint32_t lane = lane_id(), lt = (1u << lane) - 1u;
    const uint32_t rb = stage_addr, eb = stage_addr + cap * 4u;  // ranks, then ends (8-byte aligned: cap is even)
    const uint32_t c0 = __shfl_sync(0xffffffffu, c, 0), hi0 = __shfl_sync(0xffffffffu, hi, 0), hi31 = __shfl_sync(0xffffffffu, hi, 31);
    const C qs0 = __shfl_sync(0xffffffffu, qs, 0);
    const uint32_t ch0 = __shfl_sync(0xffffffffu, chrom, 0);
    const uint32_t cb = ca.cslot[ch0];
    uint32_t len, prev_hi;
    C min_end;
    if (carry.valid && carry.chrom == ch0 && hi0 >= carry.prev_hi && !(qs0 < carry.last_qs) &&
        static_cast<uint64_t>(carry.len) + (hi31 - carry.prev_hi) <= cap) {
        len = carry.len, prev_hi = carry.prev_hi, min_end = carry.min_end;  // the stage still holds the last query's list
    } else {
        // the list of query 0, by the whole warp: ranks ascending at rb, their ends at eb
        coop_walk<C, true>(ca, qs0, cb + hi0, cb, rb + 4u * c0, c0, nullptr, 0, eb);
        __syncwarp();
        len = c0, prev_hi = hi0;
        C my_min = coord_max<C>::value;
        for (uint32_t i = lane; i < len; i += 32) {
            C e;
            lds_coord(eb + i * kSz, e);
            my_min = e < my_min ? e : my_min;
        }
        min_end = warp_min<C>(my_min);
    }
    // the tile's lists are consecutive in the CSR and each has exactly as many entries as the list kept here: the output
    // position runs along with it (no per-query offset shuffle); the capacity is tested once per tile when everything fits
    uint64_t oj = wbase + __shfl_sync(0xffffffffu, excl, 0);
    const uint64_t tile_end = wbase + __shfl_sync(0xffffffffu, excl, 31) + __shfl_sync(0xffffffffu, c, 31);
    const bool fits = tile_end <= capacity;
    for (int j = 0; j < 32; ++j, oj += len) {
        const uint32_t hj = __shfl_sync(0xffffffffu, hi, j);
        const C qj = __shfl_sync(0xffffffffu, qs, j);
        if (hj != prev_hi) {  // new members: the ranks whose start lies in (q_{j-1}.end, q_j.end]
            const uint32_t add = hj - prev_hi;
            C m = coord_max<C>::value;
            for (uint32_t i = lane; i < add; i += 32) {
                const C e = ca.leaf[cb + prev_hi + i];
                sts32(rb + 4u * (len + i), prev_hi + i);
                sts_coord(eb + (len + i) * kSz, e);
                m = e < m ? e : m;
            }
            m = warp_min<C>(m);
            min_end = m < min_end ? m : min_end;
            len += add;
            prev_hi = hj;
            __syncwarp();
        }
        const uint64_t room64 = fits ? 0xFFFFFFFFull : (capacity > oj ? capacity - oj : 0ull);
        if (min_end < qj) {
            // at least one member ended before q_j.start: ONE pass filters the list, writes the survivors out and compacts
            // the list in place (queries sparser than the index lose a member at almost every step; a separate compaction
            // pass in front of the copy made those batches slower than filtering the whole candidate list every time)
            const uint32_t room = room64 > 0xFFFFFFFFull ? 0xFFFFFFFFu : static_cast<uint32_t>(room64);
            uint32_t* __restrict__ out = hits + oj;
            uint32_t kept = 0;
            C m = coord_max<C>::value;
            for (uint32_t k = 0; k < len; k += 32) {
                const uint32_t i = k + lane;
                C e = 0;
                uint32_t r = 0;
                if (i < len) {
                    lds_coord(eb + i * kSz, e);
                    r = lds32(rb + 4u * i);
                }
                const bool alive = i < len && !(e < qj);
                const uint32_t mask = __ballot_sync(0xffffffffu, alive);
                const uint32_t pos = kept + __popc(mask & lt);
                __syncwarp();  // every lane has read its slot before slots at or below it are rewritten
                if (alive) {
                    sts32(rb + 4u * pos, r);
                    sts_coord(eb + pos * kSz, e);
                    m = e < m ? e : m;
                    if (pos < room) __stcs(out + pos, r);
                }
                kept += __popc(mask);
            }
            len = kept;
            min_end = warp_min<C>(m);
            __syncwarp();
        } else {
            // nobody left: the list of query j is a plain copy
            const uint32_t n_out = room64 < len ? static_cast<uint32_t>(room64) : len;
            uint32_t* __restrict__ out = hits + oj + lane;
            uint32_t sa = rb + 4u * lane;
            uint32_t i = lane;
            for (; i + 96 < n_out; i += 128, out += 128, sa += 512u) {
                const uint32_t v0 = lds32(sa), v1 = lds32(sa + 128u), v2 = lds32(sa + 256u), v3 = lds32(sa + 384u);
                __stcs(out, v0);
                __stcs(out + 32, v1);
                __stcs(out + 64, v2);
                __stcs(out + 96, v3);
            }
            for (; i < n_out; i += 32, out += 32, sa += 128u) __stcs(out, lds32(sa));
        }
    }
    __syncwarp();
    carry.valid = true;
    carry.chrom = ch0;
    carry.len = len;
    carry.prev_hi = prev_hi;
    carry.min_end = min_end;
    carry.last_qs = __shfl_sync(0xffffffffu, qs, 31);
}

// emit_walk_staged8 (intersect.cuh) for the dense-batch variant: a group of 8 ranks is first tested as a whole (max end
// against q.start).  On nested annotations most groups of a visited node hold no hit (cfg3: 4.5 % of the walked ranks are
// hits, 7 in 10 groups have none) and the 8 compare + predicated store triples are skipped.
// budget > 0: at most that many loop steps (a skip or a group test each); the walk's state comes back through r_out / sp_out
// (sp_out > s_begin: unfinished, the warp takes over from rank r_out).
template <typename C>
__device__ __forceinline__ void emit_walk_staged8m(const IndexView<C>& ix, const uint32_t* __restrict__ tabp, uint32_t stride,
                                                   int n_levels, uint32_t chrom, C qs, uint32_t hi, uint32_t s_begin,
                                                   uint32_t s_end, int budget, uint32_t& r_out, uint32_t& sp_out) {
    const C* __restrict__ leaf_end = ix.M + tabp[chrom];
    const bool has_levels = n_levels > 1;
    const C* __restrict__ node_max = ix.M + tabp[(has_levels ? stride : 0u) + chrom];  // level 1: max end per leaf node
    uint32_t r_hi = hi, sp = s_end;
    C ea[8];
    if ((r_hi & 7u) != 0) {  // partial top group: ranks g8 .. hi-1
        const uint32_t g8 = r_hi & ~7u, top = r_hi & 7u;
        load8(leaf_end + g8, ea);
#pragma unroll
        for (int k = 6; k >= 1; --k) {
            if (top > static_cast<uint32_t>(k) && ea[k] >= qs) { sp -= 4; sts32(sp, g8 + k); }
        }
        if (ea[0] >= qs) { sp -= 4; sts32(sp, g8); }
        r_hi = g8;
    }
    int left = budget > 0 ? budget : 0x7fffffff;
    while (sp > s_begin && r_hi != 0 && left-- > 0) {
        if ((r_hi & (kFanout - 1)) == 0 && has_levels) {
            const uint32_t b = (r_hi >> kFanoutLog2) - 1;  // leaf node about to be entered from its last key
            if (node_max[b] < qs) {  // nothing in it reaches q.start; at a parent boundary try to skip whole subtrees
                r_hi -= (b & (kFanout - 1)) == (kFanout - 1) ? emit_skip_span<C>(ix, tabp, stride, n_levels, chrom, qs, r_hi) : kFanout;
                continue;
            }
        }
        const uint32_t g8 = r_hi - 8;
        load8(leaf_end + g8, ea);
        if (max_of8(ea) >= qs) {
#pragma unroll
            for (int k = 7; k >= 0; --k) {
                if (ea[k] >= qs) { sp -= 4; sts32(sp, g8 + k); }
            }
        }
        r_hi = g8;
    }
    r_out = r_hi;
    sp_out = sp;
}

// kVecOut: the stage is laid out with the phase (address mod 16) of its global destination, so the copy-out moves 16 bytes
// per lane and instruction (one LDS.128 + one STG.128 per 4 hits) instead of 4; the <= 3 hits at either end go out as
// scalar stores (the per-hit LDS + STG.32 loop was 10-13 % of this kernel's instructions).
constexpr uint32_t kVecOutTailWords = 4;  // room behind the stage for the phase shift
// kTileList (coordinate-sorted batches on ordinary indexes, cfg4): when a tile's 32 queries are on one chromosome with
// q.start and hi non-decreasing, every hit of query j is a hit of query 0 or a rank in [hi_0, hi_j) — the observation behind
// emit_tile_delta, used here with the lanes still working in parallel: the warp builds the short shared candidate list once
// (one cooperative walk for query 0 instead of 32 walks over the same ~30 ranks), then every lane filters that list for its
// own query: ~11 candidates tested per lane where the walk tested ~30 ranks.  Other tiles take the walk as before.
constexpr uint32_t kTileListCap = 64;  // candidates (ranks + ends) per warp
template <typename C>
__host__ __device__ constexpr uint32_t tile_list_words() { return kTileListCap + kTileListCap * static_cast<uint32_t>(sizeof(C)) / 4u; }
template <typename C, typename Q, int kWalk, bool kStrand = false, bool kDense = false, bool kVecOut = false, bool kTileList = false>
__global__ void __launch_bounds__(kQueryThreads)
intersect_emit_scan_kernel(IndexView<C> ix, const uint32_t* __restrict__ qchrom, const Q* __restrict__ qstart, size_t nq,
                           const uint32_t* __restrict__ cnt_in, const uint32_t* __restrict__ hi_in,
                           const uint64_t* __restrict__ block_prefix, uint32_t blk_q, uint64_t* __restrict__ offsets,
                           uint32_t* __restrict__ hits, uint64_t capacity, uint64_t* __restrict__ next_base_out,
                           uint64_t* __restrict__ total_out, uint32_t* __restrict__ err_flag, int tab_in_smem,
                           StrandArgs sa = StrandArgs{nullptr, nullptr, 0}, CoopArgs<C> ca = CoopArgs<C>{}) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr uint32_t kStage = emit_stage_entries(kStrand, kDense);
    constexpr uint32_t kStageWords = kStage + kEmitScanSlack + (kVecOut ? kVecOutTailWords : 0u) + (kTileList ? tile_list_words<C>() : 0u);
    constexpr uint32_t kWarps = kQueryThreads / 32;
    uint32_t* stab = reinterpret_cast<uint32_t*>(smem_raw) + kWarps * kStageWords;
    const uint32_t lane = lane_id(), warp = threadIdx.x >> 5;
    const uint32_t stage_addr = smem_u32(smem_raw) + (warp * kStageWords + kEmitScanSlack) * 4u;
    const uint32_t list_addr = stage_addr + (kStage + (kVecOut ? kVecOutTailWords : 0u)) * 4u;  // (kTileList) ranks, then ends
    const uint32_t stride = ix.n_chrom + 1;
    const int n_levels = ix.n_levels;
    const uint32_t* __restrict__ tabp = ix.tab;
    if (tab_in_smem) {
        const uint32_t tab_elems = static_cast<uint32_t>(n_levels) * stride;
        for (uint32_t i = threadIdx.x; i < tab_elems; i += blockDim.x) stab[i] = ix.tab[i];
        tabp = stab;
        __syncthreads();
    }
    const uint64_t base = block_prefix[0];
    if (nq == 0) {
        if (blockIdx.x == 0 && threadIdx.x == 0) {
            offsets[0] = base;
            if (next_base_out) *next_base_out = base;
            if (total_out) *total_out = 0;
        }
        return;
    }
    const size_t n_blocks = (nq + blk_q - 1) / blk_q;
    const size_t warps_total = static_cast<size_t>(gridDim.x) * kWarps;
    for (size_t blk = static_cast<size_t>(blockIdx.x) * kWarps + warp; blk < n_blocks; blk += warps_total) {
        uint64_t run = block_prefix[blk];
        const size_t qb = blk * blk_q;
        for (uint32_t it = 0; it < blk_q && qb + it < nq; it += 32) {
            const size_t q = qb + it + lane;
            const bool in = q < nq;
            const uint32_t c = in ? cnt_in[q] : 0u;
            // exclusive warp scan of the 32 counts: 32-bit when the sum cannot overflow, else 64-bit
            uint64_t excl, wtotal;
            const uint32_t cmax = __reduce_max_sync(0xffffffffu, c);
            if (cmax < (1u << 27)) {
                uint32_t inc = c;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
                    if (lane >= static_cast<uint32_t>(d)) inc += o;
                }
                excl = inc - c;
                wtotal = __shfl_sync(0xffffffffu, inc, 31);
            } else {
                uint64_t inc = c;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const uint64_t o = __shfl_up_sync(0xffffffffu, inc, d);
                    if (lane >= static_cast<uint32_t>(d)) inc += o;
                }
                excl = inc - c;
                wtotal = __shfl_sync(0xffffffffu, inc, 31);
            }
            const uint64_t o0 = run + excl;
            if (in) __stcs(offsets + q, o0);
            if (q + 1 == nq) {
                const uint64_t end_off = o0 + c;  // = base + total
                offsets[nq] = end_off;
                if (next_base_out) *next_base_out = end_off;
                if (total_out) *total_out = end_off - base;
                if (end_off - base > capacity) atomicOr(err_flag, 2u);
            }
            const uint64_t wbase = run - base;  // chunk-local index of the warp's first hit
            run += wtotal;
            if (wtotal == 0) continue;
            uint32_t chrom = 0, hi = 0, qst = 0;
            C qs = 0;
            if (c != 0) {
                chrom = qchrom[q];
                qs = static_cast<C>(qstart[q]);  // c > 0 implies q.start fits the index domain
                hi = hi_in[q];
                if (kStrand) qst = sa.qstrand[q];
            }
            if (wtotal <= kStage) {
                // (kVecOut) bytes by which the tile's stage image is shifted: shared and global addresses agree modulo 16
                const uint32_t phase = kVecOut ? static_cast<uint32_t>(reinterpret_cast<uintptr_t>(hits + wbase)) & 15u : 0u;
                if (kDense && !kStrand) {
                    // per-lane walks for a bounded number of steps, then the warp finishes the unfinished lists one by one
                    const uint32_t s_begin = stage_addr + phase + static_cast<uint32_t>(excl) * 4u;
                    uint32_t r_stop = 0, sp_stop = s_begin;
                    if (c != 0) emit_walk_staged8m<C>(ix, tabp, stride, n_levels, chrom, qs, hi, s_begin, s_begin + c * 4u, ca.budget, r_stop, sp_stop);
                    uint32_t pend = ca.budget > 0 ? __ballot_sync(0xffffffffu, sp_stop > s_begin && r_stop != 0) : 0u;
                    while (pend != 0) {
                        const int j = __ffs(pend) - 1;
                        pend &= pend - 1;
                        const uint32_t cb = ca.cslot[__shfl_sync(0xffffffffu, chrom, j)];
                        const uint32_t sp_j = __shfl_sync(0xffffffffu, sp_stop, j), sb_j = __shfl_sync(0xffffffffu, s_begin, j);
                        coop_walk<C, true>(ca, static_cast<C>(__shfl_sync(0xffffffffu, qs, j)), cb + __shfl_sync(0xffffffffu, r_stop, j),
                                           cb, sp_j, (sp_j - sb_j) >> 2, nullptr, 0);
                    }
                } else if (kTileList && tile_delta_ok<C>(c, hi, qs, chrom, wtotal, kTileListCap, 1u)) {
                    constexpr uint32_t kSz = static_cast<uint32_t>(sizeof(C));
                    const uint32_t ur = list_addr, ue = list_addr + kTileListCap * 4u;
                    const uint32_t c0 = __shfl_sync(0xffffffffu, c, 0), hi0 = __shfl_sync(0xffffffffu, hi, 0), hi31 = __shfl_sync(0xffffffffu, hi, 31);
                    const uint32_t cb = ca.cslot[__shfl_sync(0xffffffffu, chrom, 0)];
                    coop_walk<C, true>(ca, static_cast<C>(__shfl_sync(0xffffffffu, qs, 0)), cb + hi0, cb, ur + 4u * c0, c0, nullptr, 0, ue);
                    if (lane < hi31 - hi0) {  // (at most kTileListCap - c0 <= 64 ranks: two rounds)
                        for (uint32_t i = lane; i < hi31 - hi0; i += 32) {
                            sts32(ur + 4u * (c0 + i), hi0 + i);
                            sts_coord(ue + (c0 + i) * kSz, ca.leaf[cb + hi0 + i]);
                        }
                    }
                    __syncwarp();
                    const uint32_t p_max = c0 + (hi31 - hi0), p = c0 + (hi - hi0);
                    uint32_t sp = stage_addr + phase + static_cast<uint32_t>(excl) * 4u;
                    for (uint32_t i = 0; i < p_max; ++i) {  // every lane reads the same candidate (a broadcast), keeps its own hits
                        C e;
                        lds_coord(ue + i * kSz, e);
                        const uint32_t r = lds32(ur + 4u * i);
                        if (i < p && !(e < qs)) {
                            sts32(sp, r);
                            sp += 4u;
                        }
                    }
                } else if (c != 0) {
                    const uint32_t s_begin = stage_addr + phase + static_cast<uint32_t>(excl) * 4u;
                    if (kStrand)
                        emit_walk_staged8_strand<C>(ix, tabp, stride, n_levels, chrom, qs, hi,
                                                    sa.strands + (tabp[chrom] - sa.slot_base), qst, s_begin, s_begin + c * 4u);
                    else if (kWalk == 16)
                        emit_walk_staged16<C>(ix, tabp, stride, n_levels, chrom, qs, hi, s_begin, s_begin + c * 4u);
                    else
                        emit_walk_staged8<C>(ix, tabp, stride, n_levels, chrom, qs, hi, s_begin, s_begin + c * 4u);
                }
                __syncwarp();
                const uint32_t n = static_cast<uint32_t>(wtotal);
                uint32_t sa = stage_addr + phase + lane * 4u;
                if (wbase + wtotal <= capacity) {
                    if (kVecOut) {
                        const uint32_t n_bytes = n * 4u;
                        const uint32_t head = min(n_bytes, (16u - phase) & 15u);  // bytes up to the first 16-byte boundary
                        const uint32_t mid = (n_bytes - head) & ~15u;             // whole 16-byte vectors
                        const uint32_t tail = n_bytes - head - mid;
                        unsigned char* g = reinterpret_cast<unsigned char*>(hits + wbase);
                        const uint32_t s0 = stage_addr + phase;
                        for (uint32_t v = lane * 16u; v < mid; v += 512u)
                            __stcs(reinterpret_cast<uint4*>(g + head + v), lds128(s0 + head + v));
                        const uint32_t e = (lane & 3u) * 4u;  // lanes 0-3: the head words, lanes 4-7: the tail words
                        if (lane < 4) {
                            if (e < head) __stcs(reinterpret_cast<uint32_t*>(g + e), lds32(s0 + e));
                        } else if (lane < 8) {
                            if (e < tail) __stcs(reinterpret_cast<uint32_t*>(g + head + mid + e), lds32(s0 + head + mid + e));
                        }
                    } else {
                        uint32_t* out = hits + wbase + lane;
                        for (uint32_t i = lane; i < n; i += 32, out += 32, sa += 128) __stcs(out, lds32(sa));
                    }
                } else {
                    for (uint32_t i = lane; i < n; i += 32, sa += 128)
                        if (wbase + i < capacity) hits[wbase + i] = lds32(sa);
                }
                __syncwarp();
            } else if (kDense && (dense_tile(wtotal, cmax, kStage) ||
                                  (!kStrand && tile_delta_ok<C>(c, hi, qs, chrom, wtotal, delta_cap<C>(kStage), ca.delta_min)))) {
                // a dense tile, or a sorted tile that shares one candidate list: intersect_emit_dense_kernel finds it from the
                // per-tile totals of the count pass and writes it
            } else if (kDense && !kStrand && ca.budget > 0) {
                // too large for the stage, one or a few long lists among short ones: every list by the whole warp, straight to
                // global memory (a single lane walking a 100 k-hit list alone was the worst case of the per-lane form)
                uint32_t todo = __ballot_sync(0xffffffffu, c != 0);
                while (todo != 0) {
                    const int j = __ffs(todo) - 1;
                    todo &= todo - 1;
                    const uint32_t cb = ca.cslot[__shfl_sync(0xffffffffu, chrom, j)];
                    const uint64_t o_j = wbase + __shfl_sync(0xffffffffu, excl, j);
                    const uint64_t room = capacity > o_j ? capacity - o_j : 0ull;
                    coop_walk<C, false>(ca, static_cast<C>(__shfl_sync(0xffffffffu, qs, j)), cb + __shfl_sync(0xffffffffu, hi, j), cb, 0u,
                                        __shfl_sync(0xffffffffu, c, j), hits + o_j, room);
                }
            } else if (c != 0 && wbase + excl + c <= capacity) {  // too large for the stage but not dense: straight to global memory
                if (kStrand)
                    strand_walk<C, true>(ix, stride, n_levels, chrom, qs, hi, c, sa.strands + (tabp[chrom] - sa.slot_base), qst,
                                         hits + wbase + excl);
                else
                    emit_walk_global<C>(ix, tabp, stride, n_levels, chrom, qs, hi, hits + wbase + excl, c);
            }
        }
    }
}

// Second half of pass 2, enqueued right after intersect_emit_scan_kernel: writes the hit lists of the warp tiles that
// kernel skipped (at least kDenseMin hits in 32 consecutive queries).  tile_total[t] (u32, saturating; written by the
// count pass) says which tiles those are, so a batch without any costs one coalesced read of 4 bytes per 32 queries.
// A warp takes a dense tile, rebuilds its CSR base from block_prefix and the tile totals before it in the block,
// scans the 32 counts and writes the lists through emit_dense_tile.  Offsets were stored by the first kernel.
template <typename C, typename Q, bool kStrand>
__global__ void __launch_bounds__(kQueryThreads, 5)
intersect_emit_dense_kernel(IndexView<C> ix, const uint32_t* __restrict__ qchrom, const Q* __restrict__ qstart, size_t nq,
                            const uint32_t* __restrict__ cnt_in, const uint32_t* __restrict__ hi_in,
                            const uint64_t* __restrict__ block_prefix, const uint32_t* __restrict__ tile_total,
                            uint32_t blk_q, uint32_t* __restrict__ hits, uint64_t capacity, StrandArgs sa,
                            uint32_t stage_entries /* of the intersect_emit_scan_kernel build that ran before */,
                            CoopArgs<C> ca /* delta_min: sorted tiles from this many hits share one candidate list (0 = off) */) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr uint32_t kStageWords = dense_stage_entries(kStrand) + kEmitScanSlack;
    constexpr uint32_t kWarps = kQueryThreads / 32;
    const uint32_t lane = lane_id(), warp = threadIdx.x >> 5;
    const uint32_t stage_addr = smem_u32(smem_raw) + (warp * kStageWords + kEmitScanSlack) * 4u;
    const uint32_t stride = ix.n_chrom + 1;
    const int n_levels = ix.n_levels;
    const uint32_t* __restrict__ tabp = ix.tab;
    const size_t n_tiles = (nq + 31) / 32;
    const uint32_t tiles_per_blk = blk_q / 32;
    const size_t warps_total = static_cast<size_t>(gridDim.x) * kWarps;
    // 32 tiles per warp step: one coalesced read of their totals, then the dense ones (rare) one after the other
    for (size_t t0 = (static_cast<size_t>(blockIdx.x) * kWarps + warp) * 32; t0 < n_tiles; t0 += warps_total * 32) {
        const uint32_t my_total = t0 + lane < n_tiles ? tile_total[t0 + lane] : 0u;
        uint32_t dense = __ballot_sync(0xffffffffu, my_total > stage_entries);  // candidates; dense_tile() below decides
        DeltaCarry<C> carry;  // (sorted tiles) the candidate list left in the stage by the previous tile of this step
        while (dense != 0) {
            const uint32_t j = __ffs(dense) - 1;
            dense &= dense - 1;
            const size_t t = t0 + j;
            const size_t blk = t / tiles_per_blk;
            uint64_t tile_base = block_prefix[blk];
            bool exact = true;  // a saturated total in front of this tile: fall back to summing the counts
            for (size_t u = blk * tiles_per_blk; u < t; ++u) {
                const uint32_t v = tile_total[u];
                exact &= v != 0xFFFFFFFFu;
                tile_base += v;
            }
            if (!exact) {
                tile_base = block_prefix[blk];
                for (size_t qq = blk * blk_q + lane; qq < t * 32; qq += 32) tile_base += cnt_in[qq];  // per-lane partial sums
                uint64_t part = tile_base - block_prefix[blk];
#pragma unroll
                for (int d = 16; d >= 1; d >>= 1) part += __shfl_xor_sync(0xffffffffu, part, d);
                tile_base = block_prefix[blk] + part;
            }
            const size_t q = t * 32 + lane;
            const uint32_t c = q < nq ? cnt_in[q] : 0u;
            uint64_t inc = c;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint64_t o = __shfl_up_sync(0xffffffffu, inc, d);
                if (lane >= static_cast<uint32_t>(d)) inc += o;
            }
            const uint64_t excl = inc - c, wtotal = __shfl_sync(0xffffffffu, inc, 31);
            const bool is_dense = dense_tile(wtotal, __reduce_max_sync(0xffffffffu, c), stage_entries);
            if (!is_dense && (kStrand || ca.delta_min == 0 || wtotal < ca.delta_min)) continue;  // written by intersect_emit_scan_kernel itself
            const uint64_t wbase = tile_base - block_prefix[0];
            uint32_t chrom = 0, hi = 0, qst = 0;
            C qs = 0;
            if (c != 0) {
                chrom = qchrom[q];
                qs = static_cast<C>(qstart[q]);
                hi = hi_in[q];
                if (kStrand) qst = sa.qstrand[q];
            }
            // a coordinate-sorted tile: one shared candidate list instead of 32 walks (same predicate and values as in
            // intersect_emit_scan_kernel, which left the tile alone)
            if (!kStrand && tile_delta_ok<C>(c, hi, qs, chrom, wtotal, delta_cap<C>(stage_entries), ca.delta_min)) {
                emit_tile_delta<C>(ca, c, hi, qs, chrom, excl, wbase, stage_addr, delta_cap<C>(stage_entries), hits, capacity, carry);
                continue;
            }
            if (!is_dense) continue;
            carry.valid = false;  // the segment flush below reuses the stage
            const uint8_t* strands = kStrand && c != 0 ? sa.strands + (tabp[chrom] - sa.slot_base) : nullptr;
            const uint64_t room = capacity > wbase ? capacity - wbase : 0ull;  // positions of this tile that fit the buffer
            emit_dense_tile<C, kStrand>(ix, tabp, stride, n_levels, chrom, qs, hi, c, excl, stage_addr, hits + wbase, room, strands, qst);
        }
    }
}

}  // namespace gg
