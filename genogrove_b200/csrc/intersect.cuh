// intersect.cuh — batched overlap query kernels: pass 1 (count) and pass 2 (emit).
//
// Replaces detail::search_overlaps (include/genogrove/structure/grove/query_engine.hpp:76-134) called
// once per query by grove::intersect (grove_query.ipp:33-43).  The reference's descent + leaf walk emits
// exactly { r : start[r] <= q.end && end[r] >= q.start } in ascending leaf-chain rank; on the device:
//
//   count(q) = #{start <= q.end} - #{end < q.start}          (valid because start <= end <= ... )
//            = ub(S, q.end) - ub(E, q.start - 1)             two B+ tree descents, no leaf walk
//   emit(q)  : walk ranks hi-1, hi-2, ... (hi = ub(S, q.end), saved by pass 1) testing end[r] >= q.start,
//              writing hits back to front at offsets[q+1]-1 ... offsets[q]; at a node boundary the
//              max-end augmentation M[l][node] < q.start skips the whole node (and, climbing while
//              aligned, whole subtrees), so long nested intervals cost O(32 * levels) instead of a linear
//              leaf walk.  The walk stops as soon as count(q) hits are written: no atomics, output in
//              the reference's order.
//
// One query per thread, CTAs persistent over query tiles (grid = SMs x resident CTAs).  Coordinate-
// sorted batches make the lanes of a warp walk the same nodes, so a node is one L1/L2 transaction
// for the warp; the upper levels of S and E are copied into shared memory once per CTA with a TMA
// bulk copy (cp.async.bulk + mbarrier), the rest is read through L1/L2.
#pragma once
#include "common.cuh"
#include "scan.cuh"

namespace gg {

template <typename C>
struct IndexView {
    const C* S;            // start pivots / leaf starts (chain order)
    const C* E;            // sorted-end pivots / sorted ends
    const C* M;            // max-end levels / leaf ends (chain order)
    const uint32_t* tab;   // [n_levels][n_chrom + 1] absolute element offsets into S/E/M
    const uint32_t* cnt;   // [n_chrom] real interval count per chromosome
    uint32_t n_chrom;
    int n_levels;
    int stage_level;       // levels >= stage_level are staged in shared memory (== n_levels: none)
    uint32_t staged_elems; // number of leading elements of S and E staged (multiple of 32)
};

constexpr int kQueryThreads = 256;
constexpr uint32_t GG_NO_CHROM_DEV = 0xFFFFFFFFu;

// number of keys <= x in one sorted 32-key node (0..32): 6 dependent loads, branch-free
template <typename C>
__device__ __forceinline__ uint32_t node_count_le(const C* __restrict__ a, C x) {
    uint32_t c = 0;
#pragma unroll
    for (int step = 16; step >= 1; step >>= 1) {
        if (a[c + step - 1] <= x) c += step;
    }
    if (a[c] <= x) c += 1;
    return c;
}

// #keys <= x among the n_c keys of chromosome `chrom` in the family whose global buffer is g and whose
// staged prefix is s.  Descends the implicit tree from the top level; see build.cuh for the layout.
template <typename C>
__device__ __forceinline__ uint32_t tree_upper_bound(const C* __restrict__ g, const C* __restrict__ s,
                                                     const uint32_t* __restrict__ tab, uint32_t stride,
                                                     int n_levels, int stage_level, uint32_t chrom, C x,
                                                     uint32_t n_c) {
    uint32_t node = 0;
    int l = n_levels - 1;
    // staged upper levels: shared memory
    for (; l >= stage_level && l > 0; --l) {
        uint32_t off = tab[l * stride + chrom];
        uint32_t c = node_count_le(s + off + (node << kFanoutLog2), x);
        uint32_t below = (tab[(l - 1) * stride + chrom + 1] - tab[(l - 1) * stride + chrom]) >> kFanoutLog2;
        node = min((node << kFanoutLog2) + c, below - 1);
    }
    for (; l > 0; --l) {
        uint32_t off = tab[l * stride + chrom];
        uint32_t c = node_count_le(g + off + (node << kFanoutLog2), x);
        uint32_t below = (tab[(l - 1) * stride + chrom + 1] - tab[(l - 1) * stride + chrom]) >> kFanoutLog2;
        node = min((node << kFanoutLog2) + c, below - 1);
    }
    const C* leaf = ((stage_level == 0) ? s : g) + tab[chrom] + (node << kFanoutLog2);
    uint32_t c = node_count_le(leaf, x);
    return min((node << kFanoutLog2) + c, n_c);
}

// query coordinate (type Q) -> index coordinate domain (type C), saturating
template <typename C, typename Q>
__device__ __forceinline__ C clamp_coord(Q v) {
    if (sizeof(Q) > sizeof(C)) {
        return v > static_cast<Q>(coord_max<C>::value) ? coord_max<C>::value : static_cast<C>(v);
    }
    return static_cast<C>(v);
}

// ---- warp-cooperative search (coordinate-sorted batches) ----------------------------------------------
__device__ __forceinline__ uint32_t warp_min(uint32_t v) { return __reduce_min_sync(0xffffffffu, v); }
__device__ __forceinline__ uint32_t warp_max(uint32_t v) { return __reduce_max_sync(0xffffffffu, v); }
__device__ __forceinline__ uint64_t warp_min(uint64_t v) {
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
        uint64_t o = __shfl_xor_sync(0xffffffffu, v, d);
        v = o < v ? o : v;
    }
    return v;
}
__device__ __forceinline__ uint64_t warp_max(uint64_t v) {
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
        uint64_t o = __shfl_xor_sync(0xffffffffu, v, d);
        v = o > v ? o : v;
    }
    return v;
}

// v holds one sorted 32-key node, one key per lane.  Returns, per lane, #keys <= x (0..32): a binary search
// whose "loads" are register shuffles.
template <typename C>
__device__ __forceinline__ uint32_t shfl_count_le(C v, C x) {
    uint32_t c = 0;
#pragma unroll
    for (int step = 16; step >= 1; step >>= 1) {
        C t = __shfl_sync(0xffffffffu, v, c + step - 1);
        if (t <= x) c += step;
    }
    C t = __shfl_sync(0xffffffffu, v, c);
    if (t <= x) c += 1;
    return c;
}

// Warp-cooperative upper bound for 32 queries on the SAME chromosome (warp-uniform chrom, n_c).  The warp
// descends the tree twice at once — for min(x) and max(x) over its lanes — one coalesced node load and one
// ballot per level (the in-node compare is a ballot + popc, not a loop); every lane's answer then lies in
// leaf nodes [n_lo, n_hi].  When these are the same or adjacent nodes (always, for coordinate-sorted
// batches) each lane finishes with shuffles on the keys already in registers.  Returns false (warp-uniform)
// when the bracket spans more nodes: the caller falls back to the per-lane descent.
template <typename C>
__device__ __forceinline__ bool warp_upper_bound(const C* __restrict__ g, const C* __restrict__ s,
                                                 const uint32_t* __restrict__ tab, uint32_t stride, int n_levels,
                                                 int stage_level, uint32_t chrom, C x, uint32_t n_c, uint32_t& result) {
    const uint32_t lane = lane_id();
    const C xmin = warp_min(x), xmax = warp_max(x);
    uint32_t n_lo = 0, n_hi = 0;
    C a_lo = 0, a_hi = 0;
    int l = n_levels - 1;
    for (; l >= stage_level && l >= 0; --l) {  // staged levels: shared memory
        const C* base = s + tab[l * stride + chrom];
        a_lo = base[(n_lo << kFanoutLog2) + lane];
        a_hi = (n_hi == n_lo) ? a_lo : base[(n_hi << kFanoutLog2) + lane];
        const uint32_t c_lo = __popc(__ballot_sync(0xffffffffu, a_lo <= xmin));
        const uint32_t c_hi = __popc(__ballot_sync(0xffffffffu, a_hi <= xmax));
        if (l > 0) {
            const uint32_t below = (tab[(l - 1) * stride + chrom + 1] - tab[(l - 1) * stride + chrom]) >> kFanoutLog2;
            n_lo = min((n_lo << kFanoutLog2) + c_lo, below - 1);
            n_hi = min((n_hi << kFanoutLog2) + c_hi, below - 1);
        }
    }
    for (; l >= 0; --l) {  // global levels (L1/L2)
        const C* base = g + tab[l * stride + chrom];
        a_lo = base[(n_lo << kFanoutLog2) + lane];
        a_hi = (n_hi == n_lo) ? a_lo : base[(n_hi << kFanoutLog2) + lane];
        const uint32_t c_lo = __popc(__ballot_sync(0xffffffffu, a_lo <= xmin));
        const uint32_t c_hi = __popc(__ballot_sync(0xffffffffu, a_hi <= xmax));
        if (l > 0) {
            const uint32_t below = (tab[(l - 1) * stride + chrom + 1] - tab[(l - 1) * stride + chrom]) >> kFanoutLog2;
            n_lo = min((n_lo << kFanoutLog2) + c_lo, below - 1);
            n_hi = min((n_hi << kFanoutLog2) + c_hi, below - 1);
        }
    }
    if (n_hi - n_lo > 1) return false;
    const uint32_t c1 = shfl_count_le<C>(a_lo, x);
    uint32_t r = (n_lo << kFanoutLog2) + c1;
    if (n_hi != n_lo) {
        const uint32_t c2 = shfl_count_le<C>(a_hi, x);
        if (c1 == kFanout) r += c2;
    }
    result = min(r, n_c);
    return true;
}

// Pass 1.  cnt_out[q] = number of hits, hi_out[q] = ub(S, q.end) (first rank starting past q.end) for
// pass 2, and tile_total[q / 4096] += hits (one 64-bit reduction per warp) so that the exclusive scan that
// follows needs no look-back chain: it reads each count once and writes each offset once (scan.cuh).
// err_flag bit 0 is set when a query has start > end (the reference cannot even construct such an
// interval: interval.hpp:70-74); such queries get zero hits.
template <typename C, typename Q>
__global__ void __launch_bounds__(1024)
intersect_count_kernel(IndexView<C> ix, const uint32_t* __restrict__ qchrom, const Q* __restrict__ qstart,
                       const Q* __restrict__ qend, size_t nq, uint32_t* __restrict__ cnt_out,
                       uint32_t* __restrict__ hi_out, unsigned long long* __restrict__ tile_total,
                       uint32_t* __restrict__ err_flag) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    // layout: [S prefix][E prefix][tab][mbarrier]
    C* sS = reinterpret_cast<C*>(smem_raw);
    C* sE = sS + ix.staged_elems;
    uint32_t* stab = reinterpret_cast<uint32_t*>(sE + ix.staged_elems);
    const uint32_t stride = ix.n_chrom + 1;
    const uint32_t tab_elems = static_cast<uint32_t>(ix.n_levels) * stride;
    uint64_t* bar = reinterpret_cast<uint64_t*>(stab + ((tab_elems + 3u) & ~3u));

    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const uint32_t bytes = ix.staged_elems * static_cast<uint32_t>(sizeof(C));
        mbar_arrive_expect_tx(bar, 2u * bytes);
        constexpr uint32_t kChunk = 32768;
        for (uint32_t o = 0; o < bytes; o += kChunk) {
            uint32_t len = min(kChunk, bytes - o);
            tma_bulk_g2s(reinterpret_cast<char*>(sS) + o, reinterpret_cast<const char*>(ix.S) + o, len, bar);
            tma_bulk_g2s(reinterpret_cast<char*>(sE) + o, reinterpret_cast<const char*>(ix.E) + o, len, bar);
        }
    }
    for (uint32_t i = threadIdx.x; i < tab_elems; i += blockDim.x) stab[i] = ix.tab[i];
    __syncthreads();
    mbar_wait(bar, 0);

    const size_t n_tiles = (nq + blockDim.x - 1) / blockDim.x;
    for (size_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const size_t q = tile * blockDim.x + threadIdx.x;
        const bool in_range = q < nq;
        uint32_t chrom = GG_NO_CHROM_DEV;
        Q qs_raw = 0, qe_raw = 0;
        if (in_range) {
            chrom = qchrom[q];
            qs_raw = qstart[q];
            qe_raw = qend[q];
        }
        const bool bad = qs_raw > qe_raw;
        if (bad) atomicOr(err_flag, 1u);
        // answerable: known chromosome, valid interval, start inside the index coordinate domain
        const bool ok = in_range && !bad && chrom < ix.n_chrom &&
                        !(sizeof(Q) > sizeof(C) && qs_raw > static_cast<Q>(coord_max<C>::value));
        const C qe = clamp_coord<C, Q>(qe_raw);
        const C qs = ok ? static_cast<C>(qs_raw) : C(0);
        const C xe = qs > 0 ? static_cast<C>(qs - 1) : C(0);  // #{end < q.start} = ub(E, q.start - 1)
        uint32_t cnt = 0, hi = 0, lo = 0;

        const uint32_t chrom0 = __shfl_sync(0xffffffffu, chrom, 0);
        bool done = false;
        if (__all_sync(0xffffffffu, ok && chrom == chrom0)) {
            const uint32_t n_c = ix.cnt[chrom0];
            done = warp_upper_bound<C>(ix.S, sS, stab, stride, ix.n_levels, ix.stage_level, chrom0, qe, n_c, hi) &&
                   warp_upper_bound<C>(ix.E, sE, stab, stride, ix.n_levels, ix.stage_level, chrom0, xe, n_c, lo);
        }
        if (!done && ok) {  // per-lane descent: mixed chromosomes, invalid lanes or scattered coordinates
            const uint32_t n_c = ix.cnt[chrom];
            hi = tree_upper_bound<C>(ix.S, sS, stab, stride, ix.n_levels, ix.stage_level, chrom, qe, n_c);
            lo = tree_upper_bound<C>(ix.E, sE, stab, stride, ix.n_levels, ix.stage_level, chrom, xe, n_c);
        }
        if (ok) {
            if (qs == 0) lo = 0;
            cnt = hi - lo;
        }
        if (in_range) {
            cnt_out[q] = cnt;
            hi_out[q] = hi;
        }
        // warp total -> its 4096-query scan tile (the warp's 32 queries never straddle a tile)
        unsigned long long wsum = cnt;
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) wsum += __shfl_xor_sync(0xffffffffu, wsum, d);
        if (lane_id() == 0 && wsum != 0) atomicAdd(&tile_total[(tile * blockDim.x + (threadIdx.x & ~31u)) / kScanTile], wsum);
    }
}

// 4 consecutive keys with one (u32) or two (u64) 16-byte loads; p must be 16-byte aligned
__device__ __forceinline__ void load4(const uint32_t* __restrict__ p, uint32_t& a, uint32_t& b, uint32_t& c, uint32_t& d) {
    uint4 v = *reinterpret_cast<const uint4*>(p);
    a = v.x; b = v.y; c = v.z; d = v.w;
}
__device__ __forceinline__ void load4(const uint64_t* __restrict__ p, uint64_t& a, uint64_t& b, uint64_t& c, uint64_t& d) {
    ulonglong2 v0 = *reinterpret_cast<const ulonglong2*>(p);
    ulonglong2 v1 = *reinterpret_cast<const ulonglong2*>(p + 2);
    a = v0.x; b = v0.y; c = v1.x; d = v1.y;
}

constexpr int kEmitStage = 512;  // staged hits per warp (2 KB): tiles of 32 queries with <= 512 hits are written coalesced

// Backward walk of one query: ranks hi-1, hi-2, ... in aligned groups of 4 (16-byte vector loads of the leaf
// ends), store(k, rank) for the k-th hit counted from the END of the query's list (k = n_hits-1 ... 0).
// At a node boundary the max-end levels skip whole nodes / subtrees without a hit.
template <typename C, typename Store>
__device__ __forceinline__ void emit_walk(const IndexView<C>& ix, const uint32_t* __restrict__ stab, uint32_t stride,
                                          int n_levels, uint32_t chrom, C qs, uint32_t hi, uint32_t n_hits, Store store) {
    const C* __restrict__ leaf_end = ix.M + stab[chrom];
    uint32_t r_hi = hi;   // ranks >= r_hi are done
    uint32_t rem = n_hits;
    C e0, e1, e2, e3;
    if ((r_hi & 3u) != 0 && rem != 0) {  // partial top group
        const uint32_t g4 = r_hi & ~3u, top = r_hi & 3u;
        load4(leaf_end + g4, e0, e1, e2, e3);
        if (top > 2 && e2 >= qs && rem != 0) store(--rem, g4 + 2);
        if (top > 1 && e1 >= qs && rem != 0) store(--rem, g4 + 1);
        if (e0 >= qs && rem != 0) store(--rem, g4);
        r_hi = g4;
    }
    while (rem != 0 && r_hi != 0) {
        if ((r_hi & (kFanout - 1)) == 0 && n_levels > 1) {
            // about to enter leaf node b from its last key: skip it (and, climbing while b is a last child,
            // whole subtrees) when the max-end augmentation says nothing in there reaches q.start
            uint32_t b = (r_hi >> kFanoutLog2) - 1;
            if (ix.M[stab[stride + chrom] + b] < qs) {
                int l = 1;
                uint32_t span = kFanout;
                while (l + 1 < n_levels && (b & (kFanout - 1)) == (kFanout - 1) &&
                       ix.M[stab[(l + 1) * stride + chrom] + (b >> kFanoutLog2)] < qs) {
                    b >>= kFanoutLog2;
                    ++l;
                    span <<= kFanoutLog2;
                }
                r_hi = (span > r_hi) ? 0u : r_hi - span;
                continue;
            }
        }
        const uint32_t g4 = r_hi - 4;
        load4(leaf_end + g4, e0, e1, e2, e3);
        if (e3 >= qs && rem != 0) store(--rem, g4 + 3);
        if (e2 >= qs && rem != 0) store(--rem, g4 + 2);
        if (e1 >= qs && rem != 0) store(--rem, g4 + 1);
        if (e0 >= qs && rem != 0) store(--rem, g4);
        r_hi = g4;
    }
}

// Pass 2.  hits[offsets[q] - hits_base ... offsets[q+1] - hits_base) = per-chromosome ranks, ascending.
// A warp owns 32 consecutive queries, whose output ranges are contiguous in the CSR: lanes drop their hits
// into a per-warp shared-memory stage and the warp then writes the stage out with coalesced stores.
template <typename C, typename Q>
__global__ void __launch_bounds__(kQueryThreads)
intersect_emit_kernel(IndexView<C> ix, const uint32_t* __restrict__ qchrom, const Q* __restrict__ qstart,
                      size_t nq, const uint32_t* __restrict__ hi_in, const uint64_t* __restrict__ offsets,
                      uint32_t* __restrict__ hits, uint64_t hits_base) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    uint32_t* stage = reinterpret_cast<uint32_t*>(smem_raw) + (threadIdx.x >> 5) * kEmitStage;
    uint32_t* stab = reinterpret_cast<uint32_t*>(smem_raw) + (kQueryThreads / 32) * kEmitStage;
    const uint32_t stride = ix.n_chrom + 1;
    const uint32_t tab_elems = static_cast<uint32_t>(ix.n_levels) * stride;
    for (uint32_t i = threadIdx.x; i < tab_elems; i += blockDim.x) stab[i] = ix.tab[i];
    __syncthreads();

    const int n_levels = ix.n_levels;
    const uint32_t lane = lane_id();
    const size_t n_tiles = (nq + kQueryThreads - 1) / kQueryThreads;
    for (size_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const size_t q = tile * kQueryThreads + threadIdx.x;
        const size_t qa = q < nq ? q : nq, qb = q + 1 < nq ? q + 1 : nq;
        const uint64_t o0 = offsets[qa], o1 = offsets[qb];
        const uint32_t n_hits = static_cast<uint32_t>(o1 - o0);
        const uint64_t wbase = __shfl_sync(0xffffffffu, o0, 0);
        const uint64_t wtotal = __shfl_sync(0xffffffffu, o1, 31) - wbase;
        if (wtotal == 0) continue;
        uint32_t chrom = 0, hi = 0;
        C qs = 0;
        if (n_hits != 0) {
            chrom = qchrom[q];
            qs = static_cast<C>(qstart[q]);  // n_hits > 0 implies q.start fits the index domain
            hi = hi_in[q];
        }
        if (wtotal <= kEmitStage) {
            uint32_t* dst = stage + static_cast<uint32_t>(o0 - wbase);
            if (n_hits != 0)
                emit_walk<C>(ix, stab, stride, n_levels, chrom, qs, hi, n_hits,
                             [dst](uint32_t k, uint32_t rank) { dst[k] = rank; });
            __syncwarp();
            uint32_t* out = hits + (wbase - hits_base);
            for (uint32_t i = lane; i < static_cast<uint32_t>(wtotal); i += 32) out[i] = stage[i];
            __syncwarp();
        } else if (n_hits != 0) {
            uint32_t* dst = hits + (o0 - hits_base);
            emit_walk<C>(ix, stab, stride, n_levels, chrom, qs, hi, n_hits,
                         [dst](uint32_t k, uint32_t rank) { dst[k] = rank; });
        }
    }
}

}  // namespace gg
