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
    // root directory (direct-address top level, see build_dir_kernel): for chromosome c and bin b,
    // dirS[dir_off[c] + b] = #{start < b << dir_shift}, dirE likewise over the sorted ends
    const uint32_t* dirS;
    const uint32_t* dirE;
    const uint32_t* dir_off;  // [n_chrom + 1]; chromosome c has dir_off[c+1] - dir_off[c] - 1 bins
    int dir_shift;
    uint32_t base0;           // = tab[0][0]: first element of the leaf level (level-0 "slot" = element - base0)
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

// ---- root directory ---------------------------------------------------------------------------------------
// The top of the flattened tree for bounded coordinate domains: instead of walking pivot levels from the root,
// a query key x indexes a direct-address table by x >> dir_shift and gets the two ranks that bracket its
// answer (bins hold ~8 keys on average), then finishes with a 2-4 step binary search on the leaf level.
// Built on the device from the leaf levels by the same tree descent it replaces.
template <typename C>
__global__ void build_dir_kernel(IndexView<C> ix, uint32_t* __restrict__ dirS, uint32_t* __restrict__ dirE) {
    const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= ix.dir_off[ix.n_chrom]) return;
    uint32_t lo = 0, hi = ix.n_chrom;  // chromosome owning directory entry e
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (ix.dir_off[mid] <= e) lo = mid; else hi = mid;
    }
    const uint32_t c = lo, b = e - ix.dir_off[c], n_c = ix.cnt[c];
    uint32_t vs = 0, ve = 0;
    if (b != 0) {
        const C top = coord_max<C>::value >> ix.dir_shift;  // largest bin id whose lower edge is representable
        if (static_cast<C>(b) > top) {
            vs = ve = n_c;
        } else {
            const C x = (static_cast<C>(b) << ix.dir_shift) - 1;  // #{key < b << shift} = ub(key, (b << shift) - 1)
            const uint32_t stride = ix.n_chrom + 1;
            vs = tree_upper_bound<C>(ix.S, ix.S, ix.tab, stride, ix.n_levels, ix.n_levels, c, x, n_c);
            ve = tree_upper_bound<C>(ix.E, ix.E, ix.tab, stride, ix.n_levels, ix.n_levels, c, x, n_c);
        }
    }
    dirS[e] = vs;
    dirE[e] = ve;
}

// 8 consecutive keys from a 32-byte (u32) / 64-byte (u64) aligned address
__device__ __forceinline__ void load8(const uint32_t* __restrict__ p, uint32_t (&e)[8]) {
    const uint4 a = *reinterpret_cast<const uint4*>(p);
    const uint4 b = *reinterpret_cast<const uint4*>(p + 4);
    e[0] = a.x; e[1] = a.y; e[2] = a.z; e[3] = a.w;
    e[4] = b.x; e[5] = b.y; e[6] = b.z; e[7] = b.w;
}
__device__ __forceinline__ void load8(const uint64_t* __restrict__ p, uint64_t (&e)[8]) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(p + 2 * k);
        e[2 * k] = v.x;
        e[2 * k + 1] = v.y;
    }
}
template <typename C>
__device__ __forceinline__ uint32_t count_le8(const C (&e)[8], C x) {
    uint32_t c = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) c += (e[k] <= x) ? 1u : 0u;
    return c;
}

// #keys <= x in the sorted leaf array a (32-key aligned, padded segment of one chromosome), given directory
// brackets: the answer lies in [lo, hi].  Two flavours, same result.  grouped = true: directory bins hold ~8 keys,
// so instead of the 3-4 DEPENDENT loads of a bisection the bracket is read as aligned groups of 8 keys (two 16-byte
// loads in flight, 1-2 groups on average); wider brackets are first narrowed by bisection.  This wins when every
// load is an L2 round trip (scattered batches: count pass 2.9 vs 3.7 ms on cfg4 shuffled) and loses when the loads
// hit L1 (coordinate-sorted batches: 0.77 vs 0.69 ms), so callers choose per warp.  Keys below lo are <= x and keys at or above hi are > x (or padding).
template <typename C>
__device__ __forceinline__ uint32_t bracket_upper_bound(const C* __restrict__ a, uint32_t lo, uint32_t hi, C x,
                                                        bool grouped = false) {
    if (!grouped) {  // plain bisection: fewest instructions; the loads hit L1 when neighbouring lanes search nearby
        while (lo < hi) {
            const uint32_t mid = (lo + hi) >> 1;
            if (a[mid] <= x) lo = mid + 1; else hi = mid;
        }
        return lo;
    }
    while (hi - lo > 16) {
        const uint32_t mid = (lo + hi) >> 1;
        if (a[mid] <= x) lo = mid + 1; else hi = mid;
    }
    uint32_t r = lo & ~7u;
    for (uint32_t g = r; g < hi && r == g; g += 8) {  // stops at the first group that holds a key > x
        C e[8];
        load8(a + g, e);
        r += count_le8<C>(e, x);
    }
    r = r > lo ? r : lo;  // (lo == hi with lo not a multiple of 8: no group was read)
    return r < hi ? r : hi;
}

// Pass 1, directory flavour (default): one query per thread, no shared memory, no barriers.
//   hi = ub(S, q.end), lo = ub(E, q.start - 1), each = directory lookup (2 adjacent u32) + short bracketed binary
//   search on the leaf level; count = hi - lo.  Coordinate-sorted batches make neighbouring lanes hit the same
//   directory sector and the same leaf lines (L1); scattered batches cost one L2 access per step.
template <typename C, typename Q>
__global__ void __launch_bounds__(kQueryThreads)
intersect_count_dir_kernel(IndexView<C> ix, const uint32_t* __restrict__ qchrom, const Q* __restrict__ qstart,
                           const Q* __restrict__ qend, size_t nq, uint32_t* __restrict__ cnt_out,
                           uint32_t* __restrict__ hi_out, unsigned long long* __restrict__ tile_total,
                           uint32_t* __restrict__ err_flag) {
    const size_t n_tiles = (nq + kQueryThreads - 1) / kQueryThreads;
    for (size_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const size_t q = tile * kQueryThreads + threadIdx.x;
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
        const bool ok = in_range && !bad && chrom < ix.n_chrom &&
                        !(sizeof(Q) > sizeof(C) && qs_raw > static_cast<Q>(coord_max<C>::value));
        uint32_t cnt = 0, hi = 0;
        if (ok) {
            const C qe = clamp_coord<C, Q>(qe_raw);
            const C qs = static_cast<C>(qs_raw);
            const uint32_t d0 = ix.dir_off[chrom], last_bin = ix.dir_off[chrom + 1] - d0 - 2;
            const uint32_t leaf = ix.tab[chrom];
            const C be = qe >> ix.dir_shift;
            const uint32_t b1 = be > static_cast<C>(last_bin) ? last_bin : static_cast<uint32_t>(be);
            hi = bracket_upper_bound<C>(ix.S + leaf, ix.dirS[d0 + b1], ix.dirS[d0 + b1 + 1], qe);
            uint32_t lo = 0;
            if (qs > 0) {
                const C xe = static_cast<C>(qs - 1);
                const C bs = xe >> ix.dir_shift;
                const uint32_t b2 = bs > static_cast<C>(last_bin) ? last_bin : static_cast<uint32_t>(bs);
                lo = bracket_upper_bound<C>(ix.E + leaf, ix.dirE[d0 + b2], ix.dirE[d0 + b2 + 1], xe);
            }
            cnt = hi - lo;
        }
        if (in_range) {
            cnt_out[q] = cnt;
            hi_out[q] = hi;
        }
        unsigned long long wsum = cnt;
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) wsum += __shfl_xor_sync(0xffffffffu, wsum, d);
        if (lane_id() == 0 && wsum != 0) atomicAdd(&tile_total[(tile * kQueryThreads + (threadIdx.x & ~31u)) / kScanTile], wsum);
    }
}

// ---- several consecutive queries per thread -------------------------------------------------------------
// N consecutive elements from a 16-byte aligned address (N * sizeof(T) is 8 or a multiple of 16); streaming loads:
// a query is read exactly once and must not push index lines out of L1/L2
template <typename T, int N>
__device__ __forceinline__ void load_vec_stream(const T* __restrict__ p, T (&v)[N]) {
    constexpr int kBytes = N * static_cast<int>(sizeof(T));
    if constexpr (kBytes == 8) {
        union { uint2 u; T t[N]; } x;
        x.u = __ldcs(reinterpret_cast<const uint2*>(p));
#pragma unroll
        for (int i = 0; i < N; ++i) v[i] = x.t[i];
    } else {
        static_assert(kBytes % 16 == 0, "vector load needs a multiple of 16 bytes");
        union { uint4 u[kBytes / 16]; T t[N]; } x;
#pragma unroll
        for (int k = 0; k < kBytes / 16; ++k) x.u[k] = __ldcs(reinterpret_cast<const uint4*>(p) + k);
#pragma unroll
        for (int i = 0; i < N; ++i) v[i] = x.t[i];
    }
}

template <int N>
__device__ __forceinline__ void store_vec_stream(uint64_t* __restrict__ p, const uint64_t (&v)[N]) {
    static_assert(N % 2 == 0, "pairs of u64");
#pragma unroll
    for (int k = 0; k < N / 2; ++k) {
        ulonglong2 x;
        x.x = v[2 * k];
        x.y = v[2 * k + 1];
        __stcs(reinterpret_cast<ulonglong2*>(p) + k, x);
    }
}

// Search state carried from one query of a thread to the next (see the header comment)
template <typename C>
struct Finger {
    uint32_t pos = 0, bhi = 0, bin = 0;
    C x = 0;
    bool valid = false;
};

// #keys <= x in the sorted leaf array a of one chromosome; dir = that chromosome's directory row
template <typename C>
__device__ __forceinline__ uint32_t finger_upper_bound(const C* __restrict__ a, const uint32_t* __restrict__ dir,
                                                       uint32_t last_bin, int shift, C x, Finger<C>& f, bool grouped) {
    const C bx = x >> shift;
    const uint32_t b = bx > static_cast<C>(last_bin) ? last_bin : static_cast<uint32_t>(bx);
    uint32_t h;
    if (f.valid && b == f.bin && x >= f.x) {
        h = f.pos;  // ub(x) lies in [ub(previous x), bracket end]; neighbouring queries advance by 0-2 keys
        int probes = 0;
        while (h < f.bhi && a[h] <= x) {
            ++h;
            if (++probes == 4) {
                h = bracket_upper_bound<C>(a, h, f.bhi, x, grouped);
                break;
            }
        }
    } else {
        const uint32_t blo = dir[b];
        f.bhi = dir[b + 1];
        f.bin = b;
        h = bracket_upper_bound<C>(a, blo, f.bhi, x, grouped);
    }
    f.pos = h;
    f.x = x;
    f.valid = true;
    return h;
}

// Count pass for QPT consecutive queries of one thread: hi[j] = ub(S, q.end), cnt[j] = hi[j] - ub(E, q.start - 1),
// qs[j] = q.start in index coordinates; queries that cannot hit (unknown chromosome, start beyond the index
// domain, start > end) get zero.  Returns true when some query had start > end (interval.hpp:70-74).
template <typename C, typename Q, int QPT>
__device__ __forceinline__ bool count_group(const IndexView<C>& ix, const uint32_t* __restrict__ tabp,
                                            const uint32_t (&chrom)[QPT], const Q (&qs_raw)[QPT], const Q (&qe_raw)[QPT],
                                            uint32_t (&cnt)[QPT], uint32_t (&hi)[QPT], C (&qs)[QPT], uint32_t (&slot)[QPT]) {
    Finger<C> fs, fe;
    uint32_t cur = GG_NO_CHROM_DEV, d0 = 0, last_bin = 0, leaf = 0;
    bool any_bad = false;
    const int shift = ix.dir_shift;
    // Do the warp's lanes search the same neighbourhood (same chromosome, within 32 directory bins of lane 0)?
    // Then the search loads hit L1 and plain bisection is cheapest; otherwise read the bracket as 8-key groups.
    const Q e0 = qe_raw[0], e_lane0 = __shfl_sync(0xffffffffu, e0, 0);
    const uint32_t c_lane0 = __shfl_sync(0xffffffffu, chrom[0], 0);  // (both shuffles before the &&: every lane takes part)
    const bool same_area = (((e0 ^ e_lane0) >> (shift + 5)) == 0) & (chrom[0] == c_lane0);
    const bool grouped = !__all_sync(0xffffffffu, same_area);
#pragma unroll
    for (int j = 0; j < QPT; ++j) {
        const bool bad = qs_raw[j] > qe_raw[j];
        any_bad |= bad;
        const bool ok = !bad && chrom[j] < ix.n_chrom &&
                        !(sizeof(Q) > sizeof(C) && qs_raw[j] > static_cast<Q>(coord_max<C>::value));
        cnt[j] = 0;
        hi[j] = 0;
        qs[j] = 0;
        slot[j] = 0;
        if (ok) {
            if (chrom[j] != cur) {
                cur = chrom[j];
                d0 = ix.dir_off[cur];
                last_bin = ix.dir_off[cur + 1] - d0 - 2;
                leaf = tabp[cur];
                fs.valid = fe.valid = false;
            }
            const C qe = clamp_coord<C, Q>(qe_raw[j]);
            qs[j] = static_cast<C>(qs_raw[j]);
            hi[j] = finger_upper_bound<C>(ix.S + leaf, ix.dirS + d0, last_bin, shift, qe, fs, grouped);
            uint32_t lo = 0;
            if (qs[j] > 0)
                lo = finger_upper_bound<C>(ix.E + leaf, ix.dirE + d0, last_bin, shift, static_cast<C>(qs[j] - 1), fe, grouped);
            cnt[j] = hi[j] - lo;
            slot[j] = leaf - ix.base0 + hi[j];  // level-0 slot of hi over all chromosomes (emit_tma.cuh walks on slots)
        }
    }
    return any_bad;
}

template <int N>
__device__ __forceinline__ void store_vec_u32(uint32_t* __restrict__ p, const uint32_t (&v)[N]) {
    if constexpr (N == 2) {
        *reinterpret_cast<uint2*>(p) = make_uint2(v[0], v[1]);
    } else {
        static_assert(N % 4 == 0, "groups of four u32");
#pragma unroll
        for (int k = 0; k < N / 4; ++k)
            reinterpret_cast<uint4*>(p)[k] = make_uint4(v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3]);
    }
}

// Pass 1, QPT consecutive queries per thread (directory + finger search, see count_group): 16-byte vector loads of
// the queries, 16-byte stores of cnt / hi, no shared memory, no barriers.  Same outputs as
// intersect_count_dir_kernel (tile_total, for the two-phase API's scan), and / or block_total[b] = hits of the
// b-th block of 32 * QPT consecutive queries (one warp's share; plain stores, no atomics) for the batch path,
// whose emit kernel turns scanned block totals into offsets itself.  A warp's queries never straddle a scan tile.
// kRec: instead of cnt_out / hi_out the kernel writes one record {count, level-0 slot of hi} per query into
// cnt_out viewed as uint2[nq] — the input of intersect_emit_tma_kernel (emit_tma.cuh), one TMA copy per 32 queries.
template <typename C, typename Q, int QPT, bool kTiles = false, bool kRec = false>
__global__ void __launch_bounds__(kQueryThreads)
intersect_count_group_kernel(IndexView<C> ix, const uint32_t* __restrict__ qchrom, const Q* __restrict__ qstart,
                             const Q* __restrict__ qend, size_t nq, uint32_t* __restrict__ cnt_out,
                             uint32_t* __restrict__ hi_out, unsigned long long* __restrict__ tile_total,
                             unsigned long long* __restrict__ block_total, uint32_t* __restrict__ err_flag, int vec_ok,
                             uint32_t* __restrict__ tile_total32 = nullptr) {
    static_assert(kScanTile % (32 * QPT) == 0, "a warp's queries must stay inside one scan tile");
    constexpr size_t kTileQ = static_cast<size_t>(kQueryThreads) * QPT;
    const size_t n_tiles = (nq + kTileQ - 1) / kTileQ;
    for (size_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const size_t q0 = tile * kTileQ + static_cast<size_t>(threadIdx.x) * QPT;
        const bool vec = vec_ok && q0 + QPT <= nq;
        uint32_t chrom[QPT];
        Q qs_raw[QPT], qe_raw[QPT];
        if (vec) {
            load_vec_stream<uint32_t, QPT>(qchrom + q0, chrom);
            load_vec_stream<Q, QPT>(qstart + q0, qs_raw);
            load_vec_stream<Q, QPT>(qend + q0, qe_raw);
        } else {
#pragma unroll
            for (int j = 0; j < QPT; ++j) {
                const bool in = q0 + j < nq;
                chrom[j] = in ? qchrom[q0 + j] : GG_NO_CHROM_DEV;
                qs_raw[j] = in ? qstart[q0 + j] : Q(0);
                qe_raw[j] = in ? qend[q0 + j] : Q(0);
            }
        }
        uint32_t cnt[QPT], hi[QPT], slot[QPT];
        C qs[QPT];
        if (count_group<C, Q, QPT>(ix, ix.tab, chrom, qs_raw, qe_raw, cnt, hi, qs, slot)) atomicOr(err_flag, 1u);
        if constexpr (kRec) {
            uint2* rec = reinterpret_cast<uint2*>(cnt_out);
            if (q0 + QPT <= nq) {  // (the record array is library scratch: always 16-byte aligned)
#pragma unroll
                for (int j = 0; j < QPT; j += 2)
                    *reinterpret_cast<uint4*>(rec + q0 + j) = make_uint4(cnt[j], slot[j], cnt[j + 1], slot[j + 1]);
            } else {
#pragma unroll
                for (int j = 0; j < QPT; ++j)
                    if (q0 + j < nq) rec[q0 + j] = make_uint2(cnt[j], slot[j]);
            }
        } else if (vec) {  // (cnt_out / hi_out are 16-byte aligned whenever vec_ok: the host checks)
            store_vec_u32<QPT>(cnt_out + q0, cnt);
            store_vec_u32<QPT>(hi_out + q0, hi);
        } else {
#pragma unroll
            for (int j = 0; j < QPT; ++j) {
                if (q0 + j < nq) {
                    cnt_out[q0 + j] = cnt[j];
                    hi_out[q0 + j] = hi[j];
                }
            }
        }
        unsigned long long wsum = 0;
#pragma unroll
        for (int j = 0; j < QPT; ++j) wsum += cnt[j];
        if constexpr (kTiles) {
            // sums over the 32 / QPT lanes of one 32-query tile first (per-tile totals for the dense emit), then the warp
#pragma unroll
            for (int d = 1; d < 32 / QPT; d <<= 1) wsum += __shfl_xor_sync(0xffffffffu, wsum, d);
            if ((lane_id() & (32 / QPT - 1)) == 0 && q0 < nq)
                tile_total32[q0 / 32] = wsum > 0xFFFFFFFFull ? 0xFFFFFFFFu : static_cast<uint32_t>(wsum);
#pragma unroll
            for (int d = 32 / QPT; d < 32; d <<= 1) wsum += __shfl_xor_sync(0xffffffffu, wsum, d);
        } else {
#pragma unroll
            for (int d = 16; d >= 1; d >>= 1) wsum += __shfl_xor_sync(0xffffffffu, wsum, d);
        }
        if (lane_id() == 0) {
            const size_t wq0 = tile * kTileQ + static_cast<size_t>(threadIdx.x & ~31u) * QPT;  // the warp's first query
            if (tile_total && wsum != 0) atomicAdd(&tile_total[wq0 / kScanTile], wsum);
            if (block_total && wq0 < nq) block_total[wq0 / (32 * QPT)] = wsum;  // hits of block = 32 * QPT queries
        }
    }
}

// ---- CTA-cooperative search window (coordinate-sorted batches) ------------------------------------------
constexpr int kWin = 512;  // keys per staged window (per family, per CTA)

// Warp-cooperative descents for FOUR search keys at once (ILP across the four dependent chains): ub(S, x[0]),
// ub(S, x[1]), ub(E, x[2]), ub(E, x[3]) on one chromosome.  One coalesced node load + one ballot per level
// and key.  All arguments warp-uniform; results warp-uniform.
template <typename C>
__device__ __forceinline__ void coop_upper_bound4(const IndexView<C>& ix, const C* __restrict__ sS, const C* __restrict__ sE,
                                                  const uint32_t* __restrict__ tab, uint32_t stride, uint32_t chrom,
                                                  uint32_t n_c, const C (&x)[4], uint32_t (&res)[4]) {
    const uint32_t lane = lane_id();
    uint32_t node[4] = {0, 0, 0, 0};
    for (int l = ix.n_levels - 1; l >= 0; --l) {
        const uint32_t off = tab[l * stride + chrom];
        const bool staged = l >= ix.stage_level;
        const C* bS = (staged ? sS : ix.S) + off;
        const C* bE = (staged ? sE : ix.E) + off;
        C a[4];
        a[0] = bS[(node[0] << kFanoutLog2) + lane];
        a[1] = bS[(node[1] << kFanoutLog2) + lane];
        a[2] = bE[(node[2] << kFanoutLog2) + lane];
        a[3] = bE[(node[3] << kFanoutLog2) + lane];
        const uint32_t below = l > 0 ? (tab[(l - 1) * stride + chrom + 1] - tab[(l - 1) * stride + chrom]) >> kFanoutLog2 : 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint32_t c = __popc(__ballot_sync(0xffffffffu, a[k] <= x[k]));
            const uint32_t child = (node[k] << kFanoutLog2) + c;
            if (l > 0) node[k] = min(child, below - 1);
            else res[k] = min(child, n_c);
        }
    }
}

// #keys <= x among win[0..ws) (sorted, in shared memory); pow2 = largest power of two <= ws (0 when ws == 0)
template <typename C>
__device__ __forceinline__ uint32_t window_count_le(const C* __restrict__ win, uint32_t ws, uint32_t pow2, C x) {
    uint32_t c = 0;
    for (uint32_t step = pow2; step != 0; step >>= 1) {
        const uint32_t t = c + step;
        if (t <= ws && win[t - 1] <= x) c = t;
    }
    return c;
}

// Pass 1.  cnt_out[q] = number of hits, hi_out[q] = ub(S, q.end) (first rank starting past q.end) for
// pass 2, and tile_total[q / 4096] += hits (one 64-bit reduction per warp) so that the exclusive scan that
// follows needs no look-back chain (scan.cuh).
//
// Three tiers, chosen per query tile (one CTA = blockDim.x consecutive queries), all giving identical results:
//   1. CTA window  — every query of the tile is on one chromosome: warp 0 descends the tree for the tile's
//      min/max search keys (4 cooperative descents, ballot + popc per node), which brackets every query's
//      answer inside S[s_lo, s_hi) and E[e_lo, e_hi).  If both brackets fit kWin keys they are copied into
//      shared memory by TMA bulk copies (cp.async.bulk + mbarrier) and each thread finishes with a short
//      binary search in shared memory.  This is the path coordinate-sorted batches take.
//   2. warp bracket — same idea per warp with register shuffles (warp_upper_bound).
//   3. per-lane descent of the tree (tree_upper_bound) — scattered queries.
// err_flag bit 0 is set when a query has start > end (the reference cannot even construct such an
// interval: interval.hpp:70-74); such queries get zero hits.
template <typename C, typename Q>
__global__ void __launch_bounds__(1024)
intersect_count_kernel(IndexView<C> ix, const uint32_t* __restrict__ qchrom, const Q* __restrict__ qstart,
                       const Q* __restrict__ qend, size_t nq, uint32_t* __restrict__ cnt_out,
                       uint32_t* __restrict__ hi_out, unsigned long long* __restrict__ tile_total,
                       uint32_t* __restrict__ err_flag) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    // layout: [S prefix][E prefix][winS][winE][tab][barriers]
    C* sS = reinterpret_cast<C*>(smem_raw);
    C* sE = sS + ix.staged_elems;
    C* winS = sE + ix.staged_elems;
    C* winE = winS + kWin;
    uint32_t* stab = reinterpret_cast<uint32_t*>(winE + kWin);
    const uint32_t stride = ix.n_chrom + 1;
    const uint32_t tab_elems = static_cast<uint32_t>(ix.n_levels) * stride;
    uint64_t* bar = reinterpret_cast<uint64_t*>(stab + ((tab_elems + 3u) & ~3u));  // bar[0]: staging, bar[1]: windows
    __shared__ uint32_t s_wchrom[32];
    __shared__ C s_wkey[4][32];   // per warp: min q.end, max q.end, min (q.start-1), max (q.start-1)
    __shared__ uint32_t s_bounds[8];

    if (threadIdx.x == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const uint32_t bytes = ix.staged_elems * static_cast<uint32_t>(sizeof(C));
        mbar_arrive_expect_tx(&bar[0], 2u * bytes);
        constexpr uint32_t kChunk = 32768;
        for (uint32_t o = 0; o < bytes; o += kChunk) {
            uint32_t len = min(kChunk, bytes - o);
            tma_bulk_g2s(reinterpret_cast<char*>(sS) + o, reinterpret_cast<const char*>(ix.S) + o, len, &bar[0]);
            tma_bulk_g2s(reinterpret_cast<char*>(sE) + o, reinterpret_cast<const char*>(ix.E) + o, len, &bar[0]);
        }
    }
    for (uint32_t i = threadIdx.x; i < tab_elems; i += blockDim.x) stab[i] = ix.tab[i];
    __syncthreads();
    mbar_wait(&bar[0], 0);

    constexpr uint32_t kAlign = 16 / sizeof(C);  // TMA source alignment in keys
    const uint32_t lane = lane_id(), warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    uint32_t phase = 0;
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

        // ---- tier 1: tile-wide bracket + TMA-staged windows
        const uint32_t chrom0 = __shfl_sync(0xffffffffu, chrom, 0);
        const bool warp_uniform = __all_sync(0xffffffffu, ok && chrom == chrom0);
        {
            const C e_min = warp_min(qe), e_max = warp_max(qe), s_min = warp_min(xe), s_max = warp_max(xe);
            if (lane == 0) {
                s_wchrom[warp] = warp_uniform ? chrom0 : GG_NO_CHROM_DEV;
                s_wkey[0][warp] = e_min;
                s_wkey[1][warp] = e_max;
                s_wkey[2][warp] = s_min;
                s_wkey[3][warp] = s_max;
            }
        }
        __syncthreads();
        if (warp == 0) {
            const uint32_t wl = lane < n_warps ? lane : 0;
            const uint32_t chrom_t = s_wchrom[0];
            const bool uni = __all_sync(0xffffffffu, s_wchrom[wl] == chrom_t) && chrom_t != GG_NO_CHROM_DEV;
            uint32_t mode = 0, r[4] = {0, 0, 0, 0}, a_s = 0, a_e = 0, len_s = 0, len_e = 0;
            if (uni) {
                C x[4];
                x[0] = warp_min(s_wkey[0][wl]);
                x[1] = warp_max(s_wkey[1][wl]);
                x[2] = warp_min(s_wkey[2][wl]);
                x[3] = warp_max(s_wkey[3][wl]);
                const uint32_t n_c = ix.cnt[chrom_t];
                coop_upper_bound4<C>(ix, sS, sE, stab, stride, chrom_t, n_c, x, r);
                a_s = r[0] & ~(kAlign - 1);
                a_e = r[2] & ~(kAlign - 1);
                len_s = (r[1] - a_s + kAlign - 1) & ~(kAlign - 1);
                len_e = (r[3] - a_e + kAlign - 1) & ~(kAlign - 1);
                mode = (len_s <= kWin && len_e <= kWin) ? 1u : 0u;
            }
            if (lane == 0) {
                s_bounds[0] = mode;
                s_bounds[1] = r[0];            // s_lo
                s_bounds[2] = r[1] - r[0];     // keys of S in the bracket
                s_bounds[3] = r[0] - a_s;      // offset of s_lo inside the staged window
                s_bounds[4] = r[2];            // e_lo
                s_bounds[5] = r[3] - r[2];
                s_bounds[6] = r[2] - a_e;
                if (mode) {
                    const uint32_t leaf = stab[chrom_t];
                    mbar_arrive_expect_tx(&bar[1], (len_s + len_e) * static_cast<uint32_t>(sizeof(C)));
                    if (len_s) tma_bulk_g2s(winS, ix.S + leaf + a_s, len_s * static_cast<uint32_t>(sizeof(C)), &bar[1]);
                    if (len_e) tma_bulk_g2s(winE, ix.E + leaf + a_e, len_e * static_cast<uint32_t>(sizeof(C)), &bar[1]);
                } else {
                    mbar_arrive(&bar[1]);
                }
            }
        }
        mbar_wait(&bar[1], phase);
        phase ^= 1u;
        if (s_bounds[0]) {
            const uint32_t ws = s_bounds[2], we = s_bounds[5];
            const uint32_t p2s = ws ? 1u << (31 - __clz(ws)) : 0u, p2e = we ? 1u << (31 - __clz(we)) : 0u;
            hi = s_bounds[1] + window_count_le<C>(winS + s_bounds[3], ws, p2s, qe);
            lo = s_bounds[4] + window_count_le<C>(winE + s_bounds[6], we, p2e, xe);
        } else {
            // ---- tier 2: warp bracket; tier 3: per-lane descent
            bool done = false;
            if (warp_uniform) {
                const uint32_t n_c = ix.cnt[chrom0];
                done = warp_upper_bound<C>(ix.S, sS, stab, stride, ix.n_levels, ix.stage_level, chrom0, qe, n_c, hi) &&
                       warp_upper_bound<C>(ix.E, sE, stab, stride, ix.n_levels, ix.stage_level, chrom0, xe, n_c, lo);
            }
            if (!done && ok) {
                const uint32_t n_c = ix.cnt[chrom];
                hi = tree_upper_bound<C>(ix.S, sS, stab, stride, ix.n_levels, ix.stage_level, chrom, qe, n_c);
                lo = tree_upper_bound<C>(ix.E, sE, stab, stride, ix.n_levels, ix.stage_level, chrom, xe, n_c);
            }
        }
        if (ok) {
            if (qs == 0) lo = 0;
            cnt = hi - lo;
        } else {
            hi = 0;
        }
        if (in_range) {
            cnt_out[q] = cnt;
            hi_out[q] = hi;
        }
        // warp total -> its 4096-query scan tile (the warp's 32 queries never straddle a tile)
        unsigned long long wsum = cnt;
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) wsum += __shfl_xor_sync(0xffffffffu, wsum, d);
        if (lane == 0 && wsum != 0) atomicAdd(&tile_total[(tile * blockDim.x + (threadIdx.x & ~31u)) / kScanTile], wsum);
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
constexpr int kEmitSlack = 4;    // spare slots in front of each warp's stage (see emit_walk_staged)

// Skip test at a leaf-node boundary (r_hi % 32 == 0): how many ranks below r_hi can be skipped because the
// max-end augmentation says no key in them reaches q.start (whole node, or whole subtrees while climbing).
template <typename C>
__device__ __forceinline__ uint32_t emit_skip_span(const IndexView<C>& ix, const uint32_t* __restrict__ stab, uint32_t stride,
                                                   int n_levels, uint32_t chrom, C qs, uint32_t r_hi) {
    uint32_t b = (r_hi >> kFanoutLog2) - 1;  // leaf node about to be entered from its last key
    if (!(ix.M[stab[stride + chrom] + b] < qs)) return 0;
    int l = 1;
    uint32_t span = kFanout;
    while (l + 1 < n_levels && (b & (kFanout - 1)) == (kFanout - 1) &&
           ix.M[stab[(l + 1) * stride + chrom] + (b >> kFanoutLog2)] < qs) {
        b >>= kFanoutLog2;
        ++l;
        span <<= kFanoutLog2;
    }
    return span > r_hi ? r_hi : span;
}

// Backward walk of one query: ranks hi-1, hi-2, ... in aligned groups of 4 (16-byte vector loads of the leaf
// ends).  Hits go to the shared-memory stage back to front: s_end = shared address one past the query's last
// slot, s_begin = address of its first slot; the walk ends when the list is full.  Stores are not individually
// bounds-checked: with offsets that match the counts the list fills exactly; with inconsistent caller input at
// most 3 stray slots below s_begin are touched, which stay inside the stage (kEmitSlack).
template <typename C>
__device__ __forceinline__ void emit_walk_staged(const IndexView<C>& ix, const uint32_t* __restrict__ stab, uint32_t stride,
                                                 int n_levels, uint32_t chrom, C qs, uint32_t hi, uint32_t s_begin,
                                                 uint32_t s_end) {
    const C* __restrict__ leaf_end = ix.M + stab[chrom];
    uint32_t r_hi = hi, sp = s_end;
    C e0, e1, e2, e3;
    if ((r_hi & 3u) != 0) {  // partial top group
        const uint32_t g4 = r_hi & ~3u, top = r_hi & 3u;
        load4(leaf_end + g4, e0, e1, e2, e3);
        if (top > 2 && e2 >= qs) { sp -= 4; sts32(sp, g4 + 2); }
        if (top > 1 && e1 >= qs) { sp -= 4; sts32(sp, g4 + 1); }
        if (e0 >= qs) { sp -= 4; sts32(sp, g4); }
        r_hi = g4;
    }
    const bool has_levels = n_levels > 1;
    while (sp > s_begin && r_hi != 0) {
        if ((r_hi & (kFanout - 1)) == 0 && has_levels) {
            const uint32_t span = emit_skip_span<C>(ix, stab, stride, n_levels, chrom, qs, r_hi);
            if (span) {
                r_hi -= span;
                continue;
            }
        }
        const uint32_t g4 = r_hi - 4;
        load4(leaf_end + g4, e0, e1, e2, e3);
        if (e3 >= qs) { sp -= 4; sts32(sp, g4 + 3); }
        if (e2 >= qs) { sp -= 4; sts32(sp, g4 + 2); }
        if (e1 >= qs) { sp -= 4; sts32(sp, g4 + 1); }
        if (e0 >= qs) { sp -= 4; sts32(sp, g4); }
        r_hi = g4;
    }
}

// Backward walk of one query, up to 16 ranks per step: the aligned group of 8 that holds rank r_hi-1 and, when it
// lies in the same 32-key node, the group below it — their four 16-byte loads are in flight together, which halves
// the dependent-load chain of the walk (a typical cfg4 query walks ~32 ranks for its ~7 hits).  Same contract as
// emit_walk_staged: hits go to shared memory back to front between s_begin and s_end; the walk ends when the list
// is full.  Exactly count(q) ranks below hi pass the test, so the list fills exactly.
template <typename C>
__device__ __forceinline__ void emit_walk_staged16(const IndexView<C>& ix, const uint32_t* __restrict__ tabp, uint32_t stride,
                                                  int n_levels, uint32_t chrom, C qs, uint32_t hi, uint32_t s_begin,
                                                  uint32_t s_end) {
    const C* __restrict__ leaf_end = ix.M + tabp[chrom];
    uint32_t r_hi = hi, sp = s_end;
    C ea[8], eb[8];
    const bool has_levels = n_levels > 1;
    while (sp > s_begin && r_hi != 0) {
        if ((r_hi & (kFanout - 1)) == 0 && has_levels) {
            const uint32_t span = emit_skip_span<C>(ix, tabp, stride, n_levels, chrom, qs, r_hi);
            if (span) {
                r_hi -= span;
                continue;
            }
        }
        const uint32_t ga = (r_hi - 1) & ~7u;          // group holding rank r_hi - 1
        const uint32_t top = r_hi - ga;                // 1..8: ranks of that group below r_hi
        const bool two = (ga & (kFanout - 1)) != 0;    // the group below is in the same node
        load8(leaf_end + ga, ea);
        if (two) load8(leaf_end + ga - 8, eb);
#pragma unroll
        for (int k = 7; k >= 0; --k) {
            if (static_cast<uint32_t>(k) < top && ea[k] >= qs) { sp -= 4; sts32(sp, ga + k); }
        }
        r_hi = ga;
        if (two && sp > s_begin) {
#pragma unroll
            for (int k = 7; k >= 0; --k) {
                if (eb[k] >= qs) { sp -= 4; sts32(sp, ga - 8 + k); }
            }
            r_hi = ga - 8;
        }
    }
}

// Backward walk of one query in aligned groups of 8 ranks (two independent 16-byte loads in flight per step); same
// contract as emit_walk_staged16.  The faster of the two when neighbouring lanes walk the same lines (L1 hits:
// coordinate-sorted batches); the 16-wide walk wins when every step is an L2 round trip (scattered batches).
template <typename C>
__device__ __forceinline__ void emit_walk_staged8(const IndexView<C>& ix, const uint32_t* __restrict__ tabp, uint32_t stride,
                                                  int n_levels, uint32_t chrom, C qs, uint32_t hi, uint32_t s_begin,
                                                  uint32_t s_end) {
    const C* __restrict__ leaf_end = ix.M + tabp[chrom];
    uint32_t r_hi = hi, sp = s_end;
    C ea[8];
    const bool has_levels = n_levels > 1;
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
    while (sp > s_begin && r_hi != 0) {
        if ((r_hi & (kFanout - 1)) == 0 && has_levels) {
            const uint32_t span = emit_skip_span<C>(ix, tabp, stride, n_levels, chrom, qs, r_hi);
            if (span) {
                r_hi -= span;
                continue;
            }
        }
        const uint32_t g8 = r_hi - 8;
        load8(leaf_end + g8, ea);
#pragma unroll
        for (int k = 7; k >= 0; --k) {
            if (ea[k] >= qs) { sp -= 4; sts32(sp, g8 + k); }
        }
        r_hi = g8;
    }
}

// Same walk writing straight to global memory with every store bounds-checked (warp tiles whose hit lists do
// not fit the stage: very long lists).
template <typename C>
__device__ __forceinline__ void emit_walk_global(const IndexView<C>& ix, const uint32_t* __restrict__ stab, uint32_t stride,
                                                 int n_levels, uint32_t chrom, C qs, uint32_t hi, uint32_t* __restrict__ dst,
                                                 uint32_t n_hits) {
    const C* __restrict__ leaf_end = ix.M + stab[chrom];
    uint32_t r_hi = hi, rem = n_hits;
    C e0, e1, e2, e3;
    if ((r_hi & 3u) != 0 && rem != 0) {
        const uint32_t g4 = r_hi & ~3u, top = r_hi & 3u;
        load4(leaf_end + g4, e0, e1, e2, e3);
        if (top > 2 && e2 >= qs && rem != 0) dst[--rem] = g4 + 2;
        if (top > 1 && e1 >= qs && rem != 0) dst[--rem] = g4 + 1;
        if (e0 >= qs && rem != 0) dst[--rem] = g4;
        r_hi = g4;
    }
    while (rem != 0 && r_hi != 0) {
        if ((r_hi & (kFanout - 1)) == 0 && n_levels > 1) {
            const uint32_t span = emit_skip_span<C>(ix, stab, stride, n_levels, chrom, qs, r_hi);
            if (span) {
                r_hi -= span;
                continue;
            }
        }
        const uint32_t g4 = r_hi - 4;
        load4(leaf_end + g4, e0, e1, e2, e3);
        if (e3 >= qs && rem != 0) dst[--rem] = g4 + 3;
        if (e2 >= qs && rem != 0) dst[--rem] = g4 + 2;
        if (e1 >= qs && rem != 0) dst[--rem] = g4 + 1;
        if (e0 >= qs && rem != 0) dst[--rem] = g4;
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
                      uint32_t* __restrict__ hits, uint64_t hits_base, int tab_in_smem,
                      const uint32_t* __restrict__ cnt_check = nullptr, uint32_t* __restrict__ err_flag = nullptr) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr uint32_t kStageWords = kEmitStage + kEmitSlack;
    const uint32_t* __restrict__ stab = ix.tab;  // many-contig groves: the offset table stays in global memory (L1)
    const uint32_t stage_addr = smem_u32(smem_raw) + ((threadIdx.x >> 5) * kStageWords + kEmitSlack) * 4u;
    const uint32_t stride = ix.n_chrom + 1;
    if (tab_in_smem) {
        uint32_t* st = reinterpret_cast<uint32_t*>(smem_raw) + (kQueryThreads / 32) * kStageWords;
        const uint32_t tab_elems = static_cast<uint32_t>(ix.n_levels) * stride;
        for (uint32_t i = threadIdx.x; i < tab_elems; i += blockDim.x) st[i] = ix.tab[i];
        stab = st;
        __syncthreads();
    }

    const int n_levels = ix.n_levels;
    const uint32_t lane = lane_id();
    const size_t n_tiles = (nq + kQueryThreads - 1) / kQueryThreads;
    for (size_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const size_t q = tile * kQueryThreads + threadIdx.x;
        const size_t qa = q < nq ? q : nq, qb = q + 1 < nq ? q + 1 : nq;
        const uint64_t o0 = offsets[qa], o1 = offsets[qb];
        const uint32_t n_hits = static_cast<uint32_t>(o1 - o0);
        // The offsets of the device two-phase API come back from the caller.  They must still be the ones the count pass
        // wrote: a list length that differs from the count kept on the handle would make the walk fill a wrong part of the
        // stage (or leave stale entries in it).  Such a tile is not written and bit 32 of the error flag reports it.
        if (cnt_check != nullptr && __any_sync(0xffffffffu, q < nq && (o1 < o0 || o1 - o0 != cnt_check[q]))) {
            if (lane == 0 && err_flag != nullptr) atomicOr(err_flag, 32u);
            continue;
        }
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
            if (n_hits != 0) {
                const uint32_t s_begin = stage_addr + static_cast<uint32_t>(o0 - wbase) * 4u;
                emit_walk_staged<C>(ix, stab, stride, n_levels, chrom, qs, hi, s_begin, s_begin + n_hits * 4u);
            }
            __syncwarp();
            uint32_t* out = hits + (wbase - hits_base) + lane;
            uint32_t sa = stage_addr + lane * 4u;
            for (uint32_t i = lane; i < static_cast<uint32_t>(wtotal); i += 32, out += 32, sa += 128) *out = lds32(sa);
            __syncwarp();
        } else if (n_hits != 0) {
            emit_walk_global<C>(ix, stab, stride, n_levels, chrom, qs, hi, hits + (o0 - hits_base), n_hits);
        }
    }
}

}  // namespace gg
