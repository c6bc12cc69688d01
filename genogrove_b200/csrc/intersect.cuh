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

// Pass 1.  cnt_out[q] = number of hits, hi_out[q] = ub(S, q.end) (first rank starting past q.end).
// err_flag bit 0 is set when a query has start > end (the reference cannot even construct such an
// interval: interval.hpp:70-74); such queries get zero hits.
template <typename C, typename Q>
__global__ void __launch_bounds__(1024)
intersect_count_kernel(IndexView<C> ix, const uint32_t* __restrict__ qchrom, const Q* __restrict__ qstart,
                       const Q* __restrict__ qend, size_t nq, uint32_t* __restrict__ cnt_out,
                       uint32_t* __restrict__ hi_out, uint32_t* __restrict__ err_flag) {
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
        if (q >= nq) continue;
        const uint32_t chrom = qchrom[q];
        const Q qs_raw = qstart[q];
        const Q qe_raw = qend[q];
        uint32_t cnt = 0, hi = 0;
        if (qs_raw > qe_raw) {
            atomicOr(err_flag, 1u);
        } else if (chrom < ix.n_chrom && !(sizeof(Q) > sizeof(C) && qs_raw > static_cast<Q>(coord_max<C>::value))) {
            const uint32_t n_c = ix.cnt[chrom];
            const C qe = clamp_coord<C, Q>(qe_raw);
            const C qs = static_cast<C>(qs_raw);
            hi = tree_upper_bound<C>(ix.S, sS, stab, stride, ix.n_levels, ix.stage_level, chrom, qe, n_c);
            uint32_t lo = 0;
            if (qs > 0) lo = tree_upper_bound<C>(ix.E, sE, stab, stride, ix.n_levels, ix.stage_level, chrom, static_cast<C>(qs - 1), n_c);
            cnt = hi - lo;
        }
        cnt_out[q] = cnt;
        hi_out[q] = hi;
    }
}

// Pass 2.  hits[offsets[q] - hits_base ... offsets[q+1] - hits_base) = per-chromosome ranks, ascending.
template <typename C, typename Q>
__global__ void __launch_bounds__(kQueryThreads)
intersect_emit_kernel(IndexView<C> ix, const uint32_t* __restrict__ qchrom, const Q* __restrict__ qstart,
                      size_t nq, const uint32_t* __restrict__ hi_in, const uint64_t* __restrict__ offsets,
                      uint32_t* __restrict__ hits, uint64_t hits_base) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    uint32_t* stab = reinterpret_cast<uint32_t*>(smem_raw);
    const uint32_t stride = ix.n_chrom + 1;
    const uint32_t tab_elems = static_cast<uint32_t>(ix.n_levels) * stride;
    for (uint32_t i = threadIdx.x; i < tab_elems; i += blockDim.x) stab[i] = ix.tab[i];
    __syncthreads();

    const int n_levels = ix.n_levels;
    const size_t n_tiles = (nq + kQueryThreads - 1) / kQueryThreads;
    for (size_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const size_t q = tile * kQueryThreads + threadIdx.x;
        if (q >= nq) continue;
        const uint64_t o0 = offsets[q], o1 = offsets[q + 1];
        uint32_t rem = static_cast<uint32_t>(o1 - o0);
        if (rem == 0) continue;
        const uint32_t chrom = qchrom[q];
        const C qs = static_cast<C>(qstart[q]);  // rem > 0 implies q.start fits the index domain
        const C* __restrict__ leaf_end = ix.M + stab[chrom];
        uint32_t* out = hits + (o1 - hits_base);
        int64_t r = static_cast<int64_t>(hi_in[q]) - 1;
        while (rem != 0 && r >= 0) {
            const uint32_t ru = static_cast<uint32_t>(r);
            if ((ru & (kFanout - 1)) == (kFanout - 1) && n_levels > 1) {
                // entering node b of level 0 from its last key: can the whole node be skipped?
                uint32_t b = ru >> kFanoutLog2;
                if (ix.M[stab[stride + chrom] + b] < qs) {
                    int l = 1;
                    int64_t span = kFanout;
                    // climb while b is the last child of its parent and the parent holds no hit either
                    while (l + 1 < n_levels && (b & (kFanout - 1)) == (kFanout - 1) &&
                           ix.M[stab[(l + 1) * stride + chrom] + (b >> kFanoutLog2)] < qs) {
                        b >>= kFanoutLog2;
                        ++l;
                        span <<= kFanoutLog2;
                    }
                    r -= span;
                    continue;
                }
            }
            if (leaf_end[ru] >= qs) {
                *--out = ru;
                --rem;
            }
            --r;
        }
    }
}

}  // namespace gg
