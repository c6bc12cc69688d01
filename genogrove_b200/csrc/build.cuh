// build.cuh — device-side index construction kernels (after the radix sort): input validation and
// statistics, chain-order check, padding into the SoA leaf level, bottom-up level construction with
// per-node max-end augmentation.
//
// Replaces build_tree_bottom_up (include/genogrove/structure/grove/grove_insert.ipp:545-669) and the
// bounding-box bookkeeping of node::calc_keys_aggregate / calc_subtree_range (node.hpp:290-324).
// Node SHAPE is free on the device (hit lists do not depend on it, only on the key ORDER), so the
// device tree is implicit: fan-out 32, node j of level l covers ranks [j*32^l, (j+1)*32^l) of its
// chromosome, no pointers.  Three structure-of-arrays families share one offset table tab[l][c]:
//   S  level 0: start[] in chain order           level l>0: last start of each child  (pivots for
//   E  level 0: end[] sorted ascending           level l>0: last sorted end of each child   searches)
//   M  level 0: end[] in chain order             level l>0: max end of each child (max-end augmentation)
// Every chromosome segment of every level is padded to a whole number of 32-key nodes
// (pads: S,E = MAX, M = 0), so a node is one aligned 128-byte line (u32) and TMA sizes are 16-byte
// multiples.  Levels are stored top level first, so the hot upper levels are one contiguous prefix.
#pragma once
#include "common.cuh"

namespace gg {

struct BuildStats {
    unsigned long long max_coord;   // max over start/end
    unsigned long long max_tie;
    unsigned int invalid_interval;  // start > end seen
    unsigned int invalid_chrom;     // chrom >= n_chrom seen
    unsigned int unsorted;          // chain-order check failed
    unsigned int pad;
};

// max coordinate, validity, per-chromosome histogram.  Persistent CTAs; up to kStatBins chromosomes are counted in a
// shared-memory histogram first (one global atomic per CTA and chromosome: the per-warp global atomics of the first version
// took 0.64 ms of a 10 M-interval build on 24 hot addresses), more chromosomes than that go straight to global memory.
constexpr uint32_t kStatBins = 4096;
__global__ void __launch_bounds__(256)
index_stats_kernel(const uint32_t* __restrict__ chrom, const uint64_t* __restrict__ start, const uint64_t* __restrict__ end,
                   const uint64_t* __restrict__ tie, size_t n, uint32_t n_chrom, uint32_t* __restrict__ chrom_cnt,
                   BuildStats* __restrict__ st) {
    __shared__ uint32_t h[kStatBins];
    const bool local = n_chrom <= kStatBins;
    if (local) {
        for (uint32_t i = threadIdx.x; i < n_chrom; i += blockDim.x) h[i] = 0;
        __syncthreads();
    }
    unsigned long long mx = 0, mt = 0;
    bool bad_iv = false, bad_c = false;
    const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    const size_t n_round = round_up(n, 32);  // whole warps stay in the loop (match_any below)
    for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n_round; i += stride) {
        uint32_t c = 0xFFFFFFFFu;
        if (i < n) {
            const uint64_t s = start[i], e = end[i];
            const uint64_t m = e > s ? e : s;
            mx = m > mx ? m : mx;
            if (tie) {
                const uint64_t t = tie[i];
                mt = t > mt ? t : mt;
            }
            bad_iv |= s > e;
            c = chrom[i];
            if (c >= n_chrom) {
                bad_c = true;
                c = 0xFFFFFFFFu;
            }
        }
        const uint32_t peers = __match_any_sync(0xffffffffu, c);
        if (c != 0xFFFFFFFFu && lane_id() == static_cast<uint32_t>(__ffs(peers) - 1))
            atomicAdd(local ? &h[c] : &chrom_cnt[c], static_cast<uint32_t>(__popc(peers)));
    }
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
        unsigned long long o = __shfl_xor_sync(0xffffffffu, mx, d);
        mx = o > mx ? o : mx;
        unsigned long long t = __shfl_xor_sync(0xffffffffu, mt, d);
        mt = t > mt ? t : mt;
    }
    const bool any_iv = __any_sync(0xffffffffu, bad_iv), any_c = __any_sync(0xffffffffu, bad_c);
    if (lane_id() == 0) {
        if (mx > ld_volatile(&st->max_coord)) atomicMax(&st->max_coord, mx);
        if (mt > ld_volatile(&st->max_tie)) atomicMax(&st->max_tie, mt);
        if (any_iv) atomicOr(&st->invalid_interval, 1u);
        if (any_c) atomicOr(&st->invalid_chrom, 1u);
    }
    if (local) {
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < n_chrom; i += blockDim.x)
            if (h[i]) atomicAdd(&chrom_cnt[i], h[i]);
    }
}

// chain-order check for gg_index_upload_sorted: chrom non-decreasing, start non-decreasing inside a chrom
__global__ void check_chain_kernel(const uint32_t* __restrict__ chrom, const uint64_t* __restrict__ start, size_t n,
                                   BuildStats* __restrict__ st) {
    size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i + 1 < n) {
        uint32_t a = chrom[i], b = chrom[i + 1];
        if (a > b || (a == b && start[i] > start[i + 1])) atomicOr(&st->unsorted, 1u);
    }
}

template <typename C>
__global__ void narrow_kernel(const uint64_t* __restrict__ in, C* __restrict__ out, size_t n) {
    size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) out[i] = static_cast<C>(in[i]);
}

__global__ void iota_kernel(uint32_t* __restrict__ p, size_t n, int reverse) {
    size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) p[i] = reverse ? static_cast<uint32_t>(n - 1 - i) : static_cast<uint32_t>(i);
}

// chrom index owning absolute element offset p at one level: largest c with row[c] <= p
__device__ __forceinline__ uint32_t find_chrom(const uint32_t* __restrict__ row, uint32_t n_chrom, uint32_t p) {
    uint32_t lo = 0, hi = n_chrom;  // invariant: row[lo] <= p < row[hi]
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (row[mid] <= p) lo = mid; else hi = mid;
    }
    return lo;
}

// Leaf level: scatter the sorted chain into the padded layout.
//   perm_chain : chain order (rank -> input index), all chromosomes concatenated in id order
//   perm_end   : same ranks but ordered by (chrom, end) — for the E family
//   src_s/src_e: the coordinate arrays in INPUT order
//   cbegin[c]  : first concatenated rank of chromosome c (exclusive prefix of counts)
//   row0       : tab[0][*] (absolute offsets of the leaf level); level0_base = row0[0]
template <typename C>
__global__ void pad_leaf_kernel(const uint32_t* __restrict__ perm_chain, const uint32_t* __restrict__ perm_end,
                                const C* __restrict__ src_s, const C* __restrict__ src_e,
                                const uint32_t* __restrict__ cbegin, const uint32_t* __restrict__ cnt,
                                const uint32_t* __restrict__ row0, uint32_t n_chrom, C* __restrict__ S,
                                C* __restrict__ E, C* __restrict__ M) {
    uint32_t p = row0[0] + blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= row0[n_chrom]) return;
    uint32_t c = find_chrom(row0, n_chrom, p);
    uint32_t r = p - row0[c];
    if (r < cnt[c]) {
        uint32_t src = cbegin[c] + r;
        uint32_t i = perm_chain ? perm_chain[src] : src;
        S[p] = src_s[i];
        M[p] = src_e[i];
        E[p] = src_e[perm_end[src]];
    } else {
        S[p] = coord_max<C>::value;
        E[p] = coord_max<C>::value;
        M[p] = 0;
    }
}

// Level l >= 1, one warp per entry: entry j of chromosome c summarises child node j of level l-1
// (32 keys, one coalesced 128/256-byte line): S,E pivot = the child's last key, M = max over the child.
template <typename C>
__global__ void build_level_kernel(C* __restrict__ S, C* __restrict__ E, C* __restrict__ M,
                                   const uint32_t* __restrict__ row, const uint32_t* __restrict__ row_below,
                                   uint32_t n_chrom) {
    const uint32_t warps_per_block = blockDim.x >> 5;
    uint32_t entry = blockIdx.x * warps_per_block + (threadIdx.x >> 5);
    uint32_t p = row[0] + entry;
    if (p >= row[n_chrom]) return;
    uint32_t c = find_chrom(row, n_chrom, p);
    uint32_t j = p - row[c];
    uint32_t nodes_below = (row_below[c + 1] - row_below[c]) >> kFanoutLog2;
    C s = coord_max<C>::value, e = coord_max<C>::value, m = 0;
    if (j < nodes_below) {
        uint32_t child = row_below[c] + (j << kFanoutLog2) + lane_id();
        s = S[child];
        e = E[child];
        m = M[child];
    }
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
        C o = __shfl_xor_sync(0xffffffffu, m, d);
        m = o > m ? o : m;
    }
    s = __shfl_sync(0xffffffffu, s, 31);
    e = __shfl_sync(0xffffffffu, e, 31);
    if (lane_id() == 0) {
        S[p] = s;
        E[p] = e;
        M[p] = m;
    }
}

}  // namespace gg
