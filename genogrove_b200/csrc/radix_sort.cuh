// radix_sort.cuh — hand-written stable LSD radix sort (8-bit digits) of (key, u32 payload) pairs.
//
// Replaces the std::sort of (interval, data) pairs in the reference's bulk insert
// (include/genogrove/structure/grove/grove_insert.ipp:167-168).  The index sort key is
// (chrom, start, end[, tie]); it is sorted field by field, least significant field first, each
// field by its significant bytes only (constant high digits are skipped), carrying the permutation.
// Being stable, equal (chrom,start,end) keep input order — the order the reference's sorted-bulk path
// leaves in the leaf chain.
//
// Per pass: (1) per-CTA digit histogram, (2) exclusive scan of the digit-major histogram
// (scan.cuh), (3) stable scatter.  In-CTA ranks come from warp match-any ballots: items are taken in
// (warp, round, lane) order, so the rank of an item among equal digits is
//   [items of that digit in earlier warps] + [in earlier rounds of this warp] + popc(peers below me).
#pragma once
#include "common.cuh"
#include "scan.cuh"

namespace gg {

constexpr int kSortThreads = 256;
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kSortItems = 8;                             // per thread
constexpr int kSortTile = kSortThreads * kSortItems;      // 2048 keys per CTA
constexpr int kRadix = 256;

inline size_t radix_blocks(size_t n) { return div_up(n ? n : 1, kSortTile); }
// u32 entries needed for the histogram/offset table (+1 for the scan's total slot)
inline size_t radix_hist_entries(size_t n) { return radix_blocks(n) * kRadix + 1; }

template <typename Key>
__global__ void __launch_bounds__(kSortThreads)
radix_hist_kernel(const Key* __restrict__ keys, size_t n, int shift, uint32_t* __restrict__ hist, uint32_t nblocks) {
    __shared__ uint32_t h[kRadix];
    h[threadIdx.x] = 0;
    __syncthreads();
    const size_t base = static_cast<size_t>(blockIdx.x) * kSortTile;
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        size_t idx = base + static_cast<size_t>(i) * kSortThreads + threadIdx.x;
        if (idx < n) atomicAdd(&h[static_cast<uint32_t>(keys[idx] >> shift) & 0xFFu], 1u);
    }
    __syncthreads();
    hist[static_cast<size_t>(threadIdx.x) * nblocks + blockIdx.x] = h[threadIdx.x];
}

template <typename Key>
__global__ void __launch_bounds__(kSortThreads)
radix_scatter_kernel(const Key* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                     Key* __restrict__ keys_out, uint32_t* __restrict__ vals_out, size_t n, int shift,
                     const uint32_t* __restrict__ offs, uint32_t nblocks) {
    __shared__ uint32_t cnt[kSortWarps][kRadix];
    __shared__ uint32_t base[kRadix];
    for (int i = threadIdx.x; i < kSortWarps * kRadix; i += kSortThreads) (&cnt[0][0])[i] = 0;
    __syncthreads();

    const uint32_t warp = threadIdx.x >> 5, lane = lane_id();
    const size_t wbase = static_cast<size_t>(blockIdx.x) * kSortTile + static_cast<size_t>(warp) * (32 * kSortItems);
    const uint32_t lt_mask = (1u << lane) - 1u;
    Key k[kSortItems];
    uint32_t rank[kSortItems];
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        size_t idx = wbase + static_cast<size_t>(i) * 32 + lane;
        bool valid = idx < n;
        k[i] = valid ? keys_in[idx] : Key(0);
        uint32_t d = valid ? (static_cast<uint32_t>(k[i] >> shift) & 0xFFu) : 0xFFFFFFFFu;
        uint32_t peers = __match_any_sync(0xffffffffu, d);
        uint32_t prev = valid ? cnt[warp][d] : 0u;
        __syncwarp();
        if (valid && lane == static_cast<uint32_t>(__ffs(peers) - 1)) cnt[warp][d] = prev + __popc(peers);
        __syncwarp();
        rank[i] = prev + __popc(peers & lt_mask);
    }
    __syncthreads();
    {
        // digit t: turn per-warp counts into exclusive offsets across warps; fetch the global base
        uint32_t t = threadIdx.x, run = 0;
#pragma unroll
        for (int w = 0; w < kSortWarps; ++w) {
            uint32_t c = cnt[w][t];
            cnt[w][t] = run;
            run += c;
        }
        base[t] = offs[static_cast<size_t>(t) * nblocks + blockIdx.x];
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        size_t idx = wbase + static_cast<size_t>(i) * 32 + lane;
        if (idx < n) {
            uint32_t d = static_cast<uint32_t>(k[i] >> shift) & 0xFFu;
            uint32_t dst = base[d] + cnt[warp][d] + rank[i];
            keys_out[dst] = k[i];
            if (vals_in) vals_out[dst] = vals_in[idx];
        }
    }
}

// keys[i] = field[perm[i]] (perm == nullptr: identity), widened/narrowed to Key
template <typename Field, typename Key>
__global__ void gather_field_kernel(const Field* __restrict__ field, const uint32_t* __restrict__ perm,
                                    Key* __restrict__ keys, size_t n) {
    size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) keys[i] = static_cast<Key>(field[perm ? perm[i] : static_cast<uint32_t>(i)]);
}

struct RadixWorkspace {
    void* keys_a = nullptr;   // n * sizeof(Key) (sized for u64)
    void* keys_b = nullptr;
    uint32_t* vals_b = nullptr;  // n
    uint32_t* hist = nullptr;    // radix_hist_entries(n)
    void* scan_scratch = nullptr;
    int launches = 0;
};

// Stable sort of (keys, vals) by the low `nbits` bits of keys.  Ping-pongs between (keys, vals) and
// (keys_alt, vals_alt); returns through *in_alt whether the result ended up in the alt buffers.
template <typename Key>
inline cudaError_t radix_sort_pairs(Key* keys, Key* keys_alt, uint32_t* vals, uint32_t* vals_alt, size_t n,
                                    int nbits, RadixWorkspace& ws, cudaStream_t stream, bool* in_alt) {
    *in_alt = false;
    if (n == 0) return cudaSuccess;
    const uint32_t nblocks = static_cast<uint32_t>(radix_blocks(n));
    Key* kin = keys;
    Key* kout = keys_alt;
    uint32_t* vin = vals;
    uint32_t* vout = vals_alt;
    for (int shift = 0; shift < nbits; shift += 8) {
        radix_hist_kernel<Key><<<nblocks, kSortThreads, 0, stream>>>(kin, n, shift, ws.hist, nblocks);
        cudaError_t e = exclusive_scan_async<uint32_t, uint32_t>(ws.hist, ws.hist, static_cast<size_t>(nblocks) * kRadix,
                                                                 ws.scan_scratch, nullptr, nullptr, stream);
        if (e != cudaSuccess) return e;
        radix_scatter_kernel<Key><<<nblocks, kSortThreads, 0, stream>>>(kin, vin, kout, vout, n, shift, ws.hist, nblocks);
        ws.launches += 3;
        Key* tk = kin; kin = kout; kout = tk;
        uint32_t* tv = vin; vin = vout; vout = tv;
        *in_alt = !*in_alt;
    }
    return cudaGetLastError();
}

inline int bits_needed(uint64_t max_value) {
    int b = 0;
    while (max_value) {
        ++b;
        max_value >>= 1;
    }
    return b;
}

}  // namespace gg
