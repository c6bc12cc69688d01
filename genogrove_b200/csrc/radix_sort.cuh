// radix_sort.cuh — hand-written stable LSD radix sort (8-bit digits) of (key, u32 payload) pairs.
//
// Replaces the std::sort of (interval, data) pairs in the reference's bulk insert
// (include/genogrove/structure/grove/grove_insert.ipp:167-168).  The index sort key is
// (chrom, start, end[, tie]); it is sorted field by field, least significant field first, each
// field by its significant bytes only (constant high digits are skipped), carrying the permutation.
// Being stable, equal (chrom,start,end) keep input order — the order the reference's sorted-bulk path
// leaves in the leaf chain.
//
// Default form ("onesweep"): ONE histogram kernel counts all digits of the key at once; each pass is then a single kernel
// that reads a tile of (key, payload) pairs once, ranks them inside the CTA, learns where each of its 256 digit runs
// starts from a decoupled look-back over the tiles before it (one thread per digit; tiles are taken in ticket order, so
// the tiles a CTA waits for are always running or done), reorders the tile in shared memory and writes every digit run
// out contiguously.  Per pass the pairs cross DRAM once in each direction; the per-pass histogram read, the scan
// kernel over a tiles x 256 table and the 4-byte scattered stores of the first version are gone (10 M intervals,
// 14 passes: 2.76 ms -> see profiles/r2_summary.md).
// In-CTA ranks come from warp match-any ballots: items are taken in (warp, round, lane) order, so the rank of an item
// among equal digits is [items of that digit in earlier warps] + [in earlier rounds of this warp] + popc(peers below me).
// The three-kernel form (histogram, scan of the digit-major table, scatter) is kept for n >= 2^30, where a running digit
// count no longer fits beside the two status bits of a 32-bit look-back word.
#pragma once
#include <algorithm>
#include "common.cuh"
#include "scan.cuh"

namespace gg {

constexpr int kSortThreads = 256;
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kSortItems = 8;                             // per thread
constexpr int kSortTile = kSortThreads * kSortItems;      // 2048 keys per CTA
constexpr int kRadix = 256;

inline size_t radix_blocks(size_t n) { return div_up(n ? n : 1, kSortTile); }

// (in-warp ranks use match.any: a ballot-per-bit replacement was measured on B200 and is slower — 10 M keys, 14 passes:
// 3.0 ms against 2.14 ms.)
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

// ---------------------------------------------------------------------------------------------- onesweep form
constexpr int kOsThreads = 256;
constexpr int kOsWarps = kOsThreads / 32;
constexpr int kOsMaxPasses = 8;  // 64-bit keys, 8-bit digits
template <typename Key>
struct os_items {  // keys per thread: 48 bytes of keys + payloads either way
    static constexpr int value = sizeof(Key) == 8 ? 12 : 12;  // (16 for u32 keys spills at the 64 registers that 4 CTAs per SM allow)
};
constexpr int kOsTileMin = kOsThreads * 8;  // smallest tile of the instantiations (sizes the look-back table)
constexpr uint32_t kOsFlagAgg = 1u << 30, kOsFlagIncl = 2u << 30, kOsValMask = (1u << 30) - 1u;
constexpr int kOsWindow = 8;      // look-back words in flight per thread
constexpr size_t kOsMaxN = 1ull << 30;  // above: three-kernel form (a digit's running count needs 30 bits here)

inline size_t radix_os_tiles(size_t n) { return div_up(n ? n : 1, kOsTileMin); }
// layout of RadixWorkspace::hist for this form: [digit bases: 8 x 256][tile tickets: 8, padded to 64][state: tiles x 256]
constexpr size_t kOsHeadEntries = kOsMaxPasses * kRadix + 64;
// u32 entries needed for the histogram / look-back table of either form
inline size_t radix_hist_entries(size_t n) {
    const size_t legacy = radix_blocks(n) * kRadix + 1, os = kOsHeadEntries + radix_os_tiles(n) * kRadix;
    return legacy > os ? legacy : os;
}

// ghist[p][d] += number of keys whose p-th byte is d, for the first n_pass bytes.  A warp whose 32 keys share the byte
// (the high bytes of bounded coordinates, chromosome ids of a sorted file) adds once instead of 32 times to one word.
template <typename Key>
__global__ void __launch_bounds__(kOsThreads)
onesweep_hist_kernel(const Key* __restrict__ keys, size_t n, int n_pass, uint32_t* __restrict__ ghist) {
    __shared__ uint32_t h[kOsMaxPasses][kRadix];
    for (int i = threadIdx.x; i < kOsMaxPasses * kRadix; i += kOsThreads) (&h[0][0])[i] = 0;
    __syncthreads();
    const size_t stride = static_cast<size_t>(gridDim.x) * kOsThreads;
    const size_t n_round = round_up(n, 32);
    for (size_t i = static_cast<size_t>(blockIdx.x) * kOsThreads + threadIdx.x; i < n_round; i += stride) {
        const bool valid = i < n;
        const Key k = valid ? keys[i] : Key(0);
        const bool whole = __all_sync(0xffffffffu, valid);
        for (int p = 0; p < n_pass; ++p) {
            const uint32_t d = static_cast<uint32_t>(k >> (8 * p)) & 0xFFu;
            int same = 0;
            if (whole) __match_all_sync(0xffffffffu, d, &same);
            if (same) {
                if (lane_id() == 0) atomicAdd(&h[p][d], 32u);
            } else if (valid) {
                atomicAdd(&h[p][d], 1u);
            }
        }
    }
    __syncthreads();
    for (int p = 0; p < n_pass; ++p) {
        const uint32_t c = h[p][threadIdx.x];
        if (c) atomicAdd(&ghist[p * kRadix + threadIdx.x], c);
    }
}

// keys[i] = field[perm[i]] AND the digit counts of those keys in one pass over them (saves the histogram kernel's read
// of every field the index sort goes through: 5 x 35 us of a 10 M-interval build)
template <typename Field, typename Key>
__global__ void __launch_bounds__(kOsThreads)
gather_field_hist_kernel(const Field* __restrict__ field, const uint32_t* __restrict__ perm, Key* __restrict__ keys, size_t n,
                         int n_pass, uint32_t* __restrict__ ghist) {
    __shared__ uint32_t h[kOsMaxPasses][kRadix];
    for (int i = threadIdx.x; i < kOsMaxPasses * kRadix; i += kOsThreads) (&h[0][0])[i] = 0;
    __syncthreads();
    const size_t stride = static_cast<size_t>(gridDim.x) * kOsThreads;
    const size_t n_round = round_up(n, 32);
    for (size_t i = static_cast<size_t>(blockIdx.x) * kOsThreads + threadIdx.x; i < n_round; i += stride) {
        const bool valid = i < n;
        const Key k = valid ? static_cast<Key>(field[perm ? perm[i] : static_cast<uint32_t>(i)]) : Key(0);
        if (valid) keys[i] = k;
        const bool whole = __all_sync(0xffffffffu, valid);
        for (int p = 0; p < n_pass; ++p) {
            const uint32_t d = static_cast<uint32_t>(k >> (8 * p)) & 0xFFu;
            int same = 0;
            if (whole) __match_all_sync(0xffffffffu, d, &same);
            if (same) {
                if (lane_id() == 0) atomicAdd(&h[p][d], 32u);
            } else if (valid) {
                atomicAdd(&h[p][d], 1u);
            }
        }
    }
    __syncthreads();
    for (int p = 0; p < n_pass; ++p) {
        const uint32_t c = h[p][threadIdx.x];
        if (c) atomicAdd(&ghist[p * kRadix + threadIdx.x], c);
    }
}

// per pass: counts -> exclusive digit bases (one CTA of 256 threads per pass)
static __global__ void __launch_bounds__(kRadix) onesweep_bases_kernel(uint32_t* __restrict__ ghist) {
    __shared__ uint32_t wsum[kRadix / 32];
    uint32_t* row = ghist + blockIdx.x * kRadix;
    const uint32_t c = row[threadIdx.x];
    uint32_t inc = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane_id() >= static_cast<uint32_t>(d)) inc += o;
    }
    if (lane_id() == 31) wsum[threadIdx.x >> 5] = inc;
    __syncthreads();
    uint32_t before = 0;
    for (uint32_t w = 0; w < (threadIdx.x >> 5); ++w) before += wsum[w];
    row[threadIdx.x] = before + inc - c;
}

template <typename Key, int kThreads, int kItems>
constexpr size_t onesweep_smem_bytes() {
    return static_cast<size_t>(kThreads) * kItems * (sizeof(Key) + 4) + (kThreads / 32) * kRadix * 4 + 2 * kRadix * 4 + 128;
}

// One LSD pass over digit (key >> shift) & 255.  bases[d] = number of keys with a smaller digit (whole array);
// state[tile][d] = look-back word of tile `tile` for digit d (zeroed before the launch); ticket zeroed likewise.
template <typename Key, int kThreads, int kItems, int kMinBlocks>
__global__ void __launch_bounds__(kThreads, kMinBlocks)
onesweep_pass_kernel(const Key* __restrict__ keys_in, const uint32_t* __restrict__ vals_in, Key* __restrict__ keys_out,
                     uint32_t* __restrict__ vals_out, size_t n, int shift, const uint32_t* __restrict__ bases,
                     uint32_t* __restrict__ ticket, volatile uint32_t* state) {
    constexpr int kWarps = kThreads / 32;
    constexpr uint32_t kTile = kThreads * kItems;
    extern __shared__ __align__(16) unsigned char os_smem[];
    Key* skeys = reinterpret_cast<Key*>(os_smem);
    uint32_t* svals = reinterpret_cast<uint32_t*>(os_smem + static_cast<size_t>(kTile) * sizeof(Key));
    uint32_t(*cnt)[kRadix] = reinterpret_cast<uint32_t(*)[kRadix]>(svals + kTile);
    uint32_t* lstart = &cnt[0][0] + kWarps * kRadix;  // first position of the digit's run in the sorted tile
    uint32_t* gdst = lstart + kRadix;                   // global position of that run minus lstart (mod 2^32)
    uint32_t* misc = gdst + kRadix;                     // [0] ticket, [1..8] warp sums of the digit scan

    if (threadIdx.x == 0) misc[0] = atomicAdd(ticket, 1u);
    for (int i = threadIdx.x; i < kWarps * kRadix; i += kThreads) (&cnt[0][0])[i] = 0;
    __syncthreads();
    const uint32_t tile = misc[0];
    const uint32_t warp = threadIdx.x >> 5, lane = lane_id();
    const size_t tbase = static_cast<size_t>(tile) * kTile;
    const size_t wbase = tbase + static_cast<size_t>(warp) * (32 * kItems);
    const uint32_t lt_mask = (1u << lane) - 1u;
    Key k[kItems];
    uint32_t rank[kItems];
#pragma unroll
    for (int i = 0; i < kItems; ++i) {
        const size_t idx = wbase + static_cast<size_t>(i) * 32 + lane;
        k[i] = idx < n ? keys_in[idx] : Key(0);
    }
#pragma unroll
    for (int i = 0; i < kItems; ++i) {
        const size_t idx = wbase + static_cast<size_t>(i) * 32 + lane;
        const bool valid = idx < n;
        const uint32_t d = valid ? (static_cast<uint32_t>(k[i] >> shift) & 0xFFu) : 0xFFFFFFFFu;
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        // the lowest peer adds the group to the warp's counter of the digit and hands the old value round: one atomic in
        // flight per item instead of a load -> add -> store round trip (the ncu capture of the first version had 27 % of
        // its stall samples on that store); __syncwarp orders the rounds of one warp on the same counter
        const uint32_t leader = static_cast<uint32_t>(__ffs(peers) - 1);
        uint32_t prev = 0;
        if (valid && lane == leader) prev = atomicAdd(&cnt[warp][d], static_cast<uint32_t>(__popc(peers)));
        __syncwarp();
        prev = __shfl_sync(0xffffffffu, prev, leader);
        rank[i] = prev + __popc(peers & lt_mask);
    }
    __syncthreads();
    // thread t < 256 owns digit t: per-warp counts -> exclusive offsets across warps, tile count of the digit
    const uint32_t t = threadIdx.x;
    const bool owner = kThreads == kRadix || t < kRadix;
    uint32_t agg = 0;
    if (owner) {
#pragma unroll
        for (int w = 0; w < kWarps; ++w) {
            const uint32_t c = cnt[w][t];
            cnt[w][t] = agg;
            agg += c;
        }
    }
    // publish the tile's count of digit t right away; the tiles before this one are added up later, after the tile has
    // been reordered in shared memory — by then most of them have published too and the look-back rarely has to poll
    volatile uint32_t* my_state = state + static_cast<size_t>(tile) * kRadix + t;
    if (owner) *my_state = (tile == 0 ? kOsFlagIncl : kOsFlagAgg) | agg;
    // exclusive scan of the tile's digit counts -> lstart
    uint32_t inc = agg;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= static_cast<uint32_t>(d)) inc += o;
    }
    if (lane == 31 && owner) misc[1 + warp] = inc;
    __syncthreads();
    uint32_t before = 0;
#pragma unroll
    for (int w = 0; w < kRadix / 32; ++w)
        if (static_cast<uint32_t>(w) < warp) before += misc[1 + w];
    const uint32_t ls = before + inc - agg;
    if (owner) lstart[t] = ls;
    __syncthreads();
    // the tile in sorted order in shared memory
#pragma unroll
    for (int i = 0; i < kItems; ++i) {
        const size_t idx = wbase + static_cast<size_t>(i) * 32 + lane;
        if (idx < n) {
            const uint32_t d = static_cast<uint32_t>(k[i] >> shift) & 0xFFu;
            const uint32_t pos = lstart[d] + cnt[warp][d] + rank[i];
            skeys[pos] = k[i];
            rank[i] = pos;
        }
    }
    if (vals_in) {
#pragma unroll
        for (int i = 0; i < kItems; ++i) {
            const size_t idx = wbase + static_cast<size_t>(i) * 32 + lane;
            if (idx < n) svals[rank[i]] = vals_in[idx];
        }
    }
    // decoupled look-back, one thread per digit.  The words of kOsWindow predecessors are requested together: walking them
    // one dependent L2 round trip at a time costs ~30 round trips per tile (the chain back to the nearest finished tile
    // grows like sqrt(tiles in flight)).
    uint32_t excl = 0;
    if (tile != 0 && owner) {
        int64_t p = static_cast<int64_t>(tile) - 1;  // next tile to account for
        bool done = false;
        while (!done) {
            uint32_t w[kOsWindow];
#pragma unroll
            for (int j = 0; j < kOsWindow; ++j)
                w[j] = p - j >= 0 ? *(state + static_cast<size_t>(p - j) * kRadix + t) : (2u << 30);  // before tile 0: "inclusive, 0"
            bool stop = false;
            int used = 0;
#pragma unroll
            for (int j = 0; j < kOsWindow; ++j) {
                const uint32_t f = w[j] >> 30;
                if (!stop) {
                    if (f == 0u) {
                        stop = true;  // not published yet: ask again from here
                    } else {
                        excl += w[j] & kOsValMask;
                        ++used;
                        if (f == 2u) done = stop = true;
                    }
                }
            }
            p -= used;
        }
        *my_state = kOsFlagIncl | (excl + agg);
    }
    if (owner) gdst[t] = bases[t] + excl - ls;
    __syncthreads();
    const uint32_t count = n - tbase < kTile ? static_cast<uint32_t>(n - tbase) : kTile;
    for (uint32_t j = threadIdx.x; j < count; j += kThreads) {
        const Key key = skeys[j];
        const uint32_t dst = gdst[static_cast<uint32_t>(key >> shift) & 0xFFu] + j;
        keys_out[dst] = key;
        if (vals_in) vals_out[dst] = svals[j];
    }
}

struct RadixWorkspace {
    void* keys_a = nullptr;   // n * sizeof(Key) (sized for u64)
    void* keys_b = nullptr;
    uint32_t* vals_b = nullptr;  // n
    uint32_t* hist = nullptr;    // radix_hist_entries(n)
    void* scan_scratch = nullptr;  // scan_state_bytes(radix_hist_entries(n)) (three-kernel form only)
    int launches = 0;
    int force_legacy = 0;        // tests / A-B runs: the three-kernel form for any n
    int geom = 0;                // onesweep tile: 0 = 256 threads x 12 keys (4 CTAs/SM), 1 = 512 x 12 (2), 2 = 256 x 8 (5)
    bool hist_ready = false;     // the caller filled the digit counts at hist[0 .. 8 x 256) (gather_field_hist_kernel); one sort only
};
template <typename Key>
inline bool radix_uses_onesweep(size_t n, const RadixWorkspace& ws) { return n != 0 && n < kOsMaxN && !ws.force_legacy; }
inline unsigned onesweep_hist_blocks(size_t n) { return static_cast<unsigned>(std::min<size_t>(div_up(n ? n : 1, kOsThreads * 16), 148 * 8)); }

template <typename Key>
inline cudaError_t radix_sort_pairs_legacy(Key* keys, Key* keys_alt, uint32_t* vals, uint32_t* vals_alt, size_t n,
                                           int nbits, RadixWorkspace& ws, cudaStream_t stream, bool* in_alt) {
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

template <typename Key, int kThreads, int kItems, int kMinBlocks>
inline cudaError_t onesweep_passes(Key* keys, Key* keys_alt, uint32_t* vals, uint32_t* vals_alt, size_t n, int n_pass,
                                   RadixWorkspace& ws, cudaStream_t stream, bool* in_alt) {
    constexpr uint32_t kTile = kThreads * kItems;
    constexpr size_t kSmem = onesweep_smem_bytes<Key, kThreads, kItems>();
    auto kern = onesweep_pass_kernel<Key, kThreads, kItems, kMinBlocks>;
    // (per device, and a group handle may drive several devices from one process: set on every call, it costs microseconds)
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kSmem));
    if (e != cudaSuccess) return e;
    const uint32_t tiles = static_cast<uint32_t>(div_up(n, kTile));
    uint32_t* ghist = ws.hist;
    uint32_t* tickets = ws.hist + kOsMaxPasses * kRadix;
    uint32_t* state = ws.hist + kOsHeadEntries;
    Key* kin = keys;
    Key* kout = keys_alt;
    uint32_t* vin = vals;
    uint32_t* vout = vals_alt;
    for (int p = 0; p < n_pass; ++p) {
        e = cudaMemsetAsync(state, 0, static_cast<size_t>(tiles) * kRadix * 4, stream);
        if (e != cudaSuccess) return e;
        kern<<<tiles, kThreads, kSmem, stream>>>(kin, vin, kout, vout, n, 8 * p, ghist + p * kRadix, tickets + p, state);
        ws.launches += 1;
        Key* tk = kin; kin = kout; kout = tk;
        uint32_t* tv = vin; vin = vout; vout = tv;
        *in_alt = !*in_alt;
    }
    return cudaGetLastError();
}

// Stable sort of (keys, vals) by the low `nbits` bits of keys.  Ping-pongs between (keys, vals) and
// (keys_alt, vals_alt); returns through *in_alt whether the result ended up in the alt buffers.
template <typename Key>
inline cudaError_t radix_sort_pairs(Key* keys, Key* keys_alt, uint32_t* vals, uint32_t* vals_alt, size_t n,
                                    int nbits, RadixWorkspace& ws, cudaStream_t stream, bool* in_alt) {
    *in_alt = false;
    if (n == 0 || nbits <= 0) return cudaSuccess;
    if (n >= kOsMaxN || ws.force_legacy) return radix_sort_pairs_legacy<Key>(keys, keys_alt, vals, vals_alt, n, nbits, ws, stream, in_alt);
    const int n_pass = (nbits + 7) / 8;
    uint32_t* ghist = ws.hist;
    uint32_t* tickets = ws.hist + kOsMaxPasses * kRadix;
    cudaError_t e;
    if (ws.hist_ready) {  // counts are in place; the tickets behind them still need zeroing
        e = cudaMemsetAsync(tickets, 0, 64 * 4, stream);
        if (e != cudaSuccess) return e;
        ws.hist_ready = false;
    } else {
        e = cudaMemsetAsync(ws.hist, 0, kOsHeadEntries * 4, stream);
        if (e != cudaSuccess) return e;
        onesweep_hist_kernel<Key><<<onesweep_hist_blocks(n), kOsThreads, 0, stream>>>(keys, n, n_pass, ghist);
        ws.launches += 1;
    }
    onesweep_bases_kernel<<<n_pass, kRadix, 0, stream>>>(ghist);
    ws.launches += 1;
    if (ws.geom == 1) return onesweep_passes<Key, 512, 12, 2>(keys, keys_alt, vals, vals_alt, n, n_pass, ws, stream, in_alt);
    if (ws.geom == 2) return onesweep_passes<Key, 256, 8, 5>(keys, keys_alt, vals, vals_alt, n, n_pass, ws, stream, in_alt);
    return onesweep_passes<Key, 256, 12, 4>(keys, keys_alt, vals, vals_alt, n, n_pass, ws, stream, in_alt);
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
