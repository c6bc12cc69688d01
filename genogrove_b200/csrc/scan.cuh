// scan.cuh — single-pass exclusive prefix sum (chained scan with decoupled look-back), hand-written.
//
// Replaces the growth of query_result's vector (include/genogrove/data_type/query_result.hpp:108-113)
// in the two-pass count -> scan -> emit scheme: offsets[i] = base + sum(counts[0..i)), offsets[n] = total.
// Also used by the radix sort for the digit-histogram offsets.
//
// One pass over the data: each 4096-item tile publishes its aggregate, then a warp looks back over
// predecessor tiles (32 at a time) until it meets a tile whose inclusive prefix is known.  Tiles are
// handed out by an atomic ticket so a tile only ever waits on tiles that already started.
// HBM traffic = sizeof(In) read + sizeof(Out) written per item: the algorithmic minimum.
#pragma once
#include "common.cuh"

namespace gg {

constexpr int kScanThreads = 256;
constexpr int kScanItems = 16;
constexpr int kScanTile = kScanThreads * kScanItems;

// per-tile descriptor: flag 0 = not ready, 1 = aggregate available, 2 = inclusive prefix available
struct ScanTileState {
    uint32_t* flag;   // [n_tiles]
    uint64_t* agg;    // [n_tiles]
    uint64_t* incl;   // [n_tiles]
    uint32_t* ticket; // [1]
};

// generic versions in tiles (used by the fused count+scan kernel, whose tile is one CTA of queries)
inline size_t tile_state_bytes(size_t tiles) { return round_up(tiles * 4 + 16, 16) + tiles * 16; }
inline size_t tile_state_zero_bytes(size_t tiles) { return round_up(tiles * 4 + 16, 16); }
inline ScanTileState tile_state_carve(void* scratch, size_t tiles) {
    ScanTileState s;
    char* p = static_cast<char*>(scratch);
    s.ticket = reinterpret_cast<uint32_t*>(p);
    s.flag = reinterpret_cast<uint32_t*>(p + 16);
    size_t head = round_up(tiles * 4 + 16, 16);
    s.agg = reinterpret_cast<uint64_t*>(p + head);
    s.incl = s.agg + tiles;
    return s;
}

// Decoupled look-back, executed by ONE WARP of the CTA that owns `tile`: publishes the tile aggregate and
// returns (to every lane) the exclusive prefix = sum of the aggregates of all earlier tiles.
__device__ __forceinline__ uint64_t tile_lookback(const ScanTileState& st, uint32_t tile, uint64_t block_total) {
    const uint32_t lane = threadIdx.x & 31u;
    uint64_t excl = 0;
    if (tile == 0) {
        if (lane == 0) {
            st.incl[0] = block_total;
            __threadfence();
            *reinterpret_cast<volatile uint32_t*>(&st.flag[0]) = 2u;
        }
        return 0;
    }
    if (lane == 0) {
        st.agg[tile] = block_total;
        __threadfence();
        *reinterpret_cast<volatile uint32_t*>(&st.flag[tile]) = 1u;
    }
    int64_t j = static_cast<int64_t>(tile) - 1;
    while (true) {
        int64_t idx = j - static_cast<int64_t>(lane);
        uint32_t f = 2u;  // tiles before tile 0 behave as "inclusive prefix = 0"
        if (idx >= 0) {
            do {
                f = ld_volatile(&st.flag[idx]);
            } while (f == 0u);
        }
        __threadfence();  // acquire: values below were published before the flag
        uint32_t incl_mask = __ballot_sync(0xffffffffu, f == 2u);
        uint32_t first_incl = incl_mask ? static_cast<uint32_t>(__ffs(incl_mask) - 1) : 32u;
        uint64_t val = 0;
        if (idx >= 0 && lane <= first_incl) val = (f == 2u) ? ld_volatile(&st.incl[idx]) : ld_volatile(&st.agg[idx]);
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) val += __shfl_xor_sync(0xffffffffu, val, d);
        excl += val;
        if (incl_mask) break;
        j -= 32;
    }
    if (lane == 0) {
        st.incl[tile] = excl + block_total;
        __threadfence();
        *reinterpret_cast<volatile uint32_t*>(&st.flag[tile]) = 2u;
    }
    return excl;
}

inline size_t scan_state_bytes(size_t n_items) {
    size_t tiles = div_up(n_items ? n_items : 1, kScanTile);
    return round_up(tiles * 4 + 16, 16) + tiles * 16;
}
// carve a ScanTileState out of a zeroed scratch buffer (zero it with cudaMemsetAsync before each launch)
inline ScanTileState scan_state_carve(void* scratch, size_t n_items) {
    size_t tiles = div_up(n_items ? n_items : 1, kScanTile);
    ScanTileState s;
    char* p = static_cast<char*>(scratch);
    s.ticket = reinterpret_cast<uint32_t*>(p);
    s.flag = reinterpret_cast<uint32_t*>(p + 16);
    // only flags + ticket need zeroing; agg/incl follow
    size_t head = round_up(tiles * 4 + 16, 16);
    s.agg = reinterpret_cast<uint64_t*>(p + head);
    s.incl = s.agg + tiles;
    return s;
}
inline size_t scan_state_zero_bytes(size_t n_items) {
    size_t tiles = div_up(n_items ? n_items : 1, kScanTile);
    return round_up(tiles * 4 + 16, 16);
}

// kLookback = true : single pass, tile prefixes by decoupled look-back (tile ids from an atomic ticket).
// kLookback = false: tile prefixes are given (st.incl[tile] = exclusive prefix of tile `tile`, e.g. the scanned
//                    per-tile totals the count kernel accumulated); no inter-CTA dependency at all.
template <typename In, typename Out, bool kLookback = true>
__global__ void __launch_bounds__(kScanThreads)
exclusive_scan_kernel(const In* __restrict__ in, Out* __restrict__ out, size_t n, ScanTileState st,
                      const uint64_t* __restrict__ base_ptr, uint64_t* __restrict__ total_out) {
    __shared__ uint64_t warp_sums[kScanThreads / 32];
    __shared__ uint64_t tile_excl;
    __shared__ uint32_t tile_id_s;

    if (kLookback) {
        if (threadIdx.x == 0) tile_id_s = atomicAdd(st.ticket, 1u);
        __syncthreads();
    }
    const uint32_t tile = kLookback ? tile_id_s : blockIdx.x;
    const size_t tile_base = static_cast<size_t>(tile) * kScanTile;
    const uint64_t base = base_ptr ? *base_ptr : 0ull;

    // blocked arrangement: thread t owns items [t*16, t*16+16) of the tile
    const size_t first = tile_base + static_cast<size_t>(threadIdx.x) * kScanItems;
    uint64_t v[kScanItems];
    uint64_t local = 0;
    const bool full = first + kScanItems <= n;
    if (full && sizeof(In) == 4 && (reinterpret_cast<uintptr_t>(in) & 15u) == 0) {
        // 16-byte vector loads (in + first is 64-byte aligned: cudaMalloc base, first % 16 == 0)
        const uint4* p4 = reinterpret_cast<const uint4*>(in + first);
#pragma unroll
        for (int k = 0; k < kScanItems / 4; ++k) {
            uint4 q = p4[k];
            v[4 * k + 0] = q.x;
            v[4 * k + 1] = q.y;
            v[4 * k + 2] = q.z;
            v[4 * k + 3] = q.w;
        }
    } else {
#pragma unroll
        for (int i = 0; i < kScanItems; ++i) {
            size_t idx = first + i;
            v[i] = (idx < n) ? static_cast<uint64_t>(in[idx]) : 0ull;
        }
    }
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) local += v[i];
    // warp inclusive scan of thread sums
    uint64_t inc = local;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint64_t o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane_id() >= static_cast<uint32_t>(d)) inc += o;
    }
    const uint32_t warp = threadIdx.x >> 5;
    if (lane_id() == 31) warp_sums[warp] = inc;
    __syncthreads();
    uint64_t warp_excl = 0, block_total = 0;
#pragma unroll
    for (int w = 0; w < kScanThreads / 32; ++w) {
        uint64_t s = warp_sums[w];
        if (static_cast<uint32_t>(w) < warp) warp_excl += s;
        block_total += s;
    }
    uint64_t thread_excl = warp_excl + inc - local;

    // decoupled look-back by warp 0
    if (warp == 0) {
        uint64_t excl = kLookback ? tile_lookback(st, tile, block_total) : st.incl[tile];
        if (lane_id() == 0) tile_excl = excl;
    }
    __syncthreads();
    uint64_t run = base + tile_excl + thread_excl;
    if (full && sizeof(Out) == 8 && (reinterpret_cast<uintptr_t>(out) & 15u) == 0) {
        ulonglong2* o2 = reinterpret_cast<ulonglong2*>(out + first);
#pragma unroll
        for (int k = 0; k < kScanItems / 2; ++k) {
            ulonglong2 w;
            w.x = run;
            run += v[2 * k];
            w.y = run;
            run += v[2 * k + 1];
            o2[k] = w;
        }
    } else {
#pragma unroll
        for (int i = 0; i < kScanItems; ++i) {
            size_t idx = first + i;
            if (idx < n) out[idx] = static_cast<Out>(run);
            run += v[i];
        }
    }
    if (n == 0 && tile == 0 && threadIdx.x == 0) {
        out[0] = static_cast<Out>(base);
        if (total_out) *total_out = 0;
    }
    // the thread owning the last item also writes the grand total at out[n]
    if (n > 0 && first <= n - 1 && n - 1 < first + kScanItems) {
        out[n] = static_cast<Out>(run);
        if (total_out) *total_out = run - base;
    }
}

// Enqueue: out[0..n] = exclusive prefix of in[0..n) (+ *base_ptr), total (without base) to *total_out.
template <typename In, typename Out>
inline cudaError_t exclusive_scan_async(const In* in, Out* out, size_t n, void* scratch,
                                        const uint64_t* base_ptr, uint64_t* total_out, cudaStream_t stream) {
    cudaError_t e = cudaMemsetAsync(scratch, 0, scan_state_zero_bytes(n), stream);
    if (e != cudaSuccess) return e;
    ScanTileState st = scan_state_carve(scratch, n);
    size_t tiles = div_up(n ? n : 1, kScanTile);
    exclusive_scan_kernel<In, Out><<<static_cast<unsigned>(tiles), kScanThreads, 0, stream>>>(in, out, n, st,
                                                                                             base_ptr, total_out);
    return cudaGetLastError();
}

// Two-level scan for the query counts.  tile_total[t] (u64, one per kScanTile counts, accumulated by the count
// kernel) is scanned first (a few CTAs), then every tile scans its own counts from its known prefix:
// counts are read once, offsets written once, and no CTA waits on another.
// scratch layout: [tile_total u64 x tiles][tile_prefix u64 x (tiles+1)][look-back state for the small scan]
inline size_t scan2_tiles(size_t n) { return div_up(n ? n : 1, kScanTile); }
inline size_t scan2_scratch_bytes(size_t n) {
    size_t t = scan2_tiles(n);
    return round_up(t * 8, 256) + round_up((t + 1) * 8, 256) + scan_state_bytes(t);
}
inline unsigned long long* scan2_tile_total(void* scratch) { return static_cast<unsigned long long*>(scratch); }
inline cudaError_t scan2_prepare_async(void* scratch, size_t n, cudaStream_t stream) {
    return cudaMemsetAsync(scratch, 0, scan2_tiles(n) * 8, stream);
}
inline cudaError_t scan2_finish_async(const uint32_t* counts, uint64_t* offsets, size_t n, void* scratch,
                                      const uint64_t* base_ptr, cudaStream_t stream) {
    size_t t = scan2_tiles(n);
    char* p = static_cast<char*>(scratch);
    uint64_t* tile_total = reinterpret_cast<uint64_t*>(p);
    uint64_t* tile_prefix = reinterpret_cast<uint64_t*>(p + round_up(t * 8, 256));
    void* state = p + round_up(t * 8, 256) + round_up((t + 1) * 8, 256);
    cudaError_t e = exclusive_scan_async<uint64_t, uint64_t>(tile_total, tile_prefix, t, state, nullptr, nullptr, stream);
    if (e != cudaSuccess) return e;
    ScanTileState st{};
    st.incl = tile_prefix;
    exclusive_scan_kernel<uint32_t, uint64_t, false><<<static_cast<unsigned>(t), kScanThreads, 0, stream>>>(
        counts, offsets, n, st, base_ptr, nullptr);
    return cudaGetLastError();
}

}  // namespace gg
