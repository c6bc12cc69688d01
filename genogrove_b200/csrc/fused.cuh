// fused.cuh — the whole batched intersect in ONE kernel: count -> (decoupled look-back scan) -> emit.
//
// Replaces, per query, grove::intersect -> detail::search_overlaps
// (include/genogrove/structure/grove/grove_query.ipp:33-43, query_engine.hpp:76-134) and the growth of
// query_result's vector (data_type/query_result.hpp:108-113); results are identical to the two-pass
// count/scan/emit kernels in intersect.cuh (which stay as the two-phase API), but every query is read once,
// its CSR offset written once and its hits written once: HBM traffic = SURVEY 8(d)'s algorithmic bytes
// B_q = 12*Q + 8*(Q+1) + 4*H and nothing else (no count / hi scratch round trip, no scan kernels).
//
// One CTA owns a tile of 256 consecutive queries (one per thread); tiles are handed out by an atomic ticket
// so a tile only ever waits on tiles that already started (forward progress without co-residency assumptions):
//   1. hi = ub(S, q.end), lo = ub(E, q.start - 1) through the root directory + bracketed leaf search,
//      count = hi - lo (no leaf walk).
//   2. warp scan + CTA scan of the counts; the tile aggregate is published at once (packed 64-bit word:
//      2 flag bits + 62 value bits, so one store publishes flag and value together, no fence pairs).
//   3. backward leaf walk with max-end skipping (emit_walk_staged8) into the warp's shared-memory stage —
//      this needs only the warp-local prefix, so it runs while predecessors are still counting.
//   4. warp 0 resolves the tile's global prefix by looking back over predecessor tiles 32 at a time.
//   5. offsets[q] = base + prefix (coalesced 8-byte stores); the stage is written out with coalesced stores
//      at hits[prefix ...] (chunk-local, i.e. without base).  No atomics on the output, reference order.
// Warp tiles with more than kEmitStage hits walk straight to global memory after step 4.
#pragma once
#include "intersect.cuh"

namespace gg {

constexpr uint64_t kFlagAgg = 1ull << 62;   // value = aggregate of this tile
constexpr uint64_t kFlagIncl = 2ull << 62;  // value = inclusive prefix up to and including this tile
constexpr uint64_t kFlagMask = 3ull << 62;
constexpr int kFusedSlack = 8;

// scratch layout: [ticket u32, pad to 16 B][state u64 x tiles]; all of it zeroed before each launch
inline size_t fused_tiles(size_t nq) { return div_up(nq ? nq : 1, kQueryThreads); }
inline size_t fused_scratch_bytes(size_t nq) { return 16 + fused_tiles(nq) * 8; }

__device__ __forceinline__ void st_relaxed_u64(uint64_t* p, uint64_t v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_relaxed_u64(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// kMinCtas: resident CTAs per SM the register allocation targets (8 -> 32 registers, 6 -> 40)
template <typename C, typename Q, int kMinCtas>
__global__ void __launch_bounds__(kQueryThreads, kMinCtas)
intersect_fused_kernel(IndexView<C> ix, const uint32_t* __restrict__ qchrom, const Q* __restrict__ qstart,
                       const Q* __restrict__ qend, size_t nq, uint64_t* __restrict__ offsets,
                       uint32_t* __restrict__ hits, uint64_t capacity, const uint64_t* __restrict__ base_ptr,
                       uint64_t* __restrict__ next_base_out, uint64_t* __restrict__ total_out,
                       uint32_t* __restrict__ ticket, uint64_t* __restrict__ state, uint32_t* __restrict__ err_flag,
                       int tab_in_smem) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr uint32_t kStageWords = kEmitStage + kFusedSlack;
    constexpr uint32_t kWarps = kQueryThreads / 32;
    uint32_t* stab = reinterpret_cast<uint32_t*>(smem_raw) + kWarps * kStageWords;
    __shared__ uint64_t s_wsum[kWarps];
    __shared__ uint64_t s_prefix;
    __shared__ uint32_t s_tile;

    const uint32_t lane = lane_id(), warp = threadIdx.x >> 5;
    const uint32_t stage_addr = smem_u32(smem_raw) + (warp * kStageWords + kFusedSlack) * 4u;
    const uint32_t stride = ix.n_chrom + 1;
    const int n_levels = ix.n_levels;
    const uint32_t* __restrict__ tabp = ix.tab;
    if (tab_in_smem) {
        const uint32_t tab_elems = static_cast<uint32_t>(n_levels) * stride;
        for (uint32_t i = threadIdx.x; i < tab_elems; i += blockDim.x) stab[i] = ix.tab[i];
        tabp = stab;
    }
    const uint64_t base = base_ptr ? *base_ptr : 0ull;
    const size_t n_tiles = (nq + kQueryThreads - 1) / kQueryThreads;

    while (true) {
        if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
        __syncthreads();  // ticket visible; also orders the stab fill and the previous tile's shared reads
        const uint32_t tile = s_tile;
        if (tile >= n_tiles) break;

        // ---- 1. count
        const size_t q = static_cast<size_t>(tile) * kQueryThreads + threadIdx.x;
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
        C qs = 0;
        if (ok) {
            const C qe = clamp_coord<C, Q>(qe_raw);
            qs = static_cast<C>(qs_raw);
            const uint32_t d0 = ix.dir_off[chrom], last_bin = ix.dir_off[chrom + 1] - d0 - 2;
            const uint32_t leaf = tabp[chrom];
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

        // ---- 2. warp scan (64-bit: 32 counts of up to 2^32-1 each), CTA scan, publish the tile aggregate
        uint64_t inc = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint64_t o = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= static_cast<uint32_t>(d)) inc += o;
        }
        const uint64_t wtotal = __shfl_sync(0xffffffffu, inc, 31);
        const uint64_t lane_excl = inc - cnt;  // prefix inside the warp
        if (lane == 31) s_wsum[warp] = inc;
        __syncthreads();
        uint64_t warp_excl = 0, tile_total = 0;
#pragma unroll
        for (uint32_t w = 0; w < kWarps; ++w) {
            const uint64_t s = s_wsum[w];
            if (w < warp) warp_excl += s;
            tile_total += s;
        }
        if (threadIdx.x == 0) st_relaxed_u64(&state[tile], (tile == 0 ? kFlagIncl : kFlagAgg) | tile_total);

        // ---- 3. staged walk (needs only warp-local prefixes)
        const bool staged = wtotal <= kEmitStage;
        if (staged && cnt != 0) {
            const uint32_t s_begin = stage_addr + static_cast<uint32_t>(lane_excl) * 4u;
            emit_walk_staged8<C>(ix, tabp, stride, n_levels, chrom, qs, hi, s_begin, s_begin + cnt * 4u);
        }

        // ---- 4. look-back (warp 0): exclusive prefix of this tile over all earlier tiles
        if (warp == 0) {
            uint64_t excl = 0;
            if (tile != 0) {
                int64_t j = static_cast<int64_t>(tile) - 1;
                while (true) {
                    const int64_t t = j - static_cast<int64_t>(lane);
                    uint64_t v = kFlagIncl;  // tiles before tile 0: inclusive prefix 0
                    if (t >= 0) {
                        do {
                            v = ld_relaxed_u64(&state[t]);
                        } while ((v & kFlagMask) == 0);
                    }
                    const uint32_t incl_mask = __ballot_sync(0xffffffffu, (v & kFlagMask) == kFlagIncl);
                    const uint32_t first_incl = incl_mask ? static_cast<uint32_t>(__ffs(incl_mask) - 1) : 32u;
                    uint64_t val = (lane <= first_incl) ? (v & ~kFlagMask) : 0ull;
#pragma unroll
                    for (int d = 16; d >= 1; d >>= 1) val += __shfl_xor_sync(0xffffffffu, val, d);
                    excl += val;
                    if (incl_mask) break;
                    j -= 32;
                }
                if (lane == 0) st_relaxed_u64(&state[tile], kFlagIncl | (excl + tile_total));
            }
            if (lane == 0) s_prefix = excl;
        }
        __syncthreads();

        // ---- 5. offsets + copy-out
        const uint64_t wbase = s_prefix + warp_excl;  // chunk-local index of the warp's first hit
        const uint64_t o0 = wbase + lane_excl;
        if (in_range) offsets[q] = base + o0;
        if (q + 1 == nq) {
            const uint64_t total = o0 + cnt;
            offsets[nq] = base + total;
            if (next_base_out) *next_base_out = base + total;
            if (total_out) *total_out = total;
            if (total > capacity) atomicOr(err_flag, 2u);
        }
        if (wtotal == 0) continue;
        if (staged) {
            __syncwarp();
            const uint32_t n = static_cast<uint32_t>(wtotal);
            uint32_t sa = stage_addr + lane * 4u;
            if (wbase + wtotal <= capacity) {
                uint32_t* out = hits + wbase + lane;
                for (uint32_t i = lane; i < n; i += 32, out += 32, sa += 128) *out = lds32(sa);
            } else {
                for (uint32_t i = lane; i < n; i += 32, sa += 128)
                    if (wbase + i < capacity) hits[wbase + i] = lds32(sa);
            }
            __syncwarp();
        } else if (cnt != 0 && o0 + cnt <= capacity) {
            emit_walk_global<C>(ix, tabp, stride, n_levels, chrom, qs, hi, hits + o0, cnt);
        }
    }
    if (nq == 0 && blockIdx.x == 0 && threadIdx.x == 0) {
        offsets[0] = base;
        if (next_base_out) *next_base_out = base;
        if (total_out) *total_out = 0;
    }
}

}  // namespace gg
