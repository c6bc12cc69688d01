// emit_tma.cuh — pass 2 of the batch path with both ends on the async proxy (TMA in, TMA out).
//
// Same job as intersect_emit_scan_kernel (emit_scan.cuh): the leaf walk of detail::search_overlaps
// (include/genogrove/structure/grove/query_engine.hpp:83-112) and query_result::add_key
// (data_type/query_result.hpp:108-113) for a whole batch — offsets AND hits from the counts of pass 1, no atomics, hits in
// the reference's order.  What changed against emit_scan.cuh, and why (profiles/r1_summary.md: that kernel was
// instruction-issue bound at 75 % with 61 % of its stall samples on the first use of cnt / chrom / start / hi):
//
//   * inputs arrive by TMA: a warp's next two 32-query tiles ({cnt, hi-slot} records written by pass 1, chrom ids, query
//     starts) are in flight as cp.async.bulk copies into a three-slot shared-memory ring while it walks the current tile;
//     the DRAM round trips that used to sit at the head of every tile are gone and cost no registers (a register
//     prefetch was measured slower: occupancy).
//   * the walk runs on ABSOLUTE level-0 slots and a flat max-end hierarchy (FlatMax: one max per 32 / 1024 / 32 K / 1 M
//     slots over all chromosomes concatenated) — no per-chromosome offset table, no level arithmetic in the skip test.
//   * hits leave by TMA: the tile's stage is laid out so that shared and global addresses agree modulo 16; the aligned
//     middle goes out as ONE cp.async.bulk shared -> global issued by one lane (the per-hit LDS + STG.32 copy-out loop
//     was 13 % of the instructions), the <= 15 bytes at either end as scalar stores.
//   * output forms: per-chromosome ranks (u32) with global u64 CSR offsets — what gg_intersect_batch returns — or the
//     COMPACT form for the host path: per query {chunk-local u32 offset, hi slot} and per hit a back-delta
//     (hi_slot - 1 - slot) in 1, 2 or 4 bytes; rank = hi_slot - chrom_base - 1 - delta.  PCIe carries 15-23 B per
//     query instead of 38.
#pragma once
#include "intersect.cuh"

namespace gg {

constexpr int kFlatLevels = 4;
// Max-end hierarchy over the level-0 slots of ALL chromosomes (slot = absolute element index - base0; chromosome
// segments are padded to whole 32-slot nodes, pads hold end = 0).  A group may straddle a chromosome boundary: its max
// is then merely conservative (a walk never leaves its chromosome before its list is full, see emit_walk_flat).
template <typename C>
struct FlatMax {
    const C* leaf;             // ends in chain order, one per slot (= M level 0)
    const C* lv0;              // lvK[j] = max end over slots [j << 5(K+1), (j+1) << 5(K+1)); named members, not an array:
    const C* lv1;              // a dynamically indexed pointer array in kernel parameters would be copied to local memory
    const C* lv2;
    const C* lv3;
    const uint32_t* cbase;     // [n_chrom + 1] first slot of each chromosome
    uint32_t n_chrom;
    int n_lv;
};

// dst[j] = max(src[32 j .. 32 j + 32) within n_src); one warp per entry, one coalesced line per warp
template <typename C>
__global__ void flat_max_kernel(const C* __restrict__ src, C* __restrict__ dst, uint32_t n_src, uint32_t n_dst) {
    const uint32_t j = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (j >= n_dst) return;
    const uint32_t i = (j << 5) + lane_id();
    C m = i < n_src ? src[i] : C(0);
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
        const C o = __shfl_xor_sync(0xffffffffu, m, d);
        m = o > m ? o : m;
    }
    if (lane_id() == 0) dst[j] = m;
}

// ---- mbarrier / TMA by 32-bit shared address
__device__ __forceinline__ void mbar_init_a(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx_a(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait_a(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "LAB_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra DONE;\n"
        "bra LAB_WAIT;\n"
        "DONE:\n"
        "}\n" ::"r"(bar),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_g2s_a(uint32_t smem_dst, const void* gmem_src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_dst),
                 "l"(gmem_src), "r"(bytes), "r"(bar)
                 : "memory");
}

template <typename T>
__device__ __forceinline__ void sts_t(uint32_t a, uint32_t v) {
    if constexpr (sizeof(T) == 4) sts32(a, v);
    else if constexpr (sizeof(T) == 2) sts16(a, v);
    else sts8(a, v);
}
template <typename T>
__device__ __forceinline__ T lds_t(uint32_t a) {
    if constexpr (sizeof(T) == 4) return static_cast<T>(lds32(a));
    else if constexpr (sizeof(T) == 2) return static_cast<T>(lds16(a));
    else return static_cast<T>(lds8(a));
}

// How many slots below r_hi (a multiple of 32) hold no end >= qs, by the flat max-end hierarchy: 0, one node, or — while
// the boundary is also aligned one level up — 1024, 32 K, 1 M slots.
template <typename C>
__device__ __forceinline__ uint32_t flat_skip_span(const FlatMax<C>& fm, C qs, uint32_t r_hi) {
    uint32_t b = (r_hi >> kFanoutLog2) - 1;
    if (!(fm.lv0[b] < qs)) return 0;
    uint32_t span = kFanout;
    if ((b & (kFanout - 1)) == (kFanout - 1) && fm.n_lv > 1 && fm.lv1[b >> 5] < qs) {
        span = 1u << 10;
        if ((b & 1023u) == 1023u && fm.n_lv > 2 && fm.lv2[b >> 10] < qs) {
            span = 1u << 15;
            if ((b & 32767u) == 32767u && fm.n_lv > 3 && fm.lv3[b >> 15] < qs) span = 1u << 20;
        }
    }
    return span > r_hi ? r_hi : span;
}

// Backward walk of one query over absolute slots hi-1, hi-2, ... in aligned groups of 8 (two 16-byte loads per step).
// Output value per hit: kDelta ? hi - 1 - slot : slot - cb (the per-chromosome rank).  kShared: values go to shared
// memory back to front between s_begin and s_end (byte addresses, elements of OutT); else to global memory, dst[rem-1]
// downwards.  Exactly count(q) slots below hi pass the test (pass 1 counted them) and all of them lie in the query's
// chromosome, so the list is full before the walk could leave it; pads (end = 0) sit above hi or in other chromosomes.
// Returns the slot where the walk stopped (the group base of the lowest hit).
template <typename C, typename OutT, bool kDelta, bool kShared>
__device__ __forceinline__ uint32_t emit_walk_flat(const FlatMax<C>& fm, C qs, uint32_t hi, uint32_t cb, uint32_t s_begin,
                                                   uint32_t s_end, OutT* __restrict__ dst, uint32_t n_hits) {
    constexpr uint32_t kSz = sizeof(OutT);
    const C* __restrict__ leaf = fm.leaf;
    uint32_t r_hi = hi, sp = s_end, rem = n_hits;
    C ea[8];
    auto put = [&](uint32_t v) {
        if constexpr (kShared) {
            sp -= kSz;
            sts_t<OutT>(sp, v);
        } else {
            if (rem != 0) dst[--rem] = static_cast<OutT>(v);
        }
    };
    auto more = [&]() { return kShared ? sp > s_begin : rem != 0; };
    if ((r_hi & 7u) != 0) {  // partial top group: slots g8 .. hi-1
        const uint32_t g8 = r_hi & ~7u, top = r_hi & 7u;
        const uint32_t vb = kDelta ? hi - 1 - g8 : g8 - cb;
        load8(leaf + g8, ea);
#pragma unroll
        for (int k = 6; k >= 1; --k) {
            if (top > static_cast<uint32_t>(k) && ea[k] >= qs) put(kDelta ? vb - k : vb + k);
        }
        if (ea[0] >= qs) put(vb);
        r_hi = g8;
    }
    while (more() && r_hi != 0) {
        if ((r_hi & (kFanout - 1)) == 0) {
            const uint32_t span = flat_skip_span<C>(fm, qs, r_hi);
            if (span) {
                r_hi -= span;
                continue;
            }
        }
        const uint32_t g8 = r_hi - 8;
        const uint32_t vb = kDelta ? hi - 1 - g8 : g8 - cb;
        load8(leaf + g8, ea);
#pragma unroll
        for (int k = 7; k >= 0; --k) {
            if (ea[k] >= qs) put(kDelta ? vb - k : vb + k);
        }
        r_hi = g8;
    }
    return r_hi;
}

// per query of the compact form: where its hit list starts in the chunk's delta array, and the slot all deltas count back from
struct CompactQuery {
    uint32_t off;      // chunk-local index of the first hit (the list ends where the next query's begins)
    uint32_t hi_slot;  // first level-0 slot starting past q.end: rank j of the list = hi_slot - chrom_base - 1 - delta[off + j]
};

constexpr uint32_t kTmaStageBytes = 2048;  // hits of one 32-query tile staged per warp (512 ranks / 1024 u16 / 2048 u8 deltas)
constexpr uint32_t kTmaStageSlack = 32;    // in front of the stage
// Ring slots per warp.  The prefetch distance is two tiles; a slot is refilled ONE WHOLE TILE after it was read, not right
// after: with two slots the refill was issued straight behind the reads of the same slot, and nothing orders a generic-proxy
// read that is still in flight (__syncwarp() compiles to nothing on a converged warp) against the async-proxy write of the
// TMA copy — measured on B200: a handful of stale counts per 10 M queries under load.  With three slots every value read
// from a slot has been consumed (warp scan, walk) a tile ago when the refill is issued.
constexpr uint32_t kTmaSlots = 3;
// bytes of one ring slot: 32 {cnt, hi-slot} records + 32 chrom ids + 32 query starts
template <typename Q>
__host__ __device__ constexpr uint32_t tma_slot_bytes() { return 256u + 128u + 32u * static_cast<uint32_t>(sizeof(Q)); }
template <typename Q>
__host__ __device__ constexpr uint32_t tma_warp_bytes() {
    return kTmaStageSlack + kTmaStageBytes + 16u + 80u /* pad to 128 */ + kTmaSlots * tma_slot_bytes<Q>();
}
template <typename Q>
inline size_t emit_tma_smem_bytes(uint32_t cbase_entries_in_smem) {
    return static_cast<size_t>(kQueryThreads / 32) * (tma_warp_bytes<Q>() + 32) + cbase_entries_in_smem * 4u + 16;
}

// err_flag bits: 1 invalid query (pass 1), 2 hit buffer too small, 8 chunk has >= 2^32 hits (compact offsets), 16 a
// back-delta does not fit OutT (re-run the chunk with a wider delta)
template <typename C, typename Q, typename OutT, bool kDelta, bool kBulk>
__global__ void __launch_bounds__(kQueryThreads)
intersect_emit_tma_kernel(FlatMax<C> fm, const uint2* __restrict__ rec, const uint32_t* __restrict__ qchrom,
                          const Q* __restrict__ qstart, size_t nq, const uint64_t* __restrict__ block_prefix, uint32_t blk_q,
                          uint64_t* __restrict__ offsets, CompactQuery* __restrict__ qres, OutT* __restrict__ hits,
                          uint64_t capacity, uint64_t* __restrict__ next_base_out, uint64_t* __restrict__ total_out,
                          uint32_t* __restrict__ err_flag, int cbase_in_smem) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr uint32_t kWarps = kQueryThreads / 32;
    constexpr uint32_t kSz = sizeof(OutT);
    constexpr uint32_t kSlot = tma_slot_bytes<Q>();
    constexpr uint32_t kWarpBytes = tma_warp_bytes<Q>();
    constexpr uint32_t kStageElems = kTmaStageBytes / kSz;
    static_assert(kWarpBytes % 128 == 0, "ring slots must stay 128-byte aligned");
    const uint32_t lane = lane_id(), warp = threadIdx.x >> 5;
    const uint32_t wsm = smem_u32(smem_raw) + warp * kWarpBytes;
    const uint32_t stage_addr = wsm + kTmaStageSlack;  // 16-byte aligned (32 mod 128)
    const uint32_t ring_addr = wsm + kTmaStageSlack + kTmaStageBytes + 16u + 80u;
    const uint32_t bar_addr = smem_u32(smem_raw) + kWarps * kWarpBytes + warp * 32u;  // kTmaSlots mbarriers per warp
    const uint32_t* __restrict__ cbase = fm.cbase;
    if (!kDelta && cbase_in_smem) {
        uint32_t* sc = reinterpret_cast<uint32_t*>(smem_raw + kWarps * (kWarpBytes + 32u));
        for (uint32_t i = threadIdx.x; i <= fm.n_chrom; i += blockDim.x) sc[i] = fm.cbase[i];
        cbase = sc;
        __syncthreads();
    }
    const uint64_t base = block_prefix[0];
    if (nq == 0) {
        if (blockIdx.x == 0 && threadIdx.x == 0) {
            if (offsets) offsets[0] = base;
            if (next_base_out) *next_base_out = base;
            if (total_out) *total_out = 0;
        }
        return;
    }
    if (lane == 0) {
#pragma unroll
        for (uint32_t i = 0; i < kTmaSlots; ++i) mbar_init_a(bar_addr + i * 8u, 1);
        mbar_fence_init();
    }
    __syncwarp();

    // (tile indices fit 32 bits: the host sends batches of 2^32 queries or more to intersect_emit_scan_kernel)
    const uint32_t n_tiles = static_cast<uint32_t>((nq + 31) / 32);
    const uint32_t tpb = blk_q / 32;  // tiles per block (one block = one warp's share of pass 1), a power of two
    const uint32_t stride_tiles = gridDim.x * kWarps * tpb;
    // the warp's tile sequence: the tiles of blocks w, w + warps_total, ...
    auto next_tile = [&](uint32_t t) -> uint32_t {
        const uint32_t t1 = t + 1;
        if ((t1 & (tpb - 1)) != 0) return t1 < n_tiles ? t1 : n_tiles;
        const uint32_t nb = t1 - tpb + stride_tiles;  // first tile of the warp's next block
        return (nb < n_tiles && nb >= t1) ? nb : n_tiles;
    };
    // a tile's inputs -> ring slot s (full tiles only: TMA sizes are multiples of 16 bytes; the batch's last, partial
    // tile is read with plain loads)
    auto issue = [&](uint32_t t, uint32_t s) {
        if (lane == 0 && t < n_tiles && (static_cast<size_t>(t) + 1) * 32 <= nq) {
            const uint32_t dst = ring_addr + s * kSlot, bar = bar_addr + s * 8u;
            mbar_expect_tx_a(bar, kDelta ? kSlot - 128u : kSlot);
            const size_t q0 = static_cast<size_t>(t) * 32;
            tma_g2s_a(dst, rec + q0, 256u, bar);
            if (!kDelta) tma_g2s_a(dst + 256u, qchrom + q0, 128u, bar);
            tma_g2s_a(dst + 384u, qstart + q0, 32u * static_cast<uint32_t>(sizeof(Q)), bar);
        }
    };

    if (static_cast<size_t>(blockIdx.x * kWarps + warp) * tpb >= n_tiles) return;
    uint32_t t_cur = (blockIdx.x * kWarps + warp) * tpb;
    uint32_t t_nxt = next_tile(t_cur);
    issue(t_cur, 0);
    issue(t_nxt, 1);
    uint32_t slot = 0, phases = 0;
    uint64_t run = 0;
    bool stage_busy = false;  // a bulk store of this warp may still be reading the stage
    for (; t_cur < n_tiles; t_cur = t_nxt, t_nxt = next_tile(t_nxt), slot = slot == kTmaSlots - 1 ? 0u : slot + 1u) {
        if ((t_cur & (tpb - 1)) == 0) run = block_prefix[t_cur / tpb];
        const size_t q = static_cast<size_t>(t_cur) * 32 + lane;
        const bool in = q < nq;
        uint32_t c = 0, hi = 0, chrom = 0;
        Q qs_raw = 0;
        if (q - lane + 32 <= nq) {
            mbar_wait_a(bar_addr + slot * 8u, (phases >> slot) & 1u);
            phases ^= 1u << slot;
            const uint32_t sl = ring_addr + slot * kSlot;
            const uint2 r = lds64(sl + lane * 8u);
            c = r.x;
            hi = r.y;
            if (!kDelta) chrom = lds32(sl + 256u + lane * 4u);
            if constexpr (sizeof(Q) == 4) {
                qs_raw = lds32(sl + 384u + lane * 4u);
            } else {
                const uint2 v = lds64(sl + 384u + lane * 8u);
                qs_raw = (static_cast<uint64_t>(v.y) << 32) | v.x;
            }
            // the tile after next goes into the slot read one tile AGO (see kTmaSlots)
            issue(next_tile(t_nxt), slot >= 1 ? slot - 1u : kTmaSlots - 1u);
        } else if (in) {
            const uint2 r = rec[q];
            c = r.x;
            hi = r.y;
            if (!kDelta) chrom = qchrom[q];
            qs_raw = qstart[q];
        }
        // exclusive warp scan of the 32 counts: 32-bit when the sum cannot overflow, else 64-bit
        uint64_t excl, wtotal;
        if (__reduce_max_sync(0xffffffffu, c) < (1u << 27)) {
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
        if (in) {
            if constexpr (kDelta) {
                const uint64_t loc = o0 - base;
                if (loc > 0xFFFFFFFFull) atomicOr(err_flag, 8u);
                CompactQuery cq;
                cq.off = static_cast<uint32_t>(loc);
                cq.hi_slot = hi;
                __stcs(reinterpret_cast<uint2*>(qres + q), make_uint2(cq.off, cq.hi_slot));
            } else {
                __stcs(offsets + q, o0);
            }
        }
        if (q + 1 == nq) {
            const uint64_t end_off = o0 + c;  // = base + total
            if (offsets) offsets[nq] = end_off;
            if (next_base_out) *next_base_out = end_off;
            if (total_out) *total_out = end_off - base;
            if (end_off - base > capacity) atomicOr(err_flag, 2u);
            if (kDelta && end_off - base > 0xFFFFFFFFull) atomicOr(err_flag, 8u);
        }
        const uint64_t wbase = run - base;  // chunk-local index of the tile's first hit
        run += wtotal;
        if (wtotal == 0) continue;
        const C qs = static_cast<C>(qs_raw);  // c > 0 implies q.start fits the index domain
        const uint32_t cb = (!kDelta && c != 0) ? cbase[chrom] : 0u;
        if (wtotal <= kStageElems) {
            // stage layout: element j of the tile at byte a + j * kSz, a = the global destination's address mod 16
            const uintptr_t gaddr = reinterpret_cast<uintptr_t>(hits + wbase);
            const uint32_t a = static_cast<uint32_t>(gaddr) & 15u;
            const uint32_t n_bytes = static_cast<uint32_t>(wtotal) * kSz;
            if (kBulk && stage_busy) {  // the previous tile's bulk store must have read the stage before it is refilled
                if (lane == 0) tma_store_wait_read();
                __syncwarp();
                stage_busy = false;
            }
            uint32_t r_end = hi;
            if (c != 0) {
                const uint32_t s_begin = stage_addr + a + static_cast<uint32_t>(excl) * kSz;
                r_end = emit_walk_flat<C, OutT, kDelta, true>(fm, qs, hi, cb, s_begin, s_begin + c * kSz, nullptr, 0);
            }
            if (kDelta && kSz < 4 && c != 0 && hi - 1 - r_end > (kSz == 1 ? 0xFFu : 0xFFFFu)) atomicOr(err_flag, 16u);
            if (wbase + wtotal <= capacity) {
                const uint32_t head = min(n_bytes, (16u - a) & 15u);      // bytes up to the first 16-byte boundary
                const uint32_t mid = (n_bytes - head) & ~15u;             // whole 16-byte vectors
                const uint32_t tail = n_bytes - head - mid;
                unsigned char* g = reinterpret_cast<unsigned char*>(hits + wbase);
                if (kBulk) fence_proxy_async_smem();  // this lane's stage writes -> visible to the async proxy
                __syncwarp();
                if (kBulk) {
                    if (lane == 0 && mid != 0) {
                        tma_bulk_s2g(g + head, stage_addr + a + head, mid);
                        tma_store_commit();
                    }
                    stage_busy = mid != 0;
                } else {
                    for (uint32_t v = lane * 16u; v < mid; v += 512u)
                        __stcs(reinterpret_cast<uint4*>(g + head + v), lds128(stage_addr + a + head + v));
                }
                // the ends: lanes 0..15 the head elements, lanes 16..31 the tail elements
                const uint32_t e = (lane & 15u) * kSz;
                if (lane < 16) {
                    if (e < head) __stcs(reinterpret_cast<OutT*>(g + e), lds_t<OutT>(stage_addr + a + e));
                } else {
                    if (e < tail) __stcs(reinterpret_cast<OutT*>(g + head + mid + e), lds_t<OutT>(stage_addr + a + head + mid + e));
                }
            } else {
                __syncwarp();
                for (uint32_t i = lane; i < static_cast<uint32_t>(wtotal); i += 32)
                    if (wbase + i < capacity) hits[wbase + i] = lds_t<OutT>(stage_addr + a + i * kSz);
            }
            if (!kBulk) __syncwarp();
        } else if (c != 0 && wbase + excl + c <= capacity) {  // too large for the stage: straight to global memory
            const uint32_t r_end = emit_walk_flat<C, OutT, kDelta, false>(fm, qs, hi, cb, 0, 0, hits + wbase + excl, c);
            if (kDelta && kSz < 4 && hi - 1 - r_end > (kSz == 1 ? 0xFFu : 0xFFFFu)) atomicOr(err_flag, 16u);
        }
    }
    if (kBulk && stage_busy && lane == 0) tma_store_wait_read();
}

}  // namespace gg
