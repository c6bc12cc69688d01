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

// kStrand (genomic_coordinate keys, strand.cuh): cnt_in holds the strand-compatible counts and the walk also tests
// the strand byte of every rank against qstrand[q]; strands = padded level-0 strand bytes, slot_base = tab[0][0].
struct StrandArgs {
    const uint8_t* strands;
    const uint8_t* qstrand;
    uint32_t slot_base;
};

template <typename C, typename Q, int kWalk, bool kStrand = false>
__global__ void __launch_bounds__(kQueryThreads)
intersect_emit_scan_kernel(IndexView<C> ix, const uint32_t* __restrict__ qchrom, const Q* __restrict__ qstart, size_t nq,
                           const uint32_t* __restrict__ cnt_in, const uint32_t* __restrict__ hi_in,
                           const uint64_t* __restrict__ block_prefix, uint32_t blk_q, uint64_t* __restrict__ offsets,
                           uint32_t* __restrict__ hits, uint64_t capacity, uint64_t* __restrict__ next_base_out,
                           uint64_t* __restrict__ total_out, uint32_t* __restrict__ err_flag, int tab_in_smem,
                           StrandArgs sa = StrandArgs{nullptr, nullptr, 0}) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr uint32_t kStageWords = kEmitStage + kEmitScanSlack;
    constexpr uint32_t kWarps = kQueryThreads / 32;
    uint32_t* stab = reinterpret_cast<uint32_t*>(smem_raw) + kWarps * kStageWords;
    const uint32_t lane = lane_id(), warp = threadIdx.x >> 5;
    const uint32_t stage_addr = smem_u32(smem_raw) + (warp * kStageWords + kEmitScanSlack) * 4u;
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
            if (wtotal <= kEmitStage) {
                if (c != 0) {
                    const uint32_t s_begin = stage_addr + static_cast<uint32_t>(excl) * 4u;
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
                uint32_t sa = stage_addr + lane * 4u;
                if (wbase + wtotal <= capacity) {
                    uint32_t* out = hits + wbase + lane;
                    for (uint32_t i = lane; i < n; i += 32, out += 32, sa += 128) __stcs(out, lds32(sa));
                } else {
                    for (uint32_t i = lane; i < n; i += 32, sa += 128)
                        if (wbase + i < capacity) hits[wbase + i] = lds32(sa);
                }
                __syncwarp();
            } else if (c != 0 && wbase + excl + c <= capacity) {
                if (kStrand)
                    strand_walk<C, true>(ix, stride, n_levels, chrom, qs, hi, c, sa.strands + (tabp[chrom] - sa.slot_base), qst,
                                         hits + wbase + excl);
                else
                    emit_walk_global<C>(ix, tabp, stride, n_levels, chrom, qs, hi, hits + wbase + excl, c);
            }
        }
    }
}

}  // namespace gg
