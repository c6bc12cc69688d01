// strand.cuh — strand-aware batched intersect for genomic_coordinate keys.
//
// Replaces grove<genomic_coordinate, T>::intersect(query, index) -> detail::search_overlaps
// (include/genogrove/structure/grove/query_engine.hpp:76-134) with genomic_coordinate::overlaps
// (include/genogrove/data_type/genomic_coordinate.hpp:156-160): coordinates overlap (closed) AND the strands are equal
// or either one is the wildcard '*'.  The descent skips a child only when it lies before the query or when its
// aggregate strand (genomic_coordinate.hpp:178-182: common strand, '*' when mixed) cannot match, and the leaf walk
// runs to the first leaf starting past q.end, so the hit list is exactly
//     { r : start[r] <= q.end  &&  end[r] >= q.start  &&  compatible(strand[r], q.strand) }   ascending in chain rank,
// the chain being ordered by (start, end, strand with '*' < '.' < '+' < '-') (genomic_coordinate.hpp:97-110).
//
// Same flattened index as the interval path plus, per level-0 node (32 slots), a 32-byte strand record in chain
// order (SN) and one in sorted-end order (EN): the number of keys of each strand class ('*', '.', '+', '-') before
// the node and a 32-bit membership mask per class.  With them the strand-specific count is a closed form, like the
// interval count — no walk in pass 1:
//   #{r : start[r] <= q.end, end[r] >= q.start, class(r) in K} = rank_S(K, hi) - rank_E(K, lo),   K = {q.strand, '*'}
//   (hi = ub(S, q.end), lo = ub(E, q.start - 1): every key with end < q.start also has start <= q.end)
//   rank_X(K, i) = sum over k in K of pre_X[k][i >> 5] + popc(bits_X[k][i >> 5] & ((1 << (i & 31)) - 1))
// Three kernels per batch, no atomics, reference order — the shape of the interval batch path:
//   strand_count_group_kernel (4 consecutive queries per thread, finger search) -> scan of the per-128-query block
//   totals -> intersect_emit_scan_kernel<.., kStrand = true> (offsets + hits; the walk tests end >= q.start AND the
//   strand byte of each rank, 8 ranks per step: 32 bytes of ends + 8 strand bytes, SIMD byte compare)
#pragma once
#include "intersect.cuh"

namespace gg {

__device__ __forceinline__ bool strand_valid(uint32_t s) { return s == '+' || s == '-' || s == '.' || s == '*'; }
__device__ __forceinline__ bool strand_compatible(uint32_t key_strand, uint32_t q_strand) {
    return key_strand == q_strand || key_strand == '*' || q_strand == '*';
}

// strands of one chromosome in the padded level-0 layout (slot = tab[0][chrom] - tab[0][0] + rank, pads 0):
// 32-slot aligned segments, so 8 ranks = one aligned 8-byte load next to the 32-byte load of their ends
__global__ void pad_strand_kernel(const uint8_t* __restrict__ chain_strand, const uint32_t* __restrict__ cbegin,
                                  const uint32_t* __restrict__ cnt, const uint32_t* __restrict__ row0, uint32_t n_chrom,
                                  uint8_t* __restrict__ out) {
    const uint32_t p = row0[0] + blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= row0[n_chrom]) return;
    uint32_t lo = 0, hi = n_chrom;
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (row0[mid] <= p) lo = mid; else hi = mid;
    }
    const uint32_t c = lo, r = p - row0[c];
    out[p - row0[0]] = r < cnt[c] ? chain_strand[cbegin[c] + r] : uint8_t(0);
}

// per-byte compatibility of 4 key strands with the query strand: 0xFF where equal or the key is '*'
__device__ __forceinline__ uint32_t strand_compat4(uint32_t w, uint32_t q4) {
    return __vcmpeq4(w, q4) | __vcmpeq4(w, 0x2A2A2A2Au);  // '*' = 0x2A
}

// Walk ranks hi-1, hi-2, ... of one chromosome in aligned groups of 8 (one 32-byte load of ends + one 8-byte load
// of strands per group; the max-end levels skip whole nodes / subtrees at node boundaries).
// COUNT: n_stop = number of SPATIAL hits below hi (exact), returns how many of them are strand-compatible.
// EMIT : n_stop = number of COMPATIBLE hits (exact), stored at dst[n_stop-1], dst[n_stop-2], ...
// Either way the walk may finish its last group after the stop count is reached: no further hit exists below.
template <typename C, bool kEmit>
__device__ __forceinline__ uint32_t strand_walk(const IndexView<C>& ix, uint32_t stride, int n_levels, uint32_t chrom, C qs,
                                                uint32_t hi, uint32_t n_stop, const uint8_t* __restrict__ strands,
                                                uint32_t q_strand, uint32_t* __restrict__ dst) {
    const C* __restrict__ leaf_end = ix.M + ix.tab[chrom];
    const uint32_t q4 = q_strand * 0x01010101u;
    const bool wild = q_strand == '*';
    uint32_t r_hi = hi, rem = n_stop, compat = 0;
    C ea[8];
    while (rem != 0 && r_hi != 0) {
        if ((r_hi & (kFanout - 1)) == 0 && n_levels > 1) {
            const uint32_t span = emit_skip_span<C>(ix, ix.tab, stride, n_levels, chrom, qs, r_hi);
            if (span) {
                r_hi -= span;
                continue;
            }
        }
        const uint32_t g8 = (r_hi - 1) & ~7u, top = r_hi - g8;  // ranks g8 .. g8 + top - 1 (top = 8 except in the first group)
        load8(leaf_end + g8, ea);
        const uint2 sw = *reinterpret_cast<const uint2*>(strands + g8);
        const uint32_t m_lo = wild ? 0xFFFFFFFFu : strand_compat4(sw.x, q4), m_hi = wild ? 0xFFFFFFFFu : strand_compat4(sw.y, q4);
#pragma unroll
        for (int k = 7; k >= 0; --k) {
            const bool spatial = static_cast<uint32_t>(k) < top && ea[k] >= qs;
            const bool ok = spatial && (((k < 4 ? m_lo : m_hi) >> (8 * (k & 3))) & 1u);
            if (kEmit) {
                if (ok && rem != 0) dst[--rem] = g8 + k;
            } else {
                rem -= (spatial && rem != 0) ? 1u : 0u;
                compat += ok ? 1u : 0u;
            }
        }
        r_hi = g8;
    }
    return compat;
}

// ---- per-node strand records ---------------------------------------------------------------------------------
struct StrandNode {
    uint32_t pre[4];   // keys of class k ('*', '.', '+', '-') in all level-0 slots before this node (global, all chromosomes)
    uint32_t bits[4];  // bit j: slot 32 * node + j holds a key of class k
};
static_assert(sizeof(StrandNode) == 32, "one 32-byte sector per node");

__device__ __forceinline__ uint32_t strand_class(uint32_t s) { return s == '*' ? 0u : s == '.' ? 1u : s == '+' ? 2u : 3u; }

// One warp per level-0 node: class masks of the node in chain order (from the padded strand bytes) and in
// sorted-end order (strand of the interval whose end sits in each slot, through ER); per-class key counts go to
// cnt_S / cnt_E [class][node] for the prefix scans.  Node n_nodes is the sentinel behind the last chromosome.
__global__ void strand_node_kernel(const uint8_t* __restrict__ strandP, const uint32_t* __restrict__ ER,
                                   const uint32_t* __restrict__ row0, uint32_t n_chrom, uint32_t n_nodes,
                                   StrandNode* __restrict__ SN, StrandNode* __restrict__ EN, uint32_t* __restrict__ cnt_S,
                                   uint32_t* __restrict__ cnt_E) {
    const uint32_t node = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (node > n_nodes) return;
    const uint32_t lane = lane_id();
    uint32_t bs = 0, be = 0;  // strand byte of this lane's slot in either order (0 = pad)
    if (node < n_nodes) {
        const uint32_t slot = (node << kFanoutLog2) + lane;
        bs = strandP[slot];
        const uint32_t rank = ER[slot];
        if (rank != 0xFFFFFFFFu) {
            uint32_t lo = 0, hi = n_chrom;  // chromosome owning this node
            const uint32_t p = row0[0] + slot;
            while (hi - lo > 1) {
                const uint32_t mid = (lo + hi) >> 1;
                if (row0[mid] <= p) lo = mid; else hi = mid;
            }
            be = strandP[row0[lo] - row0[0] + rank];
        }
    }
#pragma unroll
    for (uint32_t k = 0; k < 4; ++k) {
        const uint32_t sym = k == 0 ? '*' : k == 1 ? '.' : k == 2 ? '+' : '-';
        const uint32_t ms = __ballot_sync(0xffffffffu, bs == sym), me = __ballot_sync(0xffffffffu, be == sym);
        if (lane == 0) {
            SN[node].bits[k] = ms;
            EN[node].bits[k] = me;
            cnt_S[k * (n_nodes + 1) + node] = __popc(ms);
            cnt_E[k * (n_nodes + 1) + node] = __popc(me);
        }
    }
}

// scanned per-class counts [class][node] -> the records' pre[] fields
__global__ void strand_pack_kernel(const uint32_t* __restrict__ pre_S, const uint32_t* __restrict__ pre_E, uint32_t n_nodes,
                                   StrandNode* __restrict__ SN, StrandNode* __restrict__ EN) {
    const uint32_t node = blockIdx.x * blockDim.x + threadIdx.x;
    if (node > n_nodes) return;
#pragma unroll
    for (uint32_t k = 0; k < 4; ++k) {
        SN[node].pre[k] = pre_S[k * (n_nodes + 1) + node];
        EN[node].pre[k] = pre_E[k * (n_nodes + 1) + node];
    }
}

// keys of class `cls` or '*' among the first i slots of the chromosome whose first node is node0 (global count)
__device__ __forceinline__ uint32_t strand_rank(const StrandNode* __restrict__ N, uint32_t node0, uint32_t i, uint32_t cls) {
    const uint4* rec = reinterpret_cast<const uint4*>(N + node0 + (i >> kFanoutLog2));
    const uint4 pre = rec[0], bits = rec[1];
    const uint32_t below = (1u << (i & 31u)) - 1u;
    const uint32_t pc = cls == 1 ? pre.y : cls == 2 ? pre.z : pre.w, bc = cls == 1 ? bits.y : cls == 2 ? bits.z : bits.w;
    return pre.x + pc + __popc((bits.x | bc) & below);
}

// Pass 1: QPT consecutive queries per thread — the interval count pass (count_group: directory + finger search)
// followed by the closed-form strand count.  Outputs as intersect_count_group_kernel: cnt / hi per query and
// block_total[b] = hits of the b-th block of 32 * QPT queries.  err_flag bit 0: start > end; bit 2: a strand byte
// outside "+-.*" (genomic_coordinate's constructor check, genomic_coordinate.hpp:78-81).
template <typename C, typename Q, int QPT, bool kTiles = false>
__global__ void __launch_bounds__(kQueryThreads)
strand_count_group_kernel(IndexView<C> ix, const StrandNode* __restrict__ SN, const StrandNode* __restrict__ EN,
                          uint32_t slot_base, const uint32_t* __restrict__ qchrom, const Q* __restrict__ qstart,
                          const Q* __restrict__ qend, const uint8_t* __restrict__ qstrand, size_t nq,
                          uint32_t* __restrict__ cnt_out, uint32_t* __restrict__ hi_out,
                          unsigned long long* __restrict__ block_total, uint32_t* __restrict__ err_flag, int vec_ok,
                             uint32_t* __restrict__ tile_total32 = nullptr) {
    constexpr size_t kTileQ = static_cast<size_t>(kQueryThreads) * QPT;
    const size_t n_tiles = (nq + kTileQ - 1) / kTileQ;
    for (size_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const size_t q0 = tile * kTileQ + static_cast<size_t>(threadIdx.x) * QPT;
        const bool vec = vec_ok && q0 + QPT <= nq;
        uint32_t chrom[QPT], qst[QPT];
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
        bool bad_strand = false;
#pragma unroll
        for (int j = 0; j < QPT; ++j) {
            qst[j] = q0 + j < nq ? qstrand[q0 + j] : uint32_t('*');
            if (!strand_valid(qst[j])) {
                bad_strand = true;
                chrom[j] = GG_NO_CHROM_DEV;  // no hits for this query; the call fails with the flag below
            }
        }
        uint32_t cnt[QPT], hi[QPT], slot[QPT];
        C qs[QPT];
        uint32_t flags = count_group<C, Q, QPT>(ix, ix.tab, chrom, qs_raw, qe_raw, cnt, hi, qs, slot) ? 1u : 0u;
        if (bad_strand) flags |= 4u;
        if (flags) atomicOr(err_flag, flags);
#pragma unroll
        for (int j = 0; j < QPT; ++j) {
            if (cnt[j] != 0 && qst[j] != '*') {  // the wildcard query matches every strand: cnt = spatial count
                const uint32_t node0 = (ix.tab[chrom[j]] - slot_base) >> kFanoutLog2, cls = strand_class(qst[j]);
                cnt[j] = strand_rank(SN, node0, hi[j], cls) - strand_rank(EN, node0, hi[j] - cnt[j], cls);
            }
        }
        if (vec) {
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
            if (wq0 < nq) block_total[wq0 / (32 * QPT)] = wsum;
        }
    }
}

}  // namespace gg

namespace gg {

// bit k set: rank g8 + k of a walk group overlaps the query (end >= q.start) and, with strand masks (0xFF per
// compatible byte, strand_compat4), has a compatible strand; ranks at or above `top` are masked off
template <typename C>
__device__ __forceinline__ uint32_t group_hit_mask(const C (&ea)[8], C qs, uint32_t top, uint32_t m_lo, uint32_t m_hi) {
    uint32_t m = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) m |= (ea[k] >= qs ? 1u : 0u) << k;
    // one bit per byte of the compat masks -> 4 + 4 bits (multiply gathers bits 0, 8, 16, 24 into the top nibble)
    const uint32_t c_lo = ((m_lo & 0x01010101u) * 0x10204080u) >> 28, c_hi = ((m_hi & 0x01010101u) * 0x10204080u) >> 28;
    return m & (c_lo | (c_hi << 4)) & ((1u << top) - 1u);
}

// The staged backward walk of emit_scan (emit_walk_staged8 in intersect.cuh) with the strand test: ranks hi-1,
// hi-2, ... in aligned groups of 8; a hit must have end >= q.start AND a compatible strand byte.  Hits go to the
// shared-memory stage back to front until the query's list [s_begin, s_end) is full (its length is the exact
// number of compatible hits, so nothing below the last one written can still match).
template <typename C>
__device__ __forceinline__ void emit_walk_staged8_strand(const IndexView<C>& ix, const uint32_t* __restrict__ tabp, uint32_t stride,
                                                         int n_levels, uint32_t chrom, C qs, uint32_t hi,
                                                         const uint8_t* __restrict__ strands, uint32_t q_strand,
                                                         uint32_t s_begin, uint32_t s_end) {
    const C* __restrict__ leaf_end = ix.M + tabp[chrom];
    const uint32_t q4 = q_strand * 0x01010101u;
    const bool wild = q_strand == '*';
    const bool has_levels = n_levels > 1;
    uint32_t r_hi = hi, sp = s_end;
    C ea[8];
    while (sp > s_begin && r_hi != 0) {
        if ((r_hi & (kFanout - 1)) == 0 && has_levels) {
            const uint32_t span = emit_skip_span<C>(ix, tabp, stride, n_levels, chrom, qs, r_hi);
            if (span) {
                r_hi -= span;
                continue;
            }
        }
        const uint32_t g8 = (r_hi - 1) & ~7u, top = r_hi - g8;
        load8(leaf_end + g8, ea);
        const uint2 sw = *reinterpret_cast<const uint2*>(strands + g8);
        const uint32_t m_lo = wild ? 0xFFFFFFFFu : strand_compat4(sw.x, q4), m_hi = wild ? 0xFFFFFFFFu : strand_compat4(sw.y, q4);
        uint32_t m = group_hit_mask<C>(ea, qs, top, m_lo, m_hi);
        while (m != 0 && sp > s_begin) {  // hits of the group, highest rank first
            const uint32_t k = 31u - __clz(m);
            m ^= 1u << k;
            sp -= 4;
            sts32(sp, g8 + k);
        }
        r_hi = g8;
    }
}

}  // namespace gg
