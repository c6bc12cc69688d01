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
// Same flattened index as the interval path plus one byte per level-0 slot (the strand, padded like the leaf level).  Two passes, no atomics:
//   pass 1  hi = ub(S, q.end), n_sp = hi - ub(E, q.start - 1) spatial hits (directory + bracketed search, as the
//           interval count pass); walk back from hi - 1 until n_sp spatial hits were seen (max-end levels skip
//           subtrees), count the strand-compatible ones -> cnt[q]; hi[q] saved
//   scan    offsets = exclusive scan of cnt
//   pass 2  the same walk, writing compatible ranks back to front at offsets[q]
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

// Pass 1: one query per thread.  err_flag bit 0: start > end; bit 2: strand byte not one of + - . *
template <typename C, typename Q>
__global__ void __launch_bounds__(kQueryThreads)
strand_count_kernel(IndexView<C> ix, const uint8_t* __restrict__ strands, uint32_t slot_base,
                    const uint32_t* __restrict__ qchrom, const Q* __restrict__ qstart, const Q* __restrict__ qend,
                    const uint8_t* __restrict__ qstrand, size_t nq, uint32_t* __restrict__ cnt_out,
                    uint32_t* __restrict__ hi_out, uint32_t* __restrict__ err_flag) {
    const uint32_t stride = ix.n_chrom + 1;
    const size_t n_tiles = (nq + kQueryThreads - 1) / kQueryThreads;
    for (size_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const size_t q = tile * kQueryThreads + threadIdx.x;
        if (q >= nq) continue;
        const uint32_t chrom = qchrom[q];
        const Q qs_raw = qstart[q], qe_raw = qend[q];
        const uint32_t qst = qstrand[q];
        const bool bad = qs_raw > qe_raw, bad_strand = !strand_valid(qst);
        if (bad) atomicOr(err_flag, 1u);
        if (bad_strand) atomicOr(err_flag, 4u);
        const bool ok = !bad && !bad_strand && chrom < ix.n_chrom &&
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
            const uint32_t n_sp = hi - lo;
            if (n_sp != 0)
                cnt = qst == '*' ? n_sp  // the wildcard query matches every strand (genomic_coordinate.hpp:158)
                                 : strand_walk<C, false>(ix, stride, ix.n_levels, chrom, qs, hi, n_sp, strands + (leaf - slot_base), qst, nullptr);
        }
        cnt_out[q] = cnt;
        hi_out[q] = hi;
    }
}

// Pass 2: hits[offsets[q] - hits_base ...) = compatible chain ranks, ascending.  Ranks at or beyond `capacity` are
// not written (the caller sized the buffer from offsets[nq]).
template <typename C, typename Q>
__global__ void __launch_bounds__(kQueryThreads)
strand_emit_kernel(IndexView<C> ix, const uint8_t* __restrict__ strands, uint32_t slot_base,
                   const uint32_t* __restrict__ qchrom, const Q* __restrict__ qstart, const uint8_t* __restrict__ qstrand,
                   size_t nq, const uint32_t* __restrict__ hi_in, const uint64_t* __restrict__ offsets,
                   uint32_t* __restrict__ hits, uint64_t hits_base, uint64_t capacity) {
    const uint32_t stride = ix.n_chrom + 1;
    const size_t n_tiles = (nq + kQueryThreads - 1) / kQueryThreads;
    for (size_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const size_t q = tile * kQueryThreads + threadIdx.x;
        if (q >= nq) continue;
        const uint64_t o0 = offsets[q], o1 = offsets[q + 1];
        const uint32_t n = static_cast<uint32_t>(o1 - o0);
        if (n == 0 || o1 - hits_base > capacity) continue;
        const uint32_t chrom = qchrom[q];
        strand_walk<C, true>(ix, stride, ix.n_levels, chrom, static_cast<C>(qstart[q]), hi_in[q], n,
                             strands + (ix.tab[chrom] - slot_base), qstrand[q], hits + (o0 - hits_base));
    }
}

}  // namespace gg
