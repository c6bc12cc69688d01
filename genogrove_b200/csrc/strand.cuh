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
// Same flattened index as the interval path plus one byte per chain rank (the strand).  Two passes, no atomics:
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

// Walk ranks hi-1, hi-2, ... of one chromosome.  COUNT: stop after n_stop SPATIAL hits, return the number of
// strand-compatible ones.  EMIT: stop after n_stop COMPATIBLE hits, storing them at dst[n_stop-1], dst[n_stop-2], ...
template <typename C, bool kEmit>
__device__ __forceinline__ uint32_t strand_walk(const IndexView<C>& ix, uint32_t stride, int n_levels, uint32_t chrom, C qs,
                                                uint32_t hi, uint32_t n_stop, const uint8_t* __restrict__ strands,
                                                uint32_t q_strand, uint32_t* __restrict__ dst) {
    const C* __restrict__ leaf_end = ix.M + ix.tab[chrom];
    uint32_t r = hi, rem = n_stop, compat = 0;
    while (rem != 0 && r != 0) {
        if ((r & (kFanout - 1)) == 0 && n_levels > 1) {
            const uint32_t span = emit_skip_span<C>(ix, ix.tab, stride, n_levels, chrom, qs, r);
            if (span) {
                r -= span;
                continue;
            }
        }
        --r;
        if (leaf_end[r] >= qs) {
            const bool ok = strand_compatible(strands[r], q_strand);
            if (kEmit) {
                if (ok) dst[--rem] = r;
            } else {
                --rem;
                compat += ok ? 1u : 0u;
            }
        }
    }
    return compat;
}

// Pass 1: one query per thread.  err_flag bit 0: start > end; bit 2: strand byte not one of + - . *
template <typename C, typename Q>
__global__ void __launch_bounds__(kQueryThreads)
strand_count_kernel(IndexView<C> ix, const uint8_t* __restrict__ strands, const uint32_t* __restrict__ cbegin,
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
                                 : strand_walk<C, false>(ix, stride, ix.n_levels, chrom, qs, hi, n_sp, strands + cbegin[chrom], qst, nullptr);
        }
        cnt_out[q] = cnt;
        hi_out[q] = hi;
    }
}

// Pass 2: hits[offsets[q] - hits_base ...) = compatible chain ranks, ascending.  Ranks at or beyond `capacity` are
// not written (the caller sized the buffer from offsets[nq]).
template <typename C, typename Q>
__global__ void __launch_bounds__(kQueryThreads)
strand_emit_kernel(IndexView<C> ix, const uint8_t* __restrict__ strands, const uint32_t* __restrict__ cbegin,
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
        strand_walk<C, true>(ix, stride, ix.n_levels, chrom, static_cast<C>(qstart[q]), hi_in[q], n, strands + cbegin[chrom],
                             qstrand[q], hits + (o0 - hits_base));
    }
}

}  // namespace gg
