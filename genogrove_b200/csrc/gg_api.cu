// gg_api.cu — C ABI (include/gg_b200.h) and host orchestration of libgg_b200.so.
//
// No torch, no C++ types across the boundary.  Everything that computes launches the hand-written
// sm_100a kernels in radix_sort.cuh / scan.cuh / build.cuh / intersect.cuh; there is no CPU path.
#include "../../include/gg_b200.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "build.cuh"
#include "common.cuh"
#include "device_buf.h"
#include "emit_scan.cuh"
#include "emit_tma.cuh"
#include "flanking.cuh"
#include "fused.cuh"
#include "intersect.cuh"
#include "radix_sort.cuh"
#include "scan.cuh"
#include "strand.cuh"

using namespace gg;

namespace gg {
std::atomic<long long> g_alloc_fail_countdown{0};
}

namespace {

#ifndef GG_BATCH_QPT
#define GG_BATCH_QPT 4
#endif
constexpr int kBatchQpt = GG_BATCH_QPT;  // consecutive queries per thread of the grouped count kernel
constexpr size_t kBatchBlockQ = 32 * kBatchQpt;  // = one warp's share: the granularity of the batch path's block totals
constexpr int kSlots = 3;        // pipeline depth of the host-buffer path
constexpr int kCountThreadsDefault = 256;

struct Slot {
    cudaStream_t st = nullptr;
    DeviceBuf qc, qs, qe, cnt, hi, off, hits, scan, fscratch, err;
    uint64_t launch_cap = 0;  // hit capacity the fused kernel of the current chunk was launched with
    cudaEvent_t ev_scan = nullptr, ev_total = nullptr, ev_done = nullptr;
    uint64_t* h_next_base = nullptr;  // pinned
    bool busy = false;
};

}  // namespace

struct gg_index {
    int device = 0;
    int sm_count = 148;
    std::string err;
    cudaStream_t stream = nullptr;  // build + two-phase host API

    // ---- index snapshot
    bool built = false;
    int coord_bytes = 4;
    uint64_t n = 0;
    uint32_t n_chrom = 0;
    int n_levels = 0;
    int stage_level = 0;
    uint32_t staged_elems = 0;
    uint64_t total_elems = 0;
    DeviceBuf S, E, M, tab, cnt, dirS, dirE, dir_off;
    DeviceBuf ER;           // chain rank of each sorted end (flanking.cuh), indexed by level-0 slot - er_base
    uint32_t er_base = 0;   // = tab[0][0]
    DeviceBuf cbegin;       // [n_chrom + 1] first concatenated chain rank of each chromosome
    // build workspace, kept between builds (a facade re-syncs after every mutation: cudaMalloc / cudaFree of ~600 MB of sort
    // buffers cost 60 of the 100 ms of a 10 M-interval rebuild); option "keep_build_workspace" = 0 frees it after each build
    struct BuildWs {
        DeviceBuf chrom, start, end, tie, stats, perm, perm_alt, perm_end, perm_end_alt, keys_a, keys_b, hist, scan, s32, e32, inv;
        std::vector<DeviceBuf*> all() {
            return {&chrom, &start, &end, &tie, &stats, &perm, &perm_alt, &perm_end, &perm_end_alt, &keys_a, &keys_b, &hist, &scan, &s32, &e32, &inv};
        }
    } bw;
    int64_t keep_build_workspace = 1;
    // what the kept workspace says about the snapshot's leaf chains (gg_index_update_chrom splices one chromosome into them):
    // 0 nothing usable, 1 bw.chrom/start/end ARE the chains (upload_sorted / update_chrom), 2 they are the build input and
    // chain_perm (inside bw.perm or bw.perm_alt) maps chain rank -> input index
    int chain_src = 0;
    const uint32_t* chain_perm = nullptr;
    int64_t debug_checks = 0;    // 1: asynchronous device calls that can detect misuse synchronise and report it (gg_intersect_emit_device)
    int64_t emit_tile_list = 0;  // pass 2 on sorted tiles: 1 = one shared candidate list per 32-query tile, filtered per lane (emit_scan.cuh, kTileList)
    int64_t emit_vec_out = 0;  // pass 2 copy-out: 1 = 16-byte vector stores from a phase-aligned stage (emit_scan.cuh, kVecOut)
    int64_t delta_min_hits = 1024;  // dense-batch emit: sorted 32-query tiles with at least this many hits share one candidate list (0 = off)
    int64_t coop_budget = 32;  // dense-batch emit: per-lane walk steps before the warp finishes a query together (0 = never)
    int64_t sort_geom = 0;     // onesweep tile geometry (radix_sort.cuh)
    int64_t sort_legacy = 0;   // 1 = three-kernel radix passes (histogram, scan, scatter) instead of the onesweep form
    // pinned staging ring for host -> device copies of pageable caller memory (parallel memcpy into it, see staged_h2d)
    static constexpr int kStageSlots = 4;
    static constexpr size_t kStageBytes = 16u << 20;
    char* h_stage = nullptr;
    cudaEvent_t ev_stage[kStageSlots] = {nullptr, nullptr, nullptr, nullptr};
    DeviceBuf flat[kFlatLevels];  // flat max-end hierarchy over the level-0 slots (emit_tma.cuh)
    DeviceBuf m8;           // max end per aligned group of 8 level-0 slots (warp-cooperative walk, emit_scan.cuh)
    uint32_t n_m8 = 0;
    DeviceBuf cslot;        // [n_chrom + 1] first level-0 slot of each chromosome (= tab[0][c] - tab[0][0])
    int n_flat = 0;
    int dir_shift = 0;
    uint64_t dir_entries = 0;
    std::vector<uint64_t> h_cnt;
    std::vector<uint32_t> h_tab;   // [n_levels][n_chrom+1]
    std::vector<uint64_t> h_lvbase;
    gg_index_stats istats{};

    // ---- device-API scratch
    DeviceBuf q_cnt, q_hi, scan_scratch, err_flag, fused_scratch, fused_out;
    size_t counted_nq = 0;
    bool timing = false;
    bool in_device_call = false;  // gg_intersect_fused_device is running: its kernels are bracketed by ev[] when timing is on
    cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    bool ev_count_valid = false, ev_emit_valid = false;

    // ---- two-phase host API state
    DeviceBuf h2_qc, h2_qs, h2_qe, h2_off, h2_hits;
    DeviceBuf strand;        // u8 per level-0 slot ('+', '-', '.', '*'; pads 0), gg_index_set_strands
    DeviceBuf SN, EN;        // StrandNode per level-0 node (+1 sentinel) in chain order / sorted-end order (strand.cuh)
    bool has_strand = false;
    DeviceBuf sq_qc, sq_qs, sq_qe, sq_str, sq_off, sq_hits;  // gg_intersect_batch_stranded staging
    DeviceBuf s_cnt, s_hi, s_tile, s_scan, s_total;                   // stranded path scratch (device calls too)
    size_t h2_nq = 0;
    uint64_t h2_total = 0;
    bool h2_valid = false;

    // ---- pipelined host path
    Slot slots[kSlots];
    bool slots_ready = false;
    DeviceBuf d_base;
    uint32_t* pinned_hits = nullptr;
    size_t pinned_cap = 0;  // elements
    uint32_t* h_err = nullptr;  // pinned

    // ---- options
    int64_t chunk_queries = 8388608;
    int64_t stage_bytes = 16384;
    int64_t count_threads = kCountThreadsDefault;
    int64_t count_ctas_per_sm = 0;  // 0 = occupancy API
    int64_t emit_ctas_per_sm = 0;
    int64_t fused_ctas_per_sm = 0;
    int64_t fused_kernel = -1;  // single-call path: 0 = count -> block scan -> emit_scan kernels, 1 = the one-kernel look-back form,
                                // -1 = by batch size (one kernel up to kOneKernelMaxQ queries: launch latency dominates there)
    bool last_one_kernel = false;  // what the last single-call dispatch ran (which kernel times are valid)
    int64_t dense_emit = -1;   // segment-staged emit of dense warp tiles (emit_scan.cuh): -1 = by the density seen so far, 0 off, 1 on
    bool s_dense = false;      // stranded path: the count phase's choice, kept for the emit phase of the same batch
    double seen_hits_per_query = 0.0;  // of the most recent batch whose hit total the host has read
    double seen_hits_per_query_strand = 0.0;  // same for the stranded path (its dense threshold differs: 32 vs 12)
    // fully asynchronous device calls (total_hits == NULL) still feed the density heuristic: the call's hit total is copied
    // to this pinned mailbox behind the kernels and picked up — if it has landed — at the start of the next call
    uint32_t* h_lazy = nullptr;
    cudaEvent_t ev_lazy = nullptr;
    bool lazy_pending = false, lazy_stranded = false;
    size_t lazy_nq = 0;
    int64_t emit_walk = 8;     // ranks per walk step of emit_scan: 8 (sorted batches) or 16 (scattered batches)
    int64_t emit_tma = 0;      // pass 2 of the single-call path: 0 = intersect_emit_scan_kernel [default: measured faster, the
                               // kernel is issue-bound and the TMA plumbing costs instructions], 1 = intersect_emit_tma_kernel
                               // (TMA ring in, TMA bulk store out), 2 = same with 16-byte vector stores instead of the bulk store
    // ---- compact host path (gg_intersect_batch_compact)
    CompactQuery* pinned_cq = nullptr;   // pinned, nq entries
    size_t pinned_cq_cap = 0;
    uint8_t* pinned_delta = nullptr;     // pinned byte stream of back-deltas
    size_t pinned_delta_cap = 0;
    std::vector<gg_compact_chunk> cchunks;
    std::vector<uint32_t> h_cslot;       // host copy of the first level-0 slot of each chromosome
    int compact_delta_bytes = 1;         // width the next chunk starts with (sticky: widened after an overflow)
    int64_t fused_regs40 = 0;  // 1: the 40-register build of the fused kernel (6 CTAs/SM) instead of 32 registers (8 CTAs/SM)
    int64_t host_mode = 0;   // host-buffer pipeline: 0 = one fused kernel per chunk, 1 = count -> scan -> emit kernels
    int64_t count_mode = 2;  // 2: directory + finger search, 4 consecutive queries per thread; 0: directory, one query
                             // per thread; 1: tiered tree kernel (TMA windows / warp / per-lane)

    gg_batch_stats bstats{};
};

namespace {

gg_status fail(gg_index* idx, gg_status st, const char* what, cudaError_t e = cudaSuccess) {
    if (idx) {
        idx->err = what;
        if (e != cudaSuccess) {
            idx->err += ": ";
            idx->err += cudaGetErrorString(e);
        }
    }
    if (e == cudaErrorMemoryAllocation) {
        cudaGetLastError();
        return GG_ERR_OOM;
    }
    return st;
}

#define GG_CUDA(call)                                                    \
    do {                                                                 \
        cudaError_t e_ = (call);                                         \
        if (e_ != cudaSuccess) return fail(idx, GG_ERR_CUDA, #call, e_); \
    } while (0)

// dense warp tiles (>= kDenseMin hits in 32 queries) only exist when queries average tens of hits; the dense pass (and the
// per-tile totals it needs from the count pass) is switched on from what earlier batches / chunks on this handle showed
inline bool dense_pass_on(const gg_index* idx, bool stranded = false) {
    if (idx->dense_emit >= 0) return idx->dense_emit != 0;
    return stranded ? idx->seen_hits_per_query_strand >= 32.0 : idx->seen_hits_per_query >= 12.0;
}
// (batches of fewer than 64 k queries say little about the next large one and do not move the estimate)
inline void note_density(gg_index* idx, uint64_t hits, size_t nq, bool stranded = false) {
    if (nq < 65536) return;
    (stranded ? idx->seen_hits_per_query_strand : idx->seen_hits_per_query) = static_cast<double>(hits) / static_cast<double>(nq);
}
// pick up the hit total an asynchronous device call left in the pinned mailbox (no wait: skipped if not landed yet)
inline void poll_lazy_density(gg_index* idx) {
    if (!idx->lazy_pending || cudaEventQuery(idx->ev_lazy) != cudaSuccess) {
        cudaGetLastError();
        return;
    }
    uint64_t total = 0;
    std::memcpy(&total, idx->h_lazy + 2, 8);
    note_density(idx, total, idx->lazy_nq, idx->lazy_stranded);
    idx->lazy_pending = false;
}
inline cudaError_t post_lazy_density(gg_index* idx, const void* d_total8, size_t nq, bool stranded, cudaStream_t st) {
    if (!idx->h_lazy) {
        cudaError_t e = cudaHostAlloc(reinterpret_cast<void**>(&idx->h_lazy), 16, cudaHostAllocDefault);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&idx->ev_lazy, cudaEventDisableTiming);
        if (e != cudaSuccess) return e;
    }
    if (idx->lazy_pending && cudaEventQuery(idx->ev_lazy) != cudaSuccess) {  // previous total still in flight: keep that one
        cudaGetLastError();
        return cudaSuccess;
    }
    cudaError_t e = cudaMemcpyAsync(idx->h_lazy + 2, d_total8, 8, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaEventRecord(idx->ev_lazy, st);
    idx->lazy_pending = e == cudaSuccess;
    idx->lazy_nq = nq;
    idx->lazy_stranded = stranded;
    return e;
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
inline unsigned blocks_for(size_t n, unsigned threads) { return static_cast<unsigned>(div_up(n ? n : 1, threads)); }

template <typename C>
IndexView<C> make_view(const gg_index* idx) {
    IndexView<C> v;
    v.S = idx->S.as<C>();
    v.E = idx->E.as<C>();
    v.M = idx->M.as<C>();
    v.tab = idx->tab.as<uint32_t>();
    v.cnt = idx->cnt.as<uint32_t>();
    v.n_chrom = idx->n_chrom;
    v.n_levels = idx->n_levels;
    v.stage_level = idx->stage_level;
    v.staged_elems = idx->staged_elems;
    v.dirS = idx->dirS.as<uint32_t>();
    v.dirE = idx->dirE.as<uint32_t>();
    v.dir_off = idx->dir_off.as<uint32_t>();
    v.dir_shift = idx->dir_shift;
    v.base0 = idx->h_tab.empty() ? 0u : idx->h_tab[0];
    return v;
}

template <typename C>
CoopArgs<C> make_coop(const gg_index* idx) {
    CoopArgs<C> a;
    a.leaf = idx->M.as<C>() + (idx->h_tab.empty() ? 0u : idx->h_tab[0]);
    a.m8 = idx->m8.as<C>();
    a.lv1 = idx->flat[1].as<C>();
    a.cslot = idx->cslot.as<uint32_t>();
    a.n_groups = idx->n_m8;
    a.n_lv = idx->n_flat;
    a.budget = static_cast<int>(idx->coop_budget);
    // (only tiles beyond the dense-batch stage reach the second pass-2 kernel, which owns the sorted-tile path)
    a.delta_min = idx->delta_min_hits > 0 ? static_cast<uint32_t>(std::max<int64_t>(idx->delta_min_hits, emit_stage_entries(false, true) + 1)) : 0u;
    return a;
}

template <typename C>
FlatMax<C> make_flat(const gg_index* idx) {
    FlatMax<C> f;
    f.leaf = idx->M.as<C>() + (idx->h_tab.empty() ? 0u : idx->h_tab[0]);
    f.lv0 = idx->flat[0].as<C>();
    f.lv1 = idx->flat[1].as<C>();
    f.lv2 = idx->flat[2].as<C>();
    f.lv3 = idx->flat[3].as<C>();
    f.cbase = idx->cslot.as<uint32_t>();
    f.n_chrom = idx->n_chrom;
    f.n_lv = idx->n_flat;
    return f;
}

size_t count_smem_bytes(const gg_index* idx) {
    size_t tab_elems = static_cast<size_t>(idx->n_levels) * (idx->n_chrom + 1);
    return 2 * (static_cast<size_t>(idx->staged_elems) + kWin) * idx->coord_bytes + ((tab_elems + 3) & ~size_t(3)) * 4 + 32;
}
// the [n_levels][n_chrom + 1] offset table is copied into shared memory only while it is small; scaffold-level
// assemblies (tens of thousands of contigs) read it through L1 instead
inline bool tab_fits_smem(const gg_index* idx) { return static_cast<size_t>(idx->n_levels) * (idx->n_chrom + 1) * 4 <= 8 * 1024; }
size_t emit_smem_bytes(const gg_index* idx) {
    size_t tab_elems = tab_fits_smem(idx) ? static_cast<size_t>(idx->n_levels) * (idx->n_chrom + 1) : 0;
    return static_cast<size_t>(kQueryThreads / 32) * (kEmitStage + kEmitSlack) * 4 + tab_elems * 4 + 16;
}

// Pass 1 + offsets: count kernel (also accumulates per-4096-query tile totals), scan of the tile totals,
// per-tile scan of the counts.  3 kernel launches; offsets[0..nq] are final (+ *d_base) afterwards.
template <typename C, typename Q>
gg_status launch_count(gg_index* idx, const uint32_t* d_chrom, const Q* d_start, const Q* d_end, size_t nq,
                       uint64_t* d_off, uint32_t* d_cnt, uint32_t* d_hi, uint32_t* d_err, void* scratch,
                       const uint64_t* d_base, cudaStream_t st, cudaEvent_t ev_after_count) {
    GG_CUDA(scan2_prepare_async(scratch, nq, st));
    // (the TMA-window kernel of count_mode 1 keeps the whole offset table in shared memory: many-contig groves take the
    // directory kernel instead — same results)
    if (idx->count_mode == 2 || (idx->count_mode == 1 && !tab_fits_smem(idx))) {
        // 4 consecutive queries per thread (finger search): the default
        const int vec_ok = aligned16(d_chrom) && aligned16(d_start) && aligned16(d_end) && aligned16(d_cnt) && aligned16(d_hi);
        auto kern = intersect_count_group_kernel<C, Q, kBatchQpt>;
        int per_sm = static_cast<int>(idx->count_ctas_per_sm);
        if (per_sm <= 0) {
            GG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kQueryThreads, 0));
            if (per_sm < 1) return fail(idx, GG_ERR_CUDA, "count kernel does not fit on an SM");
        }
        size_t tiles = div_up(nq ? nq : 1, static_cast<size_t>(kQueryThreads) * kBatchQpt);
        unsigned grid = static_cast<unsigned>(std::min<size_t>(tiles, static_cast<size_t>(idx->sm_count) * per_sm));
        kern<<<grid, kQueryThreads, 0, st>>>(make_view<C>(idx), d_chrom, d_start, d_end, nq, d_cnt, d_hi,
                                             scan2_tile_total(scratch), nullptr, d_err, vec_ok, nullptr);
    } else if (idx->count_mode == 0) {
        auto kern = intersect_count_dir_kernel<C, Q>;
        int per_sm = static_cast<int>(idx->count_ctas_per_sm);
        if (per_sm <= 0) {
            GG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kQueryThreads, 0));
            if (per_sm < 1) return fail(idx, GG_ERR_CUDA, "count kernel does not fit on an SM");
        }
        size_t tiles = div_up(nq ? nq : 1, static_cast<size_t>(kQueryThreads));
        unsigned grid = static_cast<unsigned>(std::min<size_t>(tiles, static_cast<size_t>(idx->sm_count) * per_sm));
        kern<<<grid, kQueryThreads, 0, st>>>(make_view<C>(idx), d_chrom, d_start, d_end, nq, d_cnt, d_hi,
                                             scan2_tile_total(scratch), d_err);
    } else {
        auto kern = intersect_count_kernel<C, Q>;
        size_t smem = count_smem_bytes(idx);
        GG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
        int threads = static_cast<int>(idx->count_threads);
        int per_sm = static_cast<int>(idx->count_ctas_per_sm);
        if (per_sm <= 0) {
            GG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, threads, smem));
            if (per_sm < 1) return fail(idx, GG_ERR_CUDA, "count kernel does not fit on an SM");
        }
        size_t tiles = div_up(nq ? nq : 1, static_cast<size_t>(threads));
        unsigned grid = static_cast<unsigned>(std::min<size_t>(tiles, static_cast<size_t>(idx->sm_count) * per_sm));
        kern<<<grid, threads, smem, st>>>(make_view<C>(idx), d_chrom, d_start, d_end, nq, d_cnt, d_hi,
                                          scan2_tile_total(scratch), d_err);
    }
    GG_CUDA(cudaGetLastError());
    if (ev_after_count) GG_CUDA(cudaEventRecord(ev_after_count, st));
    GG_CUDA(scan2_finish_async(d_cnt, d_off, nq, scratch, d_base, st));
    idx->bstats.kernel_launches += 3;
    return GG_OK;
}

inline size_t count_scratch_bytes(size_t nq) { return scan2_scratch_bytes(nq); }

template <typename C, typename Q>
gg_status launch_emit(gg_index* idx, const uint32_t* d_chrom, const Q* d_start, size_t nq, const uint32_t* d_hi,
                      const uint64_t* d_off, uint32_t* d_hits, uint64_t hits_base, cudaStream_t st,
                      const uint32_t* d_cnt_check = nullptr) {
    auto kern = intersect_emit_kernel<C, Q>;
    size_t smem = emit_smem_bytes(idx);
    if (smem > 48 * 1024)
        GG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    int per_sm = static_cast<int>(idx->emit_ctas_per_sm);
    if (per_sm <= 0) {
        GG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kQueryThreads, smem));
        if (per_sm < 1) return fail(idx, GG_ERR_CUDA, "emit kernel does not fit on an SM");
    }
    size_t tiles = div_up(nq ? nq : 1, static_cast<size_t>(kQueryThreads));
    unsigned grid = static_cast<unsigned>(std::min<size_t>(tiles, static_cast<size_t>(idx->sm_count) * per_sm));
    kern<<<grid, kQueryThreads, smem, st>>>(make_view<C>(idx), d_chrom, d_start, nq, d_hi, d_off, d_hits, hits_base,
                                            tab_fits_smem(idx) ? 1 : 0, d_cnt_check, d_cnt_check ? idx->err_flag.as<uint32_t>() : nullptr);
    GG_CUDA(cudaGetLastError());
    idx->bstats.kernel_launches++;
    return GG_OK;
}

// dispatch on (index coord width, query coord width)
gg_status dispatch_count(gg_index* idx, const uint32_t* d_chrom, const void* d_start, const void* d_end,
                         int q_bytes, size_t nq, uint64_t* d_off, uint32_t* d_cnt, uint32_t* d_hi, uint32_t* d_err,
                         void* scratch, const uint64_t* d_base, cudaStream_t st, cudaEvent_t ev = nullptr) {
    if (idx->coord_bytes == 4 && q_bytes == 4)
        return launch_count<uint32_t, uint32_t>(idx, d_chrom, (const uint32_t*)d_start, (const uint32_t*)d_end, nq, d_off, d_cnt, d_hi, d_err, scratch, d_base, st, ev);
    if (idx->coord_bytes == 4 && q_bytes == 8)
        return launch_count<uint32_t, uint64_t>(idx, d_chrom, (const uint64_t*)d_start, (const uint64_t*)d_end, nq, d_off, d_cnt, d_hi, d_err, scratch, d_base, st, ev);
    if (idx->coord_bytes == 8 && q_bytes == 4)
        return launch_count<uint64_t, uint32_t>(idx, d_chrom, (const uint32_t*)d_start, (const uint32_t*)d_end, nq, d_off, d_cnt, d_hi, d_err, scratch, d_base, st, ev);
    return launch_count<uint64_t, uint64_t>(idx, d_chrom, (const uint64_t*)d_start, (const uint64_t*)d_end, nq, d_off, d_cnt, d_hi, d_err, scratch, d_base, st, ev);
}
gg_status dispatch_emit(gg_index* idx, const uint32_t* d_chrom, const void* d_start, int q_bytes, size_t nq,
                        const uint32_t* d_hi, const uint64_t* d_off, uint32_t* d_hits, uint64_t hits_base,
                        cudaStream_t st, const uint32_t* d_cnt_check = nullptr) {
    if (idx->coord_bytes == 4 && q_bytes == 4)
        return launch_emit<uint32_t, uint32_t>(idx, d_chrom, (const uint32_t*)d_start, nq, d_hi, d_off, d_hits, hits_base, st, d_cnt_check);
    if (idx->coord_bytes == 4 && q_bytes == 8)
        return launch_emit<uint32_t, uint64_t>(idx, d_chrom, (const uint64_t*)d_start, nq, d_hi, d_off, d_hits, hits_base, st, d_cnt_check);
    if (idx->coord_bytes == 8 && q_bytes == 4)
        return launch_emit<uint64_t, uint32_t>(idx, d_chrom, (const uint32_t*)d_start, nq, d_hi, d_off, d_hits, hits_base, st, d_cnt_check);
    return launch_emit<uint64_t, uint64_t>(idx, d_chrom, (const uint64_t*)d_start, nq, d_hi, d_off, d_hits, hits_base, st, d_cnt_check);
}

// The one-kernel path (fused.cuh): scratch = [ticket][tile state], zeroed per launch.
size_t fused_smem_bytes(const gg_index* idx, bool* tab_in_smem) {
    size_t tab_bytes = static_cast<size_t>(idx->n_levels) * (idx->n_chrom + 1) * 4;
    *tab_in_smem = tab_bytes <= 8 * 1024;  // many-contig groves read the offset table through L1 instead
    return static_cast<size_t>(kQueryThreads / 32) * (kEmitStage + kFusedSlack) * 4 + (*tab_in_smem ? tab_bytes : 0) + 16;
}

template <typename C, typename Q>
gg_status launch_fused(gg_index* idx, const uint32_t* d_chrom, const Q* d_start, const Q* d_end, size_t nq,
                       uint64_t* d_off, uint32_t* d_hits, uint64_t capacity, const uint64_t* d_base,
                       uint64_t* d_next_base, uint64_t* d_total, void* scratch, uint32_t* d_err, cudaStream_t st) {
    auto kern = idx->fused_regs40 ? intersect_fused_kernel<C, Q, 6> : intersect_fused_kernel<C, Q, 8>;
    bool tab_in_smem = true;
    const size_t smem = fused_smem_bytes(idx, &tab_in_smem);
    if (smem > 48 * 1024)
        GG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    int per_sm = static_cast<int>(idx->fused_ctas_per_sm);
    if (per_sm <= 0) {
        GG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kQueryThreads, smem));
        if (per_sm < 1) return fail(idx, GG_ERR_CUDA, "fused kernel does not fit on an SM");
    }
    GG_CUDA(cudaMemsetAsync(scratch, 0, fused_scratch_bytes(nq), st));
    const size_t tiles = fused_tiles(nq);
    const unsigned grid = static_cast<unsigned>(std::min<size_t>(tiles, static_cast<size_t>(idx->sm_count) * per_sm));
    uint32_t* ticket = static_cast<uint32_t*>(scratch);
    uint64_t* state = reinterpret_cast<uint64_t*>(static_cast<char*>(scratch) + 16);
    kern<<<grid, kQueryThreads, smem, st>>>(make_view<C>(idx), d_chrom, d_start, d_end, nq, d_off, d_hits, capacity, d_base,
                                            d_next_base, d_total, ticket, state, d_err, tab_in_smem ? 1 : 0);
    GG_CUDA(cudaGetLastError());
    idx->bstats.kernel_launches++;
    return GG_OK;
}

// The single-call batch path (default): count (4 consecutive queries per thread, finger search) -> exclusive scan
// of the per-block hit totals (one block = 128 queries, ~0.8 % of the batch) -> emit_scan (offsets + hits).
// scratch layout: [block_total u64 x nb][block_prefix u64 x (nb + 1)][look-back state of the small scan]
//                 [cnt u32 x nq][hi u32 x nq][tile_total u32 x nq/32]
inline size_t batch_blocks(size_t nq) { return div_up(nq, kBatchBlockQ); }
inline size_t batch_ctl_bytes(size_t nq) {
    const size_t nb = batch_blocks(nq) + 1;
    return round_up(nb * 8, 256) + round_up((nb + 1) * 8, 256) + round_up(scan_state_bytes(nb), 256);
}
// one scratch buffer serves either flavour of the single-call path
inline size_t single_scratch_bytes(size_t nq) {
    return std::max(round_up(fused_scratch_bytes(nq), 256), batch_ctl_bytes(nq)) + 2 * round_up((nq ? nq : 1) * 4, 256) +
           round_up(div_up(nq ? nq : 1, 32) * 4, 256);  // + per-tile totals (u32 per 32 queries) for the dense emit
}

// Second half of pass 2: the warp tiles with more hits than the stage holds (found from the count pass's per-tile totals).
template <typename C, typename Q, bool kStrand>
gg_status launch_emit_dense(gg_index* idx, const uint32_t* d_chrom, const Q* d_start, size_t nq, const uint32_t* d_cnt,
                            const uint32_t* d_hi, const uint64_t* block_prefix, const uint32_t* d_tile,
                            uint32_t* d_hits, uint64_t capacity, StrandArgs sa, cudaStream_t st, bool delta = false) {
    if (nq == 0) return GG_OK;
    auto kern = intersect_emit_dense_kernel<C, Q, kStrand>;
    const size_t smem = static_cast<size_t>(kQueryThreads / 32) * (dense_stage_entries(kStrand) + kEmitScanSlack) * 4 + 16;
    CoopArgs<C> ca = delta ? make_coop<C>(idx) : CoopArgs<C>{};
    if (!delta) ca.delta_min = 0;
    int per_sm = 0;
    GG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kQueryThreads, smem));
    if (per_sm < 1) return fail(idx, GG_ERR_CUDA, "dense emit kernel does not fit on an SM");
    const size_t ctas = div_up(div_up(nq, 32 * 32), kQueryThreads / 32);  // a warp step covers 32 tiles
    const unsigned grid = static_cast<unsigned>(std::min<size_t>(std::max<size_t>(ctas, 1), static_cast<size_t>(idx->sm_count) * per_sm));
    kern<<<grid, kQueryThreads, smem, st>>>(make_view<C>(idx), d_chrom, d_start, nq, d_cnt, d_hi, block_prefix, d_tile,
                                            static_cast<uint32_t>(kBatchBlockQ), d_hits, capacity, sa, emit_stage_entries(kStrand, true), ca);
    GG_CUDA(cudaGetLastError());
    idx->bstats.kernel_launches++;
    return GG_OK;
}

template <typename C, typename Q>
gg_status launch_batch(gg_index* idx, const uint32_t* d_chrom, const Q* d_start, const Q* d_end, size_t nq,
                       uint64_t* d_off, uint32_t* d_hits, uint64_t capacity, const uint64_t* d_base,
                       uint64_t* d_next_base, uint64_t* d_total, void* scratch, uint32_t* d_err, cudaStream_t st) {
    char* p = static_cast<char*>(scratch);
    const size_t nb = batch_blocks(nq), nb1 = nb + 1;
    unsigned long long* block_total = reinterpret_cast<unsigned long long*>(p);
    uint64_t* block_prefix = reinterpret_cast<uint64_t*>(p + round_up(nb1 * 8, 256));
    void* scan_state = p + round_up(nb1 * 8, 256) + round_up((nb1 + 1) * 8, 256);
    char* tail = p + std::max(round_up(fused_scratch_bytes(nq), 256), batch_ctl_bytes(nq));
    uint32_t* d_cnt = reinterpret_cast<uint32_t*>(tail);
    uint32_t* d_hi = reinterpret_cast<uint32_t*>(tail + round_up((nq ? nq : 1) * 4, 256));
    uint32_t* d_tile = reinterpret_cast<uint32_t*>(tail + 2 * round_up((nq ? nq : 1) * 4, 256));
    const bool tab_in_smem = static_cast<size_t>(idx->n_levels) * (idx->n_chrom + 1) * 4 <= 8 * 1024;
    const bool dense = dense_pass_on(idx);  // the dense variants also carry the hybrid walk for tiles above the stage
    // pass 2 through the async proxy (emit_tma.cuh): TMA needs 16-byte aligned query arrays and 32-bit tile indices
    const bool tma = !dense && idx->emit_tma != 0 && idx->emit_walk == 8 && aligned16(d_chrom) && aligned16(d_start) &&
                     nq < 0xFFFFFFFFull;
    if (nq) {
        auto kern = dense ? intersect_count_group_kernel<C, Q, kBatchQpt, true>
                          : (tma ? intersect_count_group_kernel<C, Q, kBatchQpt, false, true> : intersect_count_group_kernel<C, Q, kBatchQpt, false>);
        int per_sm = static_cast<int>(idx->count_ctas_per_sm);
        if (per_sm <= 0) {
            GG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kQueryThreads, 0));
            if (per_sm < 1) return fail(idx, GG_ERR_CUDA, "count kernel does not fit on an SM");
        }
        const size_t tiles = div_up(nq, static_cast<size_t>(kQueryThreads) * kBatchQpt);
        const unsigned grid = static_cast<unsigned>(std::min<size_t>(tiles, static_cast<size_t>(idx->sm_count) * per_sm));
        const int vec_ok = aligned16(d_chrom) && aligned16(d_start) && aligned16(d_end);
        kern<<<grid, kQueryThreads, 0, st>>>(make_view<C>(idx), d_chrom, d_start, d_end, nq, d_cnt, d_hi, nullptr, block_total,
                                             d_err, vec_ok, dense ? d_tile : nullptr);
        GG_CUDA(cudaGetLastError());
        idx->bstats.kernel_launches++;
    }
    const bool timed = idx->timing && idx->in_device_call;
    if (timed) GG_CUDA(cudaEventRecord(idx->ev[1], st));
    GG_CUDA((exclusive_scan_async<uint64_t, uint64_t>(reinterpret_cast<const uint64_t*>(block_total), block_prefix, nb, scan_state,
                                                       d_base, nullptr, st)));
    idx->bstats.kernel_launches++;
    if (timed) {
        GG_CUDA(cudaEventRecord(idx->ev[2], st));
        GG_CUDA(cudaEventRecord(idx->ev[3], st));
    }
    if (tma) {
        const bool cb_smem = (static_cast<size_t>(idx->n_chrom) + 1) * 4 <= 8 * 1024;
        const size_t smem = emit_tma_smem_bytes<Q>(cb_smem ? idx->n_chrom + 1 : 0);
        auto kern = idx->emit_tma == 2 ? intersect_emit_tma_kernel<C, Q, uint32_t, false, false> : intersect_emit_tma_kernel<C, Q, uint32_t, false, true>;
        GG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
        int per_sm = static_cast<int>(idx->emit_ctas_per_sm);
        if (per_sm <= 0) {
            GG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kQueryThreads, smem));
            if (per_sm < 1) return fail(idx, GG_ERR_CUDA, "emit kernel does not fit on an SM");
        }
        const size_t ctas = div_up(std::max<size_t>(nb, 1), kQueryThreads / 32);
        const unsigned grid = static_cast<unsigned>(std::min<size_t>(ctas, static_cast<size_t>(idx->sm_count) * per_sm));
        kern<<<grid, kQueryThreads, smem, st>>>(make_flat<C>(idx), reinterpret_cast<const uint2*>(d_cnt), d_chrom, d_start, nq, block_prefix,
                                                static_cast<uint32_t>(kBatchBlockQ), d_off, nullptr, d_hits, capacity, d_next_base, d_total,
                                                d_err, cb_smem ? 1 : 0);
        GG_CUDA(cudaGetLastError());
        idx->bstats.kernel_launches++;
        return GG_OK;
    }
    {
        const size_t tab_bytes = static_cast<size_t>(idx->n_levels) * (idx->n_chrom + 1) * 4;
        const bool vec_out = !dense && idx->emit_walk == 8 && idx->emit_vec_out != 0;
        const bool tile_list = !dense && !vec_out && idx->emit_walk == 8 && idx->emit_tile_list != 0;
        const size_t smem = static_cast<size_t>(kQueryThreads / 32) * (emit_stage_entries(false, dense) + kEmitScanSlack + (vec_out ? kVecOutTailWords : 0u) +
                                                                       (tile_list ? tile_list_words<C>() : 0u)) * 4 +
                            (tab_in_smem ? tab_bytes : 0) + 16;
        void (*kern)(IndexView<C>, const uint32_t*, const Q*, size_t, const uint32_t*, const uint32_t*, const uint64_t*, uint32_t,
                     uint64_t*, uint32_t*, uint64_t, uint64_t*, uint64_t*, uint32_t*, int, StrandArgs, CoopArgs<C>) =
            dense ? intersect_emit_scan_kernel<C, Q, 8, false, true>
                  : (idx->emit_walk == 16 ? intersect_emit_scan_kernel<C, Q, 16>
                                          : (vec_out ? intersect_emit_scan_kernel<C, Q, 8, false, false, true>
                                                     : (tile_list ? intersect_emit_scan_kernel<C, Q, 8, false, false, false, true> : intersect_emit_scan_kernel<C, Q, 8>)));
        if (smem > 48 * 1024)
            GG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
        int per_sm = static_cast<int>(idx->emit_ctas_per_sm);
        if (per_sm <= 0) {
            GG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kQueryThreads, smem));
            if (per_sm < 1) return fail(idx, GG_ERR_CUDA, "emit kernel does not fit on an SM");
        }
        const size_t ctas = div_up(std::max<size_t>(nb, 1), kQueryThreads / 32);
        const unsigned grid = static_cast<unsigned>(std::min<size_t>(ctas, static_cast<size_t>(idx->sm_count) * per_sm));
        kern<<<grid, kQueryThreads, smem, st>>>(make_view<C>(idx), d_chrom, d_start, nq, d_cnt, d_hi, block_prefix,
                                                static_cast<uint32_t>(kBatchBlockQ), d_off, d_hits, capacity, d_next_base, d_total,
                                                d_err, tab_in_smem ? 1 : 0, StrandArgs{nullptr, nullptr, 0}, (dense || tile_list) ? make_coop<C>(idx) : CoopArgs<C>{});
        GG_CUDA(cudaGetLastError());
        idx->bstats.kernel_launches++;
    }
    if (!dense) return GG_OK;
    return launch_emit_dense<C, Q, false>(idx, d_chrom, d_start, nq, d_cnt, d_hi, block_prefix, d_tile, d_hits, capacity,
                                          StrandArgs{nullptr, nullptr, 0}, st, /*delta=*/true);
}

// measured on B200 (tools/small_batch_probe.py, cfg2 index): 1 k .. 100 k queries 30-33 us in one kernel against 39-44 us for
// the three launches; 56 vs 63 us at 1 M queries; from 4 M on the three-kernel form wins (121 vs 135 us, 234 vs 295 us at 10 M)
constexpr size_t kOneKernelMaxQ = 512 * 1024;
inline bool use_one_kernel(const gg_index* idx, size_t nq) {
    return idx->fused_kernel < 0 ? nq <= kOneKernelMaxQ : idx->fused_kernel != 0;
}

gg_status dispatch_fused(gg_index* idx, const uint32_t* d_chrom, const void* d_start, const void* d_end, int q_bytes,
                         size_t nq, uint64_t* d_off, uint32_t* d_hits, uint64_t capacity, const uint64_t* d_base,
                         uint64_t* d_next_base, uint64_t* d_total, void* scratch, uint32_t* d_err, cudaStream_t st) {
    const bool one = use_one_kernel(idx, nq);
    idx->last_one_kernel = one;
#define GG_SINGLE(C, Q)                                                                                                      \
    (one ? launch_fused<C, Q>(idx, d_chrom, (const Q*)d_start, (const Q*)d_end, nq, d_off, d_hits, capacity,    \
                                            d_base, d_next_base, d_total, scratch, d_err, st)                                \
                       : launch_batch<C, Q>(idx, d_chrom, (const Q*)d_start, (const Q*)d_end, nq, d_off, d_hits, capacity,    \
                                            d_base, d_next_base, d_total, scratch, d_err, st))
    if (idx->coord_bytes == 4 && q_bytes == 4) return GG_SINGLE(uint32_t, uint32_t);
    if (idx->coord_bytes == 4 && q_bytes == 8) return GG_SINGLE(uint32_t, uint64_t);
    if (idx->coord_bytes == 8 && q_bytes == 4) return GG_SINGLE(uint64_t, uint32_t);
    return GG_SINGLE(uint64_t, uint64_t);
#undef GG_SINGLE
}

// ------------------------------------------------------------------------------------------ build

// GG_TRACE_BUILD=1: host timestamps of the build stages on stderr (each after a stream synchronise), for profiling the
// host-side cost of a build (allocations, pageable copies) that CUDA events on the stream do not attribute
struct BuildTrace {
    bool on;
    cudaStream_t st;
    std::chrono::steady_clock::time_point t0;
    explicit BuildTrace(cudaStream_t s) : on(std::getenv("GG_TRACE_BUILD") != nullptr), st(s), t0(std::chrono::steady_clock::now()) {}
    void mark(const char* what) {
        if (!on) return;
        cudaStreamSynchronize(st);
        const auto t1 = std::chrono::steady_clock::now();
        std::fprintf(stderr, "[gg build] %-28s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(t1 - t0).count());
        t0 = t1;
    }
};

// Level geometry from per-chromosome counts (host; C is small).  Fills h_tab / h_lvbase / n_levels.
gg_status plan_levels(gg_index* idx) {
    const uint32_t C = idx->n_chrom;
    int depth = 1;
    for (uint32_t c = 0; c < C; ++c) {
        uint64_t m = idx->h_cnt[c];
        int d = 1;
        while (m > kFanout) {
            m = div_up(m, kFanout);
            ++d;
        }
        depth = std::max(depth, d);
    }
    if (depth > kMaxLevels) return fail(idx, GG_ERR_CAPACITY, "too many levels");
    idx->n_levels = depth;
    std::vector<std::vector<uint64_t>> sz(depth, std::vector<uint64_t>(C));
    std::vector<uint64_t> level_total(depth, 0);
    for (uint32_t c = 0; c < C; ++c) {
        uint64_t m = idx->h_cnt[c];
        for (int l = 0; l < depth; ++l) {
            sz[l][c] = round_up(std::max<uint64_t>(m, 1), kFanout);
            level_total[l] += sz[l][c];
            m = div_up(m, kFanout);
        }
    }
    idx->h_lvbase.assign(depth + 1, 0);
    uint64_t run = 0;
    for (int l = depth - 1; l >= 0; --l) {  // top level first
        idx->h_lvbase[l] = run;
        run += level_total[l];
    }
    idx->total_elems = run;
    if (run >= 0xFFFFFFFFull) return fail(idx, GG_ERR_CAPACITY, "index exceeds 2^32-1 padded ranks per handle");
    idx->h_tab.assign(static_cast<size_t>(depth) * (C + 1), 0);
    for (int l = 0; l < depth; ++l) {
        uint64_t off = idx->h_lvbase[l];
        for (uint32_t c = 0; c < C; ++c) {
            idx->h_tab[static_cast<size_t>(l) * (C + 1) + c] = static_cast<uint32_t>(off);
            off += sz[l][c];
        }
        idx->h_tab[static_cast<size_t>(l) * (C + 1) + C] = static_cast<uint32_t>(off);
    }
    // staging: the longest top-first prefix of whole levels with 2 * bytes <= stage_bytes
    idx->stage_level = depth;
    idx->staged_elems = 0;
    for (int l = depth - 1; l >= 0; --l) {
        uint64_t prefix = idx->h_lvbase[l] + level_total[l];
        if (2 * prefix * idx->coord_bytes > static_cast<uint64_t>(idx->stage_bytes)) break;
        idx->stage_level = l;
        idx->staged_elems = static_cast<uint32_t>(prefix);
    }
    return GG_OK;
}

template <typename Field, typename Key>
gg_status sort_by_field(gg_index* idx, const Field* d_field, int nbits, Key* keys_a, Key* keys_b, uint32_t*& perm,
                        uint32_t*& perm_alt, size_t n, RadixWorkspace& ws, cudaStream_t st) {
    if (nbits <= 0 || n == 0) return GG_OK;
    if (radix_uses_onesweep<Key>(n, ws)) {  // gather and count the digits of the gathered keys in one kernel
        GG_CUDA(cudaMemsetAsync(ws.hist, 0, kOsMaxPasses * kRadix * 4, st));
        gather_field_hist_kernel<Field, Key><<<onesweep_hist_blocks(n), kOsThreads, 0, st>>>(d_field, perm, keys_a, n, (nbits + 7) / 8, ws.hist);
        ws.hist_ready = true;
    } else {
        gather_field_kernel<Field, Key><<<blocks_for(n, 256), 256, 0, st>>>(d_field, perm, keys_a, n);
    }
    ws.launches++;
    bool in_alt = false;
    GG_CUDA(radix_sort_pairs<Key>(keys_a, keys_b, perm, perm_alt, n, nbits, ws, st, &in_alt));
    if (in_alt) std::swap(perm, perm_alt);
    return GG_OK;
}

// Host -> device copy of PAGEABLE caller memory through a ring of pinned 16 MB slots: a few host threads memcpy the next
// slot while the previous one is on the wire (a plain cudaMemcpyAsync from pageable memory runs at ~7 GB/s on this box: 27
// of the 35 ms of a warm 10 M-interval build).  Small copies go straight through the driver.
gg_status staged_h2d(gg_index* idx, void* dst, const void* src, size_t bytes, cudaStream_t st) {
    constexpr size_t kSlot = gg_index::kStageBytes;
    constexpr int kSlots_ = gg_index::kStageSlots;
    if (bytes < (4u << 20)) {
        GG_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, st));
        return GG_OK;
    }
    if (!idx->h_stage) {
        cudaError_t e = cudaHostAlloc(reinterpret_cast<void**>(&idx->h_stage), kSlot * kSlots_, cudaHostAllocDefault);
        if (e != cudaSuccess) {  // no pinned memory to be had: the driver's own staging still works
            cudaGetLastError();
            idx->h_stage = nullptr;
            GG_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, st));
            return GG_OK;
        }
        for (auto& ev : idx->ev_stage) GG_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    }
    const unsigned T = std::max(1u, std::min(8u, std::thread::hardware_concurrency()));
    const char* s8 = static_cast<const char*>(src);
    char* d8 = static_cast<char*>(dst);
    size_t k = 0;
    for (size_t off = 0; off < bytes; off += kSlot, ++k) {
        const size_t m = std::min(kSlot, bytes - off);
        const int slot = static_cast<int>(k % kSlots_);
        char* h = idx->h_stage + static_cast<size_t>(slot) * kSlot;
        if (k >= static_cast<size_t>(kSlots_)) GG_CUDA(cudaEventSynchronize(idx->ev_stage[slot]));  // its previous copy has left
        const size_t part = round_up(div_up(m, T), 4096);
        std::vector<std::thread> th;
        for (unsigned t = 1; t < T; ++t) {
            const size_t a = std::min(m, t * part), b = std::min(m, (t + 1) * part);
            if (b > a) th.emplace_back([=] { std::memcpy(h + a, s8 + off + a, b - a); });
        }
        std::memcpy(h, s8 + off, std::min(m, part));
        for (auto& x : th) x.join();
        GG_CUDA(cudaMemcpyAsync(d8 + off, h, m, cudaMemcpyHostToDevice, st));
        GG_CUDA(cudaEventRecord(idx->ev_stage[slot], st));
    }
    // the ring may be reused by the next call right away: wait until this call's copies have read it
    for (int i = 0; i < kSlots_; ++i) GG_CUDA(cudaEventSynchronize(idx->ev_stage[i]));
    return GG_OK;
}

// Shared by gg_index_build (do_sort = true) and gg_index_upload_sorted (do_sort = false).
// dev_input: chrom / start / end / tie_key / rank_to_input are DEVICE pointers (gg_index_build_device).
gg_status build_impl(gg_index* idx, const uint32_t* chrom, const uint64_t* start, const uint64_t* end,
                     const uint64_t* tie_key, int tie_mode, size_t n, uint32_t n_chrom, uint32_t* rank_to_input,
                     bool do_sort, bool dev_input = false, bool in_place = false) {
    // in_place: the chains already sit in the handle's own workspace buffers (gg_index_update_chrom)
    if (!idx) return GG_ERR_INVALID_ARG;
    idx->err.clear();
    idx->built = false;
    idx->chain_src = 0;
    idx->has_strand = false;
    idx->h2_valid = false;
    if (n_chrom == 0 || n_chrom == GG_NO_CHROM) return fail(idx, GG_ERR_INVALID_ARG, "n_chrom must be in [1, 2^32-2]");
    if (n > 0 && (!chrom || !start || !end)) return fail(idx, GG_ERR_INVALID_ARG, "null input array");
    if (tie_mode == GG_TIE_EXPLICIT && !tie_key && n > 0) return fail(idx, GG_ERR_INVALID_ARG, "GG_TIE_EXPLICIT needs tie_key");
    if (tie_mode < 0 || tie_mode > 2) return fail(idx, GG_ERR_INVALID_ARG, "bad tie_mode");
    if (n >= 0xFFFFFFF0ull) return fail(idx, GG_ERR_CAPACITY, "more than 2^32-16 intervals per handle");
    GG_CUDA(cudaSetDevice(idx->device));
    cudaStream_t st = idx->stream;
    cudaEvent_t e0, e1, e2, e3;
    GG_CUDA(cudaEventCreate(&e0));
    GG_CUDA(cudaEventCreate(&e1));
    GG_CUDA(cudaEventCreate(&e2));
    GG_CUDA(cudaEventCreate(&e3));

    // workspace of the build (grow-only, kept in the handle unless keep_build_workspace = 0; borrowed device inputs are
    // always dropped on exit)
    gg_index::BuildWs& w = idx->bw;
    DeviceBuf &d_chrom = w.chrom, &d_start = w.start, &d_end = w.end, &d_tie = w.tie, &d_stats = w.stats, &d_perm = w.perm,
              &d_perm_alt = w.perm_alt, &d_perm_end = w.perm_end, &d_perm_end_alt = w.perm_end_alt, &d_keys_a = w.keys_a,
              &d_keys_b = w.keys_b, &d_hist = w.hist, &d_scan = w.scan, &d_s32 = w.s32, &d_e32 = w.e32, &d_inv = w.inv;
    DeviceBuf& d_cbegin = idx->cbegin;  // kept: the flanking / stranded paths index per-rank side arrays with it
    struct Guard {
        gg_index* idx;
        std::vector<cudaEvent_t> evs;
        ~Guard() {
            for (auto* b : idx->bw.all())
                if (!b->own || !idx->keep_build_workspace) b->release();
            for (auto e : evs) cudaEventDestroy(e);
        }
    } guard{idx, {e0, e1, e2, e3}};
    for (auto* b : w.all())
        if (!b->own) b->release();  // (a borrowed view left by an earlier failed call)

    const size_t nn = n ? n : 1;
    BuildTrace trace(st);
    GG_CUDA(cudaEventRecord(e0, st));
    if (in_place) {
        // nothing to fetch
    } else if (dev_input && n) {
        d_chrom.borrow(chrom, n * 4);
        d_start.borrow(start, n * 8);
        d_end.borrow(end, n * 8);
        if (tie_mode == GG_TIE_EXPLICIT) d_tie.borrow(tie_key, n * 8);
    } else {
        GG_CUDA(d_chrom.ensure(nn * 4));
        GG_CUDA(d_start.ensure(nn * 8));
        GG_CUDA(d_end.ensure(nn * 8));
    }
    GG_CUDA(d_stats.ensure(sizeof(BuildStats)));
    GG_CUDA(idx->cnt.ensure(static_cast<size_t>(n_chrom) * 4));
    GG_CUDA(cudaMemsetAsync(d_stats.p, 0, sizeof(BuildStats), st));
    GG_CUDA(cudaMemsetAsync(idx->cnt.p, 0, static_cast<size_t>(n_chrom) * 4, st));
    if (n) {
        if (!dev_input && !in_place) {
            idx->istats.h2d_bytes = n * 20 + (tie_mode == GG_TIE_EXPLICIT ? n * 8 : 0);
            gg_status hs_ = staged_h2d(idx, d_chrom.p, chrom, n * 4, st);
            if (hs_ == GG_OK) hs_ = staged_h2d(idx, d_start.p, start, n * 8, st);
            if (hs_ == GG_OK) hs_ = staged_h2d(idx, d_end.p, end, n * 8, st);
            if (hs_ == GG_OK && tie_mode == GG_TIE_EXPLICIT) {
                GG_CUDA(d_tie.ensure(n * 8));
                hs_ = staged_h2d(idx, d_tie.p, tie_key, n * 8, st);
            }
            if (hs_ != GG_OK) return hs_;
        }
        index_stats_kernel<<<std::min<unsigned>(blocks_for(n, 256), 148 * 8), 256, 0, st>>>(d_chrom.as<uint32_t>(), d_start.as<uint64_t>(),
                                                              d_end.as<uint64_t>(), d_tie.as<uint64_t>(), n, n_chrom,
                                                              idx->cnt.as<uint32_t>(), d_stats.as<BuildStats>());
        if (!do_sort)
            check_chain_kernel<<<blocks_for(n, 256), 256, 0, st>>>(d_chrom.as<uint32_t>(), d_start.as<uint64_t>(), n,
                                                                  d_stats.as<BuildStats>());
    }
    trace.mark("alloc + H2D + stats");
    BuildStats hs{};
    std::vector<uint32_t> h_cnt32(n_chrom);
    GG_CUDA(cudaMemcpyAsync(&hs, d_stats.p, sizeof(BuildStats), cudaMemcpyDeviceToHost, st));
    GG_CUDA(cudaMemcpyAsync(h_cnt32.data(), idx->cnt.p, static_cast<size_t>(n_chrom) * 4, cudaMemcpyDeviceToHost, st));
    GG_CUDA(cudaStreamSynchronize(st));
    if (hs.invalid_chrom) return fail(idx, GG_ERR_INVALID_ARG, "chrom id >= n_chrom in index input");
    if (hs.invalid_interval) return fail(idx, GG_ERR_INVALID_INTERVAL, "interval: start must be <= end");
    if (hs.unsorted) return fail(idx, GG_ERR_UNSORTED, "upload_sorted input is not in (chrom, start) order");

    idx->n = n;
    idx->n_chrom = n_chrom;
    idx->coord_bytes = (hs.max_coord > 0xFFFFFFFFull) ? 8 : 4;
    idx->h_cnt.assign(h_cnt32.begin(), h_cnt32.end());
    gg_status ps = plan_levels(idx);
    if (ps != GG_OK) return ps;
    const uint32_t C = n_chrom;

    std::vector<uint32_t> h_cbegin(C + 1, 0);
    for (uint32_t c = 0; c < C; ++c) h_cbegin[c + 1] = h_cbegin[c] + h_cnt32[c];
    GG_CUDA(d_cbegin.ensure((C + 1) * 4));
    GG_CUDA(cudaMemcpyAsync(d_cbegin.p, h_cbegin.data(), (C + 1) * 4, cudaMemcpyHostToDevice, st));
    GG_CUDA(idx->tab.ensure(idx->h_tab.size() * 4));
    GG_CUDA(cudaMemcpyAsync(idx->tab.p, idx->h_tab.data(), idx->h_tab.size() * 4, cudaMemcpyHostToDevice, st));

    // ---- sort: permutation of input indices in leaf-chain order, then in (chrom, end) order
    const bool wide = idx->coord_bytes == 8;
    const int coord_bits = bits_needed(hs.max_coord);
    const int chrom_bits = bits_needed(C - 1);
    RadixWorkspace ws;
    GG_CUDA(d_perm.ensure(nn * 4));
    GG_CUDA(d_perm_alt.ensure(nn * 4));
    GG_CUDA(d_perm_end.ensure(nn * 4));
    GG_CUDA(d_perm_end_alt.ensure(nn * 4));
    GG_CUDA(d_keys_a.ensure(nn * 8));
    GG_CUDA(d_keys_b.ensure(nn * 8));
    GG_CUDA(d_hist.ensure(radix_hist_entries(n) * 4));
    GG_CUDA(d_scan.ensure(scan_state_bytes(radix_hist_entries(n))));
    ws.hist = d_hist.as<uint32_t>();
    ws.scan_scratch = d_scan.p;
    ws.force_legacy = static_cast<int>(idx->sort_legacy);
    ws.geom = static_cast<int>(idx->sort_geom);

    // narrowed copies of the coordinates (u32 layout) in input order
    const void* src_s = d_start.p;
    const void* src_e = d_end.p;
    if (!wide) {
        GG_CUDA(d_s32.ensure(nn * 4));
        GG_CUDA(d_e32.ensure(nn * 4));
        if (n) {
            narrow_kernel<uint32_t><<<blocks_for(n, 256), 256, 0, st>>>(d_start.as<uint64_t>(), d_s32.as<uint32_t>(), n);
            narrow_kernel<uint32_t><<<blocks_for(n, 256), 256, 0, st>>>(d_end.as<uint64_t>(), d_e32.as<uint32_t>(), n);
        }
        src_s = d_s32.p;
        src_e = d_e32.p;
    }

    trace.mark("plan + sort buffers + narrow");
    GG_CUDA(cudaEventRecord(e1, st));
    uint32_t* perm = d_perm.as<uint32_t>();
    uint32_t* perm_alt = d_perm_alt.as<uint32_t>();
    uint32_t* perm_end = d_perm_end.as<uint32_t>();
    uint32_t* perm_end_alt = d_perm_end_alt.as<uint32_t>();
    gg_status s = GG_OK;
    if (n) {
        if (do_sort) {
            iota_kernel<<<blocks_for(n, 256), 256, 0, st>>>(perm, n, tie_mode == GG_TIE_REVERSE_INPUT ? 1 : 0);
            // least significant field first: tie, end, start, chrom
            if (tie_mode == GG_TIE_EXPLICIT)
                s = sort_by_field<uint64_t, uint64_t>(idx, d_tie.as<uint64_t>(), bits_needed(hs.max_tie), d_keys_a.as<uint64_t>(), d_keys_b.as<uint64_t>(), perm, perm_alt, n, ws, st);
            if (s != GG_OK) return s;
            if (wide) {
                s = sort_by_field<uint64_t, uint64_t>(idx, d_end.as<uint64_t>(), coord_bits, d_keys_a.as<uint64_t>(), d_keys_b.as<uint64_t>(), perm, perm_alt, n, ws, st);
                if (s == GG_OK) s = sort_by_field<uint64_t, uint64_t>(idx, d_start.as<uint64_t>(), coord_bits, d_keys_a.as<uint64_t>(), d_keys_b.as<uint64_t>(), perm, perm_alt, n, ws, st);
            } else {
                s = sort_by_field<uint32_t, uint32_t>(idx, d_e32.as<uint32_t>(), coord_bits, d_keys_a.as<uint32_t>(), d_keys_b.as<uint32_t>(), perm, perm_alt, n, ws, st);
                if (s == GG_OK) s = sort_by_field<uint32_t, uint32_t>(idx, d_s32.as<uint32_t>(), coord_bits, d_keys_a.as<uint32_t>(), d_keys_b.as<uint32_t>(), perm, perm_alt, n, ws, st);
            }
            if (s == GG_OK) s = sort_by_field<uint32_t, uint32_t>(idx, d_chrom.as<uint32_t>(), chrom_bits, d_keys_a.as<uint32_t>(), d_keys_b.as<uint32_t>(), perm, perm_alt, n, ws, st);
            if (s != GG_OK) return s;
        } else {
            iota_kernel<<<blocks_for(n, 256), 256, 0, st>>>(perm, n, 0);
        }
        // ends sorted per chromosome (E family): sort input indices by (chrom, end), starting from the chain order so
        // that equal ends stay in chain order (the flanking predecessor is the FIRST key in the chain with that end)
        GG_CUDA(cudaMemcpyAsync(perm_end, perm, n * 4, cudaMemcpyDeviceToDevice, st));
        if (wide)
            s = sort_by_field<uint64_t, uint64_t>(idx, d_end.as<uint64_t>(), coord_bits, d_keys_a.as<uint64_t>(), d_keys_b.as<uint64_t>(), perm_end, perm_end_alt, n, ws, st);
        else
            s = sort_by_field<uint32_t, uint32_t>(idx, d_e32.as<uint32_t>(), coord_bits, d_keys_a.as<uint32_t>(), d_keys_b.as<uint32_t>(), perm_end, perm_end_alt, n, ws, st);
        if (s == GG_OK) s = sort_by_field<uint32_t, uint32_t>(idx, d_chrom.as<uint32_t>(), chrom_bits, d_keys_a.as<uint32_t>(), d_keys_b.as<uint32_t>(), perm_end, perm_end_alt, n, ws, st);
        if (s != GG_OK) return s;
    }
    trace.mark("sorts");
    GG_CUDA(cudaEventRecord(e2, st));

    // ---- padded SoA leaf level + upper levels
    const size_t fam_bytes = static_cast<size_t>(idx->total_elems) * idx->coord_bytes;
    GG_CUDA(idx->S.ensure(fam_bytes));
    GG_CUDA(idx->E.ensure(fam_bytes));
    GG_CUDA(idx->M.ensure(fam_bytes));
    const uint32_t* tab0 = idx->tab.as<uint32_t>();
    const uint32_t level0_elems = idx->h_tab[C] - idx->h_tab[0];
    if (wide)
        pad_leaf_kernel<uint64_t><<<blocks_for(level0_elems, 256), 256, 0, st>>>(
            perm, perm_end, (const uint64_t*)src_s, (const uint64_t*)src_e, d_cbegin.as<uint32_t>(), idx->cnt.as<uint32_t>(),
            tab0, C, idx->S.as<uint64_t>(), idx->E.as<uint64_t>(), idx->M.as<uint64_t>());
    else
        pad_leaf_kernel<uint32_t><<<blocks_for(level0_elems, 256), 256, 0, st>>>(
            perm, perm_end, (const uint32_t*)src_s, (const uint32_t*)src_e, d_cbegin.as<uint32_t>(), idx->cnt.as<uint32_t>(),
            tab0, C, idx->S.as<uint32_t>(), idx->E.as<uint32_t>(), idx->M.as<uint32_t>());
    // chain rank of every sorted end (ER) for the flanking path
    idx->er_base = idx->h_tab[0];
    GG_CUDA(idx->ER.ensure(static_cast<size_t>(level0_elems ? level0_elems : 1) * 4));
    GG_CUDA(d_inv.ensure(nn * 4));
    if (n) invert_perm_kernel<<<blocks_for(n, 256), 256, 0, st>>>(perm, d_inv.as<uint32_t>(), n);
    end_rank_kernel<<<blocks_for(level0_elems, 256), 256, 0, st>>>(perm_end, d_inv.as<uint32_t>(), d_cbegin.as<uint32_t>(),
                                                                   idx->cnt.as<uint32_t>(), tab0, C, idx->ER.as<uint32_t>());
    for (int l = 1; l < idx->n_levels; ++l) {
        const uint32_t* row = tab0 + static_cast<size_t>(l) * (C + 1);
        const uint32_t* row_below = tab0 + static_cast<size_t>(l - 1) * (C + 1);
        const uint32_t entries = idx->h_tab[static_cast<size_t>(l) * (C + 1) + C] - idx->h_tab[static_cast<size_t>(l) * (C + 1)];
        const unsigned warps_per_block = 8;
        if (wide)
            build_level_kernel<uint64_t><<<blocks_for(entries, warps_per_block), warps_per_block * 32, 0, st>>>(
                idx->S.as<uint64_t>(), idx->E.as<uint64_t>(), idx->M.as<uint64_t>(), row, row_below, C);
        else
            build_level_kernel<uint32_t><<<blocks_for(entries, warps_per_block), warps_per_block * 32, 0, st>>>(
                idx->S.as<uint32_t>(), idx->E.as<uint32_t>(), idx->M.as<uint32_t>(), row, row_below, C);
    }
    GG_CUDA(cudaGetLastError());
    trace.mark("leaf + levels");
    {
        // flat max-end hierarchy over the level-0 slots of all chromosomes (emit_tma.cuh) + first slot per chromosome
        std::vector<uint32_t>& h_cslot = idx->h_cslot;
        h_cslot.resize(C + 1);
        for (uint32_t c = 0; c <= C; ++c) h_cslot[c] = idx->h_tab[c] - idx->h_tab[0];
        GG_CUDA(idx->cslot.ensure((C + 1) * 4));
        GG_CUDA(cudaMemcpyAsync(idx->cslot.p, h_cslot.data(), (C + 1) * 4, cudaMemcpyHostToDevice, st));
        GG_CUDA(cudaStreamSynchronize(st));  // (h_cslot is pageable host data)
        uint32_t n_src = level0_elems;
        const char* src = idx->M.as<char>() + static_cast<size_t>(idx->h_tab[0]) * idx->coord_bytes;
        idx->n_flat = 0;
        for (int k = 0; k < kFlatLevels; ++k) {
            const uint32_t n_dst = static_cast<uint32_t>(div_up(n_src, kFanout));
            GG_CUDA(idx->flat[k].ensure(static_cast<size_t>(n_dst ? n_dst : 1) * idx->coord_bytes));
            if (n_dst) {
                if (wide)
                    flat_max_kernel<uint64_t><<<blocks_for(n_dst, 8), 256, 0, st>>>(reinterpret_cast<const uint64_t*>(src), idx->flat[k].as<uint64_t>(), n_src, n_dst);
                else
                    flat_max_kernel<uint32_t><<<blocks_for(n_dst, 8), 256, 0, st>>>(reinterpret_cast<const uint32_t*>(src), idx->flat[k].as<uint32_t>(), n_src, n_dst);
            }
            idx->n_flat = k + 1;
            src = idx->flat[k].as<char>();
            n_src = n_dst;
            if (n_dst <= 1) break;
        }
        // max end per group of 8 slots
        idx->n_m8 = static_cast<uint32_t>(div_up(level0_elems, 8));
        GG_CUDA(idx->m8.ensure(static_cast<size_t>(idx->n_m8 ? idx->n_m8 : 1) * idx->coord_bytes));
        if (idx->n_m8) {
            const char* leaf0 = idx->M.as<char>() + static_cast<size_t>(idx->h_tab[0]) * idx->coord_bytes;
            if (wide)
                group8_max_kernel<uint64_t><<<blocks_for(idx->n_m8, 256), 256, 0, st>>>(reinterpret_cast<const uint64_t*>(leaf0), idx->m8.as<uint64_t>(), idx->n_m8);
            else
                group8_max_kernel<uint32_t><<<blocks_for(idx->n_m8, 256), 256, 0, st>>>(reinterpret_cast<const uint32_t*>(leaf0), idx->m8.as<uint32_t>(), idx->n_m8);
        }
        GG_CUDA(cudaGetLastError());
    }
    trace.mark("flat max hierarchy");
    if (rank_to_input && n)
        GG_CUDA(cudaMemcpyAsync(rank_to_input, perm, n * 4, dev_input ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, st));
    {
        // ---- root directory: per-chromosome max coordinate = max over the (single) top node of M
        const size_t top_elems = static_cast<size_t>(C) * kFanout;
        std::vector<uint64_t> top64(top_elems);
        std::vector<uint32_t> top32(wide ? 0 : top_elems);
        const size_t top_off = idx->h_tab[static_cast<size_t>(idx->n_levels - 1) * (C + 1)];
        if (wide)
            GG_CUDA(cudaMemcpyAsync(top64.data(), idx->M.as<uint64_t>() + top_off, top_elems * 8, cudaMemcpyDeviceToHost, st));
        else
            GG_CUDA(cudaMemcpyAsync(top32.data(), idx->M.as<uint32_t>() + top_off, top_elems * 4, cudaMemcpyDeviceToHost, st));
        GG_CUDA(cudaStreamSynchronize(st));
        std::vector<uint64_t> cmax(C, 0);
        long double span = 0;
        for (uint32_t c = 0; c < C; ++c) {
            for (int j = 0; j < kFanout; ++j) {
                uint64_t v = wide ? top64[static_cast<size_t>(c) * kFanout + j] : top32[static_cast<size_t>(c) * kFanout + j];
                cmax[c] = std::max(cmax[c], v);
            }
            span += static_cast<long double>(cmax[c]);
        }
        // bins of 2^shift positions, ~8 keys per bin on average, at most ~N/8 + 2C entries per family
        const long double target_bins = std::max<long double>(static_cast<long double>(n) / 8.0L, 1024.0L);
        int shift = 0;
        const int max_shift = idx->coord_bytes * 8 - 1;
        while (shift < max_shift && span / static_cast<long double>(1ull << shift) > target_bins) ++shift;
        std::vector<uint32_t> h_doff(C + 1, 0);
        uint64_t entries = 0;
        for (uint32_t c = 0; c < C; ++c) {
            h_doff[c] = static_cast<uint32_t>(entries);
            entries += (cmax[c] >> shift) + 2;
        }
        h_doff[C] = static_cast<uint32_t>(entries);
        if (entries >= 0xFFFFFFF0ull) return fail(idx, GG_ERR_CAPACITY, "root directory too large");
        idx->dir_shift = shift;
        idx->dir_entries = entries;
        GG_CUDA(idx->dirS.ensure(entries * 4));
        GG_CUDA(idx->dirE.ensure(entries * 4));
        GG_CUDA(idx->dir_off.ensure((C + 1) * 4));
        GG_CUDA(cudaMemcpyAsync(idx->dir_off.p, h_doff.data(), (C + 1) * 4, cudaMemcpyHostToDevice, st));
        if (wide)
            build_dir_kernel<uint64_t><<<blocks_for(entries, 256), 256, 0, st>>>(make_view<uint64_t>(idx), idx->dirS.as<uint32_t>(), idx->dirE.as<uint32_t>());
        else
            build_dir_kernel<uint32_t><<<blocks_for(entries, 256), 256, 0, st>>>(make_view<uint32_t>(idx), idx->dirS.as<uint32_t>(), idx->dirE.as<uint32_t>());
        GG_CUDA(cudaGetLastError());
    }
    GG_CUDA(cudaEventRecord(e3, st));
    GG_CUDA(cudaStreamSynchronize(st));
    trace.mark("perm D2H + directory");

    float ms_sort = 0, ms_levels = 0, ms_total = 0;
    cudaEventElapsedTime(&ms_sort, e1, e2);
    cudaEventElapsedTime(&ms_levels, e2, e3);
    cudaEventElapsedTime(&ms_total, e0, e3);
    idx->istats.n_intervals = n;
    idx->istats.n_chrom = C;
    idx->istats.coord_bytes = idx->coord_bytes;
    idx->istats.n_levels = idx->n_levels;
    idx->istats.fanout = kFanout;
    idx->istats.staged_levels = idx->n_levels - idx->stage_level;
    idx->istats.device_bytes = 3 * fam_bytes + idx->h_tab.size() * 4 + static_cast<size_t>(C) * 4 + idx->dir_entries * 8 +
                               static_cast<size_t>(level0_elems) * 4;
    idx->istats.sort_ms = ms_sort;
    idx->istats.levels_ms = ms_levels;
    idx->istats.total_ms = ms_total;
    if (!dev_input && !in_place) idx->istats.h2d_bytes = n * 20 + (tie_mode == GG_TIE_EXPLICIT ? n * 8 : 0);
    idx->built = true;
    if (idx->keep_build_workspace && (!dev_input || in_place)) {
        idx->chain_src = do_sort ? 2 : 1;
        idx->chain_perm = do_sort ? perm : nullptr;
    }
    return GG_OK;
}

template <typename T>
__global__ void chain_gather_kernel(const uint32_t* __restrict__ perm, const T* __restrict__ in, size_t n, T* __restrict__ out) {
    const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) out[i] = in[perm[i]];
}
__global__ void fill_u32_kernel(uint32_t* __restrict__ p, size_t n, uint32_t v) {
    const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

// gg_index_update_chrom: splice the n_c records of one chromosome into the chains kept on the device and rebuild the levels
// there.  Only those records cross PCIe (16 B each); everything else is device-to-device at HBM speed.
gg_status update_chrom_impl(gg_index* idx, uint32_t chrom, const uint64_t* start, const uint64_t* end, size_t n_c) {
    if (!idx) return GG_ERR_INVALID_ARG;
    idx->err.clear();
    if (!idx->built || idx->chain_src == 0)
        return fail(idx, GG_ERR_NOT_BUILT, "gg_index_update_chrom needs a snapshot whose chains are still on the device "
                                           "(gg_index_upload_sorted / gg_index_build with keep_build_workspace = 1)");
    if (chrom == GG_NO_CHROM || chrom > idx->n_chrom) return fail(idx, GG_ERR_INVALID_ARG, "chrom must be an existing id or n_chrom (append)");
    if (n_c > 0 && (!start || !end)) return fail(idx, GG_ERR_INVALID_ARG, "null input array");
    GG_CUDA(cudaSetDevice(idx->device));
    cudaStream_t st = idx->stream;
    const uint32_t C_old = idx->n_chrom, C_new = std::max(C_old, chrom + 1);
    const size_t n_old = idx->n;
    std::vector<uint64_t> cb(C_old + 1, 0);
    for (uint32_t c = 0; c < C_old; ++c) cb[c + 1] = cb[c] + idx->h_cnt[c];
    const size_t before = chrom < C_old ? cb[chrom] : n_old;
    const size_t old_c = chrom < C_old ? idx->h_cnt[chrom] : 0;
    const size_t after = n_old - before - old_c;
    const size_t n_new = before + n_c + after;
    if (n_new >= 0xFFFFFFF0ull) return fail(idx, GG_ERR_CAPACITY, "more than 2^32-16 intervals per handle");
    gg_index::BuildWs& w = idx->bw;
    DeviceBuf t_chrom, t_start, t_end;
    const size_t nn = n_new ? n_new : 1;
    GG_CUDA(t_chrom.ensure(nn * 4));
    GG_CUDA(t_start.ensure(nn * 8));
    GG_CUDA(t_end.ensure(nn * 8));
    auto fail_free = [&](cudaError_t e) {
        t_chrom.release(), t_start.release(), t_end.release();
        return fail(idx, GG_ERR_CUDA, "gg_index_update_chrom", e);
    };
    cudaError_t e = cudaSuccess;
    auto move = [&](size_t dst, size_t src, size_t m) {  // m records of the old chains -> new position
        if (m == 0 || e != cudaSuccess) return;
        if (idx->chain_src == 1) {
            e = cudaMemcpyAsync(t_chrom.as<uint32_t>() + dst, w.chrom.as<uint32_t>() + src, m * 4, cudaMemcpyDeviceToDevice, st);
            if (e == cudaSuccess) e = cudaMemcpyAsync(t_start.as<uint64_t>() + dst, w.start.as<uint64_t>() + src, m * 8, cudaMemcpyDeviceToDevice, st);
            if (e == cudaSuccess) e = cudaMemcpyAsync(t_end.as<uint64_t>() + dst, w.end.as<uint64_t>() + src, m * 8, cudaMemcpyDeviceToDevice, st);
        } else {  // chains = build input gathered through the chain permutation
            chain_gather_kernel<uint32_t><<<blocks_for(m, 256), 256, 0, st>>>(idx->chain_perm + src, w.chrom.as<uint32_t>(), m, t_chrom.as<uint32_t>() + dst);
            chain_gather_kernel<uint64_t><<<blocks_for(m, 256), 256, 0, st>>>(idx->chain_perm + src, w.start.as<uint64_t>(), m, t_start.as<uint64_t>() + dst);
            chain_gather_kernel<uint64_t><<<blocks_for(m, 256), 256, 0, st>>>(idx->chain_perm + src, w.end.as<uint64_t>(), m, t_end.as<uint64_t>() + dst);
            e = cudaGetLastError();
        }
    };
    move(0, 0, before);
    move(before + n_c, before + old_c, after);
    if (e != cudaSuccess) return fail_free(e);
    if (n_c) {
        gg_status hs = staged_h2d(idx, t_start.as<uint64_t>() + before, start, n_c * 8, st);
        if (hs == GG_OK) hs = staged_h2d(idx, t_end.as<uint64_t>() + before, end, n_c * 8, st);
        if (hs != GG_OK) {
            t_chrom.release(), t_start.release(), t_end.release();
            return hs;
        }
        fill_u32_kernel<<<blocks_for(n_c, 256), 256, 0, st>>>(t_chrom.as<uint32_t>() + before, n_c, chrom);
    }
    e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return fail_free(e);
    // the new chains become the workspace's input arrays, then the ordinary sorted-upload build runs on them in place
    w.chrom.release(), w.start.release(), w.end.release();
    w.chrom = t_chrom, w.start = t_start, w.end = t_end;  // (ownership moves: the temporaries are not released)
    gg_status bs = build_impl(idx, w.chrom.as<uint32_t>(), w.start.as<uint64_t>(), w.end.as<uint64_t>(), nullptr, GG_TIE_INPUT_ORDER,
                              n_new, C_new, nullptr, /*do_sort=*/false, /*dev_input=*/false, /*in_place=*/true);
    if (bs == GG_OK) idx->istats.h2d_bytes = n_c * 16;
    return bs;
}

// ------------------------------------------------------------------------------------------ device path

gg_status ensure_query_scratch(gg_index* idx, size_t nq) {
    const size_t m = nq ? nq : 1;
    GG_CUDA(idx->q_cnt.ensure(m * 4));
    GG_CUDA(idx->q_hi.ensure(m * 4));
    GG_CUDA(idx->scan_scratch.ensure(count_scratch_bytes(m)));
    GG_CUDA(idx->err_flag.ensure(16));
    return GG_OK;
}

gg_status check_query_error(gg_index* idx, cudaStream_t st) {
    uint32_t flag = 0;
    GG_CUDA(cudaMemcpyAsync(&flag, idx->err_flag.p, 4, cudaMemcpyDeviceToHost, st));
    GG_CUDA(cudaStreamSynchronize(st));
    if (flag & 1u) return fail(idx, GG_ERR_INVALID_INTERVAL, "query interval: start must be <= end");
    if (flag & 32u) return fail(idx, GG_ERR_INVALID_ARG, "offsets passed to gg_intersect_emit_device are not the ones gg_intersect_count_device wrote");
    return GG_OK;
}

gg_status ensure_slots(gg_index* idx) {
    if (idx->slots_ready) return GG_OK;
    for (int i = 0; i < kSlots; ++i) {
        Slot& s = idx->slots[i];
        GG_CUDA(cudaStreamCreateWithFlags(&s.st, cudaStreamNonBlocking));
        GG_CUDA(cudaEventCreateWithFlags(&s.ev_scan, cudaEventDisableTiming));
        GG_CUDA(cudaEventCreateWithFlags(&s.ev_total, cudaEventDisableTiming));
        GG_CUDA(cudaEventCreateWithFlags(&s.ev_done, cudaEventDisableTiming));
        GG_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&s.h_next_base), 16, cudaHostAllocDefault));
    }
    if (!idx->h_err) GG_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&idx->h_err), 16, cudaHostAllocDefault));
    idx->slots_ready = true;
    return GG_OK;
}

gg_status ensure_pinned_hits(gg_index* idx, size_t need_elems, size_t keep_elems) {
    if (need_elems <= idx->pinned_cap && idx->pinned_hits) return GG_OK;
    size_t want = std::max<size_t>(need_elems, 1024);
    uint32_t* np = nullptr;
    cudaError_t e = cudaHostAlloc(reinterpret_cast<void**>(&np), want * 4, cudaHostAllocDefault);
    if (e != cudaSuccess) return fail(idx, GG_ERR_OOM, "pinned hit buffer", e);
    if (idx->pinned_hits) {
        if (keep_elems) std::memcpy(np, idx->pinned_hits, keep_elems * 4);
        cudaFreeHost(idx->pinned_hits);
    }
    idx->pinned_hits = np;
    idx->pinned_cap = want;
    return GG_OK;
}

// Error exits of the pipelined host-buffer calls must not leave chunks in flight: their async D2H copies would keep
// writing into the caller's buffers after the call has failed.  Drains every slot stream unless dismissed.
struct SlotDrain {
    gg_index* idx;
    bool armed = true;
    ~SlotDrain() {
        if (!armed || !idx->slots_ready) return;
        for (int i = 0; i < kSlots; ++i) {
            if (idx->slots[i].st) cudaStreamSynchronize(idx->slots[i].st);
            idx->slots[i].busy = false;
        }
        cudaGetLastError();
    }
};

// The reference-facing host-buffer call: chunked count -> scan -> emit pipeline over kSlots streams.
template <typename Q>
gg_status intersect_batch_host(gg_index* idx, const uint32_t* q_chrom, const Q* q_start, const Q* q_end, size_t nq,
                               uint64_t* offsets, const uint32_t** hits, uint64_t* n_hits) {
    if (!idx) return GG_ERR_INVALID_ARG;
    idx->err.clear();
    if (!idx->built) return fail(idx, GG_ERR_NOT_BUILT, "index not built");
    if (!offsets || !hits || !n_hits || (nq && (!q_chrom || !q_start || !q_end)))
        return fail(idx, GG_ERR_INVALID_ARG, "null argument");
    if (sizeof(Q) == 4 && idx->coord_bytes != 4) return fail(idx, GG_ERR_INVALID_ARG, "u32 queries need a u32-layout index");
    GG_CUDA(cudaSetDevice(idx->device));
    gg_status s = ensure_slots(idx);
    if (s != GG_OK) return s;
    idx->h2_valid = false;
    idx->bstats = gg_batch_stats{};
    idx->bstats.n_queries = nq;

    const size_t chunk = static_cast<size_t>(std::max<int64_t>(idx->chunk_queries, kScanTile) / kScanTile * kScanTile);
    const size_t n_chunks = nq ? div_up(nq, chunk) : 0;
    idx->bstats.n_chunks = static_cast<uint32_t>(n_chunks);
    *hits = nullptr;
    *n_hits = 0;
    if (nq == 0) {
        offsets[0] = 0;
        s = ensure_pinned_hits(idx, 1, 0);
        *hits = idx->pinned_hits;
        return s;
    }
    GG_CUDA(idx->d_base.ensure((n_chunks + 1) * 8));
    GG_CUDA(idx->err_flag.ensure(16));
    cudaStream_t st0 = idx->slots[0].st;
    GG_CUDA(cudaMemsetAsync(idx->d_base.p, 0, 8, st0));
    GG_CUDA(cudaMemsetAsync(idx->err_flag.p, 0, 4, st0));
    GG_CUDA(cudaStreamSynchronize(st0));
    for (int i = 0; i < kSlots; ++i) idx->slots[i].busy = false;  // (every exit of the previous call drained the slots)
    SlotDrain drain{idx};

    std::vector<uint64_t> base(n_chunks + 1, 0);
    cudaEvent_t prev_scan = nullptr;

    // stage A: H2D + count + scan for chunk k
    auto stage_a = [&](size_t k) -> gg_status {
        Slot& sl = idx->slots[k % kSlots];
        if (sl.busy) {
            GG_CUDA(cudaEventSynchronize(sl.ev_done));
            sl.busy = false;
        }
        const size_t q0 = k * chunk, m = std::min(chunk, nq - q0);
        GG_CUDA(sl.qc.ensure(m * 4));
        GG_CUDA(sl.qs.ensure(m * sizeof(Q)));
        GG_CUDA(sl.qe.ensure(m * sizeof(Q)));
        GG_CUDA(sl.cnt.ensure(m * 4));
        GG_CUDA(sl.hi.ensure(m * 4));
        GG_CUDA(sl.off.ensure((m + 1) * 8));
        GG_CUDA(sl.scan.ensure(count_scratch_bytes(m)));
        GG_CUDA(cudaMemcpyAsync(sl.qc.p, q_chrom + q0, m * 4, cudaMemcpyHostToDevice, sl.st));
        GG_CUDA(cudaMemcpyAsync(sl.qs.p, q_start + q0, m * sizeof(Q), cudaMemcpyHostToDevice, sl.st));
        GG_CUDA(cudaMemcpyAsync(sl.qe.p, q_end + q0, m * sizeof(Q), cudaMemcpyHostToDevice, sl.st));
        // the scan starts from the running hit total left by the previous chunk
        if (prev_scan) GG_CUDA(cudaStreamWaitEvent(sl.st, prev_scan, 0));
        uint64_t* d_base = idx->d_base.as<uint64_t>();
        gg_status r = dispatch_count(idx, sl.qc.as<uint32_t>(), sl.qs.p, sl.qe.p, sizeof(Q), m, sl.off.as<uint64_t>(),
                                     sl.cnt.as<uint32_t>(), sl.hi.as<uint32_t>(), idx->err_flag.as<uint32_t>(), sl.scan.p,
                                     d_base + k, sl.st);
        if (r != GG_OK) return r;
        // next base = last offset of this chunk (off[m] = base + total)
        GG_CUDA(cudaMemcpyAsync(d_base + k + 1, sl.off.as<uint64_t>() + m, 8, cudaMemcpyDeviceToDevice, sl.st));
        GG_CUDA(cudaEventRecord(sl.ev_scan, sl.st));
        prev_scan = sl.ev_scan;
        GG_CUDA(cudaMemcpyAsync(sl.h_next_base, sl.off.as<uint64_t>() + m, 8, cudaMemcpyDeviceToHost, sl.st));
        GG_CUDA(cudaEventRecord(sl.ev_total, sl.st));
        const size_t n_off = (k + 1 == n_chunks) ? m + 1 : m;
        GG_CUDA(cudaMemcpyAsync(offsets + q0, sl.off.p, n_off * 8, cudaMemcpyDeviceToHost, sl.st));
        sl.busy = true;
        return GG_OK;
    };
    // stage B: learn the chunk's hit total, emit, D2H the hits
    auto stage_b = [&](size_t k) -> gg_status {
        Slot& sl = idx->slots[k % kSlots];
        const size_t q0 = k * chunk, m = std::min(chunk, nq - q0);
        GG_CUDA(cudaEventSynchronize(sl.ev_total));
        base[k + 1] = *sl.h_next_base;
        const uint64_t total = base[k + 1] - base[k];
        if (total) {
            GG_CUDA(sl.hits.ensure(total * 4));
            if (base[k + 1] > idx->pinned_cap) {
                // grow with an estimate of the final size; earlier chunks' copies must have landed first
                for (int i = 0; i < kSlots; ++i)
                    if (i != static_cast<int>(k % kSlots) && idx->slots[i].busy) GG_CUDA(cudaStreamSynchronize(idx->slots[i].st));
                GG_CUDA(cudaStreamSynchronize(sl.st));
                double per_q = static_cast<double>(base[k + 1]) / static_cast<double>(q0 + m);
                size_t est = static_cast<size_t>(per_q * 1.15 * static_cast<double>(nq)) + 4096;
                gg_status r = ensure_pinned_hits(idx, std::max<size_t>(est, base[k + 1]), base[k]);
                if (r != GG_OK) return r;
            }
            gg_status r = dispatch_emit(idx, sl.qc.as<uint32_t>(), sl.qs.p, sizeof(Q), m, sl.hi.as<uint32_t>(),
                                        sl.off.as<uint64_t>(), sl.hits.as<uint32_t>(), base[k], sl.st);
            if (r != GG_OK) return r;
            GG_CUDA(cudaMemcpyAsync(idx->pinned_hits + base[k], sl.hits.p, total * 4, cudaMemcpyDeviceToHost, sl.st));
        }
        GG_CUDA(cudaEventRecord(sl.ev_done, sl.st));
        return GG_OK;
    };

    // ---- one fused kernel per chunk (default): the hit capacity of a chunk is an estimate from the chunks before
    // it; a chunk that overflows is re-run with the exact size (offsets are complete either way).
    bool any_invalid = false;
    double est_per_q = 8.0;
    auto fstage_a = [&](size_t k) -> gg_status {
        Slot& sl = idx->slots[k % kSlots];
        if (sl.busy) {
            GG_CUDA(cudaEventSynchronize(sl.ev_done));
            sl.busy = false;
        }
        const size_t q0 = k * chunk, m = std::min(chunk, nq - q0);
        GG_CUDA(sl.qc.ensure(m * 4));
        GG_CUDA(sl.qs.ensure(m * sizeof(Q)));
        GG_CUDA(sl.qe.ensure(m * sizeof(Q)));
        GG_CUDA(sl.off.ensure((m + 1) * 8));
        GG_CUDA(sl.fscratch.ensure(single_scratch_bytes(m)));
        GG_CUDA(sl.err.ensure(16));
        const size_t want = static_cast<size_t>(est_per_q * 1.25 * static_cast<double>(m)) + 4096;
        if (sl.hits.cap / 4 < want) GG_CUDA(sl.hits.ensure(want * 4));
        sl.launch_cap = sl.hits.cap / 4;
        GG_CUDA(cudaMemcpyAsync(sl.qc.p, q_chrom + q0, m * 4, cudaMemcpyHostToDevice, sl.st));
        GG_CUDA(cudaMemcpyAsync(sl.qs.p, q_start + q0, m * sizeof(Q), cudaMemcpyHostToDevice, sl.st));
        GG_CUDA(cudaMemcpyAsync(sl.qe.p, q_end + q0, m * sizeof(Q), cudaMemcpyHostToDevice, sl.st));
        GG_CUDA(cudaMemsetAsync(sl.err.p, 0, 4, sl.st));
        if (prev_scan) GG_CUDA(cudaStreamWaitEvent(sl.st, prev_scan, 0));  // d_base[k] is written by chunk k-1's kernel
        uint64_t* d_base = idx->d_base.as<uint64_t>();
        gg_status r = dispatch_fused(idx, sl.qc.as<uint32_t>(), sl.qs.p, sl.qe.p, sizeof(Q), m, sl.off.as<uint64_t>(),
                                     sl.hits.as<uint32_t>(), sl.launch_cap, d_base + k, d_base + k + 1, nullptr,
                                     sl.fscratch.p, sl.err.as<uint32_t>(), sl.st);
        if (r != GG_OK) return r;
        GG_CUDA(cudaEventRecord(sl.ev_scan, sl.st));
        prev_scan = sl.ev_scan;
        GG_CUDA(cudaMemcpyAsync(sl.h_next_base, d_base + k + 1, 8, cudaMemcpyDeviceToHost, sl.st));
        GG_CUDA(cudaMemcpyAsync(sl.h_next_base + 1, sl.err.p, 4, cudaMemcpyDeviceToHost, sl.st));
        GG_CUDA(cudaEventRecord(sl.ev_total, sl.st));
        const size_t n_off = (k + 1 == n_chunks) ? m + 1 : m;
        GG_CUDA(cudaMemcpyAsync(offsets + q0, sl.off.p, n_off * 8, cudaMemcpyDeviceToHost, sl.st));
        sl.busy = true;
        return GG_OK;
    };
    auto fstage_b = [&](size_t k) -> gg_status {
        Slot& sl = idx->slots[k % kSlots];
        const size_t q0 = k * chunk, m = std::min(chunk, nq - q0);
        GG_CUDA(cudaEventSynchronize(sl.ev_total));
        base[k + 1] = sl.h_next_base[0];
        if (static_cast<uint32_t>(sl.h_next_base[1]) & 1u) any_invalid = true;
        const uint64_t total = base[k + 1] - base[k];
        est_per_q = std::max(1.0, static_cast<double>(base[k + 1]) / static_cast<double>(q0 + m));
        note_density(idx, base[k + 1], q0 + m);
        if (total) {
            if (total > sl.launch_cap) {  // estimate too small: exact re-run of this chunk (same base, same offsets)
                GG_CUDA(cudaStreamSynchronize(sl.st));
                GG_CUDA(sl.hits.ensure(total * 4));
                sl.launch_cap = sl.hits.cap / 4;
                gg_status r = dispatch_fused(idx, sl.qc.as<uint32_t>(), sl.qs.p, sl.qe.p, sizeof(Q), m, sl.off.as<uint64_t>(),
                                             sl.hits.as<uint32_t>(), sl.launch_cap, idx->d_base.as<uint64_t>() + k, nullptr,
                                             nullptr, sl.fscratch.p, sl.err.as<uint32_t>(), sl.st);
                if (r != GG_OK) return r;
            }
            if (base[k + 1] > idx->pinned_cap) {
                for (int i = 0; i < kSlots; ++i)
                    if (i != static_cast<int>(k % kSlots) && idx->slots[i].busy) GG_CUDA(cudaStreamSynchronize(idx->slots[i].st));
                GG_CUDA(cudaStreamSynchronize(sl.st));
                size_t est = static_cast<size_t>(est_per_q * 1.15 * static_cast<double>(nq)) + 4096;
                gg_status r = ensure_pinned_hits(idx, std::max<size_t>(est, base[k + 1]), base[k]);
                if (r != GG_OK) return r;
            }
            GG_CUDA(cudaMemcpyAsync(idx->pinned_hits + base[k], sl.hits.p, total * 4, cudaMemcpyDeviceToHost, sl.st));
        }
        GG_CUDA(cudaEventRecord(sl.ev_done, sl.st));
        return GG_OK;
    };
    const bool fused_mode = idx->host_mode == 0;

    // software pipeline: A(0) A(1) B(0) A(2) B(1) ...
    const size_t lead = std::min<size_t>(kSlots - 1, n_chunks);
    for (size_t k = 0; k < lead; ++k) {
        s = fused_mode ? fstage_a(k) : stage_a(k);
        if (s != GG_OK) return s;
    }
    for (size_t k = 0; k < n_chunks; ++k) {
        if (k + lead < n_chunks) {
            // the slot of chunk k+lead was last used by chunk k+lead-kSlots = k-1: done by now or waited in stage_a
            s = fused_mode ? fstage_a(k + lead) : stage_a(k + lead);
            if (s != GG_OK) return s;
        }
        s = fused_mode ? fstage_b(k) : stage_b(k);
        if (s != GG_OK) return s;
    }
    for (int i = 0; i < kSlots; ++i) {
        GG_CUDA(cudaStreamSynchronize(idx->slots[i].st));
        idx->slots[i].busy = false;
    }
    drain.armed = false;
    s = check_query_error(idx, st0);
    if (s != GG_OK) return s;
    if (any_invalid) return fail(idx, GG_ERR_INVALID_INTERVAL, "query interval: start must be <= end");
    if (!idx->pinned_hits) {
        s = ensure_pinned_hits(idx, 1, 0);
        if (s != GG_OK) return s;
    }
    *hits = idx->pinned_hits;
    *n_hits = base[n_chunks];
    idx->bstats.n_hits = base[n_chunks];
    return GG_OK;
}

// ------------------------------------------------------------------------------------------ compact host path
// count (records) -> block-total scan -> intersect_emit_tma_kernel writing {chunk-local offset, hi slot} per query and a
// back-delta of OutT per hit (emit_tma.cuh).  u32 index layout, u32 queries.
template <typename OutT>
gg_status launch_batch_compact(gg_index* idx, const uint32_t* d_chrom, const uint32_t* d_start, const uint32_t* d_end,
                               size_t nq, CompactQuery* d_qres, OutT* d_delta, uint64_t capacity, uint64_t* d_total,
                               void* scratch, uint32_t* d_err, cudaStream_t st) {
    using C = uint32_t;
    using Q = uint32_t;
    char* p = static_cast<char*>(scratch);
    const size_t nb = batch_blocks(nq), nb1 = nb + 1;
    unsigned long long* block_total = reinterpret_cast<unsigned long long*>(p);
    uint64_t* block_prefix = reinterpret_cast<uint64_t*>(p + round_up(nb1 * 8, 256));
    void* scan_state = p + round_up(nb1 * 8, 256) + round_up((nb1 + 1) * 8, 256);
    char* tail = p + std::max(round_up(fused_scratch_bytes(nq), 256), batch_ctl_bytes(nq));
    uint32_t* d_rec = reinterpret_cast<uint32_t*>(tail);  // uint2[nq] over the cnt + hi regions of the scratch
    if (nq) {
        auto kern = intersect_count_group_kernel<C, Q, kBatchQpt, false, true>;
        int per_sm = 0;
        GG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kQueryThreads, 0));
        if (per_sm < 1) return fail(idx, GG_ERR_CUDA, "count kernel does not fit on an SM");
        const size_t tiles = div_up(nq, static_cast<size_t>(kQueryThreads) * kBatchQpt);
        const unsigned grid = static_cast<unsigned>(std::min<size_t>(tiles, static_cast<size_t>(idx->sm_count) * per_sm));
        kern<<<grid, kQueryThreads, 0, st>>>(make_view<C>(idx), d_chrom, d_start, d_end, nq, d_rec, nullptr, nullptr, block_total,
                                             d_err, 1, nullptr);
        GG_CUDA(cudaGetLastError());
        idx->bstats.kernel_launches++;
    }
    GG_CUDA((exclusive_scan_async<uint64_t, uint64_t>(reinterpret_cast<const uint64_t*>(block_total), block_prefix, nb, scan_state,
                                                       nullptr, nullptr, st)));
    idx->bstats.kernel_launches++;
    const size_t smem = emit_tma_smem_bytes<Q>(0);
    auto kern = intersect_emit_tma_kernel<C, Q, OutT, true, false>;
    GG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    int per_sm = 0;
    GG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kQueryThreads, smem));
    if (per_sm < 1) return fail(idx, GG_ERR_CUDA, "emit kernel does not fit on an SM");
    const size_t ctas = div_up(std::max<size_t>(nb, 1), kQueryThreads / 32);
    const unsigned grid = static_cast<unsigned>(std::min<size_t>(ctas, static_cast<size_t>(idx->sm_count) * per_sm));
    kern<<<grid, kQueryThreads, smem, st>>>(make_flat<C>(idx), reinterpret_cast<const uint2*>(d_rec), nullptr, d_start, nq, block_prefix,
                                            static_cast<uint32_t>(kBatchBlockQ), nullptr, d_qres, d_delta, capacity, nullptr, d_total,
                                            d_err, 0);
    GG_CUDA(cudaGetLastError());
    idx->bstats.kernel_launches++;
    return GG_OK;
}

gg_status dispatch_batch_compact(gg_index* idx, int width, const uint32_t* d_chrom, const uint32_t* d_start,
                                 const uint32_t* d_end, size_t nq, CompactQuery* d_qres, void* d_delta, uint64_t capacity,
                                 uint64_t* d_total, void* scratch, uint32_t* d_err, cudaStream_t st) {
    if (width == 1) return launch_batch_compact<uint8_t>(idx, d_chrom, d_start, d_end, nq, d_qres, static_cast<uint8_t*>(d_delta), capacity, d_total, scratch, d_err, st);
    if (width == 2) return launch_batch_compact<uint16_t>(idx, d_chrom, d_start, d_end, nq, d_qres, static_cast<uint16_t*>(d_delta), capacity, d_total, scratch, d_err, st);
    return launch_batch_compact<uint32_t>(idx, d_chrom, d_start, d_end, nq, d_qres, static_cast<uint32_t*>(d_delta), capacity, d_total, scratch, d_err, st);
}

template <typename T>
gg_status ensure_pinned(gg_index* idx, T*& ptr, size_t& cap, size_t need, size_t keep, const char* what) {
    if (need <= cap && ptr) return GG_OK;
    const size_t want = std::max<size_t>(need, 4096);
    T* np = nullptr;
    cudaError_t e = cudaHostAlloc(reinterpret_cast<void**>(&np), want * sizeof(T), cudaHostAllocDefault);
    if (e != cudaSuccess) return fail(idx, GG_ERR_OOM, what, e);
    if (ptr) {
        if (keep) std::memcpy(np, ptr, keep * sizeof(T));
        cudaFreeHost(ptr);
    }
    ptr = np;
    cap = want;
    return GG_OK;
}

// Chunked pipeline over the kSlots streams, like intersect_batch_host, for the compact result form: per chunk H2D of 12 B
// per query, the three kernels, D2H of 8 B per query + delta_bytes per hit.  A chunk is launched with an estimated delta
// capacity and the handle's current delta width; overflow of either re-runs the chunk (exact size / next width).
gg_status intersect_batch_compact_host(gg_index* idx, const uint32_t* q_chrom, const uint32_t* q_start, const uint32_t* q_end,
                                       size_t nq, gg_compact_result* out) {
    if (!idx) return GG_ERR_INVALID_ARG;
    idx->err.clear();
    if (!idx->built) return fail(idx, GG_ERR_NOT_BUILT, "index not built");
    if (!out || (nq && (!q_chrom || !q_start || !q_end))) return fail(idx, GG_ERR_INVALID_ARG, "null argument");
    if (idx->coord_bytes != 4) return fail(idx, GG_ERR_INVALID_ARG, "the compact call needs a u32-layout index");
    GG_CUDA(cudaSetDevice(idx->device));
    gg_status s = ensure_slots(idx);
    if (s != GG_OK) return s;
    idx->h2_valid = false;
    idx->bstats = gg_batch_stats{};
    idx->bstats.n_queries = nq;
    // chunks stay below 2^32 hits (chunk-local u32 offsets) for any realistic density; a chunk that does not is reported
    const size_t chunk = static_cast<size_t>(std::max<int64_t>(idx->chunk_queries, kScanTile) / kScanTile * kScanTile);
    const size_t n_chunks = nq ? div_up(nq, chunk) : 0;
    idx->bstats.n_chunks = static_cast<uint32_t>(n_chunks);
    idx->cchunks.assign(n_chunks, gg_compact_chunk{});
    *out = gg_compact_result{};
    out->chunk_queries = chunk;
    out->n_chrom = idx->n_chrom;
    out->chrom_slot_base = idx->h_cslot.data();
    s = ensure_pinned(idx, idx->pinned_cq, idx->pinned_cq_cap, std::max<size_t>(nq, 1), 0, "pinned query records");
    if (s != GG_OK) return s;
    if (!idx->pinned_delta) {
        s = ensure_pinned(idx, idx->pinned_delta, idx->pinned_delta_cap, 4096, 0, "pinned delta stream");
        if (s != GG_OK) return s;
    }
    out->queries = reinterpret_cast<const gg_compact_query*>(idx->pinned_cq);
    out->deltas = idx->pinned_delta;
    if (nq == 0) return GG_OK;
    for (int i = 0; i < kSlots; ++i) idx->slots[i].busy = false;
    SlotDrain drain{idx};

    bool any_invalid = false;
    double est_per_q = std::max(8.0, idx->seen_hits_per_query * 1.1);
    uint64_t hit_base = 0, byte_off = 0;
    std::vector<int> width_of(n_chunks, 0);
    auto stage_a = [&](size_t k) -> gg_status {
        Slot& sl = idx->slots[k % kSlots];
        if (sl.busy) {
            GG_CUDA(cudaEventSynchronize(sl.ev_done));
            sl.busy = false;
        }
        const size_t q0 = k * chunk, m = std::min(chunk, nq - q0);
        GG_CUDA(sl.qc.ensure(m * 4));
        GG_CUDA(sl.qs.ensure(m * 4));
        GG_CUDA(sl.qe.ensure(m * 4));
        GG_CUDA(sl.off.ensure(m * 8));  // CompactQuery records
        GG_CUDA(sl.fscratch.ensure(single_scratch_bytes(m)));
        GG_CUDA(sl.err.ensure(16));
        const int width = idx->compact_delta_bytes;
        width_of[k] = width;
        const size_t want = (static_cast<size_t>(est_per_q * 1.25 * static_cast<double>(m)) + 4096) * width;
        if (sl.hits.cap < want) GG_CUDA(sl.hits.ensure(want));
        sl.launch_cap = sl.hits.cap / width;
        GG_CUDA(cudaMemcpyAsync(sl.qc.p, q_chrom + q0, m * 4, cudaMemcpyHostToDevice, sl.st));
        GG_CUDA(cudaMemcpyAsync(sl.qs.p, q_start + q0, m * 4, cudaMemcpyHostToDevice, sl.st));
        GG_CUDA(cudaMemcpyAsync(sl.qe.p, q_end + q0, m * 4, cudaMemcpyHostToDevice, sl.st));
        GG_CUDA(cudaMemsetAsync(sl.err.p, 0, 16, sl.st));
        uint64_t* d_total = reinterpret_cast<uint64_t*>(sl.err.as<char>() + 8);
        gg_status r = dispatch_batch_compact(idx, width, sl.qc.as<uint32_t>(), sl.qs.as<uint32_t>(), sl.qe.as<uint32_t>(), m,
                                             sl.off.as<CompactQuery>(), sl.hits.p, sl.launch_cap, d_total, sl.fscratch.p,
                                             sl.err.as<uint32_t>(), sl.st);
        if (r != GG_OK) return r;
        GG_CUDA(cudaMemcpyAsync(sl.h_next_base, sl.err.p, 16, cudaMemcpyDeviceToHost, sl.st));  // [flag, -, total]
        GG_CUDA(cudaEventRecord(sl.ev_total, sl.st));
        GG_CUDA(cudaMemcpyAsync(idx->pinned_cq + q0, sl.off.p, m * 8, cudaMemcpyDeviceToHost, sl.st));
        sl.busy = true;
        return GG_OK;
    };
    auto stage_b = [&](size_t k) -> gg_status {
        Slot& sl = idx->slots[k % kSlots];
        const size_t q0 = k * chunk, m = std::min(chunk, nq - q0);
        for (;;) {
            GG_CUDA(cudaEventSynchronize(sl.ev_total));
            const uint32_t flag = static_cast<uint32_t>(sl.h_next_base[0]);
            const uint64_t total = sl.h_next_base[1];
            if (flag & 1u) any_invalid = true;
            if (flag & 8u) return fail(idx, GG_ERR_CAPACITY, "a chunk of the compact call has 2^32 hits or more: lower chunk_queries");
            int width = width_of[k];
            const bool too_narrow = (flag & 16u) != 0 && width < 4, too_small = total > sl.launch_cap;
            if (!too_narrow && !too_small) {
                est_per_q = std::max(1.0, static_cast<double>(hit_base + total) / static_cast<double>(q0 + m));
                note_density(idx, hit_base + total, q0 + m);
                gg_compact_chunk& ck = idx->cchunks[k];
                ck.q_begin = q0;
                ck.n_queries = m;
                ck.n_hits = total;
                ck.hit_base = hit_base;
                ck.delta_byte_offset = byte_off;
                ck.delta_bytes = static_cast<uint32_t>(width);
                const uint64_t bytes = total * width;
                if (byte_off + bytes > idx->pinned_delta_cap) {
                    // grow with an estimate of the final size; copies of earlier chunks must have landed first
                    for (int i = 0; i < kSlots; ++i)
                        if (idx->slots[i].busy) GG_CUDA(cudaStreamSynchronize(idx->slots[i].st));
                    const double per_q = static_cast<double>(byte_off + bytes) / static_cast<double>(q0 + m);
                    const size_t est = static_cast<size_t>(per_q * 1.15 * static_cast<double>(nq)) + 65536;
                    gg_status r = ensure_pinned(idx, idx->pinned_delta, idx->pinned_delta_cap, std::max<size_t>(est, byte_off + bytes),
                                                byte_off, "pinned delta stream");
                    if (r != GG_OK) return r;
                }
                if (bytes) GG_CUDA(cudaMemcpyAsync(idx->pinned_delta + byte_off, sl.hits.p, bytes, cudaMemcpyDeviceToHost, sl.st));
                hit_base += total;
                byte_off = round_up(byte_off + bytes, 16);
                GG_CUDA(cudaEventRecord(sl.ev_done, sl.st));
                return GG_OK;
            }
            // re-run this chunk: one width up and / or with the exact capacity (same queries, already on the device)
            GG_CUDA(cudaStreamSynchronize(sl.st));
            if (too_narrow) {
                width *= 2;
                idx->compact_delta_bytes = std::max(idx->compact_delta_bytes, width);
                width_of[k] = width;
            }
            const size_t want = std::max<size_t>(static_cast<size_t>(total), 1) * width;
            if (sl.hits.cap < want) GG_CUDA(sl.hits.ensure(want));
            sl.launch_cap = sl.hits.cap / width;
            GG_CUDA(cudaMemsetAsync(sl.err.p, 0, 16, sl.st));
            uint64_t* d_total = reinterpret_cast<uint64_t*>(sl.err.as<char>() + 8);
            gg_status r = dispatch_batch_compact(idx, width, sl.qc.as<uint32_t>(), sl.qs.as<uint32_t>(), sl.qe.as<uint32_t>(), m,
                                                 sl.off.as<CompactQuery>(), sl.hits.p, sl.launch_cap, d_total, sl.fscratch.p,
                                                 sl.err.as<uint32_t>(), sl.st);
            if (r != GG_OK) return r;
            GG_CUDA(cudaMemcpyAsync(sl.h_next_base, sl.err.p, 16, cudaMemcpyDeviceToHost, sl.st));
            GG_CUDA(cudaEventRecord(sl.ev_total, sl.st));
            GG_CUDA(cudaMemcpyAsync(idx->pinned_cq + q0, sl.off.p, m * 8, cudaMemcpyDeviceToHost, sl.st));
        }
    };
    const size_t lead = std::min<size_t>(kSlots - 1, n_chunks);
    for (size_t k = 0; k < lead; ++k) {
        s = stage_a(k);
        if (s != GG_OK) return s;
    }
    for (size_t k = 0; k < n_chunks; ++k) {
        if (k + lead < n_chunks) {
            s = stage_a(k + lead);
            if (s != GG_OK) return s;
        }
        s = stage_b(k);
        if (s != GG_OK) return s;
    }
    for (int i = 0; i < kSlots; ++i) {
        GG_CUDA(cudaStreamSynchronize(idx->slots[i].st));
        idx->slots[i].busy = false;
    }
    drain.armed = false;
    if (any_invalid) return fail(idx, GG_ERR_INVALID_INTERVAL, "query interval: start must be <= end");
    out->n_queries = nq;
    out->n_hits = hit_base;
    out->queries = reinterpret_cast<const gg_compact_query*>(idx->pinned_cq);
    out->deltas = idx->pinned_delta;
    out->delta_bytes_total = byte_off;
    out->chunks = idx->cchunks.data();
    out->n_chunks = static_cast<uint32_t>(n_chunks);
    idx->bstats.n_hits = hit_base;
    return GG_OK;
}

}  // namespace

// ============================================================================================== C ABI

extern "C" {

int gg_version(void) { return GG_B200_VERSION; }

int gg_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

gg_status gg_index_create(gg_index** out, int device) {
    if (!out) return GG_ERR_INVALID_ARG;
    *out = nullptr;
    int n = gg_device_count();
    if (n <= 0 || device < 0 || device >= n) return GG_ERR_CUDA;
    gg_index* idx = new (std::nothrow) gg_index();
    if (!idx) return GG_ERR_OOM;
    idx->device = device;
    cudaDeviceProp prop;
    if (cudaSetDevice(device) != cudaSuccess || cudaGetDeviceProperties(&prop, device) != cudaSuccess ||
        cudaStreamCreateWithFlags(&idx->stream, cudaStreamNonBlocking) != cudaSuccess) {
        delete idx;
        cudaGetLastError();
        return GG_ERR_CUDA;
    }
    idx->sm_count = prop.multiProcessorCount;
    for (auto& e : idx->ev) cudaEventCreate(&e);
    *out = idx;
    return GG_OK;
}

void gg_index_destroy(gg_index* idx) {
    if (!idx) return;
    cudaSetDevice(idx->device);
    cudaDeviceSynchronize();
    for (DeviceBuf* b : {&idx->S, &idx->E, &idx->M, &idx->tab, &idx->cnt, &idx->dirS, &idx->dirE, &idx->dir_off, &idx->fused_scratch, &idx->fused_out, &idx->q_cnt, &idx->q_hi, &idx->scan_scratch,
                         &idx->err_flag, &idx->h2_qc, &idx->h2_qs, &idx->h2_qe, &idx->h2_off, &idx->h2_hits, &idx->d_base,
                         &idx->ER, &idx->cbegin, &idx->cslot, &idx->flat[0], &idx->flat[1], &idx->flat[2], &idx->flat[3], &idx->m8, &idx->strand, &idx->SN, &idx->EN, &idx->sq_qc, &idx->sq_qs, &idx->sq_qe, &idx->sq_str, &idx->sq_off, &idx->sq_hits,
                         &idx->s_cnt, &idx->s_hi, &idx->s_tile, &idx->s_scan, &idx->s_total})
        b->release();
    for (auto& s : idx->slots) {
        for (DeviceBuf* b : {&s.qc, &s.qs, &s.qe, &s.cnt, &s.hi, &s.off, &s.hits, &s.scan, &s.fscratch, &s.err}) b->release();
        if (s.st) cudaStreamDestroy(s.st);
        if (s.ev_scan) cudaEventDestroy(s.ev_scan);
        if (s.ev_total) cudaEventDestroy(s.ev_total);
        if (s.ev_done) cudaEventDestroy(s.ev_done);
        if (s.h_next_base) cudaFreeHost(s.h_next_base);
    }
    if (idx->pinned_hits) cudaFreeHost(idx->pinned_hits);
    if (idx->pinned_cq) cudaFreeHost(idx->pinned_cq);
    if (idx->pinned_delta) cudaFreeHost(idx->pinned_delta);
    if (idx->h_err) cudaFreeHost(idx->h_err);
    if (idx->h_lazy) cudaFreeHost(idx->h_lazy);
    if (idx->h_stage) cudaFreeHost(idx->h_stage);
    for (auto& ev : idx->ev_stage)
        if (ev) cudaEventDestroy(ev);
    for (auto* b : idx->bw.all()) b->release();
    if (idx->ev_lazy) cudaEventDestroy(idx->ev_lazy);
    for (auto& e : idx->ev)
        if (e) cudaEventDestroy(e);
    if (idx->stream) cudaStreamDestroy(idx->stream);
    delete idx;
}

const char* gg_last_error(const gg_index* idx) { return idx ? idx->err.c_str() : "null handle"; }

gg_status gg_index_build(gg_index* idx, const uint32_t* chrom, const uint64_t* start, const uint64_t* end,
                         const uint64_t* tie_key, int tie_mode, size_t n, uint32_t n_chrom, uint32_t* rank_to_input) {
    return build_impl(idx, chrom, start, end, tie_key, tie_mode, n, n_chrom, rank_to_input, true);
}

gg_status gg_index_build_device(gg_index* idx, const uint32_t* d_chrom, const uint64_t* d_start, const uint64_t* d_end,
                                const uint64_t* d_tie_key, int tie_mode, size_t n, uint32_t n_chrom,
                                uint32_t* d_rank_to_input) {
    return build_impl(idx, d_chrom, d_start, d_end, d_tie_key, tie_mode, n, n_chrom, d_rank_to_input, true, true);
}

gg_status gg_index_upload_sorted(gg_index* idx, const uint32_t* chrom, const uint64_t* start, const uint64_t* end,
                                 size_t n, uint32_t n_chrom) {
    return build_impl(idx, chrom, start, end, nullptr, GG_TIE_INPUT_ORDER, n, n_chrom, nullptr, false);
}

gg_status gg_index_update_chrom(gg_index* idx, uint32_t chrom, const uint64_t* start, const uint64_t* end, size_t n) {
    return update_chrom_impl(idx, chrom, start, end, n);
}

gg_status gg_index_get_stats(const gg_index* idx, gg_index_stats* out) {
    if (!idx || !out) return GG_ERR_INVALID_ARG;
    if (!idx->built) return GG_ERR_NOT_BUILT;
    *out = idx->istats;
    return GG_OK;
}

gg_status gg_index_chrom_counts(const gg_index* idx, uint64_t* counts) {
    if (!idx || !counts) return GG_ERR_INVALID_ARG;
    if (!idx->built) return GG_ERR_NOT_BUILT;
    for (uint32_t c = 0; c < idx->n_chrom; ++c) counts[c] = idx->h_cnt[c];
    return GG_OK;
}

gg_status gg_intersect_batch(gg_index* idx, const uint32_t* q_chrom, const uint64_t* q_start, const uint64_t* q_end,
                             size_t nq, uint64_t* offsets, const uint32_t** hits, uint64_t* n_hits) {
    return intersect_batch_host<uint64_t>(idx, q_chrom, q_start, q_end, nq, offsets, hits, n_hits);
}

gg_status gg_intersect_batch_u32(gg_index* idx, const uint32_t* q_chrom, const uint32_t* q_start,
                                 const uint32_t* q_end, size_t nq, uint64_t* offsets, const uint32_t** hits,
                                 uint64_t* n_hits) {
    return intersect_batch_host<uint32_t>(idx, q_chrom, q_start, q_end, nq, offsets, hits, n_hits);
}

gg_status gg_intersect_batch_compact(gg_index* idx, const uint32_t* q_chrom, const uint32_t* q_start, const uint32_t* q_end,
                                     size_t nq, gg_compact_result* out) {
    return intersect_batch_compact_host(idx, q_chrom, q_start, q_end, nq, out);
}

uint64_t gg_compact_count(const gg_compact_result* r, uint64_t q) {
    if (!r || q >= r->n_queries) return 0;
    const gg_compact_chunk& ck = r->chunks[q / r->chunk_queries];
    const uint64_t end = q + 1 < ck.q_begin + ck.n_queries ? r->queries[q + 1].off : ck.n_hits;
    return end - r->queries[q].off;
}

uint32_t gg_compact_rank(const gg_compact_result* r, const uint32_t* q_chrom, uint64_t q, uint64_t j) {
    const gg_compact_chunk& ck = r->chunks[q / r->chunk_queries];
    const gg_compact_query& cq = r->queries[q];
    const uint8_t* d = r->deltas + ck.delta_byte_offset + (static_cast<uint64_t>(cq.off) + j) * ck.delta_bytes;
    uint32_t delta;
    if (ck.delta_bytes == 1) delta = *d;
    else if (ck.delta_bytes == 2) { uint16_t v; std::memcpy(&v, d, 2); delta = v; }
    else std::memcpy(&delta, d, 4);
    return cq.hi_slot - r->chrom_slot_base[q_chrom[q]] - 1u - delta;
}

gg_status gg_compact_expand(const gg_compact_result* r, const uint32_t* q_chrom, uint64_t* offsets, uint32_t* hits,
                            int n_threads) {
    if (!r || !offsets || (r->n_queries && !q_chrom) || (r->n_hits && !hits)) return GG_ERR_INVALID_ARG;
    unsigned T = n_threads > 0 ? static_cast<unsigned>(n_threads) : std::max(1u, std::thread::hardware_concurrency());
    T = std::max(1u, std::min<unsigned>(T, std::max<uint32_t>(r->n_chunks, 1u)));
    auto work = [&](unsigned t) {
        for (uint32_t k = t; k < r->n_chunks; k += T) {
            const gg_compact_chunk& ck = r->chunks[k];
            const uint8_t* d = r->deltas + ck.delta_byte_offset;
            for (uint64_t i = 0; i < ck.n_queries; ++i) {
                const uint64_t q = ck.q_begin + i;
                const gg_compact_query& cq = r->queries[q];
                const uint64_t end = i + 1 < ck.n_queries ? r->queries[q + 1].off : ck.n_hits;
                offsets[q] = ck.hit_base + cq.off;
                if (end == cq.off) continue;
                const uint32_t top = cq.hi_slot - r->chrom_slot_base[q_chrom[q]] - 1u;
                uint32_t* o = hits + ck.hit_base;
                if (ck.delta_bytes == 1) {
                    for (uint64_t h = cq.off; h < end; ++h) o[h] = top - d[h];
                } else if (ck.delta_bytes == 2) {
                    const uint16_t* d2 = reinterpret_cast<const uint16_t*>(d);
                    for (uint64_t h = cq.off; h < end; ++h) o[h] = top - d2[h];
                } else {
                    const uint32_t* d4 = reinterpret_cast<const uint32_t*>(d);
                    for (uint64_t h = cq.off; h < end; ++h) o[h] = top - d4[h];
                }
            }
        }
    };
    if (T == 1) {
        work(0);
    } else {
        std::vector<std::thread> th;
        for (unsigned t = 0; t < T; ++t) th.emplace_back(work, t);
        for (auto& x : th) x.join();
    }
    offsets[r->n_queries] = r->n_hits;
    return GG_OK;
}

gg_status gg_intersect_count_device(gg_index* idx, const uint32_t* d_chrom, const void* d_start, const void* d_end,
                                    int coord_bytes, size_t nq, uint64_t* d_offsets, uint64_t* total_hits,
                                    void* stream) {
    if (!idx) return GG_ERR_INVALID_ARG;
    idx->err.clear();
    if (!idx->built) return fail(idx, GG_ERR_NOT_BUILT, "index not built");
    if (!d_offsets || (nq && (!d_chrom || !d_start || !d_end))) return fail(idx, GG_ERR_INVALID_ARG, "null argument");
    if (coord_bytes != 4 && coord_bytes != 8) return fail(idx, GG_ERR_INVALID_ARG, "coord_bytes must be 4 or 8");
    if (coord_bytes == 4 && idx->coord_bytes != 4) return fail(idx, GG_ERR_INVALID_ARG, "u32 queries need a u32-layout index");
    GG_CUDA(cudaSetDevice(idx->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    gg_status s = ensure_query_scratch(idx, nq);
    if (s != GG_OK) return s;
    idx->bstats = gg_batch_stats{};
    idx->bstats.n_queries = nq;
    idx->bstats.n_chunks = 1;
    idx->ev_count_valid = idx->ev_emit_valid = false;
    GG_CUDA(cudaMemsetAsync(idx->err_flag.p, 0, 4, st));
    if (idx->timing) GG_CUDA(cudaEventRecord(idx->ev[0], st));
    s = dispatch_count(idx, d_chrom, d_start, d_end, coord_bytes, nq, d_offsets, idx->q_cnt.as<uint32_t>(),
                       idx->q_hi.as<uint32_t>(), idx->err_flag.as<uint32_t>(), idx->scan_scratch.p, nullptr, st,
                       idx->timing ? idx->ev[1] : nullptr);
    if (s != GG_OK) return s;
    if (idx->timing) {
        GG_CUDA(cudaEventRecord(idx->ev[2], st));
        idx->ev_count_valid = true;
    }
    idx->counted_nq = nq;
    if (total_hits) {  // both words land in pinned memory (no staging copies), one synchronisation
        if (!idx->h_err) GG_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&idx->h_err), 16, cudaHostAllocDefault));
        GG_CUDA(cudaMemcpyAsync(idx->h_err + 2, d_offsets + nq, 8, cudaMemcpyDeviceToHost, st));
        GG_CUDA(cudaMemcpyAsync(idx->h_err, idx->err_flag.p, 4, cudaMemcpyDeviceToHost, st));
        GG_CUDA(cudaStreamSynchronize(st));
        std::memcpy(total_hits, idx->h_err + 2, 8);
        if (idx->h_err[0] & 1u) return fail(idx, GG_ERR_INVALID_INTERVAL, "query interval: start must be <= end");
        idx->bstats.n_hits = *total_hits;
    }
    return GG_OK;
}

gg_status gg_intersect_emit_device(gg_index* idx, const uint32_t* d_chrom, const void* d_start, const void* d_end,
                                   int coord_bytes, size_t nq, const uint64_t* d_offsets, uint32_t* d_hits,
                                   void* stream) {
    (void)d_end;
    if (!idx) return GG_ERR_INVALID_ARG;
    idx->err.clear();
    if (!idx->built) return fail(idx, GG_ERR_NOT_BUILT, "index not built");
    if (nq != idx->counted_nq) return fail(idx, GG_ERR_NOT_BUILT, "emit must follow count on the same batch");
    if (!d_offsets || (nq && (!d_chrom || !d_start))) return fail(idx, GG_ERR_INVALID_ARG, "null argument");
    if (coord_bytes != 4 && coord_bytes != 8) return fail(idx, GG_ERR_INVALID_ARG, "coord_bytes must be 4 or 8");
    if (coord_bytes == 4 && idx->coord_bytes != 4) return fail(idx, GG_ERR_INVALID_ARG, "u32 queries need a u32-layout index");
    GG_CUDA(cudaSetDevice(idx->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (idx->timing) GG_CUDA(cudaEventRecord(idx->ev[3], st));
    // (the offsets went through the caller's hands: the kernel checks them against the counts kept on the handle)
    gg_status s = dispatch_emit(idx, d_chrom, d_start, coord_bytes, nq, idx->q_hi.as<uint32_t>(), d_offsets, d_hits, 0, st,
                                idx->q_cnt.as<uint32_t>());
    if (s != GG_OK) return s;
    if (idx->timing) {
        GG_CUDA(cudaEventRecord(idx->ev[4], st));
        idx->ev_emit_valid = true;
    }
    if (idx->debug_checks) return check_query_error(idx, st);  // synchronises: a debugging aid, the call is asynchronous otherwise
    return GG_OK;
}

gg_status gg_intersect_fused_device(gg_index* idx, const uint32_t* d_chrom, const void* d_start, const void* d_end,
                                    int coord_bytes, size_t nq, uint64_t* d_offsets, uint32_t* d_hits,
                                    uint64_t capacity, uint64_t* total_hits, void* stream) {
    if (!idx) return GG_ERR_INVALID_ARG;
    idx->err.clear();
    if (!idx->built) return fail(idx, GG_ERR_NOT_BUILT, "index not built");
    if (!d_offsets || (capacity && !d_hits) || (nq && (!d_chrom || !d_start || !d_end)))
        return fail(idx, GG_ERR_INVALID_ARG, "null argument");
    if (coord_bytes != 4 && coord_bytes != 8) return fail(idx, GG_ERR_INVALID_ARG, "coord_bytes must be 4 or 8");
    if (coord_bytes == 4 && idx->coord_bytes != 4) return fail(idx, GG_ERR_INVALID_ARG, "u32 queries need a u32-layout index");
    GG_CUDA(cudaSetDevice(idx->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    GG_CUDA(idx->fused_scratch.ensure(single_scratch_bytes(nq)));
    GG_CUDA(idx->err_flag.ensure(16));  // one 16-byte mailbox: [0, 4) error flag, [8, 16) hit total — one memset, one read
    if (!idx->h_err) GG_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&idx->h_err), 16, cudaHostAllocDefault));
    uint64_t* d_total = reinterpret_cast<uint64_t*>(idx->err_flag.as<char>() + 8);
    idx->bstats = gg_batch_stats{};
    idx->bstats.n_queries = nq;
    idx->bstats.n_chunks = 1;
    idx->ev_count_valid = idx->ev_emit_valid = false;
    idx->counted_nq = 0;  // the two-phase emit needs its own count pass
    poll_lazy_density(idx);
    GG_CUDA(cudaMemsetAsync(idx->err_flag.p, 0, 16, st));
    if (idx->timing) {
        GG_CUDA(cudaEventRecord(idx->ev[0], st));
        GG_CUDA(cudaEventRecord(idx->ev[3], st));
    }
    idx->in_device_call = true;
    gg_status s = dispatch_fused(idx, d_chrom, d_start, d_end, coord_bytes, nq, d_offsets, d_hits, capacity, nullptr, nullptr,
                                 d_total, idx->fused_scratch.p, idx->err_flag.as<uint32_t>(), st);
    idx->in_device_call = false;
    if (s != GG_OK) return s;
    if (idx->timing) {
        GG_CUDA(cudaEventRecord(idx->ev[4], st));
        idx->ev_emit_valid = true;
        idx->ev_count_valid = !idx->last_one_kernel;  // count / block-scan / emit_scan times of the three-kernel form
    }
    if (total_hits) {
        GG_CUDA(cudaMemcpyAsync(idx->h_err, idx->err_flag.p, 16, cudaMemcpyDeviceToHost, st));  // pinned: no staging copy
        GG_CUDA(cudaStreamSynchronize(st));
        const uint32_t flag = idx->h_err[0];
        std::memcpy(total_hits, idx->h_err + 2, 8);
        idx->bstats.n_hits = *total_hits;
        note_density(idx, *total_hits, nq);
        if (flag & 1u) return fail(idx, GG_ERR_INVALID_INTERVAL, "query interval: start must be <= end");
        if (flag & 2u) return fail(idx, GG_ERR_CAPACITY, "hit buffer smaller than offsets[nq]; offsets are complete");
    } else {
        GG_CUDA(post_lazy_density(idx, d_total, nq, false, st));
    }
    return GG_OK;
}

gg_status gg_intersect_count(gg_index* idx, const uint32_t* q_chrom, const uint64_t* q_start, const uint64_t* q_end,
                             size_t nq, uint64_t* offsets) {
    if (!idx) return GG_ERR_INVALID_ARG;
    idx->err.clear();
    idx->h2_valid = false;
    if (!idx->built) return fail(idx, GG_ERR_NOT_BUILT, "index not built");
    if (!offsets || (nq && (!q_chrom || !q_start || !q_end))) return fail(idx, GG_ERR_INVALID_ARG, "null argument");
    GG_CUDA(cudaSetDevice(idx->device));
    cudaStream_t st = idx->stream;
    const size_t m = nq ? nq : 1;
    GG_CUDA(idx->h2_qc.ensure(m * 4));
    GG_CUDA(idx->h2_qs.ensure(m * 8));
    GG_CUDA(idx->h2_qe.ensure(m * 8));
    GG_CUDA(idx->h2_off.ensure((m + 1) * 8));
    if (nq) {
        GG_CUDA(cudaMemcpyAsync(idx->h2_qc.p, q_chrom, nq * 4, cudaMemcpyHostToDevice, st));
        GG_CUDA(cudaMemcpyAsync(idx->h2_qs.p, q_start, nq * 8, cudaMemcpyHostToDevice, st));
        GG_CUDA(cudaMemcpyAsync(idx->h2_qe.p, q_end, nq * 8, cudaMemcpyHostToDevice, st));
    }
    uint64_t total = 0;
    gg_status s = gg_intersect_count_device(idx, idx->h2_qc.as<uint32_t>(), idx->h2_qs.p, idx->h2_qe.p, 8, nq,
                                            idx->h2_off.as<uint64_t>(), &total, st);
    if (s != GG_OK) return s;
    GG_CUDA(cudaMemcpyAsync(offsets, idx->h2_off.p, (nq + 1) * 8, cudaMemcpyDeviceToHost, st));
    GG_CUDA(cudaStreamSynchronize(st));
    idx->h2_nq = nq;
    idx->h2_total = total;
    idx->h2_valid = true;
    return GG_OK;
}

gg_status gg_intersect_emit(gg_index* idx, uint32_t* hits, uint64_t capacity) {
    if (!idx) return GG_ERR_INVALID_ARG;
    idx->err.clear();
    if (!idx->built || !idx->h2_valid) return fail(idx, GG_ERR_NOT_BUILT, "gg_intersect_emit must follow gg_intersect_count");
    if (capacity < idx->h2_total) return fail(idx, GG_ERR_CAPACITY, "hit buffer smaller than offsets[nq]");
    if (idx->h2_total == 0) return GG_OK;
    if (!hits) return fail(idx, GG_ERR_INVALID_ARG, "null hit buffer");
    GG_CUDA(cudaSetDevice(idx->device));
    cudaStream_t st = idx->stream;
    GG_CUDA(idx->h2_hits.ensure(idx->h2_total * 4));
    gg_status s = gg_intersect_emit_device(idx, idx->h2_qc.as<uint32_t>(), idx->h2_qs.p, idx->h2_qe.p, 8, idx->h2_nq,
                                           idx->h2_off.as<uint64_t>(), idx->h2_hits.as<uint32_t>(), st);
    if (s != GG_OK) return s;
    GG_CUDA(cudaMemcpyAsync(hits, idx->h2_hits.p, idx->h2_total * 4, cudaMemcpyDeviceToHost, st));
    GG_CUDA(cudaStreamSynchronize(st));
    return GG_OK;
}

gg_status gg_last_batch_stats(const gg_index* cidx, gg_batch_stats* out) {
    if (!cidx || !out) return GG_ERR_INVALID_ARG;
    gg_index* idx = const_cast<gg_index*>(cidx);
    if (idx->timing && idx->ev_count_valid) {
        float a = 0, b = 0;
        if (cudaEventSynchronize(idx->ev[2]) == cudaSuccess) {
            cudaEventElapsedTime(&a, idx->ev[0], idx->ev[1]);
            cudaEventElapsedTime(&b, idx->ev[1], idx->ev[2]);
            idx->bstats.count_ms = a;
            idx->bstats.scan_ms = b;
        }
    }
    if (idx->timing && idx->ev_emit_valid) {
        float c = 0;
        if (cudaEventSynchronize(idx->ev[4]) == cudaSuccess) {
            cudaEventElapsedTime(&c, idx->ev[3], idx->ev[4]);
            idx->bstats.emit_ms = c;
        }
    }
    idx->bstats.total_ms = idx->bstats.count_ms + idx->bstats.scan_ms + idx->bstats.emit_ms;
    *out = idx->bstats;
    return GG_OK;
}

gg_status gg_host_alloc(void** out, size_t bytes) {
    if (!out) return GG_ERR_INVALID_ARG;
    *out = nullptr;
    cudaError_t e = cudaHostAlloc(out, bytes ? bytes : 16, cudaHostAllocDefault);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return e == cudaErrorMemoryAllocation ? GG_ERR_OOM : GG_ERR_CUDA;
    }
    return GG_OK;
}

void gg_host_free(void* p) {
    if (p) cudaFreeHost(p);
}

gg_status gg_set_option(gg_index* idx, const char* key, int64_t value) {
    if (!idx || !key) return GG_ERR_INVALID_ARG;
    std::string k(key);
    if (k == "chunk_queries" && value >= 1) idx->chunk_queries = value;
    else if (k == "stage_bytes" && value >= 0 && value <= 200 * 1024) {
        idx->stage_bytes = value;
        if (idx->built) {  // re-plan staging for the existing snapshot (geometry itself is unchanged)
            gg_status s = plan_levels(idx);
            if (s != GG_OK) return s;
        }
    } else if (k == "count_threads" && (value == 128 || value == 256 || value == 512 || value == 1024)) idx->count_threads = value;
    else if (k == "count_ctas_per_sm" && value >= 0 && value <= 32) idx->count_ctas_per_sm = value;
    else if (k == "emit_ctas_per_sm" && value >= 0 && value <= 32) idx->emit_ctas_per_sm = value;
    else if (k == "fused_ctas_per_sm" && value >= 0 && value <= 32) idx->fused_ctas_per_sm = value;
    else if (k == "fused_kernel" && value >= -1 && value <= 1) idx->fused_kernel = value;
    else if (k == "dense_emit" && value >= -1 && value <= 1) idx->dense_emit = value;
    else if (k == "emit_walk" && (value == 8 || value == 16)) idx->emit_walk = value;
    else if (k == "emit_tma" && value >= 0 && value <= 2) idx->emit_tma = value;
    else if (k == "keep_build_workspace" && (value == 0 || value == 1)) {
        idx->keep_build_workspace = value;
        if (!value) {
            for (auto* b : idx->bw.all()) b->release();
            idx->chain_src = 0;
        }
    }
    else if (k == "fused_regs40" && (value == 0 || value == 1)) idx->fused_regs40 = value;
    else if (k == "sort_legacy" && (value == 0 || value == 1)) idx->sort_legacy = value;
    else if (k == "sort_geom" && value >= 0 && value <= 2) idx->sort_geom = value;
    else if (k == "coop_budget" && value >= 0 && value <= 1000000) idx->coop_budget = value;
    else if (k == "delta_min_hits" && value >= 0 && value <= (1ll << 30)) idx->delta_min_hits = value;
    else if (k == "emit_vec_out" && (value == 0 || value == 1)) idx->emit_vec_out = value;
    else if (k == "emit_tile_list" && (value == 0 || value == 1)) idx->emit_tile_list = value;
    else if (k == "debug_checks" && (value == 0 || value == 1)) idx->debug_checks = value;
    else if (k == "host_mode" && (value == 0 || value == 1)) idx->host_mode = value;
    else if (k == "count_mode" && value >= 0 && value <= 2) idx->count_mode = value;
    else if (k == "timing") idx->timing = value != 0;
    else if (k == "inject_alloc_failure" && value >= 0) g_alloc_fail_countdown.store(value);
    else if (k == "compact_delta_bytes" && (value == 1 || value == 2 || value == 4)) idx->compact_delta_bytes = static_cast<int>(value);
    else return fail(idx, GG_ERR_INVALID_ARG, "unknown option or value out of range");
    return GG_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------ flanking

namespace {
template <typename C, typename Q>
gg_status launch_flanking(gg_index* idx, const uint32_t* d_chrom, const Q* d_start, const Q* d_end, size_t nq,
                          uint32_t* d_pred, uint32_t* d_succ, uint32_t* d_err, cudaStream_t st) {
    if (nq == 0) return GG_OK;
    constexpr int QPT = 4;
    const int vec_ok = aligned16(d_chrom) && aligned16(d_start) && aligned16(d_end) && aligned16(d_pred) && aligned16(d_succ);
    int per_sm = 0;
    GG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, flanking_kernel<C, Q, QPT>, kQueryThreads, 0));
    const size_t tiles = div_up(nq, static_cast<size_t>(kQueryThreads) * QPT);
    const unsigned grid = static_cast<unsigned>(std::min<size_t>(tiles, static_cast<size_t>(idx->sm_count) * std::max(per_sm, 1)));
    flanking_kernel<C, Q, QPT><<<grid, kQueryThreads, 0, st>>>(make_view<C>(idx), idx->ER.as<uint32_t>(), idx->er_base, d_chrom,
                                                              d_start, d_end, nq, d_pred, d_succ, d_err, vec_ok);
    GG_CUDA(cudaGetLastError());
    return GG_OK;
}

template <typename C, typename Q>
gg_status launch_flanking_strand(gg_index* idx, const uint32_t* d_chrom, const Q* d_start, const Q* d_end, const uint8_t* d_strand,
                                 size_t nq, uint32_t* d_pred, uint32_t* d_succ, uint32_t* d_err, cudaStream_t st) {
    if (nq == 0) return GG_OK;
    int per_sm = 0;
    GG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, flanking_strand_kernel<C, Q>, kQueryThreads, 0));
    const unsigned grid = static_cast<unsigned>(std::min<size_t>(div_up(nq, kQueryThreads), static_cast<size_t>(idx->sm_count) * std::max(per_sm, 1)));
    flanking_strand_kernel<C, Q><<<grid, kQueryThreads, 0, st>>>(make_view<C>(idx), idx->ER.as<uint32_t>(), idx->er_base, idx->SN.as<StrandNode>(),
                                                                idx->EN.as<StrandNode>(), d_chrom, d_start, d_end, d_strand, nq, d_pred, d_succ, d_err);
    GG_CUDA(cudaGetLastError());
    return GG_OK;
}

// d_strand != nullptr: genomic_coordinate keys with the strand-compatible predicate (flanking_strand_kernel)
gg_status dispatch_flanking(gg_index* idx, const uint32_t* d_chrom, const void* d_start, const void* d_end, int q_bytes,
                            size_t nq, uint32_t* d_pred, uint32_t* d_succ, uint32_t* d_err, cudaStream_t st,
                            const uint8_t* d_strand = nullptr) {
    if (d_strand) {
        if (idx->coord_bytes == 8)
            return launch_flanking_strand<uint64_t, uint64_t>(idx, d_chrom, static_cast<const uint64_t*>(d_start),
                                                              static_cast<const uint64_t*>(d_end), d_strand, nq, d_pred, d_succ, d_err, st);
        if (q_bytes == 8)
            return launch_flanking_strand<uint32_t, uint64_t>(idx, d_chrom, static_cast<const uint64_t*>(d_start),
                                                              static_cast<const uint64_t*>(d_end), d_strand, nq, d_pred, d_succ, d_err, st);
        return launch_flanking_strand<uint32_t, uint32_t>(idx, d_chrom, static_cast<const uint32_t*>(d_start),
                                                          static_cast<const uint32_t*>(d_end), d_strand, nq, d_pred, d_succ, d_err, st);
    }
    if (idx->coord_bytes == 8)
        return launch_flanking<uint64_t, uint64_t>(idx, d_chrom, static_cast<const uint64_t*>(d_start),
                                                   static_cast<const uint64_t*>(d_end), nq, d_pred, d_succ, d_err, st);
    if (q_bytes == 8)
        return launch_flanking<uint32_t, uint64_t>(idx, d_chrom, static_cast<const uint64_t*>(d_start),
                                                   static_cast<const uint64_t*>(d_end), nq, d_pred, d_succ, d_err, st);
    return launch_flanking<uint32_t, uint32_t>(idx, d_chrom, static_cast<const uint32_t*>(d_start),
                                               static_cast<const uint32_t*>(d_end), nq, d_pred, d_succ, d_err, st);
}
}  // namespace

extern "C" {

static gg_status flanking_check(gg_index* idx, cudaStream_t st) {
    uint32_t flag = 0;
    GG_CUDA(cudaMemcpyAsync(&flag, idx->err_flag.p, 4, cudaMemcpyDeviceToHost, st));
    GG_CUDA(cudaStreamSynchronize(st));
    if (flag & 4u) return fail(idx, GG_ERR_INVALID_ARG, "genomic_coordinate: strand must be one of '+', '-', '.', '*'");
    if (flag & 1u) return fail(idx, GG_ERR_INVALID_INTERVAL, "query interval: start must be <= end");
    return GG_OK;
}

static gg_status flanking_device_impl(gg_index* idx, const uint32_t* d_chrom, const void* d_start, const void* d_end, const uint8_t* d_strand,
                                      bool stranded, int coord_bytes, size_t nq, uint32_t* d_pred, uint32_t* d_succ, int sync, void* stream) {
    if (!idx) return GG_ERR_INVALID_ARG;
    idx->err.clear();
    if (!idx->built) return fail(idx, GG_ERR_NOT_BUILT, "index not built");
    if (stranded && !idx->has_strand) return fail(idx, GG_ERR_NOT_BUILT, "index has no strands (gg_index_set_strands)");
    if (nq && (!d_chrom || !d_start || !d_end || !d_pred || !d_succ || (stranded && !d_strand))) return fail(idx, GG_ERR_INVALID_ARG, "null argument");
    if (coord_bytes != 4 && coord_bytes != 8) return fail(idx, GG_ERR_INVALID_ARG, "coord_bytes must be 4 or 8");
    if (coord_bytes == 4 && idx->coord_bytes != 4) return fail(idx, GG_ERR_INVALID_ARG, "u32 queries need a u32-layout index");
    GG_CUDA(cudaSetDevice(idx->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    GG_CUDA(idx->err_flag.ensure(16));
    GG_CUDA(cudaMemsetAsync(idx->err_flag.p, 0, 4, st));
    gg_status s = dispatch_flanking(idx, d_chrom, d_start, d_end, coord_bytes, nq, d_pred, d_succ, idx->err_flag.as<uint32_t>(), st,
                                    stranded ? d_strand : nullptr);
    if (s != GG_OK) return s;
    return sync ? flanking_check(idx, st) : GG_OK;
}

gg_status gg_flanking_device(gg_index* idx, const uint32_t* d_chrom, const void* d_start, const void* d_end,
                             int coord_bytes, size_t nq, uint32_t* d_pred, uint32_t* d_succ, int sync, void* stream) {
    return flanking_device_impl(idx, d_chrom, d_start, d_end, nullptr, false, coord_bytes, nq, d_pred, d_succ, sync, stream);
}
gg_status gg_flanking_stranded_device(gg_index* idx, const uint32_t* d_chrom, const void* d_start, const void* d_end,
                                      const uint8_t* d_strand, int coord_bytes, size_t nq, uint32_t* d_pred, uint32_t* d_succ,
                                      int sync, void* stream) {
    return flanking_device_impl(idx, d_chrom, d_start, d_end, d_strand, true, coord_bytes, nq, d_pred, d_succ, sync, stream);
}

static gg_status flanking_batch_impl(gg_index* idx, const uint32_t* q_chrom, const uint64_t* q_start, const uint64_t* q_end,
                                     const uint8_t* q_strand, bool stranded, size_t nq, uint32_t* pred, uint32_t* succ) {
    if (!idx) return GG_ERR_INVALID_ARG;
    idx->err.clear();
    if (!idx->built) return fail(idx, GG_ERR_NOT_BUILT, "index not built");
    if (stranded && !idx->has_strand) return fail(idx, GG_ERR_NOT_BUILT, "index has no strands (gg_index_set_strands)");
    if (nq && (!q_chrom || !q_start || !q_end || !pred || !succ || (stranded && !q_strand))) return fail(idx, GG_ERR_INVALID_ARG, "null argument");
    GG_CUDA(cudaSetDevice(idx->device));
    cudaStream_t st = idx->stream;
    GG_CUDA(idx->err_flag.ensure(16));
    GG_CUDA(cudaMemsetAsync(idx->err_flag.p, 0, 4, st));
    GG_CUDA(cudaStreamSynchronize(st));  // the flag is clear before any slot stream touches it
    gg_status s = ensure_slots(idx);
    if (s != GG_OK) return s;
    // chunks round-robin over the pipeline slots' streams: H2D of chunk k+1 and D2H of chunk k-1 overlap kernel k
    SlotDrain drain{idx};  // (error exits below leave no copy into pred / succ in flight)
    const size_t chunk = static_cast<size_t>(std::max<int64_t>(idx->chunk_queries, 1024));
    const size_t cap = std::min(chunk, nq ? nq : 1);
    size_t k = 0;
    for (size_t q0 = 0; q0 < nq; q0 += chunk, ++k) {
        Slot& sl = idx->slots[k % kSlots];
        const size_t m = std::min(chunk, nq - q0);
        if (k >= kSlots) GG_CUDA(cudaStreamSynchronize(sl.st));  // (buffers below may be regrown only when idle)
        GG_CUDA(sl.qc.ensure(cap * 4));
        GG_CUDA(sl.qs.ensure(cap * 8));
        GG_CUDA(sl.qe.ensure(cap * 8));
        GG_CUDA(sl.cnt.ensure(cap * 4));
        GG_CUDA(sl.hi.ensure(cap * 4));
        GG_CUDA(cudaMemcpyAsync(sl.qc.p, q_chrom + q0, m * 4, cudaMemcpyHostToDevice, sl.st));
        GG_CUDA(cudaMemcpyAsync(sl.qs.p, q_start + q0, m * 8, cudaMemcpyHostToDevice, sl.st));
        GG_CUDA(cudaMemcpyAsync(sl.qe.p, q_end + q0, m * 8, cudaMemcpyHostToDevice, sl.st));
        if (stranded) {  // (the slot's scan buffer is free in this call: it stages the strand bytes)
            GG_CUDA(sl.scan.ensure(cap));
            GG_CUDA(cudaMemcpyAsync(sl.scan.p, q_strand + q0, m, cudaMemcpyHostToDevice, sl.st));
        }
        s = dispatch_flanking(idx, sl.qc.as<uint32_t>(), sl.qs.p, sl.qe.p, 8, m, sl.cnt.as<uint32_t>(), sl.hi.as<uint32_t>(),
                              idx->err_flag.as<uint32_t>(), sl.st, stranded ? sl.scan.as<uint8_t>() : nullptr);
        if (s != GG_OK) return s;
        GG_CUDA(cudaMemcpyAsync(pred + q0, sl.cnt.p, m * 4, cudaMemcpyDeviceToHost, sl.st));
        GG_CUDA(cudaMemcpyAsync(succ + q0, sl.hi.p, m * 4, cudaMemcpyDeviceToHost, sl.st));
    }
    for (int i = 0; i < kSlots; ++i) GG_CUDA(cudaStreamSynchronize(idx->slots[i].st));
    return flanking_check(idx, st);  // synchronises the stream
}

gg_status gg_flanking_batch(gg_index* idx, const uint32_t* q_chrom, const uint64_t* q_start, const uint64_t* q_end,
                            size_t nq, uint32_t* pred, uint32_t* succ) {
    return flanking_batch_impl(idx, q_chrom, q_start, q_end, nullptr, false, nq, pred, succ);
}
gg_status gg_flanking_batch_stranded(gg_index* idx, const uint32_t* q_chrom, const uint64_t* q_start, const uint64_t* q_end,
                                     const uint8_t* q_strand, size_t nq, uint32_t* pred, uint32_t* succ) {
    return flanking_batch_impl(idx, q_chrom, q_start, q_end, q_strand, true, nq, pred, succ);
}

}  // extern "C"

// ------------------------------------------------------------------------------------------ stranded intersect

namespace {
// Count phase (emit_only = false): strand_count_group_kernel -> scan of the per-128-query block totals (+ *d_base),
// *d_total = hits of the batch.  Emit phase (emit_only = true): intersect_emit_scan_kernel<.., kStrand> writes
// d_off[0..nq] (global offsets) and the chunk-local hit ranks.  Both use idx->s_cnt / s_hi / s_scan.
template <typename C, typename Q>
gg_status launch_stranded(gg_index* idx, const uint32_t* d_chrom, const Q* d_start, const Q* d_end, const uint8_t* d_strand,
                          size_t nq, uint64_t* d_off, uint32_t* d_hits, uint64_t capacity, const uint64_t* d_base,
                          uint64_t* d_total, uint32_t* d_err, bool emit_only, cudaStream_t st) {
    char* p = static_cast<char*>(idx->s_scan.p);
    const size_t nb = batch_blocks(nq), nb1 = nb + 1;
    unsigned long long* block_total = reinterpret_cast<unsigned long long*>(p);
    uint64_t* block_prefix = reinterpret_cast<uint64_t*>(p + round_up(nb1 * 8, 256));
    void* scan_state = p + round_up(nb1 * 8, 256) + round_up((nb1 + 1) * 8, 256);
    uint32_t* d_cnt = idx->s_cnt.as<uint32_t>();
    uint32_t* d_hi = idx->s_hi.as<uint32_t>();
    if (!emit_only) {
        if (nq) {
            idx->s_dense = dense_pass_on(idx, true);  // the emit phase of this batch follows the same choice
            auto kern = idx->s_dense ? strand_count_group_kernel<C, Q, kBatchQpt, true> : strand_count_group_kernel<C, Q, kBatchQpt, false>;
            int per_sm = 0;
            GG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kQueryThreads, 0));
            const size_t tiles = div_up(nq, static_cast<size_t>(kQueryThreads) * kBatchQpt);
            const unsigned grid = static_cast<unsigned>(std::min<size_t>(tiles, static_cast<size_t>(idx->sm_count) * std::max(per_sm, 1)));
            const int vec_ok = aligned16(d_chrom) && aligned16(d_start) && aligned16(d_end);
            kern<<<grid, kQueryThreads, 0, st>>>(make_view<C>(idx), idx->SN.as<StrandNode>(), idx->EN.as<StrandNode>(), idx->er_base,
                                                 d_chrom, d_start, d_end, d_strand, nq, d_cnt, d_hi, block_total, d_err, vec_ok,
                                                 idx->s_dense ? idx->s_tile.as<uint32_t>() : nullptr);
            GG_CUDA(cudaGetLastError());
            idx->bstats.kernel_launches++;
        }
        GG_CUDA((exclusive_scan_async<uint64_t, uint64_t>(reinterpret_cast<const uint64_t*>(block_total), block_prefix, nb, scan_state,
                                                           d_base, d_total, st)));
        idx->bstats.kernel_launches++;
        return GG_OK;
    }
    const bool tab_in_smem = static_cast<size_t>(idx->n_levels) * (idx->n_chrom + 1) * 4 <= 8 * 1024;
    const size_t tab_bytes = static_cast<size_t>(idx->n_levels) * (idx->n_chrom + 1) * 4;
    const size_t smem = static_cast<size_t>(kQueryThreads / 32) * (kEmitStage + kEmitScanSlack) * 4 + (tab_in_smem ? tab_bytes : 0) + 16;
    auto kern = idx->s_dense ? intersect_emit_scan_kernel<C, Q, 8, true, true> : intersect_emit_scan_kernel<C, Q, 8, true, false>;
    if (smem > 48 * 1024) GG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    int per_sm = 0;
    GG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kQueryThreads, smem));
    if (per_sm < 1) return fail(idx, GG_ERR_CUDA, "emit kernel does not fit on an SM");
    const size_t ctas = div_up(std::max<size_t>(nb, 1), kQueryThreads / 32);
    const unsigned grid = static_cast<unsigned>(std::min<size_t>(ctas, static_cast<size_t>(idx->sm_count) * per_sm));
    kern<<<grid, kQueryThreads, smem, st>>>(make_view<C>(idx), d_chrom, d_start, nq, d_cnt, d_hi, block_prefix,
                                            static_cast<uint32_t>(kBatchBlockQ), d_off, d_hits, capacity, nullptr, nullptr, d_err,
                                            tab_in_smem ? 1 : 0, StrandArgs{idx->strand.as<uint8_t>(), d_strand, idx->er_base}, CoopArgs<C>{});
    GG_CUDA(cudaGetLastError());
    idx->bstats.kernel_launches++;
    if (!idx->s_dense) return GG_OK;
    return launch_emit_dense<C, Q, true>(idx, d_chrom, d_start, nq, d_cnt, d_hi, block_prefix, idx->s_tile.as<uint32_t>(), d_hits,
                                         capacity, StrandArgs{idx->strand.as<uint8_t>(), d_strand, idx->er_base}, st);
}

gg_status dispatch_stranded(gg_index* idx, const uint32_t* d_chrom, const void* d_start, const void* d_end,
                            const uint8_t* d_strand, int q_bytes, size_t nq, uint64_t* d_off, uint32_t* d_hits,
                            uint64_t capacity, const uint64_t* d_base, uint64_t* d_total, uint32_t* d_err, bool emit_only,
                            cudaStream_t st) {
    if (idx->coord_bytes == 8)
        return launch_stranded<uint64_t, uint64_t>(idx, d_chrom, static_cast<const uint64_t*>(d_start), static_cast<const uint64_t*>(d_end),
                                                   d_strand, nq, d_off, d_hits, capacity, d_base, d_total, d_err, emit_only, st);
    if (q_bytes == 8)
        return launch_stranded<uint32_t, uint64_t>(idx, d_chrom, static_cast<const uint64_t*>(d_start), static_cast<const uint64_t*>(d_end),
                                                   d_strand, nq, d_off, d_hits, capacity, d_base, d_total, d_err, emit_only, st);
    return launch_stranded<uint32_t, uint32_t>(idx, d_chrom, static_cast<const uint32_t*>(d_start), static_cast<const uint32_t*>(d_end),
                                               d_strand, nq, d_off, d_hits, capacity, d_base, d_total, d_err, emit_only, st);
}

gg_status ensure_strand_scratch(gg_index* idx, size_t nq) {
    const size_t m = nq ? nq : 1;
    GG_CUDA(idx->s_cnt.ensure(round_up(m * 4, 256)));
    GG_CUDA(idx->s_hi.ensure(round_up(m * 4, 256)));
    GG_CUDA(idx->s_tile.ensure(round_up(div_up(m, 32) * 4, 256)));
    GG_CUDA(idx->s_scan.ensure(batch_ctl_bytes(m)));
    GG_CUDA(idx->s_total.ensure(16));
    GG_CUDA(idx->err_flag.ensure(16));
    return GG_OK;
}

gg_status strand_query_error(gg_index* idx, uint32_t flag) {
    if (flag & 1u) return fail(idx, GG_ERR_INVALID_INTERVAL, "query interval: start must be <= end");
    if (flag & 4u) return fail(idx, GG_ERR_INVALID_ARG, "genomic_coordinate: strand must be one of '+', '-', '.', '*'");
    return GG_OK;
}
}  // namespace

extern "C" {

gg_status gg_index_set_strands(gg_index* idx, const uint8_t* strand, size_t n) {
    if (!idx) return GG_ERR_INVALID_ARG;
    idx->err.clear();
    if (!idx->built) return fail(idx, GG_ERR_NOT_BUILT, "index not built");
    if (n != idx->n || (n && !strand)) return fail(idx, GG_ERR_INVALID_ARG, "strand array must hold one byte per indexed interval");
    for (size_t i = 0; i < n; ++i) {
        const uint8_t c = strand[i];
        if (c != '+' && c != '-' && c != '.' && c != '*')
            return fail(idx, GG_ERR_INVALID_ARG, "genomic_coordinate: strand must be one of '+', '-', '.', '*'");
    }
    GG_CUDA(cudaSetDevice(idx->device));
    idx->has_strand = false;  // until every step below has succeeded
    const size_t C = idx->n_chrom;
    const uint32_t level0_elems = idx->h_tab[C] - idx->h_tab[0];
    DeviceBuf tmp;
    GG_CUDA(tmp.ensure(n ? n : 1));
    cudaError_t e = idx->strand.ensure(level0_elems ? level0_elems : 1);
    if (e == cudaSuccess && n) e = cudaMemcpyAsync(tmp.p, strand, n, cudaMemcpyHostToDevice, idx->stream);
    if (e == cudaSuccess) {
        pad_strand_kernel<<<blocks_for(level0_elems, 256), 256, 0, idx->stream>>>(tmp.as<uint8_t>(), idx->cbegin.as<uint32_t>(),
                                                                                 idx->cnt.as<uint32_t>(), idx->tab.as<uint32_t>(),
                                                                                 idx->n_chrom, idx->strand.as<uint8_t>());
        e = cudaGetLastError();
    }
    // per-node strand records (class prefix counts + masks) in chain order and sorted-end order
    const uint32_t n_nodes = level0_elems >> kFanoutLog2, nn1 = n_nodes + 1;
    DeviceBuf cntS, cntE, preS, preE, scan_state;
    if (e == cudaSuccess) e = idx->SN.ensure(static_cast<size_t>(nn1) * sizeof(StrandNode));
    if (e == cudaSuccess) e = idx->EN.ensure(static_cast<size_t>(nn1) * sizeof(StrandNode));
    if (e == cudaSuccess) e = cntS.ensure(static_cast<size_t>(nn1) * 16);
    if (e == cudaSuccess) e = cntE.ensure(static_cast<size_t>(nn1) * 16);
    if (e == cudaSuccess) e = preS.ensure(static_cast<size_t>(nn1) * 16 + 16);
    if (e == cudaSuccess) e = preE.ensure(static_cast<size_t>(nn1) * 16 + 16);
    if (e == cudaSuccess) e = scan_state.ensure(scan_state_bytes(static_cast<size_t>(nn1) * 4));
    if (e == cudaSuccess) {
        const unsigned wpb = 8;
        strand_node_kernel<<<blocks_for(nn1, wpb), wpb * 32, 0, idx->stream>>>(
            idx->strand.as<uint8_t>(), idx->ER.as<uint32_t>(), idx->tab.as<uint32_t>(), idx->n_chrom, n_nodes,
            idx->SN.as<StrandNode>(), idx->EN.as<StrandNode>(), cntS.as<uint32_t>(), cntE.as<uint32_t>());
        e = cudaGetLastError();
    }
    // one scan per family over [class][node]: the running count carries across classes, so subtract each class's
    // start (= pre[k][0]) when packing — done by scanning each class row separately instead
    for (uint32_t k = 0; k < 4 && e == cudaSuccess; ++k) {
        e = exclusive_scan_async<uint32_t, uint32_t>(cntS.as<uint32_t>() + static_cast<size_t>(k) * nn1,
                                                     preS.as<uint32_t>() + static_cast<size_t>(k) * nn1, nn1 - 1, scan_state.p, nullptr,
                                                     nullptr, idx->stream);
        if (e == cudaSuccess)
            e = exclusive_scan_async<uint32_t, uint32_t>(cntE.as<uint32_t>() + static_cast<size_t>(k) * nn1,
                                                         preE.as<uint32_t>() + static_cast<size_t>(k) * nn1, nn1 - 1, scan_state.p, nullptr,
                                                         nullptr, idx->stream);
    }
    if (e == cudaSuccess) {
        strand_pack_kernel<<<blocks_for(nn1, 256), 256, 0, idx->stream>>>(preS.as<uint32_t>(), preE.as<uint32_t>(), n_nodes,
                                                                         idx->SN.as<StrandNode>(), idx->EN.as<StrandNode>());
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(idx->stream);
    tmp.release();
    for (DeviceBuf* b : {&cntS, &cntE, &preS, &preE, &scan_state}) b->release();
    if (e != cudaSuccess) {  // the stranded entry points must not run on missing / half-built strand records
        idx->has_strand = false;
        return fail(idx, GG_ERR_CUDA, "gg_index_set_strands", e);
    }
    idx->has_strand = true;
    return GG_OK;
}

gg_status gg_intersect_stranded_device(gg_index* idx, const uint32_t* d_chrom, const void* d_start, const void* d_end,
                                       const uint8_t* d_strand, int coord_bytes, size_t nq, uint64_t* d_offsets,
                                       uint32_t* d_hits, uint64_t capacity, uint64_t* total_hits, void* stream) {
    if (!idx) return GG_ERR_INVALID_ARG;
    idx->err.clear();
    if (!idx->built) return fail(idx, GG_ERR_NOT_BUILT, "index not built");
    if (!idx->has_strand) return fail(idx, GG_ERR_NOT_BUILT, "index has no strands (gg_index_set_strands)");
    if (!d_offsets || (capacity && !d_hits) || (nq && (!d_chrom || !d_start || !d_end || !d_strand)))
        return fail(idx, GG_ERR_INVALID_ARG, "null argument");
    if (coord_bytes != 4 && coord_bytes != 8) return fail(idx, GG_ERR_INVALID_ARG, "coord_bytes must be 4 or 8");
    if (coord_bytes == 4 && idx->coord_bytes != 4) return fail(idx, GG_ERR_INVALID_ARG, "u32 queries need a u32-layout index");
    GG_CUDA(cudaSetDevice(idx->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    gg_status s = ensure_strand_scratch(idx, nq);
    if (s != GG_OK) return s;
    idx->bstats = gg_batch_stats{};
    idx->bstats.n_queries = nq;
    idx->bstats.n_chunks = 1;
    idx->counted_nq = 0;
    poll_lazy_density(idx);
    GG_CUDA(cudaMemsetAsync(idx->err_flag.p, 0, 4, st));
    s = dispatch_stranded(idx, d_chrom, d_start, d_end, d_strand, coord_bytes, nq, d_offsets, d_hits, capacity, nullptr,
                          idx->s_total.as<uint64_t>(), idx->err_flag.as<uint32_t>(), false, st);
    if (s != GG_OK) return s;
    s = dispatch_stranded(idx, d_chrom, d_start, d_end, d_strand, coord_bytes, nq, d_offsets, d_hits, capacity, nullptr,
                          nullptr, idx->err_flag.as<uint32_t>(), true, st);
    if (s != GG_OK) return s;
    if (total_hits) {
        uint32_t flag = 0;
        if (!idx->h_err) GG_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&idx->h_err), 16, cudaHostAllocDefault));
        GG_CUDA(cudaMemcpyAsync(idx->h_err + 2, idx->s_total.p, 8, cudaMemcpyDeviceToHost, st));  // pinned: no staging copies
        GG_CUDA(cudaMemcpyAsync(idx->h_err, idx->err_flag.p, 4, cudaMemcpyDeviceToHost, st));
        GG_CUDA(cudaStreamSynchronize(st));
        std::memcpy(total_hits, idx->h_err + 2, 8);
        flag = idx->h_err[0];
        idx->bstats.n_hits = *total_hits;
        note_density(idx, *total_hits, nq, true);
        s = strand_query_error(idx, flag);
        if (s != GG_OK) return s;
        if (*total_hits > capacity) return fail(idx, GG_ERR_CAPACITY, "hit buffer smaller than offsets[nq]; offsets are complete");
    } else {
        GG_CUDA(post_lazy_density(idx, idx->s_total.p, nq, true, st));
    }
    return GG_OK;
}

gg_status gg_intersect_batch_stranded(gg_index* idx, const uint32_t* q_chrom, const uint64_t* q_start, const uint64_t* q_end,
                                      const uint8_t* q_strand, size_t nq, uint64_t* offsets, const uint32_t** hits,
                                      uint64_t* n_hits) {
    if (!idx) return GG_ERR_INVALID_ARG;
    idx->err.clear();
    if (!idx->built) return fail(idx, GG_ERR_NOT_BUILT, "index not built");
    if (!idx->has_strand) return fail(idx, GG_ERR_NOT_BUILT, "index has no strands (gg_index_set_strands)");
    if (!offsets || !hits || !n_hits || (nq && (!q_chrom || !q_start || !q_end || !q_strand)))
        return fail(idx, GG_ERR_INVALID_ARG, "null argument");
    GG_CUDA(cudaSetDevice(idx->device));
    cudaStream_t st = idx->stream;
    idx->h2_valid = false;
    idx->bstats = gg_batch_stats{};
    idx->bstats.n_queries = nq;
    const size_t chunk = static_cast<size_t>(std::max<int64_t>(idx->chunk_queries, 1024));
    const size_t cap_q = std::min(chunk, nq ? nq : 1);
    gg_status s = ensure_strand_scratch(idx, cap_q);
    if (s != GG_OK) return s;
    GG_CUDA(idx->sq_qc.ensure(cap_q * 4));
    GG_CUDA(idx->sq_qs.ensure(cap_q * 8));
    GG_CUDA(idx->sq_qe.ensure(cap_q * 8));
    GG_CUDA(idx->sq_str.ensure(cap_q));
    GG_CUDA(idx->sq_off.ensure((cap_q + 1) * 8));
    GG_CUDA(idx->d_base.ensure(16));
    GG_CUDA(cudaMemsetAsync(idx->err_flag.p, 0, 4, st));
    GG_CUDA(cudaMemsetAsync(idx->d_base.p, 0, 8, st));
    uint64_t base = 0;
    offsets[0] = 0;
    for (size_t q0 = 0; q0 < nq; q0 += chunk) {
        const size_t m = std::min(chunk, nq - q0);
        GG_CUDA(cudaMemcpyAsync(idx->sq_qc.p, q_chrom + q0, m * 4, cudaMemcpyHostToDevice, st));
        GG_CUDA(cudaMemcpyAsync(idx->sq_qs.p, q_start + q0, m * 8, cudaMemcpyHostToDevice, st));
        GG_CUDA(cudaMemcpyAsync(idx->sq_qe.p, q_end + q0, m * 8, cudaMemcpyHostToDevice, st));
        GG_CUDA(cudaMemcpyAsync(idx->sq_str.p, q_strand + q0, m, cudaMemcpyHostToDevice, st));
        s = dispatch_stranded(idx, idx->sq_qc.as<uint32_t>(), idx->sq_qs.p, idx->sq_qe.p, idx->sq_str.as<uint8_t>(), 8, m,
                              idx->sq_off.as<uint64_t>(), nullptr, 0, idx->d_base.as<uint64_t>(), idx->s_total.as<uint64_t>(),
                              idx->err_flag.as<uint32_t>(), false, st);
        if (s != GG_OK) return s;
        uint64_t total = 0;
        uint32_t flag = 0;
        GG_CUDA(cudaMemcpyAsync(&total, idx->s_total.p, 8, cudaMemcpyDeviceToHost, st));
        GG_CUDA(cudaMemcpyAsync(&flag, idx->err_flag.p, 4, cudaMemcpyDeviceToHost, st));
        GG_CUDA(cudaStreamSynchronize(st));
        s = strand_query_error(idx, flag);
        if (s != GG_OK) return s;
        note_density(idx, total, m, true);
        GG_CUDA(idx->sq_hits.ensure((total ? total : 1) * 4));
        s = ensure_pinned_hits(idx, base + total, base);
        if (s != GG_OK) return s;
        s = dispatch_stranded(idx, idx->sq_qc.as<uint32_t>(), idx->sq_qs.p, idx->sq_qe.p, idx->sq_str.as<uint8_t>(), 8, m,
                              idx->sq_off.as<uint64_t>(), idx->sq_hits.as<uint32_t>(), total, nullptr, nullptr, idx->err_flag.as<uint32_t>(), true, st);
        if (s != GG_OK) return s;
        GG_CUDA(cudaMemcpyAsync(offsets + q0, idx->sq_off.p, (m + 1) * 8, cudaMemcpyDeviceToHost, st));
        if (total) GG_CUDA(cudaMemcpyAsync(idx->pinned_hits + base, idx->sq_hits.p, total * 4, cudaMemcpyDeviceToHost, st));
        base += total;
        GG_CUDA(cudaMemcpyAsync(idx->d_base.p, &base, 8, cudaMemcpyHostToDevice, st));
        GG_CUDA(cudaStreamSynchronize(st));
        idx->bstats.n_chunks++;
    }
    if (nq == 0) {
        s = ensure_pinned_hits(idx, 1, 0);
        if (s != GG_OK) return s;
    }
    *hits = idx->pinned_hits;
    *n_hits = base;
    idx->bstats.n_hits = base;
    return GG_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------ allocation-free single call
//
// gg_intersect_fused_device_static: the one-kernel single-call path (fused.cuh) for callers that must not allocate
// device memory at query time.  Three kinds of caller are in that position: a service that handed all of its HBM to its
// own pool at start-up, so a cudaMalloc from this library would simply fail; a process that shares the GPU with a
// framework whose caching allocator already reserved the free memory; and a latency-bound loop, where cudaMalloc's
// implicit device synchronisation and unbounded duration are not acceptable on the query path.  The regular entry point
// grows a scratch buffer on the handle whenever it meets a larger batch; this one never does.  Its scratch — the tile
// ticket and one 64-bit look-back word per 256 queries, plus the 16-byte error / hit-total mailbox — sits in a
// fixed-size array in the library's device data segment instead.
namespace {

constexpr size_t kStaticMailboxBytes = 16;  // [0, 4) error flag, [8, 16) hit total
constexpr size_t kStaticTiles = (static_cast<size_t>(GG_STATIC_MAX_QUERIES) + kQueryThreads - 1) / kQueryThreads;
// mailbox + what launch_fused lays out ([ticket, padded to 16 B][state u64 x tiles]), rounded to the 256-byte alignment
constexpr size_t kStaticScratchBytes = ((kStaticMailboxBytes + 16 + kStaticTiles * 8 + 255) / 256) * 256;

// One copy per GPU that loads the library (33 KB each), held until the process exits.  In host code this name is only a
// handle for the runtime's symbol table, NOT a device address: the pointer the kernel gets comes from
// cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

constexpr int kStaticMaxDevices = 64;
std::mutex g_static_mu;
unsigned char* g_static_addr[kStaticMaxDevices] = {};

// Device address of g_static_scratch on `device`, which must be the current device.  Resolved once per device.
cudaError_t static_scratch_on(int device, unsigned char** out) {
    if (device < 0 || device >= kStaticMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_addr[device]) {
        int current = -1;
        cudaError_t e = cudaGetDevice(&current);
        if (e != cudaSuccess) return e;
        if (current != device) return cudaErrorInvalidDevice;
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        g_static_addr[device] = static_cast<unsigned char*>(p);
    }
    *out = g_static_addr[device];
    return cudaSuccess;
}

}  // namespace

extern "C" {

gg_status gg_intersect_fused_device_static(gg_index* idx, const uint32_t* d_chrom, const void* d_start, const void* d_end,
                                           int coord_bytes, size_t nq, uint64_t* d_offsets, uint32_t* d_hits,
                                           uint64_t capacity, uint64_t* total_hits, void* stream) {
    if (!idx) return GG_ERR_INVALID_ARG;
    idx->err.clear();
    if (!idx->built) return fail(idx, GG_ERR_NOT_BUILT, "index not built");
    if (!d_offsets || (capacity && !d_hits) || (nq && (!d_chrom || !d_start || !d_end)))
        return fail(idx, GG_ERR_INVALID_ARG, "null argument");
    if (coord_bytes != 4 && coord_bytes != 8) return fail(idx, GG_ERR_INVALID_ARG, "coord_bytes must be 4 or 8");
    if (coord_bytes == 4 && idx->coord_bytes != 4) return fail(idx, GG_ERR_INVALID_ARG, "u32 queries need a u32-layout index");
    if (nq > static_cast<size_t>(GG_STATIC_MAX_QUERIES))
        return fail(idx, GG_ERR_CAPACITY,
                    "gg_intersect_fused_device_static: more than GG_STATIC_MAX_QUERIES (1048576) queries in one call; split the "
                    "batch or use gg_intersect_fused_device");
    GG_CUDA(cudaSetDevice(idx->device));
    unsigned char* scratch = nullptr;
    GG_CUDA(static_scratch_on(idx->device, &scratch));  // first call on this device: cudaGetSymbolAddress
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    uint32_t* d_err = reinterpret_cast<uint32_t*>(scratch);
    uint64_t* d_total = reinterpret_cast<uint64_t*>(scratch + 8);
    idx->bstats = gg_batch_stats{};
    idx->bstats.n_queries = nq;
    idx->bstats.n_chunks = 1;
    idx->ev_count_valid = idx->ev_emit_valid = false;
    // the array keeps what the previous call left: the mailbox is cleared here, ticket + look-back words by launch_fused
    GG_CUDA(cudaMemsetAsync(scratch, 0, kStaticMailboxBytes, st));
    void* lookback = scratch + kStaticMailboxBytes;
    gg_status s;
    if (idx->coord_bytes == 4 && coord_bytes == 4)
        s = launch_fused<uint32_t, uint32_t>(idx, d_chrom, static_cast<const uint32_t*>(d_start), static_cast<const uint32_t*>(d_end), nq,
                                             d_offsets, d_hits, capacity, nullptr, nullptr, d_total, lookback, d_err, st);
    else if (idx->coord_bytes == 4)
        s = launch_fused<uint32_t, uint64_t>(idx, d_chrom, static_cast<const uint64_t*>(d_start), static_cast<const uint64_t*>(d_end), nq,
                                             d_offsets, d_hits, capacity, nullptr, nullptr, d_total, lookback, d_err, st);
    else if (coord_bytes == 4)
        s = launch_fused<uint64_t, uint32_t>(idx, d_chrom, static_cast<const uint32_t*>(d_start), static_cast<const uint32_t*>(d_end), nq,
                                             d_offsets, d_hits, capacity, nullptr, nullptr, d_total, lookback, d_err, st);
    else
        s = launch_fused<uint64_t, uint64_t>(idx, d_chrom, static_cast<const uint64_t*>(d_start), static_cast<const uint64_t*>(d_end), nq,
                                             d_offsets, d_hits, capacity, nullptr, nullptr, d_total, lookback, d_err, st);
    if (s != GG_OK) return s;
    if (total_hits) {
        uint32_t mailbox[4] = {0, 0, 0, 0};  // on the stack: the handle's pinned mailbox may not exist yet and is not created here
        GG_CUDA(cudaMemcpyAsync(mailbox, scratch, kStaticMailboxBytes, cudaMemcpyDeviceToHost, st));
        GG_CUDA(cudaStreamSynchronize(st));
        std::memcpy(total_hits, mailbox + 2, 8);
        idx->bstats.n_hits = *total_hits;
        if (mailbox[0] & 1u) return fail(idx, GG_ERR_INVALID_INTERVAL, "query interval: start must be <= end");
        if (mailbox[0] & 2u) return fail(idx, GG_ERR_CAPACITY, "hit buffer smaller than offsets[nq]; offsets are complete");
    }
    return GG_OK;
}

}  // extern "C"
