/*
 * gg_b200.h — C ABI of the B200-native batched interval-overlap engine (libgg_b200.so).
 *
 * This is the drop-in boundary for genogrove's grove<interval, T>::intersect hot path.  Plain C
 * types, caller-owned buffers unless stated, `gg_status` return codes, no exceptions and no torch /
 * C++ types across the boundary.  One opaque handle per grove snapshot; calls on one handle must be
 * serialised by the caller (the reference is single-threaded with no internal locking,
 * CHANGELOG.md:46).  There is NO CPU fallback: every entry point that computes runs hand-written
 * sm_100a kernels and fails with GG_ERR_CUDA when no device is usable.
 *
 * Reference interfaces replaced (paths relative to the reference checkout):
 *
 *   gg_index_build            <- grove::insert_data(idx, vec, bulk): std::sort + build_tree_bottom_up
 *                                include/genogrove/structure/grove/grove_insert.ipp:160-172, 545-669
 *                                (one call covers every chromosome; the reference is called per index)
 *   gg_index_upload_sorted    <- the leaf chain left by insert / insert_sorted / remove_key / compact
 *                                grove_insert.ipp:182-274, 501-531; grove_remove.ipp:24-98
 *   gg_intersect_*            <- grove::intersect(query, index) -> detail::search_overlaps
 *   (batch, fused_device,        [the host-buffer batch call and gg_intersect_fused_device are single calls that run
 *    count/emit[_device])         count -> block-total scan -> emit_scan; count/emit are the two-phase form for
 *                                 callers that size the output first]
 *                                include/genogrove/structure/grove/grove_query.ipp:33-43
 *                                include/genogrove/structure/grove/query_engine.hpp:76-134
 *                                with query_result::add_key (data_type/query_result.hpp:108-113)
 *                                replaced by CSR output: offsets[Q+1] (u64) + hits[H] (u32 ranks)
 *   gg_flanking_*             <- grove::flanking(query, index) -> detail::search_flanking
 *                                include/genogrove/structure/grove/grove_flanking.ipp
 *                                include/genogrove/structure/grove/query_engine.hpp:176-199, 238-311
 *                                flanking_query_result (data_type/flanking_query_result.hpp) replaced by two
 *                                u32 chain ranks per query (GG_NO_RANK = null pointer)
 *   gg_index_set_strands,     <- grove<genomic_coordinate, T>::intersect: genomic_coordinate::overlaps / operator<
 *   gg_intersect_*stranded*      include/genogrove/data_type/genomic_coordinate.hpp:97-110, 156-160
 *   q_chrom (u32 id)          <- grove::get_root(index) hash lookup, grove.hpp:234-236; an id of
 *                                GG_NO_CHROM (or >= n_chrom) yields zero hits, like an unknown index
 *                                name does in grove_query.ipp:37-39
 *   GG_ERR_INVALID_INTERVAL   <- interval ctor's std::invalid_argument on start > end, interval.hpp:70-74
 *
 * Semantics (bit-exact with the reference): for query q on chromosome c the hits are exactly
 * { r : start_c[r] <= q.end && end_c[r] >= q.start } (closed intervals, adjacency overlaps,
 * interval.hpp:140-142), reported as per-chromosome LEAF-CHAIN RANKS r in ascending order — the order
 * search_overlaps appends keys in.  The host keeps rank -> key* per chromosome.
 */
#ifndef GG_B200_H
#define GG_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GG_B200_VERSION 100 /* 0.1.0 */
#define GG_NO_CHROM 0xFFFFFFFFu
#define GG_NO_RANK 0xFFFFFFFFu /* "no such key": absent predecessor / successor of a flanking query */

typedef struct gg_index gg_index;

typedef enum gg_status {
    GG_OK = 0,
    GG_ERR_INVALID_ARG = 1,      /* null pointer, bad size, chrom id >= n_chrom in the index input */
    GG_ERR_CUDA = 2,             /* a CUDA call or kernel failed; see gg_last_error */
    GG_ERR_OOM = 3,              /* device or pinned-host allocation failed */
    GG_ERR_INVALID_INTERVAL = 4, /* start > end in an index interval or a query */
    GG_ERR_NOT_BUILT = 5,        /* intersect before build/upload, or emit before count */
    GG_ERR_UNSORTED = 6,         /* gg_index_upload_sorted input not ordered by (chrom, start) */
    GG_ERR_CAPACITY = 7          /* caller's hit buffer too small / more than 2^32-1 intervals */
} gg_status;

/* Order among intervals with equal (chrom, start, end) — the only freedom in hit order (the leaf chain):
 *   GG_TIE_INPUT_ORDER    stable: equal keys keep input order  == insert_data(idx, vec, sorted, bulk)
 *   GG_TIE_REVERSE_INPUT  equal keys in reverse input order    == one insert_data(idx, key, data) per
 *                         record (lower_bound insertion, node.hpp:247-254)
 *   GG_TIE_EXPLICIT       ascending caller-supplied tie_key[i] (then input order) */
typedef enum gg_tie_mode { GG_TIE_INPUT_ORDER = 0, GG_TIE_REVERSE_INPUT = 1, GG_TIE_EXPLICIT = 2 } gg_tie_mode;

typedef struct gg_index_stats {
    uint64_t n_intervals;
    uint32_t n_chrom;
    uint32_t coord_bytes;   /* 4: u32 device layout (all coordinates < 2^32), 8: u64 layout */
    uint32_t n_levels;      /* levels of the implicit B+ tree, leaf level included */
    uint32_t fanout;        /* keys per node (32) */
    uint32_t staged_levels; /* upper levels copied into shared memory by TMA in the count kernel */
    uint32_t reserved;
    uint64_t device_bytes;  /* HBM held by the index */
    double sort_ms;         /* device time of the radix sorts of the last build */
    double levels_ms;       /* device time of padding + level construction of the last build */
    double total_ms;        /* device time of the last build/upload incl. host<->device copies */
    uint64_t h2d_bytes;     /* interval records copied host -> device by the last build / upload / update_chrom */
} gg_index_stats;

typedef struct gg_batch_stats {
    uint64_t n_queries;
    uint64_t n_hits;
    uint32_t n_chunks;       /* pipeline chunks used by the host-buffer entry points */
    uint32_t kernel_launches;/* kernels launched by the last intersect call */
    double count_ms;         /* device time: count kernel (device-buffer calls with option "timing" on) */
    double scan_ms;          /* device time: exclusive scan (per-query scan of the two-phase API / block-total scan) */
    double emit_ms;          /* device time: emit / emit_scan kernel (the whole kernel for "fused_kernel" = 1) */
    double total_ms;         /* device time first launch/copy -> last, host buffers included */
} gg_batch_stats;

int gg_version(void);
/* number of visible CUDA devices (0 when none or no driver) */
int gg_device_count(void);

/* ---- handle ------------------------------------------------------------------------------------ */
gg_status gg_index_create(gg_index** out, int device);
void gg_index_destroy(gg_index* idx);
/* message of the last failure on this handle ("" if none); valid until the next call on the handle */
const char* gg_last_error(const gg_index* idx);

/* ---- index build (host buffers in) -------------------------------------------------------------- */
/* Sort (chrom, start, end[, tie]) with the hand-written LSD radix sort, build the flattened B+ tree
 * (pivot levels + per-node max-end augmentation) on the device.  rank_to_input (nullable, n entries)
 * receives, chromosome by chromosome in id order, the input index of each chain rank:
 *   rank_to_input[chrom_begin[c] + r] = i  where chrom_begin = exclusive prefix of per-chrom counts. */
gg_status gg_index_build(gg_index* idx, const uint32_t* chrom, const uint64_t* start, const uint64_t* end,
                         const uint64_t* tie_key, int tie_mode, size_t n, uint32_t n_chrom,
                         uint32_t* rank_to_input);
/* Same, for a chain the host already ordered: input must be grouped by ascending chrom id and, inside
 * a chromosome, in leaf-chain order (non-decreasing start).  No sort of (start,end) is performed. */
gg_status gg_index_upload_sorted(gg_index* idx, const uint32_t* chrom, const uint64_t* start,
                                 const uint64_t* end, size_t n, uint32_t n_chrom);
/* gg_index_build with DEVICE buffers in and out (intervals that already live in HBM: the chromosome-sharded group build
 * receives its records over NVLink).  d_rank_to_input: device buffer of n u32, nullable.  The call synchronises. */
gg_status gg_index_build_device(gg_index* idx, const uint32_t* d_chrom, const uint64_t* d_start, const uint64_t* d_end,
                                const uint64_t* d_tie_key, int tie_mode, size_t n, uint32_t n_chrom,
                                uint32_t* d_rank_to_input);
/* Replace ONE chromosome of the snapshot (replaces the facade's "any mutation re-uploads everything": the reference's
 * inserts and removals touch one tree, include/genogrove/structure/grove/grove_insert.ipp:182-274, grove_remove.ipp:24-62).
 * start / end: the chromosome's n records in leaf-chain order (n = 0 empties it; ids of the other chromosomes do not
 * move).  chrom may equal the current n_chrom: a new chromosome id is appended.  Only these n records cross PCIe; the
 * handle splices them into the chain arrays it kept on the device from the last gg_index_upload_sorted / gg_index_build
 * ("keep_build_workspace" = 1, the default) and rebuilds the levels there.  GG_ERR_NOT_BUILT when no such arrays exist
 * (workspace dropped, or the snapshot came from gg_index_build_device).  Strands must be set again afterwards. */
gg_status gg_index_update_chrom(gg_index* idx, uint32_t chrom, const uint64_t* start, const uint64_t* end, size_t n);
gg_status gg_index_get_stats(const gg_index* idx, gg_index_stats* out);
/* counts[c] = number of intervals of chromosome c (n_chrom entries) */
gg_status gg_index_chrom_counts(const gg_index* idx, uint64_t* counts);

/* ---- batched intersect, HOST buffers (the reference-facing call) -------------------------------- */
/* One call = count -> exclusive scan -> emit for nq queries, chunked and pipelined over CUDA streams
 * (H2D of chunk k+1 and D2H of chunk k-1 overlap the kernels of chunk k).  offsets: caller buffer of
 * nq+1 u64.  *hits: library-owned pinned host buffer of *n_hits u32 ranks, valid until the next
 * intersect call on this handle or gg_index_destroy.  Buffers from gg_host_alloc are pinned. */
gg_status gg_intersect_batch(gg_index* idx, const uint32_t* q_chrom, const uint64_t* q_start,
                             const uint64_t* q_end, size_t nq, uint64_t* offsets, const uint32_t** hits,
                             uint64_t* n_hits);
/* u32-coordinate variant (BED/SAM-sized coordinates; needs a coord_bytes == 4 index) */
gg_status gg_intersect_batch_u32(gg_index* idx, const uint32_t* q_chrom, const uint32_t* q_start,
                                 const uint32_t* q_end, size_t nq, uint64_t* offsets, const uint32_t** hits,
                                 uint64_t* n_hits);
/* ---- compact result form of the host-buffer call ------------------------------------------------- */
/* Same hits as gg_intersect_batch_u32, encoded for the PCIe link (the host path is transfer-bound: 8 B of offset + 4 B
 * per hit down, 20 B per query up).  Per query one 8-byte record, per hit a 1-, 2- or 4-byte BACK-DELTA from the query's
 * hi slot (hits of BED/BAM-shaped data sit within a few dozen ranks of it):
 *     k         = q / chunk_queries                       (chunks are uniform except the last)
 *     n_q       = (q + 1 < chunks[k].q_begin + chunks[k].n_queries ? queries[q + 1].off : chunks[k].n_hits) - queries[q].off
 *     delta_j   = the (queries[q].off + j)-th element (chunks[k].delta_bytes wide) of deltas + chunks[k].delta_byte_offset
 *     rank_j    = queries[q].hi_slot - chrom_slot_base[q_chrom[q]] - 1 - delta_j          ascending in j, as always
 * gg_compact_rank / gg_compact_count below are that arithmetic.  The delta width is chosen per chunk: a chunk whose
 * largest delta does not fit is re-run one width up (and the handle keeps the wider width).  All arrays are library-owned
 * pinned memory, valid until the next intersect call on the handle.  Needs a coord_bytes == 4 index. */
typedef struct gg_compact_query {
    uint32_t off;     /* index of the query's first hit inside its chunk */
    uint32_t hi_slot; /* level-0 slot of the first interval starting past q.end */
} gg_compact_query;
typedef struct gg_compact_chunk {
    uint64_t q_begin;
    uint64_t n_queries;
    uint64_t n_hits;
    uint64_t hit_base;          /* hits of all chunks before this one */
    uint64_t delta_byte_offset; /* of this chunk's deltas inside `deltas` (16-byte aligned) */
    uint32_t delta_bytes;       /* 1, 2 or 4 */
    uint32_t reserved;
} gg_compact_chunk;
typedef struct gg_compact_result {
    uint64_t n_queries, n_hits;
    uint64_t chunk_queries;              /* queries per chunk (the last chunk may be shorter) */
    const gg_compact_query* queries;     /* n_queries records */
    const uint8_t* deltas;
    uint64_t delta_bytes_total;
    const gg_compact_chunk* chunks;
    uint32_t n_chunks;
    uint32_t n_chrom;
    const uint32_t* chrom_slot_base;     /* [n_chrom + 1] first level-0 slot of each chromosome */
} gg_compact_result;
gg_status gg_intersect_batch_compact(gg_index* idx, const uint32_t* q_chrom, const uint32_t* q_start,
                                     const uint32_t* q_end, size_t nq, gg_compact_result* out);
/* decode helpers (host arithmetic on the arrays above; no device work) */
uint64_t gg_compact_count(const gg_compact_result* r, uint64_t q);
uint32_t gg_compact_rank(const gg_compact_result* r, const uint32_t* q_chrom, uint64_t q, uint64_t j);
/* expand into the plain CSR of gg_intersect_batch (offsets: nq + 1 u64, hits: n_hits u32); n_threads <= 0: all cores */
gg_status gg_compact_expand(const gg_compact_result* r, const uint32_t* q_chrom, uint64_t* offsets, uint32_t* hits,
                            int n_threads);

/* Two-phase variant with a caller-owned hit buffer: count fills offsets[nq+1] and keeps the batch on the
 * device; emit writes offsets[nq] ranks into hits (capacity in elements). */
gg_status gg_intersect_count(gg_index* idx, const uint32_t* q_chrom, const uint64_t* q_start,
                             const uint64_t* q_end, size_t nq, uint64_t* offsets);
gg_status gg_intersect_emit(gg_index* idx, uint32_t* hits, uint64_t capacity);

/* ---- batched intersect, DEVICE buffers (queries and results stay in HBM) ------------------------ */
/* d_start/d_end element width = coord_bytes (4 or 8; 4 needs a coord_bytes == 4 index).  d_offsets: nq+1
 * u64.  stream: a cudaStream_t (NULL = default stream); kernels are enqueued, not synchronised, unless
 * total_hits (host pointer, nullable) is given, in which case the call waits and stores offsets[nq]. */
gg_status gg_intersect_count_device(gg_index* idx, const uint32_t* d_chrom, const void* d_start,
                                    const void* d_end, int coord_bytes, size_t nq, uint64_t* d_offsets,
                                    uint64_t* total_hits, void* stream);
/* must follow gg_intersect_count_device with the same queries and offsets */
gg_status gg_intersect_emit_device(gg_index* idx, const uint32_t* d_chrom, const void* d_start,
                                   const void* d_end, int coord_bytes, size_t nq, const uint64_t* d_offsets,
                                   uint32_t* d_hits, void* stream);
/* Single-call variant: offsets and hits from one call, every query read once by the count pass (4 consecutive
 * queries per thread), no per-query scan pass: count -> scan of the per-128-query block totals -> emit_scan (offsets +
 * hits).  (Option "fused_kernel" = 1 runs the same thing as ONE kernel with a decoupled look-back scan.)
 * d_hits has room for `capacity` ranks — the caller's estimate, e.g. from the previous batch.  d_offsets[0..nq] is
 * always complete; ranks that would land at or beyond `capacity` are not written.  With total_hits (host pointer) the
 * call waits, stores offsets[nq] and returns GG_ERR_CAPACITY when the batch produced more than `capacity` hits (re-run
 * with a buffer of *total_hits ranks).  Without it the call only enqueues; compare d_offsets[nq] with the capacity
 * yourself. */
gg_status gg_intersect_fused_device(gg_index* idx, const uint32_t* d_chrom, const void* d_start,
                                    const void* d_end, int coord_bytes, size_t nq, uint64_t* d_offsets,
                                    uint32_t* d_hits, uint64_t capacity, uint64_t* total_hits, void* stream);
gg_status gg_last_batch_stats(const gg_index* idx, gg_batch_stats* out);
/* ---- single call without any allocation (no counterpart in the reference) ------------------------------------------ */
/* Nothing in genogrove corresponds to this: it exists for callers of THIS library that may not allocate device memory
 * while they query — a service that gave all of its HBM to its own pool at start-up (a cudaMalloc from here would just
 * fail), a process whose GPU memory is already reserved by another framework's caching allocator, or a latency-bound loop
 * that cannot take cudaMalloc's implicit device synchronisation.  gg_intersect_fused_device grows a scratch buffer on the
 * handle when a batch is larger than any before it; this variant calls no cudaMalloc / cudaFree / cudaHostAlloc and no
 * library routine that does, ever.
 *   - Same arguments, results and status codes as gg_intersect_fused_device (always the one-kernel form, whatever the
 *     "fused_kernel" option says); the index must have been built beforehand, the caller owns every input and output.
 *   - Scratch (tile ticket, one look-back word per 256 queries, the 16-byte error / total mailbox) is a fixed array in the
 *     library's device data segment: 33 KB on every GPU that loads the library, held until the process exits.
 *   - Limit: nq <= GG_STATIC_MAX_QUERIES per call.  More -> GG_ERR_CAPACITY, nothing is launched; split the batch or use
 *     the regular call.
 *   - The first call on each device looks the array's device address up (cudaGetSymbolAddress, with that device
 *     current).  Make that first call during set-up, before any stream capture is started on the device.
 *   - All calls on one GPU — across handles and threads — share the one array, which still holds whatever the previous
 *     call left; each call clears the parts that must start at zero on `stream` before its kernel.  Two calls whose
 *     kernels can overlap on the GPU (different streams with no ordering between them) race on it and return garbage:
 *     order them on one stream or with events.  The host side may return before the kernel ran unless total_hits is given. */
#define GG_STATIC_MAX_QUERIES 1048576
gg_status gg_intersect_fused_device_static(gg_index* idx, const uint32_t* d_chrom, const void* d_start,
                                           const void* d_end, int coord_bytes, size_t nq, uint64_t* d_offsets,
                                           uint32_t* d_hits, uint64_t capacity, uint64_t* total_hits, void* stream);

/* ---- batched flanking (nearest predecessor / successor) ------------------------------------------ */
/* Replaces grove::flanking(query, index) -> detail::search_flanking for interval keys with the default
 * predicate (include/genogrove/structure/grove/grove_flanking.ipp; query_engine.hpp:238-311):
 *   succ[i] = chain rank of the first key (in leaf-chain order) whose start > q.end,
 *   pred[i] = chain rank of the first key in the chain among those with the largest end < q.start
 *             (nested intervals: the largest END wins, not the sort-order maximum, query_engine.hpp:256-272),
 * GG_NO_RANK when there is none (also for q_chrom == GG_NO_CHROM / unknown id: empty result, like the
 * reference's null root).  Overlapping keys are never reported.  Host buffers, nq entries each. */
gg_status gg_flanking_batch(gg_index* idx, const uint32_t* q_chrom, const uint64_t* q_start,
                            const uint64_t* q_end, size_t nq, uint32_t* pred, uint32_t* succ);
/* Device buffers (d_start / d_end element width = coord_bytes as for the intersect calls).  sync != 0: wait and
 * report GG_ERR_INVALID_INTERVAL for a query with start > end; sync == 0: enqueue only. */
gg_status gg_flanking_device(gg_index* idx, const uint32_t* d_chrom, const void* d_start, const void* d_end,
                             int coord_bytes, size_t nq, uint32_t* d_pred, uint32_t* d_succ, int sync,
                             void* stream);

/* Flanking on genomic_coordinate keys with the strand-compatible predicate — grove<genomic_coordinate, T>::flanking(query,
 * index, pred) with the predicate the reference documents (include/genogrove/structure/grove/grove_flanking.ipp:47-54: the
 * candidate's strand equals the query's, or either is '*'; applied at the leaves, query_engine.hpp:251):
 *   succ[i] = chain rank of the first compatible key whose start > q.end,
 *   pred[i] = chain rank of the first key in the chain among the compatible ones with the largest end < q.start.
 * Needs gg_index_set_strands.  (The reference's DEFAULT predicate on genomic_coordinate keys has no flat equivalent: keys
 * that overlap in coordinates but not in strand become candidates and whether the pruning rule, query_engine.hpp:176-199,
 * reaches them depends on the tree's shape — two groves with the same keys disagree, tests/test_flanking.py.)
 * A strand byte outside "+-.*" -> GG_ERR_INVALID_ARG. */
gg_status gg_flanking_batch_stranded(gg_index* idx, const uint32_t* q_chrom, const uint64_t* q_start,
                                     const uint64_t* q_end, const uint8_t* q_strand, size_t nq, uint32_t* pred,
                                     uint32_t* succ);
gg_status gg_flanking_stranded_device(gg_index* idx, const uint32_t* d_chrom, const void* d_start, const void* d_end,
                                      const uint8_t* d_strand, int coord_bytes, size_t nq, uint32_t* d_pred,
                                      uint32_t* d_succ, int sync, void* stream);

/* ---- strand-aware intersect (genomic_coordinate keys) --------------------------------------------- */
/* Replaces grove<genomic_coordinate, T>::intersect(query, index): hits must overlap in coordinates AND have an equal
 * strand, '*' on either side matching anything (include/genogrove/data_type/genomic_coordinate.hpp:156-160); same
 * traversal as the interval path (query_engine.hpp:76-134), so the hit list is the interval hit list filtered by
 * strand, in leaf-chain order.  The chain of genomic_coordinate keys is ordered by (start, end, strand) with
 * '*' < '.' < '+' < '-' (genomic_coordinate.hpp:97-110): build the index with GG_TIE_EXPLICIT and the strand's order
 * (0..3) in the high bits of tie_key, or upload a chain the host ordered.
 * gg_index_set_strands attaches one strand byte ('+', '-', '.', '*') per indexed interval, in CHAIN order with the
 * chromosomes concatenated in id order (the order rank_to_input lists); it must follow every build / upload. */
gg_status gg_index_set_strands(gg_index* idx, const uint8_t* strand, size_t n);
/* Host buffers; as gg_intersect_batch plus q_strand[nq].  A strand byte outside "+-.*" -> GG_ERR_INVALID_ARG (the
 * reference's std::invalid_argument, genomic_coordinate.hpp:78-81). */
gg_status gg_intersect_batch_stranded(gg_index* idx, const uint32_t* q_chrom, const uint64_t* q_start,
                                      const uint64_t* q_end, const uint8_t* q_strand, size_t nq, uint64_t* offsets,
                                      const uint32_t** hits, uint64_t* n_hits);
/* Device buffers; contract of gg_intersect_fused_device (capacity / total_hits / GG_ERR_CAPACITY) plus d_strand. */
gg_status gg_intersect_stranded_device(gg_index* idx, const uint32_t* d_chrom, const void* d_start,
                                       const void* d_end, const uint8_t* d_strand, int coord_bytes, size_t nq,
                                       uint64_t* d_offsets, uint32_t* d_hits, uint64_t capacity,
                                       uint64_t* total_hits, void* stream);

/* ---- multi-GPU: one grove over several B200s ------------------------------------------ */
/* A gg_group owns one gg_index per GPU and the communication between them.  Two forms, same entry points:
 *   single-process  gg_group_create(devices[], n): this process drives all listed GPUs (the C++ facade, gg_isec); data
 *                   moves between them with peer copies over NVLink (cudaMemcpyPeerAsync), no communicator needed.
 *   one process per GPU  gg_group_create_rank(device, rank, world, id): rank `rank` of `world` (torchrun, MPI); the
 *                   handle owns an NCCL communicator created from the 128-byte unique id that rank 0 obtained from
 *                   gg_group_unique_id and the launcher distributed.  libnccl.so.2 is loaded on first use (dlopen), so
 *                   single-GPU users of this library do not need NCCL.
 * Sharding (what the reference calls once per chromosome / once per query — grove_insert.ipp:160-172,
 * grove_query.ipp:33-43 — spread over the GPUs):
 *   GG_SHARD_REPLICATE  every GPU holds the whole index; a query batch is cut into contiguous slices.  No data-path
 *                       exchange; the only collective is the all-gather of the per-rank hit totals (8 bytes per rank).
 *   GG_SHARD_CHROM      chromosomes are bin-packed onto GPUs by interval count (LPT, gg_shard_assign); each GPU sorts /
 *                       builds only its chromosomes; queries are partitioned by owner ON THE DEVICE (stable radix
 *                       partition on the owner id carrying the original query index), exchanged all-to-all over
 *                       NVLink (grouped ncclSend / ncclRecv, or peer copies), answered by the owner, and the hit lists
 *                       travel back the same way and are put back in the original query order on the source GPU.
 * Calls on one group are serialised by the caller; in the rank form every rank makes the same sequence of calls. */
typedef struct gg_group gg_group;
#define GG_SHARD_REPLICATE 0
#define GG_SHARD_CHROM 1

gg_status gg_group_create(gg_group** out, const int* devices, int n_devices);
/* id128: 128 bytes (ncclUniqueId) */
gg_status gg_group_unique_id(void* id128);
gg_status gg_group_create_rank(gg_group** out, int device, int rank, int world, const void* id128);
void gg_group_destroy(gg_group* g);
const char* gg_group_last_error(const gg_group* g);
int gg_group_world(const gg_group* g);
/* the gg_index of local GPU i (single-process form: 0 .. n_devices-1; rank form: 0) — for gg_set_option / stats */
gg_index* gg_group_index(gg_group* g, int i);

/* owner[c] = rank holding chromosome c: longest-processing-time packing of the interval counts (largest chromosome first
 * onto the least loaded rank; ties: lower chromosome id first, lower rank first).  Pure host arithmetic. */
gg_status gg_shard_assign(const uint64_t* counts, uint32_t n_chrom, int world, uint32_t* owner);

/* Build.  Single-process form: the arrays hold ALL intervals.  Rank form: GG_SHARD_REPLICATE: every rank passes all
 * intervals; GG_SHARD_CHROM: every rank passes ITS PART of the input (any chromosomes — e.g. its slice of the BED file);
 * records are routed to the owner of their chromosome over NVLink and each rank builds what it owns.
 * rank_to_input (nullable): for each chain rank, the index of the interval in the global input (rank form: the
 * concatenation of the ranks' inputs in rank order).  Single-process form: n entries, chromosome by chromosome in id
 * order (the layout gg_index_build documents).  Rank form: the chromosomes this rank owns, in id order; *n_owned receives
 * how many (the buffer must hold n_owned_capacity entries; GG_ERR_CAPACITY if it does not). */
gg_status gg_group_build(gg_group* g, int shard_mode, const uint32_t* chrom, const uint64_t* start, const uint64_t* end,
                         const uint64_t* tie_key, int tie_mode, size_t n, uint32_t n_chrom, uint64_t* rank_to_input,
                         size_t n_owned_capacity, size_t* n_owned);
/* owner[c] of the last GG_SHARD_CHROM build (n_chrom entries; all zero for GG_SHARD_REPLICATE) */
gg_status gg_group_owners(const gg_group* g, uint32_t* owner);

typedef struct gg_group_stats {
    uint64_t n_queries;       /* queries this process submitted in the last call */
    uint64_t n_answered;      /* queries its GPUs answered (GG_SHARD_CHROM: the ones they own) */
    uint64_t n_hits;          /* hits in this process's result */
    uint64_t bytes_sent;      /* over NVLink, all exchanges of the call, summed over this process's GPUs */
    uint64_t bytes_received;
    double route_ms;          /* device time, max over this process's GPUs: owner lookup + stable partition + gather */
    double exchange_ms;       /* queries to the owners */
    double intersect_ms;      /* count -> scan -> emit on the owner */
    double return_ms;         /* hit lists back to the source + re-ordering into the original query order */
    double build_ms;          /* last gg_group_build: wall time of the slowest local GPU incl. routing of the records */
} gg_group_stats;
gg_status gg_group_get_stats(const gg_group* g, gg_group_stats* out);

/* Batched intersect, HOST buffers.  Single-process form: the whole batch in, one CSR in original query order out
 * (offsets[nq + 1] global, *hits = library-owned pinned buffer).  Rank form: every rank passes its slice of the batch
 * and gets its slice's CSR (offsets local, starting at 0; *hit_base = hits of the ranks before it — the all-gather of
 * totals).  coord_bytes 4 (u32 arrays) or 8 (u64). */
gg_status gg_group_intersect_batch(gg_group* g, const uint32_t* q_chrom, const void* q_start, const void* q_end,
                                   int coord_bytes, size_t nq, uint64_t* offsets, const uint32_t** hits, uint64_t* n_hits,
                                   uint64_t* hit_base);
/* Batched intersect, DEVICE buffers, rank form (one GPU per process; queries of this rank's slice resident on its GPU,
 * global chromosome ids).  return_to_source != 0: the result is the slice's own CSR in original order, as above.
 * return_to_source == 0 (GG_SHARD_CHROM): hit lists stay on the GPU that computed them — what a downstream kernel on the
 * owner wants, and the only option when the hits of a batch exceed one GPU's memory share (cfg5: 1 B queries x 140 hits):
 * the result describes the queries THIS rank answered, in arrival order, d_origin[i] = (source rank << 40) | index of the
 * query in the source's slice.  Result buffers are library-owned device memory, valid until the next call. */
typedef struct gg_group_device_result {
    uint64_t n_queries;
    uint64_t n_hits;
    uint64_t hit_base;         /* return_to_source: hits of the ranks before this one; else 0 */
    const uint64_t* d_offsets; /* n_queries + 1 */
    const uint32_t* d_hits;
    const uint64_t* d_origin;  /* n_queries entries when return_to_source == 0, else NULL */
} gg_group_device_result;
gg_status gg_group_intersect_device(gg_group* g, const uint32_t* d_chrom, const void* d_start, const void* d_end,
                                    int coord_bytes, size_t nq, int return_to_source, gg_group_device_result* out);

/* ---- pinned host memory helpers ------------------------------------------------------------------ */
gg_status gg_host_alloc(void** out, size_t bytes);
void gg_host_free(void* p);

/* ---- tuning knobs (defaults are chosen for B200; used by tests/bench to pin behaviour) ----------- */
/* key: "chunk_queries" (host pipeline chunk, default 8388608); "host_mode" (host-buffer calls: 0 = the single-call
 * path per chunk with an estimated hit capacity [default], 1 = count, read the total, emit); "fused_kernel" (single-call
 * path: 0 = count -> block scan -> emit_scan kernels, 1 = one kernel with a look-back scan, -1 = by batch size [default]:
 * one kernel up to 524288 queries, where launch latency dominates (10 k queries: 46 us against 66 us); its knobs:
 * "fused_ctas_per_sm", "fused_regs40"); "emit_walk" (ranks per walk step of emit_scan: 8 [default] or 16);
 * "emit_tma" (pass 2 of the single-call path: 0 = intersect_emit_scan_kernel [default], 1 = intersect_emit_tma_kernel: TMA
 * ring of the per-tile inputs + TMA bulk store of the hits, 2 = the same with 16-byte vector stores; the compact call
 * always runs the TMA kernel);
 * "dense_emit" (dense-batch variants of pass 2: a 1024-hit stage per 32-query tile instead of 512, and a second emit kernel
 * for tiles with >= 2048 hits, per-lane stage segments and sector-sized stores: -1 = on once a batch / chunk on this
 * handle averaged >= 12 hits per query (stranded path: >= 32) [default], 0 = off, 1 = on);
 * "count_mode" (two-phase API: 2 = directory + finger search, 4 consecutive queries per thread [default], 0 = directory,
 * one query per thread, 1 = tiered tree kernel: TMA-staged windows / warp bracket / per-lane descent; for mode 1:
 * "stage_bytes" (shared-memory budget for TMA-staged upper levels, default 16384) and "count_threads" (128..1024));
 * "count_ctas_per_sm", "emit_ctas_per_sm" (0 = occupancy API); "timing" (0/1: record CUDA events around the kernels of
 * the device-buffer calls for gg_last_batch_stats); "compact_delta_bytes" (1, 2, 4: width the compact call starts with);
 * "keep_build_workspace" (1 [default]: the sort buffers of gg_index_build stay allocated between builds — a facade that
 * re-syncs after every mutation saves 60 % of a rebuild; 0: freed after each build);
 * "coop_budget" (dense-batch pass 2: per-lane walk steps before the warp finishes a query's list together — 256 slots per
 * step through the per-8-slot maxima; 32 [default], 0 = per-lane walks only); "sort_geom" (onesweep tile: 0 = 256 threads
 * x 12 keys [default], 1 = 512 x 12, 2 = 256 x 8);
 * "delta_min_hits" (dense-batch pass 2: coordinate-sorted 32-query tiles with at least this many hits share one
 * incrementally kept candidate list instead of 32 walks; 1024 [default], 0 = off); "emit_vec_out", "emit_tile_list" (0
 * [default] / 1: measured variants of pass 2); "debug_checks" (1: gg_intersect_emit_device synchronises and
 * returns GG_ERR_INVALID_ARG when the offsets it was given are not the ones gg_intersect_count_device wrote — the kernel
 * always skips such tiles instead of trusting them; 0 [default]: the call stays asynchronous);
 * "sort_legacy" (0 [default]: onesweep radix passes — one kernel per 8-bit digit with a decoupled look-back; 1: the
 * three-kernel passes of round 1 (per-CTA histogram, scan, scatter), which builds of 2^30 intervals or more always use);
 * "inject_alloc_failure" (debug: the k-th device allocation from now on fails, to drive the error paths). */
gg_status gg_set_option(gg_index* idx, const char* key, int64_t value);

#ifdef __cplusplus
}
#endif
#endif /* GG_B200_H */
