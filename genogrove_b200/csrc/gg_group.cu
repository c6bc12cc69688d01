// gg_group.cu — one grove over several B200s: the multi-GPU part of the C ABI (include/gg_b200.h, "gg_group_*").
//
// What it parallelises in the reference: one insert_data(idx, vec, bulk) per chromosome
// (include/genogrove/structure/grove/grove_insert.ipp:160-172) and one intersect(q, index) per query
// (grove_query.ipp:33-43) — chromosomes are independent trees, queries are independent.  SURVEY §8e.
//
// Layering: per GPU this file only uses the public single-GPU C ABI (gg_index_*, gg_intersect_fused_device) plus its own
// routing kernels below.  Communication is either peer copies over NVLink between the GPUs of one process
// (cudaMemcpyPeerAsync) or grouped ncclSend / ncclRecv between processes; NCCL is bound at run time (dlopen), so
// libgg_b200.so has no link-time dependency on it.
//
// The all-to-all of GG_SHARD_CHROM, per call and per GPU r (W = world size):
//   route    dest[i] = owner[chrom[i]], local chromosome id; stable radix partition of the query indices by dest (the LSD
//            radix sort of radix_sort.cuh on the owner id, carrying the original index); gather (chrom, start, end) into
//            send buffers grouped by destination; bucket bounds -> host.
//   counts   every rank learns how much each peer sends it (W x W matrix: host memory / ncclAllGather of W u64).
//   exchange all-to-all-v of the three columns (+ the original index when results stay on the owner).
//   answer   gg_intersect_fused_device on the received queries.
//   return   per-query counts and the hit lists travel back grouped by source; the source scatters the counts through its
//            partition permutation, scans them into offsets and gathers the lists into original query order.
// Bytes over NVLink per query: 12 (u32 coordinates) or 20 out, 4 + 4 x hits back.
#include "../../include/gg_b200.h"

#include <dlfcn.h>
#include <nccl.h>  // types and prototypes only; the entry points are resolved with dlsym

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <new>
#include <numeric>
#include <string>
#include <thread>
#include <vector>

#include "common.cuh"
#include "device_buf.h"
#include "radix_sort.cuh"
#include "scan.cuh"

using namespace gg;

namespace {

// ------------------------------------------------------------------------------------------ NCCL, bound at run time
struct NcclApi {
    void* lib = nullptr;
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclGroupStart) GroupStart = nullptr;
    decltype(&ncclGroupEnd) GroupEnd = nullptr;
    decltype(&ncclSend) Send = nullptr;
    decltype(&ncclRecv) Recv = nullptr;
    decltype(&ncclAllGather) AllGather = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
    bool ok = false;
};

NcclApi& nccl() {
    static NcclApi a = [] {
        NcclApi x;
        // a process that already loaded an NCCL (torch bundles one under the same soname) gets that copy back
        x.lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!x.lib) x.lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!x.lib) return x;
#define GG_NCCL_SYM(name) x.name = reinterpret_cast<decltype(x.name)>(dlsym(x.lib, "nccl" #name))
        GG_NCCL_SYM(GetUniqueId);
        GG_NCCL_SYM(CommInitRank);
        GG_NCCL_SYM(CommDestroy);
        GG_NCCL_SYM(GroupStart);
        GG_NCCL_SYM(GroupEnd);
        GG_NCCL_SYM(Send);
        GG_NCCL_SYM(Recv);
        GG_NCCL_SYM(AllGather);
        GG_NCCL_SYM(GetErrorString);
#undef GG_NCCL_SYM
        x.ok = x.GetUniqueId && x.CommInitRank && x.CommDestroy && x.GroupStart && x.GroupEnd && x.Send && x.Recv && x.AllGather;
        return x;
    }();
    return a;
}

// ------------------------------------------------------------------------------------------ routing kernels
constexpr uint32_t kNoChrom = 0xFFFFFFFFu;

// dest[i] = rank that answers query / owns record i, lchrom[i] = its chromosome id on that rank.  Unknown chromosomes
// (id >= n_chrom, GG_NO_CHROM) have no owner: they stay on `self` with local id GG_NO_CHROM and get zero hits there, like
// an unknown index name does in grove_query.ipp:37-39.
__global__ void route_kernel(const uint32_t* __restrict__ chrom, size_t n, const uint32_t* __restrict__ owner,
                             const uint32_t* __restrict__ local_id, uint32_t n_chrom, uint32_t self,
                             uint32_t* __restrict__ dest, uint32_t* __restrict__ lchrom, uint32_t* __restrict__ iota) {
    const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t c = chrom[i];
    const bool known = c < n_chrom;
    dest[i] = known ? owner[c] : self;
    lchrom[i] = known ? local_id[c] : kNoChrom;
    if (iota) iota[i] = static_cast<uint32_t>(i);
}

// bounds[b] = first position of the sorted destination ids holding a value >= b (b = 0 .. W)
__global__ void bucket_bounds_kernel(const uint32_t* __restrict__ sorted_dest, size_t n, uint32_t W, uint64_t* __restrict__ bounds) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b > W) return;
    size_t lo = 0, hi = n;
    while (lo < hi) {
        const size_t mid = (lo + hi) >> 1;
        if (sorted_dest[mid] < b) lo = mid + 1; else hi = mid;
    }
    bounds[b] = lo;
}

template <typename T>
__global__ void gather_kernel(const uint32_t* __restrict__ perm, const T* __restrict__ in, size_t n, T* __restrict__ out) {
    const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) out[i] = in[perm ? perm[i] : static_cast<uint32_t>(i)];
}

__global__ void diff_kernel(const uint64_t* __restrict__ off, size_t n, uint32_t* __restrict__ cnt) {
    const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) cnt[i] = static_cast<uint32_t>(off[i + 1] - off[i]);
}

// out[perm[j]] = in[j]
__global__ void scatter_u32_kernel(const uint32_t* __restrict__ perm, const uint32_t* __restrict__ in, size_t n, uint32_t* __restrict__ out) {
    const size_t j = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (j < n) out[perm ? perm[j] : static_cast<uint32_t>(j)] = in[j];
}

__global__ void add_base_kernel(uint64_t* __restrict__ off, size_t n, uint64_t base) {
    const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) off[i] += base;
}

// origin[i] = (source rank << 40) | index of the query in the source's slice; seg[p] = first received query of source p
__global__ void origin_kernel(const uint32_t* __restrict__ orig, size_t n, const uint64_t* __restrict__ seg, uint32_t W,
                              uint64_t* __restrict__ origin) {
    const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t lo = 0, hi = W;  // last p with seg[p] <= i (seg[0] = 0, non-decreasing; empty sources share a value)
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (seg[mid] <= i) lo = mid; else hi = mid;
    }
    origin[i] = (static_cast<uint64_t>(lo) << 40) | orig[i];
}

// Hit lists back into original query order.  Source lists are in partition order j (b_off = scan of the counts that came
// back), destination lists in original order (f_off = scan of the scattered counts).  One warp takes 32 consecutive j: their
// source ranges are one contiguous run, read coalesced; each element finds its list by a 5-step search over the warp's 33
// offsets held in registers.
__global__ void __launch_bounds__(256)
gather_lists_kernel(const uint32_t* __restrict__ perm, const uint64_t* __restrict__ b_off, const uint64_t* __restrict__ f_off,
                    const uint32_t* __restrict__ b_hits, size_t n, uint32_t* __restrict__ f_hits) {
    const uint32_t lane = lane_id();
    const size_t warps = (static_cast<size_t>(gridDim.x) * blockDim.x) >> 5;
    for (size_t j0 = ((static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5) * 32; j0 < n; j0 += warps * 32) {
        const size_t j = j0 + lane;
        const bool in = j < n;
        const uint64_t s0 = b_off[j0];
        const uint32_t rel = in ? static_cast<uint32_t>(b_off[j] - s0) : 0xFFFFFFFFu;  // (a 32-list run has < 2^32 hits)
        const uint64_t dst = in ? f_off[perm ? perm[j] : static_cast<uint32_t>(j)] : 0;
        const size_t jl = j0 + 32 < n ? j0 + 32 : n;
        const uint32_t total = static_cast<uint32_t>(b_off[jl] - s0);
        for (uint32_t p = lane; p < ((total + 31u) & ~31u); p += 32) {
            // k = last list whose relative start is <= p
            uint32_t k = 0;
#pragma unroll
            for (int step = 16; step >= 1; step >>= 1) {
                const uint32_t r = __shfl_sync(0xffffffffu, rel, k + step);
                if (r <= p) k += step;
            }
            const uint32_t rk = __shfl_sync(0xffffffffu, rel, k);
            const uint64_t dk = __shfl_sync(0xffffffffu, dst, k);
            if (p < total) f_hits[dk + (p - rk)] = b_hits[s0 + p];
        }
    }
}

inline unsigned blocks_for(size_t n, unsigned threads) { return static_cast<unsigned>(div_up(n ? n : 1, threads)); }

// ------------------------------------------------------------------------------------------ per-GPU context
struct Rk {
    int dev = 0;
    int rank = 0;
    gg_index* idx = nullptr;
    cudaStream_t st = nullptr;
    cudaEvent_t ev_ready = nullptr, ev_done = nullptr;
    cudaEvent_t ev_t[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    DeviceBuf d_owner, d_local;                                     // [n_chrom] routing tables
    DeviceBuf in_c, in_s, in_e;                                     // host-buffer call: the slice on the device
    DeviceBuf dest, dest_alt, perm, perm_alt, lchrom, hist, scanst, d_bounds;
    DeviceBuf send_c, send_s, send_e, send_o;                       // partition order
    DeviceBuf recv_c, recv_s, recv_e, recv_o;                       // grouped by source rank
    DeviceBuf r_off, r_hits, r_cnt, origin, d_seg;                  // owner-side result
    DeviceBuf b_cnt, b_off, b_hits, f_cnt, f_off, f_hits, scan2;    // source side
    DeviceBuf d_mat;                                                // W x W count matrix (rank form)
    DeviceBuf d_agree;                                              // W status words (rank form), allocated at creation
    uint64_t* h_box = nullptr;                                      // pinned: [W + 1] bounds + W x W matrix + scratch
    std::vector<uint64_t> soff, roff, hsoff, hroff;                 // W + 1 each
    size_t nq = 0, n_recv = 0;
    uint64_t n_hits_owner = 0, n_hits_final = 0;
    double hits_per_query = 8.0;
    uint64_t bytes_sent = 0, bytes_received = 0;
    const uint32_t* q_c = nullptr;                                  // current call's device inputs
    const void* q_s = nullptr;
    const void* q_e = nullptr;
    const uint32_t* perm_ptr = nullptr;                             // nullptr = identity (W == 1)
};

}  // namespace

struct gg_group {
    int world = 1;
    bool rank_form = false;
    std::vector<Rk> rk;
    ncclComm_t comm = nullptr;
    std::string err;
    int shard_mode = GG_SHARD_REPLICATE;
    bool built = false;
    uint32_t n_chrom = 0;
    std::vector<uint32_t> owner, local_id;
    gg_group_stats stats{};
    uint32_t* pinned_hits = nullptr;  // single-process host-buffer call
    size_t pinned_cap = 0;
};

namespace {

gg_status gfail(gg_group* g, gg_status st, const std::string& what, cudaError_t e = cudaSuccess) {
    if (g) {
        g->err = what;
        if (e != cudaSuccess) {
            g->err += ": ";
            g->err += cudaGetErrorString(e);
        }
    }
    if (e == cudaErrorMemoryAllocation) {
        cudaGetLastError();
        return GG_ERR_OOM;
    }
    return st;
}

#define GGG_CUDA(call)                                                   \
    do {                                                                 \
        cudaError_t e_ = (call);                                         \
        if (e_ != cudaSuccess) return gfail(g, GG_ERR_CUDA, #call, e_);  \
    } while (0)
#define GGG_NCCL(call)                                                                                       \
    do {                                                                                                     \
        ncclResult_t r_ = (call);                                                                            \
        if (r_ != ncclSuccess)                                                                               \
            return gfail(g, GG_ERR_CUDA, std::string(#call ": ") + (nccl().GetErrorString ? nccl().GetErrorString(r_) : "nccl error")); \
    } while (0)
#define GGG_IDX(r, call)                                                                                  \
    do {                                                                                                  \
        gg_status s_ = (call);                                                                            \
        if (s_ != GG_OK) return gfail(g, s_, std::string(#call ": ") + gg_last_error((r).idx));           \
    } while (0)

gg_status init_rank_ctx(gg_group* g, Rk& r, int dev, int rank) {
    r.dev = dev;
    r.rank = rank;
    gg_status s = gg_index_create(&r.idx, dev);
    if (s != GG_OK) return gfail(g, s, "gg_index_create: no usable CUDA device");
    GGG_CUDA(cudaSetDevice(dev));
    GGG_CUDA(cudaStreamCreateWithFlags(&r.st, cudaStreamNonBlocking));
    GGG_CUDA(cudaEventCreateWithFlags(&r.ev_ready, cudaEventDisableTiming));
    GGG_CUDA(cudaEventCreateWithFlags(&r.ev_done, cudaEventDisableTiming));
    for (auto& e : r.ev_t) GGG_CUDA(cudaEventCreate(&e));
    const size_t W = static_cast<size_t>(g->world);
    GGG_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&r.h_box), (2 * (W + 1) + W * W + 16 + W + 2) * 8, cudaHostAllocDefault));
    if (g->rank_form) GGG_CUDA(r.d_agree.ensure((W + 1) * 8));  // agree() must not allocate: it runs when allocations fail
    r.soff.assign(W + 1, 0);
    r.roff.assign(W + 1, 0);
    r.hsoff.assign(W + 1, 0);
    r.hroff.assign(W + 1, 0);
    return GG_OK;
}

// all-gather of W u64 per rank -> mat[p * W + k] = value k of rank p.  Single-process form: host memory.
gg_status gather_matrix(gg_group* g, const std::vector<std::vector<uint64_t>>& mine /* per local rank: W values */,
                        std::vector<uint64_t>& mat) {
    const size_t W = static_cast<size_t>(g->world);
    mat.assign(W * W, 0);
    if (!g->rank_form) {
        for (size_t p = 0; p < W; ++p) std::copy(mine[p].begin(), mine[p].end(), mat.begin() + p * W);
        return GG_OK;
    }
    Rk& r = g->rk[0];
    GGG_CUDA(cudaSetDevice(r.dev));
    GGG_CUDA(r.d_mat.ensure(W * W * 8));
    uint64_t* h = r.h_box + 2 * (W + 1);
    std::copy(mine[0].begin(), mine[0].end(), h);
    uint64_t* d = r.d_mat.as<uint64_t>();
    GGG_CUDA(cudaMemcpyAsync(d + static_cast<size_t>(r.rank) * W, h, W * 8, cudaMemcpyHostToDevice, r.st));
    GGG_NCCL(nccl().AllGather(d + static_cast<size_t>(r.rank) * W, d, W, ncclUint64, g->comm, r.st));
    GGG_CUDA(cudaMemcpyAsync(h, d, W * W * 8, cudaMemcpyDeviceToHost, r.st));
    GGG_CUDA(cudaStreamSynchronize(r.st));
    std::copy(h, h + W * W, mat.begin());
    return GG_OK;
}

// Rank form: every rank learns the worst status any rank holds, so that all of them leave a call together.  Without it a
// rank that fails a local check or an allocation returns while the others wait for it in the next collective, forever.
// Called by every rank at the same points — before each data-path collective — with what the rank did since the last
// one; costs one all-gather of 8 bytes per rank.  Single-process form: nothing to agree on.
gg_status agree(gg_group* g, gg_status local) {
    if (!g->rank_form || g->world == 1) return local;
    Rk& r = g->rk[0];
    const size_t W = static_cast<size_t>(g->world);
    const std::string local_msg = g->err;
    GGG_CUDA(cudaSetDevice(r.dev));
    uint64_t* h = r.h_box + 2 * (W + 1) + W * W + 16;
    uint64_t* d = r.d_agree.as<uint64_t>();
    h[W] = static_cast<uint64_t>(local);
    GGG_CUDA(cudaMemcpyAsync(d + r.rank, h + W, 8, cudaMemcpyHostToDevice, r.st));
    GGG_NCCL(nccl().AllGather(d + r.rank, d, 1, ncclUint64, g->comm, r.st));
    GGG_CUDA(cudaMemcpyAsync(h, d, W * 8, cudaMemcpyDeviceToHost, r.st));
    GGG_CUDA(cudaStreamSynchronize(r.st));
    if (local != GG_OK) {
        g->err = local_msg;
        return local;
    }
    for (size_t p = 0; p < W; ++p)
        if (h[p] != GG_OK) return gfail(g, static_cast<gg_status>(h[p]), "rank " + std::to_string(p) + " failed this call (see its gg_group_last_error)");
    return GG_OK;
}

// One column of an all-to-all-v.  send(r) / recv(r): device pointers of rank r; so(r) / ro(r): its W + 1 element offsets by
// destination / by source.
struct Col {
    void* (*send)(Rk&);
    void* (*recv)(Rk&);
    size_t elem;
};
using OffFn = const std::vector<uint64_t>& (*)(Rk&);

gg_status all_to_all(gg_group* g, const std::vector<Col>& cols, OffFn so, OffFn ro) {
    const int W = g->world;
    if (g->rank_form) {
        Rk& r = g->rk[0];
        GGG_CUDA(cudaSetDevice(r.dev));
        const auto& s = so(r);
        const auto& o = ro(r);
        GGG_NCCL(nccl().GroupStart());
        for (const Col& c : cols) {
            char* sp = static_cast<char*>(c.send(r));
            char* rp = static_cast<char*>(c.recv(r));
            for (int p = 0; p < W; ++p) {
                const uint64_t ns = s[p + 1] - s[p], nr = o[p + 1] - o[p];
                if (ns) {
                    GGG_NCCL(nccl().Send(sp + s[p] * c.elem, ns * c.elem, ncclUint8, p, g->comm, r.st));
                    if (p != r.rank) r.bytes_sent += ns * c.elem;
                }
                if (nr) {
                    GGG_NCCL(nccl().Recv(rp + o[p] * c.elem, nr * c.elem, ncclUint8, p, g->comm, r.st));
                    if (p != r.rank) r.bytes_received += nr * c.elem;
                }
            }
        }
        GGG_NCCL(nccl().GroupEnd());
        return GG_OK;
    }
    // single process: the destination GPU pulls every segment with a peer copy on its own stream
    for (Rk& s : g->rk) {
        GGG_CUDA(cudaSetDevice(s.dev));
        GGG_CUDA(cudaEventRecord(s.ev_ready, s.st));
    }
    for (Rk& d : g->rk) {
        GGG_CUDA(cudaSetDevice(d.dev));
        const auto& o = ro(d);
        for (Rk& s : g->rk) {
            const auto& so_s = so(s);
            const uint64_t n = so_s[d.rank + 1] - so_s[d.rank];
            if (n == 0) continue;
            if (n != o[s.rank + 1] - o[s.rank]) return gfail(g, GG_ERR_CUDA, "all_to_all: send / receive counts disagree");
            GGG_CUDA(cudaStreamWaitEvent(d.st, s.ev_ready, 0));
            for (const Col& c : cols) {
                char* sp = static_cast<char*>(c.send(s)) + so_s[d.rank] * c.elem;
                char* rp = static_cast<char*>(c.recv(d)) + o[s.rank] * c.elem;
                if (s.dev == d.dev)
                    GGG_CUDA(cudaMemcpyAsync(rp, sp, n * c.elem, cudaMemcpyDeviceToDevice, d.st));
                else
                    GGG_CUDA(cudaMemcpyPeerAsync(rp, d.dev, sp, s.dev, n * c.elem, d.st));
                if (s.rank != d.rank) {
                    s.bytes_sent += n * c.elem;
                    d.bytes_received += n * c.elem;
                }
            }
        }
        GGG_CUDA(cudaEventRecord(d.ev_done, d.st));
    }
    for (Rk& s : g->rk) {  // a sender may reuse its buffers once every receiver has pulled
        GGG_CUDA(cudaSetDevice(s.dev));
        for (Rk& d : g->rk) GGG_CUDA(cudaStreamWaitEvent(s.st, d.ev_done, 0));
    }
    return GG_OK;
}

void lpt_assign(const uint64_t* counts, uint32_t n_chrom, int world, uint32_t* owner) {
    std::vector<uint32_t> order(n_chrom);
    std::iota(order.begin(), order.end(), 0u);
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return counts[a] > counts[b]; });
    std::vector<uint64_t> load(static_cast<size_t>(world), 0);
    for (uint32_t c : order) {
        int best = 0;
        for (int r = 1; r < world; ++r)
            if (load[r] < load[best]) best = r;
        owner[c] = static_cast<uint32_t>(best);
        load[best] += counts[c];
    }
}

// owner / local-id tables of a GG_SHARD_CHROM build onto every local GPU
gg_status upload_routing(gg_group* g) {
    const uint32_t C = g->n_chrom;
    g->local_id.assign(C, 0);
    std::vector<uint32_t> next(static_cast<size_t>(g->world), 0);
    for (uint32_t c = 0; c < C; ++c) g->local_id[c] = next[g->owner[c]]++;
    for (Rk& r : g->rk) {
        GGG_CUDA(cudaSetDevice(r.dev));
        GGG_CUDA(r.d_owner.ensure(static_cast<size_t>(C) * 4));
        GGG_CUDA(r.d_local.ensure(static_cast<size_t>(C) * 4));
        GGG_CUDA(cudaMemcpyAsync(r.d_owner.p, g->owner.data(), static_cast<size_t>(C) * 4, cudaMemcpyHostToDevice, r.st));
        GGG_CUDA(cudaMemcpyAsync(r.d_local.p, g->local_id.data(), static_cast<size_t>(C) * 4, cudaMemcpyHostToDevice, r.st));
        GGG_CUDA(cudaStreamSynchronize(r.st));
    }
    return GG_OK;
}

// route: dest / local chromosome per element, stable partition of the element indices by dest, bucket bounds -> r.soff.
// Leaves r.perm_ptr (partition order -> original index; nullptr = identity when W == 1) and r.lchrom (original order).
gg_status route_phase(gg_group* g, Rk& r, const uint32_t* d_chrom, size_t n) {
    const uint32_t W = static_cast<uint32_t>(g->world);
    GGG_CUDA(cudaSetDevice(r.dev));
    const size_t nn = n ? n : 1;
    GGG_CUDA(r.dest.ensure(nn * 4));
    GGG_CUDA(r.lchrom.ensure(nn * 4));
    GGG_CUDA(r.d_bounds.ensure((W + 1) * 8));
    r.perm_ptr = nullptr;
    if (W > 1) {
        GGG_CUDA(r.dest_alt.ensure(nn * 4));
        GGG_CUDA(r.perm.ensure(nn * 4));
        GGG_CUDA(r.perm_alt.ensure(nn * 4));
        GGG_CUDA(r.hist.ensure(radix_hist_entries(n) * 4));
        GGG_CUDA(r.scanst.ensure(scan_state_bytes(radix_hist_entries(n))));
    }
    if (n) {
        route_kernel<<<blocks_for(n, 256), 256, 0, r.st>>>(d_chrom, n, r.d_owner.as<uint32_t>(), r.d_local.as<uint32_t>(), g->n_chrom,
                                                           static_cast<uint32_t>(r.rank), r.dest.as<uint32_t>(),
                                                           r.lchrom.as<uint32_t>(), W > 1 ? r.perm.as<uint32_t>() : nullptr);
        GGG_CUDA(cudaGetLastError());
    }
    const uint32_t* sorted_dest = r.dest.as<uint32_t>();
    if (W > 1 && n) {
        RadixWorkspace ws;
        ws.hist = r.hist.as<uint32_t>();
        ws.scan_scratch = r.scanst.p;
        bool in_alt = false;
        GGG_CUDA(radix_sort_pairs<uint32_t>(r.dest.as<uint32_t>(), r.dest_alt.as<uint32_t>(), r.perm.as<uint32_t>(),
                                            r.perm_alt.as<uint32_t>(), n, std::max(1, bits_needed(W - 1)), ws, r.st, &in_alt));
        sorted_dest = in_alt ? r.dest_alt.as<uint32_t>() : r.dest.as<uint32_t>();
        r.perm_ptr = in_alt ? r.perm_alt.as<uint32_t>() : r.perm.as<uint32_t>();
    }
    bucket_bounds_kernel<<<blocks_for(static_cast<size_t>(W) + 1, 256), 256, 0, r.st>>>(sorted_dest, n, W, r.d_bounds.as<uint64_t>());
    GGG_CUDA(cudaGetLastError());
    GGG_CUDA(cudaMemcpyAsync(r.h_box, r.d_bounds.p, (W + 1) * 8, cudaMemcpyDeviceToHost, r.st));
    return GG_OK;
}

template <typename T>
gg_status gather_col(gg_group* g, Rk& r, const T* in, size_t n, DeviceBuf& out) {
    GGG_CUDA(out.ensure((n ? n : 1) * sizeof(T)));
    if (n) {
        gather_kernel<T><<<blocks_for(n, 256), 256, 0, r.st>>>(r.perm_ptr, in, n, out.as<T>());
        GGG_CUDA(cudaGetLastError());
    }
    return GG_OK;
}

// receive offsets of every local rank from the W x W send-count matrix (mat[p * W + d] = elements p sends to d)
void recv_offsets_from(const std::vector<uint64_t>& mat, int W, int rank, std::vector<uint64_t>& roff) {
    roff.assign(static_cast<size_t>(W) + 1, 0);
    for (int p = 0; p < W; ++p) roff[p + 1] = roff[p] + mat[static_cast<size_t>(p) * W + rank];
}

double ms_between(cudaEvent_t a, cudaEvent_t b) {
    float ms = 0;
    if (cudaEventElapsedTime(&ms, a, b) != cudaSuccess) {
        cudaGetLastError();
        return 0.0;
    }
    return ms;
}

// The device-side core shared by the host-buffer and the device-buffer calls: every local rank r has its slice on its GPU
// (r.q_c / r.q_s / r.q_e / r.nq).  On return r.f_off / r.f_hits (return_to_source) or r.r_off / r.r_hits / r.origin describe
// the result; streams are synchronised.
// `pre`: what the caller's own preparation (argument checks, upload of the slice) ended with; it takes part in the first
// agreement, so a rank that failed there does not leave the others waiting.  Each phase between two collectives runs as a
// lambda whose status is agreed on (agree()) before the collective that follows it.
gg_status intersect_core(gg_group* g, int coord_bytes, bool return_to_source, gg_status pre = GG_OK) {
    const int W = g->world;
    const size_t cb = static_cast<size_t>(coord_bytes);
    const bool chrom_mode = g->shard_mode == GG_SHARD_CHROM;
    gg_status ph = pre;
    if (ph == GG_OK) ph = [&]() -> gg_status {
        for (Rk& r : g->rk) {
            r.bytes_sent = r.bytes_received = 0;
            GGG_CUDA(cudaSetDevice(r.dev));
            GGG_CUDA(cudaEventRecord(r.ev_t[0], r.st));
        }
        return GG_OK;
    }();
    if (!chrom_mode) {
      if (ph == GG_OK) ph = [&]() -> gg_status {
        // ---- replicated index: answer the own slice, no exchange
        for (Rk& r : g->rk) {
            GGG_CUDA(cudaSetDevice(r.dev));
            GGG_CUDA(r.r_off.ensure((r.nq + 1) * 8));
            GGG_CUDA(cudaEventRecord(r.ev_t[1], r.st));
            GGG_CUDA(cudaEventRecord(r.ev_t[2], r.st));
        }
        // enqueue on every GPU first (no host wait), then collect the totals
        for (int attempt = 0; attempt < 2; ++attempt) {
            bool again = false;
            for (Rk& r : g->rk) {
                if (attempt == 1 && r.n_hits_owner <= r.r_hits.cap / 4) continue;
                GGG_CUDA(cudaSetDevice(r.dev));
                const size_t want = attempt == 0 ? static_cast<size_t>(r.hits_per_query * static_cast<double>(r.nq)) + 4096
                                                 : static_cast<size_t>(r.n_hits_owner);
                if (r.r_hits.cap / 4 < want) GGG_CUDA(r.r_hits.ensure((want + want / 4) * 4));  // (slack only when growing, see below)
                GGG_IDX(r, gg_intersect_fused_device(r.idx, r.q_c, r.q_s, r.q_e, coord_bytes, r.nq, r.r_off.as<uint64_t>(),
                                                     r.r_hits.as<uint32_t>(), r.r_hits.cap / 4, nullptr, r.st));
            }
            for (Rk& r : g->rk) {
                if (attempt == 1 && r.n_hits_owner <= r.r_hits.cap / 4 && r.n_hits_final == r.n_hits_owner) continue;
                GGG_CUDA(cudaSetDevice(r.dev));
                GGG_CUDA(cudaMemcpyAsync(r.h_box, r.r_off.as<uint64_t>() + r.nq, 8, cudaMemcpyDeviceToHost, r.st));
                GGG_CUDA(cudaStreamSynchronize(r.st));
                r.n_hits_owner = r.h_box[0];
                if (r.n_hits_owner > r.r_hits.cap / 4) again = true; else r.n_hits_final = r.n_hits_owner;
            }
            if (!again) break;
        }
        for (Rk& r : g->rk) {
            GGG_CUDA(cudaSetDevice(r.dev));
            GGG_CUDA(cudaEventRecord(r.ev_t[3], r.st));
            GGG_CUDA(cudaEventRecord(r.ev_t[4], r.st));
            r.n_recv = r.nq;
            r.n_hits_final = r.n_hits_owner;
            if (r.nq >= 65536) r.hits_per_query = std::max(1.0, static_cast<double>(r.n_hits_owner) / static_cast<double>(r.nq));
        }
        return GG_OK;
      }();
        ph = agree(g, ph);
        if (ph != GG_OK) return ph;
        // the path's one collective: hit totals of every rank (8 bytes each) -> global CSR base of each slice
        std::vector<std::vector<uint64_t>> mine(g->rk.size(), std::vector<uint64_t>(static_cast<size_t>(W), 0));
        for (size_t i = 0; i < g->rk.size(); ++i) mine[i][0] = g->rk[i].n_hits_owner;
        std::vector<uint64_t> mat;
        gg_status s = gather_matrix(g, mine, mat);
        if (s != GG_OK) return s;
        for (Rk& r : g->rk) {
            r.hroff.assign(static_cast<size_t>(W) + 1, 0);
            uint64_t base = 0;
            for (int p = 0; p < r.rank; ++p) base += mat[static_cast<size_t>(p) * W];
            r.hroff[0] = base;  // (slot 0 doubles as the slice's hit base in this mode)
        }
        return GG_OK;
    }

    // ---- chromosome-sharded: route -> counts -> exchange -> answer -> (return)
    std::vector<std::vector<uint64_t>> mine(g->rk.size(), std::vector<uint64_t>(static_cast<size_t>(W), 0));
    if (ph == GG_OK) ph = [&]() -> gg_status {
    for (Rk& r : g->rk) {
        gg_status s = route_phase(g, r, r.q_c, r.nq);
        if (s != GG_OK) return s;
        // the columns in partition order (enqueued behind the partition; the bounds copy is already in flight)
        s = gather_col<uint32_t>(g, r, r.lchrom.as<uint32_t>(), r.nq, r.send_c);
        if (s == GG_OK) s = cb == 4 ? gather_col<uint32_t>(g, r, static_cast<const uint32_t*>(r.q_s), r.nq, r.send_s)
                                    : gather_col<uint64_t>(g, r, static_cast<const uint64_t*>(r.q_s), r.nq, r.send_s);
        if (s == GG_OK) s = cb == 4 ? gather_col<uint32_t>(g, r, static_cast<const uint32_t*>(r.q_e), r.nq, r.send_e)
                                    : gather_col<uint64_t>(g, r, static_cast<const uint64_t*>(r.q_e), r.nq, r.send_e);
        if (s != GG_OK) return s;
        if (!return_to_source) {  // the original index rides along: results stay on the owner
            GGG_CUDA(r.send_o.ensure((r.nq ? r.nq : 1) * 4));
            if (r.perm_ptr)
                GGG_CUDA(cudaMemcpyAsync(r.send_o.p, r.perm_ptr, r.nq * 4, cudaMemcpyDeviceToDevice, r.st));
            else if (r.nq) {
                route_kernel<<<blocks_for(r.nq, 256), 256, 0, r.st>>>(r.q_c, r.nq, r.d_owner.as<uint32_t>(), r.d_local.as<uint32_t>(),
                                                                     g->n_chrom, static_cast<uint32_t>(r.rank), r.dest.as<uint32_t>(),
                                                                     r.lchrom.as<uint32_t>(), r.send_o.as<uint32_t>());
                GGG_CUDA(cudaGetLastError());
            }
        }
        GGG_CUDA(cudaEventRecord(r.ev_t[1], r.st));
    }
    for (size_t i = 0; i < g->rk.size(); ++i) {
        Rk& r = g->rk[i];
        GGG_CUDA(cudaSetDevice(r.dev));
        GGG_CUDA(cudaStreamSynchronize(r.st));
        for (int p = 0; p <= W; ++p) r.soff[p] = r.h_box[p];
        for (int p = 0; p < W; ++p) mine[i][p] = r.soff[p + 1] - r.soff[p];
    }
    return GG_OK;
    }();
    ph = agree(g, ph);
    if (ph != GG_OK) return ph;
    std::vector<uint64_t> mat;
    gg_status s = gather_matrix(g, mine, mat);
    if (s != GG_OK) return s;
    ph = [&]() -> gg_status {
    for (Rk& r : g->rk) {
        recv_offsets_from(mat, W, r.rank, r.roff);
        r.n_recv = r.roff[W];
        GGG_CUDA(cudaSetDevice(r.dev));
        const size_t m = r.n_recv ? r.n_recv : 1;
        GGG_CUDA(r.recv_c.ensure(m * 4));
        GGG_CUDA(r.recv_s.ensure(m * cb));
        GGG_CUDA(r.recv_e.ensure(m * cb));
        if (!return_to_source) GGG_CUDA(r.recv_o.ensure(m * 4));
        GGG_CUDA(r.r_off.ensure((m + 1) * 8));
    }
    return GG_OK;
    }();
    ph = agree(g, ph);
    if (ph != GG_OK) return ph;
    {
        std::vector<Col> cols = {{[](Rk& r) { return r.send_c.p; }, [](Rk& r) { return r.recv_c.p; }, 4},
                                 {[](Rk& r) { return r.send_s.p; }, [](Rk& r) { return r.recv_s.p; }, cb},
                                 {[](Rk& r) { return r.send_e.p; }, [](Rk& r) { return r.recv_e.p; }, cb}};
        if (!return_to_source) cols.push_back({[](Rk& r) { return r.send_o.p; }, [](Rk& r) { return r.recv_o.p; }, 4});
        s = all_to_all(g, cols, [](Rk& r) -> const std::vector<uint64_t>& { return r.soff; },
                       [](Rk& r) -> const std::vector<uint64_t>& { return r.roff; });
        if (s != GG_OK) return s;
    }
    ph = [&]() -> gg_status {
    for (Rk& r : g->rk) {
        GGG_CUDA(cudaSetDevice(r.dev));
        GGG_CUDA(cudaEventRecord(r.ev_t[2], r.st));
    }
    // ---- answer on the owner: enqueue everywhere with an estimated hit capacity, then read the totals
    for (int attempt = 0; attempt < 2; ++attempt) {
        bool again = false;
        for (Rk& r : g->rk) {
            if (attempt == 1 && r.n_hits_owner <= r.r_hits.cap / 4) continue;
            GGG_CUDA(cudaSetDevice(r.dev));
            // grow only when the expected hits do not fit, and then with slack: re-allocating tens of GB because a chunk is a
            // little larger than the last one cost more than the intersect itself (cfg5 on 8 GPUs: 64 M-query chunks ran at
            // 1.9 ms per M queries against 0.3 on one GPU)
            const size_t want = attempt == 0 ? static_cast<size_t>(r.hits_per_query * static_cast<double>(r.n_recv)) + 4096
                                             : static_cast<size_t>(r.n_hits_owner);
            if (r.r_hits.cap / 4 < want) GGG_CUDA(r.r_hits.ensure((want + want / 4) * 4));
            GGG_IDX(r, gg_intersect_fused_device(r.idx, r.recv_c.as<uint32_t>(), r.recv_s.p, r.recv_e.p, coord_bytes, r.n_recv,
                                                 r.r_off.as<uint64_t>(), r.r_hits.as<uint32_t>(), r.r_hits.cap / 4, nullptr, r.st));
            // hit offsets at the source-segment boundaries (what each source gets back) ride along with the total
            GGG_CUDA(r.d_seg.ensure((static_cast<size_t>(W) + 1) * 8));
            GGG_CUDA(cudaMemcpyAsync(r.d_seg.p, r.roff.data(), (static_cast<size_t>(W) + 1) * 8, cudaMemcpyHostToDevice, r.st));
            for (int p = 0; p <= W; ++p)
                GGG_CUDA(cudaMemcpyAsync(r.h_box + p, r.r_off.as<uint64_t>() + r.roff[p], 8, cudaMemcpyDeviceToHost, r.st));
        }
        for (Rk& r : g->rk) {
            GGG_CUDA(cudaSetDevice(r.dev));
            GGG_CUDA(cudaStreamSynchronize(r.st));
            for (int p = 0; p <= W; ++p) r.hsoff[p] = r.h_box[p];
            r.n_hits_owner = r.hsoff[W];
            if (r.n_hits_owner > r.r_hits.cap / 4) again = true;
        }
        if (!again) break;
    }
    for (Rk& r : g->rk) {
        GGG_CUDA(cudaSetDevice(r.dev));
        GGG_CUDA(cudaEventRecord(r.ev_t[3], r.st));
        if (r.n_recv >= 65536) r.hits_per_query = std::max(1.0, static_cast<double>(r.n_hits_owner) / static_cast<double>(r.n_recv));
    }
    return GG_OK;
    }();
    ph = agree(g, ph);  // (also when the results stay on the owner: every rank reports the same outcome)
    if (ph != GG_OK) return ph;
    if (!return_to_source) {
        for (Rk& r : g->rk) {
            GGG_CUDA(cudaSetDevice(r.dev));
            GGG_CUDA(r.origin.ensure((r.n_recv ? r.n_recv : 1) * 8));
            if (r.n_recv) {
                origin_kernel<<<blocks_for(r.n_recv, 256), 256, 0, r.st>>>(r.recv_o.as<uint32_t>(), r.n_recv, r.d_seg.as<uint64_t>(),
                                                                           static_cast<uint32_t>(W), r.origin.as<uint64_t>());
                GGG_CUDA(cudaGetLastError());
            }
            GGG_CUDA(cudaEventRecord(r.ev_t[4], r.st));
            GGG_CUDA(cudaStreamSynchronize(r.st));
            r.n_hits_final = r.n_hits_owner;
        }
        return GG_OK;
    }
    // ---- return: counts and lists back to the source, then original query order
    for (size_t i = 0; i < g->rk.size(); ++i) {
        Rk& r = g->rk[i];
        for (int p = 0; p < W; ++p) mine[i][p] = r.hsoff[p + 1] - r.hsoff[p];  // hits going back to source p
    }
    s = gather_matrix(g, mine, mat);
    if (s != GG_OK) return s;
    ph = [&]() -> gg_status {
    for (Rk& r : g->rk) {
        // hits coming back to r from owner p = mat[p * W + r.rank]
        recv_offsets_from(mat, W, r.rank, r.hroff);
        r.n_hits_final = r.hroff[W];
        GGG_CUDA(cudaSetDevice(r.dev));
        GGG_CUDA(r.r_cnt.ensure((r.n_recv ? r.n_recv : 1) * 4));
        if (r.n_recv) {
            diff_kernel<<<blocks_for(r.n_recv, 256), 256, 0, r.st>>>(r.r_off.as<uint64_t>(), r.n_recv, r.r_cnt.as<uint32_t>());
            GGG_CUDA(cudaGetLastError());
        }
        const size_t m = r.nq ? r.nq : 1;
        GGG_CUDA(r.b_cnt.ensure(m * 4));
        GGG_CUDA(r.b_hits.ensure((r.n_hits_final ? r.n_hits_final : 1) * 4));
        GGG_CUDA(r.f_cnt.ensure(m * 4));
        GGG_CUDA(r.b_off.ensure((m + 1) * 8));
        GGG_CUDA(r.f_off.ensure((m + 1) * 8));
        GGG_CUDA(r.f_hits.ensure((r.n_hits_final ? r.n_hits_final : 1) * 4));
        GGG_CUDA(r.scan2.ensure(scan_state_bytes(m)));
    }
    return GG_OK;
    }();
    ph = agree(g, ph);
    if (ph != GG_OK) return ph;
    s = all_to_all(g, {{[](Rk& r) { return r.r_cnt.p; }, [](Rk& r) { return r.b_cnt.p; }, 4}},
                   [](Rk& r) -> const std::vector<uint64_t>& { return r.roff; },
                   [](Rk& r) -> const std::vector<uint64_t>& { return r.soff; });
    if (s != GG_OK) return s;
    s = all_to_all(g, {{[](Rk& r) { return r.r_hits.p; }, [](Rk& r) { return r.b_hits.p; }, 4}},
                   [](Rk& r) -> const std::vector<uint64_t>& { return r.hsoff; },
                   [](Rk& r) -> const std::vector<uint64_t>& { return r.hroff; });
    if (s != GG_OK) return s;
    for (Rk& r : g->rk) {
        GGG_CUDA(cudaSetDevice(r.dev));
        const size_t n = r.nq;
        if (n) {
            scatter_u32_kernel<<<blocks_for(n, 256), 256, 0, r.st>>>(r.perm_ptr, r.b_cnt.as<uint32_t>(), n, r.f_cnt.as<uint32_t>());
            GGG_CUDA(cudaGetLastError());
        }
        GGG_CUDA((exclusive_scan_async<uint32_t, uint64_t>(r.b_cnt.as<uint32_t>(), r.b_off.as<uint64_t>(), n, r.scan2.p, nullptr, nullptr, r.st)));
        GGG_CUDA((exclusive_scan_async<uint32_t, uint64_t>(r.f_cnt.as<uint32_t>(), r.f_off.as<uint64_t>(), n, r.scan2.p, nullptr, nullptr, r.st)));
        if (n && r.n_hits_final) {
            const unsigned grid = static_cast<unsigned>(std::min<size_t>(div_up(div_up(n, 32), 8), 148 * 8));
            gather_lists_kernel<<<grid, 256, 0, r.st>>>(r.perm_ptr, r.b_off.as<uint64_t>(), r.f_off.as<uint64_t>(), r.b_hits.as<uint32_t>(),
                                                        n, r.f_hits.as<uint32_t>());
            GGG_CUDA(cudaGetLastError());
        }
        GGG_CUDA(cudaEventRecord(r.ev_t[4], r.st));
    }
    // global CSR base of each slice: every rank's final total is a column sum of the same matrix — no further collective
    for (Rk& r : g->rk) {
        uint64_t base = 0;
        for (int q = 0; q < r.rank; ++q)
            for (int p = 0; p < W; ++p) base += mat[static_cast<size_t>(p) * W + q];
        r.hsoff[0] = base;  // (reused as the slice's hit base from here on)
        GGG_CUDA(cudaSetDevice(r.dev));
        GGG_CUDA(cudaStreamSynchronize(r.st));
    }
    return GG_OK;
}

void collect_stats(gg_group* g, bool return_to_source) {
    gg_group_stats& st = g->stats;
    const double build_ms = st.build_ms;
    st = gg_group_stats{};
    st.build_ms = build_ms;
    for (Rk& r : g->rk) {
        st.n_queries += r.nq;
        st.n_answered += r.n_recv;
        st.n_hits += return_to_source || g->shard_mode == GG_SHARD_REPLICATE ? r.n_hits_final : r.n_hits_owner;
        st.bytes_sent += r.bytes_sent;
        st.bytes_received += r.bytes_received;
        cudaSetDevice(r.dev);
        st.route_ms = std::max(st.route_ms, ms_between(r.ev_t[0], r.ev_t[1]));
        st.exchange_ms = std::max(st.exchange_ms, ms_between(r.ev_t[1], r.ev_t[2]));
        st.intersect_ms = std::max(st.intersect_ms, ms_between(r.ev_t[2], r.ev_t[3]));
        st.return_ms = std::max(st.return_ms, ms_between(r.ev_t[3], r.ev_t[4]));
    }
}

}  // namespace

// ============================================================================================== C ABI
extern "C" {

gg_status gg_shard_assign(const uint64_t* counts, uint32_t n_chrom, int world, uint32_t* owner) {
    if (!counts || !owner || world < 1) return GG_ERR_INVALID_ARG;
    lpt_assign(counts, n_chrom, world, owner);
    return GG_OK;
}

gg_status gg_group_create(gg_group** out, const int* devices, int n_devices) {
    if (!out || !devices || n_devices < 1 || n_devices > 64) return GG_ERR_INVALID_ARG;
    *out = nullptr;
    gg_group* g = new (std::nothrow) gg_group();
    if (!g) return GG_ERR_OOM;
    g->world = n_devices;
    g->rank_form = false;
    g->rk.resize(static_cast<size_t>(n_devices));
    for (int i = 0; i < n_devices; ++i) {
        gg_status s = init_rank_ctx(g, g->rk[i], devices[i], i);
        if (s != GG_OK) {
            gg_group_destroy(g);
            return s;
        }
    }
    // peer access both ways between distinct devices (NVLink); a refusal only means the copies are staged
    for (int i = 0; i < n_devices; ++i)
        for (int j = 0; j < n_devices; ++j) {
            if (devices[i] == devices[j]) continue;
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, devices[i], devices[j]) == cudaSuccess && can) {
                cudaSetDevice(devices[i]);
                cudaDeviceEnablePeerAccess(devices[j], 0);
            }
            cudaGetLastError();
        }
    *out = g;
    return GG_OK;
}

gg_status gg_group_unique_id(void* id128) {
    if (!id128) return GG_ERR_INVALID_ARG;
    if (!nccl().ok) return GG_ERR_CUDA;
    ncclUniqueId id;
    if (nccl().GetUniqueId(&id) != ncclSuccess) return GG_ERR_CUDA;
    static_assert(sizeof(id) == 128, "ncclUniqueId is 128 bytes");
    std::memcpy(id128, &id, sizeof(id));
    return GG_OK;
}

gg_status gg_group_create_rank(gg_group** out, int device, int rank, int world, const void* id128) {
    if (!out || !id128 || world < 1 || world > 4096 || rank < 0 || rank >= world) return GG_ERR_INVALID_ARG;
    *out = nullptr;
    gg_group* g = new (std::nothrow) gg_group();
    if (!g) return GG_ERR_OOM;
    g->world = world;
    g->rank_form = true;
    g->rk.resize(1);
    gg_status s = init_rank_ctx(g, g->rk[0], device, rank);
    if (s != GG_OK) {
        gg_group_destroy(g);
        return s;
    }
    if (!nccl().ok) {
        gg_group_destroy(g);
        return GG_ERR_CUDA;
    }
    ncclUniqueId id;
    std::memcpy(&id, id128, sizeof(id));
    cudaSetDevice(device);
    if (nccl().CommInitRank(&g->comm, world, id, rank) != ncclSuccess) {
        g->comm = nullptr;
        gg_group_destroy(g);
        return GG_ERR_CUDA;
    }
    *out = g;
    return GG_OK;
}

void gg_group_destroy(gg_group* g) {
    if (!g) return;
    for (Rk& r : g->rk) {
        cudaSetDevice(r.dev);
        if (r.st) cudaStreamSynchronize(r.st);
    }
    if (g->comm && nccl().ok) nccl().CommDestroy(g->comm);
    for (Rk& r : g->rk) {
        cudaSetDevice(r.dev);
        for (DeviceBuf* b : {&r.d_owner, &r.d_local, &r.in_c, &r.in_s, &r.in_e, &r.dest, &r.dest_alt, &r.perm, &r.perm_alt, &r.lchrom,
                             &r.hist, &r.scanst, &r.d_bounds, &r.send_c, &r.send_s, &r.send_e, &r.send_o, &r.recv_c, &r.recv_s,
                             &r.recv_e, &r.recv_o, &r.r_off, &r.r_hits, &r.r_cnt, &r.origin, &r.d_seg, &r.b_cnt, &r.b_off, &r.b_hits,
                             &r.f_cnt, &r.f_off, &r.f_hits, &r.scan2, &r.d_mat, &r.d_agree})
            b->release();
        if (r.idx) gg_index_destroy(r.idx);
        if (r.h_box) cudaFreeHost(r.h_box);
        if (r.ev_ready) cudaEventDestroy(r.ev_ready);
        if (r.ev_done) cudaEventDestroy(r.ev_done);
        for (auto& e : r.ev_t)
            if (e) cudaEventDestroy(e);
        if (r.st) cudaStreamDestroy(r.st);
    }
    if (g->pinned_hits) cudaFreeHost(g->pinned_hits);
    delete g;
}

const char* gg_group_last_error(const gg_group* g) { return g ? g->err.c_str() : "null group"; }
int gg_group_world(const gg_group* g) { return g ? g->world : 0; }
gg_index* gg_group_index(gg_group* g, int i) {
    return g && i >= 0 && i < static_cast<int>(g->rk.size()) ? g->rk[static_cast<size_t>(i)].idx : nullptr;
}

gg_status gg_group_owners(const gg_group* g, uint32_t* owner) {
    if (!g || !owner) return GG_ERR_INVALID_ARG;
    if (!g->built) return GG_ERR_NOT_BUILT;
    for (uint32_t c = 0; c < g->n_chrom; ++c) owner[c] = g->shard_mode == GG_SHARD_CHROM ? g->owner[c] : 0u;
    return GG_OK;
}

gg_status gg_group_get_stats(const gg_group* g, gg_group_stats* out) {
    if (!g || !out) return GG_ERR_INVALID_ARG;
    *out = g->stats;
    return GG_OK;
}

gg_status gg_group_build(gg_group* g, int shard_mode, const uint32_t* chrom, const uint64_t* start, const uint64_t* end,
                         const uint64_t* tie_key, int tie_mode, size_t n, uint32_t n_chrom, uint64_t* rank_to_input,
                         size_t n_owned_capacity, size_t* n_owned) {
    if (!g) return GG_ERR_INVALID_ARG;
    g->err.clear();
    g->built = false;
    if (shard_mode != GG_SHARD_REPLICATE && shard_mode != GG_SHARD_CHROM) return gfail(g, GG_ERR_INVALID_ARG, "bad shard_mode");
    if (n_chrom == 0 || n_chrom == GG_NO_CHROM) return gfail(g, GG_ERR_INVALID_ARG, "n_chrom must be in [1, 2^32-2]");
    if (n > 0 && (!chrom || !start || !end)) return gfail(g, GG_ERR_INVALID_ARG, "null input array");
    if (tie_mode == GG_TIE_EXPLICIT && !tie_key && n > 0) return gfail(g, GG_ERR_INVALID_ARG, "GG_TIE_EXPLICIT needs tie_key");
    const auto t0 = std::chrono::steady_clock::now();
    const int W = g->world;
    g->shard_mode = shard_mode;
    g->n_chrom = n_chrom;
    g->owner.assign(n_chrom, 0);
    if (n_owned) *n_owned = 0;

    if (shard_mode == GG_SHARD_REPLICATE) {
        // every GPU builds the whole index (the device build is cheaper than shipping 130 MB of levels around)
        std::vector<gg_status> st(g->rk.size(), GG_OK);
        std::vector<uint32_t> perm(rank_to_input ? n : 0);
        std::vector<std::thread> th;
        for (size_t i = 0; i < g->rk.size(); ++i)
            th.emplace_back([&, i] {
                st[i] = gg_index_build(g->rk[i].idx, chrom, start, end, tie_key, tie_mode, n, n_chrom,
                                       i == 0 && rank_to_input ? perm.data() : nullptr);
            });
        for (auto& t : th) t.join();
        for (size_t i = 0; i < g->rk.size(); ++i)
            if (st[i] != GG_OK) return gfail(g, st[i], std::string("gg_index_build: ") + gg_last_error(g->rk[i].idx));
        if (rank_to_input) {
            if (n_owned_capacity < n) return gfail(g, GG_ERR_CAPACITY, "rank_to_input buffer too small");
            for (size_t i = 0; i < n; ++i) rank_to_input[i] = perm[i];
        }
        if (n_owned) *n_owned = n;
        g->built = true;
        g->stats.build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        return GG_OK;
    }

    // ---- GG_SHARD_CHROM: per-chromosome interval counts of the whole input -> LPT owners (identical on every rank)
    std::vector<uint64_t> counts(n_chrom, 0);
    gg_status ph = [&]() -> gg_status {
        // (rank form: n is this rank's part; the routing below carries the record's index in it as a u32)
        if (g->rank_form && n >= 0xFFFFFFFFull) return gfail(g, GG_ERR_CAPACITY, "more than 2^32-2 input records on one rank");
        for (size_t i = 0; i < n; ++i) {
            if (chrom[i] >= n_chrom) return gfail(g, GG_ERR_INVALID_ARG, "chrom id >= n_chrom in index input");
            ++counts[chrom[i]];
        }
        return GG_OK;
    }();
    ph = agree(g, ph);  // one rank's bad record fails the build on every rank instead of leaving the others in the gather
    if (ph != GG_OK) return ph;
    std::vector<uint64_t> n_per_rank(static_cast<size_t>(W), 0);  // input records held by each rank (rank form)
    if (g->rank_form) {
        // all-gather of the per-rank histograms in slabs of W values (the same 8-byte-per-entry collective), summed on the host
        std::vector<uint64_t> total(n_chrom, 0), mat;
        std::vector<std::vector<uint64_t>> mine(1, std::vector<uint64_t>(static_cast<size_t>(W), 0));
        for (uint32_t c0 = 0; c0 < n_chrom; c0 += static_cast<uint32_t>(W)) {
            for (int k = 0; k < W; ++k) mine[0][k] = c0 + k < n_chrom ? counts[c0 + k] : 0;
            gg_status s = gather_matrix(g, mine, mat);
            if (s != GG_OK) return s;
            for (int p = 0; p < W; ++p)
                for (int k = 0; k < W; ++k)
                    if (c0 + k < n_chrom) total[c0 + k] += mat[static_cast<size_t>(p) * W + k];
        }
        std::fill(mine[0].begin(), mine[0].end(), 0);
        mine[0][0] = n;
        gg_status s = gather_matrix(g, mine, mat);
        if (s != GG_OK) return s;
        for (int p = 0; p < W; ++p) n_per_rank[p] = mat[static_cast<size_t>(p) * W];
        counts.swap(total);
    }
    lpt_assign(counts.data(), n_chrom, W, g->owner.data());
    gg_status s = upload_routing(g);
    if (s != GG_OK && !g->rank_form) return s;
    ph = s;  // (rank form: agreed on below, together with the upload and the partition)
    std::vector<uint32_t> n_local_chrom(static_cast<size_t>(W), 0);
    for (uint32_t c = 0; c < n_chrom; ++c) ++n_local_chrom[g->owner[c]];

    if (!g->rank_form) {
        // single process: stable host partition of the records by owner, then one device build per GPU, in parallel
        std::vector<std::vector<uint32_t>> lc(W);
        std::vector<std::vector<uint64_t>> ls(W), le(W), lt(W), src(W);  // src: global input index (the total may pass 2^32)
        std::vector<uint64_t> per_rank(static_cast<size_t>(W), 0);
        for (uint32_t c = 0; c < n_chrom; ++c) per_rank[g->owner[c]] += counts[c];
        for (int r = 0; r < W; ++r) {
            lc[r].reserve(per_rank[r]);
            src[r].reserve(per_rank[r]);
            ls[r].reserve(per_rank[r]);
            le[r].reserve(per_rank[r]);
            if (tie_mode == GG_TIE_EXPLICIT) lt[r].reserve(per_rank[r]);
        }
        for (size_t i = 0; i < n; ++i) {
            const uint32_t o = g->owner[chrom[i]];
            lc[o].push_back(g->local_id[chrom[i]]);
            src[o].push_back(i);
            ls[o].push_back(start[i]);
            le[o].push_back(end[i]);
            if (tie_mode == GG_TIE_EXPLICIT) lt[o].push_back(tie_key[i]);
        }
        std::vector<gg_status> st(static_cast<size_t>(W), GG_OK);
        std::vector<std::vector<uint32_t>> perm(W);
        std::vector<std::thread> th;
        for (int r = 0; r < W; ++r)
            th.emplace_back([&, r] {
                perm[r].resize(lc[r].size());
                st[r] = gg_index_build(g->rk[r].idx, lc[r].data(), ls[r].data(), le[r].data(),
                                       tie_mode == GG_TIE_EXPLICIT ? lt[r].data() : nullptr, tie_mode, lc[r].size(),
                                       std::max<uint32_t>(n_local_chrom[r], 1u), perm[r].data());
            });
        for (auto& t : th) t.join();
        for (int r = 0; r < W; ++r)
            if (st[r] != GG_OK) return gfail(g, st[r], std::string("gg_index_build: ") + gg_last_error(g->rk[r].idx));
        if (rank_to_input) {
            if (n_owned_capacity < n) return gfail(g, GG_ERR_CAPACITY, "rank_to_input buffer too small");
            // chromosome by chromosome in global id order; a rank's permutation lists its chromosomes in local id order
            std::vector<uint64_t> gbegin(static_cast<size_t>(n_chrom) + 1, 0);
            for (uint32_t c = 0; c < n_chrom; ++c) gbegin[c + 1] = gbegin[c] + counts[c];
            std::vector<uint64_t> cursor(static_cast<size_t>(W), 0);
            for (uint32_t c = 0; c < n_chrom; ++c) {
                const uint32_t o = g->owner[c];
                for (uint64_t k = 0; k < counts[c]; ++k) rank_to_input[gbegin[c] + k] = src[o][perm[o][cursor[o] + k]];
                cursor[o] += counts[c];
            }
        }
        if (n_owned) *n_owned = n;
    } else {
        // rank form: this rank's records -> device -> routed to their owners over NVLink -> device build of what arrived
        Rk& r = g->rk[0];
        GGG_CUDA(cudaSetDevice(r.dev));
        DeviceBuf in_c, in_s, in_e, in_t, s_c, s_s, s_e, s_t, s_o, v_c, v_s, v_e, v_t, v_o, d_perm, d_r2i;
        struct Guard {
            std::vector<DeviceBuf*> b;
            ~Guard() {
                for (auto* x : b) x->release();
            }
        } guard{{&in_c, &in_s, &in_e, &in_t, &s_c, &s_s, &s_e, &s_t, &s_o, &v_c, &v_s, &v_e, &v_t, &v_o, &d_perm, &d_r2i}};
        const size_t nn = n ? n : 1;
        const bool tie = tie_mode == GG_TIE_EXPLICIT;
        std::vector<std::vector<uint64_t>> mine(1, std::vector<uint64_t>(static_cast<size_t>(W), 0));
        if (ph == GG_OK) ph = [&]() -> gg_status {
        GGG_CUDA(in_c.ensure(nn * 4));
        GGG_CUDA(in_s.ensure(nn * 8));
        GGG_CUDA(in_e.ensure(nn * 8));
        if (tie) GGG_CUDA(in_t.ensure(nn * 8));
        if (n) {
            GGG_CUDA(cudaMemcpyAsync(in_c.p, chrom, n * 4, cudaMemcpyHostToDevice, r.st));
            GGG_CUDA(cudaMemcpyAsync(in_s.p, start, n * 8, cudaMemcpyHostToDevice, r.st));
            GGG_CUDA(cudaMemcpyAsync(in_e.p, end, n * 8, cudaMemcpyHostToDevice, r.st));
            if (tie) GGG_CUDA(cudaMemcpyAsync(in_t.p, tie_key, n * 8, cudaMemcpyHostToDevice, r.st));
        }
        s = route_phase(g, r, in_c.as<uint32_t>(), n);
        if (s != GG_OK) return s;
        s = gather_col<uint32_t>(g, r, r.lchrom.as<uint32_t>(), n, s_c);
        if (s == GG_OK) s = gather_col<uint64_t>(g, r, in_s.as<uint64_t>(), n, s_s);
        if (s == GG_OK) s = gather_col<uint64_t>(g, r, in_e.as<uint64_t>(), n, s_e);
        if (s == GG_OK && tie) s = gather_col<uint64_t>(g, r, in_t.as<uint64_t>(), n, s_t);
        if (s != GG_OK) return s;
        GGG_CUDA(s_o.ensure(nn * 4));
        if (r.perm_ptr) {
            GGG_CUDA(cudaMemcpyAsync(s_o.p, r.perm_ptr, n * 4, cudaMemcpyDeviceToDevice, r.st));
        } else if (n) {
            route_kernel<<<blocks_for(n, 256), 256, 0, r.st>>>(in_c.as<uint32_t>(), n, r.d_owner.as<uint32_t>(), r.d_local.as<uint32_t>(),
                                                               n_chrom, static_cast<uint32_t>(r.rank), r.dest.as<uint32_t>(),
                                                               r.lchrom.as<uint32_t>(), s_o.as<uint32_t>());
            GGG_CUDA(cudaGetLastError());
        }
        GGG_CUDA(cudaStreamSynchronize(r.st));
        for (int p = 0; p <= W; ++p) r.soff[p] = r.h_box[p];
        for (int p = 0; p < W; ++p) mine[0][p] = r.soff[p + 1] - r.soff[p];
        return GG_OK;
        }();
        ph = agree(g, ph);
        if (ph != GG_OK) return ph;
        std::vector<uint64_t> mat;
        s = gather_matrix(g, mine, mat);
        if (s != GG_OK) return s;
        recv_offsets_from(mat, W, r.rank, r.roff);
        const size_t m = r.roff[W], mm = m ? m : 1;
        ph = [&]() -> gg_status {
            // known as soon as the matrix is: checked here, before the exchange, so that every rank hears of it
            if (rank_to_input && n_owned_capacity < m) {
                if (n_owned) *n_owned = m;
                return gfail(g, GG_ERR_CAPACITY, "rank_to_input buffer too small for the intervals this rank owns");
            }
            GGG_CUDA(v_c.ensure(mm * 4));
            GGG_CUDA(v_s.ensure(mm * 8));
            GGG_CUDA(v_e.ensure(mm * 8));
            GGG_CUDA(v_o.ensure(mm * 4));
            if (tie) GGG_CUDA(v_t.ensure(mm * 8));
            return GG_OK;
        }();
        ph = agree(g, ph);
        if (ph != GG_OK) return ph;
        // (the column buffers are locals of this call: pass them through the rank context's generic slots)
        r.send_c.borrow(s_c.p, s_c.cap);
        r.send_s.borrow(s_s.p, s_s.cap);
        r.send_e.borrow(s_e.p, s_e.cap);
        r.send_o.borrow(s_o.p, s_o.cap);
        r.recv_c.borrow(v_c.p, v_c.cap);
        r.recv_s.borrow(v_s.p, v_s.cap);
        r.recv_e.borrow(v_e.p, v_e.cap);
        r.recv_o.borrow(v_o.p, v_o.cap);
        std::vector<Col> cols = {{[](Rk& x) { return x.send_c.p; }, [](Rk& x) { return x.recv_c.p; }, 4},
                                 {[](Rk& x) { return x.send_s.p; }, [](Rk& x) { return x.recv_s.p; }, 8},
                                 {[](Rk& x) { return x.send_e.p; }, [](Rk& x) { return x.recv_e.p; }, 8},
                                 {[](Rk& x) { return x.send_o.p; }, [](Rk& x) { return x.recv_o.p; }, 4}};
        s = all_to_all(g, cols, [](Rk& x) -> const std::vector<uint64_t>& { return x.soff; },
                       [](Rk& x) -> const std::vector<uint64_t>& { return x.roff; });
        if (s == GG_OK && tie) {
            r.send_s.borrow(s_t.p, s_t.cap);
            r.recv_s.borrow(v_t.p, v_t.cap);
            s = all_to_all(g, {{[](Rk& x) { return x.send_s.p; }, [](Rk& x) { return x.recv_s.p; }, 8}},
                           [](Rk& x) -> const std::vector<uint64_t>& { return x.soff; },
                           [](Rk& x) -> const std::vector<uint64_t>& { return x.roff; });
        }
        for (DeviceBuf* b : {&r.send_c, &r.send_s, &r.send_e, &r.send_o, &r.recv_c, &r.recv_s, &r.recv_e, &r.recv_o}) b->release();
        if (s != GG_OK) return s;
        GGG_CUDA(cudaStreamSynchronize(r.st));
        GGG_CUDA(d_perm.ensure(mm * 4));
        GGG_IDX(r, gg_index_build_device(r.idx, v_c.as<uint32_t>(), v_s.as<uint64_t>(), v_e.as<uint64_t>(),
                                         tie ? v_t.as<uint64_t>() : nullptr, tie_mode, m, std::max<uint32_t>(n_local_chrom[r.rank], 1u),
                                         rank_to_input ? d_perm.as<uint32_t>() : nullptr));
        if (n_owned) *n_owned = m;
        if (rank_to_input) {
            // chain rank -> received position -> (source rank, index in the source's input) -> global input index
            std::vector<uint32_t> h_perm(m), h_orig(m);
            if (m) {
                GGG_CUDA(cudaMemcpy(h_perm.data(), d_perm.p, m * 4, cudaMemcpyDeviceToHost));
                GGG_CUDA(cudaMemcpy(h_orig.data(), v_o.p, m * 4, cudaMemcpyDeviceToHost));
            }
            std::vector<uint64_t> in_base(static_cast<size_t>(W) + 1, 0);
            for (int p = 0; p < W; ++p) in_base[p + 1] = in_base[p] + n_per_rank[p];
            for (size_t k = 0; k < m; ++k) {
                const uint32_t pos = h_perm[k];
                int p = 0;
                while (p + 1 < W && r.roff[p + 1] <= pos) ++p;
                rank_to_input[k] = in_base[p] + h_orig[pos];
            }
        }
    }
    g->built = true;
    g->stats.build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return GG_OK;
}

gg_status gg_group_intersect_device(gg_group* g, const uint32_t* d_chrom, const void* d_start, const void* d_end,
                                    int coord_bytes, size_t nq, int return_to_source, gg_group_device_result* out) {
    if (!g || !out) return GG_ERR_INVALID_ARG;
    g->err.clear();
    if (!g->built) return gfail(g, GG_ERR_NOT_BUILT, "group not built");
    if (g->rk.size() != 1) return gfail(g, GG_ERR_INVALID_ARG, "the device-buffer call drives one GPU per process (rank form, or a group of one)");
    if (coord_bytes != 4 && coord_bytes != 8) return gfail(g, GG_ERR_INVALID_ARG, "coord_bytes must be 4 or 8");
    // what depends on this rank's slice is not returned at once: the other ranks are about to enter the call's
    // collectives, so the failure goes through intersect_core's first agreement and every rank returns it
    gg_status pre = GG_OK;
    if (nq && (!d_chrom || !d_start || !d_end)) pre = gfail(g, GG_ERR_INVALID_ARG, "null argument");
    else if (nq >= 0xFFFFFFFFull) pre = gfail(g, GG_ERR_CAPACITY, "more than 2^32-2 queries per rank and call");
    Rk& r = g->rk[0];
    r.q_c = d_chrom;
    r.q_s = d_start;
    r.q_e = d_end;
    r.nq = pre == GG_OK ? nq : 0;
    const bool back = return_to_source != 0 || g->shard_mode == GG_SHARD_REPLICATE;
    gg_status s = intersect_core(g, coord_bytes, back, pre);
    if (s != GG_OK) return s;
    collect_stats(g, back);
    *out = gg_group_device_result{};
    if (g->shard_mode == GG_SHARD_REPLICATE) {
        out->n_queries = nq;
        out->n_hits = r.n_hits_owner;
        out->hit_base = r.hroff[0];
        out->d_offsets = r.r_off.as<uint64_t>();
        out->d_hits = r.r_hits.as<uint32_t>();
    } else if (back) {
        out->n_queries = nq;
        out->n_hits = r.n_hits_final;
        out->hit_base = r.hsoff[0];
        out->d_offsets = r.f_off.as<uint64_t>();
        out->d_hits = r.f_hits.as<uint32_t>();
    } else {
        out->n_queries = r.n_recv;
        out->n_hits = r.n_hits_owner;
        out->d_offsets = r.r_off.as<uint64_t>();
        out->d_hits = r.r_hits.as<uint32_t>();
        out->d_origin = r.origin.as<uint64_t>();
    }
    return GG_OK;
}

gg_status gg_group_intersect_batch(gg_group* g, const uint32_t* q_chrom, const void* q_start, const void* q_end,
                                   int coord_bytes, size_t nq, uint64_t* offsets, const uint32_t** hits, uint64_t* n_hits,
                                   uint64_t* hit_base) {
    if (!g) return GG_ERR_INVALID_ARG;
    g->err.clear();
    if (!g->built) return gfail(g, GG_ERR_NOT_BUILT, "group not built");
    if (coord_bytes != 4 && coord_bytes != 8) return gfail(g, GG_ERR_INVALID_ARG, "coord_bytes must be 4 or 8");
    const size_t L = g->rk.size();
    const size_t cb = static_cast<size_t>(coord_bytes);
    // contiguous slices of the batch, one per local GPU (rank form: the caller's batch IS this rank's slice)
    std::vector<size_t> lo(L + 1, 0);
    for (size_t i = 0; i < L; ++i) lo[i + 1] = lo[i] + nq / L + (i < nq % L ? 1 : 0);
    // checks and uploads that depend on this rank's slice: a failure is agreed on inside intersect_core (rank form), not
    // returned while the other ranks enter the collectives
    gg_status pre = [&]() -> gg_status {
    if (!offsets || !hits || !n_hits || (nq && (!q_chrom || !q_start || !q_end))) return gfail(g, GG_ERR_INVALID_ARG, "null argument");
    if (nq / L >= 0xFFFFFFF0ull) return gfail(g, GG_ERR_CAPACITY, "more than 2^32-16 queries per GPU and call");
    for (size_t i = 0; i < L; ++i) {
        Rk& r = g->rk[i];
        const size_t m = lo[i + 1] - lo[i], mm = m ? m : 1;
        GGG_CUDA(cudaSetDevice(r.dev));
        GGG_CUDA(r.in_c.ensure(mm * 4));
        GGG_CUDA(r.in_s.ensure(mm * cb));
        GGG_CUDA(r.in_e.ensure(mm * cb));
        if (m) {
            GGG_CUDA(cudaMemcpyAsync(r.in_c.p, q_chrom + lo[i], m * 4, cudaMemcpyHostToDevice, r.st));
            GGG_CUDA(cudaMemcpyAsync(r.in_s.p, static_cast<const char*>(q_start) + lo[i] * cb, m * cb, cudaMemcpyHostToDevice, r.st));
            GGG_CUDA(cudaMemcpyAsync(r.in_e.p, static_cast<const char*>(q_end) + lo[i] * cb, m * cb, cudaMemcpyHostToDevice, r.st));
        }
        r.q_c = r.in_c.as<uint32_t>();
        r.q_s = r.in_s.p;
        r.q_e = r.in_e.p;
        r.nq = m;
    }
    return GG_OK;
    }();
    if (pre != GG_OK && !g->rank_form) return pre;
    gg_status s = intersect_core(g, coord_bytes, true, pre);
    if (s != GG_OK) return s;
    collect_stats(g, true);
    const bool rep = g->shard_mode == GG_SHARD_REPLICATE;
    uint64_t total = 0;
    for (Rk& r : g->rk) total += r.n_hits_final;
    if (total + 1 > g->pinned_cap) {
        if (g->pinned_hits) cudaFreeHost(g->pinned_hits);
        g->pinned_hits = nullptr;
        g->pinned_cap = 0;
        const size_t want = static_cast<size_t>(total + total / 8) + 4096;
        GGG_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&g->pinned_hits), want * 4, cudaHostAllocDefault));
        g->pinned_cap = want;
    }
    // slices come back concatenated: offsets of slice i shifted by the hits of the slices before it (single-process form)
    const uint64_t first_base = rep ? g->rk[0].hroff[0] : g->rk[0].hsoff[0];
    for (size_t i = 0; i < L; ++i) {
        Rk& r = g->rk[i];
        GGG_CUDA(cudaSetDevice(r.dev));
        uint64_t* d_off = rep ? r.r_off.as<uint64_t>() : r.f_off.as<uint64_t>();
        const uint32_t* d_hits = rep ? r.r_hits.as<uint32_t>() : r.f_hits.as<uint32_t>();
        const uint64_t base = (rep ? r.hroff[0] : r.hsoff[0]) - first_base;  // relative to this process's first slice
        if (base) {  // (also for an empty slice: its single offset is the end of the slices before it)
            add_base_kernel<<<blocks_for(r.nq + 1, 256), 256, 0, r.st>>>(d_off, r.nq + 1, base);
            GGG_CUDA(cudaGetLastError());
        }
        const size_t n_off = i + 1 == L ? r.nq + 1 : r.nq;
        if (n_off) GGG_CUDA(cudaMemcpyAsync(offsets + lo[i], d_off, n_off * 8, cudaMemcpyDeviceToHost, r.st));
        if (r.n_hits_final) GGG_CUDA(cudaMemcpyAsync(g->pinned_hits + base, d_hits, r.n_hits_final * 4, cudaMemcpyDeviceToHost, r.st));
    }
    for (Rk& r : g->rk) {
        GGG_CUDA(cudaSetDevice(r.dev));
        GGG_CUDA(cudaStreamSynchronize(r.st));
    }
    if (nq == 0) offsets[0] = 0;
    *hits = g->pinned_hits;
    *n_hits = total;
    if (hit_base) *hit_base = first_base;
    return GG_OK;
}

}  // extern "C"
