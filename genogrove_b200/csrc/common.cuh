// common.cuh — shared device/host helpers for libgg_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace gg {

constexpr int kFanout = 32;      // keys per node of the implicit B+ tree (one 128-byte line of u32)
constexpr int kFanoutLog2 = 5;
constexpr int kMaxLevels = 8;    // 32^7 > 2^32 ranks: 7 levels suffice; 8 leaves slack

template <typename T>
struct coord_max;
template <>
struct coord_max<uint32_t> {
    static constexpr uint32_t value = 0xFFFFFFFFu;
};
template <>
struct coord_max<uint64_t> {
    static constexpr uint64_t value = 0xFFFFFFFFFFFFFFFFull;
};

__host__ __device__ inline uint64_t div_up(uint64_t a, uint64_t b) { return (a + b - 1) / b; }
__host__ __device__ inline uint64_t round_up(uint64_t a, uint64_t b) { return div_up(a, b) * b; }

// ---------------------------------------------------------------- mbarrier + TMA bulk copy (1-D)
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "LAB_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra DONE;\n"
        "bra LAB_WAIT;\n"
        "DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// Same wait for long waits (another warp's look-back, a ticket): the try_wait carries a suspend-time hint so the
// warp sleeps in hardware instead of re-issuing the poll, leaving the issue slots to the warps it waits for.
__device__ __forceinline__ void mbar_wait_long(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "LAB_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1, %2;\n"
        "@P1 bra DONE;\n"
        "bra LAB_WAIT;\n"
        "DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity), "r"(0x989680u)
        : "memory");
}
// TMA bulk copy global -> shared (SASS: UBLKCP).  dst/src 16-byte aligned, bytes a multiple of 16.
__device__ __forceinline__ void tma_bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// TMA bulk copy shared -> global (SASS: UBLKCP again, the store direction), tracked by the thread's bulk async-group.
// Both addresses 16-byte aligned, bytes a multiple of 16.  The shared source was written through the generic proxy:
// every writing thread runs fence_proxy_async_smem() and the warp / CTA synchronises before one thread issues the copy.
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tma_bulk_s2g(void* gmem_dst, uint32_t smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(smem_src), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// all of this thread's bulk stores have finished READING shared memory (the stage may be overwritten)
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// explicit shared-space accesses by 32-bit shared address (keeps address math out of the generic space)
__device__ __forceinline__ void sts32(uint32_t saddr, uint32_t v) {
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(saddr), "r"(v) : "memory");
}
__device__ __forceinline__ void sts16(uint32_t saddr, uint32_t v) {
    asm volatile("st.shared.u16 [%0], %1;" ::"r"(saddr), "h"(static_cast<uint16_t>(v)) : "memory");
}
__device__ __forceinline__ void sts8(uint32_t saddr, uint32_t v) {
    asm volatile("st.shared.u8 [%0], %1;" ::"r"(saddr), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t lds16(uint32_t saddr) {
    uint16_t v;
    asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(saddr) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t lds8(uint32_t saddr) {
    uint32_t v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(saddr) : "memory");
    return v;
}
__device__ __forceinline__ uint2 lds64(uint32_t saddr) {
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(saddr) : "memory");
    return v;
}
__device__ __forceinline__ uint4 lds128(uint32_t saddr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(saddr) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t lds32(uint32_t saddr) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(saddr) : "memory");
    return v;
}

// ---------------------------------------------------------------- loads
template <typename T>
__device__ __forceinline__ T ld_volatile(const T* p) {
    return *reinterpret_cast<const volatile T*>(p);
}

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }

}  // namespace gg
