// device_buf.h — owning device allocation used by the host orchestration (gg_api.cu, gg_group.cu).
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <cstddef>

namespace gg {

// debug facility (gg_set_option "inject_alloc_failure" = k): the k-th device allocation from now on fails as if the
// device were out of memory — the only way to drive the error paths on a box with 180 GB of HBM
extern std::atomic<long long> g_alloc_fail_countdown;

struct DeviceBuf {
    void* p = nullptr;
    size_t cap = 0;
    bool own = true;
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap && p) return cudaSuccess;
        release();
        size_t want = bytes ? bytes : 16;
        if (g_alloc_fail_countdown.load(std::memory_order_relaxed) > 0 && g_alloc_fail_countdown.fetch_sub(1) == 1)
            return cudaErrorMemoryAllocation;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    // non-owning view of memory the caller keeps alive (device-input entry points)
    void borrow(const void* ptr, size_t bytes) {
        release();
        p = const_cast<void*>(ptr);
        cap = bytes;
        own = false;
    }
    void release() {
        if (p && own) cudaFree(p);
        p = nullptr;
        cap = 0;
        own = true;
    }
    template <typename T>
    T* as() const {
        return static_cast<T*>(p);
    }
};

}  // namespace gg
