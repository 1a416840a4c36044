// mc_launch.cuh -- per-process caches for the two runtime queries every launch used to repeat
// (cudaFuncSetAttribute for > 48 KB of dynamic shared memory, cudaOccupancyMaxActiveBlocksPerMultiprocessor).
// The reference's own call sizes (1,000 x 20 ... 10,000 x 252, app.py:1189-1195) are launch-latency bound:
// the device needs ~0.2 ms, so host microseconds per launch matter.
#pragma once
#include <cuda_runtime.h>

#include <map>
#include <mutex>
#include <tuple>

namespace mc {

struct LaunchCache {
    std::mutex mu;
    std::map<std::tuple<int, const void*>, size_t> smem_limit;                 // (device, kernel) -> bytes enabled
    std::map<std::tuple<int, const void*>, size_t> static_smem;                // (device, kernel) -> static bytes
    std::map<std::tuple<int, const void*, int, size_t>, int> blocks_per_sm;   // (device, kernel, block, smem)
    static LaunchCache& get() { static LaunchCache c; return c; }
};

// make `smem` bytes of dynamic shared memory launchable for `func` on the current device: the opt-in is needed
// as soon as dynamic + STATIC shared memory exceed 48 KB (the static size is asked of the runtime once)
template <typename K>
inline cudaError_t ensure_dynamic_smem(K func, size_t smem, size_t /*static_hint*/ = 0) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    LaunchCache& c = LaunchCache::get();
    std::lock_guard<std::mutex> lock(c.mu);
    const auto key = std::make_tuple(dev, reinterpret_cast<const void*>(func));
    auto st = c.static_smem.find(key);
    if (st == c.static_smem.end()) {
        cudaFuncAttributes attr{};
        e = cudaFuncGetAttributes(&attr, func);
        if (e != cudaSuccess) return e;
        st = c.static_smem.emplace(key, attr.sharedSizeBytes).first;
    }
    if (smem + st->second <= 48 * 1024) return cudaSuccess;
    size_t& have = c.smem_limit[key];
    if (have >= smem) return cudaSuccess;
    e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess) have = smem;
    return e;
}

template <typename K>
inline cudaError_t resident_blocks_per_sm(K func, int block, size_t smem, int* per_sm) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    LaunchCache& c = LaunchCache::get();
    const auto key = std::make_tuple(dev, reinterpret_cast<const void*>(func), block, smem);
    {
        std::lock_guard<std::mutex> lock(c.mu);
        auto it = c.blocks_per_sm.find(key);
        if (it != c.blocks_per_sm.end()) { *per_sm = it->second; return cudaSuccess; }
    }
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(per_sm, func, block, smem);
    if (e == cudaSuccess) {
        std::lock_guard<std::mutex> lock(c.mu);
        c.blocks_per_sm[key] = *per_sm;
    }
    return e;
}

}  // namespace mc
