// launch_util.hpp -- per-device opt-in to more than 48 KB of dynamic shared memory.
// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is a per-device property of a kernel: a process that holds contexts on
// several GPUs (one batch_phy_t per GPU) must opt in on each of them.  One optin_t per kernel (or per
// group of kernels that are always configured together); thread safe.
#pragma once
#include <cuda_runtime.h>
#include <cstddef>
#include <mutex>

namespace dnrp {

struct smem_optin_t {
    static constexpr int kMaxDevices = 64;
    size_t configured[kMaxDevices] = {};
    std::mutex m;

    // calls set(bytes) when the current device has not yet been configured for at least `bytes`
    template <typename F>
    cudaError_t ensure(size_t bytes, F&& set) {
        int dev = 0;
        cudaError_t e = cudaGetDevice(&dev);
        if (e != cudaSuccess) return e;
        if (dev < 0 || dev >= kMaxDevices) return set(bytes);   // no cache for exotic device ordinals
        std::lock_guard<std::mutex> lk(m);
        if (bytes <= configured[dev]) return cudaSuccess;
        e = set(bytes);
        if (e == cudaSuccess) configured[dev] = bytes;
        return e;
    }
};

}  // namespace dnrp
